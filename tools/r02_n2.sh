#!/bin/bash
# Round-2 check of the data-parallel path on 2 GPUs of one box: the two-GPU tests, the bench line as the driver launches
# it, and the rank-0 timeline of the data-parallel step.
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
T0=$(date +%s); lap() { echo "[$(( $(date +%s) - T0 )) s] $*"; }
timeout 300 python -m pytest tests -m gpu -q -k "two_gpus" > $O/n2_pytest.log 2>&1; lap "two-GPU tests rc=$? $(tail -1 $O/n2_pytest.log)"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus 2 --steps 20 --warmup 5 --headline-only > $O/r02_bench_n2_final.json 2> $O/r02_bench_n2_final.err; lap "bench n2 rc=$?"
grep -o '"ms_per_step": [0-9.]*\|"replica_param_max_abs_diff": [0-9.e-]*\|"value": [0-9.]*' $O/r02_bench_n2_final.json | head -6
timeout 200 python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 tools/timeline_dp.py > $O/r02_timeline_dp_n2_final.txt 2>/dev/null; lap "timeline rc=$?"
