#!/bin/bash
# session A/B at N = 2: CTA shape of the all-reduce kernel; timeline of the data-parallel step
cd "$(dirname "$0")/.."
run() {
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 500)) \
    bench.py --gpus 2 --headline-only --steps 40 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.readlines() if l.startswith('{')][-1]);print('$*', round(d['ms_per_step'],4), round(d['e2e']['ms_per_step'],4), d.get('replica_param_max_abs_diff'))"
}
run CPC_AR_THREADS=512 CPC_AR_BLOCKS=48
run CPC_AR_THREADS=128 CPC_AR_BLOCKS=48
run CPC_AR_THREADS=128 CPC_AR_BLOCKS=96
run CPC_AR_THREADS=256 CPC_AR_BLOCKS=64
run CPC_AR_THREADS=512 CPC_AR_BLOCKS=24
run CPC_AR_THREADS=512 CPC_AR_BLOCKS=48 CPC_DIAG_NO_ALLREDUCE=1
python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 tools/timeline_dp.py > gpurun_out/s4_timeline_dp_n2.txt 2>/dev/null
python tools/timeline_dp.py > gpurun_out/s4_timeline_dp_n1.txt 2>/dev/null
