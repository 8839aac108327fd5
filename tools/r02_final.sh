#!/bin/bash
# Round-2 closing run on one B200 (gpurun): GPU tests, the bench line, its ncu launch list, one full capture of the GEMM
# launches of one eager step, per-launch event profile and the graph-replay timeline.  Bounded by per-step timeouts.
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
T0=$(date +%s); lap() { echo "[$(( $(date +%s) - T0 )) s] $*"; }
timeout 600 python -m pytest tests -m gpu -x -q > $O/final_pytest.log 2>&1; lap "pytest rc=$? $(tail -1 $O/final_pytest.log)"
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" > $O/final_smoke.log 2>&1; lap "smoke rc=$? $(tail -1 $O/final_smoke.log)"
timeout 420 python bench.py > $O/r02_bench_n1_final.json 2> $O/r02_bench_n1_final.err; lap "bench rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $O/r02_launches_final.csv \
    python bench.py --steps 2 --warmup 3 --headline-only > $O/ncu_launches.log 2>&1; lap "launch list rc=$?"
python tools/ncu_launches.py $O/r02_launches_final.csv "ncu --metrics gpu__time_duration.sum --clock-control none -c 900 python bench.py --steps 2 --warmup 3 --headline-only" > $O/r02_ncu_launches_final.txt 2>&1
timeout 120 python tools/step_profile.py > $O/r02_step_profile_final.txt 2>&1; lap "step_profile rc=$?"
timeout 120 python tools/timeline.py > $O/r02_timeline_final.txt 2>&1; lap "timeline rc=$?"
if [ "$1" != "nofull" ]; then
timeout 420 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -s 48 -c 48 -o $O/r02_gemm_full_final \
    python tools/step_once.py > $O/ncu_gemm.log 2>&1; lap "gemm capture rc=$?"
python tools/ncu_summary.py $O/r02_gemm_full_final.ncu-rep > $O/r02_gemm_full_final_summary.txt 2>&1
python tools/traffic_from_ncu.py $O/r02_gemm_full_final.ncu-rep > $O/r02_traffic_final.json 2> $O/traffic.err
mkdir -p /tmp/ncu_reports && mv $O/*.ncu-rep /tmp/ncu_reports/
fi
lap done; du -sh $O
