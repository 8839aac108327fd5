#!/bin/bash
# session A/B: (1) which tile form moves conv0.weight in the past/future split test, (2) capture-stream priority
cd "$(dirname "$0")/.."
O=gpurun_out
CPC_NORM_CLUSTER_BELOW=0 python -m pytest tests/test_step_gpu.py -k "past_future" -q > $O/s4_pf_nocl.log 2>&1; echo "no-cluster rc=$?"; tail -2 $O/s4_pf_nocl.log
python -m pytest tests/test_step_gpu.py -k "past_future" -q > $O/s4_pf_cl.log 2>&1; echo "cluster rc=$?"; tail -2 $O/s4_pf_cl.log
for i in 1 2; do
  for pr in 0 -1; do
    CPC_MAIN_PRIORITY=$pr python bench.py --headline-only --steps 40 --warmup 5 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.readlines()[-1]);print('prio $pr run $i', round(d['ms_per_step'],4), round(d['e2e']['ms_per_step'],4))"
  done
done
