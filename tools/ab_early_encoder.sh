#!/bin/bash
# same-box A/B of ops.EARLY_ENCODER (1 / 2 / 3) at N = 2 and N = 1: ms per step (device value, e2e), replica check
cd "$(dirname "$0")/.."
run() {  # run <n_gpus> ENV=...
  n=$1; shift
  if [ $n -gt 1 ]; then L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400))"; else L=python; fi
  env "$@" $L bench.py --gpus $n --headline-only --steps 40 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.readlines() if l.startswith('{')][-1]);print('N=$n $*', round(d['ms_per_step'],4), round(d['e2e']['ms_per_step'],4), d.get('replica_param_max_abs_diff'), round(d['loss_mean'],4))"
}
for rep in 1 2; do
  for m in 1 3 2; do run 2 CPC_EARLY_ENCODER=$m; done
done
for m in 1 3 2; do run 1 CPC_EARLY_ENCODER=$m; done
