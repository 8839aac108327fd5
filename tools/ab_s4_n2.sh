#!/bin/bash
# session A/B at N = 2: register footprint of the all-reduce CTAs (CPC_AR_THREADS) and the deferred early-bucket update
cd "$(dirname "$0")/.."
run() {
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 500)) \
    bench.py --gpus 2 --headline-only --steps 40 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.readlines() if l.startswith('{')][-1]);print('$*', round(d['ms_per_step'],4), round(d['e2e']['ms_per_step'],4), d.get('replica_param_max_abs_diff'), d['config'].get('grad_allreduce'))"
}
python -m pytest tests -m gpu -q -k "two_gpus" 2>&1 | tail -2
for i in 1 2; do
  run CPC_AR_THREADS=512 CPC_ADAM_DEFER=0
  run CPC_AR_THREADS=128 CPC_ADAM_DEFER=0
  run CPC_AR_THREADS=128 CPC_ADAM_DEFER=1
done
