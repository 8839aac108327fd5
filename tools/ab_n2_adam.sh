#!/bin/bash
# same-box A/B at N = 2: throttle / deferral of the staged Adam passes and the early FFN announcement
cd "$(dirname "$0")/.."
run() {
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
    bench.py --gpus 2 --headline-only --steps 40 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.readlines() if l.startswith('{')][-1]);print('N=2 $*', round(d['ms_per_step'],4), round(d['e2e']['ms_per_step'],4), d.get('replica_param_max_abs_diff'), round(d['loss_mean'],4))"
}
run CPC_X=baseline
run CPC_ADAM_CTAS=128
run CPC_ADAM_CTAS=512
run CPC_ADAM_CTAS=0
run CPC_ADAM_DEFER=2
run CPC_ADAM_DEFER=0
run CPC_EARLY_FFN=0
run CPC_X=baseline
