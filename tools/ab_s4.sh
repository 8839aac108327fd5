#!/bin/bash
# session A/B at N = 1: CTA count of the throttled staged Adam
cd "$(dirname "$0")/.."
run() {
  env "$@" python bench.py --headline-only --steps 40 --warmup 5 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.readlines()[-1]);print('$*', round(d['ms_per_step'],4), round(d['e2e']['ms_per_step'],4), round(d['loss_mean'],4))"
}
for i in 1 2; do
  for c in 192 256 296 384 592 1184; do run CPC_ADAM_CTAS=$c; done
done
