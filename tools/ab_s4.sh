#!/bin/bash
# session A/B at N = 1: waveform read in place; CTA cap of the side-stream weight-gradient GEMMs
cd "$(dirname "$0")/.."
run() {
  env "$@" python bench.py --headline-only --steps 40 --warmup 5 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.readlines()[-1]);print('$*', round(d['ms_per_step'],4), round(d['e2e']['ms_per_step'],4), round(d['loss_mean'],4))"
}
for i in 1 2; do
  run CPC_INPLACE_X=1
  run CPC_INPLACE_X=0
  run CPC_WGRAD_CTAS=112
  run CPC_WGRAD_CTAS=96
  run CPC_WGRAD_CTAS=74
  run CPC_WGRAD_CTAS=48
done
