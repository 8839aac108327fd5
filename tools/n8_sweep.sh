#!/bin/bash
# N-rank diagnosis: the same bench under NCCL settings / with the collective removed.  Usage: tools/n8_sweep.sh 8
N=${1:-8}
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port $((29600 + RANDOM % 300)) bench.py --gpus $N --steps 30 --warmup 5 2>/dev/null \
    | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$name', d['ms_per_step'], d['e2e']['ms_per_step'], d['value'])"
}
run base X=1
run no_allreduce CPC_DIAG_NO_ALLREDUCE=1
run ctas8 NCCL_MAX_CTAS=8
run ctas4 NCCL_MAX_CTAS=4
run early0 CPC_EARLY_FFN=0
