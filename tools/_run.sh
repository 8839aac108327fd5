python -m pytest tests/test_step_gpu.py -m gpu -q -x -k "not oracle and not golden" > gpurun_out/t13.log 2>&1; echo rc=$?; tail -3 gpurun_out/t13.log
run() { tag=$1; shift; env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --headline-only 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$tag', round(d['ms_per_step'],4), round(d['e2e']['ms_per_step'],4), d['config']['grad_allreduce'][:20], d.get('replica_param_max_abs_diff'), d['gpu_launches_per_step'])"; }
run n2_early A=1
run n2_noearly CPC_EARLY_ENCODER=0
run n2_early A=1
run n2_noearly CPC_EARLY_ENCODER=0
CUDA_VISIBLE_DEVICES=0 python bench.py --headline-only 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('n1_early', d['ms_per_step'], d['e2e']['ms_per_step'])"
CPC_EARLY_ENCODER=0 CUDA_VISIBLE_DEVICES=0 python bench.py --headline-only 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('n1_noearly', d['ms_per_step'], d['e2e']['ms_per_step'])"
