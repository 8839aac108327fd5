"""One eager training step of the hot path after a warm-up step -- the program ncu / compute-sanitizer are pointed at.

    python tools/step_once.py [key=value ...] [--steps N] [--fwd-only]
    (same overrides as cpc2_b200.train.defaultArgs, e.g. arMode=transformer hiddenEncoder=512 nLevelsGRU=2 sizeWindow=40960)

ncu launch list :  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/l.csv python tools/step_once.py
ncu full capture:  ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -o gpurun_out/gemm python tools/step_once.py
The warm-up step runs under `cudaProfilerStop` semantics only in the sense that ncu's `--launch-skip` can skip it:
the script prints how many C-ABI launches one step issues (use it as --launch-skip)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cpc2_b200 import kernels as K  # noqa: E402
from cpc2_b200 import train as T  # noqa: E402

over, steps, fwd_only = {}, 1, False
argv = sys.argv[1:]
i = 0
while i < len(argv):
    a = argv[i]
    if a == "--steps":
        steps = int(argv[i + 1])
        i += 1
    elif a == "--fwd-only":
        fwd_only = True
    else:
        k, v = a.split("=")
        over[k] = int(v) if v.lstrip("-").isdigit() else (v == "True" if v in ("True", "False") else v)
    i += 1
args = T.defaultArgs(**over)
torch.manual_seed(0)
model, crit = T.buildModelAndCriterion(args)
model, crit = model.cuda().train(), crit.cuda().train()
B = args.batchSizeGPU
g = torch.Generator().manual_seed(1234)
seq = (0.1 * torch.randn(B, 2, 1, args.sizeWindow, generator=g)).cuda()
label = torch.zeros(B, dtype=torch.long, device="cuda")
n0 = K.LAUNCHES
for it in range(1 + steps):
    for p in list(model.parameters()) + list(crit.parameters()):
        p.grad = None
    with torch.set_grad_enabled(not fwd_only):
        losses, acc = T.forwardBackward(model, crit, seq, label, backward=not fwd_only)
    torch.cuda.synchronize()
    if it == 0:
        print(f"C-ABI launches per step: {K.LAUNCHES - n0}", file=sys.stderr)
print("loss", [round(float(x), 4) for x in losses.flatten()])
