"""Per-launch CUDA-event timing of one eager training step (default config), printed call by call.
Usage: python tools/step_profile.py [key=value ...]   (same overrides as cpc2_b200.train.defaultArgs)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from cpc2_b200 import train as T  # noqa: E402

over = {}
for kv in sys.argv[1:]:
    k, v = kv.split("=")
    over[k] = int(v) if v.lstrip("-").isdigit() else v
args = T.defaultArgs(**over)
torch.manual_seed(0)
model, crit = T.buildModelAndCriterion(args)
model, crit = model.cuda().train(), crit.cuda().train()
B = args.batchSizeGPU
seq = bench.synthetic_batch(B, args.sizeWindow, 0).cuda()
label = torch.zeros(B, dtype=torch.long, device="cuda")
rows, eager_ms, prof = bench.kernel_profile(lambda: T.forwardBackward(model, crit, seq, label),
                                            list(model.parameters()) + list(crit.parameters()))
tot = 0.0
for name, work, e0, e1 in prof:
    us = 1e3 * e0.elapsed_time(e1)
    tot += us
    extra = ""
    if work.get("flops"):
        extra += f" {work['flops'] / us / 1e6:8.1f} TF/s"
    if work.get("bytes"):
        extra += f" {work['bytes'] / us / 1e3:8.1f} GB/s"
    print(f"{us:9.1f} us  {name:24s} {work.get('shape', '')}{extra}")
print(f"sum {tot:.1f} us over {len(prof)} launches; eager wall (GPU) {eager_ms * 1e3:.1f} us")
for r in rows:
    print(f"{r['name']:24s} n={r['launches']:3d} {1e3 * r['ms']:9.1f} us")
