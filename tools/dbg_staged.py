"""Debug: staged vs single-launch FusedAdam (tests/test_step_gpu.py::test_staged_adam_equals_single_launch), per-parameter
differences.  Usage: [CPC_DEFER_JOINS=0] python tools/dbg_staged.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cpc2_b200 import ops  # noqa: E402
from cpc2_b200 import train as T  # noqa: E402
from cpc2_b200.optim import FusedAdam  # noqa: E402

args = T.defaultArgs(hiddenEncoder=64, hiddenGar=64, nPredicts=4, negativeSamplingExt=16, sizeWindow=5120)
g = torch.Generator().manual_seed(3)
seq = (0.1 * torch.randn(4, 2, 1, args.sizeWindow, generator=g)).cuda()
label = torch.zeros(4, dtype=torch.long, device="cuda")
finals, names = [], None
for staged in (False, False, False, True, "early"):
    torch.manual_seed(0)
    model, crit = T.buildModelAndCriterion(args)
    model, crit = model.cuda().eval(), crit.cuda().eval()
    named = list(crit.named_parameters()) + list(model.named_parameters())
    names = [n for n, _ in named]
    opt = FusedAdam([p for _, p in named], lr=1e-3)
    if staged:
        opt.attach()
    ops.EARLY_FFN = "1" if staged == "early" else "0"
    grads = []
    for it in range(3):
        torch.manual_seed(50 + it)
        T.forwardBackward(model, crit, seq, label)
        torch.cuda.synchronize()
        grads.append([p.grad.detach().clone() if p.grad is not None else None for _, p in named])
        opt.step()
        opt.zero_grad()
        torch.cuda.synchronize()
        grads.append([p.detach().clone() for _, p in named])
        ar = crit.wPrediction.engine().arena
        sh = getattr(ar, "_shadow", None)
        if sh is not None:
            stale = (sh.float() - ar.flat.to(torch.bfloat16).float()).abs().max().item()
            print(f"  [staged={staged} it={it}] shadow_valid={ar.shadow_valid()} |shadow - bf16(flat)|max={stale:.3e} "
                  f"gflat ptr={named[0][1].grad.data_ptr() if named[0][1].grad is not None else None} tables={ {k: (v['n_spans'], v['blocks']) for k, v in opt._groups[0]['tables'].items()} }")
    opt.detach()
    torch.cuda.synchronize()
    finals.append(([p.detach().clone() for _, p in named], grads))
for vi, (fin, grads) in enumerate(finals[1:], 1):
    print(f"variant {vi}")
    for k, (ga, gb) in enumerate(zip(finals[0][1], grads)):
        what = "grads" if k % 2 == 0 else "params"
        worst = sorted(((x - y).abs().max().item() / (x.abs().max().item() + 1e-30), n) for n, x, y in zip(names, ga, gb)
                       if x is not None)[-3:]
        print(f"  iter {k // 2} {what}: worst rel-to-max diffs {worst}")
nbad = 0
for vi, (fin, grads) in enumerate(finals[1:], 1):
    ga, gb = finals[0][1][4], grads[4]
    w = max((x - y).abs().max().item() / (x.abs().max().item() + 1e-30) for x, y in zip(ga, gb) if x is not None)
    nbad += w > 1e-4
print(f"SUMMARY defer={os.environ.get('CPC_DEFER_JOINS', '1')} variants whose iter-2 gradients left the 1e-4 band: {nbad} of {len(finals) - 1}")
for vi, (fin, grads) in enumerate(finals[1:], 1):
    fr, mx = 0.0, 0.0
    for a, b in zip(finals[0][0], fin):
        fr = max(fr, ((a - b).abs() > 1e-6 + 1e-4 * b.abs()).float().mean().item())
        mx = max(mx, (a - b).abs().max().item())
    print(f"STATS variant {vi}: worst fraction of entries outside the tight band {fr:.4f}, max abs diff {mx:.2e}")
