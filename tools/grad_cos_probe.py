"""Per-parameter gradient cosine of the GPU step against the fp32 oracle for a few AR variants (debug aid).
Usage: python tools/grad_cos_probe.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import test_step_gpu as ts  # noqa: E402
from cpc2_b200 import train as T  # noqa: E402
from oracle import cpc_oracle as oc  # noqa: E402

base = dict(hiddenEncoder=64, hiddenGar=128, nPredicts=4, negativeSamplingExt=16, sizeWindow=5120, rnnMode="linear")
for over in (dict(arMode="GRU", nLevelsGRU=2), dict(arMode="GRU", nLevelsGRU=1), dict(arMode="LSTM", nLevelsGRU=2),
             dict(arMode="LSTM", nLevelsGRU=1), dict(arMode="RNN", nLevelsGRU=1)):
    args = T.defaultArgs(**{**base, **over})
    B = 8
    model, crit = ts.build(args)
    seq = 0.1 * torch.randn(B, 2, 1, args.sizeWindow, generator=torch.Generator().manual_seed(1234))
    out = ts.run_step(model, crit, seq.cuda())
    S = args.sizeWindow // 160
    W = S - args.nPredicts
    ext = out["ext"]
    t = torch.arange(W).view(1, 1, W).expand(B, args.negativeSamplingExt, W).reshape(-1)
    bi = ext // S
    si = torch.remainder(ext % S - t, S)
    si = torch.where(si == 0, torch.full_like(si, S), si)
    torch.set_num_threads(os.cpu_count())
    ref = oc.train_step(ts._state_for_oracle(model, crit), seq, bi, si, **ts._oracle_kwargs(args))
    print(over, "loss err", (out["loss"] - ref["loss"]).abs().max().item(),
          "c err", (out["c"] - ref["c"]).abs().max().item())
    for k, g in ref["grads"].items():
        cs = ts.cosine(out["grads"][k], g)
        flag = "  <--" if cs < 0.99 else ""
        print(f"   {k:40s} {cs:.4f}  |g| {g.norm().item():.3e}{flag}")
