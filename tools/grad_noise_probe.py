"""Prints per-parameter gradient cosines (GPU bf16 path vs CPU oracle) for a config, with the fp32-gradient
switch on and off.  Usage: python tools/grad_noise_probe.py [key=value ...]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from cpc2_b200 import ops, train as T  # noqa: E402
from oracle import cpc_oracle as oc  # noqa: E402
import test_step_gpu as ts  # noqa: E402

over = {}
for kv in sys.argv[1:]:
    k, v = kv.split("=")
    over[k] = int(v) if v.lstrip("-").isdigit() else v
args = T.defaultArgs(**over)
B = 8
model, crit = ts.build(args)
seq = 0.1 * torch.randn(B, 2, 1, args.sizeWindow, generator=torch.Generator().manual_seed(1234))
def rounded_state(model, crit):
    """Oracle weights with every tensor the GPU path consumes in bf16 rounded to bf16 (same function, fp32 math)."""
    st = ts._state_for_oracle(model, crit)
    for k, v in st.items():
        is_gemm_w = v.dim() >= 2 and (os.environ.get("ROUND_KREL") or not k.endswith("Krelpos")) and "conv0" not in k and "batchNorm" not in k
        if is_gemm_w:
            st[k] = v.to(torch.bfloat16).float()
    return st


ref = None
for flag in (True, False):
    ops.GRAD_F32 = flag
    out = ts.run_step(model, crit, seq.cuda())
    if ref is None:
        S = args.sizeWindow // 160
        W = S - args.nPredicts
        ext = out["ext"]
        t = torch.arange(W).view(1, 1, W).expand(B, args.negativeSamplingExt, W).reshape(-1)
        torch.set_num_threads(os.cpu_count())
        ref = oc.train_step(ts._state_for_oracle(model, crit), seq, ext // S, torch.remainder(ext % S - t, S),
                            **ts._oracle_kwargs(args))
        ext0 = ext
    fixed = ts.run_step(model, crit, seq.cuda(), ext=ext0)
    cs = {k: ts.cosine(fixed["grads"][k], g) for k, g in ref["grads"].items() if g.abs().max() > 0}
    worst = sorted(cs.items(), key=lambda kv: kv[1])[:12]
    print(f"GRAD_F32={flag}: min {min(cs.values()):.4f} mean {sum(cs.values()) / len(cs):.4f} "
          f"below0.99: {sum(v < 0.99 for v in cs.values())}/{len(cs)}")
    for k, v in worst:
        print(f"   {v:.4f} {k}")
    print("   loss diff", (fixed["loss"] - ref["loss"]).abs().max().item(), flush=True)
refq = oc.train_step(rounded_state(model, crit), seq, ext0 // S, torch.remainder(ext0 % S - t, S), **ts._oracle_kwargs(args))
cs = {k: ts.cosine(fixed["grads"][k], g) for k, g in refq["grads"].items() if g.abs().max() > 0}
print(f"vs oracle with bf16-rounded GEMM weights: min {min(cs.values()):.4f} mean {sum(cs.values()) / len(cs):.4f}")
for k, v in sorted(cs.items(), key=lambda kv: kv[1])[:8]:
    print(f"   {v:.4f} {k}")
cs = {k: ts.cosine(ref["grads"][k], g) for k, g in refq["grads"].items() if g.abs().max() > 0}
print(f"oracle fp32 vs oracle rounded weights: min {min(cs.values()):.4f} mean {sum(cs.values()) / len(cs):.4f}")
for k, v in sorted(cs.items(), key=lambda kv: kv[1])[:8]:
    print(f"   {v:.4f} {k}")
