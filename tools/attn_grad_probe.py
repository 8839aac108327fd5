"""Which rounding site costs the q / k gradients of the transformer heads their digits, away from init?
Runs the default config through the test's warm-up, then compares the CUDA path's gradients and several bf16-emulation
variants of the CPU oracle against the exact fp32 oracle.  Usage: python tools/attn_grad_probe.py [key=value ...]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_step_gpu as ts  # noqa: E402
from cpc2_b200 import train as T  # noqa: E402
from oracle import cpc_oracle as oc  # noqa: E402

over = {}
for kv in sys.argv[1:]:
    k, v = kv.split("=")
    over[k] = int(v) if v.lstrip("-").isdigit() else v
args = T.defaultArgs(**over)
B = 8 if args.sizeWindow <= 20480 else 4
model, crit = ts.build(args)
warm = ts.warm_up(model, crit, args, B)
print("warm-up", warm, flush=True)
seq = ts.predictable_sequence(B, args.sizeWindow, 1234)
out = ts.run_step(model, crit, seq.cuda())
S = args.sizeWindow // 160
W = S - args.nPredicts
ext = out["ext"]
t = torch.arange(W).view(1, 1, W).expand(B, args.negativeSamplingExt, W).reshape(-1)
bi = ext // S
si = torch.remainder(ext % S - t, S)
si = torch.where(si == 0, torch.full_like(si, S), si)
torch.set_num_threads(os.cpu_count())
st = ts._state_for_oracle(model, crit)
kw = ts._oracle_kwargs(args)
exact = oc.train_step(st, seq, bi, si, **kw)


def variant(emulate, extra):
    oc.EMULATE_BF16, oc.EMULATE_EXTRA = emulate, set(extra)
    try:
        return oc.train_step(st, seq, bi, si, **kw)["grads"]
    finally:
        oc.EMULATE_BF16, oc.EMULATE_EXTRA = False, set()


runs = {"gpu": out["grads"], "emu": variant(True, ()), "emu+ds": variant(True, ("ds",)), "emu+p": variant(True, ("p",)),
        "ds only": variant(False, ("ds",)), "p only": variant(False, ("p",))}
keys = [k for k in exact["grads"] if exact["grads"][k].abs().max() > 0]
bad = sorted(keys, key=lambda k: ts.cosine(runs["gpu"][k], exact["grads"][k]))[:14]
print(f"{'tensor':58s}" + "".join(f"{n:>9s}" for n in runs))
for k in bad:
    print(f"{k:58s}" + "".join(f"{ts.cosine(g[k], exact['grads'][k]):9.4f}" for g in runs.values()))
# common-mode content of the keys of the worst head
k0 = bad[0]
if "multihead" in k0:
    pre = k0[:k0.index("multihead")]
    c = out["c"][:B].float()
    for nm in ("Wq", "Wk", "Wv"):
        y = c @ st[pre + f"multihead.{nm}.weight"].t()
        print(nm, "common-mode norm / residual rms norm over time:",
              float(y.mean(1).norm(dim=-1).mean() / (y - y.mean(1, keepdim=True)).norm(dim=-1).mean()))
    print("|c mean over t| / |c - mean|:", float(c.mean(1).norm(dim=-1).mean() / (c - c.mean(1, keepdim=True)).norm(dim=-1).mean()))
