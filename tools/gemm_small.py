"""One small grouped GEMM of the heads' chain (g12 928x256x256, K-major operands, bf16 out + residual), launched a
few times: the case whose fixed per-launch cost dominates the step.  Usage: python tools/gemm_small.py [--once]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cpc2_b200 import ops  # noqa: E402

BF = torch.bfloat16
G, M, N, Kd = 12, 928, 256, 256
a = (torch.randn(G, M, Kd, device="cuda") * 0.1).to(BF)
b = (torch.randn(G, N, Kd, device="cuda") * 0.1).to(BF)
res = (torch.randn(G, M, N, device="cuda") * 0.1).to(BF)
o = torch.empty(G, M, N, device="cuda", dtype=BF)


def run():
    ops.linear_fwd(a, b, M, N, Kd, groups=G, out_bf16=o, res=res)


run()
torch.cuda.synchronize()
if "--once" in sys.argv:
    run()
    torch.cuda.synchronize()
else:
    for n in (1, 20):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            run()
        e1.record()
        torch.cuda.synchronize()
        print(f"{n} launches: {e0.elapsed_time(e1) * 1e3 / n:.1f} us each")
