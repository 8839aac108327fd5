"""Times cpc_attn_fwd / cpc_attn_bwd at the shapes of the CPC step (default heads: G=12, Bt=8, Hh=8, S=116, dk=32;
transformer AR: G=1, Bt=16, S=128).  Usage: python tools/attn_bench.py [--once]   (--once: single launches, for ncu)"""
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cpc2_b200 import kernels as K  # noqa: E402

BF = torch.bfloat16
dev = "cuda"


def case(G, Bt, Hh, S, dk, drop, fast_d=True):
    d = Hh * dk
    qkv = (torch.randn(G, Bt * S, 3 * d, device=dev) * 0.5).to(BF)
    krel = (torch.rand(G, dk, S, device=dev) - 0.5) * (2 / math.sqrt(dk))
    out = torch.empty(G, Bt * S, d, device=dev, dtype=BF)
    probs = torch.zeros(G, Bt, Hh, S, S, device=dev, dtype=BF)
    dout = (torch.randn(G, Bt * S, d, device=dev) * 0.5).to(BF)
    dqkv = torch.empty_like(qkv)
    dscore = torch.empty_like(probs)
    dkrel = torch.zeros(G, dk, S, device=dev)
    rng = torch.tensor([5, 1], device=dev, dtype=torch.int64) if drop else None
    out32 = torch.empty(G, Bt * S, d, device=dev)
    rowsum = torch.empty(G, Bt, Hh, S, device=dev)
    extra = dict(out_f32=out32, rowsum=rowsum) if fast_d else {}

    def fwd():
        K.attn_fwd(qkv, krel, out, probs, G, Bt, Hh, S, dk, rng=rng, site=1, drop_p=drop, **extra)

    def bwd():
        K.attn_bwd(qkv, krel, out, probs, dout, dqkv, dscore, dkrel, G, Bt, Hh, S, dk, rng=rng, site=1, drop_p=drop, **extra)
    return fwd, bwd


def timeit(fn, n=20):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


if __name__ == "__main__":
    once = "--once" in sys.argv
    for name, cfg in [("heads g12 b8 h8 S116 dk32 drop0.1", (12, 8, 8, 116, 32, 0.1)),
                      ("heads g12 b8 h8 S116 dk32 eval", (12, 8, 8, 116, 32, 0.0)),
                      ("AR g1 b16 h8 S128 dk32", (1, 16, 8, 128, 32, 0.1)),
                      ("big heads g12 b4 h8 S244 dk64", (12, 4, 8, 244, 64, 0.1))]:
        if not once:  # D from the forward's fp32 O / rowsum (a) against the in-kernel dP sweep (b), CpcAttn
            f2, b2 = case(*cfg, fast_d=False)
            print(f"{name + ' [D in-kernel]':40s} fwd {timeit(f2):8.1f} us   bwd {timeit(b2):8.1f} us", flush=True)
        fwd, bwd = case(*cfg)
        if once:
            fwd()
            bwd()
            torch.cuda.synchronize()
            print("ran", name)
            continue
        print(f"{name:40s} fwd {timeit(fwd):8.1f} us   bwd {timeit(bwd):8.1f} us", flush=True)
