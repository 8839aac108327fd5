"""Times cpc_colnorm_fwd / _bwd and cpc_conv0_plain_fwd / _bwd alone at the default shape of encoder layer 0 (16 windows
x 4096 frames x 256 channels) with the L2 flushed between launches, against the measured HBM peak.  Algorithmic bytes:
forward = activation read twice (statistics, apply) + written once; backward = activation and gradient read twice +
gradient written once.   python tools/colnorm_bench.py [batchNorm|instanceNorm|ID] [--once]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cpc2_b200 import kernels as K  # noqa: E402

mode = next((a for a in sys.argv[1:] if not a.startswith("-")), "batchNorm")
once = "--once" in sys.argv
N, W, C = 16, 20480, 256
L = W // 5
rows = N * L
kind = "ID" if mode == "ID" else "stats"
rpg = L if mode == "instanceNorm" else 0
groups = N if rpg else 1
x = 0.1 * torch.randn(N, W, device="cuda")
w, b = torch.randn(C, 1, 10, device="cuda") * 0.3, torch.randn(C, device="cuda") * 0.1
y = torch.empty(rows, C, device="cuda", dtype=torch.bfloat16)
g, be = (None, None) if mode == "ID" else (torch.ones(C, device="cuda"), torch.zeros(C, device="cuda"))
mean = rstd = None
if kind == "stats":
    mean, rstd = torch.empty(groups * C, device="cuda"), torch.empty(groups * C, device="cuda")
partial, sums = K.colnorm_scratch(rows, C, rpg, "cuda")
out = torch.empty(N, L + 4, C, device="cuda", dtype=torch.bfloat16)
dout = torch.randn(N, L + 4, C, device="cuda").bfloat16()
dx32 = torch.empty(rows, C, device="cuda")
dg, db, ds, dw = (torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda"),
                  torch.zeros(C, 10, device="cuda"))
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def fwd():
    K.colnorm_fwd(y, rows, C, kind, rows_per_group=rpg, gamma=g, beta=be, mean=mean, rstd=rstd, partial=partial,
                  out_bf16=out, out_map=K.rowmap(L, L + 4, 2))


def bwd():
    K.colnorm_bwd(dout, y, rows, C, kind, partial, sums, rows_per_group=rpg, gamma=g, beta=be, mean=mean, rstd=rstd,
                  dout_map=K.rowmap(L, L + 4, 2), dx_f32=dx32, dgamma=None if mode == "ID" else dg,
                  dbeta=None if mode == "ID" else db, dxsum=ds)


def timeit(fn, iters=10):
    ts = []
    for _ in range(iters + 2):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return sorted(ts[2:])[len(ts[2:]) // 2]


K.conv0_plain_fwd(x, w, b, y, N, W, C)
if once:  # ncu target: one launch of each
    fwd(); bwd(); K.conv0_plain_bwd(x, dx32, dw, N, W, C)
    torch.cuda.synchronize()
    sys.exit(0)
peak = 6500.0
p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
if os.path.exists(p):
    d = json.load(open(p))
    peak = float(d.get("hbm_gbs_burst", d.get("hbm_gbs", peak)))
act = rows * C * 2
cases = (("conv0 plain fwd", lambda: K.conv0_plain_fwd(x, w, b, y, N, W, C), act + 4 * N * W),
         (f"colnorm {mode} fwd", fwd, (3 if kind == "stats" else 2) * act),
         (f"colnorm {mode} bwd", bwd, 2 * 2 * act + rows * C * 4),
         ("conv0 plain bwd", lambda: K.conv0_plain_bwd(x, dx32, dw, N, W, C), rows * C * 4 + 4 * N * W))
for name, fn, nbytes in cases:
    us = timeit(fn)
    print(f"{name:26s} {us:8.1f} us  {nbytes / us / 1e3:8.1f} GB/s  {nbytes / us / 1e3 / peak:.3f} of measured HBM peak ({peak:.0f} GB/s)")
