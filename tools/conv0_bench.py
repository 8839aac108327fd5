"""Times cpc_conv0_cn_relu_fwd/bwd alone at the default shape (16 windows x 20480 samples, C = 256) with the L2
flushed between launches, against the HBM roofline (bytes: bf16 activation written / its gradient read + audio)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cpc2_b200 import kernels as K  # noqa: E402

N, W, C, halo = 16, 20480, int(sys.argv[1]) if len(sys.argv) > 1 else 256, 2
L0 = W // 5
x = 0.1 * torch.randn(N, W, device="cuda")
conv = torch.nn.Conv1d(1, C, 10, stride=5, padding=3).cuda()
g, b = torch.ones(C, device="cuda"), torch.zeros(C, device="cuda")
out = torch.empty(N, L0 + 2 * halo, C, device="cuda", dtype=torch.bfloat16)
mean, rstd = torch.empty(N * L0, device="cuda"), torch.empty(N * L0, device="cuda")
dout = torch.randn(N, L0 + 2 * halo, C, device="cuda").bfloat16()
dw, db, dg, dbe = (torch.zeros(C, 10, device="cuda"), torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda"),
                   torch.zeros(C, device="cuda"))
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
w, bias = conv.weight.detach(), conv.bias.detach()


def timeit(fn, iters=20):
    ts = []
    for _ in range(iters + 3):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts = sorted(ts[3:])
    return ts[len(ts) // 2]


peak = 6454.9
p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
if os.path.exists(p):
    peak = json.load(open(p))["hbm_gbs"]
nbytes = N * L0 * C * 2 + N * W * 4
tf = timeit(lambda: K.conv0_fwd(x, w, bias, g, b, out, mean, rstd, N, W, C, halo))
tb = timeit(lambda: K.conv0_bwd(x, w, bias, g, b, mean, rstd, dout, dw, db, dg, dbe, N, W, C, halo))
print(f"conv0 fwd {tf:7.1f} us  {nbytes / tf / 1e3:7.1f} GB/s  {nbytes / tf / 1e3 / peak:.3f} of measured HBM peak")
print(f"conv0 bwd {tb:7.1f} us  {nbytes / tb / 1e3:7.1f} GB/s  {nbytes / tb / 1e3 / peak:.3f} of measured HBM peak")
