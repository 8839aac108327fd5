"""Times cpc_gemm_bf16 on the GEMM shapes of the default CPC step for each tile width (auto / 64 / 128 / 256).
Usage: python tools/gemm_bench.py   (warm: mean of 20 back-to-back launches; --cold: L2 flushed before each launch, median of 7)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cpc2_b200 import kernels as K  # noqa: E402
from cpc2_b200 import ops  # noqa: E402

dev = "cuda"
BF = torch.bfloat16
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


COLD = "--cold" in sys.argv


def timeit(fn, n=7):
    if not COLD:  # warm: 20 back-to-back launches
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / 20
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def rnd(*s):
    return (torch.randn(*s, device=dev) * 0.1).to(BF)


def nt(G, M, N, Kd, shared=False, out="bf16", **epi):
    """forward-style: A [G,M,K] (or shared [M,K]) x B [G,N,K]."""
    a = rnd(M, Kd) if shared else rnd(G, M, Kd)
    b = rnd(G, N, Kd)
    o = torch.empty(G, M, N, device=dev, dtype=BF if out == "bf16" else torch.float32)

    def run(bn):
        kw = dict(out_bf16=o) if out == "bf16" else dict(out_f32=o)
        ops.linear_fwd(a, b, M, N, Kd, groups=G, x_shared=shared, force_bn=bn, **kw, **epi)
    return run


def dx(G, M, N, Kd, **epi):
    """dx[g] = dy[g] [M,N] . w[g] [N,K]  (B MN-major)."""
    dy, w = rnd(G, M, N), rnd(G, N, Kd)
    o = torch.empty(G, M, Kd, device=dev, dtype=BF)
    return lambda bn: ops.linear_dx(dy, w, M, N, Kd, groups=G, out_bf16=o, force_bn=bn, **epi)


def dw(G, M, N, Kd):
    """dw[g] [N,K] += dy[g]^T . x[g]  (both MN-major, split-K)."""
    dy, x = rnd(G, M, N), rnd(G, M, Kd)
    o = torch.zeros(G, N, Kd, device=dev)

    def run(bn):
        A, B = K.mnmajor(dy), K.mnmajor(x)
        K.gemm(A, B, N, Kd, M, groups=G, k_splits=ops.auto_splits(N, Kd, M, G), out_f32=o, accumulate=True, force_bn=bn)
    return run


def conv_fwd(N, Lout, C, k, s):
    pad = {8: 2, 4: 1}[k]
    Lp = Lout * s + 2 * pad
    x, wf = rnd(N, Lp, C), rnd(C, k * C)
    bias = torch.randn(C, device=dev)
    y = torch.empty(N * Lout, C, device=dev, dtype=BF)

    def run(bn):
        A = K.View(x, k * C, Lout, N, s * C, Lp * C, major=0, gmul=1)
        K.gemm(A, K.kmajor(wf), Lout, C, k * C, groups=N, out_bf16=(y, C, Lout * C), bias=bias, force_bn=bn)
    return run


def conv_wgrad(N, Lout, C, k, s):
    pad = {8: 2, 4: 1}[k]
    Lp = Lout * s + 2 * pad
    x, dy = rnd(N, Lp, C), rnd(N, Lout + 2, C)
    dwf = torch.zeros(C, k * C, device=dev)

    def run(bn):
        A = K.View(dy, C, Lout + 2, N, C, (Lout + 2) * C, major=1, row_off=1)
        B = K.View(x, k * C, Lout, N, s * C, Lp * C, major=1)
        K.gemm(A, B, C, k * C, N * Lout, k_per_slab=Lout, k_splits=ops.auto_splits(C, k * C, N * Lout, 1, Lout),
               out_f32=dwf, accumulate=True, force_bn=bn)
    return run


def conv_dgrad(N, Lout, C, k, s):
    pad = {8: 2, 4: 1}[k]
    Lp = Lout * s + 2 * pad
    dy, wd = rnd(N, Lout + 2, C), rnd(s * C, 2 * C)
    dxp = torch.empty(N, Lp, C, device=dev, dtype=BF)

    def run(bn):
        A = K.View(dy, 2 * C, Lout + 1, N, C, (Lout + 2) * C, major=0, gmul=1)
        K.gemm(A, K.kmajor(wd), Lout + 1, s * C, 2 * C, groups=N, out_bf16=(dxp, s * C, Lp * C), force_bn=bn)
    return run


if __name__ == "__main__":
    T, d, dff, G = 928, 256, 2048, 12
    rng = torch.tensor([1, 0], dtype=torch.int64, device=dev)
    bias_ff = torch.randn(G, dff, device=dev)
    h1 = rnd(G, T, dff)
    cases = [
        ("conv1 fwd 16x1024 K2048", conv_fwd(16, 1024, 256, 8, 4), 2.0 * 16 * 1024 * 256 * 2048),
        ("conv2 fwd 16x512 K1024", conv_fwd(16, 512, 256, 4, 2), 2.0 * 16 * 512 * 256 * 1024),
        ("conv4 fwd 16x128 K1024", conv_fwd(16, 128, 256, 4, 2), 2.0 * 16 * 128 * 256 * 1024),
        ("conv1 wgrad", conv_wgrad(16, 1024, 256, 8, 4), 2.0 * 16 * 1024 * 256 * 2048),
        ("conv1 dgrad", conv_dgrad(16, 1024, 256, 8, 4), 2.0 * 16 * 1025 * 1024 * 512),
        ("conv4 wgrad", conv_wgrad(16, 128, 256, 4, 2), 2.0 * 16 * 128 * 256 * 1024),
        ("lstm proj 2048x1024x256", nt(1, 2048, 1024, 256, out="f32"), 2.0 * 2048 * 1024 * 256),
        ("heads qkv g12 928x768x256", nt(G, T, 768, d, shared=True), 2.0 * G * T * 768 * d),
        ("heads Wo g12 928x256x256", nt(G, T, d, d), 2.0 * G * T * d * d),
        ("heads ff1 g12 928x2048x256 +bias relu drop", nt(G, T, dff, d, bias=bias_ff, bias_gs=dff, relu=True, rng=rng,
                                                         rng_site=2, drop_p=0.1), 2.0 * G * T * dff * d),
        ("heads ff1 g12 928x2048x256 plain", nt(G, T, dff, d), 2.0 * G * T * dff * d),
        ("heads ff2 g12 928x256x2048", nt(G, T, d, dff), 2.0 * G * T * dff * d),
        ("logits g12 928x1024x256 f32", nt(G, T, 1024, d, out="f32"), 2.0 * G * T * 1024 * d),
        ("heads dx ff2 g12 928x2048<-256 gate", dx(G, T, d, dff, gate=h1, gate_scale=1.1), 2.0 * G * T * dff * d),
        ("heads dx ff1 g12 928x256<-2048", dx(G, T, dff, d), 2.0 * G * T * dff * d),
        ("heads dw ff1 g12 2048x256 K928", dw(G, T, dff, d), 2.0 * G * T * dff * d),
        ("heads dw ff2 g12 256x2048 K928", dw(G, T, d, dff), 2.0 * G * T * dff * d),
        ("heads dw Wo g12 256x256 K928", dw(G, T, d, d), 2.0 * G * T * d * d),
    ]
    only = [a.split("=", 1)[1] for a in sys.argv[1:] if a.startswith("--only=")]
    if only:
        cases = [c for c in cases if any(o in c[0] for o in only)]
    if "--once" in sys.argv:  # for ncu: two launches of each selected case with the automatic tile width
        for name, run, flops in cases:
            run(0)
            run(0)
            torch.cuda.synchronize()
            print("ran", name)
        sys.exit(0)
    print(torch.cuda.get_device_name(0))
    print(f"{'case':46s} {'auto':>9s} {'bn64':>9s} {'bn128':>9s} {'bn256':>9s}   (us; TF/s for auto)")
    for name, run, flops in cases:
        row = []
        for bn in (0, 64, 128, 256):
            try:
                run(bn)
                torch.cuda.synchronize()
                row.append(timeit(lambda: run(bn)))
            except Exception as e:  # noqa: BLE001
                row.append(float("nan"))
        print(f"{name:46s} " + " ".join(f"{t:9.1f}" for t in row) + f"   {flops / row[0] / 1e6:7.1f} TF/s", flush=True)
