"""GPU probe for cpc_gemm_bf16: runs every operand-layout / epilogue variant against torch and prints a
table (does not stop at the first failure).  Usage: python tools/gemm_probe.py"""
import os
import sys
import traceback

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cpc2_b200 import kernels as K  # noqa: E402

dev = "cuda"
torch.manual_seed(0)
results = []


def rnd(*shape):
    return (torch.randn(*shape, device=dev) * 0.5).to(torch.bfloat16)


def report(name, got, ref, tol=2e-2):
    got, ref = got.float(), ref.float()
    err = (got - ref).abs().max().item()
    scale = ref.abs().max().item() + 1e-6
    ok = err <= tol * scale and torch.isfinite(got).all().item()
    results.append((name, ok, err, scale))
    print(f"{'PASS' if ok else 'FAIL'} {name}: max_abs_err={err:.4g} ref_max={scale:.4g}", flush=True)


def case(name, fn):
    try:
        fn()
        torch.cuda.synchronize()
    except Exception as e:  # noqa: BLE001
        results.append((name, False, float("nan"), 0))
        print(f"FAIL {name}: EXC {e}", flush=True)
        traceback.print_exc()


def nt(M, N, Kd, bn=0):
    a, b = rnd(M, Kd), rnd(N, Kd)
    out = torch.empty(M, N, device=dev)
    K.gemm(K.kmajor(a), K.kmajor(b), M, N, Kd, out_f32=out, force_bn=bn)
    report(f"nt M{M} N{N} K{Kd} bn{bn}", out, a.float() @ b.float().t())


def grouped_shared_a(G=3, M=200, N=256, Kd=128):
    a, b = rnd(M, Kd), rnd(G, N, Kd)
    out = torch.empty(G, M, N, device=dev)
    K.gemm(K.kmajor(a), K.kmajor(b), M, N, Kd, groups=G, out_f32=out)
    report("grouped shared A", out, torch.einsum("mk,gnk->gmn", a.float(), b.float()))


def a_mn(M=192, N=128, Kd=200, bn=0):
    at, b = rnd(Kd, M), rnd(N, Kd)  # A stored transposed: [K, M]
    out = torch.empty(M, N, device=dev)
    K.gemm(K.mnmajor(at), K.kmajor(b), M, N, Kd, out_f32=out, force_bn=bn)
    report(f"A MN-major bn{bn}", out, at.float().t() @ b.float().t())


def b_mn(M=160, N=256, Kd=136, bn=0):
    a, bt = rnd(M, Kd), rnd(Kd, N)  # B stored [K, N]
    out = torch.empty(M, N, device=dev)
    K.gemm(K.kmajor(a), K.mnmajor(bt), M, N, Kd, out_f32=out, force_bn=bn)
    report(f"B MN-major bn{bn}", out, a.float() @ bt.float())


def both_mn_slabs_splitk(S=4, R=100, M=256, N=320, splits=3):
    # reduction over (slab, row): A^T stored [S, R, M], B^T stored [S, R, N]
    at, bt = rnd(S, R, M), rnd(S, R, N)
    out = torch.full((M, N), 1.0, device=dev)
    A = K.View(at, M, R, S, M, R * M, major=1, gmul=0)
    B = K.View(bt, N, R, S, N, R * N, major=1, gmul=0)
    K.gemm(A, B, M, N, S * R, k_per_slab=R, k_splits=splits, out_f32=out, accumulate=True)
    report(f"both MN + slabs + splitK{splits}", out, 1.0 + torch.einsum("srm,srn->mn", at.float(), bt.float()))


def conv_view(Nb=3, L=64, C=32, Cout=64, k=4, s=2, p=1):
    # channels-last padded input [Nb, L+2p, C]; im2col row t = contiguous k*C slice at padded row t*s
    x = torch.zeros(Nb, L + 2 * p, C, device=dev, dtype=torch.bfloat16)
    x[:, p:p + L] = rnd(Nb, L, C)
    w = rnd(Cout, C, k)
    wr = w.permute(0, 2, 1).contiguous().view(Cout, k * C)  # [co, kk*C+ci]
    Lout = (L + 2 * p - k) // s + 1
    out = torch.empty(Nb, Lout, Cout, device=dev)
    A = K.View(x, k * C, Lout, Nb, s * C, (L + 2 * p) * C, major=0, gmul=1)
    K.gemm(A, K.kmajor(wr), Lout, Cout, k * C, groups=Nb, out_f32=out)
    ref = torch.nn.functional.conv1d(x[:, p:p + L].float().permute(0, 2, 1), w.float(), stride=s, padding=p)
    report("conv overlapping view", out, ref.permute(0, 2, 1))


def epilogue():
    M, N, Kd, G = 130, 96, 64, 2
    a, b = rnd(G, M, Kd), rnd(G, N, Kd)
    bias = torch.randn(G, N, device=dev)
    res = rnd(G, M, N)
    o32 = torch.empty(G, M, N, device=dev)
    o16 = torch.empty(G, M, N, device=dev, dtype=torch.bfloat16)
    K.gemm(K.kmajor(a), K.kmajor(b), M, N, Kd, groups=G, alpha=0.5, out_f32=o32, out_bf16=o16, bias=bias, bias_gs=N,
           res=res, relu=True)
    ref = torch.relu(0.5 * torch.einsum("gmk,gnk->gmn", a.float(), b.float()) + bias[:, None] + res.float())
    report("epilogue alpha+bias+res+relu f32", o32, ref)
    report("epilogue bf16 out", o16, ref, tol=3e-2)
    gate = rnd(G, M, N)
    o = torch.empty(G, M, N, device=dev)
    K.gemm(K.kmajor(a), K.kmajor(b), M, N, Kd, groups=G, out_f32=o, gate=gate, gate_scale=2.0)
    ref = torch.einsum("gmk,gnk->gmn", a.float(), b.float()) * (gate.float() > 0) * 2.0
    report("epilogue gate", o, ref)
    rng = torch.tensor([1234, 7], device=dev, dtype=torch.int64)
    o = torch.empty(G, M, N, device=dev)
    K.gemm(K.kmajor(a), K.kmajor(b), M, N, Kd, groups=G, out_f32=o, rng=rng, rng_site=3, drop_p=0.25)
    full = torch.einsum("gmk,gnk->gmn", a.float(), b.float())
    kept = (o != 0)
    frac = kept.float().mean().item()
    okv = torch.allclose(o[kept], full[kept] / 0.75, rtol=2e-2, atol=2e-2)
    results.append(("epilogue dropout", okv and abs(frac - 0.75) < 0.02, frac, 0.75))
    print(f"{'PASS' if okv and abs(frac - 0.75) < 0.02 else 'FAIL'} epilogue dropout keep_frac={frac:.4f}", flush=True)
    o2 = torch.empty(G, M, N, device=dev)
    K.gemm(K.kmajor(a), K.kmajor(b), M, N, Kd, groups=G, out_f32=o2, rng=rng, rng_site=3, drop_p=0.25)
    results.append(("dropout deterministic", torch.equal(o, o2), 0, 0))
    print(f"{'PASS' if torch.equal(o, o2) else 'FAIL'} dropout deterministic", flush=True)


def big():
    M, N, Kd = 4096, 2048, 2048
    a, b = rnd(M, Kd), rnd(N, Kd)
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    K.gemm(K.kmajor(a), K.kmajor(b), M, N, Kd, out_bf16=out)
    report("big 4096x2048x2048", out, (a @ b.t()), tol=2e-2)
    torch.cuda.synchronize()
    st, en = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st.record()
    for _ in range(20):
        K.gemm(K.kmajor(a), K.kmajor(b), M, N, Kd, out_bf16=out)
    en.record()
    torch.cuda.synchronize()
    ms = st.elapsed_time(en) / 20
    print(f"INFO big gemm {ms * 1e3:.1f} us  {2 * M * N * Kd / ms / 1e9:.1f} TFLOP/s", flush=True)


if __name__ == "__main__":
    print(torch.cuda.get_device_name(0), flush=True)
    case("nt basic", lambda: nt(128, 64, 64))
    case("nt ragged", lambda: nt(300, 256, 192))
    for bn in (64, 128, 256):
        case(f"nt bn{bn}", lambda bn=bn: nt(257, 512, 320, bn))
    case("nt small K", lambda: nt(100, 32, 32))
    case("grouped", grouped_shared_a)
    for bn in (64, 128):
        case("a_mn", lambda bn=bn: a_mn(bn=bn))
    for bn in (64, 128, 256):
        case("b_mn", lambda bn=bn: b_mn(bn=bn))
    case("both_mn", both_mn_slabs_splitk)
    case("both_mn nosplit", lambda: both_mn_slabs_splitk(splits=1))
    case("conv", conv_view)
    case("conv k8s4", lambda: conv_view(2, 128, 64, 64, 8, 4, 2))
    case("epilogue", epilogue)
    case("big", big)
    bad = [r for r in results if not r[1]]
    print(f"SUMMARY {len(results) - len(bad)}/{len(results)} passed", flush=True)
    sys.exit(1 if bad else 0)
