"""Kernel-level parity tests (need a B200): every C-ABI op against a plain fp32 PyTorch statement of
the same arithmetic on identical inputs (bf16-rounded where the kernel consumes bf16), and against the
CPU oracle (oracle/cpc_oracle.py) for the criterion math.  Tolerances are stated per test; integer /
index outputs are bit exact."""
import math
import os

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

BF = torch.bfloat16


@pytest.fixture(scope="module")
def K():
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    from cpc2_b200 import kernels
    return kernels


def bf(x):
    return x.to(BF)


def close(got, ref, rtol, atol, name=""):
    got, ref = got.float(), ref.float()
    err = (got - ref).abs()
    bound = atol + rtol * ref.abs()
    assert torch.isfinite(got).all(), f"{name}: non-finite"
    bad = (err > bound)
    assert not bad.any(), f"{name}: max err {err.max().item():.4g} (ref max {ref.abs().max().item():.4g}), " \
                          f"{bad.sum().item()} / {bad.numel()} out of tolerance"


def close_but(got, ref, rtol, atol, name="", allow=3):
    """`close` that tolerates a few elements: a ReLU mask recomputed in another summation order can flip where the
    pre-activation is within rounding of zero (that element's gradient then differs by its full size)."""
    got, ref = got.float(), ref.float()
    assert torch.isfinite(got).all(), f"{name}: non-finite"
    err = (got - ref).abs()
    bad = err > atol + rtol * ref.abs()
    assert bad.sum().item() <= allow, f"{name}: max err {err.max().item():.4g} (ref max {ref.abs().max().item():.4g}), " \
                                      f"{bad.sum().item()} / {bad.numel()} out of tolerance"


def cos(a, b):
    return F.cosine_similarity(a.float().flatten(), b.float().flatten(), dim=0).item()


# ------------------------------------------------------------------------------------------ GEMM
@pytest.mark.parametrize("N,G,M,Kd,unbiased,relu,with_res,halo", [
    # fewer than 65 row tiles: the row is split over a 4-CTA cluster, partial statistics exchanged through DSMEM (EPI_NORM_CL)
    (256, 3, 300, 192, True, True, False, 2), (256, 2, 928, 256, False, False, True, 0),
    (256, 1, 128, 2048, True, True, False, 1), (256, 5, 37, 64, False, False, True, 0), (256, 8, 1024, 2048, True, True, False, 2),
    # 65 or more row tiles: one BN = 256 accumulator tile per row (EPI_NORM)
    (256, 12, 928, 256, False, False, True, 0), (256, 9, 1000, 128, True, True, False, 1),
    # 512-wide rows (hidden 512): always the cluster form, 128 columns per CTA
    (512, 2, 300, 192, True, True, False, 2), (512, 12, 488, 512, False, False, True, 0), (512, 1, 50, 64, True, False, True, 1)])
def test_gemm_fused_row_norm(K, N, G, M, Kd, unbiased, relu, with_res, halo):
    """EPI_NORM / EPI_NORM_CL: ChannelNorm (unbiased variance, cpc/model.py:52-60) / LayerNorm (cpc/transformers.py:124,130)
    of the 256- or 512-wide output rows inside the GEMM epilogue, against fp32 torch on the same bf16 operands: pre-norm
    rows, the normalised rows (bf16 and fp32), mean / rstd, and the zero halo rows around every group."""
    torch.manual_seed(2)
    a = bf(torch.randn(G, M, Kd, device="cuda"))
    w = bf(torch.randn(G, N, Kd, device="cuda") / Kd ** 0.5)
    bias = torch.randn(G, N, device="cuda")
    gamma = 1 + 0.1 * torch.randn(G, N, device="cuda")
    beta = 0.1 * torch.randn(G, N, device="cuda")
    res = bf(torch.randn(G, M, N, device="cuda")) if with_res else None
    pre = torch.empty(G, M, N, device="cuda", dtype=BF)
    yb = torch.full((G, M + 2 * halo, N), 7.0, device="cuda", dtype=BF)
    y32 = torch.empty(G, M, N, device="cuda")
    mean = torch.empty(G * M, device="cuda")
    rstd = torch.empty(G * M, device="cuda")
    K.gemm(K.kmajor(a), K.kmajor(w), M, N, Kd, groups=G, out_bf16=pre, bias=bias, bias_gs=N, res=res, relu=relu,
           norm=dict(gamma=gamma, beta=beta, gs_param=N, out_bf16=(yb, N, (M + 2 * halo) * N), offset_bf16=halo * N,
                     out_f32=y32, mean=mean, rstd=rstd, gs_stat=M, unbiased=unbiased, halo=halo))
    x = torch.einsum("gmk,gnk->gmn", a.float(), w.float()) + bias[:, None, :]
    if with_res:
        x = x + res.float()
    mu = x.mean(-1, keepdim=True)
    var = x.var(-1, unbiased=unbiased, keepdim=True)
    ref = (x - mu) * torch.rsqrt(var + 1e-5) * gamma[:, None, :] + beta[:, None, :]
    if relu:
        ref = ref.relu()
    close(pre, x, 1e-2, 1e-2, "pre-norm")
    close(mean.view(G, M), mu.squeeze(-1), 1e-4, 1e-4, "mean")
    close(rstd.view(G, M), torch.rsqrt(var + 1e-5).squeeze(-1), 1e-3, 1e-5, "rstd")
    close(y32, ref, 1e-3, 2e-3, "norm f32")
    close(yb[:, halo:halo + M], ref, 1e-2, 1e-2, "norm bf16")
    if halo:
        assert float(yb[:, :halo].float().abs().max()) == 0 and float(yb[:, halo + M:].float().abs().max()) == 0


@pytest.mark.parametrize("M,N,Kd", [(128, 64, 64), (300, 256, 192), (928, 768, 256), (100, 32, 32)])
def test_gemm_nt(K, M, N, Kd):
    torch.manual_seed(0)
    a, b = bf(torch.randn(M, Kd, device="cuda")), bf(torch.randn(N, Kd, device="cuda"))
    out = torch.empty(M, N, device="cuda")
    K.gemm(K.kmajor(a), K.kmajor(b), M, N, Kd, out_f32=out)
    close(out, a.float() @ b.float().t(), 1e-4, 1e-3, "gemm nt")  # fp32 accumulate of exact bf16 products


def test_gemm_transposed_operands_and_slabs(K):
    torch.manual_seed(1)
    S_, R, M, N = 4, 100, 256, 320
    at, bt = bf(torch.randn(S_, R, M, device="cuda")), bf(torch.randn(S_, R, N, device="cuda"))
    out = torch.ones(M, N, device="cuda")
    A = K.View(at, M, R, S_, M, R * M, major=1)
    B = K.View(bt, N, R, S_, N, R * N, major=1)
    K.gemm(A, B, M, N, S_ * R, k_per_slab=R, k_splits=3, out_f32=out, accumulate=True)
    close(out, 1 + torch.einsum("srm,srn->mn", at.float(), bt.float()), 1e-4, 1e-3, "gemm tn slabs splitk")


# ------------------------------------------------------------------------------------------ row norms
@pytest.mark.parametrize("C", [32, 256, 512])
def test_channelnorm_relu_fwd_bwd(K, C):
    torch.manual_seed(2)
    N, L, halo = 3, 40, 2
    y = bf(torch.randn(N * L, C, device="cuda") * 2 + 0.3)
    gamma = torch.rand(C, device="cuda") + 0.5
    beta = torch.randn(C, device="cuda") * 0.1
    out = torch.full((N, L + 2 * halo, C), 7.0, device="cuda", dtype=BF)  # the kernel must clear the halo rows
    out32 = torch.zeros(N * L, C, device="cuda")
    mean, rstd = torch.empty(N * L, device="cuda"), torch.empty(N * L, device="cuda")
    K.rownorm_fwd(y, N * L, C, gamma, beta, out_bf16=out, mean=mean, rstd=rstd,
                  out_map=K.rowmap(L, L + 2 * halo, halo), unbiased=True, relu=True)
    K.rownorm_fwd(y, N * L, C, gamma, beta, out_f32=out32, unbiased=True, relu=True)
    yr = y.float().requires_grad_(True)
    g_, b_ = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    m = yr.mean(1, keepdim=True)
    v = yr.var(1, keepdim=True)  # unbiased, cpc/model.py:55
    ref = torch.relu((yr - m) * torch.rsqrt(v + 1e-5) * g_ + b_)
    close(out32, ref, 1e-4, 1e-4, "channelnorm fwd f32")
    close(out[:, halo:halo + L].reshape(N * L, C), ref, 1e-2, 1e-2, "channelnorm fwd bf16")
    assert out[:, :halo].abs().max() == 0 and out[:, halo + L:].abs().max() == 0  # halos cleared
    dout = torch.randn(N, L + 2 * halo, C, device="cuda")
    ref.backward(dout[:, halo:halo + L].reshape(N * L, C).to(BF).float())
    dx = torch.full((N, L + 2, C), 7.0, device="cuda", dtype=BF)
    dg, db, dsum = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    K.rownorm_bwd(bf(dout), y, mean, rstd, N * L, C, gamma, beta, dx_bf16=dx, dgamma=dg, dbeta=db, dxsum=dsum,
                  dout_map=K.rowmap(L, L + 2 * halo, halo), dx_map=K.rowmap(L, L + 2, 1), unbiased=True, relu=True)
    close(dx[:, 1:1 + L].reshape(N * L, C), yr.grad, 2e-2, 2e-2, "channelnorm dx")
    assert dx[:, 0].abs().max() == 0 and dx[:, L + 1].abs().max() == 0
    close(dg, g_.grad, 1e-3, 1e-2, "channelnorm dgamma")
    close(db, b_.grad, 1e-3, 1e-2, "channelnorm dbeta")
    close(dsum, yr.grad.sum(0), 1e-2, 3e-2, "channelnorm dxsum")


# ------------------------------------------------------------------------------------------ column norms
@pytest.mark.parametrize("mode,C,N,L", [("batchNorm", 256, 3, 700), ("batchNorm", 32, 2, 16), ("instanceNorm", 256, 5, 300),
                                        ("instanceNorm", 64, 3, 513), ("batchNormEval", 128, 2, 260), ("ID", 512, 2, 100)])
def test_colnorm_relu_fwd_bwd(K, mode, C, N, L):
    """cpc_colnorm_fwd / _bwd (encoder normMode batchNorm / instanceNorm / ID + ReLU, cpc/model.py:71-82,102-107) against
    torch's F.batch_norm / F.instance_norm on the same bf16-rounded input, incl. halo maps, ragged chunk counts and the
    running-estimate update."""
    torch.manual_seed(5)
    halo = 2
    y = bf(torch.randn(N * L, C, device="cuda") * 1.5 + 0.4 + torch.randn(C, device="cuda"))
    gamma = torch.rand(C, device="cuda") + 0.5
    beta = torch.randn(C, device="cuda") * 0.2
    rm0, rv0 = torch.randn(C, device="cuda") * 0.1 + 0.4, torch.rand(C, device="cuda") + 1.5
    rm, rv = rm0.clone(), rv0.clone()
    kind = {"batchNorm": "stats", "instanceNorm": "stats", "batchNormEval": "given", "ID": "ID"}[mode]
    rpg = L if mode == "instanceNorm" else 0
    groups = N if rpg else 1
    g, b = (None, None) if mode == "ID" else (gamma, beta)
    mean = rstd = None
    if kind == "stats":
        mean, rstd = torch.empty(groups * C, device="cuda"), torch.empty(groups * C, device="cuda")
    elif kind == "given":
        mean, rstd = rm0.clone(), torch.rsqrt(rv0 + 1e-5)
    partial, sums = K.colnorm_scratch(N * L, C, rpg, "cuda")
    out = torch.full((N, L + 2 * halo, C), 7.0, device="cuda", dtype=BF)
    out32 = torch.zeros(N * L, C, device="cuda")
    K.colnorm_fwd(y, N * L, C, kind, rows_per_group=rpg, gamma=g, beta=b, mean=mean, rstd=rstd, partial=partial,
                  running_mean=rm if mode == "batchNorm" else None, running_var=rv if mode == "batchNorm" else None,
                  momentum=0.1, out_bf16=out, out_map=K.rowmap(L, L + 2 * halo, halo), out_f32=out32)
    yr = y.float().view(N, L, C).permute(0, 2, 1).contiguous().requires_grad_(True)  # [N, C, L] as the reference sees it
    g_, b_ = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    rm_t, rv_t = rm0.clone(), rv0.clone()
    if mode == "batchNorm":
        pre = F.batch_norm(yr, rm_t, rv_t, g_, b_, training=True, momentum=0.1, eps=1e-5)
    elif mode == "batchNormEval":
        pre = F.batch_norm(yr, rm_t, rv_t, g_, b_, training=False, eps=1e-5)
    elif mode == "instanceNorm":
        pre = F.instance_norm(yr, weight=g_, bias=b_, eps=1e-5)
    else:
        pre = yr
    ref = torch.relu(pre).permute(0, 2, 1).reshape(N * L, C)
    close(out32, ref, 1e-4, 2e-4, f"colnorm {mode} fwd f32")
    close(out[:, halo:halo + L].reshape(N * L, C), ref, 1e-2, 1e-2, f"colnorm {mode} fwd bf16")
    assert out[:, :halo].abs().max() == 0 and out[:, halo + L:].abs().max() == 0
    if mode == "batchNorm":
        close(rm, rm_t, 1e-4, 1e-5, "running_mean")
        close(rv, rv_t, 1e-4, 1e-5, "running_var")
    dout = torch.randn(N, L + 2 * halo, C, device="cuda")
    ref.backward(dout[:, halo:halo + L].reshape(N * L, C).to(BF).float())
    dx = torch.full((N, L + 2, C), 7.0, device="cuda", dtype=BF)
    dx32 = torch.zeros(N * L, C, device="cuda")
    dg, db, dsum = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    K.colnorm_bwd(bf(dout), y, N * L, C, kind, partial, sums, rows_per_group=rpg, gamma=g, beta=b, mean=mean, rstd=rstd,
                  dout_map=K.rowmap(L, L + 2 * halo, halo), dx_bf16=dx, dx_map=K.rowmap(L, L + 2, 1), dx_f32=dx32,
                  dgamma=None if mode == "ID" else dg, dbeta=None if mode == "ID" else db, dxsum=dsum)
    dref = yr.grad.permute(0, 2, 1).reshape(N * L, C)
    close_but(dx32, dref, 1e-3, 1e-3 * float(dref.abs().max()) + 1e-5, f"colnorm {mode} dx f32")
    close_but(dx[:, 1:1 + L].reshape(N * L, C), dref, 2e-2, 2e-2 * float(dref.abs().max()), f"colnorm {mode} dx bf16")
    assert dx[:, 0].abs().max() == 0 and dx[:, L + 1].abs().max() == 0
    if mode != "ID":
        close(dg, g_.grad, 1e-3, 1e-2, f"colnorm {mode} dgamma")
        close(db, b_.grad, 1e-3, 1e-2, f"colnorm {mode} dbeta")
    # bias gradient of the convolution in front: zero under computed statistics (torch holds rounding noise there)
    want = dref.sum(0)
    if kind == "stats":
        assert dsum.abs().max() == 0 and want.abs().max() <= 1e-3 * dref.abs().sum(0).max()
    else:
        close(dsum, want, 1e-3, 1e-2, f"colnorm {mode} dxsum")
    # a second call with another launch shape gives the SAME bits (fixed-order chunk combination)
    if kind == "stats":
        mean2, rstd2 = torch.empty_like(mean), torch.empty_like(rstd)
        K.colnorm_fwd(y, N * L, C, kind, rows_per_group=rpg, gamma=g, beta=b, mean=mean2, rstd=rstd2, partial=partial,
                      out_f32=out32)
        assert torch.equal(mean, mean2) and torch.equal(rstd, rstd2)


@pytest.mark.parametrize("C,W,N", [(32, 1600, 2), (256, 2560, 3), (512, 1765, 2), (64, 20480, 2)])
def test_conv0_plain_fwd_bwd(K, C, W, N):
    """conv0 without a fused norm (cpc/model.py:85: Conv1d(1, C, 10, stride 5, padding 3)) against F.conv1d."""
    torch.manual_seed(6)
    x = torch.randn(N, W, device="cuda") * 0.3
    w = (torch.randn(C, 1, 10, device="cuda") * 0.3).requires_grad_(True)
    b = (torch.randn(C, device="cuda") * 0.1).requires_grad_(True)
    L0 = (W - 4) // 5 + 1
    y = torch.empty(N * L0, C, device="cuda", dtype=BF)
    K.conv0_plain_fwd(x, w.detach(), b.detach(), y, N, W, C)
    ref = F.conv1d(x.view(N, 1, W), w, b, stride=5, padding=3)  # [N, C, L0]
    assert ref.shape[-1] == L0
    close(y.view(N, L0, C), ref.permute(0, 2, 1), 1e-2, 1e-2, "conv0 plain fwd")
    dy = torch.randn(N * L0, C, device="cuda")
    ref.backward(dy.view(N, L0, C).permute(0, 2, 1))
    dw = torch.zeros(C, 10, device="cuda")
    K.conv0_plain_bwd(x, dy, dw, N, W, C)
    close(dw, w.grad.view(C, 10), 1e-3, 1e-3 * float(w.grad.abs().max()), "conv0 plain dw")


def test_layernorm_groups_fwd_bwd(K):
    torch.manual_seed(3)
    G, T, C = 3, 50, 256
    x = bf(torch.randn(G * T, C, device="cuda"))
    gamma, beta = torch.rand(G, C, device="cuda") + 0.5, torch.randn(G, C, device="cuda") * 0.1
    out = torch.empty(G * T, C, device="cuda")
    mean, rstd = torch.empty(G * T, device="cuda"), torch.empty(G * T, device="cuda")
    K.rownorm_fwd(x, G * T, C, gamma, beta, out_f32=out, mean=mean, rstd=rstd, rows_per_group=T)
    xr = x.float().view(G, T, C).requires_grad_(True)
    g_, b_ = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    ref = torch.stack([F.layer_norm(xr[g], (C,), g_[g], b_[g]) for g in range(G)])
    close(out.view(G, T, C), ref, 1e-4, 1e-4, "layernorm fwd")
    dout = torch.randn(G, T, C, device="cuda")
    ref.backward(dout)
    dx = torch.empty(G * T, C, device="cuda")
    dg, db = torch.zeros(G, C, device="cuda"), torch.zeros(G, C, device="cuda")
    K.rownorm_bwd(dout.view(G * T, C), x, mean, rstd, G * T, C, gamma, beta, dx_f32=dx, dgamma=dg, dbeta=db,
                  rows_per_group=T)
    close(dx.view(G, T, C), xr.grad, 1e-3, 1e-3, "layernorm dx")
    close(dg, g_.grad, 1e-3, 1e-3, "layernorm dgamma")
    close(db, b_.grad, 1e-3, 1e-3, "layernorm dbeta")


# ------------------------------------------------------------------------------------------ conv0
@pytest.mark.parametrize("C,W", [(32, 1600), (256, 1600), (64, 1760), (96, 480), (512, 1760), (256, 20480)])
def test_conv0_fused_fwd_bwd(K, C, W):
    from oracle import cpc_oracle as oc
    torch.manual_seed(4)
    N, halo = 3, 2
    L0 = (W - 4) // 5 + 1
    x = 0.1 * torch.randn(N, W, device="cuda")
    conv = torch.nn.Conv1d(1, C, 10, stride=5, padding=3).cuda()
    gamma = (torch.rand(C, device="cuda") + 0.5).requires_grad_(True)
    beta = (torch.randn(C, device="cuda") * 0.1).requires_grad_(True)
    out = torch.full((N, L0 + 2 * halo, C), 7.0, device="cuda", dtype=BF)  # the kernel must zero the halo rows itself
    mean, rstd = torch.empty(N * L0, device="cuda"), torch.empty(N * L0, device="cuda")
    K.conv0_fwd(x, conv.weight, conv.bias, gamma, beta, out, mean, rstd, N, W, C, halo)
    assert float(out[:, :halo].abs().max()) == 0.0 and float(out[:, halo + L0:].abs().max()) == 0.0
    ref = F.relu(oc.channel_norm(conv(x.unsqueeze(1)), gamma.view(1, C, 1), beta.view(1, C, 1)))  # [N,C,L0]
    close(out[:, halo:halo + L0], ref.permute(0, 2, 1), 1e-2, 1e-2, "conv0 fwd")
    dout = bf(torch.randn(N, L0 + 2 * halo, C, device="cuda"))
    ref.backward(dout[:, halo:halo + L0].float().permute(0, 2, 1))
    dw, db = torch.zeros(C, 10, device="cuda"), torch.zeros(C, device="cuda")
    dg, dbe = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    K.conv0_bwd(x, conv.weight, conv.bias, gamma, beta, mean, rstd, dout, dw, db, dg, dbe, N, W, C, halo)
    close(dw, conv.weight.grad.view(C, 10), 2e-3, 2e-3 * conv.weight.grad.abs().max().item(), "conv0 dw")
    close(db, conv.bias.grad, 2e-3, 2e-3 * conv.bias.grad.abs().max().item(), "conv0 db")
    close(dg, gamma.grad, 2e-3, 2e-3 * gamma.grad.abs().max().item(), "conv0 dgamma")
    close(dbe, beta.grad, 2e-3, 2e-3 * beta.grad.abs().max().item(), "conv0 dbeta")


# ------------------------------------------------------------------------------------------ conv1..4 as GEMMs
@pytest.mark.parametrize("C,k,s,p,L", [(64, 8, 4, 2, 256), (256, 4, 2, 1, 128), (32, 4, 2, 1, 32)])
def test_conv_implicit_gemm_fwd_dgrad_wgrad(K, C, k, s, p, L):
    torch.manual_seed(5)
    Nb = 3
    Lout = L // s
    Lp = L + 2 * p
    assert Lp == s * (Lout + 1)
    conv = torch.nn.Conv1d(C, C, k, stride=s, padding=p).cuda()
    x = bf(torch.randn(Nb, L, C, device="cuda"))
    xp = torch.zeros(Nb, Lp, C, device="cuda", dtype=BF)
    xp[:, p:p + L] = x
    wf = torch.empty(C, k * C, device="cuda", dtype=BF)
    wd = torch.empty(s * C, 2 * C, device="cuda", dtype=BF)
    K.conv_weight_pack(conv.weight, wf, wd, C, C, k, s)
    assert torch.equal(wf, bf(conv.weight.permute(0, 2, 1).reshape(C, k * C)))
    # forward
    y = torch.empty(Nb, Lout, C, device="cuda")
    A = K.View(xp, k * C, Lout, Nb, s * C, Lp * C, major=0, gmul=1)
    K.gemm(A, K.kmajor(wf), Lout, C, k * C, groups=Nb, out_f32=y, bias=conv.bias.detach())
    xr = x.float().permute(0, 2, 1).requires_grad_(True)
    wq = conv.weight.detach().to(BF).float().requires_grad_(True)
    ref = F.conv1d(xr, wq, conv.bias, stride=s, padding=p)
    close(y, ref.permute(0, 2, 1), 1e-3, 1e-3, "conv fwd")
    dy = bf(torch.randn(Nb, Lout, C, device="cuda"))
    ref.backward(dy.float().permute(0, 2, 1))
    # dgrad: rows [dy[t-1], dy[t]] . [W_hi; W_lo] -> padded dx
    dyp = torch.zeros(Nb, Lout + 2, C, device="cuda", dtype=BF)
    dyp[:, 1:1 + Lout] = dy
    dxp = torch.empty(Nb, Lp, C, device="cuda", dtype=BF)
    A = K.View(dyp, 2 * C, Lout + 1, Nb, C, (Lout + 2) * C, major=0, gmul=1)
    K.gemm(A, K.kmajor(wd), Lout + 1, s * C, 2 * C, groups=Nb, out_bf16=(dxp, s * C, Lp * C))
    close(dxp[:, p:p + L], xr.grad.permute(0, 2, 1), 2e-2, 2e-2 * xr.grad.abs().max().item(), "conv dgrad")
    # wgrad: dWf[co, j] = sum_{n,t} dy[n,t,co] * im2col[n,t,j]; both operands MN-major, slabs = sequences
    dwf = torch.zeros(C, k * C, device="cuda")
    A = K.View(dyp, C, Lout + 2, Nb, C, (Lout + 2) * C, major=1, row_off=1)
    B = K.View(xp, k * C, Lout, Nb, s * C, Lp * C, major=1)
    K.gemm(A, B, C, k * C, Nb * Lout, k_per_slab=Lout, k_splits=2, out_f32=dwf, accumulate=True)
    dw = torch.zeros_like(conv.weight)  # the unpack ADDS (two half-batches share one gradient)
    K.conv_wgrad_unpack(dwf, dw, C, C, k)
    close(dw, wq.grad, 1e-3, 1e-3 * wq.grad.abs().max().item(), "conv wgrad")


# ------------------------------------------------------------------------------------------ helpers
def test_streaming_helpers(K):
    torch.manual_seed(6)
    x = torch.randn(100003, device="cuda")
    assert torch.equal(K.cast_bf16(x), x.to(BF))
    assert torch.equal(K.cast_f32(x.to(BF)), x.to(BF).float())
    G, R, C = 3, 500, 96
    m = torch.randn(G, R, C, device="cuda")
    out = torch.zeros(G, C, device="cuda")
    K.colsum(m, out, R, C, C, R * C, G)
    close(out, m.sum(1), 1e-4, 1e-3, "colsum f32")
    out = torch.zeros(G, C, device="cuda")
    K.colsum(bf(m), out, R, C, C, R * C, G)
    close(out, bf(m).float().sum(1), 1e-4, 1e-3, "colsum bf16")
    o32 = torch.ones(R * C, device="cuda")
    K.groupsum(m, R * C, G, out_f32=o32, accumulate=True)
    close(o32, 1 + m.sum(0).flatten(), 1e-5, 1e-5, "groupsum")


@pytest.mark.parametrize("B,S,Kp,NEG", [(8, 128, 12, 128), (2, 16, 4, 8), (3, 256, 12, 64)])
def test_neg_ext_index_bit_exact(K, B, S, Kp, NEG):
    from oracle import cpc_oracle as oc
    W = S - Kp
    g = torch.Generator(device="cuda").manual_seed(11)
    n = B * NEG * W
    bi = torch.randint(0, B, (n,), device="cuda", generator=g)
    si = torch.randint(1, S, (n,), device="cuda", generator=g)
    ext = K.neg_ext_index(bi, si, B, NEG, W, S)
    ref = oc.negative_ext_index(bi.cpu(), si.cpu(), B, S, W, NEG)
    assert ext.dtype == torch.int64 and torch.equal(ext.cpu(), ref)


# ------------------------------------------------------------------------------------------ InfoNCE tail
@pytest.mark.parametrize("B,S,Kp,NEG,weighted", [(8, 128, 12, 128, False), (2, 16, 4, 8, True), (4, 64, 6, 200, False)])
def test_nce_fwd_bwd(K, B, S, Kp, NEG, weighted):
    from oracle import cpc_oracle as oc
    torch.manual_seed(7)
    W = S - Kp
    T, BS = B * W, B * S
    logits = torch.randn(Kp, T, BS, device="cuda") * 2
    # force exact ties between the positive and a negative on some rows: torch.max keeps index 0
    n = B * NEG * W
    bi = torch.randint(0, B, (n,), device="cuda")
    si = torch.randint(1, S, (n,), device="cuda")
    ext = K.neg_ext_index(bi, si, B, NEG, W, S)
    weights = (torch.rand(T, device="cuda") + 0.5) if weighted else None
    lse = torch.empty(Kp, T, device="cuda")
    loss, hits = torch.zeros(Kp, device="cuda"), torch.zeros(Kp, device="cuda")
    K.nce_fwd(logits, ext, weights, lse, loss, hits, Kp, B, W, S, NEG, BS)
    # reference: gather the NEG+1 scalars exactly as criterion.py:267-284 would see them
    lg = logits.clone().requires_grad_(True)
    e = ext.view(B, NEG, W)
    ref_loss, ref_hits, total = [], [], 0
    for k in range(Kp):
        rows = lg[k].view(B, W, BS)
        pos_col = (torch.arange(B, device="cuda").view(B, 1) * S + torch.arange(W, device="cuda").view(1, W) + k + 1)
        pos = rows.gather(2, pos_col.unsqueeze(2))                       # [B,W,1]
        neg = rows.gather(2, e.permute(0, 2, 1))                         # [B,W,NEG]
        full = torch.cat([pos, neg], dim=2).reshape(T, NEG + 1)
        ce = F.cross_entropy(full, torch.zeros(T, dtype=torch.long, device="cuda"), reduction="none")
        if weights is not None:
            ce = ce * weights
        ref_loss.append(ce.mean())
        ref_hits.append((full.max(1)[1] == 0).sum())
        total = total + ce.mean() * (k + 1)  # distinct upstream gradient per k
    close(loss / T, torch.stack(ref_loss), 1e-5, 1e-5, "nce loss")
    assert torch.equal(hits.long(), torch.stack(ref_hits)), "argmax hits must be bit exact"
    total.backward()
    gscale = torch.arange(1, Kp + 1, device="cuda", dtype=torch.float32)
    dl = torch.empty(Kp, T, BS, device="cuda", dtype=BF)
    K.nce_bwd(logits, ext, weights, lse, gscale, dl, Kp, B, W, S, NEG, BS)
    close(dl, lg.grad, 1e-2, 1e-2 * lg.grad.abs().max().item(), "nce dlogits")


def test_nce_tie_goes_to_positive(K):
    B, S, Kp, NEG = 1, 8, 2, 4
    W = S - Kp
    logits = torch.zeros(Kp, W, S, device="cuda")  # all equal -> argmax must be class 0 for every row
    bi = torch.zeros(B * NEG * W, dtype=torch.long, device="cuda")
    si = torch.randint(1, S, (B * NEG * W,), device="cuda")
    ext = K.neg_ext_index(bi, si, B, NEG, W, S)
    lse = torch.empty(Kp, W, device="cuda")
    loss, hits = torch.zeros(Kp, device="cuda"), torch.zeros(Kp, device="cuda")
    K.nce_fwd(logits, ext, None, lse, loss, hits, Kp, B, W, S, NEG, S)
    assert hits.tolist() == [W] * Kp
    close(loss / W, torch.full((Kp,), math.log(NEG + 1), device="cuda"), 1e-6, 1e-6, "uniform loss = ln(NEG+1)")


# ------------------------------------------------------------------------------------------ LSTM
@pytest.mark.parametrize("H,N,S", [(32, 5, 12), (64, 16, 9), (256, 16, 128), (256, 20, 16), (128, 3, 7), (512, 4, 6)])
def test_lstm_recurrence_fwd_bwd(K, H, N, S):
    torch.manual_seed(8)
    lstm = torch.nn.LSTM(H, H, batch_first=True).cuda()
    with torch.no_grad():
        lstm.weight_hh_l0.copy_(lstm.weight_hh_l0.to(BF).float())  # kernel consumes bf16 W_hh
    x = torch.randn(N, S, H, device="cuda")
    gx = (x @ lstm.weight_ih_l0.t() + lstm.bias_ih_l0 + lstm.bias_hh_l0).detach().contiguous()
    w_hh = lstm.weight_hh_l0.detach().to(BF).contiguous()
    out = torch.empty(N, S, H, device="cuda")
    h_pad = torch.empty(N, S + 1, H, device="cuda", dtype=BF)
    gates = torch.empty(N, S, 4 * H, device="cuda", dtype=BF)
    c_all = torch.empty(N, S, H, device="cuda")
    h_n, c_n = torch.empty(N, H, device="cuda"), torch.empty(N, H, device="cuda")
    K.lstm_fwd(gx, w_hh, N, S, H, h_pad=h_pad, out=out, gates=gates, c_all=c_all, h_n=h_n, c_n=c_n)
    xr = x.clone().requires_grad_(True)
    ref, (rh, rc) = lstm(xr)
    # h is carried in bf16 between steps: tolerance is bf16 epsilon accumulated over the chain
    close(out, ref, 2e-2, 2e-2, "lstm out")
    close(h_n, rh[0], 2e-2, 2e-2, "lstm h_n")
    close(c_n, rc[0], 2e-2, 3e-2, "lstm c_n")
    close(h_pad[:, 1:], ref, 2e-2, 2e-2, "lstm h_pad")
    assert h_pad[:, 0].abs().max() == 0
    dout = torch.randn(N, S, H, device="cuda")
    ref.backward(dout)
    dgates = torch.empty(N, S, 4 * H, device="cuda", dtype=BF)
    K.lstm_bwd(w_hh, gates, c_all, dout, dgates, N, S, H)
    # finish the backward with plain torch on the kernel's gate gradients
    dg = dgates.float()
    dx = dg @ lstm.weight_ih_l0
    dw_ih = torch.einsum("nsg,nsh->gh", dg, x)
    hprev = torch.cat([torch.zeros(N, 1, H, device="cuda"), ref.detach()[:, :-1]], dim=1)
    dw_hh = torch.einsum("nsg,nsh->gh", dg, hprev)
    assert cos(dx, xr.grad) > 0.999, cos(dx, xr.grad)
    assert cos(dw_ih, lstm.weight_ih_l0.grad) > 0.999
    assert cos(dw_hh, lstm.weight_hh_l0.grad) > 0.999
    assert cos(dg.sum((0, 1)), lstm.bias_ih_l0.grad) > 0.999
    close(dx, xr.grad, 5e-2, 5e-2 * xr.grad.abs().max().item(), "lstm dx")


@pytest.mark.parametrize("cell,H,Cin,N,S,nl", [("GRU", 64, 32, 5, 20, 2), ("GRU", 256, 256, 16, 128, 1), ("RNN", 32, 64, 3, 9, 2),
                                               ("LSTM", 64, 32, 9, 11, 2), ("GRU", 512, 128, 4, 6, 1)])
def test_cpcar_module_matches_torch_rnn(K, cell, H, Cin, N, S, nl):
    """cpc2_b200.model.CPCAR (cluster kernels, four-slot GRU / RNN layout) against torch's own nn.GRU / nn.RNN /
    nn.LSTM on identical parameters: outputs, carried hidden state (keepHidden, cpc/model.py:196-201) and all
    parameter / input gradients."""
    from cpc2_b200.model import CPCAR
    torch.manual_seed(11)
    ar = CPCAR(Cin, H, True, nl, mode=cell).cuda()
    ref = {"GRU": torch.nn.GRU, "RNN": torch.nn.RNN, "LSTM": torch.nn.LSTM}[cell](Cin, H, num_layers=nl, batch_first=True).cuda()
    ref.load_state_dict(ar.baseNet.state_dict())
    x1 = torch.randn(N, S, Cin, device="cuda")
    x2 = torch.randn(N, S, Cin, device="cuda", requires_grad=True)
    x2r = x2.detach().clone().requires_grad_(True)
    with torch.no_grad():
        o1 = ar(x1)
        r1, hid = ref(x1)
    close(o1, r1, 2e-2, 2e-2, f"{cell} out (first chunk)")
    o2 = ar(x2)          # starts from the kept hidden state
    r2, _ = ref(x2r, hid)
    close(o2, r2, 2e-2, 2e-2, f"{cell} out (second chunk, carried state)")
    dout = torch.randn_like(r2)
    o2.backward(dout)
    r2.backward(dout)
    assert cos(x2.grad, x2r.grad) > 0.998, cos(x2.grad, x2r.grad)
    for (n, g), (_, gr) in zip(ar.baseNet.named_parameters(), ref.named_parameters()):
        assert g.grad.shape == gr.grad.shape, n
        assert cos(g.grad, gr.grad) > 0.998, (n, cos(g.grad, gr.grad))
        close(g.grad, gr.grad, 5e-2, 5e-2 * gr.grad.abs().max().item(), f"{cell} d{n}")


def test_dropout_kernel_mask_is_a_function_of_rng_site_index(K):
    """cpc_dropout: in place, kept elements scaled by 1/(1-p), the same (rng, site) reproduces the mask on another
    tensor (that is the backward), a different site or step does not."""
    torch.manual_seed(4)
    n = 12 * 928 * 256 + 1
    x = torch.randn(n, device="cuda") + 3.0
    rng = torch.tensor([77, 5], device="cuda", dtype=torch.int64)
    y = K.dropout_(x.clone(), rng, 7, 0.5)
    keep = y != 0
    assert abs(keep.float().mean().item() - 0.5) < 2e-3
    assert torch.equal(y[keep], x[keep] * 2.0)
    g = K.dropout_(bf(torch.ones(n, device="cuda")), rng, 7, 0.5)  # bf16 tensor, same mask
    assert torch.equal(g != 0, keep) and torch.equal(g[keep].float(), torch.full_like(g[keep].float(), 2.0))
    other = K.dropout_(x.clone(), rng, 8, 0.5) != 0
    assert (other != keep).float().mean().item() > 0.4
    rng2 = torch.tensor([77, 6], device="cuda", dtype=torch.int64)
    assert ((K.dropout_(x.clone(), rng2, 7, 0.5) != 0) != keep).float().mean().item() > 0.4
    q = K.dropout_(x.clone(), rng, 7, 0.1) != 0
    assert abs(q.float().mean().item() - 0.9) < 2e-3


# ------------------------------------------------------------------------------------------ attention
@pytest.mark.parametrize("G,Bt,Hh,S,dk,relpos", [(2, 3, 8, 116, 32, True), (1, 2, 8, 12, 4, True),
                                                 (1, 2, 4, 128, 64, False), (1, 1, 8, 244, 64, True),
                                                 (1, 2, 8, 117, 32, True), (2, 1, 4, 256, 64, True),
                                                 (1, 2, 8, 200, 32, True), (1, 1, 2, 131, 32, False)])
def test_attention_fwd_bwd(K, G, Bt, Hh, S, dk, relpos):
    from oracle import cpc_oracle as oc
    torch.manual_seed(9)
    d = Hh * dk
    qkv = bf(torch.randn(G, Bt * S, 3 * d, device="cuda"))
    krel = (torch.rand(G, dk, S, device="cuda") - 0.5) * (2 / math.sqrt(dk)) if relpos else None
    out = torch.empty(G, Bt * S, d, device="cuda", dtype=BF)
    out32 = torch.empty(G, Bt * S, d, device="cuda")
    rowsum = torch.full((G, Bt, Hh, S), float("nan"), device="cuda")
    probs = torch.zeros(G, Bt, Hh, S, S, device="cuda", dtype=BF)
    K.attn_fwd(qkv, krel, out, probs, G, Bt, Hh, S, dk, out_f32=out32, rowsum=rowsum)
    if dk in (32, 64):  # tensor-core path: row sums of the stored bf16 probabilities (CpcAttn.rowsum)
        close(rowsum, probs.float().sum(-1), 1e-5, 1e-5, "row sums of the stored probabilities")
    qr = qkv.float().requires_grad_(True)
    kr = krel.clone().requires_grad_(True) if relpos else None
    refs = []
    for g in range(G):
        x = qr[g].view(Bt, S, 3, Hh, dk)
        def heads(i):
            return x[:, :, i].transpose(1, 2).reshape(Bt * Hh, S, dk)
        o = oc.relpos_causal_attention(heads(0).cpu(), heads(1).cpu(), heads(2).cpu(),
                                       None if kr is None else kr[g].cpu(), S)
        refs.append(o.view(Bt, Hh, S, dk).transpose(1, 2).reshape(Bt * S, d))
    ref = torch.stack(refs).cuda()
    close(out, ref, 2e-2, 2e-2, "attention out")
    dout = bf(torch.randn(G, Bt * S, d, device="cuda"))
    ref.backward(dout.float())
    # both ways of forming the row statistic D (CpcAttn): (a) dO.O / rowsum from the forward's fp32 O, (b) in-kernel sweep
    for extra in (dict(out_f32=out32, rowsum=rowsum), dict()):
        dqkv = torch.empty_like(qkv)
        dscore = torch.empty_like(probs)
        dkrel = torch.zeros(G, dk, S, device="cuda") if relpos else None
        K.attn_bwd(qkv, krel, out, probs, dout, dqkv, dscore, dkrel, G, Bt, Hh, S, dk, **extra)
        assert cos(dqkv, qr.grad) > 0.999, cos(dqkv, qr.grad)
        close(dqkv, qr.grad, 5e-2, 3e-2 * qr.grad.abs().max().item(), "attention dqkv")
        if relpos:
            assert cos(dkrel, kr.grad) > 0.999
            close(dkrel, kr.grad, 5e-2, 3e-2 * kr.grad.abs().max().item(), "attention dkrel")


@pytest.mark.parametrize("S,dk,Hh", [(116, 32, 8), (244, 64, 8), (24, 8, 8)])
def test_attention_backward_with_peaked_softmax(K, S, dk, Hh):
    """Trained models have peaked attention rows AND keys / values with a large component common to all positions:
    dP_ij is then nearly constant along a row, dS_ij = P_ij (dP_ij - D_i) is a small difference of large numbers, and
    any row-coherent error of D_i is multiplied by the common component of K in dQ (this is what cost the Wq gradients
    of the step two digits away from init, profiles/r02_attn_grad_probe_before_fix.txt).  Both exact forms of D
    (CpcAttn: dO.O / rowsum from the forward, or the in-kernel sweep) keep cosine >= 0.999 against fp32 autograd."""
    from oracle import cpc_oracle as oc
    torch.manual_seed(13)
    G, Bt = 1, 2
    d = Hh * dk
    qkv = torch.randn(G, Bt * S, 3 * d, device="cuda")
    qkv[..., :2 * d] *= 3.0  # q, k three times larger: logits ~ N(0, 9^2 / ...) -> one or two keys dominate a row
    qkv[..., d:] += 5.0 * torch.randn(1, 1, 2 * d, device="cuda")  # k, v: common component 5x the per-position part
    qkv = bf(qkv)
    krel = (torch.rand(G, dk, S, device="cuda") - 0.5) * (2 / math.sqrt(dk))
    out = torch.empty(G, Bt * S, d, device="cuda", dtype=BF)
    out32 = torch.empty(G, Bt * S, d, device="cuda")
    rowsum = torch.empty(G, Bt, Hh, S, device="cuda")
    probs = torch.zeros(G, Bt, Hh, S, S, device="cuda", dtype=BF)
    K.attn_fwd(qkv, krel, out, probs, G, Bt, Hh, S, dk, out_f32=out32, rowsum=rowsum)
    close(out32, out, 1e-2, 1e-2 * float(out32.abs().max()), "fp32 copy of the attention output")
    qr = qkv.float().requires_grad_(True)
    kr = krel.clone().requires_grad_(True)
    x = qr[0].view(Bt, S, 3, Hh, dk)
    heads = [x[:, :, i].transpose(1, 2).reshape(Bt * Hh, S, dk) for i in range(3)]
    o = oc.relpos_causal_attention(heads[0].cpu(), heads[1].cpu(), heads[2].cpu(), kr[0].cpu(), S)
    ref = o.view(Bt, Hh, S, dk).transpose(1, 2).reshape(1, Bt * S, d).cuda()
    assert float(probs.float().amax(-1).mean()) > 0.6  # the rows really are peaked
    dout = bf(torch.randn(G, Bt * S, d, device="cuda"))
    ref.backward(dout.float())
    res = {}
    for name, extra in (("fwd_rowsum", dict(out_f32=out32, rowsum=rowsum)), ("in_kernel", dict())):
        dqkv = torch.empty_like(qkv)
        dkrel = torch.zeros(G, dk, S, device="cuda")
        K.attn_bwd(qkv, krel, out, probs, dout, dqkv, torch.empty_like(probs), dkrel, G, Bt, Hh, S, dk, **extra)
        res[name] = (cos(dqkv[..., :d], qr.grad[..., :d]), cos(dqkv[..., d:2 * d], qr.grad[..., d:2 * d]), cos(dkrel, kr.grad))
    for name, (cq, ck, cr_) in res.items():
        assert cq >= 0.999 and ck >= 0.999 and cr_ >= 0.999, res


@pytest.mark.parametrize("S,dk", [(32, 32), (64, 64), (160, 64)])
def test_attention_dropout_backward_matches_masked_reference(K, S, dk):
    """Dropout on the probabilities (transformers.py:69): the kernel's mask depends only on (seed, element index),
    so it is recovered exactly by a forward with V = identity (out = P~) and fed to an fp32 autograd reference."""
    torch.manual_seed(11)
    G, Bt, Hh, p = 1, 2, 2, 0.25
    if S > dk:  # recover the mask block by block: V = shifted identities
        nblk = (S + dk - 1) // dk
    else:
        nblk = 1
    d = Hh * dk
    rng = torch.tensor([1234, 7], device="cuda", dtype=torch.int64)
    krel = (torch.rand(G, dk, S, device="cuda") - 0.5) * (2 / math.sqrt(dk))
    qkv = bf(torch.randn(G, Bt * S, 3 * d, device="cuda") * 0.5)
    probs = torch.zeros(G, Bt, Hh, S, S, device="cuda", dtype=BF)
    keep = torch.zeros(Bt, Hh, S, S, device="cuda", dtype=torch.bool)
    for blk in range(nblk):
        probe = qkv.clone().view(G, Bt, S, 3, Hh, dk)
        probe[:, :, :, 2] = 0
        for j in range(blk * dk, min(S, (blk + 1) * dk)):
            probe[:, :, j, 2, :, j - blk * dk] = 1
        o = torch.empty(G, Bt * S, d, device="cuda", dtype=BF)
        K.attn_fwd(probe.view(G, Bt * S, 3 * d), krel, o, probs, G, Bt, Hh, S, dk, rng=rng, site=3, drop_p=p)
        o = o.view(Bt, S, Hh, dk).permute(0, 2, 1, 3).float()  # [Bt, Hh, i, j - blk*dk] = P~[i, j]
        w = min(S, (blk + 1) * dk) - blk * dk
        keep[:, :, :, blk * dk:blk * dk + w] = o[..., :w] != 0
    tri = torch.tril(torch.ones(S, S, device="cuda", dtype=torch.bool))
    frac = keep[:, :, tri].float().mean().item()
    assert abs(frac - (1 - p)) < 0.03, frac
    out = torch.empty(G, Bt * S, d, device="cuda", dtype=BF)
    out32 = torch.empty(G, Bt * S, d, device="cuda")
    rowsum = torch.empty(G, Bt, Hh, S, device="cuda")
    K.attn_fwd(qkv, krel, out, probs, G, Bt, Hh, S, dk, rng=rng, site=3, drop_p=p, out_f32=out32, rowsum=rowsum)
    qr = qkv.float().requires_grad_(True)
    kr = krel.clone().requires_grad_(True)
    x = qr[0].view(Bt, S, 3, Hh, dk)
    q, k, v = (x[:, :, i].transpose(1, 2) for i in range(3))  # [Bt, Hh, S, dk]
    idx = (S - 1 - torch.arange(S, device="cuda").view(S, 1) + torch.arange(S, device="cuda").view(1, S)).clamp(0, S - 1)
    rel = torch.gather(q @ kr[0], 3, idx.expand(Bt, Hh, S, S))
    sc = ((q @ k.transpose(2, 3)) + rel) / math.sqrt(dk)
    sc = sc.masked_fill(~tri, float("-inf"))
    P = torch.softmax(sc, dim=-1)
    ref = ((P * keep / (1 - p)) @ v).transpose(1, 2).reshape(1, Bt * S, d)
    close(out, ref, 2e-2, 2e-2, "dropout attention out")
    dout = bf(torch.randn(G, Bt * S, d, device="cuda"))
    ref.backward(dout.float())
    for extra in (dict(out_f32=out32, rowsum=rowsum), dict()):  # both forms of the row statistic D (CpcAttn)
        dqkv = torch.empty_like(qkv)
        dkrel = torch.zeros(G, dk, S, device="cuda")
        K.attn_bwd(qkv, krel, out, probs, dout, dqkv, torch.empty_like(probs), dkrel, G, Bt, Hh, S, dk, rng=rng, site=3,
                   drop_p=p, **extra)
        assert cos(dqkv, qr.grad) > 0.999, cos(dqkv, qr.grad)
        assert cos(dkrel, kr.grad) > 0.999, cos(dkrel, kr.grad)
        close(dqkv, qr.grad, 5e-2, 3e-2 * qr.grad.abs().max().item(), "dropout attention dqkv")


def test_attention_dropout_consistent(K):
    torch.manual_seed(10)
    G, Bt, Hh, S, dk = 1, 2, 8, 32, 8
    d = Hh * dk
    qkv = bf(torch.randn(G, Bt * S, 3 * d, device="cuda"))
    rng = torch.tensor([99, 3], device="cuda", dtype=torch.int64)
    out0 = torch.empty(G, Bt * S, d, device="cuda", dtype=BF)
    out1 = torch.empty_like(out0)
    probs = torch.zeros(G, Bt, Hh, S, S, device="cuda", dtype=BF)
    K.attn_fwd(qkv, None, out0, probs, G, Bt, Hh, S, dk)
    K.attn_fwd(qkv, None, out1, probs, G, Bt, Hh, S, dk, rng=rng, site=1, drop_p=0.5)
    assert not torch.equal(out0, out1)
    out2 = torch.empty_like(out0)
    K.attn_fwd(qkv, None, out2, probs, G, Bt, Hh, S, dk, rng=rng, site=1, drop_p=0.5)
    assert torch.equal(out1, out2)  # same (seed, step, site) -> same mask
    # first row attends only to itself: either kept (scaled by 2) or dropped
    v0 = qkv.view(Bt, S, 3, Hh, dk)[:, 0, 2].float()
    o = out1.view(Bt, S, Hh, dk)[:, 0].float()
    ok = ((o - 2 * v0).abs().amax(-1) < 0.1) | (o.abs().amax(-1) == 0)
    assert ok.all()


@pytest.mark.gpu
def test_allreduce_mc_matches_nccl_on_two_gpus():
    """cpc_allreduce_mc (NVSwitch multimem all-reduce, in place on a symmetric buffer) against NCCL AVG over bucket
    sizes of the step (tools/allreduce_bench.py --check-only under torchrun).  Needs two GPUs of one NVSwitch box."""
    import subprocess
    import sys
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29641",
                          os.path.join(root, "tools", "allreduce_bench.py"), "--check-only"],
                         capture_output=True, text=True, timeout=300, cwd=root)
    text = out.stdout + out.stderr
    if "multicast unavailable" in text:
        pytest.skip("no NVLink multicast on this box")
    assert out.returncode == 0 and "ALLREDUCE_MC OK" in text, text[-2000:]
