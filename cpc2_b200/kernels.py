"""Python-side launchers for the C ABI (include/cpc_b200.h): one function per entry point.

Everything here takes CUDA tensors, borrows their pointers and launches on the current torch stream.
No arithmetic happens in Python; shapes / strides are only described to the kernels.
"""
import ctypes

import torch

from . import _lib
from ._lib import GemmDesc, check, lib, ptr, stream_ptr

BF16 = torch.bfloat16


# ---- launch accounting / optional per-call profiling (used by bench.py) -------------------------------
LAUNCHES = 0        # C-ABI launches issued by this process (bench.py reports the per-step count)
PROFILE = None      # when a list: [name, work dict, start event, end event] per call
_WORK = {}          # work model of the call being issued (set by the launcher just before _run)


def set_work(**kw):
    """Algorithmic work of the next call: flops=..., bytes=... (only read when profiling)."""
    global _WORK
    _WORK = kw


def _run(name, args):
    global LAUNCHES, _WORK
    LAUNCHES += 1
    if PROFILE is None:
        check(getattr(lib(), name)(*args), name)
        return
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    check(getattr(lib(), name)(*args), name)
    e1.record()
    PROFILE.append([name, _WORK, e0, e1])
    _WORK = {}


class View:
    """3-D operand view (dim0 contiguous) over a bf16 CUDA tensor, in elements.
    major 0: K-major (dim0 = reduction index); major 1: MN-major (dim0 = m / n index)."""
    __slots__ = ("t", "major", "dim0", "dim1", "dim2", "stride1", "stride2", "row_off", "gmul", "offset", "smul")

    def __init__(self, t, dim0, dim1, dim2=1, stride1=None, stride2=None, major=0, row_off=0, gmul=0, offset=0, smul=1):
        """smul: step of the dim2 coordinate per reduction slab (1 = consecutive planes, 0 = the same plane for every
        slab)."""
        self.smul = int(smul)
        assert t.dtype == BF16 and t.is_cuda
        self.t, self.major = t, major
        self.dim0, self.dim1, self.dim2 = int(dim0), int(dim1), int(dim2)
        self.stride1 = int(stride1 if stride1 is not None else dim0)
        self.stride2 = int(stride2 if stride2 is not None else self.stride1 * self.dim1)
        self.row_off, self.gmul, self.offset = int(row_off), int(gmul), int(offset)

    def addr(self):
        return self.t.data_ptr() + 2 * self.offset


def kmajor(t, groups_in_dim2=False, gmul=None):
    """[rows, k] or [g, rows, k] contiguous tensor as a K-major operand."""
    if t.dim() == 2:
        return View(t, t.shape[1], t.shape[0], 1, t.stride(0), None, 0, 0, 0 if gmul is None else gmul)
    assert t.dim() == 3
    return View(t, t.shape[2], t.shape[1], t.shape[0], t.stride(1), t.stride(0), 0, 0, 1 if gmul is None else gmul)


def mnmajor(t, gmul=None):
    """[k, mn] or [g, k, mn] contiguous tensor as an MN-major operand (reduction over rows)."""
    if t.dim() == 2:
        return View(t, t.shape[1], t.shape[0], 1, t.stride(0), None, 1, 0, 0 if gmul is None else gmul)
    assert t.dim() == 3
    return View(t, t.shape[2], t.shape[1], t.shape[0], t.stride(1), t.stride(0), 1, 0, 1 if gmul is None else gmul)


def _out(t):
    """(tensor, row stride, group stride) from a [rows, n] / [g, rows, n] tensor."""
    if t is None:
        return None, 0, 0
    if isinstance(t, tuple):
        return t
    if t.dim() == 2:
        return t, t.stride(0), 0
    return t, t.stride(1), t.stride(0)


def gemm(A, B, M, N, K, groups=1, k_per_slab=0, k_splits=1, alpha=1.0, out_f32=None, out_bf16=None, bias=None,
         bias_gs=0, res=None, gate=None, gate_scale=1.0, relu=False, accumulate=False, rng=None, rng_site=0,
         drop_p=0.0, force_bn=0, algo_flops=None, norm=None):
    """cpc_gemm_bf16: D[g] = alpha * A_g . B_g^T (+ epilogue).  See include/cpc_b200.h.
    algo_flops: ALGORITHMIC work of the call when it differs from the executed 2*M*N*K*groups (the all-pairs logits
    products compute every (token, frame) pair where the algorithm needs NEG + 1 of them); profiling only."""
    d = GemmDesc()
    d.M, d.N, d.K, d.groups, d.k_per_slab, d.k_splits, d.force_bn = M, N, K, groups, k_per_slab, k_splits, force_bn
    d.a, d.a_major = A.addr(), A.major
    d.a_dim0, d.a_dim1, d.a_dim2, d.a_stride1, d.a_stride2 = A.dim0, A.dim1, A.dim2, A.stride1, A.stride2
    d.a_row_off, d.a_gmul, d.a_slab_mul = A.row_off, A.gmul, (-1 if A.smul == 0 else A.smul)
    d.b, d.b_major = B.addr(), B.major
    d.b_dim0, d.b_dim1, d.b_dim2, d.b_stride1, d.b_stride2 = B.dim0, B.dim1, B.dim2, B.stride1, B.stride2
    d.b_row_off, d.b_gmul, d.b_slab_mul = B.row_off, B.gmul, (-1 if B.smul == 0 else B.smul)
    d.alpha = alpha
    t, ld, gs = _out(out_f32)
    if t is not None:
        assert t.dtype == torch.float32
        d.out_f32, d.ldd_f32, d.gs_f32 = t.data_ptr(), ld, gs
    t, ld, gs = _out(out_bf16)
    if t is not None:
        assert t.dtype == BF16
        d.out_bf16, d.ldd_bf16, d.gs_bf16 = t.data_ptr(), ld, gs
    if bias is not None:
        assert bias.dtype == torch.float32
        d.bias, d.gs_bias = bias.data_ptr(), bias_gs
    t, ld, gs = _out(res)
    if t is not None:
        assert t.dtype == BF16
        d.res, d.ld_res, d.gs_res = t.data_ptr(), ld, gs
    t, ld, gs = _out(gate)
    if t is not None:
        assert t.dtype == BF16
        d.gate, d.ld_gate, d.gs_gate = t.data_ptr(), ld, gs
    d.gate_scale = gate_scale
    if rng is not None and drop_p > 0:
        d.rng, d.rng_site, d.drop_p = rng.data_ptr(), rng_site, drop_p
    d.relu, d.accumulate = int(relu), int(accumulate)
    if norm is not None:  # fused ChannelNorm / LayerNorm of the output rows (N == 256), see CpcGemmDesc
        d.norm_gamma, d.norm_beta, d.gs_norm_param = norm["gamma"].data_ptr(), norm["beta"].data_ptr(), norm.get("gs_param", 0)
        t, ld, gs = _out(norm.get("out_bf16"))
        if t is not None:
            assert t.dtype == BF16
            d.norm_bf16, d.ld_norm_bf16, d.gs_norm_bf16 = t.data_ptr() + 2 * norm.get("offset_bf16", 0), ld, gs
        t, ld, gs = _out(norm.get("out_f32"))
        if t is not None:
            assert t.dtype == torch.float32
            d.norm_f32, d.ld_norm_f32, d.gs_norm_f32 = t.data_ptr(), ld, gs
        if norm.get("mean") is not None:
            d.norm_mean, d.norm_rstd, d.gs_norm_stat = norm["mean"].data_ptr(), norm["rstd"].data_ptr(), norm.get("gs_stat", 0)
        d.norm_eps, d.norm_unbiased, d.norm_halo = norm.get("eps", 1e-5), int(norm.get("unbiased", False)), norm.get("halo", 0)
    if PROFILE is not None:
        ob = (4 if out_f32 is not None else 0) + (2 if out_bf16 is not None else 0) + (2 if norm is not None else 0)
        set_work(kind="tensor", flops=2.0 * M * N * K * groups if algo_flops is None else float(algo_flops),
                 executed_flops=2.0 * M * N * K * groups, bytes=2.0 * groups * (M * K + N * K) + ob * M * N * groups,
                 shape=f"g{groups} {M}x{N}x{K} a{A.major}b{B.major} s{k_splits}")
    _run("cpc_gemm_bf16", (ctypes.byref(d), stream_ptr(),))


# --------------------------------------------------------------------------------------------------
def _dp(t):
    return t.data_ptr() if t is not None else None


def rowmap(seq_len=0, seq_stride=0, row_off=0):
    return _lib.RowMap(int(seq_len), int(seq_stride), int(row_off))


def rownorm_fwd(inp, rows, C, gamma=None, beta=None, out_bf16=None, out_f32=None, mean=None, rstd=None,
                in_map=None, out_map=None, rows_per_group=0, eps=1e-5, unbiased=False, relu=False):
    d = _lib.RowNormFwd()
    d.rows, d.C, d.rows_per_group = rows, C, rows_per_group
    d.inp, d.in_map = _dp(inp), in_map or rowmap()
    d.gamma, d.beta = _dp(gamma), _dp(beta)
    d.out_bf16, d.out_f32, d.out_map = _dp(out_bf16), _dp(out_f32), out_map or rowmap()
    d.mean, d.rstd = _dp(mean), _dp(rstd)
    d.eps, d.unbiased, d.relu = eps, int(unbiased), int(relu)
    if PROFILE is not None:
        set_work(kind="hbm", bytes=float(rows) * C * (2 + (2 if out_bf16 is not None else 0) + (4 if out_f32 is not None else 0)))
    _run("cpc_rownorm_fwd", (ctypes.byref(d), stream_ptr(),))


def rownorm_bwd(dout, inp, mean, rstd, rows, C, gamma=None, beta=None, dx_bf16=None, dx_f32=None, dgamma=None,
                dbeta=None, dxsum=None, dout2=None, dout_map=None, dout2_map=None, in_map=None, dx_map=None,
                rows_per_group=0, unbiased=False, relu=False, dx_accumulate=False):
    d = _lib.RowNormBwd()
    d.rows, d.C, d.rows_per_group = rows, C, rows_per_group
    d.dout, d.dout_f32, d.dout_map = _dp(dout), int(dout.dtype == torch.float32), dout_map or rowmap()
    d.dout2, d.dout2_map = _dp(dout2), dout2_map or rowmap()
    d.inp, d.in_map = _dp(inp), in_map or rowmap()
    d.mean, d.rstd, d.gamma, d.beta = _dp(mean), _dp(rstd), _dp(gamma), _dp(beta)
    d.dx_bf16, d.dx_f32, d.dx_map, d.dx_accumulate = _dp(dx_bf16), _dp(dx_f32), dx_map or rowmap(), int(dx_accumulate)
    d.dgamma, d.dbeta, d.dxsum = _dp(dgamma), _dp(dbeta), _dp(dxsum)
    d.unbiased, d.relu = int(unbiased), int(relu)
    if PROFILE is not None:
        set_work(kind="hbm", bytes=float(rows) * C * (2 + (4 if dout.dtype == torch.float32 else 2)
                                                      + (2 if dx_bf16 is not None else 0) + (4 if dx_f32 is not None else 0)))
    _run("cpc_rownorm_bwd", (ctypes.byref(d), stream_ptr(),))


def conv0_fwd(x, w, b, gamma, beta, out, mean, rstd, N, W, C, halo, eps=1e-5, x_stride=0):
    d = _lib.Conv0()
    d.N, d.W, d.C, d.halo, d.eps, d.x_stride = N, W, C, halo, eps, x_stride
    d.x, d.w, d.b, d.gamma, d.beta = _dp(x), _dp(w), _dp(b), _dp(gamma), _dp(beta)
    d.out, d.mean, d.rstd = _dp(out), _dp(mean), _dp(rstd)
    if PROFILE is not None:
        set_work(kind="hbm", bytes=float(N) * (W // 5) * C * 2 + 4.0 * N * W)
    _run("cpc_conv0_cn_relu_fwd", (ctypes.byref(d), stream_ptr(),))


def conv0_bwd(x, w, b, gamma, beta, mean, rstd, dout, dw, db, dgamma, dbeta, N, W, C, halo, eps=1e-5, x_stride=0):
    d = _lib.Conv0()
    d.N, d.W, d.C, d.halo, d.eps, d.x_stride = N, W, C, halo, eps, x_stride
    d.x, d.w, d.b, d.gamma, d.beta = _dp(x), _dp(w), _dp(b), _dp(gamma), _dp(beta)
    d.mean, d.rstd = _dp(mean), _dp(rstd)
    d.dout, d.dw, d.db, d.dgamma, d.dbeta = _dp(dout), _dp(dw), _dp(db), _dp(dgamma), _dp(dbeta)
    if PROFILE is not None:
        set_work(kind="hbm", bytes=float(N) * (W // 5) * C * 2 + 4.0 * N * W)
    _run("cpc_conv0_cn_relu_bwd", (ctypes.byref(d), stream_ptr(),))


COLNORM_MODES = {"ID": _lib.COLNORM_ID, "stats": _lib.COLNORM_STATS, "given": _lib.COLNORM_GIVEN}


def colnorm_scratch(rows, C, rows_per_group, device):
    """(partial, sums) scratch of cpc_colnorm_fwd / _bwd (include/cpc_b200.h)."""
    rpg = rows_per_group or rows
    groups = rows // rpg
    chunks = (rpg + _lib.COLNORM_CHUNK_ROWS - 1) // _lib.COLNORM_CHUNK_ROWS
    return (torch.empty(2 * groups * chunks * C, device=device, dtype=torch.float32),
            torch.empty(2 * groups * C, device=device, dtype=torch.float32))


def _colnorm_desc(inp, rows, C, rows_per_group, mode, relu, gamma, beta, mean, rstd, partial, eps):
    d = _lib.ColNorm()
    d.rows, d.C, d.rows_per_group = rows, C, rows_per_group
    d.mode, d.relu, d.eps = COLNORM_MODES[mode], int(relu), eps
    d.inp, d.gamma, d.beta, d.mean, d.rstd, d.partial = _dp(inp), _dp(gamma), _dp(beta), _dp(mean), _dp(rstd), _dp(partial)
    d.out_map, d.dout_map, d.dx_map = rowmap(), rowmap(), rowmap()
    return d


def colnorm_fwd(inp, rows, C, mode, rows_per_group=0, gamma=None, beta=None, mean=None, rstd=None, partial=None,
                running_mean=None, running_var=None, momentum=0.1, eps=1e-5, out_bf16=None, out_map=None, out_f32=None,
                relu=True):
    d = _colnorm_desc(inp, rows, C, rows_per_group, mode, relu, gamma, beta, mean, rstd, partial, eps)
    d.running_mean, d.running_var, d.momentum = _dp(running_mean), _dp(running_var), momentum
    d.out_bf16, d.out_f32, d.out_map = _dp(out_bf16), _dp(out_f32), out_map or rowmap()
    if PROFILE is not None:
        set_work(kind="hbm", bytes=float(rows) * C * ((4 if mode == "stats" else 2) + (2 if out_bf16 is not None else 0)
                                                      + (4 if out_f32 is not None else 0)))
    _run("cpc_colnorm_fwd", (ctypes.byref(d), stream_ptr(),))


def colnorm_bwd(dout, inp, rows, C, mode, partial, sums, rows_per_group=0, gamma=None, beta=None, mean=None, rstd=None,
                eps=1e-5, dout_map=None, dx_bf16=None, dx_map=None, dx_f32=None, dgamma=None, dbeta=None, dxsum=None,
                relu=True):
    d = _colnorm_desc(inp, rows, C, rows_per_group, mode, relu, gamma, beta, mean, rstd, partial, eps)
    d.dout, d.dout_f32, d.dout_map = _dp(dout), int(dout.dtype == torch.float32), dout_map or rowmap()
    d.sums = _dp(sums)
    d.dx_bf16, d.dx_f32, d.dx_map = _dp(dx_bf16), _dp(dx_f32), dx_map or rowmap()
    d.dgamma, d.dbeta, d.dxsum = _dp(dgamma), _dp(dbeta), _dp(dxsum)
    if PROFILE is not None:
        nb = 4 if dout.dtype == torch.float32 else 2
        set_work(kind="hbm", bytes=float(rows) * C * (2 * (2 + nb) + (2 if dx_bf16 is not None else 0)
                                                      + (4 if dx_f32 is not None else 0)))
    _run("cpc_colnorm_bwd", (ctypes.byref(d), stream_ptr(),))


def conv0_plain_fwd(x, w, b, out, N, W, C, x_stride=0):
    d = _lib.Conv0()
    d.N, d.W, d.C, d.x_stride = N, W, C, x_stride
    d.x, d.w, d.b, d.out = _dp(x), _dp(w), _dp(b), _dp(out)
    if PROFILE is not None:
        set_work(kind="hbm", bytes=float(N) * ((W - 4) // 5 + 1) * C * 2 + 4.0 * N * W)
    _run("cpc_conv0_plain_fwd", (ctypes.byref(d), stream_ptr(),))


def conv0_plain_bwd(x, dout, dw, N, W, C, x_stride=0):
    d = _lib.Conv0()
    d.N, d.W, d.C, d.x_stride = N, W, C, x_stride
    d.x, d.dout, d.dw = _dp(x), _dp(dout), _dp(dw)
    if PROFILE is not None:
        set_work(kind="hbm", bytes=float(N) * ((W - 4) // 5 + 1) * C * 4 + 4.0 * N * W)
    _run("cpc_conv0_plain_bwd", (ctypes.byref(d), stream_ptr(),))


def conv_weight_pack(w, wf, wd, Co, Ci, k, s):
    _run("cpc_conv_weight_pack", (ptr(w), ptr(wf), ptr(wd), Co, Ci, k, s, stream_ptr(),))


def conv_wgrad_unpack(dwf, dw, Co, Ci, k):
    _run("cpc_conv_wgrad_unpack", (ptr(dwf), ptr(dw), Co, Ci, k, stream_ptr(),))


def cast_bf16(src, dst=None):
    """fp32 -> bf16 copy of a contiguous tensor."""
    assert src.dtype == torch.float32 and src.is_contiguous()
    if dst is None:
        dst = torch.empty(src.shape, dtype=BF16, device=src.device)
    if PROFILE is not None:
        set_work(kind="hbm", bytes=6.0 * src.numel())
    _run("cpc_cast_f32_to_bf16", (ptr(src), ptr(dst), ctypes.c_int64(src.numel()), stream_ptr(),))
    return dst


def cast_f32(src, dst=None):
    assert src.dtype == BF16 and src.is_contiguous()
    if dst is None:
        dst = torch.empty(src.shape, dtype=torch.float32, device=src.device)
    _run("cpc_cast_bf16_to_f32", (ptr(src), ptr(dst), ctypes.c_int64(src.numel()), stream_ptr(),))
    return dst


def dropout_(x, rng, site, p):
    """In-place inverted dropout with the counter-hash mask of (rng, site); same call on the gradient = backward."""
    assert x.is_contiguous() and x.dtype in (BF16, torch.float32)
    _run("cpc_dropout", (ptr(x), int(x.dtype == BF16), ctypes.c_int64(x.numel()), ptr(rng), ctypes.c_uint32(site),
                         ctypes.c_float(p), stream_ptr(),))
    return x


def allreduce_mc(mc_addr, n, signal_pads_dev, rank, world, scale, max_blocks=0):
    """In-switch all-reduce (multimem.ld_reduce / multimem.st) of n floats at multicast address mc_addr, in place."""
    _run("cpc_allreduce_mc", (ctypes.c_void_p(mc_addr), ctypes.c_int64(n), ctypes.c_void_p(signal_pads_dev), rank, world,
                               ctypes.c_float(scale), max_blocks, stream_ptr(),))


def colsum(inp, out, rows, C, ld, gs_in=0, groups=1):
    _run("cpc_colsum", (ptr(inp), int(inp.dtype == BF16), ptr(out), ctypes.c_int64(rows), C, ctypes.c_int64(ld),
                           ctypes.c_int64(gs_in), groups, stream_ptr(),))


def groupsum(inp, n, groups, out_f32=None, out_bf16=None, accumulate=False):
    _run("cpc_groupsum", (ptr(inp), int(inp.dtype == BF16), ptr(out_f32), ptr(out_bf16), ctypes.c_int64(n), groups,
                             int(accumulate), stream_ptr(),))


def neg_ext_index(batch_idx, seq_idx, B, NEG, W, S, out=None):
    ext = torch.empty_like(batch_idx) if out is None else out
    _run("cpc_neg_ext_index", (ptr(batch_idx), ptr(seq_idx), ptr(ext), B, NEG, W, S, stream_ptr(),))
    return ext


def _nce(logits, ext, weights, lse, K, B, W, S, NEG, ld):
    d = _lib.Nce()
    d.K, d.B, d.W, d.S, d.NEG, d.ld = K, B, W, S, NEG, ld
    d.logits, d.ext, d.weights, d.lse = _dp(logits), _dp(ext), _dp(weights), _dp(lse)
    return d


def nce_fwd(logits, ext, weights, lse, loss_sum, hits, K, B, W, S, NEG, ld, out_scale=1.0):
    """loss_sum / hits (pre-zeroed) receive out_scale * their sums: out_scale = 1 / (B*W) yields the per-step mean
    loss and accuracy of criterion.py:350-357 directly."""
    d = _nce(logits, ext, weights, lse, K, B, W, S, NEG, ld)
    d.loss_sum, d.hits, d.out_scale = _dp(loss_sum), _dp(hits), out_scale
    if PROFILE is not None:
        set_work(kind="hbm", bytes=4.0 * K * B * W * (NEG + 1) + 8.0 * B * NEG * W)
    _run("cpc_nce_fwd", (ctypes.byref(d), stream_ptr(),))


def nce_bwd(logits, ext, weights, lse, gscale, dlogits, K, B, W, S, NEG, ld, dlogits_lo=None):
    d = _nce(logits, ext, weights, lse, K, B, W, S, NEG, ld)
    d.gscale, d.dlogits, d.dlogits_lo = _dp(gscale), _dp(dlogits), _dp(dlogits_lo)
    if PROFILE is not None:
        set_work(kind="hbm", bytes=(4.0 if dlogits_lo is not None else 2.0) * K * B * W * ld + 4.0 * K * B * W * (NEG + 1))
    _run("cpc_nce_bwd", (ctypes.byref(d), stream_ptr(),))


def adam_step(table, n_spans, blocks, state, advance=True, max_ctas=0):
    """cpc_adam_step: `table` = device bytes of a CpcAdamSpan array, `state` = device double[8]; max_ctas > 0: a
    throttled pass on that many persistent CTAs."""
    if PROFILE is not None:
        set_work(kind="hbm", bytes=28.0 * blocks * lib().cpc_adam_chunk())
    _run("cpc_adam_step", (ctypes.c_void_p(table.data_ptr()), n_spans, ctypes.c_int64(blocks),
                           ctypes.c_void_p(state.data_ptr()), int(advance), int(max_ctas), stream_ptr(),))


def attn_bwd_noalloc(qkv, krel, out, probs, dout, dqkv, dkrel, G, Bt, Hh, S, dk, rng=None, site=0, drop_p=0.0):
    """cpc_attn_bwd_noalloc: attn_bwd on the general kernel with the library's static scratch (no `dscore` buffer, no
    allocation; G*Bt*Hh*S*S <= 4 Mi elements; calls that overlap on the GPU race on the shared scratch)."""
    d = _attn(qkv, krel, out, probs, G, Bt, Hh, S, dk, rng, site, drop_p)
    d.dout, d.dqkv, d.dkrel = _dp(dout), _dp(dqkv), _dp(dkrel)
    _run("cpc_attn_bwd_noalloc", (ctypes.byref(d), stream_ptr(),))


CELL = {"LSTM": 0, "GRU": 1, "RNN": 2}


def lstm_fwd(gx, w_hh, N, S, H, h0=None, c0=None, h_pad=None, out=None, gates=None, c_all=None, h_n=None, c_n=None,
             cell="LSTM"):
    d = _lib.Lstm()
    d.N, d.S, d.H, d.cell = N, S, H, CELL[cell]
    d.gx, d.w_hh, d.h0, d.c0 = _dp(gx), _dp(w_hh), _dp(h0), _dp(c0)
    d.h_pad, d.out, d.gates, d.c_all, d.h_n, d.c_n = _dp(h_pad), _dp(out), _dp(gates), _dp(c_all), _dp(h_n), _dp(c_n)
    if PROFILE is not None:
        set_work(kind="latency", flops=2.0 * N * S * H * 4 * H, steps=S)
    _run("cpc_lstm_fwd", (ctypes.byref(d), stream_ptr(),))


def lstm_bwd(w_hh, gates, c_all, dout, dgates, N, S, H, c0=None, h0=None, cell="LSTM"):
    d = _lib.Lstm()
    d.N, d.S, d.H, d.cell = N, S, H, CELL[cell]
    d.h0 = _dp(h0)
    d.w_hh, d.c0, d.gates, d.c_all, d.dout, d.dgates = _dp(w_hh), _dp(c0), _dp(gates), _dp(c_all), _dp(dout), _dp(dgates)
    if PROFILE is not None:
        set_work(kind="latency", flops=2.0 * N * S * H * 4 * H, steps=S)
    _run("cpc_lstm_bwd", (ctypes.byref(d), stream_ptr(),))


def _attn(qkv, krel, out, probs, G, Bt, Hh, S, dk, rng, site, drop_p, out_f32=None, rowsum=None):
    d = _lib.Attn()
    if out_f32 is not None:
        assert out_f32.dtype == torch.float32 and out_f32.shape == out.shape
        d.out_f32 = out_f32.data_ptr()
    if rowsum is not None:
        assert rowsum.dtype == torch.float32 and rowsum.numel() == G * Bt * Hh * S
        d.rowsum = rowsum.data_ptr()
    d.G, d.Bt, d.Hh, d.S, d.dk = G, Bt, Hh, S, dk
    d.scale = 1.0 / (dk ** 0.5)
    d.qkv, d.gs_qkv, d.krel = _dp(qkv), Bt * S * 3 * Hh * dk, _dp(krel)
    d.out, d.gs_out, d.probs = _dp(out), Bt * S * Hh * dk, _dp(probs)
    if rng is not None and drop_p > 0:
        d.rng, d.rng_site, d.drop_p = _dp(rng), site, drop_p
    return d


def attn_fwd(qkv, krel, out, probs, G, Bt, Hh, S, dk, rng=None, site=0, drop_p=0.0, out_f32=None, rowsum=None):
    d = _attn(qkv, krel, out, probs, G, Bt, Hh, S, dk, rng, site, drop_p, out_f32, rowsum)
    if PROFILE is not None:
        set_work(kind="tensor", flops=3.0 * G * Bt * Hh * S * S * dk)  # (mma.sync tensor cores, causal: ~half executed)
    _run("cpc_attn_fwd", (ctypes.byref(d), stream_ptr(),))


def attn_bwd(qkv, krel, out, probs, dout, dqkv, dscore, dkrel, G, Bt, Hh, S, dk, rng=None, site=0, drop_p=0.0,
             out_f32=None, rowsum=None):
    d = _attn(qkv, krel, out, probs, G, Bt, Hh, S, dk, rng, site, drop_p, out_f32, rowsum)
    d.dout, d.dqkv, d.dscore, d.dkrel = _dp(dout), _dp(dqkv), _dp(dscore), _dp(dkrel)
    if PROFILE is not None:
        set_work(kind="tensor", flops=7.0 * G * Bt * Hh * S * S * dk)
    _run("cpc_attn_bwd", (ctypes.byref(d), stream_ptr(),))
