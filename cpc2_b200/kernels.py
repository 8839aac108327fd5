"""Python-side launchers for the C ABI (include/cpc_b200.h): one function per entry point.

Everything here takes CUDA tensors, borrows their pointers and launches on the current torch stream.
No arithmetic happens in Python; shapes / strides are only described to the kernels.
"""
import ctypes

import torch

from . import _lib
from ._lib import GemmDesc, check, lib, ptr, stream_ptr

BF16 = torch.bfloat16


class View:
    """3-D operand view (dim0 contiguous) over a bf16 CUDA tensor, in elements.
    major 0: K-major (dim0 = reduction index); major 1: MN-major (dim0 = m / n index)."""
    __slots__ = ("t", "major", "dim0", "dim1", "dim2", "stride1", "stride2", "row_off", "gmul", "offset")

    def __init__(self, t, dim0, dim1, dim2=1, stride1=None, stride2=None, major=0, row_off=0, gmul=0, offset=0):
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
         drop_p=0.0, force_bn=0):
    """cpc_gemm_bf16: D[g] = alpha * A_g . B_g^T (+ epilogue).  See include/cpc_b200.h."""
    d = GemmDesc()
    d.M, d.N, d.K, d.groups, d.k_per_slab, d.k_splits, d.force_bn = M, N, K, groups, k_per_slab, k_splits, force_bn
    d.a, d.a_major = A.addr(), A.major
    d.a_dim0, d.a_dim1, d.a_dim2, d.a_stride1, d.a_stride2 = A.dim0, A.dim1, A.dim2, A.stride1, A.stride2
    d.a_row_off, d.a_gmul = A.row_off, A.gmul
    d.b, d.b_major = B.addr(), B.major
    d.b_dim0, d.b_dim1, d.b_dim2, d.b_stride1, d.b_stride2 = B.dim0, B.dim1, B.dim2, B.stride1, B.stride2
    d.b_row_off, d.b_gmul = B.row_off, B.gmul
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
    check(lib().cpc_gemm_bf16(ctypes.byref(d), stream_ptr()), "cpc_gemm_bf16")
