"""ctypes binding of libcpc_b200.so (the C ABI declared in include/cpc_b200.h).

There is NO fallback: if the shared library is missing or a call fails the op raises.
`build()` compiles the library in-tree with nvcc for sm_100a (cross-compiles without a GPU).
"""
import ctypes
import glob
import os
import subprocess

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(_HERE, "libcpc_b200.so")


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def needs_build():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "*.h")) + \
        [os.path.join(os.path.dirname(_HERE), "include", "cpc_b200.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """nvcc -gencode arch=compute_100a,code=sm_100a ... -> cpc2_b200/libcpc_b200.so (one object per .cu, parallel)."""
    if not force and not needs_build():
        return LIB_PATH
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objdir = os.path.join(_HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    procs = []
    objs = []
    for src in sources():
        obj = os.path.join(objdir, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
               "-Xcompiler", "-fPIC", "-c", src, "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for src, p in procs:
        out, _ = p.communicate()
        if verbose and out:
            print(out)
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{out}")
    subprocess.check_call([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "--shared", "-o", LIB_PATH]
                          + objs + ["-lcudart"])
    return LIB_PATH


class GemmDesc(ctypes.Structure):
    _fields_ = [
        ("M", ctypes.c_int), ("N", ctypes.c_int), ("K", ctypes.c_int),
        ("groups", ctypes.c_int), ("k_per_slab", ctypes.c_int), ("k_splits", ctypes.c_int),
        ("force_bn", ctypes.c_int),
        ("a", ctypes.c_void_p), ("a_major", ctypes.c_int),
        ("a_dim0", ctypes.c_int64), ("a_dim1", ctypes.c_int64), ("a_dim2", ctypes.c_int64),
        ("a_stride1", ctypes.c_int64), ("a_stride2", ctypes.c_int64),
        ("a_row_off", ctypes.c_int), ("a_gmul", ctypes.c_int),
        ("b", ctypes.c_void_p), ("b_major", ctypes.c_int),
        ("b_dim0", ctypes.c_int64), ("b_dim1", ctypes.c_int64), ("b_dim2", ctypes.c_int64),
        ("b_stride1", ctypes.c_int64), ("b_stride2", ctypes.c_int64),
        ("b_row_off", ctypes.c_int), ("b_gmul", ctypes.c_int),
        ("alpha", ctypes.c_float),
        ("out_f32", ctypes.c_void_p), ("ldd_f32", ctypes.c_int64), ("gs_f32", ctypes.c_int64),
        ("out_bf16", ctypes.c_void_p), ("ldd_bf16", ctypes.c_int64), ("gs_bf16", ctypes.c_int64),
        ("bias", ctypes.c_void_p), ("gs_bias", ctypes.c_int64),
        ("res", ctypes.c_void_p), ("ld_res", ctypes.c_int64), ("gs_res", ctypes.c_int64),
        ("gate", ctypes.c_void_p), ("ld_gate", ctypes.c_int64), ("gs_gate", ctypes.c_int64),
        ("gate_scale", ctypes.c_float),
        ("rng", ctypes.c_void_p), ("rng_site", ctypes.c_uint32), ("drop_p", ctypes.c_float),
        ("relu", ctypes.c_int), ("accumulate", ctypes.c_int),
        ("norm_gamma", ctypes.c_void_p), ("norm_beta", ctypes.c_void_p), ("gs_norm_param", ctypes.c_int64),
        ("norm_bf16", ctypes.c_void_p), ("ld_norm_bf16", ctypes.c_int64), ("gs_norm_bf16", ctypes.c_int64),
        ("norm_f32", ctypes.c_void_p), ("ld_norm_f32", ctypes.c_int64), ("gs_norm_f32", ctypes.c_int64),
        ("norm_mean", ctypes.c_void_p), ("norm_rstd", ctypes.c_void_p), ("gs_norm_stat", ctypes.c_int64),
        ("norm_eps", ctypes.c_float), ("norm_unbiased", ctypes.c_int), ("norm_halo", ctypes.c_int),
        ("a_slab_mul", ctypes.c_int), ("b_slab_mul", ctypes.c_int),
    ]


_lib = None


def lib():
    """Loads the library (fails loudly when it has not been built)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(cpc2_b200 has no CPU / eager fallback)")
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.cpc_strerror.restype = ctypes.c_char_p
        _lib.cpc_strerror.argtypes = [ctypes.c_int]
    return _lib


def check(status, what=""):
    if status != 0:
        msg = lib().cpc_strerror(int(status)).decode()
        raise RuntimeError(f"libcpc_b200 {what}: {msg} (status {status})")


def stream_ptr():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


class RowMap(ctypes.Structure):
    _fields_ = [("seq_len", ctypes.c_int64), ("seq_stride", ctypes.c_int64), ("row_off", ctypes.c_int64)]


class RowNormFwd(ctypes.Structure):
    _fields_ = [
        ("rows", ctypes.c_int64), ("C", ctypes.c_int), ("rows_per_group", ctypes.c_int64),
        ("inp", ctypes.c_void_p), ("in_map", RowMap),
        ("gamma", ctypes.c_void_p), ("beta", ctypes.c_void_p),
        ("out_bf16", ctypes.c_void_p), ("out_f32", ctypes.c_void_p), ("out_map", RowMap),
        ("mean", ctypes.c_void_p), ("rstd", ctypes.c_void_p),
        ("eps", ctypes.c_float), ("unbiased", ctypes.c_int), ("relu", ctypes.c_int),
    ]


class RowNormBwd(ctypes.Structure):
    _fields_ = [
        ("rows", ctypes.c_int64), ("C", ctypes.c_int), ("rows_per_group", ctypes.c_int64),
        ("dout", ctypes.c_void_p), ("dout_f32", ctypes.c_int), ("dout_map", RowMap),
        ("dout2", ctypes.c_void_p), ("dout2_map", RowMap),
        ("inp", ctypes.c_void_p), ("in_map", RowMap),
        ("mean", ctypes.c_void_p), ("rstd", ctypes.c_void_p), ("gamma", ctypes.c_void_p), ("beta", ctypes.c_void_p),
        ("dx_bf16", ctypes.c_void_p), ("dx_f32", ctypes.c_void_p), ("dx_map", RowMap), ("dx_accumulate", ctypes.c_int),
        ("dgamma", ctypes.c_void_p), ("dbeta", ctypes.c_void_p), ("dxsum", ctypes.c_void_p),
        ("unbiased", ctypes.c_int), ("relu", ctypes.c_int),
    ]


class Conv0(ctypes.Structure):
    _fields_ = [
        ("N", ctypes.c_int), ("W", ctypes.c_int), ("C", ctypes.c_int), ("halo", ctypes.c_int), ("eps", ctypes.c_float),
        ("x", ctypes.c_void_p), ("w", ctypes.c_void_p), ("b", ctypes.c_void_p), ("gamma", ctypes.c_void_p),
        ("beta", ctypes.c_void_p),
        ("out", ctypes.c_void_p), ("mean", ctypes.c_void_p), ("rstd", ctypes.c_void_p),
        ("dout", ctypes.c_void_p), ("dw", ctypes.c_void_p), ("db", ctypes.c_void_p), ("dgamma", ctypes.c_void_p),
        ("dbeta", ctypes.c_void_p),
        ("x_stride", ctypes.c_int64),
    ]


class ColNorm(ctypes.Structure):
    _fields_ = [
        ("rows", ctypes.c_int64), ("C", ctypes.c_int), ("rows_per_group", ctypes.c_int64),
        ("mode", ctypes.c_int), ("relu", ctypes.c_int),
        ("inp", ctypes.c_void_p), ("gamma", ctypes.c_void_p), ("beta", ctypes.c_void_p),
        ("mean", ctypes.c_void_p), ("rstd", ctypes.c_void_p),
        ("running_mean", ctypes.c_void_p), ("running_var", ctypes.c_void_p), ("momentum", ctypes.c_float),
        ("eps", ctypes.c_float),
        ("partial", ctypes.c_void_p),
        ("out_bf16", ctypes.c_void_p), ("out_map", RowMap), ("out_f32", ctypes.c_void_p),
        ("dout", ctypes.c_void_p), ("dout_f32", ctypes.c_int), ("dout_map", RowMap),
        ("sums", ctypes.c_void_p),
        ("dx_bf16", ctypes.c_void_p), ("dx_map", RowMap), ("dx_f32", ctypes.c_void_p),
        ("dgamma", ctypes.c_void_p), ("dbeta", ctypes.c_void_p), ("dxsum", ctypes.c_void_p),
    ]


COLNORM_ID, COLNORM_STATS, COLNORM_GIVEN = 0, 1, 2
COLNORM_CHUNK_ROWS = 256


class Nce(ctypes.Structure):
    _fields_ = [
        ("K", ctypes.c_int), ("B", ctypes.c_int), ("W", ctypes.c_int), ("S", ctypes.c_int), ("NEG", ctypes.c_int),
        ("ld", ctypes.c_int64),
        ("logits", ctypes.c_void_p), ("ext", ctypes.c_void_p), ("weights", ctypes.c_void_p),
        ("lse", ctypes.c_void_p), ("loss_sum", ctypes.c_void_p), ("hits", ctypes.c_void_p),
        ("gscale", ctypes.c_void_p), ("dlogits", ctypes.c_void_p), ("dlogits_lo", ctypes.c_void_p),
        ("out_scale", ctypes.c_float),
    ]


class Lstm(ctypes.Structure):
    _fields_ = [
        ("N", ctypes.c_int), ("S", ctypes.c_int), ("H", ctypes.c_int), ("cell", ctypes.c_int),
        ("gx", ctypes.c_void_p), ("w_hh", ctypes.c_void_p), ("h0", ctypes.c_void_p), ("c0", ctypes.c_void_p),
        ("h_pad", ctypes.c_void_p), ("out", ctypes.c_void_p), ("gates", ctypes.c_void_p), ("c_all", ctypes.c_void_p),
        ("h_n", ctypes.c_void_p), ("c_n", ctypes.c_void_p),
        ("dout", ctypes.c_void_p), ("dgates", ctypes.c_void_p),
    ]


class Attn(ctypes.Structure):
    _fields_ = [
        ("G", ctypes.c_int), ("Bt", ctypes.c_int), ("Hh", ctypes.c_int), ("S", ctypes.c_int), ("dk", ctypes.c_int),
        ("scale", ctypes.c_float),
        ("qkv", ctypes.c_void_p), ("gs_qkv", ctypes.c_int64), ("krel", ctypes.c_void_p),
        ("out", ctypes.c_void_p), ("gs_out", ctypes.c_int64), ("probs", ctypes.c_void_p),
        ("rng", ctypes.c_void_p), ("rng_site", ctypes.c_uint32), ("drop_p", ctypes.c_float),
        ("dout", ctypes.c_void_p), ("dqkv", ctypes.c_void_p), ("dscore", ctypes.c_void_p), ("dkrel", ctypes.c_void_p),
        ("out_f32", ctypes.c_void_p),
        ("rowsum", ctypes.c_void_p),
    ]


class AdamSpan(ctypes.Structure):
    _fields_ = [
        ("p", ctypes.c_void_p), ("g", ctypes.c_void_p), ("m", ctypes.c_void_p), ("v", ctypes.c_void_p),
        ("p16", ctypes.c_void_p),
        ("n", ctypes.c_int64), ("first_block", ctypes.c_int64),
    ]


# every symbol include/cpc_b200.h declares (checked by tests/test_abi.py)
EXPORTED = [
    "cpc_strerror", "cpc_version", "cpc_device_arch", "cpc_set_pdl", "cpc_gemm_bf16", "cpc_rownorm_fwd", "cpc_rownorm_bwd",
    "cpc_conv0_cn_relu_fwd", "cpc_conv0_cn_relu_bwd", "cpc_colnorm_fwd", "cpc_colnorm_bwd", "cpc_conv0_plain_fwd",
    "cpc_conv0_plain_bwd", "cpc_conv_weight_pack", "cpc_conv_wgrad_unpack",
    "cpc_cast_f32_to_bf16", "cpc_cast_bf16_to_f32", "cpc_colsum", "cpc_groupsum", "cpc_neg_ext_index",
    "cpc_nce_fwd", "cpc_nce_bwd", "cpc_lstm_fwd", "cpc_lstm_bwd", "cpc_attn_fwd", "cpc_attn_bwd",
    "cpc_attn_bwd_noalloc",
    "cpc_adam_step", "cpc_adam_chunk", "cpc_dropout", "cpc_allreduce_mc", "cpc_allreduce_pad_bytes",
]
