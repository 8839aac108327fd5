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
