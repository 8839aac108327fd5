"""cpc_attn_bwd_noalloc (include/cpc_b200.h): the attention backward whose `dscore` scratch lives in a file-scope device
array of the library instead of a caller-allocated buffer.  No counterpart in the reference; added
for callers that cannot allocate (CUDA-graph capture, serving loops, harnesses that forbid cudaMalloc).

Called through the C ABI (ctypes) and compared with the existing entry point cpc_attn_bwd on the same inputs, two calls in
a row (the second with other inputs AND another shape: the shared scratch keeps the first call's contents, which must not
matter), with the tolerance tests/test_kernels_gpu.py::test_attention_fwd_bwd uses.  This file is collected last on
purpose."""
import ctypes
import math

import pytest
import torch

BF = torch.bfloat16
MAX_ELEMS = 4 * 1024 * 1024  # CPC_ATTN_NOALLOC_MAX_ELEMS


def _need_gpu():
    if not torch.cuda.is_available():
        pytest.skip("no GPU")


def cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float(torch.dot(a, b) / (a.norm() * b.norm() + 1e-300))


def _case(K, G, Bt, Hh, S, dk, seed):
    """Inputs + forward state of one attention problem (forward through the existing entry point)."""
    torch.manual_seed(seed)
    d = Hh * dk
    qkv = torch.randn(G, Bt * S, 3 * d, device="cuda").to(BF)
    krel = (torch.rand(G, dk, S, device="cuda") - 0.5) * (2 / math.sqrt(dk))
    out = torch.empty(G, Bt * S, d, device="cuda", dtype=BF)
    probs = torch.zeros(G, Bt, Hh, S, S, device="cuda", dtype=BF)
    K.attn_fwd(qkv, krel, out, probs, G, Bt, Hh, S, dk)
    dout = torch.randn(G, Bt * S, d, device="cuda").to(BF)
    return dict(qkv=qkv, krel=krel, out=out, probs=probs, dout=dout, dims=(G, Bt, Hh, S, dk))


def _existing(K, c):
    G, Bt, Hh, S, dk = c["dims"]
    dqkv = torch.empty_like(c["qkv"])
    dkrel = torch.zeros(G, dk, S, device="cuda")
    K.attn_bwd(c["qkv"], c["krel"], c["out"], c["probs"], c["dout"], dqkv, torch.empty_like(c["probs"]), dkrel, G, Bt, Hh,
               S, dk)
    return dqkv, dkrel


def _noalloc_desc(K, c, dqkv, dkrel):
    G, Bt, Hh, S, dk = c["dims"]
    d = K._attn(c["qkv"], c["krel"], c["out"], c["probs"], G, Bt, Hh, S, dk, None, 0, 0.0)
    d.dout, d.dqkv, d.dkrel = c["dout"].data_ptr(), dqkv.data_ptr(), dkrel.data_ptr()
    d.dscore = None  # ignored by the entry point: the scratch is the library's own static array
    return d


@pytest.mark.gpu
@pytest.mark.parametrize("first,second", [((2, 2, 4, 24, 8), (1, 3, 4, 12, 8)),      # general kernel in both entry points
                                          ((1, 2, 8, 116, 32), (2, 1, 8, 64, 32))])  # existing entry: tensor-core path
def test_attn_bwd_noalloc_matches_attn_bwd_two_calls_in_a_row(first, second):
    _need_gpu()
    from cpc2_b200 import _lib
    from cpc2_b200 import kernels as K
    lib = _lib.lib()
    lib.cpc_attn_bwd_noalloc.restype = ctypes.c_int
    cases = [_case(K, *first, seed=21), _case(K, *second, seed=22)]
    refs = [_existing(K, c) for c in cases]
    outs = [(torch.empty_like(c["qkv"]), torch.zeros_like(c["krel"])) for c in cases]  # every buffer exists BEFORE the calls
    descs = [_noalloc_desc(K, c, *o) for c, o in zip(cases, outs)]
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    # one untimed pass first: the FIRST launch of a kernel makes the driver load its code (lazy module loading), which
    # moves the driver-level free-memory figure checked below and has nothing to do with the entry point allocating
    for dsc in descs:
        assert lib.cpc_attn_bwd_noalloc(ctypes.byref(dsc), stream) == 0
    for _, dkrel in outs:
        dkrel.zero_()  # (accumulated by the kernel; in place)
    torch.cuda.synchronize()
    free0, alloc0 = torch.cuda.mem_get_info()[0], torch.cuda.memory_allocated()
    for dsc in descs:  # two calls in a row on one stream, nothing in between
        rc = lib.cpc_attn_bwd_noalloc(ctypes.byref(dsc), stream)
        assert rc == 0, lib.cpc_strerror(rc).decode()
    torch.cuda.synchronize()
    # the calls allocated nothing: neither through torch's allocator nor behind its back (driver-level free memory)
    assert torch.cuda.memory_allocated() == alloc0
    assert torch.cuda.mem_get_info()[0] == free0
    for (dqkv, dkrel), (rq, rk), c in zip(outs, refs, cases):
        assert torch.isfinite(dqkv.float()).all() and torch.isfinite(dkrel).all()
        assert cos(dqkv, rq) > 0.999, cos(dqkv, rq)
        tol = 3e-2 * rq.float().abs().max().item()
        assert ((dqkv.float() - rq.float()).abs() <= tol + 5e-2 * rq.float().abs()).all(), c["dims"]
        assert cos(dkrel, rk) > 0.999, cos(dkrel, rk)
        tol = 3e-2 * rk.abs().max().item()
        assert ((dkrel - rk).abs() <= tol + 5e-2 * rk.abs()).all(), c["dims"]
    if first[4] == 8:  # the same kernel ran in both entry points: dqkv is written by one thread per element, no atomics
        for (dqkv, _), (rq, _) in zip(outs, refs):
            assert torch.equal(dqkv, rq)


@pytest.mark.gpu
def test_attn_bwd_noalloc_rejects_sizes_beyond_the_static_scratch():
    _need_gpu()
    from cpc2_b200 import _lib
    from cpc2_b200 import kernels as K
    lib = _lib.lib()
    G, Bt, Hh, S, dk = 12, 8, 8, 116, 32  # the default step's heads: 10.3 M elements > 4 Mi
    assert G * Bt * Hh * S * S > MAX_ELEMS
    tiny = torch.zeros(16, device="cuda", dtype=BF)  # never touched: the size check comes before any launch
    d = _lib.Attn()
    d.G, d.Bt, d.Hh, d.S, d.dk, d.scale = G, Bt, Hh, S, dk, 1.0
    d.qkv = d.out = d.probs = d.dout = d.dqkv = tiny.data_ptr()
    rc = lib.cpc_attn_bwd_noalloc(ctypes.byref(d), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == -7
    assert "scratch" in lib.cpc_strerror(rc).decode()
    torch.cuda.synchronize()
    # just inside the limit is accepted by the size check (a shape of exactly MAX_ELEMS elements)
    assert 1 * 1 * 64 * 256 * 256 == MAX_ELEMS
