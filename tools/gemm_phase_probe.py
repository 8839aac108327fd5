"""Where does a GEMM launch spend its time?  Builds a second copy of the library with -DCPC_GEMM_TIMING (clock64 stamps
at the phase boundaries of every CTA, csrc/gemm.cu) and prints, per shape of the training step, the median over CTAs of

  setup      kernel entry -> barriers initialised, TMEM allocated, tensor maps prefetched
  1st data   -> first k-block has landed in shared memory (TMA latency)
  main loop  -> last MMA of the first tile issued + committed     (k-blocks, per-k-block time)
  acc wait   -> the epilogue sees the accumulator
  epilogue   -> first tile drained / stored
  tail       -> CTA exit

    python tools/gemm_phase_probe.py build      (in the build container: nvcc, no GPU needed)
    python tools/gemm_phase_probe.py            (on the GPU box)
"""
import ctypes
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
SO = os.path.join(ROOT, "cpc2_b200", "build", "libcpc_b200_timing.so")


def build():
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    srcs = sorted(glob.glob(os.path.join(ROOT, "cpc2_b200", "csrc", "*.cu")))
    os.makedirs(os.path.dirname(SO), exist_ok=True)
    subprocess.check_call([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
                           "-DCPC_GEMM_TIMING", "-Xcompiler", "-fPIC", "--shared", "-o", SO] + srcs + ["-lcudart"])
    print("built", SO)


if len(sys.argv) > 1 and sys.argv[1] == "build":
    build()
    sys.exit(0)

import torch  # noqa: E402

from cpc2_b200 import _lib  # noqa: E402

_lib.LIB_PATH = SO
_lib.needs_build = lambda: False
from cpc2_b200 import kernels as K  # noqa: E402
from cpc2_b200 import train as T  # noqa: E402

lib = _lib.lib()
lib.cpc_gemm_set_timing.argtypes = [ctypes.c_void_p]
buf = torch.zeros(148 * 16, dtype=torch.int64, device="cuda")
over = {}
for kv in sys.argv[1:]:
    k, v = kv.split("=")
    over[k] = int(v) if v.lstrip("-").isdigit() else v
args = T.defaultArgs(**over)
torch.manual_seed(0)
model, crit = T.buildModelAndCriterion(args)
model, crit = model.cuda().train(), crit.cuda().train()
B = args.batchSizeGPU
seq = (0.1 * torch.randn(B, 2, 1, args.sizeWindow)).cuda()
label = torch.zeros(B, dtype=torch.long, device="cuda")
T.forwardBackward(model, crit, seq, label)  # warm-up
torch.cuda.synchronize()
rows = []
orig = K._run


def run(name, a):
    if name != "cpc_gemm_bf16":
        return orig(name, a)
    torch.cuda.synchronize()
    buf.zero_()
    lib.cpc_gemm_set_timing(ctypes.c_void_p(buf.data_ptr()))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    orig(name, a)
    e1.record()
    torch.cuda.synchronize()
    lib.cpc_gemm_set_timing(ctypes.c_void_p(0))
    rows.append((dict(LAST), buf.view(148, 16).cpu().clone(), 1e3 * e0.elapsed_time(e1)))


K._run = run
LAST = {}
orig_gemm = K.gemm


def gemm(A, Bv, M, N, Kd, groups=1, k_per_slab=0, k_splits=1, **kw):
    kps = k_per_slab if k_per_slab > 0 else Kd
    total_kb = ((kps + 63) // 64) * (Kd // kps)
    mode = "norm" if kw.get("norm") is not None else ("splitk" if k_splits > 1 else ("acc" if kw.get("accumulate") else "plain"))
    LAST.clear()
    LAST.update(shape=f"g{groups} {M}x{N}x{Kd} a{A.major}b{Bv.major} s{k_splits} {mode}", nkb=(total_kb + k_splits - 1) // max(k_splits, 1))
    return orig_gemm(A, Bv, M, N, Kd, groups=groups, k_per_slab=k_per_slab, k_splits=k_splits, **kw)


K.gemm = gemm
import cpc2_b200.ops as ops  # noqa: E402
ops.CONCURRENT = False
for p in list(model.parameters()) + list(crit.parameters()):
    p.grad = None
T.forwardBackward(model, crit, seq, label)
torch.cuda.synchronize()
mhz = 1.0e-3 * 1900.0
print(f"{'shape':44s} {'ctas':>4s} {'nkb':>4s} {'evt us':>7s} | {'setup':>6s} {'1stdat':>6s} {'loop':>7s} {'/kb':>5s} {'accw':>5s} {'epi':>6s} {'tail':>6s} {'total':>7s}  (us at 1.9 GHz, median over CTAs; loop/epi of the FIRST tile)")
for info, t, evt in rows:
    act = t[:, 0] > 0
    t = t[act].double()
    n = int(act.sum())
    med = lambda x: float(x.median()) / 1900.0
    setup = med(t[:, 1] - t[:, 0])
    first = med(t[:, 4] - t[:, 1])
    loop = med(t[:, 5] - t[:, 4])
    accw = med(t[:, 7] - t[:, 5])
    epi = med(t[:, 8] - t[:, 7])
    tail = med(t[:, 10] - t[:, 9])
    total = float((t[:, 10] - t[:, 0]).max()) / 1900.0
    nkb = info.get("nkb", 0)
    print(f"{info.get('shape', '?'):44s} {n:4d} {nkb:4d} {evt:7.1f} | {setup:6.2f} {first:6.2f} {loop:7.2f} {loop / max(nkb, 1):5.2f} {accw:5.2f} {epi:6.2f} {tail:6.2f} {total:7.2f}")
