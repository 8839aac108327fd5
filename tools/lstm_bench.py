"""Times cpc_lstm_fwd / cpc_lstm_bwd at the shapes of the CPC step (H = 256, S = 128, N = 8 or 16 sequences).
Usage: python tools/lstm_bench.py [--once]   (--once: single launches, for ncu)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cpc2_b200 import kernels as K  # noqa: E402

BF = torch.bfloat16
dev = "cuda"


def case(N, S, H, cell="LSTM"):
    gx = torch.randn(N * S, 4 * H, device=dev) * 0.5
    w_hh = (torch.randn(4 * H, H, device=dev) / H ** 0.5).to(BF)
    out = torch.empty(N, S, H, device=dev)
    h_pad = torch.empty(N, S + 1, H, device=dev, dtype=BF)
    gates = torch.empty(N, S, 4 * H, device=dev, dtype=BF)
    c_all = torch.empty(N, S, H, device=dev)
    h_n, c_n = torch.empty(N, H, device=dev), torch.empty(N, H, device=dev)
    dout = torch.randn(N, S, H, device=dev)
    dgates = torch.empty(N, S, 4 * H, device=dev, dtype=BF)

    def fwd():
        K.lstm_fwd(gx, w_hh, N, S, H, h_pad=h_pad, out=out, gates=gates, c_all=c_all, h_n=h_n, c_n=c_n, cell=cell)

    def bwd():
        K.lstm_bwd(w_hh, gates, c_all, dout, dgates, N, S, H, cell=cell)
    return fwd, bwd


def timeit(fn, n=10):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


if __name__ == "__main__":
    once = "--once" in sys.argv
    for name, cfg in [("LSTM N8 S128 H256", (8, 128, 256)), ("LSTM N16 S128 H256", (16, 128, 256)),
                      ("GRU N8 S128 H256", (8, 128, 256, "GRU")), ("LSTM N8 S256 H512", (8, 256, 512))]:
        fwd, bwd = case(*cfg)
        if once:
            fwd()
            bwd()
            torch.cuda.synchronize()
            print("ran", name)
            if name.startswith("LSTM N8 S128"):
                continue
            break
        f, b = timeit(fwd), timeit(bwd)
        S = cfg[1]
        print(f"{name:24s} fwd {f:8.1f} us ({1e3 * f / S:6.0f} ns/step)   bwd {b:8.1f} us ({1e3 * b / S:6.0f} ns/step)", flush=True)
