"""BASELINE.json configs[4]: criterion-only sweep -- CPCUnsupersivedCriterion forward + backward on random
c, z [8, 128, 256] for negatives {32..1024} x prediction steps {4..24}, linear and transformer heads; the step is captured in
a CUDA graph (as train.GraphedTrainStep does) and 20 back-to-back replays are timed with CUDA events.  Usage: python tools/criterion_sweep.py > profiles/<name>.txt"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cpc2_b200 import criterion as cr  # noqa: E402

B, S, D = 8, 128, 256
print(f"# criterion only, c/z [{B},{S},{D}] ~ N(0,1), fwd+bwd, us per step (CUDA-graph replay, warm L2)")
print(f"# {'heads':12s} {'NEG':>5s} " + " ".join(f"K={k:<6d}" for k in (4, 8, 12, 16, 24)))
for rnn in ("linear", "transformer"):
    for NEG in (32, 64, 128, 256, 512, 1024):
        row = []
        for Kp in (4, 8, 12, 16, 24):
            torch.manual_seed(0)
            crit = cr.CPCUnsupersivedCriterion(Kp, D, D, NEG, rnnMode=rnn, sizeInputSeq=S).cuda().train()
            c = torch.randn(B, S, D, device="cuda", requires_grad=True)
            z = torch.randn(B, S, D, device="cuda", requires_grad=True)
            label = torch.zeros(B, dtype=torch.long, device="cuda")

            def step():
                for p in crit.parameters():
                    p.grad = None
                c.grad = z.grad = None
                losses, _ = crit(c, z, label, None)
                losses.sum().backward()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(3):
                    step()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                step()
            graph.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            n = 20
            for _ in range(n):
                graph.replay()
            e1.record()
            torch.cuda.synchronize()
            row.append(e0.elapsed_time(e1) * 1e3 / n)
            del crit, graph
        print(f"  {rnn:12s} {NEG:5d} " + " ".join(f"{v:8.1f}" for v in row), flush=True)
