"""Start time of every C-ABI launch inside ONE real CUDA-graph replay of the training step (no profiler attached):
an external timing event is recorded on the launching stream right before each launch during capture; after a replay
the elapsed times from the first mark give the step's timeline as the GPU actually ran it (idle gaps included).
Usage: python tools/graph_marks.py [key=value ...]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from cpc2_b200 import kernels as K  # noqa: E402
from cpc2_b200 import train as T  # noqa: E402
from cpc2_b200.optim import FusedAdam  # noqa: E402

over = {}
for kv in sys.argv[1:]:
    k, v = kv.split("=")
    over[k] = int(v) if v.lstrip("-").isdigit() else v
args = T.defaultArgs(**over)
torch.manual_seed(0)
model, crit = T.buildModelAndCriterion(args)
model, crit = model.cuda(), crit.cuda()
B = args.batchSizeGPU
opt = FusedAdam(list(crit.parameters()) + list(model.parameters()), lr=2e-4)
opt.attach()
marks = []
orig = K._run
capturing = False


def run(name, a):
    if capturing:
        ev = torch.cuda.Event(enable_timing=True, external=True)
        ev.record()
        marks.append((name, torch.cuda.current_stream().cuda_stream, ev))
    orig(name, a)


K._run = run
orig_init_graph = torch.cuda.graph.__enter__


def enter(self):
    global capturing
    r = orig_init_graph(self)
    capturing = True
    return r


orig_exit = torch.cuda.graph.__exit__


def exit_(self, *a):
    global capturing
    end = torch.cuda.Event(enable_timing=True, external=True)
    end.record()
    marks.append(("END(main stream)", torch.cuda.current_stream().cuda_stream, end))
    capturing = False
    return orig_exit(self, *a)


torch.cuda.graph.__enter__, torch.cuda.graph.__exit__ = enter, exit_
step = T.GraphedTrainStep(model, crit, B, args.sizeWindow, warmup=3, optimizer=opt)
step.static_seq.copy_(bench.synthetic_batch(B, args.sizeWindow, 0))
for _ in range(5):
    step.step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
step.step()
e1.record()
torch.cuda.synchronize()
print(f"# replay {1e3 * e0.elapsed_time(e1):.1f} us (events around graph.replay()), {len(marks)} marks")
streams = {}
first = marks[0][2]
rows = sorted(((1e3 * first.elapsed_time(ev), name, streams.setdefault(s, len(streams))) for name, s, ev in marks))
t0 = 1e3 * e0.elapsed_time(first)
print(f"# first mark {t0:.1f} us after the event before replay")
prev = 0.0
for t, name, s in rows:
    print(f"{t:9.1f} us  (+{t - prev:6.1f})  s{s:<2d} {name}")
    prev = t
