"""Kernel timeline of one CUDA-graph replay of the training step (torch.profiler / CUPTI): start, duration,
stream and the idle gaps, to see what actually overlaps.  Usage: python tools/timeline.py [key=value ...]"""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from cpc2_b200 import ops  # noqa: E402
from cpc2_b200 import train as T  # noqa: E402

over = {}
for kv in sys.argv[1:]:
    k, v = kv.split("=")
    over[k] = int(v) if v.lstrip("-").isdigit() else v
conc = bool(over.pop("concurrent", 1))
ops.CONCURRENT = conc
args = T.defaultArgs(**over)
torch.manual_seed(0)
model, crit = T.buildModelAndCriterion(args)
model, crit = model.cuda(), crit.cuda()
B = args.batchSizeGPU
step = T.GraphedTrainStep(model, crit, B, args.sizeWindow, warmup=3)
step.static_seq.copy_(bench.synthetic_batch(B, args.sizeWindow, 0))
for _ in range(5):
    step.step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    step.step()
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
t0 = evs[0].time_range.start
end_max = t0
busy = 0.0
print(f"# concurrent={conc}  {len(evs)} device activities")
for e in evs:
    s, d = e.time_range.start - t0, e.time_range.end - e.time_range.start
    gap = e.time_range.start - end_max
    print(f"{s:9.1f} us +{d:7.1f}  gap {gap:7.1f}  {e.name[:70]}")
    if e.time_range.end > end_max:
        busy += e.time_range.end - max(end_max, e.time_range.start)
        end_max = e.time_range.end
print(f"# span {end_max - t0:.1f} us, union of kernel time {busy:.1f} us, sum of durations "
      f"{sum(e.time_range.end - e.time_range.start for e in evs):.1f} us")
