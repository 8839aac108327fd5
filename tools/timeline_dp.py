"""Kernel timeline of one CUDA-graph replay of the DATA-PARALLEL step (bench.TrainRig: bucketed all-reduce + staged
fused Adam), rank 0, torch.profiler / CUPTI.  Launch under torchrun:
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 tools/timeline_dp.py
(also runs at world size 1: the same rig without a collective)."""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from cpc2_b200 import train as T  # noqa: E402

ctx = bench.Ctx()
args = T.defaultArgs()
rig = bench.TrainRig(ctx, args, args.batchSizeGPU)
for _ in range(8):
    rig.step.step()
ctx.barrier()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    rig.step.step()
    torch.cuda.synchronize()
ctx.barrier()
if ctx.rank == 0:
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    evs.sort(key=lambda e: e.time_range.start)
    t0 = evs[0].time_range.start
    end_max, busy = t0, 0.0
    print(f"# world {ctx.world}  {len(evs)} device activities")
    for e in evs:
        s, d = e.time_range.start - t0, e.time_range.end - e.time_range.start
        print(f"{s:9.1f} us +{d:7.1f}  gap {e.time_range.start - end_max:7.1f}  {e.name[:70]}")
        if e.time_range.end > end_max:
            busy += e.time_range.end - max(end_max, e.time_range.start)
            end_max = e.time_range.end
    print(f"# span {end_max - t0:.1f} us, union of kernel time {busy:.1f} us")
rig.close()
bench.finish(ctx)
