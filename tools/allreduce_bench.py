"""cpc_allreduce_mc (in-switch multimem all-reduce) against NCCL: same result, and the time of both per bucket size.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29577 \
        tools/allreduce_bench.py [--check-only]

Rank 0 prints one line per size: floats, max |mc - nccl|, mc us, nccl us, mc GB/s (algorithmic: 4n bytes / time).
Exits 0 with "multicast unavailable" when the box cannot map symmetric memory (the step then stays on NCCL)."""
import argparse
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--check-only", action="store_true")
    ap.add_argument("--blocks", type=int, default=0)
    opt = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import torch.distributed._symmetric_memory as symm
    from cpc2_b200 import kernels as K
    sizes = [4096, 262144 + 8, 2 * 1024 * 1024, 12_600_000 // 4 * 4, 16_600_000]
    cap = max(sizes) + 64
    try:
        buf = symm.empty(cap, dtype=torch.float32, device=dev)
        hdl = symm.rendezvous(buf, dist.group.WORLD)
        mc = int(hdl.multicast_ptr)
        if mc == 0:
            raise RuntimeError("multicast_ptr == 0")
    except Exception as e:  # noqa: BLE001
        if rank == 0:
            print(f"multicast unavailable: {type(e).__name__}: {e}")
        dist.destroy_process_group()
        return
    pads = int(hdl.signal_pad_ptrs_dev)
    ok = True
    for n in sizes:
        g = torch.Generator(device=dev).manual_seed(100 + rank)
        x = torch.randn(n, device=dev, generator=g)
        ref = x.clone()
        dist.all_reduce(ref, op=dist.ReduceOp.AVG)
        off = 32  # a span that does not start at the buffer base
        buf[off:off + n].copy_(x)
        buf[:off].fill_(7.0)
        buf[off + n:off + n + 32].fill_(7.0)
        torch.cuda.synchronize()
        dist.barrier()
        K.allreduce_mc(mc + 4 * off, n, pads, rank, world, 1.0 / world, opt.blocks)
        torch.cuda.synchronize()
        err = float((buf[off:off + n] - ref).abs().max())
        guard = bool((buf[off + n:off + n + 32] == 7.0).all()) and bool((buf[:off] == 7.0).all())
        line = f"n={n} max|mc-nccl|={err:.3e} guard_ok={guard}"
        ok = ok and err < 1e-5 and guard
        if not opt.check_only:
            times = []
            for fn in (lambda: K.allreduce_mc(mc + 4 * off, n, pads, rank, world, 1.0 / world, opt.blocks),
                       lambda: dist.all_reduce(ref, op=dist.ReduceOp.AVG)):
                for _ in range(5):
                    fn()
                torch.cuda.synchronize()
                dist.barrier()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(20):
                    fn()
                e1.record()
                torch.cuda.synchronize()
                t = torch.tensor([e0.elapsed_time(e1) / 20 * 1e3], device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                times.append(float(t))
            line += f" mc={times[0]:.1f}us nccl={times[1]:.1f}us mc_algbw={4 * n / times[0] / 1e3:.1f}GB/s"
        if rank == 0:
            print(line, flush=True)
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    if rank == 0:
        print("ALLREDUCE_MC", "OK" if int(flag) == 0 else "MISMATCH", f"world={world}", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(flag) == 0 else 1)


if __name__ == "__main__":
    main()
