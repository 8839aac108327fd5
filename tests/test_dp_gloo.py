"""World-size-2 check of the data-parallel host logic on CPU (gloo): bucket order (criterion heads, AR, encoder),
one flat all-reduce per bucket when a module hands its gradients out as views of one arena, mean reduction,
initial parameter broadcast, and the per-rank file shard (cpc/train.py:389-393, 523-527)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


class _ArenaGrads(torch.autograd.Function):
    """Stands in for the sm_100a modules on CPU: the parameter gradients come back as views of ONE flat
    buffer (value = rank + 1 everywhere), exactly how model.py / transformers.py produce them."""

    @staticmethod
    def forward(ctx, x, fill, stage, *params):
        ctx.shapes = [p.shape for p in params]
        ctx.fill, ctx.stage = fill, stage
        return x.sum() * 0.0

    @staticmethod
    def backward(ctx, g):
        n = sum(s.numel() for s in ctx.shapes)
        flat = torch.full((n,), float(ctx.fill))
        out, off = [], 0
        for s in ctx.shapes:
            out.append(flat[off:off + s.numel()].view(s))
            off += s.numel()
        from cpc2_b200 import ops
        ops.grad_ready(ctx.stage, flat)
        return (None, None, None, *out)


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from cpc2_b200 import train as T
        from cpc2_b200.dp import BucketedAllReduce, filter_distributed
        args = T.defaultArgs(hiddenEncoder=32, hiddenGar=32, nPredicts=2, negativeSamplingExt=4, sizeWindow=2560)
        torch.manual_seed(100 + rank)  # ranks start from DIFFERENT weights; the broadcast must fix that
        model, crit = T.buildModelAndCriterion(args)
        dp = BucketedAllReduce(model, crit, symmetric="auto")  # CPU parameters: no multicast mapping, gloo path
        assert dp.symmetric is None
        dp.broadcast_parameters()
        w = model.gEncoder.conv1.weight.detach().clone()
        ws = [torch.empty_like(w) for _ in range(world)]
        dist.all_gather(ws, w)
        same_after_broadcast = all(torch.equal(ws[0], x) for x in ws)
        x = torch.ones(3, requires_grad=True)
        loss = 0
        stages = list(reversed(list(zip(dp.STAGES, dp.buckets))))  # encoder, AR, criterion in forward order
        foreign = stages[0][1].pop()  # one encoder parameter whose gradient does NOT come from an arena
        dp.buckets[2].append(foreign)
        for stage, bucket in stages:
            loss = loss + _ArenaGrads.apply(x, rank + 1, stage, *[p for p in bucket if p is not foreign])
        loss = loss + (foreign * 0).sum() + (rank + 1) * foreign.sum()
        loss.backward()
        dp.finish()
        mean = sum(range(1, world + 1)) / world
        ok = all(torch.allclose(p.grad, torch.full_like(p.grad, mean)) for b in dp.buckets for p in b)
        files = list(range(11))
        q.put((rank, same_after_broadcast, ok, [st for st, _ in dp.reduced] + [dp.stragglers],
               [len(b) for b in dp.buckets],
               filter_distributed(files, rank, world), dp.bytes_per_step))
    finally:
        dist.destroy_process_group()


def test_bucketed_allreduce_world2_gloo():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    shards = []
    for rank, same, ok, reduced, sizes, shard, nbytes in res:
        assert same, "parameters differ after the initial broadcast"
        assert ok, "gradients are not the mean over ranks"
        assert reduced == ["criterion", "ar", "encoder", 1], reduced  # one flat all-reduce per bucket, 1 straggler
        assert len(sizes) == 3 and all(s > 0 for s in sizes)
        assert nbytes > 0
        shards += shard
    assert shards == list(range(11))  # the shards partition the file list, in order
