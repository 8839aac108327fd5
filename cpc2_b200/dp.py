"""Data-parallel wrap of the CPC step (the reference's cpc/train.py:523-527: DistributedDataParallel on
model and criterion, NCCL, mean-reduced gradients; cpc/distributed_training/distributed_mode.py:139-142).

One process per GPU, batch sharded per rank, negatives sampled inside the local shard only (as in the
reference), so forward/backward exchange nothing.  The single collective is the gradient all-reduce, issued
as THREE buckets in autograd readiness order -- criterion heads (90 % of the bytes, ready first), AR network,
encoder -- on a side stream as soon as the bucket's gradients exist, so the NVLink transfer of the heads'
66 MB overlaps the LSTM and encoder backward.  Each bucket is one flat fp32 arena (the modules produce their
parameter gradients as views of one arena per module and announce it through `ops.grad_ready` before autograd
sees the views), all-reduced in place with NCCL AVG; `.grad` tensors alias the reduced arena, nothing is copied.
Everything is CUDA-graph capturable (see train.GraphedTrainStep).
"""
import torch
import torch.distributed as dist

from . import ops


class BucketedAllReduce:
    STAGES = ("criterion", "ar", "encoder")

    def __init__(self, cpcModel, cpcCriterion, process_group=None):
        self.group = process_group
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.on_gpu = any(p.is_cuda for p in cpcModel.parameters())
        # the CPU (gloo) form exists for the world-size-2 host-logic tests only: same buckets, same order
        self.stream = torch.cuda.Stream() if self.on_gpu else None
        self.buckets = [
            [p for p in cpcCriterion.parameters() if p.requires_grad],
            [p for p in cpcModel.gAR.parameters() if p.requires_grad],
            [p for p in cpcModel.gEncoder.parameters() if p.requires_grad],
        ]
        self.reduced = []      # (stage, numel) per arena all-reduce of the current step (test / bench evidence)
        self.stragglers = 0    # gradients that had to be reduced one by one in finish()
        self._arenas = []      # storages reduced in the current step
        self._pending = False  # side stream carries work of the current step
        self.bytes_per_step = 4 * sum(p.numel() for b in self.buckets for p in b)
        if self.world > 1:
            ops.GRAD_READY.append(self._on_grad_ready)

    def broadcast_parameters(self, src=0):
        """DDP's initial parameter broadcast (train.py:524-527 wraps after construction)."""
        if self.world > 1:
            for b in self.buckets:
                for p in b:
                    dist.broadcast(p.data, src, group=self.group)

    def _all_reduce_mean(self, t):
        if self.on_gpu:
            dist.all_reduce(t, op=dist.ReduceOp.AVG, group=self.group)
        else:  # gloo has no AVG
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
            t.div_(self.world)

    def _on_grad_ready(self, stage, gflat, params=None, grads=None):
        self.reduced.append((stage, gflat.numel()))
        self._arenas.append(gflat.untyped_storage().data_ptr())
        if not self.on_gpu:
            self._all_reduce_mean(gflat)
            return
        self.stream.wait_stream(torch.cuda.current_stream())
        self._pending = True
        with torch.cuda.stream(self.stream):
            gflat.record_stream(self.stream)
            self._all_reduce_mean(gflat)

    def finish(self):
        """Join the side stream: call after backward, before the optimizer reads .grad.  Any gradient that does
        not alias an arena reduced this step (a foreign module, or a copy autograd made) is reduced on its own."""
        if self.world > 1:
            arenas = set(self._arenas)
            for b in self.buckets:
                for p in b:
                    if p.grad is not None and p.grad.untyped_storage().data_ptr() not in arenas:
                        self.stragglers += 1
                        self._all_reduce_mean(p.grad)
        self._arenas = []
        if self._pending:
            torch.cuda.current_stream().wait_stream(self.stream)
            self._pending = False

    def remove(self):
        if self._on_grad_ready in ops.GRAD_READY:
            ops.GRAD_READY.remove(self._on_grad_ready)


def filter_distributed(files, global_rank, world_size):
    """Rank's contiguous slice of the file list (cpc/train.py:389-393): the data-parallel shard assignment."""
    start = len(files) * global_rank // world_size
    end = len(files) * (global_rank + 1) // world_size
    return files[start:end]
