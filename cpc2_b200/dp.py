"""Data-parallel wrap of the CPC step (the reference's cpc/train.py:523-527: DistributedDataParallel on
model and criterion, NCCL, mean-reduced gradients; cpc/distributed_training/distributed_mode.py:139-142).

One process per GPU, batch sharded per rank, negatives sampled inside the local shard only (as in the
reference), so forward/backward exchange nothing.  The single collective is the gradient all-reduce, issued
as THREE buckets in autograd readiness order -- criterion heads (90 % of the bytes, ready first), AR network,
encoder -- on a side stream as soon as the bucket's gradients exist, so the NVLink transfer of the heads'
66 MB overlaps the LSTM and encoder backward.  Each bucket is one flat fp32 buffer (the modules produce
their parameter gradients as views of one arena per module), all-reduced in place with NCCL AVG.
Everything is CUDA-graph capturable (see train.GraphedTrainStep).
"""
import torch
import torch.distributed as dist


class BucketedAllReduce:
    def __init__(self, cpcModel, cpcCriterion, process_group=None):
        self.group = process_group
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.stream = torch.cuda.Stream()
        self.buckets = [
            [p for p in cpcCriterion.parameters() if p.requires_grad],
            [p for p in cpcModel.gAR.parameters() if p.requires_grad],
            [p for p in cpcModel.gEncoder.parameters() if p.requires_grad],
        ]
        self.handles = []
        self._pending = False  # side stream carries work of the current step
        self.bytes_per_step = 4 * sum(p.numel() for b in self.buckets for p in b)
        if self.world > 1:
            for b in self.buckets:
                if b:
                    self.handles.append(torch.autograd.graph.register_multi_grad_hook(tuple(b), self._make_hook()))

    def broadcast_parameters(self, src=0):
        """DDP's initial parameter broadcast (train.py:524-527 wraps after construction)."""
        if self.world > 1:
            for b in self.buckets:
                for p in b:
                    dist.broadcast(p.data, src, group=self.group)

    def _make_hook(self):
        def hook(grads):
            gs = [g for g in grads if g is not None]
            if not gs:
                return
            base = gs[0]._base
            shared = base is not None and all(g._base is base for g in gs)
            main = torch.cuda.current_stream()
            self.stream.wait_stream(main)
            self._pending = True
            with torch.cuda.stream(self.stream):
                if shared:
                    base.record_stream(self.stream)
                    dist.all_reduce(base, op=dist.ReduceOp.AVG, group=self.group)
                else:  # gradients not produced by one arena (foreign module): reduce them one by one
                    for g in gs:
                        g.record_stream(self.stream)
                        dist.all_reduce(g, op=dist.ReduceOp.AVG, group=self.group)
        return hook

    def finish(self):
        """Join the side stream: call after backward, before the optimizer reads .grad."""
        if self._pending:
            torch.cuda.current_stream().wait_stream(self.stream)
            self._pending = False

    def remove(self):
        for h in self.handles:
            h.remove()
        self.handles = []
