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

`symmetric=True` (NVSwitch boxes): the gradient arenas live in ONE symmetric-memory buffer mapped through an NVLink
multicast address (torch.distributed._symmetric_memory provides the mapping, nothing else), and each bucket is reduced
by cpc_allreduce_mc -- one kernel, the switch does the sum (multimem.ld_reduce) and the broadcast (multimem.st) --
instead of NCCL's ring: half the NVLink bytes, two barriers instead of 2(N-1) hops, and no shared memory, so its
CTAs co-reside with the backward's persistent GEMM CTAs.  The arenas then persist across steps: use
zero_grad(set_to_none=True) (torch's default) -- a gradient left in .grad would alias the next step's arena.
"""
import torch
import torch.distributed as dist

from . import kernels as K
from . import ops


class BucketedAllReduce:
    STAGES = ("criterion", "ar", "encoder")

    def __init__(self, cpcModel, cpcCriterion, process_group=None, symmetric=False, max_blocks=0):
        self.group = process_group
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        # cpcModel None: the criterion alone (BASELINE.json's criterion-only sweep: the heads' gradients are the bucket)
        mods = [cpcCriterion] + ([cpcModel.gAR, cpcModel.gEncoder] if cpcModel is not None else [])
        self.on_gpu = any(p.is_cuda for m in mods for p in m.parameters())
        # the CPU (gloo) form exists for the world-size-2 host-logic tests only: same buckets, same order
        self.stream = torch.cuda.Stream() if self.on_gpu else None
        self.buckets = [[p for p in m.parameters() if p.requires_grad] for m in mods]
        self.reduced = []      # (stage, numel) per arena all-reduce of the current step (test / bench evidence)
        self.stragglers = 0    # gradients that had to be reduced one by one in finish()
        self._arenas = []      # storages reduced in the current step
        self._pending = False  # side stream carries work of the current step
        self.bytes_per_step = 4 * sum(p.numel() for b in self.buckets for p in b)
        self.symmetric, self.max_blocks, self.mc_launches = None, max_blocks, 0
        if self.world > 1:
            ops.GRAD_READY.append(self._on_grad_ready)
            if symmetric and self.on_gpu:
                self._map_symmetric(next(p for b in self.buckets for p in b if p.is_cuda).device, symmetric)

    def _map_symmetric(self, dev, mode):
        """One symmetric fp32 buffer for every gradient arena of the step + its multicast mapping.  mode "auto":
        stay on NCCL when the box has no multicast support; True: raise."""
        import torch.distributed._symmetric_memory as symm
        from . import _lib
        group = self.group if self.group is not None else dist.group.WORLD
        numel = sum((p.numel() + 7) // 8 * 8 for b in self.buckets for p in b)
        numel += 4 * 1024 * 1024  # padded gate slots of the recurrent arenas, per-arena alignment
        try:
            if symm.get_signal_pad_size() < _lib.lib().cpc_allreduce_pad_bytes(self.world):
                raise RuntimeError("symmetric-memory signal pad too small")
            buf = symm.empty(numel, dtype=torch.float32, device=dev)
            hdl = symm.rendezvous(buf, group)
            mc = int(hdl.multicast_ptr)
            if mc == 0:
                raise RuntimeError("no NVLink multicast support on this box")
        except Exception:
            if mode == "auto":
                return
            raise
        buf.zero_()
        self.symmetric = dict(buf=buf, hdl=hdl, mc=mc, base=buf.data_ptr(), bytes=4 * numel, pads=int(hdl.signal_pad_ptrs_dev),
                              rank=int(hdl.rank), used=0, slots={})
        ops.GRAD_ARENA = self._arena

    def _arena(self, owner, numel, device, zero_on=None):
        sy = self.symmetric
        slot = sy["slots"].get(owner)
        if slot is None:  # first backward: carve (same order on every rank, so the offsets are symmetric)
            n = (numel + 31) // 32 * 32
            if sy["used"] + n > sy["buf"].numel() or device != sy["buf"].device:
                return None
            slot = sy["slots"][owner] = (sy["used"], numel)
            sy["used"] += n
        if slot[1] != numel:
            return None
        t = sy["buf"][slot[0]:slot[0] + numel]
        if zero_on is None:
            return t.zero_()
        with ops.side(zero_on):
            t.zero_()
        return t

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
        sy = self.symmetric
        off = gflat.data_ptr() - sy["base"] if sy is not None else -1
        with torch.cuda.stream(self.stream):
            skip = getattr(self, "_skip_collective", False)  # bench.py diagnosis (CPC_DIAG_NO_ALLREDUCE=1 / =<stage prefixes>)
            if skip is True or (isinstance(skip, tuple) and stage.startswith(skip)):
                pass  # same plumbing, no exchange: NOT a valid training step
            elif 0 <= off < sy["bytes"] if sy is not None else False:
                n = (gflat.numel() + 3) // 4 * 4  # arenas are padded to 32 floats inside the symmetric buffer
                K.allreduce_mc(sy["mc"] + off, n, sy["pads"], sy["rank"], self.world, 1.0 / self.world, self.max_blocks)
                self.mc_launches += 1
            else:
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
        """Unhook, and give the symmetric buffer back NOW: gradients that alias it are dropped and the buffer, its
        rendezvous handle and the carved slots are released here, synchronously, on every rank.  Left to the garbage
        collector the last reference can die in the middle of a later CUDA-graph capture, and freeing
        symmetric memory is not permitted while a stream is capturing (the process aborts)."""
        if self._on_grad_ready in ops.GRAD_READY:
            ops.GRAD_READY.remove(self._on_grad_ready)
        if self.symmetric is not None:
            if ops.GRAD_ARENA == self._arena:
                ops.GRAD_ARENA = None
            base, nbytes = self.symmetric["base"], self.symmetric["bytes"]
            for b in self.buckets:
                for p in b:
                    if p.grad is not None and base <= p.grad.data_ptr() < base + nbytes:
                        p.grad = None
            if self.on_gpu:
                torch.cuda.synchronize()
            self.symmetric.clear()
            self.symmetric = None
            self._arenas = []
            import gc
            gc.collect()
            if self.on_gpu:
                torch.cuda.synchronize()


def filter_distributed(files, global_rank, world_size):
    """Rank's contiguous slice of the file list (cpc/train.py:389-393): the data-parallel shard assignment."""
    start = len(files) * global_rank // world_size
    end = len(files) * (global_rank + 1) // world_size
    return files[start:end]
