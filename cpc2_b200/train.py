"""Drop-in for the hot-path entry points of the reference's cpc/train.py and cpc/feature_loader.py:
getEncoder / getAR (feature_loader.py:202-235), getCriterion (train.py:27-59), trainStep (train.py:72-142),
valStep (train.py:145-187) -- same signatures and batch handling, built on the sm_100a modules.

`GraphedTrainStep` is the B200-native way to run the step: the whole forward + backward (every kernel launch,
the RNG draws of the negative sampler, the gradient buffers) is captured once in a CUDA graph and replayed,
which removes the per-launch host overhead that otherwise dominates a ~1 ms step.
"""
import argparse
import time
import weakref

import numpy as np
import torch

from . import criterion as cr
from . import model as model_
from . import transformers as tr


def defaultArgs(**over):
    """The architecture defaults of cpc/cpc_default_config.py:13-162 that the hot path reads."""
    a = argparse.Namespace(
        hiddenEncoder=256, hiddenGar=256, nPredicts=12, negativeSamplingExt=128, sizeWindow=20480,
        samplingType="samespeaker", cpc_mode=None, encoder_type="cpc", normMode="layerNorm", arMode="LSTM",
        nLevelsGRU=1, rnnMode="transformer", dropout=False, abspos=False, multihead_rnn=False,
        transformer_pruning=0, n_skipped=0, growth_rate=None, inflection_point_x=None, supervised=False,
        onEncoder=False, pathPhone=None, mask_prob=0.0, mask_length=10, learningRate=2e-4, beta1=0.9, beta2=0.999,
        epsilon=1e-08, batchSizeGPU=8)
    for k, v in over.items():
        setattr(a, k, v)
    return a


def getEncoder(args):
    """cpc/feature_loader.py:202-212 (CPC waveform encoder only)."""
    if getattr(args, "encoder_type", "cpc") != "cpc":
        raise NotImplementedError("cpc2_b200: only encoder_type='cpc' is on the accelerated path")
    return model_.CPCEncoder(args.hiddenEncoder, args.normMode)


def getAR(args):
    """cpc/feature_loader.py:215-235."""
    if args.arMode == 'transformer':
        arNet = tr.buildTransformerAR(args.hiddenEncoder, args.hiddenGar, args.nLevelsGRU, args.sizeWindow // 160,
                                      args.abspos)
        args.hiddenGar = args.hiddenEncoder
    elif args.cpc_mode == "bert" or args.arMode == 'no_ar':
        raise NotImplementedError("cpc2_b200: bert / no_ar variants are outside the hot path")
    else:
        arNet = model_.CPCAR(args.hiddenEncoder, args.hiddenGar, args.samplingType == "sequential",
                             args.nLevelsGRU, mode=args.arMode, reverse=args.cpc_mode == "reverse")
    return arNet


def getCriterion(args, downsampling, nSpeakers, nPhones):
    """cpc/train.py:27-59 (unsupervised CPC criterion only)."""
    if args.supervised or args.cpc_mode == 'none':
        raise NotImplementedError("cpc2_b200: supervised / 'none' criteria are outside the hot path")
    sizeInputSeq = (args.sizeWindow // downsampling)
    return cr.CPCUnsupersivedCriterion(args.nPredicts, args.hiddenGar, args.hiddenEncoder,
                                       args.negativeSamplingExt, mode=args.cpc_mode, rnnMode=args.rnnMode,
                                       dropout=args.dropout, nSpeakers=nSpeakers, sizeInputSeq=sizeInputSeq,
                                       multihead_rnn=args.multihead_rnn,
                                       transformer_pruning=args.transformer_pruning, n_skipped=args.n_skipped,
                                       growth_rate=args.growth_rate, inflection_point_x=args.inflection_point_x)


def buildModelAndCriterion(args):
    """Construction order of cpc/train.py:442-461, so one seed gives the reference's initial weights."""
    encoderNet = getEncoder(args)
    arNet = getAR(args)
    cpcModel = model_.CPCModel(encoderNet, arNet, getattr(args, "mask_prob", 0.0), getattr(args, "mask_length", 10))
    cpcCriterion = getCriterion(args, encoderNet.DOWNSAMPLING, 0, 0)
    return cpcModel, cpcCriterion


# forwardBackward runs an unwrapped cpc2_b200 CPCModel as two halves (CPCModel.forwardPastFuture); False restores
# the literal concatenated call of cpc/train.py:95-104 (wrapped models -- DataParallel / DDP -- always use it)
SPLIT_PAST_FUTURE = True
_ONES = {}  # seeds of the backward, per loss shape


def prefetchWeights(cpcModel, cpcCriterion):
    """Issue every weight-only preparation of the step (conv weight packing, bf16 shadows of the AR and head
    weights) on side streams now, so it overlaps the encoder instead of sitting in front of each consumer.
    Works through DataParallel / DistributedDataParallel wrappers (`.module`)."""
    m = getattr(cpcModel, "module", cpcModel)
    c = getattr(cpcCriterion, "module", cpcCriterion)
    mods = [getattr(m, "gEncoder", None), getattr(c, "wPrediction", None)]
    ar = getattr(m, "gAR", None)
    mods += list(ar) if isinstance(ar, torch.nn.Sequential) else [ar]
    for mod in mods:
        if mod is not None and hasattr(mod, "prefetch") and next(mod.parameters()).is_cuda:
            mod.prefetch()


def forwardBackward(cpcModel, cpcCriterion, sequence, label, signal_quality=None, backward=True):
    """The body of the reference's batch loop, cpc/train.py:95-109."""
    prefetchWeights(cpcModel, cpcCriterion)
    past, future = sequence[:, 0, ...], sequence[:, 1, ...]
    b = past.size(0)
    crit = getattr(cpcCriterion, "module", cpcCriterion)
    if hasattr(crit, "prefetchIndices") and sequence.is_cuda:
        crit.prefetchIndices(b, model_._geometry(b, sequence.shape[-1])[0][4], sequence.device)
    if SPLIT_PAST_FUTURE and isinstance(cpcModel, model_.CPCModel) and cpcModel.splittable():
        # same outputs and gradients as the concatenated call below, without its dead half (the AR over the
        # future windows, whose output train.py:102 drops) and with the two encoder halves overlapped
        c_feature, encoded_data = cpcModel.forwardPastFuture(past, future)
    else:
        combined = torch.cat([past, future], dim=0)
        label2 = torch.cat([label, label])
        c_feature, encoded_data, label2 = cpcModel(combined, label2)
        c_feature = c_feature[:b, :, :]
        encoded_data = encoded_data[b:, :, :]
        label = label2[:b]
    allLosses, allAcc = cpcCriterion(c_feature, encoded_data, label, signal_quality)
    if backward:
        # totLoss = allLosses.sum(); totLoss.backward()  (cpc/train.py:107-109): the gradient of a sum with respect to
        # every summand is one, so the backward is seeded with a cached tensor of ones -- the same gradients without
        # the sum, the fill and the expand kernels autograd would put in front of the criterion's backward
        from . import ops
        key = (tuple(allLosses.shape), allLosses.device, allLosses.dtype)
        seed = _ONES.get(key)
        if seed is None:
            seed = _ONES[key] = torch.ones(allLosses.shape, device=allLosses.device, dtype=allLosses.dtype)
        # the modules' backward passes may leave their weight-gradient streams to the step's final join (ops.step_scope)
        # -- unless somebody reads gradients on the autograd stream as they are accumulated: the reducer of a
        # DistributedDataParallel / DataParallel wrapper, or a hook registered on a parameter
        bare = isinstance(cpcModel, model_.CPCModel) and isinstance(cpcCriterion, cr.CPCUnsupersivedCriterion) and not any(
            p._backward_hooks or getattr(p, "_post_accumulate_grad_hooks", None)
            for m in (cpcModel, cpcCriterion) for p in m.parameters())
        if bare:
            with ops.step_scope():
                torch.autograd.backward(allLosses, grad_tensors=seed)
        else:
            torch.autograd.backward(allLosses, grad_tensors=seed)
    return allLosses, allAcc


# trainStep replays the whole step (forward + backward + FusedAdam update) as one CUDA graph when it can: an unwrapped
# cpc2_b200 CPCModel without keepHidden state, a cpc2_b200.optim.FusedAdam optimizer, no signal-quality weights, a
# batch shape seen before.  Anything else -- DataParallel / DDP wrappers, torch.optim.Adam, the ragged last batch of
# an epoch -- takes the eager body below (still the two-halves forward and the fused kernels).  False: always eager.
GRAPH_TRAIN_STEP = True
_MAX_GRAPHS = 2
_GRAPHED = weakref.WeakKeyDictionary()


def _graph_runner(cpcModel, cpcCriterion, optimizer, sequence, signal_quality):
    from . import ops
    from .optim import FusedAdam
    if not (GRAPH_TRAIN_STEP and signal_quality is None and isinstance(optimizer, FusedAdam)
            and isinstance(cpcModel, model_.CPCModel) and isinstance(cpcCriterion, cr.CPCUnsupersivedCriterion)
            and cpcModel.splittable() and sequence.is_cuda and sequence.dim() == 4 and torch.is_grad_enabled()):
        return None
    if any(h for h in ops.GRAD_READY if getattr(h, "__self__", None) is not optimizer):
        return None  # somebody else (a data-parallel wrapper) hooks the gradients: its owner drives the graph itself
    cache = _GRAPHED.setdefault(cpcModel, {})  # (not an attribute: the module stays deep-copyable / picklable)
    key = (tuple(sequence.shape), id(cpcCriterion), id(optimizer), cpcModel.training, cpcCriterion.training)
    runner = cache.get(key)
    if runner is None:
        if len(cache) >= _MAX_GRAPHS:
            return None
        attached = optimizer._on_stage in ops.GRAD_READY
        optimizer.attach()  # each module's parameters are stepped as soon as its gradients exist
        try:
            runner = GraphedTrainStep(cpcModel, cpcCriterion, sequence.shape[0], sequence.shape[-1],
                                      device=sequence.device, warmup=3, train_mode=cpcModel.training, optimizer=optimizer)
        finally:
            if not attached:
                optimizer.detach()
        cache[key] = runner
    return runner


def trainStep(dataLoader, cpcModel, cpcCriterion, optimizer, scheduler, loggingStep):
    """cpc/train.py:72-142: same signature, same batch handling, same returned logs.  The per-step
    `.cpu().numpy()` log reads of the reference (train.py:121-122, one host sync per step) become float64 sums kept
    on the device and read once per call (and at every logging step, where the reference prints timings)."""
    cpcModel.train()
    cpcCriterion.train()
    start_time = time.perf_counter()
    n_examples = 0
    logs = {}
    it = 0
    sums = None  # [2, K] float64 on the device: per-step losses / accuracies summed
    for step, full_data in enumerate(dataLoader):
        sequence, label, *signal_quality = full_data
        sequence = sequence.cuda(non_blocking=True)
        label = label.cuda(non_blocking=True)
        signal_quality = signal_quality[0].cuda(non_blocking=True) if len(signal_quality) else None
        n_examples += sequence.size(0)
        runner = _graph_runner(cpcModel, cpcCriterion, optimizer, sequence, signal_quality)
        if runner is not None:
            allLosses, allAcc = runner.step(sequence)  # forward + backward + optimizer.step() in one replay
        else:
            if cpcModel in _GRAPHED:
                optimizer.zero_grad()  # .grad may still point at the static gradients of an earlier graph replay
            allLosses, allAcc = forwardBackward(cpcModel, cpcCriterion, sequence, label, signal_quality)
            optimizer.step()
            optimizer.zero_grad()
        if allLosses.nelement() > 0:
            if sums is None:
                sums = torch.zeros(2, allLosses.size(1), dtype=torch.float64, device=allLosses.device)
            it += 1
            sums[0] += allLosses.detach().mean(dim=0)
            sums[1] += allAcc.detach().mean(dim=0)
            if (step + 1) % loggingStep == 0:
                torch.cuda.synchronize()
                new_time = time.perf_counter()
                elapsed = new_time - start_time
                print(f"Update {step + 1}")
                print(f"elapsed: {elapsed:.1f} s")
                print(f"{1000.0 * elapsed / loggingStep:.1f} ms per batch, "
                      f"{1000.0 * elapsed / n_examples:.1f} ms / example")
                start_time, n_examples = new_time, 0
    if scheduler is not None:
        scheduler.step()
    if sums is not None:
        host = sums.cpu().numpy()
        logs["locLoss_train"], logs["locAcc_train"] = host[0], host[1]
    logs = {k: v / max(it, 1) for k, v in logs.items()}
    logs["iter"] = it
    return logs


def valStep(dataLoader, cpcModel, cpcCriterion):
    """cpc/train.py:145-187."""
    cpcCriterion.eval()
    cpcModel.eval()
    logs = {}
    it = 0
    for step, full_data in enumerate(dataLoader):
        sequence, label, *signal_quality = full_data
        sequence = sequence.cuda(non_blocking=True)
        label = label.cuda(non_blocking=True)
        with torch.no_grad():
            allLosses, allAcc = forwardBackward(cpcModel, cpcCriterion, sequence, label, None, backward=False)
        if "locLoss_val" not in logs:
            logs["locLoss_val"] = np.zeros(allLosses.size(1))
            logs["locAcc_val"] = np.zeros(allLosses.size(1))
        it += 1
        logs["locLoss_val"] += allLosses.mean(dim=0).cpu().numpy()
        logs["locAcc_val"] += allAcc.mean(dim=0).cpu().numpy()
    logs = {k: v / max(it, 1) for k, v in logs.items()}
    logs["iter"] = it
    return logs


class GraphedTrainStep:
    """Forward + backward of one batch captured in a CUDA graph (static input / output buffers).

    step(sequence_host_or_device) copies the batch into the static input, replays the graph and returns the
    static (losses, acc) tensors; parameter .grad tensors are static too (rewritten by every replay), so an
    optimizer step can follow as usual.  `grad_hook(stage)` (optional) is invoked during capture right after
    the criterion / AR / encoder gradients exist -- the data-parallel wrapper uses it to enqueue its bucketed
    all-reduces on a side stream inside the same graph."""

    def __init__(self, cpcModel, cpcCriterion, batch, sizeWindow, device="cuda", warmup=3, train_mode=True,
                 after_backward=None, optimizer=None):
        """`optimizer` (cpc2_b200.optim.FusedAdam, or any capturable optimizer): its step() becomes part of the
        graph, after `after_backward` (the data-parallel all-reduce) -- the whole of train.py:95-113 in one replay.

        Construction has NO side effects on the training state: the `warmup` eager passes that precede the capture
        (allocator warm-up, lazy kernel attributes, NCCL channel setup) run real optimizer steps on an all-zero batch,
        so parameters, the optimizer's moments and step counter (FusedAdam.snapshot / restore), the dropout counter and
        the CUDA RNG state are saved first and put back afterwards -- a freshly built or checkpoint-loaded model starts
        its first replay from exactly its loaded state (an optimizer without snapshot()/restore() is warmed up with
        its real step(): pass a FusedAdam, or accept `warmup` updates on silence)."""
        from . import ops
        self.model, self.crit, self.optimizer = cpcModel, cpcCriterion, optimizer
        self._copy_stream, self._staged, self._staged_ready, self._picked = None, None, False, None
        self.static_seq = torch.zeros(batch, 2, 1, sizeWindow, device=device)
        self.static_label = torch.zeros(batch, dtype=torch.long, device=device)
        self.params = [p for p in list(cpcModel.parameters()) + list(cpcCriterion.parameters()) if p.requires_grad]
        cpcModel.train(train_mode)
        cpcCriterion.train(train_mode)
        dev = self.static_seq.device
        torch.cuda.synchronize(dev)
        # ---- snapshot of everything the warm-up passes mutate
        keep = optimizer is not None and warmup > 0
        snap_p = [p.detach().clone() for p in self.params] if keep else None
        snap_o = optimizer.snapshot() if keep and hasattr(optimizer, "snapshot") else None
        # nn.BatchNorm1d (encoder normMode "batchNorm") moves its running estimates and step count in every training-mode
        # forward, optimizer or not
        bufs = [b for m in (cpcModel, cpcCriterion) for n, b in m.named_buffers()
                if ".running_" in n or n.endswith("num_batches_tracked")] if warmup > 0 else []
        snap_b = [b.detach().clone() for b in bufs]
        rng_state = torch.cuda.get_rng_state(dev)
        drop_state = ops.DropoutState._states.get((dev.type, dev.index))
        drop_snap = drop_state.clone() if drop_state is not None else None
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):
                self._zero()
                forwardBackward(cpcModel, cpcCriterion, self.static_seq, self.static_label)
                if after_backward is not None:
                    after_backward()
                if optimizer is not None:
                    optimizer.step()
            if keep and snap_o is not None:
                with torch.no_grad():
                    for p, q in zip(self.params, snap_p):
                        p.copy_(q)  # in place: addresses (and the arenas the parameters live in) are unchanged
                    optimizer.restore(snap_o)
            with torch.no_grad():
                for b, q in zip(bufs, snap_b):
                    b.copy_(q)
            if drop_snap is not None:
                drop_state.copy_(drop_snap)
            elif (dev.type, dev.index) in ops.DropoutState._states:
                ops.DropoutState._states[(dev.type, dev.index)][1:2].zero_()
            # the restore (or any edit since the last forward) left the persistent bf16 weight shadows stale: bring them
            # up to date NOW, so that the capture below records no cast -- inside the graph cpc_adam_step keeps them
            # current (CpcAdamSpan.p16)
            ops.refresh_shadows(self.params)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize(dev)
        torch.cuda.set_rng_state(rng_state, dev)
        del snap_p, snap_o, snap_b
        self.graph = torch.cuda.CUDAGraph()
        self._zero()
        # the capture stream carries the dependent chain of the step; with ops.MAIN_PRIORITY < 0 its kernels are
        # scheduled ahead of the forked branches (weight gradients, bias sums, all-reduce, Adam: ops.side streams have
        # priority 0), which otherwise take SMs from the next link of the chain whenever both are ready
        cap = torch.cuda.Stream(device=dev, priority=ops.MAIN_PRIORITY) if ops.MAIN_PRIORITY < 0 else None
        with torch.cuda.graph(self.graph, **({"stream": cap} if cap is not None else {})):
            self.losses, self.acc = forwardBackward(cpcModel, cpcCriterion, self.static_seq, self.static_label)
            if after_backward is not None:
                after_backward()
            if optimizer is not None:
                optimizer.step()
        self._versions_seen = sum(p._version for p in self.params)

    def close(self):
        """Drop the captured graph (and the static outputs that live in its memory pool).  A graph that holds captured
        NCCL launches must be destroyed BEFORE the process group: ncclCommDestroy otherwise waits on it forever.
        Call on every rank, then dist.destroy_process_group()."""
        if self.graph is not None:
            torch.cuda.synchronize(self.static_seq.device)
            self.losses = self.acc = None
            self._zero()
            self.graph.reset()
            self.graph = None

    def _zero(self):
        for p in self.params:
            p.grad = None

    def prefetch(self, sequence):
        """Input staging: start the host -> device copy of the NEXT batch (pinned host memory) on a
        copy stream while the current replay runs; the following step() picks it up with a device-side copy."""
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.static_seq.device)
            self._staged = torch.empty_like(self.static_seq)
        cs = self._copy_stream
        if self._picked is not None:
            cs.wait_event(self._picked)  # the previous pick-up (not the replay that follows it) has read the buffer
        with torch.cuda.stream(cs):
            self._staged.copy_(sequence, non_blocking=True)
        self._staged_ready = True

    def pipelined(self, batches):
        """The batch loop of trainStep (cpc/train.py:84-122) as a three-deep pipeline: while replay i runs, batch i+1
        travels host -> device on the copy stream and the host reads the (losses, acc) of step i-1 from pinned memory
        -- every step's result still reaches the host, but the GPU never waits for the host to look at it.
        `batches`: iterable of pinned [B,2,1,W] host tensors.  Yields (losses [K], acc [K]) host tensors per step,
        in order (valid until the next-but-one yield)."""
        nK = self.losses.numel()
        out = [torch.empty(2, nK).pin_memory() for _ in range(2)]
        evs = [torch.cuda.Event(), torch.cuda.Event()]
        it = iter(batches)
        nxt = next(it, None)
        if nxt is None:
            return
        self.prefetch(nxt)
        i = 0
        while nxt is not None:
            losses, acc = self.step()
            nxt = next(it, None)
            if nxt is not None:
                self.prefetch(nxt)
            o = out[i & 1]
            o[0].copy_(losses.detach().view(-1), non_blocking=True)
            o[1].copy_(acc.detach().view(-1), non_blocking=True)
            evs[i & 1].record()
            if i > 0:
                evs[(i - 1) & 1].synchronize()
                yield out[(i - 1) & 1][0], out[(i - 1) & 1][1]
            i += 1
        evs[(i - 1) & 1].synchronize()
        yield out[(i - 1) & 1][0], out[(i - 1) & 1][1]

    def step(self, sequence=None, non_blocking=True):
        if sequence is not None:
            self.static_seq.copy_(sequence, non_blocking=non_blocking)
        elif self._staged_ready:
            torch.cuda.current_stream().wait_stream(self._copy_stream)
            self.static_seq.copy_(self._staged, non_blocking=True)
            self._picked = torch.cuda.Event()
            self._picked.record()
            self._staged_ready = False
        if self.optimizer is not None and hasattr(self.optimizer, "sync_hyper"):
            self.optimizer.sync_hyper()  # a scheduler may have changed lr since the last replay
        if self.graph is None:
            raise RuntimeError("GraphedTrainStep.step() after close()")
        if self.optimizer is not None:
            # parameters edited since the last replay (load_state_dict, a manual in-place change): the graph holds no
            # cast of the arena-homed weights, so refresh their bf16 shadows first.  One sum over the autograd version
            # counters per step (~10 us of host time, hidden behind the previous replay) when nothing changed.
            ver = sum(p._version for p in self.params)
            if ver != self._versions_seen:
                from . import ops
                ops.refresh_shadows(self.params)
                self._versions_seen = ver
        self.graph.replay()
        return self.losses, self.acc
