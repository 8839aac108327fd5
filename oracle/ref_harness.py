"""Import the UNMODIFIED reference (MarvinLvn/CPC2) from /root/reference -- TEST INFRASTRUCTURE.

Used by `tests/golden/make_golden.py`, in the build container, to generate the
committed golden vectors that pin `oracle/cpc_oracle.py`, and by the reference
arm of `bench.py` (`--impl reference`, and the `stock_gpu` comparator).
/root/reference does not exist on the GPU box: there the same files are imported
from `oracle/_ref/cpc_reference.zip`, the byte-for-byte archive that
`oracle/make_ref.py` (run by `build()`) packs from /root/reference -- git-ignored,
so the history holds no reference source.  Nothing under `cpc2_b200/` and no
`-m gpu` test imports this file.

No reference file is edited or copied.  Three things are shimmed in-process:
  * empty stub modules for `progressbar`, `augment`, `torch_audiomentations`
    (imported at module top by cpc/criterion/seq_alignment.py:5 and
    cpc/data_augmentation.py:8-9, never used on this path);
  * on a host without CUDA, `torch.ones(..., device='cuda')` is rewritten to
    'cpu' (hard-coded at cpc/criterion/criterion.py:340);
  * the four lines cpc/train.py:99-109 are driven directly because `trainStep`
    itself calls `.cuda()` (train.py:89-92).
"""
import argparse
import os
import sys
import types

import torch

REFERENCE_ROOT = os.environ.get("CPC2_REFERENCE_ROOT", "/root/reference")
ARCHIVE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "cpc_reference.zip")
ONES_ON_CPU = False


def reference_location():
    """The reference tree if it is on this machine, else the archive packed by oracle/make_ref.py, else None."""
    if os.path.isdir(os.path.join(REFERENCE_ROOT, "cpc")):
        return REFERENCE_ROOT
    if os.path.exists(ARCHIVE):
        return ARCHIVE
    return None


def import_reference():
    global REFERENCE_ROOT
    loc = reference_location()
    if loc is None:
        raise RuntimeError(f"reference not found: neither {REFERENCE_ROOT} nor {ARCHIVE} (python oracle/make_ref.py)")
    REFERENCE_ROOT = loc
    for name in ("progressbar", "augment", "augment.effects", "torch_audiomentations"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    ta = sys.modules["torch_audiomentations"]
    for cls in ("Compose", "AddBackgroundNoise", "ApplyImpulseResponse"):
        if not hasattr(ta, cls):
            setattr(ta, cls, type(cls, (), {}))
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    if not getattr(torch.ones, "_cpc_shim", False):
        _ones = torch.ones

        def ones(*a, **kw):
            # criterion.py:340 hard-codes device='cuda': on a host without CUDA, or while the reference is being run
            # on the CPU of a GPU box (ONES_ON_CPU, set by timed_reference_steps(device="cpu")), that means 'cpu'
            if str(kw.get("device", "")) == "cuda" and (ONES_ON_CPU or not torch.cuda.is_available()):
                kw["device"] = "cpu"
            return _ones(*a, **kw)

        ones._cpc_shim = True
        torch.ones = ones
    import cpc.model, cpc.criterion, cpc.feature_loader, cpc.train  # noqa: E401,F401
    from cpc.cpc_default_config import get_default_cpc_config
    return get_default_cpc_config


def reference_args(**over):
    get_default = import_reference()
    args = get_default()
    args.supervised, args.onEncoder, args.pathPhone, args.CTC = False, False, None, False
    for k, v in over.items():
        setattr(args, k, v)
    return args


def build_reference(args, seed=0):
    """Builds (CPCModel, CPCUnsupersivedCriterion) exactly as cpc/train.py:442-461 does."""
    import cpc.feature_loader as fl
    import cpc.model as model
    import cpc.train as train
    torch.manual_seed(seed)
    enc = fl.getEncoder(args)
    ar = fl.getAR(args)
    m = model.CPCModel(enc, ar)
    crit = train.getCriterion(args, enc.DOWNSAMPLING, 0, 0)
    return m, crit


def reference_step(m, crit, sequence, crit_seed=4321, train_mode=False, signal_quality=None):
    """cpc/train.py:95-109 on whatever device the modules live on. Returns dict.
    signal_quality [B, n]: per-window quality scores (train.py:104-106 passes them to the criterion)."""
    m.train(train_mode)
    crit.train(train_mode)
    for q in list(m.parameters()) + list(crit.parameters()):
        q.grad = None
    B = sequence.shape[0]
    past, future = sequence[:, 0, ...], sequence[:, 1, ...]
    combined = torch.cat([past, future], dim=0)
    label = torch.zeros(2 * B, dtype=torch.long)
    c, z, label = m(combined, label)
    c_p, z_f = c[:B], z[B:]
    torch.manual_seed(crit_seed)
    losses, acc = crit(c_p, z_f, label[:B], signal_quality)
    losses.sum().backward()
    grads = {k: v.grad.detach().clone() for k, v in m.named_parameters() if v.grad is not None}
    grads.update({k: v.grad.detach().clone() for k, v in crit.named_parameters() if v.grad is not None})
    return {"loss": losses.detach(), "acc": acc.detach(), "c": c.detach(), "z": z.detach(),
            "grads": grads}


def flat_state(m, crit):
    """Flat parameter dict with the key convention used by oracle/cpc_oracle.py:
    model keys as in CPCModel.state_dict() ('gEncoder.*', 'gAR.*'), criterion keys
    as in the criterion's state_dict ('wPrediction.*')."""
    out = {k: v.detach().clone() for k, v in m.state_dict().items()}
    out.update({k: v.detach().clone() for k, v in crit.state_dict().items()})
    return out


def timed_reference_steps(over, B, steps, warmup, device="cpu", autocast=None, tf32=True, seed=0):
    """Time `steps` iterations of the reference's own batch-loop body (cpc/train.py:95-109: concatenate past and
    future windows, CPCModel.forward, slice, criterion, `allLosses.sum().backward()`) in TRAIN mode (dropout on,
    as trainStep runs it), after `warmup` untimed ones, on the unmodified reference modules.  Returns
    (seconds per step, last mean loss).  device 'cuda': CUDA-event time; `autocast` = torch.bfloat16 wraps the
    forward in torch.autocast like mixed-precision users of the reference would; `tf32` sets the cuBLAS / cuDNN
    TF32 switches (the reference pins torch 1.3 where they did not exist; on Ampere+ TF32 convs are the default)."""
    import time
    global ONES_ON_CPU
    args = reference_args(**over)
    m, crit = build_reference(args, seed=seed)
    dev = torch.device(device)
    ONES_ON_CPU = dev.type == "cpu"
    m, crit = m.to(dev).train(), crit.to(dev).train()
    g = torch.Generator().manual_seed(1234)
    seq = (0.1 * torch.randn(B, 2, 1, args.sizeWindow, generator=g)).to(dev)
    label = torch.zeros(2 * B, dtype=torch.long, device=dev)
    params = list(m.parameters()) + list(crit.parameters())
    if dev.type == "cuda":
        torch.backends.cuda.matmul.allow_tf32 = bool(tf32)
        torch.backends.cudnn.allow_tf32 = bool(tf32)

    def one():
        for q in params:
            q.grad = None
        combined = torch.cat([seq[:, 0, ...], seq[:, 1, ...]], dim=0)
        if autocast is not None:
            with torch.autocast(dev.type, dtype=autocast):
                c, z, lab = m(combined, label)
                losses, acc = crit(c[:B], z[B:], lab[:B], None)
        else:
            c, z, lab = m(combined, label)
            losses, acc = crit(c[:B], z[B:], lab[:B], None)
        losses.sum().backward()
        return losses

    for _ in range(warmup):
        one()
    if dev.type == "cuda":
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(steps):
            losses = one()
        e1.record()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) / steps
        return e0.elapsed_time(e1) * 1e-3 / steps, float(losses.detach().mean()), wall
    try:
        t0 = time.perf_counter()
        for _ in range(steps):
            losses = one()
        sec = (time.perf_counter() - t0) / steps
    finally:
        ONES_ON_CPU = False
    return sec, float(losses.detach().mean()), sec
