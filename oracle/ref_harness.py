"""Import the UNMODIFIED reference (MarvinLvn/CPC2) from /root/reference -- TEST INFRASTRUCTURE.

Used only by `tests/golden/make_golden.py`, in the build container, to generate
the committed golden vectors that pin `oracle/cpc_oracle.py`.  /root/reference
does not exist on the GPU box, so nothing under `tests/ -m gpu`, `smoke()` or
`bench.py` imports this file.

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


def import_reference():
    if not os.path.isdir(os.path.join(REFERENCE_ROOT, "cpc")):
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    for name in ("progressbar", "augment", "augment.effects", "torch_audiomentations"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    ta = sys.modules["torch_audiomentations"]
    for cls in ("Compose", "AddBackgroundNoise", "ApplyImpulseResponse"):
        if not hasattr(ta, cls):
            setattr(ta, cls, type(cls, (), {}))
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    if not torch.cuda.is_available() and not getattr(torch.ones, "_cpc_shim", False):
        _ones = torch.ones

        def ones(*a, **kw):
            if str(kw.get("device", "")) == "cuda":
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
