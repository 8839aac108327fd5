"""The CUDA path against the UNMODIFIED reference running live on the same B200 ("on the GPU box the same
import runs the unmodified reference on CUDA: that is the parity oracle").

The reference travels as oracle/_ref/cpc_reference.zip, the byte-for-byte archive packed by oracle/make_ref.py in the
build container (`__graft_entry__.build()`); without it these tests skip -- the committed fixtures of
tests/golden/ (generated from the same reference) and the CPU oracle carry the parity claim on their own.

What this adds over the fixtures: the FULL default configuration (18.4 M parameters: too big to commit as a fixture)
and BASELINE.json's configs[2] / [3], each first trained by the reference's own optimiser on the GPU until the loss has
left the ln(NEG+1) plateau, then one step of both implementations on identical weights (load_state_dict), identical
input and -- same `torch.manual_seed` before the criterion, same two `torch.randint` calls on the same device --
identical sampled negatives.  The reference runs in fp32 with TF32 off.  Tolerances as everywhere
(tests/test_step_gpu.py): loss 2e-2, accuracy 1 %, activations relative L2 3e-2, gradient cosine >= 0.99 per tensor,
relaxed only to the bf16-storage noise floor of that very state where the floor is lower (the CPU oracle is evaluated
on the reference's weights with and without bf16 emulation, exactly as for the oracle cases).
"""
import os
import sys

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _reference():
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    from oracle import ref_harness as rh
    if rh.reference_location() is None:
        pytest.skip("reference archive oracle/_ref/cpc_reference.zip not packed (python oracle/make_ref.py)")
    return rh


def cosine(a, b):
    return F.cosine_similarity(a.double().flatten(), b.double().flatten(), dim=0).item()


def report(tag, **kw):
    import json
    d = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "parity_report.jsonl"), "a") as fh:
            fh.write(json.dumps(dict(case=tag, **kw)) + "\n")


def _predictable(B, W, seed):
    """Tonal synthetic audio: per window a few amplitude-modulated sinusoids (80-1600 Hz) + a little noise, the future
    window continuing the past one.  Unlike white noise its encoder features ARE predictable from their own past, so
    Adam at the reference's learning-rate scale (cpc_default_config: 2e-4) leaves the ln(NEG+1) plateau in tens of
    steps instead of collapsing (measured with the unmodified reference: lr 1e-3 / 5e-3 sits on the plateau)."""
    g = torch.Generator().manual_seed(seed)
    t = torch.arange(2 * W).float() / 16000.0
    M = 6
    f = 80.0 + 1500.0 * torch.rand(B, M, 1, generator=g)
    ph = 6.2832 * torch.rand(B, M, 1, generator=g)
    a = 0.02 + 0.05 * torch.rand(B, M, 1, generator=g)
    fm = 1.0 + 4.0 * torch.rand(B, M, 1, generator=g)  # slow amplitude modulation, Hz
    x = (a * (1 + 0.5 * torch.sin(6.2832 * fm * t)) * torch.sin(6.2832 * f * t + ph)).sum(1)
    x = x + 0.005 * torch.randn(B, 2 * W, generator=g)
    return x.view(B, 2, 1, W).contiguous()


def _ref_step(m, crit, seq, seed):
    """cpc/train.py:95-109 on the reference modules (eval mode), returning everything on the host."""
    for q in list(m.parameters()) + list(crit.parameters()):
        q.grad = None
    B = seq.shape[0]
    combined = torch.cat([seq[:, 0, ...], seq[:, 1, ...]], dim=0)
    label = torch.zeros(2 * B, dtype=torch.long, device=seq.device)
    c, z, label = m(combined, label)
    torch.manual_seed(seed)
    losses, acc = crit(c[:B], z[B:], label[:B], None)
    losses.sum().backward()
    grads = {k: v.grad.detach().float().cpu() for k, v in list(m.named_parameters()) + list(crit.named_parameters())
             if v.grad is not None}
    return dict(loss=losses.detach().cpu(), acc=acc.detach().cpu(), c=c.detach().cpu(), z=z.detach().cpu(), grads=grads)


CASES = {
    "default": (dict(), 8, 150),
    "config3_transformer_ar": (dict(arMode="transformer"), 8, 150),
    "config4_hidden512_lstm2_w40960": (dict(hiddenEncoder=512, hiddenGar=512, nLevelsGRU=2, sizeWindow=40960), 4, 150),
}


@pytest.mark.parametrize("name", list(CASES))
def test_step_matches_live_reference_away_from_init(name):
    rh = _reference()
    from cpc2_b200 import train as T
    over, B, max_steps = CASES[name]
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    rargs = rh.reference_args(**over)
    m, crit = rh.build_reference(rargs, seed=0)
    m, crit = m.cuda().eval(), crit.cuda().eval()
    if rargs.arMode != "transformer":
        m.train()  # cuDNN refuses an RNN backward in eval mode; the encoder / LSTM have no dropout or running statistics
    import math
    opt = torch.optim.Adam(list(m.parameters()) + list(crit.parameters()), lr=3e-4)
    train_seq = _predictable(B, rargs.sizeWindow, 777).cuda()
    target = 0.85 * math.log(rargs.negativeSamplingExt + 1)
    steps, loss = 0, float("nan")
    for steps in range(1, max_steps + 1):
        out = _ref_step(m, crit, train_seq, 1000 + steps) if steps % 10 == 0 else None
        if out is None:  # cheap path: no host copies
            for q in list(m.parameters()) + list(crit.parameters()):
                q.grad = None
            combined = torch.cat([train_seq[:, 0, ...], train_seq[:, 1, ...]], dim=0)
            c, z, lab = m(combined, torch.zeros(2 * B, dtype=torch.long, device="cuda"))
            torch.manual_seed(1000 + steps)
            losses, _ = crit(c[:B], z[B:], lab[:B], None)
            losses.sum().backward()
        else:
            loss = float(out["loss"].mean())
            if loss < target and steps >= 20:
                opt.step()
                break
        opt.step()
    # ---- identical weights into the drop-in modules
    args = T.defaultArgs(**over)
    torch.manual_seed(0)
    model, ours = T.buildModelAndCriterion(args)
    model, ours = model.cuda().eval(), ours.cuda().eval()
    missing = model.load_state_dict(m.state_dict(), strict=True)
    missing = ours.load_state_dict(crit.state_dict(), strict=True)
    seq = _predictable(B, args.sizeWindow, 1234).cuda()
    ref = _ref_step(m, crit, seq, 4321)
    for q in list(model.parameters()) + list(ours.parameters()):
        q.grad = None
    combined = torch.cat([seq[:, 0], seq[:, 1]], dim=0)
    c, z, label = model(combined, torch.zeros(2 * B, dtype=torch.long, device="cuda"))
    torch.manual_seed(4321)  # the same two torch.randint draws on the same device -> the same negatives
    losses, acc = ours(c[:B], z[B:], label[:B], None)
    losses.sum().backward()
    torch.cuda.synchronize()
    grads = {k: v.grad.detach().float().cpu() for k, v in list(model.named_parameters()) + list(ours.named_parameters())
             if v.grad is not None}
    loss_err = (losses.detach().cpu() - ref["loss"]).abs().max().item()
    acc_err = (acc.cpu() - ref["acc"]).abs().max().item()
    # ---- bf16-storage noise floor of this state: the CPU oracle on the reference's weights and the SAME negatives
    from oracle import cpc_oracle as oc
    S = args.sizeWindow // 160
    W = S - args.nPredicts
    torch.manual_seed(4321)
    n = args.negativeSamplingExt * W * B
    bi = torch.randint(low=0, high=B, size=(n,), device="cuda").cpu()
    si = torch.randint(low=1, high=S, size=(n,), device="cuda").cpu()
    state = {k: v.detach().float().cpu() for k, v in list(m.state_dict().items()) + list(crit.state_dict().items())}
    kw = dict(K=args.nPredicts, NEG=args.negativeSamplingExt, ar_mode=args.arMode, n_layers=args.nLevelsGRU,
              rnn_mode=args.rnnMode)
    torch.set_num_threads(os.cpu_count())
    exact = oc.train_step(state, seq.cpu(), bi, si, **kw)
    oc.EMULATE_BF16 = True
    try:
        emu = oc.train_step(state, seq.cpu(), bi, si, **kw)
    finally:
        oc.EMULATE_BF16 = False
    # (the oracle on these weights reproduces the live reference: same loss to fp32 rounding)
    assert (exact["loss"] - ref["loss"]).abs().max() <= 1e-3, (exact["loss"], ref["loss"])
    rows, bad = [], []
    for k, g in ref["grads"].items():
        assert k in grads, f"missing grad {k}"
        if g.abs().max() == 0:
            continue
        cs = cosine(grads[k], g)
        lim = min(0.99, cosine(emu["grads"][k], exact["grads"][k]) - 0.015)
        rows.append((k, round(cs, 5), round(lim, 4)))
        if cs < lim:
            bad.append(rows[-1])
    worst = min(rows, key=lambda r: r[1] - r[2])
    act = {}
    for nm, got, rf in (("z", z.detach().cpu(), ref["z"]), ("c", c.detach().cpu(), ref["c"])):
        dd = got.double() - rf.double()
        act[nm] = dict(rel_l2=float(dd.norm() / (rf.double().norm() + 1e-30)),
                       max_err_over_max=float(dd.abs().max() / (rf.abs().max() + 1e-30)))
    report("live_reference/" + name, trained_steps=steps, train_loss=loss, ref_loss=[round(float(x), 4) for x in ref["loss"].flatten()],
           loss_err=loss_err, acc_err=acc_err, worst=worst, n_below_099=sum(r[1] < 0.99 for r in rows), n_tensors=len(rows),
           failed=bad, act=act, min_cos=min(r[1] for r in rows))
    assert loss_err <= 2e-2, (losses, ref["loss"])
    assert acc_err <= 0.0101 + 1.0 / c.shape[1]
    assert act["z"]["rel_l2"] <= 3e-2 and act["c"]["rel_l2"] <= 3e-2, act
    assert not bad, f"gradient cosine vs the live reference below its limit (tensor, cosine, limit): {bad}"
