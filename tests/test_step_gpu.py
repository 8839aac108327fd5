"""Whole-path parity on a B200: the drop-in modules (cpc2_b200.model / .criterion), driven exactly like
cpc/train.py:95-109, against
  (a) the committed golden vectors produced by the UNMODIFIED reference (tests/golden/*.pt), and
  (b) the CPU oracle (oracle/cpc_oracle.py) on identical weights, inputs and sampled negative indices.

ONE stated tolerance (bf16 operands, fp32 accumulation / statistics): per-step loss abs <= 2e-2,
accuracy abs <= 1 % (argmax can flip on near-ties in bf16), activations z and c: relative L2 error <= 3e-2 (the
encoder output z, a feed-forward quantity, additionally max-abs error <= 3e-2 of its max; the recurrent c of a trained,
saturating LSTM can flip single elements by O(1) when h is fed back in bf16, so its max-abs error is only reported),
gradient cosine per parameter tensor >= MIN_COS = 0.99 against the fp32 reference / oracle gradients.  The only exception
is DERIVED, not typed in: where rounding GEMM operands and inter-layer gradients to bf16 -- which any bf16
implementation must do -- already pushes the exact algorithm below 0.99, the limit is that noise floor minus
FLOOR_MARGIN.  The floor is the cosine between the oracle evaluated with bf16 storage emulation
(oracle.cpc_oracle.EMULATE_BF16) and the fp32 gradients: read from tests/golden/grad_floor.json
(tools/grad_floor_probe.py) for the committed fixtures, computed on the spot for the oracle cases.  It only bites
on fixtures taken AT RANDOM INITIALISATION (every logit ~0, encoder gradients are cancelling sums); the `*_trained`
fixtures and the warmed-up oracle cases sit on the 0.99 line.  Sampled indices are bit exact
(tests/test_kernels_gpu.py checks the index kernel).  Every comparison is appended to
gpurun_out/parity_report.jsonl (worst cosine, its limit, loss / accuracy errors) so the numbers behind a green run
can be read afterwards.
"""
import json
import os

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MIN_COS = 0.99
FLOOR_MARGIN = 0.015
with open(os.path.join(GOLD, "grad_floor.json")) as _fh:
    GRAD_FLOOR = json.load(_fh)


def report(tag, **kw):
    d = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "parity_report.jsonl"), "a") as fh:
            fh.write(json.dumps(dict(case=tag, **kw)) + "\n")


def _need_gpu():
    if not torch.cuda.is_available():
        pytest.skip("no GPU")


def build(args, device="cuda"):
    from cpc2_b200 import train as T
    torch.manual_seed(0)
    model, crit = T.buildModelAndCriterion(args)
    return model.to(device), crit.to(device)


def cosine(a, b):
    return F.cosine_similarity(a.double().flatten(), b.double().flatten(), dim=0).item()


def run_step(model, crit, seq, ext=None, train_mode=False, seed=4321, signal_quality=None, collect=True):
    """cpc/train.py:95-109.  If ext is given the sampler is bypassed so both sides use identical indices.
    collect=False: leave everything on the device (no host copies of c / z / gradients)."""
    model.train(train_mode)
    crit.train(train_mode)
    for p in list(model.parameters()) + list(crit.parameters()):
        p.grad = None
    B = seq.shape[0]
    captured = {}
    orig = crit.sampleIndices

    def sampler(*a):
        e = ext.to(seq.device) if ext is not None else orig(*a)
        captured["ext"] = e
        return e

    crit.sampleIndices = sampler
    try:
        combined = torch.cat([seq[:, 0], seq[:, 1]], dim=0)
        label = torch.zeros(2 * B, dtype=torch.long, device=seq.device)
        c, z, label = model(combined, label)
        torch.manual_seed(seed)
        losses, acc = crit(c[:B], z[B:], label[:B], signal_quality)
        losses.sum().backward()
    finally:
        crit.sampleIndices = orig
    if not collect:
        return dict(loss=losses.detach())
    grads = {k: v.grad.detach().float().cpu() for k, v in model.named_parameters() if v.grad is not None}
    grads.update({k: v.grad.detach().float().cpu() for k, v in crit.named_parameters() if v.grad is not None})
    return dict(loss=losses.detach().cpu(), acc=acc.detach().cpu(), c=c.detach().cpu(), z=z.detach().cpu(),
                grads=grads, ext=captured["ext"].cpu())


def zero_gradient_biases(norm_mode, bn_training=True):
    """Conv biases in front of a norm layer that removes the column mean (batch statistics, instance statistics) have a
    mathematically ZERO gradient: the reference's autograd leaves fp32 rounding noise there (|g| ~ 1e-9, cosine floor ~ 0,
    tests/golden/grad_floor.json) and cpc_colnorm_bwd writes exact zeros.  Nothing to compare."""
    if norm_mode == "instanceNorm" or (norm_mode == "batchNorm" and bn_training):
        return {f"gEncoder.conv{i}.bias" for i in range(5)}
    return set()


def check_against(out, ref_loss, ref_acc, ref_c, ref_z, ref_grads, floors=None, strides=None, tag="", loss_floor=0.0,
                  skip=()):
    """`floors`: per-tensor bf16-storage noise floor (see the module docstring); None = the plain 0.99 line.
    `strides`: fixtures that store big gradients as g.flatten()[::stride].
    `loss_floor`: the same floor for the loss -- |loss of the exact oracle - loss of the oracle with bf16 operands| at
    THIS state.  The warm-up that precedes the oracle cases is chaotic (fp32 atomics order -> a weight crossing a bf16
    rounding boundary -> another trajectory) and now and then ends in a state whose loss itself moves by 0.03 under
    bf16 storage (measured: conv8 heads, gradient floors down to 0.70 in the same state); the limit is 2e-2 unless
    twice that state's own floor is larger, derived at run time like the gradient limits."""
    loss_err = (out["loss"] - ref_loss).abs().max().item()
    acc_err = (out["acc"] - ref_acc).abs().max().item()
    rows, bad = [], []
    for k, g in ref_grads.items():
        assert k in out["grads"], f"missing grad {k}"
        if g.abs().max() == 0:
            continue
        if k in skip:
            wk = k.replace(".bias", ".weight")
            assert out["grads"][k].abs().max() <= 1e-4 * out["grads"][wk].abs().max(), f"{k} should vanish"
            continue
        st = (strides or {}).get(k, 1)
        got = out["grads"][k].flatten()[::st] if st > 1 else out["grads"][k]
        cs = cosine(got, g)
        lim = MIN_COS if floors is None else min(MIN_COS, floors.get(k, 1.0) - FLOOR_MARGIN)
        rows.append((k, round(cs, 5), round(lim, 4)))
        if cs < lim:
            bad.append((k, round(cs, 4), round(lim, 4)))
    worst = min(rows, key=lambda r: r[1] - r[2]) if rows else None
    act = {}
    for name, got, ref in (("z", out["z"], ref_z), ("c", out["c"], ref_c)):
        d = (got.double() - ref.double())
        act[name] = dict(rel_l2=float(d.norm() / (ref.double().norm() + 1e-30)),
                         max_err_over_max=float(d.abs().max() / (ref.abs().max() + 1e-30)))
    report(tag, loss_err=loss_err, acc_err=acc_err, worst=worst, min_cos=min(r[1] for r in rows) if rows else None,
           n_below_099=sum(r[1] < MIN_COS for r in rows), n_tensors=len(rows), failed=bad, act=act)
    assert loss_err <= max(2e-2, 2.0 * loss_floor), (out["loss"], ref_loss, loss_floor)
    assert acc_err <= 0.0101 + 1.0 / ref_c.shape[1], (out["acc"], ref_acc)
    assert act["z"]["rel_l2"] <= 3e-2 and act["z"]["max_err_over_max"] <= 3e-2 + 1e-3, act
    assert act["c"]["rel_l2"] <= 3e-2, act
    assert not bad, f"gradient cosine below its limit (tensor, cosine, limit): {bad}"


GOLDEN_CASES = ["tiny_lstm_linear", "tiny_lstm_reverse_linear", "tiny_transformer_ar_linear", "tiny_gru2_transformer",
                "tiny_lstm_multihead", "tiny_lstm_ffd", "tiny_lstm_conv4", "tiny_lstm_headsLSTM", "tiny_lstm_headsRNN",
                "tiny_lstm_quality_skip1"]
# (tiny_rnn_hneq_linear pins the oracle only: its hidden size 24 is not one the recurrence kernels are built for)
# the same configurations after the reference's Adam has left the ln(NEG+1) plateau, plus the default width
TRAINED_CASES = [c + "_trained" for c in GOLDEN_CASES if c != "tiny_lstm_quality_skip1"] + ["mid256_lstm_transformer_trained"]
# the encoder's other norm layers (cpc/model.py:71-82): batchNorm in train and eval mode, instanceNorm, ID
NORM_CASES = ["tiny_batchnorm_train_linear", "tiny_batchnorm_eval_linear", "tiny_instancenorm_linear", "tiny_idnorm_linear"]


@pytest.mark.parametrize("name", GOLDEN_CASES + TRAINED_CASES + NORM_CASES)
def test_against_reference_golden(name):
    """CUDA path vs vectors of the UNMODIFIED reference: identical weights, inputs and sampled indices."""
    _need_gpu()
    from cpc2_b200 import train as T
    fx = torch.load(os.path.join(GOLD, name + ".pt"), weights_only=False)
    args = T.defaultArgs(**fx["args"])
    model, crit = build(args)
    ms = {k: v.float() for k, v in fx["state"].items() if k.startswith("g")}
    cs = {k: v.float() for k, v in fx["state"].items() if k.startswith("wPrediction")}
    model.load_state_dict(ms, strict=False)
    missing = crit.load_state_dict(cs, strict=False)
    assert not missing.unexpected_keys
    quality = fx["signal_quality"].cuda() if "signal_quality" in fx else None
    train = fx.get("train_mode", False)
    out = run_step(model, crit, fx["sequence"].cuda(), ext=fx["ext_idx"].long(), signal_quality=quality, train_mode=train)
    check_against(out, fx["loss"], fx["acc"], fx["c"], fx["z"], fx["grads"], floors=GRAD_FLOOR[name],
                  strides=fx.get("grad_stride"), tag="golden/" + name,
                  skip=zero_gradient_biases(fx["args"].get("normMode", "layerNorm"), train))
    if train:  # nn.BatchNorm1d's running estimates after the step (momentum 0.1, unbiased variance) and its step count
        st = model.state_dict()
        for k, v in fx["running_after"].items():
            if k.endswith("num_batches_tracked"):
                assert int(st[k]) == int(v), k
            else:
                err = (st[k].float().cpu() - v).abs().max().item()
                assert err <= 2e-2 * max(v.abs().max().item(), 1e-3), (k, err)


def _oracle_kwargs(args):
    return dict(K=args.nPredicts, NEG=args.negativeSamplingExt, ar_mode=args.arMode, n_layers=args.nLevelsGRU,
                rnn_mode=args.rnnMode, cpc_mode=args.cpc_mode, abspos=args.abspos, n_skipped=args.n_skipped,
                multihead=args.multihead_rnn, norm_mode=args.normMode)


def _state_for_oracle(model, crit):
    st = {k: v.detach().float().cpu() for k, v in model.state_dict().items()}
    st.update({k: v.detach().float().cpu() for k, v in crit.state_dict().items()})
    return st


ORACLE_CASES = {
    # BASELINE.json configs[1]: the CPC default
    "default_lstm_transformer_heads": dict(),
    "default_linear_heads": dict(rnnMode="linear"),
    # configs[2] as BASELINE.json states it: transformer AR (feature_loader.py:216-220) AND the default transformer heads
    "config3_transformer_ar_transformer_heads": dict(arMode="transformer"),
    "transformer_ar": dict(arMode="transformer", rnnMode="linear"),
    "transformer_ar_abspos": dict(arMode="transformer", abspos=True, rnnMode="linear"),
    # configs[3]: hidden 512, 2-layer LSTM, window 40960 with the default transformer heads (dk = 64, W' = 244)
    "config4_hidden512_lstm2_window40960_transformer_heads": dict(hiddenEncoder=512, hiddenGar=512, nLevelsGRU=2,
                                                                  sizeWindow=40960),
    "lstm2_hidden512_window40960": dict(hiddenEncoder=512, hiddenGar=512, nLevelsGRU=2, sizeWindow=40960,
                                        rnnMode="linear"),
    "reverse_mode_default_size": dict(cpc_mode="reverse"),
    "gru_ar_default_size": dict(arMode="GRU", rnnMode="linear"),
    "gru2_hidden_neq_encoder": dict(arMode="GRU", nLevelsGRU=2, hiddenEncoder=64, hiddenGar=128, nPredicts=4,
                                    negativeSamplingExt=16, sizeWindow=5120, rnnMode="linear"),
    "rnn_ar": dict(arMode="RNN", hiddenEncoder=64, hiddenGar=64, nPredicts=4, negativeSamplingExt=16,
                   sizeWindow=5120, rnnMode="linear"),
    "small_transformer_heads": dict(hiddenEncoder=64, hiddenGar=64, nPredicts=4, negativeSamplingExt=16,
                                    sizeWindow=5120),
    "multihead_rnn_default_size": dict(multihead_rnn=True),
    "ffd_heads_default_size": dict(rnnMode="ffd"),
    "conv8_heads": dict(rnnMode="conv8", hiddenEncoder=128, hiddenGar=128, nPredicts=6, negativeSamplingExt=32,
                        sizeWindow=10240),
    "conv4_heads_default_size": dict(rnnMode="conv4"),
    "lstm_heads": dict(rnnMode="LSTM", hiddenEncoder=128, hiddenGar=128, nPredicts=4, negativeSamplingExt=32,
                       sizeWindow=10240),
    "rnn_heads": dict(rnnMode="RNN", hiddenEncoder=128, hiddenGar=128, nPredicts=4, negativeSamplingExt=32,
                      sizeWindow=10240),
    # the encoder's other norm layers at the default width (cpc/model.py:71-82).  `_train`: the compared step runs in
    # TRAIN mode (batch statistics; linear heads and an LSTM have no dropout, so nothing else changes); every batchNorm
    # case is warmed up in train mode, so its running estimates have left (0, 1) before the eval-mode comparison
    "batchnorm_encoder_train_mode": dict(normMode="batchNorm", rnnMode="linear", _train=True),
    "batchnorm_encoder_eval_mode": dict(normMode="batchNorm", rnnMode="linear"),
    "instancenorm_encoder": dict(normMode="instanceNorm", rnnMode="linear"),
    "instancenorm_encoder_hidden512_window40960": dict(normMode="instanceNorm", rnnMode="linear", hiddenEncoder=512,
                                                       hiddenGar=512, sizeWindow=40960),
    # no normalisation: with the reference's default init every conv + ReLU shrinks the activations 0.41x, z ends up at
    # 1e-2 of the waveform and the loss never leaves ln(NEG+1) (unmodified reference on this batch, Adam 3e-4 ... 2e-3,
    # 150 steps: 4.86 -> 4.82) -- every gradient stays a cancelling sum.  `_conv_gain` rescales the five conv weights to
    # variance-preserving (He) size before the warm-up, after which the reference leaves the plateau within 70 steps.
    "id_encoder": dict(normMode="ID", rnnMode="linear", _conv_gain=6 ** 0.5),
}
# cases that are ALSO compared at random initialisation (the regime round 1 tested); every case is compared after the
# warm-up below
AT_INIT = ("default_lstm_transformer_heads", "default_linear_heads", "transformer_ar", "gru2_hidden_neq_encoder")
WARM_STEPS, WARM_LR = 150, 3e-4


def predictable_sequence(B, W, seed):
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


def warm_up(model, crit, args, B, train_mode=False):
    """Move the model off the ln(NEG+1) plateau with torch.optim.Adam on a predictable batch (the regime the
    reference trains in: logits no longer ~0, gradients no longer cancelling sums).  Runs on the CUDA path itself --
    the comparison that follows is against the oracle evaluated at whatever state this leaves."""
    import math
    opt = torch.optim.Adam(list(model.parameters()) + list(crit.parameters()), lr=WARM_LR)
    seq = predictable_sequence(B, args.sizeWindow, 777).cuda()
    target = 0.85 * math.log(args.negativeSamplingExt + 1)
    loss = float("nan")
    for it in range(WARM_STEPS):
        out = run_step(model, crit, seq, seed=1000 + it, collect=False, train_mode=train_mode)
        opt.step()
        if it % 10 == 9:
            loss = float(out["loss"].mean())
            assert loss == loss, "NaN during warm-up"
            if loss < target and it >= 20:
                break
    return it + 1, loss


@pytest.mark.parametrize("name", [n + "@init" for n in AT_INIT] + list(ORACLE_CASES))
def test_against_oracle(name):
    _need_gpu()
    from cpc2_b200 import train as T
    from oracle import cpc_oracle as oc
    at_init = name.endswith("@init")
    over = dict(ORACLE_CASES[name.split("@")[0]])
    train = over.pop("_train", False)
    conv_gain = over.pop("_conv_gain", None)
    args = T.defaultArgs(**over)
    B = 8 if args.sizeWindow <= 20480 else 4
    model, crit = build(args)
    if conv_gain is not None:
        with torch.no_grad():
            for i in range(5):
                getattr(model.gEncoder, f"conv{i}").weight.mul_(conv_gain)
    warm = (0, None) if at_init else warm_up(model, crit, args, B, train_mode=args.normMode == "batchNorm")
    seq = predictable_sequence(B, args.sizeWindow, 1234) if not at_init else \
        0.1 * torch.randn(B, 2, 1, args.sizeWindow, generator=torch.Generator().manual_seed(1234))
    running = {k: v.detach().float().cpu().clone() for k, v in model.state_dict().items() if ".running_" in k}
    out = run_step(model, crit, seq.cuda(), train_mode=train)
    S = args.sizeWindow // 160
    W = S - args.nPredicts
    # the oracle consumes the very indices the GPU criterion drew: feed ext through (batch_idx, seq_idx)
    ext = out["ext"]
    t = torch.arange(W).view(1, 1, W).expand(B, args.negativeSamplingExt, W).reshape(-1)
    batch_idx = ext // S
    seq_idx = torch.remainder(ext % S - t, S)
    seq_idx = torch.where(seq_idx == 0, torch.full_like(seq_idx, S), seq_idx)  # any representative works
    assert torch.equal(oc.negative_ext_index(batch_idx, seq_idx, B, S, W, args.negativeSamplingExt), ext)
    torch.set_num_threads(os.cpu_count())
    st = _state_for_oracle(model, crit)
    ref = oc.train_step(st, seq, batch_idx, seq_idx, bn_training=train, **_oracle_kwargs(args))
    # bf16-storage noise floor of THIS state, per tensor (module docstring): the exact algorithm with bf16 operands
    oc.EMULATE_BF16 = True
    try:
        emu = oc.train_step(st, seq, batch_idx, seq_idx, bn_training=train, **_oracle_kwargs(args))
    finally:
        oc.EMULATE_BF16 = False
    floors = {k: cosine(emu["grads"][k], g) for k, g in ref["grads"].items() if g.abs().max() > 0}
    loss_floor = float((emu["loss"] - ref["loss"]).abs().max())
    report("oracle/" + name + "/setup", warm_steps=warm[0], warm_loss=warm[1], loss=[round(float(x), 4) for x in ref["loss"].flatten()],
           loss_floor=loss_floor, min_floor=min(floors.values()), floors_below_099=sorted((round(v, 4), k) for k, v in floors.items() if v < MIN_COS))
    check_against(out, ref["loss"], ref["acc"], ref["c"], ref["z"], ref["grads"], floors=floors, tag="oracle/" + name,
                  loss_floor=loss_floor, skip=zero_gradient_biases(args.normMode, train))
    if train and args.normMode == "batchNorm":  # the running estimates moved as nn.BatchNorm1d moves them
        for i, (mean, var) in enumerate(ref["bn_stats"]):
            for key, stat in ((f"gEncoder.batchNorm{i}.running_mean", mean), (f"gEncoder.batchNorm{i}.running_var", var)):
                want = 0.9 * running[key] + 0.1 * stat
                err = (st[key] - want).abs().max().item()
                assert err <= 2e-2 * max(want.abs().max().item(), 1e-3), (key, err)


def test_default_init_matches_reference_fingerprint():
    """Same seed -> same initial weights as the reference constructors (golden fingerprint), so the
    full-size golden losses of the reference apply to this build."""
    _need_gpu()
    from cpc2_b200 import train as T
    fx = torch.load(os.path.join(GOLD, "default_b8_summary.pt"), weights_only=False)
    torch.manual_seed(0)
    model, crit = T.buildModelAndCriterion(T.defaultArgs())
    st = dict(model.state_dict())
    st.update(crit.state_dict())
    for k, (s, a) in fx["fingerprint"].items():
        v = st[k].double()
        assert abs(float(v.sum()) - s) <= 1e-6 * max(1.0, a) and abs(float(v.abs().sum()) - a) <= 1e-6 * max(1.0, a), k
    model, crit = model.cuda(), crit.cuda()
    g = torch.Generator().manual_seed(1234)
    seq = 0.1 * torch.randn(8, 2, 1, 20480, generator=g)
    out = run_step(model, crit, seq.cuda())
    # different device RNG -> different negatives than the CPU golden, but at init every logit is ~0, so the
    # per-step losses of the reference (ln 129 = 4.8598) must be reproduced to 1e-2
    assert (out["loss"] - fx["loss"]).abs().max() < 1e-2, out["loss"]
    err = (out["z"][:, ::16, ::32] - fx["z_slice"]).abs().max().item()
    assert err <= 3e-2 * fx["z_slice"].abs().max().item()
    err = (out["c"][:, ::16, ::32] - fx["c_slice"]).abs().max().item()
    assert err <= 3e-2 * fx["c_slice"].abs().max().item() + 1e-3


def test_train_mode_dropout_runs_and_learns():
    """Train mode (dropout 0.1 in the heads): Adam steps on a fixed batch reduce the loss.  The loss sits on a
    plateau near ln(17) for ~50 steps before it drops (the escape time depends on the dropout masks), hence 160."""
    _need_gpu()
    from cpc2_b200 import train as T
    args = T.defaultArgs(hiddenEncoder=64, hiddenGar=64, nPredicts=4, negativeSamplingExt=16, sizeWindow=5120)
    model, crit = build(args)
    opt = torch.optim.Adam(list(model.parameters()) + list(crit.parameters()), lr=2e-3)
    g = torch.Generator().manual_seed(7)
    base = torch.randn(8, 1, 1, args.sizeWindow, generator=g) * 0.1
    seq = torch.cat([base, base], dim=1).cuda()  # past == future: predictable
    hist = []
    for it in range(160):
        out = run_step(model, crit, seq, train_mode=True, seed=it)
        opt.step()
        hist.append(out["loss"].mean().item())
    first, last = sum(hist[:5]) / 5, sum(hist[-5:]) / 5
    assert all(h == h for h in hist) and last < first - 0.05, (first, last)


@pytest.mark.parametrize("over", [dict(), dict(arMode="transformer", rnnMode="linear"),
                                  dict(arMode="GRU", nLevelsGRU=2, hiddenEncoder=64, hiddenGar=128, nPredicts=4,
                                       negativeSamplingExt=16, sizeWindow=5120, rnnMode="linear")])
def test_past_future_split_matches_concatenated_call(over):
    """train.forwardBackward runs the model as two halves (CPCModel.forwardPastFuture: past -> encoder -> AR,
    future -> encoder on a side stream, no AR over the future rows).  Same losses, accuracies and parameter
    gradients as the literal concatenated call of cpc/train.py:95-104 (differences: fp32 summation order only)."""
    _need_gpu()
    from cpc2_b200 import train as T
    args = T.defaultArgs(**over)
    model, crit = build(args)
    model.eval()
    crit.eval()
    g = torch.Generator().manual_seed(5)
    B = 8
    seq = (0.1 * torch.randn(B, 2, 1, args.sizeWindow, generator=g)).cuda()
    label = torch.zeros(B, dtype=torch.long, device="cuda")
    from cpc2_b200 import ops

    def both(force_bn):
        res = {}
        ops.NORM_FORCE_BN = force_bn
        try:
            for split in (False, True):
                T.SPLIT_PAST_FUTURE = split
                for p in list(model.parameters()) + list(crit.parameters()):
                    p.grad = None
                torch.manual_seed(99)
                losses, acc = T.forwardBackward(model, crit, seq, label)
                torch.cuda.synchronize()
                grads = {k: v.grad.detach().float().clone()
                         for k, v in list(model.named_parameters()) + list(crit.named_parameters()) if v.grad is not None}
                res[split] = (losses.detach().clone(), acc.detach().clone(), grads)
        finally:
            T.SPLIT_PAST_FUTURE = True
            ops.NORM_FORCE_BN = 0
        assert set(res[True][2]) == set(res[False][2])
        worst = []
        for k, gr in res[False][2].items():
            gs = res[True][2][k]
            if gr.abs().max() == 0:
                assert gs.abs().max() == 0, k
                continue
            worst.append((cosine(gs, gr), float((gs - gr).abs().max() / gr.abs().max()), k))
        return res, sorted(worst)

    # (1) The same arithmetic in the same order.  The fused conv + ChannelNorm GEMM picks its tile form by the number
    # of row tiles of the launch (one 256-column tile per row, or the row split over a 4-CTA cluster that exchanges
    # partial sums, csrc/gemm.cu), and 2 x 8 windows cross that threshold where 16 windows do not; with one form pinned
    # for both calls (ops.NORM_FORCE_BN) the two halves differ from the concatenated call by the fp32 summation order
    # of the atomics only.
    res, worst = both(256)
    assert torch.allclose(res[True][0], res[False][0], atol=1e-5), (res[True][0], res[False][0])
    # (argmax ties at init can flip with the fp32 summation order of differently tiled launches: a few tokens of B*W')
    assert (res[True][1] - res[False][1]).abs().max() <= 0.01, (res[True][1], res[False][1])
    bad = [w for w in worst if not (w[0] > 0.99999 and w[1] <= 5e-3)]
    assert not bad, bad[:5]
    # (2) The shipped tile forms: row statistics summed in a different order flip the bf16 rounding of a few
    # activations, and with them a few ReLU gates downstream (the FFN bias gradients, conv0's cancelling sum show it
    # most: measured cosine 0.99990, largest entry error 6 %).  A sanity bound, far inside the parity tolerance
    # against the reference (cosine 0.99), and the measured numbers go to the parity report.
    res, worst = both(0)
    assert torch.allclose(res[True][0], res[False][0], atol=1e-3), (res[True][0], res[False][0])
    report("past_future_split_shipped_tiles", over=over, min_cos=worst[0][0], max_rel_err=max(w[1] for w in worst),
           worst=[w[2] for w in worst[:3]])
    assert worst[0][0] > 0.9995, worst[:5]


def test_fused_adam_matches_torch_adam():
    """cpc2_b200.optim.FusedAdam against torch.optim.Adam (cpc/train.py:472-479) over 5 steps on the same gradients:
    parameters and moments agree to fp32 rounding; the state_dict round-trips into torch.optim.Adam and back."""
    _need_gpu()
    from cpc2_b200.optim import FusedAdam
    torch.manual_seed(3)
    shapes = [(256, 256, 8), (256,), (1, 256, 1), (37,), (3, 5, 7), (4099,), (1024, 768)]
    ours = [torch.nn.Parameter(torch.randn(*s, device="cuda")) for s in shapes]
    ref = [torch.nn.Parameter(p.detach().clone()) for p in ours]
    oa = FusedAdam(ours, lr=2e-4, betas=(0.9, 0.999), eps=1e-8)
    ob = torch.optim.Adam(ref, lr=2e-4, betas=(0.9, 0.999), eps=1e-8)
    sched_a = torch.optim.lr_scheduler.StepLR(oa, 2, gamma=0.5)
    sched_b = torch.optim.lr_scheduler.StepLR(ob, 2, gamma=0.5)
    for it in range(5):
        for p, q in zip(ours, ref):
            g = torch.randn_like(p) * (10.0 ** (it - 2))
            p.grad, q.grad = g.clone(), g.clone()
        oa.step()
        ob.step()
        oa.zero_grad()
        ob.zero_grad()
        sched_a.step()
        sched_b.step()
        for p, q in zip(ours, ref):
            assert torch.allclose(p, q, rtol=2e-6, atol=1e-7), (it, (p - q).abs().max())
    sa, sb = oa.state_dict(), ob.state_dict()
    for k in sb["state"]:
        assert float(sa["state"][k]["step"]) == float(sb["state"][k]["step"]) == 5
        assert torch.allclose(sa["state"][k]["exp_avg"], sb["state"][k]["exp_avg"], rtol=1e-5, atol=1e-9)
        assert torch.allclose(sa["state"][k]["exp_avg_sq"], sb["state"][k]["exp_avg_sq"], rtol=1e-5, atol=1e-12)
    # round trip: torch state -> FusedAdam, one more step on both
    oc = FusedAdam(ours, lr=2e-4)
    oc.load_state_dict(sb)
    for p, q in zip(ours, ref):
        p.data.copy_(q.data)
        g = torch.randn_like(p)
        p.grad, q.grad = g.clone(), g.clone()
    oc.step()
    ob.step()
    for p, q in zip(ours, ref):
        assert torch.allclose(p, q, rtol=2e-6, atol=1e-7)


def test_staged_adam_equals_single_launch():
    """FusedAdam.attach(): stepping each gradient bucket from the ops.grad_ready hook (side stream, one counter
    advance per iteration) gives bit-identical parameters to one launch after backward -- also when the heads announce
    their feed-forward gradients early (ops.EARLY_FFN: stages criterion.ffn / .front / .tail)."""
    _need_gpu()
    from cpc2_b200 import ops
    from cpc2_b200 import train as T
    from cpc2_b200.optim import FusedAdam
    args = T.defaultArgs(hiddenEncoder=64, hiddenGar=64, nPredicts=4, negativeSamplingExt=16, sizeWindow=5120)
    g = torch.Generator().manual_seed(3)
    seq = (0.1 * torch.randn(4, 2, 1, args.sizeWindow, generator=g)).cuda()
    label = torch.zeros(4, dtype=torch.long, device="cuda")
    finals, after2 = [], []
    seen = []
    spy = lambda stage, gflat, params=None, grads=None: seen.append(stage)
    for staged in (False, True, "early"):
        model, crit = build(args)
        model.eval()
        crit.eval()
        opt = FusedAdam(list(crit.parameters()) + list(model.parameters()), lr=1e-3)
        if staged:
            opt.attach()
        early0 = ops.EARLY_FFN
        ops.EARLY_FFN = "1" if staged == "early" else "0"
        ops.GRAD_READY.append(spy)
        try:
            for it in range(3):
                torch.manual_seed(50 + it)
                del seen[:]
                T.forwardBackward(model, crit, seq, label)
                opt.step()
                opt.zero_grad()
                if it == 1:
                    torch.cuda.synchronize()
                    after2.append([p.detach().clone() for p in list(crit.parameters()) + list(model.parameters())])
        finally:
            opt.detach()
            ops.GRAD_READY.remove(spy)
            ops.EARLY_FFN = early0
        torch.cuda.synchronize()
        assert float(opt._groups[0]["state"][0]) == 3
        if staged == "early":
            assert seen[:3] == ["criterion.ffn", "criterion.front", "criterion.tail"], seen
        else:
            assert seen[0] == "criterion", seen
        finals.append([p.detach().clone() for p in list(crit.parameters()) + list(model.parameters())])
    # Two iterations: the gradients of the three runs differ by the fp32 order of the atomics only (measured <= 1e-6 of a
    # tensor's largest entry), and so do the parameters -- a second counter advance, a bucket stepped twice or not at all
    # would show here as a different bias correction (1e-3 per element).
    for other in after2[1:]:
        for a, b in zip(after2[0], other):
            assert torch.allclose(a, b, rtol=1e-4, atol=1e-6)
    # Third iteration: that 1e-6 noise can move a parameter across a bf16 rounding boundary of its weight shadow; one
    # flipped weight (2^-8 relative) changes the step's activations and gradients by up to 0.5 % in a few entries, in
    # the single-launch run as well as in the staged ones (tools/dbg_staged.py: two repeatable outcomes, each run lands
    # on either; 9 of 16 single-launch repeats left the 1e-4 band against another single-launch run).  An Adam step moves
    # an entry by at most ~lr, so the runs stay within 2 lr of each other and equal in all but a few entries.
    for other in finals[1:]:
        for a, b in zip(finals[0], other):
            assert (a - b).abs().max().item() <= 2e-3
            assert ((a - b).abs() > 1e-6 + 1e-4 * b.abs()).float().mean().item() <= 0.2  # (measured: 2 of 64 entries)


@pytest.mark.parametrize("norm", ["layerNorm", "batchNorm", "instanceNorm"])
def test_graphed_step_with_fused_adam_learns(norm):
    """GraphedTrainStep(optimizer=FusedAdam): forward + backward + Adam in one CUDA-graph replay; the loss on a
    predictable batch (past == future) falls, and the device-side step counter advances once per replay.  With a
    batchNorm encoder (not splittable: the literal concatenated call is captured) the running estimates are untouched by
    the construction and move once per replay."""
    _need_gpu()
    from cpc2_b200 import train as T
    from cpc2_b200.optim import FusedAdam
    args = T.defaultArgs(hiddenEncoder=64, hiddenGar=64, nPredicts=4, negativeSamplingExt=16, sizeWindow=5120, normMode=norm)
    model, crit = build(args)
    assert model.splittable() == (norm == "layerNorm")
    opt = FusedAdam(list(crit.parameters()) + list(model.parameters()), lr=2e-3)
    g = torch.Generator().manual_seed(7)
    base = torch.randn(8, 1, 1, args.sizeWindow, generator=g) * 0.1
    seq = torch.cat([base, base], dim=1).cuda()
    opt.attach()  # staged: each module's parameters are stepped as soon as its gradients exist
    try:
        step = T.GraphedTrainStep(model, crit, 8, args.sizeWindow, warmup=2, optimizer=opt)
    finally:
        opt.detach()
    n0 = float(opt._groups[0]["state"][0])
    if norm == "batchNorm":
        bn = model.gEncoder.batchNorm2
        assert int(bn.num_batches_tracked) == 0 and bn.running_mean.abs().max() == 0 and (bn.running_var == 1).all()
    hist = []
    for it in range(160):
        losses, _ = step.step(seq)
        hist.append(losses.mean().item())
    assert float(opt._groups[0]["state"][0]) == n0 + 160
    if norm == "batchNorm":
        assert int(bn.num_batches_tracked) == 160 and bn.running_mean.abs().max() > 0
    first, last = sum(hist[:5]) / 5, sum(hist[-5:]) / 5
    assert all(h == h for h in hist) and last < first - 0.05, (first, last)


def test_prediction_dropout_flag():
    """--dropout (nn.Dropout(0.5) on every prediction, criterion.py:113,168-169): eval mode is unaffected; in train
    mode the step runs, the loss moves away from the eval loss (half of each prediction is zeroed, the rest doubled)
    and every gradient stays finite; the mask changes from step to step."""
    _need_gpu()
    from cpc2_b200 import train as T
    args = T.defaultArgs(hiddenEncoder=64, hiddenGar=64, nPredicts=4, negativeSamplingExt=16, sizeWindow=5120,
                         rnnMode="linear", dropout=True)
    model, crit = build(args)
    plain = T.defaultArgs(**{**vars(args), "dropout": False})
    model0, crit0 = build(plain)
    crit0.load_state_dict(crit.state_dict())
    model0.load_state_dict(model.state_dict())
    g = torch.Generator().manual_seed(11)
    seq = (0.1 * torch.randn(8, 2, 1, args.sizeWindow, generator=g)).cuda()
    ev = run_step(model, crit, seq, train_mode=False, seed=5)
    ev0 = run_step(model0, crit0, seq, train_mode=False, seed=5)
    assert torch.allclose(ev["loss"], ev0["loss"], atol=1e-5)  # (loss sums are atomics: fp32 order only)
    a = run_step(model, crit, seq, train_mode=True, seed=5)
    b = run_step(model, crit, seq, train_mode=True, seed=5)
    assert torch.isfinite(a["loss"]).all() and all(torch.isfinite(v).all() for v in a["grads"].values())
    assert (a["loss"] - ev["loss"]).abs().max() > 1e-4 and (a["loss"] - b["loss"]).abs().max() > 1e-5


def test_prefetched_batches_reach_the_graph():
    """GraphedTrainStep.prefetch (input staging on a copy stream): each step() consumes the batch staged before it,
    so alternating two different batches reproduces the losses of passing them directly."""
    _need_gpu()
    from cpc2_b200 import train as T
    args = T.defaultArgs(hiddenEncoder=64, hiddenGar=64, nPredicts=4, negativeSamplingExt=16, sizeWindow=5120,
                         rnnMode="linear")
    model, crit = build(args)
    g = torch.Generator().manual_seed(21)
    a = (0.1 * torch.randn(4, 2, 1, args.sizeWindow, generator=g)).pin_memory()
    b = (0.3 * torch.randn(4, 2, 1, args.sizeWindow, generator=g)).pin_memory()
    step = T.GraphedTrainStep(model, crit, 4, args.sizeWindow, warmup=2, train_mode=False)
    direct = []
    for x in (a, b, a, b):
        torch.manual_seed(1)
        direct.append(step.step(x)[0].clone())
    staged = []
    step.prefetch(a)
    for nxt in (b, a, b, None):
        torch.manual_seed(1)
        out = step.step()[0]
        if nxt is not None:
            step.prefetch(nxt)
        staged.append(out.clone())
    torch.cuda.synchronize()
    assert not torch.allclose(direct[0], direct[1])
    for d, s in zip(direct, staged):
        assert torch.allclose(d, s, atol=1e-5)


def test_pipelined_loop_yields_every_step_in_order():
    """GraphedTrainStep.pipelined: one (losses, acc) pair per batch, in order, equal to stepping batch by batch."""
    _need_gpu()
    from cpc2_b200 import train as T
    args = T.defaultArgs(hiddenEncoder=64, hiddenGar=64, nPredicts=4, negativeSamplingExt=16, sizeWindow=5120,
                         rnnMode="linear")
    model, crit = build(args)
    g = torch.Generator().manual_seed(22)
    batches = [((0.1 + 0.1 * i) * torch.randn(4, 2, 1, args.sizeWindow, generator=g)).pin_memory() for i in range(5)]
    step = T.GraphedTrainStep(model, crit, 4, args.sizeWindow, warmup=2, train_mode=False)
    direct = []
    for x in batches:
        torch.manual_seed(1)  # (graph replays draw the negatives from the graph's own Philox offset: seed is moot)
        l, a = step.step(x)
        direct.append((l.detach().cpu().view(-1).clone(), a.detach().cpu().view(-1).clone()))
    got = [(l.clone(), a.clone()) for l, a in step.pipelined(batches)]
    assert len(got) == len(batches)
    for (l, a), (dl, da) in zip(got, direct):
        assert l.shape == dl.shape and torch.isfinite(l).all()
        assert (l - dl).abs().max() < 0.05  # same batch, different negative draws per replay


def test_val_step_and_keep_hidden_streaming():
    """Forward-only consumers: valStep (cpc/train.py:145-187) gives the eval losses of trainStep's
    forward without touching .grad; with keepHidden (samplingType 'sequential', feature_loader.py:229) two
    consecutive half-windows through CPCAR reproduce one full window (the hidden state streams across calls) and
    the two-halves fast path is not taken."""
    _need_gpu()
    from cpc2_b200 import train as T
    args = T.defaultArgs(hiddenEncoder=64, hiddenGar=64, nPredicts=4, negativeSamplingExt=16, sizeWindow=5120,
                         rnnMode="linear")
    model, crit = build(args)
    g = torch.Generator().manual_seed(31)
    seq = (0.1 * torch.randn(4, 2, 1, args.sizeWindow, generator=g)).cuda()
    label = torch.zeros(4, dtype=torch.long, device="cuda")
    torch.manual_seed(3)
    logs = T.valStep([(seq.cpu(), label.cpu())], model, crit)
    assert logs["iter"] == 1 and logs["locLoss_val"].shape == (4,)
    assert all(p.grad is None for p in model.parameters())
    model.eval()
    crit.eval()
    torch.manual_seed(3)
    losses, _ = T.forwardBackward(model, crit, seq, label, backward=False)
    assert abs(float(losses.mean()) - float(logs["locLoss_val"].mean())) < 1e-4
    # keepHidden streaming
    ar = model.gAR
    x = torch.randn(3, 32, 64, device="cuda")
    with torch.no_grad():
        full = ar(x)
        ar.keepHidden = True
        ar.hidden = None
        a = ar(x[:, :16])
        b = ar(x[:, 16:])
        assert ar.hidden is not None
    assert (torch.cat([a, b], dim=1) - full).abs().max() < 2e-2
    assert not model.splittable()
    ar.keepHidden, ar.hidden = False, None
    assert model.splittable()


@pytest.mark.gpu
def test_reference_wrappers_dataparallel_and_ddp():
    """cpc/train.py:523-532 wraps model and criterion in nn.DataParallel (single process) or DistributedDataParallel
    (--distributed): both wrappers must drive the sm_100a modules to the same losses and gradients as the bare modules
    (the wrappers take the literal concatenated call of train.py:95-104; DDP's own bucket all-reduce replaces ours)."""
    _need_gpu()
    import torch.distributed as dist
    from cpc2_b200 import train as T
    args = T.defaultArgs(hiddenEncoder=64, hiddenGar=64, nPredicts=4, negativeSamplingExt=16, sizeWindow=5120)
    g = torch.Generator().manual_seed(5)
    seq = (0.1 * torch.randn(4, 2, 1, args.sizeWindow, generator=g)).cuda()
    label = torch.zeros(4, dtype=torch.long, device="cuda")

    def one(wrap):
        model, crit = build(args)
        model.eval()
        crit.eval()
        wm, wc = wrap(model), wrap(crit)
        torch.manual_seed(9)
        losses, _ = T.forwardBackward(wm, wc, seq, label)
        torch.cuda.synchronize()
        return losses.detach().clone(), {k: p.grad.detach().clone() for k, p in
                                         list(model.named_parameters()) + list(crit.named_parameters())}

    base_l, base_g = one(lambda m: m)
    dp_l, dp_g = one(lambda m: torch.nn.DataParallel(m, device_ids=[0]))
    assert torch.allclose(base_l, dp_l, atol=1e-5)
    for k, v in base_g.items():
        assert cosine(v, dp_g[k]) > 0.9999, k
    import socket
    with socket.socket() as sock:  # a free rendezvous port for the single-rank group
        sock.bind(("127.0.0.1", 0))
        port = sock.getsockname()[1]
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    created = not dist.is_initialized()
    if created:
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
    try:
        ddp_l, ddp_g = one(lambda m: torch.nn.parallel.DistributedDataParallel(m, device_ids=[0]))
    finally:
        if created:
            dist.destroy_process_group()
    assert torch.allclose(base_l, ddp_l, atol=1e-5)
    for k, v in base_g.items():
        assert cosine(v, ddp_g[k]) > 0.9999, k


@pytest.mark.gpu
@pytest.mark.parametrize("ar_mode", ["LSTM", "transformer"])
def test_build_feature_chunks_of_any_length(ar_mode):
    """Feature extraction as cpc/feature_loader.py:323-367 (buildFeature) drives the model: one file, batch 1, no
    grad, chunks of maxSizeSeq = 64000 samples plus a last chunk whose length is not a multiple of the 160-sample
    hop -- with and without the AR state carried across chunks (keepHidden) for the LSTM, and through the transformer
    AR's block-local padding (400 frames = 12.5 blocks of sizeSeq 32, transformers.py:38-50) -- against the oracle
    (SURVEY 8f.4)."""
    _need_gpu()
    from cpc2_b200 import train as T
    from oracle import cpc_oracle as oc
    args = T.defaultArgs(hiddenEncoder=64, hiddenGar=64, nPredicts=4, negativeSamplingExt=16, sizeWindow=5120,
                         rnnMode="linear", arMode=ar_mode)
    model, _ = build(args)
    model.eval()
    g = torch.Generator().manual_seed(77)
    wav = 0.1 * torch.randn(1, 64000 * 2 + 23457, generator=g)
    st = {k: v.detach().float().cpu() for k, v in model.state_dict().items()}
    for keep in ((False, True) if ar_mode == "LSTM" else (False,)):
        if ar_mode == "LSTM":
            model.gAR.keepHidden, model.gAR.hidden = keep, None
        outs, refs, state = [], [], None
        for start in range(0, wav.shape[1], 64000):
            chunk = wav[:, start:start + 64000].view(1, 1, -1)
            with torch.no_grad():
                c, z, _ = model(chunk.cuda(), None)
            outs.append((c.float().cpu(), z.float().cpu()))
            zr = oc.encoder_forward(st, chunk).permute(0, 2, 1)
            if ar_mode == "LSTM":
                cr, state = oc.lstm_forward(st, zr, 1, state=state if keep else None)
            else:
                cr = oc.transformer_ar_forward(st, zr, 1, args.sizeWindow // 160)
            refs.append((cr, zr))
        for (c, z), (cr, zr) in zip(outs, refs):
            assert c.shape == cr.shape and z.shape == zr.shape, (c.shape, cr.shape)
            assert (z - zr).abs().max() <= 3e-2 * zr.abs().max() + 1e-3
            assert (c - cr).abs().max() <= 3e-2 * cr.abs().max() + 1e-3
        assert sum(c.shape[1] for c, _ in outs) == 400 + 400 + 146  # floor arithmetic of the five strided convs
    if ar_mode == "LSTM":
        model.gAR.keepHidden, model.gAR.hidden = False, None


@pytest.mark.gpu
def test_symmetric_allreduce_step_equals_nccl_step_on_two_gpus():
    """The data-parallel step with its gradient arenas in multicast symmetric memory + cpc_allreduce_mc gives the same
    losses after 8 Adam steps as the same step over NCCL (bench.py at N = 2, CPC_SYMMETRIC=1 vs 0)."""
    import json
    import subprocess
    import sys
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lines = {}
    for mode, port in (("0", 29651), ("1", 29652)):
        env = dict(os.environ, CPC_SYMMETRIC=mode)
        out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                              "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(root, "bench.py"),
                              "--gpus", "2", "--steps", "5", "--warmup", "3"],
                             capture_output=True, text=True, timeout=400, cwd=root, env=env)
        if mode == "1" and out.returncode != 0 and "multicast" in (out.stderr + out.stdout):
            pytest.skip("no NVLink multicast on this box")
        assert out.returncode == 0, out.stderr[-2000:]
        lines[mode] = json.loads(out.stdout.strip().splitlines()[-1])
    assert lines["1"]["config"]["grad_allreduce"].startswith("cpc_allreduce_mc")
    assert lines["0"]["config"]["grad_allreduce"].startswith("NCCL")
    # rank 0's loss after the same 16 optimizer steps on the same batches (bf16 step, fp32 atomics order differs)
    assert abs(lines["0"]["loss_mean"] - lines["1"]["loss_mean"]) < 1e-2, (lines["0"]["loss_mean"], lines["1"]["loss_mean"])


# ------------------------------------------------------------------------------------------------------------------
# public surface of the criterion (cpc/criterion/criterion.py:237-327) and of train.py that round 1 left untested
# ------------------------------------------------------------------------------------------------------------------
def _criterion_fixture(rnn="transformer", mode=None, B=4, S=32, D=64, Kp=4, NEG=16):
    from cpc2_b200 import criterion as cr
    torch.manual_seed(0)
    crit = cr.CPCUnsupersivedCriterion(Kp, D, D, NEG, mode=mode, rnnMode=rnn, sizeInputSeq=S).cuda().eval()
    g = torch.Generator().manual_seed(17)
    c = torch.randn(B, S, D, generator=g)
    z = torch.randn(B, S, D, generator=g)
    return crit, c, z


def test_sample_clean_matches_reference_semantics():
    """sampleClean (criterion.py:237-286): K tensors [B, NEG+1, W, D] whose slot 0 is the positive z[:, k:k+W] and whose
    other slots are z gathered at extIdx = ((seqIdx + t) mod S) + batchIdx*S, the two randint draws issued in the
    reference's order -- redrawn here with the same seed on the same device and pushed through the ORACLE's integer
    arithmetic (bit exact); labels are zeros."""
    _need_gpu()
    from oracle import cpc_oracle as oc
    crit, c, z = _criterion_fixture("linear")
    B, S, D = z.shape
    Kp, NEG = crit.nPredicts, crit.negativeSamplingExt
    W = S - Kp
    zg = z.cuda()
    torch.manual_seed(99)
    outs, labels = crit.sampleClean(zg, W)
    torch.manual_seed(99)
    n = NEG * W * B
    bi = torch.randint(low=0, high=B, size=(n,), device="cuda")
    si = torch.randint(low=1, high=S, size=(n,), device="cuda")
    ext = oc.negative_ext_index(bi.cpu(), si.cpu(), B, S, W, NEG)
    assert len(outs) == Kp and labels.shape == (B * W,) and labels.dtype == torch.long and int(labels.abs().sum()) == 0
    neg = z.view(-1, D)[ext].view(B, NEG, W, D)
    for k in range(1, Kp + 1):
        o = outs[k - 1].cpu()
        assert o.shape == (B, NEG + 1, W, D)
        assert torch.equal(o[:, 0], z[:, k:k + W])
        assert torch.equal(o[:, 1:], neg)


@pytest.mark.parametrize("rnn,mode", [("transformer", None), ("linear", "reverse"), ("ffd", None)])
def test_get_prediction_and_cosine_distances_match_oracle(rnn, mode):
    """getPrediction (criterion.py:291-302) -> K score tensors [B, NEG+1, W] + labels, and getCosineDistances
    (:304-327) -> the positives' scores [B, 1, W], through PredictionNetwork.forward (:152-173, the materialised
    `(locC[:, None] * candidates).mean(3)` path feature_loader.py:78-80 uses), against the oracle on the same sampled
    indices; then the backward of PredictionNetwork.forward (_HeadsFn): gradients of a cross entropy over those scores
    w.r.t. c and every head parameter against the oracle's autograd (cosine >= 0.99)."""
    _need_gpu()
    from oracle import cpc_oracle as oc
    crit, c, z = _criterion_fixture(rnn, mode)
    B, S, D = z.shape
    Kp, NEG = crit.nPredicts, crit.negativeSamplingExt
    W = S - Kp
    torch.manual_seed(5)
    bi, si = oc.draw_negative_indices(B, S, W, NEG)
    ext = oc.negative_ext_index(bi, si, B, S, W, NEG)
    crit.sampleIndices = lambda *a: ext.cuda()
    cg = c.cuda().requires_grad_(True)
    zg = z.cuda()
    preds, labels = crit.getPrediction(cg, zg, None)
    state = {k: v.detach().float().cpu() for k, v in crit.state_dict().items()}
    params = {k: v.clone().requires_grad_(True) for k, v in state.items() if v.is_floating_point()}
    full = dict(state)
    full.update(params)
    cr_ = c.clone().requires_grad_(True)
    rl, ra, logits, _ = oc.criterion_forward(full, cr_, z, bi, si, Kp, NEG, rnn_mode=rnn, mode=mode, return_logits=True)
    assert len(preds) == Kp and labels.shape == (B * W,)
    scale = max(float(l.abs().max()) for l in logits)
    for k in range(Kp):
        got = preds[k].detach().cpu().permute(0, 2, 1).reshape(B * W, NEG + 1)  # criterion.py:346-348
        assert preds[k].shape == (B, NEG + 1, W)
        assert (got - logits[k].detach()).abs().max() <= 3e-2 * scale + 1e-3, k
    # the reference's loss on top of the materialised scores, and its gradients through PredictionNetwork.forward
    loss = sum(torch.nn.functional.cross_entropy(p_.permute(0, 2, 1).reshape(B * W, NEG + 1), labels) for p_ in preds)
    loss.backward()
    rl.sum().backward()
    assert abs(float(loss) - float(rl.sum())) <= 2e-2
    assert cosine(cg.grad.cpu(), cr_.grad) >= 0.99
    for name, p_ in crit.named_parameters():
        ref = params[name].grad
        if ref is None or ref.abs().max() == 0:
            continue
        assert cosine(p_.grad.cpu(), ref) >= 0.99, (name, cosine(p_.grad.cpu(), ref))
    with torch.no_grad():
        cosd = crit.getCosineDistances(c.cuda(), zg)
    assert len(cosd) == Kp
    for k in range(Kp):
        assert cosd[k].shape == (B, 1, W)
        ref_pos = logits[k].detach()[:, 0].view(B, W)
        assert (cosd[k][:, 0].cpu() - ref_pos).abs().max() <= 3e-2 * scale + 1e-3, k


def test_train_step_entry_point():
    """cpc2_b200.train.trainStep (cpc/train.py:72-142, unchanged signature): with a FusedAdam optimizer and an
    unwrapped model the batch loop replays ONE CUDA graph per step (forward + backward + Adam); the ragged last batch of
    the loader and the GRAPH_TRAIN_STEP = False switch take the eager body.  Both give the reference's `logs`
    (per-prediction-step means over the batches, `iter`), advance the optimizer once per batch, step the scheduler
    once per call, and agree with each other up to the different negative draws."""
    _need_gpu()
    from cpc2_b200 import train as T
    from cpc2_b200.optim import FusedAdam
    args = T.defaultArgs(hiddenEncoder=64, hiddenGar=64, nPredicts=4, negativeSamplingExt=16, sizeWindow=5120)
    g = torch.Generator().manual_seed(41)
    loader = [(predictable_sequence(4, args.sizeWindow, 100 + i).pin_memory(), torch.zeros(4, dtype=torch.long))
              for i in range(5)]
    loader.append((predictable_sequence(3, args.sizeWindow, 200).pin_memory(), torch.zeros(3, dtype=torch.long)))
    res = {}
    for graphed in (True, False):
        model, crit = build(args)
        opt = FusedAdam(list(crit.parameters()) + list(model.parameters()), lr=1e-3)
        sched = torch.optim.lr_scheduler.StepLR(opt, 1, gamma=0.5)
        before = [p.detach().clone() for p in model.parameters()]
        T.GRAPH_TRAIN_STEP = graphed
        try:
            torch.manual_seed(7)
            logs = T.trainStep(loader, model, crit, opt, sched, loggingStep=2)
            logs2 = T.trainStep(loader[:2], model, crit, opt, None, loggingStep=10)  # graph reused, lr halved
        finally:
            T.GRAPH_TRAIN_STEP = True
        torch.cuda.synchronize()
        assert logs["iter"] == 6 and logs2["iter"] == 2
        assert logs["locLoss_train"].shape == (4,) and logs["locAcc_train"].shape == (4,)
        assert all(v == v for v in logs["locLoss_train"]) and all(0 <= v <= 1 for v in logs["locAcc_train"])
        assert float(opt._groups[0]["state"][0]) == 8
        assert abs(opt.param_groups[0]["lr"] - 5e-4) < 1e-12 and abs(float(opt._groups[0]["state"][1]) - 5e-4) < 1e-12
        assert model.training and crit.training
        assert any((p.detach() - b).abs().max() > 0 for p, b in zip(model.parameters(), before))
        assert len(T._GRAPHED.get(model, {})) == (2 if graphed else 0)  # one graph per batch shape (4 and the ragged 3)
        res[graphed] = logs
    assert abs(res[True]["locLoss_train"] - res[False]["locLoss_train"]).max() < 0.05


def test_graphed_step_construction_has_no_side_effects():
    """GraphedTrainStep(optimizer=FusedAdam) warms up with real optimizer steps on an all-zero batch before
    the capture; parameters, Adam moments and the device step counter must come out of the constructor exactly as they
    went in (a checkpoint-loaded model starts from its loaded state), and the first replay is update number 1."""
    _need_gpu()
    from cpc2_b200 import train as T
    from cpc2_b200.optim import FusedAdam
    args = T.defaultArgs(hiddenEncoder=64, hiddenGar=64, nPredicts=4, negativeSamplingExt=16, sizeWindow=5120)
    model, crit = build(args)
    params = list(crit.parameters()) + list(model.parameters())
    opt = FusedAdam(params, lr=1e-3)
    before = [p.detach().clone() for p in params]
    step = T.GraphedTrainStep(model, crit, 4, args.sizeWindow, warmup=3, optimizer=opt)
    torch.cuda.synchronize()
    for p, b in zip(params, before):
        assert torch.equal(p.detach(), b)
    g0 = opt._groups[0]
    assert float(g0["state"][0]) == 0 and float(g0["m"].abs().max()) == 0 and float(g0["v"].abs().max()) == 0
    step.step(predictable_sequence(4, args.sizeWindow, 3).cuda())
    torch.cuda.synchronize()
    assert float(g0["state"][0]) == 1
    assert any((p.detach() - b).abs().max() > 0 for p, b in zip(params, before))
    # Adam's first update moves every touched weight by ~lr: the warm-up steps really were rolled back
    worst = max(float((p.detach() - b).abs().max()) for p, b in zip(params, before))
    assert worst <= 1.01e-3, worst
    step.close()
    with pytest.raises(RuntimeError):
        step.step()


@pytest.mark.parametrize("ar_mode", ["LSTM", "transformer"])
def test_adam_written_bf16_shadow_tracks_the_parameters(ar_mode):
    """The arena-homed weights (prediction heads, transformer AR) have ONE persistent bf16 shadow that cpc_adam_step writes
    together with the fp32 parameter (CpcAdamSpan.p16), so the captured step holds no weight cast.  After every replay the
    shadow must equal a fresh round-to-nearest cast of the parameters, bit for bit; a parameter edited behind the
    optimizer's back (in-place op -> autograd version bump) must be picked up by the next replay; and the same steps with
    torch.optim.Adam (which never writes the shadow: it is re-cast every step) must give the same losses."""
    _need_gpu()
    from cpc2_b200 import kernels as K
    from cpc2_b200 import train as T
    from cpc2_b200.optim import FusedAdam
    over = dict(hiddenEncoder=64, hiddenGar=64, nPredicts=4, negativeSamplingExt=16, sizeWindow=5120, arMode=ar_mode)
    args = T.defaultArgs(**over)
    model, crit = build(args)
    params = list(crit.parameters()) + list(model.parameters())
    opt = FusedAdam(params, lr=1e-3)
    step = T.GraphedTrainStep(model, crit, 4, args.sizeWindow, warmup=3, optimizer=opt, train_mode=False)
    arenas = [crit.wPrediction.engine().arena]
    if ar_mode == "transformer":
        arenas += [m.stack().arena for m in model.gAR if hasattr(m, "stack")]
    seq = predictable_sequence(4, args.sizeWindow, 3).cuda()
    n0 = K.LAUNCHES

    def check():
        torch.cuda.synchronize()
        for a in arenas:
            assert a.shadow_valid()
            assert torch.equal(a._shadow, a.flat.to(torch.bfloat16)), "bf16 shadow differs from the parameters"

    losses = []
    for _ in range(3):
        losses.append(step.step(seq)[0].clone())
        check()
    assert K.LAUNCHES == n0, "a replay must not launch anything eagerly when nothing was edited"
    with torch.no_grad():
        arenas[0].params[0].mul_(0.5)  # behind the optimizer's back
    assert not arenas[0].shadow_valid()
    step.step(seq)
    check()
    step.close()
    # the same three updates with torch.optim.Adam on a fresh copy (other negatives are drawn, so only the loss level
    # is comparable: a stale or half-written shadow would send it off by far more than 3e-2)
    model2, crit2 = build(args)
    opt2 = torch.optim.Adam(list(crit2.parameters()) + list(model2.parameters()), lr=1e-3)
    for i in range(3):
        out = run_step(model2, crit2, seq, collect=False, seed=0)
        opt2.step()
        assert (out["loss"].cuda() - losses[i]).abs().max() < 3e-2, (i, out["loss"], losses[i])


def test_multi_device_dataparallel_replicas_are_rejected():
    """nn.DataParallel over several devices replicates the modules (replicas carry `_is_replica`); the engines
    and arenas of this package are per process, so a replica's forward raises instead of launching on device-0 pointers
    (device_ids=[0] never replicates and is covered by test_reference_wrappers_dataparallel_and_ddp)."""
    _need_gpu()
    from cpc2_b200 import train as T
    args = T.defaultArgs(hiddenEncoder=64, hiddenGar=64, nPredicts=4, negativeSamplingExt=16, sizeWindow=5120)
    model, crit = build(args)
    x = torch.zeros(2, 1, 5120, device="cuda")
    for mod, inp in ((model.gEncoder, (x,)), (model.gAR, (torch.zeros(2, 32, 64, device="cuda"),)),
                     (crit, (torch.zeros(2, 32, 64, device="cuda"), torch.zeros(2, 32, 64, device="cuda"), None))):
        mod._is_replica = True
        try:
            with pytest.raises(RuntimeError, match="DataParallel"):
                mod(*inp)
        finally:
            del mod._is_replica


def test_shape_limits_fail_with_a_message():
    """B*S above the cpc_nce_bwd shared-memory row buffer and hidden sizes the recurrence kernels are not
    built for are rejected in Python with the limit in the message (not by a CPC_ERR_UNSUPPORTED in the backward)."""
    _need_gpu()
    from cpc2_b200 import criterion as cr
    from cpc2_b200 import model as M
    crit = cr.CPCUnsupersivedCriterion(4, 64, 64, 16, rnnMode="linear", sizeInputSeq=128).cuda()
    big = torch.zeros(51, 128, 64, device="cuda", requires_grad=True)
    with pytest.raises(ValueError, match="6400"):
        crit(big, big, None)
    ar = M.CPCAR(64, 48, False, 1, mode="LSTM").cuda()
    with pytest.raises(NotImplementedError, match="hidden sizes"):
        ar(torch.zeros(2, 8, 64, device="cuda"))


@pytest.mark.parametrize("S_in", [48, 70, 20])
def test_transformer_layer_pads_ragged_sequences(S_in):
    """cpc/transformers.py:38-50,69: sequences that are not a whole number of sizeSeq blocks are zero padded per
    sequence, folded into the batch as independent causal blocks and cropped again.  TransformerLayer forward AND
    backward on [N, S_in, d] with sizeSeq 32 against the oracle (which restates `prepare`)."""
    _need_gpu()
    from cpc2_b200 import transformers as tr
    from oracle import cpc_oracle as oc
    torch.manual_seed(0)
    d = 256
    layer = tr.TransformerLayer(sizeSeq=32, dmodel=d, dout=d, abspos=False).cuda().eval()
    g = torch.Generator().manual_seed(3)
    x = torch.randn(3, S_in, d, generator=g)
    dy = torch.randn(3, S_in, d, generator=g)
    xg = x.cuda().requires_grad_(True)
    y = layer(xg)
    y.backward(dy.cuda())
    state = {"l." + k: v.detach().float().cpu() for k, v in layer.state_dict().items()}
    params = {k: v.clone().requires_grad_(True) for k, v in state.items() if v.is_floating_point()}
    full = dict(state)
    full.update(params)
    xr = x.clone().requires_grad_(True)
    yr = oc.transformer_layer(full, xr, "l.", 32)
    yr.backward(dy)
    assert y.shape == yr.shape
    assert (y.detach().cpu() - yr.detach()).abs().max() <= 3e-2 * float(yr.abs().max()) + 1e-3
    assert cosine(xg.grad.cpu(), xr.grad) >= 0.99
    for name, p_ in layer.named_parameters():
        ref = params["l." + name].grad
        if ref is None or ref.abs().max() == 0:
            continue
        assert cosine(p_.grad.cpu(), ref) >= 0.99, (name, cosine(p_.grad.cpu(), ref))
