"""Pins oracle/cpc_oracle.py against vectors produced by the UNMODIFIED reference
(tests/golden/make_golden.py).  CPU only.  fp32 tolerance: rel 1e-4 / abs 1e-5 on
activations and grads, indices bit exact."""
import os

import pytest
import torch

from oracle import cpc_oracle as oc

FULL_CASES = ["tiny_lstm_linear", "tiny_gru2_transformer", "tiny_transformer_ar_linear",
              "tiny_lstm_reverse_linear", "tiny_rnn_hneq_linear", "tiny_lstm_multihead", "tiny_lstm_ffd",
              "tiny_lstm_conv4", "tiny_lstm_headsLSTM", "tiny_lstm_headsRNN", "tiny_lstm_quality_skip1"]


def _load(golden_dir, name):
    return torch.load(os.path.join(golden_dir, name + ".pt"), weights_only=False)


@pytest.mark.parametrize("name", FULL_CASES)
def test_indices_bit_exact(golden_dir, name):
    fx = _load(golden_dir, name)
    a = fx["args"]
    S = a["sizeWindow"] // 160
    W = S - a["nPredicts"]
    torch.manual_seed(fx["crit_seed"])
    b, s = oc.draw_negative_indices(fx["B"], S, W, a["negativeSamplingExt"])
    assert torch.equal(b.int(), fx["batch_idx"])
    assert torch.equal(s.int(), fx["seq_idx"])
    ext = oc.negative_ext_index(b, s, fx["B"], S, W, a["negativeSamplingExt"])
    assert torch.equal(ext.int(), fx["ext_idx"])
    assert int(ext.min()) >= 0 and int(ext.max()) < fx["B"] * S


@pytest.mark.parametrize("name", FULL_CASES)
def test_step_matches_reference(golden_dir, name):
    fx = _load(golden_dir, name)
    a = fx["args"]
    out = oc.train_step(fx["state"], fx["sequence"], fx["batch_idx"].long(), fx["seq_idx"].long(),
                        K=a["nPredicts"], NEG=a["negativeSamplingExt"], ar_mode=a["arMode"],
                        n_layers=a["nLevelsGRU"], rnn_mode=a["rnnMode"], cpc_mode=a["cpc_mode"],
                        abspos=a["abspos"], n_skipped=a["n_skipped"], multihead=a.get("multihead_rnn", False),
                        signal_quality=fx.get("signal_quality"), growth_rate=a.get("growth_rate"),
                        inflection_point_x=a.get("inflection_point_x"))
    torch.testing.assert_close(out["z"], fx["z"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(out["c"], fx["c"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(out["loss"], fx["loss"], rtol=1e-5, atol=1e-6)
    assert torch.equal(out["acc"], fx["acc"])  # argmax predictions identical
    assert set(out["grads"]) == set(fx["grads"])
    for k, g in fx["grads"].items():
        torch.testing.assert_close(out["grads"][k], g, rtol=2e-4, atol=2e-6, msg=lambda m: f"{k}: {m}")


NORM_CASES = ["tiny_batchnorm_train_linear", "tiny_batchnorm_eval_linear", "tiny_instancenorm_linear", "tiny_idnorm_linear"]


@pytest.mark.parametrize("name", NORM_CASES)
def test_encoder_norm_modes_match_reference(golden_dir, name):
    """normMode batchNorm (training mode: batch statistics and the running-estimate update; eval mode: running
    estimates), instanceNorm and ID (cpc/model.py:71-82) against the unmodified reference."""
    fx = _load(golden_dir, name)
    a = fx["args"]
    out = oc.train_step(fx["state"], fx["sequence"], fx["batch_idx"].long(), fx["seq_idx"].long(),
                        K=a["nPredicts"], NEG=a["negativeSamplingExt"], ar_mode=a["arMode"], n_layers=a["nLevelsGRU"],
                        rnn_mode=a["rnnMode"], norm_mode=a["normMode"], bn_training=fx["train_mode"])
    torch.testing.assert_close(out["z"], fx["z"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(out["c"], fx["c"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(out["loss"], fx["loss"], rtol=1e-5, atol=1e-6)
    assert torch.equal(out["acc"], fx["acc"])
    assert set(out["grads"]) == set(fx["grads"])
    for k, g in fx["grads"].items():
        # (conv biases in front of a batch / instance norm have a mathematically zero gradient: both sides hold noise)
        torch.testing.assert_close(out["grads"][k], g, rtol=5e-4, atol=2e-6 + 2e-5 * float(g.abs().max()), msg=lambda m: f"{k}: {m}")
    if fx["train_mode"]:  # running estimates as nn.BatchNorm1d updates them (momentum 0.1, unbiased variance)
        for i, (mean, var) in enumerate(out["bn_stats"]):
            k = f"gEncoder.batchNorm{i}."
            torch.testing.assert_close(0.9 * fx["state"][k + "running_mean"] + 0.1 * mean, fx["running_after"][k + "running_mean"],
                                       rtol=1e-5, atol=1e-7)
            torch.testing.assert_close(0.9 * fx["state"][k + "running_var"] + 0.1 * var, fx["running_after"][k + "running_var"],
                                       rtol=1e-5, atol=1e-7)


TRAINED_CASES = [c + "_trained" for c in FULL_CASES if c != "tiny_lstm_quality_skip1"] + ["mid256_lstm_transformer_trained"]


@pytest.mark.parametrize("name", TRAINED_CASES)
def test_step_matches_reference_away_from_init(golden_dir, name):
    """The same pin on fixtures taken after the reference's own Adam has moved the model off the ln(NEG+1) plateau
    (tests/golden/make_golden.py, `*_trained`): logits are no longer ~0, so the gradients are not cancelling sums.
    Big gradients are stored as strided samples (fx["grad_stride"])."""
    fx = _load(golden_dir, name)
    a = fx["args"]
    state = {k: v.float() for k, v in fx["state"].items()}
    out = oc.train_step(state, fx["sequence"], fx["batch_idx"].long(), fx["seq_idx"].long(),
                        K=a["nPredicts"], NEG=a["negativeSamplingExt"], ar_mode=a["arMode"],
                        n_layers=a["nLevelsGRU"], rnn_mode=a["rnnMode"], cpc_mode=a["cpc_mode"],
                        abspos=a["abspos"], n_skipped=a["n_skipped"], multihead=a.get("multihead_rnn", False))
    import math
    assert float((fx["loss"] - math.log(a["negativeSamplingExt"] + 1)).abs().max()) > 0.01  # really off the plateau
    torch.testing.assert_close(out["z"], fx["z"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(out["c"], fx["c"], rtol=1e-4, atol=2e-5)
    torch.testing.assert_close(out["loss"], fx["loss"], rtol=1e-5, atol=2e-6)
    assert torch.equal(out["acc"], fx["acc"])
    assert set(out["grads"]) == set(fx["grads"])
    for k, g in fx["grads"].items():
        st = fx["grad_stride"][k]
        got = out["grads"][k].flatten()[::st] if st > 1 else out["grads"][k]
        torch.testing.assert_close(got, g, rtol=5e-4, atol=1e-6 + 1e-5 * float(g.abs().max()), msg=lambda m: f"{k}: {m}")
        assert abs(float(out["grads"][k].double().norm()) - fx["grad_norm"][k]) <= 1e-4 * fx["grad_norm"][k] + 1e-12, k


@pytest.mark.parametrize("name", ["tiny_instancenorm_linear_trained", "tiny_batchnorm_eval_linear_trained"])
def test_encoder_norm_modes_match_reference_away_from_init(golden_dir, name):
    """The norm-mode pin after the reference's own Adam has moved the weights (instance statistics; BatchNorm1d on its
    running estimates, where nothing normalises and the activations have grown)."""
    fx = _load(golden_dir, name)
    a = fx["args"]
    state = {k: v.float() for k, v in fx["state"].items()}
    out = oc.train_step(state, fx["sequence"], fx["batch_idx"].long(), fx["seq_idx"].long(),
                        K=a["nPredicts"], NEG=a["negativeSamplingExt"], ar_mode=a["arMode"], n_layers=a["nLevelsGRU"],
                        rnn_mode=a["rnnMode"], norm_mode=a["normMode"], bn_training=False)
    torch.testing.assert_close(out["z"], fx["z"], rtol=1e-4, atol=1e-5 * max(1.0, float(fx["z"].abs().max())))
    torch.testing.assert_close(out["c"], fx["c"], rtol=1e-4, atol=2e-5)
    torch.testing.assert_close(out["loss"], fx["loss"], rtol=1e-5, atol=2e-6)
    assert torch.equal(out["acc"], fx["acc"])
    assert set(out["grads"]) == set(fx["grads"])
    for k, g in fx["grads"].items():
        torch.testing.assert_close(out["grads"][k], g, rtol=5e-4, atol=2e-6 + 2e-5 * float(g.abs().max()), msg=lambda m: f"{k}: {m}")


def test_full_size_default_step_matches_reference_summary(golden_dir):
    """BASELINE.json configs[0] at its real size (encoder 256, LSTM 256, 12 transformer heads, 128 negatives, window
    20480, batch 8): the oracle's losses, activations, sampled indices and EVERY gradient (norm + a 64-element strided
    sample per tensor) against what the unmodified reference produced (default_b8_summary.pt).  Weights come from the
    same seeded constructors (the fingerprint proves they are the reference's)."""
    from cpc2_b200 import train as T
    fx = _load(golden_dir, "default_b8_summary")
    a = fx["args"]
    torch.manual_seed(fx["init_seed"])
    model, crit = T.buildModelAndCriterion(T.defaultArgs())
    state = {k: v.detach().float() for k, v in list(model.state_dict().items()) + list(crit.state_dict().items())}
    for k, (sm, ab) in fx["fingerprint"].items():
        v = state[k].double()
        assert abs(float(v.sum()) - sm) <= 1e-6 * max(1.0, ab) and abs(float(v.abs().sum()) - ab) <= 1e-6 * max(1.0, ab), k
    B, S = fx["B"], a["sizeWindow"] // 160
    W = S - a["nPredicts"]
    seq = 0.1 * torch.randn(B, 2, 1, a["sizeWindow"], generator=torch.Generator().manual_seed(fx["seq_seed"]))
    torch.manual_seed(fx["crit_seed"])
    bi, si = oc.draw_negative_indices(B, S, W, a["negativeSamplingExt"])
    ext = oc.negative_ext_index(bi, si, B, S, W, a["negativeSamplingExt"])
    assert torch.equal(ext[:2048].int(), fx["ext_head"]) and int(ext.sum()) == fx["ext_sum"]
    assert int((ext * (torch.arange(ext.numel()) % 1009)).sum()) == fx["ext_weighted_sum"]
    import os as _os
    torch.set_num_threads(_os.cpu_count())
    out = oc.train_step(state, seq, bi, si, K=a["nPredicts"], NEG=a["negativeSamplingExt"], ar_mode=a["arMode"],
                        n_layers=a["nLevelsGRU"], rnn_mode=a["rnnMode"])
    torch.testing.assert_close(out["loss"], fx["loss"], rtol=1e-5, atol=1e-5)
    assert (out["acc"] - fx["acc"]).abs().max() <= 1.5 / (B * W)  # near-ties at init may flip one argmax in fp32
    torch.testing.assert_close(out["z"][:, ::16, ::32], fx["z_slice"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(out["c"][:, ::16, ::32], fx["c_slice"], rtol=1e-4, atol=1e-5)
    assert set(out["grads"]) == set(fx["grad_norm"])
    for k, n in fx["grad_norm"].items():
        g = out["grads"][k]
        assert abs(float(g.double().norm()) - n) <= 2e-3 * n + 1e-12, (k, float(g.double().norm()), n)
        sl = g.flatten()[:: max(1, g.numel() // 64)][:64]
        ref = fx["grad_slice"][k]
        assert (sl - ref).abs().max() <= 2e-3 * float(ref.abs().max()) + 1e-3 * n / max(1.0, g.numel() ** 0.5), k


def test_channel_norm_unbiased_variance():
    x = torch.randn(2, 8, 5)
    y = oc.channel_norm(x)
    ref = (x - x.mean(1, keepdim=True)) / torch.sqrt(x.var(1, unbiased=True, keepdim=True) + 1e-5)
    torch.testing.assert_close(y, ref)


def test_init_loss_is_log_negatives(golden_dir):
    fx = _load(golden_dir, "default_b8_summary")
    import math
    assert abs(float(fx["loss"].mean()) - math.log(129)) < 5e-3
