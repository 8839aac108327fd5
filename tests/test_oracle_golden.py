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


def test_channel_norm_unbiased_variance():
    x = torch.randn(2, 8, 5)
    y = oc.channel_norm(x)
    ref = (x - x.mean(1, keepdim=True)) / torch.sqrt(x.var(1, unbiased=True, keepdim=True) + 1e-5)
    torch.testing.assert_close(y, ref)


def test_init_loss_is_log_negatives(golden_dir):
    fx = _load(golden_dir, "default_b8_summary")
    import math
    assert abs(float(fx["loss"].mean()) - math.log(129)) < 5e-3
