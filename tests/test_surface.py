"""CPU-side checks of the reference's module surface -- no kernel is launched: the drop-in modules
expose the state_dict keys / shapes of the UNMODIFIED reference (read from the reference-generated golden fixtures),
the reference's constructor errors, and the same initial weights for the same seed."""
import os

import pytest
import torch

from cpc2_b200 import criterion as cr
from cpc2_b200 import model as M
from cpc2_b200 import train as T
from cpc2_b200 import transformers as tr

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["tiny_lstm_linear", "tiny_gru2_transformer", "tiny_transformer_ar_linear", "tiny_lstm_reverse_linear",
         "tiny_rnn_hneq_linear", "tiny_lstm_multihead", "tiny_lstm_ffd", "tiny_lstm_conv4", "tiny_lstm_headsLSTM", "tiny_lstm_headsRNN", "tiny_lstm_quality_skip1",
         # encoder normMode batchNorm / instanceNorm / ID (cpc/model.py:71-82): BatchNorm1d's running estimates are keys too
         "tiny_batchnorm_train_linear", "tiny_instancenorm_linear", "tiny_idnorm_linear"]


@pytest.mark.parametrize("name", CASES)
def test_state_dict_keys_shapes_and_init_match_reference(name):
    fx = torch.load(os.path.join(GOLD, name + ".pt"), weights_only=False)
    args = T.defaultArgs(**fx["args"])
    torch.manual_seed(fx["init_seed"])
    model, crit = T.buildModelAndCriterion(args)
    ours = dict(model.state_dict())
    ours.update(crit.state_dict())
    ours = {k: v for k, v in ours.items() if v.dtype.is_floating_point and not k.endswith(".mask") and not k.endswith(".z")}
    ref = fx["state"]
    assert set(ours) == set(ref), (sorted(set(ours) - set(ref)), sorted(set(ref) - set(ours)))
    for k, v in ref.items():
        assert tuple(ours[k].shape) == tuple(v.shape), k
        # same construction order and RNG draws as cpc/train.py:442-461 -> identical initial weights
        assert torch.equal(ours[k].float(), v), k
    # reference state loads strictly (buffers `mask` / `z` of the attention modules are not in the fixture)
    missing = crit.load_state_dict({k: v for k, v in ref.items() if k.startswith("wPrediction")}, strict=False)
    assert not missing.unexpected_keys
    assert all(k.endswith(".mask") or k.endswith(".z") for k in missing.missing_keys), missing.missing_keys


def test_constructor_errors_of_the_reference():
    with pytest.raises(ValueError):
        M.CPCEncoder(32, normMode="nope")                      # cpc/model.py:71-73
    with pytest.raises(ValueError, match="Invalid mode"):
        cr.CPCUnsupersivedCriterion(4, 32, 32, 8, mode="sideways")   # criterion.py:232-233
    with pytest.raises(ValueError, match="unknown mode"):
        cr.MultiHeadPredictionNetwork(4, 32, 32, rnnMode="linear")   # criterion.py:77-78
    c = cr.CPCUnsupersivedCriterion(4, 32, 32, 8, rnnMode="linear", sizeInputSeq=16)
    with pytest.raises(AssertionError):
        c.wPrediction(torch.zeros(1, 12, 32), [torch.zeros(1, 9, 12, 32)] * 3)   # criterion.py:154


def test_attribute_surface():
    args = T.defaultArgs(hiddenEncoder=32, hiddenGar=32, nPredicts=4, negativeSamplingExt=8, sizeWindow=2560)
    model, crit = T.buildModelAndCriterion(args)
    enc, ar = model.gEncoder, model.gAR
    assert enc.DOWNSAMPLING == 160 and enc.dimEncoded == 32 and enc.getDimOutput() == 32
    assert ar.getDimOutput() == 32 and ar.hidden is None and ar.keepHidden is False and ar.reverse is False
    assert all(hasattr(ar.baseNet, n) for n in ("weight_ih_l0", "weight_hh_l0", "bias_ih_l0", "bias_hh_l0"))
    assert crit.nPredicts == 4 and crit.negativeSamplingExt == 8 and crit.nSkipped == 0 and crit.mode is None
    assert len(crit.wPrediction.predictors) == 4 and crit.warmUp() is False and crit.update() is None
    assert isinstance(tr.buildTransformerAR(32, 32, 2, 16, True)[0], tr.StaticPositionEmbedding)
    seq_args = T.defaultArgs(samplingType="sequential")
    assert T.getAR(seq_args).keepHidden is True               # feature_loader.py:229


def test_fused_adam_has_no_cpu_path():
    from cpc2_b200.optim import FusedAdam
    p = torch.nn.Parameter(torch.zeros(4))
    p.grad = torch.ones(4)
    with pytest.raises(RuntimeError, match="CUDA parameters only"):
        FusedAdam([p]).step()
    with pytest.raises(NotImplementedError):
        FusedAdam([p], weight_decay=0.1)


def test_checkpoint_layout_round_trip(tmp_path):
    """Checkpoints in the reference's layout (feature_loader.save_checkpoint, feature_loader.py:296-304:
    {"gEncoder": model_state, "cpcCriterion": criterion_state, "optimizer": ..., "best": ...}) written from these
    modules load back into fresh ones, and the model half is what loadModel (feature_loader.py:238-283) reads."""
    args = T.defaultArgs(hiddenEncoder=32, hiddenGar=32, nPredicts=4, negativeSamplingExt=8, sizeWindow=2560)
    torch.manual_seed(1)
    model, crit = T.buildModelAndCriterion(args)
    path = os.path.join(tmp_path, "checkpoint_0.pt")
    torch.save({"gEncoder": model.state_dict(), "cpcCriterion": crit.state_dict(), "optimizer": None, "best": None},
               path)
    torch.manual_seed(2)  # different initial weights
    model2, crit2 = T.buildModelAndCriterion(args)
    assert not torch.equal(model2.gEncoder.conv1.weight, model.gEncoder.conv1.weight)
    state = torch.load(path, map_location="cpu", weights_only=False)
    assert set(state) == {"gEncoder", "cpcCriterion", "optimizer", "best"}
    model2.load_state_dict(state["gEncoder"], strict=False)      # loadModel: strict=False
    crit2.load_state_dict(state["cpcCriterion"])                 # loadCriterion (train.py:62-69): strict
    for (k, a), (_, b) in zip(model.state_dict().items(), model2.state_dict().items()):
        assert torch.equal(a, b), k
    for (k, a), (_, b) in zip(crit.state_dict().items(), crit2.state_dict().items()):
        assert torch.equal(a, b), k
    assert all(k.startswith(("gEncoder.", "gAR.")) for k in state["gEncoder"])


def test_encoder_norm_modes_surface_and_plan():
    """normMode batchNorm / instanceNorm / ID (cpc/model.py:71-82): the reference's norm modules are the parameter
    containers, the step is not splittable into past / future halves, and the launch plan follows nn.BatchNorm1d's
    semantics (training: batch statistics + one step of the running estimates; eval: the running estimates)."""
    for mode, cls in (("batchNorm", torch.nn.BatchNorm1d), ("instanceNorm", torch.nn.InstanceNorm1d), ("ID", M.IDModule)):
        enc = M.CPCEncoder(32, normMode=mode)
        assert all(isinstance(getattr(enc, f"batchNorm{i}"), cls) for i in range(5))
        assert len(enc._params()) == (10 if mode == "ID" else 20)
        model = M.CPCModel(enc, M.CPCAR(32, 32, False, 1, mode="LSTM"))
        assert not model.splittable()
    assert M.CPCModel(M.CPCEncoder(32), M.CPCAR(32, 32, False, 1, mode="LSTM")).splittable()
    enc = M.CPCEncoder(32, normMode="batchNorm")
    enc.train()
    plan = M._ColPlan(enc, None)
    assert [d["kind"] for d in plan.layers] == ["stats"] * 5 and plan.layers[0]["rm"] is enc.batchNorm0.running_mean
    assert plan.layers[0]["momentum"] == 0.1 and all(int(getattr(enc, f"batchNorm{i}").num_batches_tracked) == 1 for i in range(5))
    enc.eval()
    plan = M._ColPlan(enc, None)
    assert [d["kind"] for d in plan.layers] == ["given"] * 5 and int(enc.batchNorm0.num_batches_tracked) == 1
    enc.batchNorm2.momentum = None
    enc.train()
    with pytest.raises(NotImplementedError):
        M._ColPlan(enc, None)
    inst = M._ColPlan(M.CPCEncoder(32, normMode="instanceNorm").eval(), None)   # instance statistics in eval mode too
    assert [d["kind"] for d in inst.layers] == ["stats"] * 5 and inst.layers[0]["rm"] is None
    assert [d["kind"] for d in M._ColPlan(M.CPCEncoder(32, normMode="ID"), None).layers] == ["ID"] * 5
