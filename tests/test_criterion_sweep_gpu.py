"""BASELINE.json configs[4]: the InfoNCE criterion alone (CPCUnsupersivedCriterion.forward + backward,
cpc/criterion/criterion.py:329-363) on random c, z ~ N(0, 1) [8, 128, 256], over negatives x prediction steps,
against the CPU oracle on the same sampled indices.  Tolerances: loss abs <= 2e-2 (logits are O(1/sqrt(D)) here, the
loss sits near ln(NEG + 1)), accuracy abs <= 1 % + one token, gradient cosine >= 0.99 (bf16 operands)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

CASES = [(32, 4, "linear"), (64, 8, "linear"), (256, 16, "linear"), (1024, 24, "linear"), (512, 12, "linear"),
         (32, 24, "linear"), (1024, 4, "linear"), (128, 12, "transformer"), (64, 4, "transformer")]


@pytest.mark.parametrize("NEG,Kp,rnn", CASES)
def test_criterion_only_matches_oracle(NEG, Kp, rnn):
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    from cpc2_b200 import criterion as cr
    from oracle import cpc_oracle as oc
    B, S, D = 8, 128, 256
    torch.manual_seed(0)
    crit = cr.CPCUnsupersivedCriterion(Kp, D, D, NEG, rnnMode=rnn, sizeInputSeq=S).cuda().eval()
    g = torch.Generator().manual_seed(17)
    c = torch.randn(B, S, D, generator=g)
    z = torch.randn(B, S, D, generator=g)
    cg, zg = c.cuda().requires_grad_(True), z.cuda().requires_grad_(True)
    captured = {}
    orig = crit.sampleIndices
    crit.sampleIndices = lambda *a: captured.setdefault("ext", orig(*a))
    torch.manual_seed(4321)
    losses, acc = crit(cg, zg, torch.zeros(B, dtype=torch.long, device="cuda"), None)
    losses.sum().backward()
    crit.sampleIndices = orig
    W = S - Kp
    ext = captured["ext"].cpu()
    t = torch.arange(W).view(1, 1, W).expand(B, NEG, W).reshape(-1)
    batch_idx, seq_idx = ext // S, torch.remainder(ext % S - t, S)
    seq_idx = torch.where(seq_idx == 0, torch.full_like(seq_idx, S), seq_idx)
    assert torch.equal(oc.negative_ext_index(batch_idx, seq_idx, B, S, W, NEG), ext)  # indices bit exact
    state = {k: v.detach().float().cpu() for k, v in crit.state_dict().items()}
    cr_, zr_ = c.clone().requires_grad_(True), z.clone().requires_grad_(True)
    params = {k: v.clone().requires_grad_(True) for k, v in state.items() if v.is_floating_point()}
    full = dict(state)
    full.update(params)
    rl, ra = oc.criterion_forward(full, cr_, zr_, batch_idx, seq_idx, Kp, NEG, rnn_mode=rnn)
    rl.sum().backward()
    assert losses.shape == rl.shape == (1, Kp)
    assert (losses.detach().cpu() - rl.detach()).abs().max() <= 2e-2, (losses, rl)
    assert (acc.cpu() - ra).abs().max() <= 0.01 + 1.0 / (B * W)

    def cosine(a, b):
        return F.cosine_similarity(a.double().flatten(), b.double().flatten(), dim=0).item()
    assert cosine(cg.grad.cpu(), cr_.grad) > 0.99
    assert cosine(zg.grad.cpu(), zr_.grad) > 0.99
    for name, p in crit.named_parameters():
        ref = params[name].grad
        if ref is None or ref.abs().max() == 0:
            continue
        assert cosine(p.grad.cpu(), ref) > 0.985, (name, cosine(p.grad.cpu(), ref))
