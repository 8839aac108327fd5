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


def test_train_mode_dropout_is_pinned_end_to_end():
    """the default heads run with dropout 0.1 on the attention probabilities and on the feed-forward hidden
    layer in TRAIN mode (cpc/transformers.py:69,115); every reference-generated fixture is eval mode.  Here the criterion
    runs forward + backward in train mode, the very keep-masks its kernels drew are recovered from the same device
    (seed, step) pair -- the mask is a pure function of (rng, site, element index): attention with V = identity returns
    the masked probabilities, a GEMM with zero weights and bias 1 returns the feed-forward mask -- and fed to the
    oracle's transformer_layer(masks=..., drop_p=0.1).  Pins mask placement, the 1/(1-p) scaling and its backward."""
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    from cpc2_b200 import criterion as cr
    from cpc2_b200 import kernels as K
    from cpc2_b200 import ops
    from oracle import cpc_oracle as oc
    B, S, D, Kp, NEG, p = 4, 34, 256, 2, 16, 0.1
    W = S - Kp
    Hh, dk, dff = 8, D // 8, 2048
    torch.manual_seed(0)
    crit = cr.CPCUnsupersivedCriterion(Kp, D, D, NEG, rnnMode="transformer", sizeInputSeq=S).cuda().train()
    g = torch.Generator().manual_seed(23)
    c = torch.randn(B, S, D, generator=g)
    z = torch.randn(B, S, D, generator=g)
    torch.manual_seed(5)
    bi, si = oc.draw_negative_indices(B, S, W, NEG)
    ext = oc.negative_ext_index(bi, si, B, S, W, NEG)
    crit.sampleIndices = lambda *a: ext.cuda()
    dev = torch.device("cuda", torch.cuda.current_device())
    ops.DropoutState.next(dev)  # make sure the state exists
    rng = ops.DropoutState._states[(dev.type, dev.index)].clone()  # the (seed, step) the next forward will use
    cg, zg = c.cuda().requires_grad_(True), z.cuda().requires_grad_(True)
    losses, acc = crit(cg, zg, torch.zeros(B, dtype=torch.long, device="cuda"), None)
    losses.sum().backward()
    # a second train-mode call draws a different mask (the step counter advanced)
    with torch.no_grad():
        losses2, _ = crit(cg, zg, torch.zeros(B, dtype=torch.long, device="cuda"), None)
    assert (losses2 - losses).abs().max() > 1e-6
    # ---- recover the masks of the FIRST call
    T = B * W
    qkv = torch.zeros(Kp, B, W, 3, Hh, dk, device="cuda", dtype=torch.bfloat16)
    qkv[:, :, :, 0] = 0.01  # any q, k: only which probabilities survive matters
    for j in range(W):
        qkv[:, :, j, 2, :, j] = 1  # V = identity (W <= dk)
    out = torch.empty(Kp, T, D, device="cuda", dtype=torch.bfloat16)
    probs = torch.zeros(Kp, B, Hh, W, W, device="cuda", dtype=torch.bfloat16)
    K.attn_fwd(qkv.view(Kp, T, 3 * D), None, out, probs, Kp, B, Hh, W, dk, rng=rng, site=1, drop_p=p)
    att_keep = (out.view(Kp, B, W, Hh, dk).permute(0, 1, 3, 2, 4)[..., :W] != 0).float()  # [K, B, h, i, j]
    tri = torch.tril(torch.ones(W, W, device="cuda"))
    att_keep = att_keep * tri + (1 - tri)  # (masked-out future positions: irrelevant, keep them 1)
    frac = float((att_keep * tri).sum() / (tri.sum() * Kp * B * Hh))
    assert abs(frac - (1 - p)) < 0.02, frac
    x0 = torch.zeros(T, D, device="cuda", dtype=torch.bfloat16)
    w0 = torch.zeros(Kp, dff, D, device="cuda", dtype=torch.bfloat16)
    h = torch.empty(Kp, T, dff, device="cuda", dtype=torch.bfloat16)
    ops.linear_fwd(x0, w0, T, dff, D, groups=Kp, x_shared=True, out_bf16=h, bias=torch.ones(Kp, dff, device="cuda"),
                   bias_gs=dff, relu=True, rng=rng, rng_site=2, drop_p=p)
    ffn_keep = (h != 0).float().view(Kp, B, W, dff)
    assert abs(float(ffn_keep.mean()) - (1 - p)) < 0.01
    masks = [{"att": att_keep[k].reshape(B * Hh, W, W).cpu(), "ffn": ffn_keep[k].cpu()} for k in range(Kp)]
    # ---- the oracle with those masks
    state = {k: v.detach().float().cpu() for k, v in crit.state_dict().items()}
    params = {k: v.clone().requires_grad_(True) for k, v in state.items() if v.is_floating_point()}
    full = dict(state)
    full.update(params)
    cr_, zr_ = c.clone().requires_grad_(True), z.clone().requires_grad_(True)
    rl, ra = oc.criterion_forward(full, cr_, zr_, bi, si, Kp, NEG, rnn_mode="transformer", masks=masks, drop_p=p)
    rl.sum().backward()
    # and without masks (eval): must be visibly different from the train-mode loss, so the check above is not vacuous
    el, _ = oc.criterion_forward(state, c, z, bi, si, Kp, NEG, rnn_mode="transformer")
    assert (losses.detach().cpu() - rl.detach()).abs().max() <= 2e-2, (losses, rl)
    assert (el - rl.detach()).abs().max() > 1e-4

    def cosine(a, b):
        return F.cosine_similarity(a.double().flatten(), b.double().flatten(), dim=0).item()
    assert cosine(cg.grad.cpu(), cr_.grad) >= 0.99
    assert cosine(zg.grad.cpu(), zr_.grad) >= 0.99
    for name, q in crit.named_parameters():
        ref = params[name].grad
        if ref is None or ref.abs().max() == 0:
            continue
        assert cosine(q.grad.cpu(), ref) >= 0.985, (name, cosine(q.grad.cpu(), ref))
