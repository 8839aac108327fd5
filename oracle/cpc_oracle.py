"""CPU oracle for the CPC2 training step -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

A plain PyTorch fp32 CPU restatement of the reference algorithm for the one hot
path this repo accelerates (encoder -> AR -> InfoNCE criterion, forward and,
through torch autograd, backward).  Only `tests/`, `__graft_entry__.smoke()` and
`bench.py`'s `cpu_baseline` / `--impl reference` legs may import this module, and
only as the checker / the timed CPU baseline.  The product path
(`cpc2_b200/`) never imports it and has no CPU fallback.

Parity status: the reference's own tests pin no number on this path
(SURVEY.md section 4 / 8c), so the oracle is pinned against outputs of the
reference itself, generated in the build container by
`tests/golden/make_golden.py` (imports /root/reference unmodified) and committed
as `tests/golden/*.pt`.  `tests/test_oracle_golden.py` checks every function
here against those vectors.

Every function cites the reference file:line it restates (paths relative to
the reference root).  Parameters are passed as a flat dict keyed with the
reference's state_dict names, e.g. "gEncoder.conv1.weight",
"gAR.baseNet.weight_hh_l0", "wPrediction.predictors.3.0.multihead.Wq.weight".
"""
import math

import torch
import torch.nn.functional as F

# --------------------------------------------------------------------------- #
# bf16-storage emulation (tools/grad_floor_probe.py only)
# --------------------------------------------------------------------------- #
# With EMULATE_BF16 = True every tensor that ANY bf16-operand implementation of this path has to round -- GEMM
# operands (weights and the activations between layers) and the gradients that travel between layers -- is rounded
# to bf16 in the forward AND in the backward, everything else (accumulation, statistics, softmax, losses) stays fp32.
# The cosine between the gradients of this emulation and the exact fp32 gradients is the noise floor of bf16
# storage for a given fixture; the GPU parity tests derive their per-tensor limits from it instead of typing them in.
EMULATE_BF16 = False


class _RoundBoth(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return x.to(torch.bfloat16).float()

    @staticmethod
    def backward(ctx, g):
        return g.to(torch.bfloat16).float()


class _RoundFwd(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return x.to(torch.bfloat16).float()

    @staticmethod
    def backward(ctx, g):
        return g


class _RoundBwd(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return x

    @staticmethod
    def backward(ctx, g):
        return g.to(torch.bfloat16).float()


# extra rounding sites of one particular kernel design, for tools/attn_grad_probe.py (which rounding costs what):
#   "ds"  the gradient of the attention scores is rounded to bf16 before dS.K and dS^T.Q
#   "p"   the attention probabilities are rounded to bf16
EMULATE_EXTRA = set()


def _q(x):
    """Activation stored in bf16 (its gradient travels in bf16 too)."""
    return _RoundBoth.apply(x) if EMULATE_BF16 else x


def _w(x):
    """Weight consumed as a bf16 GEMM operand (its gradient is accumulated in fp32)."""
    return _RoundFwd.apply(x) if EMULATE_BF16 else x


# (kernel, stride, padding) of conv0..conv4 -- cpc/model.py:85-94
ENCODER_GEOMETRY = ((10, 5, 3), (8, 4, 2), (4, 2, 1), (4, 2, 1), (4, 2, 1))
DOWNSAMPLING = 160  # cpc/model.py:96


# --------------------------------------------------------------------------- #
# encoder
# --------------------------------------------------------------------------- #
def channel_norm(x, weight=None, bias=None, eps=1e-5):
    """cpc/model.py:52-60.  x [N,C,L]; mean / UNBIASED variance over C."""
    mean = x.mean(dim=1, keepdim=True)
    var = x.var(dim=1, keepdim=True)  # unbiased (torch default), /(C-1)
    y = (x - mean) * torch.rsqrt(var + eps)
    if weight is not None:
        y = y * weight + bias
    return y


def encoder_forward(p, x, prefix="gEncoder.", norm_mode="layerNorm", bn_training=False, stats_out=None):
    """cpc/model.py:102-108.  x [N,1,W] -> [N,C,W/160].  norm_mode as cpc/model.py:71-82: "layerNorm" (ChannelNorm),
    "batchNorm" (nn.BatchNorm1d: batch statistics when bn_training, else the running estimates of the state),
    "instanceNorm" (nn.InstanceNorm1d(affine=True): always instance statistics), "ID" (no norm).  `stats_out`: list that
    receives per layer the (batch mean, UNBIASED batch variance) BatchNorm1d folds into its running estimates."""
    for i, (_, s, pad) in enumerate(ENCODER_GEOMETRY):
        w = p[f"{prefix}conv{i}.weight"]
        x = F.conv1d(x, _w(w) if i else w, p[f"{prefix}conv{i}.bias"], stride=s, padding=pad)
        if i or norm_mode != "layerNorm":  # (conv0 + ChannelNorm is one kernel; the other modes store conv0's output)
            x = _q(x)
        if norm_mode == "layerNorm":
            x = channel_norm(x, p[f"{prefix}batchNorm{i}.weight"], p[f"{prefix}batchNorm{i}.bias"])
        elif norm_mode == "batchNorm":
            g, b = p[f"{prefix}batchNorm{i}.weight"], p[f"{prefix}batchNorm{i}.bias"]
            if bn_training:
                if stats_out is not None:
                    stats_out.append((x.detach().mean(dim=(0, 2)), x.detach().var(dim=(0, 2), unbiased=True)))
                x = F.batch_norm(x, None, None, g, b, training=True, eps=1e-5)
            else:
                x = F.batch_norm(x, p[f"{prefix}batchNorm{i}.running_mean"].detach(),
                                 p[f"{prefix}batchNorm{i}.running_var"].detach(), g, b, training=False, eps=1e-5)
        elif norm_mode == "instanceNorm":
            x = F.instance_norm(x, weight=p[f"{prefix}batchNorm{i}.weight"], bias=p[f"{prefix}batchNorm{i}.bias"], eps=1e-5)
        elif norm_mode != "ID":
            raise ValueError(f"Norm mode must be in ['batchNorm', 'instanceNorm', 'ID', 'layerNorm']")
        x = _q(F.relu(x))
    return x


# --------------------------------------------------------------------------- #
# AR networks
# --------------------------------------------------------------------------- #
def lstm_forward(p, x, n_layers, prefix="gAR.baseNet.", state=None):
    """nn.LSTM(batch_first) as built at cpc/model.py:171-173, gate order i,f,g,o.
    x [N,S,C] -> [N,S,H]; returns (out, (h_n, c_n))."""
    hs, cs = [], []
    for l in range(n_layers):
        w_ih, w_hh = p[f"{prefix}weight_ih_l{l}"], p[f"{prefix}weight_hh_l{l}"]
        b = p[f"{prefix}bias_ih_l{l}"] + p[f"{prefix}bias_hh_l{l}"]
        H = w_hh.shape[1]
        N, S, _ = x.shape
        if state is None:
            h = x.new_zeros(N, H)
            c = x.new_zeros(N, H)
        else:
            h, c = state[0][l], state[1][l]
        w_ih, w_hh = _w(w_ih), _w(w_hh)
        gx = _q(x) @ w_ih.t() + b
        out = []
        for t in range(S):
            g = gx[:, t] + _q(h) @ w_hh.t()
            i, f, gg, o = g.split(H, dim=1)
            c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(gg)
            h = torch.sigmoid(o) * torch.tanh(c)
            out.append(h)
        x = torch.stack(out, dim=1)
        hs.append(h)
        cs.append(c)
    return x, (torch.stack(hs), torch.stack(cs))


def gru_forward(p, x, n_layers, prefix="gAR.baseNet.", state=None):
    """nn.GRU(batch_first) as built at cpc/model.py:177-179, gate order r,z,n."""
    hs = []
    for l in range(n_layers):
        w_ih, w_hh = p[f"{prefix}weight_ih_l{l}"], p[f"{prefix}weight_hh_l{l}"]
        b_ih, b_hh = p[f"{prefix}bias_ih_l{l}"], p[f"{prefix}bias_hh_l{l}"]
        H = w_hh.shape[1]
        N, S, _ = x.shape
        h = x.new_zeros(N, H) if state is None else state[l]
        w_ih, w_hh = _w(w_ih), _w(w_hh)
        gx = _q(x) @ w_ih.t() + b_ih
        out = []
        for t in range(S):
            gh = _q(h) @ w_hh.t() + b_hh
            xr, xz, xn = gx[:, t].split(H, dim=1)
            hr, hz, hn = gh.split(H, dim=1)
            r = torch.sigmoid(xr + hr)
            z = torch.sigmoid(xz + hz)
            n = torch.tanh(xn + r * hn)
            h = (1 - z) * n + z * h
            out.append(h)
        x = torch.stack(out, dim=1)
        hs.append(h)
    return x, torch.stack(hs)


def rnn_forward(p, x, n_layers, prefix="gAR.baseNet.", state=None):
    """nn.RNN(tanh, batch_first) as built at cpc/model.py:174-176."""
    hs = []
    for l in range(n_layers):
        w_ih, w_hh = p[f"{prefix}weight_ih_l{l}"], p[f"{prefix}weight_hh_l{l}"]
        b = p[f"{prefix}bias_ih_l{l}"] + p[f"{prefix}bias_hh_l{l}"]
        N, S, _ = x.shape
        h = x.new_zeros(N, w_hh.shape[1]) if state is None else state[l]
        w_ih, w_hh = _w(w_ih), _w(w_hh)
        gx = _q(x) @ w_ih.t() + b
        out = []
        for t in range(S):
            h = torch.tanh(gx[:, t] + _q(h) @ w_hh.t())
            out.append(h)
        x = torch.stack(out, dim=1)
        hs.append(h)
    return x, torch.stack(hs)


def cpc_ar_forward(p, x, mode="LSTM", n_layers=1, reverse=False, prefix="gAR.baseNet.",
                   state=None):
    """cpc/model.py:188-207 (CPCAR.forward)."""
    if reverse:
        x = torch.flip(x, [1])
    fn = {"LSTM": lstm_forward, "GRU": gru_forward, "RNN": rnn_forward}[mode]
    out, h = fn(p, x, n_layers, prefix, state)
    if reverse:
        out = torch.flip(out, [1])
    return out, h


# --------------------------------------------------------------------------- #
# transformer blocks
# --------------------------------------------------------------------------- #
def relpos_causal_attention(q, k, v, krelpos, size_seq, drop_mask=None, drop_p=0.0):
    """cpc/transformers.py:52-70 (ScaledDotProductAttention.forward), written
    without the concatenate/view skew trick:
        score[i,j] = (q_i.k_j + q_i.Krelpos[:, S-1-(i-j)]) / sqrt(dk)   j <= i
                   = -inf                                               j >  i
    q,k,v [Nb, S_in, dk]; inputs longer than size_seq are zero padded and folded
    into the batch (transformers.py:38-50)."""
    N, S_in, dk = q.shape

    def prepare(x):
        r = S_in % size_seq
        if r > 0:
            x = torch.cat([x, x.new_zeros(N, size_seq - r, dk)], dim=1)
        return x.reshape(-1, size_seq, dk)

    q, k, v = prepare(q), prepare(k), prepare(v)
    S = size_seq
    score = q @ k.transpose(1, 2)
    if krelpos is not None:
        qp = q @ krelpos  # [Nb,S,S], column r <-> relative offset S-1-r
        i = torch.arange(S).view(S, 1)
        j = torch.arange(S).view(1, S)
        r = (S - 1 - (i - j)).clamp(0, S - 1)
        rel = torch.gather(qp, 2, r.unsqueeze(0).expand(q.shape[0], S, S))
        rel = torch.where((j <= i).unsqueeze(0), rel, torch.zeros((), dtype=q.dtype))
        score = score + rel
    if "ds" in EMULATE_EXTRA:
        score = _RoundBwd.apply(score)
    score = score / math.sqrt(dk)
    causal = torch.full((S, S), float("-inf")).triu(1)
    att = torch.softmax(score + causal, dim=2)
    if "p" in EMULATE_EXTRA:
        att = _RoundFwd.apply(att)
    if drop_mask is not None:
        att = att * drop_mask / (1.0 - drop_p)
    out = att @ v
    return out.reshape(N, -1, dk)[:, :S_in]


def transformer_layer(p, x, prefix, size_seq, nheads=8, relpos=True, masks=None, drop_p=0.0):
    """cpc/transformers.py:119-134 (TransformerLayer.forward), post-LN:
        y   = LN(x + Wo(MHA(x)))
        out = LN(last_linear(y + lin2(drop(relu(lin1(y))))))
    masks: optional dict {"att": [N*h,S,S], "ffn": [N,S,dff]} keep-masks so a
    train-mode run can be reproduced; None == eval mode."""
    N, S, d = x.shape
    dk = d // nheads

    def split(t):  # transformers.py:88-90
        return t.view(N, S, nheads, dk).transpose(1, 2).reshape(N * nheads, S, dk)

    x = _q(x)
    q = split(_q(x @ _w(p[prefix + "multihead.Wq.weight"]).t()))
    k = split(_q(x @ _w(p[prefix + "multihead.Wk.weight"]).t()))
    v = split(_q(x @ _w(p[prefix + "multihead.Wv.weight"]).t()))
    kr = p[prefix + "multihead.Att.Krelpos"] if relpos else None
    a = relpos_causal_attention(q, k, v, kr, size_seq,
                                None if masks is None else masks["att"], drop_p)
    a = _q(a.view(N, nheads, S, dk).transpose(1, 2).reshape(N, S, d))  # transformers.py:92-95
    a = a @ _w(p[prefix + "multihead.Wo.weight"]).t()
    y = _q(F.layer_norm(_q(x + a), (d,), p[prefix + "ln_multihead.weight"], p[prefix + "ln_multihead.bias"]))
    h = F.relu(y @ _w(p[prefix + "ffnetwork.lin1.weight"]).t() + p[prefix + "ffnetwork.lin1.bias"])
    if masks is not None:
        h = h * masks["ffn"] / (1.0 - drop_p)
    h = _q(h)
    f = h @ _w(p[prefix + "ffnetwork.lin2.weight"]).t() + p[prefix + "ffnetwork.lin2.bias"]
    u = _q(_q(y + f) @ _w(p[prefix + "last_linear.weight"]).t() + p[prefix + "last_linear.bias"])
    dout = u.shape[-1]
    return F.layer_norm(u, (dout,), p[prefix + "ln_ffnetwork.weight"], p[prefix + "ln_ffnetwork.bias"])


def multi_classifier_transformer_head(p, x, prefix, size_seq, nclassifiers, nheads=8, relpos=True):
    """cpc/transformers.py:137-158 (MultiClassifierTransformerHead.forward), eval mode: one attention block, an FFN
    that emits nclassifiers x dmodel features, the SAME last_linear + LayerNorm applied to every classifier slice:
        y   = LN(x + Wo(MHA(x)))
        out = LN(last_linear(lin2(relu(lin1(y))).view(N, S, K, d) + y[:, :, None]))     -> [N, S, K, dout]"""
    N, S, d = x.shape
    dk = d // nheads

    def split(t):
        return t.view(N, S, nheads, dk).transpose(1, 2).reshape(N * nheads, S, dk)

    q = split(x @ p[prefix + "multihead.Wq.weight"].t())
    k = split(x @ p[prefix + "multihead.Wk.weight"].t())
    v = split(x @ p[prefix + "multihead.Wv.weight"].t())
    kr = p[prefix + "multihead.Att.Krelpos"] if relpos else None
    a = relpos_causal_attention(q, k, v, kr, size_seq)
    a = a.view(N, nheads, S, dk).transpose(1, 2).reshape(N, S, d)
    a = a @ p[prefix + "multihead.Wo.weight"].t()
    y = F.layer_norm(x + a, (d,), p[prefix + "ln_multihead.weight"], p[prefix + "ln_multihead.bias"])
    h = F.relu(y @ p[prefix + "ffnetwork.lin1.weight"].t() + p[prefix + "ffnetwork.lin1.bias"])
    f = h @ p[prefix + "ffnetwork.lin2.weight"].t() + p[prefix + "ffnetwork.lin2.bias"]
    u = f.view(N, S, nclassifiers, d) + y.view(N, S, 1, d)
    u = u @ p[prefix + "last_linear.weight"].t() + p[prefix + "last_linear.bias"]
    dout = u.shape[-1]
    return F.layer_norm(u, (dout,), p[prefix + "ln_ffnetwork.weight"], p[prefix + "ln_ffnetwork.bias"])


def static_position_embedding(seqlen, dmodel):
    """cpc/transformers.py:161-173."""
    pos = torch.arange(0., seqlen).unsqueeze(1).repeat(1, dmodel)
    dim = torch.arange(0., dmodel).unsqueeze(0).repeat(seqlen, 1)
    pos = pos * torch.exp(-math.log(10000) * (2 * (dim // 2) / dmodel))
    pe = pos.clone()
    pe[:, 0::2] = torch.sin(pos[:, 0::2])
    pe[:, 1::2] = torch.cos(pos[:, 1::2])
    return pe.unsqueeze(0)


def transformer_ar_forward(p, x, n_layers, size_seq, abspos=False, prefix="gAR."):
    """cpc/transformers.py:176-187 (buildTransformerAR as nn.Sequential)."""
    idx = 0
    if abspos:
        x = x + static_position_embedding(size_seq, x.shape[-1])[:, :x.shape[1]]
        idx = 1
    for l in range(n_layers):
        x = transformer_layer(p, x, f"{prefix}{idx + l}.", size_seq, relpos=not abspos)
    return x


# --------------------------------------------------------------------------- #
# model container
# --------------------------------------------------------------------------- #
def cpc_model_forward(p, batch, ar_mode="LSTM", n_layers=1, reverse=False, size_seq=None,
                      abspos=False, norm_mode="layerNorm", bn_training=False, stats_out=None):
    """cpc/model.py:381-390 (mask_prob == 0).  batch [N,1,W] -> (c [N,S,H], z [N,S,C])."""
    z = encoder_forward(p, batch, norm_mode=norm_mode, bn_training=bn_training, stats_out=stats_out).permute(0, 2, 1)
    if ar_mode == "transformer":
        c = transformer_ar_forward(p, z, n_layers, size_seq, abspos)
    else:
        c, _ = cpc_ar_forward(p, z, ar_mode, n_layers, reverse)
    return c, z


# --------------------------------------------------------------------------- #
# criterion
# --------------------------------------------------------------------------- #
def negative_ext_index(batch_idx, seq_idx, B, S, W, NEG):
    """Integer part of sampleClean, cpc/criterion/criterion.py:256-266.
    batch_idx in [0,B), seq_idx in [1,S), both flat [B*NEG*W] in (b, n, t) order.
    Returns extIdx = ((seq_idx + t) mod S) + batch_idx*S, int64, bit exact."""
    t = torch.arange(W, dtype=torch.int64).view(1, 1, W).expand(B, NEG, W).reshape(-1)
    return torch.remainder(seq_idx + t, S) + batch_idx * S


def draw_negative_indices(B, S, W, NEG, generator=None, device="cpu"):
    """The two RNG draws of sampleClean in the reference's order
    (criterion.py:247-255): first batchIdx = randint(0,B), then seqIdx = randint(1,S)."""
    n = NEG * W * B
    batch_idx = torch.randint(low=0, high=B, size=(n,), generator=generator, device=device)
    seq_idx = torch.randint(low=1, high=S, size=(n,), generator=generator, device=device)
    return batch_idx, seq_idx


def prediction_heads(p, c, K, rnn_mode, size_seq, prefix="wPrediction.predictors.",
                     masks=None, drop_p=0.0, multihead=False):
    """cpc/criterion/criterion.py:111-150,161-163: K independent heads on the same c; multihead
    (--multihead_rnn, criterion.py:44-94): one MultiClassifierTransformerHead, prediction[:, :, k] per step."""
    if multihead:
        pred = multi_classifier_transformer_head(p, c, "wPrediction.predictor.0.", size_seq, K)
        return [pred[:, :, k] for k in range(K)]
    out = []
    for k in range(K):
        if rnn_mode == "transformer":
            m = None if masks is None else masks[k]
            out.append(transformer_layer(p, c, f"{prefix}{k}.0.", size_seq, masks=m, drop_p=drop_p))
        elif rnn_mode == "ffd":
            # criterion.py:10-19,125-128: FFNetwork of two EqualizedLinear layers (custom_layers.py:44-76,134-151):
            # y = (x W^T + b) * sqrt(2 / fan_in), weights N(0, 1), no dropout (p = 0)
            w1, b1 = p[f"{prefix}{k}.lin1.module.weight"], p[f"{prefix}{k}.lin1.module.bias"]
            w2, b2 = p[f"{prefix}{k}.lin2.module.weight"], p[f"{prefix}{k}.lin2.module.bias"]
            h = F.relu((c @ w1.t() + b1) * math.sqrt(2.0 / w1.shape[1]))
            out.append((h @ w2.t() + b2) * math.sqrt(2.0 / w2.shape[1]))
        elif rnn_mode == "LSTM":
            # criterion.py:120-123: nn.LSTM(dimAR, dimEncoder, batch_first=True), zero initial state, output sequence
            out.append(lstm_forward(p, c, 1, prefix=f"{prefix}{k}.")[0])
        elif rnn_mode == "RNN":
            # criterion.py:116-119: nn.RNN(dimAR, dimEncoder) WITHOUT batch_first, fed c [B, W, H] as is: the
            # recurrence runs over dim 0 (the B windows of the batch), the W frames are its "batch"
            out.append(rnn_forward(p, c.transpose(0, 1), 1, prefix=f"{prefix}{k}.")[0].transpose(0, 1))
        elif rnn_mode in ("conv4", "conv8", "conv12"):
            # criterion.py:23-41,129-137: ShiftedConv = ks-1 zero frames in front, then EqualizedConv1d
            # (custom_layers.py:44-105): (conv1d(x) + b) * sqrt(2 / (Cin * ks)), weights N(0, 1)
            w, b = p[f"{prefix}{k}.module.module.weight"], p[f"{prefix}{k}.module.module.bias"]
            ks = w.shape[2]
            x = F.pad(c.permute(0, 2, 1), (ks - 1, 0))
            y = F.conv1d(x, w, b) * math.sqrt(2.0 / (w.shape[1] * ks))
            out.append(y.permute(0, 2, 1))
        else:  # any other string falls through to the bias-free Linear (criterion.py:145-146)
            out.append(c @ _w(p[f"{prefix}{k}.weight"]).t())
    return out


def criterion_forward(p, c, z, batch_idx, seq_idx, K, NEG, rnn_mode="transformer", mode=None,
                      n_skipped=0, weights=None, masks=None, drop_p=0.0, return_logits=False, multihead=False):
    """cpc/criterion/criterion.py:329-363 (forward) with getPrediction :291-302,
    sampleClean :237-286 and PredictionNetwork.forward :152-173.

    c [B,S,H], z [B,S,D]; batch_idx/seq_idx are the two randint draws.
    weights: optional [B*W] per-token quality weights (criterion.py:334-338);
    None == ones (:340).  Returns (losses [1,K-n_skipped], acc [1,K-n_skipped],
    [logits list of [B*W, NEG+1], argmax list])."""
    if mode == "reverse":
        z = torch.flip(z, [1])
        c = torch.flip(c, [1])
    B, S, D = z.shape
    W = S - K
    cw = _q(c[:, :W])
    z = _q(z)
    ext = negative_ext_index(batch_idx, seq_idx, B, S, W, NEG)
    neg = z.reshape(-1, D)[ext].view(B, NEG, W, D)
    preds = prediction_heads(p, cw, K, rnn_mode, W, masks=masks, drop_p=drop_p, multihead=multihead)
    if EMULATE_BF16:
        preds = [_q(t) for t in preds]
    losses, accs, logits_all, arg_all = [], [], [], []
    for k in range(1, K + 1):
        pos = z[:, k:k + W].unsqueeze(1)                       # criterion.py:276-282
        cand = torch.cat([pos, neg], dim=1)                    # [B,NEG+1,W,D]
        logit = (preds[k - 1].unsqueeze(1) * cand).mean(dim=3)  # criterion.py:170-171 (scale 1/D)
        logit = logit.permute(0, 2, 1).reshape(B * W, NEG + 1)  # criterion.py:346-348
        ce = torch.logsumexp(logit, dim=1) - logit[:, 0]        # CE vs class 0, reduction none
        if weights is not None:
            ce = weights * ce
        losses.append(ce.mean().view(1, 1))
        arg = logit.max(1)[1]                                   # first maximal index
        accs.append((arg == 0).sum().float().view(1, 1))
        if return_logits:
            logits_all.append(logit)
            arg_all.append(arg)
    losses = torch.cat(losses[n_skipped:], dim=1)
    accs = torch.cat(accs[n_skipped:], dim=1) / (W * B)
    if return_logits:
        return losses, accs, logits_all, arg_all
    return losses, accs


def quality_weighting(signal_quality, W, growth_rate, inflection_point_x):
    """cpc/criterion/criterion.py:230,334-338."""
    q = signal_quality.mean(dim=1)
    w = 0.00001 + 1 / (1 + torch.exp(-growth_rate * (q - inflection_point_x)))
    return w.unsqueeze(1).repeat(1, W).reshape(-1)


# --------------------------------------------------------------------------- #
# the training step (forward + backward), cpc/train.py:95-109
# --------------------------------------------------------------------------- #
def train_step(p, sequence, batch_idx, seq_idx, K=12, NEG=128, ar_mode="LSTM", n_layers=1,
               rnn_mode="transformer", cpc_mode=None, abspos=False, n_skipped=0, with_grad=True, multihead=False,
               signal_quality=None, growth_rate=None, inflection_point_x=None, norm_mode="layerNorm", bn_training=False):
    """sequence [B,2,1,W] fp32.  Returns dict(loss [1,K], acc [1,K], c, z, grads{name: tensor}).
    signal_quality [B, n] (train.py:104-106): per-token loss weights through quality_weighting."""
    B = sequence.shape[0]
    if with_grad:
        p = {k: v.detach().clone().requires_grad_(True) for k, v in p.items()}
    combined = torch.cat([sequence[:, 0], sequence[:, 1]], dim=0)       # train.py:95-99
    S = combined.shape[-1] // DOWNSAMPLING
    bn_stats = []
    c, z = cpc_model_forward(p, combined, ar_mode, n_layers, cpc_mode == "reverse", S, abspos, norm_mode, bn_training,
                             bn_stats)
    c_past, z_future = c[:B], z[B:]                                      # train.py:102-103
    weights = None
    if signal_quality is not None:
        weights = quality_weighting(signal_quality, S - K, growth_rate, inflection_point_x)
    loss, acc = criterion_forward(p, c_past, z_future, batch_idx, seq_idx, K, NEG, rnn_mode,
                                  mode=cpc_mode, n_skipped=n_skipped, multihead=multihead, weights=weights)
    out = {"loss": loss.detach(), "acc": acc.detach(), "c": c.detach(), "z": z.detach(), "bn_stats": bn_stats}
    if with_grad:
        loss.sum().backward()                                            # train.py:108-109
        out["grads"] = {k: v.grad for k, v in p.items() if v.grad is not None}
    return out
