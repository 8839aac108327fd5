"""Drop-in for the reference's cpc/transformers.py (causal block-local attention with relative-position
keys, post-LN transformer block, builders) -- same classes, constructor signatures, parameter AND buffer
names, so reference checkpoints load with strict=True.  The modules are parameter containers; the math runs
in `TransformerStack`, which executes G independent blocks (the K prediction heads of the criterion, or one
AR layer with G = 1) as grouped tcgen05 GEMMs + fused attention / LayerNorm kernels.
"""
import math

import torch
import torch.nn as nn

from . import kernels as K
from . import ops
from .ops import BF16


class ScaledDotProductAttention(nn.Module):
    """cpc/transformers.py:10-36 (parameters / buffers only; compute is cpc_attn_fwd/bwd)."""

    def __init__(self, sizeSeq, dk, dropout, relpos=False):
        super().__init__()
        self.drop = nn.Dropout(dropout)
        self.softmax = nn.Softmax(dim=2)
        self.relpos = relpos
        self.sizeSeq = sizeSeq
        if relpos:
            self.Krelpos = nn.Parameter(torch.Tensor(dk, sizeSeq))
            self.initmat_(self.Krelpos)
            self.register_buffer('z', torch.zeros(1, sizeSeq, 1))
        mask = torch.tril(torch.ones(sizeSeq, sizeSeq), diagonal=0)
        mask = 1 - mask
        mask[mask == 1] = -float('inf')
        self.register_buffer('mask', mask.unsqueeze(0))

    def initmat_(self, mat, dim=0):
        stdv = 1. / math.sqrt(mat.size(dim))
        mat.data.uniform_(-stdv, stdv)


class MultiHeadAttention(nn.Module):
    """cpc/transformers.py:73-87."""

    def __init__(self, sizeSeq, dropout, dmodel, nheads, abspos):
        super().__init__()
        self.Wo = nn.Linear(dmodel, dmodel, bias=False)
        self.Wk = nn.Linear(dmodel, dmodel, bias=False)
        self.Wq = nn.Linear(dmodel, dmodel, bias=False)
        self.Wv = nn.Linear(dmodel, dmodel, bias=False)
        self.nheads = nheads
        self.dk = dmodel // nheads
        self.Att = ScaledDotProductAttention(sizeSeq, self.dk, dropout, not abspos)


class FFNetwork(nn.Module):
    """cpc/transformers.py:107-113."""

    def __init__(self, din, dout, dff, dropout):
        super().__init__()
        self.lin1 = nn.Linear(din, dff, bias=True)
        self.lin2 = nn.Linear(dff, dout, bias=True)
        self.relu = nn.ReLU()
        self.drop = nn.Dropout(dropout)


class TransformerLayer(nn.Module):
    """cpc/transformers.py:119-134: y = LN(x + MHA(x)); out = LN(last_linear(y + FFN(y)))."""

    def __init__(self, sizeSeq=32, dmodel=512, dout=512, dff=2048, dropout=0.1, nheads=8, abspos=False):
        super().__init__()
        self.multihead = MultiHeadAttention(sizeSeq, dropout, dmodel, nheads, abspos)
        self.ln_multihead = nn.LayerNorm(dmodel)
        self.ffnetwork = FFNetwork(dmodel, dmodel, dff, dropout)
        self.last_linear = nn.Linear(dmodel, dout)
        self.ln_ffnetwork = nn.LayerNorm(dout)
        self.sizeSeq, self.dmodel, self.dout, self.dff = sizeSeq, dmodel, dout, dff
        self.dropout_p, self.nheads = dropout, nheads
        self._stack = None

    def stack(self):
        if self._stack is None:
            object.__setattr__(self, "_stack", TransformerStack([self]))
        return self._stack

    def prefetch(self):
        self.stack().arena.prefetch(4)

    def forward(self, x):
        """x [N, S, dmodel] fp32 -> [N, S, dout] fp32 (any S: block-local attention over zero-padded blocks of sizeSeq)."""
        ops.require_cuda(x)
        ops.reject_replica(self)
        st = self.stack()
        return _TransformerFn.apply(x, st, self.training, *st.arena.params)


class StaticPositionEmbedding(nn.Module):
    """cpc/transformers.py:161-173."""

    def __init__(self, seqlen, dmodel):
        super().__init__()
        pos = torch.arange(0., seqlen).unsqueeze(1).repeat(1, dmodel)
        dim = torch.arange(0., dmodel).unsqueeze(0).repeat(seqlen, 1)
        div = torch.exp(- math.log(10000) * (2 * (dim // 2) / dmodel))
        pos *= div
        pos[:, 0::2] = torch.sin(pos[:, 0::2])
        pos[:, 1::2] = torch.cos(pos[:, 1::2])
        self.register_buffer('pe', pos.unsqueeze(0))

    def forward(self, x):
        return x + self.pe[:, :x.size(1), :]


def buildTransformerAR(dimEncoded, dimAR, nLayers, sizeSeq, abspos):
    """cpc/transformers.py:176-187."""
    layerSequence = []
    if abspos:
        layerSequence += [StaticPositionEmbedding(sizeSeq, dimAR)]
    layerSequence += [TransformerLayer(sizeSeq=sizeSeq, dmodel=dimAR, dout=dimEncoded, abspos=abspos)
                      for i in range(nLayers)]
    return nn.Sequential(*layerSequence)


class MultiClassifierTransformerHead(nn.Module):
    """cpc/transformers.py:137-158 (--multihead_rnn): one attention block whose feed-forward network emits
    nclassifiers x dmodel features; last_linear and the final LayerNorm are shared by the classifiers.
    forward(x [N, S, dmodel]) -> [N, S, nclassifiers, dout]."""

    def __init__(self, nclassifiers, sizeSeq=32, dmodel=512, dout=512, dff=2048, dropout=0.1, nheads=8, abspos=False):
        super().__init__()
        self.multihead = MultiHeadAttention(sizeSeq, dropout, dmodel, nheads, abspos)
        self.ln_multihead = nn.LayerNorm(dmodel)
        self.ffnetwork = FFNetwork(dmodel, dmodel * nclassifiers, dff, dropout)
        self.last_linear = nn.Linear(dmodel, dout)
        self.ln_ffnetwork = nn.LayerNorm(dout)
        self.nclassifiers = nclassifiers
        self.dout = dmodel  # as in the reference (:151): the width of one classifier slice
        self.sizeSeq, self.dmodel, self.dff = sizeSeq, dmodel, dff
        self.dropout_p, self.nheads = dropout, nheads
        self._stack = None

    def stack(self):
        if self._stack is None:
            object.__setattr__(self, "_stack", MultiHeadStack(self))
        return self._stack

    def prefetch(self):
        self.stack().arena.prefetch(4)

    def forward(self, x):
        ops.require_cuda(x)
        ops.reject_replica(self)
        st = self.stack()
        N, S_in, _ = x.shape
        out = _MultiHeadFn.apply(x, st, self.training, *st.arena.params)  # [K, N*S, dout]
        return out.view(st.n_out, N, S_in, st.dout).permute(1, 2, 0, 3)


def buildMultHeadTransformerAR(dimEncoded, dimAR, nLayers, sizeSeq, abspos, nHeads):
    """cpc/transformers.py:190-212."""
    layerSequence = []
    if abspos:
        layerSequence += [StaticPositionEmbedding(sizeSeq, dimAR)]
    layerSequence += [TransformerLayer(sizeSeq=sizeSeq, dmodel=dimAR, dout=dimEncoded, abspos=abspos)
                      for i in range(nLayers - 1)]
    layerSequence += [MultiClassifierTransformerHead(nHeads, dmodel=dimAR, dout=dimEncoded, sizeSeq=sizeSeq,
                                                     abspos=abspos)]
    return nn.Sequential(*layerSequence)


class TransformerStack:
    """G independent TransformerLayers (identical shapes) executed together.  Owns the parameter arena:
    stacked operands [G][rows][cols] are contiguous spans of it, their bf16 shadow is one cast per step."""

    KINDS = ("Wo", "W1", "b1", "W2", "b2", "Wl", "bl", "ln1w", "ln1b", "ln2w", "ln2b")

    def __init__(self, layers):
        self.layers = list(layers)
        l0 = self.layers[0]
        self.G = len(self.layers)
        self.S, self.d, self.dout, self.dff = l0.sizeSeq, l0.dmodel, l0.last_linear.out_features, l0.dff
        self.n_out = self.G  # prediction groups this engine emits
        self.Hh, self.dk = l0.nheads, l0.dmodel // l0.nheads
        self.p_drop = l0.dropout_p
        self.relpos = l0.multihead.Att.relpos
        if self.d % 32 or self.dout % 32 or self.dff % 8 or self.S > 256 or self.dk > 64:
            raise NotImplementedError("TransformerStack: dmodel/dout must be multiples of 32, sizeSeq <= 256, dk <= 64")
        params = []
        for l in self.layers:
            params += [l.multihead.Wq.weight, l.multihead.Wk.weight, l.multihead.Wv.weight]
        getters = {
            "Wo": lambda l: l.multihead.Wo.weight, "W1": lambda l: l.ffnetwork.lin1.weight,
            "b1": lambda l: l.ffnetwork.lin1.bias, "W2": lambda l: l.ffnetwork.lin2.weight,
            "b2": lambda l: l.ffnetwork.lin2.bias, "Wl": lambda l: l.last_linear.weight,
            "bl": lambda l: l.last_linear.bias, "ln1w": lambda l: l.ln_multihead.weight,
            "ln1b": lambda l: l.ln_multihead.bias, "ln2w": lambda l: l.ln_ffnetwork.weight,
            "ln2b": lambda l: l.ln_ffnetwork.bias,
        }
        self.first = {"Wqkv": 0}
        for kind in self.KINDS:
            self.first[kind] = len(params)
            params += [getters[kind](l) for l in self.layers]
        if self.relpos:
            self.first["krel"] = len(params)
            params += [l.multihead.Att.Krelpos for l in self.layers]
        self.arena = ops.Arena(params)

    def shapes(self):
        G, d, dout, dff = self.G, self.d, self.dout, self.dff
        sh = {"Wqkv": (G, 3 * d, d), "Wo": (G, d, d), "W1": (G, dff, d), "b1": (G, dff), "W2": (G, d, dff),
              "b2": (G, d), "Wl": (G, dout, d), "bl": (G, dout), "ln1w": (G, d), "ln1b": (G, d),
              "ln2w": (G, dout), "ln2b": (G, dout)}
        if self.relpos:
            sh["krel"] = (G, self.dk, self.S)
        return sh

    def spans(self, flat):
        out = {}
        for name, shape in self.shapes().items():
            count = 3 * self.G if name == "Wqkv" else self.G
            out[name] = self.arena.span(flat, self.first[name], count, shape)
        return out

    # ------------------------------------------------------------------------------------------ attention
    def _blocks(self, T, seq_len):
        """(tokens per sequence L, sequences N, blocks per sequence nb).  cpc/transformers.py:38-50: a sequence longer
        than sizeSeq is zero-padded to a multiple of it and folded into the batch as independent causal blocks."""
        S = self.S
        L = S if seq_len is None else int(seq_len)
        if L <= 0 or T % L != 0:
            raise ValueError(f"token count {T} is not a multiple of the sequence length {L}")
        return L, T // L, (L + S - 1) // S

    def _attn_fwd(self, qkv, krel, G, T, seq_len, rng, drop):
        """qkv bf16 [G, T, 3d] -> att bf16 [G, T, d] + what the backward needs.  When the sequences are not a whole
        number of blocks the padded copies (pure data movement: zero rows appended per sequence, transformers.py:41-48,
        cropped again after the product, :69) are what the attention kernels see."""
        S, d = self.S, self.d
        dev = qkv.device
        L, N, nb = self._blocks(T, seq_len)
        Bt = N * nb
        padded = nb * S != L
        if padded:
            qkv_k = torch.zeros(G, N, nb * S, 3 * d, device=dev, dtype=BF16)
            qkv_k[:, :, :L].copy_(qkv.view(G, N, L, 3 * d))
        else:
            qkv_k = qkv
        att_k = torch.empty(G, N * nb * S, d, device=dev, dtype=BF16)
        probs = torch.empty(G, Bt, self.Hh, S, S, device=dev, dtype=BF16)
        # fp32 copy of the attention output + the row sums of the stored (bf16) probabilities: the backward's exact row
        # statistic D_i = dO_i . O_i / rowsum_i without a second dP sweep (CpcAttn.out_f32 / rowsum)
        att32 = torch.empty(G, N * nb * S, d, device=dev) if ops.will_backward() else None
        rowsum = torch.empty(G, Bt, self.Hh, S, device=dev) if ops.will_backward() else None
        K.attn_fwd(qkv_k, krel, att_k, probs, G, Bt, self.Hh, S, self.dk, rng=rng, site=1, drop_p=drop, out_f32=att32,
                   rowsum=rowsum)
        att = att_k.view(G, N, nb * S, d)[:, :, :L].reshape(G, T, d) if padded else att_k
        return att, dict(qkv=qkv_k, att=att_k, att32=att32, rowsum=rowsum, probs=probs, Bt=Bt, geom=(L, N, nb, padded))

    def _attn_bwd(self, sa, krel, dkrel, datt, G, T, rng, drop):
        """datt bf16 [G, T, d] -> dqkv bf16 [G, T, 3d] (dKrelpos accumulated into `dkrel`)."""
        S, d = self.S, self.d
        dev = datt.device
        L, N, nb, padded = sa["geom"]
        if padded:
            datt_k = torch.zeros(G, N, nb * S, d, device=dev, dtype=BF16)
            datt_k[:, :, :L].copy_(datt.view(G, N, L, d))
        else:
            datt_k = datt
        dqkv_k = torch.empty(G, N * nb * S, 3 * d, device=dev, dtype=BF16)
        dscore = torch.empty_like(sa["probs"])
        K.attn_bwd(sa["qkv"], krel, sa["att"], sa["probs"], datt_k, dqkv_k, dscore, dkrel, G, sa["Bt"], self.Hh, S,
                   self.dk, rng=rng, site=1, drop_p=drop, out_f32=sa["att32"], rowsum=sa["rowsum"])
        return dqkv_k.view(G, N, nb * S, 3 * d)[:, :, :L].reshape(G, T, 3 * d) if padded else dqkv_k

    # ------------------------------------------------------------------------------------------ forward
    def forward(self, x, T, shared, training, want_f32=False, seq_len=None):
        """x bf16 [T, d] (shared by all groups) or [G, T, d]; `seq_len`: tokens per sequence (default sizeSeq).
        Returns (out bf16 [G,T,dout], out_f32|None, saved)."""
        G, d, dout, dff, S = self.G, self.d, self.dout, self.dff, self.S
        dev = x.device
        w32 = self.arena.ensure()
        P32 = self.spans(w32)
        P16 = self.spans(self.arena.bf16())
        drop = self.p_drop if training else 0.0
        rng = ops.DropoutState.next(dev, side_index=5) if drop > 0 else None  # first read by the attention kernel
        qkv = torch.empty(G, T, 3 * d, device=dev, dtype=BF16)
        ops.linear_fwd(x, P16["Wqkv"], T, 3 * d, d, groups=G, x_shared=shared, out_bf16=qkv)
        if rng is not None:
            ops.join(5)
        att, sa = self._attn_fwd(qkv, P32.get("krel"), G, T, seq_len, rng, drop)
        pre1 = torch.empty(G, T, d, device=dev, dtype=BF16)
        res = (x, d, 0) if shared else x
        y = torch.empty(G, T, d, device=dev, dtype=BF16)
        mean1 = torch.empty(G * T, device=dev)
        rstd1 = torch.empty(G * T, device=dev)
        if ops.can_fuse_norm(d, T, G):  # y = LN(x + att . Wo^T) in the epilogue of the Wo product
            ops.linear_fwd(att, P16["Wo"], T, d, d, groups=G, out_bf16=pre1, res=res,
                           force_bn=ops.NORM_FORCE_BN, norm=dict(gamma=P32["ln1w"], beta=P32["ln1b"], gs_param=d, out_bf16=y, mean=mean1, rstd=rstd1,
                                     gs_stat=T))
        else:
            ops.linear_fwd(att, P16["Wo"], T, d, d, groups=G, out_bf16=pre1, res=res)
            K.rownorm_fwd(pre1, G * T, d, P32["ln1w"], P32["ln1b"], out_bf16=y, mean=mean1, rstd=rstd1, rows_per_group=T)
        h1 = torch.empty(G, T, dff, device=dev, dtype=BF16)
        ops.linear_fwd(y, P16["W1"], T, dff, d, groups=G, out_bf16=h1, bias=P32["b1"], bias_gs=dff, relu=True,
                       rng=rng, rng_site=2, drop_p=drop)
        u = torch.empty(G, T, d, device=dev, dtype=BF16)
        ops.linear_fwd(h1, P16["W2"], T, d, dff, groups=G, out_bf16=u, bias=P32["b2"], bias_gs=d, res=y)
        pre2 = torch.empty(G, T, dout, device=dev, dtype=BF16)
        out = torch.empty(G, T, dout, device=dev, dtype=BF16)
        out32 = torch.empty(G, T, dout, device=dev) if want_f32 else None
        mean2 = torch.empty(G * T, device=dev)
        rstd2 = torch.empty(G * T, device=dev)
        if ops.can_fuse_norm(dout, T, G):  # out = LN(u . Wl^T + bl) in the epilogue of the last_linear product
            ops.linear_fwd(u, P16["Wl"], T, dout, d, groups=G, out_bf16=pre2, bias=P32["bl"], bias_gs=dout,
                           force_bn=ops.NORM_FORCE_BN, norm=dict(gamma=P32["ln2w"], beta=P32["ln2b"], gs_param=dout, out_bf16=out, out_f32=out32,
                                     mean=mean2, rstd=rstd2, gs_stat=T))
        else:
            ops.linear_fwd(u, P16["Wl"], T, dout, d, groups=G, out_bf16=pre2, bias=P32["bl"], bias_gs=dout)
            K.rownorm_fwd(pre2, G * T, dout, P32["ln2w"], P32["ln2b"], out_bf16=out, out_f32=out32, mean=mean2,
                          rstd=rstd2, rows_per_group=T)
        saved = dict(x=x, shared=shared, T=T, P16=P16, P32=P32, rng=rng, drop=drop, sa=sa, att=att,
                     pre1=pre1, y=y, mean1=mean1, rstd1=rstd1, h1=h1, u=u, pre2=pre2, mean2=mean2,
                     rstd2=rstd2)
        return out, out32, saved

    # ------------------------------------------------------------------------------------------ backward
    def backward(self, sv, dout, early_stage=None, defer=False):
        """dout bf16 or fp32 [G, T, dout].  Returns (dx fp32 [T,d] if shared else [G,T,d], flat fp32 grad arena).
        early_stage: announce the feed-forward weights' gradients (W1, b1, W2, b2: 3/4 of the parameters) through
        ops.grad_ready(early_stage + ".ffn") as soon as they exist, a third of the way into this backward, so their
        all-reduce / optimizer step overlaps the rest of it; sv["announced"] tells the caller what is left."""
        G, d, do, dff, S = self.G, self.d, self.dout, self.dff, self.S
        T, P16, P32 = sv["T"], sv["P16"], sv["P32"]
        dev = dout.device
        gflat = sv.pop("gflat", None)  # cleared ahead of time by the caller (criterion forward), or now
        if gflat is None:
            gflat = ops.grad_arena(self.arena, self.arena.numel, dev)
        else:
            ops.join(10)
        Gd = self.spans(gflat)
        x, shared = sv["x"], sv["shared"]
        # Buffers first (main stream), then the chain of input gradients on the main stream while the weight
        # gradients and bias column sums run on side streams (ops.side): they are off the critical path.
        dpre2 = torch.empty(G, T, do, device=dev, dtype=BF16)
        du = torch.empty(G, T, d, device=dev, dtype=BF16)
        dh1 = torch.empty(G, T, dff, device=dev, dtype=BF16)
        dy = torch.empty(G, T, d, device=dev, dtype=torch.float32 if ops.GRAD_F32 else BF16)
        dpre1 = torch.empty(G, T, d, device=dev, dtype=BF16)
        datt = torch.empty(G, T, d, device=dev, dtype=BF16)
        dx = torch.zeros(T, d, device=dev) if shared else torch.empty(G, T, d, device=dev)  # shared: split-K target
        # out = LN2(pre2)
        K.rownorm_bwd(dout.reshape(G * T, do), sv["pre2"], sv["mean2"], sv["rstd2"], G * T, do, P32["ln2w"], P32["ln2b"],
                      dx_bf16=dpre2, dgamma=Gd["ln2w"], dbeta=Gd["ln2b"], rows_per_group=T)
        # pre2 = u . Wl^T + bl
        with ops.side(0):
            ops.linear_dw(dpre2, sv["u"], Gd["Wl"], T, do, d, groups=G)
        with ops.side(1):
            K.colsum(dpre2, Gd["bl"], T, do, do, T * do, G)
        ops.linear_dx(dpre2, P16["Wl"], T, do, d, groups=G, out_bf16=du)
        # u = y + h1 . W2^T + b2
        with ops.side(0):
            ops.linear_dw(du, sv["h1"], Gd["W2"], T, d, dff, groups=G)
        with ops.side(1):
            K.colsum(du, Gd["b2"], T, d, d, T * d, G)
        ops.linear_dx(du, P16["W2"], T, d, dff, groups=G, out_bf16=dh1, gate=sv["h1"],
                      gate_scale=1.0 / (1.0 - sv["drop"]))
        # h1 = dropout(relu(y . W1^T + b1))
        with ops.side(0):
            ops.linear_dw(dh1, sv["y"], Gd["W1"], T, dff, d, groups=G)
        with ops.side(1):
            K.colsum(dh1, Gd["b1"], T, dff, dff, T * dff, G)
        if ops.GRAD_F32:
            ops.linear_dx(dh1, P16["W1"], T, dff, d, groups=G, out_f32=dy, res=du)
        else:
            ops.linear_dx(dh1, P16["W1"], T, dff, d, groups=G, out_bf16=dy, res=du)
        if early_stage is not None and ops.GRAD_READY:
            i0, i1 = self.first["W1"], self.first["b2"] + G
            off0 = self.arena.offsets[i0]
            off1 = self.arena.offsets[i1] if i1 < len(self.arena.params) else self.arena.numel
            views = self.arena.grad_views(gflat)
            # AFTER the W1 input-gradient product above: the optimizer step this announces rewrites W1 / W2 in BOTH
            # precisions (fp32 and the persistent bf16 shadow, CpcAdamSpan.p16), so every reader of their bf16 copies
            # has to be queued first; side(0) waits for the main stream up to here, for dW2 / dW1 (its own work) and
            # for the bias column sums (stream 1)
            with ops.side(0):
                ops.wait_side(1)
                ops.grad_ready(early_stage + ".ffn", gflat[off0:off1], self.arena.params[i0:i1], views[i0:i1])
            sv["announced"] = (i0, i1, off0, off1)
        # y = LN1(pre1)
        K.rownorm_bwd(dy.view(G * T, d), sv["pre1"], sv["mean1"], sv["rstd1"], G * T, d, P32["ln1w"], P32["ln1b"],
                      dx_bf16=dpre1, dgamma=Gd["ln1w"], dbeta=Gd["ln1b"], rows_per_group=T)
        # pre1 = x + att . Wo^T
        with ops.side(0):
            ops.linear_dw(dpre1, sv["att"], Gd["Wo"], T, d, d, groups=G)
        ops.linear_dx(dpre1, P16["Wo"], T, d, d, groups=G, out_bf16=datt)
        dqkv = self._attn_bwd(sv["sa"], P32.get("krel"), Gd.get("krel"), datt, G, T, sv["rng"], sv["drop"])
        # qkv = x . Wqkv^T
        with ops.side(0):
            ops.linear_dw(dqkv, x, Gd["Wqkv"], T, 3 * d, d, groups=G, x_shared=shared)
        if shared:
            # dx = sum_g (dqkv_g . Wqkv_g + dpre1_g): one GEMM whose reduction runs over (group, 3d), then a group sum
            A = K.View(dqkv, 3 * d, T, G, 3 * d, T * 3 * d, major=0, gmul=0)
            B = K.View(P16["Wqkv"], d, 3 * d, G, d, 3 * d * d, major=1, gmul=0)
            K.gemm(A, B, T, d, G * 3 * d, k_per_slab=3 * d, k_splits=ops.auto_splits(T, d, G * 3 * d, 1, 3 * d),
                   out_f32=dx, accumulate=True)
            K.groupsum(dpre1, T * d, G, out_f32=dx, accumulate=True)
        else:
            ops.linear_dx(dqkv, P16["Wqkv"], T, 3 * d, d, groups=G, out_f32=dx, res=dpre1)
        if defer:  # the caller announces from `with ops.announce()` (ops.module_join: no join inside a step)
            ops.module_join(sv, dout, dpre2, du, dh1, dy, dpre1, datt, dqkv)
        else:
            ops.join()
        return dx, gflat


class _TransformerFn(torch.autograd.Function):
    """One TransformerLayer (G = 1) as used by buildTransformerAR."""

    @staticmethod
    def forward(ctx, x, stack, training, *params):
        N, S_in, d = x.shape
        xb = K.cast_bf16(x.detach().contiguous().float()).view(N * S_in, d)
        with ops.saving(any(ctx.needs_input_grad)):
            out, out32, saved = stack.forward(xb, N * S_in, True, training, want_f32=True, seq_len=S_in)
        ctx.stack, ctx.saved, ctx.shape = stack, saved, (N, S_in, d)
        return out32.view(N, S_in, stack.dout)

    @staticmethod
    def backward(ctx, dout):
        st = ctx.stack
        N, S_in, d = ctx.shape
        dx, gflat = st.backward(ctx.saved, dout.contiguous().float().view(1, N * S_in, st.dout))
        views = st.arena.grad_views(gflat)
        ops.grad_ready("ar", gflat, st.arena.params, views)
        return (dx.view(N, S_in, d), None, None, *views)


class MultiHeadStack(TransformerStack):
    """Engine of MultiClassifierTransformerHead: the attention half of one TransformerLayer (G = 1), then
        h1 = dropout(relu(y . W1^T + b1))                                   [T, dff]
        u_k = y + h1 . W2[k]^T + b2[k]      k = 0..K-1   (one grouped GEMM, h1 and y shared)   [K, T, d]
        out = LN(u . Wl^T + bl)             (one GEMM / one LayerNorm over the K*T rows)       [K, T, dout]
    which is exactly the [K, T, dout] layout the criterion's logits GEMM consumes."""

    def __init__(self, layer):
        super().__init__([layer])
        self.K = layer.nclassifiers
        self.n_out = self.K

    def shapes(self):
        sh = super().shapes()
        sh["W2"] = (1, self.K * self.d, self.dff)
        sh["b2"] = (1, self.K * self.d)
        return sh

    def forward(self, x, T, shared, training, want_f32=False, seq_len=None):
        d, dout, dff, S, Kc = self.d, self.dout, self.dff, self.S, self.K
        if x.dim() != 2:
            raise ValueError("MultiHeadStack takes one [T, d] input")
        dev = x.device
        P32 = self.spans(self.arena.ensure())
        P16 = self.spans(self.arena.bf16())
        drop = self.p_drop if training else 0.0
        rng = ops.DropoutState.next(dev) if drop > 0 else None
        qkv = torch.empty(1, T, 3 * d, device=dev, dtype=BF16)
        ops.linear_fwd(x, P16["Wqkv"], T, 3 * d, d, groups=1, x_shared=True, out_bf16=qkv)
        att, sa = self._attn_fwd(qkv, P32.get("krel"), 1, T, seq_len, rng, drop)
        pre1 = torch.empty(1, T, d, device=dev, dtype=BF16)
        y = torch.empty(T, d, device=dev, dtype=BF16)
        mean1 = torch.empty(T, device=dev)
        rstd1 = torch.empty(T, device=dev)
        if ops.can_fuse_norm(d, T):
            ops.linear_fwd(att, P16["Wo"], T, d, d, groups=1, out_bf16=pre1, res=(x, d, 0),
                           force_bn=ops.NORM_FORCE_BN, norm=dict(gamma=P32["ln1w"], beta=P32["ln1b"], out_bf16=y, mean=mean1, rstd=rstd1))
        else:
            ops.linear_fwd(att, P16["Wo"], T, d, d, groups=1, out_bf16=pre1, res=(x, d, 0))
            K.rownorm_fwd(pre1, T, d, P32["ln1w"], P32["ln1b"], out_bf16=y, mean=mean1, rstd=rstd1, rows_per_group=T)
        h1 = torch.empty(T, dff, device=dev, dtype=BF16)
        ops.linear_fwd(y, P16["W1"].view(dff, d), T, dff, d, out_bf16=h1, bias=P32["b1"], relu=True, rng=rng,
                       rng_site=2, drop_p=drop)
        u = torch.empty(Kc, T, d, device=dev, dtype=BF16)
        ops.linear_fwd(h1, P16["W2"].view(Kc, d, dff), T, d, dff, groups=Kc, x_shared=True, out_bf16=u,
                       bias=P32["b2"], bias_gs=d, res=(y, d, 0))
        pre2 = torch.empty(Kc * T, dout, device=dev, dtype=BF16)
        out = torch.empty(Kc, T, dout, device=dev, dtype=BF16)
        out32 = torch.empty(Kc, T, dout, device=dev) if want_f32 else None
        mean2 = torch.empty(Kc * T, device=dev)
        rstd2 = torch.empty(Kc * T, device=dev)
        if ops.can_fuse_norm(dout, Kc * T):
            ops.linear_fwd(u.view(Kc * T, d), P16["Wl"].view(dout, d), Kc * T, dout, d, out_bf16=pre2, bias=P32["bl"],
                           force_bn=ops.NORM_FORCE_BN, norm=dict(gamma=P32["ln2w"], beta=P32["ln2b"], out_bf16=out.view(Kc * T, dout),
                                     out_f32=None if out32 is None else out32.view(Kc * T, dout), mean=mean2, rstd=rstd2))
        else:
            ops.linear_fwd(u.view(Kc * T, d), P16["Wl"].view(dout, d), Kc * T, dout, d, out_bf16=pre2, bias=P32["bl"])
            K.rownorm_fwd(pre2, Kc * T, dout, P32["ln2w"], P32["ln2b"], out_bf16=out, out_f32=out32, mean=mean2,
                          rstd=rstd2, rows_per_group=Kc * T)
        saved = dict(x=x, T=T, P16=P16, P32=P32, rng=rng, drop=drop, sa=sa, att=att, pre1=pre1,
                     y=y, mean1=mean1, rstd1=rstd1, h1=h1, u=u, pre2=pre2, mean2=mean2, rstd2=rstd2)
        return out, out32, saved

    def backward(self, sv, dout):
        """dout bf16 or fp32 [K, T, dout] -> (dx fp32 [T, d], flat fp32 grad arena)."""
        d, do, dff, S, Kc = self.d, self.dout, self.dff, self.S, self.K
        T, P16, P32 = sv["T"], sv["P16"], sv["P32"]
        dev = dout.device
        gflat = sv.pop("gflat", None)  # cleared ahead of time by the caller (criterion forward), or now
        if gflat is None:
            gflat = ops.grad_arena(self.arena, self.arena.numel, dev)
        else:
            ops.join(10)
        Gd = self.spans(gflat)
        x = sv["x"]
        dpre2 = torch.empty(Kc * T, do, device=dev, dtype=BF16)
        du = torch.empty(Kc, T, d, device=dev, dtype=BF16)
        dh1 = torch.empty(T, dff, device=dev, dtype=BF16)
        dy = torch.zeros(T, d, device=dev)
        dpre1 = torch.empty(1, T, d, device=dev, dtype=BF16)
        datt = torch.empty(1, T, d, device=dev, dtype=BF16)
        dx = torch.empty(T, d, device=dev)
        # out = LN2(pre2)
        K.rownorm_bwd(dout.reshape(Kc * T, do), sv["pre2"], sv["mean2"], sv["rstd2"], Kc * T, do, P32["ln2w"], P32["ln2b"],
                      dx_bf16=dpre2, dgamma=Gd["ln2w"], dbeta=Gd["ln2b"], rows_per_group=Kc * T)
        # pre2 = u . Wl^T + bl   (Wl shared by the K slices)
        with ops.side(0):
            ops.linear_dw(dpre2, sv["u"].view(Kc * T, d), Gd["Wl"].view(do, d), Kc * T, do, d)
        with ops.side(1):
            K.colsum(dpre2, Gd["bl"], Kc * T, do, do)
        ops.linear_dx(dpre2, P16["Wl"].view(do, d), Kc * T, do, d, out_bf16=du.view(Kc * T, d))
        # u_k = y + h1 . W2[k]^T + b2[k]
        with ops.side(0):
            ops.linear_dw(du, sv["h1"], Gd["W2"].view(Kc, d, dff), T, d, dff, groups=Kc, x_shared=True)
        with ops.side(1):
            K.colsum(du, Gd["b2"], T, d, d, T * d, Kc)
        # dh1 = gate o sum_k du_k . W2[k]: one GEMM whose reduction runs over (slice, feature)
        A = K.View(du, d, T, Kc, d, T * d, major=0, gmul=0)
        B = K.View(P16["W2"], dff, d, Kc, dff, d * dff, major=1, gmul=0)
        K.gemm(A, B, T, dff, Kc * d, k_per_slab=d, out_bf16=dh1, gate=sv["h1"], gate_scale=1.0 / (1.0 - sv["drop"]))
        # h1 = dropout(relu(y . W1^T + b1))
        with ops.side(0):
            ops.linear_dw(dh1, sv["y"], Gd["W1"].view(dff, d), T, dff, d)
        with ops.side(1):
            K.colsum(dh1, Gd["b1"], T, dff, dff)
        # dy = dh1 . W1 + sum_k du_k
        K.groupsum(du, T * d, Kc, out_f32=dy, accumulate=True)
        ops.linear_dx(dh1, P16["W1"].view(dff, d), T, dff, d, out_f32=dy, accumulate=True)
        # y = LN1(pre1)
        K.rownorm_bwd(dy, sv["pre1"], sv["mean1"], sv["rstd1"], T, d, P32["ln1w"], P32["ln1b"], dx_bf16=dpre1,
                      dgamma=Gd["ln1w"], dbeta=Gd["ln1b"], rows_per_group=T)
        # pre1 = x + att . Wo^T
        with ops.side(0):
            ops.linear_dw(dpre1, sv["att"], Gd["Wo"], T, d, d, groups=1)
        ops.linear_dx(dpre1, P16["Wo"], T, d, d, groups=1, out_bf16=datt)
        dqkv = self._attn_bwd(sv["sa"], P32.get("krel"), Gd.get("krel"), datt, 1, T, sv["rng"], sv["drop"])
        # qkv = x . Wqkv^T
        with ops.side(0):
            ops.linear_dw(dqkv, x, Gd["Wqkv"], T, 3 * d, d, groups=1, x_shared=True)
        ops.linear_dx(dqkv, P16["Wqkv"], T, 3 * d, d, groups=1, out_f32=dx.view(1, T, d), res=dpre1)
        ops.join()
        return dx, gflat


class _MultiHeadFn(torch.autograd.Function):
    """MultiClassifierTransformerHead as a module (x [N, S, d] fp32 -> [K, N*S, dout] fp32)."""

    @staticmethod
    def forward(ctx, x, stack, training, *params):
        N, S_in, d = x.shape
        xb = K.cast_bf16(x.detach().contiguous().float()).view(N * S_in, d)
        with ops.saving(any(ctx.needs_input_grad)):
            out, out32, saved = stack.forward(xb, N * S_in, True, training, want_f32=True, seq_len=S_in)
        ctx.stack, ctx.saved, ctx.shape = stack, saved, (N, S_in, d)
        return out32

    @staticmethod
    def backward(ctx, dout):
        st = ctx.stack
        N, S_in, d = ctx.shape
        dx, gflat = st.backward(ctx.saved, dout.contiguous().float())
        views = st.arena.grad_views(gflat)
        ops.grad_ready("criterion", gflat, st.arena.params, views)
        return (dx.view(N, S_in, d), None, None, *views)
