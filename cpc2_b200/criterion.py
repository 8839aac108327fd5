"""Drop-in for the reference's cpc/criterion/criterion.py on the training hot path:
PredictionNetwork and CPCUnsupersivedCriterion (criterion.py:97-173, 193-363) with the same constructor
signatures, attribute names, public methods and state_dict keys.

What changes is the formulation, not the result.  The reference gathers [B, NEG+1, W, D] candidates per
prediction step and reduces a broadcast product of that size (~15 GB of HBM traffic per step at the default
config).  Here the K heads run as grouped tcgen05 GEMMs, the logits of ALL (token, frame) pairs are one
batched GEMM pred_k . z^T (tensor cores, ~46 MB, L2 resident) and a gather kernel picks the NEG+1 scalars
each row needs -- same sampled indices (the two torch.randint draws are issued in the reference's order on
the same generator), same cross entropy, same first-max argmax.
"""
import math

import torch
import torch.nn as nn

from . import kernels as K
from . import ops
from .ops import BF16
from .transformers import TransformerStack, buildMultHeadTransformerAR, buildTransformerAR


class EqualizedLinear(nn.Module):
    """cpc/criterion/custom_layers.py:44-76,134-151: nn.Linear with N(0, 1) weights, zero bias and the He constant
    sqrt(2 / fan_in) applied to the OUTPUT at run time (parameter container; `weight` is that python float)."""

    def __init__(self, nChannelsPrevious, nChannels, bias=True, equalized=True, lrMul=1.0, initBiasToZero=True):
        super().__init__()
        self.module = nn.Linear(nChannelsPrevious, nChannels, bias=bias)
        self.equalized = equalized
        if initBiasToZero and self.module.bias is not None:
            self.module.bias.data.fill_(0)
        self.weight = 1.0
        if equalized:
            self.module.weight.data.normal_(0, 1)
            self.weight = math.sqrt(2.0 / nChannelsPrevious) * lrMul


class FFNetwork(nn.Module):
    """cpc/criterion/criterion.py:10-19 (the criterion's own FFNetwork: two EqualizedLinear layers)."""

    def __init__(self, din, dout, dff, dropout):
        super().__init__()
        if dropout:
            raise NotImplementedError("cpc2_b200: criterion.FFNetwork is only built with dropout 0 by the reference")
        self.lin1 = EqualizedLinear(din, dff, bias=True, equalized=True)
        self.lin2 = EqualizedLinear(dff, dout, bias=True, equalized=True)
        self.relu = nn.ReLU()
        self.drop = nn.Dropout(dropout)


class EqualizedConv1d(nn.Module):
    """cpc/criterion/custom_layers.py:44-105: nn.Conv1d with N(0, 1) weights, zero bias, output scaled by
    sqrt(2 / (Cin * kernelSize)) at run time (parameter container)."""

    def __init__(self, nChannelsPrevious, nChannels, kernelSize, padding=0, bias=True, stride=1, equalized=True,
                 lrMul=1.0, initBiasToZero=True):
        super().__init__()
        self.module = nn.Conv1d(nChannelsPrevious, nChannels, kernelSize, padding=padding, bias=bias, stride=stride)
        self.equalized = equalized
        if initBiasToZero and self.module.bias is not None:
            self.module.bias.data.fill_(0)
        self.weight = 1.0
        if equalized:
            self.module.weight.data.normal_(0, 1)
            self.weight = math.sqrt(2.0 / (nChannelsPrevious * kernelSize)) * lrMul


class ShiftedConv(nn.Module):
    """cpc/criterion/criterion.py:23-41: causal convolution head (kernelSize - 1 zero frames in front of the
    sequence, then EqualizedConv1d); parameter container, run by ConvHeads."""

    def __init__(self, dimOutputAR, dimOutputEncoder, kernelSize):
        super().__init__()
        self.module = EqualizedConv1d(dimOutputAR, dimOutputEncoder, kernelSize, equalized=True, padding=0)
        self.kernelSize = kernelSize


class PredictionNetwork(nn.Module):
    """cpc/criterion/criterion.py:97-173.  rnnMode 'transformer' (default) or the bias-free Linear that any
    other string selects in the reference ('linear', None, False); 'ffd', 'conv4/8/12', 'RNN', 'LSTM' heads."""

    def __init__(self, nPredicts, dimOutputAR, dimOutputEncoder, rnnMode=None, dropout=False, sizeInputSeq=116,
                 transformer_pruning=0):
        super().__init__()
        self.predictors = nn.ModuleList()
        self.RESIDUAL_STD = 0.01
        self.dimOutputAR = dimOutputAR
        self.dimOutputEncoder = dimOutputEncoder
        self.sizeInputSeq = sizeInputSeq
        self.dropout = nn.Dropout(p=0.5) if dropout else None  # criterion.py:113; fused path: cpc_dropout
        self.transformer = rnnMode == 'transformer'
        self.ffd = rnnMode == 'ffd'
        self.recurrent = rnnMode if rnnMode in ("RNN", "LSTM") else None
        self.conv = {"conv4": 4, "conv8": 8, "conv12": 12}.get(rnnMode)
        for i in range(nPredicts):
            if self.recurrent == "RNN":  # criterion.py:116-119 (NOT batch_first: see RecurrentHeads)
                self.predictors.append(nn.RNN(dimOutputAR, dimOutputEncoder))
            elif self.recurrent == "LSTM":  # criterion.py:120-124
                self.predictors.append(nn.LSTM(dimOutputAR, dimOutputEncoder, batch_first=True))
            elif self.conv:  # criterion.py:129-140
                self.predictors.append(ShiftedConv(dimOutputAR, dimOutputEncoder, self.conv))
            elif self.transformer:
                self.predictors.append(buildTransformerAR(dimOutputEncoder, dimOutputAR, nLayers=1,
                                                          sizeSeq=sizeInputSeq, abspos=False))
            elif self.ffd:
                self.predictors.append(FFNetwork(dimOutputAR, dimOutputEncoder, dimOutputEncoder, 0))
            else:
                self.predictors.append(nn.Linear(dimOutputAR, dimOutputEncoder, bias=False))
                if dimOutputEncoder > dimOutputAR:
                    residual = dimOutputEncoder - dimOutputAR
                    self.predictors[-1].weight.data.copy_(torch.cat([
                        torch.randn(dimOutputAR, dimOutputAR),
                        self.RESIDUAL_STD * torch.randn(residual, dimOutputAR)], dim=0))
        self._engine = None

    def engine(self):
        if self._engine is None:
            if self.transformer:
                eng = TransformerStack([p[0] for p in self.predictors])
            elif self.ffd:
                eng = FfdHeads([p for p in self.predictors])
            elif self.recurrent:
                eng = RecurrentHeads([p for p in self.predictors], self.recurrent, self.sizeInputSeq)
            elif self.conv:
                eng = ConvHeads([p for p in self.predictors], self.sizeInputSeq)
            else:
                eng = LinearHeads([p for p in self.predictors])
            object.__setattr__(self, "_engine", eng)
        return self._engine

    def prefetch(self):
        """bf16 shadow of all heads' weights (16.6 M at the default config) on a side stream, early in the step."""
        self.engine().arena.prefetch(5)

    def forward(self, c, candidates):
        """Reference signature (criterion.py:152-173): list of [B, n_cand, W, D] candidates -> list of
        [B, n_cand, W] scores.  Used by getCosineDistances; the training path goes through the fused
        criterion function instead."""
        assert (len(candidates) == len(self.predictors))
        eng = self.engine()
        preds = _HeadsFn.apply(c, eng, self.training, *eng.arena.params)  # [K, B, W, D] fp32
        drop = self.dropout if self.dropout is not None else (lambda t: t)
        return [(drop(preds[k]).unsqueeze(1) * candidates[k]).mean(dim=3) for k in range(len(self.predictors))]


class MultiHeadPredictionNetwork(nn.Module):
    """cpc/criterion/criterion.py:44-94 (--multihead_rnn): ONE transformer block whose feed-forward network emits all
    nPredicts predictions (MultiClassifierTransformerHead); prediction[:, :, k] scores the candidates of step k."""

    def __init__(self, nPredicts, dimOutputAR, dimOutputEncoder, rnnMode='transformer_multi', dropout=False,
                 sizeInputSeq=116, transformer_pruning=0):
        super().__init__()
        self.dimOutputAR = dimOutputAR
        self.dropout = nn.Dropout(p=0.5) if dropout else None  # criterion.py:59; fused path: cpc_dropout
        self.nPredicts = nPredicts
        if rnnMode == 'transformer':
            self.predictor = buildMultHeadTransformerAR(dimOutputEncoder, dimOutputAR, nLayers=1, sizeSeq=sizeInputSeq,
                                                        abspos=False, nHeads=nPredicts)
        elif rnnMode == 'transformer_adaptive_span':
            raise NotImplementedError("cpc2_b200: transformer_adaptive_span is broken in the reference itself "
                                      "(missing import, criterion.py:70-76)")
        else:
            raise ValueError(f"unknown mode {rnnMode}")

    def engine(self):
        return self.predictor[0].stack()

    def prefetch(self):
        self.engine().arena.prefetch(5)

    def forward(self, c, candidates):
        assert (len(candidates) == self.nPredicts)
        eng = self.engine()
        preds = _HeadsFn.apply(c, eng, self.training, *eng.arena.params)  # [K, B, W, D] fp32
        drop = self.dropout if self.dropout is not None else (lambda t: t)
        return [(drop(preds[k]).unsqueeze(1) * candidates[k]).mean(dim=3) for k in range(self.nPredicts)]


class LinearHeads:
    """K bias-free Linear heads on the same input: one grouped GEMM."""

    def __init__(self, linears):
        self.G = self.n_out = len(linears)
        self.dout, self.d = linears[0].weight.shape
        if self.d % 8 or self.dout % 8:
            raise NotImplementedError("LinearHeads: dimensions must be multiples of 8")
        self.arena = ops.Arena([l.weight for l in linears])

    def forward(self, x, T, shared, training, want_f32=False, seq_len=None):
        G, d, do = self.G, self.d, self.dout
        w16 = self.arena.span(self.arena.bf16(), 0, G, (G, do, d))
        out = torch.empty(G, T, do, device=x.device, dtype=BF16)
        out32 = torch.empty(G, T, do, device=x.device) if want_f32 else None
        ops.linear_fwd(x, w16, T, do, d, groups=G, x_shared=True, out_bf16=out, out_f32=out32)
        return out, out32, dict(x=x, w16=w16, T=T)

    def backward(self, sv, dout):
        G, d, do, T = self.G, self.d, self.dout, sv["T"]
        dev = dout.device
        if dout.dtype != BF16:
            dout = K.cast_bf16(dout.contiguous().float())
        dout = dout.view(G, T, do)
        gflat = ops.grad_arena(self.arena, self.arena.numel, dev)
        dw = self.arena.span(gflat, 0, G, (G, do, d))
        ops.linear_dw(dout, sv["x"], dw, T, do, d, groups=G, x_shared=True)
        dx = torch.zeros(T, d, device=dev)
        A = K.View(dout, do, T, G, do, T * do, major=0, gmul=0)
        B = K.View(sv["w16"], d, do, G, d, do * d, major=1, gmul=0)
        K.gemm(A, B, T, d, G * do, k_per_slab=do, k_splits=ops.auto_splits(T, d, G * do, 1, do), out_f32=dx,
               accumulate=True)
        return dx, gflat


class FfdHeads:
    """K feed-forward heads (rnnMode 'ffd': lin2(relu(lin1(c))) with EqualizedLinear layers, criterion.py:10-19,
    125-128) on the same input: two grouped GEMMs forward (the He constants ride on alpha and a pre-scaled bias),
    the relu mask applied in the epilogue of the dh GEMM backward."""

    def __init__(self, nets):
        self.G = self.n_out = len(nets)
        self.dff, self.d = nets[0].lin1.module.weight.shape
        self.dout = nets[0].lin2.module.weight.shape[0]
        self.c1, self.c2 = float(nets[0].lin1.weight), float(nets[0].lin2.weight)
        if self.d % 8 or self.dout % 8 or self.dff % 8:
            raise NotImplementedError("FfdHeads: dimensions must be multiples of 8")
        G = self.G
        self.arena = ops.Arena([n.lin1.module.weight for n in nets] + [n.lin1.module.bias for n in nets] +
                               [n.lin2.module.weight for n in nets] + [n.lin2.module.bias for n in nets])
        self.first = dict(W1=0, b1=G, W2=2 * G, b2=3 * G)

    def spans(self, flat):
        G, d, dff, do = self.G, self.d, self.dff, self.dout
        sh = dict(W1=(G, dff, d), b1=(G, dff), W2=(G, do, dff), b2=(G, do))
        return {k: self.arena.span(flat, self.first[k], G, v) for k, v in sh.items()}

    def forward(self, x, T, shared, training, want_f32=False, seq_len=None):
        G, d, dff, do = self.G, self.d, self.dff, self.dout
        P32 = self.spans(self.arena.ensure())
        P16 = self.spans(self.arena.bf16())
        b1s, b2s = P32["b1"] * self.c1, P32["b2"] * self.c2  # (x W^T + b) * c = c * x W^T + (c b)
        h = torch.empty(G, T, dff, device=x.device, dtype=BF16)
        ops.linear_fwd(x, P16["W1"], T, dff, d, groups=G, x_shared=True, out_bf16=h, bias=b1s, bias_gs=dff, relu=True,
                       alpha=self.c1)
        out = torch.empty(G, T, do, device=x.device, dtype=BF16)
        out32 = torch.empty(G, T, do, device=x.device) if want_f32 else None
        ops.linear_fwd(h, P16["W2"], T, do, dff, groups=G, out_bf16=out, out_f32=out32, bias=b2s, bias_gs=do,
                       alpha=self.c2)
        return out, out32, dict(x=x, h=h, P16=P16, T=T)

    def backward(self, sv, dout):
        G, d, dff, do, T = self.G, self.d, self.dff, self.dout, sv["T"]
        dev = dout.device
        if dout.dtype != BF16:
            dout = K.cast_bf16(dout.contiguous().float())
        dout = dout.view(G, T, do)
        gflat = ops.grad_arena(self.arena, self.arena.numel, dev)
        Gd = self.spans(gflat)
        dh = torch.empty(G, T, dff, device=dev, dtype=BF16)
        dx = torch.zeros(T, d, device=dev)
        with ops.side(0):  # dW2 = c2 * dout^T . h
            K.gemm(K.mnmajor(dout), K.mnmajor(sv["h"]), do, dff, T, groups=G, k_splits=ops.auto_splits(do, dff, T, G),
                   out_f32=Gd["W2"], accumulate=True, alpha=self.c2)
        with ops.side(1):
            K.colsum(dout, Gd["b2"], T, do, do, T * do, G)
            Gd["b2"].mul_(self.c2)
        # dh = relu' o (c2 * dout . W2)
        ops.linear_dx(dout, sv["P16"]["W2"], T, do, dff, groups=G, out_bf16=dh, gate=sv["h"], gate_scale=1.0,
                      alpha=self.c2)
        with ops.side(0):  # dW1 = c1 * dh^T . x
            K.gemm(K.mnmajor(dh), K.mnmajor(sv["x"], gmul=0), dff, d, T, groups=G,
                   k_splits=ops.auto_splits(dff, d, T, G), out_f32=Gd["W1"], accumulate=True, alpha=self.c1)
        with ops.side(1):
            K.colsum(dh, Gd["b1"], T, dff, dff, T * dff, G)
            Gd["b1"].mul_(self.c1)
        A = K.View(dh, dff, T, G, dff, T * dff, major=0, gmul=0)
        B = K.View(sv["P16"]["W1"], d, dff, G, d, dff * d, major=1, gmul=0)
        K.gemm(A, B, T, d, G * dff, k_per_slab=dff, k_splits=ops.auto_splits(T, d, G * dff, 1, dff), out_f32=dx,
               accumulate=True, alpha=self.c1)
        ops.join()
        return dx, gflat


class ConvHeads:
    """K ShiftedConv heads (criterion.py:23-41, rnnMode conv4 / conv8 / conv12) on the same input.  The ks frames
    ending at t are unfolded ONCE into X [T, ks*H] (zeros before the start of each window) and the K causal
    convolutions become one grouped GEMM with reduction length ks*H; (conv + b) * c = c * conv + (c b)."""

    def __init__(self, convs, sizeSeq):
        self.G = self.n_out = len(convs)
        self.S = sizeSeq
        self.dout, self.d, self.ks = convs[0].module.module.weight.shape
        self.c = float(convs[0].module.weight)
        if self.d % 8 or self.dout % 8:
            raise NotImplementedError("ConvHeads: dimensions must be multiples of 8")
        self.arena = ops.Arena([c.module.module.weight for c in convs] + [c.module.module.bias for c in convs])

    def spans(self, flat):
        G = self.G
        return dict(W=self.arena.span(flat, 0, G, (G, self.dout, self.d, self.ks)),
                    b=self.arena.span(flat, G, G, (G, self.dout)))

    def forward(self, x, T, shared, training, want_f32=False, seq_len=None):
        G, d, do, ks, S = self.G, self.d, self.dout, self.ks, int(seq_len or self.S)
        if T % S != 0:
            raise ValueError(f"token count {T} is not a multiple of the sequence length {S}")
        P32 = self.spans(self.arena.ensure())
        P16 = self.spans(self.arena.bf16())
        # X[b, t, j, :] = c[b, t - (ks - 1) + j, :]
        xu = torch.nn.functional.pad(x.view(T // S, S, d), (0, 0, ks - 1, 0)).unfold(1, ks, 1)  # [Bt, S, d, ks]
        xu = xu.permute(0, 1, 3, 2).contiguous().view(T, ks * d)
        w2 = P16["W"].permute(0, 1, 3, 2).contiguous().view(G, do, ks * d)  # W2[g, co, j*d + ci] = W[g, co, ci, j]
        out = torch.empty(G, T, do, device=x.device, dtype=BF16)
        out32 = torch.empty(G, T, do, device=x.device) if want_f32 else None
        ops.linear_fwd(xu, w2, T, do, ks * d, groups=G, x_shared=True, out_bf16=out, out_f32=out32,
                       bias=P32["b"] * self.c, bias_gs=do, alpha=self.c)
        return out, out32, dict(xu=xu, w2=w2, T=T, S=S)

    def backward(self, sv, dout):
        G, d, do, ks, S, T = self.G, self.d, self.dout, self.ks, sv["S"], sv["T"]
        dev = dout.device
        if dout.dtype != BF16:
            dout = K.cast_bf16(dout.contiguous().float())
        dout = dout.view(G, T, do)
        gflat = ops.grad_arena(self.arena, self.arena.numel, dev)
        Gd = self.spans(gflat)
        dw2 = torch.zeros(G, do, ks * d, device=dev)
        dxu = torch.zeros(T, ks * d, device=dev)
        with ops.side(0):  # dW = c * dout^T . X, back to the Conv1d layout [co, ci, j]
            K.gemm(K.mnmajor(dout), K.mnmajor(sv["xu"], gmul=0), do, ks * d, T, groups=G,
                   k_splits=ops.auto_splits(do, ks * d, T, G), out_f32=dw2, accumulate=True, alpha=self.c)
            Gd["W"].copy_(dw2.view(G, do, ks, d).permute(0, 1, 3, 2))
        with ops.side(1):
            K.colsum(dout, Gd["b"], T, do, do, T * do, G)
            Gd["b"].mul_(self.c)
        # dX = c * sum_g dout_g . W2_g, then folded back onto the frames it was unfolded from
        A = K.View(dout, do, T, G, do, T * do, major=0, gmul=0)
        B = K.View(sv["w2"], ks * d, do, G, ks * d, do * ks * d, major=1, gmul=0)
        K.gemm(A, B, T, ks * d, G * do, k_per_slab=do, k_splits=ops.auto_splits(T, ks * d, G * do, 1, do), out_f32=dxu,
               accumulate=True, alpha=self.c)
        Bt = T // S
        dpad = torch.zeros(Bt, S + ks - 1, d, device=dev)
        dxu = dxu.view(Bt, S, ks, d)
        for j in range(ks):
            dpad[:, j:j + S] += dxu[:, :, j]
        dx = dpad[:, ks - 1:].reshape(T, d)
        ops.join()
        return dx, gflat


class RecurrentHeads:
    """K recurrent heads (criterion.py:116-124): nn.LSTM(dimAR, dimEncoder, batch_first=True), or nn.RNN(dimAR,
    dimEncoder) -- which the reference builds WITHOUT batch_first and feeds c [B, W, H] as is, so its recurrence
    runs over the B windows of the batch with the W frames as independent rows; reproduced as such.  Each head is
    the cluster recurrence kernel of the AR network (cpc_lstm_fwd / bwd through model._RecurrentFn); the heads'
    gradients are gathered into one arena so the criterion announces one bucket like every other head type."""

    def __init__(self, rnns, cell, sizeSeq):
        self.G = self.n_out = len(rnns)
        self.cell, self.S = cell, sizeSeq
        self.d, self.dout = rnns[0].input_size, rnns[0].hidden_size
        self.per = [[r.weight_ih_l0, r.weight_hh_l0, r.bias_ih_l0, r.bias_hh_l0] for r in rnns]
        self.arena = ops.Arena([p for ps in self.per for p in ps])

    def forward(self, x, T, shared, training, want_f32=False, seq_len=None):
        from .model import _RecurrentFn
        G, d, do, S = self.G, self.d, self.dout, int(seq_len or self.S)
        if T % S != 0:
            raise ValueError(f"token count {T} is not a multiple of the sequence length {S}")
        self.arena.ensure()
        xin = x.float().view(T // S, S, d)
        if self.cell == "RNN":
            xin = xin.transpose(0, 1).contiguous()  # rows = frames, time = the windows of the batch
        with torch.enable_grad():
            xin = xin.detach().requires_grad_(True)
            outs = []
            for ps in self.per:
                o = _RecurrentFn.apply(xin, None, None, 1, None, self.cell, *ps)[0]
                outs.append(o.transpose(0, 1) if self.cell == "RNN" else o)
            out32 = torch.stack(outs).reshape(G, T, do)
        out = K.cast_bf16(out32.detach())
        return out.view(G, T, do), (out32.detach() if want_f32 else None), dict(xin=xin, out32=out32, T=T)

    def backward(self, sv, dout):
        G, d, do, T = self.G, self.d, self.dout, sv["T"]
        gflat = ops.grad_arena(self.arena, self.arena.numel, dout.device)
        views = self.arena.grad_views(gflat)
        hooks = list(ops.GRAD_READY)
        ops.GRAD_READY[:] = []  # the per-head arenas of _RecurrentFn are not buckets of their own
        try:
            grads = torch.autograd.grad([sv["out32"]], [sv["xin"]] + list(self.arena.params),
                                        [dout.float().reshape(G, T, do)])
        finally:
            ops.GRAD_READY[:] = hooks
        dx = grads[0].transpose(0, 1) if self.cell == "RNN" else grads[0]
        for v, g in zip(views, grads[1:]):
            v.copy_(g)
        return dx.reshape(T, d).contiguous(), gflat


class _HeadsFn(torch.autograd.Function):
    """All K heads on c [B, W, H] -> predictions [K, B, W, D] fp32 (public PredictionNetwork.forward path)."""

    @staticmethod
    def forward(ctx, c, eng, training, *params):
        B, W, H = c.shape
        xb = K.cast_bf16(c.detach().contiguous().float()).view(B * W, H)
        with ops.saving(any(ctx.needs_input_grad)):
            out, out32, saved = eng.forward(xb, B * W, True, training, want_f32=True, seq_len=W)
        ctx.eng, ctx.saved, ctx.shape = eng, saved, (B, W, H)
        return out32.view(eng.n_out, B, W, eng.dout)

    @staticmethod
    def backward(ctx, dout):
        eng = ctx.eng
        B, W, H = ctx.shape
        dx, gflat = eng.backward(ctx.saved, dout.contiguous().float().view(eng.n_out, B * W, eng.dout))
        views = eng.arena.grad_views(gflat)
        ops.grad_ready("criterion", gflat, eng.arena.params, views)
        return (dx.view(B, W, H), None, None, *views)


class BaseCriterion(nn.Module):
    """cpc/criterion/criterion.py:176-182."""

    def warmUp(self):
        return False

    def update(self):
        return


class CPCUnsupersivedCriterion(BaseCriterion):
    """cpc/criterion/criterion.py:193-363."""

    def __init__(self, nPredicts, dimOutputAR, dimOutputEncoder, negativeSamplingExt, mode=None, rnnMode=False,
                 dropout=False, nSpeakers=0, sizeInputSeq=116, multihead_rnn=False, transformer_pruning=0,
                 n_skipped=0, growth_rate=None, inflection_point_x=None):
        super().__init__()
        if multihead_rnn:
            self.wPrediction = MultiHeadPredictionNetwork(nPredicts, dimOutputAR, dimOutputEncoder, rnnMode=rnnMode,
                                                          dropout=dropout, sizeInputSeq=sizeInputSeq - nPredicts,
                                                          transformer_pruning=transformer_pruning)
        else:
            self.wPrediction = PredictionNetwork(nPredicts, dimOutputAR, dimOutputEncoder, rnnMode=rnnMode,
                                                 dropout=dropout, sizeInputSeq=sizeInputSeq - nPredicts)
        self.nSkipped = n_skipped
        self.nPredicts = nPredicts
        self.negativeSamplingExt = negativeSamplingExt
        self.lossCriterion = nn.CrossEntropyLoss(reduction='none')
        self.growth_rate = growth_rate
        self.inflection_point_x = inflection_point_x
        self.weighting_function = lambda x: 0.00001 + 1 / (1 + torch.exp(-growth_rate * (x - inflection_point_x)))
        if mode not in [None, "reverse"]:
            raise ValueError("Invalid mode")
        self.mode = mode

    # -------------------------------------------------------------------------------------- sampling
    def sampleIndices(self, batchSize, nNegativeExt, windowSize, device, out=None):
        """The two RNG draws of sampleClean (criterion.py:247-255) in the reference's order, then the
        integer index arithmetic (:256-266) on the GPU.  Returns extIdx int64 [B*NEG*W], flat (b, n, t).
        `out`: three preallocated int64 [n] buffers (batchIdx, seqIdx, extIdx)."""
        n = self.negativeSamplingExt * windowSize * batchSize
        if out is None:
            batchIdx = torch.randint(low=0, high=batchSize, size=(n,), device=device)
            seqIdx = torch.randint(low=1, high=nNegativeExt, size=(n,), device=device)
            return K.neg_ext_index(batchIdx, seqIdx, batchSize, self.negativeSamplingExt, windowSize, nNegativeExt)
        batchIdx = torch.randint(low=0, high=batchSize, size=(n,), out=out[0])
        seqIdx = torch.randint(low=1, high=nNegativeExt, size=(n,), out=out[1])
        return K.neg_ext_index(batchIdx, seqIdx, batchSize, self.negativeSamplingExt, windowSize, nNegativeExt, out=out[2])

    def prefetchIndices(self, batchSize, seqSize, device):
        """Draw this step's negatives NOW, on a side stream (train.forwardBackward calls it at the top of the step):
        the sampler depends on nothing the model computes, and its three small kernels otherwise sit between the AR
        network and the prediction heads on the critical path.  Same two randint draws in the same order on the same
        generator as forward() would issue (no other consumer of the generator runs in between: dropout uses its own
        counter RNG), so seeded runs sample the same negatives either way.  Buffers are allocated on the current
        stream (ops.side rule); forward() joins the side stream before it reads them."""
        windowSize = seqSize - self.nPredicts
        if windowSize <= 0 or "sampleIndices" in self.__dict__:  # (an instance-level sampler override: leave it alone)
            return
        n = self.negativeSamplingExt * windowSize * batchSize
        bufs = [torch.empty(n, dtype=torch.int64, device=device) for _ in range(3)]
        with ops.side(8):
            ext = self.sampleIndices(batchSize, seqSize, windowSize, device, out=bufs)
        self.__dict__["_pref_ext"] = ((batchSize, seqSize, windowSize, torch.device(device)), ext, bufs)

    def sampleClean(self, encodedData, windowSize):
        """Reference-compatible materialised sampling (criterion.py:237-286): list of K tensors
        [B, NEG+1, W, D] and the zero labels.  Kept for API parity; the training path never calls it."""
        batchSize, nNegativeExt, dimEncoded = encodedData.size()
        extIdx = self.sampleIndices(batchSize, nNegativeExt, windowSize, encodedData.device)
        negExt = encodedData.contiguous().view(-1, dimEncoded)[extIdx].view(
            batchSize, self.negativeSamplingExt, windowSize, dimEncoded)
        labelLoss = torch.zeros((batchSize * windowSize), dtype=torch.long, device=encodedData.device)
        outputs = []
        for k in range(1, self.nPredicts + 1):
            posSeq = encodedData[:, k:k + windowSize].reshape(batchSize, 1, windowSize, dimEncoded)
            outputs.append(torch.cat((posSeq, negExt), dim=1))
        return outputs, labelLoss

    def getPrediction(self, cFeature, encodedData, label):
        """criterion.py:291-302 (materialised; API parity)."""
        if self.mode == "reverse":
            encodedData = torch.flip(encodedData, [1])
            cFeature = torch.flip(cFeature, [1])
        windowSize = cFeature.size(1) - self.nPredicts
        cFeature = cFeature[:, :windowSize]
        sampledData, labelLoss = self.sampleClean(encodedData, windowSize)
        return self.wPrediction(cFeature, sampledData), labelLoss

    def getCosineDistances(self, cFeature, encodedData):
        """criterion.py:304-327: scores of the positives only."""
        if self.mode == "reverse":
            encodedData = torch.flip(encodedData, [1])
            cFeature = torch.flip(cFeature, [1])
        batchSize, seqSize, _ = cFeature.size()
        dimEncoded = encodedData.size(2)
        windowSize = seqSize - self.nPredicts
        cFeature = cFeature[:, :windowSize]
        out = [encodedData[:, k:k + windowSize].reshape(batchSize, 1, windowSize, dimEncoded)
               for k in range(1, self.nPredicts + 1)]
        return self.wPrediction(cFeature, out)

    # -------------------------------------------------------------------------------------- forward
    def forward(self, cFeature, encodedData, label, signal_quality=None):
        """criterion.py:329-363 -> (losses [1, K - n_skipped], acc [1, K - n_skipped])."""
        ops.require_cuda(cFeature, encodedData)
        ops.reject_replica(self)
        if self.mode == "reverse":
            encodedData = torch.flip(encodedData, [1])
            cFeature = torch.flip(cFeature, [1])
        batchSize, seqSize, _ = cFeature.size()
        windowSize = seqSize - self.nPredicts
        if batchSize * encodedData.size(1) > ops.NCE_BWD_MAX_BS and torch.is_grad_enabled():
            raise ValueError(f"cpc2_b200: batch x frames = {batchSize * encodedData.size(1)} exceeds {ops.NCE_BWD_MAX_BS} "
                             "(cpc_nce_bwd keeps one row of d(logits) per warp in shared memory; e.g. batchSizeGPU <= 50 at "
                             "sizeWindow 20480): lower batchSizeGPU, the all-pairs logits also grow as (batch x frames)^2")
        pref = self.__dict__.pop("_pref_ext", None)
        if pref is not None:
            ops.join(8)
        if pref is not None and pref[0] == (batchSize, encodedData.size(1), windowSize, encodedData.device):
            extIdx = pref[1]
        else:
            extIdx = self.sampleIndices(batchSize, encodedData.size(1), windowSize, encodedData.device)
        weights = None
        if signal_quality is not None:
            q = self.weighting_function(signal_quality.mean(dim=1))
            weights = q.unsqueeze(1).repeat(1, windowSize).contiguous().view(-1).float()
        eng = self.wPrediction.engine()
        # --dropout: nn.Dropout(0.5) on every prediction in train mode (criterion.py:168-169)
        drop_pred = self.wPrediction.dropout.p if (self.wPrediction.dropout is not None and self.training) else 0.0
        losses, acc = _CriterionFn.apply(cFeature, encodedData, extIdx, weights, eng, self.nPredicts,
                                         self.negativeSamplingExt, (self.training, drop_pred, torch.is_grad_enabled()),
                                         *eng.arena.params)
        return losses[:, self.nSkipped:], acc[:, self.nSkipped:]


class _CriterionFn(torch.autograd.Function):
    """heads -> all-pairs logits GEMM -> gather + CE + argmax, and the matching backward."""

    @staticmethod
    def forward(ctx, c, z, ext, weights, eng, Kp, NEG, mode, *params):
        training, drop_pred, will_backward = mode
        B, S, H = c.shape
        D = z.shape[2]
        W = S - Kp
        T, BS = B * W, B * S
        dev = c.device
        if BS % 8 != 0:
            raise ValueError("B*S must be a multiple of 8")
        cw = K.cast_bf16(c.detach()[:, :W].contiguous().float()).view(T, H)
        # z is first read by the logits product, behind the heads: its cast runs on a side stream beside them
        zf = z.detach().contiguous().float()
        zb = torch.empty(BS, D, device=dev, dtype=BF16)
        with ops.side(11):
            K.cast_bf16(zf, zb)
        with ops.saving(will_backward and any(ctx.needs_input_grad)):
            preds, _, saved = eng.forward(cw, T, True, training, seq_len=W)
        ops.join(11)
        if isinstance(eng, TransformerStack) and will_backward and any(ctx.needs_input_grad):
            # the heads' gradient arena (66 MB at the default config) is cleared NOW on an idle side stream: the same
            # memset at the top of the backward sat on the critical path for ~13 us
            # (allocated on THIS stream, cleared on side stream 10; the backward joins 10 before the first write)
            saved["gflat"] = ops.grad_arena(eng.arena, eng.arena.numel, dev, zero_on=10)
        ctx.pred_drop = None
        if drop_pred > 0:
            ctx.pred_drop = (ops.DropoutState.next(dev), drop_pred)
            K.dropout_(preds, ctx.pred_drop[0], 7, drop_pred)
        logits = torch.empty(Kp, T, BS, device=dev)
        # (algorithmic work of the logits: 2 T (NEG+1) D per prediction step -- SURVEY 8d -- the all-pairs product
        # executes 2 T (B S) D)
        algo = 2.0 * T * (NEG + 1) * D * Kp
        ops.linear_fwd(preds, zb, T, BS, D, groups=Kp, out_f32=logits, alpha=1.0 / D, algo_flops=algo)
        lse = torch.empty(Kp, T, device=dev)
        # mean loss / accuracy per prediction step (criterion.py:350-357) accumulate pre-scaled by 1/T: no divisions after
        losses, acc = torch.zeros(2, 1, Kp, device=dev).unbind(0)  # (one memset)
        K.nce_fwd(logits, ext, weights, lse, losses, acc, Kp, B, W, S, NEG, BS, out_scale=1.0 / T)
        ctx.eng, ctx.saved = eng, saved
        ctx.t = (preds, zb, logits, lse, ext, weights)
        ctx.dims = (B, S, H, D, W, Kp, NEG)
        ctx.algo = algo
        ctx.mark_non_differentiable(acc)
        return losses, acc

    @staticmethod
    def backward(ctx, dloss, dacc):
        B, S, H, D, W, Kp, NEG = ctx.dims
        T, BS = B * W, B * S
        preds, zb, logits, lse, ext, weights = ctx.t
        eng = ctx.eng
        dev = dloss.device
        gscale = dloss.reshape(Kp).contiguous().float()
        # d loss / d logits as a hi + lo pair of bf16 matrices (see nce.cu: the rounding error of the hi part is
        # coherent across rows); each backward GEMM runs once per part, the second pass accumulating.
        dlog = torch.empty(2, Kp, T, BS, device=dev, dtype=BF16)
        K.nce_bwd(logits, ext, weights, lse, gscale, dlog[0], Kp, B, W, S, NEG, BS, dlogits_lo=dlog[1])
        # dpred_k = dlog_k . z / D  (main stream: the heads' backward waits for it)
        dpred = torch.empty(Kp, T, D, device=dev)
        acc0 = torch.zeros(BS * D + B * S * H, device=dev)  # dz | dc, one memset
        dz, dc = acc0[:BS * D].view(BS, D), acc0[BS * D:].view(B, S, H)
        # dz = sum_k dlog_k^T . pred_k / D   (reduction over (k, token)); nothing downstream of the criterion
        # needs it before the encoder backward, so it runs on a side stream next to the heads' backward
        with ops.side(2):
            Bv = K.View(preds, D, T, Kp, D, T * D, major=1, gmul=0)
            for part in range(2):
                A = K.View(dlog[part], BS, T, Kp, BS, T * BS, major=1, gmul=0)
                K.gemm(A, Bv, BS, D, Kp * T, k_per_slab=T, k_splits=ops.auto_splits(BS, D, Kp * T, 1, T), out_f32=dz,
                       accumulate=True, alpha=1.0 / D, algo_flops=ctx.algo / 2)
            # dz is complete here, long before the heads' backward below: an encoder backward that runs on its
            # own side stream (the future half of CPCModel.forwardPastFuture) may start from this event
            ops._READY.clear()
            ops.set_ready(dz, ops.mark())
        # one launch over both parts: the reduction runs over [hi | lo] (two slabs of BS, Kp planes apart in dlog)
        # against the same z (slab step 0) -- one prologue / epilogue on the dependent chain instead of two
        K.gemm(K.View(dlog, BS, T, 2 * Kp, BS, T * BS, major=0, gmul=1, smul=Kp), K.View(zb, D, BS, 1, D, major=1, smul=0),
               T, D, 2 * BS, groups=Kp, k_per_slab=BS, out_f32=dpred, alpha=1.0 / D, algo_flops=ctx.algo)
        if ctx.pred_drop is not None:
            K.dropout_(dpred, ctx.pred_drop[0], 7, ctx.pred_drop[1])
        if type(eng) is TransformerStack:
            dcw, gflat = eng.backward(ctx.saved, dpred, early_stage="criterion" if ops.early_ffn() else None, defer=True)
        else:
            dcw, gflat = eng.backward(ctx.saved, dpred)
        dc[:, :W] = dcw.view(B, W, H)
        ops.join(2)  # dz: whoever consumes it on this stream (the future half of a pair starts from its own event)
        # inside a step the chain does not wait for the heads' weight gradients: the announcement does (ops.announce)
        ops.module_join(ctx.saved, ctx.t, dlog, dpred, acc0, dcw)
        views = eng.arena.grad_views(gflat)
        ann = ctx.saved.pop("announced", None)
        with ops.announce():
            if ann is None:
                ops.grad_ready("criterion", gflat, eng.arena.params, views)
            else:  # the feed-forward weights went out early (criterion.ffn): what is left, as two contiguous spans
                i0, i1, off0, off1 = ann
                ops.grad_ready("criterion.front", gflat[:off0], eng.arena.params[:i0], views[:i0])
                ops.grad_ready("criterion.tail", gflat[off1:], eng.arena.params[i1:], views[i1:])
        return (dc, dz.view(B, S, D), None, None, None, None, None, None, *views)
