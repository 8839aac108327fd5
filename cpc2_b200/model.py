"""Drop-in for the reference's cpc/model.py on the training hot path: ChannelNorm, CPCEncoder, CPCAR,
CPCModel -- same constructor signatures, attribute names and state_dict keys (cpc/model.py:27-108,
158-207, 279-390), compute in libcpc_b200.so (sm_100a).  There is no eager / CPU fallback.

Data layout in HBM: activations are bf16 CHANNELS-LAST [N, L + halo, C]; each layer's output buffer carries
the zero halo rows the next strided conv needs, so the im2col matrix of a conv is an overlapping-row TMA
view of its input and `encodedData = gEncoder(x).permute(0, 2, 1)` (model.py:382) is a free view.
"""
import torch
import torch.nn as nn

from . import kernels as K
from . import ops
from .ops import BF16

# (kernel, stride, padding) of conv0..4, cpc/model.py:85-94
GEOMETRY = ((10, 5, 3), (8, 4, 2), (4, 2, 1), (4, 2, 1), (4, 2, 1))


class ChannelNorm(nn.Module):
    """cpc/model.py:27-60.  Parameter container + standalone forward on [N, C, L] fp32."""

    def __init__(self, numFeatures, epsilon=1e-05, affine=True):
        super().__init__()
        if affine:
            self.weight = nn.parameter.Parameter(torch.Tensor(1, numFeatures, 1))
            self.bias = nn.parameter.Parameter(torch.Tensor(1, numFeatures, 1))
        else:
            self.weight = None
            self.bias = None
        self.epsilon = epsilon
        self.p = 0
        self.affine = affine
        self.reset_parameters()

    def reset_parameters(self):
        if self.affine:
            torch.nn.init.ones_(self.weight)
            torch.nn.init.zeros_(self.bias)

    def forward(self, x):
        ops.require_cuda(x)
        return _ChannelNormFn.apply(x, self.weight, self.bias, self.epsilon)


class _ChannelNormFn(torch.autograd.Function):
    """Standalone ChannelNorm (no ReLU) on NCL input; the encoder uses the fused kernels instead."""

    @staticmethod
    def forward(ctx, x, weight, bias, eps):
        N, C, L = x.shape
        xcl = K.cast_bf16(x.permute(0, 2, 1).contiguous())
        out = torch.empty(N, L, C, device=x.device, dtype=torch.float32)
        mean = torch.empty(N * L, device=x.device)
        rstd = torch.empty(N * L, device=x.device)
        g = weight.detach().reshape(C).contiguous() if weight is not None else None
        b = bias.detach().reshape(C).contiguous() if bias is not None else None
        K.rownorm_fwd(xcl, N * L, C, g, b, out_f32=out, mean=mean, rstd=rstd, eps=eps, unbiased=True)
        ctx.save_for_backward(xcl, mean, rstd, g, b)
        ctx.shape = (N, C, L)
        return out.permute(0, 2, 1)

    @staticmethod
    def backward(ctx, dout):
        xcl, mean, rstd, g, b = ctx.saved_tensors
        N, C, L = ctx.shape
        d = dout.permute(0, 2, 1).contiguous().float()
        dx = torch.empty(N, L, C, device=d.device)
        dg = torch.zeros(C, device=d.device) if g is not None else None
        db = torch.zeros(C, device=d.device) if g is not None else None
        K.rownorm_bwd(d, xcl, mean, rstd, N * L, C, g, b, dx_f32=dx, dgamma=dg, dbeta=db, unbiased=True)
        return (dx.permute(0, 2, 1), None if dg is None else dg.view(1, C, 1),
                None if db is None else db.view(1, C, 1), None)


class IDModule(nn.Module):
    """cpc/model.py:18-24 (normMode "ID")."""

    def __init__(self, *args, **kwargs):
        super().__init__()

    def forward(self, x):
        return x


class CPCEncoder(nn.Module):
    """cpc/model.py:63-108: five strided Conv1d, each followed by a norm layer and ReLU.  normMode "layerNorm" (the
    default, ChannelNorm) runs the fused kernels (_EncoderFn); "batchNorm" / "instanceNorm" / "ID" keep the reference's
    nn.BatchNorm1d / nn.InstanceNorm1d(affine=True) / IDModule as PARAMETER CONTAINERS (same state_dict keys, running
    estimates included) and run the column-statistics kernels (_EncoderColFn, csrc/colnorm.cu)."""

    def __init__(self, sizeHidden=512, normMode="layerNorm"):
        super().__init__()
        validModes = ["batchNorm", "instanceNorm", "ID", "layerNorm"]
        if normMode not in validModes:
            raise ValueError(f"Norm mode must be in {validModes}")
        if normMode == "instanceNorm":
            def normLayer(x): return nn.InstanceNorm1d(x, affine=True)
        elif normMode == "ID":
            normLayer = IDModule
        elif normMode == "layerNorm":
            normLayer = ChannelNorm
        else:
            normLayer = nn.BatchNorm1d
        self.normMode = normMode
        self.dimEncoded = sizeHidden
        self.conv0 = nn.Conv1d(1, sizeHidden, 10, stride=5, padding=3)
        self.batchNorm0 = normLayer(sizeHidden)
        self.conv1 = nn.Conv1d(sizeHidden, sizeHidden, 8, stride=4, padding=2)
        self.batchNorm1 = normLayer(sizeHidden)
        self.conv2 = nn.Conv1d(sizeHidden, sizeHidden, 4, stride=2, padding=1)
        self.batchNorm2 = normLayer(sizeHidden)
        self.conv3 = nn.Conv1d(sizeHidden, sizeHidden, 4, stride=2, padding=1)
        self.batchNorm3 = normLayer(sizeHidden)
        self.conv4 = nn.Conv1d(sizeHidden, sizeHidden, 4, stride=2, padding=1)
        self.batchNorm4 = normLayer(sizeHidden)
        self.DOWNSAMPLING = 160

    def getDimOutput(self):
        return self.conv4.out_channels

    def _params(self):
        ps = []
        for i in range(5):
            conv, norm = getattr(self, f"conv{i}"), getattr(self, f"batchNorm{i}")
            ps += [conv.weight, conv.bias]
            if self.normMode != "ID":
                ps += [norm.weight, norm.bias]
        return ps

    def _conv_weights(self):
        return [getattr(self, f"conv{i}").weight for i in range(1, 5)]

    def prefetch(self):
        """Pack conv1..4 weights into their GEMM operand layouts on a side stream (train.prefetchWeights calls
        this at the top of the step, so the packing overlaps conv0 instead of preceding every conv)."""
        ws = self._conv_weights()
        bufs = _alloc_packed(ws[0].shape[0], ws[0].device)
        with ops.side(3):
            _pack_conv_weights([w.detach() for w in ws], bufs)
        self._pref = ops.Prefetched(ws, bufs, 3)

    def forward(self, x):
        """x [N, 1, W] fp32 -> [N, C, S] fp32 (a view of channels-last memory), S = W / 160 for the training windows;
        other lengths (the last chunk of a file at inference) run forward-only."""
        ops.require_cuda(x)
        ops.reject_replica(self)
        if x.shape[-1] % 160 != 0 and torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()):
            raise NotImplementedError("cpc2_b200: training windows must be a multiple of 160 samples (the reference's "
                                      "sizeWindow, default 20480); any length runs forward-only under torch.no_grad()")
        pref, self._pref = getattr(self, "_pref", None), None
        if self.normMode != "layerNorm":
            return _EncoderColFn.apply(x, _ColPlan(self, pref), *self._params())
        return _EncoderFn.apply(x, _EncPlan(pref=pref), *self._params())


def _alloc_packed(C, dev):
    out = [None]
    for l in range(1, 5):
        k, s, _ = GEOMETRY[l]
        out.append((torch.empty(C, k * C, device=dev, dtype=BF16), torch.empty(s * C, 2 * C, device=dev, dtype=BF16)))
    return out


def _pack_conv_weights(ws, bufs):
    C = ws[0].shape[0]
    for l in range(1, 5):
        k, s, _ = GEOMETRY[l]
        K.conv_weight_pack(ws[l - 1], bufs[l][0], bufs[l][1], C, C, k, s)


def _geometry(N, W):
    """Frames after each of the five strided convolutions (floor arithmetic of nn.Conv1d, cpc/model.py:85-94): W / 5,
    W / 20 ... W / 160 when W is a multiple of 160, and whatever the floor gives for the last chunk of a file
    (feature_loader.py:323-367 feeds chunks of any length)."""
    L, n = [], W
    for k, s, pd in GEOMETRY:
        n = (n + 2 * pd - k) // s + 1
        L.append(n)
    pad_next = [GEOMETRY[i + 1][2] for i in range(4)] + [0]
    return L, pad_next


def _grad_sizes(C):
    sizes = []
    for l in range(5):
        kl = GEOMETRY[l][0]
        sizes += [C * (1 if l == 0 else C) * kl, C, C, C]
    return sizes


def _alloc_fwd(N, W, C, dev):
    """Every buffer the encoder forward writes (activations carry the next conv's zero halo rows)."""
    L, pad_next = _geometry(N, W)
    b = dict(acts=[], ys=[None], means=[], rstds=[])
    for l in range(5):
        b["acts"].append(torch.empty(N, L[l] + 2 * pad_next[l], C, device=dev, dtype=BF16))
        b["means"].append(torch.empty(N * L[l], device=dev))
        b["rstds"].append(torch.empty(N * L[l], device=dev))
        if l:
            b["ys"].append(torch.empty(N * L[l], C, device=dev, dtype=BF16))
    b["z32"] = torch.empty(N, L[4], C, device=dev, dtype=torch.float32)
    return b


def _alloc_bwd(N, W, C, dev, zero_on=None):
    """Workspace of the encoder backward: per conv layer dy (one zero row either side), the fp32 weight gradient
    in GEMM layout (zeroed: split-K accumulates into it) and the padded input gradient."""
    L, _ = _geometry(N, W)
    w = [None]
    n_dwf = sum(C * GEOMETRY[l][0] * C for l in range(1, 5))  # one memset for the four layers
    if zero_on is None:
        dwf = torch.zeros(n_dwf, device=dev)
    else:  # allocated on the current stream, cleared on a side stream (the caller joins it before the first write)
        dwf = torch.empty(n_dwf, device=dev)
        with ops.side(zero_on):
            dwf.zero_()
    off = 0
    for l in range(1, 5):
        k, s, pd = GEOMETRY[l]
        w.append(dict(dy=torch.empty(N, L[l] + 2, C, device=dev, dtype=BF16),
                      dwf=dwf[off:off + C * k * C].view(C, k * C),
                      dxp=torch.empty(N, L[l - 1] + 2 * pd, C, device=dev, dtype=BF16)))
        off += C * k * C
    return w


class _EncPlan:
    """How one _EncoderFn node runs.  Default: everything on the current stream, buffers allocated on the fly.
    As one half of a past/future pair (CPCModel.forwardPastFuture): `role` 'future' runs on side stream 6 from
    `fwd_after` on, with buffers the driver allocated on the main stream BEFORE that event; both halves reduce
    their parameter gradients into `pair.gflat` and the half whose backward is issued last returns them."""

    def __init__(self, pref=None, packed=None, role=None, pair=None, fwd_after=None, fwd_bufs=None, bwd_bufs=None):
        self.pref, self.packed, self.role, self.pair = pref, packed, role, pair
        self.fwd_after, self.fwd_bufs, self.bwd_bufs = fwd_after, fwd_bufs, bwd_bufs


class _PairCtx:
    def __init__(self, gflat):
        self.gflat = gflat
        self.pending = 2
        self.first_done = None  # event: the half whose backward was issued first has finished


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


class _EncoderFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, plan, *params):
        N, one, W = x.shape
        if one != 1:
            raise ValueError("CPCEncoder expects [N, 1, W]")
        C = params[0].shape[0]
        dev = x.device
        p = [t.detach() for t in params]
        L, pad_next = _geometry(N, W)
        if L[4] < 1:
            raise ValueError(f"CPCEncoder: {W} samples are shorter than one output frame")
        future = plan.role == "future"
        # the waveform is read IN PLACE when it is an fp32 view with unit sample stride (the past / future windows of a
        # [B,2,1,W] batch, cpc/train.py:95-96, are rows 2W apart): no copy in front of conv0, forward or backward
        xd = x.detach()
        if ops.INPLACE_X and xd.dtype == torch.float32 and xd.stride(2) == 1 and (N == 1 or xd.stride(0) >= W):
            x2, xs = xd, (xd.stride(0) if N > 1 else W)
        else:
            x2, xs = xd.reshape(N, W).contiguous().float(), W
        bufs = plan.fwd_bufs if plan.fwd_bufs is not None else _alloc_fwd(N, W, C, dev)
        acts, ys, means, rstds, z32 = bufs["acts"], bufs["ys"], bufs["means"], bufs["rstds"], bufs["z32"]
        # conv1..4 weights in GEMM layout: handed in by the pair driver, prefetched on side stream 3, or packed now
        packed, sync = plan.packed, None
        if packed is None and plan.pref is not None:
            packed = plan.pref.peek([params[4 * l] for l in range(1, 5)])
        if packed is None:
            packed = _alloc_packed(C, dev)
            with ops.side(3):
                _pack_conv_weights([p[4 * l] for l in range(1, 5)], packed)
        with (ops.side(6, after=plan.fwd_after) if future else _null()):
            # layer 0: fused conv + ChannelNorm + ReLU straight from the waveform (halo rows zeroed by the kernel)
            K.conv0_fwd(x2, p[0], p[1], p[2], p[3], acts[0], means[0], rstds[0], N, W, C, pad_next[0], x_stride=xs)
            if future:
                ops.wait_side(3)  # the past half (main stream) consumes the pending join
            else:
                ops.join(3)
            for l in range(1, 5):
                k, s, pd = GEOMETRY[l]
                w, b, g, be = p[4 * l:4 * l + 4]
                wf, wd = packed[l]
                Lin, Lout = L[l - 1], L[l]
                Lp = Lin + 2 * pd
                y, a = ys[l], acts[l]
                A = K.View(acts[l - 1], k * C, Lout, N, s * C, Lp * C, major=0, gmul=1)
                hn = pad_next[l]  # halo rows of `a` are zeroed by the kernel that writes it
                if ops.can_fuse_norm(C, Lout, N):
                    # conv + bias -> y (kept for the backward), ChannelNorm + ReLU -> a, all in the GEMM epilogue
                    K.gemm(A, K.kmajor(wf), Lout, C, k * C, groups=N, out_bf16=(y, C, Lout * C), bias=b,
                           relu=True, force_bn=ops.NORM_FORCE_BN, norm=dict(gamma=g.reshape(C), beta=be.reshape(C), out_bf16=(a, C, (Lout + 2 * hn) * C),
                                                offset_bf16=hn * C, out_f32=(z32, C, Lout * C) if l == 4 else None,
                                                mean=means[l], rstd=rstds[l], gs_stat=Lout, unbiased=True, halo=hn))
                    continue
                K.gemm(A, K.kmajor(wf), Lout, C, k * C, groups=N, out_bf16=(y, C, Lout * C), bias=b)
                K.rownorm_fwd(y, N * Lout, C, g.reshape(C), be.reshape(C), out_bf16=a, out_f32=z32 if l == 4 else None,
                              mean=means[l], rstd=rstds[l], out_map=K.rowmap(Lout, Lout + 2 * hn, hn), unbiased=True,
                              relu=True)
        ctx.geom = (N, W, C, L, pad_next)
        ctx.x2, ctx.xs, ctx.x_version = x2, xs, x._version
        ctx.acts, ctx.ys, ctx.means, ctx.rstds, ctx.packed = acts, ys, means, rstds, packed
        ctx.p = p
        ctx.params = list(params)
        ctx.plan = plan
        return z32.permute(0, 2, 1)

    @staticmethod
    def backward(ctx, dz):
        N, W, C, L, pad_next = ctx.geom
        p, plan = ctx.p, ctx.plan
        dev = dz.device
        pair, future = plan.pair, plan.role == "future"
        # one flat fp32 buffer for every parameter gradient of the encoder (a single memset, and a single
        # all-reduce bucket for the data-parallel wrapper); per-parameter gradients are views of it.  Every kernel
        # below ADDS into it atomically, so the two halves of a pair share one arena.
        sizes = _grad_sizes(C)
        gflat = pair.gflat if pair is not None else ops.grad_arena("encoder", sum(sizes), dev)
        gv, off = [], 0
        for n_ in sizes:
            gv.append(gflat[off:off + n_])
            off += n_
        ws = plan.bwd_bufs if plan.bwd_bufs is not None else _alloc_bwd(N, W, C, dev)
        wg = 7 if future else 0  # weight-gradient side stream (forked from the stream this half runs on)
        after = ops.pop_ready(dz) if future else None
        last = True
        grads = []
        for l in range(5):
            k = GEOMETRY[l][0]
            dw, dcb, dg, db = gv[4 * l:4 * l + 4]
            grads += [dw.view(C, 1 if l == 0 else C, k), dcb, dg.view(1, C, 1), db.view(1, C, 1)]
        # Early announcements (ops.EARLY_ENCODER): the gradients of layer l are complete once THIS half's weight gradient
        # of layer l is (the other half of a pair has finished or is waited for) -- they are announced from the
        # weight-gradient side stream, so that their all-reduce and optimizer step run next to the rest of the
        # encoder's backward instead of trailing the step.  Mode 1: conv1-4 as one bucket behind conv1's weight
        # gradient; mode 2: one bucket per layer (conv4 first; only conv1's 2 MB and conv0's 13 KB are reduced late);
        # mode 3: mode 1 with conv0's bucket announced FIRST, straight behind conv0's backward (which ends before conv1's
        # weight gradient does): its reduction and update then hide behind that GEMM instead of trailing conv1-4's.
        early = ops.EARLY_ENCODER if (bool(ops.GRAD_READY) and (pair is None or pair.pending == 1)) else 0
        lay_off = [sum(sizes[:4 * l]) for l in range(6)]
        with (ops.side(6, after=after) if future else _null()):
            ops.wait_side(14)  # the arena / staging memsets issued beside the forward (forwardPastFuture)
            d = dz.permute(0, 2, 1).contiguous().float()  # [N, S, C]; free when it came from our own layout
            dout, dout_map = d, None
            for l in range(4, 0, -1):
                k, s, pd = GEOMETRY[l]
                w, b, g, be = p[4 * l:4 * l + 4]
                wf, wd = ctx.packed[l]
                Lin, Lout = L[l - 1], L[l]
                Lp = Lin + 2 * pd
                dy, dwf, dxp = ws[l]["dy"], ws[l]["dwf"], ws[l]["dxp"]
                dw, dcb, dg, db = gv[4 * l:4 * l + 4]
                K.rownorm_bwd(dout, ctx.ys[l], ctx.means[l], ctx.rstds[l], N * Lout, C, g.reshape(C), be.reshape(C),
                              dx_bf16=dy, dgamma=dg, dbeta=db, dxsum=dcb, dout_map=dout_map,
                              dx_map=K.rowmap(Lout, Lout + 2, 1), unbiased=True, relu=True)
                # wgrad: dWf[co, kk*C+ci] = sum_{n,t} dy[n,t,co] * a_pad[n, t*s+kk, ci]   (side stream: only the
                # optimizer needs it; the input-gradient chain continues on this stream).  Split-K >= 2 keeps the
                # accumulation atomic (red.global), which is what lets two halves share dw.
                with ops.side(wg):
                    A = K.View(dy, C, Lout + 2, N, C, (Lout + 2) * C, major=1, row_off=1)
                    B = K.View(ctx.acts[l - 1], k * C, Lout, N, s * C, Lp * C, major=1)
                    K.gemm(A, B, C, k * C, N * Lout, k_per_slab=Lout,
                           k_splits=max(2, ops.auto_splits(C, k * C, N * Lout, 1, Lout)), out_f32=dwf, accumulate=True)
                    K.conv_wgrad_unpack(dwf, dw, C, C, k)
                    if early == 2:
                        if pair is not None and pair.first_done is not None:
                            torch.cuda.current_stream().wait_event(pair.first_done)
                        ops.grad_ready(f"encoder.conv{l}", gflat[lay_off[l]:lay_off[l + 1]], ctx.params[4 * l:4 * l + 4],
                                       grads[4 * l:4 * l + 4])
                # dgrad: padded dx rows [t*s, t*s+s) = [dy[t-1], dy[t]] . [W_hi; W_lo]
                A = K.View(dy, 2 * C, Lout + 1, N, C, (Lout + 2) * C, major=0, gmul=1)
                K.gemm(A, K.kmajor(wd), Lout + 1, s * C, 2 * C, groups=N, out_bf16=(dxp, s * C, Lp * C))
                dout, dout_map = dxp, K.rowmap(Lin, Lp, pd)
            n0 = sum(sizes[:4])

            def announce_main():
                with ops.side(wg):  # behind every weight gradient of this half (this stream) and the main chain so far
                    if pair is not None and pair.first_done is not None:
                        torch.cuda.current_stream().wait_event(pair.first_done)
                    ops.grad_ready("encoder.main", gflat[n0:], ctx.params[4:], grads[4:])

            if early == 1:
                announce_main()
            dw0, db0, dg0, dbe0 = gv[0:4]
            K.conv0_bwd(ctx.x2, p[0], p[1], p[2], p[3], ctx.means[0], ctx.rstds[0], dout, dw0, db0, dg0, dbe0, N, W, C,
                        pad_next[0], x_stride=ctx.xs)
            if early == 3:
                if pair is not None and pair.first_done is not None:
                    torch.cuda.current_stream().wait_event(pair.first_done)  # the other half's conv0 gradients
                ops.grad_ready("encoder.conv0", gflat[:n0], ctx.params[:4], grads[:4])
                announce_main()
            ops.join()
            if pair is not None:
                pair.pending -= 1
                last = pair.pending == 0
                if last:
                    if pair.first_done is not None:
                        torch.cuda.current_stream().wait_event(pair.first_done)
                else:
                    pair.first_done = ops.mark()
        if future:
            if last:
                ops.join(6)
        del ws
        if not last:
            return (None, None) + (None,) * 20
        if early == 3:
            pass  # both buckets are announced
        elif early:
            ops.grad_ready("encoder.conv0", gflat[:n0], ctx.params[:4], grads[:4])
        else:
            ops.grad_ready("encoder", gflat, ctx.params, grads)
        return (None, None, *grads)


class _ColPlan:
    """What _EncoderColFn needs from the module besides its parameters: the norm mode and, for nn.BatchNorm1d, whether
    batch statistics are used (training mode or track_running_stats off), the running estimates and the momentum."""

    def __init__(self, enc, pref):
        self.mode, self.pref = enc.normMode, pref
        self.layers = []
        for i in range(5):
            n = getattr(enc, f"batchNorm{i}")
            d = dict(kind="ID", eps=1e-5, momentum=0.0, rm=None, rv=None)
            if enc.normMode == "instanceNorm":
                if n.track_running_stats:
                    raise NotImplementedError("cpc2_b200: InstanceNorm1d with running statistics is never built by the reference")
                d.update(kind="stats", eps=n.eps)
            elif enc.normMode == "batchNorm":
                batch_stats = n.training or n.running_mean is None
                d.update(kind="stats" if batch_stats else "given", eps=n.eps)
                if n.training and n.track_running_stats:
                    # nn.BatchNorm1d.forward: num_batches_tracked += 1; momentum None = cumulative average
                    n.num_batches_tracked.add_(1)
                    if n.momentum is None:
                        raise NotImplementedError("cpc2_b200: BatchNorm1d(momentum=None) is never built by the reference")
                    d.update(rm=n.running_mean, rv=n.running_var, momentum=n.momentum)
                elif not batch_stats:
                    d.update(rm=n.running_mean, rv=n.running_var)
            self.layers.append(d)


class _EncoderColFn(torch.autograd.Function):
    """The encoder under normMode batchNorm / instanceNorm / ID (cpc/model.py:71-82,102-107): conv0 on CUDA cores, conv1-4
    as the same implicit GEMMs as _EncoderFn, statistics down the columns (csrc/colnorm.cu).  Single stream, literal
    concatenated batch (batch statistics couple every row)."""

    @staticmethod
    def forward(ctx, x, plan, *params):
        N, one, W = x.shape
        if one != 1:
            raise ValueError("CPCEncoder expects [N, 1, W]")
        ident = plan.mode == "ID"
        npl = 2 if ident else 4
        C = params[0].shape[0]
        dev = x.device
        p = [t.detach() for t in params]
        L, pad_next = _geometry(N, W)
        if L[4] < 1:
            raise ValueError(f"CPCEncoder: {W} samples are shorter than one output frame")
        x2 = x.detach().reshape(N, W).contiguous().float()
        packed = plan.pref.peek([params[npl * l] for l in range(1, 5)]) if plan.pref is not None else None
        if packed is None:
            packed = _alloc_packed(C, dev)
            with ops.side(3):
                _pack_conv_weights([p[npl * l] for l in range(1, 5)], packed)
        acts, ys, means, rstds, kinds = [], [], [], [], []
        z32 = torch.empty(N, L[4], C, device=dev, dtype=torch.float32)
        for l in range(5):
            k, s, pd = GEOMETRY[l]
            lay = plan.layers[l]
            w, b = p[npl * l], p[npl * l + 1]
            g, be = (None, None) if ident else (p[npl * l + 2].reshape(C), p[npl * l + 3].reshape(C))
            Lout, hn = L[l], pad_next[l]
            y = torch.empty(N * Lout, C, device=dev, dtype=BF16)
            a = torch.empty(N, Lout + 2 * hn, C, device=dev, dtype=BF16) if l < 4 else None  # (conv4's output: fp32 only)
            if l == 0:
                K.conv0_plain_fwd(x2, w, b, y, N, W, C)
                ops.join(3)
            else:
                Lp = L[l - 1] + 2 * pd
                A = K.View(acts[l - 1], k * C, Lout, N, s * C, Lp * C, major=0, gmul=1)
                K.gemm(A, K.kmajor(packed[l][0]), Lout, C, k * C, groups=N, out_bf16=(y, C, Lout * C), bias=b)
            rpg = Lout if plan.mode == "instanceNorm" else 0
            mean = rstd = partial = None
            if lay["kind"] == "stats":
                groups = N if rpg else 1
                mean, rstd = torch.empty(groups * C, device=dev), torch.empty(groups * C, device=dev)
                partial, _ = K.colnorm_scratch(N * Lout, C, rpg, dev)
            elif lay["kind"] == "given":
                mean = lay["rm"].detach().float().contiguous()
                rstd = torch.rsqrt(lay["rv"].detach().float() + lay["eps"]).contiguous()
            K.colnorm_fwd(y, N * Lout, C, lay["kind"], rows_per_group=rpg, gamma=g, beta=be, mean=mean, rstd=rstd,
                          partial=partial, running_mean=lay["rm"] if lay["kind"] == "stats" else None,
                          running_var=lay["rv"] if lay["kind"] == "stats" else None, momentum=lay["momentum"],
                          eps=lay["eps"], out_bf16=a, out_map=K.rowmap(Lout, Lout + 2 * hn, hn) if l < 4 else None,
                          out_f32=z32 if l == 4 else None, relu=True)
            acts.append(a); ys.append(y); means.append(mean); rstds.append(rstd); kinds.append((lay["kind"], rpg))
        ctx.geom = (N, W, C, L, pad_next)
        ctx.x2, ctx.acts, ctx.ys, ctx.means, ctx.rstds, ctx.kinds, ctx.packed = x2, acts, ys, means, rstds, kinds, packed
        ctx.p, ctx.params, ctx.ident = p, list(params), ident
        return z32.permute(0, 2, 1)

    @staticmethod
    def backward(ctx, dz):
        N, W, C, L, pad_next = ctx.geom
        p, ident = ctx.p, ctx.ident
        npl = 2 if ident else 4
        dev = dz.device
        sizes = []
        for l in range(5):
            sizes += [C * (1 if l == 0 else C) * GEOMETRY[l][0], C] + ([] if ident else [C, C])
        gflat = ops.grad_arena("encoder", sum(sizes), dev)
        gv, off = [], 0
        for n_ in sizes:
            gv.append(gflat[off:off + n_])
            off += n_
        grads = []
        for l in range(5):
            k = GEOMETRY[l][0]
            grads += [gv[npl * l].view(C, 1 if l == 0 else C, k), gv[npl * l + 1]]
            if not ident:
                shape = ctx.params[npl * l + 2].shape  # [C] (BatchNorm1d / InstanceNorm1d)
                grads += [gv[npl * l + 2].view(shape), gv[npl * l + 3].view(shape)]
        dout = dz.permute(0, 2, 1).contiguous().float()  # [N, S, C]
        dout_map = None
        for l in range(4, -1, -1):
            k, s, pd = GEOMETRY[l]
            kind, rpg = ctx.kinds[l]
            g, be = (None, None) if ident else (p[npl * l + 2].reshape(C), p[npl * l + 3].reshape(C))
            dw, dcb = gv[npl * l], gv[npl * l + 1]
            dg, db = (None, None) if ident else (gv[npl * l + 2], gv[npl * l + 3])
            Lout = L[l]
            partial, sums = K.colnorm_scratch(N * Lout, C, rpg, dev)
            if l == 0:
                dy0 = torch.empty(N * Lout, C, device=dev, dtype=torch.float32)
                K.colnorm_bwd(dout, ctx.ys[0], N * Lout, C, kind, partial, sums, rows_per_group=rpg, gamma=g, beta=be,
                              mean=ctx.means[0], rstd=ctx.rstds[0], dout_map=dout_map, dx_f32=dy0, dgamma=dg, dbeta=db,
                              dxsum=dcb, relu=True)
                K.conv0_plain_bwd(ctx.x2, dy0, dw, N, W, C)
                break
            Lin = L[l - 1]
            Lp = Lin + 2 * pd
            dy = torch.empty(N, Lout + 2, C, device=dev, dtype=BF16)
            K.colnorm_bwd(dout, ctx.ys[l], N * Lout, C, kind, partial, sums, rows_per_group=rpg, gamma=g, beta=be,
                          mean=ctx.means[l], rstd=ctx.rstds[l], dout_map=dout_map, dx_bf16=dy,
                          dx_map=K.rowmap(Lout, Lout + 2, 1), dgamma=dg, dbeta=db, dxsum=dcb, relu=True)
            # weight and input gradients: the same two GEMMs as _EncoderFn.backward
            dwf = torch.zeros(C, k * C, device=dev)
            A = K.View(dy, C, Lout + 2, N, C, (Lout + 2) * C, major=1, row_off=1)
            B = K.View(ctx.acts[l - 1], k * C, Lout, N, s * C, Lp * C, major=1)
            K.gemm(A, B, C, k * C, N * Lout, k_per_slab=Lout,
                   k_splits=max(2, ops.auto_splits(C, k * C, N * Lout, 1, Lout)), out_f32=dwf, accumulate=True)
            K.conv_wgrad_unpack(dwf, dw, C, C, k)
            dxp = torch.empty(N, Lp, C, device=dev, dtype=BF16)
            A = K.View(dy, 2 * C, Lout + 1, N, C, (Lout + 2) * C, major=0, gmul=1)
            K.gemm(A, K.kmajor(ctx.packed[l][1]), Lout + 1, s * C, 2 * C, groups=N, out_bf16=(dxp, s * C, Lp * C))
            dout, dout_map = dxp, K.rowmap(Lin, Lp, pd)
        ops.grad_ready("encoder", gflat, ctx.params, grads)
        return (None, None, *grads)


class CPCAR(nn.Module):
    """cpc/model.py:158-207.  `baseNet` is an nn.LSTM/GRU/RNN used as the PARAMETER CONTAINER (cuDNN-layout
    weight_ih_l*, weight_hh_l*, bias_ih_l*, bias_hh_l*, so reference checkpoints load unchanged); the
    recurrence itself runs in the cluster-resident LSTM kernels."""

    def __init__(self, dimEncoded, dimOutput, keepHidden, nLevelsGRU, mode="GRU", reverse=False):
        super().__init__()
        self.RESIDUAL_STD = 0.1
        if mode == "LSTM":
            self.baseNet = nn.LSTM(dimEncoded, dimOutput, num_layers=nLevelsGRU, batch_first=True)
        elif mode == "RNN":
            self.baseNet = nn.RNN(dimEncoded, dimOutput, num_layers=nLevelsGRU, batch_first=True)
        else:
            self.baseNet = nn.GRU(dimEncoded, dimOutput, num_layers=nLevelsGRU, batch_first=True)
        self.hidden = None
        self.keepHidden = keepHidden
        self.reverse = reverse

    def getDimOutput(self):
        return self.baseNet.hidden_size

    def _rnn_params(self):
        params = []
        for l in range(self.baseNet.num_layers):
            params += [getattr(self.baseNet, f"{n}_l{l}") for n in ("weight_ih", "weight_hh", "bias_ih", "bias_hh")]
        return params

    def _cell(self):
        return "LSTM" if isinstance(self.baseNet, nn.LSTM) else "GRU" if isinstance(self.baseNet, nn.GRU) else "RNN"

    def prefetch(self):
        """bf16 shadows of the recurrent weights and the summed bias, issued on a side stream at the top of the
        step (train.prefetchWeights) so they overlap the encoder."""
        params = self._rnn_params()
        cell = self._cell()
        bufs = _alloc_rnn_shadows(params, cell)
        with ops.side(4):
            _cast_rnn_weights([q.detach() for q in params], bufs, cell)
        self._pref = ops.Prefetched(params, bufs, 4)

    def forward(self, x):
        ops.require_cuda(x)
        ops.reject_replica(self)
        cell = self._cell()
        if cell == "RNN" and self.baseNet.nonlinearity != "tanh":
            raise NotImplementedError("cpc2_b200: nn.RNN with relu nonlinearity is never built by the reference")
        if self.baseNet.hidden_size not in ops.RECURRENT_H:
            raise NotImplementedError(f"cpc2_b200: the recurrence kernels are built for hidden sizes {ops.RECURRENT_H} "
                                      f"(csrc/lstm.cu), got {self.baseNet.hidden_size}")
        if self.reverse:
            x = torch.flip(x, [1])
        nl = self.baseNet.num_layers
        params = self._rnn_params()
        h0 = c0 = None
        if self.hidden is not None:
            if cell == "LSTM":
                h0, c0 = self.hidden
            else:
                h0 = self.hidden
        pref, self._pref = getattr(self, "_pref", None), None
        out, h_n, c_n = _RecurrentFn.apply(x, h0, c0, nl, pref, cell, *params)
        if self.keepHidden:
            self.hidden = (h_n.detach(), c_n.detach()) if cell == "LSTM" else h_n.detach()
        if self.reverse:
            out = torch.flip(out, [1])
        return out


_GATES = {"LSTM": 4, "GRU": 3, "RNN": 1}


def _alloc_rnn_shadows(params, cell="LSTM"):
    """Per layer: (w_ih bf16 [4H,Kin], w_hh bf16 [4H,H], bias fp32 [4H]) in the kernel's four-slot layout
    (include/cpc_b200.h, CpcLstm), allocated on the current stream.  GRU / RNN leave some slots zero."""
    out = []
    for l in range(len(params) // 4):
        w_ih, w_hh = params[4 * l:4 * l + 2]
        H, dev = w_hh.shape[1], w_ih.device
        mk = torch.empty if cell == "LSTM" else torch.zeros
        out.append((mk(4 * H, w_ih.shape[1], device=dev, dtype=BF16), mk(4 * H, H, device=dev, dtype=BF16),
                    mk(4 * H, device=dev, dtype=torch.float32)))
    return out


def _cast_rnn_weights(p, bufs, cell="LSTM"):
    for l, (w_ih16, w_hh16, bias) in enumerate(bufs):
        w_ih, w_hh, b_ih, b_hh = p[4 * l:4 * l + 4]
        H = w_hh.shape[1]
        w_ih, w_hh = w_ih.contiguous(), w_hh.contiguous()
        if cell == "LSTM":
            K.cast_bf16(w_ih, w_ih16)
            K.cast_bf16(w_hh, w_hh16)
            torch.add(b_ih, b_hh, out=bias)
        elif cell == "GRU":  # slots (r, z, n_x, hn): cpc_b200.h
            K.cast_bf16(w_ih, w_ih16[:3 * H])
            K.cast_bf16(w_hh[:2 * H], w_hh16[:2 * H])
            K.cast_bf16(w_hh[2 * H:], w_hh16[3 * H:])
            torch.add(b_ih[:2 * H], b_hh[:2 * H], out=bias[:2 * H])
            bias[2 * H:3 * H].copy_(b_ih[2 * H:])
            bias[3 * H:].copy_(b_hh[2 * H:])
        else:
            K.cast_bf16(w_ih, w_ih16[:H])
            K.cast_bf16(w_hh, w_hh16[:H])
            torch.add(b_ih, b_hh, out=bias[:H])


class _RecurrentFn(torch.autograd.Function):
    """x [N,S,C] fp32 -> out [N,S,H] fp32 (+ final h, c [layers,N,H]) for nn.LSTM / nn.GRU / nn.RNN(tanh)."""

    @staticmethod
    def forward(ctx, x, h0, c0, nl, pref, cell, *params):
        N, S, Cin = x.shape
        dev = x.device
        p = [t.detach() for t in params]
        H = p[1].shape[1]
        shadows = pref.take(list(params)) if pref is not None else None
        if shadows is None:
            shadows = _alloc_rnn_shadows(p, cell)
            _cast_rnn_weights(p, shadows, cell)
        inp = K.cast_bf16(x.detach().contiguous().float()).view(N * S, Cin)
        saved = []
        h_ns = torch.empty(nl, N, H, device=dev)  # final states written in place by the kernels (no torch.stack)
        c_ns = torch.empty(nl, N, H, device=dev)
        out = None
        prev_pad = None  # h_pad of the previous layer (input of layers > 0)
        for l in range(nl):
            Kin = p[4 * l].shape[1]
            w_ih16, w_hh16, bias = shadows[l]
            gx = torch.empty(N * S, 4 * H, device=dev)
            if l == 0:
                K.gemm(K.kmajor(inp), K.kmajor(w_ih16), N * S, 4 * H, Kin, out_f32=gx, bias=bias)
            else:  # rows 1..S of every sequence of the previous layer's padded bf16 output, one group per sequence
                A = K.View(prev_pad, H, S + 1, N, H, (S + 1) * H, major=0, row_off=1, gmul=1)
                K.gemm(A, K.kmajor(w_ih16), S, 4 * H, H, groups=N, out_f32=(gx, 4 * H, S * 4 * H), bias=bias)
            out = torch.empty(N, S, H, device=dev)
            h_pad = torch.empty(N, S + 1, H, device=dev, dtype=BF16)
            gates = torch.empty(N, S, 4 * H, device=dev, dtype=BF16)
            c_all = torch.empty(N, S, H, device=dev)
            h_n, c_n = h_ns[l], c_ns[l]
            h0l = h0[l].contiguous().float() if h0 is not None else None
            c0l = c0[l].contiguous().float() if c0 is not None else None
            K.lstm_fwd(gx, w_hh16, N, S, H, h0=h0l, c0=c0l, h_pad=h_pad, out=out, gates=gates, c_all=c_all,
                       h_n=h_n, c_n=c_n, cell=cell)
            saved.append((inp if l == 0 else prev_pad, Kin, w_ih16, w_hh16, h_pad, gates, c_all, c0l, h0l))
            prev_pad = h_pad
        ctx.saved = saved
        ctx.params = list(params)
        ctx.dims = (N, S, Cin, H, nl, cell)
        ctx.set_materialize_grads(False)  # h_n / c_n carry no gradient on this path: no zero tensors made for them
        return out, h_ns, c_ns

    @staticmethod
    def backward(ctx, dout, dh_n, dc_n):
        N, S, Cin, H, nl, cell = ctx.dims
        dev = ctx.saved[0][4].device
        G = _GATES[cell]
        if dh_n is not None or dc_n is not None:
            raise NotImplementedError("cpc2_b200: gradients through the final hidden state of CPCAR are not on the path")
        d = dout.contiguous().float() if dout is not None else torch.zeros(N, S, H, device=dev)
        grads = [None] * (4 * nl)
        sizes = []
        for l in range(nl):  # four-slot (padded) gradient spans; the returned gradients are their leading G*H rows
            sizes += [4 * H * ctx.saved[l][1], 4 * H * H, 4 * H, 4 * H]
        gflat = ops.grad_arena(("rnn", id(ctx.params[0])), sum(sizes), dev)
        gv, off = [], 0
        for n_ in sizes:
            gv.append(gflat[off:off + n_])
            off += n_
        keep = []
        for l in range(nl - 1, -1, -1):
            x_in, Kin, w_ih16, w_hh16, h_pad, gates, c_all, c0l, h0l = ctx.saved[l]
            dgates = torch.empty(N, S, 4 * H, device=dev, dtype=BF16)
            K.lstm_bwd(w_hh16, gates, c_all, d, dgates, N, S, H, c0=c0l, h0=h0l, cell=cell)
            dg2 = dgates.view(N * S, 4 * H)
            dw_ih, dw_hh = gv[4 * l].view(4 * H, Kin), gv[4 * l + 1].view(4 * H, H)
            dbias, dbias2 = gv[4 * l + 2], gv[4 * l + 3]
            dx = torch.empty(N * S, Kin, device=dev)
            with ops.side(1):
                K.colsum(dg2, dbias, N * S, 4 * H, 4 * H)
                if cell == "GRU":  # b_hh = (r, z, hn) slots 0, 1, 3
                    dbias2[:2 * H].copy_(dbias[:2 * H])
                    dbias2[2 * H:3 * H].copy_(dbias[3 * H:])
                else:
                    dbias2.copy_(dbias)
            with ops.side(0):
                # dW_hh = sum_{n,t} dgates[n,t]^T h_{t-1}[n]  (slabs = sequences, h_pad row t is h_{t-1})
                Ag = K.View(dgates, 4 * H, S, N, 4 * H, S * 4 * H, major=1)
                Bh = K.View(h_pad, H, S + 1, N, H, (S + 1) * H, major=1)
                K.gemm(Ag, Bh, 4 * H, H, N * S, k_per_slab=S, k_splits=ops.auto_splits(4 * H, H, N * S, 1, S),
                       out_f32=dw_hh, accumulate=True)
                if cell == "GRU":  # W_hh rows (r, z, n) <- slots 0, 1, 3
                    dw_hh[2 * H:3 * H].copy_(dw_hh[3 * H:])
                # dW_ih = dgates^T . input
                if l == 0:
                    K.gemm(K.mnmajor(dg2), K.mnmajor(x_in), 4 * H, Kin, N * S,
                           k_splits=ops.auto_splits(4 * H, Kin, N * S), out_f32=dw_ih, accumulate=True)
                else:
                    Bx = K.View(x_in, H, S + 1, N, H, (S + 1) * H, major=1, row_off=1)
                    K.gemm(Ag, Bx, 4 * H, Kin, N * S, k_per_slab=S,
                           k_splits=ops.auto_splits(4 * H, Kin, N * S, 1, S), out_f32=dw_ih, accumulate=True)
            keep.append(dgates)
            grads[4 * l:4 * l + 4] = [dw_ih[:G * H], dw_hh[:G * H], dbias[:G * H], dbias2[:G * H]]
            # dx = dgates . W_ih (zero slots of the padded W_ih contribute nothing)
            K.gemm(K.kmajor(dg2), K.mnmajor(w_ih16), N * S, Kin, 4 * H, out_f32=dx)
            d = dx.view(N, S, Kin)
        # inside a step the chain (the encoder's backward comes next) does not wait for dW_ih / dW_hh / the bias sums
        ops.module_join(keep, ctx.saved, dout)
        del keep
        with ops.announce():
            ops.grad_ready("ar", gflat, ctx.params, grads)
        return (d, None, None, None, None, None, *grads)


class CPCModel(nn.Module):
    """cpc/model.py:279-390 (mask_prob must stay 0: the reference's masking path is broken on current torch
    and off by default)."""

    def __init__(self, encoder, AR, mask_prob=0.0, mask_length=10):
        super().__init__()
        self.gEncoder = encoder
        self.gAR = AR
        self.mask_prob = mask_prob
        self.mask_length = mask_length
        if mask_prob > 0.0:
            raise NotImplementedError("cpc2_b200: span masking (mask_prob > 0) is outside the hot path")

    def forward(self, batchData, label):
        encodedData = self.gEncoder(batchData).permute(0, 2, 1)
        cFeature = self.gAR(encodedData)
        return cFeature, encodedData, label

    def splittable(self):
        """True when the training step may run as two halves (forwardPastFuture): a CPCEncoder front end and an
        AR that carries no state between calls (keepHidden feeds the hidden state of ALL 2B rows forward)."""
        return isinstance(self.gEncoder, CPCEncoder) and self.gEncoder.normMode == "layerNorm" and \
            not getattr(self.gAR, "keepHidden", False)  # (batch statistics span all 2B rows: no halves)

    def forwardPastFuture(self, past, future):
        """The two consumers of cpc/train.py:95-104 without the dead work: trainStep concatenates past and future
        windows, runs the model on all 2B rows and keeps `cFeature[:B]` and `encodedData[B:]` only.  Here the
        past half goes encoder -> AR on the current stream while the future half is encoded on a side stream
        (its backward likewise overlaps the heads' and the AR's), and the AR never sees the future rows.
        Returns (cFeature of the past windows [B,S,H], encodedData of the future windows [B,S,C]): the same
        values and the same parameter gradients as the concatenated call."""
        ops.require_cuda(past, future)
        enc = self.gEncoder
        params = enc._params()
        N, _, W = future.shape
        C, dev = enc.dimEncoded, future.device
        pref, enc._pref = getattr(enc, "_pref", None), None
        ws = enc._conv_weights()
        packed = pref.peek(ws) if pref is not None else None
        if packed is None:
            packed = _alloc_packed(C, dev)
            with ops.side(3):
                _pack_conv_weights([w.detach() for w in ws], packed)
        # everything the future half touches on its side stream exists BEFORE the fork event (ops.side(after=))
        fut = future.detach()
        if not (ops.INPLACE_X and fut.dtype == torch.float32 and fut.dim() == 3 and fut.stride(2) == 1):  # (else in place)
            fut = fut.reshape(N, W).contiguous().float().view(N, 1, W)
        # the gradient arena and the weight-gradient staging are cleared on an idle side stream (first written in the
        # backward; both halves' backward start behind the criterion, which joins everything)
        grad = torch.is_grad_enabled()
        pair = _PairCtx(ops.grad_arena("encoder", sum(_grad_sizes(C)), dev, zero_on=14) if grad else None)
        plan_f = _EncPlan(packed=packed, role="future", pair=pair, fwd_bufs=_alloc_fwd(N, W, C, dev),
                          bwd_bufs=_alloc_bwd(N, W, C, dev, zero_on=14) if grad else None)
        plan_f.fwd_after = ops.mark()
        plan_p = _EncPlan(packed=packed, role="past", pair=pair)
        if not torch.is_grad_enabled():
            plan_f.pair = plan_p.pair = None
        zPast = _EncoderFn.apply(past, plan_p, *params).permute(0, 2, 1)
        cPast = self.gAR(zPast)
        zFuture = _EncoderFn.apply(fut, plan_f, *params).permute(0, 2, 1)
        ops.join(6)
        return cPast, zFuture
