"""Host-side glue shared by the drop-in modules: parameter arenas, dropout RNG state, GEMM call shapes.

Nothing here computes on tensors' values: it allocates buffers, describes views and launches the C-ABI
kernels of libcpc_b200.so through `cpc2_b200.kernels`.
"""
import os

import weakref

import torch

from . import kernels as K

BF16 = torch.bfloat16
NUM_SMS = 148
# Debug switch: keep the gradients that feed a normalisation backward in fp32 between kernels.  Measured to
# make no difference to gradient parity (profiles/r01_grad_noise.log), so the default is bf16 (half the bytes).
GRAD_F32 = False
# ChannelNorm / LayerNorm of 256- and 512-wide rows inside the epilogue of the GEMM that produces them (one accumulator
# tile is one whole row, or the row is split over a 4-CTA cluster that exchanges partial statistics through DSMEM): no
# separate cpc_rownorm_fwd launch.  False restores the two-kernel form (also used for other widths).
FUSE_NORM = os.environ.get("CPC_FUSE_NORM", "1") != "0"


# Inside torch.autograd.Function.forward grad mode is always off, so "will a backward follow?" cannot be read from
# torch.is_grad_enabled() there: the Function says it (`with ops.saving(any(ctx.needs_input_grad))`), everything
# below it asks ops.will_backward() before allocating buffers only a backward reads.
_SAVING = []


class saving:
    def __init__(self, flag):
        self.flag = bool(flag)

    def __enter__(self):
        _SAVING.append(self.flag)
        return self

    def __exit__(self, *exc):
        _SAVING.pop()
        return False


def will_backward():
    return _SAVING[-1] if _SAVING else torch.is_grad_enabled()


# conv0 reads fp32 waveform views in place (row stride) instead of a contiguous copy; CPC_INPLACE_X=0: copy first
INPLACE_X = os.environ.get("CPC_INPLACE_X", "1") != "0"
# force_bn of the fused-norm GEMM launches: 0 = the kernel picks the tile form by launch size (one 256-column tile per
# row, or the row split over a 4-CTA cluster); 256 = always one tile per row.  Tests that compare two DIFFERENTLY SIZED
# launches of the same arithmetic set 256, so that both sum the row statistics in the same order.
NORM_FORCE_BN = 0


def can_fuse_norm(C, rows=None, groups=1):
    """256-wide rows: always (the kernel picks one BN = 256 tile per row or the 4-CTA cluster split by the number of row
    tiles).  512-wide rows only exist in the cluster form, which streams the A tile four times: worth it for small
    launches only (rows x groups up to 16 row tiles of 128), otherwise GEMM + cpc_rownorm_fwd is faster (measured at
    hidden 512 / window 40960: 2.29 ms of GEMM time fused everywhere against 1.95 + 0.18 ms in two kernels)."""
    if not FUSE_NORM:
        return False
    if C == 256:
        return True
    return C == 512 and rows is not None and ((rows + 127) // 128) * groups <= 16


# Callbacks invoked by the modules' backward right after every kernel writing a module's flat fp32 gradient
# arena has been enqueued: fn(stage, gflat) with stage in {"criterion", "ar", "encoder"}.  The data-parallel
# wrapper (dp.BucketedAllReduce) hooks its in-place all-reduce here, BEFORE autograd hands the views to
# AccumulateGrad, so .grad aliases the reduced arena (no copies, CUDA-graph capturable).
GRAD_READY = []


def grad_ready(stage, gflat, params=None, grads=None):
    """`params` / `grads`: the parameters whose gradients live in `gflat` and their views of it (what an optimizer
    that steps a stage as soon as it is reduced needs: autograd has not assigned .grad yet at this point)."""
    for fn in GRAD_READY:
        fn(stage, gflat, params, grads)


# Where a module's flat gradient arena lives: None = a fresh zeroed tensor per backward; the data-parallel wrapper
# installs fn(owner, numel, device) -> zeroed fp32 [numel] (or None) to place the arenas in NVSwitch-multicast
# symmetric memory, so that its all-reduce is one in-switch kernel (dp.BucketedAllReduce(symmetric=True)).
# `owner` identifies the arena across steps: the module's ops.Arena, or "encoder" / "ar" for the model's own.
GRAD_ARENA = None


def grad_arena(owner, numel, device, zero_on=None):
    """Zeroed fp32 [numel] gradient arena.  `zero_on`: side-stream index on which the clearing memset is issued
    (the ALLOCATION always happens on the current stream, so the block belongs to the pool of the stream whose
    kernels write it -- the rule above -- and only the memset overlaps other work; the caller joins that stream)."""
    if GRAD_ARENA is not None:
        t = GRAD_ARENA(owner, numel, device, zero_on) if zero_on is not None else GRAD_ARENA(owner, numel, device)
        if t is not None:
            return t
    if zero_on is None:
        return torch.zeros(numel, dtype=torch.float32, device=device)
    t = torch.empty(numel, dtype=torch.float32, device=device)
    with side(zero_on):
        t.zero_()
    return t


# ---- independent kernels on forked streams -----------------------------------------------------------
# Most launches of the step are 10-40 us and fill a fraction of the 148 SMs; weight-gradient GEMMs, bias
# column sums and weight casts do not sit on the critical path, so they are issued on side streams that fork
# from / join back into the current stream.  Inside a CUDA-graph capture this becomes parallel graph branches.
# Rule: every tensor a side stream touches is allocated on the main stream BEFORE the fork and kept alive past
# the join (the launchers only borrow pointers), so the caching allocator never sees cross-stream reuse.
CONCURRENT = True
_SIDE = {}
# priority of the stream GraphedTrainStep captures on (-1 = above the side streams, 0 = same; CPC_MAIN_PRIORITY)
MAIN_PRIORITY = int(os.environ.get("CPC_MAIN_PRIORITY", "-1"))
# announce the heads' feed-forward gradients (3/4 of all parameters) a third of the way into their backward:
# "auto" = when gradients are all-reduced across MORE THAN TWO ranks (+1 % at N = 8 in round 1, neutral at N = 1; at N = 2
# the extra bucket costs more than its early start buys: 1.229 ms per step without against 1.243 / 1.246 with it, same
# box, profiles/r02_ab_n2_adam.txt), "1" / "0" force it
EARLY_FFN = os.environ.get("CPC_EARLY_FFN", "auto")


# announce the encoder's gradients before its backward has finished (model._EncoderFn.backward): 1 = conv1-4 as one
# bucket before conv0's backward runs, 2 = one bucket per layer as each weight gradient completes; their all-reduce and
# optimizer step then overlap the rest of the backward instead of trailing the step; 3 = as 1, with conv0's small bucket
# announced first, straight behind conv0's backward.  CPC_EARLY_ENCODER=0 restores the single "encoder" announcement.
# Same-box A/B (tools/ab_early_encoder.sh, two repeats): N = 2: 1.246 / 1.242 / 1.315 ms per step for modes 1 / 3 / 2,
# N = 1: 1.079 / 1.077 / 1.082.
EARLY_ENCODER = int(os.environ.get("CPC_EARLY_ENCODER", "3"))


def early_ffn():
    if EARLY_FFN == "auto":
        import torch.distributed as dist
        return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 2
    return EARLY_FFN not in ("0", False)

_PENDING = []


class side:
    """`with ops.side(i):` launches the body on side stream i (ordered after everything already queued on the
    current stream, or -- `after=event` -- only after that event: the body must then touch nothing that was
    allocated or written on the current stream later than the event).  `ops.join()` makes the current stream wait
    for every side stream forked FROM IT since; `ops.join(i)` for stream i only.  A side stream has one parent at
    a time: code that itself runs on a side stream forks with indices nobody else uses concurrently."""

    def __init__(self, index, after=None):
        self.index = index
        self.after = after
        self.ctx = None

    def __enter__(self):
        if not CONCURRENT or K.PROFILE is not None:
            return self
        cur = torch.cuda.current_stream()
        key = (cur.device.index, self.index)
        s = _SIDE.get(key)
        if s is None:
            s = _SIDE[key] = torch.cuda.Stream(device=cur.device)
        if all(k != key for k, _, _ in _PENDING):
            _PENDING.append((key, s, cur))
        if self.after is not None:
            s.wait_event(self.after)
        else:
            s.wait_stream(cur)
        self.ctx = torch.cuda.stream(s)
        self.ctx.__enter__()
        return self

    def __exit__(self, *exc):
        if self.ctx is not None:
            self.ctx.__exit__(*exc)
            self.ctx = None
        return False


# side streams that only an explicit join(i) waits for (9: the staged optimizer step, which must not be pulled onto
# the critical path by the blanket join() at the end of a module's backward)
DETACHED = {9, 12}


def join(index=None):
    cur = torch.cuda.current_stream()
    for item in list(_PENDING):
        key, s, parent = item
        if index is None and key[1] in DETACHED:
            continue
        # blanket join: only the streams forked from the current one; join(i): stream i whoever forked it last
        if key[0] == cur.device.index and ((index is None and parent.cuda_stream == cur.cuda_stream) or key[1] == index):
            cur.wait_stream(s)
            _PENDING.remove(item)


# ---- deferred joins --------------------------------------------------------------------------------------------
# A module's backward used to end with join(): the dependent chain (the stream autograd runs on) waited for the
# module's weight-gradient GEMMs and bias sums before the next module's backward could start, although only the
# all-reduce / optimizer step of THAT module needs them -- and since the chain's stream outranks the side streams
# (MAIN_PRIORITY) those launches finish late.  Inside train.forwardBackward (step_scope) a module instead
#   * hands every buffer its side streams may still be reading to module_join(...) (kept alive until the step's final
#     join: the caching allocator must not recycle them for later launches of the chain), and
#   * announces its gradients from `with ops.announce():` -- side stream 13, which waits for the chain up to here AND for
#     every side stream forked from it, so dp's all-reduce and the staged Adam launch are ordered behind the weight
#     gradients without the chain ever waiting.
# Outside a step_scope (a module used on its own) module_join() is the plain join() and announce() does nothing.
DEFER_JOINS = os.environ.get("CPC_DEFER_JOINS", "1") != "0"
ANNOUNCE_STREAM = 13  # (8 is the sampler's: criterion.prefetchIndices)
_STEP_DEPTH = 0
_KEEP = []


def _deferring():
    return _STEP_DEPTH > 0 and DEFER_JOINS and CONCURRENT and K.PROFILE is None


class step_scope:
    """train.forwardBackward: backward passes of the modules inside may defer their joins to the end of the step."""

    def __enter__(self):
        global _STEP_DEPTH
        _STEP_DEPTH += 1
        return self

    def __exit__(self, *exc):
        global _STEP_DEPTH
        _STEP_DEPTH -= 1
        if _STEP_DEPTH == 0:
            join()
            _KEEP.clear()
        return False


def module_join(*keep):
    """join() at the end of a module's backward -- or, inside a step_scope, nothing but keeping `keep` alive until the
    step's final join.  Returns True when the current stream really joined."""
    if _deferring():
        _KEEP.append(keep)
        return False
    join()
    return True


class announce:
    def __init__(self):
        self.ctx = None

    def __enter__(self):
        if not _deferring():
            return self
        cur = torch.cuda.current_stream()
        pend = [s for key, s, parent in _PENDING
                if key[0] == cur.device.index and parent.cuda_stream == cur.cuda_stream and key[1] not in DETACHED
                and key[1] != ANNOUNCE_STREAM]
        self.ctx = side(ANNOUNCE_STREAM)
        self.ctx.__enter__()
        x = torch.cuda.current_stream()
        for s in pend:
            x.wait_stream(s)
        return self

    def __exit__(self, *exc):
        if self.ctx is not None:
            self.ctx.__exit__(*exc)
            self.ctx = None
        return False


def wait_side(index):
    """The current stream waits for side stream `index` WITHOUT consuming the pending join of its parent."""
    cur = torch.cuda.current_stream()
    s = _SIDE.get((cur.device.index, index))
    if s is not None and s.cuda_stream != cur.cuda_stream:
        cur.wait_stream(s)


def mark():
    """Event at the current point of the current stream (None when side streams are disabled)."""
    if not CONCURRENT or K.PROFILE is not None:
        return None
    ev = torch.cuda.Event()
    ev.record(torch.cuda.current_stream())
    return ev


# events attached to tensors handed across autograd nodes: "this gradient is complete at <event>", so that a
# consumer running on a side stream can start from the event instead of from the producer's whole stream
_READY = {}


def set_ready(t, ev):
    if ev is not None:
        _READY[(t.device.index, t.data_ptr())] = ev


def pop_ready(t):
    return _READY.pop((t.device.index, t.data_ptr()), None)


class Prefetched:
    """Result of work issued early on a side stream (bf16 weight shadows, packed conv weights), valid only while
    the parameters it was computed from are unchanged (checked through their autograd version counters)."""

    def __init__(self, params, value, stream_index):
        self.versions = [p._version for p in params]
        self.ptrs = [p.data_ptr() for p in params]
        self.value, self.stream_index = value, stream_index

    def peek(self, params):
        """The value if still valid, WITHOUT synchronising: the caller orders its stream after side stream
        `stream_index` itself (ops.join / ops.wait_side)."""
        if [p._version for p in params] == self.versions and [p.data_ptr() for p in params] == self.ptrs:
            return self.value
        return None

    def take(self, params):
        join(self.stream_index)
        if [p._version for p in params] == self.versions and [p.data_ptr() for p in params] == self.ptrs:
            return self.value
        return None


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("cpc2_b200 runs on CUDA tensors only (sm_100a kernels, no CPU fallback)")


def reject_replica(module):
    """nn.DataParallel over SEVERAL devices (cpc/train.py:529-532 with nGPU > 1) replicates the module per device and
    calls the replicas from one thread each; the engines / arenas / side-stream state of this package are per process
    (one process per GPU).  device_ids of length 1 never replicates and works.  Fail loudly instead of launching on
    another device's pointers."""
    if getattr(module, "_is_replica", False):
        raise RuntimeError("cpc2_b200: nn.DataParallel across several devices is not supported -- run one process per "
                           "GPU (--distributed / DistributedDataParallel, or cpc2_b200.dp.BucketedAllReduce)")


# shapes the recurrence kernels are instantiated for (csrc/lstm.cu dispatch_h)
RECURRENT_H = (32, 64, 128, 256, 512)
# cpc_nce_bwd stages one row of d(logits) per warp in shared memory: 8 warps x B*S floats <= 200 KB (csrc/nce.cu)
NCE_BWD_MAX_BS = 6400


def auto_splits(M, N, K_total, groups=1, k_per_slab=0):
    """Split-K factor for skinny weight-gradient GEMMs so that ~2 CTAs per SM are in flight."""
    kps = k_per_slab or K_total
    kblocks = ((kps + 63) // 64) * (K_total // kps)
    bn = 256 if N >= 256 else (128 if N > 64 else 64)
    tiles = ((M + 127) // 128) * ((N + bn - 1) // bn) * groups
    if tiles >= NUM_SMS or kblocks <= 4:
        return 1
    return max(1, min(kblocks // 4, (2 * NUM_SMS + tiles - 1) // tiles))


# parameter id -> (arena, offset, parameter): where cpc_adam_step stores the bf16 copy of an arena-homed parameter
_SHADOW_OWNER = {}


def shadow_target(p):
    """(arena, address of p's slot in the arena's persistent bf16 shadow) or (None, None)."""
    ent = _SHADOW_OWNER.get(id(p))
    if ent is None:
        return None, None
    arena, q = ent[0](), ent[2]()
    if arena is None or q is not p or getattr(arena, "_shadow", None) is None or not arena.valid():
        return None, None
    return arena, arena._shadow.data_ptr() + 2 * ent[1]


def refresh_shadows(params):
    """Bring the bf16 shadow of every arena that homes one of `params` up to date (eager cast on the current stream)."""
    seen = set()
    for p in params:
        ent = _SHADOW_OWNER.get(id(p))
        arena = ent[0]() if ent is not None else None
        if arena is not None and id(arena) not in seen:
            seen.add(id(arena))
            arena.refresh_shadow()


class Arena:
    """A flat fp32 buffer that OWNS the storage of a list of parameters, in a layout chosen so that
    groups of parameters form the stacked GEMM operands directly ([group][rows][cols]); one cast kernel
    produces the bf16 shadow of all of them per step.  Parameter names/shapes (state_dict) are untouched:
    each nn.Parameter's .data becomes a view into the arena."""

    def __init__(self, params):
        self.params = list(params)
        self.offsets = []
        off = 0
        for p in self.params:
            self.offsets.append(off)
            off += (p.numel() + 7) // 8 * 8  # keep every fp32 and bf16-shadow view 16-byte aligned
        self.numel = off
        self.flat = None

    def valid(self):
        if self.flat is None or not self.params:
            return False
        base = self.flat.data_ptr()
        dev = self.flat.device
        for p, off in zip(self.params, self.offsets):  # every pointer (~110): a re-pointed .data anywhere rebuilds
            if p.device != dev or p.data_ptr() != base + 4 * off:
                return False
        return True

    def ensure(self):
        """(Re)build the arena if the module was moved (.cuda()/.to()) since the last call."""
        if self.valid():
            return self.flat
        dev = self.params[0].device
        flat = torch.zeros(self.numel, dtype=torch.float32, device=dev)
        with torch.no_grad():
            for p, off in zip(self.params, self.offsets):
                view = flat[off:off + p.numel()].view(p.shape)
                view.copy_(p.data)
                p.data = view
        self.flat = flat
        return flat

    # ---- persistent bf16 shadow -------------------------------------------------------------------------------
    # ONE bf16 copy of the arena, kept across steps and valid while the parameters' autograd versions are the ones it
    # was written at.  Writers: the cast kernel (any time the parameters changed behind our back: torch.optim.Adam,
    # load_state_dict, an in-place edit) and cpc_adam_step itself (CpcAdamSpan.p16: the fused optimizer stores the
    # updated parameter in both precisions, so a training loop on FusedAdam never launches the 16.6 M-element cast
    # that used to sit next to conv0 at the top of every step).
    def shadow_buffer(self):
        flat = self.ensure()
        sh = getattr(self, "_shadow", None)
        if sh is None or sh.device != flat.device or sh.numel() != flat.numel():
            sh = torch.empty(flat.shape, dtype=BF16, device=flat.device)
            self._shadow, self._shadow_key = sh, None
            for q, off in zip(self.params, self.offsets):
                _SHADOW_OWNER[id(q)] = (weakref.ref(self), off, weakref.ref(q))
        return sh

    def _key(self):
        return ([q._version for q in self.params], self.flat.data_ptr() if self.flat is not None else 0)

    def shadow_valid(self):
        return getattr(self, "_shadow", None) is not None and self.valid() and self._shadow_key == self._key()

    def mark_shadow_written(self):
        """The shadow now equals the parameters (called by whoever just wrote it)."""
        self._shadow_key = self._key()

    def refresh_shadow(self):
        """Cast now, on the current stream, if the shadow is stale."""
        if not self.shadow_valid():
            sh = self.shadow_buffer()
            K.cast_bf16(self.flat, sh)
            self.mark_shadow_written()
        return self._shadow

    def prefetch(self, stream_index):
        """Make the bf16 shadow current: nothing to do when the optimizer wrote it, else the cast is issued on a
        side stream now and the next bf16() call joins it."""
        if self.shadow_valid():
            return
        sh = self.shadow_buffer()
        with side(stream_index):
            K.cast_bf16(self.flat, sh)
        self.mark_shadow_written()
        self._pref_stream = stream_index

    def bf16(self):
        st, self._pref_stream = getattr(self, "_pref_stream", None), None
        if st is not None:
            join(st)
        return self.refresh_shadow()

    def view(self, flat, index, shape=None):
        """View of parameter `index` inside another flat tensor with the same layout (bf16 shadow, grads)."""
        p = self.params[index]
        off = self.offsets[index]
        return flat[off:off + p.numel()].view(p.shape if shape is None else shape)

    def span(self, flat, first, count, shape):
        """Contiguous view over `count` consecutive equally sized parameters starting at `first`."""
        off = self.offsets[first]
        n = 1
        for s in shape:
            n *= s
        assert self.offsets[first + count - 1] + self.params[first + count - 1].numel() - off == n, \
            "parameters are not densely packed"
        return flat[off:off + n].view(shape)

    def grad_views(self, gflat):
        return [self.view(gflat, i) for i in range(len(self.params))]


class DropoutState:
    """Device-side (seed, step) pair feeding the counter RNG of the kernels; CUDA-graph safe."""
    _states = {}

    @classmethod
    def next(cls, device, side_index=None):
        """Snapshot of (seed, step), then step += 1.  side_index: the two tiny kernels are issued on that side stream
        (the caller joins it -- ops.join(side_index) -- before the first kernel that reads the snapshot and before the
        next call): they then run beside the launch that precedes the first dropout site instead of in front of it."""
        key = (device.type, device.index)
        st = cls._states.get(key)
        if st is None:
            st = torch.tensor([torch.initial_seed() & 0x7FFFFFFFFFFFFFFF, 0], dtype=torch.int64, device=device)
            cls._states[key] = st
        if side_index is None:
            snap = st.clone()
            st[1:2].add_(1)
            return snap
        snap = torch.empty_like(st)  # allocated on the current stream (the rule above), written on the side stream
        with side(side_index):
            snap.copy_(st)
            st[1:2].add_(1)
        return snap


def linear_fwd(x, w, M, N, Kd, groups=1, x_shared=False, **epi):
    """y[g] = x[g] . w[g]^T ; x bf16 [M,K] (shared) or [G,M,K]; w bf16 [G,N,K] or [N,K]."""
    A = K.kmajor(x, gmul=0 if (x_shared or x.dim() == 2) else 1)
    B = K.kmajor(w)
    K.gemm(A, B, M, N, Kd, groups=groups, **epi)


def linear_dx(dy, w, M, N, Kd, groups=1, **epi):
    """dx[g] = dy[g] . w[g] ; dy bf16 [G,M,N]; w bf16 [G,N,K] consumed MN-major (reduction over its rows)."""
    A = K.kmajor(dy)
    B = K.mnmajor(w)
    K.gemm(A, B, M, Kd, N, groups=groups, **epi)


def linear_dw(dy, x, dw, M, N, Kd, groups=1, x_shared=False):
    """dw[g] (fp32 [G,N,K], pre-zeroed) += dy[g]^T . x[g]; reduction over the M rows, split-K."""
    A = K.mnmajor(dy)
    B = K.mnmajor(x, gmul=0 if (x_shared or x.dim() == 2) else 1)
    K.gemm(A, B, N, Kd, M, groups=groups, k_splits=auto_splits(N, Kd, M, groups), out_f32=dw, accumulate=True)
