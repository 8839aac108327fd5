"""Drop-in for the optimizer of cpc/train.py:472-479 (`torch.optim.Adam(g_params, lr, betas, eps)`), SURVEY 8f.1:
one multi-tensor sm_100a launch (cpc_adam_step) updates every parameter of the model and the criterion.

`FusedAdam` is a `torch.optim.Optimizer`: `param_groups` (so the reference's LR schedulers, train.py:502-517, drive
`lr` as usual), `step()`, `zero_grad()`, and a `state_dict()` in torch.optim.Adam's format (`step`, `exp_avg`,
`exp_avg_sq` per parameter), so optimizer states of reference checkpoints load and ours load back into
torch.optim.Adam (train.py:486-490).  The step counter, learning rate and bias corrections live in device memory:
a `step()` captured in a CUDA graph replays correctly (`sync_hyper()` pushes a changed lr before the replay).
"""
import os

import torch

from . import _lib
from . import kernels as K
from . import ops

# update an early-announced bucket only when the next bucket is announced (see FusedAdam._on_stage); CPC_ADAM_DEFER=0
# restores update-right-behind-the-reduction
DEFER_UPDATE = int(os.environ.get("CPC_ADAM_DEFER", "1"))
# persistent CTAs of an update launched from a gradient announcement, i.e. next to the rest of the backward (0: one CTA
# per 4096-element chunk, which saturates HBM and slowed the recurrent backward beside it from 100 to 132 us).
# Same-box sweep at N = 1 (ms/step): 64: 1.287, 96: 1.17, 148: 1.097, 192: 1.084, 256: 1.084, 296: 1.083, 384: 1.10,
# 592: 1.10, unthrottled: 1.106.
STAGED_CTAS = int(os.environ.get("CPC_ADAM_CTAS", "256"))
SMALL_BUCKET = int(os.environ.get("CPC_ADAM_SMALL", "65536"))  # parameters; -1: every staged update on one stream


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False):
        if weight_decay != 0 or amsgrad:
            raise NotImplementedError("cpc2_b200.FusedAdam: weight_decay / amsgrad are not used by cpc/train.py")
        if not 0.0 <= lr or not 0.0 <= eps or not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0:
            raise ValueError("Invalid Adam hyper-parameter")
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=0, amsgrad=False))
        self._groups = [None] * len(self.param_groups)  # per group: device state, span tables, flat moments
        self._dp, self._done, self._advanced = None, set(), False
        self._shadowed = []  # arenas whose bf16 shadow the launches of this iteration wrote
        self._first_done = None  # event behind the launch that advanced the step counter of this iteration
        self._deferred = []  # (stage, pairs, event) of early buckets whose update waits for the next announcement

    # ------------------------------------------------------------------------------------------ internals
    def _init_group(self, gi):
        group = self.param_groups[gi]
        params = [p for p in group["params"] if p.requires_grad]
        if not params:
            return None
        dev = params[0].device
        if dev.type != "cuda":
            raise RuntimeError("cpc2_b200.FusedAdam runs on CUDA parameters only (no CPU fallback)")
        total = sum((p.numel() + 3) // 4 * 4 for p in params)
        m_flat = torch.zeros(total, device=dev)
        v_flat = torch.zeros(total, device=dev)
        off, step0 = 0, 0.0
        for p in params:
            n = p.numel()
            st = self.state[p]
            m, v = m_flat[off:off + n].view(p.shape), v_flat[off:off + n].view(p.shape)
            if "exp_avg" in st:  # loaded state (load_state_dict) or a previous table: keep the values
                m.copy_(st["exp_avg"])
                v.copy_(st["exp_avg_sq"])
                step0 = max(step0, float(st.get("step", 0.0)))
            st["exp_avg"], st["exp_avg_sq"] = m, v
            off += (n + 3) // 4 * 4
        b1, b2 = group["betas"]
        hyper = (group["lr"], b1, b2, group["eps"])
        state = torch.tensor([step0, hyper[0], b1, b2, hyper[3], 0.0, 0.0, 0.0], dtype=torch.float64, device=dev)
        g = dict(params=params, ids={id(p) for p in params}, state=state, hyper=hyper, tables={}, staging=[], m=m_flat,
                 v=v_flat, pinned=torch.empty(8, dtype=torch.float64).pin_memory())
        self._groups[gi] = g
        return g

    def _table(self, g, name, pairs):
        """Device span table `name` over (parameter, gradient) pairs; rebuilt only when a pointer changed."""
        chunk = _lib.lib().cpc_adam_chunk()
        # arena-homed parameters (the prediction heads, a transformer AR): the kernel also stores the updated value
        # into the arena's persistent bf16 shadow, so the next step finds its GEMM operands ready (ops.Arena)
        shadow = [ops.shadow_target(p) for p, _ in pairs]
        keys = [(p.data_ptr(), gr.data_ptr(), p.numel(), sh[1]) for (p, gr), sh in zip(pairs, shadow)]
        t = g["tables"].setdefault(name, dict(keys=None, table=None, n_spans=0, blocks=0, arenas=[]))
        if keys == t["keys"]:
            return t
        t["arenas"] = list({id(a): a for a, _ in shadow if a is not None}.values())
        spans = (_lib.AdamSpan * max(len(pairs), 1))()
        blocks = 0
        for i, (p, gr) in enumerate(pairs):
            if gr.dtype != torch.float32 or not gr.is_contiguous() or not p.is_contiguous() or p.dtype != torch.float32 \
                    or gr.numel() != p.numel():
                raise RuntimeError("cpc2_b200.FusedAdam: parameters and gradients must be contiguous fp32 of equal size")
            st = self.state[p]
            spans[i].p, spans[i].g = p.data_ptr(), gr.data_ptr()
            spans[i].m, spans[i].v = st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
            spans[i].p16 = shadow[i][1]
            spans[i].n, spans[i].first_block = p.numel(), blocks
            blocks += (p.numel() + chunk - 1) // chunk
        # pinned staging buffer + async copy.  While a CUDA graph is being captured the copy becomes a graph node that
        # is replayed, so that buffer (and the device table it fills) must stay alive and untouched for the lifetime
        # of the optimizer: one per captured table, kept in g["staging"].  Eager rebuilds (gradient pointers that move
        # from step to step) reuse ONE pinned buffer per table, guarded by the event of its previous copy -- the list
        # no longer grows with the number of steps.
        raw = torch.frombuffer(bytearray(bytes(spans)), dtype=torch.uint8)
        capturing = torch.cuda.is_current_stream_capturing()
        host = None if capturing else t.get("host")
        if host is None or host.numel() != raw.numel():
            host = torch.empty(raw.numel(), dtype=torch.uint8).pin_memory()
            if capturing:
                g["staging"].append(host)
            else:
                t["host"], t["host_ev"] = host, torch.cuda.Event()
        elif t.get("host_ev") is not None:
            t["host_ev"].synchronize()  # the previous async copy out of this buffer has completed
        host.copy_(raw)
        if t["table"] is None or t["table"].numel() != host.numel() or capturing:
            t["table"] = torch.empty(host.numel(), dtype=torch.uint8, device=g["state"].device)
        t["table"].copy_(host, non_blocking=True)
        if not capturing:
            t["host_ev"].record()
        t["keys"], t["n_spans"], t["blocks"] = keys, len(pairs), blocks
        return t

    # ------------------------------------------------------------------------------------------ staged stepping
    def attach(self, dp=None):
        """Step each gradient bucket as soon as its module announces it (ops.grad_ready: criterion heads first,
        90 % of the parameters), on a side stream -- behind the event that marks the bucket's all-reduce done when
        `dp` (BucketedAllReduce) is given -- so the update overlaps the rest of the backward and the next bucket's
        all-reduce.  step() then only joins and handles leftovers."""
        if len(self.param_groups) != 1:
            raise NotImplementedError("FusedAdam.attach: staged stepping supports a single parameter group")
        self._dp = dp
        if self._groups[0] is None:
            self._init_group(0)
        if self._on_stage not in ops.GRAD_READY:
            ops.GRAD_READY.append(self._on_stage)  # after dp's hook: the all-reduce is queued first
        return self

    def detach(self):
        if self._on_stage in ops.GRAD_READY:
            ops.GRAD_READY.remove(self._on_stage)

    def _on_stage(self, stage, gflat, params=None, grads=None):
        g = self._groups[0]
        if params is None or g is None:
            return
        mine = g["ids"]
        pairs = [(p, gr) for p, gr in zip(params, grads) if gr is not None and id(p) in mine and id(p) not in self._done]
        if not pairs:
            return
        if not torch.cuda.is_current_stream_capturing():
            self.sync_hyper()
        dp = self._dp
        after = None
        if dp is not None and dp.world > 1 and dp.on_gpu:
            # dp's hook ran first: this bucket's all-reduce is the last thing queued on dp.stream.  The update waits
            # for exactly that, on its OWN stream -- on dp.stream the next bucket's all-reduce would queue behind the
            # 16.6 M-parameter update of the heads and the last (encoder) reduction would trail the backward
            after = torch.cuda.Event()
            after.record(dp.stream)
        self._done.update(id(p) for p, _ in pairs)
        heads = stage.startswith("criterion")
        if (DEFER_UPDATE == 2 and heads) or (DEFER_UPDATE == 1 and after is not None and stage.endswith(".ffn")):
            # Reduced now, UPDATED later.  Mode 1 (N > 1): the early bucket (the heads' feed-forward weights, announced a
            # third of the way into the heads' backward so that its all-reduce starts early) waits for the next
            # announcement.  Mode 2: every bucket of the heads waits for the AR's announcement -- their 16.6 M-parameter
            # pass saturates HBM for 100 us and, launched at the end of the heads' backward, ran beside the
            # latency-bound recurrent backward (100 -> 132 us); behind it, it runs beside the encoder's backward.
            self._deferred.append((stage, pairs, after))
            return
        if not (DEFER_UPDATE == 2 and heads):
            self._flush_deferred()
        self._launch(g, stage, pairs, after)

    def _flush_deferred(self):
        if not self._deferred:
            return
        here = ops.mark()  # not before this point of the announcing stream
        for stage, pairs, after in self._deferred:
            self._launch(self._groups[0], stage, pairs, after, also=here)
        self._deferred = []

    def _launch(self, g, stage, pairs, after, also=None):
        # small buckets (conv0's 3.3 K parameters, announced last) go to a second stream: on the first they queued behind
        # the update of the bucket announced just before and the step ended that much later
        small = sum(p.numel() for p, _ in pairs) <= SMALL_BUCKET
        with ops.side(12 if small else 9, after=after):
            cur = torch.cuda.current_stream()
            if after is not None:
                cur.wait_event(after)  # (also when side streams are disabled)
            if also is not None:
                cur.wait_event(also)
            if self._advanced and self._first_done is not None:
                cur.wait_event(self._first_done)  # the launch that advanced the step counter (maybe on the other stream)
            t = self._table(g, stage, pairs)
            # staged launches run next to the backward: large ones as a throttled pass (STAGED_CTAS persistent CTAs)
            K.adam_step(t["table"], t["n_spans"], t["blocks"], g["state"], advance=not self._advanced,
                        max_ctas=STAGED_CTAS if t["blocks"] > 4 * max(STAGED_CTAS, 1) else 0)
            self._shadowed.extend(t["arenas"])
            if not self._advanced:
                self._first_done = ops.mark()
        self._advanced = True

    # ------------------------------------------------------------------------------------------ snapshot / restore
    def snapshot(self):
        """Copies of everything step() mutates besides the parameters (moments, device step counter), so that a
        caller can run throw-away steps (GraphedTrainStep's warm-up before capture) and put the optimizer back.
        Buffers keep their addresses: tables captured in a CUDA graph stay valid after restore()."""
        out = []
        for gi in range(len(self.param_groups)):
            if self._groups[gi] is None:
                self._init_group(gi)
            g = self._groups[gi]
            out.append(None if g is None else (g["m"].clone(), g["v"].clone(), g["state"].clone()))
        return out

    def restore(self, snap):
        for g, sn in zip(self._groups, snap):
            if g is not None and sn is not None:
                g["m"].copy_(sn[0])
                g["v"].copy_(sn[1])
                g["state"].copy_(sn[2])

    def sync_hyper(self):
        """Pushes lr / betas / eps of param_groups to the device if a scheduler changed them (call before replaying
        a CUDA graph that contains step(); step() itself calls it when it runs eagerly)."""
        for gi, group in enumerate(self.param_groups):
            g = self._groups[gi]
            if g is None:
                continue
            b1, b2 = group["betas"]
            hyper = (group["lr"], b1, b2, group["eps"])
            if hyper != g["hyper"]:
                g["pinned"][:4] = torch.tensor(hyper, dtype=torch.float64)
                g["state"][1:5].copy_(g["pinned"][:4], non_blocking=True)
                g["hyper"] = hyper

    # ------------------------------------------------------------------------------------------ Optimizer API
    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for gi in range(len(self.param_groups)):  # (raises on CPU parameters before anything touches CUDA)
            if self._groups[gi] is None:
                self._init_group(gi)
        capturing = torch.cuda.is_current_stream_capturing()
        self._flush_deferred()
        ops.join(12)
        ops.join(9)  # staged launches of this iteration (FusedAdam.attach); dp.finish() has joined dp.stream
        for gi in range(len(self.param_groups)):
            g = self._groups[gi]
            if g is None:
                continue
            if not capturing:
                self.sync_hyper()
            pairs = [(p, p.grad) for p in g["params"] if p.grad is not None and id(p) not in self._done]
            if pairs:
                t = self._table(g, "rest", pairs)
                K.adam_step(t["table"], t["n_spans"], t["blocks"], g["state"], advance=not (self._advanced and gi == 0))
                self._shadowed.extend(t["arenas"])
            # arenas whose bf16 shadow matched the parameters before this update still match: the kernel stored both
            current = [a for a in {id(a): a for a in self._shadowed}.values() if a.shadow_valid()]
            # the kernel wrote through raw pointers: tell autograd (and ops.Prefetched / ops.Arena, which validate
            # bf16 weight shadows by version counter) that the parameters changed
            torch.autograd.graph.increment_version(g["params"])
            for a in current:
                a.mark_shadow_written()
        self._shadowed = []
        self._done.clear()
        self._advanced = False
        self._first_done = None
        return loss

    def state_dict(self):
        for gi, g in enumerate(self._groups):
            if g is None:
                continue
            step = g["state"][0:1].to(torch.float32).cpu()[0]
            for p in g["params"]:
                self.state[p]["step"] = step.clone()
        return super().state_dict()

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._groups = [None] * len(self.param_groups)  # rebuilt (moments re-homed, step taken over) at the next step
