"""bench.py -- training throughput of the CPC2 hot path (forward + backward [+ gradient all-reduce] + Adam).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--quick]
    (N > 1: launched by torchrun, one rank per GPU, NCCL)

Metric (BASELINE.json): training audio-seconds per second = N * B * sizeWindow / 16000 / step_time on the CPC
default config (encoder 256, 1-layer LSTM-256 AR, 12 transformer prediction heads, 128 negatives, window
20480, dropout on), B = 8 windows per GPU (weak scaling), synthetic audio, bf16 operands / fp32 accumulation.

Prints ONE JSON line (rank 0).
  value         inputs resident in HBM, CUDA-event time per step summed over K steps, max over ranks, L2 flushed
                between steps.
  e2e           the same step through the public API (cpc2_b200.train.GraphedTrainStep.pipelined) with a pinned HOST
                batch copied in and the losses / accuracies copied out and read by the host every step.
  roofline      the dominant kernel family of the step, timed per launch with CUDA events in a separate eager pass
                behind a GPU-side delay; `achieved` on ALGORITHMIC FLOPs (the all-pairs logits products count
                2*T*(NEG+1)*D*K, not the executed 2*T*(B*S)*D*K).
  cpu_baseline  the reference's own training step on the host cores: the UNMODIFIED reference modules imported from
                oracle/_ref/cpc_reference.zip (kind "reference"; packed by build() from /root/reference) or, when that
                archive is absent, the CPU oracle port (kind "port").
  configs       (timed AFTER the headline region, so `value` is untouched) BASELINE.json configs[2..4]:
                transformer AR, hidden 512 / 2-layer LSTM / window 40960, and the criterion-only sweep corners.
  dropin        cpc2_b200.train.trainStep (the reference's entry point, cpc/train.py:72-142) over a 20-batch loader,
                host-timed: with FusedAdam (one CUDA-graph replay per batch) and with torch.optim.Adam (eager).
  stock_gpu     the UNMODIFIED reference modules on the same B200 (eager PyTorch: cuDNN + cuBLAS), fp32/TF32 and under
                torch.autocast(bfloat16), same batch -- the stock kernels this build replaces.
  replica_param_max_abs_diff (N > 1)  max |parameter - rank 0's parameter| over all ranks after the timed loop.
`--impl reference` times the reference CPU implementation alone with the same metric.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SAMPLE_RATE = 16000
WORKLOAD = ("CPC default: encoder 256, LSTM-256 AR (1 layer), nPredicts 12 transformer heads, 128 negatives, "
            "sizeWindow 20480, batch 8 per GPU")


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(hbm=p["hbm_gbs"], tensor=p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                    tensor_burst=p["bf16_tflops"], source="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, tensor=1400.0, tensor_burst=1590.0, source="fallback (B200_PROFILING.md)")


def traffic_of(kernel):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture (profiles/traffic.json)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(kernel, {}).get("bytes_per_launch")
    except (OSError, ValueError):
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.samples, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.proc is not None:
            time.sleep(0.25)
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        sm = sorted(int(s[0]) for s in self.samples if s and s[0].isdigit())
        mx = [int(s[1]) for s in self.samples if len(s) > 1 and s[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for s in self.samples if len(s) >= 6 for i in range(4) if s[2 + i] == "Active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def synthetic_batch(B, W, rank):
    g = torch.Generator().manual_seed(1234 + rank)
    return (0.1 * torch.randn(B, 2, 1, W, generator=g)).pin_memory() if torch.cuda.is_available() else \
        0.1 * torch.randn(B, 2, 1, W, generator=g)


# ------------------------------------------------------------------------------------------------------ CPU arms
def reference_available():
    from oracle import ref_harness as rh
    return rh.reference_location() is not None


def cpu_reference_seconds(over, B, steps, warmup):
    """The UNMODIFIED reference's batch-loop body (cpc/train.py:95-109, train mode) on all host threads."""
    from oracle import ref_harness as rh
    torch.set_num_threads(os.cpu_count())
    sec, loss, _ = rh.timed_reference_steps(over, B, steps, warmup, device="cpu")
    return sec, torch.get_num_threads()


def cpu_port_seconds(args, B, steps, warmup):
    """The CPU oracle's training step (fwd + bwd) on all host threads; returns (seconds per step, threads)."""
    from cpc2_b200 import train as T
    from oracle import cpc_oracle as oc
    torch.set_num_threads(os.cpu_count())
    torch.manual_seed(0)
    model, crit = T.buildModelAndCriterion(args)
    state = {k: v.detach().float() for k, v in list(model.state_dict().items()) + list(crit.state_dict().items())}
    S = args.sizeWindow // 160
    W = S - args.nPredicts
    seq = synthetic_batch(B, args.sizeWindow, 0).clone()
    times = []
    for it in range(warmup + steps):
        torch.manual_seed(4321 + it)
        bi, si = oc.draw_negative_indices(B, S, W, args.negativeSamplingExt)
        t0 = time.perf_counter()
        oc.train_step(state, seq, bi, si, K=args.nPredicts, NEG=args.negativeSamplingExt, ar_mode=args.arMode,
                      n_layers=args.nLevelsGRU, rnn_mode=args.rnnMode)
        if it >= warmup:
            times.append(time.perf_counter() - t0)
    return sum(times) / len(times), torch.get_num_threads()


def cpu_baseline(args, B, steps, warmup):
    """-> dict for the `cpu_baseline` key: the reference itself when its archive travelled, else the oracle port."""
    W = args.sizeWindow
    if reference_available():
        sec, threads = cpu_reference_seconds({}, B, steps, warmup)
        kind, what = "reference", "unmodified reference modules (oracle/_ref/cpc_reference.zip), train mode, fp32"
    else:
        sec, threads = cpu_port_seconds(args, B, steps, warmup)
        kind, what = "port", "oracle/cpc_oracle.py, fp32"
    return {"value": B * W / SAMPLE_RATE / sec, "unit": "audio-s/s", "cores": threads, "kind": kind,
            "sample": f"{steps} full fwd+bwd steps of the same batch ({B} x {W} samples) after {warmup} warm-up, {what}, "
                      f"{threads} threads, {sec:.2f} s/step"}


def run_reference(opt, out):
    """--impl reference: the reference's CPU implementation of the step with every host thread."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from cpc2_b200 import train as T
    args = T.defaultArgs()
    B = 8
    ref = reference_available()
    timer = (lambda b, k, w: cpu_reference_seconds({}, b, k, w)) if ref else (lambda b, k, w: cpu_port_seconds(args, b, k, w))
    # bound the run: probe one step, shrink the per-step sample if K+W full steps would take too long
    probe, threads = timer(B, 1, 0)
    if probe * (opt.steps + opt.warmup) > 240:
        B = 2
    sec, threads = timer(B, opt.steps, opt.warmup)
    value = B * args.sizeWindow / SAMPLE_RATE / sec
    what = "unmodified reference, train mode" if ref else "oracle port"
    sample = f"{opt.steps} full fwd+bwd steps of batch {B} x {args.sizeWindow} samples ({what}, fp32, {threads} threads)"
    line = {
        "impl": "reference", "metric": "train_audio_sec_per_sec", "value": value, "unit": "audio-s/s",
        "n_gpus": opt.gpus, "steps": opt.steps, "warmup": opt.warmup, "ms_per_step": sec * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD.replace("batch 8 per GPU", f"batch {B} on the host CPU"), "global_batch": B},
        "cpu_baseline": {"value": value, "unit": "audio-s/s", "cores": threads, "kind": "reference" if ref else "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    out.emit(json.dumps(line))


# ------------------------------------------------------------------------------------------------------ GPU arm
def kernel_profile(fwd_bwd, params):
    """One eager forward+backward with a CUDA-event pair around every C-ABI launch.  A GPU-side delay is queued
    first so the launches pile up behind it and the events bracket kernel time, not host launch latency."""
    from cpc2_b200 import kernels as K
    for p in params:
        p.grad = None
    fwd_bwd()  # allocator warm-up at these shapes
    for p in params:
        p.grad = None
    torch.cuda.synchronize()
    K.PROFILE = []
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda._sleep(int(60e6))  # ~30-40 ms of GPU-side delay
    t0.record()
    fwd_bwd()
    t1.record()
    torch.cuda.synchronize()
    prof, K.PROFILE = K.PROFILE, None
    for p in params:
        p.grad = None
    rows = {}
    for name, work, e0, e1 in prof:
        r = rows.setdefault(name, dict(name=name, launches=0, ms=0.0, flops=0.0, bytes=0.0, kind=work.get("kind", "hbm")))
        r["launches"] += 1
        r["ms"] += e0.elapsed_time(e1)
        r["flops"] += work.get("flops", 0.0)
        r["bytes"] += work.get("bytes", 0.0)
    return sorted(rows.values(), key=lambda r: -r["ms"]), t0.elapsed_time(t1), prof


def roofline_of(rows, pk):
    total = sum(r["ms"] for r in rows)
    top = rows[0]
    if top["kind"] == "hbm":
        ach = top["bytes"] / (top["ms"] * 1e-3) / 1e9
        roof = {"bound": "hbm", "achieved": ach, "peak": pk["hbm"], "unit": "GB/s", "frac": ach / pk["hbm"]}
    else:  # tensor, or the latency-bound recurrence reported against the tensor peak as SURVEY 8d asks
        ach = top["flops"] / (top["ms"] * 1e-3) / 1e12
        roof = {"bound": "tensor", "achieved": ach, "peak": pk["tensor"], "unit": "TFLOP/s", "frac": ach / pk["tensor"]}
        if top["kind"] != "tensor":
            roof["note"] = f"{top['kind']}-bound kernel; see DESIGN.md"
    roof.update({"kernel": top["name"], "launches_per_step": top["launches"], "share_of_kernel_time": top["ms"] / total,
                 "avg_launch_us": 1e3 * top["ms"] / top["launches"], "flops_counted": "algorithmic",
                 "peak_source": pk["source"]})
    return roof, total


def kernel_table(rows, total):
    return [{"name": r["name"], "launches": r["launches"], "us": round(1e3 * r["ms"], 1), "share": round(r["ms"] / total, 4),
             "tflops": round(r["flops"] / (r["ms"] * 1e-3) / 1e12, 1) if r["flops"] else None,
             "gbs": round(r["bytes"] / (r["ms"] * 1e-3) / 1e9, 1) if r["bytes"] else None, "kind": r["kind"]}
            for r in rows]


class Ctx:
    """Process-wide facts of this bench run."""

    def __init__(self):
        import torch.distributed as dist
        self.dist = dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        # gradient arenas in NVSwitch-multicast symmetric memory + cpc_allreduce_mc when the box supports it
        # (CPC_SYMMETRIC=0: NCCL all-reduce; =1: fail instead of falling back to NCCL)
        self.sym = {"0": False, "1": True}.get(os.environ.get("CPC_SYMMETRIC", "auto"), "auto")
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)  # > 126 MB L2

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, *vals):
        if self.world == 1:
            return list(vals)
        t = torch.tensor(vals, device=self.dev, dtype=torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.tolist()


class TrainRig:
    """Model + criterion + data-parallel buckets + fused Adam + the captured step, for one configuration."""

    def __init__(self, ctx, args, B):
        from cpc2_b200 import kernels as K
        from cpc2_b200 import train as T
        from cpc2_b200.dp import BucketedAllReduce
        from cpc2_b200.optim import FusedAdam
        self.ctx, self.args, self.B = ctx, args, B
        torch.manual_seed(0)
        model, crit = T.buildModelAndCriterion(args)
        self.model, self.crit = model.to(ctx.dev), crit.to(ctx.dev)
        torch.manual_seed(4321 + ctx.rank)  # every rank draws its own negatives, as in the reference
        self.dp = BucketedAllReduce(self.model, self.crit, symmetric=ctx.sym)
        self.dp.broadcast_parameters()
        diag = os.environ.get("CPC_DIAG_NO_ALLREDUCE", "")
        if diag == "1":  # diagnosis only (the printed line says so)
            self.dp._all_reduce_mean = lambda t: None
            self.dp._skip_collective = True
        elif diag:  # comma-separated stage prefixes whose reduction is skipped (symmetric path only)
            self.dp._skip_collective = tuple(diag.split(","))
        n0 = K.LAUNCHES
        self.adam = FusedAdam(list(self.crit.parameters()) + list(self.model.parameters()), lr=args.learningRate,
                              betas=(args.beta1, args.beta2), eps=args.epsilon)  # cpc/train.py:472-479
        self.adam.attach(self.dp if ctx.world > 1 else None)  # each bucket is stepped as soon as it is (all-)reduced
        self.step = T.GraphedTrainStep(self.model, self.crit, B, args.sizeWindow, device=ctx.dev, warmup=3,
                                       train_mode=True, after_backward=self.dp.finish, optimizer=self.adam)
        self.launches_per_step = (K.LAUNCHES - n0) // 4  # 3 eager warm-up passes + 1 capture pass
        self.host_seq = synthetic_batch(B, args.sizeWindow, ctx.rank)
        self.step.static_seq.copy_(self.host_seq)

    def timed(self, steps, warmup, clocks=False):
        ctx = self.ctx
        for _ in range(max(warmup, 3)):
            self.step.step()
        ctx.barrier()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        sampler = ClockSampler(ctx.local) if clocks else None
        if sampler is not None:
            sampler.__enter__()
        ctx.barrier()
        for e0, e1 in evs:
            ctx.flush.zero_()
            e0.record()
            self.step.step()
            e1.record()
        ctx.barrier()
        if sampler is not None:
            sampler.__exit__()
        return sum(e0.elapsed_time(e1) for e0, e1 in evs), sampler

    def profile(self):
        from cpc2_b200 import train as T
        params = list(self.model.parameters()) + list(self.crit.parameters())
        return kernel_profile(lambda: T.forwardBackward(self.model, self.crit, self.step.static_seq, self.step.static_label),
                              params)

    def close(self):
        import gc
        self.step.close()
        self.adam.detach()
        for p in list(self.model.parameters()) + list(self.crit.parameters()):
            p.grad = None
        self.dp.remove()  # (releases the symmetric gradient buffer synchronously: never inside a later capture)
        self.step = self.adam = self.dp = self.model = self.crit = None
        gc.collect()
        torch.cuda.synchronize()


def measure_extra_config(ctx, name, over, B, steps, pk):
    """One of BASELINE.json configs[2..3] through the same rig as the headline: ms/step, audio-s/s, dominant kernel."""
    from cpc2_b200 import train as T
    args = T.defaultArgs(**over)
    rig = TrainRig(ctx, args, B)
    try:
        ms_total, _ = rig.timed(steps, 3)
        (ms_total,) = ctx.max_over_ranks(ms_total)
        audio_s = ctx.world * B * args.sizeWindow / SAMPLE_RATE
        res = {"workload": name, "batch_per_gpu": B, "sizeWindow": args.sizeWindow, "steps": steps,
               "ms_per_step": ms_total / steps, "value": audio_s * steps / (ms_total * 1e-3), "unit": "audio-s/s",
               "gpu_launches_per_step": rig.launches_per_step,
               "params": sum(p.numel() for p in list(rig.model.parameters()) + list(rig.crit.parameters()))}
        if ctx.world == 1:
            rows, _, _ = rig.profile()
            roof, total = roofline_of(rows, pk)
            res["roofline"] = roof
            res["kernel_time_sum_ms"] = total
            res["kernels"] = kernel_table(rows, total)[:8]
        return res
    finally:
        rig.close()


def criterion_sweep(ctx, steps, pk, quick):
    """BASELINE.json configs[4]: CPCUnsupersivedCriterion forward + backward (+ the heads' gradient all-reduce and
    their fused Adam update) alone, on random c, z ~ N(0,1) [8, 128, 256] per GPU, as one CUDA-graph replay per step."""
    from cpc2_b200 import criterion as cr
    from cpc2_b200 import kernels as K
    from cpc2_b200 import ops
    from cpc2_b200.dp import BucketedAllReduce
    from cpc2_b200.optim import FusedAdam
    B, S, D = 8, 128, 256
    corners = [(32, 4), (128, 12), (1024, 24)] if quick else [(n, k) for n in (32, 128, 1024) for k in (4, 12, 24)]
    out = []
    for rnn in ("transformer", "linear"):
        for NEG, Kp in corners:
            torch.manual_seed(0)
            crit = cr.CPCUnsupersivedCriterion(Kp, D, D, NEG, rnnMode=rnn, sizeInputSeq=S).to(ctx.dev).train()
            g = torch.Generator().manual_seed(17 + ctx.rank)
            c = torch.randn(B, S, D, generator=g).to(ctx.dev).requires_grad_(True)
            z = torch.randn(B, S, D, generator=g).to(ctx.dev).requires_grad_(True)
            label = torch.zeros(B, dtype=torch.long, device=ctx.dev)
            dp = BucketedAllReduce(None, crit, symmetric=ctx.sym)
            dp.broadcast_parameters()
            adam = FusedAdam(list(crit.parameters()), lr=2e-4)
            adam.attach(dp if ctx.world > 1 else None)
            params = list(crit.parameters())

            def body():
                for p in params:
                    p.grad = None
                c.grad = z.grad = None
                crit.wPrediction.prefetch()
                losses, acc = crit(c, z, label, None)
                losses.sum().backward()
                dp.finish()
                adam.step()
                return losses

            n0 = K.LAUNCHES
            snap_p = [p.detach().clone() for p in params]
            snap_o = adam.snapshot()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(3):
                    body()
                with torch.no_grad():
                    for p, q in zip(params, snap_p):
                        p.copy_(q)
                    adam.restore(snap_o)
                ops.refresh_shadows(params)  # (as GraphedTrainStep does: the capture then holds no weight cast)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            for p in params:
                p.grad = None
            with torch.cuda.graph(graph):
                losses = body()
            launches = (K.LAUNCHES - n0) // 4
            for _ in range(3):
                graph.replay()
            ctx.barrier()
            evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
            for e0, e1 in evs:
                ctx.flush.zero_()
                e0.record()
                graph.replay()
                e1.record()
            ctx.barrier()
            (ms,) = ctx.max_over_ranks(sum(e0.elapsed_time(e1) for e0, e1 in evs))
            W = S - Kp
            algo = 3 * 2.0 * B * W * (NEG + 1) * D * Kp  # logits fwd + its two backward products
            out.append({"rnnMode": rnn, "negatives": NEG, "nPredicts": Kp, "us_per_step": round(1e3 * ms / steps, 1),
                        "windows_per_s": round(ctx.world * B * steps / (ms * 1e-3), 1),
                        "logit_algorithmic_gflop": round(algo / 1e9, 3), "gpu_launches_per_step": launches,
                        "loss_mean": round(float(losses.detach().mean()), 4),
                        "head_params": sum(p.numel() for p in params)})
            torch.cuda.synchronize()
            graph.reset()
            del graph, losses
            c.grad = z.grad = None
            adam.detach()
            dp.remove()
    return {"shape": "c, z ~ N(0,1) [8, 128, 256] per GPU; step = criterion fwd + bwd + heads' all-reduce + fused Adam, "
                     "one CUDA-graph replay, L2 flushed between steps", "n_gpus": ctx.world, "cases": out}


def dropin_numbers(ctx, args, B, n_batches):
    """cpc2_b200.train.trainStep -- the reference's own entry point (cpc/train.py:72-142) -- over a synthetic loader,
    timed on the HOST clock around the call (H2D of every batch and the log read-back included)."""
    from cpc2_b200 import train as T
    from cpc2_b200.optim import FusedAdam
    W = args.sizeWindow
    distinct = [synthetic_batch(B, W, 100 + i) for i in range(4)]
    label = torch.zeros(B, dtype=torch.long)
    loader = [(distinct[i % 4], label) for i in range(n_batches)]
    res = {}
    for key, make in (("trainStep_fused_adam_graph", lambda ps: FusedAdam(ps, lr=args.learningRate)),
                      ("trainStep_torch_adam_eager", lambda ps: torch.optim.Adam(ps, lr=args.learningRate))):
        torch.manual_seed(0)
        model, crit = T.buildModelAndCriterion(args)
        model, crit = model.to(ctx.dev), crit.to(ctx.dev)
        opt = make(list(crit.parameters()) + list(model.parameters()))
        T.trainStep(loader[:3], model, crit, opt, None, 10 ** 9)  # warm-up (captures the graph on the fused path)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        logs = T.trainStep(loader, model, crit, opt, None, 10 ** 9)
        torch.cuda.synchronize()
        sec = time.perf_counter() - t0
        res[key] = {"value": n_batches * B * W / SAMPLE_RATE / sec, "unit": "audio-s/s", "ms_per_batch": 1e3 * sec / n_batches,
                    "batches": n_batches, "loss_mean": float(logs["locLoss_train"].mean()),
                    "graph_replay": len(T._GRAPHED.get(model, {})) > 0}
        for r in T._GRAPHED.pop(model, {}).values():
            r.close()
    return res


def stock_gpu_numbers(ctx, B, steps):
    """The UNMODIFIED reference modules (eager PyTorch on cuDNN / cuBLAS) on this GPU, same batch shape, train mode."""
    from oracle import ref_harness as rh
    if rh.reference_location() is None:
        return {"unavailable": "oracle/_ref/cpc_reference.zip not packed (python oracle/make_ref.py in the build container)"}
    res = {}
    for key, ac, tf32 in (("fp32_tf32", None, True), ("autocast_bf16", torch.bfloat16, True), ("fp32_strict", None, False)):
        try:
            sec, loss, wall = rh.timed_reference_steps({}, B, steps, 3, device=str(ctx.dev), autocast=ac, tf32=tf32)
            res[key] = {"value": B * 20480 / SAMPLE_RATE / wall, "unit": "audio-s/s", "ms_per_step_wall": 1e3 * wall,
                        "ms_per_step_gpu": 1e3 * sec, "loss_mean": loss}
        except Exception as e:  # a stock path that does not run under autocast is a fact worth recording
            res[key] = {"error": f"{type(e).__name__}: {str(e)[:200]}"}
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = True
    res["what"] = ("cpc/train.py:95-109 on the reference's own modules, eager, fwd + bwd only (no optimizer), host wall "
                   "clock per step with a final synchronize; value = audio-s/s of that")
    return res


def replica_check(ctx, rigs_params):
    """max over ranks and parameters of |p - p on rank 0| (must be 0.0: replicas stay bit-identical)."""
    flat = torch.cat([p.detach().reshape(-1) for p in rigs_params])
    ref = flat.clone()
    ctx.dist.broadcast(ref, 0)
    d = (flat - ref).abs().max().to(torch.float64)
    ctx.dist.all_reduce(d, op=ctx.dist.ReduceOp.MAX)
    return float(d)


def run_ours(opt, out):
    from cpc2_b200 import _lib
    from cpc2_b200 import train as T
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a B200; there is no CPU fallback for the product path")
    if _lib.needs_build():
        _lib.build()
    ctx = Ctx()
    world, rank = ctx.world, ctx.rank
    args = T.defaultArgs()
    B, W = args.batchSizeGPU, args.sizeWindow
    rig = TrainRig(ctx, args, B)
    host_out = torch.empty(2, args.nPredicts).pin_memory()
    # ---- value: inputs resident in HBM, per-step CUDA events, L2 flushed between steps
    ms_total, clocks = rig.timed(opt.steps, opt.warmup, clocks=True)
    # ---- e2e: pinned host batch in, losses out, every step, through the public API
    ctx.barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    n_read = 0
    for losses_h, acc_h in rig.step.pipelined(rig.host_seq for _ in range(opt.steps)):
        host_out[0].copy_(losses_h)  # the caller consumes every step's losses / accuracies on the host
        host_out[1].copy_(acc_h)     # (train.py:121-122), one step behind the GPU
        n_read += 1
    assert n_read == opt.steps
    t1.record()
    ctx.barrier()
    ms_e2e = t0.elapsed_time(t1)
    ms_total, ms_e2e = ctx.max_over_ranks(ms_total, ms_e2e)
    loss_now = float(host_out[0].mean())
    audio_s = world * B * W / SAMPLE_RATE
    dp = rig.dp
    line = {
        "metric": "train_audio_sec_per_sec", "value": audio_s * opt.steps / (ms_total * 1e-3), "unit": "audio-s/s",
        "n_gpus": world, "steps": opt.steps, "warmup": max(opt.warmup, 3), "ms_per_step": ms_total / opt.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "global_batch": B * world, "parallelism": f"dp{world}",
                   "l2_flush_between_steps": True, "cuda_graph": True, "dropout": "train mode (p=0.1 in the heads)",
                   "programmatic_dependent_launch": os.environ.get("CPC_PDL", "0") == "1",
                   "capture_stream_priority": __import__("cpc2_b200.ops", fromlist=["ops"]).MAIN_PRIORITY,
                   "capture_stream_priority": __import__("cpc2_b200.ops", fromlist=["ops"]).MAIN_PRIORITY,
                   "optimizer": "fused multi-tensor Adam (cpc_adam_step) inside the timed step, per bucket after its all-reduce",
                   "grad_allreduce_bytes": dp.bytes_per_step if world > 1 else 0,
                   "grad_allreduce": ("none" if world == 1 else "cpc_allreduce_mc (NVSwitch multimem, in place)"
                                      if dp.symmetric is not None else "NCCL all-reduce (AVG)"),
                   **({"DIAGNOSTIC": "all-reduce skipped, not a valid number"}
                      if os.environ.get("CPC_DIAG_NO_ALLREDUCE") else {})},
        "e2e": {"value": audio_s * opt.steps / (ms_e2e * 1e-3), "unit": "audio-s/s",
                "h2d_bytes_per_step": rig.host_seq.numel() * 4, "d2h_bytes_per_step": host_out.numel() * 4,
                "ms_per_step": ms_e2e / opt.steps},
        "gpu_launches": rig.launches_per_step * opt.steps, "gpu_launches_per_step": rig.launches_per_step,
        "clocks": clocks.summary(), "loss_mean": loss_now,
    }
    if world > 1:
        line["replica_param_max_abs_diff"] = replica_check(
            ctx, list(rig.crit.parameters()) + list(rig.model.parameters()))
    pk = peaks()
    if world == 1:
        rows, eager_ms, _ = rig.profile()
        roof, total = roofline_of(rows, pk)
        roof.update({"traffic": traffic_of(roof["kernel"]),
                     "traffic_unit": "bytes per launch (ncu dram read+write, profiles/traffic.json)"})
        line["roofline"] = roof
        line["kernels"] = kernel_table(rows, total)
        line["kernel_time_sum_ms"] = total
    rig.close()
    del rig
    torch.cuda.empty_cache()

    # ---- everything below is timed AFTER the headline region and never touches `value`
    extra_steps = 5 if opt.quick else 10

    def guarded(fn, *a):
        try:
            return fn(*a)
        except Exception as e:  # the headline line must survive a failing extra
            if world > 1:
                raise  # (ranks must stay in lock step: fail loudly instead of desynchronising the collectives)
            return {"error": f"{type(e).__name__}: {str(e)[:300]}"}

    if opt.headline_only:
        if rank == 0:
            out.emit(json.dumps(line))
        finish(ctx)
        return
    # At N > 1 an extra section that fails on one rank takes the job down (see `guarded`): the headline region above is
    # complete and valid at that point, so rank 0 still prints its line -- with the failure named in `configs` -- from the
    # exception path, or from SIGTERM when torchrun stops it because ANOTHER rank died.
    emitted = []

    def emit_headline(note):
        if rank == 0 and not emitted:
            emitted.append(True)
            line.setdefault("configs", {})
            line["configs"]["error"] = note
            out.emit(json.dumps(line))

    if world > 1 and rank == 0:
        # a Python-level handler would only run once the main thread is back in the interpreter, and here it may sit in a
        # collective that never completes: the C-level handler writes the signal number to a pipe a watcher thread reads
        import signal
        rfd, wfd = os.pipe()
        os.set_blocking(wfd, False)
        signal.signal(signal.SIGTERM, lambda signum, frame: None)
        signal.set_wakeup_fd(wfd, warn_on_full_buffer=False)

        def on_term():
            os.read(rfd, 1)
            emit_headline("terminated during the extra sections (another rank failed); the headline region is complete")
            os._exit(1)

        threading.Thread(target=on_term, daemon=True).start()
    try:
        run_extras(opt, out, ctx, line, args, B, pk, extra_steps, guarded)
    except BaseException as e:
        if world > 1:
            emit_headline(f"{type(e).__name__}: {str(e)[:300]}")
        raise
    if rank == 0 and not emitted:
        out.emit(json.dumps(line))
    finish(ctx)


def run_extras(opt, out, ctx, line, args, B, pk, extra_steps, guarded):
    world, rank = ctx.world, ctx.rank
    configs = {}
    configs["config3_transformer_ar"] = guarded(
        measure_extra_config, ctx, "CPC with transformer AR (1 layer, sizeSeq 128, rel-pos; feature_loader.py:216-220), "
        "transformer heads, samespeaker batches (identical criterion), batch 8 per GPU",
        dict(arMode="transformer"), 8, extra_steps, pk)
    torch.cuda.empty_cache()
    configs["config4_hidden512_lstm2_w40960"] = guarded(
        measure_extra_config, ctx, "CPC hidden 512, 2-layer LSTM AR, sizeWindow 40960, transformer heads, batch 8 per GPU",
        dict(hiddenEncoder=512, hiddenGar=512, nLevelsGRU=2, sizeWindow=40960), 8, extra_steps, pk)
    torch.cuda.empty_cache()
    configs["config5_criterion_sweep"] = guarded(criterion_sweep, ctx, extra_steps, pk, opt.quick)
    torch.cuda.empty_cache()
    line["configs"] = configs
    if rank == 0 and world == 1:
        line["dropin"] = guarded(dropin_numbers, ctx, args, B, 20)
        torch.cuda.empty_cache()
        line["stock_gpu"] = guarded(stock_gpu_numbers, ctx, B, 5 if opt.quick else 10)
        torch.cuda.empty_cache()
        line["cpu_baseline"] = guarded(cpu_baseline, args, B, 2, 1)


def finish(ctx):
    if ctx.world > 1:
        # every captured graph (they hold NCCL launches) has been closed above, so the communicator can be torn down
        # normally; a watchdog turns a teardown that still blocks into a clean exit instead of a hung job
        ctx.barrier()
        sys.stdout.flush()
        sys.stderr.flush()
        dog = threading.Timer(60.0, lambda: os._exit(0))
        dog.daemon = True
        dog.start()
        ctx.dist.destroy_process_group()


class StdoutToStderr:
    """Everything libraries print on fd 1 (NCCL's version banner, ...) goes to stderr; the JSON line alone reaches
    the real stdout (emit())."""

    def __init__(self):
        sys.stdout.flush()
        self.real = os.dup(1)
        os.dup2(2, 1)

    def emit(self, text):
        sys.stdout.flush()
        os.write(self.real, (text + "\n").encode())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--quick", action="store_true", help="fewer steps / sweep corners in the extra sections")
    ap.add_argument("--headline-only", action="store_true", help="skip configs / dropin / stock_gpu / cpu_baseline (A/B runs)")
    opt = ap.parse_args()
    out = StdoutToStderr()
    if opt.impl == "reference":
        run_reference(opt, out)
    else:
        run_ours(opt, out)


if __name__ == "__main__":
    main()
