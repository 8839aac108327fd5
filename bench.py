"""bench.py -- training throughput of the CPC2 hot path (forward + backward [+ gradient all-reduce]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    (N > 1: launched by torchrun, one rank per GPU, NCCL)

Metric (BASELINE.json): training audio-seconds per second = N * B * sizeWindow / 16000 / step_time on the CPC
default config (encoder 256, 1-layer LSTM-256 AR, 12 transformer prediction heads, 128 negatives, window
20480, dropout on), B = 8 windows per GPU (weak scaling), synthetic audio, bf16 operands / fp32 accumulation.

Prints ONE JSON line (rank 0).  `value`: inputs resident in HBM, CUDA-event time per step summed over K steps,
max over ranks, L2 flushed between steps.  `e2e`: the same step through the public API
(cpc2_b200.train.GraphedTrainStep.pipelined) with a pinned HOST batch copied in and the losses / accuracies copied
out and read by the host every step (batch i+1 in flight and result i-1 being read while replay i runs).
`roofline`: the dominant kernel family of the step, timed per launch with CUDA events in a separate eager
pass behind a GPU-side delay (so host launch latency is not measured).  `cpu_baseline`: the CPU oracle
(oracle/cpc_oracle.py, the restatement of the reference pinned by tests/golden) on the host cores.
`--impl reference` times that CPU implementation alone with the same metric.
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


def cpu_step_seconds(args, B, steps, warmup):
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


def run_reference(opt, out):
    """--impl reference: the reference algorithm's CPU implementation (oracle port) with every host thread."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from cpc2_b200 import train as T
    args = T.defaultArgs()
    B = 8
    # bound the run: probe one step, shrink the per-step sample if K+W full steps would take too long
    probe, threads = cpu_step_seconds(args, B, 1, 0)
    if probe * (opt.steps + opt.warmup) > 240:
        B = 2
    sec, threads = cpu_step_seconds(args, B, opt.steps, opt.warmup)
    value = B * args.sizeWindow / SAMPLE_RATE / sec
    sample = f"{opt.steps} full fwd+bwd steps of batch {B} x {args.sizeWindow} samples (oracle port, fp32, {threads} threads)"
    line = {
        "impl": "reference", "metric": "train_audio_sec_per_sec", "value": value, "unit": "audio-s/s",
        "n_gpus": opt.gpus, "steps": opt.steps, "warmup": opt.warmup, "ms_per_step": sec * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD.replace("batch 8 per GPU", f"batch {B} on the host CPU"), "global_batch": B},
        "cpu_baseline": {"value": value, "unit": "audio-s/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    out.emit(json.dumps(line))


def kernel_profile(model, crit, seq, label):
    """One eager forward+backward with a CUDA-event pair around every C-ABI launch.  A GPU-side delay is queued
    first so the launches pile up behind it and the events bracket kernel time, not host launch latency."""
    from cpc2_b200 import kernels as K
    from cpc2_b200 import train as T
    for p in list(model.parameters()) + list(crit.parameters()):
        p.grad = None
    T.forwardBackward(model, crit, seq, label)  # allocator warm-up at these shapes
    for p in list(model.parameters()) + list(crit.parameters()):
        p.grad = None
    torch.cuda.synchronize()
    K.PROFILE = []
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda._sleep(int(60e6))  # ~30-40 ms of GPU-side delay
    t0.record()
    T.forwardBackward(model, crit, seq, label)
    t1.record()
    torch.cuda.synchronize()
    prof, K.PROFILE = K.PROFILE, None
    rows = {}
    for name, work, e0, e1 in prof:
        r = rows.setdefault(name, dict(name=name, launches=0, ms=0.0, flops=0.0, bytes=0.0, kind=work.get("kind", "hbm")))
        r["launches"] += 1
        r["ms"] += e0.elapsed_time(e1)
        r["flops"] += work.get("flops", 0.0)
        r["bytes"] += work.get("bytes", 0.0)
    return sorted(rows.values(), key=lambda r: -r["ms"]), t0.elapsed_time(t1), prof


def run_ours(opt, out):
    import torch.distributed as dist
    from cpc2_b200 import _lib
    from cpc2_b200 import kernels as K
    from cpc2_b200 import train as T
    from cpc2_b200.dp import BucketedAllReduce
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a B200; there is no CPU fallback for the product path")
    if _lib.needs_build():
        _lib.build()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    args = T.defaultArgs()
    B, W = args.batchSizeGPU, args.sizeWindow
    torch.manual_seed(0)
    model, crit = T.buildModelAndCriterion(args)
    model, crit = model.to(dev), crit.to(dev)
    torch.manual_seed(4321 + rank)  # every rank draws its own negatives, as in the reference
    # gradient arenas in NVSwitch-multicast symmetric memory + cpc_allreduce_mc when the box supports it
    # (CPC_SYMMETRIC=0: NCCL all-reduce; =1: fail instead of falling back to NCCL)
    sym = {"0": False, "1": True}.get(os.environ.get("CPC_SYMMETRIC", "auto"), "auto")
    dp = BucketedAllReduce(model, crit, symmetric=sym)
    dp.broadcast_parameters()
    if os.environ.get("CPC_DIAG_NO_ALLREDUCE") == "1":  # diagnosis only (the printed line says so): what the step
        dp._all_reduce_mean = lambda t: None            # costs at N ranks without the collective
    host_seq = synthetic_batch(B, W, rank)
    host_out = torch.empty(2, args.nPredicts).pin_memory()

    # launches per step, counted while the step is captured into the graph
    n0 = K.LAUNCHES
    from cpc2_b200.optim import FusedAdam
    adam = FusedAdam(list(crit.parameters()) + list(model.parameters()), lr=args.learningRate,
                    betas=(args.beta1, args.beta2), eps=args.epsilon)  # cpc/train.py:472-479
    adam.attach(dp if world > 1 else None)  # each bucket is stepped as soon as it is (all-)reduced
    step = T.GraphedTrainStep(model, crit, B, W, device=dev, warmup=3, train_mode=True, after_backward=dp.finish,
                              optimizer=adam)
    launches_per_step = (K.LAUNCHES - n0) // 4  # 3 eager warm-up passes + 1 capture pass
    step.static_seq.copy_(host_seq)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(opt.warmup, 3)):
        step.step()
    barrier()
    # ---- value: inputs resident in HBM, per-step CUDA events, L2 flushed between steps
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(opt.steps)]
    with ClockSampler(local) as clocks:
        barrier()
        for e0, e1 in evs:
            flush.zero_()
            e0.record()
            step.step()
            e1.record()
        barrier()
    ms_total = sum(e0.elapsed_time(e1) for e0, e1 in evs)
    # ---- e2e: pinned host batch in, losses out, every step, through the public API
    barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    n_read = 0
    for losses_h, acc_h in step.pipelined(host_seq for _ in range(opt.steps)):
        host_out[0].copy_(losses_h)  # the caller consumes every step's losses / accuracies on the host
        host_out[1].copy_(acc_h)     # (train.py:121-122), one step behind the GPU
        n_read += 1
    assert n_read == opt.steps
    t1.record()
    barrier()
    ms_e2e = t0.elapsed_time(t1)
    if world > 1:
        t = torch.tensor([ms_total, ms_e2e], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total, ms_e2e = t.tolist()
    loss_now = float(host_out[0].mean())
    audio_s = world * B * W / SAMPLE_RATE
    value = audio_s * opt.steps / (ms_total * 1e-3)
    e2e = audio_s * opt.steps / (ms_e2e * 1e-3)
    line = {
        "metric": "train_audio_sec_per_sec", "value": value, "unit": "audio-s/s", "n_gpus": world,
        "steps": opt.steps, "warmup": max(opt.warmup, 3), "ms_per_step": ms_total / opt.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "global_batch": B * world, "parallelism": f"dp{world}",
                   "l2_flush_between_steps": True, "cuda_graph": True, "dropout": "train mode (p=0.1 in the heads)",
                   "optimizer": "fused multi-tensor Adam (cpc_adam_step) inside the timed step, per bucket after its all-reduce",
                   "grad_allreduce_bytes": dp.bytes_per_step if world > 1 else 0,
                   "grad_allreduce": ("none" if world == 1 else "cpc_allreduce_mc (NVSwitch multimem, in place)"
                                      if dp.symmetric is not None else "NCCL all-reduce (AVG)"),
                   **({"DIAGNOSTIC": "all-reduce skipped, not a valid number"}
                      if os.environ.get("CPC_DIAG_NO_ALLREDUCE") == "1" else {})},
        "e2e": {"value": e2e, "unit": "audio-s/s", "h2d_bytes_per_step": host_seq.numel() * 4,
                "d2h_bytes_per_step": host_out.numel() * 4, "ms_per_step": ms_e2e / opt.steps},
        "gpu_launches": launches_per_step * opt.steps, "gpu_launches_per_step": launches_per_step,
        "clocks": clocks.summary(), "loss_mean": loss_now,
    }
    if rank == 0 and world == 1:
        pk = peaks()
        rows, eager_ms, _ = kernel_profile(model, crit, step.static_seq, step.static_label)
        total = sum(r["ms"] for r in rows)
        top = rows[0]
        if top["kind"] == "tensor":
            ach = top["flops"] / (top["ms"] * 1e-3) / 1e12
            roof = {"bound": "tensor", "achieved": ach, "peak": pk["tensor"], "unit": "TFLOP/s", "frac": ach / pk["tensor"]}
        elif top["kind"] == "hbm":
            ach = top["bytes"] / (top["ms"] * 1e-3) / 1e9
            roof = {"bound": "hbm", "achieved": ach, "peak": pk["hbm"], "unit": "GB/s", "frac": ach / pk["hbm"]}
        else:  # latency-bound recurrence / CUDA-core attention: reported against the tensor peak as SURVEY 8d asks
            ach = top["flops"] / (top["ms"] * 1e-3) / 1e12
            roof = {"bound": "tensor", "achieved": ach, "peak": pk["tensor"], "unit": "TFLOP/s", "frac": ach / pk["tensor"],
                    "note": f"{top['kind']}-bound kernel; see DESIGN.md"}
        roof.update({"kernel": top["name"], "launches_per_step": top["launches"], "share_of_kernel_time": top["ms"] / total,
                     "avg_launch_us": 1e3 * top["ms"] / top["launches"], "traffic": traffic_of(top["name"]),
                     "traffic_unit": "bytes per launch (ncu dram read+write, profiles/traffic.json)",
                     "peak_source": pk["source"]})
        line["roofline"] = roof
        line["kernels"] = [
            {"name": r["name"], "launches": r["launches"], "us": round(1e3 * r["ms"], 1), "share": round(r["ms"] / total, 4),
             "tflops": round(r["flops"] / (r["ms"] * 1e-3) / 1e12, 1) if r["flops"] else None,
             "gbs": round(r["bytes"] / (r["ms"] * 1e-3) / 1e9, 1) if r["bytes"] else None, "kind": r["kind"]}
            for r in rows]
        line["kernel_time_sum_ms"] = total
        sec, threads = cpu_step_seconds(args, B, 2, 1)
        line["cpu_baseline"] = {"value": B * W / SAMPLE_RATE / sec, "unit": "audio-s/s", "cores": threads, "kind": "port",
                                "sample": f"2 full fwd+bwd steps of the same batch (8 x 20480 samples) after 1 warm-up, "
                                          f"oracle/cpc_oracle.py fp32, {threads} threads, {sec:.2f} s/step"}
    if rank == 0:
        out.emit(json.dumps(line))
    if world > 1:
        # The step's CUDA graph holds captured NCCL launches; tearing the communicator down while the graph is
        # alive blocks inside ncclCommDestroy (observed at N=2: JSON printed, process never exited).  All ranks
        # meet once more, the graph is dropped, and the process leaves without the communicator teardown.
        barrier()
        del step
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


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
    opt = ap.parse_args()
    out = StdoutToStderr()
    if opt.impl == "reference":
        run_reference(opt, out)
    else:
        run_ours(opt, out)


if __name__ == "__main__":
    main()
