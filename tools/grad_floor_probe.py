"""The bf16-storage noise floor of every golden fixture, per parameter tensor (CPU; run in the build container):

    python tools/grad_floor_probe.py        -> tests/golden/grad_floor.json

For each fixture the oracle is evaluated with `oracle.cpc_oracle.EMULATE_BF16 = True` -- the exact fp32 algorithm
except that every GEMM operand (weights, activations between layers) and every gradient that travels between layers
is rounded to bf16, which ANY bf16-operand implementation of the path has to do -- and the cosine between its
parameter gradients and the reference's fp32 gradients (the fixture) is recorded.  tests/test_step_gpu.py requires
cosine >= min(0.99, floor - 0.01) per tensor: the stated tolerance (0.99) wherever bf16 storage permits it, and the
probe-derived floor -- not a number typed into the test -- where it does not (the tiny 32-channel fixtures at random
initialisation, whose encoder gradients are sums of almost exactly cancelling terms).
"""
import json
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import cpc_oracle as oc  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def floors_of(fx):
    a = fx["args"]
    state = {k: v.float() for k, v in fx["state"].items()}
    oc.EMULATE_BF16 = True
    try:
        out = oc.train_step(state, fx["sequence"], fx["batch_idx"].long(), fx["seq_idx"].long(),
                            K=a["nPredicts"], NEG=a["negativeSamplingExt"], ar_mode=a["arMode"],
                            n_layers=a["nLevelsGRU"], rnn_mode=a["rnnMode"], cpc_mode=a["cpc_mode"],
                            abspos=a["abspos"], n_skipped=a["n_skipped"], multihead=a.get("multihead_rnn", False),
                            signal_quality=fx.get("signal_quality"), growth_rate=a.get("growth_rate"),
                            inflection_point_x=a.get("inflection_point_x"), norm_mode=a.get("normMode", "layerNorm"),
                            bn_training=fx.get("train_mode", False))
    finally:
        oc.EMULATE_BF16 = False
    res = {}
    strides = fx.get("grad_stride", {})
    for k, g in fx["grads"].items():
        st = strides.get(k, 1)
        got = out["grads"][k].flatten()[::st] if st > 1 else out["grads"][k]
        if float(g.abs().max()) == 0:
            continue
        res[k] = round(F.cosine_similarity(got.double().flatten(), g.double().flatten(), dim=0).item(), 5)
    return res


def main():
    torch.set_num_threads(os.cpu_count())
    table = {}
    only = sys.argv[1:]  # optional: fixture names to (re)compute, the rest of the table is kept
    if only:
        with open(os.path.join(GOLD, "grad_floor.json")) as fh:
            table = json.load(fh)
    for f in sorted(os.listdir(GOLD)):
        if not f.endswith(".pt") or (only and f[:-3] not in only):
            continue
        fx = torch.load(os.path.join(GOLD, f), weights_only=False)
        if "state" not in fx or "grads" not in fx:
            continue
        table[f[:-3]] = floors_of(fx)
        worst = min(table[f[:-3]].items(), key=lambda kv: kv[1])
        print(f"{f[:-3]:40s} min floor {worst[1]:.4f} ({worst[0]})", flush=True)
    with open(os.path.join(GOLD, "grad_floor.json"), "w") as fh:
        json.dump(table, fh, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
