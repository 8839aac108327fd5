"""profiles/traffic.json from an `ncu --set full` capture of the shipped code: DRAM read + write bytes per launch,
averaged per C-ABI entry point (kernel family).

    python tools/traffic_from_ncu.py gpurun_out/r02_step_full.ncu-rep [more.ncu-rep ...] > profiles/traffic.json
bench.py reads the file for `roofline.traffic`."""
import csv
import json
import re
import subprocess
import sys

FAMILY = [("gemm_bf16_tcgen05", "cpc_gemm_bf16"), ("attn_fwd", "cpc_attn_fwd"), ("attn_bwd", "cpc_attn_bwd"),
          ("lstm_fwd", "cpc_lstm_fwd"), ("lstm_bwd", "cpc_lstm_bwd"), ("rownorm_fwd", "cpc_rownorm_fwd"),
          ("rownorm_bwd", "cpc_rownorm_bwd"), ("conv0_cn_relu_fwd", "cpc_conv0_cn_relu_fwd"),
          ("conv0_cn_relu_bwd", "cpc_conv0_cn_relu_bwd"), ("nce_fwd", "cpc_nce_fwd"), ("nce_bwd", "cpc_nce_bwd"),
          ("adam_kernel", "cpc_adam_step"), ("cast_f32_bf16", "cpc_cast_f32_to_bf16"), ("colsum", "cpc_colsum"),
          ("conv_weight_pack", "cpc_conv_weight_pack"), ("conv_wgrad_unpack", "cpc_conv_wgrad_unpack")]
acc = {}
for rep in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, units = rows[0], rows[1]
    ix = {n: i for i, n in enumerate(h)}

    def to_bytes(r, name):
        k = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(units[ix[name]], 1)
        return float(r[ix[name]].replace(",", "")) * k

    for r in rows[2:]:
        if len(r) < len(h):
            continue
        name = r[ix["Kernel Name"]]
        fam = next((f for pat, f in FAMILY if re.search(pat, name)), None)
        if fam is None:
            continue
        a = acc.setdefault(fam, dict(bytes=0.0, n=0, src=set()))
        a["bytes"] += to_bytes(r, "dram__bytes_read.sum") + to_bytes(r, "dram__bytes_write.sum")
        a["n"] += 1
        a["src"].add(rep)
res = {"_comment": "DRAM read+write bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum), averaged over the "
                   "launches of one training step, from ncu --set full captures of the shipped code "
                   "(tools/traffic_from_ncu.py); read by bench.py for roofline.traffic"}
for fam, a in sorted(acc.items()):
    res[fam] = {"bytes_per_launch": round(a["bytes"] / a["n"]), "launches": a["n"],
                "source": ", ".join(sorted(a["src"]))}
print(json.dumps(res, indent=1))
