"""Aggregates the warp-stall samples of an ncu report's source page by SASS opcode, per kernel.
Usage: python tools/ncu_ops.py report.ncu-rep [top_n_lines]"""
import collections
import csv
import re
import subprocess
import sys

rep = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
blocks, i = [], 0
while i < len(rows):
    if rows[i] and rows[i][0] == "Kernel Name":
        name, hdr, j, body = rows[i][1], rows[i + 1], i + 2, []
        while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
            body.append(rows[j])
            j += 1
        blocks.append((name, hdr, body))
        i = j
    else:
        i += 1
for name, h, body in blocks:
    si, src, ie = h.index("Warp Stall Sampling (All Samples)"), h.index("Source"), h.index("Instructions Executed")
    body = [x for x in body if len(x) > max(si, ie)]
    tot = sum(int(x[si] or 0) for x in body) or 1
    ti = sum(int(x[ie] or 0) for x in body) or 1
    print(f"== {name[:90]}  samples {tot}  warp-instr {ti}  sass lines {len(body)}")
    agg, cnt = collections.Counter(), collections.Counter()
    for x in body:
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", x[src])
        op = m.group(2).split(".")[0] if m else "?"
        agg[op] += int(x[si] or 0)
        cnt[op] += int(x[ie] or 0)
    for op, s in agg.most_common(16):
        print(f"   {op:10s} samples {s:6d} ({100 * s / tot:4.1f}%)  executed {cnt[op]:9d} ({100 * cnt[op] / ti:4.1f}%)")
    if topn:
        for x in sorted(body, key=lambda x: -int(x[si] or 0))[:topn]:
            print(f"      {x[si]:>6s} {x[ie]:>8s}  {x[src][:100]}")
