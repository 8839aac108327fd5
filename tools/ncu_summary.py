"""Per-launch summary of an ncu report (--set full): duration, DRAM bytes, tensor-pipe and DRAM utilisation.
Usage: python tools/ncu_summary.py report.ncu-rep > profiles/<name>.txt"""
import csv
import subprocess
import sys

rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h = rows[0]
ix = {n: i for i, n in enumerate(h)}
units = rows[1]


def col(r, name, scale=1.0):
    i = ix.get(name)
    if i is None or r[i] == "":
        return float("nan")
    return float(r[i].replace(",", "")) * scale


def to_bytes(r, name):
    u = units[ix[name]]
    k = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    return col(r, name) * k


def to_us(r, name):
    u = units[ix[name]]
    k = {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(u, 1e-3)
    return col(r, name) * k


print(f"# {rep}: per-launch (ncu --set full --clock-control none; cold-cache, serialised replays)")
print(f"# {'kernel':46s} {'grid':>6s} {'us':>8s} {'dramRd MB':>10s} {'dramWr MB':>10s} {'dram%':>6s} {'tensor%':>8s} {'regs':>5s} {'smemKB':>7s}")
tot_us = tot_b = 0.0
for r in rows[2:]:
    if len(r) < len(h):
        continue
    name = r[ix["Kernel Name"]][:46]
    us = to_us(r, "gpu__time_duration.sum")
    rd, wr = to_bytes(r, "dram__bytes_read.sum"), to_bytes(r, "dram__bytes_write.sum")
    tot_us += us
    tot_b += rd + wr
    tens = col(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active")
    if tens != tens:
        tens = col(r, "sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active")
    print(f"  {name:46s} {r[ix['launch__grid_size']]:>6s} {us:8.1f} {rd / 1e6:10.2f} {wr / 1e6:10.2f} "
          f"{col(r, 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'):6.1f} {tens:8.1f} "
          f"{r[ix['launch__registers_per_thread']]:>5s} "
          f"{col(r, 'launch__shared_mem_per_block_dynamic') / 1024 if 'launch__shared_mem_per_block_dynamic' in ix else float('nan'):7.1f}")
n = len(rows) - 2
print(f"# {n} launches, {tot_us:.1f} us, DRAM read+write {tot_b / 1e6:.1f} MB total = {tot_b / max(n, 1) / 1e6:.2f} MB per launch")
