"""Family totals of an ncu launch list (--metrics gpu__time_duration.sum --csv).
Usage: python tools/ncu_launches.py launches.csv "command line" > profiles/<name>.txt"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit()]
tot = collections.Counter()
cnt = collections.Counter()
for r in rows:
    name = re.sub(r"\(.*$", "", r[4])
    ns = float(r[14].replace(",", ""))
    tot[name] += ns
    cnt[name] += 1
total = sum(tot.values())
print(f"# {sys.argv[2] if len(sys.argv) > 2 else ''}")
print(f"# total {total / 1e3:.1f} us over {len(rows)} launches; share = kernel family time / total")
for name, ns in tot.most_common():
    print(f"{cnt[name]:5d} {ns / 1e3:10.1f} us {100 * ns / total:5.1f}%  {name[:110]}")
