"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel name launches and time."""
import collections
import csv
import re
import sys

rows = list(csv.reader(l for l in open(sys.argv[1]) if not l.startswith("==")))
hdr = rows[0]
col = {n: i for i, n in enumerate(hdr)}
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[1:]:
    if len(r) < len(hdr) or r[col["Metric Name"]] != "gpu__time_duration.sum":
        continue
    v = float(r[col["Metric Value"]].replace(",", ""))
    u = r[col["Metric Unit"]]
    us = v * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}.get(u, 1e-3)
    n = re.sub(r"^void ", "", r[col["Kernel Name"]])[:110]
    tot[n] += us
    cnt[n] += 1
T = sum(tot.values())
ours = sum(v for n, v in tot.items() if "slider::" in n)
print(f"# {sum(cnt.values())} launches, {T:.1f} us; libslider kernels {ours:.1f} us ({100 * ours / T:.1f} %)")
for n, v in tot.most_common(40):
    print(f"{v:10.1f} us {100 * v / T:5.1f}% x{cnt[n]:4d} {n}")
