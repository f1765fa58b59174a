#!/usr/bin/env python
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list: python tools/launch_summary.py file.csv"""
import collections, csv, re, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]; ci = {h: i for i, h in enumerate(hdr)}
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    if r[ci["Metric Name"]] != "gpu__time_duration.sum": continue
    name = re.sub(r"\(.*", "", r[ci["Kernel Name"]]).replace("void mp::", "")
    v = float(r[ci["Metric Value"]].replace(",", "")); u = r[ci["Metric Unit"]]
    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1e-6)
    a = agg[name]; a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
print(f"{'kernel':60s} {'n':>5s} {'ms':>9s} {'share':>7s} {'avg us':>8s}")
for k, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:60]:60s} {n:5d} {ms:9.2f} {100 * ms / tot:6.1f}% {1e3 * ms / n:8.1f}")
