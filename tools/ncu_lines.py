#!/usr/bin/env python
"""Per-source-line and per-function summary of an `ncu --page source --csv` export (SASS level) of one kernel, joined
with `nvdisasm -g -c` of the same cubin (the line table).  python tools/ncu_lines.py src.csv disasm.txt kernel_substr"""
import collections, csv, re, subprocess, sys
src, dis, kname = sys.argv[1:4]
lines = open(dis).read().split("\n")
sec = None; cur = None; func = None; table = []  # per instruction (in order): (file, line, func)
for l in lines:
    m = re.match(r"^//-+ \.text\.(\S+)", l)
    if m: sec = m.group(1); cur = None; func = sec; continue
    if sec is None or kname not in sec: continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r"^(\S+):\s*$", l)
    if m and not m.group(1).startswith(".L"): func = m.group(1); continue
    if re.match(r"\s+/\*[0-9a-f]{4,6}\*/", l): table.append((cur, func))
rows = list(csv.reader(open(src)))
hdr = rows[1]; body = rows[2:]
ci = {h: i for i, h in enumerate(hdr)}
assert len(body) >= len(table), (len(body), len(table))
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
per_line = collections.defaultdict(lambda: [0, 0, collections.Counter()]); per_func = collections.defaultdict(lambda: [0, 0])
tot_s = tot_i = 0
for r, (cl, fn) in zip(body, table):
    s = int(r[ci["# Samples"]] or 0); n = int(r[ci["Instructions Executed"]] or 0)
    tot_s += s; tot_i += n
    e = per_line[cl]; e[0] += s; e[1] += n
    for h in stall_cols:
        v = int(r[ci[h]] or 0)
        if v: e[2][h[6:]] += v
    f = per_func[fn]; f[0] += s; f[1] += n
print("total samples", tot_s, "warp instructions", tot_i, "static instructions", len(table))
def dem(n):
    n = n.split("$")[-1] if "$" in n else n
    return subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()[:70]
print("--- per function: samples share | inst share")
for fn, (s, n) in sorted(per_func.items(), key=lambda kv: -kv[1][0])[:16]:
    print(f"{s / tot_s:6.3f} {n / tot_i:6.3f}  {dem(fn)}")
flt = sys.argv[4] if len(sys.argv) > 4 else None
print("--- top lines by samples" + (f" in {flt}" if flt else ""))
items = [(k, v) for k, v in per_line.items() if k and (not flt or k[0] == flt)]
for (f, ln), (s, n, st) in sorted(items, key=lambda kv: -kv[1][0])[:int(sys.argv[5]) if len(sys.argv) > 5 else 40]:
    top = " ".join(f"{k}:{v}" for k, v in st.most_common(3))
    print(f"{s / tot_s:6.3f} {n / tot_i:6.3f} {f}:{ln} [{top}]")
