#!/usr/bin/env python
"""Summarise an .ncu-rep per source line: `python tools/ncu_lines.py report.ncu-rep [top]`.
Uses `ncu --page source --print-source cuda,sass --csv` (needs -lineinfo and --import-source on)."""
import collections
import csv
import subprocess
import sys


def load(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                         capture_output=True, text=True).stdout
    cur, hdr, rows = None, None, []
    for r in csv.reader(out.splitlines()):
        if len(r) >= 2 and r[0] == "File Path":
            cur = r[1]
        elif r and r[0] == "Line No":
            hdr = r
        elif hdr and len(r) == len(hdr) and r[0] not in ("", "Line No"):
            rows.append((cur, r))
    return hdr, rows


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    hdr, rows = load(rep)
    ci = {n: i for i, n in enumerate(hdr)}
    i_s, i_inst, i_thr = ci["# Samples"], ci["Instructions Executed"], ci["Thread Instructions Executed"]
    tot_s = sum(int(r[i_s]) for _, r in rows) or 1
    tot_i = sum(int(r[i_inst]) for _, r in rows) or 1
    per_file = collections.defaultdict(lambda: [0, 0, 0])
    for f, r in rows:
        a = per_file[f.split("/")[-1]]
        a[0] += int(r[i_s]); a[1] += int(r[i_inst]); a[2] += int(r[i_thr])
    print(f"total samples {tot_s}, warp instructions {tot_i}")
    print("--- per file: samples share | inst share | threads/inst")
    for f, a in sorted(per_file.items(), key=lambda x: -x[1][0]):
        print(f"{f:28s} {a[0] / tot_s:6.3f} {a[1] / tot_i:6.3f} {a[2] / max(a[1], 1):6.1f}")
    print(f"--- top {top} lines by samples")
    stall_cols = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
    for f, r in sorted(rows, key=lambda x: -int(x[1][i_s]))[:top]:
        st = sorted(((int(r[ci[n]]), n[6:]) for n in stall_cols), reverse=True)[:2]
        sts = " ".join(f"{n}:{v}" for v, n in st if v)
        print(f"{int(r[i_s]) / tot_s:6.3f} {int(r[i_inst]) / tot_i:6.3f} thr/inst {int(r[i_thr]) / max(int(r[i_inst]), 1):5.1f} "
              f"{f.split('/')[-1]}:{r[0]:>4s} [{sts}] {r[1].strip()[:90]}")


if __name__ == "__main__":
    main()
