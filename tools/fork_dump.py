#!/usr/bin/env python
"""Parity forensics: both event traces of one pair around the first difference.
usage: python tools/fork_dump.py <config> <pair> [context=6]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
from madpose_b200 import capi  # noqa: E402
from madpose_b200.workload import problem  # noqa: E402
from oracle import oracle_lib  # noqa: E402
from tests.test_estimate_gpu import _cnt3, _parse_device_trace, _parse_oracle_trace  # noqa: E402


def main():
    config, p = int(sys.argv[1]), int(sys.argv[2])
    ctxn = int(sys.argv[3]) if len(sys.argv) > 3 else 6
    d, pair, opt, cfg, keep = problem(config, p=p, n=int(os.environ["MP_N"]) if "MP_N" in os.environ else None)
    kind = d["kind"]
    ora = oracle_lib.load(build=False)
    _, so, _, tr = ora.estimate_batch(kind, [pair], opt, cfg, trace_cap=1 << 22)
    eo = [e for e in _parse_oracle_trace(tr)]
    ctx = capi.Context(0)
    ctx.set_minimal_mode(int(os.environ.get("MP_MIN_MODE", "0")))
    ctx.set_trace(1 << 17)
    _, sg, _ = ctx.estimate_batch(kind, [pair], opt, cfg)
    raw, n = ctx.read_trace(0, cap=1 << 17)
    ed = _parse_device_trace(raw)
    print("oracle iterations", so[0].num_iterations_total, "LO", so[0].number_lo_iterations, "| device", sg[0].num_iterations_total, sg[0].number_lo_iterations)
    fmt = lambda e: f"it {e[0]:5d} kind {e[1]:3d} a {(_cnt3(e[2]) if e[1] == -2 else e[2])} b {e[3]!r}"
    oi = 0
    for di, (it, knd, a, b) in enumerate(ed):
        while oi < len(eo) and eo[oi][1] >= 0 and not (knd >= 0 and eo[oi][0] == it):
            oi += 1
        if oi == len(eo):
            print("device continues past the oracle"); break
        o = eo[oi]
        same = (o[1] == knd or (knd >= 0 and o[1] >= 0)) and (a == o[2] or knd >= 0) and (b == o[3] or abs(b - o[3]) <= 1e-9 * abs(o[3]) or knd == -2)
        if not same:
            print("first difference at device event", di, "oracle event", oi)
            for k in range(max(0, di - ctxn), min(len(ed), di + 3)):
                print("  dev", k, fmt(ed[k]))
            for k in range(max(0, oi - ctxn), min(len(eo), oi + 3)):
                if eo[k][1] < 0 or abs(k - oi) < 3:
                    print("  ora", k, fmt(eo[k]))
            return
        oi += 1
    print("no difference in the traced events; oracle events left:", len(eo) - oi)


if __name__ == "__main__":
    main()
