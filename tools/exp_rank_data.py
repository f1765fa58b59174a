#!/usr/bin/env python
"""experiment: resident step time of configs[1] on ONE GPU for the pair ranges the ranks of a weak-scaling run own
(rank r: pairs [1500 r, 1500 (r + 1))) -- separates data-dependent step time from multi-process effects"""
import os, sys
import numpy as np
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madpose_b200 import capi, workload
ctx = capi.Context(0)
for r in [int(v) for v in sys.argv[1:]] or range(8):
    kind, pairs, opt, cfg, keep = workload.make_batch(1, 1500, first=1500 * r)
    ctx.upload_batch(kind, pairs, opt)
    ms = []
    for _ in range(5):
        st = ctx.estimate_resident(opt, cfg)[1]
        ms.append(ctx.last_timing()[1])
    print("rank", r, "ms/step", round(float(np.median(ms[2:])), 1), "matches", sum(p.n for p in pairs),
          "iterations", sum(s.num_iterations_total for s in st), "LO", sum(s.number_lo_iterations for s in st), flush=True)
