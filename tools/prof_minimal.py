#!/usr/bin/env python
"""Stand-alone minimal-solver timing: `prof_minimal.py config solver n_samples` -> kernel ms, us per sample."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madpose_b200 import capi
from madpose_b200.workload import problem
config, solver = int(sys.argv[1]), int(sys.argv[2])
count = int(sys.argv[3]) if len(sys.argv) > 3 else 40000
K = {0: (3, 5), 1: (4, 6), 2: (4, 7)}
d, pair, opt, cfg, keep = problem(config, p=5)
kind = d["kind"]
rng = np.random.default_rng(1)
inl = np.nonzero(d["gt"]["inlier_mask"])[0]
S = np.array([rng.choice(inl if i % 2 == 0 else pair.n, K[kind][solver], replace=False) for i in range(count)], np.int32)
ctx = capi.Context(0)
ms = []
for _ in range(4):
    got = ctx.minimal_solve_batch(kind, pair, opt, cfg, solver, S)
    ms.append(ctx.last_timing()[0])
print({"config": config, "solver": solver, "samples": count, "kernel_ms": min(ms), "ns_per_sample": min(ms) * 1e6 / count,
       "models": int(sum(len(g) for g in got))})
