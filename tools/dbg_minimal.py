#!/usr/bin/env python
"""debug: per-sample comparison of mp_minimal_solve_batch with the oracle. usage: dbg_minimal.py config solver [count]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madpose_b200 import capi
from madpose_b200.workload import problem
from oracle import oracle_lib
config, solver = int(sys.argv[1]), int(sys.argv[2])
count = int(sys.argv[3]) if len(sys.argv) > 3 else 400
K = {0: (3, 5), 1: (4, 6), 2: (4, 7)}
d, pair, opt, cfg, keep = problem(config, p=5)
kind = d["kind"]
rng = np.random.default_rng(100 * config + solver)
inl = np.nonzero(d["gt"]["inlier_mask"])[0]
S = np.array([rng.choice(inl if i % 2 == 0 else pair.n, K[kind][solver], replace=False) for i in range(count)], np.int32)
ora = oracle_lib.load(build=False)
ctx = capi.Context(0)
ref = ora.minimal_solve(kind, pair, opt, cfg, solver, S)
got = ctx.minimal_solve_batch(kind, pair, opt, cfg, solver, S)
np.set_printoptions(precision=6, suppress=True, linewidth=200)
bad = 0
for i, (a, b) in enumerate(zip(ref, got)):
    if len(a) != len(b):
        bad += 1
        if bad <= 8:
            print("sample", i, "oracle", len(a), "device", len(b))
            print("  oracle f:", np.array([m[15] for m in a]), " scale:", np.array([m[12] for m in a]))
            print("  device f:", np.array([m[15] for m in b]), " scale:", np.array([m[12] for m in b]))
print("count mismatches", bad, "of", count)
