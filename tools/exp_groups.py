#!/usr/bin/env python
"""experiment: step time of a config vs stream groups and speculation batch cap: exp_groups.py [config]"""
import os, sys
import numpy as np
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madpose_b200 import capi, workload
config = int(sys.argv[1]) if len(sys.argv) > 1 else 1
kind, pairs, opt, cfg, keep = workload.make_batch(config, workload.CONFIG_INFO[config]["pairs"])
ctx = capi.Context(0)
ctx.upload_batch(kind, pairs, opt)
for g, b in [(0, 0), (2, 0), (3, 0), (4, 0), (6, 0), (8, 0), (12, 0), (4, 32), (4, 64), (8, 32), (2, 32)]:
    ctx.set_groups(g); ctx.set_batch(b)
    ms = []
    for _ in range(4):
        ctx.estimate_resident(opt, cfg)
        ms.append(ctx.last_timing()[1])
    print("groups", g, "batch", b, "ms/step", round(float(np.median(ms[1:])), 1), "rounds", ctx.last_breakdown()[1])
