#!/usr/bin/env python
"""experiment: step time of configs[1] vs the iteration from which the warp solver takes over"""
import os, sys
import numpy as np
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madpose_b200 import capi, workload
kind, pairs, opt, cfg, keep = workload.make_batch(1, 1500)
ctx = capi.Context(0)
ctx.upload_batch(kind, pairs, opt)
for w in [int(v) for v in sys.argv[1:]] or [100000, 512, 384, 256, 128, 0]:
    ctx.set_warp_from_iteration(w)
    ms = []
    for _ in range(5):
        ctx.estimate_resident(opt, cfg)
        ms.append(ctx.last_timing()[1])
    print(w, "ms/step", round(float(np.median(ms[2:])), 1), "rounds", ctx.last_breakdown()[1])
