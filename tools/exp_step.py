#!/usr/bin/env python
"""experiment helper: resident step time of one config with the library given in MADPOSE_B200_LIB
python tools/exp_step.py [config] [pairs ...]"""
import os, sys
import numpy as np
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madpose_b200 import capi, workload
config = int(sys.argv[1]) if len(sys.argv) > 1 else 1
sizes = [int(v) for v in sys.argv[2:]] or [workload.CONFIG_INFO[config]["pairs"]]
ctx = capi.Context(0)
if os.environ.get("MP_SET_BATCH"): ctx.set_batch(int(os.environ["MP_SET_BATCH"]))
for n in sizes:
    kind, pairs, opt, cfg, keep = workload.make_batch(config, n)
    ctx.upload_batch(kind, pairs, opt)
    ms = []
    for _ in range(6):
        ctx.estimate_resident(opt, cfg)
        ms.append(ctx.last_timing()[1])
    pc = ctx.last_phase_cycles()
    t = float(np.median(ms[2:]))
    print(os.environ.get("MADPOSE_B200_LIB", "default"), "config", config, "pairs", n, "ms/step", round(t, 1), "pairs/s", round(1e3 * n / t),
          {k: round(v / 1e9, 2) for k, v in pc.items()}, flush=True)
