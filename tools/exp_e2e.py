#!/usr/bin/env python
"""experiment: where the end-to-end time of mp_estimate_batch on host buffers goes (configs[1], 1500 pairs)"""
import os, sys, time
import numpy as np
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
from madpose_b200 import capi, workload
kind, pairs, opt, cfg, keep = workload.make_batch(1, 1500)
arr = (capi.Pair * len(pairs))(*pairs)
ctx = capi.Context(0)
for _ in range(2): ctx.estimate_batch(kind, arr, opt, cfg, split=False)
w, k, t, up = [], [], [], []
for _ in range(5):
    t0 = time.perf_counter(); ctx.estimate_batch(kind, arr, opt, cfg, split=False); w.append((time.perf_counter() - t0) * 1e3)
    a, b, _n = ctx.last_timing(); k.append(a); t.append(b)
    t0 = time.perf_counter(); ctx.lib.mp_upload_batch(ctx.ctx, kind, len(pairs), arr, C.byref(opt)); ctx.score_batch  # upload only (async tail)
    import torch; torch.cuda.synchronize(); up.append((time.perf_counter() - t0) * 1e3)
m = lambda v: round(float(np.median(v)), 2)
print("wall e2e", m(w), "ms | device loop (ev1-ev2)", m(k), "| stream total (ev0-ev3)", m(t), "| upload alone (pack + H2D + front-end, synced)", m(up))
