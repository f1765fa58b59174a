#!/usr/bin/env python
"""Library-level multi-GPU throughput: ONE process, mp_estimate_mixed_multi / mp_estimate_batch_multi over 1..G devices
on a fixed batch (strong scaling inside the library; host buffers in, host buffers out, wall clock).
usage: bench_multi.py [config=4] [pairs]"""
import json, os, sys, time
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madpose_b200 import capi, workload
config = int(sys.argv[1]) if len(sys.argv) > 1 else 4
n = int(sys.argv[2]) if len(sys.argv) > 2 else workload.CONFIG_INFO[config]["pairs"]
lib = capi.load_library()
G = lib.mp_device_count()
if config == 4:
    mixed = workload.make_mixed_batch(range(n))
    pairs, kinds, opts, cfg = [None] * n, [0] * n, {}, None
    for k, (ps, opt, c, keep, idx) in mixed.items():
        opts[k] = opt; cfg = c
        for p, i in zip(ps, idx):
            pairs[i] = p; kinds[i] = k
else:
    kind, pairs, opt, cfg, keep = workload.make_batch(config, n)
out = {"workload": workload.CONFIG_INFO[config]["name"], "pairs": n, "devices_visible": G, "runs": []}
g = 1
while g <= G:
    devs = list(range(g))
    for rep in range(3):
        t0 = time.perf_counter()
        if config == 4:
            capi.estimate_batch_multi(devs, 0, pairs, None, cfg, lib=lib, kinds=kinds, opt_by_kind=opts)
        else:
            capi.estimate_batch_multi(devs, kind, pairs, opt, cfg, lib=lib)
        dt = time.perf_counter() - t0
    out["runs"].append({"devices": g, "seconds": dt, "pairs_per_s": n / dt})
    g *= 2
print(json.dumps(out))
