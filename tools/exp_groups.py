"""Experiment: step time of configs[1] versus the number of stream groups."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from madpose_b200 import capi
kind, pairs, opt, cfg, keep = bench.make_batch(1500, 0)
ctx = capi.Context(0)
ctx.upload_batch(kind, pairs, opt)
B = int(os.environ.get("MP_BATCH", "0"))
print("MP_LO_THREADS", os.environ.get("MP_LO_THREADS"))
ctx.set_batch(B)
for G in [int(x) for x in sys.argv[1:]] or [16, 32, 64, 128, 256]:
    ctx.set_groups(G)
    ctx.estimate_resident(opt, cfg)
    ts = []
    for _ in range(3):
        t0 = time.perf_counter(); ctx.estimate_resident(opt, cfg); wall = time.perf_counter() - t0
        ts.append((ctx.last_timing()[1], wall * 1e3))
    print("batch", B, "groups", G, "device ms", [round(t[0], 1) for t in ts], "wall ms", [round(t[1], 1) for t in ts], "launches", ctx.last_timing()[2], flush=True)
