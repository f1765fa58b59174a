"""Experiment: throughput of the resident path versus the number of pairs per step (is the step bound by the longest
pair's chain of rounds or by the SMs?)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from madpose_b200 import capi
ctx = capi.Context(0)
for P in [int(x) for x in sys.argv[1:]] or [188, 375, 750, 1500, 3000, 6000]:
    kind, pairs, opt, cfg, keep = bench.make_batch(P, 0)
    ctx.upload_batch(kind, pairs, opt)
    ctx.estimate_resident(opt, cfg)
    ts = []
    for _ in range(3):
        ctx.estimate_resident(opt, cfg)
        ts.append(ctx.last_timing()[1])
    t = min(ts)
    print("pairs", P, "device ms", [round(x, 1) for x in ts], "pairs/s", round(P / t * 1e3), "launches", ctx.last_timing()[2], flush=True)
