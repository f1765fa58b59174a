"""Host-side timeline of one resident step (MP_TRACE=1)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from madpose_b200 import capi
kind, pairs, opt, cfg, keep = bench.make_batch(1500, 0)
ctx = capi.Context(0)
ctx.upload_batch(kind, pairs, opt)
ctx.estimate_resident(opt, cfg)
os.environ["MP_TRACE"] = "1"
ctx.estimate_resident(opt, cfg)
