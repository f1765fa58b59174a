"""Host-side timeline of one resident step (MP_TRACE=1): python tools/exp_trace.py [config] [pairs]"""
import sys, os
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madpose_b200 import capi, workload
config = int(sys.argv[1]) if len(sys.argv) > 1 else 1
n = int(sys.argv[2]) if len(sys.argv) > 2 else workload.CONFIG_INFO[config]["pairs"]
kind, pairs, opt, cfg, keep = workload.make_batch(config, n)
ctx = capi.Context(0)
ctx.upload_batch(kind, pairs, opt)
ctx.estimate_resident(opt, cfg)
ctx.estimate_resident(opt, cfg)
os.environ["MP_TRACE"] = "1"
ctx.estimate_resident(opt, cfg)
