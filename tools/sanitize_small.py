"""Small whole-path runs for compute-sanitizer (memcheck / racecheck): a few pairs of every estimator, two batch caps."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from madpose_b200 import capi
from tests.helpers import problem
npairs = int(sys.argv[1]) if len(sys.argv) > 1 else 4
n = int(sys.argv[2]) if len(sys.argv) > 2 else 300
ctx = capi.Context(0)
for config in (0, 1, 2):
    ps = [problem(config, p=i, n=n) for i in range(npairs)]
    kind, opt, cfg = ps[0][0]["kind"], ps[0][2], ps[0][3]
    for batch in (0, 256):
        ctx.set_batch(batch)
        m, s, inl = ctx.estimate_batch(kind, [q[1] for q in ps], opt, cfg)
        print("config", config, "batch", batch, "iterations", [x.num_iterations_total for x in s], "inliers", [x.best_num_inliers for x in s], flush=True)
z = np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "get_depths.npz"))
print("get_depths", ctx.get_depths([int(v) for v in z["c1_image_shape"]], z["c1_depth"], z["c1_mkpts"])[:3])
