#!/usr/bin/env python
"""Golden vectors for mp_get_depths from the reference's own madpose.utils.get_depths (madpose/utils.py:4-22),
loaded straight from the file (the package itself needs the compiled extension).  Run in the build container:
    python tests/golden/make_get_depths_golden.py      -> tests/golden/get_depths.npz"""
import importlib.util
import os

import numpy as np

REF = "/root/reference/madpose/utils.py"
spec = importlib.util.spec_from_file_location("ref_utils", REF)
ref_utils = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref_utils)

rng = np.random.Generator(np.random.PCG64(20260101))
cases = {}
for (ih, iw, mh, mw, dt, n) in [(120, 160, 120, 160, "float32", 400), (480, 640, 48, 64, "float32", 400),
                                (480, 640, 74, 100, "float64", 400), (1080, 1920, 37, 53, "float32", 400)]:
    depth = rng.uniform(0.2, 9.0, (mh, mw)).astype(dt)
    kp = rng.uniform(-20.0, 1.05 * max(ih, iw), (n, 2))
    # exact ties of the rounding (x * resize = k + 0.5) and the borders
    kp[:40, 0] = (np.arange(40) + 0.5) * iw / mw
    kp[40:80, 1] = (np.arange(40) + 0.5) * ih / mh
    kp[80:84] = [[0, 0], [iw - 1, ih - 1], [iw, ih], [-0.5, -0.5]]
    image = np.zeros((ih, iw, 3), np.uint8)
    out = ref_utils.get_depths(image, depth, kp)
    k = "c%d_" % (len(cases) // 5)
    cases.update({k + "image_shape": np.array([ih, iw]), k + "depth": depth, k + "mkpts": kp,
                  k + "expect": np.asarray(out, np.float64), k + "source": np.array("madpose/utils.py:4-22")})
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "get_depths.npz")
np.savez_compressed(path, **cases)
print(path, os.path.getsize(path), "bytes")
