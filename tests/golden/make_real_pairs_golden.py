"""Generates tests/golden/real_pairs.npz from the reference's own example data and its own helper:
examples/image_pairs/1_eth3d (shared-focal script, examples/shared_focal.py:36-90) and 2_2d3ds (two-focal script,
examples/two_focal.py:36-90).  Keypoints are loaded as the scripts load them, the prior depths are looked up by
madpose.utils.get_depths ITSELF (imported from /root/reference/madpose/utils.py), min depths / principal points are
computed as the scripts compute them.  Only these small per-match arrays are stored (the 2.7 MB depth maps are not);
the GT pose and intrinsics come from info.json.  Run in the build container only; the npz is committed."""
import importlib.util
import json
import os

import cv2
import numpy as np

REF = "/root/reference"
spec = importlib.util.spec_from_file_location("ref_utils", os.path.join(REF, "madpose", "utils.py"))
ref_utils = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref_utils)


def load(name):
    path = os.path.join(REF, "examples", "image_pairs", name)
    info = json.load(open(os.path.join(path, "info.json")))
    image0 = cv2.imread(os.path.join(path, "image0.png")); image1 = cv2.imread(os.path.join(path, "image1.png"))
    mk0 = np.load(os.path.join(path, info["matches_0_file"])); mk1 = np.load(os.path.join(path, info["matches_1_file"]))
    dm0 = np.load(os.path.join(path, info["depth_0_file"])); dm1 = np.load(os.path.join(path, info["depth_1_file"]))
    d0 = ref_utils.get_depths(image0, dm0, mk0); d1 = ref_utils.get_depths(image1, dm1, mk1)
    pp0 = (np.array(image0.shape[:2][::-1]) - 1) / 2; pp1 = (np.array(image1.shape[:2][::-1]) - 1) / 2
    return dict(mkpts0=mk0.astype(np.float64), mkpts1=mk1.astype(np.float64), depth0=d0.astype(np.float64),
                depth1=d1.astype(np.float64), min_depth=np.array([dm0.min(), dm1.min()], np.float64), pp0=pp0, pp1=pp1,
                K0=np.array(info["K0"]), K1=np.array(info["K1"]), T_0to1=np.array(info["T_0to1"]),
                image0_shape=np.array(image0.shape), image1_shape=np.array(image1.shape),
                depth0_shape=np.array(dm0.shape), depth1_shape=np.array(dm1.shape))


out = {}
for key, name in (("eth3d", "1_eth3d"), ("s2d3ds", "2_2d3ds")):
    for k, v in load(name).items():
        out[f"{key}_{k}"] = v
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "real_pairs.npz")
np.savez_compressed(path, **out)
print("wrote", path, {k: v.shape for k, v in out.items() if k.endswith("mkpts0")})
