#!/usr/bin/env python
"""Golden vectors for mp_pose_errors from the reference's own madpose.utils.compute_pose_error (madpose/utils.py:59-78).
Run in the build container:  python tests/golden/make_pose_error_golden.py  -> tests/golden/pose_errors.npz"""
import importlib.util
import os

import numpy as np

spec = importlib.util.spec_from_file_location("ref_utils", "/root/reference/madpose/utils.py")
ref_utils = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref_utils)

rng = np.random.Generator(np.random.PCG64(20260102))


def rot(axis, deg):
    a = axis / np.linalg.norm(axis)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    th = np.deg2rad(deg)
    return np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)


n = 256
T = np.zeros((n, 4, 4)); R = np.zeros((n, 3, 3)); t = np.zeros((n, 3))
for i in range(n):
    Rg = rot(rng.normal(size=3), rng.uniform(0, 180))
    tg = rng.normal(size=3) * rng.choice([1e-3, 0.1, 1.0, 5.0])
    T[i] = np.eye(4); T[i, :3, :3] = Rg; T[i, :3, 3] = tg
    dR = rot(rng.normal(size=3), rng.choice([0.0, 1e-4, 0.5, 5.0, 60.0, 180.0]))
    R[i] = dR @ Rg
    t[i] = rot(rng.normal(size=3), rng.choice([0.0, 1e-3, 2.0, 45.0, 120.0, 179.0])) @ tg * rng.uniform(0.2, 3.0)
t[:4] = -T[:4, :3, 3]           # exactly opposite translation: folded to 0
R[4:8] = T[4:8, :3, :3]          # identical rotation
t[8] = 0.0                       # zero-length estimate: the reference returns NaN for error_t
out = {"T": T, "R": R, "t": t}
for name, thr in (("none", None), ("thr", 0.05)):
    e = np.array([ref_utils.compute_pose_error(T[i], R[i], t[i], thr) for i in range(n)], np.float64)
    out["err_t_" + name], out["err_R_" + name] = e[:, 0], e[:, 1]
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "pose_errors.npz")
np.savez_compressed(path, **out)
print(path, os.path.getsize(path), "bytes", "nan:", int(np.isnan(out["err_t_none"]).sum()))
