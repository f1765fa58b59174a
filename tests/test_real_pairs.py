"""KAT-4: the reference's own example pairs, end to end with the example scripts' settings
(examples/shared_focal.py:12-80 on image_pairs/1_eth3d, examples/two_focal.py:12-80 on image_pairs/2_2d3ds), against
the ground-truth pose and intrinsics of info.json.  Inputs: tests/golden/real_pairs.npz, built from the reference's
data by the reference's own get_depths (tests/golden/make_real_pairs_golden.py).  Not a reference output (the reference
cannot be built here) -- a sanity bound on real matcher / depth-network data for the oracle and for the GPU path."""
import os

import numpy as np
import pytest

from madpose_b200 import capi

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "real_pairs.npz"))
CASES = [("eth3d", 1, (64.0, 4.0)), ("s2d3ds", 2, (256.0, 1.0))]
# error bounds against info.json (degrees, degrees, relative focal): the hybrid estimator is far inside them
BOUNDS = {"eth3d": (2.0, 5.0, 0.10), "s2d3ds": (3.0, 6.0, 0.15)}


def setup(key, kind, thr):
    g = {k[len(key) + 1:]: GOLD[k] for k in GOLD.files if k.startswith(key + "_")}
    opt = capi.Options(min_num_iterations=100, max_num_iterations=1000, final_least_squares=1, threshold_multiplier=5.0,
                       num_lo_steps=4, squared_inlier_thresholds=thr, data_type_weights=(1.0, 1.0), random_seed=0)
    cfg = capi.Config()
    pair, keep = capi.make_pair(g["mkpts0"], g["mkpts1"], g["depth0"], g["depth1"], g["min_depth"], pp0=g["pp0"], pp1=g["pp1"])
    return g, pair, opt, cfg, keep


def pose_errors(T, R, t):
    """madpose/utils.py:59-78 (compute_pose_error)"""
    R_gt, t_gt = T[:3, :3], T[:3, 3]
    c = np.clip(t @ t_gt / (np.linalg.norm(t) * np.linalg.norm(t_gt)), -1, 1)
    et = np.degrees(np.arccos(c)); et = min(et, 180 - et)
    er = np.degrees(abs(np.arccos(np.clip((np.trace(R.T @ R_gt) - 1) / 2, -1, 1))))
    return et, er


def check(key, kind, g, m):
    R, t = m[:9].reshape(3, 3), m[9:12]
    et, er = pose_errors(g["T_0to1"], R, t)
    if kind == 1:
        f_mean = np.mean([g["K0"][0, 0], g["K0"][1, 1], g["K1"][0, 0], g["K1"][1, 1]])   # examples/shared_focal.py:89-90
        ef = abs(m[15] - f_mean) / f_mean
    else:
        ef = max(abs(m[15] - g["K0"][0, 0]) / g["K0"][0, 0], abs(m[16] - g["K1"][0, 0]) / g["K1"][0, 0])   # two_focal.py:93
    b = BOUNDS[key]
    assert er < b[0] and et < b[1] and ef < b[2], (key, er, et, ef)
    assert abs(np.linalg.det(R) - 1) < 1e-9 and m[12] > 0
    return er, et, ef


@pytest.mark.parametrize("key,kind,thr", CASES)
def test_oracle_on_real_pairs(oracle, key, kind, thr):
    g, pair, opt, cfg, keep = setup(key, kind, thr)
    m, s, inl = oracle.estimate_batch(kind, [pair], opt, cfg)
    check(key, kind, g, m[0])
    assert s[0].best_num_inliers > 0.3 * 3 * pair.n


@pytest.mark.gpu
@pytest.mark.parametrize("key,kind,thr", CASES)
def test_gpu_on_real_pairs(oracle, key, kind, thr):
    """Through the drop-in Python API, exactly as the example scripts call it; same result as the oracle."""
    import madpose_b200 as madpose
    g, pair, opt, cfg, keep = setup(key, kind, thr)
    o = madpose.HybridLORansacOptions()
    o.min_num_iterations, o.max_num_iterations, o.final_least_squares = 100, 1000, True
    o.threshold_multiplier, o.num_lo_steps, o.random_seed = 5.0, 4, 0
    o.squared_inlier_thresholds, o.data_type_weights = list(thr), [1.0, 1.0]
    est = madpose.EstimatorConfig()
    est.min_depth_constraint, est.use_shift, est.ceres_num_threads = True, True, 8
    fn = madpose.HybridEstimatePoseScaleOffsetSharedFocal if kind == 1 else madpose.HybridEstimatePoseScaleOffsetTwoFocal
    pose, stats = fn(g["mkpts0"], g["mkpts1"], g["depth0"], g["depth1"], list(g["min_depth"]), g["pp0"], g["pp1"], o, est)
    m = np.concatenate([pose.R().reshape(9), pose.t(), [pose.scale, pose.offset0, pose.offset1],
                        [pose.focal, pose.focal] if kind == 1 else [pose.focal0, pose.focal1]])
    check(key, kind, g, m)
    mo, so, io = oracle.estimate_batch(kind, [pair], opt, cfg)
    assert stats.num_iterations_total == so[0].num_iterations_total and stats.best_num_inliers == so[0].best_num_inliers
    np.testing.assert_allclose(m, mo[0], rtol=1e-6, atol=1e-8)


@pytest.mark.gpu
def test_get_depths_then_estimate_matches_stored_depths(ctx):
    """The step before the path on the device (mp_get_depths) feeds the same numbers the reference's get_depths produced:
    checked on a synthetic depth map sampled at the real keypoints (the real maps are not stored)."""
    g = {k[6:]: GOLD[k] for k in GOLD.files if k.startswith("eth3d_")}
    rng = np.random.default_rng(0)
    dm = rng.uniform(0.5, 8.0, tuple(g["depth0_shape"])).astype(np.float32)
    shape = tuple(int(v) for v in g["image0_shape"])
    resize = np.array([dm.shape[1] / shape[1], dm.shape[0] / shape[0]])
    sc = np.round(g["mkpts0"] * resize).astype(int)   # madpose/utils.py:4-22
    sc[:, 0] = np.clip(sc[:, 0], 0, dm.shape[1] - 1); sc[:, 1] = np.clip(sc[:, 1], 0, dm.shape[0] - 1)
    np.testing.assert_array_equal(ctx.get_depths(shape, dm, g["mkpts0"]), dm[sc[:, 1], sc[:, 0]])
