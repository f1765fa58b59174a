"""The drop-in Python module (madpose_b200/_madpose, csrc/bindings.cpp): same names, fields and call signatures as
the reference's pybind11 module (/root/reference/src/bindings.cpp:33-175).  The GPU part mirrors the reference's
example scripts (examples/calibrated.py:20-74, shared_focal.py, two_focal.py) on synthetic pairs and checks the
result against the oracle."""
import numpy as np
import pytest

import madpose_b200 as madpose
from madpose_b200 import synth


def _options(config):
    o = madpose.HybridLORansacOptions()
    for k, v in synth.options_for(config).items():
        setattr(o, k, list(v) if isinstance(v, tuple) else v)
    return o


def test_api_surface_matches_reference():
    assert madpose.HAVE_BINDINGS
    o = madpose.HybridLORansacOptions()
    for f, d in dict(min_num_iterations=100, max_num_iterations=10000, max_num_iterations_per_solver=10000,
                     success_probability=0.9999, random_seed=0, num_lo_steps=10, num_lsq_iterations=4,
                     min_sample_multiplicator=7, non_min_sample_multiplier=3, lo_starting_iterations=50,
                     final_least_squares=False, use_ours=False, use_4p4d=False).items():
        assert getattr(o, f) == d                      # defaults of ext/RansacLib/RansacLib/hybrid_ransac.h:49-92
    assert o.squared_inlier_thresholds == [] and o.data_type_weights == [] and abs(o.threshold_multiplier - 2 ** 0.5) < 1e-15
    c = madpose.EstimatorConfig(solver=1, score=2, LO=0)
    assert c.min_depth_constraint and c.use_shift and c.ceres_max_num_iterations == 25 and c.ceres_num_threads == 1
    s = madpose.HybridRansacStatistics()
    for f in ("num_iterations_total", "num_iterations_per_solver", "best_num_inliers", "best_solver_type",
              "best_model_score", "inlier_ratios", "inlier_indices", "number_lo_iterations"):
        assert hasattr(s, f)
    p = madpose.PoseScaleOffsetTwoFocal(np.eye(3), np.array([1., 2, 3]), 2.0, 0.1, 0.2, 500.0, 600.0)
    assert p.pose.shape == (3, 4) and np.allclose(p.t(), [1, 2, 3]) and np.allclose(p.R(), np.eye(3))
    assert (p.scale, p.offset0, p.offset1, p.focal0, p.focal1) == (2.0, 0.1, 0.2, 500.0, 600.0)
    q = madpose.PoseScaleOffsetSharedFocal(np.hstack([np.eye(3), np.ones((3, 1))]), 1.0, 0.0, 0.0, 450.0)
    assert q.focal == 450.0 and np.allclose(q.t(), 1)
    assert np.allclose(madpose.PoseScaleOffset().pose, 0) and madpose.PoseScaleOffset().scale == 1   # pose.h:11,25
    for name in ("RansacStats", "RansacOptions", "LORansacOptions", "ostream_redirect", "HybridEstimatePoseScaleOffset",
                 "HybridEstimatePoseScaleOffsetSharedFocal", "HybridEstimatePoseScaleOffsetTwoFocal",
                 "HybridEstimatePoseAndScale", "solve_scale_and_shift", "solve_scale_shift_pose_two_focal"):
        assert hasattr(madpose, name)
    for name in ("estimate_scale_and_pose", "solve_scale_and_shift_shared_focal", "solve_scale_and_shift_two_focal",
                 "solve_scale_shift_pose", "solve_scale_shift_pose_shared_focal", "PoseAndScale"):
        assert hasattr(madpose, name)      # bindings.cpp:111-118,156-166
    ps = madpose.PoseAndScale(np.eye(3), np.array([1., 2, 3]), 2.5)
    assert ps.scale == 2.5 and np.allclose(ps.t(), [1, 2, 3]) and ps.pose.shape == (3, 4) and madpose.PoseAndScale().scale == 1
    with pytest.raises(ValueError, match="3x3"):   # argument validation happens before any device call
        madpose.solve_scale_and_shift(np.eye(4), np.eye(3), np.ones(3), np.ones(3))


def test_argument_validation():
    o = madpose.HybridLORansacOptions()
    x = np.zeros((6, 2)); d = np.ones(6)
    with pytest.raises(ValueError, match="2 entries"):   # the reference indexes [1] of an empty vector (UB)
        madpose.HybridEstimatePoseScaleOffset(x, x, d, d, [0, 0], np.eye(3), np.eye(3), o)
    o.squared_inlier_thresholds, o.data_type_weights = [64.0, 4.0], [1.0, 1.0]
    with pytest.raises(ValueError, match="same length"):
        madpose.HybridEstimatePoseScaleOffset(x, x, d[:5], d, [0, 0], np.eye(3), np.eye(3), o)
    with pytest.raises(ValueError, match="2-vectors"):
        madpose.HybridEstimatePoseScaleOffsetSharedFocal(x, x, d, d, [0, 0], np.eye(3), np.eye(3), o)


@pytest.mark.gpu
@pytest.mark.parametrize("config", [0, 1, 2])
def test_drop_in_call_matches_oracle(oracle, config):
    from madpose_b200 import capi
    from tests.helpers import problem
    d, pair, opt, cfg, keep = problem(config, p=3)
    o = _options(config)
    est = madpose.EstimatorConfig()
    if config == 0:   # float32 keypoints are accepted and upcast, as pybind11's Eigen casters do
        x0 = d["x0"].astype(np.float32)
        pose, stats = madpose.HybridEstimatePoseScaleOffset(x0, d["x1"], d["d0"], d["d1"], d["min_depth"], d["K0"], d["K1"], o, est)
        pair, keep = capi.make_pair(x0, d["x1"], d["d0"], d["d1"], d["min_depth"], d["K0"], d["K1"])
    elif config == 1:
        pose, stats = madpose.HybridEstimatePoseScaleOffsetSharedFocal(d["x0"], d["x1"], d["d0"], d["d1"], list(d["min_depth"]),
                                                                       d["pp0"], d["pp1"], o)
    else:
        pose, stats = madpose.HybridEstimatePoseScaleOffsetTwoFocal(d["x0"], d["x1"], d["d0"], d["d1"], d["min_depth"], d["pp0"],
                                                                    d["pp1"], o, est_config=est)
    mo, so, io = oracle.estimate_batch(d["kind"], [pair], opt, cfg)
    assert stats.num_iterations_total == so[0].num_iterations_total
    assert stats.num_iterations_per_solver == list(so[0].num_iterations_per_solver)
    assert stats.number_lo_iterations == so[0].number_lo_iterations and stats.best_num_inliers == so[0].best_num_inliers
    assert all(np.array_equal(np.array(a), b) for a, b in zip(stats.inlier_indices, io[0]))
    np.testing.assert_allclose(pose.R().reshape(9), mo[0][:9], atol=1e-9)
    np.testing.assert_allclose(pose.t(), mo[0][9:12], rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose([pose.scale, pose.offset0, pose.offset1], mo[0][12:15], rtol=1e-6)
    if config == 1:
        assert abs(pose.focal / mo[0][15] - 1) < 1e-6
    if config == 2:
        assert abs(pose.focal0 / mo[0][15] - 1) < 1e-6 and abs(pose.focal1 / mo[0][16] - 1) < 1e-6


@pytest.mark.gpu
def test_batch_superset_call():
    ds = [synth.make_pair(1, p, n=500) for p in range(5)]
    o = _options(1)
    out = madpose.HybridEstimatePoseScaleOffsetSharedFocalBatch([d["x0"] for d in ds], [d["x1"] for d in ds], [d["d0"] for d in ds],
                                                                [d["d1"] for d in ds], [d["min_depth"] for d in ds],
                                                                [d["pp0"] for d in ds], [d["pp1"] for d in ds], o)
    assert len(out) == 5
    for d, (pose, stats) in zip(ds, out):
        single, s1 = madpose.HybridEstimatePoseScaleOffsetSharedFocal(d["x0"], d["x1"], d["d0"], d["d1"], d["min_depth"], d["pp0"], d["pp1"], o)
        assert np.array_equal(pose.pose, single.pose) and pose.focal == single.focal   # batching does not change results
        assert stats.num_iterations_total == s1.num_iterations_total
        R = pose.R()
        assert np.degrees(np.arccos(np.clip((np.trace(R.T @ d["gt"]["R"]) - 1) / 2, -1, 1))) < 3.0


@pytest.mark.gpu
def test_solver_functions_match_reference_numpy_goldens():
    """solve_scale_and_shift{,_shared_focal,_two_focal} and solve_scale_shift_pose* (bindings.cpp:157-166) on the GPU
    against the outputs of the reference's own numpy solvers (tests/golden/affine_solvers.json): solution sets to 1e-7,
    and every returned pose maps the corrected back-projections of view 0 onto those of view 1."""
    import json
    import os
    cases = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "affine_solvers.json")))
    raw = [madpose.solve_scale_and_shift, madpose.solve_scale_and_shift_shared_focal, madpose.solve_scale_and_shift_two_focal]
    pose = [madpose.solve_scale_shift_pose, madpose.solve_scale_shift_pose_shared_focal, madpose.solve_scale_shift_pose_two_focal]
    n_ref = n_pose = 0
    for g in cases:
        k = g["kind"]
        xh, yh = np.array(g["x1"]).T.copy(), np.array(g["x2"]).T.copy()   # 3 x NP, columns = homogeneous points
        d1, d2 = np.array(g["d1"]), np.array(g["d2"])
        mine = np.array(raw[k](xh, yh, d1, d2)).reshape(-1, 4 + k)
        ref = np.array(g["sols"]).reshape(-1, 4 + k)
        assert len(mine) == len(ref)
        for r in ref:
            n_ref += 1
            err = np.max(np.abs(mine[:, 1:] - r[1:]) / np.maximum(np.abs(r[1:]), 1e-300), axis=1)
            assert err.min() < 1e-7 and np.all(mine[:, 0] == 1.0)
        for m in pose[k](xh, yh, d1, d2):
            n_pose += 1
            f0 = 1.0 if k == 0 else (m.focal if k == 1 else m.focal0)
            f1 = 1.0 if k == 0 else (m.focal if k == 1 else m.focal1)
            X = (xh / np.array([f0, f0, 1.0])[:, None]) * (d1 + m.offset0)
            Y = (yh / np.array([f1, f1, 1.0])[:, None]) * (d2 * m.scale + m.offset1)
            assert np.all(d1 + m.offset0 > 0) and np.all(d2 * m.scale + m.offset1 > 0)   # solver.cpp:503,702,1006
            res = m.R() @ X + m.t()[:, None] - Y
            if k == 0:  # 3 points: Kabsch aligns exactly (4 noisy points of the focal kinds: least squares, any residual)
                assert np.max(np.abs(res)) <= 1e-6 * max(1.0, np.max(np.abs(Y))), "pose does not align the corrected points"
            assert abs(np.linalg.det(m.R()) - 1) < 1e-9
    assert n_ref > 250 and n_pose > 100


@pytest.mark.gpu
def test_estimate_scale_and_pose_and_scale_only_estimator(oracle):
    """estimate_scale_and_pose (solver.cpp:5-33) on a synthetic similarity; HybridEstimatePoseAndScale
    (hybrid_pose_estimator.cpp:37-63) against the oracle's scale-only path."""
    from madpose_b200 import capi
    from tests.helpers import problem
    rng = np.random.default_rng(3)
    X = rng.standard_normal((3, 12))
    R = np.linalg.qr(rng.standard_normal((3, 3)))[0]; R *= np.linalg.det(R)
    t, sc = rng.standard_normal(3), 1.7
    Y = sc * R @ X + t[:, None]
    ps = madpose.estimate_scale_and_pose(X, Y, rng.uniform(0.5, 2.0, 12))
    np.testing.assert_allclose(ps.R(), R, atol=1e-10); np.testing.assert_allclose(ps.t(), t, atol=1e-10)
    assert abs(ps.scale - sc) < 1e-10
    d, pair, opt, cfg, keep = problem(0, p=9, zero_offsets=True)
    pose, stats = madpose.HybridEstimatePoseAndScale(d["x0"], d["x1"], d["d0"], d["d1"], d["K0"], d["K1"], _options(0))
    mo, so, io = oracle.estimate_batch(capi.KIND_SCALE_ONLY, [pair], opt, cfg)
    assert stats.num_iterations_total == so[0].num_iterations_total and stats.best_num_inliers == so[0].best_num_inliers
    np.testing.assert_allclose(pose.R().reshape(9), mo[0][:9], atol=1e-9)
    np.testing.assert_allclose(pose.t(), mo[0][9:12], rtol=1e-6, atol=1e-9)
    assert abs(pose.scale / mo[0][12] - 1) < 1e-6
    Rg = d["gt"]["R"]
    assert np.degrees(np.arccos(np.clip((np.trace(pose.R().T @ Rg) - 1) / 2, -1, 1))) < 2.0
