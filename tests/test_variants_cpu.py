"""CPU checks of the estimator variants beyond the default path: EstimatorConfig.use_shift = False
(hybrid_pose_estimator.cpp:87-120,146-157; LM asymmetry :206,281 / ...shared_focal....cpp:149) and the scale-only
estimator HybridEstimatePoseAndScale (hybrid_pose_estimator.cpp:37-63,297-442; optimizer.h:133-245; solver.cpp:5-33).
Oracle restatement vs the reference's own loop headers (oracle/_ref), vs the device math compiled for the host, and
against ground truth."""
import ctypes as C
import os

import numpy as np
import pytest

from madpose_b200 import capi
from tests import hostcheck_lib, oracle_lib
from tests.helpers import models_around_gt, problem

REF = os.path.join(oracle_lib.ORACLE_DIR, "_ref", "liboracle_ref.so")
VARIANTS = [  # (name, config, kind, cfg)
    ("calib_noshift", 0, 0, dict(use_shift=0)),
    ("scale_only", 0, capi.KIND_SCALE_ONLY, {}),
    ("shared_noshift", 1, 1, dict(use_shift=0)),
]


@pytest.fixture(scope="module")
def hc():
    return hostcheck_lib.load()


@pytest.mark.parametrize("name,config,kind,cfgkw", VARIANTS[:2])
@pytest.mark.parametrize("solver", [0, 1])
def test_variant_minimal_math_matches_oracle(hc, oracle, name, config, kind, cfgkw, solver):
    d, pair, opt, cfg, keep = problem(config, p=3, n=500, zero_offsets=True, cfg=cfgkw)
    rng = np.random.default_rng(7 + solver)
    inl = np.nonzero(d["gt"]["inlier_mask"])[0]
    k = (3, 5)[solver]
    S = np.array([rng.choice(inl if i % 2 == 0 else pair.n, k, replace=False) for i in range(80)], np.int32)
    a = oracle.minimal_solve(kind, pair, opt, cfg, solver, S)
    b = hc.minimal_solve(kind, pair, opt, cfg, solver, S)
    n_models, errs = 0, []
    for x, y in zip(a, b):
        assert len(x) == len(y)
        if solver == 0:
            assert len(x) <= 1  # one model per sample, no filter
        for m in x:
            assert m[13] == 0 and m[14] == 0  # offsets absent
            errs.append(min(np.max(np.abs(m - mm) / np.maximum(np.abs(m), 1e-3)) for mm in y))
            n_models += 1
    errs = np.array(errs)
    # (a sample whose 10th-degree polynomial has a near-double root may keep a different member of the pair: <= 2 %)
    assert n_models > 40 and np.mean(errs <= 1e-9) >= 0.98


def test_noshift_solver0_recovers_scale_on_clean_sample(oracle):
    """use_shift = False, solver 0 on noise-free inlier samples of priors that are off by a scale only: the
    distance-ratio scale is the ground-truth scale and Kabsch returns the ground-truth pose."""
    from madpose_b200 import synth
    d = synth.make_pair(0, 5, n=300, outlier_ratio=0.0, zero_offsets=True)
    # remove the noise the generator adds: rebuild exact observations from the ground truth
    g = d["gt"]
    K0i = np.linalg.inv(d["K0"])
    z0 = d["d0"]  # treat the (noisy) prior as the exact depth in view 0 (alpha0 scale folded in below)
    X = (K0i @ np.c_[d["x0"], np.ones(len(z0))].T) * z0 * g["alpha0"]
    Y = g["R"] @ X + g["t"][:, None]
    x1 = (d["K1"] @ Y); x1 = (x1[:2] / x1[2]).T
    d1 = Y[2] / (g["alpha0"] * g["scale"])
    pair, keep = capi.make_pair(d["x0"], x1, z0, d1, [z0.min(), d1.min()], d["K0"], d["K1"])
    opt = capi.Options(**synth.options_for(0)); cfg = capi.Config(use_shift=0)
    S = np.array([[0, 1, 2], [10, 50, 90], [7, 8, 200]], np.int32)
    for ms in oracle.minimal_solve(0, pair, opt, cfg, 0, S):
        assert len(ms) == 1
        m = ms[0]
        np.testing.assert_allclose(m[:9].reshape(3, 3), g["R"], atol=1e-9)
        np.testing.assert_allclose(m[12], g["scale"], rtol=1e-9)
        np.testing.assert_allclose(m[9:12], g["t"] / g["alpha0"], atol=1e-9)


@pytest.mark.parametrize("name,config,kind,cfgkw", VARIANTS)
def test_variant_lm_math_matches_oracle(hc, oracle, name, config, kind, cfgkw):
    d, pair, opt, cfg, keep = problem(config, p=4, n=700, zero_offsets=True, cfg=cfgkw)
    fe = oracle.frontend(kind, pair, opt, cfg)
    starts = models_around_gt(d, fe["norm_scale"], 4, seed=3, spread=0.1)
    inl = np.nonzero(d["gt"]["inlier_mask"])[0]
    rng = np.random.default_rng(5)
    samples = [[inl, inl, inl] if k % 2 == 0 else [rng.choice(inl, 60), rng.choice(inl, 50), rng.choice(inl, 70)]
               for k in range(len(starts))]
    for nonminimal in (False, True):
        a, ia = oracle.lm(kind, pair, opt, cfg, starts, samples, nonminimal=nonminimal)
        b, ib = hc.lm(kind, pair, opt, cfg, starts, samples, nonminimal=nonminimal)
        np.testing.assert_array_equal(ia[:, :2], ib[:, :2])
        conv = ia[:, 1] != 2
        err = np.max(np.abs(a - b) / np.maximum(np.abs(a), 1e-3), axis=1)
        assert np.all(err[conv] <= 1e-8), err
        frozen = name != "shared_noshift" or nonminimal  # shared focal: LeastSquares ignores use_shift (...shared....cpp:205-228)
        moved = np.abs(a[:, 13] - starts[:, 13]) + np.abs(a[:, 14] - starts[:, 14])
        if frozen:
            assert np.all(moved == 0), "offsets must stay constant"
        else:
            assert np.any(moved > 0), "LeastSquares of the shared-focal estimator optimises the offsets"


@pytest.mark.skipif(not os.path.exists(REF), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("name,config,kind,cfgkw", VARIANTS)
def test_variant_loop_equals_reference_loop(oracle, name, config, kind, cfgkw):
    ref = C.CDLL(REF)
    ps = [problem(config, p=20 + i, n=600, zero_offsets=True, cfg=cfgkw) for i in range(2)]
    opt, cfg = ps[0][2], ps[0][3]
    pairs = [q[1] for q in ps]
    mo, so, io = oracle.estimate_batch(kind, pairs, opt, cfg, threads=2)
    mr, sr, ir = oracle.estimate_batch(kind, pairs, opt, cfg, threads=2, lib=ref)
    for i in range(2):
        assert so[i].num_iterations_total == sr[i].num_iterations_total
        assert so[i].number_lo_iterations == sr[i].number_lo_iterations
        assert so[i].best_model_score == sr[i].best_model_score
        assert all(np.array_equal(x, y) for x, y in zip(io[i], ir[i]))
        assert np.array_equal(mo[i], mr[i])


@pytest.mark.parametrize("name,config,kind,cfgkw", VARIANTS[:2])
def test_variant_recovers_ground_truth(oracle, name, config, kind, cfgkw):
    d, pair, opt, cfg, keep = problem(config, p=11, zero_offsets=True, cfg=cfgkw)
    m, s, inl = oracle.estimate_batch(kind, [pair], opt, cfg)
    R = m[0][:9].reshape(3, 3)
    ang = np.degrees(np.arccos(np.clip((np.trace(R.T @ d["gt"]["R"]) - 1) / 2, -1, 1)))
    assert ang < 1.0
    t = m[0][9:12]
    assert np.degrees(np.arccos(np.clip(t @ d["gt"]["t"] / np.linalg.norm(t) / np.linalg.norm(d["gt"]["t"]), -1, 1))) < 3.0
    assert abs(m[0][12] / d["gt"]["scale"] - 1) < 0.05
    assert m[0][13] == 0 and m[0][14] == 0
    assert s[0].best_num_inliers > 0.4 * 3 * pair.n
