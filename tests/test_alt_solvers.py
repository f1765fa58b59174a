"""The alternative depth-aware minimal solvers behind HybridLORansacOptions.use_ours / use_4p4d
(solver.cpp:536-680, 818-984, 1045-1153, 1287-1406; selected at hybrid_pose_estimator.cpp:75-78,
hybrid_pose_shared_focal_estimator.cpp:62-65, hybrid_pose_two_focal_estimator.cpp:86-93).  No reference test or numpy
counterpart exists for them: the oracle restatement (oracle/om_alt.h) is checked on noise-free instances -- the ground
truth must be among the returned models and every model must align the corrected back-projections -- the device math
against the oracle (CPU, hostcheck), and the whole path with the flags on against the oracle (GPU)."""
import numpy as np
import pytest

from madpose_b200 import capi, synth
from tests import hostcheck_lib
from tests.helpers import problem

VARIANTS = [("calib_ours", 0, dict(use_ours=1)), ("shared_ours", 1, dict(use_ours=1)), ("two_ours", 2, dict(use_ours=1)),
            ("two_4p4d", 2, dict(use_4p4d=1))]


def clean_pair(kind, seed, n=12, shifts=False):
    """Noise-free correspondences of a random scene; priors off by a scale (and, if shifts, by offsets)."""
    rng = np.random.default_rng(seed)
    f0 = 577.0 if kind == 0 else rng.uniform(400, 700)
    f1 = f0 if kind < 2 else rng.uniform(400, 700)
    pp = np.array([320.0, 240.0])
    while True:
        ax = rng.standard_normal(3); ax /= np.linalg.norm(ax)
        ang = np.deg2rad(rng.uniform(5, 25))
        K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
        R = np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * K @ K
        t = rng.standard_normal(3); t *= rng.uniform(0.3, 0.8) / np.linalg.norm(t)
        x0 = np.c_[rng.uniform(40, 600, n), rng.uniform(40, 440, n)]
        z0 = rng.uniform(1.5, 5, n)
        X = np.c_[(x0 - pp) / f0, np.ones(n)] * z0[:, None]
        Y = X @ R.T + t
        if np.all(Y[:, 2] > 0.5):
            break
    x1 = Y[:, :2] / Y[:, 2:] * f1 + pp
    a0, a1 = rng.uniform(0.6, 1.8, 2)
    o0, o1 = (rng.uniform(-0.2, 0.3, 2) if shifts else (0.0, 0.0))
    d0 = z0 / a0 - o0; d1 = Y[:, 2] / a1 - o1
    K0 = np.array([[f0, 0, pp[0]], [0, f0, pp[1]], [0, 0, 1.0]]); K1 = np.array([[f1, 0, pp[0]], [0, f1, pp[1]], [0, 0, 1.0]])
    pair, keep = capi.make_pair(x0, x1, d0, d1, [d0.min(), d1.min()], K0, K1, pp, pp)
    gt = dict(R=R, t=t / a0, scale=a1 / a0, o0=o0, o1=o1, f0=f0, f1=f1)
    return pair, keep, gt


@pytest.mark.parametrize("name,kind,optkw", VARIANTS)
def test_oracle_alt_solver_recovers_ground_truth(oracle, name, kind, optkw):
    hits = total = 0
    for seed in range(12):
        pair, keep, gt = clean_pair(kind, seed, shifts=(kind == 0))
        opt = capi.Options(**{**synth.options_for(kind), **optkw}); cfg = capi.Config(min_depth_constraint=0)
        ns = oracle.frontend(kind, pair, opt, cfg)["norm_scale"]
        k = 3 if kind == 0 else 4
        S = np.array([[0, 1, 2, 3][:k], [4, 5, 6, 7][:k], [8, 9, 10, 11][:k]], np.int32)
        for models in oracle.minimal_solve(kind, pair, opt, cfg, 0, S):
            total += 1
            ok = False
            for m in models:
                Rm = m[:9].reshape(3, 3)
                assert abs(np.linalg.det(Rm) - 1) < 1e-6
                good = np.linalg.norm(Rm - gt["R"]) < 1e-6
                if name != "two_4p4d":   # the 4p4d solver returns the direction of t only (scale 1, :1398-1400)
                    good = good and abs(m[12] / gt["scale"] - 1) < 1e-6 and np.linalg.norm(m[9:12] - gt["t"]) < 1e-6 * max(1, np.linalg.norm(gt["t"]))
                    good = good and abs(m[13] - gt["o0"]) < 1e-6 and abs(m[14] - gt["o1"]) < 1e-6
                else:
                    c = m[9:12] @ gt["t"] / np.linalg.norm(m[9:12]) / np.linalg.norm(gt["t"])
                    good = good and c > 1 - 1e-9 and m[12] == 1.0
                if kind > 0:
                    good = good and abs(m[15] * ns / gt["f0"] - 1) < 1e-6 and abs(m[16] * ns / gt["f1"] - 1) < 1e-6
                ok = ok or good
            hits += ok
    # a sample can be degenerate (near-planar triplet); the 4p4d solver tests cheirality on un-calibrated bearings
    # (solver.cpp:1301-1302, a reference quirk that is kept) and rejects the correct pose in some scenes
    assert total == 36 and hits >= (25 if name == "two_4p4d" else 34), (hits, total)


@pytest.mark.parametrize("name,kind,optkw", VARIANTS)
def test_alt_solver_device_math_matches_oracle(oracle, name, kind, optkw):
    hc = hostcheck_lib.load()
    d, pair, opt, cfg, keep = problem(kind, p=3, n=500, zero_offsets=(kind != 0), **optkw)
    rng = np.random.default_rng(5)
    inl = np.nonzero(d["gt"]["inlier_mask"])[0]
    k = 3 if kind == 0 else 4
    S = np.array([rng.choice(inl if i % 2 == 0 else pair.n, k, replace=False) for i in range(150)], np.int32)
    a = oracle.minimal_solve(kind, pair, opt, cfg, 0, S)
    b = hc.minimal_solve(kind, pair, opt, cfg, 0, S)
    errs, n_models = [], 0
    assert sum(len(x) != len(y) for x, y in zip(a, b)) <= 3
    for x, y in zip(a, b):
        if len(x) == len(y):
            for m in x:
                n_models += 1
                errs.append(min(np.max(np.abs(m - mm) / np.maximum(np.abs(m), 1e-3)) for mm in y))
    errs = np.array(errs)
    # (on noisy samples the 4p4d solver rarely returns anything: Bougnoux's f^2 < 0 or its un-calibrated cheirality test)
    assert n_models > (5 if name == "two_4p4d" else 30) and np.mean(errs <= 1e-8) >= 0.95, (n_models, np.sort(errs)[-5:])


@pytest.mark.gpu
@pytest.mark.parametrize("name,kind,optkw", VARIANTS)
def test_alt_solvers_whole_path(ctx, oracle, name, kind, optkw):
    from tests.test_estimate_gpu import compare, rot_err_deg
    ps = [problem(kind, p=120 + i, n=700, zero_offsets=(kind != 0), **optkw) for i in range(4)]
    opt, cfg = ps[0][2], ps[0][3]
    pairs = [q[1] for q in ps]
    mo, so, io = oracle.estimate_batch(kind, pairs, opt, cfg, threads=4)
    mg, sg, ig = ctx.estimate_batch(kind, pairs, opt, cfg)
    compare(oracle, kind, pairs, opt, cfg, mo, so, io, mg, sg, ig, max_forks=1, ctx=ctx)
    for q, m in zip(ps, mg):
        assert rot_err_deg(q[0]["gt"]["R"].reshape(9), m[:9]) < 3.0
