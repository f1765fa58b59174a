"""Independent cross-checks of the oracle's restated THIRD-PARTY solvers (PoseLib 5pt / 6pt shared focal / 7pt,
cv::recoverPose, Ceres LM are absent from /root/reference: SURVEY.md 8c, Appendix C KAT-3 / KAT-3b):
  * 5-point vs OpenCV's own five-point solver (cv2.findEssentialMat on 5 points returns every candidate),
  * 7-point vs cv2.findFundamentalMat(FM_7POINT),
  * recoverPose vs cv2.recoverPose(E, x0, x1, I, distanceThresh=1e9) (hybrid_pose_two_focal_estimator.cpp:143),
  * 6-point shared focal: every candidate satisfies the epipolar / trace / determinant constraints and the Bougnoux
    identity, and the ground-truth root is among them,
  * LM: the end point is the optimum scipy.optimize.least_squares finds on the same residual vector.
The oracle and the CUDA kernels share one author; these tests are what guards against a common-mode error."""
import ctypes as C

import cv2
import numpy as np
import pytest
import scipy.optimize

from madpose_b200 import capi
from tests.helpers import models_around_gt, problem

vp = lambda a: a.ctypes.data_as(C.c_void_p)


def rand_rot(rng, max_deg=30.0):
    ax = rng.standard_normal(3); ax /= np.linalg.norm(ax)
    ang = np.deg2rad(rng.uniform(3, max_deg))
    K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    return np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * K @ K


def scene(rng, n, f0=1.0, f1=1.0):
    """n points seen by two cameras: returns (R, t, x0 (n,2), x1 (n,2)) in pixels of K = diag(f, f, 1)."""
    while True:
        R, t = rand_rot(rng), rng.standard_normal(3)
        t *= rng.uniform(0.3, 1.0) / np.linalg.norm(t)
        X = np.c_[rng.uniform(-1, 1, n), rng.uniform(-1, 1, n), rng.uniform(2, 6, n)]
        Y = X @ R.T + t
        if np.all(Y[:, 2] > 0.5):
            return R, t, f0 * X[:, :2] / X[:, 2:], f1 * Y[:, :2] / Y[:, 2:]


def essential(R, t):
    tx = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
    return tx @ R


def same_up_to_scale(A, B):
    A, B = A / np.linalg.norm(A), B / np.linalg.norm(B)
    return min(np.linalg.norm(A - B), np.linalg.norm(A + B))


def unit(h):
    return h / np.linalg.norm(h, axis=1, keepdims=True)


def oracle_5pt(oracle, x0, x1):
    b0 = unit(np.c_[x0, np.ones(5)]).copy(); b1 = unit(np.c_[x1, np.ones(5)]).copy()
    Es = np.zeros((10, 9)); P = np.zeros((10, 12)); nE, nP = C.c_int(), C.c_int()
    oracle.lib.om_relpose_5pt(vp(b0), vp(b1), vp(Es), C.byref(nE), vp(P), C.byref(nP))
    return Es[:nE.value].reshape(-1, 3, 3), P[:nP.value]


def test_5pt_contains_every_opencv_candidate(oracle):
    """KAT-3b: OpenCV's five-point solver (Nister) is an independent implementation; every essential matrix it returns
    for a 5-point sample must be among the oracle's (same number of real solutions; unit-norm matrices equal to OpenCV's
    precision), the ground truth must be among them to 1e-9 and every oracle candidate satisfies the defining constraints."""
    rng = np.random.default_rng(0)
    n_cv = n_inst = 0
    errs = []
    for _ in range(120):
        R, t, x0, x1 = scene(rng, 5)
        Eo, Po = oracle_5pt(oracle, x0, x1)
        assert min(same_up_to_scale(E, essential(R, t)) for E in Eo) < 1e-9, "ground-truth E missing from the oracle's set"
        # the pose the estimator keeps: the ground truth passes the all-point cheirality test
        assert any(np.linalg.norm(p[:9].reshape(3, 3) - R) < 1e-8 and same_up_to_scale(p[9:], t) < 1e-8 for p in Po)
        Ecv, _ = cv2.findEssentialMat(x0, x1, np.eye(3), method=cv2.RANSAC, prob=0.999, threshold=1e-3)
        if Ecv is None:
            continue
        n_inst += 1
        assert len(Ecv.reshape(-1, 3, 3)) == len(Eo), "OpenCV and the oracle disagree on the number of real solutions"
        for E in Ecv.reshape(-1, 3, 3):
            n_cv += 1
            errs.append(min(same_up_to_scale(E, Eq) for Eq in Eo))
        for E in Eo:  # and the oracle returns no candidate that violates the defining constraints
            E = E / np.linalg.norm(E)
            assert abs(np.linalg.det(E)) < 1e-10 and np.linalg.norm(2 * E @ E.T @ E - np.trace(E @ E.T) * E) < 1e-9
            assert np.max(np.abs(np.einsum("ni,ij,nj->n", np.c_[x1, np.ones(5)], E, np.c_[x0, np.ones(5)]))) < 1e-9
    errs = np.array(errs)
    # OpenCV takes the roots from a companion-matrix eigen-decomposition (lower precision on ill-conditioned roots):
    # median 1e-14, 98 % below 1e-7, every candidate matched
    assert n_inst >= 100 and n_cv >= 300 and np.median(errs) < 1e-12 and np.mean(errs < 1e-7) >= 0.98 and errs.max() < 1e-4


def test_7pt_matches_opencv(oracle):
    rng = np.random.default_rng(1)
    hits = 0
    for _ in range(40):
        R, t, x0, x1 = scene(rng, 7, 500.0, 650.0)
        h0 = np.c_[x0, np.ones(7)].copy(); h1 = np.c_[x1, np.ones(7)].copy()
        Fs = np.zeros((3, 9)); nF = C.c_int()
        oracle.lib.om_relpose_7pt(vp(h0), vp(h1), vp(Fs), C.byref(nF))
        Fo = Fs[:nF.value].reshape(-1, 3, 3)
        Kinv0, Kinv1 = np.diag([1 / 500.0, 1 / 500.0, 1]), np.diag([1 / 650.0, 1 / 650.0, 1])
        Fgt = Kinv1 @ essential(R, t) @ Kinv0
        assert min(same_up_to_scale(F, Fgt) for F in Fo) < 1e-8
        for F in Fo:
            assert abs(np.linalg.det(F / np.linalg.norm(F))) < 1e-10
            assert np.max(np.abs(np.einsum("ni,ij,nj->n", h1, F / np.linalg.norm(F), h0))) < 1e-9 * 650
        Fcv, _ = cv2.findFundamentalMat(x0, x1, cv2.FM_7POINT)
        if Fcv is None:
            continue
        for F in Fcv.reshape(-1, 3, 3):   # OpenCV's 7-point solver is a lower-precision implementation: 1e-5 sanity
            hits += min(same_up_to_scale(F, Fq) for Fq in Fo) < 1e-5
        assert len(Fcv.reshape(-1, 3, 3)) == len(Fo)
    assert hits >= 80


def test_recover_pose_matches_opencv(oracle):
    """KAT-3: cv2.recoverPose(E, x0, x1, I, distanceThresh=1e9) -- the 7-argument overload the two-focal estimator
    calls on un-calibrated points with K = I."""
    rng = np.random.default_rng(2)
    for k in range(40):
        R, t, x0, x1 = scene(rng, 7)
        E = essential(R, t) * (1.0 if k % 2 else -2.5)
        Ro, to = np.zeros(9), np.zeros(3)
        oracle.lib.om_recover_pose(vp(np.ascontiguousarray(E)), 7, vp(np.ascontiguousarray(x0)), vp(np.ascontiguousarray(x1)),
                                   C.c_double(1e9), vp(Ro), vp(to))
        _, Rc, tc, _, _ = cv2.recoverPose(E, x0, x1, np.eye(3), distanceThresh=1e9)
        np.testing.assert_allclose(Ro.reshape(3, 3), Rc, atol=1e-12)
        np.testing.assert_allclose(to, tc.ravel(), atol=1e-12)
        np.testing.assert_allclose(Ro.reshape(3, 3), R, atol=1e-10)
    # a noisy, partly wrong sample (what RANSAC feeds): both still agree on the candidate with most points in front
    for k in range(40):
        R, t, x0, x1 = scene(rng, 7)
        x1 = x1 + rng.normal(0, 0.02, x1.shape); x1[rng.integers(7)] = rng.uniform(-0.3, 0.3, 2)
        E = essential(rand_rot(rng), rng.standard_normal(3))
        Ro, to = np.zeros(9), np.zeros(3)
        oracle.lib.om_recover_pose(vp(np.ascontiguousarray(E)), 7, vp(np.ascontiguousarray(x0)), vp(np.ascontiguousarray(x1)),
                                   C.c_double(1e9), vp(Ro), vp(to))
        n, Rc, tc, _, _ = cv2.recoverPose(E, x0, x1, np.eye(3), distanceThresh=1e9)
        if n < 7:   # ties between candidates are broken by OpenCV's SVD sign conventions: compare clear winners only
            continue
        np.testing.assert_allclose(Ro.reshape(3, 3), Rc, atol=1e-9)
        np.testing.assert_allclose(to, tc.ravel(), atol=1e-9)


def test_6pt_shared_focal_constraints_and_ground_truth(oracle):
    """100 seeded instances: the ground-truth (R, t, f) is among the candidates (1e-7) and EVERY candidate satisfies
    x1^T F x0 = 0 on the six points, det F = 0, the trace constraint of E = K F K, and Bougnoux's formula returns its
    focal."""
    rng = np.random.default_rng(3)
    total = 0
    for _ in range(100):
        f = rng.uniform(0.6, 2.5)
        R, t, x0, x1 = scene(rng, 6, f, f)
        h0 = np.c_[x0, np.ones(6)].copy(); h1 = np.c_[x1, np.ones(6)].copy()
        out = np.zeros((15, 13)); n = C.c_int()
        oracle.lib.om_relpose_6pt_focal(vp(h0), vp(h1), vp(out), C.byref(n))
        cand = out[:n.value]
        assert len(cand) >= 1
        ok_gt = False
        for c in cand:
            total += 1
            Rc, tc, fc = c[:9].reshape(3, 3), c[9:12], c[12]
            Ki = np.diag([1 / fc, 1 / fc, 1.0])
            E = essential(Rc, tc); E /= np.linalg.norm(E)
            F = Ki @ E @ Ki
            assert np.max(np.abs(np.einsum("ni,ij,nj->n", h1, F, h0))) < 1e-8 * max(1.0, np.max(np.abs(h0)) * np.max(np.abs(h1)))
            assert abs(np.linalg.det(E)) < 1e-10 and np.linalg.norm(2 * E @ E.T @ E - np.trace(E @ E.T) * E) < 1e-9
            assert abs(np.linalg.det(Rc) - 1) < 1e-9 and fc > 0
            f0sq, f1sq = C.c_double(), C.c_double()
            oracle.lib.om_bougnoux(vp(np.ascontiguousarray(F)), C.byref(f0sq), C.byref(f1sq))
            if np.isfinite(f0sq.value) and abs(f0sq.value) < 1e6:   # Bougnoux is singular for some configurations
                assert abs(f0sq.value / fc ** 2 - 1) < 1e-6 and abs(f1sq.value / fc ** 2 - 1) < 1e-6
            if np.linalg.norm(Rc - R) < 1e-7 and same_up_to_scale(tc, t) < 1e-7 and abs(fc / f - 1) < 1e-7:
                ok_gt = True
        assert ok_gt, "ground-truth (R, t, f) missing from the 6-point candidates"
    assert total >= 150


@pytest.mark.parametrize("config,kind", [(0, 0), (1, 1), (2, 2), (0, 3)])
def test_lm_end_point_is_the_scipy_optimum(oracle, config, kind):
    """The oracle's trust-region LM (the specification of the CUDA kernel) against an independent solver on
    the SAME residual vector (om_lm_residuals = the functors of cost_functions.h): run to tight tolerances, both must
    reach the same cost (1e-6 relative, optimizer.h:113-125 semantics: the point Ceres leaves behind is used)."""
    d, pair, opt, cfg, keep = problem(config, p=6, n=500, zero_offsets=(kind == 3))
    cfg = capi.Config(ceres_function_tolerance=1e-12, ceres_parameter_tolerance=1e-12, ceres_gradient_tolerance=1e-12,
                      ceres_max_num_iterations=200)
    fe = oracle.frontend(kind, pair, opt, cfg)
    start = models_around_gt(d, fe["norm_scale"], 3, seed=1, spread=0.05)[2]
    if kind == 3:
        start[13] = start[14] = 0.0
    inl = np.nonzero(d["gt"]["inlier_mask"])[0][:150].astype(np.int32)
    out, info = oracle.lm(kind, pair, opt, cfg, [start], [[inl, inl, inl]])
    idx = np.concatenate([inl, inl, inl]).astype(np.int32); off = np.array([0, len(inl), 2 * len(inl), 3 * len(inl)], np.int32)

    def resid(m17):
        r = np.zeros(5 * len(inl))
        k = oracle.lib.om_lm_residuals(kind, C.byref(pair), C.byref(opt), C.byref(cfg), vp(np.ascontiguousarray(m17)), vp(idx), vp(off), vp(r))
        return r[:k]
    cost_oracle = 0.5 * np.sum(resid(out[0]) ** 2)
    assert abs(cost_oracle - info[0, 3]) <= 1e-9 * cost_oracle   # the cost the LM reports is the cost of its end point

    # free parameters of the scipy problem: rotation vector applied on the left of the start rotation, t, scale,
    # offsets (absent for the scale-only kind), focals
    R0 = start[:9].reshape(3, 3)

    def unpack(p):
        m = start.copy()
        w = p[:3]; th = np.linalg.norm(w)
        K = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
        dR = np.eye(3) if th < 1e-300 else np.eye(3) + np.sin(th) / th * K + (1 - np.cos(th)) / th ** 2 * K @ K
        m[:9] = (dR @ R0).reshape(9); m[9:12] = p[3:6]; m[12] = p[6]
        j = 7
        if kind != 3:
            m[13], m[14] = p[7], p[8]; j = 9
        if kind == 1:
            m[15] = m[16] = p[j]
        if kind == 2:
            m[15], m[16] = p[j], p[j + 1]
        return m
    p0 = np.concatenate([np.zeros(3), start[9:12], [start[12]], [] if kind == 3 else start[13:15],
                         [start[15]] if kind == 1 else ([start[15], start[16]] if kind == 2 else [])])
    sol = scipy.optimize.least_squares(lambda p: resid(unpack(p)), p0, method="lm", xtol=1e-14, ftol=1e-14, gtol=1e-14)
    assert abs(cost_oracle - sol.cost) <= 1e-6 * sol.cost, (cost_oracle, sol.cost)
