"""mp_pose_errors (the step after the path) against golden vectors produced by the reference's own
madpose.utils.compute_pose_error (tests/golden/make_pose_error_golden.py).

Tolerance: 2e-6 degrees absolute.  arccos near +-1 amplifies one ulp of the cosine (a different summation order of
the trace / dot product) to sqrt(2 * 2.2e-16) rad = 1.2e-6 degrees; everywhere else the agreement is ~1e-13."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "pose_errors.npz")
TOL_DEG = 2e-6


def restatement(T, R, t, t_thres=None):
    """numpy restatement of madpose/utils.py:46-78, vectorised over the poses."""
    Rg, tg = T[:, :3, :3], T[:, :3, 3]
    with np.errstate(invalid="ignore", divide="ignore"):
        c = np.sum(t * tg, 1) / (np.linalg.norm(t, axis=1) * np.linalg.norm(tg, axis=1))
        et = np.rad2deg(np.arccos(np.clip(c, -1.0, 1.0)))
    et = np.minimum(et, 180 - et)
    cr = (np.einsum("nij,nij->n", R, Rg) - 1) / 2
    er = np.rad2deg(np.abs(np.arccos(np.clip(cr, -1.0, 1.0))))
    if t_thres is not None:
        et = np.where(np.linalg.norm(tg, axis=1) < t_thres, 0.0, et)
    return et, er


def check(got, z, name):
    for g, key in zip(got, ("err_t_" + name, "err_R_" + name)):
        want = z[key]
        assert np.array_equal(np.isnan(g), np.isnan(want))
        ok = ~np.isnan(want)
        assert np.max(np.abs(g[ok] - want[ok])) <= TOL_DEG


def test_restatement_matches_reference_golden():
    z = np.load(GOLD)
    check(restatement(z["T"], z["R"], z["t"]), z, "none")
    check(restatement(z["T"], z["R"], z["t"], 0.05), z, "thr")


@pytest.mark.gpu
def test_device_matches_reference_golden(ctx):
    z = np.load(GOLD)
    check(ctx.pose_errors(z["T"], z["R"], z["t"]), z, "none")
    check(ctx.pose_errors(z["T"], z["R"], z["t"], 0.05), z, "thr")


@pytest.mark.gpu
def test_utils_module_signature():
    from madpose_b200.utils import compute_pose_error
    z = np.load(GOLD)
    for i in (0, 5, 17, 100):
        et, er = compute_pose_error(z["T"][i], z["R"][i], z["t"][i])
        assert abs(et - z["err_t_none"][i]) <= TOL_DEG and abs(er - z["err_R_none"][i]) <= TOL_DEG
