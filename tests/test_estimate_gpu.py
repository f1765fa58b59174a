"""Whole-path parity: mp_estimate_batch (device-side hybrid LO-MSAC loop) vs the CPU oracle's loop on the same
pairs, options and seed.  The device loop draws the same std::mt19937 streams, so it replays the oracle's sampled
minimal sets; north_star: inlier sets bit-exact (except residuals within 1e-6 of the threshold), R/t within 0.01 deg,
scale / shift / focal within 1e-6 relative."""
import numpy as np
import pytest

from madpose_b200 import capi
from tests.helpers import problem

pytestmark = pytest.mark.gpu


def rot_err_deg(Ra, Rb):
    if not np.any(Ra) and not np.any(Rb):
        return 0.0  # both are the default (zero) pose: no model was found
    c = (np.trace(Ra.reshape(3, 3).T @ Rb.reshape(3, 3)) - 1) / 2
    return np.degrees(np.arccos(np.clip(c, -1, 1)))


def vec_angle_deg(a, b):
    if not np.any(a) and not np.any(b):
        return 0.0
    c = a @ b / max(np.linalg.norm(a) * np.linalg.norm(b), 1e-300)
    return np.degrees(np.arccos(np.clip(c, -1, 1)))


def compare(oracle, kind, pairs, opt, cfg, mo, so, io, mg, sg, ig, min_identical):
    identical = 0
    for i, pr in enumerate(pairs):
        a, b = so[i], sg[i]
        same = (a.num_iterations_total == b.num_iterations_total
                and list(a.num_iterations_per_solver) == list(b.num_iterations_per_solver)
                and a.number_lo_iterations == b.number_lo_iterations and a.best_solver_type == b.best_solver_type)
        inl_same = all(np.array_equal(x, y) for x, y in zip(io[i], ig[i]))
        if same and not inl_same:
            # only residuals within 1e-6 (relative) of the threshold may differ
            r2, thr = oracle.residuals(kind, pr, opt, cfg, mo[i] if kind == 0 else _internal(oracle, kind, pr, opt, cfg, mo[i]))
            for t in range(3):
                diff = np.setxor1d(io[i][t], ig[i][t])
                assert np.all(np.abs(r2[t][diff] - thr[t]) <= 1e-6 * thr[t])
        if same:
            identical += 1
            assert rot_err_deg(mo[i][:9], mg[i][:9]) <= 0.01
            assert vec_angle_deg(mo[i][9:12], mg[i][9:12]) <= 0.01
            np.testing.assert_allclose(mg[i][12:17], mo[i][12:17], rtol=1e-6, atol=1e-9)
            assert abs(a.best_model_score - b.best_model_score) <= 1e-9 * abs(a.best_model_score)
            assert a.best_num_inliers == b.best_num_inliers or not inl_same
    assert identical >= min_identical, f"only {identical}/{len(pairs)} pairs followed the oracle's path"


def _internal(oracle, kind, pr, opt, cfg, m):
    ns = oracle.frontend(kind, pr, opt, cfg)["norm_scale"]
    m = m.copy(); m[15] /= ns; m[16] /= ns
    return m


@pytest.mark.parametrize("config,npairs", [(0, 6), (1, 8), (2, 4)])
def test_estimate_matches_oracle(ctx, oracle, config, npairs):
    ps = [problem(config, p=i) for i in range(npairs)]
    kind, opt, cfg = ps[0][0]["kind"], ps[0][2], ps[0][3]
    pairs = [q[1] for q in ps]
    mo, so, io = oracle.estimate_batch(kind, pairs, opt, cfg, threads=8)
    mg, sg, ig = ctx.estimate_batch(kind, pairs, opt, cfg)
    compare(oracle, kind, pairs, opt, cfg, mo, so, io, mg, sg, ig, min_identical=npairs - 1)
    # and the result is a good pose: rotation error against ground truth below 2 degrees
    for q, m in zip(ps, mg):
        assert rot_err_deg(q[0]["gt"]["R"].reshape(9), m[:9]) < 2.0


def test_large_pairs_multi_tile(ctx, oracle):
    """Pairs with more matches than one shared-memory tile of the fused scoring kernel holds (~7000): the tile size is
    what the static + dynamic shared memory of that kernel allows, and the scores are accumulated over the tiles."""
    ps = [problem(0, p=i, n=9000 + 500 * i) for i in range(2)] + [problem(0, p=5, n=400)]
    kind, opt, cfg = ps[0][0]["kind"], ps[0][2], ps[0][3]
    pairs = [q[1] for q in ps]
    mo, so, io = oracle.estimate_batch(kind, pairs, opt, cfg, threads=4)
    mg, sg, ig = ctx.estimate_batch(kind, pairs, opt, cfg)
    compare(oracle, kind, pairs, opt, cfg, mo, so, io, mg, sg, ig, min_identical=2)


@pytest.mark.parametrize("kw", [dict(final_least_squares=0), dict(random_seed=7), dict(lo_starting_iterations=0),
                                dict(num_lo_steps=1, num_lsq_iterations=2), dict(min_num_iterations=10, max_num_iterations=30),
                                dict(max_num_iterations_per_solver=60)])
def test_option_variants(ctx, oracle, kw):
    ps = [problem(1, p=20 + i, n=700, **kw) for i in range(3)]
    kind, opt, cfg = ps[0][0]["kind"], ps[0][2], ps[0][3]
    pairs = [q[1] for q in ps]
    mo, so, io = oracle.estimate_batch(kind, pairs, opt, cfg, threads=4)
    mg, sg, ig = ctx.estimate_batch(kind, pairs, opt, cfg)
    compare(oracle, kind, pairs, opt, cfg, mo, so, io, mg, sg, ig, min_identical=2)


@pytest.mark.parametrize("solver_type,score_type,lo_type", [(1, 0, 0), (2, 0, 0), (0, 1, 0), (0, 2, 0), (0, 0, 1), (0, 0, 2)])
def test_estimator_config_modes(ctx, oracle, solver_type, score_type, lo_type):
    ps = [problem(0, p=40 + i, n=600) for i in range(2)]
    kind, opt = 0, ps[0][2]
    cfg = capi.Config(solver_type=solver_type, score_type=score_type, lo_type=lo_type)
    pairs = [q[1] for q in ps]
    mo, so, io = oracle.estimate_batch(kind, pairs, opt, cfg, threads=2)
    mg, sg, ig = ctx.estimate_batch(kind, pairs, opt, cfg)
    compare(oracle, kind, pairs, opt, cfg, mo, so, io, mg, sg, ig, min_identical=1)


def test_ragged_and_degenerate_pairs(ctx, oracle):
    """tiny pairs: n below the sample sizes (default model, score DBL_MAX), n = 6 / 9 (Fisher-Yates sampling branch,
    zero-padded LO samples: quirk Q3), and a normal pair in the same batch."""
    sizes = [2, 4, 6, 9, 55, 400]
    ps = [problem(1, p=60 + i, n=n) for i, n in enumerate(sizes)]
    kind, opt, cfg = 1, ps[0][2], ps[0][3]
    pairs = [q[1] for q in ps]
    mo, so, io = oracle.estimate_batch(kind, pairs, opt, cfg, threads=2)
    mg, sg, ig = ctx.estimate_batch(kind, pairs, opt, cfg)
    # n = 2: infeasible for both solvers -> zero pose, best_solver_type -1, no iterations (hybrid_ransac.h:60-62)
    assert sg[0].best_solver_type == so[0].best_solver_type == -1
    assert sg[0].num_iterations_total == so[0].num_iterations_total == 0
    np.testing.assert_allclose(mg[0], mo[0], rtol=1e-12, atol=0)
    # n = 4: only the depth solver is feasible (prior of the 6-point solver zeroed by VerifyData)
    assert sg[1].num_iterations_per_solver[1] == so[1].num_iterations_per_solver[1] == 0
    compare(oracle, kind, pairs[1:], opt, cfg, mo[1:], so[1:], io[1:], mg[1:], sg[1:], ig[1:], min_identical=4)


def test_rejects_unimplemented_flags(ctx):
    d, pair, opt, cfg, keep = problem(0, p=0, n=100)
    opt.use_ours = 1
    with pytest.raises(RuntimeError, match="not implemented"):
        ctx.estimate_batch(0, [pair], opt, cfg)


def test_knob_and_helper_argument_checks(ctx):
    """Bad arguments come back as an error code + message, never as a launch."""
    with pytest.raises(RuntimeError, match="mp_set_batch"):
        ctx.set_batch(257)
    with pytest.raises(RuntimeError, match="mp_set_batch"):
        ctx.set_batch(-1)
    with pytest.raises(RuntimeError, match="mp_set_groups"):
        ctx.set_groups(-1)
    ctx.set_batch(0); ctx.set_groups(0)
    with pytest.raises(RuntimeError, match="mp_get_depths"):
        ctx.get_depths((0, 0), np.ones((4, 4), np.float32), np.zeros((3, 2)))
    with pytest.raises(ValueError):
        ctx.pose_errors(np.zeros((2, 4, 4)), np.zeros((3, 3, 3)), np.zeros((2, 3)))


def test_schedule_invariance(ctx):
    """The result of a pair must not depend on how the device loop is scheduled: iterations speculated per round
    (mp_set_batch), number of stream groups (mp_set_groups; 1 = everything on one stream), i.e. on which round a pair
    sits out while its local optimisation runs on a side stream.  Bit-for-bit."""
    ps = [problem(1, p=i, n=700) for i in range(24)]
    kind, opt, cfg = ps[0][0]["kind"], ps[0][2], ps[0][3]
    pairs = [q[1] for q in ps]
    ref = None
    try:
        for batch, groups in [(0, 0), (8, 1), (64, 5), (17, 24), (32, 2), (256, 1), (128, 3)]:
            ctx.set_batch(batch); ctx.set_groups(groups)
            m, s, inl = ctx.estimate_batch(kind, pairs, opt, cfg)
            sig = [(x.num_iterations_total, tuple(x.num_iterations_per_solver), x.number_lo_iterations, x.best_solver_type,
                    x.best_model_score, x.best_num_inliers) for x in s]
            if ref is None:
                ref = (m.copy(), sig, inl)
                continue
            assert sig == ref[1], f"statistics changed with batch={batch} groups={groups}"
            np.testing.assert_array_equal(m, ref[0])
            for a, b in zip(inl, ref[2]):
                for x, y in zip(a, b):
                    np.testing.assert_array_equal(x, y)
    finally:
        ctx.set_batch(0); ctx.set_groups(0)
