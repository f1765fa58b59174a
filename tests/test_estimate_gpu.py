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


def _parse_oracle_trace(tr):
    """Oracle trace records (oracle/om_ransac.h): (it, kind, a, b), kind >= 0 minimal iteration (b = local best score),
    -1 ScoreModel in LO (b), -2 GetInliers (a = count, b = min |r^2 - thr| / thr of the pass), -3 LM (a = iterations, b = cost)."""
    ev = []
    for k in range(0, len(tr) - 3, 4):
        if tr[k] < 0:
            break
        ev.append((int(tr[k]), int(tr[k + 1]), tr[k + 2], tr[k + 3]))
    return ev


def _parse_device_trace(ed):
    """Device records (mp_loop.cuh): {it, score} | {-1, score} | {-2, count} | {-3, iterations}{-4, cost} -> same tuples."""
    ev, k = [], 0
    while k < len(ed):
        tag, val = ed[k]
        if tag >= 0:
            ev.append((int(tag), 0, 0.0, val))
        elif tag == -1:
            ev.append((-1, -1, 0.0, val))
        elif tag == -2:
            ev.append((-1, -2, val, 0.0))
        elif tag == -3:
            ev.append((-1, -3, val, ed[k + 1][1])); k += 1
        k += 1
    return ev


TIE = 1e-7  # two fp64 scores this close (relative) count as the same decision input (LM end points agree to ~1e-8)


def _cnt3(v):
    v = int(v)
    return (v % 131072, (v >> 17) % 131072, v >> 34)


def explain_fork(ctx, oracle, kind, pair, opt, cfg):
    """A pair whose device run left the oracle's path: replay it alone on both sides with event traces and classify the
    FIRST event where the two runs differ.  The device's events are a chronological subsequence of the oracle's (it
    re-scores only the iterations that pass the fp32 filter).  A fork is explained when that first difference is a
    comparison between values that agree to rounding:
      "inlier tie"   a GetInliers pass whose count differs while the oracle saw a residual within 1e-9 (relative) of the
                     threshold in that pass (the match falls on the other side on the device; the sample drawn from the
                     list, and everything after it, then differs);
      "LM tie"       an LM solve on identical inputs whose iteration count differs by a convergence test at its
                     tolerance: the final costs agree to the function tolerance;
      "unconverged LM"  an LM solve on identical inputs (all earlier events agree) that used all ceres_max_num_iterations
                     iterations without converging on both sides: such a trajectory on an ill-conditioned sample (a handful
                     of reprojection blocks) amplifies rounding differences; the end points differ;
      "score tie"    a decision (new best / update) taken differently on scores that agree to 1e-7.
    Returns (ok, what)."""
    _, _, _, tr = oracle.estimate_batch(kind, [pair], opt, cfg, trace_cap=1 << 22)
    eo = _parse_oracle_trace(tr)
    ctx.set_trace(1 << 17)
    try:
        ctx.estimate_batch(kind, [pair], opt, cfg)
        raw, n = ctx.read_trace(0, cap=1 << 17)
    finally:
        ctx.set_trace(0)
    assert n <= (1 << 17), "trace capacity exceeded"
    ed = _parse_device_trace(raw)
    oi = 0
    for it, knd, a, b in ed:
        # advance the oracle to the matching event; minimal iterations the device did not re-score are skipped
        while oi < len(eo) and eo[oi][1] >= 0 and not (knd >= 0 and eo[oi][0] == it):
            oi += 1
        if oi == len(eo):
            return True, "score tie: the device continues past the oracle's last event (all earlier events agree)"
        o_it, o_k, o_a, o_b = eo[oi]
        if knd >= 0 and o_k < 0:
            return True, f"score tie: the oracle ran a local optimisation before iteration {it} that the device did not (all earlier events agree)"
        if knd < 0 and o_k != knd:
            return False, f"event kinds differ at oracle event {oi}: device {knd}, oracle {o_k}"
        if knd == -2:
            if a != o_a:
                if o_b <= 1e-8:
                    return True, f"inlier tie at oracle event {oi} (it {o_it}): counts {_cnt3(a)} vs {_cnt3(o_a)}, a residual within {o_b:.1e} of its threshold"
                return False, f"inlier counts differ at oracle event {oi} (it {o_it}): {_cnt3(a)} vs {_cnt3(o_a)}, closest residual {o_b:.2e} from its threshold"
        elif knd == -3:
            rel = abs(b - o_b) / max(abs(o_b), 1e-300)
            if a != o_a:
                ftol = max(cfg.ceres_function_tolerance, 1e-12)
                if rel <= 10 * ftol:
                    return True, f"LM tie at oracle event {oi} (it {o_it}): {int(a)} vs {int(o_a)} iterations, costs agree to {rel:.1e}"
                return False, f"LM iterations differ at oracle event {oi} (it {o_it}): {int(a)} vs {int(o_a)}, costs {b!r} vs {o_b!r}"
            if rel > 1e-7:
                if a >= cfg.ceres_max_num_iterations:
                    return True, (f"unconverged LM at oracle event {oi} (it {o_it}): both solves used all {int(a)} iterations on identical "
                                  f"inputs without converging (ill-conditioned sample, non-monotonic steps), end costs {b!r} vs {o_b!r}")
                return False, f"LM costs differ at oracle event {oi} (it {o_it}): {b!r} vs {o_b!r}"
        else:
            ok = (b == o_b) or (np.isfinite(o_b) and abs(b - o_b) <= TIE * abs(o_b))
            if not ok:
                return False, f"scores differ at oracle event {oi} (it {o_it}, {'LO' if knd < 0 else 'minimal'}): {b!r} vs {o_b!r}"
        oi += 1
    if any(e[1] < 0 for e in eo[oi:]):
        return True, "score tie: the oracle continues with a local optimisation the device did not run (all earlier events agree)"
    return True, "score tie: every traced event agrees; the runs differ only after the last traced decision"


# A pair may leave the oracle's path only through an fp64 near-tie (explain_fork); both runs then end at statistically
# equivalent models.  Bounds checked for every such pair:
FORK_INLIER_REL = 0.01   # relative difference of the final inlier counts
FORK_ROT_DEG = 0.5       # rotation between the two final models


def compare(oracle, kind, pairs, opt, cfg, mo, so, io, mg, sg, ig, max_forks, ctx=None):
    """Every pair is checked: pairs on the oracle's path to the north_star tolerances, the others (at most max_forks)
    must be explained as near-ties by explain_fork and end at equivalent models."""
    forks = []
    for i, pr in enumerate(pairs):
        a, b = so[i], sg[i]
        same = (a.num_iterations_total == b.num_iterations_total
                and list(a.num_iterations_per_solver) == list(b.num_iterations_per_solver)
                and a.number_lo_iterations == b.number_lo_iterations and a.best_solver_type == b.best_solver_type)
        inl_same = all(np.array_equal(x, y) for x, y in zip(io[i], ig[i]))
        if same and not inl_same:
            # only residuals within 1e-6 (relative) of the threshold may differ
            r2, thr = oracle.residuals(kind, pr, opt, cfg, _internal(oracle, kind, pr, opt, cfg, mo[i]))
            for t in range(3):
                diff = np.setxor1d(io[i][t], ig[i][t])
                assert np.all(np.abs(r2[t][diff] - thr[t]) <= 1e-6 * thr[t])
        if same and rot_err_deg(mo[i][:9], mg[i][:9]) <= 0.01 and vec_angle_deg(mo[i][9:12], mg[i][9:12]) <= 0.01 \
                and np.allclose(mg[i][12:17], mo[i][12:17], rtol=1e-6, atol=1e-9):
            assert abs(a.best_model_score - b.best_model_score) <= 1e-9 * abs(a.best_model_score)
            assert a.best_num_inliers == b.best_num_inliers or not inl_same
            continue
        forks.append(i)
    assert len(forks) <= max_forks, f"{len(forks)}/{len(pairs)} pairs left the oracle's path: {forks[:20]}"
    report = []
    for i in forks:
        assert ctx is not None, f"pair {i} left the oracle's path"
        ok, what = explain_fork(ctx, oracle, kind, pairs[i], opt, cfg)
        a, b = so[i], sg[i]
        d_inl = abs(a.best_num_inliers - b.best_num_inliers) / max(a.best_num_inliers, 1)
        d_rot = rot_err_deg(mo[i][:9], mg[i][:9])
        report.append((i, ok, d_inl, d_rot, what))
    bad = [r for r in report if not r[1] or r[2] > FORK_INLIER_REL or r[3] > FORK_ROT_DEG]
    assert not bad, f"forks that are not explained near-ties ending at equivalent models: {bad}"
    return report


def _internal(oracle, kind, pr, opt, cfg, m):
    if kind in (0, 3):
        return m
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
    compare(oracle, kind, pairs, opt, cfg, mo, so, io, mg, sg, ig, max_forks=1, ctx=ctx)
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
    compare(oracle, kind, pairs, opt, cfg, mo, so, io, mg, sg, ig, max_forks=1, ctx=ctx)


@pytest.mark.parametrize("kw", [dict(final_least_squares=0), dict(random_seed=7), dict(lo_starting_iterations=0),
                                dict(num_lo_steps=1, num_lsq_iterations=2), dict(min_num_iterations=10, max_num_iterations=30),
                                dict(max_num_iterations_per_solver=60)])
def test_option_variants(ctx, oracle, kw):
    ps = [problem(1, p=20 + i, n=700, **kw) for i in range(3)]
    kind, opt, cfg = ps[0][0]["kind"], ps[0][2], ps[0][3]
    pairs = [q[1] for q in ps]
    mo, so, io = oracle.estimate_batch(kind, pairs, opt, cfg, threads=4)
    mg, sg, ig = ctx.estimate_batch(kind, pairs, opt, cfg)
    compare(oracle, kind, pairs, opt, cfg, mo, so, io, mg, sg, ig, max_forks=1, ctx=ctx)


@pytest.mark.parametrize("solver_type,score_type,lo_type", [(1, 0, 0), (2, 0, 0), (0, 1, 0), (0, 2, 0), (0, 0, 1), (0, 0, 2)])
def test_estimator_config_modes(ctx, oracle, solver_type, score_type, lo_type):
    ps = [problem(0, p=40 + i, n=600) for i in range(2)]
    kind, opt = 0, ps[0][2]
    cfg = capi.Config(solver_type=solver_type, score_type=score_type, lo_type=lo_type)
    pairs = [q[1] for q in ps]
    mo, so, io = oracle.estimate_batch(kind, pairs, opt, cfg, threads=2)
    mg, sg, ig = ctx.estimate_batch(kind, pairs, opt, cfg)
    compare(oracle, kind, pairs, opt, cfg, mo, so, io, mg, sg, ig, max_forks=1, ctx=ctx)


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
    compare(oracle, kind, pairs[1:], opt, cfg, mo[1:], so[1:], io[1:], mg[1:], sg[1:], ig[1:], max_forks=1, ctx=ctx)


@pytest.mark.parametrize("name,config,kind,cfgkw", [("calib_noshift", 0, 0, dict(use_shift=0)),
                                                   ("scale_only", 0, capi.KIND_SCALE_ONLY, {}),
                                                   ("shared_noshift", 1, 1, dict(use_shift=0)),
                                                   ("two_noshift", 2, 2, dict(use_shift=0))])
def test_variants_match_oracle(ctx, oracle, name, config, kind, cfgkw):
    """EstimatorConfig.use_shift = False (calibrated: other minimal solvers + constant offsets; shared focal: constant
    offsets in NonMinimalSolver only; two focal: ignored) and the scale-only estimator, whole path."""
    ps = [problem(config, p=70 + i, n=700, zero_offsets=True, cfg=cfgkw) for i in range(4)]
    opt, cfg = ps[0][2], ps[0][3]
    pairs = [q[1] for q in ps]
    mo, so, io = oracle.estimate_batch(kind, pairs, opt, cfg, threads=4)
    mg, sg, ig = ctx.estimate_batch(kind, pairs, opt, cfg)
    compare(oracle, kind, pairs, opt, cfg, mo, so, io, mg, sg, ig, max_forks=1, ctx=ctx)
    if name in ("calib_noshift", "scale_only"):
        assert np.all(mg[:, 13:15] == 0)
        for q, m in zip(ps, mg):
            assert rot_err_deg(q[0]["gt"]["R"].reshape(9), m[:9]) < 2.0


# BASELINE.json configs at full size: every pair checked, forks bounded and explained.
# (config, pairs, share of the pairs that may fork through an fp64 near-tie)
@pytest.mark.parametrize("config,npairs,max_fork_share", [(1, 1500, 0.01), (2, 200, 0.015), (0, 300, 0.01), (3, 16, 0.07)])
def test_full_size_parity(ctx, oracle, config, npairs, max_fork_share):
    import json
    import os
    ps = [problem(config, p=i) for i in range(npairs)]
    kind, opt, cfg = ps[0][0]["kind"], ps[0][2], ps[0][3]
    pairs = [q[1] for q in ps]
    mo, so, io = oracle.estimate_batch(kind, pairs, opt, cfg, threads=os.cpu_count() or 8)
    mg, sg, ig = ctx.estimate_batch(kind, pairs, opt, cfg)
    report = compare(oracle, kind, pairs, opt, cfg, mo, so, io, mg, sg, ig, max_forks=max(1, int(np.ceil(max_fork_share * npairs))), ctx=ctx)
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if os.path.isdir(out):  # kept as evidence next to the bench lines (profiles/)
        with open(os.path.join(out, f"parity_full_config{config}.json"), "w") as f:
            json.dump({"config": config, "pairs": npairs, "on_oracle_path": npairs - len(report),
                       "forks": [dict(pair=i, explained=bool(ok), inlier_rel=di, rot_deg=dr, what=w) for i, ok, di, dr, w in report]}, f)


def test_resident_batch_guards(ctx):
    """mp_estimate_resident after a per-subsystem call (which overwrites the resident buffers) or with thresholds other
    than the uploaded ones must fail instead of reading stale / foreign data."""
    ps = [problem(1, p=i, n=300) for i in range(3)]
    kind, opt, cfg = 1, ps[0][2], ps[0][3]
    pairs = [q[1] for q in ps]
    ctx.upload_batch(kind, pairs, opt)
    m0, s0, _ = ctx.estimate_resident(opt, cfg)
    opt2 = capi.Options(**{**{k: getattr(opt, k) for k in ("min_num_iterations", "max_num_iterations", "final_least_squares")},
                           "squared_inlier_thresholds": (16.0, 4.0), "data_type_weights": (1.0, 1.0)})
    with pytest.raises(RuntimeError, match="upload again"):
        ctx.estimate_resident(opt2, cfg)
    ctx.score_batch(kind, pairs[0], opt, cfg, np.array([m0[0]]))  # overwrites the resident buffers with one pair
    with pytest.raises(RuntimeError, match="no resident batch"):
        ctx.estimate_resident(opt, cfg)
    ctx.upload_batch(kind, pairs, opt)
    m1, s1, _ = ctx.estimate_resident(opt, cfg)
    np.testing.assert_array_equal(m0, m1)


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


@pytest.mark.parametrize("mode", [1, 2])
def test_minimal_modes_whole_path(ctx, oracle, mode):
    """Both forms of the 6-point solver (one thread / one warp per sample) through the whole path against the oracle."""
    ps = [problem(1, p=100 + i, n=800) for i in range(6)]
    kind, opt, cfg = 1, ps[0][2], ps[0][3]
    pairs = [q[1] for q in ps]
    mo, so, io = oracle.estimate_batch(kind, pairs, opt, cfg, threads=6)
    ctx.set_minimal_mode(mode)
    try:
        mg, sg, ig = ctx.estimate_batch(kind, pairs, opt, cfg)
        compare(oracle, kind, pairs, opt, cfg, mo, so, io, mg, sg, ig, max_forks=1, ctx=ctx)
    finally:
        ctx.set_minimal_mode(0)


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


def test_mixed_width_batch_is_batch_invariant(ctx):
    """Local optimisation runs on 256-thread CTAs for pairs with >= 4096 matches and on 128-thread CTAs for the others
    (LP_WIDE_N, mp_loop.cuh); the width depends on the pair alone.  A batch mixing both classes must give every pair the
    bits it gets when estimated alone or with pairs of its own class only."""
    sizes = [900, 4096, 1500, 6000, 4095, 5200, 300, 4500]
    ps = [problem(0, p=i, n=n) for i, n in enumerate(sizes)]
    kind, opt, cfg = ps[0][0]["kind"], ps[0][2], ps[0][3]
    pairs = [q[1] for q in ps]

    def sig(m, s, inl):
        return [(tuple(m[i]), s[i].num_iterations_total, s[i].number_lo_iterations, s[i].best_model_score,
                 tuple(int(x) for x in s[i].num_inliers), tuple(a.tobytes() for a in inl[i])) for i in range(len(s))]

    mixed = sig(*ctx.estimate_batch(kind, pairs, opt, cfg))
    assert any(s_[2] > 0 for s_ in mixed)  # local optimisations did run
    for i, p in enumerate(pairs):  # alone
        assert sig(*ctx.estimate_batch(kind, [p], opt, cfg))[0] == mixed[i], f"pair {i} (n = {sizes[i]}) differs when estimated alone"
    for cls in (False, True):  # with its own width class only
        idx = [i for i, n in enumerate(sizes) if (n >= 4096) == cls]
        got = sig(*ctx.estimate_batch(kind, [pairs[i] for i in idx], opt, cfg))
        assert got == [mixed[i] for i in idx]
