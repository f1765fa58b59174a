"""K5 parity: the batched Levenberg-Marquardt kernel (mp_lm_batch, analytic Jacobians, one CTA per problem)
vs the oracle's LM (forward-mode Jets): same iteration count, termination, and final parameters on
inlier-only problems (the ones local optimisation issues)."""
import numpy as np
import pytest

from tests.helpers import models_around_gt, problem

pytestmark = pytest.mark.gpu


def _problems(d, pair, count, seed):
    rng = np.random.default_rng(seed)
    inl = np.nonzero(d["gt"]["inlier_mask"])[0]
    out = []
    for k in range(count):
        if k % 4 == 0:
            out.append([inl, inl, inl])                                    # all-inlier fit (LeastSquaresFit, first call)
        elif k % 4 == 1:
            out.append([rng.choice(inl, 98), rng.choice(inl, 98), rng.choice(inl, 98)])   # 294-style sample with repeats
        elif k % 4 == 2:
            out.append([rng.choice(inl, 40), np.zeros(0, int), rng.choice(inl, 30)])      # no type-1 blocks: scale frozen
        else:
            out.append([rng.choice(inl, 2), rng.choice(inl, 2), rng.choice(inl, 3)])      # guard rejects: model untouched
    return out


@pytest.mark.parametrize("config", [0, 1, 2])
def test_lm_matches_oracle(ctx, oracle, config):
    d, pair, opt, cfg, keep = problem(config, p=4)
    kind = d["kind"]
    fe = oracle.frontend(kind, pair, opt, cfg)
    starts = models_around_gt(d, fe["norm_scale"], 16, seed=3, spread=0.12)
    samples = _problems(d, pair, len(starts), 9)
    ref, iref = oracle.lm(kind, pair, opt, cfg, starts, samples)
    got, igot = ctx.lm_batch(kind, pair, opt, cfg, starts, samples)
    np.testing.assert_array_equal(iref[:, :2], igot[:, :2])        # iterations, termination type
    np.testing.assert_allclose(igot[:, 2], iref[:, 2], rtol=1e-10)  # initial cost: same residuals
    conv = iref[:, 1] != 2                                          # runs that hit max iterations are path-chaotic
    err = np.max(np.abs(ref - got) / np.maximum(np.abs(ref), 1e-3), axis=1)
    assert np.all(err[conv] <= 1e-8), err
    guard = iref[:, 1] == 0
    assert guard.sum() == 4 and np.array_equal(got[guard], np.asarray(starts)[guard])  # guard: untouched, bit for bit


@pytest.mark.parametrize("lo_type,use_shift", [(1, 1), (2, 1), (0, 0)])
def test_lm_config_variants(ctx, oracle, lo_type, use_shift):
    d, pair, opt, cfg, keep = problem(0, p=2, n=800)
    cfg.lo_type, cfg.use_shift = lo_type, use_shift
    fe = oracle.frontend(0, pair, opt, cfg)
    starts = models_around_gt(d, fe["norm_scale"], 6, seed=1, spread=0.1)
    inl = np.nonzero(d["gt"]["inlier_mask"])[0]
    samples = [[inl, inl, inl]] * len(starts)
    ref, iref = oracle.lm(0, pair, opt, cfg, starts, samples)
    got, igot = ctx.lm_batch(0, pair, opt, cfg, starts, samples)
    np.testing.assert_array_equal(iref[:, :2], igot[:, :2])
    np.testing.assert_allclose(got, ref, rtol=1e-8, atol=1e-10)


def test_lm_infeasible_start_is_noop(ctx, oracle):
    d, pair, opt, cfg, keep = problem(0, p=2, n=400)
    fe = oracle.frontend(0, pair, opt, cfg)
    m = models_around_gt(d, fe["norm_scale"], 1)[0].copy()
    m[13] = -d["min_depth"][0] + 1e-3   # offset0 inside the 1e-2 margin of its bound
    inl = np.nonzero(d["gt"]["inlier_mask"])[0]
    ref, iref = oracle.lm(0, pair, opt, cfg, [m], [[inl, inl, inl]])
    got, igot = ctx.lm_batch(0, pair, opt, cfg, [m], [[inl, inl, inl]])
    assert iref[0, 1] == 4 and igot[0, 1] == 4
    np.testing.assert_allclose(got, ref, rtol=1e-14, atol=1e-15)
