"""The oracle's control loop (oracle/om_ransac.h) against (1) the known-answer tests of SURVEY.md Appendix C that
were produced by compiling the reference's own RansacLib headers, and (2) oracle/_ref: the REFERENCE'S OWN
HybridLOMSAC loop (/root/reference/src/hybrid_ransac.h + RansacLib sampler/utils, compiled where they lie)
driving the same solver functions -- both must give identical results, iteration for iteration."""
import ctypes as C
import os

import numpy as np
import pytest

from tests import oracle_lib
from tests.helpers import problem

REF = os.path.join(oracle_lib.ORACLE_DIR, "_ref", "liboracle_ref.so")


def test_kat_num_required_iterations(oracle):
    f = oracle.lib.om_num_required_iterations
    f.restype = C.c_uint
    f.argtypes = [C.c_void_p, C.c_double, C.c_void_p, C.c_uint, C.c_uint]

    def nri(r, sizes, mn, mx):
        r = np.array(r, np.float64); s = np.array(sizes, np.int32)
        return f(r.ctypes.data, 1e-4, s.ctypes.data, mn, mx)
    p = 1.0 - 0.9999  # success_probability 0.9999 as the estimator passes it
    f2 = lambda r, s, mn, mx: f(np.array(r, np.float64).ctypes.data, p, np.array(s, np.int32).ctypes.data, mn, mx)
    assert f2([0.6, 0.6, 0.6], [3, 3, 0], 100, 10000) == 194
    assert f2([0.6, 0.6, 0.6], [0, 0, 5], 100, 10000) == 115
    assert f2([0.3, 0.3, 0.3], [3, 3, 0], 1000, 100000) == 12631
    assert f2([0.3, 0.3, 0.3], [0, 0, 5], 1000, 100000) == 3787
    assert f2([0.5, 0.45, 0.55], [4, 4, 0], 100, 10000) == 3590
    assert f2([0.5, 0.45, 0.55], [0, 0, 6], 100, 10000) == 329
    assert f2([0.5, 0.45, 0.55], [0, 0, 7], 100, 10000) == 601
    assert f2([0.0, 0.5, 0.5], [3, 3, 0], 100, 10000) == 10000
    assert f2([1.0, 1.0, 1.0], [3, 3, 0], 100, 10000) == 100


def test_kat_shuffle_resize_grows_with_zeros(oracle):
    """mt19937(0): RandomShuffleAndResize(6, {0..9}) -> 5 6 7 8 2 9; then (6, {5,6,7}) -> 6 7 5 0 0 0 (quirk Q3)."""
    a = np.arange(10, dtype=np.int32); b = np.array([5, 6, 7, 0, 0, 0], np.int32)
    oracle.lib.om_shuffle_resize(0, 6, 10, a.ctypes.data_as(C.c_void_p), 6, 3, b.ctypes.data_as(C.c_void_p))
    assert list(a[:6]) == [5, 6, 7, 8, 2, 9]
    assert list(b) == [6, 7, 5, 0, 0, 0]


@pytest.mark.skipif(not os.path.exists(REF), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("config,npairs,n", [(0, 3, None), (1, 3, 900), (2, 2, 1200), (1, 2, 40)])
def test_oracle_loop_equals_reference_loop(oracle, config, npairs, n):
    ref = C.CDLL(REF)
    ps = [problem(config, p=i, n=n) for i in range(npairs)]
    kind, opt, cfg = ps[0][0]["kind"], ps[0][2], ps[0][3]
    pairs = [q[1] for q in ps]
    mo, so, io = oracle.estimate_batch(kind, pairs, opt, cfg, threads=4)
    mr, sr, ir = oracle.estimate_batch(kind, pairs, opt, cfg, threads=4, lib=ref)
    for i in range(npairs):
        a, b = so[i], sr[i]
        assert a.num_iterations_total == b.num_iterations_total
        assert list(a.num_iterations_per_solver) == list(b.num_iterations_per_solver)
        assert a.number_lo_iterations == b.number_lo_iterations and a.best_solver_type == b.best_solver_type
        assert a.best_num_inliers == b.best_num_inliers and a.best_model_score == b.best_model_score
        assert list(a.inlier_ratios) == list(b.inlier_ratios)
        assert all(np.array_equal(x, y) for x, y in zip(io[i], ir[i]))
        assert np.array_equal(mo[i], mr[i])  # bit-identical models


@pytest.mark.parametrize("config", [0, 1, 2])
def test_oracle_recovers_ground_truth(oracle, config):
    d, pair, opt, cfg, keep = problem(config, p=11)
    m, s, inl = oracle.estimate_batch(d["kind"], [pair], opt, cfg)
    R = m[0][:9].reshape(3, 3)
    ang = np.degrees(np.arccos(np.clip((np.trace(R.T @ d["gt"]["R"]) - 1) / 2, -1, 1)))
    assert ang < 1.0
    t = m[0][9:12]
    assert np.degrees(np.arccos(np.clip(t @ d["gt"]["t"] / np.linalg.norm(t) / np.linalg.norm(d["gt"]["t"]), -1, 1))) < 3.0
    assert abs(m[0][12] / d["gt"]["scale"] - 1) < 0.05
    if config > 0:
        assert abs(m[0][15] / d["gt"]["f0"] - 1) < 0.05 and abs(m[0][16] / d["gt"]["f1"] - 1) < 0.05
    assert s[0].best_num_inliers > 0.4 * 3 * pair.n
