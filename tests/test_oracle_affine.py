"""Oracle affine minimal solvers vs the golden vectors produced by RUNNING the reference's
own numpy solvers (tests/golden/make_affine_golden.py, /root/reference/solver_py/*.py)."""
import json
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(__file__), "golden", "affine_solvers.json")


def test_affine_solvers_match_reference_numpy(oracle):
    cases = json.load(open(GOLD))
    assert len(cases) == 120
    n_ref = 0
    for g in cases:
        k = g["kind"]
        mine = oracle.solve_affine(k, g["x1"], g["x2"], g["d1"], g["d2"])
        ref = np.array(g["sols"]) if g["sols"] else np.zeros((0, 4))
        cols = 3 + k  # b1, a2, b2*a2 [, f | f1, f2]
        assert len(mine) == len(ref), "solution count differs from the reference solver"
        for r in ref:
            n_ref += 1
            r = r[1:1 + cols]
            err = np.max(np.abs(mine[:, :cols] - r) / np.maximum(np.abs(r), 1e-300), axis=1)
            # the reference's own 36x36 / 40x40 template solves carry ~1e-9 error; 1e-7 pins the root
            assert err.min() < 1e-7
    assert n_ref > 250
