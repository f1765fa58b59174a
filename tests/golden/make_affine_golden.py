"""Generates tests/golden/affine_solvers.json by RUNNING the reference's own numpy
solvers (/root/reference/solver_py/*.py) on seeded synthetic instances built with
the recipe of their test_solver() (solver_py/scale_and_shift.py:131-155,
..._shared_focal.py:308-338, ..._two_focal.py:367-400).  Run in the build
container only (the reference does not travel to the GPU box); the JSON is
committed.  `madpose` is stubbed because the numpy solvers only import it for
their print-only demo."""
import json
import os
import sys
import types

import numpy as np

sys.modules.setdefault("madpose", types.ModuleType("madpose"))
sys.path.insert(0, "/root/reference/solver_py")
import scale_and_shift as s0  # noqa: E402
import scale_and_shift_shared_focal as s1  # noqa: E402
import scale_and_shift_two_focal as s2  # noqa: E402


def instance(rng, kind):
    n = 3 if kind == 0 else 4
    while True:
        x1 = np.c_[rng.standard_normal((n, 2)), np.ones((n,))]
        f1 = 1000.0 + 2000.0 * rng.random()
        f2 = f1 if kind == 1 else 1000.0 + 2000.0 * rng.random()
        d1_gt = 1.0 + 5 * rng.random(n)
        X = x1 * d1_gt[:, None]
        R = np.linalg.qr(rng.standard_normal((3, 3)))[0]
        R = R * np.linalg.det(R)
        t = rng.standard_normal(3)
        X2 = X @ R.T + t
        d2_gt = X2[:, 2]
        x2 = X2 / d2_gt[:, None]
        if kind > 0:
            x1[:, 0:2] *= f1
            x2[:, 0:2] *= f2
            x2[:, 0:2] += 0.5 * rng.standard_normal((n, 2)) - 0.25
        if np.all(d2_gt > 0):
            break
    a1, a2 = rng.random(), rng.random()
    b1, b2 = rng.standard_normal(), rng.standard_normal()
    d1 = (d1_gt - b1) / a1
    d2 = (d2_gt - b2) / a2
    return x1, x2, d1, d2


def main():
    out = []
    fns = [s0.solve_shift_and_scale, s1.solve_shift_and_scale_shared_focal, s2.solve_shift_and_scale_two_focal]
    for kind in range(3):
        rng = np.random.default_rng(1234 + kind)
        for _ in range(40):
            x1, x2, d1, d2 = instance(rng, kind)
            sols = fns[kind](x1.copy(), x2.copy(), d1.copy(), d2.copy())
            sols = [[float(np.real(v)) for v in s] for s in sols]
            sols = [s for s in sols if all(np.isfinite(s))]
            out.append(dict(kind=kind, x1=x1.tolist(), x2=x2.tolist(), d1=d1.tolist(), d2=d2.tolist(), sols=sols))
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "affine_solvers.json")
    with open(path, "w") as f:
        json.dump(out, f)
    print("wrote", path, len(out))


if __name__ == "__main__":
    main()
