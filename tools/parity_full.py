#!/usr/bin/env python
"""Whole-path parity at BASELINE size: GPU loop vs the CPU oracle on all pairs of a config.
usage: python tools/parity_full.py [config=1] [pairs=1500]  -> one JSON line (commit it under profiles/)."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madpose_b200 import capi  # noqa: E402
from tests import oracle_lib  # noqa: E402
from tests.helpers import problem  # noqa: E402


def rot_err(Ra, Rb):
    if not np.any(Ra) and not np.any(Rb):
        return 0.0
    c = (np.trace(Ra.reshape(3, 3).T @ Rb.reshape(3, 3)) - 1) / 2
    return float(np.degrees(np.arccos(np.clip(c, -1, 1))))


def main():
    config = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    npairs = int(sys.argv[2]) if len(sys.argv) > 2 else 1500
    ps = [problem(config, p=i) for i in range(npairs)]
    kind, opt, cfg = ps[0][0]["kind"], ps[0][2], ps[0][3]
    pairs = [q[1] for q in ps]
    ora = oracle_lib.load(build=False)
    ctx = capi.Context(0)
    t0 = time.time()
    mo, so, io = ora.estimate_batch(kind, pairs, opt, cfg, threads=os.cpu_count() or 1)
    t1 = time.time()
    mg, sg, ig = ctx.estimate_batch(kind, pairs, opt, cfg)
    t2 = time.time()
    same_path = same_inl = 0
    max_rel, max_rot, off_rot, off_rel, off_inl = 0.0, 0.0, [], [], []
    for i in range(npairs):
        a, b = so[i], sg[i]
        path = (a.num_iterations_total == b.num_iterations_total
                and list(a.num_iterations_per_solver) == list(b.num_iterations_per_solver)
                and a.number_lo_iterations == b.number_lo_iterations and a.best_solver_type == b.best_solver_type)
        inl = all(np.array_equal(x, y) for x, y in zip(io[i], ig[i]))
        re = rot_err(mo[i][:9], mg[i][:9])
        rel = float(np.max(np.abs(mo[i][12:17] - mg[i][12:17]) / np.maximum(np.abs(mo[i][12:17]), 1e-3)))
        if path:
            same_path += 1
            same_inl += int(inl)
            max_rel, max_rot = max(max_rel, rel), max(max_rot, re)
        else:
            off_rot.append(re); off_rel.append(rel)
            off_inl.append(abs(a.best_num_inliers - b.best_num_inliers) / max(a.best_num_inliers, 1))
    out = {"config": config, "pairs": npairs, "same_path": same_path, "same_path_and_inlier_sets": same_inl,
           "max_rot_err_deg_same_path": max_rot, "max_rel_scale_offsets_focals_same_path": max_rel,
           "other_path": npairs - same_path,
           "other_path_max_rot_err_deg": max(off_rot) if off_rot else 0.0,
           "other_path_max_rel": max(off_rel) if off_rel else 0.0,
           "other_path_max_rel_inlier_count_diff": max(off_inl) if off_inl else 0.0,
           "oracle_s": t1 - t0, "gpu_s": t2 - t1}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
