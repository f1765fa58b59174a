import sys, time, numpy as np
sys.path.insert(0, '/root/repo')
from tests import oracle_lib
from tests.helpers import problem
from madpose_b200 import capi
ora = oracle_lib.load(build=False)
ctx = capi.Context(0)
def rot_err(Ra, Rb):
    c = (np.trace(Ra.reshape(3,3).T @ Rb.reshape(3,3)) - 1) / 2
    return np.degrees(np.arccos(np.clip(c, -1, 1)))
for config, npairs in ((0, 4), (1, 6), (2, 3)):
    ps = [problem(config, p=i) for i in range(npairs)]
    kind = ps[0][0]["kind"]; opt, cfg = ps[0][2], ps[0][3]
    pairs = [q[1] for q in ps]
    t0 = time.time(); mo, so, io = ora.estimate_batch(kind, pairs, opt, cfg, threads=8); t1 = time.time()
    mg, sg, ig = ctx.estimate_batch(kind, pairs, opt, cfg); t2 = time.time()
    print("config", config, "oracle %.2fs gpu %.3fs" % (t1 - t0, t2 - t1), ctx.last_timing())
    for i in range(npairs):
        a, b = so[i], sg[i]
        same_inl = all(np.array_equal(x, y) for x, y in zip(io[i], ig[i]))
        print("  pair", i, "iters", a.num_iterations_total, b.num_iterations_total, list(a.num_iterations_per_solver), list(b.num_iterations_per_solver),
              "lo", a.number_lo_iterations, b.number_lo_iterations, "inl", a.best_num_inliers, b.best_num_inliers, "same_inl", same_inl,
              "score %.6f %.6f" % (a.best_model_score, b.best_model_score), "rot_err %.2e" % rot_err(mo[i][:9], mg[i][:9]),
              "max rel %.2e" % np.max(np.abs(mo[i] - mg[i]) / np.maximum(np.abs(mo[i]), 1e-3)),
              "hyps", a.hypotheses_scored, b.hypotheses_scored, "lm", a.lm_solves, b.lm_solves, a.lm_iterations, b.lm_iterations)
