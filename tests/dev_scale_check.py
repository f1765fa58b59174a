import sys, time, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import oracle_lib
from tests.helpers import problem
from madpose_b200 import capi
config = int(sys.argv[1]); npairs = int(sys.argv[2]); check = int(sys.argv[3]) if len(sys.argv) > 3 else 1
ora = oracle_lib.load(build=False)
ctx = capi.Context(0)
ctx.set_groups(int(os.environ.get('MP_GROUPS', '0')))
t0 = time.time()
ps = [problem(config, p=i) for i in range(npairs)]
print("gen %.1fs" % (time.time() - t0))
kind = ps[0][0]["kind"]; opt, cfg = ps[0][2], ps[0][3]
pairs = [q[1] for q in ps]
for rep in range(int(os.environ.get("REPS", "3"))):
    t1 = time.time(); mg, sg, ig = ctx.estimate_batch(kind, pairs, opt, cfg); t2 = time.time()
    k, t, l = ctx.last_timing()
    print("gpu wall %.3fs kernel %.1f ms total %.1f ms launches %d -> %.0f pairs/s" % (t2 - t1, k, t, l, npairs / (t * 1e-3)))
print("phase Mcycles", {k: round(v / 1e6, 1) for k, v in ctx.last_phase_cycles().items()})
print("iters mean", np.mean([s.num_iterations_total for s in sg]), "lo mean", np.mean([s.number_lo_iterations for s in sg]),
      "hyps", sum(s.hypotheses_scored for s in sg), "evals %.3e" % sum(s.residual_evals for s in sg), "lm", sum(s.lm_solves for s in sg), sum(s.lm_iterations for s in sg))
if check:
    nt = os.cpu_count()
    t0 = time.time(); mo, so, io = ora.estimate_batch(kind, pairs, opt, cfg, threads=nt); t1 = time.time()
    print("oracle %d threads %.2fs -> %.1f pairs/s" % (nt, t1 - t0, npairs / (t1 - t0)))
    bad = 0; worst = 0
    for i in range(npairs):
        a, b = so[i], sg[i]
        same = (a.num_iterations_total == b.num_iterations_total and a.number_lo_iterations == b.number_lo_iterations and
                all(np.array_equal(x, y) for x, y in zip(io[i], ig[i])))
        rel = np.max(np.abs(mo[i] - mg[i]) / np.maximum(np.abs(mo[i]), 1e-3))
        worst = max(worst, rel if same else 0)
        if not same or rel > 1e-6:
            bad += 1
            if bad < 8: print("  MISMATCH pair", i, a.num_iterations_total, b.num_iterations_total, a.number_lo_iterations, b.number_lo_iterations, a.best_num_inliers, b.best_num_inliers, "%.3e" % rel, a.best_model_score, b.best_model_score)
    print("pairs identical (iters, LO, inlier sets, model 1e-6):", npairs - bad, "/", npairs, "worst rel among identical %.2e" % worst)
