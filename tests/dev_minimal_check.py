import sys, numpy as np
sys.path.insert(0, '/root/repo')
from tests import oracle_lib, hostcheck_lib
from tests.helpers import problem
ora = oracle_lib.load(); hc = hostcheck_lib.load()
def relerr(a, b):
    return np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-3))
for config in (0, 1, 2):
    d, pair, opt, cfg, keep = problem(config, p=2)
    rng = np.random.default_rng(config)
    inl = np.nonzero(d["gt"]["inlier_mask"])[0]
    for solver in (0, 1):
        k = [[3, 5], [4, 6], [4, 7]][d["kind"]][solver]
        S = np.array([rng.choice(inl if i % 2 == 0 else pair.n, k, replace=False) for i in range(300)], np.int32)
        a = ora.minimal_solve(d["kind"], pair, opt, cfg, solver, S)
        b = hc.minimal_solve(d["kind"], pair, opt, cfg, solver, S)
        nm = sum(len(x) for x in a); mism = 0; worst = []; 
        for x, y in zip(a, b):
            if len(x) != len(y): mism += 1; continue
            for m in x:
                e = [relerr(m, mm) for mm in y]
                worst.append(min(e) if e else 0)
        worst = np.array(worst)
        print(config, solver, "models", nm, "count mismatches", mism, "max err", worst.max() if len(worst) else 0, "frac>1e-9", np.mean(worst > 1e-9) if len(worst) else 0)
print("---- detail config 1 solver 0")
d, pair, opt, cfg, keep = problem(1, p=2)
rng = np.random.default_rng(1)
inl = np.nonzero(d["gt"]["inlier_mask"])[0]
S = np.array([rng.choice(inl if i % 2 == 0 else pair.n, 4, replace=False) for i in range(300)], np.int32)
a = ora.minimal_solve(1, pair, opt, cfg, 0, S); b = hc.minimal_solve(1, pair, opt, cfg, 0, S)
for i,(x, y) in enumerate(zip(a, b)):
    for m in x:
        e = [relerr(m, mm) for mm in y]
        if min(e) > 1e-9:
            j = int(np.argmin(e)); print(i, min(e)); print(m); print(y[j]); print(np.abs(m-y[j])/np.maximum(np.abs(m),1e-3))
