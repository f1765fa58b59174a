"""The synthetic workload as C-ABI records: synthetic pairs (synth.py) -> (Pair, Options, Config),
batches of the BASELINE.json configs, perturbed models for the stand-alone scoring measurements.  Pure numpy + ctypes
records: importing this module loads no native code (bench.py's CPU reference arm uses it too)."""
import numpy as np

from . import capi, synth


def problem(config, p=0, n=None, zero_offsets=False, cfg=None, **optkw):
    d = synth.make_pair(config, p, n=n, zero_offsets=zero_offsets)
    o = synth.options_for(config)
    o.update(optkw)
    opt = capi.Options(**o)
    cfg = capi.Config(**(cfg or {}))
    pair, keep = capi.make_pair(d["x0"], d["x1"], d["d0"], d["d1"], d["min_depth"], d["K0"], d["K1"], d["pp0"], d["pp1"])
    return d, pair, opt, cfg, keep


def rodrigues(v):
    th = np.linalg.norm(v)
    if th < 1e-300:
        return np.eye(3)
    k = v / th
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)


def models_around_gt(d, norm_scale, count, seed=0, spread=1.0):
    """Models in the estimator's internal units: GT perturbed by growing amounts (first one is GT)."""
    rng = np.random.default_rng(seed)
    R, t, s, o0, o1 = synth.gt_model(d)
    f0, f1 = d["gt"]["f0"] / norm_scale, d["gt"]["f1"] / norm_scale
    if d["kind"] == 0:
        f0 = f1 = 1.0
    out = []
    for i in range(count):
        a = spread * (i / max(count - 1, 1)) ** 2
        Rp = rodrigues(rng.standard_normal(3) * 0.2 * a) @ R
        tp = t + rng.standard_normal(3) * 0.3 * a * np.linalg.norm(t)
        ff = 1 + 0.3 * a * rng.standard_normal(2)
        out.append(capi.pack_model(Rp, tp, s * (1 + 0.3 * a * rng.standard_normal()), o0 + 0.2 * a * rng.standard_normal(),
                                   o1 + 0.2 * a * rng.standard_normal(), f0 * abs(ff[0]) if d["kind"] else 1.0,
                                   (f0 * abs(ff[0]) if d["kind"] == 1 else f1 * abs(ff[1])) if d["kind"] else 1.0))
    return np.array(out)


# BASELINE.json "configs" (index = list position) -> what one bench step runs on ONE GPU; SURVEY.md 8(d)
CONFIG_INFO = {
    0: dict(pairs=1, name="configs[0]: calibrated HybridEstimatePoseScaleOffset, 1 synthetic pair, 2000 matches, 40% outliers, "
                          "affine-distorted depth, options of examples/calibrated.py"),
    1: dict(pairs=1500, name="configs[1]: shared-focal HybridEstimatePoseScaleOffsetSharedFocal, 1500 synthetic ScanNet-1500-shaped "
                             "pairs (640x480, N~U{1500..2500} matches, 30-50% outliers), options of examples/shared_focal.py"),
    2: dict(pairs=1500, name="configs[2]: two-focal HybridEstimatePoseScaleOffsetTwoFocal, 1500 synthetic pairs, 5000 matches, "
                             "independent unknown focals ~U[350,900], options of examples/two_focal.py"),
    3: dict(pairs=64, name="configs[3]: calibrated dense RoMa-like, 64 synthetic pairs, 10000 matches, 70% outliers, "
                           "min 1000 / max 100000 iterations (per solver 100000)"),
    4: dict(pairs=15000, name="configs[4]: mixed-estimator batch of 15000 synthetic pairs = 5000 each of the configs[0] / [1] / [2] "
                              "styles interleaved (pair i has style i % 3), sharded over the GPUs (strong scaling)"),
}


def make_batch(config, n_pairs, first=0):
    """Pairs first .. first + n_pairs - 1 of a single-kind config: (kind, [Pair], Options, Config, keepalive)."""
    ps = [problem(config, p=first + i) for i in range(n_pairs)]
    return ps[0][0]["kind"], [q[1] for q in ps], ps[0][2], ps[0][3], ps


def make_mixed_batch(indices):
    """configs[4]: global pair i has the style of config i % 3 (pair i // 3 of that config).  Returns a dict
    kind -> (pairs, Options, Config, keepalive, global indices)."""
    out = {}
    for i in indices:
        c = i % 3
        d, pair, opt, cfg, keep = problem(c, p=i // 3)
        e = out.setdefault(d["kind"], ([], opt, cfg, [], []))
        e[0].append(pair); e[3].append((d, keep)); e[4].append(i)
    return out
