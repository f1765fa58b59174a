"""Shared helpers for the parity tests: synthetic pairs -> (Pair, Options, Config), perturbed models."""
import numpy as np

from madpose_b200 import capi, synth


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
