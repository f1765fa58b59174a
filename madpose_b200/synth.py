"""Synthetic ScanNet-1500-shaped image pairs (SURVEY.md section 8d recipe).

Used by bench.py, the tests and smoke(); pure numpy, deterministic per (config, pair).
Configs follow BASELINE.json `configs` (0-based index here = list position):
  0 calibrated, 1 pair, N=2000, 40% outliers
  1 shared focal, N ~ U{1500..2500}, outliers ~ U[0.3,0.5], f ~ U[400,700]
  2 two focal, N=5000, f0,f1 ~ U[350,900]
  3 calibrated dense, N=10000, 70% outliers
"""
import numpy as np

W, H = 640, 480


def _rot(axis, ang):
    axis = axis / np.linalg.norm(axis)
    K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    return np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * (K @ K)


def make_pair(config, p, n=None, outlier_ratio=None, zero_offsets=False):
    """Returns a dict with x0,x1 (n,2), d0,d1 (n,), min_depth, K0,K1, pp0,pp1, kind, gt."""
    rng = np.random.Generator(np.random.PCG64(10_000 * config + p))
    if config == 0:
        kind, n_, rho, f0, f1 = 0, 2000, 0.40, 577.0, 577.0
    elif config == 1:
        kind, n_, rho = 1, int(rng.integers(1500, 2501)), float(rng.uniform(0.3, 0.5))
        f0 = f1 = float(rng.uniform(400, 700))
    elif config == 2:
        kind, n_, rho = 2, 5000, float(rng.uniform(0.3, 0.5))
        f0, f1 = float(rng.uniform(350, 900)), float(rng.uniform(350, 900))
    else:
        kind, n_, rho, f0, f1 = 0, 10000, 0.70, 577.0, 577.0
    n = n or n_
    rho = rho if outlier_ratio is None else outlier_ratio
    pp = np.array([320.0, 240.0]) if kind == 0 else np.array([319.5, 239.5])
    K0 = np.array([[f0, 0, pp[0]], [0, f0, pp[1]], [0, 0, 1.0]])
    K1 = np.array([[f1, 0, pp[0]], [0, f1, pp[1]], [0, 0, 1.0]])
    R = _rot(rng.standard_normal(3), np.deg2rad(rng.uniform(5, 30)))
    t = rng.standard_normal(3)
    t = t / np.linalg.norm(t) * rng.uniform(0.2, 1.0)
    x0 = np.zeros((0, 2)); x1 = np.zeros((0, 2)); z0 = np.zeros(0); z1 = np.zeros(0)
    while x0.shape[0] < n:
        m = 2 * n
        u = np.c_[rng.uniform(8, W - 8, m), rng.uniform(8, H - 8, m)]
        z = rng.uniform(1, 5, m)
        X = (np.linalg.inv(K0) @ np.c_[u, np.ones(m)].T) * z
        Y = R @ X + t[:, None]
        v = (K1 @ Y)
        v = (v[:2] / v[2]).T
        ok = (Y[2] > 0.2) & (v[:, 0] > 0) & (v[:, 0] < W) & (v[:, 1] > 0) & (v[:, 1] < H)
        x0 = np.r_[x0, u[ok]]; x1 = np.r_[x1, v[ok]]; z0 = np.r_[z0, z[ok]]; z1 = np.r_[z1, Y[2][ok]]
    x0, x1, z0, z1 = x0[:n].copy(), x1[:n].copy(), z0[:n], z1[:n]
    x0 += rng.normal(0, 0.5, x0.shape); x1 += rng.normal(0, 0.5, x1.shape)
    al = rng.uniform(0.5, 2.0, 2); of = rng.uniform(-0.2, 0.3, 2)
    if zero_offsets:  # priors that are off by a scale only (use_shift = False / the scale-only estimator)
        of = np.zeros(2)
    d0 = (z0 / al[0] - of[0]) * (1 + rng.normal(0, 0.02, n))
    d1 = (z1 / al[1] - of[1]) * (1 + rng.normal(0, 0.02, n))
    n_out = int(round(rho * n))
    out = rng.permutation(n)[:n_out]
    inl = np.setdiff1d(np.arange(n), out)
    x1[out] = np.c_[rng.uniform(0, W, n_out), rng.uniform(0, H, n_out)]
    d1[out] = rng.choice(d1[inl], n_out)
    gt = dict(R=R, t=t, scale=al[1] / al[0], offset0=of[0], offset1=of[1], f0=f0, f1=f1, alpha0=al[0],
              inlier_mask=np.isin(np.arange(n), inl))
    return dict(kind=kind, x0=x0, x1=x1, d0=d0, d1=d1, min_depth=np.array([d0.min(), d1.min()]), K0=K0, K1=K1,
                pp0=pp.copy(), pp1=pp.copy(), gt=gt)


def options_for(config):
    """Option values of the reference's example scripts (examples/calibrated.py:20-33,
    shared_focal.py:20-33, two_focal.py:13-33) as a dict for capi.Options."""
    o = dict(min_num_iterations=100, max_num_iterations=1000, max_num_iterations_per_solver=10000,
             success_probability=0.9999, random_seed=0, final_least_squares=1, threshold_multiplier=5.0,
             num_lo_steps=4, squared_inlier_thresholds=(64.0, 4.0), data_type_weights=(1.0, 1.0))
    if config == 2:
        o["squared_inlier_thresholds"] = (256.0, 1.0)
    if config == 3:
        o.update(min_num_iterations=1000, max_num_iterations=100000, max_num_iterations_per_solver=100000)
    return o


def gt_model(pair):
    """Ground-truth model in the estimator's INTERNAL units (what scoring expects):
    t in units of 1/alpha0, focal kinds divided by nothing here (callers normalise)."""
    g = pair["gt"]
    return g["R"], g["t"] / g["alpha0"], g["scale"], g["offset0"], g["offset1"]
