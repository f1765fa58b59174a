"""ctypes mirror of include/madpose_b200.h (the C ABI of libmadpose_b200.so).

This is the thin Python view of the boundary used by the tests, bench.py and
smoke(); the user-facing drop-in API is the pybind11 module built from
csrc/bindings.cpp (same names as /root/reference/src/bindings.cpp:78-174).
There is no CPU path: loading fails loudly when the CUDA library is missing and
every call fails loudly when no sm_100 device is present.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MADPOSE_B200_LIB") or os.path.join(_HERE, "libmadpose_b200.so")  # (override: kernel experiments)

KIND_CALIBRATED, KIND_SHARED_FOCAL, KIND_TWO_FOCAL, KIND_SCALE_ONLY = 0, 1, 2, 3
MODEL_DOUBLES = 17
_dp = C.POINTER(C.c_double)


class Options(C.Structure):
    """madpose.HybridLORansacOptions (src/hybrid_ransac.h:17-25 + RansacLib defaults)."""

    _fields_ = [
        ("min_num_iterations", C.c_uint32), ("max_num_iterations", C.c_uint32),
        ("max_num_iterations_per_solver", C.c_uint32), ("random_seed", C.c_uint32),
        ("success_probability", C.c_double), ("squared_inlier_thresholds", C.c_double * 2),
        ("data_type_weights", C.c_double * 2), ("threshold_multiplier", C.c_double),
        ("num_lo_steps", C.c_int32), ("num_lsq_iterations", C.c_int32),
        ("min_sample_multiplicator", C.c_int32), ("non_min_sample_multiplier", C.c_int32),
        ("lo_starting_iterations", C.c_uint32), ("final_least_squares", C.c_int32),
        ("use_ours", C.c_int32), ("use_4p4d", C.c_int32),
    ]

    def __init__(self, **kw):
        super().__init__()
        self.min_num_iterations, self.max_num_iterations = 100, 10000
        self.max_num_iterations_per_solver, self.random_seed = 10000, 0
        self.success_probability = 0.9999
        self.threshold_multiplier = 2.0 ** 0.5
        self.num_lo_steps, self.num_lsq_iterations = 10, 4
        self.min_sample_multiplicator, self.non_min_sample_multiplier = 7, 3
        self.lo_starting_iterations, self.final_least_squares = 50, 0
        for k, v in kw.items():
            if k in ("squared_inlier_thresholds", "data_type_weights"):
                getattr(self, k)[0], getattr(self, k)[1] = float(v[0]), float(v[1])
            else:
                setattr(self, k, v)


class Config(C.Structure):
    """madpose.EstimatorConfig (src/estimator_config.h:5-33)."""

    _fields_ = [
        ("solver_type", C.c_int32), ("score_type", C.c_int32), ("lo_type", C.c_int32),
        ("min_depth_constraint", C.c_int32), ("use_shift", C.c_int32),
        ("ceres_use_nonmonotonic_steps", C.c_int32), ("ceres_num_threads", C.c_int32), ("reserved_", C.c_int32),
        ("ceres_function_tolerance", C.c_double), ("ceres_gradient_tolerance", C.c_double),
        ("ceres_parameter_tolerance", C.c_double), ("ceres_max_num_iterations", C.c_double),
    ]

    def __init__(self, **kw):
        super().__init__()
        self.min_depth_constraint, self.use_shift = 1, 1
        self.ceres_use_nonmonotonic_steps, self.ceres_num_threads = 1, 1
        self.ceres_function_tolerance, self.ceres_gradient_tolerance = 1e-6, 1e-8
        self.ceres_parameter_tolerance, self.ceres_max_num_iterations = 1e-6, 25
        for k, v in kw.items():
            setattr(self, k, v)


class Pair(C.Structure):
    _fields_ = [
        ("n", C.c_int32), ("reserved_", C.c_int32),
        ("x0", _dp), ("x1", _dp), ("d0", _dp), ("d1", _dp),
        ("min_depth", C.c_double * 2), ("K0", C.c_double * 9), ("K1", C.c_double * 9),
        ("pp0", C.c_double * 2), ("pp1", C.c_double * 2),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("num_iterations_total", C.c_uint32), ("num_iterations_per_solver", C.c_uint32 * 2),
        ("best_num_inliers", C.c_int32), ("best_solver_type", C.c_int32), ("number_lo_iterations", C.c_int32),
        ("num_inliers", C.c_int32 * 3), ("reserved_", C.c_int32),
        ("best_model_score", C.c_double), ("inlier_ratios", C.c_double * 3),
        ("hypotheses_scored", C.c_uint64), ("residual_evals", C.c_uint64),
        ("lm_solves", C.c_uint32), ("lm_iterations", C.c_uint32),
        ("evals_fp64", C.c_uint64), ("lm_block_evals", C.c_uint64), ("evals_fp32", C.c_uint64),
    ]


def make_pair(x0, x1, d0, d1, min_depth, K0=None, K1=None, pp0=None, pp1=None):
    """Builds an mp_pair over contiguous float64 copies; returns (Pair, keepalive)."""
    a = [np.ascontiguousarray(v, dtype=np.float64) for v in (x0, x1, d0, d1)]
    n = a[2].shape[0]
    assert a[0].shape == (n, 2) and a[1].shape == (n, 2) and a[3].shape == (n,), "x: (n,2), depth: (n,)"
    p = Pair()
    p.n = n
    p.x0, p.x1, p.d0, p.d1 = (v.ctypes.data_as(_dp) for v in a)
    p.min_depth[0], p.min_depth[1] = float(min_depth[0]), float(min_depth[1])
    eye = np.eye(3)
    for dst, src in ((p.K0, K0), (p.K1, K1)):
        m = np.asarray(eye if src is None else src, dtype=np.float64).reshape(9)
        for i in range(9):
            dst[i] = m[i]
    for dst, src in ((p.pp0, pp0), (p.pp1, pp1)):
        v = np.zeros(2) if src is None else np.asarray(src, dtype=np.float64).reshape(2)
        dst[0], dst[1] = v[0], v[1]
    return p, a


def models_to_array(models):
    m = np.ascontiguousarray(models, dtype=np.float64)
    assert m.ndim == 2 and m.shape[1] == MODEL_DOUBLES
    return m


def pack_model(R, t, scale=1.0, o0=0.0, o1=0.0, f0=1.0, f1=1.0):
    return np.concatenate([np.asarray(R, float).reshape(9), np.asarray(t, float).reshape(3), [scale, o0, o1, f0, f1]])


_SYMBOLS = [
    "mp_last_error", "mp_create", "mp_destroy", "mp_device_info", "mp_measure_alu_peaks", "mp_estimate_batch",
    "mp_upload_batch", "mp_estimate_resident", "mp_last_breakdown", "mp_set_groups", "mp_set_batch", "mp_last_phase_cycles", "mp_last_timing", "mp_score_batch", "mp_inliers", "mp_minimal_solve_batch", "mp_lm_batch", "mp_get_depths", "mp_pose_errors",
    "mp_solve_affine", "mp_estimate_scale_and_pose", "mp_set_trace", "mp_read_trace", "mp_set_minimal_mode", "mp_set_warp_from_iteration", "mp_estimate_batch_multi", "mp_estimate_mixed_multi", "mp_device_count",
    "mp_score_models_noalloc",
]


def load_library(path=LIB_PATH):
    if not os.path.exists(path):
        raise ImportError(f"{path} missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(madpose_b200 has no CPU fallback)")
    lib = C.CDLL(path)
    for s in _SYMBOLS:
        getattr(lib, s)  # AttributeError if the library does not export the full ABI
    lib.mp_last_error.restype = C.c_char_p
    for s in _SYMBOLS[1:]:
        getattr(lib, s).restype = C.c_int
    lib.mp_destroy.restype = None
    return lib


NOALLOC_MAX_MATCHES = 16384  # MP_NOALLOC_MAX_MATCHES (include/madpose_b200.h)


def score_models_noalloc(device, kind, pair_dev, opt, cfg, n_models, d_models, d_scores64, stream=0, lib=None):
    """mp_score_models_noalloc: fp64 scores of n_models models of one pair with NO allocation inside the library.
    pair_dev: a Pair whose x0/x1/d0/d1 hold DEVICE addresses; d_models / d_scores64: device addresses (ints);
    stream: a cudaStream_t handle (int).  Only enqueues work: synchronise the stream before reading the scores."""
    lib = lib or load_library()
    rc = lib.mp_score_models_noalloc(int(device), int(kind), C.byref(pair_dev), C.byref(opt), C.byref(cfg), int(n_models),
                                     C.c_void_p(int(d_models)), C.c_void_p(int(d_scores64)), C.c_void_p(int(stream)))
    if rc != 0:
        raise RuntimeError(f"mp_score_models_noalloc failed ({rc}): {lib.mp_last_error().decode()}")


def estimate_batch_multi(devices, kind, pairs, opt, cfg, lib=None, kinds=None, opt_by_kind=None):
    """mp_estimate_batch_multi / mp_estimate_mixed_multi (kinds = per-pair kinds, opt_by_kind = dict kind -> Options).
    Returns (models (P,17), list of Stats, list of 3 inlier index arrays per pair)."""
    lib = lib or load_library()
    n = len(pairs)
    arr = (Pair * n)(*pairs)
    dev = (C.c_int * len(devices))(*devices)
    models = np.empty((n, MODEL_DOUBLES)); stats = (Stats * n)()
    tot = sum(p.n for p in pairs)
    inl = np.zeros(3 * tot + 1, np.int32)
    if kinds is None:
        rc = lib.mp_estimate_batch_multi(len(devices), dev, kind, n, arr, C.byref(opt), C.byref(cfg), models.ctypes.data_as(C.c_void_p),
                                         stats, inl.ctypes.data_as(C.c_void_p))
    else:
        ks = (C.c_int * n)(*kinds)
        ob = (Options * 4)()
        for k, o in opt_by_kind.items():
            ob[k] = o
        rc = lib.mp_estimate_mixed_multi(len(devices), dev, n, ks, arr, ob, C.byref(cfg), models.ctypes.data_as(C.c_void_p),
                                         stats, inl.ctypes.data_as(C.c_void_p))
    if rc != 0:
        raise RuntimeError(f"madpose_b200 error {rc}: {lib.mp_last_error().decode()}")
    out, o = [], 0
    for p, s in zip(pairs, stats):
        out.append([inl[o + t * p.n: o + t * p.n + s.num_inliers[t]].copy() for t in range(3)])
        o += 3 * p.n
    return models, list(stats), out


class Context:
    """Owns an mp_ctx on one GPU."""

    def __init__(self, device=0, lib=None):
        self.lib = lib or load_library()
        self.ctx = C.c_void_p()
        self._check(self.lib.mp_create(int(device), C.byref(self.ctx)))

    def _check(self, rc):
        if rc != 0:
            raise RuntimeError(f"madpose_b200 error {rc}: {self.lib.mp_last_error().decode()}")

    def close(self):
        if self.ctx:
            self.lib.mp_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def device_info(self):
        sm, khz = C.c_int(), C.c_int()
        self._check(self.lib.mp_device_info(self.ctx, C.byref(sm), C.byref(khz)))
        return sm.value, khz.value

    def measure_alu_peaks(self):
        a, b = C.c_double(), C.c_double()
        self._check(self.lib.mp_measure_alu_peaks(self.ctx, C.byref(a), C.byref(b)))
        return a.value, b.value

    def last_timing(self):
        k, t, n = C.c_double(), C.c_double(), C.c_int()
        self._check(self.lib.mp_last_timing(self.ctx, C.byref(k), C.byref(t), C.byref(n)))
        return k.value, t.value, n.value

    def score_batch(self, kind, pair, opt, cfg, models, want32=True, want64=True):
        m = models_to_array(models)
        nm = m.shape[0]
        s32 = np.zeros(nm, np.float32) if want32 else None
        s64 = np.zeros(nm, np.float64) if want64 else None
        self._check(self.lib.mp_score_batch(
            self.ctx, kind, C.byref(pair), C.byref(opt), C.byref(cfg), nm, m.ctypes.data_as(C.c_void_p),
            s32.ctypes.data_as(C.c_void_p) if want32 else None, s64.ctypes.data_as(C.c_void_p) if want64 else None))
        return s32, s64

    def get_depths(self, image_shape, depth_map, mkpts):
        """madpose.utils.get_depths(image, depth_map, mkpts) with the image given by its shape (H, W[, C])."""
        dm = np.ascontiguousarray(depth_map)
        if dm.dtype not in (np.float32, np.float64):
            dm = dm.astype(np.float64)
        kp = np.ascontiguousarray(mkpts, np.float64).reshape(-1, 2)
        out = np.empty(kp.shape[0], np.float64)
        self._check(self.lib.mp_get_depths(self.ctx, dm.ctypes.data_as(C.c_void_p), dm.dtype.itemsize, dm.shape[0], dm.shape[1],
                                           int(image_shape[0]), int(image_shape[1]), kp.ctypes.data_as(C.c_void_p),
                                           kp.shape[0], out.ctypes.data_as(C.c_void_p)))
        return out.astype(dm.dtype)

    def pose_errors(self, T_0to1, R, t, t_thres=None):
        """madpose.utils.compute_pose_error for n poses: T_0to1 (n,4,4), R (n,3,3), t (n,3) -> (err_t, err_R) in degrees."""
        T = np.ascontiguousarray(T_0to1, np.float64).reshape(-1, 16)
        Rm = np.ascontiguousarray(R, np.float64).reshape(-1, 9)
        tv = np.ascontiguousarray(t, np.float64).reshape(-1, 3)
        n = T.shape[0]
        if Rm.shape[0] != n or tv.shape[0] != n:
            raise ValueError("pose_errors: T_0to1, R and t must describe the same number of poses")
        et, er = np.empty(n, np.float64), np.empty(n, np.float64)
        self._check(self.lib.mp_pose_errors(self.ctx, n, T.ctypes.data_as(C.c_void_p), Rm.ctypes.data_as(C.c_void_p),
                                            tv.ctypes.data_as(C.c_void_p), C.c_double(-1.0 if t_thres is None else t_thres),
                                            et.ctypes.data_as(C.c_void_p), er.ctypes.data_as(C.c_void_p)))
        return et, er

    def inliers(self, kind, pair, opt, cfg, model, thr3):
        m = np.ascontiguousarray(model, np.float64).reshape(MODEL_DOUBLES)
        thr = np.ascontiguousarray(thr3, np.float64).reshape(3)
        idx = np.zeros(3 * pair.n, np.int32)
        cnt = np.zeros(3, np.int32)
        self._check(self.lib.mp_inliers(self.ctx, kind, C.byref(pair), C.byref(opt), C.byref(cfg),
                                        m.ctypes.data_as(C.c_void_p), thr.ctypes.data_as(C.c_void_p),
                                        idx.ctypes.data_as(C.c_void_p), cnt.ctypes.data_as(C.c_void_p)))
        return [idx[t * pair.n: t * pair.n + cnt[t]].copy() for t in range(3)]

    def minimal_solve_batch(self, kind, pair, opt, cfg, solver_idx, samples):
        """samples: (n_samples, k) int32.  Returns a list of (count, 17) arrays in the front-end's internal units."""
        idx = np.ascontiguousarray(samples, np.int32)
        ns = idx.shape[0]
        mm = ([4, 8, 4, 4] if solver_idx == 0 else [10, 15, 3, 10])[kind]
        out = np.zeros((ns, mm, MODEL_DOUBLES)); cnt = np.zeros(ns, np.int32)
        self._check(self.lib.mp_minimal_solve_batch(self.ctx, kind, solver_idx, C.byref(pair), C.byref(opt), C.byref(cfg),
                                                    ns, idx.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p),
                                                    cnt.ctypes.data_as(C.c_void_p)))
        return [out[i, :cnt[i]].copy() for i in range(ns)]

    def set_minimal_mode(self, mode):
        self._check(self.lib.mp_set_minimal_mode(self.ctx, int(mode)))

    def set_warp_from_iteration(self, iteration):
        self._check(self.lib.mp_set_warp_from_iteration(self.ctx, int(iteration)))

    def set_trace(self, events_per_pair):
        self._check(self.lib.mp_set_trace(self.ctx, int(events_per_pair)))

    def read_trace(self, pair, cap=1 << 16):
        """(events (k, 2) of the last estimate call for `pair`, number of events the pair produced)."""
        out = np.zeros((cap, 2)); n = C.c_int()
        self._check(self.lib.mp_read_trace(self.ctx, int(pair), out.ctypes.data_as(C.c_void_p), cap, C.byref(n)))
        return out[:min(n.value, cap)].copy(), n.value

    def solve_affine(self, kind, x_homo, y_homo, depth_x, depth_y):
        """The stand-alone solver bindings on explicit samples: x_homo / y_homo (S, 3, NP) with COLUMNS = homogeneous
        points (NP = 3 calibrated, 4 focal kinds), depths (S, NP).  Returns (sols, models): per sample an (ns, 6) array
        of (1, b1, a2, b2 a2, f1, f2) = solve_scale_and_shift* and an (nm, 17) array = solve_scale_shift_pose*."""
        npt = 3 if kind == 0 else 4
        x = np.ascontiguousarray(x_homo, np.float64).reshape(-1, 3, npt); y = np.ascontiguousarray(y_homo, np.float64).reshape(-1, 3, npt)
        dx = np.ascontiguousarray(depth_x, np.float64).reshape(-1, npt); dy = np.ascontiguousarray(depth_y, np.float64).reshape(-1, npt)
        S = x.shape[0]
        sols = np.zeros((S, 8, 6)); ns = np.zeros(S, np.int32); models = np.zeros((S, 8, MODEL_DOUBLES)); nm = np.zeros(S, np.int32)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        self._check(self.lib.mp_solve_affine(self.ctx, kind, S, vp(x), vp(y), vp(dx), vp(dy), vp(sols), vp(ns), vp(models), vp(nm)))
        return [sols[i, :ns[i]].copy() for i in range(S)], [models[i, :nm[i]].copy() for i in range(S)]

    def estimate_scale_and_pose(self, X, Y, W):
        """estimate_scale_and_pose (solver.cpp:5-33): X, Y (3, n), W (n,) -> 17-double model (R, t, scale)."""
        X = np.ascontiguousarray(X, np.float64); Y = np.ascontiguousarray(Y, np.float64); W = np.ascontiguousarray(W, np.float64)
        out = np.zeros(MODEL_DOUBLES)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        self._check(self.lib.mp_estimate_scale_and_pose(self.ctx, X.shape[1], vp(X), vp(Y), vp(W), vp(out)))
        return out

    def lm_batch(self, kind, pair, opt, cfg, starts, samples, nonminimal=False):
        """samples: per problem a list of 3 index arrays.  Returns (models (P,17), info (P,4))."""
        st = np.ascontiguousarray(starts, np.float64).reshape(-1, MODEL_DOUBLES)
        flat, off = [], [0]
        for s3 in samples:
            for t in range(3):
                flat.extend(int(v) for v in s3[t]); off.append(len(flat))
        idx = np.array(flat + [0], np.int32); off = np.array(off, np.int32)
        out = np.zeros_like(st); info = np.zeros((st.shape[0], 4))
        self._check(self.lib.mp_lm_batch(self.ctx, kind, C.byref(pair), C.byref(opt), C.byref(cfg), st.shape[0],
                                         st.ctypes.data_as(C.c_void_p), idx.ctypes.data_as(C.c_void_p),
                                         off.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p),
                                         info.ctypes.data_as(C.c_void_p), int(nonminimal)))
        return out, info

    def estimate_batch(self, kind, pairs, opt, cfg, split=True):
        """pairs: list of Pair (or a ready `(Pair * n)` array).  Returns (models (P,17), list of Stats, list of 3 inlier
        index arrays per pair); split=False returns the flat inlier buffer of the C ABI instead (no per-pair copies)."""
        n = len(pairs)
        arr = pairs if isinstance(pairs, C.Array) else (Pair * n)(*pairs)
        models = np.empty((n, MODEL_DOUBLES)); stats = (Stats * n)()
        tot = sum(p.n for p in pairs)
        inl = np.empty(3 * tot + 1, np.int32)
        self._check(self.lib.mp_estimate_batch(self.ctx, kind, n, arr, C.byref(opt), C.byref(cfg),
                                               models.ctypes.data_as(C.c_void_p), stats, inl.ctypes.data_as(C.c_void_p)))
        if not split:
            return models, stats, inl
        out, o = [], 0
        for p, s in zip(pairs, stats):
            out.append([inl[o + t * p.n: o + t * p.n + s.num_inliers[t]].copy() for t in range(3)])
            o += 3 * p.n
        return models, list(stats), out

    def upload_batch(self, kind, pairs, opt):
        n = len(pairs)
        self._resident = ((Pair * n)(*pairs), [p.n for p in pairs])
        self._check(self.lib.mp_upload_batch(self.ctx, kind, n, self._resident[0], C.byref(opt)))

    def estimate_resident(self, opt, cfg, want_inliers=True):
        ns = self._resident[1]
        n = len(ns)
        models = np.zeros((n, MODEL_DOUBLES)); stats = (Stats * n)()
        inl = np.zeros(3 * sum(ns) + 1, np.int32) if want_inliers else None
        self._check(self.lib.mp_estimate_resident(self.ctx, C.byref(opt), C.byref(cfg), models.ctypes.data_as(C.c_void_p),
                                                  stats, inl.ctypes.data_as(C.c_void_p) if want_inliers else None))
        return models, stats, inl

    def last_breakdown(self):
        ms = (C.c_double * 4)(); ln = (C.c_int * 4)(); r = C.c_int()
        self._check(self.lib.mp_last_breakdown(self.ctx, ms, ln, C.byref(r)))
        names = ["sample", "minimal", "score", "decide"]
        return {k: {"ms": ms[i], "launches": ln[i]} for i, k in enumerate(names)}, r.value

    def last_phase_cycles(self):
        a = (C.c_uint64 * 8)()
        self._check(self.lib.mp_last_phase_cycles(self.ctx, a))
        names = ["scan", "inliers", "score64", "lm_eval", "lm_total", "shuffle", "unflatten", "lo_total"]
        return {k: int(a[i]) for i, k in enumerate(names)}

    def set_batch(self, iterations):
        self._check(self.lib.mp_set_batch(self.ctx, int(iterations)))

    def set_groups(self, groups):
        self._check(self.lib.mp_set_groups(self.ctx, int(groups)))
