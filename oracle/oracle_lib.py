"""ctypes loader of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm -- never by madpose_b200/.
Uses only the plain ctypes records of madpose_b200.capi (no native product code is loaded by importing them)."""
import ctypes as C
import os
import subprocess

import numpy as np

ORACLE_DIR = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(ORACLE_DIR)
_vp = C.c_void_p


def _ptr(a):
    return a.ctypes.data_as(_vp)


class Oracle:
    def __init__(self, lib):
        self.lib = lib

    def solve_affine(self, kind, x1, x2, d1, d2):
        x1 = np.ascontiguousarray(x1, np.float64); x2 = np.ascontiguousarray(x2, np.float64)
        d1 = np.ascontiguousarray(d1, np.float64); d2 = np.ascontiguousarray(d2, np.float64)
        out = np.zeros((8, 5))
        n = self.lib.om_solve_affine(kind, _ptr(x1), _ptr(x2), _ptr(d1), _ptr(d2), _ptr(out))
        return out[:n].copy()

    def score(self, kind, pair, opt, cfg, models):
        m = np.ascontiguousarray(models, np.float64)
        s = np.zeros(m.shape[0])
        self.lib.om_score(kind, C.byref(pair), C.byref(opt), C.byref(cfg), m.shape[0], _ptr(m), _ptr(s))
        return s

    def residuals(self, kind, pair, opt, cfg, model):
        m = np.ascontiguousarray(model, np.float64)
        r = np.zeros((3, pair.n)); thr = np.zeros(3)
        self.lib.om_residuals(kind, C.byref(pair), C.byref(opt), C.byref(cfg), _ptr(m), _ptr(r), _ptr(thr))
        return r, thr

    def inliers(self, kind, pair, opt, cfg, model, thr3=None):
        m = np.ascontiguousarray(model, np.float64)
        idx = np.zeros(3 * pair.n, np.int32); cnt = np.zeros(3, np.int32)
        t = None if thr3 is None else np.ascontiguousarray(thr3, np.float64)
        self.lib.om_inliers(kind, C.byref(pair), C.byref(opt), C.byref(cfg), _ptr(m), None if t is None else _ptr(t),
                            _ptr(idx), _ptr(cnt))
        return [idx[k * pair.n: k * pair.n + cnt[k]].copy() for k in range(3)]

    def minimal_solve(self, kind, pair, opt, cfg, solver_idx, samples, max_models=15):
        idx = np.ascontiguousarray(samples, np.int32)
        ns = idx.shape[0]
        out = np.zeros((ns, max_models, 17)); cnt = np.zeros(ns, np.int32)
        self.lib.om_minimal_solve(kind, C.byref(pair), C.byref(opt), C.byref(cfg), solver_idx, ns, _ptr(idx),
                                  max_models, _ptr(out), _ptr(cnt))
        return [out[i, :cnt[i]].copy() for i in range(ns)]

    def lm(self, kind, pair, opt, cfg, starts, samples, nonminimal=False):
        """samples: list (per problem) of 3 index arrays."""
        st = np.ascontiguousarray(starts, np.float64).reshape(-1, 17)
        flat, off = [], [0]
        for s3 in samples:
            for t in range(3):
                flat.extend(int(v) for v in s3[t]); off.append(len(flat))
        idx = np.array(flat + [0], np.int32); off = np.array(off, np.int32)
        out = np.zeros_like(st); info = np.zeros((st.shape[0], 4))
        self.lib.om_lm(kind, C.byref(pair), C.byref(opt), C.byref(cfg), st.shape[0], _ptr(st), _ptr(idx), _ptr(off),
                       _ptr(out), _ptr(info), int(nonminimal))
        return out, info

    def estimate_batch(self, kind, pairs, opt, cfg, threads=1, trace_cap=0, lib=None):
        """pairs: list of capi.Pair.  Returns (models (P,17), list of capi.Stats, list of 3 inlier arrays[, trace])."""
        from madpose_b200 import capi
        lib = lib or self.lib
        n = len(pairs)
        arr = (capi.Pair * n)(*pairs)
        models = np.zeros((n, 17)); stats = (capi.Stats * n)()
        tot = sum(p.n for p in pairs)
        inl = np.zeros(3 * tot + 1, np.int32)
        tr = np.zeros(max(trace_cap, 1))
        fn = lib.om_estimate_batch if hasattr(lib, "om_estimate_batch") else lib.omref_estimate_batch
        fn(kind, n, arr, C.byref(opt), C.byref(cfg), _ptr(models), stats, _ptr(inl), threads,
           _ptr(tr) if trace_cap else None, trace_cap)
        out, o = [], 0
        for p, s in zip(pairs, stats):
            out.append([inl[o + t * p.n: o + t * p.n + s.num_inliers[t]].copy() for t in range(3)])
            o += 3 * p.n
        return (models, list(stats), out, tr) if trace_cap else (models, list(stats), out)

    def frontend(self, kind, pair, opt, cfg):
        o = np.zeros(8)
        self.lib.om_frontend(kind, C.byref(pair), C.byref(opt), C.byref(cfg), _ptr(o))
        return dict(norm_scale=o[0], thr=o[1:4].copy(), w=o[4:7].copy(), loss_scale=o[7])


def load(build=True):
    path = os.path.join(ORACLE_DIR, "liboracle.so")
    if build:
        subprocess.run(["make", "-s", "-C", ORACLE_DIR, "liboracle.so"], check=True)
    lib = C.CDLL(path)
    for name in dir(lib):
        pass
    return Oracle(lib)
