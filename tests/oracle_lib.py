"""ctypes loader of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
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
