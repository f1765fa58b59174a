"""ctypes loader of tests/hostcheck/libmp_hostcheck.so: the MP_HD scalar math of the CUDA sources compiled
for the host.  TEST INFRASTRUCTURE ONLY -- lets the CPU-only container check the device math against the
oracle; the product never loads it."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "hostcheck")


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class HostCheck:
    def __init__(self, lib):
        self.lib = lib

    def minimal_solve(self, kind, pair, opt, cfg, solver_idx, samples, max_models=15):
        idx = np.ascontiguousarray(samples, np.int32)
        ns = idx.shape[0]
        out = np.zeros((ns, max_models, 17)); cnt = np.zeros(ns, np.int32)
        self.lib.hc_minimal_solve(kind, C.byref(pair), C.byref(opt), C.byref(cfg), solver_idx, ns, _ptr(idx),
                                  max_models, _ptr(out), _ptr(cnt))
        return [out[i, :cnt[i]].copy() for i in range(ns)]


    def lm(self, kind, pair, opt, cfg, starts, samples, nonminimal=False):
        st = np.ascontiguousarray(starts, np.float64).reshape(-1, 17)
        flat, off = [], [0]
        for s3 in samples:
            for t in range(3):
                flat.extend(int(v) for v in s3[t]); off.append(len(flat))
        idx = np.array(flat + [0], np.int32); off = np.array(off, np.int32)
        out = np.zeros_like(st); info = np.zeros((st.shape[0], 4))
        self.lib.hc_lm(kind, C.byref(pair), C.byref(opt), C.byref(cfg), st.shape[0], _ptr(st), _ptr(idx), _ptr(off),
                       _ptr(out), _ptr(info), int(nonminimal))
        return out, info


    def rng_check(self, seed, n):
        out = np.zeros(6, np.uint32); r0 = C.c_double()
        self.lib.hc_rng_check(C.c_uint(seed), int(n), _ptr(out), C.byref(r0))
        return dict(first=[int(out[0]), int(out[1])], words_equal=bool(out[2]), ints_equal=bool(out[3]),
                    reals_equal=bool(out[4]), samples_equal=bool(out[5]), real0=r0.value)


def load(build=True):
    if build:
        subprocess.run(["make", "-s", "-C", HERE], check=True)
    return HostCheck(C.CDLL(os.path.join(HERE, "libmp_hostcheck.so")))
