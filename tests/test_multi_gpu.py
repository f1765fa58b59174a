"""Batch sharding inside the library (mp_estimate_batch_multi / mp_estimate_mixed_multi): the deal over
devices and the scatter back into the caller's order must not change any result.  On a one-GPU box the same device is
listed several times (the shares then take turns on one context); with several GPUs the real devices are used too."""
import numpy as np
import pytest

from madpose_b200 import capi
from tests.helpers import problem

pytestmark = pytest.mark.gpu


def _same(a, b):
    ma, sa, ia = a
    mb, sb, ib = b
    np.testing.assert_array_equal(ma, mb)
    for x, y in zip(sa, sb):
        assert (x.num_iterations_total, x.number_lo_iterations, x.best_num_inliers, x.best_model_score) == \
               (y.num_iterations_total, y.number_lo_iterations, y.best_num_inliers, y.best_model_score)
    for x, y in zip(ia, ib):
        for u, v in zip(x, y):
            np.testing.assert_array_equal(u, v)


def test_sharded_batch_equals_single_context(ctx):
    ps = [problem(1, p=200 + i, n=300 + 97 * (i % 5)) for i in range(11)]   # ragged sizes: the deal is by match count
    kind, opt, cfg = 1, ps[0][2], ps[0][3]
    pairs = [q[1] for q in ps]
    ref = ctx.estimate_batch(kind, pairs, opt, cfg)
    lib = ctx.lib
    for devices in ([0], [0, 0], [0, 0, 0, 0]):
        _same(ref, capi.estimate_batch_multi(devices, kind, pairs, opt, cfg, lib=lib))
    n = lib.mp_device_count()
    if n > 1:
        _same(ref, capi.estimate_batch_multi(list(range(n)), kind, pairs, opt, cfg, lib=lib))


def test_mixed_kinds_batch(ctx):
    """configs[4] in small: three estimator kinds interleaved in one call."""
    items = [problem(i % 3, p=300 + i, n=400) for i in range(9)]
    kinds = [q[0]["kind"] for q in items]
    pairs = [q[1] for q in items]
    opts = {q[0]["kind"]: q[2] for q in items}
    cfg = items[0][3]
    lib = ctx.lib
    n = max(lib.mp_device_count(), 1)
    m, s, inl = capi.estimate_batch_multi(list(range(n)) if n > 1 else [0, 0], 0, pairs, None, cfg, lib=lib, kinds=kinds, opt_by_kind=opts)
    for k in range(3):
        sel = [i for i in range(9) if kinds[i] == k]
        mk, sk, ik = ctx.estimate_batch(k, [pairs[i] for i in sel], opts[k], cfg)
        _same((m[sel], [s[i] for i in sel], [inl[i] for i in sel]), (mk, sk, ik))


def test_python_batch_call_with_devices():
    import madpose_b200 as madpose
    from madpose_b200 import synth
    ds = [synth.make_pair(1, p, n=400) for p in range(6)]
    o = madpose.HybridLORansacOptions()
    for k, v in synth.options_for(1).items():
        setattr(o, k, list(v) if isinstance(v, tuple) else v)
    args = ([d["x0"] for d in ds], [d["x1"] for d in ds], [d["d0"] for d in ds], [d["d1"] for d in ds],
            [d["min_depth"] for d in ds], [d["pp0"] for d in ds], [d["pp1"] for d in ds], o)
    one = madpose.HybridEstimatePoseScaleOffsetSharedFocalBatch(*args)
    n = madpose.device_count()
    many = madpose.HybridEstimatePoseScaleOffsetSharedFocalBatch(*args, devices=list(range(n)) if n > 1 else [0, 0])
    for (pa, sa), (pb, sb) in zip(one, many):
        assert np.array_equal(pa.pose, pb.pose) and pa.focal == pb.focal and sa.num_iterations_total == sb.num_iterations_total
