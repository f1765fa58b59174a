"""mp_score_models_noalloc (no counterpart in the reference): the fp64
scoring path with its scratch in a fixed file-scope device array instead of cudaMalloc'ed buffers.  Checked through
the C ABI against the existing entry point (mp_score_batch, scores_f64) on two calls in a row -- the second call
finds the first call's contents in the shared scratch -- and against the size limit.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from tests.helpers import models_around_gt, problem

pytestmark = pytest.mark.gpu


def _device_pair(torch, pair, keep_list):
    """copy of `pair` whose data pointers are device addresses (the test owns the allocations, not the library)"""
    from madpose_b200 import capi
    n = pair.n
    host = [np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_double)), shape=(k,)).copy()
            for p, k in ((pair.x0, 2 * n), (pair.x1, 2 * n), (pair.d0, n), (pair.d1, n))]
    dev = [torch.from_numpy(h).cuda() for h in host]
    keep_list.extend(dev)
    q = capi.Pair()
    C.memmove(C.byref(q), C.byref(pair), C.sizeof(capi.Pair))
    q.x0, q.x1, q.d0, q.d1 = (C.cast(C.c_void_p(t.data_ptr()), C.POINTER(C.c_double)) for t in dev)
    return q


def test_noalloc_scores_match_score_batch_twice(ctx, oracle):
    import torch
    from madpose_b200 import capi
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    keep = []
    # two calls in a row on different pairs / kinds / sizes: the second one runs over the first one's leftovers
    # (a longer pair first, so that stale records beyond the second pair's length are really there)
    cases = [(1, 4, None, 67), (0, 2, 900, 33), (2, 5, None, 40), (1, 4, None, 67)]
    for call, (config, p, n, nm) in enumerate(cases):
        d, pair, opt, cfg, k2 = problem(config, p=p, n=n)
        assert pair.n <= capi.NOALLOC_MAX_MATCHES
        fe = oracle.frontend(d["kind"], pair, opt, cfg)
        models = capi.models_to_array(models_around_gt(d, fe["norm_scale"], nm, seed=call))
        _, want = ctx.score_batch(d["kind"], pair, opt, cfg, models, want32=False)
        pd = _device_pair(torch, pair, keep)
        dm = torch.from_numpy(np.ascontiguousarray(models)).cuda()
        ds = torch.full((nm,), float("nan"), dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        capi.score_models_noalloc(0, d["kind"], pd, opt, cfg, nm, dm.data_ptr(), ds.data_ptr(), stream, lib=ctx.lib)
        torch.cuda.synchronize()
        free1 = torch.cuda.mem_get_info()[0]
        got = ds.cpu().numpy()
        # same kernels' arithmetic on the same records: the existing tests hold the fp64 scores to rtol 1e-11
        np.testing.assert_allclose(got, want, rtol=1e-11, atol=0)
        assert np.array_equal(got, want), f"call {call}: not bit-identical (max rel {np.max(np.abs(got / want - 1))})"
        if call > 0:  # (the first call may load the module's kernels lazily)
            assert free1 == free0, f"call {call}: device memory in use changed by {free0 - free1} bytes"
        ref = oracle.score(d["kind"], pair, opt, cfg, models)
        np.testing.assert_allclose(got, ref, rtol=1e-11)


def test_noalloc_rejects_oversize_and_bad_arguments(ctx):
    import torch
    from madpose_b200 import capi
    torch.cuda.set_device(0)
    d, pair, opt, cfg, k2 = problem(0, p=0, n=64)
    keep = []
    pd = _device_pair(torch, pair, keep)
    dm = torch.zeros((2, 17), dtype=torch.float64, device="cuda")
    ds = torch.zeros(2, dtype=torch.float64, device="cuda")
    pd.n = capi.NOALLOC_MAX_MATCHES + 1  # (never dereferenced: the size check comes first)
    with pytest.raises(RuntimeError, match="exceed the fixed scratch"):
        capi.score_models_noalloc(0, 0, pd, opt, cfg, 2, dm.data_ptr(), ds.data_ptr(), 0, lib=ctx.lib)
    pd.n = 64
    with pytest.raises(RuntimeError, match="bad arguments"):
        capi.score_models_noalloc(0, 0, pd, opt, cfg, 0, dm.data_ptr(), ds.data_ptr(), 0, lib=ctx.lib)
    with pytest.raises(RuntimeError, match="kind must be"):
        capi.score_models_noalloc(0, 7, pd, opt, cfg, 2, dm.data_ptr(), ds.data_ptr(), 0, lib=ctx.lib)
