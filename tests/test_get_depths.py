"""mp_get_depths (the step before the path) against golden vectors produced by the reference's own
madpose.utils.get_depths (tests/golden/make_get_depths_golden.py)."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "get_depths.npz")


def restatement(image_shape, depth_map, mkpts):
    """numpy restatement of madpose/utils.py:4-22 (the oracle of this entry point)."""
    resize = np.array([depth_map.shape[1] / image_shape[1], depth_map.shape[0] / image_shape[0]])
    c = np.round(mkpts * resize).astype(int)
    c[:, 0] = np.clip(c[:, 0], 0, depth_map.shape[1] - 1)
    c[:, 1] = np.clip(c[:, 1], 0, depth_map.shape[0] - 1)
    return depth_map[c[:, 1], c[:, 0]]


def cases():
    z = np.load(GOLD)
    for i in range(len(z.files) // 5):
        k = "c%d_" % i
        yield [int(v) for v in z[k + "image_shape"]], z[k + "depth"], z[k + "mkpts"], z[k + "expect"]


def test_restatement_matches_reference_golden():
    for shape, depth, kp, expect in cases():
        np.testing.assert_array_equal(np.asarray(restatement(shape, depth, kp), np.float64), expect)


@pytest.mark.gpu
def test_device_matches_reference_golden(ctx):
    for shape, depth, kp, expect in cases():
        got = ctx.get_depths(shape, depth, kp)
        assert got.dtype == depth.dtype
        np.testing.assert_array_equal(np.asarray(got, np.float64), expect)  # a gather: bit-exact


@pytest.mark.gpu
def test_utils_module_signature():
    from madpose_b200.utils import get_depths
    for shape, depth, kp, expect in cases():
        image = np.zeros((shape[0], shape[1], 3), np.uint8)
        np.testing.assert_array_equal(np.asarray(get_depths(image, depth, kp), np.float64), expect)
