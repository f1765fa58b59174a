"""madpose.utils entry points that sit right before the hot path, on the device.

`get_depths(image, depth_map, mkpts)` has the reference's signature (madpose/utils.py:4) and runs the gather through the
C ABI (`mp_get_depths`); there is no CPU path.  The examples call it once per image before the estimator."""
import numpy as _np

from . import capi as _capi

_ctx = None


def _context():
    global _ctx
    if _ctx is None:
        _ctx = _capi.Context(0)
    return _ctx


def get_depths(image, depth_map, mkpts):
    """Depth prior at every keypoint: nearest map texel after scaling to the depth map's resolution (utils.py:4-22)."""
    return _context().get_depths(_np.shape(image), depth_map, mkpts)
