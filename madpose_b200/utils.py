"""madpose.utils entry points that sit right before and right after the hot path, on the device.

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


def compute_pose_errors(T_0to1, R, t, t_thres=None):
    """compute_pose_error (utils.py:59-78) for stacks of poses: (n,4,4), (n,3,3), (n,3) -> two (n,) arrays, degrees."""
    return _context().pose_errors(T_0to1, R, t, t_thres)


def compute_pose_error(T_0to1, R, t, t_thres=None):
    """The reference's signature and return value: (error_t, error_R) in degrees for one pose."""
    et, er = _context().pose_errors(_np.asarray(T_0to1)[None], _np.asarray(R)[None], _np.asarray(t)[None], t_thres)
    return et[0], er[0]
