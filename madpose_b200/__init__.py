"""madpose_b200 -- B200-native hybrid LO-RANSAC hot path of MADPose behind the reference's API.

`from madpose_b200 import *` / `madpose_b200.HybridEstimatePoseScaleOffset(...)` expose the pybind11 module (same names as
`madpose`); `madpose_b200.capi` is the ctypes view of the C ABI.  There is no CPU fallback.

The extension module is loaded on first use of one of its names (PEP 562), so that `import madpose_b200.capi` /
`madpose_b200.synth` (plain ctypes records and the numpy workload generator, used by the CPU reference arm of bench.py)
load no native code.  The library's streams want CUDA_DEVICE_MAX_CONNECTIONS=32 in the environment BEFORE the process
creates its CUDA context; that is the caller's setting -- this package does not touch the environment."""
import importlib as _importlib

__all__ = [
    "RansacStats", "RansacOptions", "LORansacOptions", "HybridRansacStatistics", "HybridLORansacOptions", "EstimatorConfig",
    "PoseAndScale", "PoseScaleOffset", "PoseScaleOffsetSharedFocal", "PoseScaleOffsetTwoFocal", "ostream_redirect",
    "estimate_scale_and_pose", "solve_scale_and_shift", "solve_scale_and_shift_shared_focal", "solve_scale_and_shift_two_focal",
    "solve_scale_shift_pose", "solve_scale_shift_pose_shared_focal", "solve_scale_shift_pose_two_focal",
    "HybridEstimatePoseAndScale", "HybridEstimatePoseScaleOffset", "HybridEstimatePoseScaleOffsetSharedFocal",
    "HybridEstimatePoseScaleOffsetTwoFocal",
    # batch superset: one call, many pairs
    "HybridEstimatePoseAndScaleBatch", "HybridEstimatePoseScaleOffsetBatch", "HybridEstimatePoseScaleOffsetSharedFocalBatch",
    "HybridEstimatePoseScaleOffsetTwoFocalBatch",
]
_ext = None


def _load():
    global _ext
    if _ext is None:
        _ext = _importlib.import_module("._madpose", __name__)  # ImportError if the extension was not built: no fallback
    return _ext


def have_bindings():
    """True when the pybind11 module (csrc/bindings.cpp) is built and loads."""
    try:
        _load()
        return True
    except ImportError:
        return False


def __getattr__(name):
    if name == "HAVE_BINDINGS":
        return have_bindings()
    if name.startswith("__"):
        raise AttributeError(name)
    try:
        return getattr(_load(), name)
    except ImportError as e:
        raise AttributeError(f"madpose_b200.{name}: the extension module is not built ({e})") from e
