"""madpose_b200 -- B200-native hybrid LO-RANSAC hot path of MADPose behind the reference's API.

`from madpose_b200 import *` exposes the pybind11 module (same names as `madpose`) once it is built;
`madpose_b200.capi` is the ctypes view of the C ABI.  There is no CPU fallback."""
import os as _os

# more hardware queues for the library's streams; must be set before the process initialises CUDA (see mp_create)
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

try:  # the drop-in module (csrc/bindings.cpp); absent until built
    from ._madpose import *  # noqa: F401,F403
    HAVE_BINDINGS = True
except ImportError:  # pragma: no cover
    HAVE_BINDINGS = False
