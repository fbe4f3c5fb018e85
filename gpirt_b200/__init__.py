"""gpirt_b200 — B200-native implementation of gpirt's per-sweep MCMC hot path.

Everything numerical lives in lib/libgpirt_b200.so (hand-written sm_100a CUDA behind the C-ABI of
include/gpirt_b200.h); this package is the host-side mirror of the reference's interface.
"""
from ._lib import LIB_PATH, NGRID, GpirtError  # noqa: F401
from .api import K, K_time, chol, delta_to_threshold, double_solve, gpirtMCMC, interp1, recover_fstar, threshold_to_delta  # noqa: F401
from .sampler import Sampler, default_thresholds  # noqa: F401
from .panel import Panel  # noqa: F401
from .constant_panel import ConstantPanel  # noqa: F401
