"""slam_b200 — B200-native FastSLAM backend: CUDA kernels behind the C ABI of include/slamgpu.h, plus the host-side
mirror of the reference's FastSLAM1/FastSLAM2 method objects.  See DESIGN.md."""
from .backend import (FASTSLAM1, FASTSLAM2, MATH_FAST, MATH_REFERENCE_ORDER, SCAN_BLOCKED, SCAN_REFERENCE, FastSLAM, SlamGpuError,  # noqa: F401
                      compute_jacobians, estimate_noalloc, NOALLOC_MAX_PARTICLES)
from .build import build  # noqa: F401
from .ekf import EKFSLAM  # noqa: F401
