"""visnav_b200 -- B200 (sm_100a) front-end hot path of visnav (yunjinli/visual-slam).

The package holds the CUDA kernels + C ABI (csrc/, include/visnav_b200.h) and a thin host-side
mirror of the reference's operator interface (keypoints.py).  The directory name contains a
hyphen; import it through the repo-root shim `visnav_b200.py`.
"""
from .build import build_native, LIB  # noqa: F401
from .keypoints import (  # noqa: F401
    KeypointsData, Context, VslamError, lib_path, load_library, default_context,
    detectKeypoints, computeAngles, computeDescriptors, detectKeypointsAndDescriptors,
    matchDescriptors, EDGE_THRESHOLD, PATCH_SIZE, HALF_PATCH_SIZE, ABI_SYMBOLS,
)
