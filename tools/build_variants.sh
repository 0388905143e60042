#!/bin/bash
# Builds the profiling variant of the library (cycle counters in match_mma_kernel) next to the product build.
set -e
cd "$(dirname "$0")/.."
VSLAM_NVCC_EXTRA=-DVSLAM_MMA_PROFILE=${MMA_PROFILE:-1} python -c "import visnav_b200; visnav_b200.build_native(force=True)"
mkdir -p tools/probes/bin
cp visual-slam_b200/csrc/libvisnav_b200.so tools/probes/bin/libvisnav_b200_prof.so
python -c "import visnav_b200; print(visnav_b200.build_native(force=True))"
