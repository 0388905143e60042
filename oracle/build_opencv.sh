#!/bin/bash
# Builds the REFERENCE's vendored OpenCV (thirdparty/opencv, 4.7.0-dev) modules core + imgproc as static
# libraries, so that oracle/_ref/libvisnav_ref.so can link the stock cv::goodFeaturesToTrack the reference
# calls (include/visnav/keypoints.h:138-139 -> modules/imgproc/src/featureselect.cpp:382-548).
# TEST INFRASTRUCTURE ONLY.  Sources are compiled where they lie under $REF; the build tree is a scratch
# directory outside the repository ($OCV_BUILD, default /tmp/visnav_ocv_build); nothing but the final .so
# (oracle/_ref, git-ignored) is kept.  The configure line is SURVEY.md section 8c's (offline, ~15 s configure +
# ~100 s build on 8 cores).  Exits 0 with the libraries in place, 1 if they cannot be built (no cmake / no
# reference): the Makefile then falls back to the C restatement and the library reports itself as "port".
set -e
REF=${REF:-/root/reference}
OCV_BUILD=${OCV_BUILD:-/tmp/visnav_ocv_build}
if [ -f "$OCV_BUILD/lib/libopencv_imgproc.a" ] && [ -f "$OCV_BUILD/lib/libopencv_core.a" ] && [ -f "$OCV_BUILD/3rdparty/lib/libzlib.a" ]; then
  exit 0
fi
[ -d "$REF/thirdparty/opencv/modules/imgproc" ] || { echo "build_opencv: $REF/thirdparty/opencv absent" >&2; exit 1; }
command -v cmake >/dev/null && command -v ninja >/dev/null || { echo "build_opencv: cmake/ninja absent" >&2; exit 1; }
mkdir -p "$OCV_BUILD"
cd "$OCV_BUILD"
OFF=""
for f in IPP OPENCL ITT TBB OPENMP EIGEN LAPACK PROTOBUF ADE PNG JPEG TIFF WEBP OPENJPEG JASPER OPENEXR FFMPEG GSTREAMER V4L GTK; do OFF="$OFF -DWITH_$f=OFF"; done
for f in TESTS PERF_TESTS EXAMPLES opencv_apps JAVA opencv_python3; do OFF="$OFF -DBUILD_$f=OFF"; done
cmake -G Ninja "$REF/thirdparty/opencv" -DCMAKE_BUILD_TYPE=Release -DBUILD_LIST=core,imgproc -DBUILD_SHARED_LIBS=OFF \
  -DBUILD_ZLIB=ON -DCMAKE_POSITION_INDEPENDENT_CODE=ON -DCMAKE_POLICY_VERSION_MINIMUM=3.5 $OFF > cmake.log 2>&1
ninja opencv_core opencv_imgproc > ninja.log 2>&1
[ -f lib/libopencv_imgproc.a ]
