// keypoints_b200.h -- the hot-path functions of visnav's include/visnav/keypoints.h, re-implemented
// as thin inline wrappers over the C ABI of libvisnav_b200.so (include/visnav_b200.h).
//
// Same namespace, names, signatures, constants and side effects as the reference
// (/root/reference/include/visnav/keypoints.h):
//   detectKeypoints :133-150, computeAngles :152-189, computeDescriptors :191-221,
//   detectKeypointsAndDescriptors :223-229, matchDescriptors :323-369 (isPQiffQP :278-313 is folded
//   into the CUDA kernel), so slam.cpp:1125-1143,1331, sim3.h:252,281 and tracking.h:283,317 compile
//   and link unchanged.  Host code stays C++; all arithmetic runs on the GPU -- there is no CPU path:
//   a failing call throws std::runtime_error (the reference's functions return void).
//
// Expects the including header to have KeypointsData (visnav/common_types.h:111-122) and
// pangolin::ManagedImage<uint8_t> (Pangolin image.h: w, h, pitch, ptr) in scope.
#pragma once

#include <bitset>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../visnav_b200.h"

namespace visnav {

const int PATCH_SIZE = 31;       // keypoints.h:49
const int HALF_PATCH_SIZE = 15;  // keypoints.h:50
const int EDGE_THRESHOLD = 19;   // keypoints.h:51

typedef std::bitset<256> Descriptor;  // keypoints.h:53
static_assert(sizeof(Descriptor) == 32, "std::bitset<256> must be 8 x 32-bit words (libstdc++ layout)");
static_assert(sizeof(std::pair<int, int>) == 8, "std::pair<int,int> must be two packed ints");

namespace b200 {

// One context per host thread (the reference calls these functions from the main thread only,
// slam.cpp:516-526).  VISNAV_B200_DEVICE selects the GPU.
inline vslam_ctx* context() {
  struct Holder {
    vslam_ctx* ctx = nullptr;
    Holder() {
      const char* dev = std::getenv("VISNAV_B200_DEVICE");
      if (vslam_create(dev ? std::atoi(dev) : 0, &ctx) != VSLAM_OK || !ctx)
        throw std::runtime_error("visnav_b200: no usable sm_100 GPU (there is no CPU fallback)");
    }
    ~Holder() { vslam_destroy(ctx); }
  };
  static thread_local Holder h;
  return h.ctx;
}

inline void check(int rc, const char* what) {
  if (rc != VSLAM_OK)
    throw std::runtime_error(std::string("visnav_b200: ") + what + " failed (" + std::to_string(rc) +
                             "): " + vslam_last_error(context()));
}

template <class Image>
inline int capacity(const Image& img, int num_features) {
  return num_features > 0 ? num_features : static_cast<int>(img.w * img.h);
}

}  // namespace b200

template <class Image, class KD>
inline void detectKeypointsImpl(const Image& img_raw, KD& kd, int num_features) {
  kd.corners.clear();
  kd.corner_angles.clear();
  kd.corner_descriptors.clear();
  const int cap = b200::capacity(img_raw, num_features);
  kd.corners.resize(cap);
  static_assert(sizeof(kd.corners[0]) == 16, "corners must be 2 packed doubles (Eigen::Vector2d)");
  int n = 0;
  b200::check(vslam_detect(b200::context(), img_raw.ptr, (int)img_raw.w, (int)img_raw.h, (int)img_raw.pitch,
                           num_features, reinterpret_cast<double*>(kd.corners.data()), cap, &n),
              "detectKeypoints");
  kd.corners.resize(n);
}

template <class Image, class KD>
inline void computeAnglesImpl(const Image& img_raw, KD& kd, bool rotate_features) {
  kd.corner_angles.resize(kd.corners.size());
  b200::check(vslam_angles(b200::context(), img_raw.ptr, (int)img_raw.w, (int)img_raw.h, (int)img_raw.pitch,
                           reinterpret_cast<const double*>(kd.corners.data()), (int)kd.corners.size(),
                           rotate_features ? 1 : 0, kd.corner_angles.data()),
              "computeAngles");
}

template <class Image, class KD>
inline void computeDescriptorsImpl(const Image& img_raw, KD& kd) {
  kd.corner_descriptors.resize(kd.corners.size());
  b200::check(vslam_describe(b200::context(), img_raw.ptr, (int)img_raw.w, (int)img_raw.h, (int)img_raw.pitch,
                             reinterpret_cast<const double*>(kd.corners.data()), kd.corner_angles.data(),
                             (int)kd.corners.size(), reinterpret_cast<uint32_t*>(kd.corner_descriptors.data())),
              "computeDescriptors");
}

template <class Image, class KD>
inline void detectKeypointsAndDescriptorsImpl(const Image& img_raw, KD& kd, int num_features, bool rotate_features) {
  // one upload and one device pass instead of the reference's three sequential calls
  const int cap = b200::capacity(img_raw, num_features);
  kd.corners.resize(cap);
  kd.corner_angles.resize(cap);
  kd.corner_descriptors.resize(cap);
  int n = 0;
  b200::check(vslam_detect_describe(b200::context(), img_raw.ptr, (int)img_raw.w, (int)img_raw.h, (int)img_raw.pitch,
                                    num_features, rotate_features ? 1 : 0,
                                    reinterpret_cast<double*>(kd.corners.data()), kd.corner_angles.data(),
                                    reinterpret_cast<uint32_t*>(kd.corner_descriptors.data()), cap, &n),
              "detectKeypointsAndDescriptors");
  kd.corners.resize(n);
  kd.corner_angles.resize(n);
  kd.corner_descriptors.resize(n);
}

inline void matchDescriptors(const std::vector<std::bitset<256>>& corner_descriptors_1,
                             const std::vector<std::bitset<256>>& corner_descriptors_2,
                             std::vector<std::pair<int, int>>& matches, int threshold, double dist_2_best) {
  matches.clear();
  const int n1 = (int)corner_descriptors_1.size(), n2 = (int)corner_descriptors_2.size();
  if (n1 == 0 || n2 == 0) return;
  matches.resize(n1 < n2 ? n1 : n2);
  int n = 0;
  b200::check(vslam_match(b200::context(), reinterpret_cast<const uint32_t*>(corner_descriptors_1.data()), n1,
                          reinterpret_cast<const uint32_t*>(corner_descriptors_2.data()), n2, threshold, dist_2_best,
                          reinterpret_cast<int32_t*>(matches.data()), (int)matches.size(), &n),
              "matchDescriptors");
  matches.resize(n);
}

}  // namespace visnav
