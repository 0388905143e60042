// keypoints.h -- drop-in replacement for visnav's include/visnav/keypoints.h: identical public
// names and signatures (reference keypoints.h:133-229, 323-369), bodies = calls into the B200
// library through include/visnav_b200.h.  See INTEGRATION.md for how it is wired into the
// reference tree; compute_bow_vector (keypoints.h:243-254, cv::ORB + DBoW2, not on the
// accelerated path) stays in the reference's own file, which INTEGRATION.md renames to
// keypoints_bow.h.
#pragma once

#include <pangolin/image/managed_image.h>
#include <visnav/common_types.h>

#include "keypoints_b200.h"

namespace visnav {

inline void detectKeypoints(const pangolin::ManagedImage<uint8_t>& img_raw, KeypointsData& kd, int num_features) {
  detectKeypointsImpl(img_raw, kd, num_features);
}
inline void computeAngles(const pangolin::ManagedImage<uint8_t>& img_raw, KeypointsData& kd, bool rotate_features) {
  computeAnglesImpl(img_raw, kd, rotate_features);
}
inline void computeDescriptors(const pangolin::ManagedImage<uint8_t>& img_raw, KeypointsData& kd) {
  computeDescriptorsImpl(img_raw, kd);
}
inline void detectKeypointsAndDescriptors(const pangolin::ManagedImage<uint8_t>& img_raw, KeypointsData& kd,
                                          int num_features, bool rotate_features) {
  detectKeypointsAndDescriptorsImpl(img_raw, kd, num_features, rotate_features);
}

}  // namespace visnav

#ifdef VISNAV_B200_WITH_BOW
#include <visnav/keypoints_bow.h>  // the reference's compute_bow_vector, untouched
#endif
