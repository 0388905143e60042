// geometry_b200.h -- visnav::findInliersEssential (reference include/visnav/matching_utils.h:63-88) and
// visnav::project_landmarks (include/visnav/vo_utils.h:48-81) as thin inline wrappers over
// vslam_find_inliers_essential / vslam_project_landmarks (include/visnav_b200.h).
//
// Same argument lists and side effects as the reference: md.inliers is cleared and refilled in match order;
// projected_points / projected_track_ids are cleared and refilled.  The camera is whatever the caller's
// std::shared_ptr<AbstractCamera<double>> points to: the wrapper reads name(), getParam(), width(), height()
// (camera_models.h:447-473) and hands them to the device -- all projection arithmetic runs on the GPU.
// Templated on the container / camera types so that it compiles against the reference's headers unchanged.
// No CPU path: a failing call throws std::runtime_error.  INTEGRATION.md shows the call-site change.
#pragma once

#include <cstdint>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "keypoints_b200.h"

namespace visnav {
namespace b200 {

// AbstractCamera -> vslam_camera (model by name(): camera_models.h:78,155,244,338)
template <class CamPtr>
inline vslam_camera camera_of(const CamPtr& cam) {
  vslam_camera c;
  const std::string n = cam->name();
  if (n == "pinhole") c.model = VSLAM_CAM_PINHOLE;
  else if (n == "eucm") c.model = VSLAM_CAM_EUCM;
  else if (n == "ds") c.model = VSLAM_CAM_DS;
  else if (n == "kb4") c.model = VSLAM_CAM_KB4;
  else throw std::runtime_error("visnav_b200: Camera model " + n + " is not implemented.");
  c.width = cam->width();
  c.height = cam->height();
  c.reserved = 0;
  const auto& p = cam->getParam();
  for (int i = 0; i < 8; ++i) c.param[i] = p[i];
  return c;
}

}  // namespace b200

// E: Eigen::Matrix3d (column-major storage) or anything with operator()(row, col)
template <class KD, class CamPtr, class Mat3, class MatchDataT>
inline void findInliersEssential_b200(const KD& kd1, const KD& kd2, const CamPtr& cam1, const CamPtr& cam2,
                                      const Mat3& E, double epipolar_error_threshold, MatchDataT& md) {
  md.inliers.clear();
  const int nm = (int)md.matches.size();
  if (nm == 0) return;
  static_assert(sizeof(kd1.corners[0]) == 16, "corners must be 2 packed doubles (Eigen::Vector2d)");
  static_assert(sizeof(md.matches[0]) == 8, "matches must be packed int pairs");
  const vslam_camera c1 = b200::camera_of(cam1), c2 = b200::camera_of(cam2);
  double e[9];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) e[3 * r + c] = E(r, c);
  std::vector<int32_t> inl((size_t)nm * 2);
  int n = 0;
  b200::check(vslam_find_inliers_essential(
                  b200::context(), reinterpret_cast<const double*>(kd1.corners.data()), (int)kd1.corners.size(),
                  reinterpret_cast<const double*>(kd2.corners.data()), (int)kd2.corners.size(), &c1, &c2, e,
                  epipolar_error_threshold, reinterpret_cast<const int32_t*>(md.matches.data()), nm, inl.data(), nm,
                  &n, nullptr),
              "findInliersEssential");
  md.inliers.reserve(n);
  for (int k = 0; k < n; ++k) md.inliers.emplace_back(inl[2 * k], inl[2 * k + 1]);
}

// current_pose: Sophus::SE3d (data() = qx qy qz qw tx ty tz) or anything with such a data()
template <class Pose, class CamPtr, class LandmarksT, class PointVec, class TrackIdVec>
inline void project_landmarks_b200(const Pose& current_pose, const CamPtr& cam, const LandmarksT& landmarks,
                                   const double cam_z_threshold, PointVec& projected_points,
                                   TrackIdVec& projected_track_ids) {
  projected_points.clear();
  projected_track_ids.clear();
  const int n = (int)landmarks.size();
  if (n == 0) return;
  std::vector<double> xyz;
  xyz.reserve((size_t)n * 3);
  TrackIdVec ids;
  ids.reserve(n);
  for (const auto& kv : landmarks) {  // same iteration order as the reference's loop (vo_utils.h:62)
    ids.push_back(kv.first);
    xyz.push_back(kv.second.p[0]);
    xyz.push_back(kv.second.p[1]);
    xyz.push_back(kv.second.p[2]);
  }
  const vslam_camera c = b200::camera_of(cam);
  std::vector<double> uv((size_t)n * 2);
  std::vector<int32_t> idx(n);
  int m = 0;
  b200::check(vslam_project_landmarks(b200::context(), current_pose.data(), &c, xyz.data(), n, cam_z_threshold,
                                      uv.data(), idx.data(), n, &m),
              "project_landmarks");
  projected_points.resize(m);
  static_assert(sizeof(projected_points[0]) == 16, "projected points must be 2 packed doubles");
  for (int k = 0; k < m; ++k) {
    projected_points[k][0] = uv[2 * k];
    projected_points[k][1] = uv[2 * k + 1];
    projected_track_ids.push_back(ids[idx[k]]);
  }
}

}  // namespace visnav
