// ref_shim_geometry.cpp -- compiles the REFERENCE's own geometry helpers that follow matching
//  into oracle/_ref/libvisnav_ref_geo.so.  TEST INFRASTRUCTURE ONLY.
//
//   * the four camera models: /root/reference/include/visnav/camera_models.h is included where it
//     lies (Eigen + Sophus are the reference's vendored copies, used in place; the one visnav header
//     it pulls in is replaced by oracle/stubs/visnav/common_types.h);
//   * computeEssential / findInliersEssential: matching_utils.h:50-88, and project_landmarks:
//     vo_utils.h:48-81 -- the Makefile extracts those line ranges into oracle/_ref/*.inc (build
//     products, never committed) because the headers themselves need opengv / TBB / OpenCV.
// The few types the extracts touch mirror include/visnav/common_types.h (:104-122 FeatureId, TrackId,
// KeypointsData; :139-149 MatchData; :176 FeatureTrack; :229-250 Landmark, Landmarks).
#include <Eigen/Core>
#include <Eigen/Dense>
#include <Eigen/StdVector>
#include <bitset>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <sophus/se3.hpp>
#include <unordered_map>
#include <utility>
#include <vector>

#include <visnav/camera_models.h>

namespace visnav {
using FeatureId = int;
using TrackId = int64_t;
struct KeypointsData {
  std::vector<Eigen::Vector2d, Eigen::aligned_allocator<Eigen::Vector2d>> corners;
  std::vector<double> corner_angles;
  std::vector<std::bitset<256>> corner_descriptors;
};
struct MatchData {
  Sophus::SE3d T_i_j;
  std::vector<std::pair<FeatureId, FeatureId>> matches;
  std::vector<std::pair<FeatureId, FeatureId>> inliers;
};
struct Landmark {
  Eigen::Vector3d p;
};
// reference: std::unordered_map<TrackId, Landmark>; an ordered map keeps the shim's output order
// equal to ascending track id so that results can be compared as lists (the reference's order is the
// hash table's iteration order, which no caller depends on: vo_utils.h:113-122 scans all of them).
using Landmarks = std::map<TrackId, Landmark>;

#include "_ref/essential_extract.inc"
#include "_ref/project_extract.inc"
}  // namespace visnav

static const char* kModelNames[4] = {"pinhole", "eucm", "ds", "kb4"};

static std::shared_ptr<visnav::AbstractCamera<double>> make_cam(int model, const double* param, int w, int h) {
  auto cam = visnav::AbstractCamera<double>::from_data(kModelNames[model], param);
  cam->width() = w;
  cam->height() = h;
  return cam;
}

extern "C" void ref_cam_project(int model, const double* param, const double* xyz, int n, double* uv) {
  auto cam = make_cam(model, param, 0, 0);
  for (int i = 0; i < n; i++) {
    Eigen::Vector2d r = cam->project(Eigen::Vector3d(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]));
    uv[2 * i] = r[0];
    uv[2 * i + 1] = r[1];
  }
}

extern "C" void ref_cam_unproject(int model, const double* param, const double* uv, int n, double* xyz) {
  auto cam = make_cam(model, param, 0, 0);
  for (int i = 0; i < n; i++) {
    Eigen::Vector3d r = cam->unproject(Eigen::Vector2d(uv[2 * i], uv[2 * i + 1]));
    xyz[3 * i] = r[0];
    xyz[3 * i + 1] = r[1];
    xyz[3 * i + 2] = r[2];
  }
}

// pose = Sophus::SE3d::data(): qx qy qz qw tx ty tz
extern "C" void ref_compute_essential(const double* pose_0_1, double* E_rowmajor) {
  Eigen::Map<const Sophus::SE3d> T(pose_0_1);
  Eigen::Matrix3d E;
  visnav::computeEssential(Sophus::SE3d(T), E);
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) E_rowmajor[3 * r + c] = E(r, c);
}

extern "C" int ref_find_inliers_essential(const double* xy1, int n1, const double* xy2, int n2, int model1,
                                          const double* param1, int model2, const double* param2,
                                          const double* E_rowmajor, double thr, const int32_t* matches, int n_matches,
                                          int32_t* inliers, double* err_out) {
  using namespace visnav;
  KeypointsData kd1, kd2;
  for (int i = 0; i < n1; i++) kd1.corners.emplace_back(xy1[2 * i], xy1[2 * i + 1]);
  for (int i = 0; i < n2; i++) kd2.corners.emplace_back(xy2[2 * i], xy2[2 * i + 1]);
  auto cam1 = make_cam(model1, param1, 0, 0), cam2 = make_cam(model2, param2, 0, 0);
  Eigen::Matrix3d E;
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) E(r, c) = E_rowmajor[3 * r + c];
  MatchData md;
  for (int j = 0; j < n_matches; j++) md.matches.emplace_back(matches[2 * j], matches[2 * j + 1]);
  findInliersEssential(kd1, kd2, cam1, cam2, E, thr, md);
  for (size_t k = 0; k < md.inliers.size(); k++) {
    inliers[2 * k] = md.inliers[k].first;
    inliers[2 * k + 1] = md.inliers[k].second;
  }
  if (err_out)  // the tested quantity itself, so a checker can tell borderline matches apart
    for (int j = 0; j < n_matches; j++) {
      Eigen::Vector3d a = cam1->unproject(kd1.corners[matches[2 * j]]);
      Eigen::Vector3d b = cam2->unproject(kd2.corners[matches[2 * j + 1]]);
      err_out[j] = a.transpose() * E * b;
    }
  return (int)md.inliers.size();
}

extern "C" int ref_project_landmarks(const double* pose_w_c, int model, const double* param, int width, int height,
                                     const double* xyz, int n, double cam_z_threshold, double* proj_xy,
                                     int32_t* proj_idx) {
  using namespace visnav;
  Eigen::Map<const Sophus::SE3d> T(pose_w_c);
  auto cam = make_cam(model, param, width, height);
  Landmarks lms;
  for (int i = 0; i < n; i++) lms[i].p = Eigen::Vector3d(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
  std::vector<Eigen::Vector2d, Eigen::aligned_allocator<Eigen::Vector2d>> pp;
  std::vector<TrackId> ids;
  project_landmarks(Sophus::SE3d(T), cam, lms, cam_z_threshold, pp, ids);
  for (size_t k = 0; k < pp.size(); k++) {
    proj_xy[2 * k] = pp[k][0];
    proj_xy[2 * k + 1] = pp[k][1];
    proj_idx[k] = (int32_t)ids[k];
  }
  return (int)pp.size();
}
