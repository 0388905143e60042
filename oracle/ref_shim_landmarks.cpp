// ref_shim_landmarks.cpp -- compiles the REFERENCE's own find_matches_landmarks
// (/root/reference/include/visnav/vo_utils.h:83-167) into oracle/_ref/libvisnav_ref_lm.so.
// TEST INFRASTRUCTURE ONLY.
//
// vo_utils.h cannot be included as is (opengv, Sophus, TBB, the camera models), so the Makefile
// extracts the function's line range into oracle/_ref/landmarks_extract.inc -- a build product that
// is never committed -- and this file supplies the few types it touches, mirroring
// include/visnav/common_types.h (:55-96 FrameCamId, :104-122 FeatureId/TrackId/KeypointsData,
// :135 Corners, :151-159 LandmarkMatchData, :176 FeatureTrack, :229-250 Landmark).  Vectors and
// norm() are the reference's vendored Eigen (thirdparty/eigen), used where it lies.
#include <Eigen/Core>
#include <Eigen/StdVector>
#include <algorithm>
#include <bitset>
#include <cstdint>
#include <cstring>
#include <map>
#include <unordered_map>
#include <utility>
#include <vector>

namespace visnav {
using FrameId = int64_t;
using CamId = std::size_t;
struct FrameCamId {
  FrameCamId() : frame_id(0), cam_id(0) {}
  FrameCamId(const FrameId& f, const CamId& c) : frame_id(f), cam_id(c) {}
  FrameId frame_id;
  CamId cam_id;
  bool operator<(const FrameCamId& o) const {
    if (frame_id == o.frame_id) return cam_id < o.cam_id;
    return frame_id < o.frame_id;
  }
};
using FeatureId = int;
using TrackId = int64_t;
struct KeypointsData {
  std::vector<Eigen::Vector2d, Eigen::aligned_allocator<Eigen::Vector2d>> corners;
  std::vector<double> corner_angles;
  std::vector<std::bitset<256>> corner_descriptors;
};
using Corners = std::map<FrameCamId, KeypointsData>;  // reference: tbb::concurrent_unordered_map; only .at() is used
using FeatureTrack = std::map<FrameCamId, FeatureId>;
struct Landmark {
  Eigen::Vector3d p;
  FeatureTrack obs;
  FeatureTrack all_obs;
};
using Landmarks = std::unordered_map<TrackId, Landmark>;
struct LandmarkMatchData {
  std::vector<std::pair<FeatureId, TrackId>> matches;
};

#include "_ref/landmarks_extract.inc"
}  // namespace visnav

static_assert(sizeof(std::bitset<256>) == 32, "bitset<256> must be 32 bytes");

// Same flat interface as the C ABI / the C restatement: landmark j = track id j, its observations
// are spread over synthetic frames (one KeypointsData per observation slot).
extern "C" int ref_match_landmarks(const double* kp_xy, const uint32_t* kp_desc, int n_kp, const double* proj_xy,
                                   int n_proj, const int32_t* obs_off, const uint32_t* obs_desc,
                                   double match_max_dist_2d, int threshold, double dist_2_best, int32_t* pairs,
                                   int cap) {
  using namespace visnav;
  KeypointsData kdl;
  for (int i = 0; i < n_kp; i++) kdl.corners.emplace_back(kp_xy[2 * i], kp_xy[2 * i + 1]);
  kdl.corner_descriptors.resize(n_kp);
  if (n_kp) std::memcpy(kdl.corner_descriptors.data(), kp_desc, (size_t)n_kp * 32);
  Landmarks landmarks;
  Corners feature_corners;
  std::vector<Eigen::Vector2d, Eigen::aligned_allocator<Eigen::Vector2d>> projected_points;
  std::vector<TrackId> projected_track_ids;
  for (int j = 0; j < n_proj; j++) {
    projected_points.emplace_back(proj_xy[2 * j], proj_xy[2 * j + 1]);
    const TrackId tid = 1000 + 7 * (TrackId)j;  // arbitrary distinct ids
    projected_track_ids.push_back(tid);
    Landmark& lm = landmarks[tid];
    for (int o = obs_off[j]; o < obs_off[j + 1]; o++) {
      // observation slot k of a landmark lives in frame k; feature id = landmark index
      const FrameCamId fcid(o - obs_off[j], 0);
      KeypointsData& kd = feature_corners[fcid];
      if ((int)kd.corner_descriptors.size() < n_proj) kd.corner_descriptors.resize(n_proj);
      std::memcpy(&kd.corner_descriptors[j], obs_desc + (size_t)o * 8, 32);
      lm.all_obs[fcid] = j;
    }
  }
  LandmarkMatchData md;
  find_matches_landmarks(kdl, landmarks, feature_corners, projected_points, projected_track_ids, match_max_dist_2d,
                         threshold, dist_2_best, md);
  const int m = (int)md.matches.size();
  for (int k = 0; k < m && k < cap; k++) {
    pairs[2 * k] = md.matches[k].first;
    pairs[2 * k + 1] = (int32_t)((md.matches[k].second - 1000) / 7);
  }
  return m;
}
