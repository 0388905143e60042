// landmark_matching_b200.h -- visnav::find_matches_landmarks (reference include/visnav/vo_utils.h:83-167)
// as a thin inline wrapper over vslam_match_landmarks (include/visnav_b200.h).
//
// Same name, argument list and result as the reference: md.matches is cleared and refilled with
// (featureId, trackId) pairs in ascending featureId; ties between equally good landmarks resolve as
// the reference's std::partial_sort does.  The host side only marshals: it gathers the observation
// descriptors feature_corners.at(fcid).corner_descriptors[f_id] of every projected landmark
// (vo_utils.h:114-121) into one array; all distance work runs on the GPU.  No CPU path: a failing
// call throws std::runtime_error.
//
// Templated on the container types so that it compiles against the reference's
// common_types.h (Landmarks = std::unordered_map<TrackId, Landmark>, Corners =
// tbb::concurrent_unordered_map<FrameCamId, KeypointsData>, LandmarkMatchData) unchanged.
// INTEGRATION.md shows the two-line change in vo_utils.h.
#pragma once

#include <bitset>
#include <cstdint>
#include <utility>
#include <vector>

#include "keypoints_b200.h"

namespace visnav {

template <class KD, class LandmarksT, class CornersT, class PointVec, class TrackIdVec, class MatchDataT>
inline void find_matches_landmarks_b200(const KD& kdl, const LandmarksT& landmarks, const CornersT& feature_corners,
                                        const PointVec& projected_points, const TrackIdVec& projected_track_ids,
                                        const double match_max_dist_2d, const int feature_match_threshold,
                                        const double feature_match_dist_2_best, MatchDataT& md) {
  md.matches.clear();
  const int n_kp = (int)kdl.corners.size(), n_proj = (int)projected_points.size();
  if (n_kp == 0 || n_proj == 0) return;
  static_assert(sizeof(kdl.corners[0]) == 16, "corners must be 2 packed doubles (Eigen::Vector2d)");
  static_assert(sizeof(projected_points[0]) == 16, "projected points must be 2 packed doubles");
  std::vector<int32_t> offsets(n_proj + 1, 0);
  std::vector<std::bitset<256>> obs;
  for (int j = 0; j < n_proj; ++j) {
    for (const auto& ob : landmarks.at(projected_track_ids[j]).all_obs)
      obs.push_back(feature_corners.at(ob.first).corner_descriptors[ob.second]);
    offsets[j + 1] = (int32_t)obs.size();
  }
  std::vector<int32_t> pairs((size_t)n_kp * 2);
  int n = 0;
  b200::check(vslam_match_landmarks(b200::context(), reinterpret_cast<const double*>(kdl.corners.data()),
                                    reinterpret_cast<const uint32_t*>(kdl.corner_descriptors.data()), n_kp,
                                    reinterpret_cast<const double*>(projected_points.data()), n_proj, offsets.data(),
                                    reinterpret_cast<const uint32_t*>(obs.data()), match_max_dist_2d,
                                    feature_match_threshold, feature_match_dist_2_best, pairs.data(), n_kp, &n),
              "find_matches_landmarks");
  md.matches.reserve(n);
  for (int k = 0; k < n; ++k) md.matches.emplace_back(pairs[2 * k], projected_track_ids[pairs[2 * k + 1]]);
}

}  // namespace visnav
