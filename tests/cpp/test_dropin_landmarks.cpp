// Compiles include/visnav/landmark_matching_b200.h against std containers shaped like the
// reference's (common_types.h: Landmarks, Corners, LandmarkMatchData) and drives it the way
// slam.cpp:1159-1162 does.  usage: test_dropin_landmarks  (prints "matches N" + pairs)
#include <visnav/common_types.h>
#include <visnav/landmark_matching_b200.h>

#include <cstdio>
#include <map>
#include <unordered_map>

namespace visnav {
using FrameCamId = std::pair<long, int>;
using TrackId = int64_t;
using FeatureId = int;
struct Landmark { std::map<FrameCamId, FeatureId> all_obs; };
using Landmarks = std::unordered_map<TrackId, Landmark>;
using Corners = std::map<FrameCamId, KeypointsData>;
struct LandmarkMatchData { std::vector<std::pair<FeatureId, TrackId>> matches; };
}  // namespace visnav

int main() {
  using namespace visnav;
  KeypointsData kdl;
  Corners corners;
  Landmarks landmarks;
  std::vector<Vec2d> projected;
  std::vector<TrackId> ids;
  for (int i = 0; i < 40; ++i) {
    kdl.corners.push_back(Vec2d{{30.0 + 15 * i, 40.0 + 7 * (i % 5)}});
    std::bitset<256> d;
    for (int b = 0; b < 256; ++b) d[b] = ((i * 2654435761u + b * 40503u) >> 7) & 1;
    kdl.corner_descriptors.push_back(d);
  }
  KeypointsData& kf = corners[FrameCamId(0, 0)];
  for (int j = 0; j < 40; j += 2) {   // one landmark near every second keypoint, descriptor = 3 flipped bits
    std::bitset<256> d = kdl.corner_descriptors[j];
    d.flip(1); d.flip(77); d.flip(200);
    kf.corner_descriptors.push_back(d);
    landmarks[500 + j].all_obs[FrameCamId(0, 0)] = (int)kf.corner_descriptors.size() - 1;
    projected.push_back(Vec2d{{kdl.corners[j][0] + 1.5, kdl.corners[j][1] - 2.0}});
    ids.push_back(500 + j);
  }
  LandmarkMatchData md;
  md.matches.emplace_back(-1, -1);
  try {
    find_matches_landmarks_b200(kdl, landmarks, corners, projected, ids, 20.0, 70, 1.2, md);
  } catch (const std::exception& e) {
    std::printf("exception: %s\n", e.what());
    return 3;
  }
  std::printf("matches %zu\n", md.matches.size());
  for (auto& m : md.matches) std::printf("m %d %ld\n", m.first, (long)m.second);
  return 0;
}
