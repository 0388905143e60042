// Compiles include/visnav/geometry_b200.h against std containers / a camera class shaped like the
// reference's (camera_models.h AbstractCamera: name(), getParam(), width(), height(); common_types.h MatchData,
// Landmarks) and drives it the way slam.cpp:1141-1143 and :1104-1113 do.
// usage: test_dropin_geometry   (prints "inliers N" + pairs, "projected M" + ids)
#include <visnav/common_types.h>
#include <visnav/geometry_b200.h>

#include <array>
#include <cstdio>
#include <map>
#include <memory>

namespace visnav {
using TrackId = int64_t;
using FeatureId = int;
struct PinholeStub {
  std::array<double, 8> param{{400, 400, 376, 240, 0, 0, 0, 0}};
  int w = 752, h = 480;
  std::string name() const { return "pinhole"; }
  const std::array<double, 8>& getParam() const { return param; }
  int width() const { return w; }
  int height() const { return h; }
};
struct MatchData {
  std::vector<std::pair<FeatureId, FeatureId>> matches, inliers;
};
struct Landmark { double p[3]; };
using Landmarks = std::map<TrackId, Landmark>;
struct Mat3 {
  double m[9];
  double operator()(int r, int c) const { return m[3 * r + c]; }
};
struct Pose {
  double d[7];
  const double* data() const { return d; }
};
}  // namespace visnav

int main() {
  using namespace visnav;
  auto cam = std::make_shared<PinholeStub>();
  KeypointsData kd1, kd2;
  MatchData md;
  for (int i = 0; i < 30; ++i) {
    kd1.corners.push_back(Vec2d{{100.0 + 20 * i, 60.0 + 12 * i}});
    // same row -> on the epipolar line of a pure x-baseline; every third partner is 10 rows off
    kd2.corners.push_back(Vec2d{{88.0 + 20 * i, 60.0 + 12 * i + (i % 3 == 2 ? 10.0 : 0.0)}});
    md.matches.emplace_back(i, i);
  }
  md.inliers.emplace_back(-1, -1);
  const Mat3 E{{0, 0, 0, 0, 0, -1, 0, 1, 0}};  // skew((1,0,0)) * I
  Landmarks lms;
  for (int j = 0; j < 12; ++j) lms[700 + j] = Landmark{{-3.0 + 0.5 * j, 0.1 * j, j < 2 ? -1.0 : 2.0 + j}};
  lms[900] = Landmark{{40.0, 0.0, 1.0}};  // far outside the image
  const Pose pose{{0, 0, 0, 1, 0, 0, 0}};
  std::vector<Vec2d> pts{Vec2d{{1, 1}}};
  std::vector<TrackId> ids{5};
  try {
    findInliersEssential_b200(kd1, kd2, cam, cam, E, 5e-3, md);
    project_landmarks_b200(pose, cam, lms, 0.1, pts, ids);
  } catch (const std::exception& e) {
    std::printf("exception: %s\n", e.what());
    return 3;
  }
  std::printf("inliers %zu\n", md.inliers.size());
  for (auto& m : md.inliers) std::printf("i %d %d\n", m.first, m.second);
  std::printf("projected %zu\n", ids.size());
  for (size_t k = 0; k < ids.size(); ++k) std::printf("p %ld %.9f %.9f\n", (long)ids[k], pts[k][0], pts[k][1]);
  return 0;
}
