// Test stub of visnav::KeypointsData (reference include/visnav/common_types.h:111-122) without Eigen:
// Eigen::Vector2d is two packed doubles.
#pragma once
#include <bitset>
#include <vector>
namespace visnav {
struct Vec2d {
  double v[2];
  double operator[](int i) const { return v[i]; }
  double& operator[](int i) { return v[i]; }
};
struct KeypointsData {
  std::vector<Vec2d> corners;
  std::vector<double> corner_angles;
  std::vector<std::bitset<256>> corner_descriptors;
};
}  // namespace visnav
