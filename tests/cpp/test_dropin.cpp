// Drives the drop-in header the way the reference does (slam.cpp:1125-1143 and
// test/src/test_loop_closure_utils.cpp:117-118): two images -> detectKeypointsAndDescriptors,
// then the three separate calls, then matchDescriptors.  Prints a digest the python test compares
// with the oracle.  usage: test_dropin left.raw right.raw w h
#include <visnav/keypoints.h>

#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <vector>

static std::vector<uint8_t> slurp(const char* p, size_t n) {
  std::vector<uint8_t> b(n);
  std::ifstream f(p, std::ios::binary);
  f.read(reinterpret_cast<char*>(b.data()), n);
  if (!f) { std::fprintf(stderr, "cannot read %s\n", p); std::exit(2); }
  return b;
}

int main(int argc, char** argv) {
  if (argc < 5) return 2;
  const size_t w = std::atoi(argv[3]), h = std::atoi(argv[4]);
  std::vector<uint8_t> l = slurp(argv[1], w * h), r = slurp(argv[2], w * h);
  pangolin::ManagedImage<uint8_t> il, ir;
  il.ptr = l.data(); il.w = w; il.h = h; il.pitch = w;
  ir.ptr = r.data(); ir.w = w; ir.h = h; ir.pitch = w;
  visnav::KeypointsData kdl, kdr, kd3;
  try {
    visnav::detectKeypointsAndDescriptors(il, kdl, 1500, true);
    visnav::detectKeypointsAndDescriptors(ir, kdr, 1500, true);
    visnav::detectKeypoints(ir, kd3, 1500);
    visnav::computeAngles(ir, kd3, true);
    visnav::computeDescriptors(ir, kd3);
    std::vector<std::pair<int, int>> matches(3, std::make_pair(-1, -1));
    visnav::matchDescriptors(kdl.corner_descriptors, kdr.corner_descriptors, matches, 70, 1.2);
    bool same = kd3.corners.size() == kdr.corners.size() && kd3.corner_descriptors == kdr.corner_descriptors;
    std::printf("nl %zu nr %zu matches %zu same %d\n", kdl.corners.size(), kdr.corners.size(), matches.size(), (int)same);
    for (size_t i = 0; i < matches.size(); ++i) std::printf("m %d %d\n", matches[i].first, matches[i].second);
    for (size_t i = 0; i < kdl.corners.size(); ++i)
      std::printf("k %g %g %.17g %s\n", kdl.corners[i][0], kdl.corners[i][1], kdl.corner_angles[i],
                  kdl.corner_descriptors[i].to_string().c_str());
  } catch (const std::exception& e) {
    std::printf("exception: %s\n", e.what());
    return 3;
  }
  return 0;
}
