// ref_shim.cpp -- compiles the REFERENCE's own hot-path functions into
// oracle/_ref/libvisnav_ref.so.  TEST INFRASTRUCTURE ONLY.
//
// The reference defines these functions in a header
// (/root/reference/include/visnav/keypoints.h) that cannot be included as is:
// it pulls in TBB, Pangolin's generated config, DBoW2 and all of OpenCV.  The
// Makefile therefore extracts the line ranges that hold the hot path
// (:47-229 constants, pattern tables, detectKeypoints, computeAngles,
// computeDescriptors, detectKeypointsAndDescriptors; :256-369 checkStillBest,
// isPQiffQP, matchDescriptors) into oracle/_ref/keypoints_extract.inc -- a build
// product that is never committed -- and this file supplies the few types the
// extract needs.  No reference source lives in this repository.
//
// cv::goodFeaturesToTrack: with -DVSO_REF_WITH_OPENCV (the default whenever oracle/build_opencv.sh
// produced the static libraries) the reference's call (keypoints.h:138-139) reaches the reference's OWN
// vendored OpenCV 4.7.0-dev (thirdparty/opencv, modules core + imgproc, built where it lies by its own
// cmake files into a scratch directory; featureselect.cpp:382-548) -- the stock code path, AVX2/FMA
// dispatch and all.  Without the define (no cmake, no /root/reference) the call is routed to the C
// restatement vso_gftt and ref_gftt_impl() says so; bench.py then labels the arm "reference+port".
#include <bitset>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <set>
#include <utility>
#include <vector>

extern "C" int vso_gftt(const uint8_t* img, int w, int h, int pitch, int max_corners,
                        double quality, double min_distance, float* xy, int cap);

// Mirrors pangolin::Image<T> (thirdparty/Pangolin/include/pangolin/image/image.h:240-262, 312-315).
namespace pangolin {
template <class T>
struct ManagedImage {
  size_t pitch; T* ptr; size_t w; size_t h;
  const T& operator()(size_t x, size_t y) const {
    return reinterpret_cast<const T*>(reinterpret_cast<const uint8_t*>(ptr) + y * pitch)[x];
  }
  bool InBounds(float x, float y, float border) const {
    return border <= x && x < (w - border) && border <= y && y < (h - border);
  }
};
}  // namespace pangolin

namespace Eigen {
struct Vector2d {
  double d[2];
  Vector2d() {}
  Vector2d(double x, double y) { d[0] = x; d[1] = y; }
  const double& operator[](int i) const { return d[i]; }
  double& operator[](int i) { return d[i]; }
};
}  // namespace Eigen

#ifdef VSO_REF_WITH_OPENCV
#include <opencv2/core.hpp>
#include <opencv2/core/utility.hpp>
#include <opencv2/imgproc.hpp>
#else
#define CV_8U 0
namespace cv {
struct Point2f { float x, y; };
struct Mat {
  int rows, cols; void* data;
  Mat(size_t r, size_t c, int, void* p) : rows((int)r), cols((int)c), data(p) {}
};
struct NoArray {};
inline NoArray noArray() { return NoArray(); }
inline void goodFeaturesToTrack(const Mat& image, std::vector<Point2f>& pts, int maxCorners,
                                double quality, double minDistance, NoArray, int blockSize,
                                bool harris) {
  (void)blockSize; (void)harris;
  int cap = maxCorners > 0 ? maxCorners : image.rows * image.cols;
  std::vector<float> xy((size_t)cap * 2 + 2);
  int n = vso_gftt((const uint8_t*)image.data, image.cols, image.rows, image.cols, maxCorners,
                   quality, minDistance, xy.data(), cap);
  pts.clear();
  for (int i = 0; i < n; i++) pts.push_back(Point2f{xy[2 * i], xy[2 * i + 1]});
}
}  // namespace cv
#endif

// Mirrors visnav::KeypointsData (include/visnav/common_types.h:111-122).
namespace visnav {
struct KeypointsData {
  std::vector<Eigen::Vector2d> corners;
  std::vector<double> corner_angles;
  std::vector<std::bitset<256>> corner_descriptors;
};
}  // namespace visnav

#include "_ref/keypoints_extract.inc"   // opens namespace visnav (its closing brace is past the extract)
}  // namespace visnav

static_assert(sizeof(std::bitset<256>) == 32, "bitset<256> must be 32 bytes");

static pangolin::ManagedImage<uint8_t> wrap(const uint8_t* img, int w, int h, int pitch) {
  pangolin::ManagedImage<uint8_t> im;
  im.ptr = const_cast<uint8_t*>(img); im.w = w; im.h = h; im.pitch = pitch;
  return im;
}

extern "C" {

int ref_detect(const uint8_t* img, int w, int h, int pitch, int num_features, double* xy, int cap) {
  visnav::KeypointsData kd;
  visnav::detectKeypoints(wrap(img, w, h, pitch), kd, num_features);
  int n = (int)kd.corners.size();
  for (int i = 0; i < n && i < cap; i++) { xy[2 * i] = kd.corners[i][0]; xy[2 * i + 1] = kd.corners[i][1]; }
  return n;
}

void ref_angles(const uint8_t* img, int w, int h, int pitch, const double* xy, int n, int rotate,
                double* angles) {
  visnav::KeypointsData kd;
  for (int i = 0; i < n; i++) kd.corners.emplace_back(xy[2 * i], xy[2 * i + 1]);
  visnav::computeAngles(wrap(img, w, h, pitch), kd, rotate != 0);
  for (int i = 0; i < n; i++) angles[i] = kd.corner_angles[i];
}

void ref_describe(const uint8_t* img, int w, int h, int pitch, const double* xy, const double* angles,
                  int n, uint32_t* desc) {
  visnav::KeypointsData kd;
  for (int i = 0; i < n; i++) kd.corners.emplace_back(xy[2 * i], xy[2 * i + 1]);
  kd.corner_angles.assign(angles, angles + n);
  visnav::computeDescriptors(wrap(img, w, h, pitch), kd);
  if (n) std::memcpy(desc, kd.corner_descriptors.data(), (size_t)n * 32);
}

int ref_detect_describe(const uint8_t* img, int w, int h, int pitch, int num_features, int rotate,
                        double* xy, double* angles, uint32_t* desc, int cap) {
  visnav::KeypointsData kd;
  visnav::detectKeypointsAndDescriptors(wrap(img, w, h, pitch), kd, num_features, rotate != 0);
  int n = (int)kd.corners.size();
  if (n > cap) return -3;
  for (int i = 0; i < n; i++) { xy[2 * i] = kd.corners[i][0]; xy[2 * i + 1] = kd.corners[i][1]; angles[i] = kd.corner_angles[i]; }
  if (n) std::memcpy(desc, kd.corner_descriptors.data(), (size_t)n * 32);
  return n;
}

int ref_match(const uint32_t* d1, int n1, const uint32_t* d2, int n2, int threshold, double dist_2_best,
              int32_t* pairs, int cap) {
  std::vector<std::bitset<256>> a(n1), b(n2);
  if (n1) std::memcpy(a.data(), d1, (size_t)n1 * 32);
  if (n2) std::memcpy(b.data(), d2, (size_t)n2 * 32);
  std::vector<std::pair<int, int>> matches;
  if (n2 > 0) visnav::matchDescriptors(a, b, matches, threshold, dist_2_best);
  int m = (int)matches.size();
  for (int i = 0; i < m && i < cap; i++) { pairs[2 * i] = matches[i].first; pairs[2 * i + 1] = matches[i].second; }
  return m;
}

// Which goodFeaturesToTrack this library runs (bench.py / tests label the CPU arm with it).
const char* ref_gftt_impl(void) {
#ifdef VSO_REF_WITH_OPENCV
  return "opencv " CV_VERSION " (the reference's vendored thirdparty/opencv: core+imgproc static, no IPP/TBB)";
#else
  return "port (C restatement vso_gftt; vendored OpenCV not built)";
#endif
}

#ifdef VSO_REF_WITH_OPENCV
// cv::setUseOptimized / cv::setNumThreads of the vendored build: the scalar dispatch is the bit-exact pin of
// the response map, the default (AVX2/FMA) dispatch is what a stock build of the reference executes.
void ref_set_use_optimized(int on) { cv::setUseOptimized(on != 0); }
void ref_set_num_threads(int n) { cv::setNumThreads(n); }

// cv::goodFeaturesToTrack alone (featureselect.cpp:382-548): xy[cap][2] float, returns the count.
int ref_gftt(const uint8_t* img, int w, int h, int pitch, int max_corners, double quality, double min_distance,
             int block_size, int use_harris, double k, float* xy, int cap) {
  cv::Mat image(h, w, CV_8U, const_cast<uint8_t*>(img), (size_t)pitch);
  std::vector<cv::Point2f> pts;
  cv::goodFeaturesToTrack(image, pts, max_corners, quality, min_distance, cv::noArray(), block_size, use_harris != 0, k);
  int n = (int)pts.size();
  for (int i = 0; i < n && i < cap; i++) { xy[2 * i] = pts[i].x; xy[2 * i + 1] = pts[i].y; }
  return n;
}

// cv::cornerMinEigenVal / cv::cornerHarris (corner.cpp:237-322, 700-760): the response map gftt thresholds.
void ref_corner_response(const uint8_t* img, int w, int h, int pitch, int block_size, int ksize, int use_harris,
                         double k, float* out) {
  cv::Mat image(h, w, CV_8U, const_cast<uint8_t*>(img), (size_t)pitch);
  cv::Mat dst(h, w, CV_32F, out);
  if (use_harris) cv::cornerHarris(image, dst, block_size, ksize, k);
  else cv::cornerMinEigenVal(image, dst, block_size, ksize);
}
#endif

}  // extern "C"
