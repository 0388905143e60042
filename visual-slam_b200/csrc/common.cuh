// common.cuh -- context, workspace and error plumbing shared by the kernels' host launchers.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <string>

#include "../../include/visnav_b200.h"

struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
  // grow-only scratch; contents are not preserved
  cudaError_t reserve(size_t need) {
    if (need <= bytes) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    bytes = 0;
    size_t want = need + need / 4 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) bytes = want;
    return e;
  }
  void release() {
    if (p && owned) cudaFree(p);
    p = nullptr;
    bytes = 0;
    owned = true;
  }
  // a window into another buffer (never freed, never grown)
  void alias(void* ptr, size_t n) {
    release();
    p = ptr;
    bytes = n;
    owned = false;
  }
  bool owned = true;
  template <class T>
  T* as() const { return reinterpret_cast<T*>(p); }
};

struct vslam_ctx {
  int device = 0;
  int num_sms = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t copy_stream[2] = {nullptr, nullptr};
  cudaEvent_t ev[8] = {};
  std::string err;
  int fp_contract = 1;
  int match_csa = 1;
  int harris_on = 0;       // goodFeaturesToTrack(useHarrisDetector): the reference passes false (keypoints.h:139)
  double harris_k = 0.04;  // cv::goodFeaturesToTrack's default k
  int resp_band = 0;  // response kernel: forced box height (0 = chosen per launch)
  int two_streams = 1; // stereo plans: large batches as two halves on two streams (VSLAM_OPT_TWO_STREAMS)
  int resp_multi = 1; // response kernel: a narrow last column strip of 2-3 images per warp (VSLAM_OPT_RESPONSE_MULTI)
  int match_split_small = 1;  // batches with fewer jobs than SMs/2: split every job's columns over CTAs
  int use_graph = 1;          // small pipeline plans of vslam_stereo_batch replay their launches as one CUDA graph
  int match_mma = 1;          // batches with at least SMs/2 jobs: Hamming matrix on the tensor cores (match_mma.cu)
  long long launches = 0;
  float last_match_ms = 0.f;
  // scratch for the single-call entry points
  DevBuf img, xy, ang, desc, cnt, lam;
  DevBuf cand, candcnt, imgmax, selws;
  DevBuf d1, d2, pairs, mcnt, rowpart, colstat, jobs;
  DevBuf d1x, d2x;  // expanded (signed byte) copies of d1 / d2 for the tensor-core matcher
  DevBuf lm;  // vslam_match_landmarks staging
  DevBuf geo;  // geometry.cu staging (camera models, essential inliers, landmark projection)
  // orb.cu: pyramid, blurred pyramid, candidates, counters, keypoints, descriptors, staging, rotations, tables, batch
  // counts | status, the second keypoint / descriptor buffers (chunks alternate)
  DevBuf orb[12];
  int orb_tab_key[4] = {0, 0, 0, 0};  // geometry the resize tables in orb[8] were built for (w, h, levels, scale bits)
  DevBuf pyr;  // vslam_pyramid_detect_describe_batch: levels 1.. of the chunk in flight
  void* pinned = nullptr;  // small pinned staging block for counts
  void* pinned_status = nullptr;  // vslam_stereo_batch: per-image detector statuses when the caller passes no n_kp
  size_t pinned_status_bytes = 0;
  void* pinned_out = nullptr;     // vslam_stereo_batch, single-chunk calls of a few pairs: the plan's whole result arena
  size_t pinned_out_bytes = 0;    // arrives here in ONE copy and is handed out to the caller's arrays by the CPU
  // pipeline plans of vslam_stereo_batch, kept between calls (same geometry -> no reallocation)
  static constexpr int kMaxBatchDepth = 4;
  struct vslam_stereo_plan* batch_plans[kMaxBatchDepth] = {nullptr, nullptr, nullptr, nullptr};
  int batch_chunk = 120;  // stereo pairs per pipeline chunk (VSLAM_OPT_BATCH_CHUNK): 240 images x 7 strips = one wave of full-height boxes
  int batch_depth = 4;    // chunks in flight: upload / compute / download overlap (VSLAM_OPT_BATCH_DEPTH)
};

#define VS_CUDA(ctx, call)                                                              \
  do {                                                                                  \
    cudaError_t e_ = (call);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      char b_[512];                                                                     \
      snprintf(b_, sizeof b_, "%s:%d: %s: %s", __FILE__, __LINE__, #call,               \
               cudaGetErrorString(e_));                                                 \
      (ctx)->err = b_;                                                                  \
      return VSLAM_E_CUDA;                                                              \
    }                                                                                   \
  } while (0)

static inline int vs_fail(vslam_ctx* ctx, int code, const char* msg) {
  if (ctx) ctx->err = msg;
  return code;
}

// ---- image set on the device: count images of h rows, pitch (multiple of 128) bytes apart ----
struct DevImages {
  const uint8_t* base;
  int w, h, pitch;
  size_t stride;  // bytes between images (>= pitch*h, multiple of 128)
};

// ---- launchers (each file's kernels stay private to it) ----
// detect.cu
struct DetectWorkspace {
  uint64_t* cand;      // [count][cand_cap] keys (float bits << 32 | y << 16 | x)
  uint32_t* cand_cnt;  // [count]
  uint32_t* img_max;   // [count] ordered-int encoding of the max response
  int cand_cap;
  uint8_t* sel_ws;     // global fallback workspace for select, [count][sel_ws_stride]
  size_t sel_ws_stride;
  // response of a FEW images (launch_response, detect.cu "small-count path"): per image row sums [3][h][w] f64,
  // box sums [3][h][w] f32, response [h][w] f32; room for small_images images (0: none)
  float* small_ws;
  int small_images;
};
constexpr int kRespSmallMaxImages = 16;  // up to this many images per launch take the small-count path
size_t detect_sel_ws_bytes(int w, int h, int cand_cap, int cell);
// 3-D TMA tensor map {w, h, count} over an image set with a box of box_w x box_h x 1 bytes
int make_image_tmap(vslam_ctx* ctx, const DevImages& im, int count, int box_w, int box_h, CUtensorMap* tmap);
int launch_response(vslam_ctx* ctx, const DevImages& im, int count, const DetectWorkspace& ws,
                    float* lam_out /*optional, [count][h][w]*/);
int launch_select(vslam_ctx* ctx, const DevImages& im, int count, const DetectWorkspace& ws,
                  int num_features, double quality, int min_distance, int border, double* xy,
                  int32_t* n_kp, int cap);
// describe.cu
int launch_orient_describe(vslam_ctx* ctx, const DevImages& im, int count, const double* xy,
                           const int32_t* n_kp, int cap, int mode_angles /*0: zero, 1: compute, 2: given*/,
                           double* angles, uint32_t* desc /*nullable*/,
                           uint8_t* descx = nullptr /*optional [count][cap][256]: descriptors as signed bytes*/);
// match.cu
struct MatchJob {
  const uint32_t* d1;
  const uint32_t* d2;
  const int32_t* n1p;  // device counts (used when non-null)
  const int32_t* n2p;
  int n1, n2;
  int32_t* pairs;  // [cap][2]
  int32_t* count;  // out
};
// rowpart / colstat: scratch of the small-batch column-split path (nullptr: the ctx's own, ctx stream only)
int launch_match_jobs(vslam_ctx* ctx, const MatchJob* jobs_dev, int n_jobs, int max_n1, int max_n2,
                      int threshold, double ratio, int cap, DevBuf* rowpart = nullptr, DevBuf* colstat = nullptr);
int launch_match_large(vslam_ctx* ctx, const uint32_t* d1, int n1, const uint32_t* d2, int n2,
                       int threshold, double ratio, int32_t* pairs, int32_t* count, int cap,
                       int32_t* best2_out /*nullable: [n1][3], skips the match predicate*/);
bool match_fits_fused(int n1, int n2);
// DB-sharded matching: part4 [n1][4] = (m1, a1 + index_offset, m2, col_ok) of one shard; merge of n_parts parts
int launch_match_shard(vslam_ctx* ctx, const uint32_t* d1, int n1, const uint32_t* d2, int n2, int index_offset,
                       int threshold, double ratio, int32_t* part4);
int launch_merge_shards(vslam_ctx* ctx, const int32_t* parts4, int n_parts, int n1, int threshold, double ratio,
                        int32_t* pairs, int32_t* count, int cap);
// jobs_i2: [n_jobs][2] set indices into (offsets, counts); fills MatchJob records on the device
int launch_build_store_jobs(vslam_ctx* ctx, const int32_t* jobs_i2, int n_jobs, const uint32_t* desc,
                            const int64_t* offsets, const int32_t* counts, int n_sets, int32_t* pairs,
                            int32_t* out_counts, int cap, MatchJob* out);
int launch_popc_peak(vslam_ctx* ctx, double* popc_per_s);
// match_mma.cu: the same jobs with the Hamming matrix on the tensor cores (tcgen05.mma kind::i8).  The jobs' d1 / d2
// must point into the packed arrays x1.packed / x2.packed; `expanded` holds those arrays' descriptors as 256 signed
// bytes each (launch_expand_desc), row r = the descriptor at packed + 8 r.
struct MatchExpanded {
  const uint32_t* packed;
  const uint8_t* expanded;
  long long rows;
};
bool match_mma_fits(int max_n1, int max_n2);
// rows_per_set / n_valid: optional (n_valid may be nullptr): only the first n_valid[set] rows of every set of
// rows_per_set rows are expanded (the describe stage's fixed-stride slabs)
int launch_expand_desc(vslam_ctx* ctx, const uint32_t* desc, long long rows, int rows_per_set, const int32_t* n_valid,
                       uint8_t* expanded);
int launch_match_jobs_mma(vslam_ctx* ctx, const MatchJob* jobs_dev, int n_jobs, int max_n1, int max_n2, int threshold,
                          double ratio, int cap, const MatchExpanded& x1, const MatchExpanded& x2);
int match_mma_split_range(const vslam_ctx* ctx, int n2);
int launch_match_split_mma(vslam_ctx* ctx, const uint8_t* x1, int n1, const uint8_t* x2, int n2, int range, uint2* rowpart,
                           uint2* colstat);
int match_mma_timeouts(int* flag);
// pyramid.cu: dst level = 2x2 rounded mean of src (dst.w = src.w / 2, dst.h = src.h / 2)
int launch_pyr_down(vslam_ctx* ctx, const DevImages& src, const DevImages& dst, int count);
// orb.cu: cv::ORB front end on images already on the device; outputs stay on the device
struct OrbDeviceOut {
  float* kp;        // [count][cap][6]
  uint8_t* desc;    // [count][cap][32]
  int32_t* n_kp;    // [count]
  int32_t* status;  // [count]
};
int launch_orb(vslam_ctx* ctx, const uint8_t* images_dev, int count, int w, int h, int pitch, size_t stride,
               const vslam_orb_params* prm, int cap, OrbDeviceOut* out);
int orb_check_status(vslam_ctx* ctx, const int32_t* n_kp, const int32_t* status, int count, int cap);
// landmarks.cu
int launch_match_landmarks(vslam_ctx* ctx, const double* kp_xy, const uint32_t* kp_desc, int n_kp,
                           const double* proj_xy, int n_proj, const int32_t* obs_off, const uint32_t* obs_desc,
                           double max_dist_2d, int threshold, double ratio, int32_t* result, int32_t* pairs,
                           int32_t* count, int cap);
// geometry.cu: findInliersEssential for n_jobs independent match lists that are already on the device
// (xy1/xy2: corner slabs, job k at + k*xy_job_stride doubles; matches: job k at + k*match_job_stride pairs;
// n_matches_dev: per-job counts on the device, or nullptr -> n_matches_max for every job).
size_t essential_ws_bytes(int n_matches_max, int n_jobs);  // keep flags + block counts
int launch_essential_jobs(vslam_ctx* ctx, const vslam_camera* cam1, const vslam_camera* cam2, const double* E,
                          double thr, const double* xy1, const double* xy2, size_t xy_job_stride,
                          const int32_t* matches, size_t match_job_stride, const int32_t* n_matches_dev,
                          int n_matches_max, int n_jobs, uint8_t* ws /*essential_ws_bytes()*/, int32_t* inliers,
                          size_t inlier_job_stride, int cap, int32_t* n_inliers, double* err_out);
