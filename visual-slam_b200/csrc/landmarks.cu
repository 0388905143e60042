// landmarks.cu -- radius-gated matching of keypoints against projected map landmarks, for sm_100a.
//
// Semantics: visnav::find_matches_landmarks (reference include/visnav/vo_utils.h:83-167).  For every
// keypoint i (corner c_i, descriptor d_i), in order:
//   * candidates = projected landmarks p, in input order, with |c_i - p| < match_max_dist_2d
//     (Eigen norm: sqrt(dx*dx + dy*dy) in double, :108-109);
//   * distance of a candidate = min over the landmark's observations of hamming(obs, d_i), starting
//     from 256 (:113-128);
//   * the two smallest are taken with std::partial_sort(first, first + 2, last) (:132-137); the match
//     (i, landmark[0]) is kept iff dist[0] < threshold and !(dist[1] < dist[0] * dist_2_best), with
//     256 standing in for dist[1] when there is a single candidate (:139-162).
// partial_sort is not stable: WHICH of several equally good candidates ends up first follows from
// libstdc++'s heap algorithm (bits/stl_algo.h __heap_select + __sort_heap, bits/stl_heap.h), restated
// in heap2_push() below so that even accepted ties (distance 0 twice) name the same landmark.
//
// Design: one warp per keypoint.  Lanes test 32 projected points per step against the radius
// (ballot keeps input order); for each hit the lanes split the landmark's observations, one
// 256-bit XOR + POPC each, and fold with redux.min.  Integer / fp64 work, a few MB per frame:
// latency-bound, no tensor cores.  A second tiny kernel compacts the per-keypoint results in
// keypoint order.
#include "common.cuh"

namespace {

constexpr int kLmWarps = 8;

struct LmParams {
  const double* kp_xy;       // [n_kp][2]
  const uint32_t* kp_desc;   // [n_kp][8]
  int n_kp;
  const double* proj_xy;     // [n_proj][2]
  int n_proj;
  const int32_t* obs_off;    // [n_proj + 1]
  const uint32_t* obs_desc;  // [total_obs][8]
  double max_dist_2d;
  int threshold;
  double ratio;
  int32_t* result;           // [n_kp]: matched projected index or -1
};

// The two-element max-heap libstdc++ keeps while partial_sort(first, first + 2, last) scans the
// candidates: h0 is the heap top (the worse of the two), h1 the other.  After the scan
// sort_heap leaves [0] = h1, [1] = h0.
struct Heap2 {
  int n;
  int id0, d0, id1, d1;
};
__device__ __forceinline__ void heap2_push(Heap2& h, int id, int d) {
  if (h.n == 0) {
    h.id1 = id; h.d1 = d;            // lone element; ordered when the second arrives
  } else if (h.n == 1) {
    // make_heap on (e0, e1): e1 < e0 keeps e0 on top, otherwise (ties included) e1 goes on top
    const int e0id = h.id1, e0d = h.d1;
    if (d < e0d) { h.id0 = e0id; h.d0 = e0d; h.id1 = id; h.d1 = d; }
    else { h.id0 = id; h.d0 = d; h.id1 = e0id; h.d1 = e0d; }
  } else if (d < h.d0) {
    // pop_heap(first, middle, i): the top leaves; the newcomer goes on top only if the other is smaller
    if (h.d1 < d) { h.id0 = id; h.d0 = d; }
    else { h.id0 = h.id1; h.d0 = h.d1; h.id1 = id; h.d1 = d; }
  }
  ++h.n;
}

__global__ void __launch_bounds__(kLmWarps * 32) match_landmarks_kernel(const LmParams p) {
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x * kLmWarps + (threadIdx.x >> 5);
  if (i >= p.n_kp) return;
  const double cx = p.kp_xy[2 * i], cy = p.kp_xy[2 * i + 1];
  uint32_t a[8];
  {
    const uint4 lo = __ldg(reinterpret_cast<const uint4*>(p.kp_desc) + 2 * (size_t)i);
    const uint4 hi = __ldg(reinterpret_cast<const uint4*>(p.kp_desc) + 2 * (size_t)i + 1);
    a[0] = lo.x; a[1] = lo.y; a[2] = lo.z; a[3] = lo.w;
    a[4] = hi.x; a[5] = hi.y; a[6] = hi.z; a[7] = hi.w;
  }
  Heap2 h;
  h.n = 0; h.id0 = h.d0 = h.id1 = h.d1 = 0;
  for (int base = 0; base < p.n_proj; base += 32) {
    const int j = base + lane;
    bool hit = false;
    if (j < p.n_proj) {
      const double dx = cx - p.proj_xy[2 * j], dy = cy - p.proj_xy[2 * j + 1];
      // Eigen's Vector2d::norm(): packet multiply, horizontal add, sqrt -- no fused multiply-add
      const double dist = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
      hit = dist < p.max_dist_2d;
    }
    unsigned hits = __ballot_sync(0xFFFFFFFFu, hit);
    while (hits) {
      const int b = __ffs(hits) - 1;
      hits &= hits - 1;
      const int lm = base + b;
      const int o0 = p.obs_off[lm], o1 = p.obs_off[lm + 1];
      uint32_t best = 256;
      for (int o = o0 + lane; o < o1; o += 32) {
        const uint4 lo = __ldg(reinterpret_cast<const uint4*>(p.obs_desc) + 2 * (size_t)o);
        const uint4 hi = __ldg(reinterpret_cast<const uint4*>(p.obs_desc) + 2 * (size_t)o + 1);
        const uint32_t d = __popc(a[0] ^ lo.x) + __popc(a[1] ^ lo.y) + __popc(a[2] ^ lo.z) + __popc(a[3] ^ lo.w) +
                           __popc(a[4] ^ hi.x) + __popc(a[5] ^ hi.y) + __popc(a[6] ^ hi.z) + __popc(a[7] ^ hi.w);
        best = min(best, d);
      }
      best = __reduce_min_sync(0xFFFFFFFFu, best);
      heap2_push(h, lm, (int)best);  // warp-uniform
    }
  }
  if (lane == 0) {
    int res = -1;
    if (h.n >= 1) {
      const int bd = h.d1, bid = h.id1;                 // [0] after sort_heap
      const int sd = h.n >= 2 ? h.d0 : 256;             // [1], or the literal 256 of vo_utils.h:146
      if (bd < p.threshold && !((double)sd < __dmul_rn((double)bd, p.ratio))) res = bid;
    }
    p.result[i] = res;
  }
}

// Ordered compaction of result[] into (keypoint, projected index) pairs; one CTA.
__global__ void __launch_bounds__(1024) compact_landmark_matches_kernel(const int32_t* __restrict__ result, int n_kp,
                                                                        int32_t* pairs, int32_t* count, int cap) {
  __shared__ int warp_sums[32];
  __shared__ int running;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) running = 0;
  __syncthreads();
  for (int base = 0; base < n_kp; base += 1024) {
    const int i = base + tid;
    const int r = i < n_kp ? result[i] : -1;
    const bool ok = r >= 0;
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, ok);
    if (lane == 0) warp_sums[warp] = __popc(bal);
    __syncthreads();
    int off = running, tot = 0;
    for (int w = 0; w < 32; ++w) {
      if (w < warp) off += warp_sums[w];
      tot += warp_sums[w];
    }
    const int pos = off + __popc(bal & ((1u << lane) - 1));
    if (ok && pos < cap) {
      pairs[2 * pos] = i;
      pairs[2 * pos + 1] = r;
    }
    __syncthreads();
    if (tid == 0) running += tot;
    __syncthreads();
  }
  if (tid == 0) *count = running;
}

}  // namespace

int launch_match_landmarks(vslam_ctx* ctx, const double* kp_xy, const uint32_t* kp_desc, int n_kp,
                           const double* proj_xy, int n_proj, const int32_t* obs_off, const uint32_t* obs_desc,
                           double max_dist_2d, int threshold, double ratio, int32_t* result, int32_t* pairs,
                           int32_t* count, int cap) {
  LmParams p;
  p.kp_xy = kp_xy; p.kp_desc = kp_desc; p.n_kp = n_kp;
  p.proj_xy = proj_xy; p.n_proj = n_proj; p.obs_off = obs_off; p.obs_desc = obs_desc;
  p.max_dist_2d = max_dist_2d; p.threshold = threshold; p.ratio = ratio; p.result = result;
  match_landmarks_kernel<<<(n_kp + kLmWarps - 1) / kLmWarps, kLmWarps * 32, 0, ctx->stream>>>(p);
  compact_landmark_matches_kernel<<<1, 1024, 0, ctx->stream>>>(result, n_kp, pairs, count, cap);
  ctx->launches += 2;
  VS_CUDA(ctx, cudaGetLastError());
  return VSLAM_OK;
}
