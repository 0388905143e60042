// orb.cu -- cv::ORB::detectAndCompute as visnav::compute_bow_vector configures it (reference keypoints.h:249:
// cv::ORB::create(num_features, 1.2, 8, 19, 0, 2, cv::ORB::FAST_SCORE); patchSize 31, fastThreshold 20), batched over
// images, for sm_100a.  SURVEY.md §8f-4, ORB half: the second detector the reference runs per keyframe, feeding the
// vocabulary transform (bow.cu).
//
// Semantics (thirdparty/opencv/modules; every stage below was first restated on the CPU -- the test-only ORB restatement under oracle/
// -- and pinned to cv2.ORB):
//   pyramid      features2d/src/orb.cpp:1063-1130: level l is cvRound(size / 1.2^l), made from level l-1 by
//                resize(INTER_LINEAR_EXACT): 8.8 fixed-point taps (imgproc/src/resize.cpp:733-771, :344-398, :691-731)
//   FAST-9/16    features2d/src/fast.cpp:56-290: corner iff 9 contiguous ring pixels are all brighter than v + t or
//                all darker than v - t; score = the largest t that keeps it a corner (fast_score.cpp:119-215);
//                kept iff the score is strictly above its 8 neighbours' scores
//   selection    keypoint.cpp:105-117 (19 px border), :69-90 retainBest: every keypoint whose score reaches the n-th
//                largest of its level (ties all stay); per-level budgets orb.cpp:843-856
//   orientation  orb.cpp:181-216 intensity centroid over the radius-15 disc, fastAtan2 (mathfuncs_core.simd.hpp:50-71)
//   blur         GaussianBlur(7x7, sigma 2) on a sub-matrix -> the float separable filter (smooth.dispatch.cpp:611-757,
//                filter.dispatch.cpp:290-375), multiply-adds fused as OpenCV's AVX2 dispatch executes them (measured
//                against cv2: 0 differing pixels in 12 M with fused adds, 57 without)
//   descriptor   orb.cpp:219-268 with bit_pattern_31_ = the reference's own rBRIEF table (rbrief_pattern.inc)
// Keypoint ORDER inside a level is std::nth_element's in the reference (unspecified); here: raster order.  The set,
// and every keypoint's fields and descriptor, are the reference's.
//
// Data layout in HBM: one pyramid slab per image (levels back to back, row pitch a multiple of 16 bytes), a second
// slab for the blurred levels; per (image, level) a candidate list of packed keys (score << 24 | y << 12 | x) with a
// 256-bin score histogram; keypoints [count][cap] x {x, y, size, angle, response, octave} floats and [count][cap][32]
// descriptor bytes.  All byte/integer work except the blur and the rotation of the sampling pattern.
#include <math.h>

#include <vector>

#include "common.cuh"

namespace {

constexpr int kMaxLevels = 8;
constexpr int kMaxLevelKp = 8192;  // keypoints one level may keep (sorted in shared memory)

struct OrbGeom {
  int nlevels;
  int w[kMaxLevels], h[kMaxLevels], pitch[kMaxLevels];
  long long off[kMaxLevels];  // byte offset of the level inside an image's slab
  long long slab;             // bytes per image
  float scale[kMaxLevels];
  int nper[kMaxLevels];
  int cand_off[kMaxLevels];   // offset of the level's candidate list inside an image's candidate block (keys)
  int cand_cap[kMaxLevels];
  int cand_total;             // keys per image
  int tab_off[kMaxLevels];    // offset of the level's resize tables ({x ofs, x c1} x w, {y ofs, y c1} x h)
};

// ---- level 0: copy the caller's images into the slabs -----------------------------------------------------------
__global__ void orb_copy_level0_kernel(const uint8_t* src, int w, int h, int pitch, size_t stride, OrbGeom g, uint8_t* pyr) {
  const int img = blockIdx.z;
  const int y = blockIdx.y;
  const uint8_t* s = src + (size_t)img * stride + (size_t)y * pitch;
  uint8_t* d = pyr + (size_t)img * g.slab + (size_t)y * g.pitch[0];
  for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < w; x += gridDim.x * blockDim.x) d[x] = s[x];
}

// ---- resize INTER_LINEAR_EXACT: one thread per destination pixel ---------------------------------------------------
// tab: [w] {ofs, c1} then [h] {ofs, c1} (uint16 pairs); c0 = 256 - c1.  Destination pixels left of / beyond the
// interpolated range carry c1 = 0 and the edge offset, which reproduces hlineResize's replicated ends and vlineSet.
__global__ void orb_resize_kernel(OrbGeom g, int l, const ushort2* tab, uint8_t* pyr) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y;
  const int img = blockIdx.z;
  if (x >= g.w[l]) return;
  const ushort2 tx = tab[x], ty = tab[g.w[l] + y];
  const int sw = g.w[l - 1], sh = g.h[l - 1];
  const uint8_t* s = pyr + (size_t)img * g.slab + g.off[l - 1];
  const int x1 = min(tx.x + 1, sw - 1), y1 = min(ty.x + 1, sh - 1);
  const uint8_t* r0 = s + (size_t)ty.x * g.pitch[l - 1];
  const uint8_t* r1 = s + (size_t)y1 * g.pitch[l - 1];
  const uint32_t cx1 = tx.y, cx0 = 256u - cx1, cy1 = ty.y, cy0 = 256u - cy1;
  const uint32_t a = cx0 * r0[tx.x] + cx1 * r0[x1];  // ufixedpoint16 row values
  const uint32_t b = cx0 * r1[tx.x] + cx1 * r1[x1];
  const uint32_t v = (a * cy0 + b * cy1 + 32768u) >> 16;
  pyr[(size_t)img * g.slab + g.off[l] + (size_t)y * g.pitch[l] + x] = (uint8_t)min(v, 255u);
}

// ---- FAST-9/16 score + 3x3 non-maximum suppression + border filter -------------------------------------------------
constexpr int kFastTW = 62, kFastTH = 32;                     // outputs per CTA
constexpr int kFastSW = kFastTW + 2, kFastSH = kFastTH + 2;   // score tile (1 px NMS halo): 64 wide = two lane passes
constexpr int kFastIH = kFastTH + 8;                          // image tile rows: + 3 (ring) on every side
constexpr int kFastIP = 80;                                   // image tile row pitch: 19 words + padding
constexpr int kFastWords = 19;                                // 62 + 8 columns + up to 3 of alignment slack

__device__ __forceinline__ int ring_offset(int k, int pitch) {
  // fast_score.cpp:52-56: (dx, dy) of ring pixel k
  const int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
  const int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
  return dx[k] + dy[k] * pitch;
}

// High-speed rejection (fast.cpp:208-222) on the raw ring values: an arc of 9 holds one pixel of each opposite pair
// (k, k + 8), k = 0, 2, 4, 6.  All-darker needs min(r_k, r_k+8) < v - t for every pair, all-brighter max(...) > v + t.
__device__ __forceinline__ bool fast_pretest(const uint8_t* c, int thr) {
  const int v = c[0];
  int lo = 0, hi = 255;  // max over pairs of the pair minimum, min over pairs of the pair maximum
#pragma unroll
  for (int k = 0; k < 8; k += 2) {
    const int r0 = c[ring_offset(k, kFastIP)], r1 = c[ring_offset(k + 8, kFastIP)];
    lo = max(lo, min(r0, r1));
    hi = min(hi, max(r0, r1));
  }
  return (v - lo > thr) | (hi - v > thr);
}

// Corner test and score in one: with A = min over the 16 arcs of (max ring value of the arc) and B = max over the arcs
// of (min ring value), the pixel is a corner at threshold t iff v - A > t (an all-darker arc) or B - v > t (an
// all-brighter arc), and cornerScore<16> (fast_score.cpp:119-215) = max(v - A, B - v) - 1.  Arc extrema by doubling:
// windows of 2, 4, 8 ring pixels, then the ninth.  The two sides are kept apart down to the last line: nvcc 12.9
// miscompiles the per-arc form max(best, max(arc_min_d, -arc_max_d)) for sm_100a (tools/probes/t_fast2.cu).
__device__ int fast_corner_score(const uint8_t* c) {
  int r[16], lo[16], hi[16], t1[16], t2[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) r[k] = c[ring_offset(k, kFastIP)];
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    lo[k] = min(r[k], r[(k + 1) & 15]);
    hi[k] = max(r[k], r[(k + 1) & 15]);
  }
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    t1[k] = min(lo[k], lo[(k + 2) & 15]);
    t2[k] = max(hi[k], hi[(k + 2) & 15]);
  }
  int a = 255, b = 0;
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    a = min(a, max(max(t2[k], t2[(k + 4) & 15]), r[(k + 8) & 15]));
    b = max(b, min(min(t1[k], t1[(k + 4) & 15]), r[(k + 8) & 15]));
  }
  const int v = c[0];
  return max(v - a, b - v) - 1;
}

struct FastParams {
  OrbGeom g;
  const uint8_t* pyr;
  int threshold, edge;
  uint32_t* cand;      // [count][cand_total]
  uint32_t* cand_cnt;  // [count][kMaxLevels]
  uint32_t* hist;      // [count][kMaxLevels][256]
  int tiles_x[kMaxLevels], tiles_y[kMaxLevels], tile_first[kMaxLevels + 1];
};

__global__ void __launch_bounds__(256) orb_fast_kernel(const FastParams p) {
  __shared__ __align__(16) uint8_t s_img[kFastIH][kFastIP];
  __shared__ uint8_t s_score[kFastSH][kFastSW + 4];
  __shared__ uint16_t s_queue[kFastSH * kFastSW];
  __shared__ int s_qn;
  const int img = blockIdx.y;
  if (threadIdx.x == 0) s_qn = 0;
  int l = 0;
  while (l + 1 < p.g.nlevels && (int)blockIdx.x >= p.tile_first[l + 1]) ++l;
  const int t = blockIdx.x - p.tile_first[l];
  const int tx0 = (t % p.tiles_x[l]) * kFastTW, ty0 = (t / p.tiles_x[l]) * kFastTH;
  const int W = p.g.w[l], H = p.g.h[l], pitch = p.g.pitch[l];
  const uint8_t* base = p.pyr + (size_t)img * p.g.slab + p.g.off[l];
  // the image tile by aligned 32-bit words (the row pitch is a multiple of 16, so a word lies entirely inside
  // [0, pitch) or entirely outside; bytes between W and pitch are never used below).  Tile column of image x:
  // x - gxb; score-tile position xx sits at image x = tx0 - 1 + xx, i.e. tile column xx + 3 + xoff.
  const int gxb = (tx0 - 4) & ~3;
  const int xoff = (tx0 - 4) - gxb;
  for (int i = threadIdx.x; i < kFastIH * 32; i += blockDim.x) {
    const int yy = i >> 5, wx = i & 31;
    if (wx >= kFastWords) continue;
    const int gx = gxb + 4 * wx, gy = ty0 - 4 + yy;
    uint32_t word = 0;
    if (gx >= 0 && gx < pitch && gy >= 0 && gy < H) word = *reinterpret_cast<const uint32_t*>(base + (size_t)gy * pitch + gx);
    *reinterpret_cast<uint32_t*>(&s_img[yy][4 * wx]) = word;
  }
  __syncthreads();
  // pass 1: the high-speed rejection for every position; survivors are queued so that pass 2 runs on dense warps
  for (int i = threadIdx.x; i < kFastSH * kFastSW; i += blockDim.x) {
    const int yy = i >> 6, xx = i & 63;
    const int gx = tx0 - 1 + xx, gy = ty0 - 1 + yy;
    s_score[yy][xx] = 0;
    if (gx >= 3 && gx < W - 3 && gy >= 3 && gy < H - 3 &&  // fast.cpp:96, :206
        fast_pretest(&s_img[yy + 3][xx + 3 + xoff], p.threshold))
      s_queue[atomicAdd(&s_qn, 1)] = (uint16_t)i;
  }
  __syncthreads();
  // pass 2: ring test and score
  for (int q = threadIdx.x; q < s_qn; q += blockDim.x) {
    const int i = s_queue[q];
    const int yy = i >> 6, xx = i & 63;
    const int sc = fast_corner_score(&s_img[yy + 3][xx + 3 + xoff]);
    if (sc >= p.threshold) s_score[yy][xx] = (uint8_t)sc;
  }
  __syncthreads();
  const int edge = p.edge;
  if (H <= 2 * edge || W <= 2 * edge) return;  // runByImageBorder clears everything (keypoint.cpp:109-110)
  for (int i = threadIdx.x; i < kFastTH * 64; i += blockDim.x) {
    const int yy = i >> 6, xx = i & 63;
    if (xx >= kFastTW) continue;
    const int gx = tx0 + xx, gy = ty0 + yy;
    const int v = s_score[yy + 1][xx + 1];
    if (v == 0 || gx < edge || gx >= W - edge || gy < edge || gy >= H - edge) continue;
    const uint8_t* s = &s_score[yy + 1][xx + 1];
    constexpr int sp = kFastSW + 4;
    if (v > s[1] && v > s[-1] && v > s[-sp - 1] && v > s[-sp] && v > s[-sp + 1] && v > s[sp - 1] && v > s[sp] &&
        v > s[sp + 1]) {
      const uint32_t pos = atomicAdd(p.cand_cnt + img * kMaxLevels + l, 1u);
      if (pos < (uint32_t)p.g.cand_cap[l])
        p.cand[(size_t)img * p.g.cand_total + p.g.cand_off[l] + pos] = ((uint32_t)v << 24) | ((uint32_t)gy << 12) | (uint32_t)gx;
      atomicAdd(p.hist + ((size_t)img * kMaxLevels + l) * 256 + v, 1u);
    }
  }
}

// ---- retainBest per level + raster order: one CTA per image -----------------------------------------------------------
struct SelectParams {
  OrbGeom g;
  const uint32_t* cand;
  const uint32_t* cand_cnt;
  const uint32_t* hist;
  float* kp;        // [count][cap][6]
  int32_t* n_kp;    // [count]
  int32_t* status;  // [count]: bit 0 = a level overflowed its candidate list, bit 1 = more than kMaxLevelKp kept
  int cap;
};

__device__ void bitonic_asc_u32_blk(uint32_t* keys, int n) {
  for (int k = 2; k <= n; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int l = i ^ j;
        if (l > i) {
          const uint32_t a = keys[i], b = keys[l];
          if (((i & k) == 0) ? (a > b) : (a < b)) {
            keys[i] = b;
            keys[l] = a;
          }
        }
      }
      __syncthreads();
    }
}

// One CTA per (level, image).  Warp j <= l derives level j's cut-off score and kept count from its histogram, so every
// CTA knows where its level starts in the image's keypoint list without waiting for the levels before it.
__global__ void __launch_bounds__(512) orb_select_kernel(const SelectParams p) {
  __shared__ uint32_t s_keys[kMaxLevelKp];
  __shared__ int s_thr[kMaxLevels], s_kept[kMaxLevels], s_cnt;
  const int l = blockIdx.x, img = blockIdx.y;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_cnt = 0;
  if (warp <= l) {
    // KeyPointsFilter::retainBest (keypoint.cpp:69-90) for level `warp`: nothing to do unless there are more than
    // n_points candidates; then everything whose response reaches the n_points-th largest stays
    const int j = warp;
    const uint32_t raw = p.cand_cnt[img * kMaxLevels + j];
    const int n = (int)min(raw, (uint32_t)p.g.cand_cap[j]);
    const int nper = p.g.nper[j];
    int thr = 0, kept = n;
    if (n > nper) {
      if (nper == 0) {
        thr = 256;
        kept = 0;
      } else {
        const uint32_t* h = p.hist + ((size_t)img * kMaxLevels + j) * 256 + 8 * lane;  // this lane: scores 8*lane .. +7
        int bins[8], mine = 0;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          bins[q] = (int)h[q];
          mine += bins[q];
        }
        int suffix = mine;  // candidates with a score in this lane's bins or above
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const int o = __shfl_down_sync(0xFFFFFFFFu, suffix, d);
          if (lane + d < 32) suffix += o;
        }
        const unsigned reach = __ballot_sync(0xFFFFFFFFu, suffix >= nper);
        const int top = 31 - __clz(reach);  // the highest lane whose suffix reaches n_points (reach != 0: n > nper)
        int acc = suffix - mine;
        int t_found = 8 * lane, k_found = suffix;
#pragma unroll
        for (int q = 7; q >= 0; --q) {
          acc += bins[q];
          if (acc >= nper) {
            t_found = 8 * lane + q;
            k_found = acc;
            break;
          }
        }
        thr = __shfl_sync(0xFFFFFFFFu, t_found, top);
        kept = __shfl_sync(0xFFFFFFFFu, k_found, top);
      }
    }
    if (lane == 0) {
      s_thr[j] = thr;
      s_kept[j] = kept;
    }
  }
  __syncthreads();
  int out = 0;
  for (int j = 0; j < l; ++j) out += min(s_kept[j], kMaxLevelKp);
  const uint32_t thr = (uint32_t)s_thr[l];
  const uint32_t raw = p.cand_cnt[img * kMaxLevels + l];
  const int n = (int)min(raw, (uint32_t)p.g.cand_cap[l]);
  int status = raw > (uint32_t)p.g.cand_cap[l] ? 1 : 0;
  const uint32_t* c = p.cand + (size_t)img * p.g.cand_total + p.g.cand_off[l];
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const uint32_t key = c[i];
    if ((key >> 24) >= thr) {
      const int pos = atomicAdd(&s_cnt, 1);
      if (pos < kMaxLevelKp) s_keys[pos] = (key << 8) | (key >> 24);  // y, x on top: raster order; score below
    }
  }
  __syncthreads();
  int m = s_cnt;
  if (m > kMaxLevelKp) {
    status |= 2;
    m = kMaxLevelKp;
  }
  int np2 = 32;
  while (np2 < m) np2 <<= 1;
  for (int i = m + threadIdx.x; i < np2; i += blockDim.x) s_keys[i] = 0xFFFFFFFFu;
  __syncthreads();
  bitonic_asc_u32_blk(s_keys, np2);
  const float sc = p.g.scale[l];
  for (int i = threadIdx.x; i < m; i += blockDim.x) {
    const uint32_t k = s_keys[i];
    const int x = (k >> 8) & 0xFFF, y = k >> 20;
    if (out + i < p.cap) {
      float* o = p.kp + ((size_t)img * p.cap + out + i) * 6;
      o[0] = __fmul_rn((float)x, sc);   // allKeypoints[i].pt *= scale (orb.cpp:993-997)
      o[1] = __fmul_rn((float)y, sc);
      o[2] = __fmul_rn(31.f, sc);       // size = patchSize * scale (:907)
      o[3] = 0.f;                       // angle: orb_angle_kernel
      o[4] = (float)(k & 0xFF);
      o[5] = (float)l;
    }
  }
  if (threadIdx.x == 0) {
    if (l == p.g.nlevels - 1) p.n_kp[img] = out + m;
    if (status) atomicOr(p.status + img, status);
  }
}

// ---- orientation: one warp per keypoint ---------------------------------------------------------------------------------
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
  // core/src/mathfuncs_core.simd.hpp:34-71, every operation rounded on its own (cv2 evaluates it unfused)
  const float p1 = 0.9997878412794807f * (float)(180 / 3.1415926535897932384626433832795);
  const float p3 = -0.3258083974640975f * (float)(180 / 3.1415926535897932384626433832795);
  const float p5 = 0.1555786518463281f * (float)(180 / 3.1415926535897932384626433832795);
  const float p7 = -0.04432655554792128f * (float)(180 / 3.1415926535897932384626433832795);
  const float eps = (float)2.2204460492503131e-16;
  const float ax = fabsf(x), ay = fabsf(y);
  float a;
  if (ax >= ay) {
    const float c = __fdiv_rn(ay, __fadd_rn(ax, eps));
    const float c2 = __fmul_rn(c, c);
    a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
  } else {
    const float c = __fdiv_rn(ax, __fadd_rn(ay, eps));
    const float c2 = __fmul_rn(c, c);
    a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
  }
  if (x < 0) a = __fsub_rn(180.f, a);
  if (y < 0) a = __fsub_rn(360.f, a);
  return a;
}

struct KpParams {
  OrbGeom g;
  const uint8_t* pyr;   // unblurred (angles) or blurred (descriptors) slabs
  float* kp;
  const int32_t* n_kp;
  uint8_t* desc;
  float2* rot;          // [count][cap] (cos, sin) of the keypoint angle, written by the angle kernel
  const char4* pattern;   // [256] sampling pairs (xa, ya, xb, yb)
  unsigned long long umax;  // orb.cpp:860-878 for halfPatchSize 15, 4 bits per row
  int cap;
};

__global__ void __launch_bounds__(256) orb_angle_kernel(const KpParams p) {
  const int lane = threadIdx.x & 31;
  const int j = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int img = blockIdx.y;
  if (j >= min(p.n_kp[img], p.cap)) return;
  float* k = p.kp + ((size_t)img * p.cap + j) * 6;
  const int l = (int)k[5];
  const float inv = __fdiv_rn(1.f, p.g.scale[l]);
  // the level coordinates: x * scale / scale rounds back to x for every integer x below 4096
  const int x = __float2int_rn(__fmul_rn(k[0], inv)), y = __float2int_rn(__fmul_rn(k[1], inv));
  const int pitch = p.g.pitch[l];
  const uint8_t* c = p.pyr + (size_t)img * p.g.slab + p.g.off[l] + (size_t)y * pitch + x;
  int m10 = 0, m01 = 0;
  if (lane < 31) {  // lane = row v + 15 of the disc
    const int v = lane - 15;
    const int d = (int)((p.umax >> (4 * abs(v))) & 15ull);
    const uint8_t* r = c + v * pitch;
    int s = 0;
    for (int u = -d; u <= d; ++u) {
      const int val = r[u];
      m10 += u * val;
      s += val;
    }
    m01 = v * s;
  }
  m10 = __reduce_add_sync(0xFFFFFFFFu, m10);
  m01 = __reduce_add_sync(0xFFFFFFFFu, m01);
  if (lane == 0) {
    const float deg = fast_atan2_deg((float)m01, (float)m10);
    k[3] = deg;
    // orb.cpp:231-232: angle *= (float)(CV_PI / 180.f); a = cosf(angle), b = sinf(angle) -- evaluated in double and
    // rounded once
    const float angle = __fmul_rn(deg, (float)(3.1415926535897932384626433832795 / 180.f));
    p.rot[(size_t)img * p.cap + j] = make_float2((float)cos((double)angle), (float)sin((double)angle));
  }
}

// ---- blur: separable 7-tap float filter, BORDER_REFLECT_101 ----------------------------------------------------------------
constexpr int kBlurTW = 64, kBlurTH = 16;

struct BlurParams {
  OrbGeom g;
  const uint8_t* pyr;
  uint8_t* blur;
  float k7[7];
  int tiles_x[kMaxLevels], tile_first[kMaxLevels + 1];
};

__device__ __forceinline__ int reflect101(int p, int len) {
  if (len == 1) return 0;
  while (p < 0 || p >= len) p = p < 0 ? -p : 2 * len - 2 - p;
  return p;
}

// u8 -> f32 on the ALU/FMA pipes: the byte in the mantissa of 2^23, minus 2^23
__device__ __forceinline__ float byte_f32(uint32_t word, int k) {
  return __fsub_rn(__uint_as_float(__byte_perm(word, 0x4B000000u, 0x7540u + k)), 8388608.f);
}

// One CTA = 64 x 16 output pixels.  The input tile (22 rows x 72 columns, image x = x0 - 4 ...) is converted to float
// once; interior tiles fetch it by aligned words, tiles that touch an image edge pixel by pixel through the reflection.
// Row pass: 4 adjacent outputs per item from three float4 loads; column pass: 4 adjacent outputs per thread from
// seven float4 loads, stored as one 32-bit word.
__global__ void __launch_bounds__(256) orb_blur_kernel(const BlurParams p) {
  __shared__ __align__(16) float s_in[kBlurTH + 6][72];
  __shared__ __align__(16) float s_row[kBlurTH + 6][kBlurTW];
  const int img = blockIdx.y;
  int l = 0;
  while (l + 1 < p.g.nlevels && (int)blockIdx.x >= p.tile_first[l + 1]) ++l;
  const int t = blockIdx.x - p.tile_first[l];
  const int tx0 = (t % p.tiles_x[l]) * kBlurTW, ty0 = (t / p.tiles_x[l]) * kBlurTH;
  const int W = p.g.w[l], H = p.g.h[l], pitch = p.g.pitch[l];
  const uint8_t* base = p.pyr + (size_t)img * p.g.slab + p.g.off[l];
  const bool interior = tx0 >= 4 && tx0 + 67 < W && ty0 >= 3 && ty0 + kBlurTH + 2 < H;  // warp-uniform
  if (interior) {
    for (int i = threadIdx.x; i < (kBlurTH + 6) * 32; i += blockDim.x) {
      const int yy = i >> 5, wx = i & 31;
      if (wx >= 18) continue;
      const uint32_t word = *reinterpret_cast<const uint32_t*>(base + (size_t)(ty0 - 3 + yy) * pitch + tx0 - 4 + 4 * wx);
      *reinterpret_cast<float4*>(&s_in[yy][4 * wx]) =
          make_float4(byte_f32(word, 0), byte_f32(word, 1), byte_f32(word, 2), byte_f32(word, 3));
    }
  } else {
    for (int i = threadIdx.x; i < (kBlurTH + 6) * 72; i += blockDim.x) {
      const int yy = i / 72, xx = i % 72;
      const uint32_t b = base[(size_t)reflect101(ty0 - 3 + yy, H) * pitch + reflect101(tx0 - 4 + xx, W)];
      s_in[yy][xx] = byte_f32(b, 0);
    }
  }
  __syncthreads();
  // RowFilter<uchar, float>: s = k[0] * p[0]; s += k[i] * p[i], left to right (fused); output x reads inputs x-3..x+3
  // = tile columns x + 1 .. x + 7
  for (int i = threadIdx.x; i < (kBlurTH + 6) * 16; i += blockDim.x) {
    const int yy = i >> 4, g = i & 15;
    const float4 a = *reinterpret_cast<const float4*>(&s_in[yy][4 * g]);
    const float4 b = *reinterpret_cast<const float4*>(&s_in[yy][4 * g + 4]);
    const float4 c = *reinterpret_cast<const float4*>(&s_in[yy][4 * g + 8]);
    const float in[12] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, c.x, c.y, c.z, c.w};
    float o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float s = __fmul_rn(p.k7[0], in[j + 1]);
#pragma unroll
      for (int q = 1; q < 7; ++q) s = __fmaf_rn(p.k7[q], in[j + 1 + q], s);
      o[j] = s;
    }
    *reinterpret_cast<float4*>(&s_row[yy][4 * g]) = make_float4(o[0], o[1], o[2], o[3]);
  }
  __syncthreads();
  // SymmColumnFilter: s = k[3] * r[0] + delta (0); s += k[3 + j] * (r[+j] + r[-j]); Cast<float, uchar> = cvRound, saturate
  {
    const int yy = threadIdx.x >> 4, g = threadIdx.x & 15;
    const int gx = tx0 + 4 * g, gy = ty0 + yy;
    if (gx < pitch && gy < H) {
      float4 r[7];
#pragma unroll
      for (int j = 0; j < 7; ++j) r[j] = *reinterpret_cast<const float4*>(&s_row[yy + j][4 * g]);
      uint32_t word = 0;
#pragma unroll
      for (int cmp = 0; cmp < 4; ++cmp) {
        const float v[7] = {(&r[0].x)[cmp], (&r[1].x)[cmp], (&r[2].x)[cmp], (&r[3].x)[cmp],
                            (&r[4].x)[cmp], (&r[5].x)[cmp], (&r[6].x)[cmp]};
        float s = __fmaf_rn(p.k7[3], v[3], 0.f);
#pragma unroll
        for (int j = 1; j <= 3; ++j) s = __fmaf_rn(p.k7[3 + j], __fadd_rn(v[3 + j], v[3 - j]), s);
        word |= (uint32_t)min(max(__float2int_rn(s), 0), 255) << (8 * cmp);
      }
      *reinterpret_cast<uint32_t*>(p.blur + (size_t)img * p.g.slab + p.g.off[l] + (size_t)gy * pitch + gx) = word;
    }
  }
}

// cvRound for |x| < 2^22 on the FMA/ALU pipes: adding 1.5 * 2^23 leaves the nearest-even integer in the mantissa
__device__ __forceinline__ int round_small(float x) {
  return __float_as_int(__fadd_rn(x, 12582912.f)) - 0x4B400000;
}

// ---- descriptors: one warp per keypoint, lane = descriptor byte ---------------------------------------------------------------
__global__ void __launch_bounds__(256) orb_describe_kernel(const KpParams p) {
  // the sampling pattern, transposed: s_pat[bit][lane] = pair 8 * lane + bit (lane-indexed reads of __constant__
  // memory serialise 32-fold: 2.8 ms per 256 images before, measured)
  __shared__ float4 s_pat[8][32];
  {
    const char4 q = p.pattern[8 * (threadIdx.x & 31) + (threadIdx.x >> 5)];
    s_pat[threadIdx.x >> 5][threadIdx.x & 31] = make_float4((float)q.x, (float)q.y, (float)q.z, (float)q.w);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int j = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int img = blockIdx.y;
  if (j >= min(p.n_kp[img], p.cap)) return;
  const float* k = p.kp + ((size_t)img * p.cap + j) * 6;
  const int l = (int)k[5];
  const float inv = __fdiv_rn(1.f, p.g.scale[l]);  // orb.cpp:230
  const int x = __float2int_rn(__fmul_rn(k[0], inv)), y = __float2int_rn(__fmul_rn(k[1], inv));
  const int pitch = p.g.pitch[l];
  const uint8_t* c = p.pyr + (size_t)img * p.g.slab + p.g.off[l] + (size_t)y * pitch + x;
  const float2 ab = p.rot[(size_t)img * p.cap + j];
  const float a = ab.x, b = ab.y;
  int val = 0;
#pragma unroll
  for (int bit = 0; bit < 8; ++bit) {
    const float4 q = s_pat[bit][lane];
    const float qx = q.x, qy = q.y, qz = q.z, qw = q.w;
    const float xa = __fsub_rn(__fmul_rn(qx, a), __fmul_rn(qy, b));
    const float ya = __fadd_rn(__fmul_rn(qx, b), __fmul_rn(qy, a));
    const float xb = __fsub_rn(__fmul_rn(qz, a), __fmul_rn(qw, b));
    const float yb = __fadd_rn(__fmul_rn(qz, b), __fmul_rn(qw, a));
    const int t0 = c[round_small(ya) * pitch + round_small(xa)];
    const int t1 = c[round_small(yb) * pitch + round_small(xb)];
    val |= (t0 < t1) << bit;
  }
  p.desc[((size_t)img * p.cap + j) * 32 + lane] = (uint8_t)val;
}

}  // namespace

// ---- host side -------------------------------------------------------------------------------------------------------------
static int cv_round_f(float v) { return (int)lrintf(v); }  // cvRound: round half to even

// interpolationLinear::getCoeffs (resize.cpp:739-765) for one axis, as {offset, c1} pairs (c0 = 256 - c1)
static void linear_table(int srcsize, int dstsize, std::vector<uint16_t>& tab) {
  const double inv_scale = (double)dstsize / srcsize;
  const double scale = 1.0 / inv_scale;
  for (int val = 0; val < dstsize; ++val) {
    const double fval = scale * ((double)val + 0.5) - 0.5;
    const int ival = (int)floor(fval);
    uint16_t ofs = 0, c1 = 0;
    if (ival >= 0 && srcsize > 1) {
      if (ival < srcsize - 1) {
        ofs = (uint16_t)ival;
        const double frac = fval - (double)ival;
        c1 = frac < 0 ? 0 : (uint16_t)lrint(frac * 256.0);
      } else {
        ofs = (uint16_t)(srcsize - 1);  // beyond max_x: the last source pixel
      }
    }  // left of min_x: source pixel 0
    tab.push_back(ofs);
    tab.push_back(c1);
  }
}

// Runs the whole ORB front end for `count` images already on the device; outputs stay on the device (ctx->orb[...]).
int launch_orb(vslam_ctx* ctx, const uint8_t* images_dev, int count, int w, int h, int pitch, size_t stride,
               const vslam_orb_params* prm, int cap, OrbDeviceOut* out) {
  if (prm->nlevels < 1 || prm->nlevels > kMaxLevels) return vs_fail(ctx, VSLAM_E_LIMIT, "ORB: 1..8 pyramid levels");
  if (w >= 4096 || h >= 4096) return vs_fail(ctx, VSLAM_E_LIMIT, "ORB: images up to 4095 x 4095");
  // edge_threshold >= 19: the rotated sampling pattern reaches 18 px (13 * sqrt(2), rounded) and the levels carry no
  // border here (OpenCV keeps 32 px of reflected, unblurred pixels around each level; the reference passes 19)
  if (prm->nfeatures < 0 || prm->edge_threshold < 19 || prm->fast_threshold < 1 || prm->fast_threshold > 254 ||
      !(prm->scale_factor > 1.f))
    return vs_fail(ctx, VSLAM_E_ARG, "ORB: nfeatures >= 0, edge_threshold >= 19 (pattern reach), 1 <= fast_threshold <= 254, scale_factor > 1");
  OrbGeom g;
  memset(&g, 0, sizeof g);
  g.nlevels = prm->nlevels;
  const double scale_factor = (double)prm->scale_factor;  // ORB_Impl keeps the float argument in a double
  {
    // per-level budgets, orb.cpp:843-856 (float arithmetic)
    const float factor = (float)(1.0 / scale_factor);
    float nd = prm->nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)g.nlevels));
    int sum = 0;
    for (int l = 0; l < g.nlevels - 1; ++l) {
      g.nper[l] = cv_round_f(nd);
      sum += g.nper[l];
      nd *= factor;
    }
    g.nper[g.nlevels - 1] = std::max(prm->nfeatures - sum, 0);
  }
  std::vector<uint16_t> tab;
  long long off = 0;
  int cand_off = 0;
  for (int l = 0; l < g.nlevels; ++l) {
    g.scale[l] = (float)pow(scale_factor, (double)l);  // getScale, orb.cpp:653-656
    const float inv = 1.0f / g.scale[l];
    g.w[l] = cv_round_f(w * inv);
    g.h[l] = cv_round_f(h * inv);
    if (g.w[l] < 1 || g.h[l] < 1) return vs_fail(ctx, VSLAM_E_LIMIT, "ORB: image too small for the pyramid");
    g.pitch[l] = (g.w[l] + 15) & ~15;
    g.off[l] = off;
    off += (long long)g.pitch[l] * g.h[l];
    off = (off + 255) & ~255LL;
    g.cand_off[l] = cand_off;
    g.cand_cap[l] = ((g.w[l] + 1) / 2) * ((g.h[l] + 1) / 2) + 16;  // strict 3x3 maxima: at most one per 2x2 block
    cand_off += g.cand_cap[l];
    g.tab_off[l] = (int)(tab.size() / 2);
    if (l > 0) {
      linear_table(g.w[l - 1], g.w[l], tab);
      linear_table(g.h[l - 1], g.h[l], tab);
    }
  }
  g.slab = off;
  g.cand_total = cand_off;

  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  DevBuf& b_pyr = ctx->orb[0];
  DevBuf& b_blur = ctx->orb[1];
  DevBuf& b_cand = ctx->orb[2];
  DevBuf& b_misc = ctx->orb[3];  // cand_cnt | hist | n_kp | status | tables
  DevBuf& b_kp = ctx->orb[4];
  DevBuf& b_desc = ctx->orb[5];
  VS_CUDA(ctx, b_pyr.reserve((size_t)g.slab * count + 64));
  VS_CUDA(ctx, b_blur.reserve((size_t)g.slab * count + 64));
  VS_CUDA(ctx, b_cand.reserve((size_t)g.cand_total * count * 4));
  const size_t o_hist = ((size_t)count * kMaxLevels * 4 + 255) & ~(size_t)255;
  const size_t o_nkp = o_hist + (size_t)count * kMaxLevels * 256 * 4;
  const size_t o_status = o_nkp + (((size_t)count * 4 + 255) & ~(size_t)255);
  VS_CUDA(ctx, b_misc.reserve(o_status + (size_t)count * 4 + 256));
  // the sampling pattern and the resize tables stay on the device between calls with the same geometry
  const size_t o_pat = 0, o_tab = 1024;
  DevBuf& b_tab = ctx->orb[8];
  int scale_bits;
  memcpy(&scale_bits, &prm->scale_factor, 4);
  const int key[4] = {w, h, g.nlevels, scale_bits};
  const bool tables_stale = memcmp(key, ctx->orb_tab_key, sizeof key) != 0 || b_tab.bytes < o_tab + tab.size() * 2;
  VS_CUDA(ctx, b_kp.reserve((size_t)count * cap * 6 * 4));
  VS_CUDA(ctx, b_desc.reserve((size_t)count * cap * 32));
  VS_CUDA(ctx, ctx->orb[7].reserve((size_t)count * cap * 8));
  uint8_t* misc = b_misc.as<uint8_t>();
  cudaStream_t st = ctx->stream;
  VS_CUDA(ctx, cudaMemsetAsync(misc, 0, o_status + (size_t)count * 4, st));  // candidate counters, histograms, counts, status
  if (tables_stale) {
    VS_CUDA(ctx, cudaStreamSynchronize(st));
    VS_CUDA(ctx, b_tab.reserve(o_tab + tab.size() * 2 + 256));
    if (!tab.empty()) VS_CUDA(ctx, cudaMemcpy(b_tab.as<uint8_t>() + o_tab, tab.data(), tab.size() * 2, cudaMemcpyHostToDevice));
  }
  const uint8_t* tabs = b_tab.as<uint8_t>();
  // umax of the radius-15 disc (orb.cpp:860-878), 4 bits per row, and the sampling pattern (behind the resize tables)
  unsigned long long umax_packed = 0;
  {
    int umax[17];
    const int half = 15;
    int v, v0;
    const int vmax = (int)floorf(half * sqrtf(2.f) / 2 + 1), vmin = (int)ceilf(half * sqrtf(2.f) / 2);
    for (v = 0; v <= vmax; ++v) umax[v] = (int)lrint(sqrt((double)half * half - v * v));
    for (v = half, v0 = 0; v >= vmin; --v) {
      while (umax[v0] == umax[v0 + 1]) ++v0;
      umax[v] = v0;
      ++v0;
    }
    for (v = 0; v <= half; ++v) umax_packed |= (unsigned long long)umax[v] << (4 * v);
  }
  static const int8_t pattern[256][4] = {
#include "rbrief_pattern.inc"
  };
  if (tables_stale) {
    VS_CUDA(ctx, cudaMemcpy(b_tab.as<uint8_t>() + o_pat, pattern, sizeof pattern, cudaMemcpyHostToDevice));
    memcpy(ctx->orb_tab_key, key, sizeof key);
  }
  uint8_t* pyr = b_pyr.as<uint8_t>();
  orb_copy_level0_kernel<<<dim3((w + 255) / 256, h, count), 256, 0, st>>>(images_dev, w, h, pitch, stride, g, pyr);
  ctx->launches += 1;
  for (int l = 1; l < g.nlevels; ++l) {
    if (g.w[l] == g.w[l - 1] && g.h[l] == g.h[l - 1])
      return vs_fail(ctx, VSLAM_E_LIMIT, "ORB: two pyramid levels of the same size");  // resize() would copy
    orb_resize_kernel<<<dim3((g.w[l] + 127) / 128, g.h[l], count), 128, 0, st>>>(
        g, l, reinterpret_cast<const ushort2*>(tabs + o_tab) + g.tab_off[l], pyr);
    ctx->launches += 1;
  }
  VS_CUDA(ctx, cudaGetLastError());
  FastParams fp;
  fp.g = g;
  fp.pyr = pyr;
  fp.threshold = prm->fast_threshold;
  fp.edge = prm->edge_threshold;
  fp.cand = b_cand.as<uint32_t>();
  fp.cand_cnt = reinterpret_cast<uint32_t*>(misc);
  fp.hist = reinterpret_cast<uint32_t*>(misc + o_hist);
  BlurParams bp;
  bp.g = g;
  bp.pyr = pyr;
  bp.blur = b_blur.as<uint8_t>();
  {
    // getGaussianKernel(7, 2, CV_32F): exp(-x^2 / (2 sigma^2)) normalised, computed in double and rounded to float
    // (smooth.dispatch.cpp getGaussianKernelBitExact; checked against cv2.getGaussianKernel bit for bit in
    // tests/test_oracle_orb.py)
    double kd[7], sum = 0;
    for (int i = 0; i < 7; ++i) { kd[i] = exp(-(double)((i - 3) * (i - 3)) / (2.0 * 2.0 * 2.0)); sum += kd[i]; }
    for (int i = 0; i < 7; ++i) bp.k7[i] = (float)(kd[i] / sum);
  }
  int tf = 0, tb = 0;
  for (int l = 0; l < g.nlevels; ++l) {
    fp.tiles_x[l] = (g.w[l] + kFastTW - 1) / kFastTW;
    fp.tiles_y[l] = (g.h[l] + kFastTH - 1) / kFastTH;
    fp.tile_first[l] = tf;
    tf += fp.tiles_x[l] * fp.tiles_y[l];
    bp.tiles_x[l] = (g.w[l] + kBlurTW - 1) / kBlurTW;
    bp.tile_first[l] = tb;
    tb += bp.tiles_x[l] * ((g.h[l] + kBlurTH - 1) / kBlurTH);
  }
  fp.tile_first[g.nlevels] = tf;
  bp.tile_first[g.nlevels] = tb;
  orb_fast_kernel<<<dim3(tf, count), 256, 0, st>>>(fp);
  SelectParams sp;
  sp.g = g;
  sp.cand = fp.cand;
  sp.cand_cnt = fp.cand_cnt;
  sp.hist = fp.hist;
  sp.kp = b_kp.as<float>();
  sp.n_kp = reinterpret_cast<int32_t*>(misc + o_nkp);
  sp.status = reinterpret_cast<int32_t*>(misc + o_status);
  sp.cap = cap;
  orb_select_kernel<<<dim3(g.nlevels, count), 512, 0, st>>>(sp);
  KpParams kp;
  kp.g = g;
  kp.pyr = pyr;
  kp.kp = sp.kp;
  kp.n_kp = sp.n_kp;
  kp.desc = b_desc.as<uint8_t>();
  kp.rot = ctx->orb[7].as<float2>();
  kp.pattern = reinterpret_cast<const char4*>(tabs + o_pat);
  kp.umax = umax_packed;
  kp.cap = cap;
  orb_angle_kernel<<<dim3((cap + 7) / 8, count), 256, 0, st>>>(kp);
  orb_blur_kernel<<<dim3(tb, count), 256, 0, st>>>(bp);
  kp.pyr = bp.blur;
  orb_describe_kernel<<<dim3((cap + 7) / 8, count), 256, 0, st>>>(kp);
  ctx->launches += 5;
  VS_CUDA(ctx, cudaGetLastError());
  out->kp = sp.kp;
  out->desc = kp.desc;
  out->n_kp = sp.n_kp;
  out->status = sp.status;
  return VSLAM_OK;
}

int orb_check_status(vslam_ctx* ctx, const int32_t* n_kp, const int32_t* status, int count, int cap) {
  for (int i = 0; i < count; ++i) {
    if (status[i] & 1) return vs_fail(ctx, VSLAM_E_LIMIT, "ORB: candidate list overflow");
    if (status[i] & 2) return vs_fail(ctx, VSLAM_E_LIMIT, "ORB: more than 8192 keypoints kept on one pyramid level");
  }
  for (int i = 0; i < count; ++i)
    if (n_kp[i] > cap) return vs_fail(ctx, VSLAM_E_CAPACITY, "ORB: keypoint capacity too small (ties at the cut-off score all stay)");
  return VSLAM_OK;
}

constexpr int kOrbChunk = 128;

extern "C" int vslam_orb_detect_compute_batch(vslam_ctx* ctx, const uint8_t* images, int count, int w, int h, int pitch,
                                              size_t image_stride, const vslam_orb_params* prm, float* kp, uint8_t* desc,
                                              int32_t* n_kp, int cap) {
  if (!ctx) return VSLAM_E_ARG;
  if (!images || !prm || count < 0 || w <= 0 || h <= 0 || pitch < w || cap <= 0 || !kp || !desc || !n_kp ||
      image_stride < (size_t)pitch * (h - 1) + w)
    return vs_fail(ctx, VSLAM_E_ARG, "bad ORB arguments");
  if (count == 0) return VSLAM_OK;
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  // Chunks of kOrbChunk images: the upload of chunk c + 1 (copy stream 0) and the download of chunk c - 1's keypoints
  // and descriptors (copy stream 1; the two output buffers alternate) overlap the kernels of chunk c.
  const size_t tail = (size_t)pitch * (h - 1) + w;  // bytes of the last image of a run
  VS_CUDA(ctx, ctx->orb[6].reserve(image_stride * (size_t)(count - 1) + tail + 64));
  VS_CUDA(ctx, ctx->orb[9].reserve((size_t)count * 8 + 256));  // keypoint counts | status of the whole batch
  uint8_t* staged = ctx->orb[6].as<uint8_t>();
  int32_t* counts_dev = ctx->orb[9].as<int32_t>();
  int32_t* status_dev = counts_dev + count;
  const int n_chunks = (count + kOrbChunk - 1) / kOrbChunk;
  std::vector<cudaEvent_t> ev((size_t)n_chunks * 3, nullptr);  // uploaded | computed | downloaded
  int rc = VSLAM_OK;
  cudaError_t e = cudaStreamSynchronize(ctx->stream);  // the staging buffer may still be read by an earlier call
  for (int c = 0; c < n_chunks && rc == VSLAM_OK && e == cudaSuccess; ++c) {
    const int c0 = c * kOrbChunk, cn = std::min(kOrbChunk, count - c0);
    cudaEvent_t* evc = &ev[(size_t)c * 3];
    for (int i = 0; i < 3 && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&evc[i], cudaEventDisableTiming);
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(staged + image_stride * c0, images + image_stride * c0, image_stride * (size_t)(cn - 1) + tail,
                          cudaMemcpyHostToDevice, ctx->copy_stream[0]);
    if (e == cudaSuccess) e = cudaEventRecord(evc[0], ctx->copy_stream[0]);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->stream, evc[0], 0);
    if (c >= 2 && e == cudaSuccess) e = cudaStreamWaitEvent(ctx->stream, ev[(size_t)(c - 2) * 3 + 2], 0);  // its buffers are free again
    if (e != cudaSuccess) break;
    std::swap(ctx->orb[4], ctx->orb[10]);
    std::swap(ctx->orb[5], ctx->orb[11]);
    OrbDeviceOut out;
    rc = launch_orb(ctx, staged + image_stride * c0, cn, w, h, pitch, image_stride, prm, cap, &out);
    if (rc) break;
    e = cudaMemcpyAsync(counts_dev + c0, out.n_kp, (size_t)cn * 4, cudaMemcpyDeviceToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(status_dev + c0, out.status, (size_t)cn * 4, cudaMemcpyDeviceToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaEventRecord(evc[1], ctx->stream);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->copy_stream[1], evc[1], 0);
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(kp + (size_t)c0 * cap * 6, out.kp, (size_t)cn * cap * 24, cudaMemcpyDeviceToHost, ctx->copy_stream[1]);
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(desc + (size_t)c0 * cap * 32, out.desc, (size_t)cn * cap * 32, cudaMemcpyDeviceToHost, ctx->copy_stream[1]);
    if (e == cudaSuccess) e = cudaEventRecord(evc[2], ctx->copy_stream[1]);
  }
  const cudaError_t s0 = cudaStreamSynchronize(ctx->copy_stream[0]);
  const cudaError_t s1 = cudaStreamSynchronize(ctx->stream);
  const cudaError_t s2 = cudaStreamSynchronize(ctx->copy_stream[1]);
  for (cudaEvent_t x : ev)
    if (x) cudaEventDestroy(x);
  if (rc) return rc;
  VS_CUDA(ctx, e);
  VS_CUDA(ctx, s0);
  VS_CUDA(ctx, s1);
  VS_CUDA(ctx, s2);
  std::vector<int32_t> status((size_t)count);
  VS_CUDA(ctx, cudaMemcpy(n_kp, counts_dev, (size_t)count * 4, cudaMemcpyDeviceToHost));
  VS_CUDA(ctx, cudaMemcpy(status.data(), status_dev, (size_t)count * 4, cudaMemcpyDeviceToHost));
  return orb_check_status(ctx, n_kp, status.data(), count, cap);
}
