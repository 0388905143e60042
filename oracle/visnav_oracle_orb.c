/* visnav_oracle_orb.c -- CPU restatement of cv::ORB::detectAndCompute as visnav::compute_bow_vector configures it
 * (keypoints.h:249: cv::ORB::create(num_features, 1.2, 8, 19, 0, 2, cv::ORB::FAST_SCORE), patchSize 31,
 * fastThreshold 20), SURVEY §8f-4, ORB half.  TEST INFRASTRUCTURE ONLY: linked into oracle/liboracle.so; only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may call it.  Parity status: PINNED against
 * cv2.ORB_create(...).detectAndCompute (cv2 4.13; the reference vendors OpenCV 4.7.0-dev, same algorithm) by
 * tests/test_oracle_orb.py and the committed fixture tests/golden/orb.npz.
 *
 * Follows (thirdparty/opencv/modules):
 *   pyramid layout and levels    features2d/src/orb.cpp:1012-1130 (getScale :653-656)
 *   resize INTER_LINEAR_EXACT    imgproc/src/resize.cpp:733-771 (coefficients), :344-398 (rows), :691-731 (columns);
 *                                imgproc/src/fixedpoint.inl.hpp:326-374 (ufixedpoint16)
 *   FAST-9/16 + score + NMS      features2d/src/fast.cpp:56-290, fast_score.cpp:50-81, :119-215
 *   border filter, retainBest    features2d/src/keypoint.cpp:69-117; per-level budget orb.cpp:843-856
 *   orientation                  orb.cpp:181-216 (umax :860-878); core/src/mathfuncs_core.simd.hpp:34-71 (fastAtan2)
 *   blur                         imgproc/src/smooth.dispatch.cpp:611-757 -> sepFilter2D, float kernels
 *                                (filter.dispatch.cpp:290-375: the 8-bit integer kernel is rejected for this
 *                                kernel); filter.simd.hpp RowFilter / SymmColumnFilter, Cast<float, uchar>
 *   descriptors                  orb.cpp:219-268 with bit_pattern_31_ (:380-638 -- the same 256 point pairs as the
 *                                reference's pattern_31_* tables, keypoints.h:55-131)
 * Keypoint ORDER inside a pyramid level is std::nth_element's + std::partition's in the reference (keypoint.cpp:80-86):
 * unspecified by the C++ standard.  The SET is determined (every keypoint whose response is >= the n-th largest);
 * this restatement lists a level's keypoints in raster order.  BowVectors do not depend on the order.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static const int8_t kPattern[256][4] = {
#include "../visual-slam_b200/csrc/rbrief_pattern.inc"
};

static int cv_round_f(float v) { return (int)lrintf(v); }   /* cvRound: round half to even */
static int cv_round_d(double v) { return (int)lrint(v); }

/* ---- resize INTER_LINEAR_EXACT, 8UC1 --------------------------------------------------------------------------- */
typedef struct { int ofs; uint16_t c0, c1; } lin_coef;

/* interpolationLinear::getCoeffs (resize.cpp:739-765).  Returns min/max through the pointers. */
static void linear_coeffs(int srcsize, int dstsize, lin_coef* co, int* pmin, int* pmax) {
  const double inv_scale = (double)dstsize / srcsize; /* resize(): inv_scale_x = dsize.width / ssize.width */
  const double scale = 1.0 / inv_scale;
  int minofst = 0, maxofst = dstsize;
  for (int val = 0; val < dstsize; val++) {
    const double fval = scale * ((double)val + 0.5) - 0.5;
    const int ival = (int)floor(fval);
    co[val].ofs = 0; co[val].c0 = 0; co[val].c1 = 0;
    if (ival >= 0 && srcsize > 1) {
      if (ival < srcsize - 1) {
        co[val].ofs = ival;
        const double frac = fval - (double)ival;
        co[val].c1 = frac < 0 ? 0 : (uint16_t)cv_round_d(frac * 256.0);
        co[val].c0 = (uint16_t)(256 > co[val].c1 ? 256 - co[val].c1 : 0);
      } else {
        co[val].ofs = srcsize - 1;
        if (val < maxofst) maxofst = val;
      }
    } else if (val + 1 > minofst) {
      minofst = val + 1;
    }
  }
  *pmin = minofst;
  *pmax = maxofst;
}

static void hline(const uint8_t* src, const lin_coef* cx, int min_x, int max_x, int dw, uint16_t* dst) {
  int i = 0;
  for (; i < min_x; i++) dst[i] = (uint16_t)(src[0] << 8);
  for (; i < max_x; i++) {
    const uint8_t* px = src + cx[i].ofs;
    dst[i] = (uint16_t)(cx[i].c0 * px[0] + cx[i].c1 * px[1]); /* c0 + c1 = 256: no saturation */
  }
  const uint16_t last = (uint16_t)(src[cx[dw - 1].ofs] << 8);
  for (; i < dw; i++) dst[i] = last;
}

void vso_resize_linear_exact(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh,
                             int dstride) {
  lin_coef* cx = (lin_coef*)malloc(sizeof(lin_coef) * (size_t)dw);
  lin_coef* cy = (lin_coef*)malloc(sizeof(lin_coef) * (size_t)dh);
  uint16_t* r0 = (uint16_t*)malloc(2 * (size_t)dw);
  uint16_t* r1 = (uint16_t*)malloc(2 * (size_t)dw);
  int min_x, max_x, min_y, max_y;
  linear_coeffs(sw, dw, cx, &min_x, &max_x);
  linear_coeffs(sh, dh, cy, &min_y, &max_y);
  for (int dy = 0; dy < dh; dy++) {
    uint8_t* d = dst + (size_t)dy * dstride;
    if (dy < min_y || dy >= max_y) { /* vlineSet of the first / last source row */
      hline(src + (size_t)(dy < min_y ? 0 : sh - 1) * sstride, cx, min_x, max_x, dw, r0);
      for (int i = 0; i < dw; i++) d[i] = (uint8_t)((r0[i] + 128) >> 8);
      continue;
    }
    const int iy = cy[dy].ofs;
    hline(src + (size_t)iy * sstride, cx, min_x, max_x, dw, r0);
    hline(src + (size_t)(iy + 1 < sh ? iy + 1 : sh - 1) * sstride, cx, min_x, max_x, dw, r1);
    for (int i = 0; i < dw; i++) {
      const uint32_t v = (uint32_t)r0[i] * cy[dy].c0 + (uint32_t)r1[i] * cy[dy].c1;
      const uint32_t q = (v + 32768u) >> 16;
      d[i] = (uint8_t)(q > 255 ? 255 : q);
    }
  }
  free(cx); free(cy); free(r0); free(r1);
}

/* ---- FAST-9/16 --------------------------------------------------------------------------------------------------- */
static const int kRing[16][2] = {{0, 3}, {1, 3}, {2, 2}, {3, 1}, {3, 0}, {3, -1}, {2, -2}, {1, -3},
                                 {0, -3}, {-1, -3}, {-2, -2}, {-3, -1}, {-3, 0}, {-3, 1}, {-2, 2}, {-1, 3}};

/* cornerScore<16> (fast_score.cpp:119-215): the largest threshold at which the pixel is still a corner =
 * max over the 16 arcs of 9 contiguous ring pixels of min(v - ring) resp. min(ring - v), minus 1. */
static int fast_score(const uint8_t* p, int stride) {
  int d[25];
  const int v = p[0];
  for (int k = 0; k < 25; k++) d[k] = v - p[kRing[k & 15][0] + kRing[k & 15][1] * stride];
  int best = -1000;
  for (int s = 0; s < 16; s++) {
    int mn = d[s], mx = d[s];
    for (int i = 1; i < 9; i++) {
      if (d[s + i] < mn) mn = d[s + i];
      if (d[s + i] > mx) mx = d[s + i];
    }
    if (mn > best) best = mn;
    if (-mx > best) best = -mx;
  }
  return best - 1;
}

/* Score map of FAST_t<16> with threshold t: score (>= t) at corners, 0 elsewhere; rows / columns 0..2 and the last
 * three stay 0 (fast.cpp:96-100, :206). */
void vso_fast_score_map(const uint8_t* img, int w, int h, int stride, int threshold, uint8_t* score) {
  memset(score, 0, (size_t)w * h);
  for (int y = 3; y < h - 3; y++)
    for (int x = 3; x < w - 3; x++) {
      const int s = fast_score(img + (size_t)y * stride + x, stride);
      if (s >= threshold) score[(size_t)y * w + x] = (uint8_t)s;
    }
}

/* ---- orientation ------------------------------------------------------------------------------------------------- */
float vso_fast_atan2(float y, float x) { /* mathfuncs_core.simd.hpp:50-71 */
  const float p1 = 0.9997878412794807f * (float)(180 / 3.1415926535897932384626433832795);
  const float p3 = -0.3258083974640975f * (float)(180 / 3.1415926535897932384626433832795);
  const float p5 = 0.1555786518463281f * (float)(180 / 3.1415926535897932384626433832795);
  const float p7 = -0.04432655554792128f * (float)(180 / 3.1415926535897932384626433832795);
  const float ax = fabsf(x), ay = fabsf(y);
  float a, c, c2;
  if (ax >= ay) {
    c = ay / (ax + (float)2.2204460492503131e-16);
    c2 = c * c;
    a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  } else {
    c = ax / (ay + (float)2.2204460492503131e-16);
    c2 = c * c;
    a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  }
  if (x < 0) a = 180.f - a;
  if (y < 0) a = 360.f - a;
  return a;
}

static void make_umax(int half, int* umax) { /* orb.cpp:860-878 */
  int v, v0;
  const int vmax = (int)floor(half * sqrt(2.f) / 2 + 1);
  const int vmin = (int)ceil(half * sqrt(2.f) / 2);
  for (v = 0; v <= vmax; ++v) umax[v] = cv_round_d(sqrt((double)half * half - v * v));
  for (v = half, v0 = 0; v >= vmin; --v) {
    while (umax[v0] == umax[v0 + 1]) ++v0;
    umax[v] = v0;
    ++v0;
  }
}

static float ic_angle(const uint8_t* center, int step, const int* umax, int half) { /* orb.cpp:181-216 */
  int m_01 = 0, m_10 = 0;
  for (int u = -half; u <= half; ++u) m_10 += u * center[u];
  for (int v = 1; v <= half; ++v) {
    int v_sum = 0;
    const int d = umax[v];
    for (int u = -d; u <= d; ++u) {
      const int val_plus = center[u + v * step], val_minus = center[u - v * step];
      v_sum += (val_plus - val_minus);
      m_10 += u * (val_plus + val_minus);
    }
    m_01 += v * v_sum;
  }
  return vso_fast_atan2((float)m_01, (float)m_10);
}

/* ---- blur: GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) on 8U through the float separable filter ------------------
 * kernel: getGaussianKernel(7, 2, CV_32F) (bit-exact softdouble computation, rounded to float) -- the caller passes
 * it.  fused != 0: multiply-adds contracted as in OpenCV's AVX2 dispatch of filter.simd.hpp. */
static int reflect101(int p, int len) {
  if (len == 1) return 0;
  while (p < 0 || p >= len) p = p < 0 ? -p : 2 * len - 2 - p;
  return p;
}

void vso_gauss7(const uint8_t* src, int w, int h, int stride, const float* k7, int fused, uint8_t* dst, int dstride) {
  float* rows = (float*)malloc(sizeof(float) * (size_t)w * h);
  for (int y = 0; y < h; y++)
    for (int x = 0; x < w; x++) { /* RowFilter: s = k0 * p0; s += k[i] * p[i], left to right */
      float s = k7[0] * (float)src[(size_t)y * stride + reflect101(x - 3, w)];
      for (int i = 1; i < 7; i++) {
        const float p = (float)src[(size_t)y * stride + reflect101(x - 3 + i, w)];
        s = fused ? fmaf(k7[i], p, s) : s + k7[i] * p;
      }
      rows[(size_t)y * w + x] = s;
    }
  for (int y = 0; y < h; y++)
    for (int x = 0; x < w; x++) { /* SymmColumnFilter: s = k3 * c + delta(0); s += k[3+j] * (r[+j] + r[-j]) */
      const float c = rows[(size_t)y * w + x];
      float s = fused ? fmaf(k7[3], c, 0.f) : k7[3] * c + 0.f;
      for (int j = 1; j <= 3; j++) {
        const float t = rows[(size_t)reflect101(y + j, h) * w + x] + rows[(size_t)reflect101(y - j, h) * w + x];
        s = fused ? fmaf(k7[3 + j], t, s) : s + k7[3 + j] * t;
      }
      int v = cv_round_f(s);
      dst[(size_t)y * dstride + x] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
    }
  free(rows);
}

/* ---- cv::ORB::detectAndCompute (orb.cpp:1012-1258, computeKeyPoints :826-998) ---------------------------------------
 * Fixed by the reference's call: firstLevel 0, WTA_K 2, FAST_SCORE, patchSize 31.  Outputs, level by level and in
 * raster order inside a level: kp [n][6] = x, y (level-0 coordinates: level coordinate * scale, float), size, angle
 * (degrees), response (FAST score), octave; desc [n][32].  level_dims (nullable) [nlevels][2] = width, height.
 * Returns n (may exceed cap: then only cap entries were written). */
int vso_orb_detect_compute(const uint8_t* img, int w, int h, int stride, int nfeatures, int nlevels,
                           double scale_factor, int edge_threshold, int fast_threshold, const float* k7, int fused,
                           float* kp, uint8_t* desc, int cap, int32_t* level_dims) {
  enum { MAXL = 32 };
  if (nlevels > MAXL) nlevels = MAXL;
  const int patch = 31, half = patch / 2;
  int umax[32];
  make_umax(half, umax);
  /* per-level budget, orb.cpp:843-856 */
  int nper[MAXL];
  {
    const float factor = (float)(1.0 / scale_factor);
    float nd = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int l = 0; l < nlevels - 1; l++) {
      nper[l] = cv_round_f(nd);
      sum += nper[l];
      nd *= factor;
    }
    nper[nlevels - 1] = nfeatures - sum > 0 ? nfeatures - sum : 0;
  }
  uint8_t* lev[MAXL];
  int lw[MAXL], lh[MAXL];
  float lscale[MAXL];
  for (int l = 0; l < nlevels; l++) {
    lscale[l] = (float)pow(scale_factor, (double)l); /* getScale */
    const float inv = 1.0f / lscale[l];
    lw[l] = cv_round_f(w * inv);
    lh[l] = cv_round_f(h * inv);
    if (level_dims) { level_dims[2 * l] = lw[l]; level_dims[2 * l + 1] = lh[l]; }
    lev[l] = (uint8_t*)malloc((size_t)lw[l] * lh[l] + 1);
    if (l == 0) {
      for (int y = 0; y < h; y++) memcpy(lev[0] + (size_t)y * w, img + (size_t)y * stride, (size_t)w);
    } else if (lw[l] == lw[l - 1] && lh[l] == lh[l - 1]) {
      memcpy(lev[l], lev[l - 1], (size_t)lw[l] * lh[l]);
    } else {
      vso_resize_linear_exact(lev[l - 1], lw[l - 1], lh[l - 1], lw[l - 1], lev[l], lw[l], lh[l], lw[l]);
    }
  }
  int n = 0;
  int first_of_level[MAXL + 1];
  for (int l = 0; l < nlevels; l++) {
    first_of_level[l] = n;
    const int W = lw[l], H = lh[l];
    uint8_t* score = (uint8_t*)malloc((size_t)W * H + 1);
    vso_fast_score_map(lev[l], W, H, W, fast_threshold, score);
    /* NMS (fast.cpp:279-285) + runByImageBorder (keypoint.cpp:105-117) */
    int hist[256];
    memset(hist, 0, sizeof hist);
    uint8_t* keep = (uint8_t*)calloc((size_t)W * H + 1, 1);
    int count = 0;
    if (!(H <= edge_threshold * 2 || W <= edge_threshold * 2))
      for (int y = 3; y < H - 3; y++)
        for (int x = 3; x < W - 3; x++) {
          const uint8_t* s = score + (size_t)y * W + x;
          const int v = s[0];
          if (v == 0) continue;
          if (!(v > s[1] && v > s[-1] && v > s[-W - 1] && v > s[-W] && v > s[-W + 1] && v > s[W - 1] && v > s[W] &&
                v > s[W + 1]))
            continue;
          if (!(x >= edge_threshold && x < W - edge_threshold && y >= edge_threshold && y < H - edge_threshold))
            continue;
          keep[(size_t)y * W + x] = 1;
          hist[v]++;
          count++;
        }
    /* retainBest(keypoints, nper[l]) (keypoint.cpp:69-90): all responses >= the n-th largest */
    int thr = 0;
    if (count > nper[l]) {
      if (nper[l] == 0) {
        thr = 256;
      } else {
        int acc = 0;
        for (thr = 255; thr > 0; thr--) {
          acc += hist[thr];
          if (acc >= nper[l]) break;
        }
      }
    }
    for (int y = 0; y < H; y++)
      for (int x = 0; x < W; x++) {
        if (!keep[(size_t)y * W + x] || score[(size_t)y * W + x] < thr) continue;
        if (n < cap) {
          float* k = kp + 6 * (size_t)n;
          k[0] = (float)x * lscale[l];
          k[1] = (float)y * lscale[l];
          k[2] = patch * lscale[l];
          k[3] = ic_angle(lev[l] + (size_t)y * W + x, W, umax, half);
          k[4] = (float)score[(size_t)y * W + x];
          k[5] = (float)l;
        }
        n++;
      }
    free(score);
    free(keep);
  }
  first_of_level[nlevels] = n;
  /* descriptors on the blurred levels (orb.cpp:1216-1221, :219-268) */
  for (int l = 0; l < nlevels; l++) {
    const int W = lw[l], H = lh[l];
    const int a0 = first_of_level[l], a1 = first_of_level[l + 1] < cap ? first_of_level[l + 1] : cap;
    if (a0 >= a1) continue;
    uint8_t* blur = (uint8_t*)malloc((size_t)W * H + 1);
    vso_gauss7(lev[l], W, H, W, k7, fused, blur, W);
    const float scale = 1.f / lscale[l];
    for (int j = a0; j < a1; j++) {
      const float* k = kp + 6 * (size_t)j;
      float angle = k[3];
      angle *= (float)(3.1415926535897932384626433832795 / 180.f);
      const float a = cosf(angle), b = sinf(angle);
      const uint8_t* center = blur + (size_t)cv_round_f(k[1] * scale) * W + cv_round_f(k[0] * scale);
      uint8_t* d = desc + 32 * (size_t)j;
      for (int i = 0; i < 32; i++) {
        int val = 0;
        for (int bit = 0; bit < 8; bit++) {
          const int8_t* p = kPattern[8 * i + bit];
          const float xa = p[0] * a - p[1] * b, ya = p[0] * b + p[1] * a;
          const float xb = p[2] * a - p[3] * b, yb = p[2] * b + p[3] * a;
          const int t0 = center[cv_round_f(ya) * W + cv_round_f(xa)];
          const int t1 = center[cv_round_f(yb) * W + cv_round_f(xb)];
          val |= (t0 < t1) << bit;
        }
        d[i] = (uint8_t)val;
      }
    }
    free(blur);
  }
  for (int l = 0; l < nlevels; l++) free(lev[l]);
  return n;
}
