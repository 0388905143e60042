/*
 * visnav_oracle.c -- CPU restatement of the visnav front-end hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (visual-slam_b200/,
 * include/) may include, link or call this file.  It is used by tests/, by
 * __graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference
 * legs, as the checker and the timed CPU baseline.
 *
 * Parity status: PINNED.
 *   - vso_response / vso_gftt are checked bit-for-bit / corner-for-corner
 *     against OpenCV (cv2 4.13.0, cv::setUseOptimized(false)) through the
 *     committed fixtures tests/golden/gftt_*.npz (tools/make_golden.py) and
 *     live against cv2 when it is importable.
 *   - vso_angles / vso_describe / vso_match are checked against the
 *     reference's own functions compiled from /root/reference into
 *     oracle/_ref/libvisnav_ref.so (oracle/ref_shim.cpp, oracle/Makefile).
 *
 * Every function cites the reference lines it restates.  "OCV" is the
 * reference's vendored OpenCV, thirdparty/opencv/modules/imgproc/src/.
 *
 * Build: gcc -O2 -ffp-contract=off (no fast-math): every float expression
 * below is meant to be evaluated one IEEE operation at a time.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

#ifdef __cplusplus
extern "C" {
#endif

static const signed char kPattern[256][4] = {
#include "../visual-slam_b200/csrc/rbrief_pattern.inc"
};

/* BORDER_REFLECT_101: -1 -> 1, n -> n-2 (OCV borderInterpolate). */
static inline int r101(int i, int n) {
  if (n == 1) return 0;
  while (i < 0 || i >= n) {
    if (i < 0) i = -i;
    else i = 2 * n - 2 - i;
  }
  return i;
}

/*
 * Min-eigenvalue corner response, blockSize=3, gradientSize=3, 8-bit input.
 * Restates OCV cornerMinEigenVal -> cornerEigenValsVecs (corner.cpp:237-322,
 * 550-572) with the scalar operation order:
 *   Sobel with the scale folded into the smoothing kernel (deriv.cpp:414-465),
 *   row/column passes of the separable filter engine (filter.simd.hpp:2477-2487
 *   scalar row tail, :2906-2937 SymmColumnSmallFilter scalar forms),
 *   cov products (corner.cpp:302-310), 3x3 un-normalised box sum accumulated in
 *   double (box_filter.simd.hpp:84-90, 225-262), calcMinEigenVal scalar tail
 *   (corner.cpp:93-99).
 * lam is h*w floats, dense.
 */
/* cv::goodFeaturesToTrack's useHarrisDetector / k for vso_response / vso_gftt / vso_detect (test knob; the
 * reference passes false, keypoints.h:138-139).  cornerHarris shares everything with cornerMinEigenVal up to the
 * box-filtered covariance (corner.cpp:237-322); calcHarris (corner.cpp:104-153) then runs its universal-intrinsics
 * loop (float k, float operations) over all but the last (w*h) % 4 pixels of the continuous image and the scalar
 * tail (double k) over the rest. */
static int g_harris_on = 0;
static double g_harris_k = 0.04;
void vso_set_harris(int on, double k) { g_harris_on = on; g_harris_k = k; }
/* 0 (default): OCV's running column sum; 1: direct three-row sum (see the column pass in vso_response). */
static int g_box_direct = 0;
void vso_set_box_direct(int on) { g_box_direct = on; }
/* Probe for the row-banded evaluation: with band > 0, vso_response counts the (column, channel) pairs at rows y = band,
 * 2 band, ... whose running SUM differs from the two-term sum R(y-1) + R(y) a band that starts there would seed it with
 * (a rounding residue of an inexact addition further up the column).  Zero for an image = a banded evaluation that carries
 * the running sum inside each band reproduces OCV bit for bit. */
static int g_residue_band = 0;
static long long g_residue_count = 0;
void vso_residue_probe(int band) { g_residue_band = band; g_residue_count = 0; }
long long vso_residue_count(void) { return g_residue_count; }

int vso_response(const uint8_t* img, int w, int h, int pitch, float* lam) {
  if (!img || !lam || w <= 0 || h <= 0 || pitch < w) return -1;
  const double s = 1.0 / (4.0 * 3.0 * 255.0); /* corner.cpp:247-253 */
  const float f1 = (float)s;
  const float f0 = (float)(2.0 * s);
  size_t npx = (size_t)w * h;
  float* rdx = (float*)malloc(npx * sizeof(float)); /* row pass of Dx: [-1 0 1] */
  float* rdy = (float*)malloc(npx * sizeof(float)); /* row pass of Dy: [f1 f0 f1] */
  float* cov = (float*)malloc(npx * 3 * sizeof(float));
  double* rs = (double*)malloc(npx * 3 * sizeof(double)); /* box row sums */
  if (!rdx || !rdy || !cov || !rs) { free(rdx); free(rdy); free(cov); free(rs); return -2; }

  for (int y = 0; y < h; y++) {
    const uint8_t* row = img + (size_t)y * pitch;
    for (int x = 0; x < w; x++) {
      float l = (float)row[r101(x - 1, w)];
      float c = (float)row[x];
      float r = (float)row[r101(x + 1, w)];
      rdx[(size_t)y * w + x] = r - l;
      float t = l * f1;
      float u = c * f0;
      t = t + u;
      u = r * f1;
      rdy[(size_t)y * w + x] = t + u;
    }
  }
  for (int y = 0; y < h; y++) {
    int ym = r101(y - 1, h), yp = r101(y + 1, h);
    for (int x = 0; x < w; x++) {
      float a = rdx[(size_t)ym * w + x] + rdx[(size_t)yp * w + x];
      a = a * f1;
      float b = rdx[(size_t)y * w + x] * f0;
      float dx = a + b;
      float dy = rdy[(size_t)yp * w + x] - rdy[(size_t)ym * w + x];
      float* c3 = cov + ((size_t)y * w + x) * 3;
      c3[0] = dx * dx;
      c3[1] = dx * dy;
      c3[2] = dy * dy;
    }
  }
  for (int y = 0; y < h; y++)
    for (int x = 0; x < w; x++) {
      int xm = r101(x - 1, w), xp = r101(x + 1, w);
      for (int k = 0; k < 3; k++) {
        double v = (double)cov[((size_t)y * w + xm) * 3 + k];
        v = v + (double)cov[((size_t)y * w + x) * 3 + k];
        v = v + (double)cov[((size_t)y * w + xp) * 3 + k];
        rs[((size_t)y * w + x) * 3 + k] = v;
      }
    }
  /* Column pass of the box filter.  OCV's ColumnSum<double,float> (box_filter.simd.hpp:189-262) keeps ONE running
   * sum per column and channel: SUM = 0 + R(-1) + R(0) before the first row (:206-216, rows reflected), then per
   * output row y:  s0 = SUM + R(y+1);  D = (float)s0;  SUM = s0 - R(y-1)  (:248-262).  Where one of these double
   * additions is inexact the residue travels down the column, so the result can differ in the last bit from the
   * direct three-row sum (R(y-1) + R(y)) + R(y+1) -- which vso_set_box_direct(1) selects instead (what a row-banded
   * evaluation computes; tests bound the difference). */
  double* SUM = (double*)malloc((size_t)w * 3 * sizeof(double));
  if (!SUM) { free(rdx); free(rdy); free(cov); free(rs); return -2; }
  {
    const double* r_m1 = rs + (size_t)r101(-1, h) * w * 3;
    for (size_t i = 0; i < (size_t)w * 3; i++) {
      double v = 0.0;
      v = v + r_m1[i];
      v = v + rs[i];
      SUM[i] = v;
    }
  }
  for (int y = 0; y < h; y++) {
    int ym = r101(y - 1, h), yp = r101(y + 1, h);
    for (int x = 0; x < w; x++) {
      float abc[3];
      for (int k = 0; k < 3; k++) {
        if (g_box_direct) {
          double v = rs[((size_t)ym * w + x) * 3 + k];
          v = v + rs[((size_t)y * w + x) * 3 + k];
          v = v + rs[((size_t)yp * w + x) * 3 + k];
          abc[k] = (float)v;
        } else {
          if (g_residue_band > 0 && y > 0 && y % g_residue_band == 0) {
            double two = rs[((size_t)ym * w + x) * 3 + k];
            two = two + rs[((size_t)y * w + x) * 3 + k];
            if (two != SUM[(size_t)x * 3 + k]) g_residue_count++;
          }
          double s0 = SUM[(size_t)x * 3 + k] + rs[((size_t)yp * w + x) * 3 + k];
          abc[k] = (float)s0;
          SUM[(size_t)x * 3 + k] = s0 - rs[((size_t)ym * w + x) * 3 + k];
        }
      }
      if (g_harris_on) {
        const size_t i = (size_t)y * w + x;
        float ac_bb = abc[0] * abc[2];
        float bb2 = abc[1] * abc[1];
        ac_bb = ac_bb - bb2;
        float ac = abc[0] + abc[2];
        if (i < (npx & ~(size_t)3)) { /* corner.cpp:129-141 */
          float t = (float)g_harris_k * ac;
          t = t * ac;
          lam[i] = ac_bb - t;
        } else { /* corner.cpp:144-150 */
          double t = g_harris_k * (double)ac;
          t = t * (double)ac;
          lam[i] = (float)((double)ac_bb - t);
        }
        continue;
      }
      float a = abc[0] * 0.5f;
      float b = abc[1];
      float c = abc[2] * 0.5f;
      float t = a - c;
      float q = t * t;
      float bb = b * b;
      q = q + bb;
      q = sqrtf(q);
      float p = a + c;
      lam[(size_t)y * w + x] = p - q;
    }
  }
  free(rdx); free(rdy); free(cov); free(rs); free(SUM);
  return 0;
}

typedef struct { float v; int idx; } cand_t;

/* greaterThanPtr, featureselect.cpp:55-60: value desc, then address desc. */
static int cand_cmp(const void* pa, const void* pb) {
  const cand_t* a = (const cand_t*)pa;
  const cand_t* b = (const cand_t*)pb;
  if (a->v > b->v) return -1;
  if (a->v < b->v) return 1;
  return (a->idx > b->idx) ? -1 : (a->idx < b->idx) ? 1 : 0;
}

/*
 * cv::goodFeaturesToTrack CPU path without mask, Harris off
 * (featureselect.cpp:382-548): global max (:414-415), THRESH_TOZERO at
 * (float)(maxVal*quality) (:416, thresh.cpp:71-74), 3x3 dilate with -FLT_MAX
 * outside (:417), interior local maxima (:424-436), sort (:449), greedy
 * minimum-distance selection on a cell grid (:451-525), cap (:521).
 * Output xy: (x, y) floats in selection order.  Returns the count or <0.
 */
int vso_gftt(const uint8_t* img, int w, int h, int pitch, int max_corners,
             double quality, double min_distance, float* xy, int cap) {
  if (!img || w <= 0 || h <= 0 || !(quality > 0) || min_distance < 0 || max_corners < 0) return -1;
  size_t npx = (size_t)w * h;
  float* eig = (float*)malloc(npx * sizeof(float));
  if (!eig) return -2;
  int rc = vso_response(img, w, h, pitch, eig);
  if (rc) { free(eig); return rc; }
  double maxVal = 0; /* minMaxLoc over the whole image */
  {
    float m = eig[0];
    for (size_t i = 1; i < npx; i++) if (eig[i] > m) m = eig[i];
    maxVal = m;
  }
  const float thr = (float)(maxVal * quality);
  for (size_t i = 0; i < npx; i++) eig[i] = eig[i] > thr ? eig[i] : 0.f;

  cand_t* cands = (cand_t*)malloc(npx * sizeof(cand_t));
  if (!cands) { free(eig); return -2; }
  size_t total = 0;
  for (int y = 1; y < h - 1; y++)
    for (int x = 1; x < w - 1; x++) {
      float v = eig[(size_t)y * w + x];
      if (v == 0) continue;
      float m = -FLT_MAX;
      for (int dy = -1; dy <= 1; dy++)
        for (int dx = -1; dx <= 1; dx++) {
          float n = eig[(size_t)(y + dy) * w + (x + dx)];
          if (n > m) m = n;
        }
      if (v == m) { cands[total].v = v; cands[total].idx = y * w + x; total++; }
    }
  free(eig);
  int ncorners = 0;
  if (total == 0) { free(cands); return 0; }
  qsort(cands, total, sizeof(cand_t), cand_cmp);

  if (min_distance >= 1) {
    const int cell = (int)lrint(min_distance); /* cvRound */
    const int gw = (w + cell - 1) / cell, gh = (h + cell - 1) / cell;
    /* per-cell singly linked lists of accepted corners */
    int* head = (int*)malloc((size_t)gw * gh * sizeof(int));
    int* next = (int*)malloc(total * sizeof(int));
    float* ax = (float*)malloc(total * 2 * sizeof(float));
    if (!head || !next || !ax) { free(head); free(next); free(ax); free(cands); return -2; }
    for (int i = 0; i < gw * gh; i++) head[i] = -1;
    const double md2 = min_distance * min_distance;
    for (size_t i = 0; i < total; i++) {
      int y = cands[i].idx / w, x = cands[i].idx - y * w;
      int xc = x / cell, yc = y / cell;
      int x1 = xc - 1 < 0 ? 0 : xc - 1, y1 = yc - 1 < 0 ? 0 : yc - 1;
      int x2 = xc + 1 > gw - 1 ? gw - 1 : xc + 1, y2 = yc + 1 > gh - 1 ? gh - 1 : yc + 1;
      int good = 1;
      for (int yy = y1; yy <= y2 && good; yy++)
        for (int xx = x1; xx <= x2 && good; xx++)
          for (int j = head[yy * gw + xx]; j >= 0; j = next[j]) {
            float dx = (float)x - ax[2 * j], dy = (float)y - ax[2 * j + 1];
            float d = dx * dx;
            float e = dy * dy;
            d = d + e;
            if ((double)d < md2) { good = 0; break; }
          }
      if (good) {
        ax[2 * ncorners] = (float)x; ax[2 * ncorners + 1] = (float)y;
        next[ncorners] = head[yc * gw + xc];
        head[yc * gw + xc] = ncorners;
        if (ncorners < cap && xy) { xy[2 * ncorners] = (float)x; xy[2 * ncorners + 1] = (float)y; }
        ncorners++;
        if (max_corners > 0 && ncorners == max_corners) break;
      }
    }
    free(head); free(next); free(ax);
  } else {
    for (size_t i = 0; i < total; i++) {
      int y = cands[i].idx / w, x = cands[i].idx - y * w;
      if (ncorners < cap && xy) { xy[2 * ncorners] = (float)x; xy[2 * ncorners + 1] = (float)y; }
      ncorners++;
      if (max_corners > 0 && ncorners == max_corners) break;
    }
  }
  free(cands);
  return ncorners;
}

/*
 * visnav::detectKeypoints (include/visnav/keypoints.h:133-150): gftt with
 * (num_features, 0.01, 8, no mask, 3, Harris off), then keep the corners with
 * pangolin InBounds(x, y, 19) (Pangolin image.h:312-315) in order.
 * xy out: doubles (x, y) as KeypointsData::corners stores them.
 */
int vso_detect(const uint8_t* img, int w, int h, int pitch, int num_features,
               double* xy, int cap) {
  if (num_features < 0) return -1;
  int tmpcap = num_features > 0 ? num_features : w * h;
  float* pts = (float*)malloc((size_t)tmpcap * 2 * sizeof(float) + 8);
  if (!pts) return -2;
  int n = vso_gftt(img, w, h, pitch, num_features, 0.01, 8.0, pts, tmpcap);
  if (n < 0) { free(pts); return n; }
  int m = 0;
  const float border = 19.f;
  for (int i = 0; i < n; i++) {
    float x = pts[2 * i], y = pts[2 * i + 1];
    /* InBounds(float x, float y, float border): border <= x < w - border ... */
    if (border <= x && x < ((float)w - border) && border <= y && y < ((float)h - border)) {
      if (m < cap && xy) { xy[2 * m] = x; xy[2 * m + 1] = y; }
      m++;
    }
  }
  free(pts);
  return m;
}

static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
/* Image read.  The reference indexes unchecked (keypoints.h:174-175, 215-216);
 * out-of-image taps are undefined there.  Here (and in the CUDA path) they
 * replicate the border pixel so that both sides stay comparable. */
static inline int px(const uint8_t* img, int w, int h, int pitch, int x, int y) {
  return img[(size_t)clampi(y, 0, h - 1) * pitch + clampi(x, 0, w - 1)];
}

/*
 * visnav::computeAngles (keypoints.h:152-189): first-order moments over the
 * disc x^2+y^2 <= 15^2, angle = atan2(m01, m10).  The reference accumulates in
 * double; all terms are integers < 2^53, so int64 accumulation is identical.
 */
void vso_angles(const uint8_t* img, int w, int h, int pitch, const double* xy,
                int n, int rotate, double* angles) {
  for (int i = 0; i < n; i++) {
    const int cx = (int)xy[2 * i], cy = (int)xy[2 * i + 1];
    double angle = 0;
    if (rotate) {
      int64_t m01 = 0, m10 = 0;
      for (int x = -15; x <= 15; x++)
        for (int y = -15; y <= 15; y++)
          if (x * x + y * y <= 225) {
            int v = px(img, w, h, pitch, x + cx, y + cy);
            m01 += y * v;
            m10 += x * v;
          }
      angle = atan2((double)m01, (double)m10);
    }
    angles[i] = angle;
  }
}

/*
 * visnav::computeDescriptors (keypoints.h:191-221).  bit i =
 *   I(cx + round(c*xa - s*ya), cy + round(s*xa + c*ya)) <
 *   I(cx + round(c*xb - s*yb), cy + round(s*xb + c*yb)),  c = cos(angle), s = sin(angle).
 * contract != 0 evaluates the rotations the way GCC contracts them at -O3 with
 * FMA hardware (x' = fma(c, x, -(s*y)), y' = fma(s, x, c*y)); contract == 0 is
 * the plain two-multiplies form.  Which one the reference binary uses is a
 * property of its compile flags (CMakeLists.txt:44-61); oracle/_ref is built with
 * contraction on, as the reference's default build would be.
 * desc out: n x 8 uint32, bit i -> word i/32, bit i%32 (std::bitset<256> layout).
 */
void vso_describe(const uint8_t* img, int w, int h, int pitch, const double* xy,
                  const double* angles, int n, int contract, uint32_t* desc) {
  for (int k = 0; k < n; k++) {
    const int cx = (int)xy[2 * k], cy = (int)xy[2 * k + 1];
    const double c = cos(angles[k]), s = sin(angles[k]);
    uint32_t* d = desc + (size_t)k * 8;
    memset(d, 0, 32);
    for (int i = 0; i < 256; i++) {
      int q[4];
      for (int p = 0; p < 2; p++) {
        const double x = kPattern[i][2 * p], y = kPattern[i][2 * p + 1];
        double xr, yr;
        if (contract) {
          double t = s * y;
          xr = fma(c, x, -t);
          t = c * y;
          yr = fma(s, x, t);
        } else {
          double t = c * x, u = s * y;
          xr = t - u;
          t = s * x; u = c * y;
          yr = t + u;
        }
        q[2 * p] = (int)round(xr);
        q[2 * p + 1] = (int)round(yr);
      }
      int va = px(img, w, h, pitch, q[0] + cx, q[1] + cy);
      int vb = px(img, w, h, pitch, q[2] + cx, q[3] + cy);
      if (va < vb) d[i >> 5] |= 1u << (i & 31);
    }
  }
}

static inline int hamming256(const uint32_t* a, const uint32_t* b) {
  const uint64_t* p = (const uint64_t*)a;
  const uint64_t* q = (const uint64_t*)b;
  return __builtin_popcountll(p[0] ^ q[0]) + __builtin_popcountll(p[1] ^ q[1]) +
         __builtin_popcountll(p[2] ^ q[2]) + __builtin_popcountll(p[3] ^ q[3]);
}

/* One direction of the scan in matchDescriptors / isPQiffQP
 * (keypoints.h:331-354, 283-304): best, lowest index attaining it, second best. */
static void scan_best2(const uint32_t* q, const uint32_t* db, int n, int* b1, int* id1, int* b2) {
  int best1 = 256, best1_id = 0, best2 = 256;
  for (int j = 0; j < n; j++) {
    int d = hamming256(q, db + (size_t)j * 8);
    if (d < best2) {
      if (d < best1) { best2 = best1; best1 = d; best1_id = j; }
      else best2 = d;
    }
  }
  *b1 = best1; *id1 = best1_id; *b2 = best2;
}

/*
 * visnav::matchDescriptors + isPQiffQP (keypoints.h:323-369, 278-313).
 * pairs out: (id1, id2) int32 pairs in ascending id1.  Returns the count
 * (which may exceed cap; only cap pairs are written) or <0.
 */
int vso_match(const uint32_t* d1, int n1, const uint32_t* d2, int n2, int threshold,
              double dist_2_best, int32_t* pairs, int cap) {
  if (n1 < 0 || n2 < 0) return -1;
  int m = 0;
  for (int i = 0; i < n1; i++) {
    int b1, id, b2;
    scan_best2(d1 + (size_t)i * 8, d2, n2, &b1, &id, &b2);
    if (b1 >= threshold) continue;
    if ((double)b2 < (double)b1 * dist_2_best) continue;
    if (n2 == 0) continue; /* reference would index desc2[0] of an empty vector; no match is the only sane reading */
    int c1, cid, c2;
    scan_best2(d2 + (size_t)id * 8, d1, n1, &c1, &cid, &c2);
    if (c1 >= threshold) continue;
    if ((double)c2 < (double)c1 * dist_2_best) continue;
    if (cid != i) continue;
    if (m < cap && pairs) { pairs[2 * m] = i; pairs[2 * m + 1] = id; }
    m++;
  }
  return m;
}

/* Row statistics only: (best, argbest, second) per query, for kernel-level tests. */
void vso_best2(const uint32_t* q, int nq, const uint32_t* db, int ndb, int32_t* out3) {
  for (int i = 0; i < nq; i++)
    scan_best2(q + (size_t)i * 8, db, ndb, &out3[3 * i], &out3[3 * i + 1], &out3[3 * i + 2]);
}

/* ---- landmark matching ---------------------------------------------------------------------
 * visnav::find_matches_landmarks (include/visnav/vo_utils.h:83-167).  kp_xy/kp_desc: the frame's
 * keypoints; proj_xy: projected landmark positions in the caller's order; landmark j owns the
 * observation descriptors obs_desc[obs_off[j] .. obs_off[j+1]).  pairs = (feature id, j).
 *
 * The reference orders the in-radius candidates with std::partial_sort(first, first + 2, last,
 * by distance) (:132-137), which is not stable.  The two-element heap below restates libstdc++'s
 * algorithm (bits/stl_algo.h __heap_select / __sort_heap, bits/stl_heap.h __make_heap, __pop_heap,
 * __adjust_heap, __push_heap) for middle = first + 2, so equally good candidates are named in the
 * same order as the reference does (pinned against the reference build in oracle/_ref). */
int vso_match_landmarks(const double* kp_xy, const uint32_t* kp_desc, int n_kp, const double* proj_xy,
                        int n_proj, const int32_t* obs_off, const uint32_t* obs_desc,
                        double match_max_dist_2d, int threshold, double dist_2_best, int32_t* pairs,
                        int cap) {
  int m = 0;
  for (int i = 0; i < n_kp; i++) {
    int n = 0, id0 = 0, d0 = 0, id1 = 0, d1 = 0; /* heap: (id0,d0) on top, (id1,d1) below */
    for (int j = 0; j < n_proj; j++) {
      const double dx = kp_xy[2 * i] - proj_xy[2 * j], dy = kp_xy[2 * i + 1] - proj_xy[2 * j + 1];
      const double dist = sqrt(dx * dx + dy * dy); /* Eigen norm(), :108 (built with -ffp-contract=off) */
      if (!(dist < match_max_dist_2d)) continue;
      int best = 256; /* :113 */
      for (int o = obs_off[j]; o < obs_off[j + 1]; o++) {
        const int d = hamming256(obs_desc + (size_t)o * 8, kp_desc + (size_t)i * 8);
        if (d < best) best = d;
      }
      if (n == 0) { id1 = j; d1 = best; }
      else if (n == 1) { /* make_heap: the later element goes on top unless it is strictly smaller */
        if (best < d1) { id0 = id1; d0 = d1; id1 = j; d1 = best; }
        else { id0 = j; d0 = best; }
      } else if (best < d0) { /* __heap_select: pop the top, sift the newcomer */
        if (d1 < best) { id0 = j; d0 = best; }
        else { id0 = id1; d0 = d1; id1 = j; d1 = best; }
      }
      n++;
    }
    if (n == 0) continue;
    if (d1 >= threshold) continue;                                    /* :142 */
    const int second = n >= 2 ? d0 : 256;                             /* :146, :154 */
    if ((double)second < (double)d1 * dist_2_best) continue;
    if (m < cap) { pairs[2 * m] = i; pairs[2 * m + 1] = id1; }
    m++;
  }
  return m;
}

/* detect + angles + describe for one image (keypoints.h:223-229). */
int vso_detect_describe(const uint8_t* img, int w, int h, int pitch, int num_features,
                        int rotate, int contract, double* xy, double* angles,
                        uint32_t* desc, int cap) {
  int n = vso_detect(img, w, h, pitch, num_features, xy, cap);
  if (n < 0) return n;
  if (n > cap) return -3;
  vso_angles(img, w, h, pitch, xy, n, rotate, angles);
  vso_describe(img, w, h, pitch, xy, angles, n, contract, desc);
  return n;
}

#ifdef __cplusplus
}
#endif
