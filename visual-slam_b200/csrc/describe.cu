// describe.cu -- intensity-centroid orientation and 256-bit rBRIEF description for sm_100a.
//
// Semantics: visnav::computeAngles (reference include/visnav/keypoints.h:152-189) and
// visnav::computeDescriptors (:191-221).
//   angle = atan2(m01, m10), m10 = sum x*I(cx+x, cy+y), m01 = sum y*I(...) over x^2+y^2 <= 15^2
//   bit i = I(cx + round(c*xa - s*ya), cy + round(s*xa + c*ya)) <
//           I(cx + round(c*xb - s*yb), cy + round(s*xb + c*yb)),  c = cos(angle), s = sin(angle)
// The moments are integers (|.| < 2^23), accumulated in int32; atan2 / sincos / the rotation are
// fp64 with explicit rounding (__dmul_rn / __fma_rn: no contraction the source does not spell).
//
// Design (B200): one warp owns 32 consecutive keypoints of one image and works in three phases.
//   1. per keypoint: the 37 x 37 patch both functions can touch (reach 18 px, < the 19 px border
//      detectKeypoints leaves) is staged into shared memory with aligned 128-bit loads, and the
//      moments are taken four pixels at a time with DP4A (weights x, and 1 for the row sums that
//      y multiplies), one (row, 4-column group) item per lane and step;
//   2. ONCE per 32 keypoints: lane i evaluates atan2 / sincos of keypoint i -- the ~250 fp64
//      instructions that a warp-per-keypoint layout executes 32 times redundantly;
//   3. per keypoint: patch staged again (L1/L2 hit), lane l computes bits 32k + l; the lane's
//      eight pattern rows stay in registers as doubles, coordinates are rounded by ONE fp64 add of
//      a fixed-point magic constant (C round() semantics proven below), descriptor word k is a ballot.
// Blocks of one image are consecutive (blockIdx.x = chunk, .y = image), so an image's patches
// are read from L2 once: the first layout had the image index fastest and re-read DRAM 11 times.
#include "common.cuh"

namespace {

constexpr int kWarpsPerCta = 4;
constexpr int kPatchRows = 37;   // dy in [-18, 18]
constexpr int kPatchPitch = 64;  // bytes; 4 aligned 16-byte chunks cover 37 px at any alignment
constexpr int kReach = 18;
constexpr int kHalfPatch = 15;   // keypoints.h:49 HALF_PATCH_SIZE

// row i = (xa, ya, xb, yb); plain global memory: lanes read 32 different rows at once
__device__ const char4 g_pattern[256] = {
#include "rbrief_pattern.inc"
};

struct DescribeParams {
  const uint8_t* base;
  int w, h, pitch;
  size_t stride;
  const double* xy;      // [count][cap][2]
  const int32_t* n_kp;   // [count]
  int cap;
  int mode_angles;       // 0: zero, 1: compute, 2: given in `angles`
  double* angles;        // [count][cap]
  uint32_t* desc;        // [count][cap][8] or null
  uint8_t* descx;        // optional [count][cap][256]: every bit b as the signed byte 2b - 1 (operands of match_mma.cu)
  int contract;
};

__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b_signed_bytes, int c) {
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b_signed_bytes), "r"(c));
  return d;
}

// Per-lane constants of the patch copy: lane = (row0, 16-byte chunk q); 5 rows per lane, 8 apart.
struct StageLane {
  const uint8_t* src;   // image base + row0 * pitch + q * 16
  uint32_t dst0;        // shared address of (row0, q) in patch buffer 0 of the warp (buffer 1: + patch bytes)
  size_t pitch8;        // 8 image rows in bytes
  int q;
  bool row4;            // the lane's fifth row (row0 + 32) exists
};

// Starts staging one keypoint's patch (cp.async, 16 bytes per item: global -> shared without passing
// through registers) and commits the group; the caller waits with patch_wait<N>() before reading.
// kp_off / kp_meta are the keypoint's precomputed byte offset of patch (row 0, byte 0) in the image
// and its (patch column of the keypoint | interior << 8).  Two buffers per warp: the next keypoint's
// patch is in flight while the current one is processed.
__device__ __forceinline__ void stage_patch(const DescribeParams& p, const uint8_t* im, const StageLane& L, int buf,
                                            int kp_off, int kp_meta, int cx, int cy, unsigned char* patch,
                                            int lane) {
  if (kp_meta & 256) {
    if (L.q * 16 <= (kp_meta & 255) + kReach) {  // chunk holds a needed pixel; it ends at or before `pitch`
      const uint8_t* src = L.src + kp_off;
      const uint32_t dst = L.dst0 + buf * (kPatchRows * kPatchPitch);
#pragma unroll
      for (int it = 0; it < 5; ++it) {
        if (it < 4 || L.row4)
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + it * 8 * kPatchPitch),
                       "l"(src + it * L.pitch8)
                       : "memory");
      }
    }
  } else {
    // Border keypoints (never produced by detectKeypoints): the reference reads out of bounds
    // here; like the oracle we replicate the edge pixel so the result is defined.
    const int x0 = cx - kReach, y0 = cy - kReach;
    for (int item = lane; item < kPatchRows * kPatchRows; item += 32) {
      const int row = item / kPatchRows, col = item - row * kPatchRows;
      const int yy = min(max(y0 + row, 0), p.h - 1), xx = min(max(x0 + col, 0), p.w - 1);
      patch[row * kPatchPitch + col] = im[(size_t)yy * p.pitch + xx];
    }
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int kPending>
__device__ __forceinline__ void patch_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(kPending) : "memory");
  __syncwarp();
}

// (int)round(v) for |v| < 2^10 from ONE fp64 add, plus a tie detector.
//   A = low word of (v + kRoundMagic), kRoundMagic = 1.5 * 2^32 + 0.5: the add rounds v + 0.5 to a
//   multiple of 2^-20 (nearest-even), so A = RNE((v + 0.5) * 2^20) as a two's-complement integer and
//   A >> 20 = floor(v + 0.5 +- 2^-21).  That equals C round(v) (halfway cases away from zero) unless
//   v lies within 2^-21 of a halfway point -- exactly the case (A & 0xFFFFF) == 0, which the caller
//   screens for and recomputes with round() itself.  (For negative halfway points floor(v + 0.5)
//   differs from round(v); they are screened the same way.)
constexpr double kRoundMagic = 6442450944.5;
__device__ __forceinline__ int round_fix20(double v) { return __double2loint(__dadd_rn(v, kRoundMagic)); }

// Rotates the two test points of one pattern row by the keypoint angle (keypoints.h:206-214).
template <bool kContract>
__device__ __forceinline__ void rotate_pair(double c, double s, char4 q, double& xar, double& yar, double& xbr,
                                            double& ybr) {
  const double xa = q.x, ya = q.y, xb = q.z, yb = q.w;
  if (kContract) {
    // x' = fma(c, x, -(s*y)); y' = fma(s, x, c*y): GCC's contraction of the reference expressions
    xar = __fma_rn(c, xa, -__dmul_rn(s, ya));
    yar = __fma_rn(s, xa, __dmul_rn(c, ya));
    xbr = __fma_rn(c, xb, -__dmul_rn(s, yb));
    ybr = __fma_rn(s, xb, __dmul_rn(c, yb));
  } else {
    xar = __dsub_rn(__dmul_rn(c, xa), __dmul_rn(s, ya));
    yar = __dadd_rn(__dmul_rn(s, xa), __dmul_rn(c, ya));
    xbr = __dsub_rn(__dmul_rn(c, xb), __dmul_rn(s, yb));
    ybr = __dadd_rn(__dmul_rn(s, xb), __dmul_rn(c, yb));
  }
}

#ifndef VSLAM_DESC_CTAS
#define VSLAM_DESC_CTAS 6
#endif
template <bool kContract>
__global__ void __launch_bounds__(kWarpsPerCta * 32, VSLAM_DESC_CTAS) orient_describe_kernel(const DescribeParams p) {
  __shared__ __align__(16) unsigned char patch_all[kWarpsPerCta][2][kPatchRows * kPatchPitch];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int img = blockIdx.y;
  const int kp0 = (blockIdx.x * kWarpsPerCta + warp) * 32;
  const int n = p.n_kp[img];
  const int nk = min(32, n - kp0);
  if (nk <= 0) return;
  const size_t slot0 = (size_t)img * p.cap + kp0;
  const uint8_t* im = p.base + (size_t)img * p.stride;
  unsigned char* const patch0 = patch_all[warp][0];
  unsigned char* const patch1 = patch_all[warp][1];

  int cx = 0, cy = 0;
  if (lane < nk) {
    const double2 pt = reinterpret_cast<const double2*>(p.xy)[slot0 + lane];
    cx = (int)pt.x;  // keypoints.h:160-161: double -> int truncation
    cy = (int)pt.y;
  }

  // patch geometry of this lane's keypoint: byte offset of patch (row 0, byte 0), patch column of the
  // keypoint (18..33 when the patch starts on a 16-byte boundary, 18 on the border path), interior flag
  int kp_off, kp_meta;
  {
    const int x0 = cx - kReach, y0 = cy - kReach;
    const bool interior = x0 >= 0 && y0 >= 0 && cx + kReach < p.w && cy + kReach < p.h && (p.pitch & 15) == 0 &&
                          ((uintptr_t)im & 15) == 0;
    const int xbase = interior ? (x0 & ~15) : x0;
    kp_off = y0 * p.pitch + xbase;  // < 2^30 (check_image)
    kp_meta = (cx - xbase) | (interior ? 256 : 0);
  }
  StageLane L;
  {
    const int q = lane & 3, row0 = lane >> 2;
    L.q = q;
    L.row4 = row0 + 32 < kPatchRows;
    L.src = im + (size_t)row0 * p.pitch + q * 16;
    L.pitch8 = (size_t)8 * p.pitch;
    L.dst0 = static_cast<uint32_t>(__cvta_generic_to_shared(patch0 + row0 * kPatchPitch + q * 16));
  }
#define STAGE(i_, buf_)                                                                                   \
  stage_patch(p, im, L, (buf_), __shfl_sync(0xFFFFFFFFu, kp_off, (i_)), __shfl_sync(0xFFFFFFFFu, kp_meta, (i_)), \
              __shfl_sync(0xFFFFFFFFu, cx, (i_)), __shfl_sync(0xFFFFFFFFu, cy, (i_)), (buf_) ? patch1 : patch0, lane)

  // ---- phase 1: moments (keypoints.h:171-183) -------------------------------------------------
  double angle = 0.0;
  if (p.mode_angles == 1) {
    // lane = (row group, 4-column group): rows rsub + 4*it, columns dx = 4g - 16 + j
    const int g = lane & 7, rsub = lane >> 3;
    uint32_t wy[8], wx[8];
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int row = rsub + 4 * it, dy = row - kHalfPatch;
      uint32_t o = 0, wv = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int dx = 4 * g - 16 + j;
        const bool in = row <= 2 * kHalfPatch && dx >= -kHalfPatch && dx * dx + dy * dy <= kHalfPatch * kHalfPatch;
        if (in) {
          o |= (uint32_t)(dy & 0xFF) << (8 * j);
          wv |= (uint32_t)(dx & 0xFF) << (8 * j);
        }
      }
      wy[it] = o;
      wx[it] = wv;
    }
    int my_m10 = 0, my_m01 = 0;
    STAGE(0, 0);
    for (int i = 0; i < nk; ++i) {
      const unsigned char* patch = (i & 1) ? patch1 : patch0;
      const int off = __shfl_sync(0xFFFFFFFFu, kp_meta, i) & 255;
      if (i + 1 < nk) {
        STAGE(i + 1, (i + 1) & 1);
        patch_wait<1>();
      } else {
        patch_wait<0>();
      }
      const int sb = off - 16;  // patch byte of column dx = -16
      const uint32_t* wrow = reinterpret_cast<const uint32_t*>(patch + (rsub + 3) * kPatchPitch) + (sb >> 2) + g;
      const int shift = (sb & 3) * 8;
      int m10 = 0, m01 = 0;
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const uint32_t lo = wrow[it * 4 * (kPatchPitch / 4)], hi = wrow[it * 4 * (kPatchPitch / 4) + 1];
        const uint32_t pix = __funnelshift_r(lo, hi, shift);
        m10 = dp4a_us(pix, wx[it], m10);
        m01 = dp4a_us(pix, wy[it], m01);
      }
      m10 = __reduce_add_sync(0xFFFFFFFFu, m10);
      m01 = __reduce_add_sync(0xFFFFFFFFu, m01);
      if (lane == i) {
        my_m10 = m10;
        my_m01 = m01;
      }
      __syncwarp();  // every lane is done with this buffer before it is refilled
    }
    // ---- phase 2: one atan2 per lane = per keypoint ----
    angle = atan2((double)my_m01, (double)my_m10);
  } else if (p.mode_angles == 2) {
    if (lane < nk) angle = p.angles[slot0 + lane];
  }
  if (p.mode_angles != 2 && lane < nk) p.angles[slot0 + lane] = angle;
  if (!p.desc) return;
  double c, s;
  sincos(angle, &s, &c);

  // ---- phase 3: rBRIEF (keypoints.h:205-217) ----------------------------------------------------
  // this lane's pattern rows 32k + lane stay packed (8 registers); the int8 -> fp64 conversions run on
  // the otherwise idle conversion unit
  char4 pat[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) pat[k] = g_pattern[k * 32 + lane];
  STAGE(0, 0);
  for (int i = 0; i < nk; ++i) {
    const double ci = __shfl_sync(0xFFFFFFFFu, c, i), si = __shfl_sync(0xFFFFFFFFu, s, i);
    const unsigned char* patch = (i & 1) ? patch1 : patch0;
    const int off = __shfl_sync(0xFFFFFFFFu, kp_meta, i) & 255;
    if (i + 1 < nk) {
      STAGE(i + 1, (i + 1) & 1);
      patch_wait<1>();
    } else {
      patch_wait<0>();
    }
    const unsigned char* centre = patch + kReach * kPatchPitch + off;
    uint32_t my_word = 0;
    int fmin = 1;  // smallest fixed-point fraction seen: 0 <=> some coordinate needs the exact round()
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      double xar, yar, xbr, ybr;
      rotate_pair<kContract>(ci, si, pat[k], xar, yar, xbr, ybr);
      const int axa = round_fix20(xar), aya = round_fix20(yar), axb = round_fix20(xbr), ayb = round_fix20(ybr);
      const int xai = axa >> 20, yai = aya >> 20, xbi = axb >> 20, ybi = ayb >> 20;
      // (a - (a >> 20 << 20) = a & 0xFFFFF: one LOP per coordinate, the minima as three-input VIMNMX3)
      fmin = __vimin3_s32(__vimin3_s32(fmin, axa & 0xFFFFF, aya & 0xFFFFF), axb & 0xFFFFF, ayb & 0xFFFFF);
      const int va = centre[yai * kPatchPitch + xai], vb = centre[ybi * kPatchPitch + xbi];
      const uint32_t word = __ballot_sync(0xFFFFFFFFu, va < vb);
      if (lane == k) my_word = word;
    }
    if (__any_sync(0xFFFFFFFFu, fmin == 0)) {
      // a coordinate within 2^-21 of a halfway point (practically never): redo the keypoint with round()
#pragma unroll 1
      for (int k = 0; k < 8; ++k) {
        double xar, yar, xbr, ybr;
        rotate_pair<kContract>(ci, si, pat[k], xar, yar, xbr, ybr);
        const int xai = (int)round(xar), yai = (int)round(yar), xbi = (int)round(xbr), ybi = (int)round(ybr);
        const int va = centre[yai * kPatchPitch + xai], vb = centre[ybi * kPatchPitch + xbi];
        const uint32_t word = __ballot_sync(0xFFFFFFFFu, va < vb);
        if (lane == k) my_word = word;
      }
    }
    if (lane < 8) p.desc[(slot0 + i) * 8 + lane] = my_word;
    if (p.descx) {
      // lane l expands bits [8l, 8l + 8) = byte (l & 3) of word l >> 2 (held by lane l >> 2: one shuffle) to eight signed
      // bytes (1 -> 0x01, 0 -> 0xFF): one coalesced 256-byte row per keypoint
      const uint32_t my_xword = __shfl_sync(0xFFFFFFFFu, my_word, lane >> 2);
      const uint32_t bits = (my_xword >> (8 * (lane & 3))) & 0xFFu;
      const uint32_t b0 = ((bits & 0xFu) * 0x00204081u) & 0x01010101u, b1 = ((bits >> 4) * 0x00204081u) & 0x01010101u;
      reinterpret_cast<uint2*>(p.descx + (slot0 + i) * 256)[lane] =
          make_uint2(b0 | ((b0 ^ 0x01010101u) * 0xFFu), b1 | ((b1 ^ 0x01010101u) * 0xFFu));
    }
    __syncwarp();  // every lane is done with this buffer before it is refilled
  }
#undef STAGE
}

}  // namespace

int launch_orient_describe(vslam_ctx* ctx, const DevImages& im, int count, const double* xy,
                           const int32_t* n_kp, int cap, int mode_angles, double* angles,
                           uint32_t* desc, uint8_t* descx) {
  if (count <= 0 || cap <= 0) return VSLAM_OK;
  DescribeParams p;
  p.base = im.base; p.w = im.w; p.h = im.h; p.pitch = im.pitch; p.stride = im.stride;
  p.xy = xy; p.n_kp = n_kp; p.cap = cap; p.mode_angles = mode_angles;
  p.angles = angles; p.desc = desc; p.descx = desc ? descx : nullptr; p.contract = ctx->fp_contract;
  const int per_cta = kWarpsPerCta * 32;
  if (count > 65535) return vs_fail(ctx, VSLAM_E_LIMIT, "more than 65535 images in one describe launch");
  dim3 grid((cap + per_cta - 1) / per_cta, count);
  if (p.contract) orient_describe_kernel<true><<<grid, kWarpsPerCta * 32, 0, ctx->stream>>>(p);
  else orient_describe_kernel<false><<<grid, kWarpsPerCta * 32, 0, ctx->stream>>>(p);
  ctx->launches++;
  VS_CUDA(ctx, cudaGetLastError());
  return VSLAM_OK;
}
