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
// Design (B200): one warp per keypoint.  The 37 x 37 patch the two functions can touch (reach
// 18 px, < the 19 px border detectKeypoints leaves) is staged once into shared memory with
// aligned 128-bit loads (37 rows x 4 x 16 B); moments and the 512 pattern taps then read shared
// memory.  Lane l produces bits 32k + l, so descriptor word k is one __ballot_sync.
#include "common.cuh"

namespace {

constexpr int kWarpsPerCta = 8;
constexpr int kPatchRows = 37;   // dy in [-18, 18]
constexpr int kPatchPitch = 64;  // bytes; 4 aligned 16-byte chunks cover 37 px at any alignment
constexpr int kReach = 18;

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
  int contract;
};

__global__ void __launch_bounds__(kWarpsPerCta * 32) orient_describe_kernel(const DescribeParams p) {
  __shared__ __align__(16) unsigned char patch_all[kWarpsPerCta][kPatchRows * kPatchPitch];
  __shared__ char4 pat[256];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  pat[tid] = g_pattern[tid];
  __syncthreads();

  const int img = blockIdx.x;
  const int kp = blockIdx.y * kWarpsPerCta + warp;
  const int n = p.n_kp[img];
  if (kp >= n) return;
  const size_t slot = (size_t)img * p.cap + kp;
  const double2 pt = reinterpret_cast<const double2*>(p.xy)[slot];
  const int cx = (int)pt.x, cy = (int)pt.y;  // keypoints.h:160-161: double -> int truncation
  const uint8_t* im = p.base + (size_t)img * p.stride;
  unsigned char* patch = patch_all[warp];

  // ---- stage the patch ------------------------------------------------------------------
  const int x0 = cx - kReach, y0 = cy - kReach;
  const bool interior = x0 >= 0 && y0 >= 0 && cx + kReach < p.w && cy + kReach < p.h;
  int xbase;  // image x of patch byte 0
  if (interior && (p.pitch & 15) == 0 && ((uintptr_t)im & 15) == 0) {
    xbase = x0 & ~15;
    for (int item = lane; item < kPatchRows * 4; item += 32) {
      const int row = item >> 2, q = item & 3;
      const int xs = xbase + q * 16;
      if (xs <= cx + kReach) {  // chunk holds a needed pixel; it ends at or before `pitch`
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(im + (size_t)(y0 + row) * p.pitch + xs));
        *reinterpret_cast<uint4*>(patch + row * kPatchPitch + q * 16) = v;
      }
    }
  } else {
    // Border keypoints (never produced by detectKeypoints): the reference reads out of bounds
    // here; like the oracle we replicate the edge pixel so the result is defined.
    xbase = x0;
    for (int item = lane; item < kPatchRows * kPatchRows; item += 32) {
      const int row = item / kPatchRows, col = item - row * kPatchRows;
      const int yy = min(max(y0 + row, 0), p.h - 1), xx = min(max(x0 + col, 0), p.w - 1);
      patch[row * kPatchPitch + col] = im[(size_t)yy * p.pitch + xx];
    }
  }
  __syncwarp();
  const int off = cx - xbase;  // patch column of the keypoint
#define PATCH(dx, dy) patch[((dy) + kReach) * kPatchPitch + (dx) + off]

  // ---- orientation -----------------------------------------------------------------------
  double angle = 0.0;
  if (p.mode_angles == 1) {
    int m10 = 0, m01 = 0;
    const int dx = lane - 15;  // lanes 0..30 <-> dx -15..15
    if (lane < 31) {
      const int lim = 225 - dx * dx;
#pragma unroll 1
      for (int dy = -15; dy <= 15; ++dy) {
        if (dy * dy <= lim) {
          const int v = PATCH(dx, dy);
          m10 += dx * v;
          m01 += dy * v;
        }
      }
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
      m10 += __shfl_xor_sync(0xFFFFFFFFu, m10, s);
      m01 += __shfl_xor_sync(0xFFFFFFFFu, m01, s);
    }
    angle = atan2((double)m01, (double)m10);
    if (lane == 0) p.angles[slot] = angle;
  } else if (p.mode_angles == 2) {
    angle = p.angles[slot];
  } else if (lane == 0) {
    p.angles[slot] = 0.0;
  }
  if (!p.desc) return;

  // ---- rBRIEF ------------------------------------------------------------------------------
  const double c = cos(angle), s = sin(angle);
  uint32_t my_word = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const char4 q = pat[k * 32 + lane];
    const double xa = q.x, ya = q.y, xb = q.z, yb = q.w;
    double xar, yar, xbr, ybr;
    if (p.contract) {
      // x' = fma(c, x, -(s*y)); y' = fma(s, x, c*y): GCC's contraction of keypoints.h:206-214
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
    const int xai = (int)round(xar), yai = (int)round(yar);
    const int xbi = (int)round(xbr), ybi = (int)round(ybr);
    const int va = PATCH(xai, yai), vb = PATCH(xbi, ybi);
    const uint32_t word = __ballot_sync(0xFFFFFFFFu, va < vb);
    if (lane == k) my_word = word;
  }
  if (lane < 8) p.desc[slot * 8 + lane] = my_word;
#undef PATCH
}

}  // namespace

int launch_orient_describe(vslam_ctx* ctx, const DevImages& im, int count, const double* xy,
                           const int32_t* n_kp, int cap, int mode_angles, double* angles,
                           uint32_t* desc) {
  if (count <= 0 || cap <= 0) return VSLAM_OK;
  DescribeParams p;
  p.base = im.base; p.w = im.w; p.h = im.h; p.pitch = im.pitch; p.stride = im.stride;
  p.xy = xy; p.n_kp = n_kp; p.cap = cap; p.mode_angles = mode_angles;
  p.angles = angles; p.desc = desc; p.contract = ctx->fp_contract;
  dim3 grid(count, (cap + kWarpsPerCta - 1) / kWarpsPerCta);
  orient_describe_kernel<<<grid, kWarpsPerCta * 32, 0, ctx->stream>>>(p);
  ctx->launches++;
  VS_CUDA(ctx, cudaGetLastError());
  return VSLAM_OK;
}
