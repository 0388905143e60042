// pyramid.cu -- factor-2 image pyramid for the multi-scale configuration (BASELINE config 5).
//
// The reference has no multi-scale behaviour on this path (cv::goodFeaturesToTrack is single-scale;
// only cv::ORB inside compute_bow_vector builds a pyramid), so this is a declared EXTENSION
// (SURVEY.md section 8d): level k+1 (x, y) = (a + b + c + d + 2) >> 2 over the 2x2 block of level k,
// integer arithmetic, bit-exact on both sides; detection / orientation / description then run per
// level with the reference's semantics (keypoints.h:133-229) on the level image.
//
// Byte work, HBM-bound: each thread averages a 16 x 2 source block into 8 output pixels (two
// 128-bit loads, one 64-bit store); every source byte is read once.
#include "common.cuh"

namespace {

struct PyrParams {
  const uint8_t* src;
  uint8_t* dst;
  int sw, sh, spitch;
  size_t sstride;
  int dw, dh, dpitch;
  size_t dstride;
  int count;
};

__device__ __forceinline__ uint32_t avg4(uint32_t a0, uint32_t a1) {
  // a0 = 4 bytes of the upper row, a1 = the 4 bytes below: two output pixels
  const uint32_t s0 = (a0 & 0xFF) + ((a0 >> 8) & 0xFF) + (a1 & 0xFF) + ((a1 >> 8) & 0xFF) + 2;
  const uint32_t s1 = ((a0 >> 16) & 0xFF) + (a0 >> 24) + ((a1 >> 16) & 0xFF) + (a1 >> 24) + 2;
  return (s0 >> 2) | ((s1 >> 2) << 8);
}

__global__ void __launch_bounds__(256) pyr_down_kernel(const PyrParams p) {
  const int groups = (p.dw + 7) / 8;  // 8 output pixels per thread
  const long long total = (long long)groups * p.dh * p.count;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(t % groups);
    const long long r = t / groups;
    const int y = (int)(r % p.dh), img = (int)(r / p.dh);
    const uint8_t* s0 = p.src + (size_t)img * p.sstride + (size_t)(2 * y) * p.spitch + 16 * g;
    const uint8_t* s1 = s0 + p.spitch;
    uint8_t* d = p.dst + (size_t)img * p.dstride + (size_t)y * p.dpitch + 8 * g;
    const int nout = min(8, p.dw - 8 * g);
    if (nout == 8 && 16 * g + 16 <= p.spitch) {
      const uint4 u = __ldg(reinterpret_cast<const uint4*>(s0));
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(s1));
      uint2 o;
      o.x = avg4(u.x, v.x) | (avg4(u.y, v.y) << 16);
      o.y = avg4(u.z, v.z) | (avg4(u.w, v.w) << 16);
      *reinterpret_cast<uint2*>(d) = o;
    } else {
      for (int k = 0; k < nout; ++k)
        d[k] = (uint8_t)((s0[2 * k] + s0[2 * k + 1] + s1[2 * k] + s1[2 * k + 1] + 2) >> 2);
    }
  }
}

}  // namespace

int launch_pyr_down(vslam_ctx* ctx, const DevImages& src, const DevImages& dst, int count) {
  if (count <= 0 || dst.w <= 0 || dst.h <= 0) return VSLAM_OK;
  PyrParams p;
  p.src = src.base; p.dst = const_cast<uint8_t*>(dst.base);
  p.sw = src.w; p.sh = src.h; p.spitch = src.pitch; p.sstride = src.stride;
  p.dw = dst.w; p.dh = dst.h; p.dpitch = dst.pitch; p.dstride = dst.stride;
  p.count = count;
  const long long total = (long long)((dst.w + 7) / 8) * dst.h * count;
  long long blocks = (total + 255) / 256;
  const long long maxb = (long long)ctx->num_sms * 16;
  if (blocks > maxb) blocks = maxb;
  pyr_down_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(p);
  ctx->launches++;
  VS_CUDA(ctx, cudaGetLastError());
  return VSLAM_OK;
}
