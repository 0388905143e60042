// Standalone probe (debugging aid for detect.cu): which u8 TMA tile loads run on sm_100a.
// usage: tma_probe2 <variant>   0: 3-D u8 box 128x66   1: 2-D u8 box 128x66   2: 3-D u8 box 128x64
//   3: 3-D u8 box 64x66   4: 3-D u32 box 32x66 (same bytes)   5: 2-D u8 box 128x32
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>
#include <stdlib.h>
#include <vector>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int RANK>
__global__ void k(const __grid_constant__ CUtensorMap tmap, int bytes, int x, int y, int z, uint8_t* out) {
  extern __shared__ __align__(128) unsigned char tile[];
  uint64_t& bar = *reinterpret_cast<uint64_t*>(tile + bytes);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bytes) : "memory");
    if (RANK == 3)
      asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                   ::"r"(smem_u32(tile)), "l"(&tmap), "r"(x), "r"(y), "r"(z), "r"(smem_u32(&bar)) : "memory");
    else
      asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                   ::"r"(smem_u32(tile)), "l"(&tmap), "r"(x), "r"(y), "r"(smem_u32(&bar)) : "memory");
  }
  __syncwarp();
  asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(smem_u32(&bar)) : "memory");
  for (int i = threadIdx.x; i < bytes; i += 32) out[i] = tile[i];
}
int main(int argc, char** argv) {
  int variant = argc > 1 ? atoi(argv[1]) : 0;
  int w = 752, h = 480, count = 2, pitch = 768;
  std::vector<uint8_t> img((size_t)pitch * h * count);
  for (size_t i = 0; i < img.size(); ++i) img[i] = (uint8_t)(i * 7 + (i >> 8));
  uint8_t *d, *o; cudaMalloc(&d, img.size()); cudaMalloc(&o, 70 * 128);
  cudaMemcpy(d, img.data(), img.size(), cudaMemcpyHostToDevice);
  void* fn = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  int rank = (variant == 1 || variant == 5) ? 2 : 3;
  int bw = variant == 3 ? 64 : 128, bh = variant == 2 ? 64 : (variant == 5 ? 32 : 66);
  bool u32 = variant == 4;
  CUtensorMap tm;
  cuuint64_t dims[3] = {(cuuint64_t)(u32 ? w / 4 : w), (cuuint64_t)h, (cuuint64_t)count};
  cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)pitch * h};
  cuuint32_t box[3] = {(cuuint32_t)(u32 ? bw / 4 : bw), (cuuint32_t)bh, 1};
  cuuint32_t es[3] = {1, 1, 1};
  CUresult r = ((Enc)fn)(&tm, u32 ? CU_TENSOR_MAP_DATA_TYPE_UINT32 : CU_TENSOR_MAP_DATA_TYPE_UINT8, rank, d, dims, strides, box, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("variant %d rank %d box %dx%d encode %d\n", variant, rank, bw, bh, (int)r);
  int x = argc > 2 ? atoi(argv[2]) : 116, y = 57, z = rank == 3 ? 1 : 0;
  int bytes = bw * bh;
  if (rank == 3) k<3><<<1, 32, bytes + 16>>>(tm, bytes, u32 ? x / 4 : x, y, z, o);
  else k<2><<<1, 32, bytes + 16>>>(tm, bytes, x, y, z, o);
  cudaError_t e = cudaDeviceSynchronize();
  printf("variant %d: %s\n", variant, cudaGetErrorString(e));
  if (e != cudaSuccess) return 1;
  std::vector<uint8_t> t(bytes);
  cudaMemcpy(t.data(), o, t.size(), cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int r2 = 0; r2 < bh; ++r2) for (int c = 0; c < bw; ++c) {
    int gx = x + c, gy = y + r2;
    uint8_t want = (gx < 0 || gx >= w || gy < 0 || gy >= h) ? 0 : img[(size_t)z * pitch * h + (size_t)gy * pitch + gx];
    if (t[r2 * bw + c] != want) ++bad;
  }
  printf("variant %d mismatches %d\n", variant, bad);
  return 0;
}
