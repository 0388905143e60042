// umma_hamming.cu -- probe: the 256-bit Hamming matrix as an int8 contraction on the 5th-gen tensor cores.
//
//   hamming(a, b) = (256 - dot(sa, sb)) / 2   with sa, sb in {-1, +1}^256 (one signed byte per bit)
//
// (1) correctness: one 128 x 256 tile through tcgen05.mma kind::i8 (operands placed by TMA with the 128-byte swizzle,
//     K-major; accumulators in TMEM, read back with tcgen05.ld 32x32b) against the CPU popcount.
// (2) rates (distances/s, whole chip, persistent CTAs):
//       mma      tcgen05.mma alone (8 x K=32 per 128 x 256 tile, two TMEM stages, no epilogue work)
//       ld       + every accumulator read back with tcgen05.ld
//       e32      + (min, second-min) of 32-bit keys: 1 IMAD + 3 VIMNMX per distance
//       e16      + two 16-bit keys per register: 2 IMAD + (VIMNMX.U16x2 min, max, min) per TWO distances
//       eh16     + the same with the keys compared as f16 bit patterns (HMNMX2 / VHMNMX)
//     and, for the record the round-1 verdict asked for, the legacy path:
//       b1       mma.sync.m16n8k256 .b1 and.popc (ptxas expands it to MOVM.U4TO8 + 8 IMMA.16832.U8)
//       u8       mma.sync.m16n8k32 .u8 on bytes expanded beforehand (8 per 16 x 8 x 256-bit tile)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o umma_hamming umma_hamming.cu
// Every wait is bounded (a spin that times out raises a flag and leaves), so a wrong descriptor cannot hang the GPU.
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#define CK(x)                                                                                  \
  do {                                                                                         \
    cudaError_t e_ = (x);                                                                      \
    if (e_ != cudaSuccess) {                                                                   \
      fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_));       \
      exit(2);                                                                                 \
    }                                                                                          \
  } while (0)

constexpr int kM = 128, kN = 256, kKB = 256;  // tile rows, tile columns, bytes of K per descriptor
constexpr int kEpiWarps = 4;
constexpr uint32_t kSpinLimit = 20000000u;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ int g_timeout;

__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok = 0;
  for (uint32_t it = 0; it < kSpinLimit; ++it) {
    asm volatile(
        "{\n.reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (ok) return true;
  }
  atomicExch(&g_timeout, 1);
  return false;
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* tm, void* dst, uint64_t* bar, int x, int y) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
          smem_u32(dst)),
      "l"(tm), "r"(x), "r"(y), "r"(smem_u32(bar))
      : "memory");
}

// K-major operand tile in shared memory, 128-byte swizzle (what TMA wrote): rows 128 B apart, 8-row groups 1024 B apart.
// Fields (cute/arch/mma_sm100_desc.hpp, SmemDescriptor): start >> 4 at [0,14), LBO >> 4 at [16,30) (1: unused by the
// swizzled K-major forms), SBO >> 4 at [32,46), version 1 at [46,48), layout SWIZZLE_128B = 2 at [61,64).
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// kind::i8 instruction descriptor (InstrDescriptor): c_format S32 = 2 at [4,6), a/b format signed 8 bit = 1 at [7,10) /
// [10,13), A and B K-major (bits 15, 16 = 0), N >> 3 at [17,23), M >> 4 at [24,29)
__host__ __device__ constexpr uint32_t umma_idesc_s8(int m, int n) {
  return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n}\n" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "r"(0u)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,"
      "%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
        "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
        "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
// 64 columns as 32 registers: register j = low 16 bits of column 2j | low 16 bits of column 2j + 1 << 16
__device__ __forceinline__ void tmem_ld32_pack16(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.pack::16b.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,"
      "%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
        "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
        "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// bits -> signed bytes: bit 1 -> +1 (0x01), bit 0 -> -1 (0xFF).  One thread per (descriptor, 16-bit group): one 16-byte store.
__global__ void expand_kernel(const uint32_t* __restrict__ desc, int n, uint8_t* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * 16) return;
  const int d = i >> 4, g = i & 15;
  const uint32_t bits = (desc[d * 8 + (g >> 1)] >> ((g & 1) * 16)) & 0xFFFFu;
  uint32_t w[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint32_t nib = (bits >> (4 * q)) & 0xFu;
    const uint32_t b = (nib * 0x00204081u) & 0x01010101u;  // bit k of the nibble -> byte k (0 / 1)
    w[q] = b | ((b ^ 0x01010101u) * 0xFFu);                // 1 -> 0x01, 0 -> 0xFF
  }
  reinterpret_cast<uint4*>(out)[i] = make_uint4(w[0], w[1], w[2], w[3]);
}

struct Smem {
  // operand tiles: two K halves each, [rows][128 B] swizzled; 1024-byte aligned
  uint8_t a[2][kM * 128];
  uint8_t b[2][kN * 128];
  uint64_t bar_load;
  uint64_t bar_full[2];
  uint64_t bar_empty[2];
  uint32_t tmem_base;
};

// mode 0: correctness (one tile, dots written to out[128][256]); modes 1..5: rates
template <int kMode>
__global__ void __launch_bounds__(32 * (kEpiWarps + 1), 1)
    umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, int tiles, int32_t* out,
                unsigned long long* sink) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    mbar_init(&s.bar_load, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s.bar_full[i], 1);
      mbar_init(&s.bar_empty[i], kEpiWarps);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kEpiWarps) {  // the MMA warp owns the TMEM allocation: all 512 columns = two 128 x 256 s32 accumulators
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s.tmem_base)), "n"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_base;

  if (warp == kEpiWarps) {
    // ---- producer + MMA issuer (one lane) ----
    if (lane == 0) {
      mbar_expect_tx(&s.bar_load, (kM + kN) * kKB);
      for (int hk = 0; hk < 2; ++hk) {
        tma_load_2d(&tmA, s.a[hk], &s.bar_load, hk * 128, 0);
        tma_load_2d(&tmB, s.b[hk], &s.bar_load, hk * 128, 0);
      }
      bool ok = mbar_wait(&s.bar_load, 0);
      tc_fence_after();
      const uint32_t idesc = umma_idesc_s8(kM, kN);
      for (int t = 0; t < tiles && ok; ++t) {
        const int st = t & 1;
        if (t >= 2) ok = mbar_wait(&s.bar_empty[st], ((t >> 1) - 1) & 1);
        tc_fence_after();
        if (!ok) break;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint64_t da = umma_desc_sw128(smem_u32(s.a[k >> 2]) + (k & 3) * 32);
          const uint64_t db = umma_desc_sw128(smem_u32(s.b[k >> 2]) + (k & 3) * 32);
          umma_i8(tmem + st * kN, da, db, idesc, k > 0);
        }
        umma_commit(&s.bar_full[st]);
      }
    }
    __syncwarp();
  } else {
    // ---- epilogue: warp w reads TMEM lanes 32w..32w+31 (= tile rows), 32 columns per load ----
    const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
    uint32_t m1 = 0xFFFFFFFFu, m2 = 0xFFFFFFFFu;       // running (min, second min) of this row
    uint32_t p1 = 0xFFFFFFFFu, p2 = 0xFFFFFFFFu;       // packed 16-bit variants
    if (kMode == 5) p1 = p2 = 0x7BFF7BFFu;             // largest finite f16 pattern
    unsigned long long acc = 0;
    bool ok = true;
    for (int t = 0; t < tiles && ok; ++t) {
      const int st = t & 1;
      ok = mbar_wait(&s.bar_full[st], (t >> 1) & 1);
      tc_fence_after();
      if (!ok) break;
      if (kMode >= 6) {
        // 16-bit packed loads (dot in [-256, 256] fits): 64 columns per load.  6: loads only; 7: keys
        // (256 - dot) << 6 | column & 63 from ONE xor (sign bits -> offset binary, no borrow between the halves) and ONE
        // IMAD per two distances, then the packed (min, second min); 8: the registers as they arrive, for the host check
#pragma unroll 1
        for (int c0 = 0; c0 < kN; c0 += 64) {
          uint32_t v[32];
          tmem_ld32_pack16(lane_base + st * kN + c0, v);
          tmem_ld_wait();
          if (kMode == 8) {
            int32_t* o = out + (size_t)(warp * 32 + lane) * kN + c0 / 2;
#pragma unroll
            for (int j = 0; j < 32; ++j) o[j] = (int32_t)v[j];
          } else if (kMode == 6) {
            uint32_t x = 0;
#pragma unroll
            for (int j = 0; j < 32; j += 4) x ^= (v[j] ^ v[j + 1]) ^ (v[j + 2] ^ v[j + 3]);
            acc += x;
          } else {
            const int mul = -64;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const uint32_t half = 64u * (32768u + 256u) + (uint32_t)(2 * j);
              const uint32_t cst = half + ((half + 1u) << 16);  // mod 2^32: per half (33024 - x) * 64 + column
              const uint32_t key = (uint32_t)((int)(v[j] ^ 0x80008000u) * mul + (int)cst);
              const uint32_t tmax = __vmaxu2(p1, key);
              p1 = __vminu2(p1, key);
              p2 = __vminu2(p2, tmax);
            }
          }
        }
      } else if (kMode != 1) {
#pragma unroll 1
        for (int c0 = 0; c0 < kN; c0 += 32) {
          uint32_t v[32];
          tmem_ld32(lane_base + st * kN + c0, v);
          tmem_ld_wait();
          if (kMode == 0) {
            int32_t* o = out + (size_t)(warp * 32 + lane) * kN + c0;
#pragma unroll
            for (int j = 0; j < 32; ++j) o[j] = (int32_t)v[j];
          } else if (kMode == 2) {
            uint32_t x = 0;
#pragma unroll
            for (int j = 0; j < 32; j += 4) x ^= (v[j] ^ v[j + 1]) ^ (v[j + 2] ^ v[j + 3]);  // 3-input LOPs: ~0.4 op per value
            acc += x;
          } else if (kMode == 3) {
            // key = (256 - dot) << 19 | column   (256 - dot = 2 * hamming): one IMAD, then (min, second min)
            const int kmul = -(1 << 19);
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const uint32_t key = (uint32_t)((int)v[j] * kmul + ((256 << 19) + c0 + j));
              const uint32_t tmax = max(m1, key);
              m1 = min(m1, key);
              m2 = min(m2, tmax);
            }
          } else if (kMode == 4 || kMode == 5) {
            // two 16-bit keys per register: key16 = (256 - dot) << 5 | (column & 63) for columns (j, j + 1)
            const int mlo = -32, mhi = -(32 << 16);
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
              const uint32_t cst = (uint32_t)((8192 + ((c0 + j) & 63)) | ((8192 + ((c0 + j + 1) & 63)) << 16));
              uint32_t key = (uint32_t)((int)v[j + 1] * mhi + (int)cst);
              key = (uint32_t)((int)v[j] * mlo + (int)key);
              if (kMode == 4) {
                const uint32_t tmax = __vmaxu2(p1, key);
                p1 = __vminu2(p1, key);
                p2 = __vminu2(p2, tmax);
              } else {
                const __half2 k2 = *reinterpret_cast<const __half2*>(&key);
                __half2 a1 = *reinterpret_cast<__half2*>(&p1), a2 = *reinterpret_cast<__half2*>(&p2);
                const __half2 tmax = __hmax2(a1, k2);
                a1 = __hmin2(a1, k2);
                a2 = __hmin2(a2, tmax);
                p1 = *reinterpret_cast<uint32_t*>(&a1);
                p2 = *reinterpret_cast<uint32_t*>(&a2);
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&s.bar_empty[st]);
    }
    if (kMode >= 2) sink[blockIdx.x * 128 + threadIdx.x] = acc + m1 + m2 + p1 + p2;
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kEpiWarps) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512) : "memory");
  }
}

// ---- MMA-only rate for other instruction shapes: N = 128 tiles, optionally as the matcher issues them (per tile pair
// eight MMAs A x B^T and eight with the operands swapped), no epilogue, two TMEM stages ----
template <int kNn, bool kSwap>
__global__ void __launch_bounds__(32 * (kEpiWarps + 1), 1)
    umma_shape_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, int tiles, int side = 0,
                      uint32_t* side_sink = nullptr) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  volatile uint32_t* done = reinterpret_cast<volatile uint32_t*>(&s.bar_empty[0]);  // (unused barrier slot as a flag)
  if (threadIdx.x == 0) {
    mbar_init(&s.bar_load, 1);
    mbar_init(&s.bar_full[0], 1);
    *done = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kEpiWarps) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s.tmem_base)), "n"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_base;
  if (warp == kEpiWarps && lane == 0) {
    mbar_expect_tx(&s.bar_load, (kM + kN) * kKB);
    for (int hk = 0; hk < 2; ++hk) {
      tma_load_2d(&tmA, s.a[hk], &s.bar_load, hk * 128, 0);
      tma_load_2d(&tmB, s.b[hk], &s.bar_load, hk * 128, 0);
    }
    bool ok = mbar_wait(&s.bar_load, 0);
    tc_fence_after();
    const uint32_t idesc = umma_idesc_s8(kM, kNn);
    for (int t = 0; t < tiles && ok; ++t) {
      const uint32_t d = tmem + (t & 1) * 256;
#pragma unroll
      for (int k = 0; k < 8; ++k)
        umma_i8(d, umma_desc_sw128(smem_u32(s.a[k >> 2]) + (k & 3) * 32), umma_desc_sw128(smem_u32(s.b[k >> 2]) + (k & 3) * 32), idesc, k > 0);
      if (kSwap) {
#pragma unroll
        for (int k = 0; k < 8; ++k)
          umma_i8(d + 128, umma_desc_sw128(smem_u32(s.b[k >> 2]) + (k & 3) * 32), umma_desc_sw128(smem_u32(s.a[k >> 2]) + (k & 3) * 32), idesc, k > 0);
      }
    }
    umma_commit(&s.bar_full[0]);
    mbar_wait(&s.bar_full[0], 0);
    *done = 1;
  } else if (warp < kEpiWarps && side) {
    // side work on the other warps (one per scheduler) while the MMAs run: does register-only arithmetic slow the tensor pipe?
    //   1: 16-bit min/max (ALU pipe)  2: IMAD (FMA pipe)  3: both  4: shared-memory loads and stores
    uint32_t x[8], y[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { x[i] = threadIdx.x * 2654435761u + i; y[i] = x[i] ^ 0x5bd1e995u; }
    uint32_t* sm = reinterpret_cast<uint32_t*>(s.b[1]) + 8192 + threadIdx.x;  // a corner the MMA's operands do not use (kN = 256 rows: all used; see below)
    while (!*done) {
#pragma unroll 1
      for (int it = 0; it < 64; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          if (side & 1) { const uint32_t t = __vmaxu2(x[i], y[i]); x[i] = __vminu2(x[i], y[i]); y[i] = __vminu2(y[i] + 1, t); }
          if (side & 2) { x[i] = x[i] * 0x9E3779B1u + y[i]; y[i] = y[i] * 0x85EBCA77u + x[i]; }
          if (side & 4) { x[i] += side_sink[(x[i] & 1023u)]; }
        }
      }
    }
    uint32_t a = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) a += x[i] + y[i];
    side_sink[1024 + blockIdx.x * 128 + threadIdx.x] = a;
    (void)sm;
  }
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == kEpiWarps) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512) : "memory");
}

template <int kNn, bool kSwap>
static double run_shape(const CUtensorMap& tmA, const CUtensorMap& tmB, int sms, int tiles, size_t smem, int side = 0) {
  static uint32_t* side_sink = nullptr;
  if (!side_sink) {
    CK(cudaMalloc(&side_sink, (1024 + 148 * 256) * 4));
    CK(cudaMemset(side_sink, 0, (1024 + 148 * 256) * 4));
  }
  CK(cudaFuncSetAttribute(umma_shape_kernel<kNn, kSwap>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  umma_shape_kernel<kNn, kSwap><<<sms, 32 * (kEpiWarps + 1), smem>>>(tmA, tmB, 64, side, side_sink);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  umma_shape_kernel<kNn, kSwap><<<sms, 32 * (kEpiWarps + 1), smem>>>(tmA, tmB, tiles, side, side_sink);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  return (double)sms * tiles * kM * kNn * (kSwap ? 2 : 1) / (ms * 1e-3);   // raw 128 x N x 256 products per second, as distances
}

// ---- legacy mma.sync probes: register-only loops, 8 warps per CTA, many CTAs ----
__global__ void mma_b1_kernel(int iters, int* sink) {
  uint32_t a0 = threadIdx.x * 2654435761u, a1 = a0 ^ 0x9E3779B9u, a2 = a0 * 3, a3 = a1 * 5, b0 = a0 + 7, b1 = a1 + 11;
  int c[4][4] = {};
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      asm volatile(
          "mma.sync.aligned.m16n8k256.row.col.s32.b1.b1.s32.and.popc {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
          : "+r"(c[u][0]), "+r"(c[u][1]), "+r"(c[u][2]), "+r"(c[u][3])
          : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0 + u), "r"(b1 + it));
    }
  }
  int s = 0;
  for (int u = 0; u < 4; ++u)
    for (int k = 0; k < 4; ++k) s += c[u][k];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void mma_u8_kernel(int iters, int* sink) {
  uint32_t a[8][4], b[8][2];
  for (int k = 0; k < 8; ++k) {
    for (int q = 0; q < 4; ++q) a[k][q] = ((threadIdx.x + k * 4 + q) * 2654435761u) & 0x01010101u;
    for (int q = 0; q < 2; ++q) b[k][q] = ((threadIdx.x * 7 + k * 2 + q) * 2246822519u) & 0x01010101u;
  }
  int c[4][4] = {};
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        asm volatile(
            "mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
            : "+r"(c[u][0]), "+r"(c[u][1]), "+r"(c[u][2]), "+r"(c[u][3])
            : "r"(a[k][0]), "r"(a[k][1]), "r"(a[k][2]), "r"(a[k][3]), "r"(b[k][0] + u), "r"(b[k][1] + it));
      }
    }
  }
  int s = 0;
  for (int u = 0; u < 4; ++u)
    for (int k = 0; k < 4; ++k) s += c[u][k];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ---- which pipe do the 16-bit min/max forms run on?  register-only chains, rates in lane-ops per clock per SM ----
template <int kKind>
__global__ void minmax_kernel(int iters, uint32_t* sink, unsigned long long* clk) {
  uint32_t m[8], x[8];
  for (int i = 0; i < 8; ++i) {
    m[i] = 0x7BFF7BFFu - threadIdx.x * 17u - i;
    x[i] = 0x12341234u + threadIdx.x * 3u + i * 977u;
  }
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (kKind == 0) {  // VIMNMX.U16x2
        m[i] = __vminu2(m[i], x[i]);
        x[i] = __vmaxu2(x[i], m[(i + 1) & 7]);
      } else if (kKind == 1) {  // HMNMX2
        __half2 a = *reinterpret_cast<__half2*>(&m[i]), b = *reinterpret_cast<__half2*>(&x[i]);
        __half2 c = *reinterpret_cast<__half2*>(&m[(i + 1) & 7]);
        a = __hmin2(a, b);
        b = __hmax2(b, c);
        m[i] = *reinterpret_cast<uint32_t*>(&a);
        x[i] = *reinterpret_cast<uint32_t*>(&b);
      } else {  // one of each: if they sit on different pipes the pair rate doubles
        m[i] = __vminu2(m[i], x[i]);
        __half2 b = *reinterpret_cast<__half2*>(&x[i]), c = *reinterpret_cast<__half2*>(&m[(i + 1) & 7]);
        b = __hmax2(b, c);
        x[i] = *reinterpret_cast<uint32_t*>(&b);
      }
    }
  }
  const long long t1 = clock64();
  uint32_t s = 0;
  for (int i = 0; i < 8; ++i) s += m[i] ^ x[i];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) clk[blockIdx.x] = (unsigned long long)(t1 - t0);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static CUtensorMap make_map(EncodeTiledFn enc, void* base, int rows, int box_rows) {
  CUtensorMap tm;
  const cuuint64_t dims[2] = {(cuuint64_t)kKB, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)kKB};
  const cuuint32_t box[2] = {128, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    fprintf(stderr, "cuTensorMapEncodeTiled failed: %d\n", (int)r);
    exit(2);
  }
  return tm;
}

template <int kMode>
static double run_rate(const CUtensorMap& tmA, const CUtensorMap& tmB, int sms, int tiles, int32_t* out, unsigned long long* sink,
                       size_t smem) {
  CK(cudaFuncSetAttribute(umma_kernel<kMode>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  umma_kernel<kMode><<<sms, 32 * (kEpiWarps + 1), smem>>>(tmA, tmB, 64, out, sink);  // warm-up
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  umma_kernel<kMode><<<sms, 32 * (kEpiWarps + 1), smem>>>(tmA, tmB, tiles, out, sink);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  return (double)sms * tiles * kM * kN / (ms * 1e-3);
}

template <int kKind>
static double run_minmax(int sms) {
  uint32_t* sink;
  unsigned long long* clk;
  CK(cudaMalloc(&sink, (size_t)sms * 512 * 4));
  CK(cudaMalloc(&clk, (size_t)sms * 8));
  const int iters = 20000;
  minmax_kernel<kKind><<<sms, 512>>>(100, sink, clk);
  minmax_kernel<kKind><<<sms, 512>>>(iters, sink, clk);
  CK(cudaDeviceSynchronize());
  std::vector<unsigned long long> h(sms);
  CK(cudaMemcpy(h.data(), clk, (size_t)sms * 8, cudaMemcpyDeviceToHost));
  double mean = 0;
  for (auto c : h) mean += (double)c;
  mean /= sms;
  CK(cudaFree(sink));
  CK(cudaFree(clk));
  return 512.0 * iters * 16 / mean;  // lane-ops per clock per SM (16 ops per iteration per thread)
}

int main() {
  int dev = 0;
  CK(cudaSetDevice(dev));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, dev));
  const int sms = prop.multiProcessorCount;
  void* fp = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
  EncodeTiledFn enc = (EncodeTiledFn)fp;

  // random descriptors
  std::vector<uint32_t> ha(kM * 8), hb(kN * 8);
  uint64_t st = 0x9E3779B97F4A7C15ull;
  auto rnd = [&]() { st ^= st << 13; st ^= st >> 7; st ^= st << 17; return (uint32_t)(st >> 16); };
  for (auto& x : ha) x = rnd();
  for (auto& x : hb) x = rnd();
  for (int i = 0; i < 8; ++i) hb[5 * 8 + i] = ha[3 * 8 + i];  // one exact match: row 3 / column 5
  uint32_t *da, *db;
  uint8_t *ea, *eb;
  int32_t* dout;
  unsigned long long* sink;
  CK(cudaMalloc(&da, ha.size() * 4));
  CK(cudaMalloc(&db, hb.size() * 4));
  CK(cudaMalloc(&ea, (size_t)kM * kKB));
  CK(cudaMalloc(&eb, (size_t)kN * kKB));
  CK(cudaMalloc(&dout, (size_t)kM * kN * 4));
  CK(cudaMalloc(&sink, (size_t)sms * 256 * 8));
  CK(cudaMemcpy(da, ha.data(), ha.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(db, hb.data(), hb.size() * 4, cudaMemcpyHostToDevice));
  expand_kernel<<<(kM * 16 + 255) / 256, 256>>>(da, kM, ea);
  expand_kernel<<<(kN * 16 + 255) / 256, 256>>>(db, kN, eb);
  CK(cudaDeviceSynchronize());
  {  // the expansion itself
    std::vector<uint8_t> he((size_t)kM * kKB);
    CK(cudaMemcpy(he.data(), ea, he.size(), cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int d = 0; d < kM; ++d)
      for (int k = 0; k < 256; ++k) {
        const int bit = (ha[d * 8 + k / 32] >> (k % 32)) & 1;
        bad += he[(size_t)d * kKB + k] != (bit ? 0x01 : 0xFF);
      }
    printf("{\"probe\": \"expand\", \"bad_bytes\": %d}\n", bad);
  }
  const CUtensorMap tmA = make_map(enc, ea, kM, kM), tmB = make_map(enc, eb, kN, kN);
  const size_t smem = sizeof(Smem) + 1024;
  const int zero = 0;
  CK(cudaMemcpyToSymbol(g_timeout, &zero, sizeof(int)));

  // (1) correctness
  CK(cudaFuncSetAttribute(umma_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaMemset(dout, 0x7F, (size_t)kM * kN * 4));
  umma_kernel<0><<<1, 32 * (kEpiWarps + 1), smem>>>(tmA, tmB, 1, dout, sink);
  CK(cudaDeviceSynchronize());
  int timeout = 0;
  CK(cudaMemcpyFromSymbol(&timeout, g_timeout, sizeof(int)));
  std::vector<int32_t> hout((size_t)kM * kN);
  CK(cudaMemcpy(hout.data(), dout, hout.size() * 4, cudaMemcpyDeviceToHost));
  long bad = 0;
  int first_bad = -1;
  for (int i = 0; i < kM; ++i)
    for (int j = 0; j < kN; ++j) {
      int h = 0;
      for (int w = 0; w < 8; ++w) h += __builtin_popcount(ha[i * 8 + w] ^ hb[j * 8 + w]);
      if (hout[(size_t)i * kN + j] != 256 - 2 * h) {
        if (first_bad < 0) first_bad = i * kN + j;
        ++bad;
      }
    }
  printf("{\"probe\": \"umma_i8_tile\", \"timeout\": %d, \"mismatches\": %ld, \"of\": %d, \"first_bad\": %d, \"d_3_5\": %d, "
         "\"d_0_0\": %d, \"d_0_1\": %d, \"d_1_0\": %d}\n",
         timeout, bad, kM * kN, first_bad, hout[3 * kN + 5], hout[0], hout[1], hout[kN]);
  fflush(stdout);
  {  // what pack::16b hands out: register j of the k-th 64-column load = columns (64k + 2j, 64k + 2j + 1) as int16 halves?
    CK(cudaFuncSetAttribute(umma_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CK(cudaMemset(dout, 0x7F, (size_t)kM * kN * 4));
    umma_kernel<8><<<1, 32 * (kEpiWarps + 1), smem>>>(tmA, tmB, 1, dout, sink);
    CK(cudaDeviceSynchronize());
    std::vector<int32_t> hp((size_t)kM * kN);
    CK(cudaMemcpy(hp.data(), dout, hp.size() * 4, cudaMemcpyDeviceToHost));
    long badp = 0, badkey = 0;
    int firstp = -1;
    for (int i = 0; i < kM; ++i)
      for (int j = 0; j < kN / 2; ++j) {
        const uint32_t r = (uint32_t)hp[(size_t)i * kN + j];
        const int lo = (int16_t)(r & 0xFFFF), hi = (int16_t)(r >> 16);
        const int wlo = hout[(size_t)i * kN + 2 * j], whi = hout[(size_t)i * kN + 2 * j + 1];
        if (lo != wlo || hi != whi) {
          if (firstp < 0) firstp = i * kN + j;
          ++badp;
        }
        // the key arithmetic of mode 7 on the host: per half (256 - dot) << 6 | column & 63, no cross-half borrow
        const uint32_t half = 64u * (32768u + 256u) + (uint32_t)((2 * j) & 63);
        const uint32_t cst = half + ((half + 1u) << 16);
        const uint32_t key = (uint32_t)((int)(r ^ 0x80008000u) * -64 + (int)cst);
        const uint32_t want = (uint32_t)(((256 - wlo) << 6) | ((2 * j) & 63)) | ((uint32_t)(((256 - whi) << 6) | ((2 * j + 1) & 63)) << 16);
        badkey += key != want;
      }
    printf("{\"probe\": \"tmem_ld_pack16\", \"mismatches\": %ld, \"key_mismatches\": %ld, \"of\": %d, \"first_bad\": %d, "
           "\"r_0_0\": \"%08x\", \"want_0_0\": [%d, %d]}\n",
           badp, badkey, kM * kN / 2, firstp, (uint32_t)hp[0], hout[0], hout[1]);
    fflush(stdout);
  }
  if (timeout || bad) {
    printf("{\"probe\": \"umma_rates\", \"skipped\": \"tile test failed\"}\n");
  } else {
    const int tiles = 4000;
    const double r1 = run_rate<1>(tmA, tmB, sms, tiles, dout, sink, smem);
    const double r2 = run_rate<2>(tmA, tmB, sms, tiles, dout, sink, smem);
    const double r3 = run_rate<3>(tmA, tmB, sms, tiles, dout, sink, smem);
    const double r4 = run_rate<4>(tmA, tmB, sms, tiles, dout, sink, smem);
    const double r5 = run_rate<5>(tmA, tmB, sms, tiles, dout, sink, smem);
    const double r6 = run_rate<6>(tmA, tmB, sms, tiles, dout, sink, smem);
    const double r7 = run_rate<7>(tmA, tmB, sms, tiles, dout, sink, smem);
    CK(cudaMemcpyFromSymbol(&timeout, g_timeout, sizeof(int)));
    printf("{\"probe\": \"umma_rates\", \"sms\": %d, \"tiles_per_cta\": %d, \"timeout\": %d, \"dist_per_s\": {\"mma\": %.4g, "
           "\"ld\": %.4g, \"e32\": %.4g, \"e16\": %.4g, \"eh16\": %.4g, \"ld_pack16\": %.4g, \"e16_pack16\": %.4g}, "
           "\"int8_tops_mma\": %.1f}\n",
           sms, tiles, timeout, r1, r2, r3, r4, r5, r6, r7, r1 * 512 / 1e12);
  }
  fflush(stdout);

  if (!timeout && !bad) {
    const double s256 = run_shape<256, false>(tmA, tmB, sms, 4000, smem);
    const double s128 = run_shape<128, false>(tmA, tmB, sms, 8000, smem);
    const double s128s = run_shape<128, true>(tmA, tmB, sms, 4000, smem);
    CK(cudaMemcpyFromSymbol(&timeout, g_timeout, sizeof(int)));
    printf("{\"probe\": \"umma_shapes\", \"timeout\": %d, \"raw_dist_per_s\": {\"m128n256\": %.4g, \"m128n128\": %.4g, "
           "\"m128n128_plus_swapped\": %.4g}, \"cycles_per_mma_at_1965\": {\"m128n256\": %.1f, \"m128n128\": %.1f, \"m128n128_plus_swapped\": %.1f}}\n",
           timeout, s256, s128, s128s, 128.0 * 256 / 8 / (s256 / sms / 1.965e9), 128.0 * 128 / 8 / (s128 / sms / 1.965e9),
           128.0 * 128 / 8 / (s128s / sms / 1.965e9));
    fflush(stdout);
    // the same MMA stream with the other four warps (one per scheduler) busy with register-only arithmetic
    double sd[5];
    for (int side = 0; side < 5; ++side) sd[side] = run_shape<128, true>(tmA, tmB, sms, 4000, smem, side);
    printf("{\"probe\": \"umma_under_side_work\", \"shape\": \"m128n128_plus_swapped\", \"raw_dist_per_s\": {\"none\": %.4g, "
           "\"minmax_alu\": %.4g, \"imad_fma\": %.4g, \"both\": %.4g, \"global_loads\": %.4g}}\n",
           sd[0], sd[1], sd[2], sd[3], sd[4]);
    fflush(stdout);
  }
  printf("{\"probe\": \"minmax_pipes\", \"lane_ops_per_clk_per_sm\": {\"vimnmx_u16x2\": %.1f, \"hmnmx2\": %.1f, \"mixed\": %.1f}}\n",
         run_minmax<0>(sms), run_minmax<1>(sms), run_minmax<2>(sms));

  // legacy mma.sync
  {
    int* isink;
    CK(cudaMalloc(&isink, (size_t)sms * 8 * 256 * 4));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const int iters = 4000, blocks = sms * 8;
    float ms;
    mma_b1_kernel<<<blocks, 256>>>(16, isink);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    mma_b1_kernel<<<blocks, 256>>>(iters, isink);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    CK(cudaEventElapsedTime(&ms, e0, e1));
    const double rb1 = (double)blocks * 8 * iters * 4 * 128 / (ms * 1e-3);
    mma_u8_kernel<<<blocks, 256>>>(16, isink);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    mma_u8_kernel<<<blocks, 256>>>(iters, isink);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    CK(cudaEventElapsedTime(&ms, e0, e1));
    const double ru8 = (double)blocks * 8 * iters * 4 * 128 / (ms * 1e-3);
    printf("{\"probe\": \"mma_sync\", \"dist_per_s\": {\"b1_and_popc_m16n8k256\": %.4g, \"u8_m16n8k32_x8\": %.4g}}\n", rb1, ru8);
  }
  return 0;
}
