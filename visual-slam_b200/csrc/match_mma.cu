// match_mma.cu -- matchDescriptors with the all-pairs Hamming matrix on the 5th-generation tensor cores (sm_100a).
//
// Semantics: visnav::matchDescriptors + isPQiffQP (reference include/visnav/keypoints.h:323-369, 278-313), the same
// predicate match.cu implements on the integer pipes (see its header): T(q, D) = (m1, a1, m2), threshold, ratio test,
// mutual check, matches in ascending first index.
//
// The distance matrix is an exact integer contraction: with every descriptor bit b stored as the signed byte 2b - 1,
//     dot(sa, sb) = #equal bits - #different bits = 256 - 2 * hamming(a, b),
// which tcgen05.mma kind::i8 (s8 x s8 -> s32, K = 256 = 8 instructions of K = 32) computes without rounding.  Measured
// (tools/probes/umma_hamming.cu, profiles/r2a_umma_hamming_probe.jsonl): 8.3e12 distances/s for the MMA alone, against
// 6.6e11 for the POPC kernel of match.cu, 2.0e12 for mma.sync u8 and 8.2e11 for mma.sync b1 and.popc.
//
// Layout: descriptors are EXPANDED once per set to 256 signed bytes (expand_desc_kernel; the describe stage's output is
// expanded right behind it, a keyframe store at append time).  One CTA owns one job (one matchDescriptors call) at a
// time, persistent over the batch, warp-specialised (20 warps):
//   warp 0  (one lane)  TMA producer: 128-descriptor operand tiles ([2 K halves][128 rows][128 B], 128-byte swizzle,
//                       K-major) -- kGroup tiles of desc1 stay resident while the tiles of desc2 stream through a ring
//   warps 1, 2          MMA issuers, one per TMEM stage: per (desc1 tile, desc2 tile) TWO 128 x 128 products,
//                       D1 = A * B^T (rows = desc1) and D2 = B * A^T (rows = desc2): both directions of the reference's
//                       test become ROW statistics, so no reduction ever crosses threads.  The whole warp walks the
//                       loops with warp-uniform values and an elected lane issues, so the descriptors live in uniform
//                       registers (3 instructions per MMA); the two issuers take turns, handing over 13 MMAs into a
//                       pair, so that pairs complete one after the other
//   warp 3              spare (keeps warp % 4 = TMEM lane quarter for the epilogue warps)
//   warps 4-19          epilogue, two groups of eight: group g takes the pairs of TMEM stage g; in a group four warps
//                       fold the rows of D1, four the rows of D2.  thread = one TMEM lane = one matrix row;
//                       tcgen05.ld pack::16b hands out two adjacent columns per register (dot fits 16 bits); one xor +
//                       one IMAD turn the register into two 16-bit keys (hamming << 7 | column); (min, second min)
//                       kept with VIMNMX.U16x2 / VIMNMX3 (ALU pipe) and, for part of the max, IMAD pairs (FMA pipe);
//                       per tile the pair is widened to the 32-bit keys of match.cu and merged into the row's
//                       running statistics in shared memory
// Two TMEM stages (2 x 256 columns = the whole TMEM) let the MMA of the next tile pair run under the epilogue of the
// current one.  The ratio test, the mutual check and the ordered compaction run in the same CTA on the epilogue warps
// while the producer / MMA warps already work on the next job.
#include <cuda.h>

#include "common.cuh"

namespace {

constexpr int kTile = 128;                 // descriptors per operand tile = UMMA M = UMMA N
constexpr int kHalfBytes = kTile * 128;    // one K half of a tile: [128 rows][128 B]
constexpr int kTileBytes = 2 * kHalfBytes; // 32 KB
constexpr int kGroup = 2;                  // desc1 tiles resident per pass over desc2
// ring of desc2 tiles: a tile serves two pairs (~2400 cycles), so two stages cover the TMA latency -- three measured the
// same (1.334 / 1.339 ms) and their 32 KB are what lets sets of 4000 descriptors (config 5) keep their statistics here
#ifndef VSLAM_MMA_BSTAGES
#define VSLAM_MMA_BSTAGES 2
#endif
constexpr int kBStages = VSLAM_MMA_BSTAGES;
constexpr int kEpiWarps = 16;              // two groups of 8: group g takes the tile pairs of TMEM stage g
constexpr int kEpiGroupWarps = 8;
constexpr int kMmaWarps = 2;               // one MMA-issuing warp per TMEM stage (see the kernel's header)
constexpr int kCtlWarps = 4;                // hardware warps 0-3: TMA producer, two MMA issuers, one spare (keeps the epilogue's
                                           // warp % 4 = TMEM lane quarter); the epilogue warps follow
constexpr int kThreads = (kCtlWarps + kEpiWarps) * 32;
constexpr uint32_t kShift = 20;            // 32-bit key = distance << 20 | index (as in match.cu)
constexpr uint32_t kIdxMask = (1u << kShift) - 1;
constexpr uint32_t kSent = 256u << kShift;
constexpr int kMaxIndex = 1 << kShift;
constexpr uint32_t kSpinLimit = 1u << 26;
// 1: a ninth K = 32 step multiplies a constant tile by itself (sixteen bytes of 4 per row: 16 * 4 * 4 = 256), so the
// accumulators are dot + 256 = 512 - 2 hamming >= 0 and the epilogue's packed 16-bit arithmetic needs no sign handling
// (one ALU instruction less per two distances, on the pipe that bounds the epilogue) for 12.5 % more tensor work,
// which has the slack.  0: eight steps, signed accumulators, one xor per register.
#ifndef VSLAM_MMA_BIAS
#define VSLAM_MMA_BIAS 1
#endif
constexpr int kMmasPerPair = VSLAM_MMA_BIAS ? 18 : 16;
#ifndef VSLAM_MMA_HANDOFF
#define VSLAM_MMA_HANDOFF (VSLAM_MMA_BIAS ? 15 : 13)
#endif
constexpr uint32_t kHandoff = VSLAM_MMA_HANDOFF;  // MMAs (of a pair's 16) an issuer has issued when the other may start its pair

#ifdef VSLAM_MMA_PROFILE
// cycle counters of CTA 0 (build with -DVSLAM_MMA_PROFILE; read with vslam_debug_mma_clocks):
// 0 MMA total, 1 MMA wait a_full, 2 wait b_full, 3 wait t_empty, 4 issue; 8 epi total, 9 wait t_full, 10 fold, 11 finalize,
// 12 tile pairs; 16.. the same epilogue counters for warp 4
__device__ long long g_mma_clk[32];
// per tile pair of CTA 0 (first 256): clock64 at 0 MMA issue start, 1 issue end (commit), 2 + w: epilogue warp w done
// (arrived on t_empty), 10: warp 0 saw t_full, 11: warp 4 saw t_full
__device__ long long g_mma_trace[128][12];
#define MMA_TRACE(pair, k) do { if ((pair) < 128) s.prof_trace[pair][k] = clock64(); } while (0)
// the counters live in shared memory while the kernel runs (a global read-modify-write per event would put ~700
// cycles on the MMA thread's critical path -- measured) and are dumped by CTA 0 at the end
#if VSLAM_MMA_PROFILE >= 2  // trace only: the accumulating counters cost a shared-memory read-modify-write per event
#define MMA_T0() do { } while (0)
#define MMA_ACC(slot) do { } while (0)
#else
#define MMA_T0() const long long t0_ = clock64()
#define MMA_ACC(slot) do { s.prof_clk[slot] += clock64() - t0_; } while (0)
#endif
#else
#define MMA_T0() do { } while (0)
#define MMA_ACC(slot) do { } while (0)
#define MMA_TRACE(pair, k) do { } while (0)
#endif
__device__ int g_mma_timeout;  // a bounded wait ran out (never in a correct run): read by vslam_debug_mma_timeouts

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// Bounded wait.  A wait that runs out (a pipeline bug; never in a correct run) raises the CTA's fail flag and the global
// one; every later wait of the CTA then returns at once, all roles keep walking their loops (so that every barrier is
// still reached by everybody) and the job's result is reported as an error instead of hanging the GPU.
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity, volatile int* fail) {
  uint32_t ok = 0;
  asm volatile(  // the common case first: a phase that has completed already costs one non-blocking test
      "{\n.reg .pred p;\n"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  if (ok) return true;
  for (uint32_t it = 0; it < kSpinLimit; ++it) {
    if ((it & 1023u) == 1023u && *fail) return false;
    asm volatile(
        "{\n.reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (ok) return true;
  }
  *fail = 1;
  atomicExch(&g_mma_timeout, 1);
  return false;
}
// bounded spin on a shared-memory counter another thread of the CTA advances (same failure protocol as mbar_wait)
__device__ __forceinline__ bool spin_until_ge(volatile uint32_t* ctr, uint32_t target, volatile int* fail) {
  for (uint32_t it = 0; it < kSpinLimit; ++it) {
    if ((int32_t)(*ctr - target) >= 0) return true;
    if ((it & 1023u) == 1023u && *fail) return false;
  }
  *fail = 1;
  atomicExch(&g_mma_timeout, 1);
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
// K-major operand in shared memory under the 128-byte swizzle TMA wrote: rows 128 B apart, 8-row groups 1024 B apart
// (SmemDescriptor: start >> 4 | LBO(unused) 1 << 16 | SBO 64 << 32 | version 1 << 46 | SWIZZLE_128B 2 << 61).
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// kind::i8: D s32 (2 << 4), A / B signed 8 bit (1 << 7, 1 << 10), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
constexpr uint32_t kIdesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kTile >> 3) << 17) | ((uint32_t)(kTile >> 4) << 24);
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t accumulate) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n}\n" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(kIdesc), "r"(accumulate), "r"(0u)
      : "memory");
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// 64 accumulator columns of this thread's TMEM lane as 32 registers: pack::16b keeps the low 16 bits of every column
// (exact here: dot is in [-256, 256]) and puts columns (2j, 2j + 1) into the halves of register j.  Half the TMEM read
// traffic of the 32-bit form: measured (tools/probes/umma_hamming.cu, profiles/r3c_umma_pack16_probe.jsonl) the MMA runs
// at its full rate under these loads (8.6e12 distances/s) and at 6.75e12 under the 32-bit ones.
__device__ __forceinline__ void tmem_ld64_pack16(uint32_t taddr, uint32_t (&v)[32]) {
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
__device__ __forceinline__ void epi_barrier() { asm volatile("bar.sync 1, %0;" ::"n"(kEpiWarps * 32) : "memory"); }

// ---- expansion: bit b of a descriptor -> signed byte 2b - 1 (0x01 / 0xFF) -------------------------------------------
// One thread per (descriptor, 16-bit group): one 16-byte store.  rows_per_set / n_valid (optional): descriptors are
// stored in sets of rows_per_set rows of which only the first n_valid[set] exist (the describe stage's slabs).
__global__ void expand_desc_kernel(const uint32_t* __restrict__ desc, long long rows, int rows_per_set,
                                   const int32_t* __restrict__ n_valid, uint8_t* __restrict__ out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * 16) return;
  const long long d = i >> 4;
  const int g = (int)(i & 15);
  if (n_valid) {
    const long long set = d / rows_per_set;
    if ((int)(d - set * rows_per_set) >= n_valid[set]) return;
  }
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

struct MmaParams {
  const MatchJob* jobs;
  int n_jobs;
  const uint32_t* base1;  // packed descriptor arrays the jobs' d1 / d2 point into; row r of the expanded array under
  const uint32_t* base2;  // tmap1 / tmap2 is the descriptor at base + 8 r
  int threshold;
  double ratio;
  int cap;
  int max_n1, max_n2;
  // -64: the multiplier that turns a register of two 16-bit accumulators into a packed key pair (fold_half).  A kernel
  // parameter rather than a literal: as a literal ptxas re-materialises it with MOVs between the IMADs
  int key_mul;
  int key_one, key_neg1;  // 1 and -1, for the same reason (fold2)
  // Column-split mode for ONE large job (split_range > 0; BASELINE config 4: a query set against a large map): CTA s
  // takes all split_n1 rows of the first array against columns [s * split_range, + split_range) of the second; its
  // row statistics are partial (rowpart[s][i], column indices relative to the range), its column statistics complete
  // (colstat[column]); match.cu's merge_rows / finalize kernels finish the job.
  int split_range, split_n1, split_n2;
  uint2* rowpart;
  uint2* colstat;
};

struct MmaSmem {
  uint8_t a[kGroup][kTileBytes];
  uint8_t b[kBStages][kTileBytes];
#if VSLAM_MMA_BIAS
  uint8_t bias[kHalfBytes];  // [128 rows][128 B], 128-byte swizzle: logical bytes 0-15 of every row = 4, 16-31 = 0
#endif
  uint64_t a_full, a_empty;
  uint64_t b_full[kBStages], b_empty[kBStages];
  uint64_t t_full[2], t_empty[2];
  uint32_t tmem_base;
  int warp_sums[kEpiWarps];
  int running;
  int fail;
  uint32_t progress[kMmaWarps];  // MMAs each issuer has issued so far (16 per tile pair): the issuers' turn-taking
#ifdef VSLAM_MMA_PROFILE
  long long prof_clk[32];
  long long prof_trace[128][12];
#endif
};

__device__ __forceinline__ bool passes(uint32_t m1, uint32_t m2, int threshold, double ratio) {
  // keypoints.h:355-357: best1_d >= threshold -> no; best2_d < best1_d * dist_2_best -> no.
  if ((int)m1 >= threshold) return false;
  return !((double)(int)m2 < __dmul_rn((double)(int)m1, ratio));
}

// Folds two packed keys into a (min, second min) pair of packed lists:
//   lo = min(x, y), hi = max(x, y);  m1' = min(m1, lo);  m2' = min3(m2, max(m1, lo), hi)
// Five min/max on the ALU pipe, which is what bounds the epilogue (16-bit min/max issue every other cycle, like all ALU
// work).  The FMA pipe idles next to it, so kFma of the two plain max are computed there instead, per 16-bit half
//   max(a, b) = (a - min(a, b)) + b        (a - min >= 0: no borrow between the halves; the sum is the max: no carry)
// as two IMADs with multipliers -1 and 1 that are kernel parameters (as literals ptxas folds them back into IADD3s on
// the ALU pipe).
template <int kFma>
__device__ __forceinline__ void fold2(uint32_t x, uint32_t y, uint32_t& m1, uint32_t& m2, int one, int neg1) {
  const uint32_t lo = __vminu2(x, y);
  const uint32_t hi = kFma >= 1 ? (uint32_t)(((int)lo * neg1 + (int)x) * one + (int)y) : __vmaxu2(x, y);
  const uint32_t n1 = __vminu2(m1, lo);
  const uint32_t t = kFma >= 2 ? (uint32_t)(((int)n1 * neg1 + (int)m1) * one + (int)lo) : __vmaxu2(m1, lo);
  m1 = n1;
  m2 = __vimin3_u16x2(m2, t, hi);
}

#ifndef VSLAM_FOLD_FMA_A
#define VSLAM_FOLD_FMA_A (VSLAM_MMA_BIAS ? 1 : 2)  // max on the FMA pipe in the even / odd fold2 of a chain (0, 1 or 2 each)
#endif
#ifndef VSLAM_FOLD_FMA_B
#define VSLAM_FOLD_FMA_B 1
#endif

// One 64-column half of a row (32 registers of two 16-bit accumulators each) into two independent (min, second min)
// chains of packed 16-bit keys (registers 0-15 -> chain a, 16-31 -> chain b: the serial min/max dependency is what
// bounds a warp here).  acc = dot = 256 - 2 hamming, key16 = (256 - dot) << 6 | column = hamming << 7 | column.
// Both halves of a register in ONE xor and ONE IMAD: flipping the sign bits turns the two's-complement halves into
// offset binary x = dot + 32768 > 0, and then in plain 32-bit arithmetic
//   (64 * 33024 + c) * 0x00010001 + 0x00010000 - 64 * (x_lo + 65536 x_hi)
//     = ((256 - dot_lo) * 64 + c) + 65536 ((256 - dot_hi) * 64 + c + 1)          (mod 2^32)
// with both parentheses in [0, 65536): no borrow crosses the halves.  (Without the xor a negative low half would
// borrow 64 from the high one.)
template <bool kMasked>
__device__ __forceinline__ void fold_half(const uint32_t (&v)[32], int c0, int nv, int mul, int one, int neg1, uint32_t& p1a,
                                          uint32_t& p2a, uint32_t& p1b, uint32_t& p2b) {
  const uint32_t cb = (uint32_t)c0 * 0x00010001u;
  uint32_t key[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) {
#if VSLAM_MMA_BIAS
    // accumulators are 512 - 2 hamming in [0, 512]: key = (512 - acc) * 64 + column per half, nothing crosses the halves
    const uint32_t half = 64u * 512u + (uint32_t)(2 * j);
    const uint32_t cst = half + ((half + 1u) << 16);
    uint32_t k = (uint32_t)((int)v[j] * mul + (int)(cst + cb));
#else
    const uint32_t half = 64u * (32768u + 256u) + (uint32_t)(2 * j);
    const uint32_t cst = half + ((half + 1u) << 16);
    uint32_t k = (uint32_t)((int)(v[j] ^ 0x80008000u) * mul + (int)(cst + cb));
#endif
    if (kMasked) k |= (2 * j >= nv ? 0x0000FFFFu : 0u) | (2 * j + 1 >= nv ? 0xFFFF0000u : 0u);  // nv = valid columns of the half
    key[j] = k;
  }
#pragma unroll
  for (int i = 0; i < 16; i += 4) {
    fold2<VSLAM_FOLD_FMA_A>(key[i], key[i + 1], p1a, p2a, one, neg1);
    fold2<VSLAM_FOLD_FMA_B>(key[i + 2], key[i + 3], p1a, p2a, one, neg1);
  }
#pragma unroll
  for (int i = 16; i < 32; i += 4) {
    fold2<VSLAM_FOLD_FMA_A>(key[i], key[i + 1], p1b, p2b, one, neg1);
    fold2<VSLAM_FOLD_FMA_B>(key[i + 2], key[i + 3], p1b, p2b, one, neg1);
  }
}

// (min, second min) of one matrix row over the nvalid first columns of a 128-column TMEM tile, as 16-bit keys
// hamming << 7 | column.  Both of the row's loads are issued before the first one is consumed: a TMEM load behind a
// running MMA takes several hundred cycles (measured: a warp that loaded chunk by chunk spent 2/3 of its time waiting).
__device__ __forceinline__ void fold_tile(uint32_t taddr, int nvalid, int mul, int one, int neg1, uint64_t* t_empty, int lane,
                                          uint32_t& k1, uint32_t& k2, long long* loaded_clk = nullptr) {
  uint32_t p1a = 0xFFFFFFFFu, p2a = 0xFFFFFFFFu, p1b = 0xFFFFFFFFu, p2b = 0xFFFFFFFFu;
  uint32_t v0[32], v1[32];
  tmem_ld64_pack16(taddr, v0);
  if (nvalid > 64) tmem_ld64_pack16(taddr + 64, v1);  // warp-uniform
  tmem_ld_wait();
#ifdef VSLAM_MMA_PROFILE
  if (loaded_clk) *loaded_clk = clock64();
#endif
  // the row now sits in registers: hand the TMEM stage back to the MMA issuer BEFORE the fold, so that the next product
  // into this stage runs under the min/max below instead of after them
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(t_empty);
  if (nvalid == kTile) {
    // the common case as ONE basic block: the scheduler can run the key IMADs (FMA pipe) of one half under the min/max
    // (ALU pipe) of the other
    fold_half<false>(v0, 0, 64, mul, one, neg1, p1a, p2a, p1b, p2b);
    fold_half<false>(v1, 64, 64, mul, one, neg1, p1a, p2a, p1b, p2b);
  } else {
    if (nvalid >= 64) fold_half<false>(v0, 0, 64, mul, one, neg1, p1a, p2a, p1b, p2b);
    else fold_half<true>(v0, 0, nvalid, mul, one, neg1, p1a, p2a, p1b, p2b);
    if (nvalid > 64) fold_half<true>(v1, 64, nvalid - 64, mul, one, neg1, p1a, p2a, p1b, p2b);
  }
  // merge the two chains, then the even-column list (low halves) with the odd-column list (high halves)
  const uint32_t p1 = __vminu2(p1a, p1b), p2 = __vminu2(__vmaxu2(p1a, p1b), __vminu2(p2a, p2b));
  const uint32_t a1 = p1 & 0xFFFFu, b1 = p1 >> 16, a2 = p2 & 0xFFFFu, b2 = p2 >> 16;
  k1 = min(a1, b1);
  k2 = min(max(a1, b1), min(a2, b2));
}

// 16-bit tile key -> 32-bit key of match.cu: distance << 20 | (tile base + column)
__device__ __forceinline__ uint32_t widen_key(uint32_t k16, uint32_t base) {
  return ((k16 >> 7) << kShift) + base + (k16 & 127u);
}

__global__ void __launch_bounds__(kThreads, 1)
    match_mma_kernel(const __grid_constant__ CUtensorMap tmap1, const __grid_constant__ CUtensorMap tmap2, const MmaParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  MmaSmem& s = *reinterpret_cast<MmaSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  // Running statistics of the job, 6 bytes per row: the best key (distance << 20 | index) and the second-best DISTANCE
  // (its index is never needed).  rk / rd: [max_n1] rows T(d1[i], desc2); ck / cd: [2][max_n2] columns T(d2[j], desc1),
  // one copy per epilogue group (both groups meet every desc2 tile), merged at the end.
  const int n1pad = (p.max_n1 + 1) & ~1, n2pad = (p.max_n2 + 1) & ~1;
  uint32_t* rk = reinterpret_cast<uint32_t*>(&s + 1);
  uint32_t* ck = rk + n1pad;
  uint16_t* rd = reinterpret_cast<uint16_t*>(ck + 2 * n2pad);
  uint16_t* cd = rd + n1pad;
  // (broadcast: ptxas then knows the warp index -- and everything derived from it -- is uniform over the warp)
  const int warp = __shfl_sync(0xFFFFFFFFu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    mbar_init(&s.a_full, 1);
    mbar_init(&s.a_empty, kMmaWarps);  // both MMA warps commit to the operand barriers, whether or not they had a tile
    for (int i = 0; i < kBStages; ++i) {
      mbar_init(&s.b_full[i], 1);
      mbar_init(&s.b_empty[i], kMmaWarps);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s.t_full[i], 1);
      mbar_init(&s.t_empty[i], kEpiGroupWarps);
    }
    s.running = 0;
    s.fail = 0;
    s.progress[0] = s.progress[1] = 0;
#ifdef VSLAM_MMA_PROFILE
    for (int i = 0; i < 32; ++i) s.prof_clk[i] = 0;
#endif
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
#if VSLAM_MMA_BIAS
  // the constant operand of the bias step, written as TMA would have (16-byte chunk c of row r sits at chunk c ^ (r & 7)),
  // then made visible to the tensor core's (async proxy) reads
  for (int i = threadIdx.x; i < kHalfBytes / 16; i += kThreads) {
    const int r = i >> 3, pc = i & 7;
    const uint32_t v = ((pc ^ (r & 7)) == 0) ? 0x04040404u : 0u;
    reinterpret_cast<uint4*>(s.bias)[i] = make_uint4(v, v, v, v);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#endif
  if (warp == 1) {  // the first MMA warp owns the TMEM allocation: 512 columns = two stages of (D1 | D2)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s.tmem_base)), "n"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_base;

  // Pipeline positions, each role advancing the ones it uses: desc1 groups, desc2 tiles, tile pairs.  After a failed
  // wait (fail flag) every role still walks the same loops -- waits return at once, work is skipped -- so that the
  // barriers below are reached by everybody.
  uint32_t a_it = 0, b_it = 0, t_it = 0;
  uint32_t seq0 = 0, seq1 = 0;  // MMA issuers: tile pairs of stage 0 / of stage 1 before the current (g0, b)
  volatile int* fail = &s.fail;
  for (int jb = blockIdx.x; jb < p.n_jobs; jb += gridDim.x) {
    const bool split = p.split_range > 0;
    MatchJob job{};
    int n1, n2, row1, row2;
    if (split) {  // job jb = column range jb of the one large job
      n1 = p.split_n1;
      n2 = min(p.split_range, p.split_n2 - jb * p.split_range);
      row1 = 0;
      row2 = jb * p.split_range;
    } else {
      job = p.jobs[jb];
      n1 = job.n1p ? *job.n1p : job.n1;
      n2 = job.n2p ? *job.n2p : job.n2;
      row1 = (int)((job.d1 - p.base1) >> 3);
      row2 = (int)((job.d2 - p.base2) >> 3);
    }
    n1 = __shfl_sync(0xFFFFFFFFu, n1, 0);  // (the same in every lane already: the broadcast tells ptxas so)
    n2 = __shfl_sync(0xFFFFFFFFu, n2, 0);
    const bool valid = n1 > 0 && n2 > 0 && n1 <= p.max_n1 && n2 <= p.max_n2;  // uniform over the CTA
    if (!valid) {
      if (threadIdx.x == 0 && !split) *job.count = (n1 > p.max_n1 || n2 > p.max_n2) ? -1 : 0;
      continue;
    }
    const int nA = (n1 + kTile - 1) / kTile, nB = (n2 + kTile - 1) / kTile;

    if (warp == 0) {
      // ================= TMA producer =================
      if (lane == 0) {
        for (int g0 = 0; g0 < nA; g0 += kGroup) {
          const int ng = min(kGroup, nA - g0);
          bool ok = true;
          if (a_it > 0) ok = mbar_wait(&s.a_empty, (a_it - 1) & 1, fail);
          if (ok) {
            mbar_expect_tx(&s.a_full, ng * kTileBytes);
            for (int t = 0; t < ng; ++t)
              for (int hk = 0; hk < 2; ++hk)
                tma_load_2d(&tmap1, s.a[t] + hk * kHalfBytes, &s.a_full, hk * 128, row1 + (g0 + t) * kTile);
          }
          ++a_it;
          for (int b = 0; b < nB; ++b) {
            const int st = b_it % kBStages;
            if (b_it >= kBStages) ok = mbar_wait(&s.b_empty[st], ((b_it / kBStages) - 1) & 1, fail);
            if (ok) {
              mbar_expect_tx(&s.b_full[st], kTileBytes);
              for (int hk = 0; hk < 2; ++hk)
                tma_load_2d(&tmap2, s.b[st] + hk * kHalfBytes, &s.b_full[st], hk * 128, row2 + b * kTile);
            }
            ++b_it;
          }
        }
      }
      __syncwarp();
    } else if (warp == 3) {
      // spare
    } else if (warp < kCtlWarps) {
      // ================= MMA issuers: warp 1 + t multiplies the group's desc1 tile t into TMEM stage t =================
      // A tcgen05.mma issue blocks while the tensor pipe's short queue is full, and everything else this thread does per
      // pair (barrier waits, commits, descriptor arithmetic) is one serial dependency chain of several hundred cycles:
      // with one issuer per stage that chain runs under the other stage's MMAs.
      // The issuers take turns in the order the pairs are consumed -- stage 1's pair (g0, b) behind stage 0's pair
      // (g0, b), stage 0's next pair behind that -- handing over kHandoff MMAs into a pair, so that the next pair's first
      // MMA queues up behind this pair's last.  The tensor pipe executes in issue order: the pairs complete one pair time
      // apart and the two epilogue groups alternate.  (Without the turns both issuers start together, their MMAs
      // interleave, both pairs complete together and both groups then load and fold at the same time while the pipe
      // idles: measured 1775 cycles per pair.)
      // The whole warp walks the loops with warp-uniform values (sizes and the TMEM base are broadcast, wait results
      // voted), one elected lane issues the tcgen05 instructions: ptxas then builds the shared-memory descriptors in
      // uniform registers (two UIADDs per MMA) instead of per-lane arithmetic followed by an ELECT / R2UR loop per operand
      // (21 instructions per MMA, which is what bounded the issue rate once four epilogue warps share the scheduler).
      const int t = warp - 1;
      const bool leader = elect_one();
      volatile uint32_t* mine = &s.progress[t];
      volatile uint32_t* other = &s.progress[1 - t];
      const uint32_t tm = __shfl_sync(0xFFFFFFFFu, tmem, 0) + t * 256;
      for (int g0 = 0; g0 < nA; g0 += kGroup) {
        const int ng = min(kGroup, nA - g0);
        bool ok;
        { MMA_T0(); ok = mbar_wait(&s.a_full, a_it & 1, fail); MMA_ACC(1); }
        for (int b = 0; b < nB; ++b) {
          const int st = b_it % kBStages;
          { MMA_T0(); ok = mbar_wait(&s.b_full[st], (b_it / kBStages) & 1, fail) && ok; MMA_ACC(2); }
          if (t < ng) {
            if (leader) MMA_TRACE(2 * t_it + t, 4);
            // descriptor low words (address >> 4 | LBO): adding (offset >> 4) moves the start address (no carry: < 2^14)
            const uint32_t a_lo = (uint32_t)umma_desc_sw128(smem_u32(s.a[t])), b_lo = (uint32_t)umma_desc_sw128(smem_u32(s.b[st]));
            constexpr uint64_t kDescHi = (64ull << 32) | (1ull << 46) | (2ull << 61);
#if VSLAM_MMA_BIAS
            const uint32_t bias_lo = (uint32_t)umma_desc_sw128(smem_u32(s.bias));
#else
            const uint32_t bias_lo = 0;
#endif
            // my turn (true long before it is needed once the pipeline runs), then the stage: what follows the stage's
            // release is on the critical path of the pair (release -> first MMA), so nothing else is left for there.
            // (Handing the two halves of a stage over separately -- D1's MMAs as soon as the D1 warps have read columns
            // 0-127, D2's release waited for between the halves -- measured slower, 1.40 against 1.33 ms: the wait in the
            // middle of a pair stalls the issuer longer than the earlier start gains.)
            if (t == 1 || seq1 > 0) ok = spin_until_ge(other, (t ? kMmasPerPair * seq0 : kMmasPerPair * (seq1 - 1)) + kHandoff, fail) && ok;
            if (leader) MMA_TRACE(2 * t_it + t, 5);
            { MMA_T0(); if (t_it >= 1) ok = mbar_wait(&s.t_empty[t], (t_it - 1) & 1, fail) && ok; MMA_ACC(3); }
            tc_fence_after();
            MMA_T0();
            if (leader) MMA_TRACE(2 * t_it + t, 0);
            if (leader) {
              if (ok) {
#pragma unroll
                for (int k = 0; k < kMmasPerPair; ++k) {
                  // first half: D1[i][j] = dot(desc1 row i, desc2 row j); second half: D2[j][i], the roles swapped; the
                  // last step of a half (VSLAM_MMA_BIAS) adds 256 to every accumulator
                  constexpr int kHalf = kMmasPerPair / 2;
                  const int kk = k % kHalf;
                  const bool second = k >= kHalf;
                  const uint32_t d = tm + (second ? kTile : 0);
                  if (kk < 8) {
                    const uint32_t off = ((kk >> 2) * kHalfBytes + (kk & 3) * 32) >> 4;
                    const uint32_t r_lo = (second ? b_lo : a_lo) + off, c_lo = (second ? a_lo : b_lo) + off;
                    umma_i8(d, kDescHi | r_lo, kDescHi | c_lo, kk > 0);
                  } else {
                    umma_i8(d, kDescHi | bias_lo, kDescHi | bias_lo, 1);
                  }
                  if (k == (int)kHandoff - 1) *mine = kMmasPerPair * t_it + kHandoff;
                }
                MMA_TRACE(2 * t_it + t, 6);
                umma_commit(&s.t_full[t]);
              } else {
                *mine = kMmasPerPair * t_it + kHandoff;
              }
              MMA_TRACE(2 * t_it + t, 1);
            }
            MMA_ACC(4);
            __syncwarp();
            ++t_it;  // uses of this warp's stage
          }
          ++seq0;
          if (ng == kGroup) ++seq1;
          ok = __all_sync(0xFFFFFFFFu, ok);
          if (ok && leader) umma_commit(&s.b_empty[st]);  // arrives when this warp's MMAs that read the desc2 tile are done
          ++b_it;
        }
        ok = __all_sync(0xFFFFFFFFu, ok);
        if (ok && leader) umma_commit(&s.a_empty);
        ++a_it;
      }
      __syncwarp();
    } else {
      // ================= epilogue: group g = warp / 8 takes the tile pairs of TMEM stage g (= desc1 tile g of the
      // resident group); in a group warps 0-3 fold the rows of D1, warps 4-7 the rows of D2.  While one group folds a
      // pair (ALU work), the other waits for / loads the next one: the TMEM load latency (several hundred cycles behind
      // a running MMA) and the barrier waits hide behind the other group's min/max =================
      const int ew = warp - kCtlWarps;           // 0..15
      const int et = threadIdx.x - kCtlWarps * 32;  // 0..511
      const int g = ew >> 3;
      for (int i = et; i < n1; i += kEpiWarps * 32) { rk[i] = kSent; rd[i] = 256; }
      for (int j = et; j < n2; j += kEpiWarps * 32) { ck[j] = ck[n2pad + j] = kSent; cd[j] = cd[n2pad + j] = 256; }
      epi_barrier();
      const bool d2 = (ew & 4) != 0;
      const int q = ew & 3;  // = warp % 4: the TMEM lane quarter this warp may read
      for (int g0 = 0; g0 < nA; g0 += kGroup) {
        const int ng = min(kGroup, nA - g0);
        if (g >= ng) continue;  // an odd last group: stage 1 is not used
        const int t = g;
        for (int b = 0; b < nB; ++b) {
          const uint32_t use = t_it;  // uses of my stage so far
          bool ok;
          { MMA_T0(); ok = mbar_wait(&s.t_full[t], use & 1, fail); if (lane == 0 && q == 0 && g == 0) MMA_ACC(d2 ? 17 : 9); }
          tc_fence_after();
          MMA_T0();
          if (lane == 0 && q == 0) MMA_TRACE(2 * use + t, d2 ? 11 : 10);
          const int i0 = (g0 + t) * kTile, j0 = b * kTile;
          // my row of the tile, the columns that exist, and where the row's statistics live
          const int row = (d2 ? j0 : i0) + q * 32 + lane;
          const int nrows = d2 ? n2 : n1;
          const int nvalid = d2 ? min(kTile, n1 - i0) : min(kTile, n2 - j0);
          const uint32_t cbase = d2 ? i0 : j0;
          if (ok && q * 32 < (d2 ? n2 - j0 : n1 - i0)) {  // warp-uniform: the warp holds at least one real row
            uint32_t k1, k2;
#ifdef VSLAM_MMA_PROFILE
            long long* lc = (lane == 0 && q == 0 && !d2 && 2 * use + t < 128) ? &s.prof_trace[2 * use + t][7] : nullptr;
#else
            long long* lc = nullptr;
#endif
            fold_tile(tmem + ((uint32_t)(q * 32) << 16) + t * 256 + (d2 ? kTile : 0), nvalid, p.key_mul, p.key_one, p.key_neg1,
                      &s.t_empty[t], lane, k1, k2, lc);
            if (row < nrows) {
              uint32_t* kp = (d2 ? ck + g * n2pad : rk) + row;
              uint16_t* dp = (d2 ? cd + g * n2pad : rd) + row;
              const uint32_t old = *kp, w1 = widen_key(k1, cbase);
              *kp = min(old, w1);
              *dp = (uint16_t)min(min((uint32_t)*dp, k2 >> 7), max(old, w1) >> kShift);  // (k2 >> 7: the tile's second distance)
            }
          } else {  // nothing to read from this stage: release it at once
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&s.t_empty[t]);
          }
          if (lane == 0 && q == 0 && g == 0) MMA_ACC(d2 ? 18 : 10);
          if (lane == 0 && q == 0) MMA_TRACE(2 * use + t, d2 ? 3 : 2);
          ++t_it;
        }
      }
      epi_barrier();
      // one list of column statistics out of the two groups' (the associative merge of the scan)
      for (int j = et; j < n2; j += kEpiWarps * 32) {
        const uint32_t a = ck[j], c = ck[n2pad + j];
        ck[j] = min(a, c);
        cd[j] = (uint16_t)min(max(a, c) >> kShift, (uint32_t)min(cd[j], cd[n2pad + j]));
      }
      epi_barrier();
      MMA_T0();
      const bool good = *fail == 0;  // read behind the barrier: the same for all epilogue warps
      if (split) {
        // partial row statistics and complete column statistics of this column range; a pipeline failure leaves the
        // "never matched" sentinel with an impossible index so that the host-visible result is empty, and the global flag
        uint2* rp = p.rowpart + (size_t)jb * p.split_n1;
        uint2* cp = p.colstat + (size_t)jb * p.split_range;
        // (match.cu's merge / finalize kernels read the second entry as a key: the distance goes to its distance field)
        for (int i = et; i < n1; i += kEpiWarps * 32) rp[i] = good ? make_uint2(rk[i], (uint32_t)rd[i] << kShift) : make_uint2(kSent, kSent);
        for (int j = et; j < n2; j += kEpiWarps * 32) cp[j] = good ? make_uint2(ck[j], (uint32_t)cd[j] << kShift) : make_uint2(kSent, kSent);
      } else
      // ---- ratio test both ways, mutual check, compaction in ascending row order (as match_fused_kernel) ----
      if (good) {
        for (int base = 0; base < n1; base += kEpiWarps * 32) {
          const int i = base + et;
          bool hit = false;
          int j = 0;
          if (i < n1) {
            const uint32_t R = rk[i];
            if (passes(R >> kShift, rd[i], p.threshold, p.ratio)) {
              j = R & kIdxMask;
              const uint32_t C = ck[j];
              hit = passes(C >> kShift, cd[j], p.threshold, p.ratio) && (int)(C & kIdxMask) == i;
            }
          }
          const unsigned bal = __ballot_sync(0xFFFFFFFFu, hit);
          if (lane == 0) s.warp_sums[ew] = __popc(bal);
          epi_barrier();
          int off = s.running, tot = 0;
          for (int w = 0; w < kEpiWarps; ++w) {
            if (w < ew) off += s.warp_sums[w];
            tot += s.warp_sums[w];
          }
          const int pos = off + __popc(bal & ((1u << lane) - 1));
          if (hit && pos < p.cap) {
            job.pairs[2 * pos] = i;
            job.pairs[2 * pos + 1] = j;
          }
          epi_barrier();
          if (et == 0) s.running += tot;
          epi_barrier();
        }
        if (et == 0) {
          *job.count = s.running;
          s.running = 0;
        }
      } else if (et == 0) {
        *job.count = -2;  // pipeline failure (see mbar_wait)
      }
      epi_barrier();
      if (et == 0) MMA_ACC(11);
    }
  }

  tc_fence_before();
  __syncthreads();
#ifdef VSLAM_MMA_PROFILE
  if (blockIdx.x == 0) {
    for (int i = threadIdx.x; i < 32; i += kThreads) g_mma_clk[i] += s.prof_clk[i];
    for (int i = threadIdx.x; i < 128 * 12; i += kThreads) g_mma_trace[i / 12][i % 12] = s.prof_trace[i / 12][i % 12];
  }
#endif
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512) : "memory");
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

// 2-D map over an expanded descriptor array [rows][256 B], box = one K half of an operand tile (128 B x 128 rows)
int make_desc_tmap(vslam_ctx* ctx, const uint8_t* expanded, long long rows, CUtensorMap* tm) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return vs_fail(ctx, VSLAM_E_CUDA, "cuTensorMapEncodeTiled not available from the driver");
  const cuuint64_t dims[2] = {256, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {256};
  const cuuint32_t box[2] = {128, (cuuint32_t)kTile};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<uint8_t*>(expanded), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[128];
    snprintf(buf, sizeof buf, "cuTensorMapEncodeTiled (descriptors) failed: %d", (int)r);
    return vs_fail(ctx, VSLAM_E_CUDA, buf);
  }
  return VSLAM_OK;
}

}  // namespace

size_t match_mma_smem_bytes(int max_n1, int max_n2) {
  return sizeof(MmaSmem) + 1024 + (size_t)(((max_n1 + 1) & ~1) + 2 * ((max_n2 + 1) & ~1)) * 6;
}

bool match_mma_fits(int max_n1, int max_n2) {
  return match_mma_smem_bytes(max_n1, max_n2) <= 227 * 1024 && max_n1 <= kMaxIndex && max_n2 <= kMaxIndex;
}

int launch_expand_desc(vslam_ctx* ctx, const uint32_t* desc, long long rows, int rows_per_set, const int32_t* n_valid,
                       uint8_t* expanded) {
  if (rows <= 0) return VSLAM_OK;
  const long long threads = rows * 16;
  const long long blocks = (threads + 255) / 256;
  if (blocks > 0x7FFFFFFFLL) return vs_fail(ctx, VSLAM_E_LIMIT, "descriptor array too large for one expansion launch");
  expand_desc_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(desc, rows, rows_per_set, n_valid, expanded);
  ctx->launches++;
  VS_CUDA(ctx, cudaGetLastError());
  return VSLAM_OK;
}

int launch_match_jobs_mma(vslam_ctx* ctx, const MatchJob* jobs_dev, int n_jobs, int max_n1, int max_n2, int threshold,
                          double ratio, int cap, const MatchExpanded& x1, const MatchExpanded& x2) {
  if (n_jobs <= 0) return VSLAM_OK;
  if (!match_mma_fits(max_n1, max_n2)) return vs_fail(ctx, VSLAM_E_LIMIT, "job too large for the tensor-core matcher");
  if (x1.rows > 0x7FFFFFFFLL || x2.rows > 0x7FFFFFFFLL) return vs_fail(ctx, VSLAM_E_LIMIT, "descriptor array beyond 2^31 rows");
  CUtensorMap tm1, tm2;
  int rc = make_desc_tmap(ctx, x1.expanded, x1.rows, &tm1);
  if (!rc) rc = make_desc_tmap(ctx, x2.expanded, x2.rows, &tm2);
  if (rc) return rc;
  MmaParams p;
  p.jobs = jobs_dev;
  p.n_jobs = n_jobs;
  p.base1 = x1.packed;
  p.base2 = x2.packed;
  p.threshold = threshold;
  p.ratio = ratio;
  p.cap = cap;
  p.max_n1 = max_n1;
  p.max_n2 = max_n2;
  p.key_mul = -64; p.key_one = 1; p.key_neg1 = -1;
  p.split_range = 0; p.split_n1 = p.split_n2 = 0;
  p.rowpart = p.colstat = nullptr;
  const size_t smem = match_mma_smem_bytes(max_n1, max_n2);
  VS_CUDA(ctx, cudaFuncSetAttribute(match_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int grid = n_jobs < ctx->num_sms ? n_jobs : ctx->num_sms;
  match_mma_kernel<<<grid, kThreads, smem, ctx->stream>>>(tm1, tm2, p);
  ctx->launches++;
  VS_CUDA(ctx, cudaGetLastError());
  return VSLAM_OK;
}

#ifdef VSLAM_MMA_PROFILE
extern "C" int vslam_debug_mma_trace(long long* out1536) {
  return cudaMemcpyFromSymbol(out1536, g_mma_trace, sizeof(long long) * 1536) == cudaSuccess ? 0 : -1;
}
extern "C" int vslam_debug_mma_clocks(long long* out32, int reset) {
  if (cudaMemcpyFromSymbol(out32, g_mma_clk, sizeof(long long) * 32) != cudaSuccess) return -1;
  if (reset) {
    long long z[32] = {};
    if (cudaMemcpyToSymbol(g_mma_clk, z, sizeof z) != cudaSuccess) return -1;
  }
  return 0;
}
#endif

// One large job with its columns split over the SMs: fills rowpart [nsplit][n1] and colstat [n2] in the layout of
// match.cu's column-split path (merge_rows_kernel / finalize_large_kernel take it from there).
int match_mma_split_range(const vslam_ctx* ctx, int n2) {
  int range = (n2 + ctx->num_sms - 1) / ctx->num_sms;
  range = ((range + kTile - 1) / kTile) * kTile;
  return range < kTile ? kTile : range;
}

int launch_match_split_mma(vslam_ctx* ctx, const uint8_t* x1, int n1, const uint8_t* x2, int n2, int range, uint2* rowpart,
                           uint2* colstat) {
  const int nsplit = (n2 + range - 1) / range;
  if (!match_mma_fits(n1, range)) return vs_fail(ctx, VSLAM_E_LIMIT, "job too large for the tensor-core matcher");
  CUtensorMap tm1, tm2;
  int rc = make_desc_tmap(ctx, x1, n1, &tm1);
  if (!rc) rc = make_desc_tmap(ctx, x2, n2, &tm2);
  if (rc) return rc;
  MmaParams p{};
  p.n_jobs = nsplit;
  p.max_n1 = n1;
  p.max_n2 = range;
  p.key_mul = -64; p.key_one = 1; p.key_neg1 = -1;
  p.split_range = range; p.split_n1 = n1; p.split_n2 = n2;
  p.rowpart = rowpart;
  p.colstat = colstat;
  const size_t smem = match_mma_smem_bytes(n1, range);
  VS_CUDA(ctx, cudaFuncSetAttribute(match_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int grid = nsplit < ctx->num_sms ? nsplit : ctx->num_sms;
  match_mma_kernel<<<grid, kThreads, smem, ctx->stream>>>(tm1, tm2, p);
  ctx->launches++;
  VS_CUDA(ctx, cudaGetLastError());
  return VSLAM_OK;
}

int match_mma_timeouts(int* flag) {
  return cudaMemcpyFromSymbol(flag, g_mma_timeout, sizeof(int)) == cudaSuccess ? 0 : -1;
}
