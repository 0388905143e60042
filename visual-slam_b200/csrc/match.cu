// match.cu -- brute-force Hamming matching with threshold, best/second-best ratio test and the
// mutual (cross) check, for sm_100a.
//
// Semantics: visnav::matchDescriptors + isPQiffQP (reference include/visnav/keypoints.h:323-369,
// 278-313).  For a query q over a set D, T(q, D) = (m1, a1, m2): m1 the minimum distance, a1 the
// lowest index attaining it, m2 the second order statistic of the distances, all starting from
// 256 / 0 / 256 (strict '<' updates, :338-354).  pass(T) := m1 < threshold and
// !((double)m2 < m1 * dist_2_best) (:355-357).  (i, j) is a match iff T(d1[i], d2) passes,
// j = a1, T(d2[j], d1) passes and its a1 == i.  Matches come out in ascending i (:331, :361).
//
// Design (B200): the job is integer work on the POPC pipe, not a tensor-core contraction.
//  * One CTA owns one job (or one column range of a large job) so that row AND column
//    statistics of the N1 x N2 distance matrix are produced by ONE pass over the matrix:
//    each distance is computed once and feeds both directions (the reference rescans
//    descriptors_1 for every accepted row).
//  * Rows live in registers: 8 warps x 32 lanes x 6 rows = 1536 rows per pass, 8 words each.
//  * Columns stream through shared memory in 64-descriptor (2 KB) stages filled by 1-D bulk
//    TMA copies (cp.async.bulk + mbarrier) and are read as warp-wide broadcasts.
//  * A (distance, index) pair is one packed key (d << 20 | idx): min() gives the first-index
//    tie-break, and (min, second min) of keys gives (m1, a1, m2) with three integer min/max.
//  * Column statistics: per column each lane folds its rows, the warp folds lanes with two
//    redux.sync.min, and 64 threads fold the 8 warps per stage into shared memory.
//  * With VSLAM_OPT_MATCH_CSA the eight XOR words are compressed by carry-save adders
//    (LOP3 on the ALU pipe) to four POPCs of weight 1,1,2,4: the POPC pipe (16 lanes/clk/SM)
//    is the bottleneck, the ALU pipe has slack.
//  * The ratio test, mutual check and ordered compaction run in the same CTA.
#include "common.cuh"

namespace {

constexpr int kWarps = 8;
constexpr int kThreads = kWarps * 32;
constexpr int kR = 6;                             // row groups (of 32 rows) per warp per pass
constexpr int kRowsPerPass = kWarps * 32 * kR;    // 1536
constexpr int kChunk = 64;                        // descriptors per shared-memory stage
constexpr int kStages = 3;
constexpr uint32_t kShift = 20;                   // key = distance << 20 | index
constexpr uint32_t kIdxMask = (1u << kShift) - 1;
constexpr uint32_t kSent = 256u << kShift;        // (256, 0): the reference's initial best
constexpr uint32_t kInvalidRow = 0x60000000u;     // added to keys of rows >= n1: above any real key
constexpr int kMaxIndex = 1 << kShift;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// 1-D bulk TMA copy global -> shared, completion on an mbarrier (SASS: UBLKCP).
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

__device__ __forceinline__ uint32_t xor3(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d;
  asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ uint32_t maj3(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d;
  asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

// 256-bit Hamming distance of a (8 words in registers) and b (two uint4).
template <bool kCsa>
__device__ __forceinline__ uint32_t hamming256(const uint32_t (&a)[8], const uint4& b0, const uint4& b1) {
  const uint32_t x0 = a[0] ^ b0.x, x1 = a[1] ^ b0.y, x2 = a[2] ^ b0.z, x3 = a[3] ^ b0.w;
  const uint32_t x4 = a[4] ^ b1.x, x5 = a[5] ^ b1.y, x6 = a[6] ^ b1.z, x7 = a[7] ^ b1.w;
  if (kCsa) {
    // Harley-Seal style compression: 8 words -> ones {s3, x7}, twos {t}, fours {f}.
    const uint32_t s1 = xor3(x0, x1, x2), c1 = maj3(x0, x1, x2);
    const uint32_t s2 = xor3(x3, x4, x5), c2 = maj3(x3, x4, x5);
    const uint32_t s3 = xor3(s1, s2, x6), c3 = maj3(s1, s2, x6);
    // (stopping one adder earlier -- 5 POPC, 14 LOP3 -- measured slower: 5.94e11 vs 6.14e11 distances/s)
    const uint32_t t = xor3(c1, c2, c3), f = maj3(c1, c2, c3);
    return (__popc(s3) + __popc(x7)) + 2u * __popc(t) + 4u * __popc(f);
  } else {
    return (__popc(x0) + __popc(x1) + __popc(x2)) + (__popc(x3) + __popc(x4) + __popc(x5)) +
           (__popc(x6) + __popc(x7));
  }
}

// One shared-memory stage of columns against the first kNr row groups of a warp.
template <bool kCsa, int kNr>
__device__ __forceinline__ void scan_chunk(const uint32_t (&a)[kR][8], uint32_t (&m1)[kR], uint32_t (&m2)[kR],
                                           const uint32_t (&rid)[kR], const uint4* bs, int nc, uint32_t jbase,
                                           uint2* part, int lane) {
#pragma unroll 2
  for (int c = 0; c < nc; ++c) {
    const uint4 b0 = bs[2 * c], b1 = bs[2 * c + 1];
    uint32_t c1 = 0xFFFFFFFFu, c2 = 0xFFFFFFFFu;
    const uint32_t j = jbase + c;
#pragma unroll
    for (int r = 0; r < kNr; ++r) {
      const uint32_t d = hamming256<kCsa>(a[r], b0, b1);
      const uint32_t kr = (d << kShift) + j;
      const uint32_t t = max(m1[r], kr);
      m1[r] = min(m1[r], kr);
      m2[r] = min(m2[r], t);
      const uint32_t kc = (d << kShift) + rid[r];
      const uint32_t u = max(c1, kc);
      c1 = min(c1, kc);
      c2 = min(c2, u);
    }
    // fold the 32 lanes: smallest key, then the smallest of (owner's second, others' first)
    const uint32_t g1 = __reduce_min_sync(0xFFFFFFFFu, c1);
    const uint32_t g2 = __reduce_min_sync(0xFFFFFFFFu, c1 == g1 ? c2 : c1);
    if (lane == 0) part[c] = make_uint2(g1, g2);
  }
}

struct ScanSmem {
  uint4* bs;        // [kStages][kChunk][2]
  uint2* part;      // [2][kWarps][kChunk]
  uint64_t* full;   // [kStages]
};

// Scans rows [0, n1) of A against columns [0, ncols) of B (B already offset to the column
// range).  rs[i] receives row i's (key1, key2) with column indices relative to the range; cs[j]
// (shared memory, pre-set to kSent) is folded with every row.  All threads of the CTA call this.
template <bool kCsa>
__device__ void scan_rows_cols(const uint4* __restrict__ A, int n1, const uint4* __restrict__ B, int ncols,
                               uint2* rs, uint2* cs, const ScanSmem& sm) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nchunks = (ncols + kChunk - 1) / kChunk;
  const int npass = (n1 + kRowsPerPass - 1) / kRowsPerPass;
  const int total = npass * nchunks;

  auto issue = [&](int it) {
    const int chunk = it % nchunks;
    const int st = it % kStages;
    const int nc = min(kChunk, ncols - chunk * kChunk);
    mbar_expect_tx(&sm.full[st], nc * 32);
    bulk_g2s(sm.bs + st * kChunk * 2, B + (size_t)chunk * kChunk * 2, nc * 32, &sm.full[st]);
  };
  if (tid == 0) {
    for (int it = 0; it < min(kStages, total); ++it) issue(it);
  }

  int it = 0;
  for (int pass = 0; pass < npass; ++pass) {
    const int pbase = pass * kRowsPerPass;
    uint32_t a[kR][8], m1[kR], m2[kR], rid[kR];
    int nr = 0;
#pragma unroll
    for (int r = 0; r < kR; ++r) {
      const int gbase = pbase + (r * kWarps + warp) * 32;
      const int row = gbase + lane;
      if (gbase < n1) nr = r + 1;
      if (row < n1) {
        const uint4 lo = __ldg(A + (size_t)row * 2), hi = __ldg(A + (size_t)row * 2 + 1);
        a[r][0] = lo.x; a[r][1] = lo.y; a[r][2] = lo.z; a[r][3] = lo.w;
        a[r][4] = hi.x; a[r][5] = hi.y; a[r][6] = hi.z; a[r][7] = hi.w;
        rid[r] = row;
      } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) a[r][k] = 0;
        rid[r] = kInvalidRow;
      }
      m1[r] = kSent;
      m2[r] = kSent;
    }

    for (int chunk = 0; chunk < nchunks; ++chunk, ++it) {
      const int st = it % kStages;
      mbar_wait(&sm.full[st], (it / kStages) & 1);
      const uint4* bs = sm.bs + st * kChunk * 2;
      const int nc = min(kChunk, ncols - chunk * kChunk);
      const uint32_t jbase = chunk * kChunk;
      uint2* part = sm.part + ((it & 1) * kWarps + warp) * kChunk;
      if (nr > 0) {
        // the row-group count is warp-uniform: dispatch once per stage to a loop whose row dimension is a
        // compile-time constant (a guard inside the loop costs a branch per distance and keeps the
        // compiler from interleaving the independent rows)
        switch (nr) {
          case 1: scan_chunk<kCsa, 1>(a, m1, m2, rid, bs, nc, jbase, part, lane); break;
          case 2: scan_chunk<kCsa, 2>(a, m1, m2, rid, bs, nc, jbase, part, lane); break;
          case 3: scan_chunk<kCsa, 3>(a, m1, m2, rid, bs, nc, jbase, part, lane); break;
          case 4: scan_chunk<kCsa, 4>(a, m1, m2, rid, bs, nc, jbase, part, lane); break;
          case 5: scan_chunk<kCsa, 5>(a, m1, m2, rid, bs, nc, jbase, part, lane); break;
          default: scan_chunk<kCsa, kR>(a, m1, m2, rid, bs, nc, jbase, part, lane); break;
        }
      } else if (lane == 0) {
        for (int c = 0; c < nc; ++c) part[c] = make_uint2(0xFFFFFFFFu, 0xFFFFFFFFu);
      }
      __syncthreads();  // stage st consumed; part[it & 1] complete
      if (tid == 0 && it + kStages < total) issue(it + kStages);
      if (tid < nc) {
        uint2 acc = cs[jbase + tid];
        const uint2* p = sm.part + (it & 1) * kWarps * kChunk + tid;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) {
          const uint2 v = p[w * kChunk];
          // merge (acc.x <= acc.y) with (v.x <= v.y)
          const uint32_t hi = max(acc.x, v.x);
          acc.x = min(acc.x, v.x);
          acc.y = min(min(acc.y, v.y), hi);
        }
        cs[jbase + tid] = acc;
      }
    }
#pragma unroll
    for (int r = 0; r < kR; ++r) {
      const int row = pbase + (r * kWarps + warp) * 32 + lane;
      if (row < n1) rs[row] = make_uint2(m1[r], m2[r]);
    }
  }
  __syncthreads();
}

__device__ __forceinline__ bool passes(uint32_t m1, uint32_t m2, int threshold, double ratio) {
  // keypoints.h:355-357: best1_d >= threshold -> no; best2_d < best1_d * dist_2_best -> no.
  if ((int)m1 >= threshold) return false;
  return !((double)(int)m2 < __dmul_rn((double)(int)m1, ratio));
}

struct MatchParams {
  const MatchJob* jobs;
  int threshold;
  double ratio;
  int cap;
  int max_n1, max_n2;
};

__device__ __forceinline__ ScanSmem carve(unsigned char* smem, unsigned char** rest) {
  ScanSmem s;
  s.bs = reinterpret_cast<uint4*>(smem);
  smem += kStages * kChunk * 32;
  s.part = reinterpret_cast<uint2*>(smem);
  smem += 2 * kWarps * kChunk * sizeof(uint2);
  s.full = reinterpret_cast<uint64_t*>(smem);
  smem += 64;
  *rest = smem;
  return s;
}
constexpr int kScanSmemBytes = kStages * kChunk * 32 + 2 * kWarps * kChunk * 8 + 64;

// One CTA = one whole matchDescriptors call.
template <bool kCsa>
__global__ void __launch_bounds__(kThreads, 2) match_fused_kernel(const MatchParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char* rest;
  const ScanSmem sm = carve(smem_raw, &rest);
  uint2* cs = reinterpret_cast<uint2*>(rest);
  uint2* rs = cs + p.max_n2;
  __shared__ int warp_sums[kWarps];
  __shared__ int running;

  const MatchJob job = p.jobs[blockIdx.x];
  const int n1 = job.n1p ? *job.n1p : job.n1;
  const int n2 = job.n2p ? *job.n2p : job.n2;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (n1 <= 0 || n2 <= 0 || n1 > p.max_n1 || n2 > p.max_n2) {
    if (tid == 0) *job.count = (n1 > p.max_n1 || n2 > p.max_n2) ? -1 : 0;
    return;
  }
  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) mbar_init(&sm.full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    running = 0;
  }
  for (int j = tid; j < n2; j += kThreads) cs[j] = make_uint2(kSent, kSent);
  __syncthreads();

  scan_rows_cols<kCsa>(reinterpret_cast<const uint4*>(job.d1), n1, reinterpret_cast<const uint4*>(job.d2), n2,
                       rs, cs, sm);

  // ratio test both ways, mutual check, compaction in ascending row order
  for (int base = 0; base < n1; base += kThreads) {
    const int i = base + tid;
    bool ok = false;
    int j = 0;
    if (i < n1) {
      const uint2 R = rs[i];
      if (passes(R.x >> kShift, R.y >> kShift, p.threshold, p.ratio)) {
        j = R.x & kIdxMask;
        const uint2 C = cs[j];
        ok = passes(C.x >> kShift, C.y >> kShift, p.threshold, p.ratio) && (int)(C.x & kIdxMask) == i;
      }
    }
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, ok);
    if (lane == 0) warp_sums[warp] = __popc(bal);
    __syncthreads();
    int off = running;
    for (int w = 0; w < warp; ++w) off += warp_sums[w];
    int tot = 0;
    for (int w = 0; w < kWarps; ++w) tot += warp_sums[w];
    const int pos = off + __popc(bal & ((1u << lane) - 1));
    if (ok && pos < p.cap) {
      job.pairs[2 * pos] = i;
      job.pairs[2 * pos + 1] = j;
    }
    __syncthreads();
    if (tid == 0) running += tot;
    __syncthreads();
  }
  if (tid == 0) *job.count = running;
}

// ---- large jobs: columns split across CTAs ----------------------------------------------------
struct LargeParams {
  const uint32_t* d1;
  const uint32_t* d2;
  int n1, n2;
  int range;        // columns per CTA (multiple of kChunk)
  uint2* rowpart;   // [nsplit][n1]
  uint2* colstat;   // [n2]
};

template <bool kCsa>
__global__ void __launch_bounds__(kThreads, 2) match_split_kernel(const LargeParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char* rest;
  const ScanSmem sm = carve(smem_raw, &rest);
  uint2* cs = reinterpret_cast<uint2*>(rest);
  const int tid = threadIdx.x;
  const int c0 = blockIdx.x * p.range;
  const int ncols = min(p.range, p.n2 - c0);
  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) mbar_init(&sm.full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int j = tid; j < ncols; j += kThreads) cs[j] = make_uint2(kSent, kSent);
  __syncthreads();
  scan_rows_cols<kCsa>(reinterpret_cast<const uint4*>(p.d1), p.n1,
                       reinterpret_cast<const uint4*>(p.d2) + (size_t)c0 * 2, ncols,
                       p.rowpart + (size_t)blockIdx.x * p.n1, cs, sm);
  for (int j = tid; j < ncols; j += kThreads) p.colstat[c0 + j] = cs[j];
}

// Merge the per-range row partials: rowstat[i] = (m1, a1 global, m2).
__global__ void merge_rows_kernel(const uint2* __restrict__ rowpart, int n1, int nsplit, int range,
                                  int3* __restrict__ rowstat) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n1) return;
  int m1 = 256, a1 = 0, m2 = 256;
  for (int s = 0; s < nsplit; ++s) {
    const uint2 v = rowpart[(size_t)s * n1 + i];
    const int d1 = v.x >> kShift, d2 = v.y >> kShift;
    if (d1 < m1) {  // strictly better: earlier ranges keep ties (lowest index wins)
      m2 = min(m1, d2);
      m1 = d1;
      a1 = s * range + (int)(v.x & kIdxMask);
    } else {
      m2 = min(m2, d1);
    }
  }
  rowstat[i] = make_int3(m1, a1, m2);
}

__global__ void __launch_bounds__(1024) finalize_large_kernel(const int3* __restrict__ rowstat, const uint2* __restrict__ colstat,
                                      int n1, int threshold, double ratio, int32_t* pairs, int32_t* count,
                                      int cap) {
  __shared__ int warp_sums[32];
  __shared__ int running;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) running = 0;
  __syncthreads();
  for (int base = 0; base < n1; base += 1024) {
    const int i = base + tid;
    bool ok = false;
    int j = 0;
    if (i < n1) {
      const int3 R = rowstat[i];
      if (passes(R.x, R.z, threshold, ratio)) {
        j = R.y;
        const uint2 C = colstat[j];
        ok = passes(C.x >> kShift, C.y >> kShift, threshold, ratio) && (int)(C.x & kIdxMask) == i;
      }
    }
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
      pairs[2 * pos + 1] = j;
    }
    __syncthreads();
    if (tid == 0) running += tot;
    __syncthreads();
  }
  if (tid == 0) *count = running;
}

// ---- few jobs: every job's columns are split across CTAs (grid = column ranges x jobs) ----------------
// One CTA per job leaves most SMs idle when a batch holds fewer jobs than the GPU has CTA slots (a single stereo
// pair: 0.44 ms on one SM).  Here job blockIdx.y hands column range blockIdx.x to a CTA; a second launch merges the
// per-range row partials and finalises each job (one CTA per job).  Counts may live on the device (MatchJob).
struct SplitJobsParams {
  const MatchJob* jobs;
  int max_n1, max_n2;
  int range, nsplit;  // columns per CTA (multiple of kChunk), ranges per job
  uint2* rowpart;     // [n_jobs][nsplit][max_n1]
  uint2* colstat;     // [n_jobs][max_n2]
  int threshold;
  double ratio;
  int cap;
};

template <bool kCsa>
__global__ void __launch_bounds__(kThreads, 2) match_split_jobs_kernel(const SplitJobsParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const MatchJob job = p.jobs[blockIdx.y];
  const int n1 = job.n1p ? *job.n1p : job.n1;
  const int n2 = job.n2p ? *job.n2p : job.n2;
  const int c0 = blockIdx.x * p.range;
  if (n1 <= 0 || n2 <= 0 || n1 > p.max_n1 || n2 > p.max_n2 || c0 >= n2) return;  // uniform per CTA
  unsigned char* rest;
  const ScanSmem sm = carve(smem_raw, &rest);
  uint2* cs = reinterpret_cast<uint2*>(rest);
  const int tid = threadIdx.x;
  const int ncols = min(p.range, n2 - c0);
  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) mbar_init(&sm.full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int j = tid; j < ncols; j += kThreads) cs[j] = make_uint2(kSent, kSent);
  __syncthreads();
  scan_rows_cols<kCsa>(reinterpret_cast<const uint4*>(job.d1), n1,
                       reinterpret_cast<const uint4*>(job.d2) + (size_t)c0 * 2, ncols,
                       p.rowpart + ((size_t)blockIdx.y * p.nsplit + blockIdx.x) * p.max_n1, cs, sm);
  uint2* colstat = p.colstat + (size_t)blockIdx.y * p.max_n2;
  for (int j = tid; j < ncols; j += kThreads) colstat[c0 + j] = cs[j];
}

__global__ void __launch_bounds__(1024) finalize_split_jobs_kernel(const SplitJobsParams p) {
  __shared__ int warp_sums[32];
  __shared__ int running;
  const MatchJob job = p.jobs[blockIdx.x];
  const int n1 = job.n1p ? *job.n1p : job.n1;
  const int n2 = job.n2p ? *job.n2p : job.n2;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (n1 <= 0 || n2 <= 0 || n1 > p.max_n1 || n2 > p.max_n2) {
    if (tid == 0) *job.count = (n1 > p.max_n1 || n2 > p.max_n2) ? -1 : 0;
    return;
  }
  const int used = (n2 + p.range - 1) / p.range;
  const uint2* rowpart = p.rowpart + (size_t)blockIdx.x * p.nsplit * p.max_n1;
  const uint2* colstat = p.colstat + (size_t)blockIdx.x * p.max_n2;
  if (tid == 0) running = 0;
  __syncthreads();
  for (int base = 0; base < n1; base += 1024) {
    const int i = base + tid;
    bool ok = false;
    int j = 0;
    if (i < n1) {
      int m1 = 256, a1 = 0, m2 = 256;
      for (int s = 0; s < used; ++s) {
        const uint2 v = rowpart[(size_t)s * p.max_n1 + i];
        const int d1 = v.x >> kShift, d2 = v.y >> kShift;
        if (d1 < m1) {  // strictly better: earlier ranges keep ties (lowest index wins)
          m2 = min(m1, d2);
          m1 = d1;
          a1 = s * p.range + (int)(v.x & kIdxMask);
        } else {
          m2 = min(m2, d1);
        }
      }
      if (passes(m1, m2, p.threshold, p.ratio)) {
        j = a1;
        const uint2 C = colstat[j];
        ok = passes(C.x >> kShift, C.y >> kShift, p.threshold, p.ratio) && (int)(C.x & kIdxMask) == i;
      }
    }
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, ok);
    if (lane == 0) warp_sums[warp] = __popc(bal);
    __syncthreads();
    int off = running, tot = 0;
    for (int w = 0; w < 32; ++w) {
      if (w < warp) off += warp_sums[w];
      tot += warp_sums[w];
    }
    const int pos = off + __popc(bal & ((1u << lane) - 1));
    if (ok && pos < p.cap) {
      job.pairs[2 * pos] = i;
      job.pairs[2 * pos + 1] = j;
    }
    __syncthreads();
    if (tid == 0) running += tot;
    __syncthreads();
  }
  if (tid == 0) *job.count = running;
}

__global__ void build_store_jobs_kernel(const int2* __restrict__ jobs, int n_jobs, const uint32_t* desc,
                                       const int64_t* __restrict__ offsets, const int32_t* __restrict__ counts,
                                       int n_sets, int32_t* pairs, int32_t* out_counts, int cap, MatchJob* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_jobs) return;
  int2 j = jobs[i];
  MatchJob m;
  const bool ok = j.x >= 0 && j.x < n_sets && j.y >= 0 && j.y < n_sets;
  if (!ok) j = make_int2(0, 0);
  m.d1 = desc + offsets[j.x] * 8;
  m.d2 = desc + offsets[j.y] * 8;
  m.n1p = nullptr;
  m.n2p = nullptr;
  m.n1 = ok ? counts[j.x] : 0;
  m.n2 = ok ? counts[j.y] : 0;
  m.pairs = pairs + (size_t)i * cap * 2;
  m.count = out_counts + i;
  out[i] = m;
}

// ---- POPC microbenchmark ----------------------------------------------------------------------
__global__ void __launch_bounds__(1024) popc_peak_kernel(uint32_t* out, int iters, uint32_t seed) {
  uint32_t acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] = seed * (threadIdx.x + 1) + k * 0x9E3779B9u;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = __popc(acc[k]) ^ seed;  // one POPC + one ALU op, 8 chains
  }
  uint32_t s = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += acc[k];
  if (s == 0xDEADBEEFu) out[0] = s;
}

}  // namespace

static size_t fused_smem_bytes(int max_n1, int max_n2) {
  return (size_t)kScanSmemBytes + (size_t)(max_n1 + max_n2) * sizeof(uint2);
}

int launch_match_jobs(vslam_ctx* ctx, const MatchJob* jobs_dev, int n_jobs, int max_n1, int max_n2,
                      int threshold, double ratio, int cap, DevBuf* rowpart_buf, DevBuf* colstat_buf) {
  if (!rowpart_buf) rowpart_buf = &ctx->rowpart;  // callers that run on their own stream bring their own scratch
  if (!colstat_buf) colstat_buf = &ctx->colstat;
  if (n_jobs <= 0) return VSLAM_OK;
  if (max_n1 > kMaxIndex || max_n2 > kMaxIndex) return vs_fail(ctx, VSLAM_E_LIMIT, "descriptor set larger than 2^20");
  const size_t smem = fused_smem_bytes(max_n1, max_n2);
  if (smem > 200 * 1024) return vs_fail(ctx, VSLAM_E_LIMIT, "job too large for the fused matcher");
  // Fewer jobs than CTA slots and jobs worth splitting: spread every job's columns over the idle SMs
  // (one stereo pair: 0.44 ms on one SM -> 0.05 ms on all of them).
  const int slots = ctx->num_sms * 2;
  if (ctx->match_split_small && n_jobs * 2 <= ctx->num_sms && (long long)max_n1 * max_n2 > 256LL * 256LL) {
    const int per_job = slots / n_jobs;  // >= 4
    int range = (max_n2 + per_job - 1) / per_job;
    range = ((range + kChunk - 1) / kChunk) * kChunk;
    if (range > 8192) range = 8192;
    SplitJobsParams sp;
    sp.jobs = jobs_dev;
    sp.max_n1 = max_n1; sp.max_n2 = max_n2;
    sp.range = range;
    sp.nsplit = (max_n2 + range - 1) / range;
    VS_CUDA(ctx, rowpart_buf->reserve((size_t)n_jobs * sp.nsplit * max_n1 * sizeof(uint2)));
    VS_CUDA(ctx, colstat_buf->reserve((size_t)n_jobs * max_n2 * sizeof(uint2)));
    sp.rowpart = rowpart_buf->as<uint2>();
    sp.colstat = colstat_buf->as<uint2>();
    sp.threshold = threshold; sp.ratio = ratio; sp.cap = cap;
    const size_t ssm = (size_t)kScanSmemBytes + (size_t)range * sizeof(uint2);
    auto sk = ctx->match_csa ? match_split_jobs_kernel<true> : match_split_jobs_kernel<false>;
    VS_CUDA(ctx, cudaFuncSetAttribute(sk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssm));
    sk<<<dim3(sp.nsplit, n_jobs), kThreads, ssm, ctx->stream>>>(sp);
    finalize_split_jobs_kernel<<<n_jobs, 1024, 0, ctx->stream>>>(sp);
    ctx->launches += 2;
    VS_CUDA(ctx, cudaGetLastError());
    return VSLAM_OK;
  }
  MatchParams p;
  p.jobs = jobs_dev;
  p.threshold = threshold;
  p.ratio = ratio;
  p.cap = cap;
  p.max_n1 = max_n1;
  p.max_n2 = max_n2;
  auto kern = ctx->match_csa ? match_fused_kernel<true> : match_fused_kernel<false>;
  VS_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<n_jobs, kThreads, smem, ctx->stream>>>(p);
  ctx->launches++;
  VS_CUDA(ctx, cudaGetLastError());
  return VSLAM_OK;
}

bool match_fits_fused(int n1, int n2) { return fused_smem_bytes(n1, n2) <= 200 * 1024 && n1 <= kMaxIndex && n2 <= kMaxIndex; }

// columns per CTA of the split path: aim at two CTAs per SM, keep ranges a multiple of the stage size
static int split_range(const vslam_ctx* ctx, int n2) {
  const int target = ctx->num_sms * 2;
  int range = (n2 + target - 1) / target;
  range = ((range + kChunk - 1) / kChunk) * kChunk;
  if (range < kChunk) range = kChunk;
  if (range > 8192) range = 8192;
  return range;
}

// Columns per CTA of a large job and which scan runs it: the tensor-core one (match_mma.cu) when there is enough work
// to fill the GPU and the query set's statistics fit one CTA's shared memory, else the POPC one.
static int large_range(const vslam_ctx* ctx, int n1, int n2, bool* mma) {
  const int mrange = match_mma_split_range(ctx, n2);
  *mma = ctx->match_mma && (long long)n1 * n2 >= (1LL << 23) && n2 >= 4 * ctx->num_sms && match_mma_fits(n1, mrange);
  return *mma ? mrange : split_range(ctx, n2);
}

int launch_match_large(vslam_ctx* ctx, const uint32_t* d1, int n1, const uint32_t* d2, int n2,
                       int threshold, double ratio, int32_t* pairs, int32_t* count, int cap,
                       int32_t* best2_out) {
  if (n1 > kMaxIndex) return vs_fail(ctx, VSLAM_E_LIMIT, "more than 2^20 query descriptors");
  // enough work to fill the GPU and a query set whose statistics fit one CTA's shared memory: the Hamming matrix goes
  // to the tensor cores (match_mma.cu, column ranges over the SMs); the merge / finalize kernels below are shared
  bool mma;
  const int range = large_range(ctx, n1, n2, &mma);
  const int nsplit = (n2 + range - 1) / range;
  VS_CUDA(ctx, ctx->rowpart.reserve((size_t)nsplit * n1 * sizeof(uint2) + (size_t)n1 * sizeof(int3)));
  VS_CUDA(ctx, ctx->colstat.reserve((size_t)n2 * sizeof(uint2)));
  uint2* rowpart = ctx->rowpart.as<uint2>();
  int3* rowstat = reinterpret_cast<int3*>(rowpart + (size_t)nsplit * n1);
  LargeParams p;
  p.d1 = d1; p.d2 = d2; p.n1 = n1; p.n2 = n2; p.range = range;
  p.rowpart = rowpart;
  p.colstat = ctx->colstat.as<uint2>();
  if (mma) {
    VS_CUDA(ctx, ctx->d1x.reserve((size_t)n1 * 256));
    VS_CUDA(ctx, ctx->d2x.reserve((size_t)n2 * 256));
    int rc = launch_expand_desc(ctx, d1, n1, 1, nullptr, ctx->d1x.as<uint8_t>());
    if (!rc) rc = launch_expand_desc(ctx, d2, n2, 1, nullptr, ctx->d2x.as<uint8_t>());
    if (!rc) rc = launch_match_split_mma(ctx, ctx->d1x.as<uint8_t>(), n1, ctx->d2x.as<uint8_t>(), n2, range, rowpart, p.colstat);
    if (rc) return rc;
  } else {
    const size_t smem = (size_t)kScanSmemBytes + (size_t)range * sizeof(uint2);
    auto kern = ctx->match_csa ? match_split_kernel<true> : match_split_kernel<false>;
    VS_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<nsplit, kThreads, smem, ctx->stream>>>(p);
  }
  merge_rows_kernel<<<(n1 + 255) / 256, 256, 0, ctx->stream>>>(rowpart, n1, nsplit, range, rowstat);
  ctx->launches += 2;
  if (best2_out) {
    VS_CUDA(ctx, cudaMemcpyAsync(best2_out, rowstat, (size_t)n1 * sizeof(int3), cudaMemcpyDeviceToDevice, ctx->stream));
  } else {
    finalize_large_kernel<<<1, 1024, 0, ctx->stream>>>(rowstat, p.colstat, n1, threshold, ratio, pairs, count, cap);
    ctx->launches++;
  }
  VS_CUDA(ctx, cudaGetLastError());
  return VSLAM_OK;
}

// ---- DB-sharded matching (BASELINE config 4 across GPUs) -------------------------------------
// One shard of desc2 per GPU, all of desc1 on every GPU: the column statistics T(d2[j], desc1) are complete
// locally; the row statistics are partial.  A shard reports per query i the int4 (m1, a1 + index_offset, m2,
// col_ok) with col_ok = "T(d2[a1], desc1) passes and names i".  Merging shards is the associative operator of
// the row triple (lowest global index wins ties); the winner's col_ok is the mutual check.
__global__ void shard_part_kernel(const int3* __restrict__ rowstat, const uint2* __restrict__ colstat, int n1,
                                  int index_offset, int threshold, double ratio, int4* __restrict__ part) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n1) return;
  const int3 R = rowstat[i];
  const uint2 C = colstat[R.y];
  const bool col_ok = passes(C.x >> kShift, C.y >> kShift, threshold, ratio) && (int)(C.x & kIdxMask) == i;
  part[i] = make_int4(R.x, R.y + index_offset, R.z, col_ok ? 1 : 0);
}

__global__ void __launch_bounds__(1024)
    merge_shards_kernel(const int4* __restrict__ parts, int n_parts, int n1, int threshold, double ratio,
                        int32_t* pairs, int32_t* count, int cap) {
  __shared__ int warp_sums[32];
  __shared__ int running;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) running = 0;
  __syncthreads();
  for (int base = 0; base < n1; base += 1024) {
    const int i = base + tid;
    bool ok = false;
    int j = 0;
    if (i < n1) {
      int4 best = parts[i];
      for (int s = 1; s < n_parts; ++s) {
        const int4 v = parts[(size_t)s * n1 + i];
        const int m2 = min(min(best.z, v.z), max(best.x, v.x));
        if (v.x < best.x || (v.x == best.x && v.y < best.y)) best = v;
        best.z = m2;
      }
      j = best.y;
      ok = passes(best.x, best.z, threshold, ratio) && best.w != 0;
    }
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
      pairs[2 * pos + 1] = j;
    }
    __syncthreads();
    if (tid == 0) running += tot;
    __syncthreads();
  }
  if (tid == 0) *count = running;
}

int launch_match_shard(vslam_ctx* ctx, const uint32_t* d1, int n1, const uint32_t* d2, int n2, int index_offset,
                       int threshold, double ratio, int32_t* part4) {
  // row statistics land in ctx->rowpart (after the partials), column statistics in ctx->colstat
  int rc = launch_match_large(ctx, d1, n1, d2, n2, threshold, ratio, nullptr, nullptr, 0, part4 /*scratch: 12 B/row*/);
  if (rc) return rc;
  bool mma;
  const int range = large_range(ctx, n1, n2, &mma);
  const int nsplit = (n2 + range - 1) / range;
  const int3* rowstat = reinterpret_cast<const int3*>(ctx->rowpart.as<uint2>() + (size_t)nsplit * n1);
  shard_part_kernel<<<(n1 + 255) / 256, 256, 0, ctx->stream>>>(rowstat, ctx->colstat.as<uint2>(), n1, index_offset,
                                                               threshold, ratio, reinterpret_cast<int4*>(part4));
  ctx->launches++;
  VS_CUDA(ctx, cudaGetLastError());
  return VSLAM_OK;
}

int launch_merge_shards(vslam_ctx* ctx, const int32_t* parts4, int n_parts, int n1, int threshold, double ratio,
                        int32_t* pairs, int32_t* count, int cap) {
  merge_shards_kernel<<<1, 1024, 0, ctx->stream>>>(reinterpret_cast<const int4*>(parts4), n_parts, n1, threshold, ratio,
                                                   pairs, count, cap);
  ctx->launches++;
  VS_CUDA(ctx, cudaGetLastError());
  return VSLAM_OK;
}

int launch_build_store_jobs(vslam_ctx* ctx, const int32_t* jobs_i2, int n_jobs, const uint32_t* desc,
                            const int64_t* offsets, const int32_t* counts, int n_sets, int32_t* pairs,
                            int32_t* out_counts, int cap, MatchJob* out) {
  if (n_jobs <= 0) return VSLAM_OK;
  build_store_jobs_kernel<<<(n_jobs + 255) / 256, 256, 0, ctx->stream>>>(
      reinterpret_cast<const int2*>(jobs_i2), n_jobs, desc, offsets, counts, n_sets, pairs, out_counts, cap, out);
  ctx->launches++;
  VS_CUDA(ctx, cudaGetLastError());
  return VSLAM_OK;
}

int launch_popc_peak(vslam_ctx* ctx, double* popc_per_s) {
  VS_CUDA(ctx, ctx->mcnt.reserve(256));
  const int iters = 1 << 14;
  const int blocks = ctx->num_sms * 2;
  cudaEvent_t a = ctx->ev[0], b = ctx->ev[1];
  double best = 0;
  for (int rep = 0; rep < 5; ++rep) {
    VS_CUDA(ctx, cudaEventRecord(a, ctx->stream));
    popc_peak_kernel<<<blocks, 1024, 0, ctx->stream>>>(ctx->mcnt.as<uint32_t>(), iters, 0x12345u + rep);
    VS_CUDA(ctx, cudaEventRecord(b, ctx->stream));
    VS_CUDA(ctx, cudaEventSynchronize(b));
    float ms = 0;
    VS_CUDA(ctx, cudaEventElapsedTime(&ms, a, b));
    const double n = (double)blocks * 1024 * iters * 8;
    if (rep > 0 && n / (ms * 1e-3) > best) best = n / (ms * 1e-3);
  }
  ctx->launches += 5;
  *popc_per_s = best;
  return VSLAM_OK;
}
