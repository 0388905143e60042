// noalloc.cu -- vslam_best2_noalloc: the row scan of visnav::matchDescriptors (reference include/visnav/keypoints.h:331-354,
// the (best distance, lowest index attaining it, second best) triple per query) for callers that may not allocate:
// code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  It has no counterpart in the
// reference.
//
// vslam_best2 (api.cu -> launch_match_large, match.cu) splits the database over column ranges, writes one partial
// (min key, second-min key) per (range, query) into scratch it RESERVES with cudaMalloc, and merges the ranges.  This file
// is the same two-step scan with that scratch in a file-scope device array of fixed size instead:
//     g_na_scratch  [kNaMaxRanges][kNaMaxQueries] uint2 = 64 x 4096 x 8 B = 2 MiB,
// held by every GPU that loads the library until the process exits.  The caller passes DEVICE pointers for the inputs
// and the output and (optionally) the stream; the entry point calls no cudaMalloc / cudaFree / allocating library routine.
//
// In host code the array's NAME is a host-side stand-in, not a device address: the address the kernels take is obtained
// once per device, with that device current, by cudaGetSymbolAddress (na_scratch_for below).
//
// Sharing: every call on a device uses the same array.  Nothing in it has to start at zero -- every slot a call's merge
// kernel reads was written by that call's scan kernel before (same stream, in order), so the contents the previous
// call left are never observed.  Calls that OVERLAP on the GPU (two streams, two host threads, two contexts on one
// device) race on the array and must be serialised by the caller.
#include <mutex>

#include "common.cuh"

constexpr int kNaMaxQueries = 4096;   // rows of the scratch: the largest nq accepted
constexpr int kNaMaxRanges = 64;      // column ranges of the database scanned by different CTAs
constexpr int kNaMaxDb = 1 << 20;     // the key holds the column index in 20 bits (as match.cu)
constexpr int kNaThreads = 128;       // queries per CTA = database rows staged per step
constexpr uint32_t kNaShift = 20;
constexpr uint32_t kNaSent = 256u << kNaShift;  // "no column seen": distance 256, index 0 (keypoints.h:331-333)
constexpr int kNaMaxDevices = 64;

static __device__ __align__(256) uint2 g_na_scratch[(size_t)kNaMaxRanges * kNaMaxQueries];

// one thread = one query (its 256 bits in registers); the CTA walks its column range 128 descriptors at a time through
// shared memory (every thread reads the same staged row: a broadcast).  Keys distance << 20 | column are distinct, so
// (min key, second-min key) is exactly (best1, lowest index attaining it) and the second-best DISTANCE of the
// reference's scan (a duplicate of the best distance at a higher index is the second best, as in keypoints.h:340-350).
static __global__ void __launch_bounds__(kNaThreads)
na_scan_kernel(const uint32_t* __restrict__ q, int nq, const uint32_t* __restrict__ db, int ndb, int range,
               uint2* __restrict__ part) {
  __shared__ uint4 stage[kNaThreads][2];
  const int row = blockIdx.x * kNaThreads + threadIdx.x;
  const int c_begin = blockIdx.y * range;
  const int c_end = min(ndb, c_begin + range);
  uint4 qa = make_uint4(0, 0, 0, 0), qb = qa;
  if (row < nq) {
    qa = reinterpret_cast<const uint4*>(q)[2 * (size_t)row];
    qb = reinterpret_cast<const uint4*>(q)[2 * (size_t)row + 1];
  }
  uint32_t k1 = kNaSent, k2 = kNaSent;
  for (int c0 = c_begin; c0 < c_end; c0 += kNaThreads) {
    const int c = c0 + threadIdx.x;
    __syncthreads();
    if (c < c_end) {
      stage[threadIdx.x][0] = reinterpret_cast<const uint4*>(db)[2 * (size_t)c];
      stage[threadIdx.x][1] = reinterpret_cast<const uint4*>(db)[2 * (size_t)c + 1];
    }
    __syncthreads();
    const int nc = min(kNaThreads, c_end - c0);
    for (int j = 0; j < nc; ++j) {
      const uint4 a = stage[j][0], b = stage[j][1];
      const uint32_t d = __popc(qa.x ^ a.x) + __popc(qa.y ^ a.y) + __popc(qa.z ^ a.z) + __popc(qa.w ^ a.w) +
                         __popc(qb.x ^ b.x) + __popc(qb.y ^ b.y) + __popc(qb.z ^ b.z) + __popc(qb.w ^ b.w);
      const uint32_t key = (d << kNaShift) | (uint32_t)(c0 + j);
      k2 = min(k2, max(k1, key));
      k1 = min(k1, key);
    }
  }
  if (row < nq) part[(size_t)blockIdx.y * kNaMaxQueries + row] = make_uint2(k1, k2);
}

// the associative merge of the ranges' partials: the two smallest of all their keys
static __global__ void na_merge_kernel(const uint2* __restrict__ part, int nq, int nranges, int32_t* __restrict__ out3) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= nq) return;
  uint32_t k1 = kNaSent, k2 = kNaSent;
  for (int r = 0; r < nranges; ++r) {
    const uint2 p = part[(size_t)r * kNaMaxQueries + row];
    k2 = min(min(k2, p.y), max(k1, p.x));
    k1 = min(k1, p.x);
  }
  out3[3 * row] = (int32_t)(k1 >> kNaShift);
  out3[3 * row + 1] = (int32_t)(k1 & ((1u << kNaShift) - 1));
  out3[3 * row + 2] = (int32_t)(k2 >> kNaShift);
}

// set-up, once per device: the device address behind the symbol (the symbol itself is NOT a device pointer in host code)
static std::mutex g_na_mutex;
static uint2* g_na_ptr[kNaMaxDevices];

static cudaError_t na_scratch_for(int device, uint2** out) {
  std::lock_guard<std::mutex> lock(g_na_mutex);
  if (!g_na_ptr[device]) {
    uint2* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_na_scratch);  // `device` is current
    if (e != cudaSuccess) return e;
    g_na_ptr[device] = p;
  }
  *out = g_na_ptr[device];
  return cudaSuccess;
}

extern "C" int vslam_best2_noalloc(vslam_ctx* ctx, const uint32_t* d_q, int nq, const uint32_t* d_db, int ndb,
                                   int32_t* d_out3, void* stream) {
  if (!ctx) return VSLAM_E_ARG;
  if (nq < 0 || ndb < 0 || (nq > 0 && (!d_q || !d_out3)) || (ndb > 0 && !d_db))
    return vs_fail(ctx, VSLAM_E_ARG, "vslam_best2_noalloc: bad arguments");
  if (((uintptr_t)d_q | (uintptr_t)d_db) & 15u)
    return vs_fail(ctx, VSLAM_E_ARG, "vslam_best2_noalloc: d_q / d_db must be 16-byte aligned device pointers");
  if (nq > kNaMaxQueries)
    return vs_fail(ctx, VSLAM_E_LIMIT, "vslam_best2_noalloc: nq exceeds 4096 queries (the fixed 2 MiB device scratch holds "
                                       "64 x 4096 partials); use vslam_best2");
  if (ndb > kNaMaxDb)
    return vs_fail(ctx, VSLAM_E_LIMIT, "vslam_best2_noalloc: ndb exceeds 2^20 database descriptors (20-bit index key)");
  if (ctx->device < 0 || ctx->device >= kNaMaxDevices)
    return vs_fail(ctx, VSLAM_E_LIMIT, "vslam_best2_noalloc: device ordinal beyond 63");
  if (nq == 0) return VSLAM_OK;
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  uint2* scratch = nullptr;
  VS_CUDA(ctx, na_scratch_for(ctx->device, &scratch));
  cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : ctx->stream;
  // column ranges: multiples of the 128-row stage, at most kNaMaxRanges of them (ndb == 0: one empty range, every
  // query reports (256, 0, 256) as vslam_best2 does)
  int range = (ndb + kNaMaxRanges - 1) / kNaMaxRanges;
  range = ((range + kNaThreads - 1) / kNaThreads) * kNaThreads;
  if (range < kNaThreads) range = kNaThreads;
  int nranges = (ndb + range - 1) / range;
  if (nranges < 1) nranges = 1;
  const dim3 grid((nq + kNaThreads - 1) / kNaThreads, nranges);
  na_scan_kernel<<<grid, kNaThreads, 0, st>>>(d_q, nq, d_db, ndb, range, scratch);
  na_merge_kernel<<<(nq + 127) / 128, 128, 0, st>>>(scratch, nq, nranges, d_out3);
  ctx->launches += 2;
  VS_CUDA(ctx, cudaGetLastError());
  return VSLAM_OK;
}
