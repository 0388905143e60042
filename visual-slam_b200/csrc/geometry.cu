// geometry.cu -- the per-match / per-landmark geometry that follows matching:
//   camera models        include/visnav/camera_models.h (pinhole :75-117, eucm :158-212, ds :246-305, kb4 :341-439)
//   findInliersEssential include/visnav/matching_utils.h:63-88
//   project_landmarks    include/visnav/vo_utils.h:48-81
// One thread per match / landmark, fp64 throughout in the reference's expression order; results that survive
// are compacted IN INPUT ORDER (the reference push_back()s while it scans).  These are latency-sized jobs
// (<= a few thousand matches, <= 1e5..1e6 landmarks): each 1024-item block evaluates its items, publishes
// its survivor count, and a second launch turns the counts into offsets and scatters.
#include "common.cuh"

namespace {

struct Cam {
  int model, width, height;
  double P[8];
};

__device__ __forceinline__ void cam_project(const Cam& c, double x, double y, double z, double& u, double& v) {
  const double fx = c.P[0], fy = c.P[1], cx = c.P[2], cy = c.P[3];
  switch (c.model) {
    case VSLAM_CAM_PINHOLE:
      u = fx * x / z + cx;
      v = fy * y / z + cy;
      break;
    case VSLAM_CAM_EUCM: {
      const double alpha = c.P[4], beta = c.P[5];
      const double d = sqrt(beta * (x * x + y * y) + z * z);
      const double den = alpha * d + (1.0 - alpha) * z;
      u = fx * x / den + cx;
      v = fy * y / den + cy;
      break;
    }
    case VSLAM_CAM_DS: {
      const double xi = c.P[4], alpha = c.P[5];
      const double d1 = sqrt(x * x + y * y + z * z);
      const double k = xi * d1 + z;
      const double d2 = sqrt(x * x + y * y + k * k);
      const double den = alpha * d2 + (1.0 - alpha) * k;
      u = fx * x / den + cx;
      v = fy * y / den + cy;
      break;
    }
    default: {  // kb4
      const double k1 = c.P[4], k2 = c.P[5], k3 = c.P[6], k4 = c.P[7];
      const double r = sqrt(x * x + y * y);
      const double t = atan2(r, z);
      const double t2 = t * t;
      const double d = t + k1 * t * t2 + k2 * t * t2 * t2 + k3 * t * t2 * t2 * t2 + k4 * t * t2 * t2 * t2 * t2;
      if (r == 0.0) {
        u = cx;
        v = cy;
      } else {
        u = fx * d * x / r + cx;
        v = fy * d * y / r + cy;
      }
    }
  }
}

__device__ __forceinline__ void cam_unproject(const Cam& c, double u, double v, double& ox, double& oy, double& oz) {
  const double fx = c.P[0], fy = c.P[1], cx = c.P[2], cy = c.P[3];
  const double mx = (u - cx) / fx, my = (v - cy) / fy;
  switch (c.model) {
    case VSLAM_CAM_PINHOLE: {
      const double s = 1.0 / sqrt(mx * mx + my * my + 1.0);
      ox = mx * s;
      oy = my * s;
      oz = s;
      break;
    }
    case VSLAM_CAM_EUCM: {
      const double alpha = c.P[4], beta = c.P[5];
      const double rr = mx * mx + my * my;
      const double mz =
          (1.0 - beta * alpha * alpha * rr) / (alpha * sqrt(1.0 - (2.0 * alpha - 1.0) * beta * rr) + (1.0 - alpha));
      const double s = 1.0 / sqrt(mx * mx + my * my + mz * mz);
      ox = mx * s;
      oy = my * s;
      oz = mz * s;
      break;
    }
    case VSLAM_CAM_DS: {
      const double xi = c.P[4], alpha = c.P[5];
      const double rr = mx * mx + my * my;
      const double mz = (1.0 - alpha * alpha * rr) / (alpha * sqrt(1.0 - (2.0 * alpha - 1.0) * rr) + 1.0 - alpha);
      const double s = (mz * xi + sqrt(mz * mz + (1.0 - xi * xi) * rr)) / (mz * mz + rr);
      ox = mx * s;
      oy = my * s;
      oz = mz * s - xi;
      break;
    }
    default: {  // kb4: five Newton steps from theta = 0 (camera_models.h:402-424)
      const double k1 = c.P[4], k2 = c.P[5], k3 = c.P[6], k4 = c.P[7];
      const double ru = sqrt(mx * mx + my * my);
      double x = 0.0;
#pragma unroll
      for (int itr = 0; itr < 5; ++itr) {
        const double x2 = x * x;
        const double d = x + k1 * x * x2 + k2 * x * x2 * x2 + k3 * x * x2 * x2 * x2 + k4 * x * x2 * x2 * x2 * x2 - ru;
        const double df = 1.0 + 3.0 * k1 * x2 + 5.0 * k2 * x2 * x2 + 7.0 * k3 * x2 * x2 * x2 + 9.0 * k4 * x2 * x2 * x2 * x2;
        x = x - d / df;
      }
      double sn, cs;
      sincos(x, &sn, &cs);
      if (ru == 0.0) {
        ox = 0.0;
        oy = 0.0;
      } else {
        ox = sn * mx / ru;
        oy = sn * my / ru;
      }
      oz = cs;
    }
  }
}

// Sophus so3.hpp:362-371: p + w*uv + qv x uv, uv = 2 (qv x p)
__device__ __forceinline__ void quat_rotate(const double* q, double px, double py, double pz, double& ox, double& oy,
                                            double& oz) {
  double a = q[1] * pz - q[2] * py, b = q[2] * px - q[0] * pz, c = q[0] * py - q[1] * px;
  a += a;
  b += b;
  c += c;
  ox = px + q[3] * a + (q[1] * c - q[2] * b);
  oy = py + q[3] * b + (q[2] * a - q[0] * c);
  oz = pz + q[3] * c + (q[0] * b - q[1] * a);
}

constexpr int kBlock = 1024;

// ordered in-block compaction helper: returns this thread's rank among the block's keepers; total via *total
__device__ __forceinline__ int block_rank(bool keep, int* warp_sums /*[32]*/, int* total) {
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned bal = __ballot_sync(0xffffffffu, keep);
  if (lane == 0) warp_sums[warp] = __popc(bal);
  __syncthreads();
  if (warp == 0) {
    int v = warp_sums[lane];
    int s = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= (unsigned)o) s += t;
    }
    warp_sums[lane] = s - v;  // exclusive
    if (lane == 31) *total = s;
  }
  __syncthreads();
  return warp_sums[warp] + __popc(bal & ((1u << lane) - 1u));
}

__global__ void __launch_bounds__(kBlock) cam_project_kernel(Cam cam, const double* __restrict__ xyz, int n,
                                                             double* __restrict__ uv) {
  const int i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= n) return;
  double u, v;
  cam_project(cam, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], u, v);
  reinterpret_cast<double2*>(uv)[i] = make_double2(u, v);
}

__global__ void __launch_bounds__(kBlock) cam_unproject_kernel(Cam cam, const double* __restrict__ uv, int n,
                                                               double* __restrict__ xyz) {
  const int i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= n) return;
  const double2 p = reinterpret_cast<const double2*>(uv)[i];
  double x, y, z;
  cam_unproject(cam, p.x, p.y, x, y, z);
  xyz[3 * i] = x;
  xyz[3 * i + 1] = y;
  xyz[3 * i + 2] = z;
}

struct EssentialArgs {
  Cam cam1, cam2;
  double E[9];
  double thr;
};

// matching_utils.h:70-86 for job `blockIdx.y`; matches / counts may come from a matcher that ran just before
// (device counts), in which case n_matches_dev != nullptr.
__global__ void __launch_bounds__(kBlock)
    essential_eval_kernel(EssentialArgs a, const double* __restrict__ xy1, const double* __restrict__ xy2,
                          size_t xy_job_stride /*doubles between jobs (0: single job)*/,
                          const int32_t* __restrict__ matches, size_t match_job_stride /*pairs between jobs*/,
                          const int32_t* __restrict__ n_matches_dev, int n_matches, uint8_t* __restrict__ keep,
                          int32_t* __restrict__ block_counts, int blocks_per_job, double* __restrict__ err_out) {
  __shared__ int warp_sums[32];
  __shared__ int total;
  const int job = blockIdx.y;
  const int nm = n_matches_dev ? n_matches_dev[job] : n_matches;
  const int j = blockIdx.x * kBlock + threadIdx.x;
  bool k = false;
  if (j < nm) {
    const int2 m = reinterpret_cast<const int2*>(matches + 2 * match_job_stride * job)[j];
    const double2 p0 = reinterpret_cast<const double2*>(xy1 + xy_job_stride * job)[m.x];
    const double2 p1 = reinterpret_cast<const double2*>(xy2 + xy_job_stride * job)[m.y];
    double ax, ay, az, bx, by, bz;
    cam_unproject(a.cam1, p0.x, p0.y, ax, ay, az);
    cam_unproject(a.cam2, p1.x, p1.y, bx, by, bz);
    const double r0 = ax * a.E[0] + ay * a.E[3] + az * a.E[6];
    const double r1 = ax * a.E[1] + ay * a.E[4] + az * a.E[7];
    const double r2 = ax * a.E[2] + ay * a.E[5] + az * a.E[8];
    const double err = r0 * bx + r1 * by + r2 * bz;
    if (err_out) err_out[(size_t)job * match_job_stride + j] = err;
    k = !(fabs(err) > a.thr);
  }
  keep[(size_t)job * blocks_per_job * kBlock + j] = k;
  block_rank(k, warp_sums, &total);
  if (threadIdx.x == 0) block_counts[job * blocks_per_job + blockIdx.x] = total;
}

__global__ void __launch_bounds__(kBlock)
    essential_scatter_kernel(const int32_t* __restrict__ matches, size_t match_job_stride,
                             const uint8_t* __restrict__ keep, const int32_t* __restrict__ block_counts,
                             int blocks_per_job, int32_t* __restrict__ inliers, size_t inlier_job_stride, int cap,
                             int32_t* __restrict__ n_inliers) {
  __shared__ int warp_sums[32];
  __shared__ int total;
  const int job = blockIdx.y;
  int base = 0;
  for (int b = 0; b < (int)blockIdx.x; ++b) base += block_counts[job * blocks_per_job + b];
  const int j = blockIdx.x * kBlock + threadIdx.x;
  const bool k = keep[(size_t)job * blocks_per_job * kBlock + j] != 0;
  const int r = base + block_rank(k, warp_sums, &total);
  if (k && r < cap)
    reinterpret_cast<int2*>(inliers + 2 * inlier_job_stride * job)[r] =
        reinterpret_cast<const int2*>(matches + 2 * match_job_stride * job)[j];
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) n_inliers[job] = base + total;
}

// se3.hpp:208-211 on one thread: the kernels below read the inverse pose from constant arguments
__global__ void pose_inverse_kernel(const double* __restrict__ pose, double* __restrict__ out7) {
  const double len = sqrt(pose[0] * pose[0] + pose[1] * pose[1] + pose[2] * pose[2] + pose[3] * pose[3]);
  double q[4] = {-pose[0] / len, -pose[1] / len, -pose[2] / len, pose[3] / len};
  double x, y, z;
  quat_rotate(q, pose[4] * -1.0, pose[5] * -1.0, pose[6] * -1.0, x, y, z);
  out7[0] = q[0]; out7[1] = q[1]; out7[2] = q[2]; out7[3] = q[3];
  out7[4] = x; out7[5] = y; out7[6] = z;
}

__global__ void __launch_bounds__(kBlock)
    project_eval_kernel(Cam cam, const double* __restrict__ inv_pose, double z_thr, const double* __restrict__ xyz,
                        int n, double* __restrict__ uv, uint8_t* __restrict__ keep, int32_t* __restrict__ block_counts) {
  __shared__ int warp_sums[32];
  __shared__ int total;
  __shared__ double T[7];
  if (threadIdx.x < 7) T[threadIdx.x] = inv_pose[threadIdx.x];
  __syncthreads();
  const int i = blockIdx.x * kBlock + threadIdx.x;
  bool k = false;
  if (i < n) {
    double x, y, z;
    quat_rotate(T, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], x, y, z);
    x += T[4];
    y += T[5];
    z += T[6];
    if (!(z < z_thr)) {  // vo_utils.h:64
      double u, v;
      cam_project(cam, x, y, z, u, v);
      k = !(u > (double)cam.width || v > (double)cam.height || u < 0.0 || v < 0.0);  // :69-70
      if (k) reinterpret_cast<double2*>(uv)[i] = make_double2(u, v);
    }
  }
  keep[i] = k;
  block_rank(k, warp_sums, &total);
  if (threadIdx.x == 0) block_counts[blockIdx.x] = total;
}

__global__ void __launch_bounds__(kBlock)
    project_scatter_kernel(const double* __restrict__ uv, const uint8_t* __restrict__ keep,
                           const int32_t* __restrict__ block_offsets, double* __restrict__ proj_xy,
                           int32_t* __restrict__ proj_idx, int cap) {
  __shared__ int warp_sums[32];
  __shared__ int total;
  const int i = blockIdx.x * kBlock + threadIdx.x;
  const bool k = keep[i] != 0;
  const int r = block_offsets[blockIdx.x] + block_rank(k, warp_sums, &total);
  if (k && r < cap) {
    reinterpret_cast<double2*>(proj_xy)[r] = reinterpret_cast<const double2*>(uv)[i];
    proj_idx[r] = i;
  }
}

// exclusive scan of the block counts (one CTA; n_blocks <= 1e6/1024 in practice, any size works)
__global__ void __launch_bounds__(kBlock) scan_counts_kernel(const int32_t* __restrict__ counts, int n_blocks,
                                                            int32_t* __restrict__ offsets, int32_t* __restrict__ total_out) {
  __shared__ int warp_sums[32];
  __shared__ int total;
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int b0 = 0; b0 < n_blocks; b0 += kBlock) {
    const int b = b0 + threadIdx.x;
    const int v = b < n_blocks ? counts[b] : 0;
    // inclusive scan by warps
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int s = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= (unsigned)o) s += t;
    }
    if (lane == 31) warp_sums[warp] = s;
    __syncthreads();
    if (warp == 0) {
      int w = warp_sums[lane], ws = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, ws, o);
        if (lane >= (unsigned)o) ws += t;
      }
      warp_sums[lane] = ws - w;
      if (lane == 31) total = ws;
    }
    __syncthreads();
    if (b < n_blocks) offsets[b] = carry + warp_sums[warp] + s - v;
    __syncthreads();
    if (threadIdx.x == 0) carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total_out = carry;
}

int to_cam(vslam_ctx* ctx, const vslam_camera* c, Cam* out) {
  if (!c || c->model < VSLAM_CAM_PINHOLE || c->model > VSLAM_CAM_KB4)
    return vs_fail(ctx, VSLAM_E_ARG, "bad camera (model must be one of VSLAM_CAM_*)");
  out->model = c->model;
  out->width = c->width;
  out->height = c->height;
  for (int i = 0; i < 8; ++i) out->P[i] = c->param[i];
  return VSLAM_OK;
}

inline size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace

// ---- launcher used by the stereo plan: inlier filtering straight from the matcher's device output ----
int launch_essential_jobs(vslam_ctx* ctx, const vslam_camera* cam1, const vslam_camera* cam2, const double* E,
                          double thr, const double* xy1, const double* xy2, size_t xy_job_stride,
                          const int32_t* matches, size_t match_job_stride, const int32_t* n_matches_dev,
                          int n_matches_max, int n_jobs, uint8_t* ws, int32_t* inliers, size_t inlier_job_stride,
                          int cap, int32_t* n_inliers, double* err_out) {
  EssentialArgs a;
  int rc = to_cam(ctx, cam1, &a.cam1);
  if (rc) return rc;
  rc = to_cam(ctx, cam2, &a.cam2);
  if (rc) return rc;
  for (int i = 0; i < 9; ++i) a.E[i] = E[i];
  a.thr = thr;
  if (n_jobs < 1 || n_jobs > 65535 || n_matches_max < 1) return vs_fail(ctx, VSLAM_E_LIMIT, "inlier jobs: 1..65535 jobs per launch");
  const int bpj = (n_matches_max + kBlock - 1) / kBlock;
  uint8_t* keep = ws;  // layout of essential_ws_bytes(): keep flags, then block counts
  int32_t* block_counts = reinterpret_cast<int32_t*>(ws + up256((size_t)n_jobs * bpj * kBlock));
  dim3 grid(bpj, n_jobs);
  essential_eval_kernel<<<grid, kBlock, 0, ctx->stream>>>(a, xy1, xy2, xy_job_stride, matches, match_job_stride,
                                                         n_matches_dev, n_matches_max, keep, block_counts, bpj,
                                                         err_out);
  essential_scatter_kernel<<<grid, kBlock, 0, ctx->stream>>>(matches, match_job_stride, keep, block_counts, bpj,
                                                            inliers, inlier_job_stride, cap, n_inliers);
  ctx->launches += 2;
  VS_CUDA(ctx, cudaGetLastError());
  return VSLAM_OK;
}

size_t essential_ws_bytes(int n_matches_max, int n_jobs) {
  const int bpj = (n_matches_max + kBlock - 1) / kBlock;
  return up256((size_t)n_jobs * bpj * kBlock) + up256((size_t)n_jobs * bpj * 4);
}

extern "C" {

int vslam_cam_project(vslam_ctx* ctx, const vslam_camera* cam, const double* xyz, int n, double* uv) {
  if (!ctx) return VSLAM_E_ARG;
  Cam c;
  int rc = to_cam(ctx, cam, &c);
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!xyz || !uv))) return vs_fail(ctx, VSLAM_E_ARG, "bad arguments");
  if (n == 0) return VSLAM_OK;
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t o_uv = up256((size_t)n * 24);
  VS_CUDA(ctx, ctx->geo.reserve(o_uv + (size_t)n * 16));
  uint8_t* base = ctx->geo.as<uint8_t>();
  VS_CUDA(ctx, cudaMemcpyAsync(base, xyz, (size_t)n * 24, cudaMemcpyHostToDevice, ctx->stream));
  cam_project_kernel<<<(n + kBlock - 1) / kBlock, kBlock, 0, ctx->stream>>>(
      c, reinterpret_cast<const double*>(base), n, reinterpret_cast<double*>(base + o_uv));
  ctx->launches += 1;
  VS_CUDA(ctx, cudaGetLastError());
  VS_CUDA(ctx, cudaMemcpyAsync(uv, base + o_uv, (size_t)n * 16, cudaMemcpyDeviceToHost, ctx->stream));
  VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return VSLAM_OK;
}

int vslam_cam_unproject(vslam_ctx* ctx, const vslam_camera* cam, const double* uv, int n, double* xyz) {
  if (!ctx) return VSLAM_E_ARG;
  Cam c;
  int rc = to_cam(ctx, cam, &c);
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!xyz || !uv))) return vs_fail(ctx, VSLAM_E_ARG, "bad arguments");
  if (n == 0) return VSLAM_OK;
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t o_xyz = up256((size_t)n * 16);
  VS_CUDA(ctx, ctx->geo.reserve(o_xyz + (size_t)n * 24));
  uint8_t* base = ctx->geo.as<uint8_t>();
  VS_CUDA(ctx, cudaMemcpyAsync(base, uv, (size_t)n * 16, cudaMemcpyHostToDevice, ctx->stream));
  cam_unproject_kernel<<<(n + kBlock - 1) / kBlock, kBlock, 0, ctx->stream>>>(
      c, reinterpret_cast<const double*>(base), n, reinterpret_cast<double*>(base + o_xyz));
  ctx->launches += 1;
  VS_CUDA(ctx, cudaGetLastError());
  VS_CUDA(ctx, cudaMemcpyAsync(xyz, base + o_xyz, (size_t)n * 24, cudaMemcpyDeviceToHost, ctx->stream));
  VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return VSLAM_OK;
}

int vslam_find_inliers_essential(vslam_ctx* ctx, const double* xy1, int n1, const double* xy2, int n2,
                                 const vslam_camera* cam1, const vslam_camera* cam2, const double* E,
                                 double epipolar_error_threshold, const int32_t* matches, int n_matches,
                                 int32_t* inliers, int cap, int* n_inliers, double* err_out) {
  if (!ctx) return VSLAM_E_ARG;
  if (!n_inliers || !E || n1 < 0 || n2 < 0 || n_matches < 0 || cap < 0 || (n1 > 0 && !xy1) || (n2 > 0 && !xy2) ||
      (n_matches > 0 && !matches) || (cap > 0 && !inliers))
    return vs_fail(ctx, VSLAM_E_ARG, "bad inlier arguments");
  *n_inliers = 0;
  Cam probe;
  int rc = to_cam(ctx, cam1, &probe);
  if (rc) return rc;
  rc = to_cam(ctx, cam2, &probe);
  if (rc) return rc;
  if (n_matches == 0) return VSLAM_OK;  // matching_utils.h:68 clear(); nothing to test
  for (int j = 0; j < n_matches; ++j)   // the reference indexes unchecked; an id outside the corner lists is a bad argument
    if (matches[2 * j] < 0 || matches[2 * j] >= n1 || matches[2 * j + 1] < 0 || matches[2 * j + 1] >= n2)
      return vs_fail(ctx, VSLAM_E_ARG, "match id outside the corner lists");
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  const int dcap = cap > 0 ? cap : 1;
  const size_t o_xy1 = 0, o_xy2 = o_xy1 + up256((size_t)n1 * 16), o_m = o_xy2 + up256((size_t)n2 * 16),
               o_keep = o_m + up256((size_t)n_matches * 8), o_inl = o_keep + essential_ws_bytes(n_matches, 1), o_cnt = o_inl + up256((size_t)dcap * 8), o_err = o_cnt + 256,
               total = o_err + up256((size_t)n_matches * 8);
  VS_CUDA(ctx, ctx->geo.reserve(total));
  uint8_t* base = ctx->geo.as<uint8_t>();
  cudaStream_t st = ctx->stream;
  VS_CUDA(ctx, cudaMemcpyAsync(base + o_xy1, xy1, (size_t)n1 * 16, cudaMemcpyHostToDevice, st));
  VS_CUDA(ctx, cudaMemcpyAsync(base + o_xy2, xy2, (size_t)n2 * 16, cudaMemcpyHostToDevice, st));
  VS_CUDA(ctx, cudaMemcpyAsync(base + o_m, matches, (size_t)n_matches * 8, cudaMemcpyHostToDevice, st));
  rc = launch_essential_jobs(ctx, cam1, cam2, E, epipolar_error_threshold, reinterpret_cast<const double*>(base + o_xy1),
                             reinterpret_cast<const double*>(base + o_xy2), 0,
                             reinterpret_cast<const int32_t*>(base + o_m), 0, nullptr, n_matches, 1, base + o_keep,
                             reinterpret_cast<int32_t*>(base + o_inl), 0,
                             dcap, reinterpret_cast<int32_t*>(base + o_cnt),
                             err_out ? reinterpret_cast<double*>(base + o_err) : nullptr);
  if (rc) return rc;
  int32_t* hn = static_cast<int32_t*>(ctx->pinned);
  VS_CUDA(ctx, cudaMemcpyAsync(hn, base + o_cnt, 4, cudaMemcpyDeviceToHost, st));
  if (err_out) VS_CUDA(ctx, cudaMemcpyAsync(err_out, base + o_err, (size_t)n_matches * 8, cudaMemcpyDeviceToHost, st));
  VS_CUDA(ctx, cudaStreamSynchronize(st));
  *n_inliers = *hn;
  const int m = *n_inliers < cap ? *n_inliers : cap;
  if (m > 0) VS_CUDA(ctx, cudaMemcpy(inliers, base + o_inl, (size_t)m * 8, cudaMemcpyDeviceToHost));
  if (*n_inliers > cap) return vs_fail(ctx, VSLAM_E_CAPACITY, "more inliers than the output capacity");
  return VSLAM_OK;
}

int vslam_project_landmarks(vslam_ctx* ctx, const double* pose_w_c, const vslam_camera* cam, const double* xyz, int n,
                            double cam_z_threshold, double* proj_xy, int32_t* proj_idx, int cap, int* n_out) {
  if (!ctx) return VSLAM_E_ARG;
  Cam c;
  int rc = to_cam(ctx, cam, &c);
  if (rc) return rc;
  if (!n_out || !pose_w_c || n < 0 || cap < 0 || (n > 0 && !xyz) || (cap > 0 && (!proj_xy || !proj_idx)))
    return vs_fail(ctx, VSLAM_E_ARG, "bad projection arguments");
  *n_out = 0;
  if (n == 0) return VSLAM_OK;  // vo_utils.h:55-56 clear()
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  const int dcap = cap > 0 ? cap : 1;
  const int nb = (n + kBlock - 1) / kBlock;
  const size_t o_xyz = 0, o_uv = o_xyz + up256((size_t)n * 24), o_keep = o_uv + up256((size_t)nb * kBlock * 16),
               o_bc = o_keep + up256((size_t)nb * kBlock), o_bo = o_bc + up256((size_t)nb * 4),
               o_pose = o_bo + up256((size_t)nb * 4), o_cnt = o_pose + 256, o_pxy = o_cnt + 256,
               o_pidx = o_pxy + up256((size_t)dcap * 16), total = o_pidx + up256((size_t)dcap * 4);
  VS_CUDA(ctx, ctx->geo.reserve(total));
  uint8_t* base = ctx->geo.as<uint8_t>();
  cudaStream_t st = ctx->stream;
  VS_CUDA(ctx, cudaMemcpyAsync(base + o_xyz, xyz, (size_t)n * 24, cudaMemcpyHostToDevice, st));
  VS_CUDA(ctx, cudaMemcpyAsync(base + o_pose, pose_w_c, 56, cudaMemcpyHostToDevice, st));
  double* inv = reinterpret_cast<double*>(base + o_pose + 64);
  pose_inverse_kernel<<<1, 1, 0, st>>>(reinterpret_cast<const double*>(base + o_pose), inv);
  project_eval_kernel<<<nb, kBlock, 0, st>>>(c, inv, cam_z_threshold, reinterpret_cast<const double*>(base + o_xyz), n,
                                             reinterpret_cast<double*>(base + o_uv), base + o_keep,
                                             reinterpret_cast<int32_t*>(base + o_bc));
  scan_counts_kernel<<<1, kBlock, 0, st>>>(reinterpret_cast<const int32_t*>(base + o_bc), nb,
                                           reinterpret_cast<int32_t*>(base + o_bo),
                                           reinterpret_cast<int32_t*>(base + o_cnt));
  project_scatter_kernel<<<nb, kBlock, 0, st>>>(reinterpret_cast<const double*>(base + o_uv), base + o_keep,
                                                reinterpret_cast<const int32_t*>(base + o_bo),
                                                reinterpret_cast<double*>(base + o_pxy),
                                                reinterpret_cast<int32_t*>(base + o_pidx), dcap);
  ctx->launches += 4;
  VS_CUDA(ctx, cudaGetLastError());
  int32_t* hn = static_cast<int32_t*>(ctx->pinned);
  VS_CUDA(ctx, cudaMemcpyAsync(hn, base + o_cnt, 4, cudaMemcpyDeviceToHost, st));
  VS_CUDA(ctx, cudaStreamSynchronize(st));
  *n_out = *hn;
  const int m = *n_out < cap ? *n_out : cap;
  if (m > 0) {
    VS_CUDA(ctx, cudaMemcpy(proj_xy, base + o_pxy, (size_t)m * 16, cudaMemcpyDeviceToHost));
    VS_CUDA(ctx, cudaMemcpy(proj_idx, base + o_pidx, (size_t)m * 4, cudaMemcpyDeviceToHost));
  }
  if (*n_out > cap) return vs_fail(ctx, VSLAM_E_CAPACITY, "more projected landmarks than the output capacity");
  return VSLAM_OK;
}

}  // extern "C"
