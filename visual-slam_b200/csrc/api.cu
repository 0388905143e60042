// api.cu -- the C ABI (include/visnav_b200.h): context, host<->device marshalling, batching.
// No computation happens on the host here; every entry point launches the CUDA kernels of
// detect.cu / describe.cu / match.cu and fails with VSLAM_E_CUDA when that is impossible.
#include <new>
#include <vector>

#include "common.cuh"

namespace {

constexpr int kMinDistance = 8;      // keypoints.h:138: goodFeaturesToTrack(..., 0.01, 8, ...)
constexpr double kQuality = 0.01;
constexpr int kEdgeThreshold = 19;   // keypoints.h:50 EDGE_THRESHOLD

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
// Device rows are `w` bytes apart when that keeps every row 16-byte aligned (TMA stride rule, 128-bit
// patch loads): a dense host image then uploads as ONE linear DMA instead of a slow strided 2-D copy.
inline int dev_pitch(int w) { return (w % 16 == 0) ? w : (int)align_up((size_t)w, 128); }
inline int cand_capacity(int w, int h) { return (w * h) / 4 + 1024; }

// Image set in ctx scratch + detection workspace for `count` images.
struct Scratch {
  DevImages im;
  DetectWorkspace ws;
};

int prepare_detect_ws(vslam_ctx* ctx, int w, int h, int count, DevBuf& cand, DevBuf& candcnt, DevBuf& imgmax,
                      DevBuf& selws, DetectWorkspace* ws) {
  const int ccap = cand_capacity(w, h);
  const size_t wsb = detect_sel_ws_bytes(w, h, ccap, kMinDistance);
  VS_CUDA(ctx, cand.reserve((size_t)count * ccap * sizeof(uint64_t)));
  VS_CUDA(ctx, candcnt.reserve((size_t)count * sizeof(uint32_t)));
  VS_CUDA(ctx, imgmax.reserve((size_t)count * sizeof(uint32_t)));
  // behind the select workspace: the planes of the small-count response path (10 floats per pixel and image)
  // (at most 1 GiB of them: 16 images of 752 x 480, 9 of 1920 x 1080, 2 of 3840 x 2160; launches of more images than the
  // planes hold use the streaming kernel)
  const size_t small_per_image = (((size_t)w * h + 1) & ~(size_t)1) * 10 * sizeof(float);
  int nsmall = count < kRespSmallMaxImages ? count : kRespSmallMaxImages;
  if ((size_t)nsmall * small_per_image > ((size_t)1 << 30)) nsmall = (int)(((size_t)1 << 30) / small_per_image);
  const size_t sel_bytes = ((size_t)count * wsb + 255) & ~(size_t)255;
  VS_CUDA(ctx, selws.reserve(sel_bytes + (size_t)nsmall * small_per_image));
  ws->small_ws = reinterpret_cast<float*>(selws.as<uint8_t>() + sel_bytes);
  ws->small_images = nsmall;
  ws->cand = cand.as<uint64_t>();
  ws->cand_cnt = candcnt.as<uint32_t>();
  ws->img_max = imgmax.as<uint32_t>();
  ws->cand_cap = ccap;
  ws->sel_ws = selws.as<uint8_t>();
  ws->sel_ws_stride = wsb;
  return VSLAM_OK;
}

int upload_images(vslam_ctx* ctx, DevBuf& buf, const uint8_t* host, int count, int w, int h, int pitch,
                  size_t image_stride, DevImages* im, cudaStream_t stream) {
  const int dp = dev_pitch(w);
  const size_t ds = (size_t)dp * h;
  VS_CUDA(ctx, buf.reserve(ds * count));
  im->base = buf.as<uint8_t>();
  im->w = w; im->h = h; im->pitch = dp; im->stride = ds;
  if (pitch == dp && (count == 1 || image_stride == ds)) {
    VS_CUDA(ctx, cudaMemcpyAsync(buf.p, host, ds * count, cudaMemcpyHostToDevice, stream));
  } else if (count > 1 && image_stride == (size_t)pitch * h) {
    // images are back to back: one strided copy of count*h rows
    VS_CUDA(ctx, cudaMemcpy2DAsync(buf.p, dp, host, pitch, w, (size_t)h * count, cudaMemcpyHostToDevice, stream));
  } else {
    for (int i = 0; i < count; ++i)
      VS_CUDA(ctx, cudaMemcpy2DAsync(buf.as<uint8_t>() + ds * i, dp, host + image_stride * i, pitch, w, h,
                                     cudaMemcpyHostToDevice, stream));
  }
  return VSLAM_OK;
}

int check_image(vslam_ctx* ctx, const uint8_t* img, int w, int h, int pitch) {
  if (!ctx) return VSLAM_E_ARG;
  if (!img || w < 2 || h < 2 || pitch < w) return vs_fail(ctx, VSLAM_E_ARG, "bad image arguments");
  if ((long long)w * h > (1LL << 30)) return vs_fail(ctx, VSLAM_E_LIMIT, "image larger than 2^30 pixels");
  return VSLAM_OK;
}

int kp_status(vslam_ctx* ctx, int n) {
  if (n == -1) return vs_fail(ctx, VSLAM_E_LIMIT, "corner candidate list overflowed (more than w*h/4 local maxima)");
  if (n == -2) return vs_fail(ctx, VSLAM_E_CAPACITY, "more corners than the output capacity");
  return VSLAM_OK;
}

}  // namespace

extern "C" {

int vslam_create(int device, vslam_ctx** out) {
  if (!out) return VSLAM_E_ARG;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return VSLAM_E_CUDA;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return VSLAM_E_CUDA;
  if (prop.major != 10) return VSLAM_E_CUDA;  // sm_100a SASS only; nothing else can run it
  if (cudaSetDevice(device) != cudaSuccess) return VSLAM_E_CUDA;
  vslam_ctx* ctx = new (std::nothrow) vslam_ctx();
  if (!ctx) return VSLAM_E_CUDA;
  ctx->device = device;
  ctx->num_sms = prop.multiProcessorCount;
  bool ok = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) == cudaSuccess;
  for (int i = 0; i < 2 && ok; ++i) ok = cudaStreamCreateWithFlags(&ctx->copy_stream[i], cudaStreamNonBlocking) == cudaSuccess;
  for (int i = 0; i < 8 && ok; ++i) ok = cudaEventCreate(&ctx->ev[i]) == cudaSuccess;
  ok = ok && cudaMallocHost(&ctx->pinned, 4096) == cudaSuccess;
  if (!ok) {
    vslam_destroy(ctx);
    return VSLAM_E_CUDA;
  }
  *out = ctx;
  return VSLAM_OK;
}

void vslam_destroy(vslam_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  for (int i = 0; i < vslam_ctx::kMaxBatchDepth; ++i) {
    if (ctx->batch_plans[i]) vslam_stereo_plan_destroy(ctx->batch_plans[i]);
    ctx->batch_plans[i] = nullptr;
  }
  DevBuf* bufs[] = {&ctx->img, &ctx->xy, &ctx->ang, &ctx->desc, &ctx->cnt, &ctx->lam, &ctx->cand, &ctx->candcnt,
                    &ctx->imgmax, &ctx->selws, &ctx->d1, &ctx->d2, &ctx->pairs, &ctx->mcnt, &ctx->rowpart,
                    &ctx->colstat, &ctx->jobs, &ctx->lm, &ctx->pyr, &ctx->geo};
  for (DevBuf* b : bufs) b->release();
  for (DevBuf& b : ctx->orb) b.release();
  for (int i = 0; i < 8; ++i)
    if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
  for (int i = 0; i < 2; ++i)
    if (ctx->copy_stream[i]) cudaStreamDestroy(ctx->copy_stream[i]);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  if (ctx->pinned) cudaFreeHost(ctx->pinned);
  if (ctx->pinned_status) cudaFreeHost(ctx->pinned_status);
  if (ctx->pinned_out) cudaFreeHost(ctx->pinned_out);
  delete ctx;
}

const char* vslam_last_error(vslam_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int vslam_set_option(vslam_ctx* ctx, int option, int value) {
  if (!ctx) return VSLAM_E_ARG;
  if (option == VSLAM_OPT_FP_CONTRACT) ctx->fp_contract = value != 0;
  else if (option == VSLAM_OPT_MATCH_CSA) ctx->match_csa = value != 0;
  else if (option == VSLAM_OPT_MATCH_SPLIT_SMALL) ctx->match_split_small = value != 0;
  else if (option == VSLAM_OPT_MATCH_MMA && value >= 0 && value <= 2) ctx->match_mma = value;
  else if (option == VSLAM_OPT_GRAPH) ctx->use_graph = value != 0;
  else if (option == VSLAM_OPT_TWO_STREAMS) ctx->two_streams = value != 0;
  else if (option == VSLAM_OPT_RESPONSE_MULTI && value >= 0 && value <= 2) ctx->resp_multi = value;
  else if (option == VSLAM_OPT_RESPONSE_BAND && (value == 0 || value == 1 || (value >= 8 && value <= 65535))) ctx->resp_band = value;
  else if (option == VSLAM_OPT_BATCH_CHUNK && value >= 1 && value <= 4096) ctx->batch_chunk = value;
  else if (option == VSLAM_OPT_BATCH_DEPTH && value >= 1 && value <= vslam_ctx::kMaxBatchDepth) ctx->batch_depth = value;
  else return vs_fail(ctx, VSLAM_E_ARG, "unknown option");
  return VSLAM_OK;
}

int vslam_set_harris(vslam_ctx* ctx, int use_harris_detector, double k) {
  if (!ctx) return VSLAM_E_ARG;
  if (!(k == k)) return vs_fail(ctx, VSLAM_E_ARG, "harris k is NaN");
  ctx->harris_on = use_harris_detector != 0;
  ctx->harris_k = k;
  return VSLAM_OK;
}

void* vslam_stream(vslam_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

int vslam_sync(vslam_ctx* ctx) {
  if (!ctx) return VSLAM_E_ARG;
  VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return VSLAM_OK;
}

long long vslam_launch_count(vslam_ctx* ctx) { return ctx ? ctx->launches : 0; }

void* vslam_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (cudaMallocHost(&p, bytes) != cudaSuccess) return nullptr;
  return p;
}
void vslam_host_free(void* p) {
  if (p) cudaFreeHost(p);
}

// ---- single image -------------------------------------------------------------------------------

static int detect_on_device(vslam_ctx* ctx, const uint8_t* img, int w, int h, int pitch, int num_features, int cap,
                            DevImages* im) {
  DetectWorkspace ws;
  int rc = upload_images(ctx, ctx->img, img, 1, w, h, pitch, (size_t)pitch * h, im, ctx->stream);
  if (rc) return rc;
  rc = prepare_detect_ws(ctx, w, h, 1, ctx->cand, ctx->candcnt, ctx->imgmax, ctx->selws, &ws);
  if (rc) return rc;
  VS_CUDA(ctx, ctx->xy.reserve((size_t)cap * 16));
  VS_CUDA(ctx, ctx->cnt.reserve(64));
  rc = launch_response(ctx, *im, 1, ws, nullptr);
  if (rc) return rc;
  return launch_select(ctx, *im, 1, ws, num_features, kQuality, kMinDistance, kEdgeThreshold, ctx->xy.as<double>(),
                       ctx->cnt.as<int32_t>(), cap);
}

int vslam_detect(vslam_ctx* ctx, const uint8_t* img, int w, int h, int pitch, int num_features, double* xy, int cap,
                 int* n) {
  int rc = check_image(ctx, img, w, h, pitch);
  if (rc) return rc;
  if (!n || cap < 0 || (cap > 0 && !xy)) return vs_fail(ctx, VSLAM_E_ARG, "bad output arguments");
  *n = 0;
  if (w < 2 * kEdgeThreshold + 1 || h < 2 * kEdgeThreshold + 1) return VSLAM_OK;  // InBounds(.,.,19) keeps nothing
  if (cap == 0) return vs_fail(ctx, VSLAM_E_CAPACITY, "cap == 0");
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  DevImages im;
  rc = detect_on_device(ctx, img, w, h, pitch, num_features, cap, &im);
  if (rc) return rc;
  int32_t* hn = static_cast<int32_t*>(ctx->pinned);
  VS_CUDA(ctx, cudaMemcpyAsync(hn, ctx->cnt.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
  VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if ((rc = kp_status(ctx, *hn))) return rc;
  *n = *hn;
  if (*n > 0) VS_CUDA(ctx, cudaMemcpy(xy, ctx->xy.p, (size_t)*n * 16, cudaMemcpyDeviceToHost));
  return VSLAM_OK;
}

static int orient_describe_host(vslam_ctx* ctx, const uint8_t* img, int w, int h, int pitch, const double* xy,
                                int n, int mode, double* angles_io, uint32_t* desc) {
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  DevImages im;
  int rc = upload_images(ctx, ctx->img, img, 1, w, h, pitch, (size_t)pitch * h, &im, ctx->stream);
  if (rc) return rc;
  VS_CUDA(ctx, ctx->xy.reserve((size_t)n * 16));
  VS_CUDA(ctx, ctx->ang.reserve((size_t)n * 8));
  VS_CUDA(ctx, ctx->cnt.reserve(64));
  if (desc) VS_CUDA(ctx, ctx->desc.reserve((size_t)n * 32));
  int32_t* hn = static_cast<int32_t*>(ctx->pinned);
  *hn = n;
  VS_CUDA(ctx, cudaMemcpyAsync(ctx->cnt.p, hn, 4, cudaMemcpyHostToDevice, ctx->stream));
  VS_CUDA(ctx, cudaMemcpyAsync(ctx->xy.p, xy, (size_t)n * 16, cudaMemcpyHostToDevice, ctx->stream));
  if (mode == 2) VS_CUDA(ctx, cudaMemcpyAsync(ctx->ang.p, angles_io, (size_t)n * 8, cudaMemcpyHostToDevice, ctx->stream));
  rc = launch_orient_describe(ctx, im, 1, ctx->xy.as<double>(), ctx->cnt.as<int32_t>(), n, mode, ctx->ang.as<double>(),
                              desc ? ctx->desc.as<uint32_t>() : nullptr);
  if (rc) return rc;
  if (mode != 2) VS_CUDA(ctx, cudaMemcpyAsync(angles_io, ctx->ang.p, (size_t)n * 8, cudaMemcpyDeviceToHost, ctx->stream));
  if (desc) VS_CUDA(ctx, cudaMemcpyAsync(desc, ctx->desc.p, (size_t)n * 32, cudaMemcpyDeviceToHost, ctx->stream));
  VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return VSLAM_OK;
}

int vslam_angles(vslam_ctx* ctx, const uint8_t* img, int w, int h, int pitch, const double* xy, int n, int rotate,
                 double* angles) {
  int rc = check_image(ctx, img, w, h, pitch);
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!xy || !angles))) return vs_fail(ctx, VSLAM_E_ARG, "bad keypoint arguments");
  if (n == 0) return VSLAM_OK;
  return orient_describe_host(ctx, img, w, h, pitch, xy, n, rotate ? 1 : 0, angles, nullptr);
}

int vslam_describe(vslam_ctx* ctx, const uint8_t* img, int w, int h, int pitch, const double* xy,
                   const double* angles, int n, uint32_t* desc) {
  int rc = check_image(ctx, img, w, h, pitch);
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!xy || !angles || !desc))) return vs_fail(ctx, VSLAM_E_ARG, "bad keypoint arguments");
  if (n == 0) return VSLAM_OK;
  return orient_describe_host(ctx, img, w, h, pitch, xy, n, 2, const_cast<double*>(angles), desc);
}

int vslam_detect_describe(vslam_ctx* ctx, const uint8_t* img, int w, int h, int pitch, int num_features, int rotate,
                          double* xy, double* angles, uint32_t* desc, int cap, int* n) {
  int rc = check_image(ctx, img, w, h, pitch);
  if (rc) return rc;
  if (!n || cap < 0 || (cap > 0 && (!xy || !angles || !desc))) return vs_fail(ctx, VSLAM_E_ARG, "bad output arguments");
  *n = 0;
  if (w < 2 * kEdgeThreshold + 1 || h < 2 * kEdgeThreshold + 1) return VSLAM_OK;
  if (cap == 0) return vs_fail(ctx, VSLAM_E_CAPACITY, "cap == 0");
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  DevImages im;
  rc = detect_on_device(ctx, img, w, h, pitch, num_features, cap, &im);
  if (rc) return rc;
  VS_CUDA(ctx, ctx->ang.reserve((size_t)cap * 8));
  VS_CUDA(ctx, ctx->desc.reserve((size_t)cap * 32));
  rc = launch_orient_describe(ctx, im, 1, ctx->xy.as<double>(), ctx->cnt.as<int32_t>(), cap, rotate ? 1 : 0,
                              ctx->ang.as<double>(), ctx->desc.as<uint32_t>());
  if (rc) return rc;
  int32_t* hn = static_cast<int32_t*>(ctx->pinned);
  VS_CUDA(ctx, cudaMemcpyAsync(hn, ctx->cnt.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
  VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if ((rc = kp_status(ctx, *hn))) return rc;
  *n = *hn;
  if (*n > 0) {
    VS_CUDA(ctx, cudaMemcpyAsync(xy, ctx->xy.p, (size_t)*n * 16, cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaMemcpyAsync(angles, ctx->ang.p, (size_t)*n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaMemcpyAsync(desc, ctx->desc.p, (size_t)*n * 32, cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  return VSLAM_OK;
}

int vslam_response(vslam_ctx* ctx, const uint8_t* img, int w, int h, int pitch, float* lam) {
  int rc = check_image(ctx, img, w, h, pitch);
  if (rc) return rc;
  if (!lam) return vs_fail(ctx, VSLAM_E_ARG, "null output");
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  DevImages im;
  DetectWorkspace ws;
  rc = upload_images(ctx, ctx->img, img, 1, w, h, pitch, (size_t)pitch * h, &im, ctx->stream);
  if (rc) return rc;
  rc = prepare_detect_ws(ctx, w, h, 1, ctx->cand, ctx->candcnt, ctx->imgmax, ctx->selws, &ws);
  if (rc) return rc;
  VS_CUDA(ctx, ctx->lam.reserve((size_t)w * h * 4));
  rc = launch_response(ctx, im, 1, ws, ctx->lam.as<float>());
  if (rc) return rc;
  VS_CUDA(ctx, cudaMemcpyAsync(lam, ctx->lam.p, (size_t)w * h * 4, cudaMemcpyDeviceToHost, ctx->stream));
  VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return VSLAM_OK;
}

// ---- matching -------------------------------------------------------------------------------------

static int match_common(vslam_ctx* ctx, const uint32_t* d1, int n1, const uint32_t* d2, int n2, int threshold,
                        double ratio, int32_t* pairs, int cap, int* n, int32_t* best2_out) {
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  VS_CUDA(ctx, ctx->d1.reserve((size_t)n1 * 32));
  VS_CUDA(ctx, ctx->d2.reserve((size_t)n2 * 32));
  const int dcap = best2_out ? 0 : (cap > 0 ? cap : 1);
  VS_CUDA(ctx, ctx->pairs.reserve(best2_out ? (size_t)n1 * 12 : (size_t)dcap * 8));
  VS_CUDA(ctx, ctx->mcnt.reserve(256));
  VS_CUDA(ctx, cudaMemcpyAsync(ctx->d1.p, d1, (size_t)n1 * 32, cudaMemcpyHostToDevice, ctx->stream));
  VS_CUDA(ctx, cudaMemcpyAsync(ctx->d2.p, d2, (size_t)n2 * 32, cudaMemcpyHostToDevice, ctx->stream));
  VS_CUDA(ctx, cudaEventRecord(ctx->ev[0], ctx->stream));
  int rc;
  // A single call is latency-bound: the fused kernel would run it on ONE SM (0.6 ms at 1300 x 1300).
  // Beyond a few thousand distances per SM the column-split path (all SMs, three short launches) wins.
  const bool small = (long long)n1 * n2 <= 256LL * 256LL;
  if (!best2_out && small && match_fits_fused(n1, n2)) {
    VS_CUDA(ctx, ctx->jobs.reserve(sizeof(MatchJob)));
    MatchJob job;
    job.d1 = ctx->d1.as<uint32_t>(); job.d2 = ctx->d2.as<uint32_t>();
    job.n1p = nullptr; job.n2p = nullptr; job.n1 = n1; job.n2 = n2;
    job.pairs = ctx->pairs.as<int32_t>();
    job.count = ctx->mcnt.as<int32_t>();
    memcpy(static_cast<char*>(ctx->pinned) + 64, &job, sizeof job);
    VS_CUDA(ctx, cudaMemcpyAsync(ctx->jobs.p, static_cast<char*>(ctx->pinned) + 64, sizeof job, cudaMemcpyHostToDevice,
                                 ctx->stream));
    rc = launch_match_jobs(ctx, ctx->jobs.as<MatchJob>(), 1, n1, n2, threshold, ratio, dcap);
  } else {
    rc = launch_match_large(ctx, ctx->d1.as<uint32_t>(), n1, ctx->d2.as<uint32_t>(), n2, threshold, ratio,
                            ctx->pairs.as<int32_t>(), ctx->mcnt.as<int32_t>(), dcap,
                            best2_out ? ctx->pairs.as<int32_t>() : nullptr);
  }
  if (rc) return rc;
  VS_CUDA(ctx, cudaEventRecord(ctx->ev[1], ctx->stream));
  if (best2_out) {
    VS_CUDA(ctx, cudaMemcpyAsync(best2_out, ctx->pairs.p, (size_t)n1 * 12, cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  } else {
    int32_t* hn = static_cast<int32_t*>(ctx->pinned);
    VS_CUDA(ctx, cudaMemcpyAsync(hn, ctx->mcnt.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *n = *hn;
    const int m = *n < cap ? *n : cap;
    if (m > 0) VS_CUDA(ctx, cudaMemcpy(pairs, ctx->pairs.p, (size_t)m * 8, cudaMemcpyDeviceToHost));
  }
  VS_CUDA(ctx, cudaEventElapsedTime(&ctx->last_match_ms, ctx->ev[0], ctx->ev[1]));
  if (!best2_out && *n > cap) return vs_fail(ctx, VSLAM_E_CAPACITY, "more matches than the output capacity");
  return VSLAM_OK;
}

int vslam_match(vslam_ctx* ctx, const uint32_t* d1, int n1, const uint32_t* d2, int n2, int threshold,
                double dist_2_best, int32_t* pairs, int cap, int* n) {
  if (!ctx) return VSLAM_E_ARG;
  if (!n || n1 < 0 || n2 < 0 || cap < 0 || (n1 > 0 && !d1) || (n2 > 0 && !d2) || (cap > 0 && !pairs))
    return vs_fail(ctx, VSLAM_E_ARG, "bad match arguments");
  *n = 0;
  if (n1 == 0 || n2 == 0) return VSLAM_OK;  // keypoints.h:327 clear(); nothing can pass
  return match_common(ctx, d1, n1, d2, n2, threshold, dist_2_best, pairs, cap, n, nullptr);
}

int vslam_match_landmarks(vslam_ctx* ctx, const double* kp_xy, const uint32_t* kp_desc, int n_kp,
                          const double* proj_xy, int n_proj, const int32_t* obs_offsets, const uint32_t* obs_desc,
                          double match_max_dist_2d, int threshold, double dist_2_best, int32_t* pairs, int cap,
                          int* n) {
  if (!ctx) return VSLAM_E_ARG;
  if (!n || n_kp < 0 || n_proj < 0 || cap < 0 || (n_kp > 0 && (!kp_xy || !kp_desc)) ||
      (n_proj > 0 && (!proj_xy || !obs_offsets)) || (cap > 0 && !pairs))
    return vs_fail(ctx, VSLAM_E_ARG, "bad landmark-match arguments");
  *n = 0;
  if (n_kp == 0 || n_proj == 0) return VSLAM_OK;  // vo_utils.h:92 clear(); no candidate can exist
  const int total_obs = obs_offsets[n_proj];
  if (obs_offsets[0] != 0 || total_obs < 0 || (total_obs > 0 && !obs_desc))
    return vs_fail(ctx, VSLAM_E_ARG, "bad observation offsets");
  for (int j = 0; j < n_proj; ++j)
    if (obs_offsets[j + 1] < obs_offsets[j]) return vs_fail(ctx, VSLAM_E_ARG, "observation offsets must ascend");
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  // one staging buffer: [kp_xy | proj_xy | kp_desc | obs_desc | obs_offsets | result | pairs | count]
  auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
  const int dcap = cap > 0 ? cap : 1;
  const size_t o_kpxy = 0, o_proj = o_kpxy + up((size_t)n_kp * 16), o_kpd = o_proj + up((size_t)n_proj * 16),
               o_obsd = o_kpd + up((size_t)n_kp * 32), o_off = o_obsd + up((size_t)total_obs * 32 + 32),
               o_res = o_off + up((size_t)(n_proj + 1) * 4), o_pairs = o_res + up((size_t)n_kp * 4),
               o_cnt = o_pairs + up((size_t)dcap * 8), total = o_cnt + 256;
  VS_CUDA(ctx, ctx->lm.reserve(total));
  uint8_t* base = ctx->lm.as<uint8_t>();
  cudaStream_t st = ctx->stream;
  VS_CUDA(ctx, cudaMemcpyAsync(base + o_kpxy, kp_xy, (size_t)n_kp * 16, cudaMemcpyHostToDevice, st));
  VS_CUDA(ctx, cudaMemcpyAsync(base + o_proj, proj_xy, (size_t)n_proj * 16, cudaMemcpyHostToDevice, st));
  VS_CUDA(ctx, cudaMemcpyAsync(base + o_kpd, kp_desc, (size_t)n_kp * 32, cudaMemcpyHostToDevice, st));
  if (total_obs > 0)
    VS_CUDA(ctx, cudaMemcpyAsync(base + o_obsd, obs_desc, (size_t)total_obs * 32, cudaMemcpyHostToDevice, st));
  VS_CUDA(ctx, cudaMemcpyAsync(base + o_off, obs_offsets, (size_t)(n_proj + 1) * 4, cudaMemcpyHostToDevice, st));
  VS_CUDA(ctx, cudaEventRecord(ctx->ev[0], st));
  int rc = launch_match_landmarks(ctx, reinterpret_cast<const double*>(base + o_kpxy),
                                  reinterpret_cast<const uint32_t*>(base + o_kpd), n_kp,
                                  reinterpret_cast<const double*>(base + o_proj), n_proj,
                                  reinterpret_cast<const int32_t*>(base + o_off),
                                  reinterpret_cast<const uint32_t*>(base + o_obsd), match_max_dist_2d, threshold,
                                  dist_2_best, reinterpret_cast<int32_t*>(base + o_res),
                                  reinterpret_cast<int32_t*>(base + o_pairs), reinterpret_cast<int32_t*>(base + o_cnt),
                                  dcap);
  if (rc) return rc;
  VS_CUDA(ctx, cudaEventRecord(ctx->ev[1], st));
  int32_t* hn = static_cast<int32_t*>(ctx->pinned);
  VS_CUDA(ctx, cudaMemcpyAsync(hn, base + o_cnt, 4, cudaMemcpyDeviceToHost, st));
  VS_CUDA(ctx, cudaStreamSynchronize(st));
  *n = *hn;
  const int m = *n < cap ? *n : cap;
  if (m > 0) VS_CUDA(ctx, cudaMemcpy(pairs, base + o_pairs, (size_t)m * 8, cudaMemcpyDeviceToHost));
  VS_CUDA(ctx, cudaEventElapsedTime(&ctx->last_match_ms, ctx->ev[0], ctx->ev[1]));
  if (*n > cap) return vs_fail(ctx, VSLAM_E_CAPACITY, "more landmark matches than the output capacity");
  return VSLAM_OK;
}

int vslam_best2(vslam_ctx* ctx, const uint32_t* q, int nq, const uint32_t* db, int ndb, int32_t* out3) {
  if (!ctx) return VSLAM_E_ARG;
  if (nq < 0 || ndb < 0 || (nq > 0 && (!q || !out3)) || (ndb > 0 && !db)) return vs_fail(ctx, VSLAM_E_ARG, "bad arguments");
  if (nq == 0) return VSLAM_OK;
  if (ndb == 0) {
    for (int i = 0; i < nq; ++i) { out3[3 * i] = 256; out3[3 * i + 1] = 0; out3[3 * i + 2] = 256; }
    return VSLAM_OK;
  }
  int dummy = 0;
  return match_common(ctx, q, nq, db, ndb, 0, 0.0, nullptr, 0, &dummy, out3);
}

int vslam_match_shard(vslam_ctx* ctx, const uint32_t* d1, int n1, const uint32_t* d2_shard, int n2_shard,
                      int index_offset, int threshold, double dist_2_best, int32_t* part4) {
  if (!ctx) return VSLAM_E_ARG;
  if (n1 < 0 || n2_shard < 0 || index_offset < 0 || (n1 > 0 && (!d1 || !part4)) || (n2_shard > 0 && !d2_shard))
    return vs_fail(ctx, VSLAM_E_ARG, "bad shard arguments");
  if (n1 == 0) return VSLAM_OK;
  if (n2_shard == 0) {  // an empty shard contributes the scan's initial values and can never win the mutual check
    for (int i = 0; i < n1; ++i) {
      part4[4 * i] = 256; part4[4 * i + 1] = index_offset; part4[4 * i + 2] = 256; part4[4 * i + 3] = 0;
    }
    return VSLAM_OK;
  }
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  VS_CUDA(ctx, ctx->d1.reserve((size_t)n1 * 32));
  VS_CUDA(ctx, ctx->d2.reserve((size_t)n2_shard * 32));
  VS_CUDA(ctx, ctx->pairs.reserve((size_t)n1 * 16));
  VS_CUDA(ctx, cudaMemcpyAsync(ctx->d1.p, d1, (size_t)n1 * 32, cudaMemcpyHostToDevice, ctx->stream));
  VS_CUDA(ctx, cudaMemcpyAsync(ctx->d2.p, d2_shard, (size_t)n2_shard * 32, cudaMemcpyHostToDevice, ctx->stream));
  VS_CUDA(ctx, cudaEventRecord(ctx->ev[0], ctx->stream));
  int rc = launch_match_shard(ctx, ctx->d1.as<uint32_t>(), n1, ctx->d2.as<uint32_t>(), n2_shard, index_offset, threshold,
                              dist_2_best, ctx->pairs.as<int32_t>());
  if (rc) return rc;
  VS_CUDA(ctx, cudaEventRecord(ctx->ev[1], ctx->stream));
  VS_CUDA(ctx, cudaMemcpyAsync(part4, ctx->pairs.p, (size_t)n1 * 16, cudaMemcpyDeviceToHost, ctx->stream));
  VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  VS_CUDA(ctx, cudaEventElapsedTime(&ctx->last_match_ms, ctx->ev[0], ctx->ev[1]));
  return VSLAM_OK;
}

int vslam_match_merge_shards(vslam_ctx* ctx, const int32_t* parts4, int n_parts, int n1, int threshold,
                             double dist_2_best, int32_t* pairs, int cap, int* n) {
  if (!ctx) return VSLAM_E_ARG;
  if (!n || n_parts < 0 || n1 < 0 || cap < 0 || (n_parts > 0 && n1 > 0 && !parts4) || (cap > 0 && !pairs))
    return vs_fail(ctx, VSLAM_E_ARG, "bad merge arguments");
  *n = 0;
  if (n_parts == 0 || n1 == 0) return VSLAM_OK;
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  const int dcap = cap > 0 ? cap : 1;
  const size_t pbytes = (size_t)n_parts * n1 * 16;
  VS_CUDA(ctx, ctx->rowpart.reserve(pbytes));
  VS_CUDA(ctx, ctx->pairs.reserve((size_t)dcap * 8));
  VS_CUDA(ctx, ctx->mcnt.reserve(256));
  VS_CUDA(ctx, cudaMemcpyAsync(ctx->rowpart.p, parts4, pbytes, cudaMemcpyHostToDevice, ctx->stream));
  int rc = launch_merge_shards(ctx, ctx->rowpart.as<int32_t>(), n_parts, n1, threshold, dist_2_best,
                               ctx->pairs.as<int32_t>(), ctx->mcnt.as<int32_t>(), dcap);
  if (rc) return rc;
  int32_t* hn = static_cast<int32_t*>(ctx->pinned);
  VS_CUDA(ctx, cudaMemcpyAsync(hn, ctx->mcnt.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
  VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *n = *hn;
  const int m = *n < cap ? *n : cap;
  if (m > 0) VS_CUDA(ctx, cudaMemcpy(pairs, ctx->pairs.p, (size_t)m * 8, cudaMemcpyDeviceToHost));
  if (*n > cap) return vs_fail(ctx, VSLAM_E_CAPACITY, "more matches than the output capacity");
  return VSLAM_OK;
}

float vslam_last_match_ms(vslam_ctx* ctx) { return ctx ? ctx->last_match_ms : 0.f; }

int vslam_event_record(vslam_ctx* ctx, int slot) {
  if (!ctx || slot < 0 || slot > 3) return VSLAM_E_ARG;
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  VS_CUDA(ctx, cudaEventRecord(ctx->ev[4 + slot], ctx->stream));
  return VSLAM_OK;
}

int vslam_event_elapsed_ms(vslam_ctx* ctx, int slot_a, int slot_b, float* ms) {
  if (!ctx || !ms || slot_a < 0 || slot_a > 3 || slot_b < 0 || slot_b > 3) return VSLAM_E_ARG;
  VS_CUDA(ctx, cudaEventSynchronize(ctx->ev[4 + slot_b]));
  VS_CUDA(ctx, cudaEventElapsedTime(ms, ctx->ev[4 + slot_a], ctx->ev[4 + slot_b]));
  return VSLAM_OK;
}

int vslam_popc_peak(vslam_ctx* ctx, double* popc_per_s) {
  if (!ctx || !popc_per_s) return VSLAM_E_ARG;
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  return launch_popc_peak(ctx, popc_per_s);
}

int vslam_debug_mma_timeouts(vslam_ctx* ctx) {
  if (!ctx) return VSLAM_E_ARG;
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  int flag = 0;
  if (match_mma_timeouts(&flag)) return vs_fail(ctx, VSLAM_E_CUDA, "cannot read the matcher's timeout flag");
  return flag;
}

}  // extern "C"

// ---- stereo plan --------------------------------------------------------------------------------------
static constexpr size_t kArenaMaxBytes = 2u << 20;  // plans whose results fit 2 MB (up to ~5 pairs of 1500) use one arena

struct vslam_stereo_plan {
  vslam_ctx* ctx = nullptr;
  int w = 0, h = 0, max_pairs = 0, cap = 0;
  DevImages im{};
  DevBuf images, xy, ang, desc, nkp, matches, nmatch, jobs;
  DevBuf descx;  // the descriptors as 256 signed bytes each: operands of the tensor-core matcher (match_mma.cu)
  // small plans (a few pairs: the latency path) keep xy | ang | desc | matches | nkp | nmatch in ONE allocation, so that a
  // call's results travel in one device-to-host copy instead of six
  DevBuf arena;
  size_t arena_used = 0;
  // ... and replay their launch sequence (init, response, select, describe, the two-launch split matcher) as ONE CUDA
  // graph: captured on the second run with the same arguments and options, replayed from the third (config 1: the
  // reference calls the path one pair at a time, slam.cpp:1125-1143, and six launches cost more host time than the
  // copies).  0: nothing yet, 1: ran once uncaptured with graph_key (allocations are warm), 2: graph_exec ready,
  // 3: capture failed once, never again.
  int graph_state = 0;
  unsigned long long graph_key = 0;
  cudaGraphExec_t graph_exec = nullptr;
  long long graph_launches = 0;
  DevBuf cand, candcnt, imgmax, selws;
  DevBuf mrow, mcol;            // scratch of the small-batch column-split matcher (plans run on their own streams)
  DevBuf inliers, ninl, epiws;  // optional fused findInliersEssential stage (vslam_stereo_plan_set_epipolar)
  bool epi_on = false;
  vslam_epipolar epi{};
  DetectWorkspace ws{};
  cudaEvent_t ev[6] = {};
  cudaStream_t stream = nullptr;  // null: ctx stream
  cudaStream_t stream2 = nullptr;  // second half of a large batch (vslam_stereo_plan_run), forked / joined with the events
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  bool timed = false;
};

extern "C" {

void vslam_stereo_plan_destroy(vslam_stereo_plan* plan) {
  if (!plan) return;
  cudaSetDevice(plan->ctx->device);
  cudaDeviceSynchronize();
  DevBuf* bufs[] = {&plan->images, &plan->xy, &plan->ang, &plan->desc, &plan->nkp, &plan->matches, &plan->nmatch,
                    &plan->jobs, &plan->cand, &plan->candcnt, &plan->imgmax, &plan->selws, &plan->inliers,
                    &plan->ninl, &plan->epiws, &plan->mrow, &plan->mcol, &plan->descx, &plan->arena};
  for (DevBuf* b : bufs) b->release();
  for (int i = 0; i < 6; ++i)
    if (plan->ev[i]) cudaEventDestroy(plan->ev[i]);
  if (plan->graph_exec) cudaGraphExecDestroy(plan->graph_exec);
  if (plan->stream) cudaStreamDestroy(plan->stream);
  if (plan->stream2) cudaStreamDestroy(plan->stream2);
  if (plan->ev_fork) cudaEventDestroy(plan->ev_fork);
  if (plan->ev_join) cudaEventDestroy(plan->ev_join);
  delete plan;
}

static int plan_create_impl(vslam_ctx* ctx, int w, int h, int max_pairs, int cap, bool own_stream,
                            vslam_stereo_plan** out) {
  if (!ctx || !out || w < 2 || h < 2 || max_pairs <= 0 || cap <= 0) return vs_fail(ctx, VSLAM_E_ARG, "bad plan arguments");
  *out = nullptr;
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  vslam_stereo_plan* p = new (std::nothrow) vslam_stereo_plan();
  if (!p) return VSLAM_E_CUDA;
  p->ctx = ctx; p->w = w; p->h = h; p->max_pairs = max_pairs; p->cap = cap;
  const int nimg = 2 * max_pairs;
  const int dp = dev_pitch(w);
  p->im.w = w; p->im.h = h; p->im.pitch = dp; p->im.stride = (size_t)dp * h;
  int rc = VSLAM_OK;
  auto fail = [&](int code) {
    vslam_stereo_plan_destroy(p);
    return code;
  };
#define PLAN_CUDA(call)                                   \
  do {                                                    \
    if ((call) != cudaSuccess) {                          \
      ctx->err = std::string("plan: ") + #call + ": " + cudaGetErrorString(cudaGetLastError()); \
      return fail(VSLAM_E_CUDA);                          \
    }                                                     \
  } while (0)
  PLAN_CUDA(p->images.reserve(p->im.stride * nimg));
  p->im.base = p->images.as<uint8_t>();
  {
    const size_t b_xy = (size_t)nimg * cap * 16, b_ang = (size_t)nimg * cap * 8, b_desc = (size_t)nimg * cap * 32;
    const size_t b_m = (size_t)max_pairs * cap * 8, b_nkp = ((size_t)nimg * 4 + 15) & ~(size_t)15, b_nm = ((size_t)max_pairs * 4 + 15) & ~(size_t)15;
    const size_t total = b_xy + b_ang + b_desc + b_m + b_nkp + b_nm;
    if (total <= kArenaMaxBytes) {
      PLAN_CUDA(p->arena.reserve(total));
      uint8_t* a = p->arena.as<uint8_t>();
      p->xy.alias(a, b_xy); a += b_xy;
      p->ang.alias(a, b_ang); a += b_ang;
      p->desc.alias(a, b_desc); a += b_desc;
      p->matches.alias(a, b_m); a += b_m;
      p->nkp.alias(a, b_nkp); a += b_nkp;
      p->nmatch.alias(a, b_nm);
      p->arena_used = total;
    } else {
      PLAN_CUDA(p->xy.reserve(b_xy));
      PLAN_CUDA(p->ang.reserve(b_ang));
      PLAN_CUDA(p->desc.reserve(b_desc));
      PLAN_CUDA(p->nkp.reserve((size_t)nimg * 4));
      PLAN_CUDA(p->matches.reserve(b_m));
      PLAN_CUDA(p->nmatch.reserve((size_t)max_pairs * 4));
    }
  }
  PLAN_CUDA(p->jobs.reserve((size_t)max_pairs * sizeof(MatchJob)));
  rc = prepare_detect_ws(ctx, w, h, nimg, p->cand, p->candcnt, p->imgmax, p->selws, &p->ws);
  if (rc) return fail(rc);
  for (int i = 0; i < 6; ++i) PLAN_CUDA(cudaEventCreate(&p->ev[i]));
  if (own_stream) PLAN_CUDA(cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking));
  // the stereo jobs only depend on the plan's buffers: build them once
  std::vector<MatchJob> jobs(max_pairs);
  for (int k = 0; k < max_pairs; ++k) {
    MatchJob& j = jobs[k];
    j.d1 = p->desc.as<uint32_t>() + (size_t)(2 * k) * cap * 8;
    j.d2 = p->desc.as<uint32_t>() + (size_t)(2 * k + 1) * cap * 8;
    j.n1p = p->nkp.as<int32_t>() + 2 * k;
    j.n2p = p->nkp.as<int32_t>() + 2 * k + 1;
    j.n1 = j.n2 = 0;
    j.pairs = p->matches.as<int32_t>() + (size_t)k * cap * 2;
    j.count = p->nmatch.as<int32_t>() + k;
  }
  PLAN_CUDA(cudaMemcpy(p->jobs.p, jobs.data(), jobs.size() * sizeof(MatchJob), cudaMemcpyHostToDevice));
#undef PLAN_CUDA
  *out = p;
  return VSLAM_OK;
}

int vslam_stereo_plan_create(vslam_ctx* ctx, int w, int h, int max_pairs, int cap, vslam_stereo_plan** out) {
  return plan_create_impl(ctx, w, h, max_pairs, cap, false, out);
}

static cudaStream_t plan_stream(vslam_stereo_plan* p) { return p->stream ? p->stream : p->ctx->stream; }

int vslam_stereo_plan_upload(vslam_stereo_plan* plan, const uint8_t* images, int count, int pitch,
                             size_t image_stride) {
  if (!plan) return VSLAM_E_ARG;
  vslam_ctx* ctx = plan->ctx;
  if (!images || count < 0 || count > plan->max_pairs || pitch < plan->w) return vs_fail(ctx, VSLAM_E_ARG, "bad upload arguments");
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  DevImages im;
  return upload_images(ctx, plan->images, images, 2 * count, plan->w, plan->h, pitch, image_stride, &im, plan_stream(plan));
}

int vslam_stereo_plan_run(vslam_stereo_plan* plan, int count, int num_features, int rotate, int match_threshold,
                          double match_dist_2_best) {
  if (!plan) return VSLAM_E_ARG;
  vslam_ctx* ctx = plan->ctx;
  if (count < 0 || count > plan->max_pairs) return vs_fail(ctx, VSLAM_E_ARG, "bad pair count");
  if (num_features <= 0 || num_features > plan->cap) return vs_fail(ctx, VSLAM_E_ARG, "num_features must be in [1, cap]");
  if (count == 0) return VSLAM_OK;
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  // latency path: a small pipeline plan replays its launches as one graph (see vslam_stereo_plan::graph_state)
  const bool want_graph = ctx->use_graph && plan->arena_used && plan->stream && !plan->epi_on && plan->graph_state != 3;
  unsigned long long key = 1469598103934665603ull;
  if (want_graph) {
    unsigned long long rbits, kbits;
    memcpy(&rbits, &match_dist_2_best, 8);
    memcpy(&kbits, &ctx->harris_k, 8);
    const unsigned long long parts[] = {(unsigned long long)count, (unsigned long long)num_features, (unsigned long long)(rotate != 0),
                                        (unsigned long long)(unsigned)match_threshold, rbits, kbits,
                                        (unsigned long long)ctx->fp_contract, (unsigned long long)ctx->match_csa,
                                        (unsigned long long)ctx->harris_on, (unsigned long long)ctx->resp_band,
                                        (unsigned long long)ctx->match_split_small, (unsigned long long)ctx->match_mma,
                                        (unsigned long long)ctx->resp_multi};
    for (unsigned long long v : parts) key = (key ^ v) * 1099511628211ull;
    if (plan->graph_state == 2 && plan->graph_key == key) {
      VS_CUDA(ctx, cudaGraphLaunch(plan->graph_exec, plan->stream));
      ctx->launches += plan->graph_launches;
      plan->timed = false;  // no per-stage events inside a replay
      return VSLAM_OK;
    }
    if (plan->graph_state == 2) {  // other arguments: start over
      cudaGraphExecDestroy(plan->graph_exec);
      plan->graph_exec = nullptr;
      plan->graph_state = 0;
    }
  }
  const bool capture = want_graph && plan->graph_state == 1 && plan->graph_key == key;
  const long long launches0 = ctx->launches;
  cudaStream_t saved = ctx->stream;
  ctx->stream = plan_stream(plan);
  if (capture && cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeRelaxed) != cudaSuccess) {
    ctx->stream = saved;
    plan->graph_state = 3;
    return vslam_stereo_plan_run(plan, count, num_features, rotate, match_threshold, match_dist_2_best);
  }
  const cudaStream_t s0 = ctx->stream;  // the plan's stream (ctx->stream is pointed at the half's stream launch by launch)
#define PLAN_EVENT(i) do { if (!capture) cudaEventRecord(plan->ev[i], s0); } while (0)
  int rc = VSLAM_OK;
  // enough jobs to fill the GPU with one CTA per SM: Hamming matrix on the tensor cores (the describe kernel then also
  // writes every descriptor as 256 signed bytes, the matcher's operand format); else the POPC kernels, whose
  // column-split path spreads a few jobs over all SMs
  const bool mma = (ctx->match_mma == 2 || (ctx->match_mma && 2 * count >= ctx->num_sms)) && match_mma_fits(plan->cap, plan->cap);
  if (mma && !plan->descx.p) {  // first tensor-core run of this plan (no early return: ctx->stream is swapped)
    if (plan->descx.reserve((size_t)2 * plan->max_pairs * plan->cap * 256) != cudaSuccess ||
        cudaMemsetAsync(plan->descx.p, 0, plan->descx.bytes, ctx->stream) != cudaSuccess)
      rc = vs_fail(ctx, VSLAM_E_CUDA, "plan: cannot allocate the signed-byte descriptor slab");
  }
  // Large resident batches run as TWO halves on two streams, their launches issued stage by stage in turn: every kernel
  // here fills the registers or the shared memory of an SM by itself, so kernels never share an SM, but the last wave of
  // one half's kernel leaves SMs idle that the other half's next kernel takes (measured with two plans on two streams:
  // +5.3 % pairs/s at 2000 pairs).  The halves only share read-only tables; the second stream is forked from and joined
  // to the plan's stream with events, so the call keeps its stream semantics.  VSLAM_OPT_TWO_STREAMS = 0: one stream,
  // whole-batch launches (what the per-stage times of vslam_stereo_plan_last_ms describe without overlap).
  // (tensor-core matcher only: the POPC matcher's column-split scratch belongs to the plan, not to a half)
  const bool two = ctx->two_streams && !capture && !plan->epi_on && mma && count >= 4 * ctx->num_sms && !rc;
  if (two && !plan->stream2) {
    if (cudaStreamCreateWithFlags(&plan->stream2, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&plan->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&plan->ev_join, cudaEventDisableTiming) != cudaSuccess)
      rc = vs_fail(ctx, VSLAM_E_CUDA, "plan: cannot create the second stream");
  }
  const int nhalf = two ? 2 : 1;
  const int first_pair[2] = {0, two ? count / 2 : 0};
  const int npairs[2] = {two ? count / 2 : count, count - count / 2};
  cudaStream_t hs[2] = {s0, plan->stream2};
  const bool tiny = plan->w < 2 * kEdgeThreshold + 1 || plan->h < 2 * kEdgeThreshold + 1;
  PLAN_EVENT(0);
  if (two && !rc) {
    cudaEventRecord(plan->ev_fork, hs[0]);
    cudaStreamWaitEvent(hs[1], plan->ev_fork, 0);
  }
  auto sub_ws = [&](int img0) {
    DetectWorkspace w = plan->ws;
    w.cand += (size_t)img0 * w.cand_cap;
    w.cand_cnt += img0;
    w.img_max += img0;
    w.sel_ws += (size_t)img0 * w.sel_ws_stride;
    return w;
  };
  auto sub_im = [&](int img0) {
    DevImages im = plan->im;
    im.base += (size_t)img0 * im.stride;
    return im;
  };
  const size_t cap = plan->cap;
  // stage 1: response
  for (int hf = 0; hf < nhalf && !rc; ++hf) {
    ctx->stream = hs[hf];
    const int img0 = 2 * first_pair[hf], n = 2 * npairs[hf];
    if (tiny) cudaMemsetAsync(plan->nkp.as<int32_t>() + img0, 0, (size_t)n * 4, ctx->stream);
    else rc = launch_response(ctx, sub_im(img0), n, sub_ws(img0), nullptr);
  }
  PLAN_EVENT(1);
  // stage 2: select
  for (int hf = 0; hf < nhalf && !rc && !tiny; ++hf) {
    ctx->stream = hs[hf];
    const int img0 = 2 * first_pair[hf], n = 2 * npairs[hf];
    rc = launch_select(ctx, sub_im(img0), n, sub_ws(img0), num_features, kQuality, kMinDistance, kEdgeThreshold,
                       plan->xy.as<double>() + (size_t)img0 * cap * 2, plan->nkp.as<int32_t>() + img0, plan->cap);
  }
  PLAN_EVENT(2);
  // stage 3: orientation + descriptors
  for (int hf = 0; hf < nhalf && !rc; ++hf) {
    ctx->stream = hs[hf];
    const int img0 = 2 * first_pair[hf], n = 2 * npairs[hf];
    rc = launch_orient_describe(ctx, sub_im(img0), n, plan->xy.as<double>() + (size_t)img0 * cap * 2,
                                plan->nkp.as<int32_t>() + img0, plan->cap, rotate ? 1 : 0,
                                plan->ang.as<double>() + (size_t)img0 * cap, plan->desc.as<uint32_t>() + (size_t)img0 * cap * 8,
                                mma ? plan->descx.as<uint8_t>() + (size_t)img0 * cap * 256 : nullptr);
  }
  PLAN_EVENT(3);
  // stage 4: stereo matching
  for (int hf = 0; hf < nhalf && !rc; ++hf) {
    ctx->stream = hs[hf];
    const MatchJob* jobs = plan->jobs.as<MatchJob>() + first_pair[hf];
    if (mma) {
      const long long rows = (long long)2 * count * plan->cap;
      const MatchExpanded x = {plan->desc.as<uint32_t>(), plan->descx.as<uint8_t>(), rows};
      rc = launch_match_jobs_mma(ctx, jobs, npairs[hf], plan->cap, plan->cap, match_threshold, match_dist_2_best, plan->cap, x, x);
    } else {
      rc = launch_match_jobs(ctx, jobs, npairs[hf], plan->cap, plan->cap, match_threshold, match_dist_2_best, plan->cap,
                             &plan->mrow, &plan->mcol);
    }
  }
  ctx->stream = hs[0];
  if (two) {  // (also after an error: the second stream must not be left forked)
    cudaEventRecord(plan->ev_join, hs[1]);
    cudaStreamWaitEvent(hs[0], plan->ev_join, 0);
  }
  PLAN_EVENT(4);
  if (!rc && plan->epi_on) {
    // matching_utils.h:63-88 on the matcher's device output: left corners of pair k are image 2k, right 2k+1
    rc = launch_essential_jobs(ctx, &plan->epi.cam1, &plan->epi.cam2, plan->epi.E, plan->epi.epipolar_error_threshold,
                               plan->xy.as<double>(), plan->xy.as<double>() + (size_t)plan->cap * 2,
                               (size_t)plan->cap * 4, plan->matches.as<int32_t>(), (size_t)plan->cap,
                               plan->nmatch.as<int32_t>(), plan->cap, count, plan->epiws.as<uint8_t>(),
                               plan->inliers.as<int32_t>(), (size_t)plan->cap, plan->cap, plan->ninl.as<int32_t>(),
                               nullptr);
  }
  PLAN_EVENT(5);
#undef PLAN_EVENT
  plan->timed = !capture;
  if (capture) {
    cudaGraph_t g = nullptr;
    const cudaError_t e = cudaStreamEndCapture(ctx->stream, &g);
    ctx->stream = saved;
    if (rc == VSLAM_OK && e == cudaSuccess && g &&
        cudaGraphInstantiate(&plan->graph_exec, g, 0) == cudaSuccess) {
      plan->graph_state = 2;
      plan->graph_launches = ctx->launches - launches0;
      cudaGraphDestroy(g);
      VS_CUDA(ctx, cudaGraphLaunch(plan->graph_exec, plan->stream));
      return VSLAM_OK;
    }
    if (g) cudaGraphDestroy(g);
    cudaGetLastError();
    plan->graph_state = 3;     // nothing has run: do it the ordinary way, and never try again on this plan
    ctx->launches = launches0;
    return vslam_stereo_plan_run(plan, count, num_features, rotate, match_threshold, match_dist_2_best);
  }
  if (want_graph && rc == VSLAM_OK) {
    plan->graph_state = 1;
    plan->graph_key = key;
  }
  ctx->stream = saved;
  return rc;
}

int vslam_stereo_plan_download(vslam_stereo_plan* plan, int count, double* xy, double* angles, uint32_t* desc,
                               int32_t* n_kp, int32_t* matches, int32_t* n_matches) {
  if (!plan) return VSLAM_E_ARG;
  vslam_ctx* ctx = plan->ctx;
  if (count < 0 || count > plan->max_pairs) return vs_fail(ctx, VSLAM_E_ARG, "bad pair count");
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t s = plan_stream(plan);
  const size_t nimg = 2 * (size_t)count, cap = plan->cap;
  if (xy) VS_CUDA(ctx, cudaMemcpyAsync(xy, plan->xy.p, nimg * cap * 16, cudaMemcpyDeviceToHost, s));
  if (angles) VS_CUDA(ctx, cudaMemcpyAsync(angles, plan->ang.p, nimg * cap * 8, cudaMemcpyDeviceToHost, s));
  if (desc) VS_CUDA(ctx, cudaMemcpyAsync(desc, plan->desc.p, nimg * cap * 32, cudaMemcpyDeviceToHost, s));
  if (n_kp) VS_CUDA(ctx, cudaMemcpyAsync(n_kp, plan->nkp.p, nimg * 4, cudaMemcpyDeviceToHost, s));
  if (matches) VS_CUDA(ctx, cudaMemcpyAsync(matches, plan->matches.p, (size_t)count * cap * 8, cudaMemcpyDeviceToHost, s));
  if (n_matches) VS_CUDA(ctx, cudaMemcpyAsync(n_matches, plan->nmatch.p, (size_t)count * 4, cudaMemcpyDeviceToHost, s));
  VS_CUDA(ctx, cudaStreamSynchronize(s));
  return VSLAM_OK;
}

int vslam_stereo_plan_last_ms(vslam_stereo_plan* plan, float* ms) {
  if (!plan || !ms) return VSLAM_E_ARG;
  vslam_ctx* ctx = plan->ctx;
  if (!plan->timed) return vs_fail(ctx, VSLAM_E_ARG, "plan has not run");
  VS_CUDA(ctx, cudaEventSynchronize(plan->ev[5]));
  VS_CUDA(ctx, cudaEventElapsedTime(&ms[0], plan->ev[0], plan->ev[5]));
  for (int i = 0; i < 4; ++i) VS_CUDA(ctx, cudaEventElapsedTime(&ms[1 + i], plan->ev[i], plan->ev[i + 1]));
  return VSLAM_OK;
}

int vslam_stereo_plan_set_epipolar(vslam_stereo_plan* plan, const vslam_epipolar* ep) {
  if (!plan) return VSLAM_E_ARG;
  vslam_ctx* ctx = plan->ctx;
  if (!ep) {
    plan->epi_on = false;
    return VSLAM_OK;
  }
  for (const vslam_camera* c : {&ep->cam1, &ep->cam2})
    if (c->model < VSLAM_CAM_PINHOLE || c->model > VSLAM_CAM_KB4) return vs_fail(ctx, VSLAM_E_ARG, "bad camera model");
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  VS_CUDA(ctx, plan->inliers.reserve((size_t)plan->max_pairs * plan->cap * 8));
  VS_CUDA(ctx, plan->ninl.reserve((size_t)plan->max_pairs * 4));
  VS_CUDA(ctx, plan->epiws.reserve(essential_ws_bytes(plan->cap, plan->max_pairs)));  // sized for max_pairs jobs: run(count <= max_pairs) uses a prefix
  plan->epi = *ep;
  plan->epi_on = true;
  return VSLAM_OK;
}

int vslam_stereo_plan_download_inliers(vslam_stereo_plan* plan, int count, int32_t* inliers, int32_t* n_inliers) {
  if (!plan) return VSLAM_E_ARG;
  vslam_ctx* ctx = plan->ctx;
  if (count < 0 || count > plan->max_pairs) return vs_fail(ctx, VSLAM_E_ARG, "bad pair count");
  if (!plan->epi_on) return vs_fail(ctx, VSLAM_E_ARG, "the plan has no epipolar stage (vslam_stereo_plan_set_epipolar)");
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t s = plan_stream(plan);
  if (inliers) VS_CUDA(ctx, cudaMemcpyAsync(inliers, plan->inliers.p, (size_t)count * plan->cap * 8, cudaMemcpyDeviceToHost, s));
  if (n_inliers) VS_CUDA(ctx, cudaMemcpyAsync(n_inliers, plan->ninl.p, (size_t)count * 4, cudaMemcpyDeviceToHost, s));
  VS_CUDA(ctx, cudaStreamSynchronize(s));
  return VSLAM_OK;
}

int vslam_stereo_plan_last_epipolar_ms(vslam_stereo_plan* plan, float* ms) {
  if (!plan || !ms) return VSLAM_E_ARG;
  vslam_ctx* ctx = plan->ctx;
  if (!plan->timed) return vs_fail(ctx, VSLAM_E_ARG, "plan has not run");
  VS_CUDA(ctx, cudaEventSynchronize(plan->ev[5]));
  VS_CUDA(ctx, cudaEventElapsedTime(ms, plan->ev[4], plan->ev[5]));
  return VSLAM_OK;
}

// ---- batched front ends on host buffers -------------------------------------------------------------

}  // extern "C"

static int stereo_batch_impl(vslam_ctx* ctx, const uint8_t* images, int count, const vslam_stereo_params* sp,
                             const vslam_epipolar* ep, double* xy, double* angles, uint32_t* desc, int32_t* n_kp,
                             int32_t* matches, int32_t* n_matches, int32_t* inliers, int32_t* n_inliers, int cap) {
  if (!ctx) return VSLAM_E_ARG;
  if (!images || !sp || count < 0 || cap <= 0) return vs_fail(ctx, VSLAM_E_ARG, "bad batch arguments");
  if (sp->w < 2 || sp->h < 2 || sp->pitch < sp->w || sp->num_features <= 0 || sp->num_features > cap)
    return vs_fail(ctx, VSLAM_E_ARG, "bad stereo parameters (num_features must be in [1, cap])");
  if (count == 0) return VSLAM_OK;
  // `batch_depth` plans, each on its own stream: while chunk k computes, chunk k+1 uploads and chunk
  // k-1 downloads.  Stream order serialises a plan's next upload after its own download, so three
  // plans are needed for the three phases to overlap freely.
  const int chunk = count < ctx->batch_chunk ? count : ctx->batch_chunk;
  vslam_stereo_plan* plans[vslam_ctx::kMaxBatchDepth] = {};
  int rc = VSLAM_OK;
  // Chunk schedule: the pipeline's fill (first upload) and drain (last compute + download) are not
  // overlapped, so the first and last chunks are short (chunk/4, chunk/2) when the batch is long enough.
  std::vector<int> sizes;
  {
    const int q = chunk / 4, hlf = chunk / 2;
    if (q >= 8 && count >= 2 * (q + hlf) + chunk) {
      int mid = count - 2 * (q + hlf);
      sizes.push_back(q);
      sizes.push_back(hlf);
      while (mid > 0) {
        const int c = mid < chunk ? mid : chunk;
        sizes.push_back(c);
        mid -= c;
      }
      sizes.push_back(hlf);
      sizes.push_back(q);
    } else {
      for (int left = count; left > 0; left -= chunk) sizes.push_back(left < chunk ? left : chunk);
    }
  }
  const int nchunks = (int)sizes.size();
  const int nplans = nchunks < ctx->batch_depth ? nchunks : ctx->batch_depth;
  for (int i = 0; i < nplans && !rc; ++i) {
    vslam_stereo_plan*& cp = ctx->batch_plans[i];
    if (cp && (cp->w != sp->w || cp->h != sp->h || cp->cap != cap || cp->max_pairs < chunk)) {
      vslam_stereo_plan_destroy(cp);
      cp = nullptr;
    }
    if (!cp) rc = plan_create_impl(ctx, sp->w, sp->h, chunk, cap, true, &cp);
    if (!rc) rc = vslam_stereo_plan_set_epipolar(cp, ep);
    plans[i] = cp;
  }
  int bad_kp = 0;
  const vslam_stereo_plan* arena_plan = nullptr;
  // keypoint counts / detector statuses land in the caller's n_kp, or in the ctx's own pinned block when the caller
  // passes none (an overflowed image must not come back as "no matches, OK")
  int32_t* kp_status_host = n_kp;
  if (!kp_status_host && !rc) {
    const size_t need = 2 * (size_t)count * sizeof(int32_t);
    if (need > ctx->pinned_status_bytes) {
      if (ctx->pinned_status) cudaFreeHost(ctx->pinned_status);
  if (ctx->pinned_out) cudaFreeHost(ctx->pinned_out);
      ctx->pinned_status = nullptr;
      ctx->pinned_status_bytes = 0;
      if (cudaMallocHost(&ctx->pinned_status, need) != cudaSuccess) rc = vs_fail(ctx, VSLAM_E_CUDA, "stereo batch: pinned status block");
      else ctx->pinned_status_bytes = need;
    }
    kp_status_host = static_cast<int32_t*>(ctx->pinned_status);
  }
  for (int c0 = 0, k = 0; k < nchunks && !rc; c0 += sizes[k], ++k) {
    vslam_stereo_plan* p = plans[k % nplans];
    const int nc = sizes[k];
    // stream order on p->stream serialises this chunk after the plan's previous download
    rc = vslam_stereo_plan_upload(p, images + sp->image_stride * 2 * (size_t)c0, nc, sp->pitch, sp->image_stride);
    if (!rc) rc = vslam_stereo_plan_run(p, nc, sp->num_features, sp->rotate_features, sp->match_threshold, sp->match_dist_2_best);
    if (rc) break;
    cudaStream_t s = p->stream;
    const size_t i0 = 2 * (size_t)c0, ni = 2 * (size_t)nc, cp = cap;
    // no early return inside the pipeline: queued copies of earlier chunks still write into the caller's buffers, so
    // every exit goes through the stream synchronisation below
    cudaError_t e = cudaSuccess;
    auto d2h = [&](void* dst, const void* src, size_t bytes) {
      if (e == cudaSuccess) e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, s);
    };
    if (nchunks == 1 && p->arena_used && nc == p->max_pairs && !ep) {
      // latency path (one chunk, a few pairs: config 1 as slam.cpp:1125-1143 issues it): ONE copy of the plan's result
      // arena into the ctx's pinned block; after the synchronisation below the CPU hands the rows that exist to the
      // caller's arrays (six small copies on the stream cost ~7 us each, and far more into pageable memory)
      if (p->arena_used > ctx->pinned_out_bytes) {
        if (ctx->pinned_out) cudaFreeHost(ctx->pinned_out);
        ctx->pinned_out = nullptr;
        ctx->pinned_out_bytes = 0;
        e = cudaMallocHost(&ctx->pinned_out, p->arena_used);
        if (e == cudaSuccess) ctx->pinned_out_bytes = p->arena_used;
      }
      d2h(ctx->pinned_out, p->arena.p, p->arena_used);
      if (e == cudaSuccess) {
        arena_plan = p;
        continue;
      }
    }
    if (xy) d2h(xy + i0 * cp * 2, p->xy.p, ni * cp * 16);
    if (angles) d2h(angles + i0 * cp, p->ang.p, ni * cp * 8);
    if (desc) d2h(desc + i0 * cp * 8, p->desc.p, ni * cp * 32);
    d2h(kp_status_host + i0, p->nkp.p, ni * 4);  // the detector's overflow statuses are checked whatever the caller asks for
    if (matches) d2h(matches + (size_t)c0 * cp * 2, p->matches.p, (size_t)nc * cp * 8);
    if (n_matches) d2h(n_matches + c0, p->nmatch.p, (size_t)nc * 4);
    if (ep && inliers) d2h(inliers + (size_t)c0 * cp * 2, p->inliers.p, (size_t)nc * cp * 8);
    if (ep && n_inliers) d2h(n_inliers + c0, p->ninl.p, (size_t)nc * 4);
    if (e != cudaSuccess) {
      ctx->err = std::string("stereo batch: cudaMemcpyAsync: ") + cudaGetErrorString(e);
      rc = VSLAM_E_CUDA;
    }
  }
  for (int i = 0; i < nplans; ++i)
    if (plans[i]) cudaStreamSynchronize(plans[i]->stream);
  if (!rc && arena_plan) {
    // hand the arena out: rows [0, n_kp) of every image, [0, n_matches) of every pair (the slabs' tails stay untouched,
    // as undefined as the device left them before)
    const vslam_stereo_plan* p = arena_plan;
    const uint8_t* a = static_cast<const uint8_t*>(ctx->pinned_out);
    const size_t nimg = 2 * (size_t)count, cp = cap;
    const double* h_xy = reinterpret_cast<const double*>(a + ((uint8_t*)p->xy.p - (uint8_t*)p->arena.p));
    const double* h_ang = reinterpret_cast<const double*>(a + ((uint8_t*)p->ang.p - (uint8_t*)p->arena.p));
    const uint32_t* h_desc = reinterpret_cast<const uint32_t*>(a + ((uint8_t*)p->desc.p - (uint8_t*)p->arena.p));
    const int32_t* h_m = reinterpret_cast<const int32_t*>(a + ((uint8_t*)p->matches.p - (uint8_t*)p->arena.p));
    const int32_t* h_nkp = reinterpret_cast<const int32_t*>(a + ((uint8_t*)p->nkp.p - (uint8_t*)p->arena.p));
    const int32_t* h_nm = reinterpret_cast<const int32_t*>(a + ((uint8_t*)p->nmatch.p - (uint8_t*)p->arena.p));
    for (size_t i = 0; i < nimg; ++i) {
      const size_t n = h_nkp[i] > 0 ? (size_t)(h_nkp[i] < cap ? h_nkp[i] : cap) : 0;
      kp_status_host[i] = h_nkp[i];
      if (xy) memcpy(xy + i * cp * 2, h_xy + i * cp * 2, n * 16);
      if (angles) memcpy(angles + i * cp, h_ang + i * cp, n * 8);
      if (desc) memcpy(desc + i * cp * 8, h_desc + i * cp * 8, n * 32);
    }
    for (size_t k = 0; k < (size_t)count; ++k) {
      const size_t n = h_nm[k] > 0 ? (size_t)(h_nm[k] < cap ? h_nm[k] : cap) : 0;
      if (n_matches) n_matches[k] = h_nm[k];
      if (matches) memcpy(matches + k * cp * 2, h_m + k * cp * 2, n * 8);
    }
  }
  if (!rc)
    for (size_t i = 0; i < 2 * (size_t)count; ++i)
      if (kp_status_host[i] < 0) bad_kp = kp_status_host[i];
  if (!rc && bad_kp) rc = kp_status(ctx, bad_kp);
  return rc;
}

extern "C" {

int vslam_stereo_batch(vslam_ctx* ctx, const uint8_t* images, int count, const vslam_stereo_params* sp, double* xy,
                       double* angles, uint32_t* desc, int32_t* n_kp, int32_t* matches, int32_t* n_matches, int cap) {
  return stereo_batch_impl(ctx, images, count, sp, nullptr, xy, angles, desc, n_kp, matches, n_matches, nullptr,
                           nullptr, cap);
}

int vslam_stereo_batch_epipolar(vslam_ctx* ctx, const uint8_t* images, int count, const vslam_stereo_params* sp,
                                const vslam_epipolar* ep, double* xy, double* angles, uint32_t* desc, int32_t* n_kp,
                                int32_t* matches, int32_t* n_matches, int32_t* inliers, int32_t* n_inliers, int cap) {
  if (ctx && !ep) return vs_fail(ctx, VSLAM_E_ARG, "epipolar parameters missing");
  return stereo_batch_impl(ctx, images, count, sp, ep, xy, angles, desc, n_kp, matches, n_matches, inliers, n_inliers,
                           cap);
}

int vslam_detect_describe_batch(vslam_ctx* ctx, const uint8_t* images, int count, int w, int h, int pitch,
                                size_t image_stride, int num_features, int rotate, double* xy, double* angles,
                                uint32_t* desc, int32_t* n_kp, int cap) {
  int rc = check_image(ctx, images, w, h, pitch);
  if (rc) return rc;
  if (count < 0 || cap <= 0 || num_features <= 0 || num_features > cap || !xy || !angles || !desc || !n_kp)
    return vs_fail(ctx, VSLAM_E_ARG, "bad batch arguments");
  if (count == 0) return VSLAM_OK;
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  if (w < 2 * kEdgeThreshold + 1 || h < 2 * kEdgeThreshold + 1) {
    for (int i = 0; i < count; ++i) n_kp[i] = 0;
    return VSLAM_OK;
  }
  const int chunk = 256;
  for (int c0 = 0; c0 < count; c0 += chunk) {
    const int nc = count - c0 < chunk ? count - c0 : chunk;
    DevImages im;
    DetectWorkspace ws;
    rc = upload_images(ctx, ctx->img, images + image_stride * (size_t)c0, nc, w, h, pitch, image_stride, &im, ctx->stream);
    if (rc) return rc;
    rc = prepare_detect_ws(ctx, w, h, nc, ctx->cand, ctx->candcnt, ctx->imgmax, ctx->selws, &ws);
    if (rc) return rc;
    VS_CUDA(ctx, ctx->xy.reserve((size_t)nc * cap * 16));
    VS_CUDA(ctx, ctx->ang.reserve((size_t)nc * cap * 8));
    VS_CUDA(ctx, ctx->desc.reserve((size_t)nc * cap * 32));
    VS_CUDA(ctx, ctx->cnt.reserve((size_t)nc * 4));
    rc = launch_response(ctx, im, nc, ws, nullptr);
    if (!rc) rc = launch_select(ctx, im, nc, ws, num_features, kQuality, kMinDistance, kEdgeThreshold, ctx->xy.as<double>(), ctx->cnt.as<int32_t>(), cap);
    if (!rc) rc = launch_orient_describe(ctx, im, nc, ctx->xy.as<double>(), ctx->cnt.as<int32_t>(), cap, rotate ? 1 : 0, ctx->ang.as<double>(), ctx->desc.as<uint32_t>());
    if (rc) return rc;
    const size_t cp = cap;
    VS_CUDA(ctx, cudaMemcpyAsync(xy + (size_t)c0 * cp * 2, ctx->xy.p, (size_t)nc * cp * 16, cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaMemcpyAsync(angles + (size_t)c0 * cp, ctx->ang.p, (size_t)nc * cp * 8, cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaMemcpyAsync(desc + (size_t)c0 * cp * 8, ctx->desc.p, (size_t)nc * cp * 32, cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaMemcpyAsync(n_kp + c0, ctx->cnt.p, (size_t)nc * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int i = 0; i < nc; ++i)
      if ((rc = kp_status(ctx, n_kp[c0 + i]))) return rc;
  }
  return VSLAM_OK;
}

int vslam_pyramid_detect_describe_batch(vslam_ctx* ctx, const uint8_t* images, int count, int w, int h, int pitch,
                                        size_t image_stride, int n_levels, const int* num_features, int rotate,
                                        double* xy, double* angles, uint32_t* desc, int32_t* n_kp, int cap) {
  int rc = check_image(ctx, images, w, h, pitch);
  if (rc) return rc;
  if (count < 0 || n_levels < 1 || n_levels > VSLAM_MAX_PYRAMID_LEVELS || !num_features || cap <= 0 || !xy ||
      !angles || !desc || !n_kp)
    return vs_fail(ctx, VSLAM_E_ARG, "bad pyramid arguments");
  long long total_feat = 0;
  for (int k = 0; k < n_levels; ++k) {
    if (num_features[k] <= 0) return vs_fail(ctx, VSLAM_E_ARG, "num_features[k] must be positive");
    total_feat += num_features[k];
  }
  if (total_feat > cap) return vs_fail(ctx, VSLAM_E_ARG, "cap must hold the sum of num_features");
  if (count == 0) return VSLAM_OK;
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  // level geometry: level k is (w >> k) x (h >> k)
  int lw[VSLAM_MAX_PYRAMID_LEVELS], lh[VSLAM_MAX_PYRAMID_LEVELS], lp[VSLAM_MAX_PYRAMID_LEVELS];
  size_t loff[VSLAM_MAX_PYRAMID_LEVELS + 1];
  const int chunk = count < 64 ? count : 64;
  loff[1] = 0;
  for (int k = 0; k < n_levels; ++k) {
    lw[k] = w >> k; lh[k] = h >> k; lp[k] = lw[k] > 0 ? dev_pitch(lw[k]) : 16;
    if (k >= 1) loff[k + 1] = loff[k] + align_up((size_t)lp[k] * (lh[k] > 0 ? lh[k] : 1) * chunk, 256);
  }
  if (n_levels > 1) VS_CUDA(ctx, ctx->pyr.reserve(loff[n_levels] + 256));
  const size_t cp = cap;
  for (int c0 = 0; c0 < count; c0 += chunk) {
    const int nc = count - c0 < chunk ? count - c0 : chunk;
    DevImages lv[VSLAM_MAX_PYRAMID_LEVELS];
    rc = upload_images(ctx, ctx->img, images + image_stride * (size_t)c0, nc, w, h, pitch, image_stride, &lv[0], ctx->stream);
    if (rc) return rc;
    VS_CUDA(ctx, ctx->xy.reserve((size_t)nc * cp * 16));
    VS_CUDA(ctx, ctx->ang.reserve((size_t)nc * cp * 8));
    VS_CUDA(ctx, ctx->desc.reserve((size_t)nc * cp * 32));
    VS_CUDA(ctx, ctx->cnt.reserve((size_t)nc * n_levels * 4));
    size_t feat_off = 0;
    for (int k = 0; k < n_levels; ++k) {
      if (k >= 1) {
        lv[k].base = ctx->pyr.as<uint8_t>() + loff[k];
        lv[k].w = lw[k]; lv[k].h = lh[k]; lv[k].pitch = lp[k]; lv[k].stride = (size_t)lp[k] * (lh[k] > 0 ? lh[k] : 1);
        rc = launch_pyr_down(ctx, lv[k - 1], lv[k], nc);
        if (rc) return rc;
      }
      int32_t* cnt_k = ctx->cnt.as<int32_t>() + (size_t)k * nc;  // level-major on the device
      if (lw[k] < 2 * kEdgeThreshold + 1 || lh[k] < 2 * kEdgeThreshold + 1) {
        VS_CUDA(ctx, cudaMemsetAsync(cnt_k, 0, (size_t)nc * 4, ctx->stream));  // nothing survives the 19 px border
      } else {
        DetectWorkspace ws;
        rc = prepare_detect_ws(ctx, lw[k], lh[k], nc, ctx->cand, ctx->candcnt, ctx->imgmax, ctx->selws, &ws);
        if (rc) return rc;
        // the level's outputs start feat_off entries into each image's slab (stride cap)
        double* xy_k = ctx->xy.as<double>() + feat_off * 2;
        rc = launch_response(ctx, lv[k], nc, ws, nullptr);
        if (!rc) rc = launch_select(ctx, lv[k], nc, ws, num_features[k], kQuality, kMinDistance, kEdgeThreshold, xy_k, cnt_k, cap);
        if (!rc) rc = launch_orient_describe(ctx, lv[k], nc, xy_k, cnt_k, cap, rotate ? 1 : 0,
                                             ctx->ang.as<double>() + feat_off, ctx->desc.as<uint32_t>() + feat_off * 8);
        if (rc) return rc;
      }
      feat_off += num_features[k];
    }
    VS_CUDA(ctx, cudaMemcpyAsync(xy + (size_t)c0 * cp * 2, ctx->xy.p, (size_t)nc * cp * 16, cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaMemcpyAsync(angles + (size_t)c0 * cp, ctx->ang.p, (size_t)nc * cp * 8, cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaMemcpyAsync(desc + (size_t)c0 * cp * 8, ctx->desc.p, (size_t)nc * cp * 32, cudaMemcpyDeviceToHost, ctx->stream));
    // n_kp is [count][n_levels] on the host: one strided copy per level
    for (int k = 0; k < n_levels; ++k)
      VS_CUDA(ctx, cudaMemcpy2DAsync(n_kp + (size_t)c0 * n_levels + k, (size_t)n_levels * 4,
                                     ctx->cnt.as<int32_t>() + (size_t)k * nc, 4, 4, nc, cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int i = 0; i < nc * n_levels; ++i)
      if ((rc = kp_status(ctx, n_kp[(size_t)c0 * n_levels + i]))) return rc;
  }
  return VSLAM_OK;
}

}  // extern "C"

// ---- descriptor store + batched matching --------------------------------------------------------------

struct vslam_desc_store {
  vslam_ctx* ctx = nullptr;
  DevBuf desc, offsets, counts;
  DevBuf descx;               // descriptors as signed bytes (tensor-core matcher), rows [0, expanded) are current
  int64_t expanded = 0;
  int n_sets = 0;
  int max_count = 0;
  // host mirrors, so that sets can be appended (vslam_desc_store_append) without reading the device back
  std::vector<int64_t> h_offsets;
  std::vector<int32_t> h_counts;
  int64_t total = 0;  // descriptors stored (== the next free slot)
};

// grow a device buffer to at least `need` bytes keeping its first `keep` bytes
static cudaError_t grow_preserve(DevBuf& b, size_t need, size_t keep) {
  if (need <= b.bytes) return cudaSuccess;
  void* np = nullptr;
  const size_t want = need * 2 + 256;
  cudaError_t e = cudaMalloc(&np, want);
  if (e != cudaSuccess) return e;
  if (b.p && keep) e = cudaMemcpy(np, b.p, keep, cudaMemcpyDeviceToDevice);
  if (e != cudaSuccess) { cudaFree(np); return e; }
  if (b.p) cudaFree(b.p);
  b.p = np;
  b.bytes = want;
  return cudaSuccess;
}

extern "C" {

void vslam_desc_store_destroy(vslam_desc_store* s) {
  if (!s) return;
  cudaSetDevice(s->ctx->device);
  cudaDeviceSynchronize();
  s->desc.release();
  s->descx.release();
  s->offsets.release();
  s->counts.release();
  delete s;
}

int vslam_desc_store_create(vslam_ctx* ctx, const uint32_t* desc, const int64_t* offsets, const int32_t* counts,
                            int n_sets, vslam_desc_store** out) {
  if (!ctx) return VSLAM_E_ARG;
  if (!out || !desc || !offsets || !counts || n_sets <= 0) return vs_fail(ctx, VSLAM_E_ARG, "bad store arguments");
  *out = nullptr;
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  int64_t total = 0;
  int maxc = 0;
  for (int i = 0; i < n_sets; ++i) {
    if (counts[i] < 0 || offsets[i] < 0) return vs_fail(ctx, VSLAM_E_ARG, "negative offset/count");
    if (offsets[i] + counts[i] > total) total = offsets[i] + counts[i];
    if (counts[i] > maxc) maxc = counts[i];
  }
  vslam_desc_store* s = new (std::nothrow) vslam_desc_store();
  if (!s) return VSLAM_E_CUDA;
  s->ctx = ctx; s->n_sets = n_sets; s->max_count = maxc;
  s->h_offsets.assign(offsets, offsets + n_sets);
  s->h_counts.assign(counts, counts + n_sets);
  s->total = total;
  cudaError_t e = s->desc.reserve((size_t)(total > 0 ? total : 1) * 32);
  if (e == cudaSuccess) e = s->offsets.reserve((size_t)n_sets * 8);
  if (e == cudaSuccess) e = s->counts.reserve((size_t)n_sets * 4);
  if (e == cudaSuccess && total > 0) e = cudaMemcpy(s->desc.p, desc, (size_t)total * 32, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(s->offsets.p, offsets, (size_t)n_sets * 8, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(s->counts.p, counts, (size_t)n_sets * 4, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    ctx->err = std::string("desc store: ") + cudaGetErrorString(e);
    vslam_desc_store_destroy(s);
    return VSLAM_E_CUDA;
  }
  *out = s;
  return VSLAM_OK;
}

int vslam_desc_store_append(vslam_desc_store* s, const uint32_t* desc, int count, int* set_index) {
  if (!s) return VSLAM_E_ARG;
  vslam_ctx* ctx = s->ctx;
  if (count < 0 || (count > 0 && !desc)) return vs_fail(ctx, VSLAM_E_ARG, "bad append arguments");
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // no matching job may still read the buffers that move
  const int idx = s->n_sets;
  VS_CUDA(ctx, grow_preserve(s->desc, (size_t)(s->total + count + 1) * 32, (size_t)s->total * 32));
  VS_CUDA(ctx, grow_preserve(s->offsets, (size_t)(idx + 1) * 8, (size_t)idx * 8));
  VS_CUDA(ctx, grow_preserve(s->counts, (size_t)(idx + 1) * 4, (size_t)idx * 4));
  const int64_t off = s->total;
  const int32_t cnt = count;
  if (count > 0)
    VS_CUDA(ctx, cudaMemcpy(s->desc.as<uint8_t>() + (size_t)off * 32, desc, (size_t)count * 32, cudaMemcpyHostToDevice));
  VS_CUDA(ctx, cudaMemcpy(s->offsets.as<int64_t>() + idx, &off, 8, cudaMemcpyHostToDevice));
  VS_CUDA(ctx, cudaMemcpy(s->counts.as<int32_t>() + idx, &cnt, 4, cudaMemcpyHostToDevice));
  s->h_offsets.push_back(off);
  s->h_counts.push_back(cnt);
  s->total += count;
  s->n_sets = idx + 1;
  if (count > s->max_count) s->max_count = count;
  if (set_index) *set_index = idx;
  return VSLAM_OK;
}

int vslam_desc_store_size(const vslam_desc_store* s) { return s ? s->n_sets : 0; }

int vslam_match_batch(vslam_ctx* ctx, const vslam_desc_store* s, const int32_t* jobs, int n_jobs, int threshold,
                      double dist_2_best, int32_t* matches, int32_t* n_matches, int cap) {
  if (!ctx) return VSLAM_E_ARG;
  if (!s || s->ctx != ctx || !jobs || n_jobs < 0 || cap <= 0) return vs_fail(ctx, VSLAM_E_ARG, "bad batch arguments");
  if (n_jobs == 0) return VSLAM_OK;
  for (int k = 0; k < 2 * n_jobs; ++k)
    if (jobs[k] < 0 || jobs[k] >= s->n_sets) return vs_fail(ctx, VSLAM_E_ARG, "match batch: set index out of range");
  if (!match_fits_fused(s->max_count, s->max_count))
    return vs_fail(ctx, VSLAM_E_LIMIT, "descriptor sets too large for batched matching; use vslam_match");
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  // bounded device footprint: jobs run in chunks whose match slabs are copied back as they finish
  const size_t per_job = (size_t)cap * 8;
  int chunk = (int)((size_t)(1u << 30) / per_job);
  if (chunk < 1) chunk = 1;
  if (chunk > n_jobs) chunk = n_jobs;
  VS_CUDA(ctx, ctx->pairs.reserve((size_t)chunk * per_job));
  VS_CUDA(ctx, ctx->mcnt.reserve((size_t)chunk * 4));
  VS_CUDA(ctx, ctx->jobs.reserve((size_t)chunk * (sizeof(MatchJob) + 8)));
  MatchJob* jobs_dev = ctx->jobs.as<MatchJob>();
  int32_t* ji_dev = reinterpret_cast<int32_t*>(jobs_dev + chunk);
  // batches that fill the GPU run on the tensor cores: bring the signed-byte copy of the store up to date first
  const bool use_mma = (ctx->match_mma == 2 || (ctx->match_mma && 2 * n_jobs >= ctx->num_sms)) && match_mma_fits(s->max_count, s->max_count) && s->total > 0;
  if (use_mma && s->expanded < s->total) {
    vslam_desc_store* ms = const_cast<vslam_desc_store*>(s);
    VS_CUDA(ctx, grow_preserve(ms->descx, (size_t)(s->total + 1) * 256, (size_t)s->expanded * 256));
    const int rc = launch_expand_desc(ctx, s->desc.as<uint32_t>() + (size_t)s->expanded * 8, s->total - s->expanded, 1, nullptr,
                                      ms->descx.as<uint8_t>() + (size_t)s->expanded * 256);
    if (rc) return rc;
    ms->expanded = s->total;
  }
  float total_ms = 0.f;
  int status = VSLAM_OK;
  for (int c0 = 0; c0 < n_jobs; c0 += chunk) {
    const int nc = n_jobs - c0 < chunk ? n_jobs - c0 : chunk;
    VS_CUDA(ctx, cudaMemcpyAsync(ji_dev, jobs + 2 * (size_t)c0, (size_t)nc * 8, cudaMemcpyHostToDevice, ctx->stream));
    VS_CUDA(ctx, cudaEventRecord(ctx->ev[0], ctx->stream));
    int rc = launch_build_store_jobs(ctx, ji_dev, nc, s->desc.as<uint32_t>(), s->offsets.as<int64_t>(),
                                     s->counts.as<int32_t>(), s->n_sets, ctx->pairs.as<int32_t>(),
                                     ctx->mcnt.as<int32_t>(), cap, jobs_dev);
    if (!rc) {
      if (use_mma) {
        const MatchExpanded x = {s->desc.as<uint32_t>(), s->descx.as<uint8_t>(), (long long)s->total};
        rc = launch_match_jobs_mma(ctx, jobs_dev, nc, s->max_count, s->max_count, threshold, dist_2_best, cap, x, x);
      } else {
        rc = launch_match_jobs(ctx, jobs_dev, nc, s->max_count, s->max_count, threshold, dist_2_best, cap);
      }
    }
    if (rc) return rc;
    VS_CUDA(ctx, cudaEventRecord(ctx->ev[1], ctx->stream));
    if (matches) VS_CUDA(ctx, cudaMemcpyAsync(matches + (size_t)c0 * cap * 2, ctx->pairs.p, (size_t)nc * per_job, cudaMemcpyDeviceToHost, ctx->stream));
    if (n_matches) VS_CUDA(ctx, cudaMemcpyAsync(n_matches + c0, ctx->mcnt.p, (size_t)nc * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0.f;
    VS_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]));
    total_ms += ms;
    // the same contract as vslam_match: a list longer than cap is an error, not a silently truncated result
    if (n_matches)
      for (int k = 0; k < nc; ++k) {
        const int32_t c = n_matches[c0 + k];
        if (c < 0) status = status ? status : VSLAM_E_LIMIT;
        else if (c > cap) status = status ? status : VSLAM_E_CAPACITY;
      }
  }
  ctx->last_match_ms = total_ms;
  if (status == VSLAM_E_CAPACITY) return vs_fail(ctx, status, "match batch: a job has more matches than cap (n_matches holds the full counts, the lists their first cap pairs)");
  if (status) return vs_fail(ctx, status, "match batch: a job failed on the device (n_matches < 0)");
  return VSLAM_OK;
}

}  // extern "C"
