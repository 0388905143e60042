/*
 * visnav_b200.h -- C ABI of the B200 (sm_100a) front-end hot path of visnav
 * (yunjinli/visual-slam).  This is the drop-in boundary: plain pointers and
 * sizes, no C++/torch types.  Everything below runs as hand-written CUDA; there
 * is no CPU fallback -- every entry point fails with VSLAM_E_CUDA when no
 * sm_100 device is usable.
 *
 * Reference interfaces replaced (all in /root/reference/include/visnav/keypoints.h):
 *   detectKeypoints                :133-150  -> vslam_detect
 *   computeAngles                  :152-189  -> vslam_angles
 *   computeDescriptors             :191-221  -> vslam_describe
 *   detectKeypointsAndDescriptors  :223-229  -> vslam_detect_describe
 *   matchDescriptors (+isPQiffQP)  :323-369 (:278-313) -> vslam_match
 * Batched forms of the same operations (frames / keyframe pairs are independent,
 * slam.cpp:1125-1143, sim3.h:252,281, tracking.h:283,317) follow below.
 *
 * Data layouts (identical to the reference's containers, so the C++ shim in
 * include/visnav/keypoints.h passes vector::data() straight through):
 *   corners      N x 2 double  (x, y)   == std::vector<Eigen::Vector2d>
 *   angles       N double               == std::vector<double>
 *   descriptors  N x 8 uint32, bit i of the std::bitset<256> at word i/32, bit i%32
 *                (libstdc++ in-memory layout, sizeof == 32)
 *   matches      M x 2 int32 (id1, id2) == std::vector<std::pair<int,int>>
 *
 * Threading: one vslam_ctx per host thread / GPU.  Calls on one ctx are
 * serialised by the caller; different ctxs are independent.  All calls are
 * synchronous at return unless stated otherwise.
 */
#ifndef VISNAV_B200_H_
#define VISNAV_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct vslam_ctx vslam_ctx;
typedef struct vslam_stereo_plan vslam_stereo_plan;
typedef struct vslam_desc_store vslam_desc_store;

/* status codes */
#define VSLAM_OK 0
#define VSLAM_E_ARG (-1)      /* bad argument */
#define VSLAM_E_CAPACITY (-2) /* caller-provided output capacity too small */
#define VSLAM_E_CUDA (-3)     /* CUDA error / no usable device; see vslam_last_error */
#define VSLAM_E_LIMIT (-4)    /* input exceeds a documented internal limit */

/* options for vslam_set_option */
#define VSLAM_OPT_FP_CONTRACT 1 /* 1 (default): rBRIEF rotation uses the fused form GCC emits for the
                                   reference at -O3 on FMA hosts (x' = fma(c,x,-(s*y))); 0: two
                                   multiplies + add (reference built with -ffp-contract=off) */
#define VSLAM_OPT_MATCH_CSA 2   /* 1 (default): Hamming kernel folds words with carry-save adders
                                   before POPC; 0: one POPC per 32-bit word.  Results identical. */
#define VSLAM_OPT_BATCH_CHUNK 3 /* stereo pairs per pipeline chunk of vslam_stereo_batch (default 120) */
#define VSLAM_OPT_BATCH_DEPTH 4 /* chunks in flight in vslam_stereo_batch, 1..4 (default 4): upload,
                                   compute and download of consecutive chunks overlap */

#define VSLAM_OPT_MATCH_SPLIT_SMALL 5 /* 1 (default): a batch with fewer matching jobs than half the SM count
                                        spreads every job's columns over all SMs (latency of a single stereo
                                        pair); 0: always one CTA per job.  Results identical. */

#define VSLAM_OPT_RESPONSE_BAND 6 /* output rows per box of the corner-response kernel.  0 (default): the whole image
                                    height -- a warp walks whole columns and carries OpenCV's running column sum
                                    (box_filter.simd.hpp:206-262): every bit of the response as cv::goodFeaturesToTrack
                                    computes it.  1: row bands chosen per launch for latency (<= 64 rows for a single
                                    image); >= 8: bands of that height.  Bands sum the three rows of the box directly: the
                                    response then differs from OpenCV's in the last bits of a few pixels per million
                                    (where a double addition was inexact), which can swap two corners of equal strength
                                    (DESIGN.md section 2). */
#define VSLAM_OPT_MATCH_MMA 7 /* 1 (default): batches of matching jobs that fill the GPU (stereo plans / vslam_stereo_batch
                                 chunks, vslam_match_batch) compute the Hamming matrix on the tensor cores (tcgen05.mma
                                 kind::i8 on descriptors expanded to signed bytes: dot = 256 - 2 * hamming, exact);
                                 0: the POPC kernels; 2: the tensor cores for every batch, however small (tests).
                                 Results identical. */
#define VSLAM_OPT_GRAPH 8 /* 1 (default): vslam_stereo_batch calls of a few pairs (one chunk whose results fit 2 MB: the
                             one-pair-at-a-time use of slam.cpp:1125-1143) replay their kernel sequence as one CUDA graph
                             from the third call with the same arguments on; 0: plain launches.  Results identical. */
#define VSLAM_OPT_RESPONSE_MULTI 9 /* 1 (default): when the last 120-column strip of the image is at most 56 columns wide
                                    * (752 = 6 x 120 + 32), launches of at least two waves of warps give that strip of two or
                                    * three consecutive images to ONE warp instead of one mostly idle warp per image; 0: one
                                    * warp per strip; 2: shared strips for any launch of two images or more (tests).
                                    * Results are identical (tests/test_gpu_detect_describe.py). */
#define VSLAM_OPT_TWO_STREAMS 10 /* 1 (default): vslam_stereo_plan_run / vslam_stereo_batch chunks of at least four pairs per SM
                                  * run as two halves on two streams (forked from and joined to the call's stream), so that the
                                  * last wave of one half's kernel overlaps the other half's next kernel; 0: one stream,
                                  * whole-batch launches (the per-stage times of vslam_stereo_plan_last_ms then describe
                                  * kernels that run alone).  Results identical. */

/* ---- context ------------------------------------------------------------------------ */
int vslam_create(int device, vslam_ctx** out);
void vslam_destroy(vslam_ctx* ctx);
const char* vslam_last_error(vslam_ctx* ctx);
int vslam_set_option(vslam_ctx* ctx, int option, int value);
/* cv::goodFeaturesToTrack's useHarrisDetector / k (featureselect.cpp:376-380 -> cornerHarris, corner.cpp:104-153)
 * for every detection call of this ctx.  The reference always passes false (keypoints.h:138-139): default off,
 * k = 0.04 is OpenCV's default.  Response = a*c - b*b - k*(a + c)^2 on the same box-filtered covariance. */
int vslam_set_harris(vslam_ctx* ctx, int use_harris_detector, double k);
/* the CUDA stream (cudaStream_t) all work of this ctx is issued on; for event timing */
void* vslam_stream(vslam_ctx* ctx);
int vslam_sync(vslam_ctx* ctx);
/* number of kernels this ctx launched so far (bench.py's gpu_launches) */
long long vslam_launch_count(vslam_ctx* ctx);
/* pinned host memory, so that batched calls overlap copies with kernels */
void* vslam_host_alloc(size_t bytes);
void vslam_host_free(void* p);

/* ---- single-image operations (host buffers in, host buffers out) --------------------- */
/* keypoints.h:133-150.  img: h rows of w bytes, `pitch` bytes apart.  num_features <= 0: no cap.
 * xy: cap x 2 doubles.  *n = number of corners written (<= cap). */
int vslam_detect(vslam_ctx* ctx, const uint8_t* img, int w, int h, int pitch, int num_features,
                 double* xy, int cap, int* n);
/* keypoints.h:152-189.  rotate == 0 -> all angles 0. */
int vslam_angles(vslam_ctx* ctx, const uint8_t* img, int w, int h, int pitch, const double* xy,
                 int n, int rotate, double* angles);
/* keypoints.h:191-221.  desc: n x 8 uint32. */
int vslam_describe(vslam_ctx* ctx, const uint8_t* img, int w, int h, int pitch, const double* xy,
                   const double* angles, int n, uint32_t* desc);
/* keypoints.h:223-229. */
int vslam_detect_describe(vslam_ctx* ctx, const uint8_t* img, int w, int h, int pitch,
                          int num_features, int rotate, double* xy, double* angles, uint32_t* desc,
                          int cap, int* n);
/* keypoints.h:323-369.  pairs: cap x 2 int32, ascending id1.  *n = number of matches; if it
 * exceeds cap only cap pairs are written and VSLAM_E_CAPACITY is returned. */
int vslam_match(vslam_ctx* ctx, const uint32_t* d1, int n1, const uint32_t* d2, int n2,
                int threshold, double dist_2_best, int32_t* pairs, int cap, int* n);
/* Row statistics of the scan in keypoints.h:331-354 for every query against `db`:
 * out3[3*i..] = (best distance, lowest index attaining it, second best).  Relocalisation-shaped
 * queries (BASELINE config 4) use this together with vslam_match. */
int vslam_best2(vslam_ctx* ctx, const uint32_t* q, int nq, const uint32_t* db, int ndb,
                int32_t* out3);
/* visnav::find_matches_landmarks (vo_utils.h:83-167): keypoints of the current frame against the
 * projected map landmarks.  kp_xy [n_kp][2] / kp_desc [n_kp][8]: kdl.corners / kdl.corner_descriptors;
 * proj_xy [n_proj][2]: projected_points (their order decides ties exactly as std::partial_sort does in
 * the reference); landmark j owns the observation descriptors obs_desc[obs_offsets[j] ..
 * obs_offsets[j+1]) (the host side gathers feature_corners.at(fcid).corner_descriptors[f_id] over
 * landmarks.at(projected_track_ids[j]).all_obs).  pairs[cap][2] = (feature id, index j into
 * proj_xy) in ascending feature id; the caller maps j to projected_track_ids[j]. */
int vslam_match_landmarks(vslam_ctx* ctx, const double* kp_xy, const uint32_t* kp_desc, int n_kp,
                          const double* proj_xy, int n_proj, const int32_t* obs_offsets,
                          const uint32_t* obs_desc, double match_max_dist_2d, int threshold,
                          double dist_2_best, int32_t* pairs, int cap, int* n);
/* matchDescriptors with desc2 SHARDED over GPUs (BASELINE config 4: one query set against a large map; SURVEY
 * 8e).  Every GPU holds all of d1 and one contiguous shard of desc2 starting at global row index_offset.
 * vslam_match_shard writes part4 [n1][4] int32 = (best distance, GLOBAL index of the lowest-index best row of this
 * shard, second best, mutual-check flag of that row) -- 16 bytes per query, the only data that has to be
 * gathered.  vslam_match_merge_shards merges n_parts such parts ([n_parts][n1][4], any order) with the
 * associative operator of the scan (lowest global index wins ties) and applies threshold / ratio / mutual check:
 * the result equals vslam_match(d1, all of desc2). */
int vslam_match_shard(vslam_ctx* ctx, const uint32_t* d1, int n1, const uint32_t* d2_shard, int n2_shard,
                      int index_offset, int threshold, double dist_2_best, int32_t* part4);
int vslam_match_merge_shards(vslam_ctx* ctx, const int32_t* parts4, int n_parts, int n1, int threshold,
                             double dist_2_best, int32_t* pairs, int cap, int* n);
/* Min-eigenvalue response map (the float image cv::goodFeaturesToTrack thresholds,
 * OCV featureselect.cpp:411) -- exported for parity tests.  lam: h x w floats, dense. */
int vslam_response(vslam_ctx* ctx, const uint8_t* img, int w, int h, int pitch, float* lam);

/* ---- batched detect+describe (independent images) -------------------------------------- */
/* images: `count` images, image k at images + k*image_stride.  Outputs are fixed-stride slabs:
 * xy [count][cap][2], angles [count][cap], desc [count][cap][8], n_kp [count].
 * cap must be >= num_features (> 0). */
int vslam_detect_describe_batch(vslam_ctx* ctx, const uint8_t* images, int count, int w, int h,
                                int pitch, size_t image_stride, int num_features, int rotate,
                                double* xy, double* angles, uint32_t* desc, int32_t* n_kp, int cap);

/* ---- multi-scale extension (BASELINE config 5; no reference behaviour) ------ */
/* Factor-2 pyramid, level k+1 (x, y) = (a + b + c + d + 2) >> 2 over the 2x2 block of level k, level k
 * being (w >> k) x (h >> k); detectKeypointsAndDescriptors semantics per level with num_features[k]
 * (coordinates are in the level's own pixels).  Per image, level k's outputs occupy the slab entries
 * [sum(num_features[0..k)), +n_kp[image][k]): xy [count][cap][2], angles [count][cap],
 * desc [count][cap][8], n_kp [count][n_levels]; cap >= sum(num_features). */
#define VSLAM_MAX_PYRAMID_LEVELS 8
int vslam_pyramid_detect_describe_batch(vslam_ctx* ctx, const uint8_t* images, int count, int w, int h,
                                        int pitch, size_t image_stride, int n_levels,
                                        const int* num_features, int rotate, double* xy, double* angles,
                                        uint32_t* desc, int32_t* n_kp, int cap);

/* ---- batched stereo front end: detect+describe both images, then matchDescriptors -------- */
/* slam.cpp:1125-1143 for `count` independent stereo pairs.  images: [count][2] images (left,
 * right) `image_stride` bytes apart.  Per image outputs as above with 2*count images; per pair:
 * matches [count][cap][2], n_matches [count].  Any output pointer may be NULL (not copied back). */
typedef struct {
  int w, h, pitch;
  size_t image_stride;
  int num_features;
  int rotate_features;
  int match_threshold;
  double match_dist_2_best;
} vslam_stereo_params;

int vslam_stereo_batch(vslam_ctx* ctx, const uint8_t* images, int count,
                       const vslam_stereo_params* p, double* xy, double* angles, uint32_t* desc,
                       int32_t* n_kp, int32_t* matches, int32_t* n_matches, int cap);

/* The same pipeline split into upload / run / download so that a caller (bench.py) can time
 * the device-resident part alone.  run() is asynchronous on vslam_stream(ctx). */
int vslam_stereo_plan_create(vslam_ctx* ctx, int w, int h, int max_pairs, int cap,
                             vslam_stereo_plan** out);
void vslam_stereo_plan_destroy(vslam_stereo_plan* plan);
int vslam_stereo_plan_upload(vslam_stereo_plan* plan, const uint8_t* images, int count, int pitch,
                             size_t image_stride);
int vslam_stereo_plan_run(vslam_stereo_plan* plan, int count, int num_features, int rotate,
                          int match_threshold, double match_dist_2_best);
int vslam_stereo_plan_download(vslam_stereo_plan* plan, int count, double* xy, double* angles,
                               uint32_t* desc, int32_t* n_kp, int32_t* matches, int32_t* n_matches);
/* time of the last run() on the device, CUDA events on the ctx stream: total (including the optional
 * epipolar stage) and per stage (response, select, describe, match); ms[5]. */
int vslam_stereo_plan_last_ms(vslam_stereo_plan* plan, float* ms);

/* ---- batched matching against a device-resident descriptor store ----------------------- */
/* Keyframe descriptors stay on the device (sim3.h:252,281 / tracking.h:283,317 match the same
 * stored keyframes again and again).  sets: n_sets descriptor sets, set k has counts[k]
 * descriptors starting at desc + offsets[k]*8 words. */
int vslam_desc_store_create(vslam_ctx* ctx, const uint32_t* desc, const int64_t* offsets,
                            const int32_t* counts, int n_sets, vslam_desc_store** out);
void vslam_desc_store_destroy(vslam_desc_store* s);
/* Appends one descriptor set (a new keyframe: slam.cpp:1153-1154 copies kd into feature_corners) and returns its
 * set index; earlier indices stay valid.  Amortised O(count): the device buffers grow geometrically. */
int vslam_desc_store_append(vslam_desc_store* s, const uint32_t* desc, int count, int* set_index);
int vslam_desc_store_size(const vslam_desc_store* s);
/* jobs: n_jobs x 2 int32 (set index of descriptors_1, set index of descriptors_2); an index outside the store is
 * VSLAM_E_ARG.  Outputs: matches [n_jobs][cap][2], n_matches [n_jobs] (host; either may be NULL).  n_matches holds
 * the FULL count of every job; when one exceeds cap its list holds the first cap pairs and the call returns
 * VSLAM_E_CAPACITY (as vslam_match does) after all jobs have run -- pass cap >= the largest set to rule that out.
 * With n_matches == NULL the counts cannot be checked: the caller vouches for cap. */
int vslam_match_batch(vslam_ctx* ctx, const vslam_desc_store* s, const int32_t* jobs, int n_jobs,
                      int threshold, double dist_2_best, int32_t* matches, int32_t* n_matches,
                      int cap);
/* device time of the last vslam_match_batch / vslam_match / vslam_best2 kernels (ms) */
float vslam_last_match_ms(vslam_ctx* ctx);

/* ---- geometry that follows matching ------------------------------------------ */
/* The four camera models of include/visnav/camera_models.h, all with the reference's 8-double parameter vector
 * (fx fy cx cy + model parameters, AbstractCamera::getParam(); width/height = AbstractCamera::width()/height()):
 *   pinhole :47-124    eucm :126-216 (alpha, beta)    ds :218-309 (xi, alpha)    kb4 :311-443 (k1..k4) */
#define VSLAM_CAM_PINHOLE 0
#define VSLAM_CAM_EUCM 1
#define VSLAM_CAM_DS 2
#define VSLAM_CAM_KB4 3
typedef struct {
  int model; /* VSLAM_CAM_* */
  int width, height;
  int reserved;
  double param[8];
} vslam_camera;
/* AbstractCamera::project / unproject for n points: xyz [n][3], uv [n][2] (doubles). */
int vslam_cam_project(vslam_ctx* ctx, const vslam_camera* cam, const double* xyz, int n, double* uv);
int vslam_cam_unproject(vslam_ctx* ctx, const vslam_camera* cam, const double* uv, int n, double* xyz);
/* visnav::findInliersEssential (matching_utils.h:63-88): for every match (id1, id2) unproject
 * kd1.corners[id1] / kd2.corners[id2] and keep the match iff !(|p0^T E p1| > epipolar_error_threshold), in
 * input order.  E: 3x3 row-major (computeEssential, matching_utils.h:55-61, stays on the host: one 3x3 product).
 * inliers [cap][2]; *n_inliers may exceed cap (then VSLAM_E_CAPACITY).  err_out (nullable): p0^T E p1 per match. */
int vslam_find_inliers_essential(vslam_ctx* ctx, const double* xy1, int n1, const double* xy2, int n2,
                                 const vslam_camera* cam1, const vslam_camera* cam2, const double* E,
                                 double epipolar_error_threshold, const int32_t* matches, int n_matches,
                                 int32_t* inliers, int cap, int* n_inliers, double* err_out);
/* visnav::project_landmarks (vo_utils.h:48-81): p_c = current_pose.inverse() * p for n landmark positions
 * xyz [n][3]; kept iff !(p_c.z < cam_z_threshold) and the projection lies in [0, width] x [0, height].
 * pose_w_c = Sophus::SE3d::data() of current_pose: qx qy qz qw tx ty tz.  Outputs in ascending landmark index
 * (the reference iterates an unordered_map; no caller depends on its order): proj_xy [cap][2], proj_idx [cap]
 * = index into xyz (the host maps it to the TrackId). */
int vslam_project_landmarks(vslam_ctx* ctx, const double* pose_w_c, const vslam_camera* cam, const double* xyz,
                            int n, double cam_z_threshold, double* proj_xy, int32_t* proj_idx, int cap, int* n_out);
/* Stereo front end with the inlier filter fused behind the matcher (slam.cpp:1141-1143: matchDescriptors, then
 * findInliersEssential on the same pair): inliers [count][cap][2], n_inliers [count] in addition to the
 * outputs of vslam_stereo_batch; corners and matches never leave the device in between. */
typedef struct {
  vslam_camera cam1, cam2;
  double E[9]; /* row-major */
  double epipolar_error_threshold;
} vslam_epipolar;
int vslam_stereo_batch_epipolar(vslam_ctx* ctx, const uint8_t* images, int count, const vslam_stereo_params* p,
                                const vslam_epipolar* ep, double* xy, double* angles, uint32_t* desc,
                                int32_t* n_kp, int32_t* matches, int32_t* n_matches, int32_t* inliers,
                                int32_t* n_inliers, int cap);
/* plan form: ep == NULL switches the stage off again */
int vslam_stereo_plan_set_epipolar(vslam_stereo_plan* plan, const vslam_epipolar* ep);
int vslam_stereo_plan_download_inliers(vslam_stereo_plan* plan, int count, int32_t* inliers, int32_t* n_inliers);
int vslam_stereo_plan_last_epipolar_ms(vslam_stereo_plan* plan, float* ms);

/* ---- bag of words: vocabulary transform and BowVector scoring (DBoW2 half) --- */
/* Replaces DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB> (= visnav::ORBVocabulary, ORBVocabulary.h:13) on
 * the two calls the reference makes: voc->transform(descriptors, bow_vector, feature_vector, 4)
 * (keypoints.h:253; TemplatedVocabulary.h:1127-1195, :1218-1260) and voc->score(v1, v2)
 * (loop_closure_utils.h:119, :201; tracking.h:208; ScoringObject.cpp).  Results are bit-identical to the
 * reference's: distances are integers, and the double sums are accumulated in the reference's order.
 * Scoring / weighting values are DBoW2's enums (BowVector.h:35-53). */
#define VSLAM_BOW_L1_NORM 0
#define VSLAM_BOW_L2_NORM 1
#define VSLAM_BOW_CHI_SQUARE 2
#define VSLAM_BOW_KL 3 /* not provided: VSLAM_E_LIMIT */
#define VSLAM_BOW_BHATTACHARYYA 4
#define VSLAM_BOW_DOT_PRODUCT 5
#define VSLAM_BOW_TF_IDF 0
#define VSLAM_BOW_TF 1
#define VSLAM_BOW_IDF 2
#define VSLAM_BOW_BINARY 3
typedef struct vslam_bow_vocab vslam_bow_vocab;
typedef struct vslam_bow_db vslam_bow_db;
/* Vocabulary from arrays: node i+1 (the root is node 0) has parent[i] (<= i), leaf flag is_leaf[i] (leaves get
 * word ids in node order), descriptor desc[i][32] (cv::Mat row of cv::ORB = 8 little-endian u32) and weight[i]
 * -- the fields of one line of the text format. */
int vslam_bow_vocab_create(vslam_ctx* ctx, int k, int L, int scoring, int weighting, int n_nodes,
                           const int32_t* parent, const uint8_t* is_leaf, const uint8_t* desc, const double* weight,
                           vslam_bow_vocab** out);
/* TemplatedVocabulary::loadFromTextFile (TemplatedVocabulary.h:1338-1425; slam.cpp:373): "k L scoring weighting",
 * then "parent leaf b0 .. b31 weight" per node.  Blank lines are skipped (the reference turns a trailing newline
 * into a phantom root child with an uninitialised descriptor). */
int vslam_bow_vocab_load_text(vslam_ctx* ctx, const char* path, vslam_bow_vocab** out);
void vslam_bow_vocab_destroy(vslam_bow_vocab* v);
/* info = {k, L, scoring, weighting, number of nodes incl. the root, number of words} */
int vslam_bow_vocab_info(const vslam_bow_vocab* v, int32_t info[6]);
/* transform for `count` images at once: image i owns descriptors [set_begin[i], set_begin[i+1]) of desc
 * [total][32].  Outputs per image: the BowVector as bow_ids / bow_vals [count][cap] in ascending word id
 * (std::map order) and n_bow [count]; per feature (all nullable): its word id, the id of the node it passes at
 * level L - levelsup (0 = root when that level is <= 0) and the word's weight -- feature f belongs to the
 * FeatureVector entry feat_node[f] iff feat_weight[f] > 0.  At most 8192 features per image. */
int vslam_bow_transform_batch(vslam_ctx* ctx, vslam_bow_vocab* v, const uint8_t* desc, const int32_t* set_begin,
                              int count, int levelsup, uint32_t* bow_ids, double* bow_vals, int32_t* n_bow, int cap,
                              uint32_t* feat_word, uint32_t* feat_node, double* feat_weight);
/* voc->score(v1, v2); word ids ascend strictly. */
int vslam_bow_score(vslam_ctx* ctx, vslam_bow_vocab* v, const uint32_t* ids1, const double* vals1, int n1,
                    const uint32_t* ids2, const double* vals2, int n2, double* score);
/* The keyframes' BowVectors (Camera::bow_vector, common_types.h:213) kept on the device; a query scores
 * voc->score(query, entry) against EVERY entry in one launch and also returns the number of shared words
 * (what detect_loop_candidates / detect_relocalization_candidate count through the inverted file,
 * loop_closure_utils.h:152-185, tracking.h:176-203). */
int vslam_bow_db_create(vslam_ctx* ctx, vslam_bow_vocab* v, vslam_bow_db** out);
void vslam_bow_db_destroy(vslam_bow_db* db);
int vslam_bow_db_add(vslam_bow_db* db, const uint32_t* ids, const double* vals, int n, int* index);
int vslam_bow_db_size(const vslam_bow_db* db);
int vslam_bow_db_query(vslam_ctx* ctx, vslam_bow_db* db, const uint32_t* ids, const double* vals, int n,
                       double* scores, int32_t* n_common);

/* ---- cv::ORB front end of the bag-of-words path (ORB half) ------------------- */
/* Replaces orb->detectAndCompute(img, cv::Mat(), keypoints, descriptors) of visnav::compute_bow_vector
 * (keypoints.h:249-250) with orb = cv::ORB::create(nfeatures, scale_factor, nlevels, edge_threshold, 0, 2,
 * cv::ORB::FAST_SCORE) (patchSize 31, fast_threshold 20 by default), for `count` images of one size at once.
 * kp [count][cap][6] = cv::KeyPoint {pt.x, pt.y, size, angle, response, octave}; desc [count][cap][32]; n_kp [count].
 * Level by level; inside a level in raster order (the reference's order there is std::nth_element's: unspecified).
 * Ties at a level's cut-off score all stay (KeyPointsFilter::retainBest), so n_kp may exceed nfeatures: give cap
 * some room (VSLAM_E_CAPACITY otherwise). */
typedef struct {
  int nfeatures;
  float scale_factor; /* 1.2f */
  int nlevels;        /* 8; at most 8 */
  int edge_threshold; /* 19; at least 19 (the reach of the rotated sampling pattern) */
  int fast_threshold; /* 20 */
} vslam_orb_params;
int vslam_orb_detect_compute_batch(vslam_ctx* ctx, const uint8_t* images, int count, int w, int h, int pitch,
                                   size_t image_stride, const vslam_orb_params* prm, float* kp, uint8_t* desc,
                                   int32_t* n_kp, int cap);
/* visnav::compute_bow_vector (keypoints.h:243-254) for `count` images: ORB as above, then
 * voc->transform(descriptors, bow_vector, feature_vector, levelsup) -- the descriptors never leave the device.
 * BowVector outputs as vslam_bow_transform_batch; kp / desc / n_kp / feat_node / feat_weight (all nullable) give the
 * keypoints and, per keypoint, its FeatureVector node and word weight ([count][cap] each). */
int vslam_compute_bow_vector_batch(vslam_ctx* ctx, vslam_bow_vocab* v, const uint8_t* images, int count, int w, int h,
                                   int pitch, size_t image_stride, const vslam_orb_params* prm, int levelsup,
                                   uint32_t* bow_ids, double* bow_vals, int32_t* n_bow, int bow_cap, float* kp,
                                   uint8_t* desc, int32_t* n_kp, uint32_t* feat_node, double* feat_weight, int cap);

/* ---- measurement helpers --------------------------------------------------------------- */
/* CUDA events on vslam_stream(ctx): record event `slot` (0..3); elapsed ms between two recorded
 * slots (synchronises on the later one). */
int vslam_event_record(vslam_ctx* ctx, int slot);
int vslam_event_elapsed_ms(vslam_ctx* ctx, int slot_a, int slot_b, float* ms);
/* Register-only POPC throughput microbenchmark: popc.b32 instructions per second (thread
 * level) sustained over all SMs -- the measured denominator of the matching roofline. */
int vslam_popc_peak(vslam_ctx* ctx, double* popc_per_s);
/* 1 if a bounded pipeline wait of the tensor-core matcher ever ran out in this process (a bug, never in a correct
 * run: the job then reports count -2 instead of hanging the GPU), 0 if not, < 0 on a CUDA error. */
int vslam_debug_mma_timeouts(vslam_ctx* ctx);

/* ---- allocation-free entry point ------------------------------------------------------- */
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as
 * vslam_best2 -- out3[3*i..] = (best distance, lowest index attaining it, second best), the row scan of
 * keypoints.h:331-354 -- but d_q [nq][8], d_db [ndb][8] and d_out3 [nq][3] are DEVICE pointers (d_q / d_db 16-byte
 * aligned) and the call allocates nothing: the per-range partials live in a fixed 2 MiB file-scope device array
 * (csrc/noalloc.cu; every GPU that loads the library holds it until exit).  Limits: nq <= 4096, ndb <= 2^20;
 * beyond them VSLAM_E_LIMIT with a message in vslam_last_error.  Work is issued on `stream` (a cudaStream_t; NULL:
 * vslam_stream(ctx)) and is NOT synchronised.  The first call on a device looks the array's address up
 * (cudaGetSymbolAddress): make it once before a capture.  All calls on a device share the array: calls that
 * overlap on the GPU race and must be serialised by the caller. */
int vslam_best2_noalloc(vslam_ctx* ctx, const uint32_t* d_q, int nq, const uint32_t* d_db, int ndb,
                        int32_t* d_out3, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VISNAV_B200_H_ */
