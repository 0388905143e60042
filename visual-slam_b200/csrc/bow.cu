// bow.cu -- bag-of-words place recognition (DBoW2 half): vocabulary-tree descent by Hamming
// distance, BowVector / FeatureVector construction and BowVector scoring, for sm_100a.
//
// Semantics: DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB> as the reference vendors it
// (thirdparty/DBoW2_ORBSLAM/DBoW2), reached through visnav::compute_bow_vector (keypoints.h:243-254:
// voc->transform(descriptors, bow_vector, feature_vector, 4)) and voc->score (loop_closure_utils.h:119,201;
// tracking.h:208):
//   * transform of one feature (TemplatedVocabulary.h:1218-1260): from the root, move to the child with the
//     smallest Hamming distance (strict '<' while scanning the children in order: the FIRST minimum wins), until
//     a node without children; report its word id and weight, and the node passed at level L - levelsup;
//   * transform of a feature set (:1127-1195): TF_IDF / TF add the word weight once per feature, in feature
//     order (BowVector::addWeight, BowVector.cpp:34-46); IDF / BINARY set it once; features whose word weight
//     is not positive are dropped ("stopped"); without normalisation TF weights are divided by the number of
//     DISTINCT words (:1162-1168); then BowVector::normalize (BowVector.cpp:62-85): the norm is accumulated in
//     ascending word order, one double addition per word, and every value is divided by it;
//   * score (ScoringObject.cpp): a sum over the words both vectors hold, in ascending word order, one addition
//     per word, then the scoring type's closing formula.
// Double additions are not associative, so the sums above are accumulated in exactly that order (by one lane);
// everything around them -- the 256-bit distances, the sort by word id, the run lengths, the intersections --
// is parallel.  Every f64 operation that could be contracted is an explicit __d*_rn intrinsic.
//
// Layout in HBM.  The tree is stored in "slot" order: breadth first, the children of a node in consecutive
// slots in the reference's child order, so that one descent step reads one contiguous run of 32-byte
// descriptors (lane c of the warp that owns the feature takes child c) and the top levels, which every feature
// visits, sit together at the start (the ORBvoc tree, k = 10, L = 6, 1.1 M nodes, is 35 MB of descriptors: it
// stays in the 126 MB L2).  Per slot: descriptor (2 x uint4), {first child slot, child count, node id, word
// id}, weight.  BowVectors are (word id u32, value f64) arrays in ascending word order.
#include <math.h>

#include <algorithm>
#include <fstream>
#include <sstream>
#include <vector>

#include "common.cuh"

struct vslam_bow_vocab {
  vslam_ctx* ctx = nullptr;
  int k = 0, L = 0, scoring = 0, weighting = 0;
  int n_nodes = 0;  // including the root
  int n_words = 0;
  int root_children = 0;  // the root's children are slots [0, root_children)
  int max_children = 0;
  DevBuf slot_desc;    // [n_nodes - 1] x 8 u32
  DevBuf slot_info;    // [n_nodes - 1] x int4 {child_begin, child_count, node_id, word_id}
  DevBuf slot_weight;  // [n_nodes - 1] f64
  DevBuf word_weight;  // [n_words] f64
  // scratch of the transform / score entry points
  DevBuf in, feat, bow, cnt, q, scores;
};

struct vslam_bow_db {
  vslam_ctx* ctx = nullptr;
  vslam_bow_vocab* voc = nullptr;
  DevBuf ids, vals, offsets, counts;
  std::vector<int64_t> h_offsets;
  std::vector<int32_t> h_counts;
  int64_t total = 0;
  int n = 0;
};

namespace {

constexpr int kMaxFeatures = 8192;  // features per image the BowVector kernel sorts in shared memory
constexpr uint32_t kStopped = 0xFFFFFFFFu;

// ---- descent: one warp per feature ---------------------------------------------------------------------------
struct DescendParams {
  const uint4* slot_desc;
  const int4* slot_info;
  const double* slot_weight;
  int root_children;
  int nid_level;  // L - levelsup
  const uint4* feat;  // [n] x 2 uint4
  int n;
  const int32_t* counts;  // nullable: features are [image][cap] slots, image i holds counts[i] of them
  int cap;
  uint32_t* word;   // [n]
  uint32_t* node;   // [n]
  double* weight;   // [n]
};

__global__ void __launch_bounds__(256) bow_descend_kernel(const DescendParams p) {
  const int lane = threadIdx.x & 31;
  const int f = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (f >= p.n) return;
  if (p.counts && f % p.cap >= p.counts[f / p.cap]) return;
  const uint4 a0 = __ldg(p.feat + 2 * (size_t)f), a1 = __ldg(p.feat + 2 * (size_t)f + 1);
  int begin = 0, count = p.root_children;
  int level = 0;
  uint32_t nid = 0;  // nid_level <= 0: the root (TemplatedVocabulary.h:1227)
  int4 info = make_int4(0, 0, 0, 0);
  int slot = 0;
  do {
    ++level;
    uint32_t best = 0xFFFFFFFFu;  // distance << 16 | child index: the minimum is the first child at the least distance
    for (int c = lane; c < count; c += 32) {
      const uint4 b0 = __ldg(p.slot_desc + 2 * (size_t)(begin + c)), b1 = __ldg(p.slot_desc + 2 * (size_t)(begin + c) + 1);
      const int d = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                    __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
      best = min(best, ((uint32_t)d << 16) | (uint32_t)c);
    }
    best = __reduce_min_sync(0xFFFFFFFFu, best);
    slot = begin + (int)(best & 0xFFFFu);
    info = __ldg(p.slot_info + slot);
    if (level == p.nid_level) nid = (uint32_t)info.z;
    begin = info.x;
    count = info.y;
  } while (count > 0);
  // a leaf above level nid_level: the reference leaves *nid unset there (undefined); the leaf itself is reported
  if (p.nid_level > 0 && level < p.nid_level) nid = (uint32_t)info.z;
  if (lane == 0) {
    p.word[f] = (uint32_t)info.w;
    p.node[f] = nid;
    p.weight[f] = __ldg(p.slot_weight + slot);
  }
}

// ---- BowVector of one image: one CTA -------------------------------------------------------------------------
struct BowVecParams {
  const uint32_t* word;    // [total] per feature
  const double* weight;    // [total]
  const int32_t* set_begin;  // [count + 1], or nullptr: image i owns features [i * feat_cap, i * feat_cap + counts[i])
  const int32_t* counts;
  int feat_cap;
  const double* word_weight;
  int weighting, must_normalize, norm_l2;
  uint32_t* ids;   // [count][cap]
  double* vals;    // [count][cap]
  int32_t* n_out;  // [count]
  int cap;
};

__device__ void bitonic_asc_u32(uint32_t* keys, int n) {
  for (int k = 2; k <= n; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int l = i ^ j;
        if (l > i) {
          const uint32_t a = keys[i], b = keys[l];
          if (((i & k) == 0) ? (a > b) : (a < b)) {
            keys[i] = b;
            keys[l] = a;
          }
        }
      }
      __syncthreads();
    }
}

constexpr int kBowThreads = 256;

__global__ void __launch_bounds__(kBowThreads) bow_vector_kernel(const BowVecParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int s_warp[kBowThreads / 32];
  __shared__ int s_total;
  __shared__ double s_norm;
  const int img = blockIdx.x;
  const int f0 = p.set_begin ? p.set_begin[img] : img * p.feat_cap;
  const int n = p.set_begin ? p.set_begin[img + 1] - f0 : min(p.counts[img], p.feat_cap);
  int np2 = 32;
  while (np2 < n) np2 <<= 1;
  uint32_t* keys = reinterpret_cast<uint32_t*>(smem_raw);
  for (int i = threadIdx.x; i < np2; i += blockDim.x) {
    uint32_t key = kStopped;
    if (i < n && p.weight[f0 + i] > 0.0) key = p.word[f0 + i];  // "if(w > 0) // not stopped", TemplatedVocabulary.h:1157
    keys[i] = key;
  }
  __syncthreads();
  bitonic_asc_u32(keys, np2);

  // heads of the runs of equal word ids, in ascending order: output position = number of heads before
  uint32_t* ids = p.ids + (size_t)img * p.cap;
  double* vals = p.vals + (size_t)img * p.cap;
  int base = 0;
  for (int i0 = 0; i0 < np2; i0 += blockDim.x) {
    const int i = i0 + threadIdx.x;
    const uint32_t key = i < np2 ? keys[i] : kStopped;
    const bool head = i < np2 && key != kStopped && (i == 0 || keys[i - 1] != key);
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, head);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    int before = base;
    for (int w2 = 0; w2 < warp; ++w2) before += s_warp[w2];
    int chunk = 0;
    for (int w2 = 0; w2 < kBowThreads / 32; ++w2) chunk += s_warp[w2];
    if (head) {
      const int pos = before + __popc(bal & ((1u << lane) - 1u));
      int run = 1;
      while (i + run < np2 && keys[i + run] == key) ++run;
      const double w = p.word_weight[key];
      double v = w;
      if (p.weighting == 0 || p.weighting == 1)  // TF_IDF, TF: addWeight once per feature, in feature order
        for (int c = 1; c < run; ++c) v = __dadd_rn(v, w);
      if (pos < p.cap) {
        ids[pos] = key;
        vals[pos] = v;
      }
    }
    base += chunk;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    s_total = base;
    p.n_out[img] = base;
  }
  __syncthreads();
  const int m = min(s_total, p.cap);
  if (m == 0) return;
  if ((p.weighting == 0 || p.weighting == 1) && !p.must_normalize) {
    // TemplatedVocabulary.h:1162-1168: divided by the number of distinct words
    const double nd = (double)s_total;
    for (int i = threadIdx.x; i < m; i += blockDim.x) vals[i] = __ddiv_rn(vals[i], nd);
    return;
  }
  if (!p.must_normalize) return;
  __threadfence_block();
  __syncthreads();
  if (threadIdx.x < 32) {
    // BowVector::normalize, BowVector.cpp:62-85: one addition per word, ascending.  Warp 0 fetches 32 values at a
    // time (coalesced) and every lane runs the same chain of additions over them, fed by shuffles.
    const int lane = threadIdx.x;
    double norm = 0.0;
    double next = lane < m ? vals[lane] : 0.0;
    for (int i0 = 0; i0 < m; i0 += 32) {
      const double v = next;
      next = i0 + 32 + lane < m ? vals[i0 + 32 + lane] : 0.0;  // in flight while this chunk's chain runs
      const int cnt = min(32, m - i0);
      if (cnt == 32) {  // full chunk: unrolled, so the 32 shuffles are issued ahead of the chain of additions
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const double x = __shfl_sync(0xFFFFFFFFu, v, j);
          norm = p.norm_l2 ? __fma_rn(x, x, norm)  // fused, as the -march=native build of DBoW2
                           : __dadd_rn(norm, fabs(x));
        }
      } else {
        for (int j = 0; j < cnt; ++j) {
          const double x = __shfl_sync(0xFFFFFFFFu, v, j);
          norm = p.norm_l2 ? __fma_rn(x, x, norm) : __dadd_rn(norm, fabs(x));
        }
      }
    }
    if (p.norm_l2) norm = __dsqrt_rn(norm);
    if (lane == 0) s_norm = norm;
  }
  __syncthreads();
  const double norm = s_norm;
  if (norm > 0.0)
    for (int i = threadIdx.x; i < m; i += blockDim.x) vals[i] = __ddiv_rn(vals[i], norm);
}

// ---- scoring: one warp per stored BowVector ------------------------------------------------------------------
// score(v1 = query, v2 = entry) for the scoring types whose sum runs over the COMMON words only
// (ScoringObject.cpp: L1 :22-66, L2 :71-118, chi-square :123-168, Bhattacharyya :226-263, dot product :268-305).
struct ScoreParams {
  const uint32_t* q_ids;
  const double* q_vals;
  int nq;
  const uint32_t* ids;      // concatenated entries
  const double* vals;
  const int64_t* offsets;   // [n]
  const int32_t* counts;    // [n]
  int n;
  int scoring;
  double* scores;     // [n]
  int32_t* n_common;  // [n], nullable
};

__device__ __forceinline__ double score_term(int scoring, double vi, double wi) {
  switch (scoring) {
    case 0:  // L1: fabs(vi - wi) - fabs(vi) - fabs(wi)
      return __dsub_rn(__dsub_rn(fabs(__dsub_rn(vi, wi)), fabs(vi)), fabs(wi));
    case 2: {  // chi-square: vi * wi / (vi + wi) unless the denominator is zero
      const double s = __dadd_rn(vi, wi);
      return s != 0.0 ? __ddiv_rn(__dmul_rn(vi, wi), s) : 0.0;
    }
    case 4:  // Bhattacharyya
      return __dsqrt_rn(__dmul_rn(vi, wi));
    default:  // L2, dot product: accumulated with a fused multiply-add below
      return 0.0;
  }
}

// The query's word ids are staged in shared memory once per CTA; every warp then walks stored vectors (lane l
// takes element i0 + l: coalesced) and looks each word up in the query by binary search.  Both vectors ascend, so
// the hits arrive in ascending word order whichever side is walked.
constexpr int kScoreQueryMax = 8192;  // query words held in shared memory (a transform yields at most 8192)

template <bool kSmemQuery>
__global__ void __launch_bounds__(256) bow_score_kernel(const ScoreParams p) {
  extern __shared__ uint32_t s_qids[];
  const int lane = threadIdx.x & 31;
  if (kSmemQuery) {
    for (int i = threadIdx.x; i < p.nq; i += blockDim.x) s_qids[i] = __ldg(p.q_ids + i);
    __syncthreads();
  }
  const uint32_t* qids = kSmemQuery ? s_qids : p.q_ids;
  // "score += vi * wi" (L2, dot product) is a fused multiply-add in the reference's -O3 -march=native build of DBoW2
  // (thirdparty/DBoW2_ORBSLAM/CMakeLists.txt:4-5): followed here, like "norm += v * v" in the normalisation
  const bool fused = p.scoring == 1 || p.scoring == 5;
  const int warps = gridDim.x * (blockDim.x >> 5);
  for (int e = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); e < p.n; e += warps) {
    const uint32_t* ids = p.ids + p.offsets[e];
    const double* vals = p.vals + p.offsets[e];
    const int m = p.counts[e];
    double score = 0.0;
    int common = 0;
    for (int i0 = 0; i0 < m; i0 += 32) {
      const int i = i0 + lane;
      bool found = false;
      double term = 0.0, other = 0.0;
      if (i < m && p.nq > 0) {
        const uint32_t id = __ldg(ids + i);
        int lo = 0, hi = p.nq;  // lower bound in the query
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (qids[mid] < id) lo = mid + 1; else hi = mid;
        }
        if (lo < p.nq && qids[lo] == id) {
          found = true;
          const double vi = __ldg(p.q_vals + lo), wi = __ldg(vals + i);  // v1 = query, v2 = stored vector
          term = fused ? vi : score_term(p.scoring, vi, wi);
          other = wi;
        }
      }
      unsigned mask = __ballot_sync(0xFFFFFFFFu, found);
      common += __popc(mask);
      while (mask) {  // ascending word order, one addition per common word (every lane keeps the same sum)
        const int src = __ffs(mask) - 1;
        const double t = __shfl_sync(0xFFFFFFFFu, term, src);
        if (fused) score = __fma_rn(t, __shfl_sync(0xFFFFFFFFu, other, src), score);
        else score = __dadd_rn(score, t);
        mask &= mask - 1;
      }
    }
    if (lane == 0) {
      double s = score;
      switch (p.scoring) {
        case 0: s = __ddiv_rn(-score, 2.0); break;                                          // ScoringObject.cpp:63
        case 1: s = score >= 1.0 ? 1.0 : __dsub_rn(1.0, __dsqrt_rn(__dsub_rn(1.0, score))); break;  // :112-115
        case 2: s = __dmul_rn(2.0, score); break;                                           // :165
        default: break;
      }
      p.scores[e] = s;
      if (p.n_common) p.n_common[e] = common;
    }
  }
}

inline size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

int check_scoring(vslam_ctx* ctx, int scoring, int weighting) {
  if (scoring < 0 || scoring > 5 || weighting < 0 || weighting > 3) return vs_fail(ctx, VSLAM_E_ARG, "bad scoring / weighting type");
  if (scoring == 3)
    return vs_fail(ctx, VSLAM_E_LIMIT, "KL scoring is not provided (its log terms are not reproducible bit for bit)");
  return VSLAM_OK;
}

// must normalize / which norm: ScoringObject.h:44-60
inline bool must_normalize(int scoring) { return scoring != 5; }
inline bool norm_is_l2(int scoring) { return scoring == 1; }

}  // namespace

// ---- host side: vocabulary ------------------------------------------------------------------------------------
static int build_vocab(vslam_ctx* ctx, int k, int L, int scoring, int weighting, int n, const int32_t* parent,
                       const uint8_t* is_leaf, const uint8_t* desc, const double* weight, vslam_bow_vocab** out) {
  int rc = check_scoring(ctx, scoring, weighting);
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!parent || !is_leaf || !desc || !weight)) || !out) return vs_fail(ctx, VSLAM_E_ARG, "bad vocabulary arguments");
  // children in the order the nodes appear (loadFromTextFile: m_nodes[pid].children.push_back(nid))
  std::vector<std::vector<int>> children((size_t)n + 1);
  for (int i = 0; i < n; ++i) {
    if (parent[i] < 0 || parent[i] > i) return vs_fail(ctx, VSLAM_E_ARG, "vocabulary node names a parent that does not precede it");
    children[parent[i]].push_back(i + 1);
  }
  std::vector<int> word_of((size_t)n + 1, 0);
  std::vector<double> h_word_weight;
  for (int i = 0; i < n; ++i)
    if (is_leaf[i]) {
      word_of[i + 1] = (int)h_word_weight.size();
      h_word_weight.push_back(weight[i]);
    }
  // breadth-first slot order
  std::vector<int> slot_node;
  slot_node.reserve(n);
  std::vector<int> first_child_slot((size_t)n + 1, 0);
  size_t max_children = children[0].size();
  for (int c : children[0]) slot_node.push_back(c);
  for (size_t s = 0; s < slot_node.size(); ++s) {
    const int node = slot_node[s];
    first_child_slot[node] = (int)slot_node.size();
    max_children = std::max(max_children, children[node].size());
    for (int c : children[node]) slot_node.push_back(c);
  }
  if (max_children > 65535) return vs_fail(ctx, VSLAM_E_LIMIT, "a vocabulary node has more than 65535 children");
  std::vector<uint8_t> h_desc((size_t)n * 32);
  std::vector<int32_t> h_info((size_t)n * 4);
  std::vector<double> h_weight((size_t)n);
  for (size_t s = 0; s < slot_node.size(); ++s) {
    const int node = slot_node[s];
    memcpy(&h_desc[s * 32], desc + (size_t)(node - 1) * 32, 32);
    h_info[4 * s + 0] = first_child_slot[node];
    h_info[4 * s + 1] = (int)children[node].size();
    h_info[4 * s + 2] = node;
    h_info[4 * s + 3] = word_of[node];
    h_weight[s] = weight[node - 1];
  }
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  vslam_bow_vocab* v = new vslam_bow_vocab;
  v->ctx = ctx;
  v->k = k; v->L = L; v->scoring = scoring; v->weighting = weighting;
  v->n_nodes = n + 1;
  v->n_words = (int)h_word_weight.size();
  v->root_children = (int)children[0].size();
  v->max_children = (int)max_children;
  cudaError_t e = v->slot_desc.reserve((size_t)n * 32 + 32);
  if (e == cudaSuccess) e = v->slot_info.reserve((size_t)n * 16 + 16);
  if (e == cudaSuccess) e = v->slot_weight.reserve((size_t)n * 8 + 8);
  if (e == cudaSuccess) e = v->word_weight.reserve(h_word_weight.size() * 8 + 8);
  if (e == cudaSuccess && n > 0) e = cudaMemcpy(v->slot_desc.p, h_desc.data(), (size_t)n * 32, cudaMemcpyHostToDevice);
  if (e == cudaSuccess && n > 0) e = cudaMemcpy(v->slot_info.p, h_info.data(), (size_t)n * 16, cudaMemcpyHostToDevice);
  if (e == cudaSuccess && n > 0) e = cudaMemcpy(v->slot_weight.p, h_weight.data(), (size_t)n * 8, cudaMemcpyHostToDevice);
  if (e == cudaSuccess && !h_word_weight.empty())
    e = cudaMemcpy(v->word_weight.p, h_word_weight.data(), h_word_weight.size() * 8, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    ctx->err = std::string("vocabulary upload: ") + cudaGetErrorString(e);
    vslam_bow_vocab_destroy(v);
    return VSLAM_E_CUDA;
  }
  *out = v;
  return VSLAM_OK;
}

extern "C" {

void vslam_bow_vocab_destroy(vslam_bow_vocab* v) {
  if (!v) return;
  cudaSetDevice(v->ctx->device);
  cudaDeviceSynchronize();
  DevBuf* bufs[] = {&v->slot_desc, &v->slot_info, &v->slot_weight, &v->word_weight, &v->in, &v->feat, &v->bow, &v->cnt, &v->q, &v->scores};
  for (DevBuf* b : bufs) b->release();
  delete v;
}

int vslam_bow_vocab_create(vslam_ctx* ctx, int k, int L, int scoring, int weighting, int n_nodes, const int32_t* parent,
                           const uint8_t* is_leaf, const uint8_t* desc, const double* weight, vslam_bow_vocab** out) {
  if (!ctx) return VSLAM_E_ARG;
  return build_vocab(ctx, k, L, scoring, weighting, n_nodes, parent, is_leaf, desc, weight, out);
}

// The text format of TemplatedVocabulary::loadFromTextFile / saveToTextFile (TemplatedVocabulary.h:1338-1450):
// "k L scoring weighting", then one node per line: parent id, leaf flag, 32 descriptor bytes, weight.
// The reference's read loop runs until eof(), so a file that ends with a newline (saveToTextFile writes one) gives
// it one more, phantom child of the root whose descriptor is uninitialised memory (cv::Mat::create) and whose
// weight is 0: undefined behaviour that cannot be mirrored.  Blank lines are skipped here.
int vslam_bow_vocab_load_text(vslam_ctx* ctx, const char* path, vslam_bow_vocab** out) {
  if (!ctx) return VSLAM_E_ARG;
  if (!path || !out) return vs_fail(ctx, VSLAM_E_ARG, "bad arguments");
  std::ifstream f(path);
  if (!f.good()) return vs_fail(ctx, VSLAM_E_ARG, "cannot open the vocabulary file");
  std::string s;
  std::getline(f, s);
  int k = -1, L = -1, n1 = -1, n2 = -1;
  {
    std::stringstream ss(s);
    ss >> k >> L >> n1 >> n2;
  }
  if (k < 0 || k > 20 || L < 1 || L > 10 || n1 < 0 || n1 > 5 || n2 < 0 || n2 > 3)  // :1360
    return vs_fail(ctx, VSLAM_E_ARG, "vocabulary loading failure: this is not a correct text file");
  std::vector<int32_t> parent;
  std::vector<uint8_t> leaf, desc;
  std::vector<double> weight;
  while (std::getline(f, s)) {
    if (s.find_first_not_of(" \t\r\n") == std::string::npos) continue;
    std::stringstream ss(s);
    int pid = 0, is_leaf = 0;
    ss >> pid >> is_leaf;
    uint8_t d[32];
    for (int i = 0; i < 32; ++i) {
      int b = 0;
      ss >> b;
      d[i] = (unsigned char)b;
    }
    double w = 0.0;
    ss >> w;
    if (ss.fail()) return vs_fail(ctx, VSLAM_E_ARG, "vocabulary loading failure: malformed node line");
    parent.push_back(pid);
    leaf.push_back(is_leaf > 0);
    desc.insert(desc.end(), d, d + 32);
    weight.push_back(w);
  }
  return build_vocab(ctx, k, L, n1, n2, (int)parent.size(), parent.data(), leaf.data(), desc.data(), weight.data(), out);
}

int vslam_bow_vocab_info(const vslam_bow_vocab* v, int32_t info[6]) {
  if (!v || !info) return VSLAM_E_ARG;
  info[0] = v->k; info[1] = v->L; info[2] = v->scoring; info[3] = v->weighting; info[4] = v->n_nodes; info[5] = v->n_words;
  return VSLAM_OK;
}

int vslam_bow_transform_batch(vslam_ctx* ctx, vslam_bow_vocab* v, const uint8_t* desc, const int32_t* set_begin,
                              int count, int levelsup, uint32_t* bow_ids, double* bow_vals, int32_t* n_bow, int cap,
                              uint32_t* feat_word, uint32_t* feat_node, double* feat_weight) {
  if (!ctx) return VSLAM_E_ARG;
  if (!v || v->ctx != ctx || count < 0 || cap <= 0 || (count > 0 && (!set_begin || !bow_ids || !bow_vals || !n_bow)))
    return vs_fail(ctx, VSLAM_E_ARG, "bad transform arguments");
  if (count == 0) return VSLAM_OK;
  int max_n = 0;
  for (int i = 0; i < count; ++i) {
    if (set_begin[i + 1] < set_begin[i] || set_begin[0] != 0) return vs_fail(ctx, VSLAM_E_ARG, "set_begin must start at 0 and not decrease");
    max_n = std::max(max_n, set_begin[i + 1] - set_begin[i]);
  }
  if (max_n > kMaxFeatures) return vs_fail(ctx, VSLAM_E_LIMIT, "more than 8192 features in one image");
  const int total = set_begin[count];
  if (total > 0 && !desc) return vs_fail(ctx, VSLAM_E_ARG, "descriptors missing");
  if (v->n_words == 0) {  // empty(): transform clears both vectors (TemplatedVocabulary.h:1135-1138)
    for (int i = 0; i < count; ++i) n_bow[i] = 0;
    if (feat_word) memset(feat_word, 0, (size_t)total * 4);
    if (feat_node) memset(feat_node, 0, (size_t)total * 4);
    if (feat_weight) memset(feat_weight, 0, (size_t)total * 8);
    return VSLAM_OK;
  }
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t o_begin = up256((size_t)total * 32);
  VS_CUDA(ctx, v->in.reserve(o_begin + (size_t)(count + 1) * 4));
  const size_t o_node = up256((size_t)total * 4), o_w = o_node + up256((size_t)total * 4);
  VS_CUDA(ctx, v->feat.reserve(o_w + (size_t)total * 8 + 8));
  const size_t o_vals = up256((size_t)count * cap * 4);
  VS_CUDA(ctx, v->bow.reserve(o_vals + (size_t)count * cap * 8));
  VS_CUDA(ctx, v->cnt.reserve((size_t)count * 4));
  uint8_t* in = v->in.as<uint8_t>();
  uint8_t* feat = v->feat.as<uint8_t>();
  uint8_t* bow = v->bow.as<uint8_t>();
  if (total > 0) VS_CUDA(ctx, cudaMemcpyAsync(in, desc, (size_t)total * 32, cudaMemcpyHostToDevice, ctx->stream));
  VS_CUDA(ctx, cudaMemcpyAsync(in + o_begin, set_begin, (size_t)(count + 1) * 4, cudaMemcpyHostToDevice, ctx->stream));
  if (total > 0) {
    DescendParams dp;
    dp.slot_desc = v->slot_desc.as<uint4>();
    dp.slot_info = v->slot_info.as<int4>();
    dp.slot_weight = v->slot_weight.as<double>();
    dp.root_children = v->root_children;
    dp.nid_level = v->L - levelsup;
    dp.feat = reinterpret_cast<const uint4*>(in);
    dp.n = total;
    dp.counts = nullptr;
    dp.cap = 0;
    dp.word = reinterpret_cast<uint32_t*>(feat);
    dp.node = reinterpret_cast<uint32_t*>(feat + o_node);
    dp.weight = reinterpret_cast<double*>(feat + o_w);
    bow_descend_kernel<<<(total + 7) / 8, 256, 0, ctx->stream>>>(dp);
    ctx->launches += 1;
    VS_CUDA(ctx, cudaGetLastError());
  }
  BowVecParams bp;
  bp.word = reinterpret_cast<const uint32_t*>(feat);
  bp.weight = reinterpret_cast<const double*>(feat + o_w);
  bp.set_begin = reinterpret_cast<const int32_t*>(in + o_begin);
  bp.counts = nullptr;
  bp.feat_cap = 0;
  bp.word_weight = v->word_weight.as<double>();
  bp.weighting = v->weighting;
  bp.must_normalize = must_normalize(v->scoring);
  bp.norm_l2 = norm_is_l2(v->scoring);
  bp.ids = reinterpret_cast<uint32_t*>(bow);
  bp.vals = reinterpret_cast<double*>(bow + o_vals);
  bp.n_out = v->cnt.as<int32_t>();
  bp.cap = cap;
  int np2 = 32;
  while (np2 < max_n) np2 <<= 1;
  bow_vector_kernel<<<count, kBowThreads, (size_t)np2 * 4, ctx->stream>>>(bp);
  ctx->launches += 1;
  VS_CUDA(ctx, cudaGetLastError());
  VS_CUDA(ctx, cudaMemcpyAsync(n_bow, v->cnt.p, (size_t)count * 4, cudaMemcpyDeviceToHost, ctx->stream));
  VS_CUDA(ctx, cudaMemcpyAsync(bow_ids, bow, (size_t)count * cap * 4, cudaMemcpyDeviceToHost, ctx->stream));
  VS_CUDA(ctx, cudaMemcpyAsync(bow_vals, bow + o_vals, (size_t)count * cap * 8, cudaMemcpyDeviceToHost, ctx->stream));
  if (feat_word && total > 0) VS_CUDA(ctx, cudaMemcpyAsync(feat_word, feat, (size_t)total * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (feat_node && total > 0) VS_CUDA(ctx, cudaMemcpyAsync(feat_node, feat + o_node, (size_t)total * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (feat_weight && total > 0) VS_CUDA(ctx, cudaMemcpyAsync(feat_weight, feat + o_w, (size_t)total * 8, cudaMemcpyDeviceToHost, ctx->stream));
  VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  for (int i = 0; i < count; ++i)
    if (n_bow[i] > cap) return vs_fail(ctx, VSLAM_E_CAPACITY, "BowVector capacity too small");
  return VSLAM_OK;
}

// One chunk of compute_bow_vector on images that are already on the device: ORB (orb.cu) -> descent -> BowVector, and
// the downloads of its results, all queued on ctx->stream without waiting.  counts / status: DEVICE arrays of the chunk.
static int bow_chunk(vslam_ctx* ctx, vslam_bow_vocab* v, const uint8_t* images_dev, int count, int w, int h, int pitch,
                     size_t image_stride, const vslam_orb_params* prm, int levelsup, uint32_t* bow_ids, double* bow_vals,
                     int32_t* n_bow, int bow_cap, float* kp, uint8_t* desc, uint32_t* feat_node, double* feat_weight,
                     int cap, int32_t* counts, int32_t* status) {
  OrbDeviceOut o;
  int rc = launch_orb(ctx, images_dev, count, w, h, pitch, image_stride, prm, cap, &o);
  if (rc) return rc;
  const int total = count * cap;
  const size_t o_node = up256((size_t)total * 4), o_w = o_node + up256((size_t)total * 4);
  VS_CUDA(ctx, v->feat.reserve(o_w + (size_t)total * 8 + 8));
  const size_t o_vals = up256((size_t)count * bow_cap * 4);
  VS_CUDA(ctx, v->bow.reserve(o_vals + (size_t)count * bow_cap * 8));
  VS_CUDA(ctx, v->cnt.reserve((size_t)count * 4));
  uint8_t* feat = v->feat.as<uint8_t>();
  uint8_t* bow = v->bow.as<uint8_t>();
  if (v->n_words > 0) {
    VS_CUDA(ctx, cudaMemsetAsync(feat, 0, o_w + (size_t)total * 8, ctx->stream));
    DescendParams dp;
    dp.slot_desc = v->slot_desc.as<uint4>();
    dp.slot_info = v->slot_info.as<int4>();
    dp.slot_weight = v->slot_weight.as<double>();
    dp.root_children = v->root_children;
    dp.nid_level = v->L - levelsup;
    dp.feat = reinterpret_cast<const uint4*>(o.desc);
    dp.n = total;
    dp.counts = o.n_kp;
    dp.cap = cap;
    dp.word = reinterpret_cast<uint32_t*>(feat);
    dp.node = reinterpret_cast<uint32_t*>(feat + o_node);
    dp.weight = reinterpret_cast<double*>(feat + o_w);
    bow_descend_kernel<<<(total + 7) / 8, 256, 0, ctx->stream>>>(dp);
    BowVecParams bp;
    bp.word = dp.word;
    bp.weight = dp.weight;
    bp.set_begin = nullptr;
    bp.counts = o.n_kp;
    bp.feat_cap = cap;
    bp.word_weight = v->word_weight.as<double>();
    bp.weighting = v->weighting;
    bp.must_normalize = must_normalize(v->scoring);
    bp.norm_l2 = norm_is_l2(v->scoring);
    bp.ids = reinterpret_cast<uint32_t*>(bow);
    bp.vals = reinterpret_cast<double*>(bow + o_vals);
    bp.n_out = v->cnt.as<int32_t>();
    bp.cap = bow_cap;
    int np2 = 32;
    while (np2 < cap) np2 <<= 1;
    bow_vector_kernel<<<count, kBowThreads, (size_t)np2 * 4, ctx->stream>>>(bp);
    ctx->launches += 2;
    VS_CUDA(ctx, cudaGetLastError());
    VS_CUDA(ctx, cudaMemcpyAsync(n_bow, v->cnt.p, (size_t)count * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaMemcpyAsync(bow_ids, bow, (size_t)count * bow_cap * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VS_CUDA(ctx, cudaMemcpyAsync(bow_vals, bow + o_vals, (size_t)count * bow_cap * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (feat_node) VS_CUDA(ctx, cudaMemcpyAsync(feat_node, feat + o_node, (size_t)total * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (feat_weight) VS_CUDA(ctx, cudaMemcpyAsync(feat_weight, feat + o_w, (size_t)total * 8, cudaMemcpyDeviceToHost, ctx->stream));
  }
  // device to device: a download into the caller's pageable bookkeeping would stall the host until the chunk is done
  VS_CUDA(ctx, cudaMemcpyAsync(counts, o.n_kp, (size_t)count * 4, cudaMemcpyDeviceToDevice, ctx->stream));
  VS_CUDA(ctx, cudaMemcpyAsync(status, o.status, (size_t)count * 4, cudaMemcpyDeviceToDevice, ctx->stream));
  if (kp) VS_CUDA(ctx, cudaMemcpyAsync(kp, o.kp, (size_t)total * 24, cudaMemcpyDeviceToHost, ctx->stream));
  if (desc) VS_CUDA(ctx, cudaMemcpyAsync(desc, o.desc, (size_t)total * 32, cudaMemcpyDeviceToHost, ctx->stream));
  return VSLAM_OK;
}

// visnav::compute_bow_vector (keypoints.h:243-254) for a batch.  Large batches go through in chunks: the upload of
// chunk c + 1 (copy stream) overlaps the kernels of chunk c (the workspaces are reused chunk after chunk in stream order).
constexpr int kBowChunk = 128;

int vslam_compute_bow_vector_batch(vslam_ctx* ctx, vslam_bow_vocab* v, const uint8_t* images, int count, int w, int h,
                                   int pitch, size_t image_stride, const vslam_orb_params* prm, int levelsup,
                                   uint32_t* bow_ids, double* bow_vals, int32_t* n_bow, int bow_cap, float* kp,
                                   uint8_t* desc, int32_t* n_kp, uint32_t* feat_node, double* feat_weight, int cap) {
  if (!ctx) return VSLAM_E_ARG;
  if (!v || v->ctx != ctx || !images || !prm || count < 0 || w <= 0 || h <= 0 || pitch < w || cap <= 0 || bow_cap <= 0 ||
      image_stride < (size_t)pitch * (h - 1) + w || (count > 0 && (!bow_ids || !bow_vals || !n_bow)))
    return vs_fail(ctx, VSLAM_E_ARG, "bad compute_bow_vector arguments");
  if (cap > kMaxFeatures) return vs_fail(ctx, VSLAM_E_LIMIT, "more than 8192 features in one image");
  if (count == 0) return VSLAM_OK;
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t tail = (size_t)pitch * (h - 1) + w;  // bytes of the last image of a run
  VS_CUDA(ctx, ctx->orb[6].reserve(image_stride * (size_t)(count - 1) + tail + 64));
  uint8_t* staged = ctx->orb[6].as<uint8_t>();
  std::vector<int32_t> status((size_t)count), counts((size_t)count);
  VS_CUDA(ctx, v->q.reserve((size_t)count * 8 + 256));  // keypoint counts | status of the whole batch
  int32_t* counts_dev = v->q.as<int32_t>();
  int32_t* status_dev = counts_dev + count;
  const int n_chunks = (count + kBowChunk - 1) / kBowChunk;
  std::vector<cudaEvent_t> ev((size_t)n_chunks, nullptr);
  int rc = VSLAM_OK;
  VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // the staging buffer may still be read by an earlier call
  for (int c = 0; c < n_chunks && rc == VSLAM_OK; ++c) {
    const int c0 = c * kBowChunk, cn = std::min(kBowChunk, count - c0);
    cudaError_t e = cudaEventCreateWithFlags(&ev[c], cudaEventDisableTiming);
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(staged + image_stride * c0, images + image_stride * c0, image_stride * (size_t)(cn - 1) + tail,
                          cudaMemcpyHostToDevice, ctx->copy_stream[0]);
    if (e == cudaSuccess) e = cudaEventRecord(ev[c], ctx->copy_stream[0]);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->stream, ev[c], 0);
    if (e != cudaSuccess) {
      ctx->err = std::string("compute_bow_vector upload: ") + cudaGetErrorString(e);
      rc = VSLAM_E_CUDA;
      break;
    }
    rc = bow_chunk(ctx, v, staged + image_stride * c0, cn, w, h, pitch, image_stride, prm, levelsup,
                   bow_ids + (size_t)c0 * bow_cap, bow_vals + (size_t)c0 * bow_cap, n_bow + c0, bow_cap,
                   kp ? kp + (size_t)c0 * cap * 6 : nullptr, desc ? desc + (size_t)c0 * cap * 32 : nullptr,
                   feat_node ? feat_node + (size_t)c0 * cap : nullptr, feat_weight ? feat_weight + (size_t)c0 * cap : nullptr,
                   cap, counts_dev + c0, status_dev + c0);
  }
  cudaError_t es = cudaStreamSynchronize(ctx->copy_stream[0]);
  cudaError_t e2 = cudaStreamSynchronize(ctx->stream);
  for (cudaEvent_t e : ev)
    if (e) cudaEventDestroy(e);
  if (rc) return rc;
  VS_CUDA(ctx, es);
  VS_CUDA(ctx, e2);
  VS_CUDA(ctx, cudaMemcpy(counts.data(), counts_dev, (size_t)count * 4, cudaMemcpyDeviceToHost));
  VS_CUDA(ctx, cudaMemcpy(status.data(), status_dev, (size_t)count * 4, cudaMemcpyDeviceToHost));
  if (v->n_words == 0) {  // empty(): transform clears both vectors
    for (int i = 0; i < count; ++i) n_bow[i] = 0;
    if (feat_node) memset(feat_node, 0, (size_t)count * cap * 4);
    if (feat_weight) memset(feat_weight, 0, (size_t)count * cap * 8);
  }
  if (n_kp) memcpy(n_kp, counts.data(), (size_t)count * 4);
  rc = orb_check_status(ctx, counts.data(), status.data(), count, cap);
  if (rc) return rc;
  for (int i = 0; i < count; ++i)
    if (n_bow[i] > bow_cap) return vs_fail(ctx, VSLAM_E_CAPACITY, "BowVector capacity too small");
  return VSLAM_OK;
}

// ---- BowVector database: the keyframes' vectors stay on the device ---------------------------------------------
static cudaError_t grow_keep(DevBuf& b, size_t need, size_t keep) {
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

static int check_bow_vector(vslam_ctx* ctx, const uint32_t* ids, const double* vals, int n) {
  if (n < 0 || (n > 0 && (!ids || !vals))) return vs_fail(ctx, VSLAM_E_ARG, "bad BowVector");
  for (int i = 1; i < n; ++i)
    if (ids[i] <= ids[i - 1]) return vs_fail(ctx, VSLAM_E_ARG, "BowVector word ids must ascend strictly (std::map order)");
  return VSLAM_OK;
}

int vslam_bow_db_create(vslam_ctx* ctx, vslam_bow_vocab* v, vslam_bow_db** out) {
  if (!ctx) return VSLAM_E_ARG;
  if (!v || v->ctx != ctx || !out) return vs_fail(ctx, VSLAM_E_ARG, "bad arguments");
  vslam_bow_db* d = new vslam_bow_db;
  d->ctx = ctx;
  d->voc = v;
  *out = d;
  return VSLAM_OK;
}

void vslam_bow_db_destroy(vslam_bow_db* d) {
  if (!d) return;
  cudaSetDevice(d->ctx->device);
  cudaDeviceSynchronize();
  d->ids.release(); d->vals.release(); d->offsets.release(); d->counts.release();
  delete d;
}

int vslam_bow_db_size(const vslam_bow_db* d) { return d ? d->n : 0; }

int vslam_bow_db_add(vslam_bow_db* d, const uint32_t* ids, const double* vals, int n, int* index) {
  if (!d) return VSLAM_E_ARG;
  vslam_ctx* ctx = d->ctx;
  int rc = check_bow_vector(ctx, ids, vals, n);
  if (rc) return rc;
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // no query may still read the buffers that move
  const int idx = d->n;
  VS_CUDA(ctx, grow_keep(d->ids, (size_t)(d->total + n + 1) * 4, (size_t)d->total * 4));
  VS_CUDA(ctx, grow_keep(d->vals, (size_t)(d->total + n + 1) * 8, (size_t)d->total * 8));
  VS_CUDA(ctx, grow_keep(d->offsets, (size_t)(idx + 1) * 8, (size_t)idx * 8));
  VS_CUDA(ctx, grow_keep(d->counts, (size_t)(idx + 1) * 4, (size_t)idx * 4));
  const int64_t off = d->total;
  const int32_t cnt = n;
  if (n > 0) {
    VS_CUDA(ctx, cudaMemcpy(d->ids.as<uint32_t>() + off, ids, (size_t)n * 4, cudaMemcpyHostToDevice));
    VS_CUDA(ctx, cudaMemcpy(d->vals.as<double>() + off, vals, (size_t)n * 8, cudaMemcpyHostToDevice));
  }
  VS_CUDA(ctx, cudaMemcpy(d->offsets.as<int64_t>() + idx, &off, 8, cudaMemcpyHostToDevice));
  VS_CUDA(ctx, cudaMemcpy(d->counts.as<int32_t>() + idx, &cnt, 4, cudaMemcpyHostToDevice));
  d->h_offsets.push_back(off);
  d->h_counts.push_back(cnt);
  d->total += n;
  d->n = idx + 1;
  if (index) *index = idx;
  return VSLAM_OK;
}

static int run_score(vslam_ctx* ctx, vslam_bow_vocab* v, const uint32_t* q_ids, const double* q_vals, int nq,
                     const uint32_t* ids, const double* vals, const int64_t* offsets, const int32_t* counts, int n,
                     double* scores, int32_t* n_common) {
  // q_*: host; ids/vals/offsets/counts: device; scores / n_common: host
  const size_t o_qv = up256((size_t)nq * 4 + 4);
  VS_CUDA(ctx, v->q.reserve(o_qv + (size_t)nq * 8 + 8));
  const size_t o_nc = up256((size_t)n * 8);
  VS_CUDA(ctx, v->scores.reserve(o_nc + (size_t)n * 4));
  uint8_t* q = v->q.as<uint8_t>();
  uint8_t* sc = v->scores.as<uint8_t>();
  if (nq > 0) {
    VS_CUDA(ctx, cudaMemcpyAsync(q, q_ids, (size_t)nq * 4, cudaMemcpyHostToDevice, ctx->stream));
    VS_CUDA(ctx, cudaMemcpyAsync(q + o_qv, q_vals, (size_t)nq * 8, cudaMemcpyHostToDevice, ctx->stream));
  }
  ScoreParams sp;
  sp.q_ids = reinterpret_cast<const uint32_t*>(q);
  sp.q_vals = reinterpret_cast<const double*>(q + o_qv);
  sp.nq = nq;
  sp.ids = ids; sp.vals = vals; sp.offsets = offsets; sp.counts = counts;
  sp.n = n;
  sp.scoring = v->scoring;
  sp.scores = reinterpret_cast<double*>(sc);
  sp.n_common = reinterpret_cast<int32_t*>(sc + o_nc);
  const int grid = std::max(1, std::min((n + 7) / 8, ctx->num_sms * 8));
  if (nq <= kScoreQueryMax)
    bow_score_kernel<true><<<grid, 256, (size_t)nq * 4, ctx->stream>>>(sp);
  else
    bow_score_kernel<false><<<grid, 256, 0, ctx->stream>>>(sp);
  ctx->launches += 1;
  VS_CUDA(ctx, cudaGetLastError());
  VS_CUDA(ctx, cudaMemcpyAsync(scores, sc, (size_t)n * 8, cudaMemcpyDeviceToHost, ctx->stream));
  if (n_common) VS_CUDA(ctx, cudaMemcpyAsync(n_common, sc + o_nc, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return VSLAM_OK;
}

int vslam_bow_db_query(vslam_ctx* ctx, vslam_bow_db* d, const uint32_t* ids, const double* vals, int n, double* scores,
                       int32_t* n_common) {
  if (!ctx) return VSLAM_E_ARG;
  if (!d || d->ctx != ctx || (d->n > 0 && !scores)) return vs_fail(ctx, VSLAM_E_ARG, "bad query arguments");
  int rc = check_bow_vector(ctx, ids, vals, n);
  if (rc) return rc;
  if (d->n == 0) return VSLAM_OK;
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  return run_score(ctx, d->voc, ids, vals, n, d->ids.as<uint32_t>(), d->vals.as<double>(), d->offsets.as<int64_t>(),
                   d->counts.as<int32_t>(), d->n, scores, n_common);
}

// voc->score(v1, v2) for one pair (TemplatedVocabulary.h:1199-1203)
int vslam_bow_score(vslam_ctx* ctx, vslam_bow_vocab* v, const uint32_t* ids1, const double* vals1, int n1,
                    const uint32_t* ids2, const double* vals2, int n2, double* score) {
  if (!ctx) return VSLAM_E_ARG;
  if (!v || v->ctx != ctx || !score) return vs_fail(ctx, VSLAM_E_ARG, "bad score arguments");
  int rc = check_bow_vector(ctx, ids1, vals1, n1);
  if (!rc) rc = check_bow_vector(ctx, ids2, vals2, n2);
  if (rc) return rc;
  VS_CUDA(ctx, cudaSetDevice(ctx->device));
  // v2 staged behind the transform scratch: {ids, vals, offset, count}
  const size_t o_vals = up256((size_t)n2 * 4 + 4), o_off = o_vals + up256((size_t)n2 * 8 + 8), o_cnt = o_off + 256;
  VS_CUDA(ctx, v->bow.reserve(o_cnt + 256));
  uint8_t* b = v->bow.as<uint8_t>();
  const int64_t off = 0;
  const int32_t cnt = n2;
  if (n2 > 0) {
    VS_CUDA(ctx, cudaMemcpyAsync(b, ids2, (size_t)n2 * 4, cudaMemcpyHostToDevice, ctx->stream));
    VS_CUDA(ctx, cudaMemcpyAsync(b + o_vals, vals2, (size_t)n2 * 8, cudaMemcpyHostToDevice, ctx->stream));
  }
  VS_CUDA(ctx, cudaMemcpyAsync(b + o_off, &off, 8, cudaMemcpyHostToDevice, ctx->stream));
  VS_CUDA(ctx, cudaMemcpyAsync(b + o_cnt, &cnt, 4, cudaMemcpyHostToDevice, ctx->stream));
  VS_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // off / cnt are stack variables
  return run_score(ctx, v, ids1, vals1, n1, reinterpret_cast<const uint32_t*>(b), reinterpret_cast<const double*>(b + o_vals),
                   reinterpret_cast<const int64_t*>(b + o_off), reinterpret_cast<const int32_t*>(b + o_cnt), 1, score, nullptr);
}

}  // extern "C"
