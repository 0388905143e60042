/* visnav_oracle_bow.c -- CPU restatement of the DBoW2 calls the reference makes for place recognition
 * (SURVEY §8f-4): vocabulary-tree transform of a descriptor set into a BowVector / FeatureVector and the
 * BowVector score.  TEST INFRASTRUCTURE ONLY: linked into oracle/liboracle.so; only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg may call it.  Parity status: PINNED against the
 * reference's own DBoW2 (oracle/_ref/libvisnav_ref_bow.so, built from thirdparty/DBoW2_ORBSLAM where it lies)
 * by tests/test_oracle_bow.py and the committed fixture tests/golden/bow.npz.
 *
 * Follows (thirdparty/DBoW2_ORBSLAM/DBoW2):
 *   loadFromTextFile          TemplatedVocabulary.h:1338-1425
 *   transform, one feature    TemplatedVocabulary.h:1218-1260   (distance: FORB.cpp:81-100)
 *   transform, feature set    TemplatedVocabulary.h:1127-1195   (addWeight / addIfNotExist / normalize:
 *                             BowVector.cpp:34-85; addFeature: FeatureVector.cpp:28-42)
 *   score                     ScoringObject.cpp:22-66 (L1), :71-118 (L2), :123-168 (chi-square),
 *                             :226-263 (Bhattacharyya), :268-305 (dot product); KL is not restated
 * Doubles, expressions and summation order as in the reference; compiled with -ffp-contract=off.  The reference
 * builds DBoW2 with -O3 -march=native (thirdparty/DBoW2_ORBSLAM/CMakeLists.txt:4-5), where GCC contracts
 * "norm += v * v" (BowVector.cpp:74) and "score += vi * wi" (ScoringObject.cpp:88, :285) into fused multiply-adds:
 * those three accumulations are written with fma() here, as the reference build executes them.  L1 scoring --
 * what ORBvoc.txt selects -- has no product and is unaffected.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
  int k, L, scoring, weighting;
  int n_nodes; /* including the root */
  int n_words;
  int* child_begin; /* [n_nodes] into children */
  int* child_count; /* [n_nodes] */
  int* children;    /* [n_nodes - 1] node ids, grouped by parent in order of appearance */
  uint8_t* desc;    /* [n_nodes][32] (root unused) */
  double* weight;   /* [n_nodes] */
  uint32_t* word_id; /* [n_nodes], 0 for nodes that are not flagged as words (Node() default) */
  double* acc;       /* [n_words] scratch of transform */
  uint8_t* present;  /* [n_words] */
} bow_vocab;

void vso_bow_vocab_free(void* p) {
  bow_vocab* v = (bow_vocab*)p;
  if (!v) return;
  free(v->child_begin); free(v->child_count); free(v->children); free(v->desc); free(v->weight);
  free(v->word_id); free(v->acc); free(v->present);
  free(v);
}

/* node i + 1 = (parent[i], is_leaf[i], desc[i], weight[i]): the fields of one line of the text format */
void* vso_bow_vocab_create(int k, int L, int scoring, int weighting, int n, const int32_t* parent,
                           const uint8_t* is_leaf, const uint8_t* desc, const double* weight) {
  bow_vocab* v = (bow_vocab*)calloc(1, sizeof(bow_vocab));
  v->k = k; v->L = L; v->scoring = scoring; v->weighting = weighting;
  v->n_nodes = n + 1;
  v->child_begin = (int*)calloc((size_t)n + 2, sizeof(int));
  v->child_count = (int*)calloc((size_t)n + 1, sizeof(int));
  v->children = (int*)calloc((size_t)n + 1, sizeof(int));
  v->desc = (uint8_t*)calloc((size_t)n + 1, 32);
  v->weight = (double*)calloc((size_t)n + 1, sizeof(double));
  v->word_id = (uint32_t*)calloc((size_t)n + 1, sizeof(uint32_t));
  for (int i = 0; i < n; i++) {
    if (parent[i] < 0 || parent[i] > i) { vso_bow_vocab_free(v); return 0; }
    v->child_count[parent[i]]++;
  }
  for (int i = 0; i < n + 1; i++) v->child_begin[i + 1] = v->child_begin[i] + v->child_count[i];
  int* fill = (int*)calloc((size_t)n + 1, sizeof(int));
  for (int i = 0; i < n; i++) { /* children.push_back(nid) in node order, :1395 */
    const int pid = parent[i];
    v->children[v->child_begin[pid] + fill[pid]++] = i + 1;
    memcpy(v->desc + (size_t)(i + 1) * 32, desc + (size_t)i * 32, 32);
    v->weight[i + 1] = weight[i];
    if (is_leaf[i]) v->word_id[i + 1] = (uint32_t)v->n_words++; /* :1412-1418 */
  }
  free(fill);
  v->acc = (double*)calloc((size_t)v->n_words + 1, sizeof(double));
  v->present = (uint8_t*)calloc((size_t)v->n_words + 1, 1);
  return v;
}

/* Blank lines are skipped: the reference's "while(!f.eof())" turns a trailing newline into a phantom child of
 * the root whose descriptor is uninitialised memory -- undefined behaviour that is not restated. */
void* vso_bow_vocab_load_text(const char* path) {
  FILE* f = fopen(path, "r");
  if (!f) return 0;
  int k = -1, L = -1, n1 = -1, n2 = -1;
  if (fscanf(f, "%d %d %d %d", &k, &L, &n1, &n2) != 4 || k < 0 || k > 20 || L < 1 || L > 10 || n1 < 0 || n1 > 5 ||
      n2 < 0 || n2 > 3) { /* :1360 */
    fclose(f);
    return 0;
  }
  size_t cap = 1024, n = 0;
  int32_t* parent = (int32_t*)malloc(cap * 4);
  uint8_t* leaf = (uint8_t*)malloc(cap);
  uint8_t* desc = (uint8_t*)malloc(cap * 32);
  double* weight = (double*)malloc(cap * 8);
  for (;;) {
    int pid, is_leaf;
    if (fscanf(f, "%d %d", &pid, &is_leaf) != 2) break;
    if (n == cap) {
      cap *= 2;
      parent = (int32_t*)realloc(parent, cap * 4);
      leaf = (uint8_t*)realloc(leaf, cap);
      desc = (uint8_t*)realloc(desc, cap * 32);
      weight = (double*)realloc(weight, cap * 8);
    }
    int ok = 1;
    for (int i = 0; i < 32; i++) {
      int b;
      if (fscanf(f, "%d", &b) != 1) { ok = 0; break; }
      desc[n * 32 + i] = (uint8_t)b;
    }
    if (!ok || fscanf(f, "%lf", &weight[n]) != 1) break;
    parent[n] = pid;
    leaf[n] = is_leaf > 0;
    n++;
  }
  fclose(f);
  void* v = vso_bow_vocab_create(k, L, n1, n2, (int)n, parent, leaf, desc, weight);
  free(parent); free(leaf); free(desc); free(weight);
  return v;
}

void vso_bow_vocab_info(void* p, int32_t* info) {
  bow_vocab* v = (bow_vocab*)p;
  info[0] = v->k; info[1] = v->L; info[2] = v->scoring; info[3] = v->weighting; info[4] = v->n_nodes; info[5] = v->n_words;
}

static int forb_distance(const uint8_t* a, const uint8_t* b) { /* FORB.cpp:81-100 */
  int dist = 0;
  for (int i = 0; i < 8; i++) {
    uint32_t x, y;
    memcpy(&x, a + 4 * i, 4);
    memcpy(&y, b + 4 * i, 4);
    uint32_t w = x ^ y;
    w = w - ((w >> 1) & 0x55555555u);
    w = (w & 0x33333333u) + ((w >> 2) & 0x33333333u);
    dist += (int)((((w + (w >> 4)) & 0xF0F0F0Fu) * 0x1010101u) >> 24);
  }
  return dist;
}

/* TemplatedVocabulary.h:1218-1260.  *nid: node at level L - levelsup; where the leaf lies above that level the
 * reference leaves it unset (undefined) -- the leaf is reported, as the CUDA path does. */
static void transform_one(const bow_vocab* v, const uint8_t* feature, int levelsup, uint32_t* word, double* weight,
                          uint32_t* nid) {
  const int nid_level = v->L - levelsup;
  *nid = 0;
  int final_id = 0, level = 0;
  do {
    ++level;
    const int* nodes = v->children + v->child_begin[final_id];
    const int cnt = v->child_count[final_id];
    final_id = nodes[0];
    double best_d = forb_distance(feature, v->desc + (size_t)final_id * 32);
    for (int c = 1; c < cnt; c++) {
      const double d = forb_distance(feature, v->desc + (size_t)nodes[c] * 32);
      if (d < best_d) {
        best_d = d;
        final_id = nodes[c];
      }
    }
    if (level == nid_level) *nid = (uint32_t)final_id;
  } while (v->child_count[final_id] > 0);
  if (nid_level > 0 && level < nid_level) *nid = (uint32_t)final_id;
  *word = v->word_id[final_id];
  *weight = v->weight[final_id];
}

static int must_normalize(int scoring) { return scoring != 5; } /* ScoringObject.h:44-60 */

/* TemplatedVocabulary.h:1127-1195.  Outputs: BowVector (ids ascending, vals), per feature word / node / weight
 * (nullable).  Returns the BowVector size. */
int vso_bow_transform(void* p, const uint8_t* desc, int n, int levelsup, uint32_t* ids, double* vals,
                      uint32_t* feat_word, uint32_t* feat_node, double* feat_weight) {
  bow_vocab* v = (bow_vocab*)p;
  if (v->n_words == 0) return 0; /* empty() */
  memset(v->present, 0, (size_t)v->n_words);
  const int tf = v->weighting == 0 || v->weighting == 1;
  for (int i = 0; i < n; i++) {
    uint32_t id, nid;
    double w;
    transform_one(v, desc + (size_t)i * 32, levelsup, &id, &w, &nid);
    if (feat_word) feat_word[i] = id;
    if (feat_node) feat_node[i] = nid;
    if (feat_weight) feat_weight[i] = w;
    if (w > 0) {
      if (!v->present[id]) {
        v->present[id] = 1;
        v->acc[id] = w;
      } else if (tf) {
        v->acc[id] += w; /* addWeight; addIfNotExist leaves the value alone */
      }
    }
  }
  int m = 0;
  for (int id = 0; id < v->n_words; id++)
    if (v->present[id]) {
      ids[m] = (uint32_t)id;
      vals[m] = v->acc[id];
      m++;
    }
  const int must = must_normalize(v->scoring);
  if (tf && m > 0 && !must) {
    const double nd = m;
    for (int i = 0; i < m; i++) vals[i] /= nd;
  }
  if (must) { /* BowVector::normalize */
    double norm = 0.0;
    if (v->scoring != 1) {
      for (int i = 0; i < m; i++) norm += fabs(vals[i]);
    } else {
      for (int i = 0; i < m; i++) norm = fma(vals[i], vals[i], norm); /* fused: see the header */
      norm = sqrt(norm);
    }
    if (norm > 0.0)
      for (int i = 0; i < m; i++) vals[i] /= norm;
  }
  return m;
}

/* ScoringObject.cpp: all provided types sum over the common words in ascending order.  Returns NaN for KL. */
double vso_bow_score(int scoring, const uint32_t* ids1, const double* vals1, int n1, const uint32_t* ids2,
                     const double* vals2, int n2) {
  if (scoring == 3) return NAN;
  double score = 0;
  int i = 0, j = 0;
  while (i < n1 && j < n2) {
    if (ids1[i] == ids2[j]) {
      const double vi = vals1[i], wi = vals2[j];
      switch (scoring) {
        case 0: score += fabs(vi - wi) - fabs(vi) - fabs(wi); break;
        case 2: if (vi + wi != 0.0) score += vi * wi / (vi + wi); break;
        case 4: score += sqrt(vi * wi); break;
        default: score = fma(vi, wi, score); break; /* L2, dot product: fused, see the header */
      }
      i++;
      j++;
    } else if (ids1[i] < ids2[j]) {
      i++; /* lower_bound(v2 id): the same position, reached one step at a time */
    } else {
      j++;
    }
  }
  switch (scoring) {
    case 0: return -score / 2.0;
    case 1: return score >= 1 ? 1.0 : 1.0 - sqrt(1.0 - score);
    case 2: return 2. * score;
    default: return score;
  }
}
