// bow_b200.h -- bag-of-words place recognition on the GPU: drop-in for the three things the reference
// does with its ORBVocabulary (include/visnav/ORBVocabulary.h:13 = DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>):
//
//   orb_voc->loadFromTextFile(path)                                       slam.cpp:373
//   compute_bow_vector(img, orb, 1500, voc, bow_vector, feature_vector)   keypoints.h:243-254, tracking.h:261
//   voc->score(v1, v2)                                                    loop_closure_utils.h:119, :201; tracking.h:208
//
//   visnav::b200::Vocabulary voc;                 voc.loadFromTextFile("ORBvoc.txt");
//   visnav::b200::compute_bow_vector(img, 1500, voc, bow_vector, feature_vector);   // cv::ORB + transform, on the GPU
//   double s = voc.score(v1, v2);
//   visnav::b200::BowDatabase db(voc);  db.add(kf.bow_vector);  db.query(v, scores, shared_words);
//
// BowVector / FeatureVector are templates over the container, so DBoW2::BowVector (std::map<WordId, WordValue>) and
// DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned>>) pass straight through, as do plain std::maps.
// Results are the reference's bit for bit (BowVector doubles included); the keypoints of a pyramid level come in
// raster order where the reference's order is std::nth_element's (unspecified) -- feature indices in the
// FeatureVector refer to that order.  Thin wrapper over vslam_bow_* / vslam_compute_bow_vector_batch
// (include/visnav_b200.h); no CPU path: failures throw.
#pragma once

#include <cstdint>
#include <map>
#include <string>
#include <vector>

#include "keypoints_b200.h"

namespace visnav {
namespace b200 {

class Vocabulary {
 public:
  Vocabulary() = default;
  Vocabulary(const Vocabulary&) = delete;
  Vocabulary& operator=(const Vocabulary&) = delete;
  ~Vocabulary() {
    if (v_) vslam_bow_vocab_destroy(v_);
  }

  // TemplatedVocabulary::loadFromTextFile (TemplatedVocabulary.h:1338-1425): false if the file is not a vocabulary
  bool loadFromTextFile(const std::string& filename) {
    if (v_) vslam_bow_vocab_destroy(v_);
    v_ = nullptr;
    const int rc = vslam_bow_vocab_load_text(context(), filename.c_str(), &v_);
    if (rc == VSLAM_E_ARG) return false;
    check(rc, "Vocabulary::loadFromTextFile");
    return true;
  }

  bool empty() const { return size() == 0; }
  unsigned size() const { return v_ ? (unsigned)info(5) : 0u; }
  int getBranchingFactor() const { return info(0); }
  int getDepthLevels() const { return info(1); }

  // voc->transform(features, v, fv, levelsup) for descriptors given as rows of 32 bytes (cv::Mat rows of cv::ORB)
  template <class BowVector, class FeatureVector>
  void transform(const uint8_t* descriptors, int n, BowVector& v, FeatureVector& fv, int levelsup) const {
    v.clear();
    fv.clear();
    if (!v_ || n <= 0) return;
    const int32_t begin[2] = {0, n};
    std::vector<uint32_t> ids((size_t)n), node((size_t)n);
    std::vector<double> vals((size_t)n), weight((size_t)n);
    int32_t m = 0;
    check(vslam_bow_transform_batch(context(), v_, descriptors, begin, 1, levelsup, ids.data(), vals.data(), &m, n, nullptr,
                                    node.data(), weight.data()),
          "Vocabulary::transform");
    for (int i = 0; i < m; ++i) v.insert(v.end(), typename BowVector::value_type(ids[i], vals[i]));
    for (int f = 0; f < n; ++f)
      if (weight[f] > 0) fv[node[f]].push_back((unsigned)f);  // FeatureVector::addFeature, FeatureVector.cpp:28-42
  }

  template <class BowVector>
  double score(const BowVector& a, const BowVector& b) const {
    std::vector<uint32_t> ia, ib;
    std::vector<double> xa, xb;
    flatten(a, ia, xa);
    flatten(b, ib, xb);
    double s = 0;
    check(vslam_bow_score(context(), v_, ia.data(), xa.data(), (int)ia.size(), ib.data(), xb.data(), (int)ib.size(), &s),
          "Vocabulary::score");
    return s;
  }

  vslam_bow_vocab* handle() const { return v_; }

  template <class BowVector>
  static void flatten(const BowVector& v, std::vector<uint32_t>& ids, std::vector<double>& vals) {
    ids.clear();
    vals.clear();
    for (const auto& kv : v) {
      ids.push_back((uint32_t)kv.first);
      vals.push_back((double)kv.second);
    }
  }

 private:
  int info(int i) const {
    int32_t x[6] = {};
    if (v_) vslam_bow_vocab_info(v_, x);
    return x[i];
  }
  vslam_bow_vocab* v_ = nullptr;
};

// visnav::compute_bow_vector (keypoints.h:243-254): cv::ORB::create(num_features, 1.2, 8, 19, 0, 2, FAST_SCORE)
// ->detectAndCompute, then voc->transform(descriptors, bow_vector, feature_vector, 4).  Image: anything with
// ptr / w / h / pitch (pangolin::ManagedImage<uint8_t>) -- for a continuous cv::Mat pass (data, cols, rows, step).
template <class BowVector, class FeatureVector>
inline void compute_bow_vector(const uint8_t* ptr, int w, int h, int pitch, int num_features, const Vocabulary& voc,
                               BowVector& bow_vector, FeatureVector& feature_vector) {
  bow_vector.clear();
  feature_vector.clear();
  vslam_orb_params prm = {num_features, 1.2f, 8, 19, 20};
  int cap = num_features + num_features / 4 + 256;  // ties at a level's cut-off score all stay (retainBest)
  if (cap > 8192) cap = 8192;
  std::vector<uint32_t> ids((size_t)cap), node((size_t)cap);
  std::vector<double> vals((size_t)cap), weight((size_t)cap);
  int32_t m = 0, n = 0;
  check(vslam_compute_bow_vector_batch(context(), voc.handle(), ptr, 1, w, h, pitch, (size_t)pitch * h, &prm, 4, ids.data(),
                                       vals.data(), &m, cap, nullptr, nullptr, &n, node.data(), weight.data(), cap),
        "compute_bow_vector");
  for (int i = 0; i < m; ++i) bow_vector.insert(bow_vector.end(), typename BowVector::value_type(ids[i], vals[i]));
  for (int f = 0; f < n; ++f)
    if (weight[f] > 0) feature_vector[node[f]].push_back((unsigned)f);
}

template <class Image, class BowVector, class FeatureVector>
inline void compute_bow_vector(const Image& img, int num_features, const Vocabulary& voc, BowVector& bow_vector,
                               FeatureVector& feature_vector) {
  compute_bow_vector(img.ptr, (int)img.w, (int)img.h, (int)img.pitch, num_features, voc, bow_vector, feature_vector);
}

// The keyframes' BowVectors (Camera::bow_vector, common_types.h:213) on the device: one query scores
// voc->score(query, entry) against every entry and counts the shared words -- what detect_loop_candidates
// (loop_closure_utils.h:152-201) and detect_relocalization_candidate (tracking.h:176-208) gather through the inverted file.
class BowDatabase {
 public:
  explicit BowDatabase(const Vocabulary& voc) { check(vslam_bow_db_create(context(), voc.handle(), &db_), "BowDatabase"); }
  BowDatabase(const BowDatabase&) = delete;
  BowDatabase& operator=(const BowDatabase&) = delete;
  ~BowDatabase() {
    if (db_) vslam_bow_db_destroy(db_);
  }
  template <class BowVector>
  int add(const BowVector& v) {
    std::vector<uint32_t> ids;
    std::vector<double> vals;
    Vocabulary::flatten(v, ids, vals);
    int idx = -1;
    check(vslam_bow_db_add(db_, ids.data(), vals.data(), (int)ids.size(), &idx), "BowDatabase::add");
    return idx;
  }
  int size() const { return vslam_bow_db_size(db_); }
  template <class BowVector>
  void query(const BowVector& v, std::vector<double>& scores, std::vector<int>& shared_words) {
    std::vector<uint32_t> ids;
    std::vector<double> vals;
    Vocabulary::flatten(v, ids, vals);
    scores.assign((size_t)size(), 0.0);
    std::vector<int32_t> common((size_t)size());
    check(vslam_bow_db_query(context(), db_, ids.data(), vals.data(), (int)ids.size(), scores.data(), common.data()),
          "BowDatabase::query");
    shared_words.assign(common.begin(), common.end());
  }

 private:
  vslam_bow_db* db_ = nullptr;
};

}  // namespace b200
}  // namespace visnav
