// ref_shim_bow.cpp -- builds the REFERENCE's own DBoW2 (thirdparty/DBoW2_ORBSLAM, compiled from the sources where
// they lie: FORB.cpp, BowVector.cpp, FeatureVector.cpp, ScoringObject.cpp, TemplatedVocabulary.h, DUtils/Random.cpp,
// DUtils/Timestamp.cpp) behind a C interface, over oracle/stubs/opencv2/core/core.hpp (a byte-row cv::Mat).
// TEST INFRASTRUCTURE ONLY: only tests/, smoke() and bench.py's CPU baseline may load it.  The vocabulary type is
// the reference's visnav::ORBVocabulary (include/visnav/ORBVocabulary.h:13).
#include <DBoW2/FORB.h>
#include <DBoW2/TemplatedVocabulary.h>

#include <cstdint>
#include <cstring>
#include <map>
#include <string>
#include <vector>

typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> ORBVocabulary;

static std::vector<cv::Mat> to_rows(const uint8_t* desc, int n) {
  std::vector<cv::Mat> v((size_t)n);
  for (int i = 0; i < n; ++i) {
    v[i].create(1, 32, CV_8U);
    memcpy(v[i].ptr<unsigned char>(), desc + (size_t)i * 32, 32);
  }
  return v;
}

static DBoW2::BowVector to_bow(const uint32_t* ids, const double* vals, int n) {
  DBoW2::BowVector v;
  for (int i = 0; i < n; ++i) v.insert(v.end(), DBoW2::BowVector::value_type(ids[i], vals[i]));
  return v;
}

extern "C" {

void* vref_bow_load_text(const char* path) {
  ORBVocabulary* voc = new ORBVocabulary();
  if (!voc->loadFromTextFile(path)) {
    delete voc;
    return nullptr;
  }
  return voc;
}

void vref_bow_free(void* voc) { delete static_cast<ORBVocabulary*>(voc); }

// TemplatedVocabulary::create (hierarchical k-means++ over the training descriptors) + saveToTextFile
int vref_bow_create_save(const uint8_t* desc, const int32_t* set_begin, int n_sets, int k, int L, int weighting,
                         int scoring, const char* path) {
  std::vector<std::vector<cv::Mat>> training((size_t)n_sets);
  for (int s = 0; s < n_sets; ++s) training[s] = to_rows(desc + (size_t)set_begin[s] * 32, set_begin[s + 1] - set_begin[s]);
  ORBVocabulary voc(k, L, (DBoW2::WeightingType)weighting, (DBoW2::ScoringType)scoring);
  voc.create(training);
  voc.saveToTextFile(path);
  return (int)voc.size();
}

void vref_bow_info(void* voc, int32_t* info) {
  ORBVocabulary* v = static_cast<ORBVocabulary*>(voc);
  info[0] = v->getBranchingFactor();
  info[1] = v->getDepthLevels();
  info[2] = (int)v->getScoringType();
  info[3] = (int)v->getWeightingType();
  info[4] = 0;
  info[5] = (int)v->size();
}

// voc->transform(features, bow_vector, feature_vector, levelsup) (keypoints.h:253).  Returns the BowVector size;
// fv_node[f] = the FeatureVector entry that holds feature f, or 0xFFFFFFFF if the feature was stopped.
int vref_bow_transform(void* voc, const uint8_t* desc, int n, int levelsup, uint32_t* ids, double* vals,
                       uint32_t* fv_node) {
  ORBVocabulary* v = static_cast<ORBVocabulary*>(voc);
  std::vector<cv::Mat> f = to_rows(desc, n);
  DBoW2::BowVector bv;
  DBoW2::FeatureVector fv;
  v->transform(f, bv, fv, levelsup);
  int m = 0;
  for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it, ++m) {
    ids[m] = it->first;
    vals[m] = it->second;
  }
  for (int i = 0; i < n; ++i) fv_node[i] = 0xFFFFFFFFu;
  for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it)
    for (size_t j = 0; j < it->second.size(); ++j) fv_node[it->second[j]] = it->first;
  return m;
}

// single-feature transform: word id (TemplatedVocabulary.h:1050-1061)
void vref_bow_words(void* voc, const uint8_t* desc, int n, uint32_t* word) {
  ORBVocabulary* v = static_cast<ORBVocabulary*>(voc);
  std::vector<cv::Mat> f = to_rows(desc, n);
  for (int i = 0; i < n; ++i) {
    word[i] = v->transform(f[i]);
  }
}

double vref_bow_score(void* voc, const uint32_t* ids1, const double* vals1, int n1, const uint32_t* ids2,
                      const double* vals2, int n2) {
  ORBVocabulary* v = static_cast<ORBVocabulary*>(voc);
  return v->score(to_bow(ids1, vals1, n1), to_bow(ids2, vals2, n2));
}

}  // extern "C"
