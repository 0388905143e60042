// keyframe_store_b200.h -- device-resident keyframe descriptors for the repeated matchDescriptors calls of loop
// closing and relocalisation.  compute_sim3 (include/visnav/sim3.h:252,281) and relocalize_camera
// (include/visnav/tracking.h:283,317) match the current frame against stored keyframes -- the same stored
// descriptor sets again and again, and relocalisation repeats identical calls in its retry loop.  Keeping
// feature_corners' descriptors on the GPU removes the upload per call and lets one launch serve all candidates.
//
//   visnav::b200::KeyframeStore store;
//   int id = store.add(kd.corner_descriptors);              // when slam.cpp:1153-1154 stores a keyframe
//   store.match({{id_query, id_cand0}, {id_query, id_cand1}}, 70, 1.2, results);   // results[k] == what
//       matchDescriptors(descriptors of jobs[k].first, descriptors of jobs[k].second, ., 70, 1.2) returns
//
// Thin wrapper over vslam_desc_store_* / vslam_match_batch (include/visnav_b200.h); no CPU path: failures throw.
#pragma once

#include <bitset>
#include <cstdint>
#include <utility>
#include <vector>

#include "keypoints_b200.h"

namespace visnav {
namespace b200 {

class KeyframeStore {
 public:
  KeyframeStore() = default;
  KeyframeStore(const KeyframeStore&) = delete;
  KeyframeStore& operator=(const KeyframeStore&) = delete;
  ~KeyframeStore() {
    if (store_) vslam_desc_store_destroy(store_);
  }

  // descriptors of one keyframe (KeypointsData::corner_descriptors); returns the set index used in match()
  int add(const std::vector<std::bitset<256>>& descriptors) {
    const uint32_t* d = reinterpret_cast<const uint32_t*>(descriptors.data());
    const int n = (int)descriptors.size();
    if ((int)descriptors.size() > max_count_) max_count_ = n;
    if (!store_) {
      const int64_t off = 0;
      const int32_t cnt = n;
      const uint32_t dummy[8] = {};
      check(vslam_desc_store_create(context(), n ? d : dummy, &off, &cnt, 1, &store_), "KeyframeStore::add");
      return 0;
    }
    int idx = -1;
    check(vslam_desc_store_append(store_, d, n, &idx), "KeyframeStore::add");
    return idx;
  }

  int size() const { return store_ ? vslam_desc_store_size(store_) : 0; }

  // jobs[k] = (set index of descriptors_1, set index of descriptors_2); results[k] = the match list of that call
  void match(const std::vector<std::pair<int, int>>& jobs, int threshold, double dist_2_best,
             std::vector<std::vector<std::pair<int, int>>>& results) {
    results.assign(jobs.size(), {});
    if (jobs.empty() || !store_) return;
    static_assert(sizeof(std::pair<int, int>) == 8, "std::pair<int,int> must be two packed ints");
    const int cap = max_count_ > 0 ? max_count_ : 1;
    std::vector<int32_t> pairs(jobs.size() * (size_t)cap * 2), counts(jobs.size());
    check(vslam_match_batch(context(), store_, reinterpret_cast<const int32_t*>(jobs.data()), (int)jobs.size(),
                            threshold, dist_2_best, pairs.data(), counts.data(), cap),
          "KeyframeStore::match");
    for (size_t k = 0; k < jobs.size(); ++k) {
      const int32_t* p = pairs.data() + k * (size_t)cap * 2;
      results[k].reserve(counts[k]);
      for (int m = 0; m < counts[k]; ++m) results[k].emplace_back(p[2 * m], p[2 * m + 1]);
    }
  }

 private:
  vslam_desc_store* store_ = nullptr;
  int max_count_ = 0;
};

}  // namespace b200
}  // namespace visnav
