// Compiles include/visnav/keyframe_store_b200.h and drives it like loop closing would: keyframes are added one by
// one, then the newest one is matched against every earlier one in ONE call; every result must equal the plain
// matchDescriptors call on the same two descriptor sets.  usage: test_dropin_store  (prints "ok N jobs M matches")
#include <visnav/keypoints.h>
#include <visnav/keyframe_store_b200.h>

#include <cstdio>

int main() {
  using namespace visnav;
  try {
    std::vector<std::vector<std::bitset<256>>> kfs;
    b200::KeyframeStore store;
    unsigned s = 12345;
    auto rnd = [&]() { s = s * 1664525u + 1013904223u; return s >> 16; };  // high bits: long period
    std::vector<std::bitset<256>> base(600);
    for (auto& d : base)
      for (int b = 0; b < 256; ++b) d[b] = rnd() & 1;
    for (int k = 0; k < 7; ++k) {
      std::vector<std::bitset<256>> kd(k == 3 ? 0 : 600 - 37 * k);  // one empty keyframe, ragged sizes
      for (size_t i = 0; i < kd.size(); ++i) {
        kd[i] = base[(i * 7 + k) % base.size()];
        for (int f = 0; f < 12; ++f) kd[i].flip(rnd() % 256);
      }
      kfs.push_back(kd);
      if (store.add(kd) != k) { std::printf("bad index\n"); return 4; }
    }
    std::vector<std::pair<int, int>> jobs;
    for (int k = 0; k < 6; ++k) jobs.emplace_back(6, k);
    jobs.emplace_back(2, 5);
    std::vector<std::vector<std::pair<int, int>>> res;
    store.match(jobs, 70, 1.2, res);
    size_t total = 0;
    for (size_t k = 0; k < jobs.size(); ++k) {
      std::vector<std::pair<int, int>> want;
      matchDescriptors(kfs[jobs[k].first], kfs[jobs[k].second], want, 70, 1.2);
      if (want != res[k]) { std::printf("job %zu differs: %zu vs %zu\n", k, res[k].size(), want.size()); return 5; }
      total += want.size();
    }
    std::printf("ok %zu jobs %zu matches\n", jobs.size(), total);
  } catch (const std::exception& e) {
    std::printf("exception: %s\n", e.what());
    return 3;
  }
  return 0;
}
