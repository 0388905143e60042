// Compiles include/visnav/bow_b200.h and drives it like the reference drives its ORBVocabulary: load the text
// vocabulary, compute_bow_vector on images, score keyframes against each other and through the database.
// usage: test_dropin_bow <vocab.txt> <w> <h> <n_images> <raw u8 images file>
// prints, per image: "v <n_words> <sum of values> <n feature-vector nodes> <n features listed>", then the score matrix rows
// "s ..." and the database rows "q ..." (scores) / "c ..." (shared words).
#include <visnav/keypoints.h>
#include <visnav/bow_b200.h>

#include <cstdio>
#include <cstdlib>
#include <map>
#include <vector>

int main(int argc, char** argv) {
  using namespace visnav;
  typedef std::map<unsigned, double> BowVector;                    // DBoW2::BowVector's base class
  typedef std::map<unsigned, std::vector<unsigned>> FeatureVector;  // DBoW2::FeatureVector's base class
  try {
    b200::Vocabulary voc;
    if (argc < 6) {  // no arguments: only touch the GPU context (fails loudly without one)
      voc.loadFromTextFile("/nonexistent");
      return 1;
    }
    if (!voc.loadFromTextFile(argv[1])) { std::printf("not a vocabulary\n"); return 4; }
    const int w = std::atoi(argv[2]), h = std::atoi(argv[3]), n = std::atoi(argv[4]);
    std::vector<unsigned char> raw((size_t)w * h * n);
    FILE* f = std::fopen(argv[5], "rb");
    if (!f || std::fread(raw.data(), 1, raw.size(), f) != raw.size()) { std::printf("bad image file\n"); return 4; }
    std::fclose(f);
    std::printf("voc %d %d %u\n", voc.getBranchingFactor(), voc.getDepthLevels(), voc.size());
    std::vector<BowVector> vecs((size_t)n);
    b200::BowDatabase db(voc);
    for (int i = 0; i < n; ++i) {
      pangolin::ManagedImage<uint8_t> img;
      img.ptr = raw.data() + (size_t)i * w * h; img.w = w; img.h = h; img.pitch = w;
      FeatureVector fv;
      b200::compute_bow_vector(img, 1500, voc, vecs[i], fv);
      double sum = 0;
      size_t listed = 0;
      for (auto& kv : vecs[i]) sum += kv.second;
      for (auto& kv : fv) listed += kv.second.size();
      std::printf("v %zu %.17g %zu %zu\n", vecs[i].size(), sum, fv.size(), listed);
      if (db.add(vecs[i]) != i) return 5;
      img.ptr = nullptr;
    }
    for (int i = 0; i < n; ++i) {
      std::printf("s");
      for (int j = 0; j < n; ++j) std::printf(" %.17g", voc.score(vecs[i], vecs[j]));
      std::printf("\n");
    }
    for (int i = 0; i < n; ++i) {
      std::vector<double> scores;
      std::vector<int> shared;
      db.query(vecs[i], scores, shared);
      std::printf("q");
      for (double s : scores) std::printf(" %.17g", s);
      std::printf("\nc");
      for (int c : shared) std::printf(" %d", c);
      std::printf("\n");
    }
  } catch (const std::exception& e) {
    std::printf("exception: %s\n", e.what());
    return 3;
  }
  return 0;
}
