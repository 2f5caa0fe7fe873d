// ORBVocabulary.h — drop-in for the reference's include/ORBVocabulary.h (typedef of
// DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>) for the members the hot path uses:
//   loadFromTextFile  (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1338-1424; System.cc loads Vocabulary/ORBvoc.txt)
//   transform(features, BowVector&, FeatureVector&, levelsup)  (:1127-1196; Frame::ComputeBoW src/Frame.cc:420-427,
//                                                               KeyFrame::ComputeBoW src/KeyFrame.cc:59-68)
// The tree lives on the device (liborbfe.so, orbfe_vocabulary_*); nothing is computed here.  score() (:1201-1205) only
// needs the two BowVectors and DBoW2's scoring object, it stays with DBoW2 (KeyFrameDatabase / LoopClosing are out of
// scope).
#ifndef ORBVOCABULARY_H
#define ORBVOCABULARY_H

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#ifdef ORBFE_HAVE_OPENCV
#include <opencv2/core/core.hpp>
#include "Thirdparty/DBoW2/DBoW2/BowVector.h"
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"
#else
#include "slam_min.h"
#endif

#include "../include/orbfe.h"

namespace ORB_SLAM2 {

class ORBVocabulary {
 public:
  explicit ORBVocabulary(int device = 0) : mDevice(device), mpVoc(NULL) {}
  ~ORBVocabulary() { orbfe_vocabulary_destroy(mpVoc); }

  bool loadFromTextFile(const std::string& filename) {
    orbfe_vocabulary_destroy(mpVoc);
    mpVoc = NULL;
    if (orbfe_vocabulary_load_text(mDevice, filename.c_str(), &mpVoc) != ORBFE_OK) {
      std::fprintf(stderr, "Vocabulary loading failure: %s\n", orbfe_last_error());  // the reference prints and returns false
      return false;
    }
    return true;
  }

  bool empty() const { return size() == 0; }
  unsigned int size() const {  // number of words
    int32_t info[6] = {0};
    return mpVoc && orbfe_vocabulary_info(mpVoc, info) == ORBFE_OK ? (unsigned)info[5] : 0u;
  }

  void transform(const std::vector<cv::Mat>& features, DBoW2::BowVector& v, DBoW2::FeatureVector& fv,
                 int levelsup) const {
    v.clear();
    fv.clear();
    const int n = (int)features.size();
    if (!mpVoc || n == 0) return;
    std::vector<uint8_t> desc((size_t)n * 32);
    for (int i = 0; i < n; ++i) std::memcpy(&desc[(size_t)i * 32], features[i].ptr(0), 32);  // Converter::toDescriptorVector rows
    std::vector<uint32_t> ids(n), nodes(n);
    std::vector<double> vals(n);
    std::vector<int32_t> off(n + 1), feat(n);
    int32_t nb = 0, nf = 0;
    if (orbfe_vocabulary_transform(mpVoc, desc.data(), n, levelsup, ids.data(), vals.data(), n, &nb, nodes.data(),
                                   off.data(), feat.data(), n, &nf, NULL, NULL, NULL) != ORBFE_OK) {
      std::fprintf(stderr, "ORBVocabulary::transform: %s\n", orbfe_last_error());
      return;
    }
    for (int j = 0; j < nb; ++j) v.insert(v.end(), std::make_pair(ids[j], vals[j]));  // ascending: O(1) hinted inserts
    for (int j = 0; j < nf; ++j) {
      std::vector<unsigned int>& dst = fv.insert(fv.end(), std::make_pair(nodes[j], std::vector<unsigned int>()))->second;
      dst.assign(feat.begin() + off[j], feat.begin() + off[j + 1]);
    }
  }

 private:
  ORBVocabulary(const ORBVocabulary&);
  ORBVocabulary& operator=(const ORBVocabulary&);
  int mDevice;
  orbfe_vocabulary* mpVoc;
};

}  // namespace ORB_SLAM2

#endif
