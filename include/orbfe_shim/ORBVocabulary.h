// ORBVocabulary.h — ORB_SLAM3::ORBVocabulary over the orbfe C ABI (include/orbfe.h, "BoW transform").
//
// The reference's type is DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB>
// (include/ORBVocabulary.h:28-29); Frame::ComputeBoW (src/Frame.cc:721-728) and KeyFrame::ComputeBoW
// (src/KeyFrame.cc:105-114) use exactly two members of it, kept here with their signatures:
//   bool loadFromTextFile(const std::string&)                                  (System.cc:109)
//   void transform(const std::vector<cv::Mat>& features, BowVector&, FeatureVector&, int levelsup)
// BowVector / FeatureVector are the DBoW2 std::map types; when DBoW2's headers are not on the include path
// (ORBFE_NO_DBOW2) equivalent std::map typedefs stand in.  Scoring (KeyFrameDatabase) is outside the path.
#ifndef ORBVOCABULARY_H
#define ORBVOCABULARY_H

#include <cstdint>
#include <map>
#include <string>
#include <vector>

#ifdef ORBFE_NO_OPENCV
#include "cv_standin.h"
#else
#include <opencv2/core/core.hpp>
#endif

#ifdef ORBFE_NO_DBOW2
namespace DBoW2 {
typedef std::map<unsigned int, double> BowVector;
typedef std::map<unsigned int, std::vector<unsigned int> > FeatureVector;
}
#else
#include "Thirdparty/DBoW2/DBoW2/BowVector.h"
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"
#endif

#include "../orbfe.h"

namespace ORB_SLAM3 {

class ORBVocabulary {
public:
    explicit ORBVocabulary(int device = 0) : device_(device) {}
    ~ORBVocabulary() { orbfe_vocabulary_destroy(v_); }
    ORBVocabulary(const ORBVocabulary&) = delete;
    ORBVocabulary& operator=(const ORBVocabulary&) = delete;

    // ORBvoc.txt is scored with L1_NORM: vectors are L1-normalised (TemplatedVocabulary.h:1201).
    bool loadFromTextFile(const std::string& filename) {
        orbfe_vocabulary_destroy(v_);
        v_ = nullptr;
        status_ = orbfe_vocabulary_load_text(filename.c_str(), ORBFE_BOW_NORM_L1, device_, &v_);
        return status_ == ORBFE_OK;
    }
    bool empty() const { return !v_ || orbfe_vocabulary_words(v_) == 0; }
    unsigned int size() const { return v_ ? (unsigned)orbfe_vocabulary_words(v_) : 0u; }

    void transform(const std::vector<cv::Mat>& features, DBoW2::BowVector& v, DBoW2::FeatureVector& fv, int levelsup) {
        v.clear();
        fv.clear();
        const int n = (int)features.size();
        if (!v_ || n == 0) return;
        std::vector<uint8_t> d(32 * (size_t)n);
        for (int i = 0; i < n; ++i) std::copy(features[i].ptr(), features[i].ptr() + 32, d.begin() + 32 * (size_t)i);
        transformRows(d.data(), n, v, fv, levelsup);
    }
    // Same on the n x 32 descriptor matrix itself (saves Converter::toDescriptorVector's per-row Mats).
    void transformRows(const uint8_t* desc, int n, DBoW2::BowVector& v, DBoW2::FeatureVector& fv, int levelsup) {
        v.clear();
        fv.clear();
        if (!v_ || n <= 0) return;
        std::vector<uint32_t> words(n), nodes(n);
        std::vector<double> values(n);
        std::vector<int32_t> offsets(n + 1), indices(n);
        int nw = 0, nn = 0;
        status_ = orbfe_bow_transform(v_, desc, n, levelsup, words.data(), values.data(), &nw, nodes.data(), offsets.data(),
                                      indices.data(), &nn);
        if (status_ != ORBFE_OK) return;
        for (int i = 0; i < nw; ++i) v.insert(v.end(), std::make_pair(words[i], values[i]));
        for (int k = 0; k < nn; ++k)
            fv.insert(fv.end(), std::make_pair(nodes[k], std::vector<unsigned int>(indices.begin() + offsets[k], indices.begin() + offsets[k + 1])));
    }
    orbfe_vocabulary* handle() const { return v_; }
    int lastStatus() const { return status_; }

private:
    orbfe_vocabulary* v_ = nullptr;
    int device_;
    int status_ = ORBFE_OK;
};

}  // namespace ORB_SLAM3
#endif
