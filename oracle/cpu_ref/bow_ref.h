// bow_ref.h — CPU restatement of the DBoW2 vocabulary-tree transform (TEST INFRASTRUCTURE).
//
// Follows /root/reference/Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h: transform(features, BowVector,
// FeatureVector, levelsup) :1137-1206, transform(feature, word, weight, nid, levelsup) :1230-1271,
// loadFromTextFile :1350-1428; FORB::distance (FORB.cpp:81-101, 256-bit Hamming); BowVector::addWeight /
// addIfNotExist / normalize (BowVector.cpp:30-98); FeatureVector::addFeature (FeatureVector.cpp:31-45).
// Called by Frame::ComputeBoW (src/Frame.cc:721-728) and KeyFrame::ComputeBoW (src/KeyFrame.cc:105-114)
// with levelsup = 4.  std::map outputs are flattened to ascending arrays, exactly as at the C ABI.
// One defined deviation: a leaf above the FeatureVector level leaves `nid` uninitialised in the reference
// (:1241-1265); here the leaf's own node id is reported.
#pragma once
#include <cstdint>
#include <vector>

namespace orbref {

enum BowWeighting { BOW_TF_IDF = 0, BOW_TF = 1, BOW_IDF = 2, BOW_BINARY = 3 };   // DBoW2::WeightingType
enum BowNorm { BOW_NORM_NONE = 0, BOW_NORM_L1 = 1, BOW_NORM_L2 = 2 };             // what mustNormalize() returns

struct VocabularyRef {
    int k = 0, L = 0, weighting = BOW_TF_IDF, norm = BOW_NORM_L1;
    // node 0 is the root; nodes in file order
    std::vector<int32_t> parent, childBase, childCount, childIdx, wordId;
    std::vector<uint8_t> desc;          // nnodes x 32
    std::vector<double> weight;
    int nwords = 0;
    // parent[i] < i for i > 0 (file order).  isLeaf[i] != 0 assigns the next word id (loadFromTextFile).
    bool build(int k, int L, int weighting, int norm, int nnodes, const int32_t* parent, const uint8_t* isLeaf,
               const uint8_t* desc, const double* weight);
    // :1230-1271
    void transformOne(const uint8_t* feature, int levelsup, uint32_t& wordId, double& w, uint32_t& nodeId) const;
};

// :1137-1206.  Outputs ascending by id: words/values (BowVector), nodeIds + CSR offsets/indices (FeatureVector).
void bowTransform(const VocabularyRef& v, const uint8_t* desc, int n, int levelsup, std::vector<uint32_t>& words,
                  std::vector<double>& values, std::vector<uint32_t>& nodeIds, std::vector<int32_t>& offsets,
                  std::vector<int32_t>& indices);

}  // namespace orbref
