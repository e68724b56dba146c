// ref_voc_capi.cc — C entry points around the reference's own vendored DBoW2 (Thirdparty/DBoW2: TemplatedVocabulary.h,
// FORB.cpp, BowVector.cpp, FeatureVector.cpp, ScoringObject.cpp), compiled unmodified against the OpenCV / boost
// stand-ins of oracle/refshim (TEST INFRASTRUCTURE).  ORBVocabulary = TemplatedVocabulary<FORB::TDescriptor, FORB>
// (include/ORBVocabulary.h:29-30); Frame::ComputeBoW (src/Frame.cc:721-728) calls transform(vCurrentDesc, mBowVec,
// mFeatVec, 4) on the descriptor rows.
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "Thirdparty/DBoW2/DBoW2/FORB.h"
#include "Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h"

typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> ORBVocabulary;

extern "C" {
void* orbslam3_voc_load_text(const char* path) {
    ORBVocabulary* v = new ORBVocabulary();
    if (!v->loadFromTextFile(path)) { delete v; return nullptr; }
    return v;
}
void orbslam3_voc_destroy(void* v) { delete static_cast<ORBVocabulary*>(v); }
int orbslam3_voc_words(void* v) { return (int)static_cast<ORBVocabulary*>(v)->size(); }

// transform(features, BowVector&, FeatureVector&, levelsup); the two maps flattened in key order.
// Capacities: n entries each (offsets: n + 1).
void orbslam3_voc_transform(void* vp, const uint8_t* desc, int n, int levelsup, uint32_t* words, double* values, int* nwords,
                            uint32_t* nodeIds, int32_t* offsets, int32_t* indices, int* nnodes) {
    ORBVocabulary& voc = *static_cast<ORBVocabulary*>(vp);
    std::vector<cv::Mat> rows;                       // Converter::toDescriptorVector: one 1 x 32 row per feature
    rows.reserve(n);
    cv::Mat all(n > 0 ? n : 1, 32, CV_8U);
    if (n > 0) memcpy(all.data, desc, 32 * (size_t)n);
    for (int i = 0; i < n; ++i) rows.push_back(all.row(i));
    DBoW2::BowVector bv;
    DBoW2::FeatureVector fv;
    voc.transform(rows, bv, fv, levelsup);
    int k = 0;
    for (const auto& e : bv) { words[k] = e.first; values[k] = e.second; ++k; }
    *nwords = k;
    int m = 0, pos = 0;
    for (const auto& e : fv) {
        nodeIds[m] = e.first;
        offsets[m] = pos;
        for (unsigned int idx : e.second) indices[pos++] = (int32_t)idx;
        ++m;
    }
    offsets[m] = pos;
    *nnodes = m;
}
}
