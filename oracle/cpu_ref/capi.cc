// capi.cc — flat C entry points over the CPU oracle so tests/bench can drive it with ctypes.
// TEST INFRASTRUCTURE: never linked into, or called from, the product library.
#include <cstring>
#include <thread>
#include <vector>

#include "cvprim.h"
#include "bow_ref.h"
#include "match_ref.h"
#include "proj_ref.h"
#include "orb_ref.h"

using namespace orbref;

extern "C" {

void* orbref_create(int nf, float sf, int nl, int ini, int mn) { return new OrbExtractorRef(nf, sf, nl, ini, mn); }
void orbref_destroy(void* h) { delete (OrbExtractorRef*)h; }
void orbref_set_tie_rule(void* h, int rule) { ((OrbExtractorRef*)h)->pointerTieRule = rule; }

void orbref_tables(void* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2, int* per_level, int* umax) {
    OrbExtractorRef* e = (OrbExtractorRef*)h;
    for (int i = 0; i < e->nlevels; ++i) {
        scale[i] = e->mvScaleFactor[i];
        inv_scale[i] = e->mvInvScaleFactor[i];
        sigma2[i] = e->mvLevelSigma2[i];
        inv_sigma2[i] = e->mvInvLevelSigma2[i];
        per_level[i] = e->mnFeaturesPerLevel[i];
    }
    for (int i = 0; i < 16; ++i) umax[i] = e->umax[i];
}

// Returns monoIndex (>= 0) or -1 (empty image) or -2 (cap too small); *n_out = keypoint count.
int orbref_extract(void* h, const uint8_t* img, int w, int hgt, long stride, int lap0, int lap1, KeyPoint* kps,
                   uint8_t* desc, int cap, int* n_out) {
    OrbExtractorRef* e = (OrbExtractorRef*)h;
    std::vector<KeyPoint> k;
    std::vector<uint8_t> d;
    int mono = e->extract(img, w, hgt, stride, lap0, lap1, k, d);
    if (mono < 0) return mono;
    *n_out = (int)k.size();
    if ((int)k.size() > cap) return -2;
    if (!k.empty()) {
        std::memcpy(kps, k.data(), k.size() * sizeof(KeyPoint));
        std::memcpy(desc, d.data(), d.size());
    }
    return mono;
}

int orbref_level_size(void* h, int level, int* w, int* hgt) {
    OrbExtractorRef* e = (OrbExtractorRef*)h;
    if (level < 0 || level >= e->nlevels) return -1;
    *w = e->pyramid[level].w;
    *hgt = e->pyramid[level].h;
    return 0;
}
// dst: dense (h+38) x (w+38) bordered level.
int orbref_get_level(void* h, int level, uint8_t* dst) {
    OrbExtractorRef* e = (OrbExtractorRef*)h;
    if (level < 0 || level >= e->nlevels) return -1;
    std::memcpy(dst, e->pyramid[level].buf.data(), e->pyramid[level].buf.size());
    return 0;
}
// FAST candidates of a level in reference push order, as (x, y, response) float triples.
int orbref_get_candidates(void* h, int level, float* xyr, int cap) {
    OrbExtractorRef* e = (OrbExtractorRef*)h;
    if (level < 0 || level >= (int)e->candidates.size()) return -1;
    const auto& c = e->candidates[level];
    int n = (int)c.size() < cap ? (int)c.size() : cap;
    for (int i = 0; i < n; ++i) { xyr[3 * i] = c[i].x; xyr[3 * i + 1] = c[i].y; xyr[3 * i + 2] = c[i].response; }
    return (int)c.size();
}

// ---- primitives (for the cv2 known-answer tests)
void orbref_resize(const uint8_t* s, int sw, int sh, long ss, uint8_t* d, int dw, int dh, long ds) {
    resizeLinearU8(s, sw, sh, ss, d, dw, dh, ds);
}
void orbref_border101(uint8_t* buf, int w, int h, int b, long step) { fillBorderReflect101(buf, w, h, b, step); }
int orbref_fast(const uint8_t* img, int w, int h, long step, int t, int* xys, int cap) {
    std::vector<FastKp> k;
    fast9_16(img, w, h, step, t, k);
    int n = (int)k.size() < cap ? (int)k.size() : cap;
    for (int i = 0; i < n; ++i) { xys[3 * i] = k[i].x; xys[3 * i + 1] = k[i].y; xys[3 * i + 2] = k[i].score; }
    return (int)k.size();
}
void orbref_blur(const uint8_t* s, int w, int h, long ss, uint8_t* d, long ds) { gaussianBlur7x7(s, w, h, ss, d, ds); }
float orbref_fast_atan2(float y, float x) { return fastAtan2(y, x); }
void orbref_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx, int32_t* dist) {
    knn2Hamming(q, nq, t, nt, idx, dist);
}
int orbref_hamming(const uint8_t* a, const uint8_t* b) { return hammingDistance256(a, b); }

// ---- timed CPU baseline: extract `nframes` frames (dense w x h each) on `nthreads` host threads,
// one extractor instance per thread (the reference runs one instance per image thread,
// Frame.cc:119-122).  Returns total keypoints so the work cannot be optimised away.
long orbref_extract_many(int nf, float sf, int nl, int ini, int mn, const uint8_t* frames, int nframes, int w, int h,
                         int lap0, int lap1, int nthreads) {
    std::vector<long> counts(nthreads, 0);
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t)
        th.emplace_back([&, t]() {
            OrbExtractorRef e(nf, sf, nl, ini, mn);
            std::vector<KeyPoint> k;
            std::vector<uint8_t> d;
            for (int f = t; f < nframes; f += nthreads) {
                e.extract(frames + (size_t)f * w * h, w, h, w, lap0, lap1, k, d);
                counts[t] += (long)k.size();
            }
        });
    for (auto& x : th) x.join();
    long total = 0;
    for (long c : counts) total += c;
    return total;
}

// ---- matchers (match_ref.cc)
int orbref_descriptor_distance(const uint8_t* a, const uint8_t* b) { return descriptorDistanceSwar(a, b); }
void orbref_three_maxima(const int* counts, int L, int* ind) {
    int a = -1, b = -1, c = -1;
    computeThreeMaxima(counts, L, a, b, c);
    ind[0] = a; ind[1] = b; ind[2] = c;
}
// Runs both extractors' pyramids must be current (call orbref_extract on each first).
void orbref_stereo_match(void* hl, void* hr, const KeyPoint* kpL, const uint8_t* descL, int nL, const KeyPoint* kpR,
                         const uint8_t* descR, int nR, float bf, float b, float* uRight, float* depth, int32_t* bestDist) {
    computeStereoMatches(*(OrbExtractorRef*)hl, *(OrbExtractorRef*)hr, kpL, descL, nL, kpR, descR, nR, bf, b, uRight, depth,
                         bestDist);
}
int orbref_search_bow(const uint8_t* descA, int nA, int nnA, const uint32_t* nodeA, const int32_t* offA, const int32_t* idxA,
                      const uint8_t* maskA, const float* angA, const uint8_t* descB, int nB, int nnB, const uint32_t* nodeB,
                      const int32_t* offB, const int32_t* idxB, const uint8_t* maskB, const float* angB, float ratio,
                      int checkOri, int th, int strictTh, int nleftB, int32_t* a2b, int32_t* b2a) {
    FeatureVectorCsr fa{nnA, nodeA, offA, idxA}, fb{nnB, nodeB, offB, idxB};
    BowParams p{ratio, checkOri, th, strictTh, nleftB};
    return searchByBoW(descA, nA, fa, maskA, angA, descB, nB, fb, maskB, angB, p, a2b, b2a);
}
int orbref_search_triangulation(const KeyPoint* kpA, const uint8_t* descA, int nA, int nnA, const uint32_t* nodeA,
                                const int32_t* offA, const int32_t* idxA, const uint8_t* mpA, const float* urA,
                                const KeyPoint* kpB, const uint8_t* descB, int nB, int nnB, const uint32_t* nodeB,
                                const int32_t* offB, const int32_t* idxB, const uint8_t* mpB, const float* urB,
                                const float* sigma2B, const float* scaleB, const float* F12, float ex, float ey,
                                int checkOri, int onlyStereo, int coarse, int32_t* m12) {
    FeatureVectorCsr fa{nnA, nodeA, offA, idxA}, fb{nnB, nodeB, offB, idxB};
    TriParams p;
    p.checkOri = checkOri; p.onlyStereo = onlyStereo; p.coarse = coarse; p.ex = ex; p.ey = ey;
    std::memcpy(p.F12, F12, sizeof(p.F12));
    return searchForTriangulation(kpA, descA, nA, fa, mpA, urA, kpB, descB, nB, fb, mpB, urB, sigma2B, scaleB, p, m12);
}
void orbref_knn2_ratio(const uint8_t* q, int nq, const uint8_t* t, int nt, double ratio, int32_t* match) {
    knn2RatioMatch(q, nq, t, nt, ratio, match);
}
// Timed CPU baseline for matching: `npairs` independent knn2 problems on `nthreads` host threads.
long orbref_knn2_many(const uint8_t* q, int nq, const uint8_t* t, int nt, int npairs, long qstride, long tstride, int nthreads) {
    std::vector<long> acc(nthreads, 0);
    std::vector<std::thread> th;
    for (int w = 0; w < nthreads; ++w)
        th.emplace_back([&, w]() {
            std::vector<int32_t> idx(2 * (size_t)nq), dist(2 * (size_t)nq);
            for (int p = w; p < npairs; p += nthreads) {
                knn2Hamming(q + p * qstride, nq, t + p * tstride, nt, idx.data(), dist.data());
                acc[w] += dist[0];
            }
        });
    for (auto& x : th) x.join();
    long s = 0;
    for (long a : acc) s += a;
    return s;
}

}  // extern "C"

extern "C" {
// ---- KB8 fisheye (kb8_ref.cc)
void orbref_kb8_project(const float* cam, const float* p, float* uv) { kb8Project(*(const Kb8*)cam, p, uv); }
void orbref_kb8_unproject(const float* cam, float u, float v, float* ray) { kb8Unproject(*(const Kb8*)cam, u, v, ray); }
// KannalaBrandt8::TriangulateMatches_ per pair (the predicate behind epipolarConstrain_, KannalaBrandt8.cpp:423-426, 409-480)
void orbref_kb8_triangulate_pairs(const float* cam1, const float* cam2, const KeyPoint* kp1, const KeyPoint* kp2, int n, const float* R12,
                                  const float* t12, const float* sigma, const float* unc, float* depth, float* p3d, float* margins) {
    for (int i = 0; i < n; ++i)
        depth[i] = kb8TriangulateMatches(*(const Kb8*)cam1, *(const Kb8*)cam2, kp1[i], kp2[i], R12, t12, sigma[i], unc[i], p3d + 3 * i, margins + i);
}
int orbref_fisheye_stereo(const float* cam1, const float* cam2, const KeyPoint* kpL, const uint8_t* descL, int nL, int monoLeft,
                          const KeyPoint* kpR, const uint8_t* descR, int nR, int monoRight, const float* R12, const float* t12,
                          const float* sigma2, int32_t* l2r, int32_t* r2l, float* depth, float* p3d, float* margins) {
    return computeStereoFishEyeMatches(*(const Kb8*)cam1, *(const Kb8*)cam2, kpL, descL, nL, monoLeft, kpR, descR, nR, monoRight, R12,
                                       t12, sigma2, l2r, r2l, depth, p3d, margins);
}
// camA / camB: 2 x 8 floats (mpCamera, mpCamera2); R12: 4 x 9, t12: 4 x 3 in the order ll, lr, rl, rr.
int orbref_search_triangulation_kb8(const KeyPoint* kpA, const uint8_t* descA, int nA, int nnA, const uint32_t* nodeA,
                                    const int32_t* offA, const int32_t* idxA, const uint8_t* mpA, const float* sigma2A,
                                    const KeyPoint* kpB, const uint8_t* descB, int nB, int nnB, const uint32_t* nodeB,
                                    const int32_t* offB, const int32_t* idxB, const uint8_t* mpB, const float* sigma2B,
                                    const float* scaleB, const float* camA, const float* camB, int nleftA, int nleftB,
                                    const float* R12, const float* t12, float ex, float ey, int checkOri, int onlyStereo,
                                    int coarse, int32_t* m12, float* points12, float* margins) {
    FeatureVectorCsr fa{nnA, nodeA, offA, idxA}, fb{nnB, nodeB, offB, idxB};
    TriKb8Params p;
    p.checkOri = checkOri; p.onlyStereo = onlyStereo; p.coarse = coarse; p.ex = ex; p.ey = ey;
    p.nleftA = nleftA; p.nleftB = nleftB;
    std::memcpy(p.camA, camA, sizeof(p.camA));
    std::memcpy(p.camB, camB, sizeof(p.camB));
    std::memcpy(p.R12, R12, sizeof(p.R12));
    std::memcpy(p.t12, t12, sizeof(p.t12));
    return searchForTriangulationKb8(kpA, descA, nA, fa, mpA, sigma2A, kpB, descB, nB, fb, mpB, sigma2B, scaleB, p, m12, points12,
                                     margins);
}
// ---- DBoW2 transform (bow_ref.cc)
void* orbref_voc_create(int k, int L, int weighting, int norm, int nnodes, const int32_t* parent, const uint8_t* isLeaf,
                        const uint8_t* desc, const double* weight) {
    VocabularyRef* v = new VocabularyRef();
    if (!v->build(k, L, weighting, norm, nnodes, parent, isLeaf, desc, weight)) { delete v; return nullptr; }
    return v;
}
void orbref_voc_destroy(void* v) { delete (VocabularyRef*)v; }
int orbref_voc_words(void* v) { return ((VocabularyRef*)v)->nwords; }
// words / values: capacity n; nodeIds: n; offsets: n + 1; indices: n.  counts[0] = words, counts[1] = nodes.
void orbref_bow_transform(void* voc, const uint8_t* desc, int n, int levelsup, uint32_t* words, double* values,
                          uint32_t* nodeIds, int32_t* offsets, int32_t* indices, int32_t* counts) {
    std::vector<uint32_t> w, nd;
    std::vector<double> val;
    std::vector<int32_t> off, idx;
    bowTransform(*(VocabularyRef*)voc, desc, n, levelsup, w, val, nd, off, idx);
    std::copy(w.begin(), w.end(), words);
    std::copy(val.begin(), val.end(), values);
    std::copy(nd.begin(), nd.end(), nodeIds);
    std::copy(off.begin(), off.end(), offsets);
    std::copy(idx.begin(), idx.end(), indices);
    counts[0] = (int32_t)w.size(); counts[1] = (int32_t)nd.size();
}
// per-feature tap: word id, weight, node id of every feature (before the map stage)
void orbref_bow_descend(void* voc, const uint8_t* desc, int n, int levelsup, uint32_t* word, double* w, uint32_t* node) {
    for (int i = 0; i < n; ++i) ((VocabularyRef*)voc)->transformOne(desc + 32 * (size_t)i, levelsup, word[i], w[i], node[i]);
}
// timed CPU baseline: `nframes` transforms of n features each on `nthreads` threads
long orbref_bow_transform_many(void* voc, const uint8_t* desc, int n, int nframes, long stride, int levelsup, int nthreads) {
    std::vector<long> acc(nthreads, 0);
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t)
        th.emplace_back([&, t]() {
            std::vector<uint32_t> w, nd;
            std::vector<double> val;
            std::vector<int32_t> off, idx;
            for (int f = t; f < nframes; f += nthreads) {
                bowTransform(*(VocabularyRef*)voc, desc + f * stride, n, levelsup, w, val, nd, off, idx);
                acc[t] += (long)w.size();
            }
        });
    for (auto& x : th) x.join();
    long s = 0;
    for (long a : acc) s += a;
    return s;
}
// ---- projection-guided window searches (proj_ref.cc)
int orbref_search_by_projection(const KeyPoint* kps, const uint8_t* desc, int n, int nleft, const uint8_t* blocked,
                                const float* uRight, const int32_t* l2r, const int32_t* r2l, float minX, float minY,
                                float gridWInv, float gridHInv, const void* queries, int nq, int mode, int th, float nnRatio,
                                int linkStereo, int rotHist, const float* invLevelSigma2, int32_t* assign, int32_t* choice) {
    GridFrame F{kps, desc, n, nleft, blocked, uRight, l2r, r2l, minX, minY, gridWInv, gridHInv, invLevelSigma2};
    ProjParams p{mode, th, nnRatio, linkStereo, rotHist};
    return searchByProjection(F, (const ProjQuery*)queries, nq, p, assign, choice);
}
int orbref_search_for_initialization(const KeyPoint* kps1, const uint8_t* desc1, int n1, const KeyPoint* kps2, const uint8_t* desc2,
                                     int n2, float minX, float minY, float gridWInv, float gridHInv, float* prevMatched,
                                     int windowSize, float nnRatio, int checkOri, int32_t* matches12) {
    GridFrame F{kps2, desc2, n2, -1, nullptr, nullptr, nullptr, nullptr, minX, minY, gridWInv, gridHInv, nullptr};
    return searchForInitialization(kps1, desc1, n1, F, prevMatched, windowSize, nnRatio, checkOri != 0, matches12);
}
int orbref_features_in_area(const KeyPoint* kps, int n, int nleft, float minX, float minY, float gridWInv, float gridHInv,
                            float x, float y, float r, int minLevel, int maxLevel, int bRight, int32_t* out) {
    GridFrame F{kps, nullptr, n, nleft, nullptr, nullptr, nullptr, nullptr, minX, minY, gridWInv, gridHInv, nullptr};
    FrameGrid g(F);
    std::vector<int> v;
    g.featuresInArea(x, y, r, minLevel, maxLevel, bRight != 0, v);
    std::copy(v.begin(), v.end(), out);
    return (int)v.size();
}
void orbref_distinctive_descriptors(const uint8_t* desc, const int32_t* offsets, int npoints, int32_t* best) {
    for (int p = 0; p < npoints; ++p) best[p] = distinctiveDescriptor(desc + 32 * (size_t)offsets[p], offsets[p + 1] - offsets[p]);
}
void orbref_undistort_keypoints(const KeyPoint* kps, int n, const float* K4, const float* dist, int ndist, KeyPoint* out) {
    undistortKeyPoints(kps, n, K4, dist, ndist, out);
}
}
