// ref_matcher_capi.cc — C entry points around the reference's own ORB_SLAM3::ORBmatcher, compiled from
// /root/reference/src/ORBmatcher.cc + include/ORBmatcher.h (unmodified) against the stand-ins of oracle/refshim
// (opencv2/opencv.hpp, orbslam3_standin.h) (TEST INFRASTRUCTURE).  Each wrapper builds the stand-in Frame / KeyFrame /
// MapPoint objects from flat arrays (the layout of include/orbfe.h) and calls the reference's member function.
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#include "ORBmatcher.h"

using namespace ORB_SLAM3;

namespace {
struct KpPod { float x, y, size, angle, response; int32_t octave, class_id; };
static_assert(sizeof(KpPod) == sizeof(cv::KeyPoint), "KeyPoint layout");

std::vector<cv::KeyPoint> keys(const KpPod* k, int n) {
    std::vector<cv::KeyPoint> v(n);
    if (n) memcpy(v.data(), k, sizeof(KpPod) * (size_t)n);
    return v;
}
cv::Mat descs(const uint8_t* d, int n) {
    cv::Mat m(n > 0 ? n : 1, 32, CV_8U);
    if (n > 0) memcpy(m.data, d, 32 * (size_t)n);
    return n > 0 ? m : m.rowRange(0, 0);
}
DBoW2::FeatureVector featvec(const uint32_t* nodeIds, const int32_t* off, const int32_t* idx, int nnodes) {
    DBoW2::FeatureVector fv;
    for (int k = 0; k < nnodes; ++k)
        for (int j = off[k]; j < off[k + 1]; ++j) fv.addFeature(nodeIds[k], (unsigned)idx[j]);
    return fv;
}
struct Matcher : ORBmatcher {
    using ORBmatcher::ORBmatcher;
    void threeMaxima(std::vector<int>* histo, int L, int& a, int& b, int& c) { ComputeThreeMaxima(histo, L, a, b, c); }
};
}  // namespace

extern "C" {
int orbslam3_descriptor_distance(const uint8_t* a, const uint8_t* b) {
    return ORBmatcher::DescriptorDistance(descs(a, 1), descs(b, 1));
}
void orbslam3_three_maxima(const int32_t* counts, int L, int32_t* ind) {
    std::vector<std::vector<int>> h(L);
    for (int i = 0; i < L; ++i) h[i].assign(counts[i], 0);
    int a = -1, b = -1, c = -1;
    Matcher m(0.6f, true);
    m.threeMaxima(h.data(), L, a, b, c);
    ind[0] = a; ind[1] = b; ind[2] = c;
}

// SearchByBoW.  form 0: (KeyFrame* A, Frame& B, vpMapPointMatches) :269-471;  form 1: (KeyFrame* A, KeyFrame* B, vpMatches12) :823-963.
// maskA / maskB: the feature holds a good MapPoint.  a2b[nA], b2a[nB]: matched partner or -1.  Returns nmatches.
int orbslam3_search_bow(int form, float nnratio, int checkOri, const KpPod* kpsA, const uint8_t* descA, int nA, const uint8_t* maskA,
                        const uint32_t* nodesA, const int32_t* offA, const int32_t* idxA, int nnodesA, const KpPod* kpsB,
                        const uint8_t* descB, int nB, const uint8_t* maskB, const uint32_t* nodesB, const int32_t* offB, const int32_t* idxB,
                        int nnodesB, int nleftB, int32_t* a2b, int32_t* b2a) {
    ORBmatcher matcher(nnratio, checkOri != 0);
    std::vector<std::unique_ptr<MapPoint>> pool;
    auto points = [&](const uint8_t* mask, int n) {
        std::vector<MapPoint*> v(n, nullptr);
        for (int i = 0; i < n; ++i)
            if (mask && mask[i]) { pool.emplace_back(new MapPoint()); v[i] = pool.back().get(); }
        return v;
    };
    KeyFrame A;
    A.N = nA; A.mvKeysUn = keys(kpsA, nA); A.mvKeys = A.mvKeysUn; A.mDescriptors = descs(descA, nA);
    A.mFeatVec = featvec(nodesA, offA, idxA, nnodesA); A.mvpMapPoints = points(maskA, nA);
    for (int i = 0; i < nA; ++i) a2b[i] = -1;
    for (int i = 0; i < nB; ++i) b2a[i] = -1;
    int n;
    Pinhole rig(1.f, 1.f, 0.f, 0.f);       // a second camera only has to exist: it switches the keypoint accessors to mvKeys / mvKeysRight
    if (form == 0) {
        Frame B;
        if (nleftB != -1) { A.mpCamera2 = &rig; A.NLeft = nA; B.mpCamera2 = &rig; }
        B.N = nB; B.Nleft = nleftB; B.mDescriptors = descs(descB, nB); B.mFeatVec = featvec(nodesB, offB, idxB, nnodesB);
        std::vector<cv::KeyPoint> all = keys(kpsB, nB);
        if (nleftB == -1) { B.mvKeys = all; B.mvKeysUn = all; }
        else { B.mvKeys.assign(all.begin(), all.begin() + nleftB); B.mvKeysRight.assign(all.begin() + nleftB, all.end()); B.mvKeysUn = B.mvKeys; }
        std::vector<MapPoint*> out;
        n = matcher.SearchByBoW(&A, B, out);
        for (int j = 0; j < (int)out.size() && j < nB; ++j)
            if (out[j])
                for (int i = 0; i < nA; ++i)
                    if (A.mvpMapPoints[i] == out[j]) { b2a[j] = i; a2b[i] = j; break; }
    } else {
        KeyFrame B;
        B.N = nB; B.mvKeysUn = keys(kpsB, nB); B.mvKeys = B.mvKeysUn; B.mDescriptors = descs(descB, nB);
        B.mFeatVec = featvec(nodesB, offB, idxB, nnodesB); B.mvpMapPoints = points(maskB, nB);
        std::vector<MapPoint*> out;
        n = matcher.SearchByBoW(&A, &B, out);
        for (int i = 0; i < (int)out.size() && i < nA; ++i)
            if (out[i])
                for (int j = 0; j < nB; ++j)
                    if (B.mvpMapPoints[j] == out[i]) { a2b[i] = j; b2a[j] = i; break; }
    }
    return n;
}

// SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) :706-821.
int orbslam3_search_for_initialization(float nnratio, int checkOri, const KpPod* kps1, const uint8_t* desc1, int n1, const KpPod* kps2,
                                       const uint8_t* desc2, int n2, float minX, float minY, float maxX, float maxY, float* prevMatched,
                                       int windowSize, int32_t* matches12) {
    ORBmatcher matcher(nnratio, checkOri != 0);
    Frame F1, F2;
    F1.N = n1; F1.mvKeysUn = keys(kps1, n1); F1.mvKeys = F1.mvKeysUn; F1.mDescriptors = descs(desc1, n1);
    F2.N = n2; F2.mvKeysUn = keys(kps2, n2); F2.mvKeys = F2.mvKeysUn; F2.mDescriptors = descs(desc2, n2);
    F2.setBounds(minX, minY, maxX, maxY);
    F2.AssignFeaturesToGrid();
    std::vector<cv::Point2f> prev(n1);
    for (int i = 0; i < n1; ++i) prev[i] = cv::Point2f(prevMatched[2 * i], prevMatched[2 * i + 1]);
    std::vector<int> m12;
    const int n = matcher.SearchForInitialization(F1, F2, prev, m12, windowSize);
    for (int i = 0; i < n1; ++i) { matches12[i] = m12[i]; prevMatched[2 * i] = prev[i].x; prevMatched[2 * i + 1] = prev[i].y; }
    return n;
}

// SearchByProjection(Frame& F, const vector<MapPoint*>& vpMapPoints, th, bFarPoints, thFarPoints) :44-236.
// Frame: kps = mvKeysUn (nleft == -1) or mvKeys ++ mvKeysRight; blocked[i]: mvpMapPoints[i] holds a MapPoint with observations.
// MapPoints: one array per member the function reads.  mpOfFeature[n]: index of the MapPoint that ends up in
// F.mvpMapPoints[i] (-1: entry untouched).  Returns nmatches.
int orbslam3_search_by_projection_mappoints(float nnratio, const KpPod* kps, const uint8_t* desc, int n, int nleft, const uint8_t* blocked,
                                            const float* uRight, const int32_t* l2r, const int32_t* r2l, float minX, float minY, float maxX,
                                            float maxY, const float* scaleFactors, int nlevels, int m, const uint8_t* inView,
                                            const uint8_t* inViewR, const uint8_t* bad, const float* depth, const float* projX,
                                            const float* projY, const float* projXR, const float* projYR, const int32_t* level,
                                            const int32_t* levelR, const float* viewCos, const float* viewCosR, const uint8_t* mpDesc,
                                            const int32_t* obs, float th, int bFarPoints, float thFarPoints, int32_t* mpOfFeature) {
    ORBmatcher matcher(nnratio, true);
    Frame F;
    F.N = n; F.Nleft = nleft; F.mDescriptors = descs(desc, n);
    std::vector<cv::KeyPoint> all = keys(kps, n);
    Pinhole rig(1.f, 1.f, 0.f, 0.f);
    if (nleft == -1) { F.mvKeysUn = all; F.mvKeys = all; }
    else { F.mvKeys.assign(all.begin(), all.begin() + nleft); F.mvKeysRight.assign(all.begin() + nleft, all.end()); F.mvKeysUn = F.mvKeys; F.mpCamera2 = &rig; F.Nright = n - nleft; }
    F.mvuRight.assign(n, -1.f);
    if (uRight) F.mvuRight.assign(uRight, uRight + n);
    if (nleft != -1) {
        F.mvLeftToRightMatch.assign(nleft, -1);
        F.mvRightToLeftMatch.assign(n - nleft, -1);
        if (l2r) F.mvLeftToRightMatch.assign(l2r, l2r + nleft);
        if (r2l) F.mvRightToLeftMatch.assign(r2l, r2l + (n - nleft));
    }
    F.mvScaleFactors.assign(scaleFactors, scaleFactors + nlevels);
    F.setBounds(minX, minY, maxX, maxY);
    F.AssignFeaturesToGrid();
    MapPoint old;
    old.nObs = 1;
    F.mvpMapPoints.assign(n, nullptr);
    for (int i = 0; i < n; ++i) if (blocked && blocked[i]) F.mvpMapPoints[i] = &old;
    std::vector<MapPoint> mps(m);
    std::vector<MapPoint*> vp(m);
    for (int i = 0; i < m; ++i) {
        MapPoint& p = mps[i];
        p.mbTrackInView = inView[i]; p.mbTrackInViewR = inViewR[i]; p.mbBad = bad[i]; p.mTrackDepth = depth[i];
        p.mTrackProjX = projX[i]; p.mTrackProjY = projY[i]; p.mTrackProjXR = projXR[i]; p.mTrackProjYR = projYR[i];
        p.mnTrackScaleLevel = level[i]; p.mnTrackScaleLevelR = levelR[i]; p.mTrackViewCos = viewCos[i]; p.mTrackViewCosR = viewCosR[i];
        p.mDescriptor = descs(mpDesc + 32 * (size_t)i, 1).clone(); p.nObs = obs[i];
        vp[i] = &p;
    }
    const int nm = matcher.SearchByProjection(F, vp, th, bFarPoints != 0, thFarPoints);
    for (int i = 0; i < n; ++i) {
        MapPoint* p = F.mvpMapPoints[i];
        mpOfFeature[i] = (p && p != &old) ? (int32_t)(p - mps.data()) : -1;
    }
    return nm;
}

// SearchForTriangulation_(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo, bCoarse) :1208-1432 with pinhole cameras.
// hasMp*: the feature already holds a MapPoint; uRight*: mvuRight; F12 row-major; (ex, ey): the epipole in image 2.
// matches12[n1]: partner in keyframe 2 or -1 (what vMatchedPairs lists in ascending idx1).  Returns nmatches.
int orbslam3_search_for_triangulation(float nnratio, int checkOri, const KpPod* kps1, const uint8_t* desc1, int n1, const uint8_t* hasMp1,
                                      const float* uRight1, const uint32_t* nodes1, const int32_t* off1, const int32_t* idx1, int nnodes1,
                                      const KpPod* kps2, const uint8_t* desc2, int n2, const uint8_t* hasMp2, const float* uRight2,
                                      const uint32_t* nodes2, const int32_t* off2, const int32_t* idx2, int nnodes2, const float* sigma2,
                                      const float* scale, int nlevels, const float* F12, float ex, float ey, int onlyStereo, int coarse,
                                      int32_t* matches12) {
    ORBmatcher matcher(nnratio, checkOri != 0);
    MapPoint some;
    Pinhole cam1(1.f, 1.f, 0.f, 0.f), cam2(1.f, 1.f, 0.f, 0.f);
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) cam1.F12(i, j) = F12[3 * i + j];
    cam2.fixedEpipole = true;
    cam2.epipole = cv::Point2f(ex, ey);
    auto fill = [&](KeyFrame& K, const KpPod* kps, const uint8_t* desc, int n, const uint8_t* hasMp, const float* ur, const uint32_t* nodes,
                    const int32_t* off, const int32_t* idx, int nnodes, Pinhole* cam) {
        K.N = n; K.mvKeysUn = keys(kps, n); K.mvKeys = K.mvKeysUn; K.mDescriptors = descs(desc, n);
        K.mFeatVec = featvec(nodes, off, idx, nnodes);
        K.mvpMapPoints.assign(n, nullptr);
        for (int i = 0; i < n; ++i) if (hasMp[i]) K.mvpMapPoints[i] = &some;
        K.mvuRight.assign(ur, ur + n);
        K.mvLevelSigma2.assign(sigma2, sigma2 + nlevels);
        K.mvScaleFactors.assign(scale, scale + nlevels);
        K.mpCamera = cam;
        K.Tcw = cv::Mat::zeros(4, 4, CV_32F);
        for (int i = 0; i < 4; ++i) K.Tcw.at<float>(i, i) = 1.f;
    };
    KeyFrame K1, K2;
    fill(K1, kps1, desc1, n1, hasMp1, uRight1, nodes1, off1, idx1, nnodes1, &cam1);
    fill(K2, kps2, desc2, n2, hasMp2, uRight2, nodes2, off2, idx2, nnodes2, &cam2);
    cv::Matx33f F;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) F(i, j) = F12[3 * i + j];
    std::vector<std::pair<size_t, size_t>> pairs;
    const int n = matcher.SearchForTriangulation_(&K1, &K2, F, pairs, onlyStereo != 0, coarse != 0);
    for (int i = 0; i < n1; ++i) matches12[i] = -1;
    for (const auto& pr : pairs) matches12[pr.first] = (int32_t)pr.second;
    return n;
}

namespace {
cv::Mat mat44(const float* t) {
    cv::Mat m(4, 4, CV_32F);
    memcpy(m.data, t, 64);
    return m;
}
cv::Mat vec3(const float* v) {
    cv::Mat m(3, 1, CV_32F);
    memcpy(m.data, v, 12);
    return m;
}
// A Frame from the arrays of orbfe_grid_frame (kps = mvKeysUn, or mvKeys ++ mvKeysRight when nleft != -1).
void fillFrame(Frame& F, Pinhole* cam, Pinhole* rig, MapPoint* old, const KpPod* kps, const uint8_t* desc, int n, int nleft, const uint8_t* blocked,
               const float* uRight, float minX, float minY, float maxX, float maxY, const float* scaleFactors, int nlevels) {
    F.N = n; F.Nleft = nleft; F.mDescriptors = descs(desc, n);
    std::vector<cv::KeyPoint> all = keys(kps, n);
    if (nleft == -1) { F.mvKeysUn = all; F.mvKeys = all; }
    else { F.mvKeys.assign(all.begin(), all.begin() + nleft); F.mvKeysRight.assign(all.begin() + nleft, all.end()); F.mvKeysUn = F.mvKeys; F.mpCamera2 = rig; F.Nright = n - nleft; }
    F.mpCamera = cam;
    F.mvuRight.assign(n, -1.f);
    if (uRight) F.mvuRight.assign(uRight, uRight + n);
    F.mvScaleFactors.assign(scaleFactors, scaleFactors + nlevels);
    F.mnScaleLevels = nlevels;
    F.setBounds(minX, minY, maxX, maxY);
    F.AssignFeaturesToGrid();
    F.mvpMapPoints.assign(n, nullptr);
    for (int i = 0; i < n; ++i) if (blocked && blocked[i]) F.mvpMapPoints[i] = old;
}
}  // namespace

extern "C" {
// SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono) :2173-2389.  cam: fx, fy, cx, cy of mpCamera;
// Tcw / Tlw: 4x4 row-major poses of the current / last frame; Trl: CurrentFrame.mTrl (3x4 used) or NULL.
// lastOfFeature[n]: LastFrame feature whose MapPoint ends up in CurrentFrame.mvpMapPoints[i], -1 otherwise.
int orbslam3_search_by_projection_lastframe(int checkOri, const KpPod* kps, const uint8_t* desc, int n, int nleft, const uint8_t* blocked,
                                            const float* uRight, float minX, float minY, float maxX, float maxY, const float* scaleFactors,
                                            int nlevels, float mbf, float mb, const float* cam, const float* Tcw, const float* Tlw,
                                            const float* Trl, int m, const uint8_t* hasMp, const uint8_t* outlier, const int32_t* octave,
                                            const float* angle, const float* worldPos, const uint8_t* mpDesc, const int32_t* obs, float th,
                                            int bMono, int32_t* lastOfFeature) {
    ORBmatcher matcher(0.9f, checkOri != 0);
    Pinhole camera(cam[0], cam[1], cam[2], cam[3]), rig(1.f, 1.f, 0.f, 0.f);
    MapPoint old;
    Frame Cur, Last;
    fillFrame(Cur, &camera, &rig, &old, kps, desc, n, nleft, blocked, uRight, minX, minY, maxX, maxY, scaleFactors, nlevels);
    Cur.mbf = mbf; Cur.mb = mb; Cur.mTcw = mat44(Tcw);
    if (Trl) Cur.mTrl = mat44(Trl).rowRange(0, 3).clone();       // Frame::mTrl is 3 x 4
    Last.N = m; Last.mTcw = mat44(Tlw);
    Last.mvKeys.resize(m);
    for (int i = 0; i < m; ++i) { Last.mvKeys[i].octave = octave[i]; Last.mvKeys[i].angle = angle[i]; }
    Last.mvKeysUn = Last.mvKeys;
    Last.mvbOutlier.resize(m);
    std::vector<MapPoint> mps(m);
    Last.mvpMapPoints.assign(m, nullptr);
    for (int i = 0; i < m; ++i) {
        Last.mvbOutlier[i] = outlier[i] != 0;
        mps[i].mWorldPos = vec3(worldPos + 3 * (size_t)i);
        mps[i].mDescriptor = descs(mpDesc + 32 * (size_t)i, 1).clone();
        mps[i].nObs = obs[i];
        if (hasMp[i]) Last.mvpMapPoints[i] = &mps[i];
    }
    const int nm = matcher.SearchByProjection(Cur, Last, th, bMono != 0);
    for (int i = 0; i < n; ++i) {
        MapPoint* p = Cur.mvpMapPoints[i];
        lastOfFeature[i] = (p && p != &old) ? (int32_t)(p - mps.data()) : -1;
    }
    return nm;
}
}
}

namespace {
struct KfInput {             // a KeyFrame in the arrays of orbfe_grid_frame + what the KeyFrame-side searches read besides
    const KpPod* kps; const uint8_t* desc; int n, nleft; const uint8_t* blocked; const float* uRight;
    float minX, minY, maxX, maxY; const float* scaleFactors; const float* invLevelSigma2; int nlevels;
    float cam[4]; float mbf; float logScaleFactor; const float* Tcw;
};
struct KfPoints {            // vpPoints as arrays
    int m; const uint8_t* null; const uint8_t* bad; const uint8_t* flagged; const float* worldPos; const float* normal;
    const float* minDist; const float* maxDist; const uint8_t* desc;
};
void fillKeyFrame(KeyFrame& K, Pinhole* cam, std::vector<MapPoint>& occupants, const KfInput& in) {
    const int n = in.n;
    K.N = n; K.NLeft = in.nleft; K.mDescriptors = descs(in.desc, n);
    std::vector<cv::KeyPoint> all = keys(in.kps, n);
    if (in.nleft == -1) { K.mvKeysUn = all; K.mvKeys = all; }
    else { K.mvKeys.assign(all.begin(), all.begin() + in.nleft); K.mvKeysRight.assign(all.begin() + in.nleft, all.end()); K.mvKeysUn = K.mvKeys; K.mpCamera2 = cam; K.NRight = n - in.nleft; }
    K.mpCamera = cam;
    K.fx = in.cam[0]; K.fy = in.cam[1]; K.cx = in.cam[2]; K.cy = in.cam[3]; K.mbf = in.mbf; K.mfLogScaleFactor = in.logScaleFactor;
    K.mvuRight.assign(n, -1.f);
    if (in.uRight) K.mvuRight.assign(in.uRight, in.uRight + n);
    K.mvScaleFactors.assign(in.scaleFactors, in.scaleFactors + in.nlevels);
    if (in.invLevelSigma2) K.mvInvLevelSigma2.assign(in.invLevelSigma2, in.invLevelSigma2 + in.nlevels);
    K.mnScaleLevels = in.nlevels;
    K.setBounds(in.minX, in.minY, in.maxX, in.maxY);
    K.AssignFeaturesToGrid();
    K.Tcw = mat44(in.Tcw);
    K.Tcr = cv::Mat::zeros(4, 4, CV_32F);
    for (int i = 0; i < 4; ++i) K.Tcr.at<float>(i, i) = 1.f;
    K.mvpMapPoints.assign(n, nullptr);
    occupants.assign(n, MapPoint());
    for (int i = 0; i < n; ++i)
        if (in.blocked && in.blocked[i]) { occupants[i].observe(&K, i); occupants[i].nObs = 1; K.mvpMapPoints[i] = &occupants[i]; }
}
void fillPoints(std::vector<MapPoint>& mps, std::vector<MapPoint*>& vp, const KfPoints& P) {
    mps.assign(P.m, MapPoint());
    vp.assign(P.m, nullptr);
    for (int i = 0; i < P.m; ++i) {
        MapPoint& p = mps[i];
        p.mbBad = P.bad && P.bad[i];
        p.mWorldPos = vec3(P.worldPos + 3 * (size_t)i);
        p.mNormalVector = vec3(P.normal + 3 * (size_t)i);
        p.mfMinDistance = P.minDist[i]; p.mfMaxDistance = P.maxDist[i];
        p.mDescriptor = descs(P.desc + 32 * (size_t)i, 1).clone();
        p.nObs = 1;
        if (!(P.null && P.null[i])) vp[i] = &p;
    }
}
}  // namespace

extern "C" {
// SearchByProjection(KeyFrame* pKF, cv::Mat Scw, vpPoints, vpMatched, th, ratioHamming) :473-585.  blocked = vpMatched[idx] != NULL;
// P.flagged = the point is in spAlreadyFound (it sits in vpMatched beyond the keyframe's own features).
// pointOfFeature[n]: index of the point written to vpMatched[idx], -1 elsewhere.
int orbslam3_search_by_projection_sim3(const KfInput* in, const float* Scw, const KfPoints* P, int th, float ratioHamming, int32_t* pointOfFeature) {
    ORBmatcher matcher(0.9f, true);
    Pinhole cam(in->cam[0], in->cam[1], in->cam[2], in->cam[3]);
    KeyFrame K;
    std::vector<MapPoint> occ, mps;
    std::vector<MapPoint*> vp;
    fillKeyFrame(K, &cam, occ, *in);
    fillPoints(mps, vp, *P);
    std::vector<MapPoint*> vpMatched(K.mvpMapPoints);
    for (int i = 0; i < P->m; ++i) if (P->flagged && P->flagged[i]) vpMatched.push_back(&mps[i]);
    const int nm = matcher.SearchByProjection(&K, mat44(Scw), vp, vpMatched, th, ratioHamming);
    for (int i = 0; i < in->n; ++i) {
        MapPoint* p = vpMatched[i];
        pointOfFeature[i] = (p && p >= mps.data() && p < mps.data() + mps.size()) ? (int32_t)(p - mps.data()) : -1;
    }
    return nm;
}

// Fuse(KeyFrame* pKF, vpMapPoints, th, bRight) :1629-1821 (scw == NULL; P.flagged = the point is already observed by pKF) and
// Fuse(KeyFrame* pKF, cv::Mat Scw, vpPoints, th, vpReplacePoint) :1823-1945 (P.flagged = the point is among pKF's MapPoints).
// featureOfPoint[m]: the feature each point was fused with (added, replaced or replacing), -1 = not fused.  Returns nFused.
int orbslam3_fuse(const KfInput* in, const float* Scw, const KfPoints* P, float th, int bRight, int32_t* featureOfPoint) {
    ORBmatcher matcher(0.9f, true);
    Pinhole cam(in->cam[0], in->cam[1], in->cam[2], in->cam[3]);
    KeyFrame K;
    std::vector<MapPoint> occ, mps;
    std::vector<MapPoint*> vp;
    fillKeyFrame(K, &cam, occ, *in);
    fillPoints(mps, vp, *P);
    MapPoint::fuseLog().clear();
    int nFused;
    std::vector<MapPoint*> vpReplace(P->m, nullptr);
    if (Scw) {
        for (int i = 0; i < P->m; ++i) if (P->flagged && P->flagged[i]) K.mspExtra.insert(&mps[i]);
        nFused = matcher.Fuse(&K, mat44(Scw), vp, th, vpReplace);
    } else {
        KeyFrame elsewhere;                  // "already in this keyframe": an observation at a feature outside the arrays
        for (int i = 0; i < P->m; ++i) if (P->flagged && P->flagged[i]) mps[i].mObservations[&K] = std::tuple<int, int>(-1, -1);
        nFused = matcher.Fuse(&K, vp, th, bRight != 0);
    }
    for (int i = 0; i < P->m; ++i) featureOfPoint[i] = -1;
    for (const auto& e : MapPoint::fuseLog())
        if (e.first >= mps.data() && e.first < mps.data() + mps.size() && e.second >= 0) featureOfPoint[e.first - mps.data()] = e.second;
    for (int i = 0; i < P->m; ++i)           // Fuse(Scw): a point that met an occupied feature is only reported through vpReplacePoint
        if (vpReplace[i] && featureOfPoint[i] < 0) {
            auto it = vpReplace[i]->mObservations.find(&K);
            if (it != vpReplace[i]->mObservations.end()) featureOfPoint[i] = std::get<0>(it->second);
        }
    return nFused;
}
}

extern "C" {
// SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th) :1947-2171.  P1 / P2: the keyframes' own MapPoints (null[i] = no MapPoint
// at feature i; flagged = vpMatches12[i] already holds a point (P1) / that point is observed by pKF2 at this feature (P2)).
// matches12[n1]: feature of pKF2 whose MapPoint the call wrote into vpMatches12[i], -1 elsewhere.  Returns nFound.
int orbslam3_search_by_sim3(const KfInput* in1, const KfInput* in2, const KfPoints* P1, const KfPoints* P2, float s12, const float* R12,
                            const float* t12, float th, int32_t* matches12) {
    ORBmatcher matcher(0.9f, true);
    Pinhole cam(in1->cam[0], in1->cam[1], in1->cam[2], in1->cam[3]);
    KeyFrame K1, K2;
    std::vector<MapPoint> occ1, occ2, mps1, mps2;
    std::vector<MapPoint*> vp1, vp2;
    fillKeyFrame(K1, &cam, occ1, *in1);
    fillKeyFrame(K2, &cam, occ2, *in2);
    fillPoints(mps1, vp1, *P1);
    fillPoints(mps2, vp2, *P2);
    K1.mvpMapPoints = vp1;
    K2.mvpMapPoints = vp2;
    std::vector<int> marked;
    for (int i = 0; i < in2->n; ++i) if (P2->flagged && P2->flagged[i]) marked.push_back(i);
    std::vector<MapPoint> holders(in1->n);
    std::vector<MapPoint*> vpMatches12(in1->n, nullptr);
    size_t k = 0;
    for (int i = 0; i < in1->n; ++i)
        if (P1->flagged && P1->flagged[i]) {
            if (k < marked.size()) holders[i].mObservations[&K2] = std::tuple<int, int>(marked[k], -1);
            ++k;
            vpMatches12[i] = &holders[i];
        }
    cv::Mat R(3, 3, CV_32F);
    memcpy(R.data, R12, 36);
    const int nFound = matcher.SearchBySim3(&K1, &K2, vpMatches12, s12, R, vec3(t12), th);
    for (int i = 0; i < in1->n; ++i) {
        MapPoint* p = vpMatches12[i];
        matches12[i] = (p && p >= mps2.data() && p < mps2.data() + mps2.size()) ? (int32_t)(p - mps2.data()) : -1;
    }
    return nFound;
}
}

extern "C" {
// SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>& sAlreadyFound, th, ORBdist) :2391-2510 (relocalisation).
// cur: the frame (blocked = mvpMapPoints[i] != NULL); P: pKF's MapPoints by feature (null, bad, flagged = in sAlreadyFound);
// kfAngle[m]: pKF->mvKeysUn[i].angle.  pointOfFeature[n]: keyframe feature whose MapPoint landed in mvpMapPoints[i], -1 elsewhere.
int orbslam3_search_by_projection_reloc(int checkOri, const KfInput* cur, const KfPoints* P, const float* kfAngle, float th, int ORBdist,
                                        int32_t* pointOfFeature) {
    ORBmatcher matcher(0.9f, checkOri != 0);
    Pinhole cam(cur->cam[0], cur->cam[1], cur->cam[2], cur->cam[3]), rig(1.f, 1.f, 0.f, 0.f);
    MapPoint old;
    Frame F;
    fillFrame(F, &cam, &rig, &old, cur->kps, cur->desc, cur->n, cur->nleft, cur->blocked, cur->uRight, cur->minX, cur->minY, cur->maxX,
              cur->maxY, cur->scaleFactors, cur->nlevels);
    F.mTcw = mat44(cur->Tcw);
    F.mfLogScaleFactor = cur->logScaleFactor;
    KeyFrame K;
    std::vector<MapPoint> mps;
    std::vector<MapPoint*> vp;
    fillPoints(mps, vp, *P);
    K.N = P->m;
    K.mvpMapPoints = vp;
    K.mvKeysUn.resize(P->m);
    for (int i = 0; i < P->m; ++i) K.mvKeysUn[i].angle = kfAngle[i];
    std::set<MapPoint*> found;
    for (int i = 0; i < P->m; ++i) if (P->flagged && P->flagged[i]) found.insert(&mps[i]);
    const int nm = matcher.SearchByProjection(F, &K, found, th, ORBdist);
    for (int i = 0; i < cur->n; ++i) {
        MapPoint* p = F.mvpMapPoints[i];
        pointOfFeature[i] = (p && p != &old) ? (int32_t)(p - mps.data()) : -1;
    }
    return nm;
}
}
