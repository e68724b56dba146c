// ref_objects_capi.cc — the reference's matcher on the reference's OWN object model (TEST INFRASTRUCTURE).
//
// liborbslam3_objects.so compiles, unmodified and together: src/ORBmatcher.cc, src/KeyFrame.cc, src/Frame.cc, src/MapPoint.cc,
// src/ORBextractor.cc, src/CameraModels/Pinhole.cpp, KannalaBrandt8.cpp with their own headers (include/ORBmatcher.h, KeyFrame.h,
// Frame.h, MapPoint.h, CameraModels/*.h).  No data-holder stand-in for Frame / KeyFrame / MapPoint is left: what stays out is what
// these classes only hold pointers to or call once (ImuTypes.h, Config.h, G2oTypes.h, Map.h, Converter.h, KeyFrameDatabase.h,
// TwoViewReconstruction.h: refshim/frame_standin.h) and Eigen / boost (name-only stand-ins).
//
// The wrappers fill a default-constructed reference Frame member by member (all public), turn it into a KeyFrame with the
// reference's constructor KeyFrame(Frame&, Map*, KeyFrameDatabase*) (src/KeyFrame.cc:48-104: copies the features, the feature
// vector, the cameras, mTlr; SetPose builds Tcw_ / Ow_), and call the reference's ORBmatcher members.
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

// the wrappers set state the reference keeps private / protected (Frame::AssignFeaturesToGrid, MapPoint::mDescriptor, mbBad,
// mObservations): refshim/access_public.h, force-included ahead of everything by the Makefile, opens the access (layouts unchanged)
#include "ORBmatcher.h"
#include "KeyFrame.h"
#include "KeyFrameDatabase.h"
#include "Map.h"
#include "Pinhole.h"
#include "KannalaBrandt8.h"

#include <cassert>
#include "../cpu_ref/match_ref.h"

using namespace ORB_SLAM3;

// cv::undistortPoints of the stand-in (called by Frame.cc, which this library links) forwards to the oracle's restatement, as in
// refshim/ref_frame_capi.cc; no wrapper of this file reaches it.
namespace cv {
void orbref_undistort_points_hook(const float* xy, int n, const float K4[4], const float* dist, int ndist, const float P4[4], float* out) {
    assert(!std::memcmp(K4, P4, 4 * sizeof(float)));
    std::vector<orbref::KeyPoint> in(n), o(n);
    for (int i = 0; i < n; ++i) { in[i].x = xy[2 * i]; in[i].y = xy[2 * i + 1]; }
    orbref::undistortKeyPoints(in.data(), n, K4, dist, ndist, o.data());
    for (int i = 0; i < n; ++i) { out[2 * i] = o[i].x; out[2 * i + 1] = o[i].y; }
}
}  // namespace cv

namespace {
struct KpPod { float x, y, size, angle, response; int32_t octave, class_id; };
static_assert(sizeof(KpPod) == sizeof(cv::KeyPoint), "KeyPoint layout");

struct KfArrays {                // the same record as refshim/ref_matcher_cam_capi.cc
    const KpPod* kps;            // mvKeysUn, or mvKeys ++ mvKeysRight when nleft >= 0
    const uint8_t* desc;
    int n, nleft;
    const uint8_t* hasMp;
    const float* uRight;         // nullable: all mono
    const uint32_t* nodes; const int32_t* off; const int32_t* idx; int nnodes;
    const float* sigma2; const float* scale; int nlevels;
    const float* cam;            // [2][8]: mpCamera, mpCamera2 (pinhole reads 4 of 8)
    const float* Tcw;            // 4x4 row-major, left camera
    const float* Trl;            // 4x4 row-major, right camera from left; used when nleft >= 0
};

cv::Mat mat44(const float* t) {
    cv::Mat m(4, 4, CV_32F);
    memcpy(m.data, t, 64);
    return m;
}

GeometricCamera* makeCamera(int fisheye, const float* p) {
    if (fisheye) return new KannalaBrandt8(std::vector<float>(p, p + 8));
    return new Pinhole(std::vector<float>(p, p + 4));
}

struct World {                   // owners of everything a test's keyframes point to
    Map map;
    KeyFrameDatabase db;
    std::vector<std::unique_ptr<GeometricCamera>> cams;
    std::vector<std::unique_ptr<KeyFrame>> kfs;
    std::unique_ptr<MapPoint> some;
};

KeyFrame* makeKeyFrame(World& w, const KfArrays& a, int fisheye) {
    Frame F;
    std::vector<cv::KeyPoint> all(a.n);
    if (a.n) memcpy((void*)all.data(), a.kps, sizeof(KpPod) * (size_t)a.n);
    F.N = a.n;
    F.Nleft = a.nleft;
    if (a.nleft < 0) { F.mvKeys = all; F.mvKeysUn = all; F.Nright = -1; }
    else {
        F.mvKeys.assign(all.begin(), all.begin() + a.nleft);
        F.mvKeysRight.assign(all.begin() + a.nleft, all.end());
        F.mvKeysUn = F.mvKeys;
        F.Nright = a.n - a.nleft;
    }
    cv::Mat d(a.n > 0 ? a.n : 1, 32, CV_8U);
    if (a.n > 0) memcpy(d.data, a.desc, 32 * (size_t)a.n);
    F.mDescriptors = a.n > 0 ? d : d.rowRange(0, 0);
    for (int k = 0; k < a.nnodes; ++k)
        for (int j = a.off[k]; j < a.off[k + 1]; ++j) F.mFeatVec.addFeature(a.nodes[k], (unsigned)a.idx[j]);
    F.mvpMapPoints.assign(a.n, static_cast<MapPoint*>(nullptr));
    F.mvuRight.assign(a.n, -1.f);
    if (a.uRight) F.mvuRight.assign(a.uRight, a.uRight + a.n);
    F.mvDepth.assign(a.n, -1.f);
    F.mnScaleLevels = a.nlevels;
    F.mvLevelSigma2.assign(a.sigma2, a.sigma2 + a.nlevels);
    F.mvScaleFactors.assign(a.scale, a.scale + a.nlevels);
    F.mvInvLevelSigma2.resize(a.nlevels);
    for (int l = 0; l < a.nlevels; ++l) F.mvInvLevelSigma2[l] = 1.f / a.sigma2[l];
    w.cams.emplace_back(makeCamera(fisheye, a.cam));
    F.mpCamera = w.cams.back().get();
    F.mpCamera2 = nullptr;
    F.mTlr = cv::Mat::eye(4, 4, CV_32F);
    if (a.nleft >= 0) {
        w.cams.emplace_back(makeCamera(fisheye, a.cam + 8));
        F.mpCamera2 = w.cams.back().get();
        // Frame keeps both directions (Frame.cc:1066-1078: mTlr from the settings, mTrl its rigid inverse); the keyframe's right pose
        // comes out of mTlr (KeyFrame.cc:1045-1110).  Trl arrives from the test; Tlr is its rigid inverse, formed as the reference does.
        cv::Mat Trl = mat44(a.Trl);
        cv::Mat Rrl = Trl.rowRange(0, 3).colRange(0, 3), trl = Trl.rowRange(0, 3).col(3);
        cv::Mat Rlr = Rrl.t();
        cv::Mat tlr = -Rlr * trl;
        cv::Mat Tlr = cv::Mat::eye(4, 4, CV_32F);
        Rlr.copyTo(Tlr.rowRange(0, 3).colRange(0, 3));
        tlr.copyTo(Tlr.rowRange(0, 3).col(3));
        F.mTlr = Tlr;
        F.mTrl = Trl;
    }
    F.mTcw = mat44(a.Tcw);
    KeyFrame* K = new KeyFrame(F, &w.map, &w.db);
    w.kfs.emplace_back(K);
    if (!w.some) w.some.reset(new MapPoint(cv::Mat::zeros(3, 1, CV_32F), K, &w.map));
    for (int i = 0; i < a.n; ++i)
        if (a.hasMp[i]) K->AddMapPoint(w.some.get(), (size_t)i);
    return K;
}

// LocalMapping::ComputeF12_ (src/LocalMapping.cc:886-903): the argument LocalMapping::CreateNewMapPoints passes (:477)
cv::Matx33f callerF12(KeyFrame* K1, KeyFrame* K2) {
    const cv::Matx33f R12 = K1->GetRotation_() * K2->GetRotation_().t();
    const cv::Matx31f t12 = -K1->GetRotation_() * K2->GetRotation_().t() * K2->GetTranslation_() + K1->GetTranslation_();
    const cv::Matx33f cross{0.f, -t12(2), t12(1), t12(2), 0.f, -t12(0), -t12(1), t12(0), 0.f};
    return K1->mpCamera->toK_().t().inv() * cross * R12 * K2->mpCamera->toK_().inv();
}
}  // namespace

extern "C" {
// ORBmatcher::SearchForTriangulation_ (src/ORBmatcher.cc:1208-1432) on two real KeyFrames.  matches12[A.n]; returns nmatches.
int orbslam3_objects_search_for_triangulation(float nnratio, int checkOri, int fisheye, const KfArrays* A, const KfArrays* B,
                                              int onlyStereo, int coarse, int32_t* matches12) {
    World w;
    KeyFrame* K1 = makeKeyFrame(w, *A, fisheye);
    KeyFrame* K2 = makeKeyFrame(w, *B, fisheye);
    ORBmatcher matcher(nnratio, checkOri != 0);
    std::vector<std::pair<size_t, size_t>> pairs;
    const int n = matcher.SearchForTriangulation_(K1, K2, callerF12(K1, K2), pairs, onlyStereo != 0, coarse != 0);
    for (int i = 0; i < A->n; ++i) matches12[i] = -1;
    for (const auto& pr : pairs) matches12[pr.first] = (int32_t)pr.second;
    return n;
}

// ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12) (src/ORBmatcher.cc:823-963) on two real KeyFrames whose features
// with hasMp hold (distinct) real MapPoints.  a2b[A.n] = index in B of the MapPoint matched to A's feature, or -1.
int orbslam3_objects_search_by_bow_kf_kf(float nnratio, int checkOri, const KfArrays* A, const KfArrays* B, int32_t* a2b) {
    World w;
    KeyFrame* K1 = makeKeyFrame(w, *A, 0);
    KeyFrame* K2 = makeKeyFrame(w, *B, 0);
    // one MapPoint per observed feature, as a map would hold them
    std::vector<std::unique_ptr<MapPoint>> pts;
    auto populate = [&](KeyFrame* K, const KfArrays& a) {
        for (int i = 0; i < a.n; ++i)
            if (a.hasMp[i]) {
                pts.emplace_back(new MapPoint(cv::Mat::zeros(3, 1, CV_32F), K, &w.map));
                K->AddMapPoint(pts.back().get(), (size_t)i);
            }
    };
    populate(K1, *A);
    const size_t firstB = pts.size();
    populate(K2, *B);
    std::vector<int> featureOfB(pts.size() - firstB);
    {
        size_t k = 0;
        for (int i = 0; i < B->n; ++i) if (B->hasMp[i]) featureOfB[k++] = i;
    }
    ORBmatcher matcher(nnratio, checkOri != 0);
    std::vector<MapPoint*> m12;
    const int n = matcher.SearchByBoW(K1, K2, m12);
    for (int i = 0; i < A->n; ++i) {
        a2b[i] = -1;
        if (i < (int)m12.size() && m12[i])
            for (size_t k = firstB; k < pts.size(); ++k)
                if (pts[k].get() == m12[i]) { a2b[i] = featureOfB[k - firstB]; break; }
    }
    return n;
}
}

// ---- the per-frame tracking matchers on real Frames and real MapPoints; same C names and argument lists as refshim/ref_matcher_capi.cc
namespace {
cv::Mat descRow(const uint8_t* d) {
    cv::Mat m(1, 32, CV_8U);
    memcpy(m.data, d, 32);
    return m;
}
cv::Mat vec3(const float* v) {
    cv::Mat m(3, 1, CV_32F);
    memcpy(m.data, v, 12);
    return m;
}
// a reference Frame from the arrays of orbfe_grid_frame: features, image bounds (the statics of Frame.cc:158-159), the 64 x 48 grid
// filled by the reference's own AssignFeaturesToGrid
void fillRealFrame(Frame& F, GeometricCamera* cam, GeometricCamera* rig, MapPoint* old, const KpPod* kps, const uint8_t* desc, int n, int nleft,
                   const uint8_t* blocked, const float* uRight, float minX, float minY, float maxX, float maxY, const float* scaleFactors,
                   int nlevels) {
    std::vector<cv::KeyPoint> all(n);
    if (n) memcpy((void*)all.data(), kps, sizeof(KpPod) * (size_t)n);
    F.N = n; F.Nleft = nleft;
    cv::Mat d(n > 0 ? n : 1, 32, CV_8U);
    if (n > 0) memcpy(d.data, desc, 32 * (size_t)n);
    F.mDescriptors = n > 0 ? d : d.rowRange(0, 0);
    if (nleft == -1) { F.mvKeysUn = all; F.mvKeys = all; F.Nright = -1; F.mpCamera2 = nullptr; }
    else {
        F.mvKeys.assign(all.begin(), all.begin() + nleft); F.mvKeysRight.assign(all.begin() + nleft, all.end());
        F.mvKeysUn = F.mvKeys; F.mpCamera2 = rig; F.Nright = n - nleft;
    }
    F.mpCamera = cam;
    F.mvuRight.assign(n, -1.f);
    if (uRight) F.mvuRight.assign(uRight, uRight + n);
    F.mvDepth.assign(n, -1.f);
    F.mvScaleFactors.assign(scaleFactors, scaleFactors + nlevels);
    F.mnScaleLevels = nlevels;
    Frame::mnMinX = minX; Frame::mnMinY = minY; Frame::mnMaxX = maxX; Frame::mnMaxY = maxY;
    Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(maxX - minX);
    Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(maxY - minY);
    F.AssignFeaturesToGrid();
    F.mvpMapPoints.assign(n, static_cast<MapPoint*>(nullptr));
    F.mvbOutlier.assign(n, false);
    for (int i = 0; i < n; ++i) if (blocked && blocked[i]) F.mvpMapPoints[i] = old;
}
}  // namespace

extern "C" {
// SearchByProjection(Frame& F, const vector<MapPoint*>& vpMapPoints, th, bFarPoints, thFarPoints) (src/ORBmatcher.cc:44-236).
int orbslam3_search_by_projection_mappoints(float nnratio, const KpPod* kps, const uint8_t* desc, int n, int nleft, const uint8_t* blocked,
                                            const float* uRight, const int32_t* l2r, const int32_t* r2l, float minX, float minY, float maxX,
                                            float maxY, const float* scaleFactors, int nlevels, int m, const uint8_t* inView,
                                            const uint8_t* inViewR, const uint8_t* bad, const float* depth, const float* projX,
                                            const float* projY, const float* projXR, const float* projYR, const int32_t* level,
                                            const int32_t* levelR, const float* viewCos, const float* viewCosR, const uint8_t* mpDesc,
                                            const int32_t* obs, float th, int bFarPoints, float thFarPoints, int32_t* mpOfFeature) {
    World w;
    ORBmatcher matcher(nnratio, true);
    const float intr[8] = {1.f, 1.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    w.cams.emplace_back(makeCamera(0, intr));
    w.cams.emplace_back(makeCamera(0, intr));
    MapPoint old;
    old.nObs = 1;
    Frame F;
    fillRealFrame(F, w.cams[0].get(), w.cams[1].get(), &old, kps, desc, n, nleft, blocked, uRight, minX, minY, maxX, maxY, scaleFactors, nlevels);
    if (nleft != -1) {
        F.mvLeftToRightMatch.assign(nleft, -1);
        F.mvRightToLeftMatch.assign(n - nleft, -1);
        if (l2r) F.mvLeftToRightMatch.assign(l2r, l2r + nleft);
        if (r2l) F.mvRightToLeftMatch.assign(r2l, r2l + (n - nleft));
    }
    std::vector<std::unique_ptr<MapPoint>> mps(m);
    std::vector<MapPoint*> vp(m);
    for (int i = 0; i < m; ++i) {
        mps[i].reset(new MapPoint());
        MapPoint& p = *mps[i];
        p.mbTrackInView = inView[i]; p.mbTrackInViewR = inViewR[i]; p.mbBad = bad[i]; p.mTrackDepth = depth[i];
        p.mTrackProjX = projX[i]; p.mTrackProjY = projY[i]; p.mTrackProjXR = projXR[i]; p.mTrackProjYR = projYR[i];
        p.mnTrackScaleLevel = level[i]; p.mnTrackScaleLevelR = levelR[i]; p.mTrackViewCos = viewCos[i]; p.mTrackViewCosR = viewCosR[i];
        p.mDescriptor = descRow(mpDesc + 32 * (size_t)i); p.nObs = obs[i];
        p.mpReplaced = nullptr;
        vp[i] = &p;
    }
    const int nm = matcher.SearchByProjection(F, vp, th, bFarPoints != 0, thFarPoints);
    for (int i = 0; i < n; ++i) {
        mpOfFeature[i] = -1;
        MapPoint* q = F.mvpMapPoints[i];
        if (q && q != &old)
            for (int k = 0; k < m; ++k) if (mps[k].get() == q) { mpOfFeature[i] = k; break; }
    }
    return nm;
}

// SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono) (src/ORBmatcher.cc:2173-2389).
int orbslam3_search_by_projection_lastframe(int checkOri, const KpPod* kps, const uint8_t* desc, int n, int nleft, const uint8_t* blocked,
                                            const float* uRight, float minX, float minY, float maxX, float maxY, const float* scaleFactors,
                                            int nlevels, float mbf, float mb, const float* cam, const float* Tcw, const float* Tlw,
                                            const float* Trl, int m, const uint8_t* hasMp, const uint8_t* outlier, const int32_t* octave,
                                            const float* angle, const float* worldPos, const uint8_t* mpDesc, const int32_t* obs, float th,
                                            int bMono, int32_t* lastOfFeature) {
    World w;
    ORBmatcher matcher(0.9f, checkOri != 0);
    const float intr[8] = {cam[0], cam[1], cam[2], cam[3], 0.f, 0.f, 0.f, 0.f}, unit[8] = {1.f, 1.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    w.cams.emplace_back(makeCamera(0, intr));
    w.cams.emplace_back(makeCamera(0, unit));
    MapPoint old;
    old.nObs = 1;
    Frame Cur, Last;
    fillRealFrame(Cur, w.cams[0].get(), w.cams[1].get(), &old, kps, desc, n, nleft, blocked, uRight, minX, minY, maxX, maxY, scaleFactors, nlevels);
    Cur.mbf = mbf; Cur.mb = mb;
    Cur.SetPose(mat44(Tcw));
    if (Trl) Cur.mTrl = mat44(Trl).rowRange(0, 3).clone();       // Frame::mTrl is 3 x 4
    Last.N = m;
    Last.SetPose(mat44(Tlw));
    Last.Nleft = -1;
    Last.mvKeys.resize(m);
    for (int i = 0; i < m; ++i) { Last.mvKeys[i].octave = octave[i]; Last.mvKeys[i].angle = angle[i]; }
    Last.mvKeysUn = Last.mvKeys;
    Last.mvbOutlier.resize(m);
    std::vector<std::unique_ptr<MapPoint>> mps(m);
    Last.mvpMapPoints.assign(m, static_cast<MapPoint*>(nullptr));
    for (int i = 0; i < m; ++i) {
        mps[i].reset(new MapPoint());
        Last.mvbOutlier[i] = outlier[i] != 0;
        mps[i]->SetWorldPos(vec3(worldPos + 3 * (size_t)i));
        mps[i]->mDescriptor = descRow(mpDesc + 32 * (size_t)i);
        mps[i]->nObs = obs[i];
        mps[i]->mbBad = false;
        mps[i]->mpReplaced = nullptr;
        if (hasMp[i]) Last.mvpMapPoints[i] = mps[i].get();
    }
    const int nm = matcher.SearchByProjection(Cur, Last, th, bMono != 0);
    for (int i = 0; i < n; ++i) {
        lastOfFeature[i] = -1;
        MapPoint* q = Cur.mvpMapPoints[i];
        if (q && q != &old)
            for (int k = 0; k < m; ++k) if (mps[k].get() == q) { lastOfFeature[i] = k; break; }
    }
    return nm;
}
}

// ---- the keyframe-side searches on real KeyFrames (the reference's constructor, its own grid copy, IsInImage, GetFeaturesInArea, pose
// accessors) and real MapPoints (their own PredictScale, distance invariance, observations); records as in refshim/ref_matcher_capi.cc
namespace {
struct KfInput {
    const KpPod* kps; const uint8_t* desc; int n, nleft; const uint8_t* blocked; const float* uRight;
    float minX, minY, maxX, maxY; const float* scaleFactors; const float* invLevelSigma2; int nlevels;
    float cam[4]; float mbf; float logScaleFactor; const float* Tcw;
};
struct KfPoints {
    int m; const uint8_t* null; const uint8_t* bad; const uint8_t* flagged; const float* worldPos; const float* normal;
    const float* minDist; const float* maxDist; const uint8_t* desc;
};
struct Points {
    std::vector<std::unique_ptr<MapPoint>> own;
    std::vector<MapPoint*> vp;               // NULL where the test says so
    int indexOf(const MapPoint* p) const {
        for (size_t k = 0; k < own.size(); ++k) if (own[k].get() == p) return (int)k;
        return -1;
    }
};
void fillRealPoints(Points& P, const KfPoints& in) {
    P.own.resize(in.m);
    P.vp.assign(in.m, static_cast<MapPoint*>(nullptr));
    for (int i = 0; i < in.m; ++i) {
        P.own[i].reset(new MapPoint());
        MapPoint& p = *P.own[i];
        p.mbBad = in.bad && in.bad[i];
        p.SetWorldPos(vec3(in.worldPos + 3 * (size_t)i));
        p.mNormalVector = vec3(in.normal + 3 * (size_t)i);
        p.mNormalVectorx = cv::Matx31f(in.normal[3 * i], in.normal[3 * i + 1], in.normal[3 * i + 2]);
        p.mfMinDistance = in.minDist[i]; p.mfMaxDistance = in.maxDist[i];
        p.mDescriptor = descRow(in.desc + 32 * (size_t)i);
        p.nObs = 1;
        if (!(in.null && in.null[i])) P.vp[i] = &p;
    }
}
// a reference Frame from a KfInput: intrinsics are statics of Frame (Frame.cc:152-157), the keyframe copies them
void fillRealFrameFromInput(Frame& F, World& w, const KfInput& in, MapPoint* old) {
    const float intr[8] = {in.cam[0], in.cam[1], in.cam[2], in.cam[3], 0.f, 0.f, 0.f, 0.f};
    w.cams.emplace_back(makeCamera(0, intr));
    GeometricCamera* cam = w.cams.back().get();
    fillRealFrame(F, cam, cam, old, in.kps, in.desc, in.n, in.nleft, in.blocked, in.uRight, in.minX, in.minY, in.maxX, in.maxY,
                  in.scaleFactors, in.nlevels);
    Frame::fx = in.cam[0]; Frame::fy = in.cam[1]; Frame::cx = in.cam[2]; Frame::cy = in.cam[3];
    Frame::invfx = 1.0f / in.cam[0]; Frame::invfy = 1.0f / in.cam[1];
    F.mbf = in.mbf; F.mb = in.mbf / in.cam[0];
    F.mfLogScaleFactor = in.logScaleFactor;
    if (in.invLevelSigma2) F.mvInvLevelSigma2.assign(in.invLevelSigma2, in.invLevelSigma2 + in.nlevels);
    F.mTlr = cv::Mat::eye(4, 4, CV_32F);
    F.SetPose(mat44(in.Tcw));
}
// KeyFrame(F, map, db); blocked features get an occupant MapPoint that observes the keyframe there
KeyFrame* makeRealKeyFrame(World& w, const KfInput& in, std::vector<std::unique_ptr<MapPoint>>& occupants) {
    Frame F;
    fillRealFrameFromInput(F, w, in, nullptr);
    KeyFrame* K = new KeyFrame(F, &w.map, &w.db);
    w.kfs.emplace_back(K);
    for (int i = 0; i < in.n; ++i)
        if (in.blocked && in.blocked[i]) {
            occupants.emplace_back(new MapPoint());
            occupants.back()->AddObservation(K, (size_t)i);
            K->AddMapPoint(occupants.back().get(), (size_t)i);
        }
    return K;
}
}  // namespace

extern "C" {
// SearchByProjection(KeyFrame* pKF, cv::Mat Scw, vpPoints, vpMatched, th, ratioHamming) (src/ORBmatcher.cc:473-585)
int orbslam3_search_by_projection_sim3(const KfInput* in, const float* Scw, const KfPoints* P, int th, float ratioHamming, int32_t* pointOfFeature) {
    World w;
    ORBmatcher matcher(0.9f, true);
    std::vector<std::unique_ptr<MapPoint>> occ;
    KeyFrame* K = makeRealKeyFrame(w, *in, occ);
    Points pts;
    fillRealPoints(pts, *P);
    std::vector<MapPoint*> vpMatched = K->GetMapPointMatches();
    for (int i = 0; i < P->m; ++i) if (P->flagged && P->flagged[i]) vpMatched.push_back(pts.own[i].get());
    const int nm = matcher.SearchByProjection(K, mat44(Scw), pts.vp, vpMatched, th, ratioHamming);
    for (int i = 0; i < in->n; ++i) pointOfFeature[i] = vpMatched[i] ? pts.indexOf(vpMatched[i]) : -1;
    return nm;
}

// SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th) (src/ORBmatcher.cc:1947-2171)
int orbslam3_search_by_sim3(const KfInput* in1, const KfInput* in2, const KfPoints* P1, const KfPoints* P2, float s12, const float* R12,
                            const float* t12, float th, int32_t* matches12) {
    World w;
    ORBmatcher matcher(0.9f, true);
    std::vector<std::unique_ptr<MapPoint>> occ;
    KfInput a = *in1, b = *in2;
    a.blocked = nullptr; b.blocked = nullptr;                  // the keyframes' MapPoints are P1 / P2 themselves
    KeyFrame* K1 = makeRealKeyFrame(w, a, occ);
    KeyFrame* K2 = makeRealKeyFrame(w, b, occ);
    Points p1, p2;
    fillRealPoints(p1, *P1);
    fillRealPoints(p2, *P2);
    for (int i = 0; i < in1->n; ++i) if (p1.vp[i]) K1->AddMapPoint(p1.vp[i], (size_t)i);
    for (int i = 0; i < in2->n; ++i) if (p2.vp[i]) K2->AddMapPoint(p2.vp[i], (size_t)i);
    std::vector<int> marked;
    for (int i = 0; i < in2->n; ++i) if (P2->flagged && P2->flagged[i]) marked.push_back(i);
    std::vector<std::unique_ptr<MapPoint>> holders(in1->n);
    std::vector<MapPoint*> vpMatches12(in1->n, static_cast<MapPoint*>(nullptr));
    size_t k = 0;
    for (int i = 0; i < in1->n; ++i)
        if (P1->flagged && P1->flagged[i]) {
            holders[i].reset(new MapPoint());
            if (k < marked.size()) holders[i]->AddObservation(K2, (size_t)marked[k]);
            ++k;
            vpMatches12[i] = holders[i].get();
        }
    cv::Mat R(3, 3, CV_32F);
    memcpy(R.data, R12, 36);
    const int nFound = matcher.SearchBySim3(K1, K2, vpMatches12, s12, R, vec3(t12), th);
    for (int i = 0; i < in1->n; ++i) matches12[i] = vpMatches12[i] ? p2.indexOf(vpMatches12[i]) : -1;
    return nFound;
}

// SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:2391-2510)
int orbslam3_search_by_projection_reloc(int checkOri, const KfInput* cur, const KfPoints* P, const float* kfAngle, float th, int ORBdist,
                                        int32_t* pointOfFeature) {
    World w;
    ORBmatcher matcher(0.9f, checkOri != 0);
    MapPoint old;
    old.nObs = 1;
    Frame F;
    fillRealFrameFromInput(F, w, *cur, &old);
    // the keyframe: P->m features that only carry an angle, and the MapPoints
    Frame G;
    G.N = P->m; G.Nleft = -1; G.Nright = -1;
    G.mvKeys.resize(P->m);
    for (int i = 0; i < P->m; ++i) G.mvKeys[i].angle = kfAngle[i];
    G.mvKeysUn = G.mvKeys;
    G.mvuRight.assign(P->m, -1.f);
    G.mvDepth.assign(P->m, -1.f);
    G.mvpMapPoints.assign(P->m, static_cast<MapPoint*>(nullptr));
    G.mDescriptors = cv::Mat(P->m > 0 ? P->m : 1, 32, CV_8U);
    G.mTlr = cv::Mat::eye(4, 4, CV_32F);
    G.mpCamera = F.mpCamera;
    G.SetPose(cv::Mat::eye(4, 4, CV_32F));
    KeyFrame* K = new KeyFrame(G, &w.map, &w.db);
    w.kfs.emplace_back(K);
    Points pts;
    fillRealPoints(pts, *P);
    for (int i = 0; i < P->m; ++i) if (pts.vp[i]) K->AddMapPoint(pts.vp[i], (size_t)i);
    std::set<MapPoint*> found;
    for (int i = 0; i < P->m; ++i) if (P->flagged && P->flagged[i]) found.insert(pts.own[i].get());
    const int nm = matcher.SearchByProjection(F, K, found, th, ORBdist);
    for (int i = 0; i < cur->n; ++i) {
        MapPoint* q = F.mvpMapPoints[i];
        pointOfFeature[i] = (q && q != &old) ? pts.indexOf(q) : -1;
    }
    return nm;
}
}

extern "C" {
// Fuse(KeyFrame* pKF, vpMapPoints, th, bRight) (src/ORBmatcher.cc:1629-1821; Scw == NULL; P.flagged = the point is already observed by
// pKF) and Fuse(KeyFrame* pKF, cv::Mat Scw, vpPoints, th, vpReplacePoint) (:1823-1945; P.flagged = the point is among pKF's MapPoints).
// The map surgery is the reference's own (MapPoint::AddObservation, Replace with its observation transfer, KeyFrame::AddMapPoint /
// ReplaceMapPointMatch); featureOfPoint[m] is read off the state it leaves: the feature where point i now sits in pKF, or where the
// MapPoint that replaced it sits, or (Scw form) where the reported vpReplacePoint sits; -1 = not fused.  Returns nFused.
int orbslam3_fuse(const KfInput* in, const float* Scw, const KfPoints* P, float th, int bRight, int32_t* featureOfPoint) {
    World w;
    ORBmatcher matcher(0.9f, true);
    std::vector<std::unique_ptr<MapPoint>> occ;
    KeyFrame* K = makeRealKeyFrame(w, *in, occ);
    long unsigned int nextId = 1;
    for (auto& o : occ) { o->nObs = 1; o->mnId = nextId++; o->mpMap = &w.map; o->mDescriptor = descRow(in->desc); }
    Points pts;
    fillRealPoints(pts, *P);
    for (auto& p : pts.own) { p->mnId = nextId++; p->mpMap = &w.map; }
    auto featureIn = [&](MapPoint* p) {
        const std::tuple<int, int> ix = p->GetIndexInKeyFrame(K);
        return std::get<0>(ix) >= 0 ? std::get<0>(ix) : std::get<1>(ix);
    };
    int nFused;
    std::vector<MapPoint*> vpReplace(P->m, static_cast<MapPoint*>(nullptr));
    if (Scw) {
        // "among pKF's MapPoints" without taking one of the test's features: entries behind the keyframe's own N
        for (int i = 0; i < P->m; ++i) if (P->flagged && P->flagged[i]) K->mvpMapPoints.push_back(pts.own[i].get());
        nFused = matcher.Fuse(K, mat44(Scw), pts.vp, th, vpReplace);
    } else {
        for (int i = 0; i < P->m; ++i)       // "already in this keyframe": an observation at a feature outside the arrays
            if (P->flagged && P->flagged[i]) pts.own[i]->mObservations[K] = std::tuple<int, int>(-1, -1);
        nFused = matcher.Fuse(K, pts.vp, th, bRight != 0);
    }
    for (int i = 0; i < P->m; ++i) {
        MapPoint* p = pts.own[i].get();
        int f = -1;
        if (!(P->flagged && P->flagged[i]) || Scw) f = p->IsInKeyFrame(K) ? featureIn(p) : -1;
        if (Scw && P->flagged && P->flagged[i]) f = -1;            // it sat behind N before the call, not at a feature
        if (f < 0 && p->isBad() && p->GetReplaced() && !(P->bad && P->bad[i])) f = featureIn(p->GetReplaced());
        if (f < 0 && vpReplace[i]) f = featureIn(vpReplace[i]);
        featureOfPoint[i] = f;
    }
    return nFused;
}
}

namespace {
// a keyframe / frame that only carries what SearchByBoW reads: features, descriptors, the feature vector, MapPoints by mask
void fillBowFrame(Frame& F, const KpPod* kps, const uint8_t* desc, int n, int nleft, const uint32_t* nodes, const int32_t* off,
                  const int32_t* idx, int nnodes, GeometricCamera* cam, GeometricCamera* rig) {
    std::vector<cv::KeyPoint> all(n);
    if (n) memcpy((void*)all.data(), kps, sizeof(KpPod) * (size_t)n);
    F.N = n; F.Nleft = nleft; F.Nright = nleft == -1 ? -1 : n - nleft;
    if (nleft == -1) { F.mvKeys = all; F.mvKeysUn = all; }
    else { F.mvKeys.assign(all.begin(), all.begin() + nleft); F.mvKeysRight.assign(all.begin() + nleft, all.end()); F.mvKeysUn = F.mvKeys; }
    cv::Mat d(n > 0 ? n : 1, 32, CV_8U);
    if (n > 0) memcpy(d.data, desc, 32 * (size_t)n);
    F.mDescriptors = n > 0 ? d : d.rowRange(0, 0);
    for (int k = 0; k < nnodes; ++k)
        for (int j = off[k]; j < off[k + 1]; ++j) F.mFeatVec.addFeature(nodes[k], (unsigned)idx[j]);
    F.mvpMapPoints.assign(n, static_cast<MapPoint*>(nullptr));
    F.mvuRight.assign(n, -1.f);
    F.mvDepth.assign(n, -1.f);
    F.mpCamera = cam;
    F.mpCamera2 = nleft == -1 ? nullptr : rig;
    F.mTlr = cv::Mat::eye(4, 4, CV_32F);
    F.SetPose(cv::Mat::eye(4, 4, CV_32F));
}
}  // namespace

extern "C" {
// SearchByBoW.  form 0: (KeyFrame* A, Frame& B, vpMapPointMatches) (src/ORBmatcher.cc:269-471);  form 1: (KeyFrame* A, KeyFrame* B,
// vpMatches12) (:823-963).  Real KeyFrames (GetMapPointMatches, mFeatVec copied by the constructor), a real Frame, real MapPoints.
int orbslam3_search_bow(int form, float nnratio, int checkOri, const KpPod* kpsA, const uint8_t* descA, int nA, const uint8_t* maskA,
                        const uint32_t* nodesA, const int32_t* offA, const int32_t* idxA, int nnodesA, const KpPod* kpsB,
                        const uint8_t* descB, int nB, const uint8_t* maskB, const uint32_t* nodesB, const int32_t* offB, const int32_t* idxB,
                        int nnodesB, int nleftB, int32_t* a2b, int32_t* b2a) {
    World w;
    ORBmatcher matcher(nnratio, checkOri != 0);
    const float unit[8] = {1.f, 1.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    w.cams.emplace_back(makeCamera(0, unit));
    w.cams.emplace_back(makeCamera(0, unit));
    GeometricCamera *cam = w.cams[0].get(), *rig = w.cams[1].get();
    std::vector<std::unique_ptr<MapPoint>> pool;
    auto populate = [&](KeyFrame* K, const uint8_t* mask, int n) {
        std::vector<MapPoint*> v(n, static_cast<MapPoint*>(nullptr));
        for (int i = 0; i < n; ++i)
            if (mask && mask[i]) { pool.emplace_back(new MapPoint()); v[i] = pool.back().get(); K->AddMapPoint(v[i], (size_t)i); }
        return v;
    };
    for (int i = 0; i < nA; ++i) a2b[i] = -1;
    for (int i = 0; i < nB; ++i) b2a[i] = -1;
    // with a two-camera frame the reference reads the keyframe's features through mvKeys / mvKeysRight as well: every feature of A
    // belongs to its left camera (NLeft = nA), as in the stand-in wrapper
    Frame FA;
    fillBowFrame(FA, kpsA, descA, nA, (form == 0 && nleftB != -1) ? nA : -1, nodesA, offA, idxA, nnodesA, cam, rig);
    KeyFrame* A = new KeyFrame(FA, &w.map, &w.db);
    w.kfs.emplace_back(A);
    const std::vector<MapPoint*> ptsA = populate(A, maskA, nA);
    int n;
    if (form == 0) {
        Frame B;
        fillBowFrame(B, kpsB, descB, nB, nleftB, nodesB, offB, idxB, nnodesB, cam, rig);
        std::vector<MapPoint*> out;
        n = matcher.SearchByBoW(A, B, out);
        for (int j = 0; j < (int)out.size() && j < nB; ++j)
            if (out[j])
                for (int i = 0; i < nA; ++i)
                    if (ptsA[i] == out[j]) { b2a[j] = i; a2b[i] = j; break; }
    } else {
        Frame FB;
        fillBowFrame(FB, kpsB, descB, nB, -1, nodesB, offB, idxB, nnodesB, cam, rig);
        KeyFrame* B = new KeyFrame(FB, &w.map, &w.db);
        w.kfs.emplace_back(B);
        const std::vector<MapPoint*> ptsB = populate(B, maskB, nB);
        std::vector<MapPoint*> out;
        n = matcher.SearchByBoW(A, B, out);
        for (int i = 0; i < (int)out.size() && i < nA; ++i)
            if (out[i])
                for (int j = 0; j < nB; ++j)
                    if (ptsB[j] == out[i]) { a2b[i] = j; b2a[j] = i; break; }
    }
    return n;
}

// SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) (src/ORBmatcher.cc:706-821) on two real Frames.
int orbslam3_search_for_initialization(float nnratio, int checkOri, const KpPod* kps1, const uint8_t* desc1, int n1, const KpPod* kps2,
                                       const uint8_t* desc2, int n2, float minX, float minY, float maxX, float maxY, float* prevMatched,
                                       int windowSize, int32_t* matches12) {
    World w;
    ORBmatcher matcher(nnratio, checkOri != 0);
    const float unit[8] = {1.f, 1.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    w.cams.emplace_back(makeCamera(0, unit));
    GeometricCamera* cam = w.cams[0].get();
    const float one = 1.f;
    Frame F1, F2;
    fillRealFrame(F1, cam, cam, nullptr, kps1, desc1, n1, -1, nullptr, nullptr, minX, minY, maxX, maxY, &one, 1);
    fillRealFrame(F2, cam, cam, nullptr, kps2, desc2, n2, -1, nullptr, nullptr, minX, minY, maxX, maxY, &one, 1);
    std::vector<cv::Point2f> prev(n1);
    for (int i = 0; i < n1; ++i) prev[i] = cv::Point2f(prevMatched[2 * i], prevMatched[2 * i + 1]);
    std::vector<int> m12;
    const int n = matcher.SearchForInitialization(F1, F2, prev, m12, windowSize);
    for (int i = 0; i < n1; ++i) { matches12[i] = m12[i]; prevMatched[2 * i] = prev[i].x; prevMatched[2 * i + 1] = prev[i].y; }
    return n;
}
}
