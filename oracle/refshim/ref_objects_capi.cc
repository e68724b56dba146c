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
