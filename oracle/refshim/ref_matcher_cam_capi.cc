// ref_matcher_cam_capi.cc — the reference's own ORBmatcher::SearchForTriangulation_ (src/ORBmatcher.cc:1208-1432) running on the
// reference's own camera models (src/CameraModels/Pinhole.cpp, KannalaBrandt8.cpp, unmodified) (TEST INFRASTRUCTURE).
//
// liborbslam3_orbmatcher.so pins the matcher's loops with a restated pinhole camera that is handed F12 and the epipole;
// liborbslam3_frame.so pins the camera models pair by pair.  This library is the composition: keyframe poses go in, the reference
// forms the epipole, R12 / t12 of the four camera pairings and (pinhole) F12 itself, and every candidate pair goes through
// GeometricCamera::epipolarConstrain_ of the real class.  Built with -DORBSLAM3_REAL_CAMERAS (orbslam3_standin.h then includes
// include/CameraModels/*.h instead of its stand-in camera).
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#include "ORBmatcher.h"

using namespace ORB_SLAM3;

namespace {
struct KpPod { float x, y, size, angle, response; int32_t octave, class_id; };
static_assert(sizeof(KpPod) == sizeof(cv::KeyPoint), "KeyPoint layout");

struct KfArrays {                // one keyframe, the layout of orbfe_bow_side + the pose
    const KpPod* kps;            // mvKeysUn, or mvKeys ++ mvKeysRight when nleft >= 0
    const uint8_t* desc;
    int n, nleft;
    const uint8_t* hasMp;
    const float* uRight;         // nullable: all mono
    const uint32_t* nodes; const int32_t* off; const int32_t* idx; int nnodes;
    const float* sigma2; const float* scale; int nlevels;
    const float* cam;            // [2][8]: mpCamera, mpCamera2 (pinhole reads 4 of 8)
    const float* Tcw;            // 4x4 row-major, left camera
    const float* Trl;            // 4x4 row-major, right camera from left (KeyFrame::GetRightPose = mTrl * Tcw); used when nleft >= 0
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

void fill(KeyFrame& K, const KfArrays& a, int fisheye, std::vector<std::unique_ptr<GeometricCamera>>& cams, MapPoint* some) {
    std::vector<cv::KeyPoint> all(a.n);
    if (a.n) memcpy((void*)all.data(), a.kps, sizeof(KpPod) * (size_t)a.n);
    K.N = a.n;
    K.NLeft = a.nleft;
    if (a.nleft < 0) { K.mvKeysUn = all; K.mvKeys = all; }
    else {
        K.mvKeys.assign(all.begin(), all.begin() + a.nleft);
        K.mvKeysRight.assign(all.begin() + a.nleft, all.end());
        K.mvKeysUn = K.mvKeys;
        K.NRight = a.n - a.nleft;
    }
    cv::Mat d(a.n > 0 ? a.n : 1, 32, CV_8U);
    if (a.n > 0) memcpy(d.data, a.desc, 32 * (size_t)a.n);
    K.mDescriptors = a.n > 0 ? d : d.rowRange(0, 0);
    for (int k = 0; k < a.nnodes; ++k)
        for (int j = a.off[k]; j < a.off[k + 1]; ++j) K.mFeatVec.addFeature(a.nodes[k], (unsigned)a.idx[j]);
    K.mvpMapPoints.assign(a.n, nullptr);
    for (int i = 0; i < a.n; ++i) if (a.hasMp[i]) K.mvpMapPoints[i] = some;
    K.mvuRight.assign(a.n, -1.f);
    if (a.uRight) K.mvuRight.assign(a.uRight, a.uRight + a.n);
    K.mvLevelSigma2.assign(a.sigma2, a.sigma2 + a.nlevels);
    K.mvScaleFactors.assign(a.scale, a.scale + a.nlevels);
    cams.emplace_back(makeCamera(fisheye, a.cam));
    K.mpCamera = cams.back().get();
    if (a.nleft >= 0) {
        cams.emplace_back(makeCamera(fisheye, a.cam + 8));
        K.mpCamera2 = cams.back().get();
        K.Tcr = mat44(a.Trl);
    }
    K.Tcw = mat44(a.Tcw);
}

// The F12 argument as the reference's caller forms it (LocalMapping::CreateNewMapPoints, src/LocalMapping.cc:477, through
// ComputeF12_ :886-903): K1^-T [t12]x R12 K2^-1 of the two left cameras.  The reference's matcher ignores the argument
// (Pinhole::epipolarConstrain_ rebuilds the same product per pair); the orbfe shim uses it.
cv::Matx33f callerF12(KeyFrame& K1, KeyFrame& K2) {
    const cv::Matx33f R12 = K1.GetRotation_() * K2.GetRotation_().t();
    const cv::Matx31f t12 = -K1.GetRotation_() * K2.GetRotation_().t() * K2.GetTranslation_() + K1.GetTranslation_();
    const cv::Matx33f cross{0.f, -t12(2), t12(1), t12(2), 0.f, -t12(0), -t12(1), t12(0), 0.f};
    return K1.mpCamera->toK_().t().inv() * cross * R12 * K2.mpCamera->toK_().inv();
}
}  // namespace

extern "C" {
// fisheye = 0: Pinhole cameras (F12 formed by Pinhole::epipolarConstrain_ from R12, t12 and both K);  1: KannalaBrandt8.
// matches12[A.n]: partner in keyframe 2 or -1.  Returns nmatches.
int orbslam3_search_for_triangulation_cam(float nnratio, int checkOri, int fisheye, const KfArrays* A, const KfArrays* B, int onlyStereo,
                                          int coarse, int32_t* matches12) {
    ORBmatcher matcher(nnratio, checkOri != 0);
    MapPoint some;
    std::vector<std::unique_ptr<GeometricCamera>> cams;
    KeyFrame K1, K2;
    fill(K1, *A, fisheye, cams, &some);
    fill(K2, *B, fisheye, cams, &some);
    std::vector<std::pair<size_t, size_t>> pairs;
    const int n = matcher.SearchForTriangulation_(&K1, &K2, callerF12(K1, K2), pairs, onlyStereo != 0, coarse != 0);
    for (int i = 0; i < A->n; ++i) matches12[i] = -1;
    for (const auto& pr : pairs) matches12[pr.first] = (int32_t)pr.second;
    return n;
}
}
