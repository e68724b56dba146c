// ORBmatcher.h — ORB_SLAM3::ORBmatcher over the orbfe C ABI (include/orbfe.h).
//
// Keeps the reference's class name, constructor, static DescriptorDistance and TH_LOW / TH_HIGH /
// HISTO_LENGTH (include/ORBmatcher.h:35-111).  The searches come in two layers:
//   * array layer (always available): the reference's inputs already flattened — descriptors,
//     keypoints, a FeatureVector as CSR, MapPoint-validity masks.  This is what the tests drive.
//   * object layer (#ifdef ORBFE_WITH_ORBSLAM3, needs the reference's Frame.h / KeyFrame.h / MapPoint.h): every
//     Search* / Fuse member of the reference's class with the reference's signature (ORBmatcherObjects.h); they flatten
//     the objects, call the C ABI and write the results back into the objects.  Compiled and run in this repository
//     against the data-holder stand-ins of oracle/refshim (tests/test_shim_objects_gpu.py: every method equals the
//     reference's own ORBmatcher.cc on the same objects).
#ifndef ORBMATCHER_H
#define ORBMATCHER_H

#include <cstdint>
#include <map>
#include <utility>
#include <vector>

#ifdef ORBFE_NO_OPENCV
#include "cv_standin.h"
#else
#include <opencv2/core/core.hpp>
#endif

#include "../orbfe.h"

#ifdef ORBFE_WITH_ORBSLAM3
#include <set>
#ifndef ORBFE_ORBSLAM3_TYPES_PROVIDED     // tests force-include stand-ins for these three (they pull in Eigen, g2o, boost)
#include "Frame.h"
#include "KeyFrame.h"
#include "MapPoint.h"
#endif
#endif

namespace ORB_SLAM3 {

namespace orbfe_shim {

// A DBoW2::FeatureVector (std::map<unsigned, std::vector<unsigned>>) flattened to the ABI's CSR.
struct FeatureVectorCsr {
    std::vector<uint32_t> nodeIds;
    std::vector<int32_t> offsets, indices;

    template <class Map>   // any std::map<NodeId, std::vector<unsigned>>-like container
    explicit FeatureVectorCsr(const Map& fv) {
        offsets.push_back(0);
        for (const auto& kv : fv) {
            nodeIds.push_back((uint32_t)kv.first);
            for (unsigned i : kv.second) indices.push_back((int32_t)i);
            offsets.push_back((int32_t)indices.size());
        }
    }
    orbfe_feature_vector view() const {
        orbfe_feature_vector v;
        v.nnodes = (int)nodeIds.size();
        v.node_ids = nodeIds.data();
        v.offsets = offsets.data();
        v.indices = indices.data();
        return v;
    }
};

// One side of a BoW-guided search, as arrays.
struct Side {
    const uint8_t* desc = nullptr;            // n x 32, continuous (cv::Mat::data of mDescriptors)
    int n = 0;
    const FeatureVectorCsr* fv = nullptr;
    const uint8_t* mask = nullptr;
    const cv::KeyPoint* kps = nullptr;        // mvKeysUn
    orbfe_bow_side view() const {
        orbfe_bow_side s;
        s.desc = desc; s.n = n; s.fv = fv->view(); s.mask = mask;
        s.kps = reinterpret_cast<const orbfe_keypoint*>(kps);
        return s;
    }
};

}  // namespace orbfe_shim

class ORBmatcher {
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true, int device = 0) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {
        status_ = orbfe_matcher_create(nnratio, checkOri ? 1 : 0, device, &m_);
        if (status_ != ORBFE_OK) m_ = nullptr;
    }
    ~ORBmatcher() { orbfe_matcher_destroy(m_); }
    ORBmatcher(const ORBmatcher&) = delete;
    ORBmatcher& operator=(const ORBmatcher&) = delete;

    // ORBmatcher.cc:2561-2577
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b) { return orbfe_descriptor_distance(a.ptr(), b.ptr()); }

    // ---- array layer ---------------------------------------------------------------------------------
    // SearchByBoW(KeyFrame*, Frame&, ...) (ORBmatcher.cc:269-471): kfOfFrameFeature[j] = keyframe feature
    // matched to frame feature j or -1.  Returns nmatches (0 on error, see lastStatus()).
    int SearchByBoWArrays(const orbfe_shim::Side& kf, const orbfe_shim::Side& frame, int frameNleft,
                          std::vector<int32_t>& kfOfFrameFeature) {
        kfOfFrameFeature.assign(frame.n, -1);
        if (!m_) return 0;
        orbfe_bow_side A = kf.view(), B = frame.view();
        B.mask = nullptr;
        int32_t* out = kfOfFrameFeature.data();
        int n = 0;
        status_ = orbfe_search_bow(m_, 0, &A, &B, 1, 0, frameNleft, nullptr, &out, &n);
        return status_ < 0 ? 0 : n;
    }
    // SearchByBoW(KeyFrame*, KeyFrame*, ...) (ORBmatcher.cc:823-963): kf2OfKf1Feature[i] = pKF2 feature or -1.
    int SearchByBoWArrays(const orbfe_shim::Side& kf1, const orbfe_shim::Side& kf2, std::vector<int32_t>& kf2OfKf1Feature) {
        kf2OfKf1Feature.assign(kf1.n, -1);
        if (!m_) return 0;
        orbfe_bow_side A = kf1.view(), B = kf2.view();
        int32_t* out = kf2OfKf1Feature.data();
        int n = 0;
        status_ = orbfe_search_bow(m_, 1, &A, &B, 1, 0, -1, &out, nullptr, &n);
        return status_ < 0 ? 0 : n;
    }
    // SearchForTriangulation_ (ORBmatcher.cc:1208-1432), pinhole: masks = "already has a MapPoint".
    int SearchForTriangulationArrays(const orbfe_shim::Side& kf1, const float* uRight1, const orbfe_shim::Side& kf2,
                                     const float* uRight2, const orbfe_epipolar& geo, bool bOnlyStereo, bool bCoarse,
                                     std::vector<std::pair<size_t, size_t>>& vMatchedPairs) {
        vMatchedPairs.clear();
        if (!m_) return 0;
        std::vector<int32_t> m12(kf1.n > 0 ? kf1.n : 1, -1);
        orbfe_bow_side A = kf1.view(), B = kf2.view();
        const int n = orbfe_search_triangulation(m_, &A, uRight1, &B, uRight2, &geo, bOnlyStereo, bCoarse, m12.data());
        status_ = n < 0 ? n : ORBFE_OK;
        if (n < 0) return 0;
        vMatchedPairs.reserve(n);
        for (int i = 0; i < kf1.n; ++i)
            if (m12[i] >= 0) vMatchedPairs.push_back(std::make_pair((size_t)i, (size_t)m12[i]));
        return n;
    }
    // Frame::ComputeStereoMatches (Frame.cc:794-964) on the device-resident results of two extractor handles.
    int ComputeStereoMatches(orbfe_handle* left, orbfe_handle* right, float mbf, float mb, std::vector<float>& mvuRight,
                             std::vector<float>& mvDepth, int N) {
        mvuRight.assign(N, -1.0f);
        mvDepth.assign(N, -1.0f);
        if (!m_ || N == 0) return 0;
        const int n = orbfe_stereo_match(m_, left, 0, right, 0, mbf, mb, mvuRight.data(), mvDepth.data(), N);
        status_ = n < 0 ? n : ORBFE_OK;
        return n < 0 ? 0 : n;
    }
    // Frame::ComputeStereoFishEyeMatches up to the ratio test (Frame.cc:1134-1141).
    int MatchLapping(const cv::Mat& descLeft, const cv::Mat& descRight, std::vector<int32_t>& match) {
        match.assign(descLeft.rows, -1);
        if (!m_ || descLeft.rows == 0) return 0;
        const int n = orbfe_match_lapping(m_, descLeft.ptr(), descLeft.rows, descRight.ptr(), descRight.rows, 0.7, match.data());
        status_ = n < 0 ? n : ORBFE_OK;
        return n < 0 ? 0 : n;
    }
    // The window-search loop of SearchByProjection (ORBmatcher.cc:44-236, :2173-2389, :2391-2510); the caller keeps
    // the reference's projection lines and emits one orbfe_proj_query per GetFeaturesInArea call.
    // assign[i]: index of the window whose MapPoint lands in mvpMapPoints[i], -1 untouched, -2 set to NULL.
    // choice (optional): the feature each window picked, -1 = none — what the KeyFrame-side callers consume (Fuse
    // :1629-1945 and SearchBySim3 :1947-2171 use windows that never claim, obs = 0; SearchByProjection(KeyFrame*, Scw,
    // ...) :473-700 uses claiming windows with blocked = vpMatched non-NULL).
    int SearchByProjectionWindows(const orbfe_grid_frame& F, const std::vector<orbfe_proj_query>& windows, int mode, int th,
                                  bool linkStereo, bool rotHist, std::vector<int32_t>& assign, std::vector<int32_t>* choice = nullptr) {
        assign.assign(F.n > 0 ? F.n : 1, -1);
        if (choice) choice->assign(windows.empty() ? 1 : windows.size(), -1);
        if (!m_) return 0;
        const int n = orbfe_search_by_projection(m_, &F, windows.data(), (int)windows.size(), mode, th, linkStereo, rotHist,
                                                 assign.data(), choice ? choice->data() : nullptr);
        assign.resize(F.n > 0 ? F.n : 0);
        if (choice) choice->resize(windows.size());
        status_ = n < 0 ? n : ORBFE_OK;
        return n < 0 ? 0 : n;
    }
    // ORBmatcher::SearchForInitialization (ORBmatcher.cc:706-821) on arrays: F1 = (mvKeysUn1, mDescriptors1), F2 as a grid
    // frame (kps = mvKeysUn, nleft = -1, mnMinX / mnMinY / grid inverses).  vbPrevMatched and vnMatches12 as in the reference.
    int SearchForInitializationArrays(const std::vector<cv::KeyPoint>& mvKeysUn1, const cv::Mat& mDescriptors1, const orbfe_grid_frame& F2,
                                      std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12, int windowSize = 10) {
        static_assert(sizeof(cv::Point2f) == 8 && sizeof(int) == sizeof(int32_t), "Point2f / int layout");
        vnMatches12.assign(mvKeysUn1.size(), -1);
        if (!m_ || mvKeysUn1.empty()) return 0;
        if (vbPrevMatched.size() != mvKeysUn1.size()) { status_ = ORBFE_ERR_BAD_ARGUMENT; return 0; }
        const int n = orbfe_search_for_initialization(m_, reinterpret_cast<const orbfe_keypoint*>(mvKeysUn1.data()), mDescriptors1.ptr(),
                                                      (int)mvKeysUn1.size(), &F2, reinterpret_cast<float*>(vbPrevMatched.data()),
                                                      windowSize, vnMatches12.data());
        status_ = n < 0 ? n : ORBFE_OK;
        return n < 0 ? 0 : n;
    }
    // Frame::UndistortKeyPoints (Frame.cc:730-763).  K4 = fx, fy, cx, cy; dist = mDistCoef (4 or 5 floats).
    bool UndistortKeyPoints(const std::vector<cv::KeyPoint>& mvKeys, const float* K4, const float* dist, int ndist,
                            std::vector<cv::KeyPoint>& mvKeysUn) {
        mvKeysUn.resize(mvKeys.size());
        if (!m_) return false;
        status_ = orbfe_undistort_keypoints(m_, reinterpret_cast<const orbfe_keypoint*>(mvKeys.data()), (int)mvKeys.size(), K4, dist,
                                            ndist, reinterpret_cast<orbfe_keypoint*>(mvKeysUn.data()));
        return status_ == ORBFE_OK;
    }
    // MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:349-423) for many map points in one launch: rows
    // offsets[p]..offsets[p+1] of `observed` are point p's descriptors; best[p] = the row to clone into mDescriptor.
    bool ComputeDistinctiveDescriptors(const cv::Mat& observed, const std::vector<int32_t>& offsets, std::vector<int32_t>& best) {
        const int np = offsets.empty() ? 0 : (int)offsets.size() - 1;
        best.assign(np > 0 ? np : 0, -1);
        if (!m_ || np <= 0) return m_ != nullptr;
        status_ = orbfe_distinctive_descriptors(m_, observed.rows ? observed.ptr() : nullptr, offsets.data(), np, best.data());
        return status_ == ORBFE_OK;
    }
    // Frame::ComputeStereoFishEyeMatches, whole body (Frame.cc:1116-1156).  cam1 / cam2: mvParameters of the two
    // KannalaBrandt8 cameras (8 floats), Rlr (row-major 3x3) / tlr: Frame::mRlr, mtlr.  Returns nMatches.
    int ComputeStereoFishEyeMatches(const std::vector<cv::KeyPoint>& mvKeys, const cv::Mat& mDescriptors, int monoLeft,
                                    const std::vector<cv::KeyPoint>& mvKeysRight, const cv::Mat& mDescriptorsRight, int monoRight,
                                    const float* cam1, const float* cam2, const float* Rlr, const float* tlr,
                                    const std::vector<float>& mvLevelSigma2, std::vector<int>& mvLeftToRightMatch,
                                    std::vector<int>& mvRightToLeftMatch, std::vector<float>& mvDepth, std::vector<float>& p3d) {
        const int nL = (int)mvKeys.size(), nR = (int)mvKeysRight.size();
        mvLeftToRightMatch.assign(nL, -1);
        mvRightToLeftMatch.assign(nR, -1);
        mvDepth.assign(nL, -1.0f);
        p3d.assign(3 * (size_t)nL, 0.0f);
        if (!m_ || nL == 0) return 0;
        static_assert(sizeof(int) == sizeof(int32_t), "int32 match vectors");
        const int n = orbfe_fisheye_stereo_match(m_, reinterpret_cast<const orbfe_keypoint*>(mvKeys.data()), mDescriptors.ptr(), nL,
                                                 monoLeft, reinterpret_cast<const orbfe_keypoint*>(mvKeysRight.data()),
                                                 nR ? mDescriptorsRight.ptr() : nullptr, nR, monoRight, cam1, cam2, Rlr, tlr,
                                                 mvLevelSigma2.data(), (int)mvLevelSigma2.size(), mvLeftToRightMatch.data(),
                                                 mvRightToLeftMatch.data(), mvDepth.data(), p3d.data());
        status_ = n < 0 ? n : ORBFE_OK;
        return n < 0 ? 0 : n;
    }
    // SearchForTriangulation_ with KannalaBrandt8 cameras (ORBmatcher.cc:1208-1432 + KannalaBrandt8.cpp:239-242).
    // points (optional): triangulated point of each pair, camera-1 frame (the matchAndtriangulate overload).
    int SearchForTriangulationKB8Arrays(const orbfe_shim::Side& kf1, const orbfe_shim::Side& kf2, const orbfe_kb8_epipolar& geo,
                                        bool bOnlyStereo, bool bCoarse, std::vector<std::pair<size_t, size_t>>& vMatchedPairs,
                                        std::vector<float>* points = nullptr) {
        vMatchedPairs.clear();
        if (points) points->clear();
        if (!m_) return 0;
        std::vector<int32_t> m12(kf1.n > 0 ? kf1.n : 1, -1);
        std::vector<float> p(points ? 3 * (size_t)(kf1.n > 0 ? kf1.n : 1) : 0);
        orbfe_bow_side A = kf1.view(), B = kf2.view();
        const int n = orbfe_search_triangulation_kb8(m_, &A, &B, &geo, bOnlyStereo, bCoarse, m12.data(), points ? p.data() : nullptr);
        status_ = n < 0 ? n : ORBFE_OK;
        if (n < 0) return 0;
        vMatchedPairs.reserve(n);
        for (int i = 0; i < kf1.n; ++i)
            if (m12[i] >= 0) {
                vMatchedPairs.push_back(std::make_pair((size_t)i, (size_t)m12[i]));
                if (points) points->insert(points->end(), p.begin() + 3 * (size_t)i, p.begin() + 3 * (size_t)i + 3);
            }
        return n;
    }

#ifdef ORBFE_WITH_ORBSLAM3
    // ---- object layer: the reference's signatures (include/ORBmatcher.h:44-91), defined in ORBmatcherObjects.h.  Each one
    // flattens the objects it is given into the arrays of the C ABI, runs the search on the GPU and writes the results
    // back into the objects exactly where the reference does.
    int SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th = 3, const bool bFarPoints = false,
                           const float thFarPoints = 50.0f);                                                         // :46
    int SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono);          // :50
    int SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound, const float th,
                           const int ORBdist);                                                                       // :54
    int SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, std::vector<MapPoint*>& vpMatched,
                           int th, float ratioHamming = 1.0);                                                        // :58
    int SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, const std::vector<KeyFrame*>& vpPointsKFs,
                           std::vector<MapPoint*>& vpMatched, std::vector<KeyFrame*>& vpMatchedKF, int th, float ratioHamming = 1.0);   // :62
    int SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches);                             // :67
    int SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12);                            // :68
    int SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12,
                                int windowSize = 10);                                                                // :71
    int SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12, std::vector<std::pair<size_t, size_t> >& vMatchedPairs,
                               const bool bOnlyStereo, const bool bCoarse = false);                                  // :74
    int SearchForTriangulation_(KeyFrame* pKF1, KeyFrame* pKF2, cv::Matx33f F12, std::vector<std::pair<size_t, size_t> >& vMatchedPairs,
                                const bool bOnlyStereo, const bool bCoarse = false);                                 // :77
    int SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12, std::vector<std::pair<size_t, size_t> >& vMatchedPairs,
                               const bool bOnlyStereo, std::vector<cv::Mat>& vMatchedPoints);                        // :80
    int SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const float& s12, const cv::Mat& R12,
                     const cv::Mat& t12, const float th);                                                            // :85
    int Fuse(KeyFrame* pKF, const std::vector<MapPoint*>& vpMapPoints, const float th = 3.0, const bool bRight = false);   // :88
    int Fuse(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, float th, std::vector<MapPoint*>& vpReplacePoint);   // :91

protected:
    float RadiusByViewingCos(const float& viewCos) { return viewCos > 0.998 ? 2.5f : 4.0f; }                        // ORBmatcher.cc:238-244
    void ComputeThreeMaxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3);                  // :2515-2556
    int triangulationSearch(KeyFrame* pKF1, KeyFrame* pKF2, const cv::Matx33f& F12, std::vector<std::pair<size_t, size_t> >& vMatchedPairs,
                            bool bOnlyStereo, bool bCoarse, std::vector<cv::Mat>* vMatchedPoints);
    int sim3Projection(KeyFrame* pKF, const cv::Mat& Scw, const std::vector<MapPoint*>& vpPoints, const std::vector<KeyFrame*>* vpPointsKFs,
                       std::vector<MapPoint*>& vpMatched, std::vector<KeyFrame*>* vpMatchedKF, int th, float ratioHamming);

public:
#endif

    static constexpr int TH_LOW = ORBFE_TH_LOW;       // constexpr: implicitly inline, so odr-uses need no definition elsewhere
    static constexpr int TH_HIGH = ORBFE_TH_HIGH;
    static constexpr int HISTO_LENGTH = ORBFE_HISTO_LENGTH;

    orbfe_matcher* handle() { return m_; }
    int lastStatus() const { return status_; }

protected:
    float mfNNratio;
    bool mbCheckOrientation;

private:
    orbfe_matcher* m_ = nullptr;
    int status_ = ORBFE_OK;
};

}  // namespace ORB_SLAM3

#ifdef ORBFE_WITH_ORBSLAM3
#include "ORBmatcherObjects.h"
#endif

#endif  // ORBMATCHER_H
