// orbslam3_standin.h — data-holder stand-ins for ORB_SLAM3::Frame, KeyFrame, MapPoint and GeometricCamera, so that the
// reference's OWN src/ORBmatcher.cc (with its own include/ORBmatcher.h) compiles here (TEST INFRASTRUCTURE).
//
// The real Frame.h / KeyFrame.h / MapPoint.h pull in Eigen, Sophus-free g2o types, boost and the whole map; the build
// pre-defines their include guards (-DFRAME_H -DKEYFRAME_H -DMAPPOINT_H) and force-includes this file instead.  Only the
// members ORBmatcher.cc touches exist, with the reference's names and types.  What is real when oracle/_ref runs the
// matcher: every line of ORBmatcher.cc (the Search* / Fuse loops, DescriptorDistance, ComputeThreeMaxima, rotation
// histograms, ratio tests).  What is restated here: Frame / KeyFrame::GetFeaturesInArea and the grid they walk
// (src/Frame.cc:380-411, 640-719; src/KeyFrame.cc:737-781), MapPoint::PredictScale (src/MapPoint.cc:534-566), the pinhole
// camera (src/CameraModels/Pinhole.cpp:34-51, 159-181) — each cited where it is defined.
#pragma once
#include <cmath>
#include <map>
#include <mutex>
#include <set>
#include <tuple>
#include <vector>

#include <opencv2/opencv.hpp>

#include "Thirdparty/DBoW2/DBoW2/BowVector.h"
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"

using namespace std;      // the real headers leak it too (e.g. include/Converter.h / ImuTypes.h); ORBmatcher.h says plain `pair`, `vector`

#define FRAME_GRID_ROWS 48
#define FRAME_GRID_COLS 64

namespace ORB_SLAM3 {

class Frame;
class KeyFrame;
class MapPoint;

#ifdef ORBSLAM3_REAL_CAMERAS
}  // namespace ORB_SLAM3
// liborbslam3_orbmatcher_cam.so: the reference's OWN include/CameraModels/{GeometricCamera,Pinhole,KannalaBrandt8}.h and their .cpp
// files instead of the stand-in camera below (TwoViewReconstruction.h stays out through its include guard: it belongs to monocular
// initialisation, not to the matcher).
namespace ORB_SLAM3 {
class TwoViewReconstruction {
public:
    TwoViewReconstruction(cv::Mat&, float sigma = 1.0, int iterations = 200) { (void)sigma; (void)iterations; }
    bool Reconstruct(const std::vector<cv::KeyPoint>&, const std::vector<cv::KeyPoint>&, const std::vector<int>&, cv::Mat&, cv::Mat&,
                     std::vector<cv::Point3f>&, std::vector<bool>&) { return false; }
};
}  // namespace ORB_SLAM3
#include "CameraModels/GeometricCamera.h"
#include "CameraModels/Pinhole.h"
#include "CameraModels/KannalaBrandt8.h"
namespace ORB_SLAM3 {
#else
class GeometricCamera {
public:
    virtual ~GeometricCamera() {}
    // include/CameraModels/GeometricCamera.h:86-93: camera kind and raw parameters (read by the orbfe shim to pick the pinhole
    // or the Kannala-Brandt search); the stand-in only has pinhole cameras
    const unsigned int CAM_PINHOLE = 0;
    const unsigned int CAM_FISHEYE = 1;
    virtual unsigned int GetType() { return CAM_PINHOLE; }
    virtual float getParameter(const int i) { return 0.f; }
    virtual cv::Point2f project(const cv::Point3f& p3D) = 0;
    virtual cv::Point2f project(const cv::Matx31f& m3D) = 0;
    virtual cv::Point2f project(const cv::Mat& m3D) = 0;
    virtual bool epipolarConstrain(GeometricCamera* otherCamera, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat& R12,
                                   const cv::Mat& t12, const float sigmaLevel, const float unc) = 0;
    virtual bool epipolarConstrain_(GeometricCamera* otherCamera, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Matx33f& R12,
                                    const cv::Matx31f& t12, const float sigmaLevel, const float unc) = 0;
    virtual bool matchAndtriangulate(const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, GeometricCamera* pOther, cv::Mat& Tcw1, cv::Mat& Tcw2,
                                     const float sigmaLevel1, const float sigmaLevel2, cv::Mat& x3Dtriangulated) { return false; }
};

// src/CameraModels/Pinhole.cpp: project :34-51; epipolarConstrain_ :159-181 from the line "Epipolar line in second image" on.
// The reference first builds F12 = K1^-T [t12]x R12 K2^-1 with cv::Matx::inv(); the C ABI (and the oracle) receive F12 from
// the caller, so this camera uses the F12 the test hands it and ignores R12 / t12.
class Pinhole : public GeometricCamera {
public:
    float fx, fy, cx, cy;
    cv::Matx33f F12;
    bool fixedEpipole = false;            // test hook: SearchForTriangulation_ projects camera 1's centre with this overload only
    cv::Point2f epipole;                  // (ORBmatcher.cc:1213-1218); the C ABI receives that point from the caller
    Pinhole(float fx_, float fy_, float cx_, float cy_) : fx(fx_), fy(fy_), cx(cx_), cy(cy_) {}
    cv::Point2f project(const cv::Point3f& p) override { return cv::Point2f(fx * p.x / p.z + cx, fy * p.y / p.z + cy); }
    cv::Point2f project(const cv::Matx31f& m) override { return fixedEpipole ? epipole : project(cv::Point3f(m(0), m(1), m(2))); }
    cv::Point2f project(const cv::Mat& m) override {
        return project(cv::Point3f(m.at<float>(0), m.at<float>(1), m.at<float>(2)));
    }
    bool epipolarConstrain_(GeometricCamera*, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Matx33f&, const cv::Matx31f&,
                            const float, const float unc) override {
        const float a = kp1.pt.x * F12(0, 0) + kp1.pt.y * F12(1, 0) + F12(2, 0);
        const float b = kp1.pt.x * F12(0, 1) + kp1.pt.y * F12(1, 1) + F12(2, 1);
        const float c = kp1.pt.x * F12(0, 2) + kp1.pt.y * F12(1, 2) + F12(2, 2);
        const float num = a * kp2.pt.x + b * kp2.pt.y + c;
        const float den = a * a + b * b;
        if (den == 0) return false;
        const float dsqr = num * num / den;
        return dsqr < 3.84 * unc;
    }
    bool epipolarConstrain(GeometricCamera* o, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat&, const cv::Mat&,
                           const float s, const float unc) override {
        return epipolarConstrain_(o, kp1, kp2, cv::Matx33f(), cv::Matx31f(), s, unc);
    }
};

#endif  // ORBSLAM3_REAL_CAMERAS

class MapPoint {
public:
    // tracking fields written by Frame::isInFrustum (src/Frame.cc:455-638), read by SearchByProjection(Frame&, vpMapPoints)
    float mTrackProjX = 0, mTrackProjY = 0, mTrackDepth = 0, mTrackDepthR = 0, mTrackProjXR = 0, mTrackProjYR = 0;
    bool mbTrackInView = false, mbTrackInViewR = false;
    int mnTrackScaleLevel = 0, mnTrackScaleLevelR = 0;
    float mTrackViewCos = 0, mTrackViewCosR = 0;
    // state the stand-in holds directly
    cv::Mat mDescriptor, mWorldPos, mNormalVector;
    float mfMinDistance = 0, mfMaxDistance = 0;
    int nObs = 1;
    bool mbBad = false;
    std::map<KeyFrame*, std::tuple<int, int>> mObservations;
    MapPoint* mpReplaced = nullptr;

    cv::Mat GetDescriptor() { return mDescriptor.clone(); }
    cv::Mat GetWorldPos() { return mWorldPos.clone(); }
    cv::Mat GetNormal() { return mNormalVector.clone(); }
    bool isBad() { return mbBad; }
    int Observations() { return nObs; }
    float GetMinDistanceInvariance() { return 0.8f * mfMinDistance; }       // src/MapPoint.cc:522-526
    float GetMaxDistanceInvariance() { return 1.2f * mfMaxDistance; }       // :528-532
    inline int PredictScale(const float& currentDist, KeyFrame* pKF);      // :534-549
    inline int PredictScale(const float& currentDist, Frame* pF);          // :551-566
    bool IsInKeyFrame(KeyFrame* pKF) { return mObservations.count(pKF) != 0; }
    std::tuple<int, int> GetIndexInKeyFrame(KeyFrame* pKF) {
        auto it = mObservations.find(pKF);
        return it != mObservations.end() ? it->second : std::tuple<int, int>(-1, -1);
    }
    // Test tap: every (MapPoint, feature index) that Fuse joins — AddObservation from the matcher, and both parties of a Replace.
    static std::vector<std::pair<MapPoint*, int>>& fuseLog() { static std::vector<std::pair<MapPoint*, int>> v; return v; }
    void AddObservation(KeyFrame* pKF, int idx) { fuseLog().push_back(std::make_pair(this, idx)); observe(pKF, idx); }
    void observe(KeyFrame* pKF, int idx) { mObservations[pKF] = std::tuple<int, int>(idx, -1); ++nObs; }
    inline void Replace(MapPoint* pMP);                                    // src/MapPoint.cc:226-283
};

// The 64 x 48 feature grid of Frame::AssignFeaturesToGrid / PosInGrid (src/Frame.cc:380-411, 708-719), shared by Frame and
// KeyFrame (KeyFrame copies the Frame's grid, src/KeyFrame.cc:60-82).
struct FeatureGrid {
    float mnMinX = 0, mnMinY = 0, mnMaxX = 0, mnMaxY = 0, mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;
    std::vector<std::size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS], mGridRight[FRAME_GRID_COLS][FRAME_GRID_ROWS];
    void setBounds(float x0, float y0, float x1, float y1) {
        mnMinX = x0; mnMinY = y0; mnMaxX = x1; mnMaxY = y1;
        mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(mnMaxX - mnMinX);     // Frame.cc:158-159
        mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(mnMaxY - mnMinY);
    }
    bool PosInGrid(const cv::KeyPoint& kp, int& posX, int& posY) const {
        posX = round((kp.pt.x - mnMinX) * mfGridElementWidthInv);
        posY = round((kp.pt.y - mnMinY) * mfGridElementHeightInv);
        return !(posX < 0 || posX >= FRAME_GRID_COLS || posY < 0 || posY >= FRAME_GRID_ROWS);
    }
};

class Frame : public FeatureGrid {
public:
    int N = 0, Nleft = -1, Nright = -1;
    std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn;
    cv::Mat mDescriptors;
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    std::vector<float> mvuRight, mvScaleFactors, mvInvLevelSigma2, mvLevelSigma2;
    std::vector<int> mvLeftToRightMatch, mvRightToLeftMatch;
    DBoW2::BowVector mBowVec;
    DBoW2::FeatureVector mFeatVec;
    cv::Mat mTcw, mTrl;
    GeometricCamera *mpCamera = nullptr, *mpCamera2 = nullptr;
    float mbf = 0, mb = 0, mfLogScaleFactor = 0;
    int mnScaleLevels = 8;

    void AssignFeaturesToGrid() {                                        // src/Frame.cc:380-411
        for (int i = 0; i < N; i++) {
            const cv::KeyPoint& kp = (Nleft == -1) ? mvKeysUn[i] : (i < Nleft) ? mvKeys[i] : mvKeysRight[i - Nleft];
            int nGridPosX, nGridPosY;
            if (PosInGrid(kp, nGridPosX, nGridPosY)) {
                if (Nleft == -1 || i < Nleft) mGrid[nGridPosX][nGridPosY].push_back(i);
                else mGridRight[nGridPosX][nGridPosY].push_back(i - Nleft);
            }
        }
    }
    // src/Frame.cc:640-706
    std::vector<std::size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel = -1, const int maxLevel = -1,
                                               const bool bRight = false) const {
        std::vector<std::size_t> vIndices;
        vIndices.reserve(N);
        float factorX = r, factorY = r;
        const int nMinCellX = std::max(0, (int)floor((x - mnMinX - factorX) * mfGridElementWidthInv));
        if (nMinCellX >= FRAME_GRID_COLS) return vIndices;
        const int nMaxCellX = std::min((int)FRAME_GRID_COLS - 1, (int)ceil((x - mnMinX + factorX) * mfGridElementWidthInv));
        if (nMaxCellX < 0) return vIndices;
        const int nMinCellY = std::max(0, (int)floor((y - mnMinY - factorY) * mfGridElementHeightInv));
        if (nMinCellY >= FRAME_GRID_ROWS) return vIndices;
        const int nMaxCellY = std::min((int)FRAME_GRID_ROWS - 1, (int)ceil((y - mnMinY + factorY) * mfGridElementHeightInv));
        if (nMaxCellY < 0) return vIndices;
        const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
        for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
            for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
                const std::vector<std::size_t>& vCell = (!bRight) ? mGrid[ix][iy] : mGridRight[ix][iy];
                for (std::size_t j = 0, jend = vCell.size(); j < jend; j++) {
                    const cv::KeyPoint& kpUn = (Nleft == -1) ? mvKeysUn[vCell[j]] : (!bRight) ? mvKeys[vCell[j]] : mvKeysRight[vCell[j]];
                    if (bCheckLevels) {
                        if (kpUn.octave < minLevel) continue;
                        if (maxLevel >= 0)
                            if (kpUn.octave > maxLevel) continue;
                    }
                    const float distx = kpUn.pt.x - x, disty = kpUn.pt.y - y;
                    if (fabs(distx) < factorX && fabs(disty) < factorY) vIndices.push_back(vCell[j]);
                }
            }
        return vIndices;
    }
};

class KeyFrame : public FeatureGrid {
public:
    int N = 0, NLeft = -1, NRight = -1;
    std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn;
    cv::Mat mDescriptors;
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<float> mvuRight, mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
    DBoW2::BowVector mBowVec;
    DBoW2::FeatureVector mFeatVec;
    GeometricCamera *mpCamera = nullptr, *mpCamera2 = nullptr;
    float fx = 0, fy = 0, cx = 0, cy = 0, mbf = 0, mfLogScaleFactor = 0;
    int mnScaleLevels = 8;
    cv::Mat Tcw, Tcr;                    // pose of the left camera; right camera: Tcr * Tcw (unused by the pinned subset)

    std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
    MapPoint* GetMapPoint(const std::size_t& idx) { return mvpMapPoints[idx]; }
    std::set<MapPoint*> mspExtra;          // test hook: MapPoints the keyframe observes at features outside the test's arrays
    std::set<MapPoint*> GetMapPoints() {
        std::set<MapPoint*> s(mspExtra);
        for (MapPoint* p : mvpMapPoints) if (p && !p->isBad()) s.insert(p);
        return s;
    }
    void AddMapPoint(MapPoint* pMP, const std::size_t& idx) { mvpMapPoints[idx] = pMP; }
    bool IsInImage(const float& x, const float& y) const { return (x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY); }   // KeyFrame.cc:783-786
    // src/KeyFrame.cc:737-781
    std::vector<std::size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const bool bRight = false) const {
        std::vector<std::size_t> vIndices;
        vIndices.reserve(N);
        float factorX = r, factorY = r;
        const int nMinCellX = std::max(0, (int)floor((x - mnMinX - factorX) * mfGridElementWidthInv));
        if (nMinCellX >= FRAME_GRID_COLS) return vIndices;
        const int nMaxCellX = std::min((int)FRAME_GRID_COLS - 1, (int)ceil((x - mnMinX + factorX) * mfGridElementWidthInv));
        if (nMaxCellX < 0) return vIndices;
        const int nMinCellY = std::max(0, (int)floor((y - mnMinY - factorY) * mfGridElementHeightInv));
        if (nMinCellY >= FRAME_GRID_ROWS) return vIndices;
        const int nMaxCellY = std::min((int)FRAME_GRID_ROWS - 1, (int)ceil((y - mnMinY + factorY) * mfGridElementHeightInv));
        if (nMaxCellY < 0) return vIndices;
        for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
            for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
                const std::vector<std::size_t>& vCell = (!bRight) ? mGrid[ix][iy] : mGridRight[ix][iy];
                for (std::size_t j = 0, jend = vCell.size(); j < jend; j++) {
                    const cv::KeyPoint& kpUn = (NLeft == -1) ? mvKeysUn[vCell[j]] : (!bRight) ? mvKeys[vCell[j]] : mvKeysRight[vCell[j]];
                    const float distx = kpUn.pt.x - x, disty = kpUn.pt.y - y;
                    if (fabs(distx) < r && fabs(disty) < r) vIndices.push_back(vCell[j]);
                }
            }
        return vIndices;
    }
    void AssignFeaturesToGrid() {
        for (int i = 0; i < N; i++) {
            const cv::KeyPoint& kp = (NLeft == -1) ? mvKeysUn[i] : (i < NLeft) ? mvKeys[i] : mvKeysRight[i - NLeft];
            int px, py;
            if (PosInGrid(kp, px, py)) {
                if (NLeft == -1 || i < NLeft) mGrid[px][py].push_back(i);
                else mGridRight[px][py].push_back(i - NLeft);
            }
        }
    }
    // poses: src/KeyFrame.cc:126-170, :1101-1170.  Float blocks of Tcw; camera centre Ow = -Rcw^T tcw.
    cv::Mat GetPose() { return Tcw.clone(); }
    cv::Mat GetPoseInverse() {                                       // src/KeyFrame.cc:160-164 (Twc), from Rwc = Rcw^T and Ow
        cv::Mat Twc = cv::Mat::zeros(4, 4, CV_32F);
        const cv::Mat Rwc = GetRotation().t(), Ow = GetCameraCenter();
        for (int i = 0; i < 3; ++i) { for (int j = 0; j < 3; ++j) Twc.at<float>(i, j) = Rwc.at<float>(i, j); Twc.at<float>(i, 3) = Ow.at<float>(i); }
        Twc.at<float>(3, 3) = 1.f;
        return Twc;
    }
    cv::Mat GetRotation() { return Tcw.rowRange(0, 3).colRange(0, 3).clone(); }
    cv::Mat GetTranslation() { return Tcw.rowRange(0, 3).col(3).clone(); }
    cv::Mat GetCameraCenter() { return -GetRotation().t() * GetTranslation(); }
    cv::Mat GetRightPose() { return Tcr * Tcw; }
    cv::Mat GetRightRotation() { return GetRightPose().rowRange(0, 3).colRange(0, 3).clone(); }
    cv::Mat GetRightTranslation() { return GetRightPose().rowRange(0, 3).col(3).clone(); }
    cv::Mat GetRightCameraCenter() { return -GetRightRotation().t() * GetRightTranslation(); }
    static cv::Matx33f m33(const cv::Mat& m) { cv::Matx33f r; for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) r(i, j) = m.at<float>(i, j); return r; }
    static cv::Matx31f m31(const cv::Mat& m) { cv::Matx31f r; for (int i = 0; i < 3; ++i) r(i) = m.at<float>(i); return r; }
    cv::Matx33f GetRotation_() { return m33(GetRotation()); }
    cv::Matx31f GetTranslation_() { return m31(GetTranslation()); }
    cv::Matx31f GetCameraCenter_() { return m31(GetCameraCenter()); }
    cv::Matx33f GetRightRotation_() { return m33(GetRightRotation()); }
    cv::Matx31f GetRightTranslation_() { return m31(GetRightTranslation()); }
    cv::Matx31f GetRightCameraCenter_() { return m31(GetRightCameraCenter()); }
};

// src/MapPoint.cc:226-283: this point goes bad, its observations move to pMP (or are erased where pMP is already seen).
inline void MapPoint::Replace(MapPoint* pMP) {
    if (pMP == this) return;
    std::map<KeyFrame*, std::tuple<int, int>> obs = mObservations;
    for (auto& o : pMP->mObservations)                                      // the feature where the two points met
        if (!obs.count(o.first)) { fuseLog().push_back(std::make_pair(this, std::get<0>(o.second))); fuseLog().push_back(std::make_pair(pMP, std::get<0>(o.second))); }
    mObservations.clear();
    mbBad = true;
    mpReplaced = pMP;
    for (auto& o : obs) {
        KeyFrame* pKF = o.first;
        const int idx = std::get<0>(o.second);
        fuseLog().push_back(std::make_pair(this, idx));
        fuseLog().push_back(std::make_pair(pMP, idx));
        if (!pMP->IsInKeyFrame(pKF)) { pKF->mvpMapPoints[idx] = pMP; pMP->observe(pKF, idx); }
        else pKF->mvpMapPoints[idx] = nullptr;
    }
}

inline int MapPoint::PredictScale(const float& currentDist, KeyFrame* pKF) {
    float ratio = mfMaxDistance / currentDist;
    int nScale = ceil(log(ratio) / pKF->mfLogScaleFactor);
    if (nScale < 0) nScale = 0;
    else if (nScale >= pKF->mnScaleLevels) nScale = pKF->mnScaleLevels - 1;
    return nScale;
}
inline int MapPoint::PredictScale(const float& currentDist, Frame* pF) {
    float ratio = mfMaxDistance / currentDist;
    int nScale = ceil(log(ratio) / pF->mfLogScaleFactor);
    if (nScale < 0) nScale = 0;
    else if (nScale >= pF->mnScaleLevels) nScale = pF->mnScaleLevels - 1;
    return nScale;
}

}  // namespace ORB_SLAM3
