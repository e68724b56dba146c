// TEST INFRASTRUCTURE (oracle/_ref).  Stand-ins that let the reference's own src/Frame.cc + include/Frame.h (and, with them,
// src/MapPoint.cc and the camera models) compile UNMODIFIED: the include guards of ImuTypes.h / Config.h / G2oTypes.h /
// KeyFrame.h / Map.h / Converter.h / TwoViewReconstruction.h are pre-defined on the command line and this header is
// force-included.  Only what Frame.cc / MapPoint.cc / the camera models name exists here; Eigen, g2o, Sophus, the IMU
// preintegration and the map stay out.
#pragma once
#include <opencv2/opencv.hpp>
#include <algorithm>
#include <chrono>
#include <climits>
#include <cmath>
#include <iostream>
#include <list>
#include <map>
#include <mutex>
#include <set>
#include <string>
#include <thread>
#include <tuple>
#include <vector>

using namespace std;   // as the reference's own headers do

namespace ORB_SLAM3 {

class MapPoint;
class GeometricCamera;
class Frame;
class KeyFrame;

namespace IMU {                  // ImuTypes.h: the three types Frame.h holds by value or by pointer
class Bias {
public:
    Bias() : bax(0), bay(0), baz(0), bwx(0), bwy(0), bwz(0) {}
    float bax, bay, baz, bwx, bwy, bwz;
};
class Calib {
public:
    Calib() {}
    cv::Mat Tcb, Tbc;
};
class Preintegrated {
public:
    void SetNewBias(const Bias&) {}
};
}  // namespace IMU

class ConstraintPoseImu {};      // G2oTypes.h: Frame only stores the pointer

class Converter {                // Converter.cc:24-32
public:
    static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& Descriptors) {
        std::vector<cv::Mat> vDesc;
        vDesc.reserve(Descriptors.rows);
        for (int j = 0; j < Descriptors.rows; j++) vDesc.push_back(Descriptors.row(j));
        return vDesc;
    }
};

class TwoViewReconstruction {    // TwoViewReconstruction.h: owned by the camera models, used only by monocular initialisation
public:
    TwoViewReconstruction(cv::Mat&, float sigma = 1.0, int iterations = 200) { (void)sigma; (void)iterations; }
    bool Reconstruct(const std::vector<cv::KeyPoint>&, const std::vector<cv::KeyPoint>&, const std::vector<int>&, cv::Mat&, cv::Mat&,
                     std::vector<cv::Point3f>&, std::vector<bool>&) { return false; }
};

class Map {                      // Map.h: what MapPoint.cc touches
public:
    std::mutex mMutexPointCreation;
    unsigned long mnId = 0;
    std::vector<MapPoint*> erased;
    unsigned long GetId() { return mnId; }
    void EraseMapPoint(MapPoint* p) { erased.push_back(p); }
    // what KeyFrame.cc touches (liborbslam3_objects.so)
    std::vector<KeyFrame*> erasedKeyFrames;
    bool isImuInitialized() { return false; }
    unsigned long GetInitKFid() { return 0; }
    void EraseKeyFrame(KeyFrame* p) { erasedKeyFrames.push_back(p); }
};

class KeyFrameDatabase {         // KeyFrameDatabase.h: KeyFrame::SetBadFlag erases itself from it
public:
    std::vector<KeyFrame*> erased;
    void erase(KeyFrame* p) { erased.push_back(p); }
};

#ifndef ORBSLAM3_REAL_KEYFRAME
class KeyFrame {                 // KeyFrame.h members read by MapPoint.cc (Frame.cc only passes the pointer around)
public:
    long unsigned int mnId = 0, mnFrameId = 0;
    int NLeft = -1;
    GeometricCamera* mpCamera2 = nullptr;
    std::vector<float> mvuRight;
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn, mvKeysRight;
    std::vector<float> mvScaleFactors;
    int mnScaleLevels = 0;
    float mfLogScaleFactor = 0;
    cv::Mat mDescriptors;
    cv::Mat Ow, OwRight;
    bool bad = false;
    std::vector<MapPoint*> mvpMapPoints;
    bool isBad() { return bad; }
    void EraseMapPointMatch(int idx) { mvpMapPoints[idx] = nullptr; }
    void ReplaceMapPointMatch(int idx, MapPoint* p) { mvpMapPoints[idx] = p; }
    cv::Mat GetCameraCenter() { return Ow.clone(); }
    cv::Mat GetRightCameraCenter() { return OwRight.clone(); }
};

#endif  // ORBSLAM3_REAL_KEYFRAME (liborbslam3_keyframe.so compiles the reference's own KeyFrame.h / KeyFrame.cc instead)

}  // namespace ORB_SLAM3
