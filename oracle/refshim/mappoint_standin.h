// TEST INFRASTRUCTURE (oracle/_ref).  Data-holder stand-ins for KeyFrame / Frame / Map so that the reference's own
// src/MapPoint.cc + include/MapPoint.h compile UNMODIFIED (their include guards KEYFRAME_H / FRAME_H / MAP_H are
// pre-defined on the command line and this header is force-included).  Only the members MapPoint.cc reads exist;
// the real classes (Eigen, g2o, DBoW2, boost archives, the whole map) stay out.
#pragma once
#include <opencv2/core/core.hpp>
#include <algorithm>
#include <climits>
#include <cmath>
#include <map>
#include <mutex>
#include <set>
#include <tuple>
#include <vector>

using namespace std;   // the reference's headers do the same (KeyFrame.h pulls it in through DBoW2)

namespace ORB_SLAM3 {

class MapPoint;
class GeometricCamera;

class Map {                      // Map.h: what MapPoint.cc touches (Map.cc:113-121 EraseMapPoint only drops the pointer from a set)
public:
    std::mutex mMutexPointCreation;
    unsigned long mnId = 0;
    std::vector<MapPoint*> erased;
    unsigned long GetId() { return mnId; }
    void EraseMapPoint(MapPoint* p) { erased.push_back(p); }
};

class KeyFrame {                 // KeyFrame.h members read at MapPoint.cc:40-41,161-186,188-222,236-258,268-320,349-423,446-513,534-549
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

class Frame {                    // Frame.h members read at MapPoint.cc:87-121,551-566
public:
    long unsigned int mnId = 0;
    int Nleft = -1;
    cv::Mat mRwc, mTlr, mOw, mDescriptors;
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn, mvKeysRight;
    std::vector<float> mvScaleFactors;
    int mnScaleLevels = 0;
    float mfLogScaleFactor = 0;
    cv::Mat GetCameraCenter() { return mOw.clone(); }
};

}  // namespace ORB_SLAM3
