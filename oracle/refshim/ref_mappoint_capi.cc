// TEST INFRASTRUCTURE (oracle/_ref): C entry points over the reference's own src/MapPoint.cc, compiled unmodified
// against refshim/mappoint_standin.h.  ORBmatcher::DescriptorDistance comes from liborbslam3_orbmatcher.so (the
// reference's own ORBmatcher.cc).  Nothing here is linked into the product.
#include "MapPoint.h"

#include <cstdint>
#include <cstring>

using namespace ORB_SLAM3;

namespace {

struct OpenMapPoint : MapPoint {       // reach the protected distance fields without touching the reference header
    OpenMapPoint(const cv::Mat& Pos, KeyFrame* kf, Map* map) : MapPoint(Pos, kf, map) {}
    void setMaxDistance(float d) { mfMaxDistance = d; }
};

cv::Mat origin() { return cv::Mat::zeros(3, 1, CV_32F); }

}  // namespace

extern "C" {

// MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:349-423) for `npoints` map points; point p is observed by rows
// offsets[p] .. offsets[p+1] of `desc`.  mode 0: one monocular KeyFrame per row.  mode 1: stereo-fisheye KeyFrames holding
// two rows each (left index, right index; an odd last row is a left-only observation).  The observation map is keyed by
// KeyFrame POINTER (MapPoint.h:164), so the KeyFrames live in one array: address order == row order.
// best[p] = row (relative to offsets[p]) whose bytes the point's descriptor is a copy of, first such row; -1 = none.
void orbslam3_distinctive_descriptors(const uint8_t* desc, const int32_t* offsets, int npoints, int mode, int32_t* best) {
    Map map;
    for (int p = 0; p < npoints; ++p) {
        const int n = offsets[p + 1] - offsets[p];
        best[p] = -1;
        if (n <= 0) continue;
        uint8_t* rows = const_cast<uint8_t*>(desc) + 32 * (size_t)offsets[p];
        const int nkf = mode ? (n + 1) / 2 : n;
        std::vector<KeyFrame> kfs(nkf);
        for (int k = 0; k < nkf; ++k) {
            KeyFrame& kf = kfs[k];
            kf.mnId = k;
            kf.mDescriptors = cv::Mat(n, 32, CV_8U, rows, 32);
            kf.mvuRight.assign(n, -1.f);
            kf.mvpMapPoints.assign(n, nullptr);
            if (mode) { kf.NLeft = 2 * k + 1; kf.mpCamera2 = reinterpret_cast<GeometricCamera*>(&map); }
        }
        OpenMapPoint mp(origin(), &kfs[0], &map);
        if (mode) for (int i = 0; i < n; ++i) mp.AddObservation(&kfs[i / 2], i);
        else      for (int i = 0; i < n; ++i) mp.AddObservation(&kfs[i], i);
        mp.ComputeDistinctiveDescriptors();
        const cv::Mat d = mp.GetDescriptor();
        if (d.empty()) continue;
        for (int i = 0; i < n; ++i)
            if (!std::memcmp(rows + 32 * (size_t)i, d.ptr(0), 32)) { best[p] = i; break; }
    }
}

// nObs after the observations above (AddObservation MapPoint.cc:161-186), for the harness's own sanity test.
int orbslam3_mappoint_nobs(int n, int mode) {
    Map map;
    const int nkf = mode ? (n + 1) / 2 : n;
    std::vector<KeyFrame> kfs(nkf);
    for (int k = 0; k < nkf; ++k) {
        kfs[k].mvuRight.assign(n, -1.f);
        if (mode) { kfs[k].NLeft = 2 * k + 1; kfs[k].mpCamera2 = reinterpret_cast<GeometricCamera*>(&map); }
    }
    OpenMapPoint mp(origin(), &kfs[0], &map);
    for (int i = 0; i < n; ++i) mp.AddObservation(&kfs[mode ? i / 2 : i], i);
    return mp.Observations();
}

// MapPoint::PredictScale (KeyFrame overload MapPoint.cc:534-549, Frame overload :551-566).
void orbslam3_mappoint_predict_scale(const float* max_distance, const float* current_dist, int n, float log_scale_factor,
                                     int n_levels, int via_frame, int32_t* level) {
    Map map;
    KeyFrame kf;
    kf.mfLogScaleFactor = log_scale_factor;
    kf.mnScaleLevels = n_levels;
    Frame fr;
    fr.mfLogScaleFactor = log_scale_factor;
    fr.mnScaleLevels = n_levels;
    OpenMapPoint mp(origin(), &kf, &map);
    for (int i = 0; i < n; ++i) {
        mp.setMaxDistance(max_distance[i]);
        level[i] = via_frame ? mp.PredictScale(current_dist[i], &fr) : mp.PredictScale(current_dist[i], &kf);
    }
}

}  // extern "C"
