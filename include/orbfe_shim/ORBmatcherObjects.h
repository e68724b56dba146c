// ORBmatcherObjects.h — the object layer of the ORB_SLAM3::ORBmatcher shim: every Search* / Fuse member of the
// reference's class (include/ORBmatcher.h:44-91) with the reference's signature, on top of the orbfe C ABI.
//
// Included by ORBmatcher.h when ORBFE_WITH_ORBSLAM3 is defined (Frame / KeyFrame / MapPoint / GeometricCamera must be
// complete types by then).  Division of labour, the same for every method:
//   host   what differs between the overloads — which MapPoints take part and where each one's search window lies in
//          the image (pose, projection, distance / viewing-angle gates, predicted pyramid level).  This is a few float
//          operations per MapPoint on objects that live in host memory; it runs here, in the float operation order of
//          OpenCV's small-matrix algebra (products summed left to right, the translation added last; cv::norm and
//          Mat::dot accumulate in double), so that windows are bit-identical to the reference's;
//   device everything that touches descriptors — Frame / KeyFrame::GetFeaturesInArea, the Hamming scans, best / second
//          best, the ordered claiming of features, rotation histograms: orbfe_search_by_projection,
//          orbfe_search_for_initialization, orbfe_search_bow, orbfe_search_triangulation[_kb8];
//   host   writing the outcome back where the reference writes it (Frame::mvpMapPoints, vpMatched, vMatchedPairs, the
//          AddObservation / AddMapPoint / Replace calls of Fuse), in the reference's order.
#ifndef ORBFE_SHIM_ORBMATCHER_OBJECTS_H
#define ORBFE_SHIM_ORBMATCHER_OBJECTS_H

#include <cmath>
#include <cstring>
#include <set>
#include <tuple>
#include <vector>

namespace ORB_SLAM3 {
namespace orbfe_shim {

// A rigid (or similarity) transform as 12 floats read once from a cv::Mat; apply() is OpenCV's 3x3 * 3x1 + 3x1:
// float products added left to right, then the translation.
struct Rt {
    float r[3][3], t[3];
    Rt() { std::memset(this, 0, sizeof(*this)); }
    Rt(const cv::Mat& R, const cv::Mat& tvec) {
        for (int i = 0; i < 3; ++i) {
            for (int j = 0; j < 3; ++j) r[i][j] = R.at<float>(i, j);
            t[i] = tvec.at<float>(i);
        }
    }
    // the top 3x4 block of a 4x4 (or 3x4) pose matrix
    explicit Rt(const cv::Mat& T) {
        for (int i = 0; i < 3; ++i) {
            for (int j = 0; j < 3; ++j) r[i][j] = T.at<float>(i, j);
            t[i] = T.at<float>(i, 3);
        }
    }
    void apply(const float p[3], float out[3]) const {
        for (int i = 0; i < 3; ++i) {
            float s = r[i][0] * p[0];
            s += r[i][1] * p[1];
            s += r[i][2] * p[2];
            out[i] = s + t[i];
        }
    }
};

struct Vec3 {
    float v[3];
    Vec3() { v[0] = v[1] = v[2] = 0.f; }
    explicit Vec3(const cv::Mat& m) { for (int i = 0; i < 3; ++i) v[i] = m.at<float>(i); }
};

// cv::norm of a float 3-vector: squares accumulated in double, the root narrowed by the caller's `const float`.
inline float norm3(const float a[3]) {
    double s = 0.0;
    for (int i = 0; i < 3; ++i) s += (double)a[i] * (double)a[i];
    return (float)std::sqrt(s);
}
// cv::Mat::dot of two float 3-vectors: a double.
inline double dot3(const float a[3], const float b[3]) {
    double s = 0.0;
    for (int i = 0; i < 3; ++i) s += (double)a[i] * (double)b[i];
    return s;
}

// The decomposition both Sim3 entry points start with (ORBmatcher.cc:483-487, 1830-1835), written with the same cv::Mat
// operators so that the library rounds exactly as it does for the reference.
struct Sim3Pose {
    Rt cw;         // Rcw, tcw (scale removed)
    Vec3 Ow;
    explicit Sim3Pose(const cv::Mat& Scw) {
        const cv::Mat sR = Scw.rowRange(0, 3).colRange(0, 3);
        const float s = (float)std::sqrt(sR.row(0).dot(sR.row(0)));
        const cv::Mat R = sR / s, t = Scw.rowRange(0, 3).col(3) / s;
        const cv::Mat O = -R.t() * t;
        cw = Rt(R, t);
        Ow = Vec3(O);
    }
};

inline void set_descriptor(orbfe_proj_query& q, const cv::Mat& d) { std::memcpy(q.desc, d.ptr(), 32); }

inline orbfe_proj_query make_window(float u, float v, float r, int minLevel, int maxLevel) {
    orbfe_proj_query q;
    std::memset(&q, 0, sizeof(q));
    q.u = u; q.v = v; q.r = r;
    q.min_level = (int16_t)minLevel; q.max_level = (int16_t)maxLevel;
    q.valid = 1;
    return q;
}

// What the device needs of the Frame / KeyFrame whose 64 x 48 grid is searched.  FrameT: ORB_SLAM3::Frame (Nleft) or
// ORB_SLAM3::KeyFrame (NLeft); the two spell the member differently, hence the explicit argument.
struct GridView {
    std::vector<cv::KeyPoint> both;      // mvKeys ++ mvKeysRight of a two-camera rig
    std::vector<uint8_t> blocked;
    std::vector<int32_t> l2r, r2l;
    orbfe_grid_frame g;

    template <class FrameT>
    GridView(const FrameT& F, int nleft) {
        std::memset(&g, 0, sizeof(g));
        if (nleft == -1) {
            g.kps = reinterpret_cast<const orbfe_keypoint*>(F.mvKeysUn.data());
            g.n = (int)F.mvKeysUn.size();
        } else {
            both = F.mvKeys;
            both.insert(both.end(), F.mvKeysRight.begin(), F.mvKeysRight.end());
            g.kps = reinterpret_cast<const orbfe_keypoint*>(both.data());
            g.n = (int)both.size();
        }
        g.nleft = nleft;
        g.desc = F.mDescriptors.data;
        if ((int)F.mvuRight.size() >= g.n && g.n > 0) g.u_right = F.mvuRight.data();
        g.min_x = (float)F.mnMinX; g.min_y = (float)F.mnMinY;
        g.grid_w_inv = F.mfGridElementWidthInv; g.grid_h_inv = F.mfGridElementHeightInv;
        blocked.assign(g.n > 0 ? g.n : 1, 0);
        g.blocked = blocked.data();
    }
    template <class FrameT>
    void stereoLinks(const FrameT& F) {      // Frame::mvLeftToRightMatch / mvRightToLeftMatch of a fisheye rig
        if (g.nleft == -1) return;
        l2r.assign(F.mvLeftToRightMatch.begin(), F.mvLeftToRightMatch.end());
        r2l.assign(F.mvRightToLeftMatch.begin(), F.mvRightToLeftMatch.end());
        if ((int)l2r.size() >= g.nleft && g.nleft > 0) g.left_to_right = l2r.data();
        if ((int)r2l.size() >= g.n - g.nleft && g.n > g.nleft) g.right_to_left = r2l.data();
    }
};

// bestDist <= TH_LOW * ratioHamming compares an int with a float product: the largest int that passes.
inline int hamming_limit(int thLow, float ratioHamming) { return (int)std::floor((float)thLow * ratioHamming); }

// One side of a BoW / triangulation search from a KeyFrame.
struct KeyFrameSide {
    FeatureVectorCsr fv;
    std::vector<uint8_t> mask;
    std::vector<cv::KeyPoint> both;
    Side side;
    template <class MaskFn>
    KeyFrameSide(KeyFrame* pKF, MaskFn hasPoint) : fv(pKF->mFeatVec) {
        const std::vector<MapPoint*> mps = pKF->GetMapPointMatches();
        mask.resize(mps.size() ? mps.size() : 1);
        for (size_t i = 0; i < mps.size(); ++i) mask[i] = hasPoint(mps[i]) ? 1 : 0;
        side.desc = pKF->mDescriptors.data;
        side.n = pKF->mDescriptors.rows;
        side.fv = &fv;
        side.mask = mask.data();
        if (pKF->NLeft == -1) side.kps = pKF->mvKeysUn.data();
        else {
            both = pKF->mvKeys;
            both.insert(both.end(), pKF->mvKeysRight.begin(), pKF->mvKeysRight.end());
            side.kps = both.data();
        }
    }
};

}  // namespace orbfe_shim

// ---- Tracking: local map points into the current frame (ORBmatcher.cc:44-236) -----------------------------------
inline int ORBmatcher::SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th, const bool bFarPoints,
                                          const float thFarPoints) {
    if (!m_) return 0;
    const bool scaleRadius = th != 1.0;
    std::vector<orbfe_proj_query> windows;
    std::vector<MapPoint*> owner;                // MapPoint of each window
    windows.reserve(vpMapPoints.size());
    owner.reserve(vpMapPoints.size());
    for (MapPoint* pMP : vpMapPoints) {
        const bool seenLeft = pMP->mbTrackInView, seenRight = pMP->mbTrackInViewR;
        if (!seenLeft && !seenRight) continue;
        if (bFarPoints && pMP->mTrackDepth > thFarPoints) continue;
        if (pMP->isBad()) continue;
        const bool hasObs = pMP->Observations() > 0;
        if (seenLeft) {
            const int level = pMP->mnTrackScaleLevel;
            float r = RadiusByViewingCos(pMP->mTrackViewCos);
            if (scaleRadius) r *= th;
            orbfe_proj_query q = orbfe_shim::make_window(pMP->mTrackProjX, pMP->mTrackProjY, r * F.mvScaleFactors[level], level - 1, level);
            q.ur = pMP->mTrackProjXR;
            q.check_ur = 1;                      // the device applies it only to frames without a second camera (:96)
            q.obs = hasObs;
            orbfe_shim::set_descriptor(q, pMP->GetDescriptor());
            windows.push_back(q);
            owner.push_back(pMP);
        }
        if (F.Nleft != -1 && seenRight) {
            const int level = pMP->mnTrackScaleLevelR;
            if (level == -1) continue;
            const float r = RadiusByViewingCos(pMP->mTrackViewCosR);     // th is not applied to the right camera (:151)
            orbfe_proj_query q = orbfe_shim::make_window(pMP->mTrackProjXR, pMP->mTrackProjYR, r * F.mvScaleFactors[level], level - 1, level);
            q.right = 1;
            q.obs = hasObs;
            q.skip_if_prev_ratio_rejected = seenLeft;   // the `continue` of :135 leaves the whole MapPoint
            orbfe_shim::set_descriptor(q, pMP->GetDescriptor());
            windows.push_back(q);
            owner.push_back(pMP);
        }
    }
    orbfe_shim::GridView G(F, F.Nleft);
    G.stereoLinks(F);
    for (int i = 0; i < G.g.n && i < (int)F.mvpMapPoints.size(); ++i)
        G.blocked[i] = F.mvpMapPoints[i] && F.mvpMapPoints[i]->Observations() > 0;
    std::vector<int32_t> assign;
    const int n = SearchByProjectionWindows(G.g, windows, /*mode*/ 1, TH_HIGH, /*linkStereo*/ true, /*rotHist*/ false, assign);
    for (size_t i = 0; i < assign.size(); ++i)
        if (assign[i] >= 0) F.mvpMapPoints[i] = owner[assign[i]];
    return n;
}

// ---- Tracking: the last frame's MapPoints into the current frame (ORBmatcher.cc:2173-2389) -------------------------
inline int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono) {
    if (!m_) return 0;
    const orbfe_shim::Rt Tcw(CurrentFrame.mTcw);
    bool forward, backward;
    {   // which way the camera moved along its optical axis, from the two poses (cv::Mat algebra as the reference writes it)
        const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3), tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
        const cv::Mat twc = -Rcw.t() * tcw;
        const cv::Mat tlc = LastFrame.mTcw.rowRange(0, 3).colRange(0, 3) * twc + LastFrame.mTcw.rowRange(0, 3).col(3);
        const float dz = tlc.at<float>(2);
        forward = dz > CurrentFrame.mb && !bMono;
        backward = -dz > CurrentFrame.mb && !bMono;
    }
    const bool rig = CurrentFrame.Nleft != -1;
    orbfe_shim::Rt Trl;
    if (rig) Trl = orbfe_shim::Rt(CurrentFrame.mTrl);
    std::vector<orbfe_proj_query> windows;
    std::vector<MapPoint*> owner;
    for (int i = 0; i < LastFrame.N; ++i) {
        MapPoint* pMP = LastFrame.mvpMapPoints[i];
        if (!pMP || LastFrame.mvbOutlier[i]) continue;
        const orbfe_shim::Vec3 Xw(pMP->GetWorldPos());
        float Xc[3];
        Tcw.apply(Xw.v, Xc);
        const float invzc = 1.0 / Xc[2];
        if (invzc < 0) continue;
        const cv::Point2f uv = CurrentFrame.mpCamera->project(cv::Point3f(Xc[0], Xc[1], Xc[2]));
        if (uv.x < CurrentFrame.mnMinX || uv.x > CurrentFrame.mnMaxX || uv.y < CurrentFrame.mnMinY || uv.y > CurrentFrame.mnMaxY) continue;
        const bool leftFeature = LastFrame.Nleft == -1 || i < LastFrame.Nleft;
        const cv::KeyPoint& kpLevel = leftFeature ? LastFrame.mvKeys[i] : LastFrame.mvKeysRight[i - LastFrame.Nleft];
        const cv::KeyPoint& kpAngle = LastFrame.Nleft == -1 ? LastFrame.mvKeysUn[i] : kpLevel;
        const int octave = kpLevel.octave;
        const float radius = th * CurrentFrame.mvScaleFactors[octave];
        const int lo = forward ? octave : backward ? 0 : octave - 1;
        const int hi = forward ? -1 : backward ? octave : octave + 1;
        const bool hasObs = pMP->Observations() > 0;
        orbfe_proj_query q = orbfe_shim::make_window(uv.x, uv.y, radius, lo, hi);
        q.ur = uv.x - CurrentFrame.mbf * invzc;
        q.check_ur = 1;
        q.angle = kpAngle.angle;
        q.obs = hasObs;
        const cv::Mat d = pMP->GetDescriptor();
        orbfe_shim::set_descriptor(q, d);
        windows.push_back(q);
        owner.push_back(pMP);
        if (rig) {                                // the same point through the right camera's pose (:2291-2353)
            float Xr[3];
            Trl.apply(Xc, Xr);
            const cv::Point2f uvr = CurrentFrame.mpCamera->project(cv::Point3f(Xr[0], Xr[1], Xr[2]));
            orbfe_proj_query qr = orbfe_shim::make_window(uvr.x, uvr.y, radius, lo, hi);
            qr.right = 1;
            qr.skip_if_prev_empty = 1;            // an empty left window leaves the MapPoint altogether (:2236-2237)
            qr.angle = kpAngle.angle;
            qr.obs = hasObs;
            orbfe_shim::set_descriptor(qr, d);
            windows.push_back(qr);
            owner.push_back(pMP);
        }
    }
    orbfe_shim::GridView G(CurrentFrame, CurrentFrame.Nleft);
    for (int i = 0; i < G.g.n && i < (int)CurrentFrame.mvpMapPoints.size(); ++i)
        G.blocked[i] = CurrentFrame.mvpMapPoints[i] && CurrentFrame.mvpMapPoints[i]->Observations() > 0;
    std::vector<int32_t> assign;
    const int n = SearchByProjectionWindows(G.g, windows, /*mode*/ 0, TH_HIGH, false, /*rotHist*/ true, assign);
    for (size_t i = 0; i < assign.size(); ++i) {
        if (assign[i] >= 0) CurrentFrame.mvpMapPoints[i] = owner[assign[i]];
        else if (assign[i] == -2) CurrentFrame.mvpMapPoints[i] = static_cast<MapPoint*>(NULL);     // dropped by the rotation check
    }
    return n;
}

// ---- Relocalisation: a candidate keyframe's MapPoints into the current frame (ORBmatcher.cc:2391-2510) ------------
inline int ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound, const float th,
                                          const int ORBdist) {
    if (!m_) return 0;
    const orbfe_shim::Rt Tcw(CurrentFrame.mTcw);
    orbfe_shim::Vec3 Ow;
    {
        const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3), tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
        const cv::Mat O = -Rcw.t() * tcw;
        Ow = orbfe_shim::Vec3(O);
    }
    const std::vector<MapPoint*> vpMPs = pKF->GetMapPointMatches();
    std::vector<orbfe_proj_query> windows;
    std::vector<MapPoint*> owner;
    for (size_t i = 0; i < vpMPs.size(); ++i) {
        MapPoint* pMP = vpMPs[i];
        if (!pMP || pMP->isBad() || sAlreadyFound.count(pMP)) continue;
        const orbfe_shim::Vec3 Xw(pMP->GetWorldPos());
        float Xc[3];
        Tcw.apply(Xw.v, Xc);
        const cv::Point2f uv = CurrentFrame.mpCamera->project(cv::Point3f(Xc[0], Xc[1], Xc[2]));
        if (uv.x < CurrentFrame.mnMinX || uv.x > CurrentFrame.mnMaxX || uv.y < CurrentFrame.mnMinY || uv.y > CurrentFrame.mnMaxY) continue;
        const float PO[3] = {Xw.v[0] - Ow.v[0], Xw.v[1] - Ow.v[1], Xw.v[2] - Ow.v[2]};
        float dist3D = orbfe_shim::norm3(PO);
        const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        const int level = pMP->PredictScale(dist3D, &CurrentFrame);
        orbfe_proj_query q = orbfe_shim::make_window(uv.x, uv.y, th * CurrentFrame.mvScaleFactors[level], level - 1, level + 1);
        q.angle = pKF->mvKeysUn[i].angle;
        q.obs = 1;                                // here ANY MapPoint in mvpMapPoints hides a feature (:2455): every match claims
        orbfe_shim::set_descriptor(q, pMP->GetDescriptor());
        windows.push_back(q);
        owner.push_back(pMP);
    }
    orbfe_shim::GridView G(CurrentFrame, CurrentFrame.Nleft);
    for (int i = 0; i < G.g.n && i < (int)CurrentFrame.mvpMapPoints.size(); ++i) G.blocked[i] = CurrentFrame.mvpMapPoints[i] != NULL;
    G.g.u_right = nullptr;                        // no stereo gate in this overload
    std::vector<int32_t> assign;
    const int n = SearchByProjectionWindows(G.g, windows, /*mode*/ 0, ORBdist, false, /*rotHist*/ true, assign);
    for (size_t i = 0; i < assign.size(); ++i) {
        if (assign[i] >= 0) CurrentFrame.mvpMapPoints[i] = owner[assign[i]];
        else if (assign[i] == -2) CurrentFrame.mvpMapPoints[i] = static_cast<MapPoint*>(NULL);
    }
    return n;
}

// ---- Loop closing / merging: MapPoints through a Sim3 into a keyframe (ORBmatcher.cc:473-585 and its twin :588-700) ---
// The twin projects with fx * (x / z) + cx instead of the camera object (one rounding more), hence `viaCamera`.
inline int ORBmatcher::sim3Projection(KeyFrame* pKF, const cv::Mat& Scw, const std::vector<MapPoint*>& vpPoints,
                                      const std::vector<KeyFrame*>* vpPointsKFs, std::vector<MapPoint*>& vpMatched,
                                      std::vector<KeyFrame*>* vpMatchedKF, int th, float ratioHamming) {
    if (!m_) return 0;
    const bool viaCamera = vpPointsKFs == nullptr;
    const orbfe_shim::Sim3Pose S(Scw);
    std::set<MapPoint*> already(vpMatched.begin(), vpMatched.end());
    already.erase(static_cast<MapPoint*>(NULL));
    std::vector<orbfe_proj_query> windows;
    std::vector<int> owner;                       // index into vpPoints
    for (int k = 0, kend = (int)vpPoints.size(); k < kend; ++k) {
        MapPoint* pMP = vpPoints[k];
        if (pMP->isBad() || already.count(pMP)) continue;
        const orbfe_shim::Vec3 Xw(pMP->GetWorldPos());
        float Xc[3];
        S.cw.apply(Xw.v, Xc);
        if (Xc[2] < 0.0) continue;
        float u, v;
        if (viaCamera) {
            const cv::Point2f uv = pKF->mpCamera->project(cv::Point3f(Xc[0], Xc[1], Xc[2]));
            u = uv.x; v = uv.y;
        } else {
            const float invz = 1 / Xc[2];
            const float x = Xc[0] * invz, y = Xc[1] * invz;
            u = pKF->fx * x + pKF->cx;
            v = pKF->fy * y + pKF->cy;
        }
        if (!pKF->IsInImage(u, v)) continue;
        const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
        const float PO[3] = {Xw.v[0] - S.Ow.v[0], Xw.v[1] - S.Ow.v[1], Xw.v[2] - S.Ow.v[2]};
        const float dist = orbfe_shim::norm3(PO);
        if (dist < minDistance || dist > maxDistance) continue;
        const orbfe_shim::Vec3 Pn(pMP->GetNormal());
        if (orbfe_shim::dot3(PO, Pn.v) < 0.5 * dist) continue;            // more than 60 degrees off the mean viewing direction
        const int level = pMP->PredictScale(dist, pKF);
        orbfe_proj_query q = orbfe_shim::make_window(u, v, th * pKF->mvScaleFactors[level], level - 1, level);
        q.obs = 1;                                // vpMatched[bestIdx] = pMP hides the feature from later points (:558)
        orbfe_shim::set_descriptor(q, pMP->GetDescriptor());
        windows.push_back(q);
        owner.push_back(k);
    }
    orbfe_shim::GridView G(*pKF, -1);             // these overloads read mvKeysUn and the left grid only
    for (int i = 0; i < G.g.n && i < (int)vpMatched.size(); ++i) G.blocked[i] = vpMatched[i] != NULL;
    G.g.u_right = nullptr;
    std::vector<int32_t> assign;
    const int n = SearchByProjectionWindows(G.g, windows, /*mode*/ 0, orbfe_shim::hamming_limit(TH_LOW, ratioHamming), false, false, assign);
    for (size_t i = 0; i < assign.size(); ++i)
        if (assign[i] >= 0) {
            vpMatched[i] = vpPoints[owner[assign[i]]];
            if (vpMatchedKF) (*vpMatchedKF)[i] = (*vpPointsKFs)[owner[assign[i]]];
        }
    return n;
}

inline int ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, std::vector<MapPoint*>& vpMatched,
                                          int th, float ratioHamming) {
    return sim3Projection(pKF, Scw, vpPoints, nullptr, vpMatched, nullptr, th, ratioHamming);
}

inline int ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints,
                                          const std::vector<KeyFrame*>& vpPointsKFs, std::vector<MapPoint*>& vpMatched,
                                          std::vector<KeyFrame*>& vpMatchedKF, int th, float ratioHamming) {
    return sim3Projection(pKF, Scw, vpPoints, &vpPointsKFs, vpMatched, &vpMatchedKF, th, ratioHamming);
}

// ---- BoW-guided searches (ORBmatcher.cc:269-471, 823-963) ------------------------------------------------------------
inline int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches) {
    const std::vector<MapPoint*> kfPoints = pKF->GetMapPointMatches();
    vpMapPointMatches = std::vector<MapPoint*>(F.N, static_cast<MapPoint*>(NULL));
    orbfe_shim::KeyFrameSide A(pKF, [](MapPoint* p) { return p && !p->isBad(); });
    orbfe_shim::FeatureVectorCsr fvF(F.mFeatVec);
    std::vector<cv::KeyPoint> both;
    orbfe_shim::Side B;
    B.desc = F.mDescriptors.data; B.n = F.mDescriptors.rows; B.fv = &fvF;
    if (F.Nleft == -1) B.kps = F.mvKeys.data();                    // the rotation check reads F.mvKeys here (:375)
    else {
        both = F.mvKeys;
        both.insert(both.end(), F.mvKeysRight.begin(), F.mvKeysRight.end());
        B.kps = both.data();
    }
    std::vector<int32_t> kfOfFrameFeature;
    const int n = SearchByBoWArrays(A.side, B, F.Nleft, kfOfFrameFeature);
    for (size_t j = 0; j < kfOfFrameFeature.size() && j < vpMapPointMatches.size(); ++j)
        if (kfOfFrameFeature[j] >= 0) vpMapPointMatches[j] = kfPoints[kfOfFrameFeature[j]];
    return n;
}

inline int ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12) {
    const std::vector<MapPoint*> points1 = pKF1->GetMapPointMatches(), points2 = pKF2->GetMapPointMatches();
    vpMatches12 = std::vector<MapPoint*>(points1.size(), static_cast<MapPoint*>(NULL));
    auto good = [](MapPoint* p) { return p && !p->isBad(); };
    orbfe_shim::KeyFrameSide A(pKF1, good), B(pKF2, good);
    A.side.kps = pKF1->mvKeysUn.data();                              // this overload reads mvKeysUn only (:906)
    B.side.kps = pKF2->mvKeysUn.data();
    std::vector<int32_t> kf2OfKf1Feature;
    const int n = SearchByBoWArrays(A.side, B.side, kf2OfKf1Feature);
    for (size_t i = 0; i < kf2OfKf1Feature.size() && i < vpMatches12.size(); ++i)
        if (kf2OfKf1Feature[i] >= 0) vpMatches12[i] = points2[kf2OfKf1Feature[i]];
    return n;
}

// ---- Monocular initialisation (ORBmatcher.cc:706-821) -------------------------------------------------------------------
inline int ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12,
                                               int windowSize) {
    orbfe_shim::GridView G(F2, -1);
    G.g.blocked = nullptr;
    G.g.u_right = nullptr;
    return SearchForInitializationArrays(F1.mvKeysUn, F1.mDescriptors, G.g, vbPrevMatched, vnMatches12, windowSize);
}

// ---- Epipolar search for new MapPoints (ORBmatcher.cc:965-1624) -----------------------------------------------------------
namespace orbfe_shim {
inline void copy33(const cv::Matx33f& M, float* out) { for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) out[3 * i + j] = M(i, j); }
inline void copy31(const cv::Matx31f& v, float* out) { for (int i = 0; i < 3; ++i) out[i] = v(i); }
template <class Camera> inline bool is_fisheye(Camera* c) { return c && c->GetType() == c->CAM_FISHEYE; }
template <class Camera> inline void camera_parameters(Camera* c, float* out8) { for (int i = 0; i < 8; ++i) out8[i] = c->getParameter(i); }
}  // namespace orbfe_shim

inline int ORBmatcher::triangulationSearch(KeyFrame* pKF1, KeyFrame* pKF2, const cv::Matx33f& F12,
                                           std::vector<std::pair<size_t, size_t> >& vMatchedPairs, bool bOnlyStereo, bool bCoarse,
                                           std::vector<cv::Mat>* vMatchedPoints) {
    vMatchedPairs.clear();
    if (vMatchedPoints) vMatchedPoints->clear();
    if (!m_) return 0;
    // where camera 1 sits in image 2 (features next to it are skipped, :1089-1097)
    const cv::Matx31f C2 = pKF2->GetRotation_() * pKF1->GetCameraCenter_() + pKF2->GetTranslation_();
    const cv::Point2f ep = pKF2->mpCamera->project(C2);
    auto hasPoint = [](MapPoint* p) { return p != NULL; };
    orbfe_shim::KeyFrameSide A(pKF1, hasPoint), B(pKF2, hasPoint);
    const int nlevels = (int)pKF2->mvScaleFactors.size();
    if (!orbfe_shim::is_fisheye(pKF1->mpCamera)) {
        // Pinhole::epipolarConstrain_ rebuilds K1^-T [t12]x R12 K2^-1 for every pair from the same poses and intrinsics the
        // caller's F12 was made of (LocalMapping::ComputeF12_, the same expression): the argument IS that matrix.
        orbfe_epipolar geo;
        orbfe_shim::copy33(F12, geo.F12);
        geo.ex = ep.x; geo.ey = ep.y;
        geo.sigma2_b = pKF2->mvLevelSigma2.data();
        geo.scale_b = pKF2->mvScaleFactors.data();
        geo.nlevels = nlevels;
        const float* ur1 = pKF1->mpCamera2 ? nullptr : pKF1->mvuRight.data();
        const float* ur2 = pKF2->mpCamera2 ? nullptr : pKF2->mvuRight.data();
        return SearchForTriangulationArrays(A.side, ur1, B.side, ur2, geo, bOnlyStereo, bCoarse, vMatchedPairs);
    }
    // Kannala-Brandt keyframes: the camera pair of every (feature 1, feature 2) combination and its relative pose
    orbfe_kb8_epipolar geo;
    std::memset(&geo, 0, sizeof(geo));
    orbfe_shim::camera_parameters(pKF1->mpCamera, geo.cam_a[0]);
    orbfe_shim::camera_parameters(pKF2->mpCamera, geo.cam_b[0]);
    if (pKF1->mpCamera2) orbfe_shim::camera_parameters(pKF1->mpCamera2, geo.cam_a[1]);
    if (pKF2->mpCamera2) orbfe_shim::camera_parameters(pKF2->mpCamera2, geo.cam_b[1]);
    geo.nleft_a = pKF1->NLeft; geo.nleft_b = pKF2->NLeft;
    auto relative = [&](const cv::Matx33f& Ra, const cv::Matx31f& ta, const cv::Matx33f& Rb, const cv::Matx31f& tb, int slot, bool nested) {
        const cv::Matx33f R = Ra * Rb.t();
        const cv::Matx31f t = nested ? Ra * (-Rb.t() * tb) + ta : -Ra * Rb.t() * tb + ta;     // the two spellings of :1234-1262
        orbfe_shim::copy33(R, geo.R12[slot]);
        orbfe_shim::copy31(t, geo.t12[slot]);
    };
    if (!pKF1->mpCamera2 && !pKF2->mpCamera2) {
        relative(pKF1->GetRotation_(), pKF1->GetTranslation_(), pKF2->GetRotation_(), pKF2->GetTranslation_(), 0, false);
    } else {
        relative(pKF1->GetRotation_(), pKF1->GetTranslation_(), pKF2->GetRotation_(), pKF2->GetTranslation_(), 0, true);
        relative(pKF1->GetRotation_(), pKF1->GetTranslation_(), pKF2->GetRightRotation_(), pKF2->GetRightTranslation_(), 1, true);
        relative(pKF1->GetRightRotation_(), pKF1->GetRightTranslation_(), pKF2->GetRotation_(), pKF2->GetTranslation_(), 2, true);
        relative(pKF1->GetRightRotation_(), pKF1->GetRightTranslation_(), pKF2->GetRightRotation_(), pKF2->GetRightTranslation_(), 3, true);
    }
    geo.ex = ep.x; geo.ey = ep.y;
    geo.sigma2_a = pKF1->mvLevelSigma2.data();
    geo.sigma2_b = pKF2->mvLevelSigma2.data();
    geo.scale_b = pKF2->mvScaleFactors.data();
    geo.nlevels = nlevels;
    std::vector<float> pts;
    const int n = SearchForTriangulationKB8Arrays(A.side, B.side, geo, bOnlyStereo, bCoarse, vMatchedPairs, vMatchedPoints ? &pts : nullptr);
    if (vMatchedPoints) {                         // camera-1 coordinates -> world, as the matchAndtriangulate overload returns them
        const cv::Mat Twc = pKF1->GetPoseInverse();
        for (size_t k = 0; k + 2 < pts.size(); k += 3) {
            cv::Mat x = (cv::Mat_<float>(3, 1) << pts[k], pts[k + 1], pts[k + 2]);
            vMatchedPoints->push_back(Twc.rowRange(0, 3).colRange(0, 3) * x + Twc.rowRange(0, 3).col(3));
        }
    }
    return n;
}

inline int ORBmatcher::SearchForTriangulation_(KeyFrame* pKF1, KeyFrame* pKF2, cv::Matx33f F12,
                                               std::vector<std::pair<size_t, size_t> >& vMatchedPairs, const bool bOnlyStereo,
                                               const bool bCoarse) {
    return triangulationSearch(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo, bCoarse, nullptr);
}

inline int ORBmatcher::SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12,
                                              std::vector<std::pair<size_t, size_t> >& vMatchedPairs, const bool bOnlyStereo,
                                              const bool bCoarse) {
    cv::Matx33f F;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) F(i, j) = F12.at<float>(i, j);
    return triangulationSearch(pKF1, pKF2, F, vMatchedPairs, bOnlyStereo, bCoarse, nullptr);
}

inline int ORBmatcher::SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12,
                                              std::vector<std::pair<size_t, size_t> >& vMatchedPairs, const bool bOnlyStereo,
                                              std::vector<cv::Mat>& vMatchedPoints) {
    cv::Matx33f F;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) F(i, j) = F12.at<float>(i, j);
    // only Kannala-Brandt cameras define matchAndtriangulate in the reference; with pinhole cameras the pairs come back
    // without points
    return triangulationSearch(pKF1, pKF2, F, vMatchedPairs, bOnlyStereo, false, &vMatchedPoints);
}

// ---- Loop closing: mutual projection of two keyframes' MapPoints through a Sim3 (ORBmatcher.cc:1947-2171) ----------------
inline int ORBmatcher::SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const float& s12,
                                    const cv::Mat& R12, const cv::Mat& t12, const float th) {
    if (!m_) return 0;
    const float fx = pKF1->fx, fy = pKF1->fy, cx = pKF1->cx, cy = pKF1->cy;      // the reference projects both ways with pKF1's intrinsics
    const orbfe_shim::Rt T1w(pKF1->GetRotation(), pKF1->GetTranslation()), T2w(pKF2->GetRotation(), pKF2->GetTranslation());
    orbfe_shim::Rt T12, T21;                      // similarity maps between the two cameras
    {
        const cv::Mat sR12 = s12 * R12;
        const cv::Mat sR21 = (1.0 / s12) * R12.t();
        const cv::Mat t21 = -sR21 * t12;
        T12 = orbfe_shim::Rt(sR12, t12);
        T21 = orbfe_shim::Rt(sR21, t21);
    }
    const std::vector<MapPoint*> points1 = pKF1->GetMapPointMatches(), points2 = pKF2->GetMapPointMatches();
    const int N1 = (int)points1.size(), N2 = (int)points2.size();
    std::vector<char> matched1(N1, 0), matched2(N2, 0);
    for (int i = 0; i < N1; ++i) {
        MapPoint* pMP = vpMatches12[i];
        if (!pMP) continue;
        matched1[i] = 1;
        const int idx2 = std::get<0>(pMP->GetIndexInKeyFrame(pKF2));
        if (idx2 >= 0 && idx2 < N2) matched2[idx2] = 1;
    }
    // one direction: the MapPoints of `src` (camera-from-world Tsw) through Tds into keyframe `dst`
    auto direction = [&](const std::vector<MapPoint*>& src, const std::vector<char>& done, const orbfe_shim::Rt& Tsw,
                         const orbfe_shim::Rt& Tds, KeyFrame* dst, std::vector<int>& match) {
        std::vector<orbfe_proj_query> windows;
        std::vector<int> owner;
        for (int i = 0; i < (int)src.size(); ++i) {
            MapPoint* pMP = src[i];
            if (!pMP || done[i] || pMP->isBad()) continue;
            const orbfe_shim::Vec3 Xw(pMP->GetWorldPos());
            float Xs[3], Xd[3];
            Tsw.apply(Xw.v, Xs);
            Tds.apply(Xs, Xd);
            if (Xd[2] < 0.0) continue;
            const float invz = 1.0 / Xd[2];
            const float x = Xd[0] * invz, y = Xd[1] * invz;
            const float u = fx * x + cx, v = fy * y + cy;
            if (!dst->IsInImage(u, v)) continue;
            const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
            const float dist3D = orbfe_shim::norm3(Xd);
            if (dist3D < minDistance || dist3D > maxDistance) continue;
            const int level = pMP->PredictScale(dist3D, dst);
            orbfe_proj_query q = orbfe_shim::make_window(u, v, th * dst->mvScaleFactors[level], level - 1, level);
            orbfe_shim::set_descriptor(q, pMP->GetDescriptor());     // obs = 0: nothing is claimed, every point searches alone
            windows.push_back(q);
            owner.push_back(i);
        }
        orbfe_shim::GridView G(*dst, -1);
        G.g.blocked = nullptr;
        G.g.u_right = nullptr;
        std::vector<int32_t> assign, choice;
        SearchByProjectionWindows(G.g, windows, /*mode*/ 0, TH_HIGH, false, false, assign, &choice);
        for (size_t k = 0; k < owner.size() && k < choice.size(); ++k) match[owner[k]] = choice[k];
    };
    std::vector<int> match1(N1, -1), match2(N2, -1);
    direction(points1, matched1, T1w, T21, pKF2, match1);
    direction(points2, matched2, T2w, T12, pKF1, match2);
    int nFound = 0;                               // keep what both directions agree on
    for (int i1 = 0; i1 < N1; ++i1) {
        const int i2 = match1[i1];
        if (i2 >= 0 && match2[i2] == i1) {
            vpMatches12[i1] = points2[i2];
            ++nFound;
        }
    }
    return nFound;
}

// ---- Fuse: duplicated MapPoints (ORBmatcher.cc:1629-1821, 1823-1945) -------------------------------------------------------
// The searches do not depend on one another (a Fuse window never hides a feature from a later one), so all of them run in
// one launch; the map surgery then happens here, one point after the other, in the reference's order — including the
// tests that can change while it runs (a point replaced a moment ago is bad by the time its own turn comes).
inline int ORBmatcher::Fuse(KeyFrame* pKF, const std::vector<MapPoint*>& vpMapPoints, const float th, const bool bRight) {
    if (!m_) return 0;
    const orbfe_shim::Rt Tcw(bRight ? pKF->GetRightRotation() : pKF->GetRotation(), bRight ? pKF->GetRightTranslation() : pKF->GetTranslation());
    const orbfe_shim::Vec3 Ow(bRight ? pKF->GetRightCameraCenter() : pKF->GetCameraCenter());
    GeometricCamera* camera = bRight ? pKF->mpCamera2 : pKF->mpCamera;
    const float bf = pKF->mbf;
    std::vector<orbfe_proj_query> windows;
    std::vector<int> owner;
    for (int i = 0, n = (int)vpMapPoints.size(); i < n; ++i) {
        MapPoint* pMP = vpMapPoints[i];
        if (!pMP || pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
        const orbfe_shim::Vec3 Xw(pMP->GetWorldPos());
        float Xc[3];
        Tcw.apply(Xw.v, Xc);
        if (Xc[2] < 0.0f) continue;
        const float invz = 1 / Xc[2];
        const cv::Point2f uv = camera->project(cv::Point3f(Xc[0], Xc[1], Xc[2]));
        if (!pKF->IsInImage(uv.x, uv.y)) continue;
        const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
        const float PO[3] = {Xw.v[0] - Ow.v[0], Xw.v[1] - Ow.v[1], Xw.v[2] - Ow.v[2]};
        const float dist3D = orbfe_shim::norm3(PO);
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        const orbfe_shim::Vec3 Pn(pMP->GetNormal());
        if (orbfe_shim::dot3(PO, Pn.v) < 0.5 * dist3D) continue;
        const int level = pMP->PredictScale(dist3D, pKF);
        orbfe_proj_query q = orbfe_shim::make_window(uv.x, uv.y, th * pKF->mvScaleFactors[level], level - 1, level);
        q.ur = uv.x - bf * invz;
        q.chi2 = 1;                               // the reprojection gate of :1744-1770 runs per candidate on the device
        q.right = bRight;
        orbfe_shim::set_descriptor(q, pMP->GetDescriptor());
        windows.push_back(q);
        owner.push_back(i);
    }
    orbfe_shim::GridView G(*pKF, pKF->NLeft);
    G.g.blocked = nullptr;
    G.g.inv_level_sigma2 = pKF->mvInvLevelSigma2.data();
    G.g.n_levels = (int)pKF->mvInvLevelSigma2.size();
    std::vector<int32_t> assign, choice;
    SearchByProjectionWindows(G.g, windows, /*mode*/ 0, TH_LOW, false, false, assign, &choice);
    int nFused = 0;
    for (size_t k = 0; k < owner.size() && k < choice.size(); ++k) {
        MapPoint* pMP = vpMapPoints[owner[k]];
        if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;        // changed by an earlier point of this very call
        const int bestIdx = choice[k];
        if (bestIdx < 0) continue;
        MapPoint* inKF = pKF->GetMapPoint(bestIdx);
        if (inKF) {
            if (!inKF->isBad()) {
                if (inKF->Observations() > pMP->Observations()) pMP->Replace(inKF);
                else inKF->Replace(pMP);
            }
        } else {
            pMP->AddObservation(pKF, bestIdx);
            pKF->AddMapPoint(pMP, bestIdx);
        }
        ++nFused;
    }
    return nFused;
}

inline int ORBmatcher::Fuse(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, float th, std::vector<MapPoint*>& vpReplacePoint) {
    if (!m_) return 0;
    const orbfe_shim::Sim3Pose S(Scw);
    const std::set<MapPoint*> already = pKF->GetMapPoints();
    std::vector<orbfe_proj_query> windows;
    std::vector<int> owner;
    for (int i = 0, n = (int)vpPoints.size(); i < n; ++i) {
        MapPoint* pMP = vpPoints[i];
        if (pMP->isBad() || already.count(pMP)) continue;
        const orbfe_shim::Vec3 Xw(pMP->GetWorldPos());
        float Xc[3];
        S.cw.apply(Xw.v, Xc);
        if (Xc[2] < 0.0f) continue;
        const cv::Point2f uv = pKF->mpCamera->project(cv::Point3f(Xc[0], Xc[1], Xc[2]));
        if (!pKF->IsInImage(uv.x, uv.y)) continue;
        const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
        const float PO[3] = {Xw.v[0] - S.Ow.v[0], Xw.v[1] - S.Ow.v[1], Xw.v[2] - S.Ow.v[2]};
        const float dist3D = orbfe_shim::norm3(PO);
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        const orbfe_shim::Vec3 Pn(pMP->GetNormal());
        if (orbfe_shim::dot3(PO, Pn.v) < 0.5 * dist3D) continue;
        const int level = pMP->PredictScale(dist3D, pKF);
        orbfe_proj_query q = orbfe_shim::make_window(uv.x, uv.y, th * pKF->mvScaleFactors[level], level - 1, level);
        orbfe_shim::set_descriptor(q, pMP->GetDescriptor());
        windows.push_back(q);
        owner.push_back(i);
    }
    orbfe_shim::GridView G(*pKF, -1);
    G.g.blocked = nullptr;
    G.g.u_right = nullptr;
    std::vector<int32_t> assign, choice;
    SearchByProjectionWindows(G.g, windows, /*mode*/ 0, TH_LOW, false, false, assign, &choice);
    int nFused = 0;
    for (size_t k = 0; k < owner.size() && k < choice.size(); ++k) {
        const int bestIdx = choice[k];
        if (bestIdx < 0) continue;
        MapPoint* pMP = vpPoints[owner[k]];
        MapPoint* inKF = pKF->GetMapPoint(bestIdx);
        if (inKF) {
            if (!inKF->isBad()) vpReplacePoint[owner[k]] = inKF;
        } else {
            pMP->AddObservation(pKF, bestIdx);
            pKF->AddMapPoint(pMP, bestIdx);
        }
        ++nFused;
    }
    return nFused;
}

// ---- ORBmatcher.cc:2515-2556: the three fullest bins of a rotation histogram, second / third dropped below 10 % of the first ---
inline void ORBmatcher::ComputeThreeMaxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3) {
    int top[3] = {0, 0, 0};
    int at[3] = {ind1, ind2, ind3};            // callers pass -1; an all-empty histogram leaves them untouched
    for (int bin = 0; bin < L; ++bin) {
        const int count = (int)histo[bin].size();
        int rank = 3;                             // strictly-greater insertion keeps the earliest bin among equals
        while (rank > 0 && count > top[rank - 1]) --rank;
        if (rank == 3) continue;
        for (int k = 2; k > rank; --k) { top[k] = top[k - 1]; at[k] = at[k - 1]; }
        top[rank] = count;
        at[rank] = bin;
    }
    if (top[1] < 0.1f * (float)top[0]) { at[1] = -1; at[2] = -1; }
    else if (top[2] < 0.1f * (float)top[0]) at[2] = -1;
    ind1 = at[0]; ind2 = at[1]; ind3 = at[2];
}

}  // namespace ORB_SLAM3

#endif  // ORBFE_SHIM_ORBMATCHER_OBJECTS_H
