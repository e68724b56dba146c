// match_ref.h — CPU restatement of the reference's Hamming matchers (TEST INFRASTRUCTURE).
//
// Follows /root/reference/src/ORBmatcher.cc (DescriptorDistance :2561-2577, ComputeThreeMaxima
// :2515-2556, SearchByBoW(KF,Frame) :269-471, SearchByBoW(KF,KF) :823-963, SearchForTriangulation_
// :1208-1432) and /root/reference/src/Frame.cc (ComputeStereoMatches :794-964,
// ComputeStereoFishEyeMatches :1116-1141 up to the Lowe-ratio test), with Pinhole::epipolarConstrain_
// (src/CameraModels/Pinhole.cpp:159-181).  Objects the reference reaches through KeyFrame / Frame /
// MapPoint pointers are flattened to arrays here exactly as at the C ABI (include/orbfe.h): a
// DBoW2::FeatureVector is a CSR (node ids ascending, feature indices per node), "has a good
// MapPoint" is a byte mask.  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs use it.
// Parity pin: tests/golden/match_*.npz, produced by oracle/py/match_py.py (a line-by-line Python
// restatement that calls cv2.BFMatcher / cv2.norm for the OpenCV parts).
#pragma once
#include <cmath>
#include <cstdint>
#include <utility>
#include <vector>

#include "orb_ref.h"

namespace orbref {

static const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;   // ORBmatcher.cc:36-38

// Rotation-histogram bin of ORBmatcher.cc:390-396 (factor = 1.0f/HISTO_LENGTH, so bins are 30 deg wide).
inline int rotBin(float angle1, float angle2) {
    const float factor = 1.0f / HISTO_LENGTH;
    float rot = angle1 - angle2;
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)std::round(rot * factor);
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
}

// The SWAR popcount of ORBmatcher.cc:2561-2577, verbatim arithmetic.
int descriptorDistanceSwar(const uint8_t* a, const uint8_t* b);

// ORBmatcher.cc:2515-2556.  counts[L] = histo[i].size().
void computeThreeMaxima(const int* counts, int L, int& ind1, int& ind2, int& ind3);

// DBoW2::FeatureVector flattened: nodeIds ascending, features of node k = indices[offsets[k] .. offsets[k+1]).
struct FeatureVectorCsr {
    int nnodes;
    const uint32_t* nodeIds;
    const int32_t* offsets;
    const int32_t* indices;
};

// Frame::ComputeStereoMatches (Frame.cc:794-964).  uRight / depth: nL floats, -1 where unmatched.
// bestDistOut (nullable): the SAD of accepted matches, -1 elsewhere (debug tap).
void computeStereoMatches(const OrbExtractorRef& left, const OrbExtractorRef& right, const KeyPoint* kpL,
                          const uint8_t* descL, int nL, const KeyPoint* kpR, const uint8_t* descR, int nR, float mbf,
                          float mb, float* uRight, float* depth, int32_t* bestDistOut);

struct BowParams {
    float nnRatio;      // mfNNratio
    int checkOri;       // mbCheckOrientation
    int th;             // TH_LOW
    int strictTh;       // 0: bestDist1 <= th (KF,Frame :373); 1: bestDist1 < th (KF,KF :904)
    int nleftB;         // F.Nleft (-1: no second camera); >= 0 selects the left/right split of :330-368
};
// SearchByBoW, both overloads.  Side A iterates in the outer loop (the KeyFrame / pKF1), side B is
// searched and claimed (the Frame / pKF2).  maskA[i] != 0: feature i of A has a good MapPoint
// (required); maskB may be null (KF,Frame form: every Frame feature is eligible) or gives pKF2's
// MapPoint validity.  matchBtoA[nB]: index in A claimed for each B feature (-1 none) — this is
// vpMapPointMatches of :269; matchAtoB[nA]: vpMatches12 of :823 (left-camera match when nleftB >= 0).
int searchByBoW(const uint8_t* descA, int nA, const FeatureVectorCsr& fvA, const uint8_t* maskA, const float* angleA,
                const uint8_t* descB, int nB, const FeatureVectorCsr& fvB, const uint8_t* maskB, const float* angleB,
                const BowParams& p, int32_t* matchAtoB, int32_t* matchBtoA);

struct TriParams {
    int checkOri;        // mbCheckOrientation
    int onlyStereo;      // bOnlyStereo
    int coarse;          // bCoarse: accept without the epipolar test (:1361)
    float F12[9];        // row-major fundamental matrix K1^-T [t12]x R12 K2^-1, as epipolarConstrain_ builds it
    float ex, ey;        // epipole of camera 1 in image 2 (:1218)
};
// SearchForTriangulation_ (pinhole cameras, no second camera).  hasMp*: feature already has a
// MapPoint (skipped); uRight*: mvuRight (>= 0 means stereo observation); kp*: undistorted keypoints;
// sigma2B / scaleB: pKF2->mvLevelSigma2 / mvScaleFactors.  Writes matches12[nA] (-1 none); the
// reference's vMatchedPairs is the ascending list of (i, matches12[i]) with matches12[i] >= 0.
int searchForTriangulation(const KeyPoint* kpA, const uint8_t* descA, int nA, const FeatureVectorCsr& fvA,
                           const uint8_t* hasMpA, const float* uRightA, const KeyPoint* kpB, const uint8_t* descB, int nB,
                           const FeatureVectorCsr& fvB, const uint8_t* hasMpB, const float* uRightB, const float* sigma2B,
                           const float* scaleB, const TriParams& p, int32_t* matches12);

// Frame::UndistortKeyPoints (src/Frame.cc:730-763): cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK) on
// every keypoint position: OpenCV's 5 fixed-point iterations in double (radial k1 k2 k3 + tangential p1 p2), back
// through K, stored as float.  dist = (k1, k2, p1, p2[, k3]); dist[0] == 0 copies the keypoints (:732-736).
void undistortKeyPoints(const KeyPoint* kps, int n, const float K4[4], const float* dist, int ndist, KeyPoint* out);

// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:349-423): among the n observed descriptors of a map
// point, the one with the least median Hamming distance to all of them (median = sorted row [0.5 * (n - 1)], the
// row includes the 0 to itself; the first index wins ties).  Returns the index, -1 for n == 0.
int distinctiveDescriptor(const uint8_t* desc, int n);

// Frame::ComputeStereoFishEyeMatches up to the descriptor test (Frame.cc:1134-1141): 2-NN brute force
// (BFMatcher NORM_HAMMING, ties to the lowest train index) + Lowe ratio d0 < 0.7 * d1.
// match[nq] = train index or -1.
void knn2RatioMatch(const uint8_t* q, int nq, const uint8_t* t, int nt, double ratio, int32_t* match);

}  // namespace orbref

// ---- Kannala-Brandt fisheye stereo (Frame::ComputeStereoFishEyeMatches, Frame.cc:1116-1156) ------------
namespace orbref {

struct Kb8 { float fx, fy, cx, cy, k0, k1, k2, k3; };   // mvParameters[0..7]

// KannalaBrandt8::project / unproject (KannalaBrandt8.cpp:24-40, 96-123), float32 like the reference.
void kb8Project(const Kb8& c, const float p[3], float uv[2]);
void kb8Unproject(const Kb8& c, float u, float v, float ray[3]);
// KannalaBrandt8::TriangulateMatches (KannalaBrandt8.cpp:338-407): returns z1 (depth in camera 1) or -1.
// The 4x4 cv::SVD of Triangulate (:514-526) is replaced by a double-precision Jacobi eigen-solve of A^T A:
// float-tolerance parity only.  margin (nullable) receives the smallest relative distance
// of any of the five accept tests to its threshold, so tests can skip knife-edge cases.
float kb8TriangulateMatches(const Kb8& c1, const Kb8& c2, const KeyPoint& kp1, const KeyPoint& kp2, const float R12[9],
                            const float t12[3], float sigmaLevel, float unc, float p3D[3], float* margin);
// The whole function: lapping-area 2-NN + Lowe ratio 0.7 + triangulation; accept depth > 0.0001.
// leftToRight[nL], rightToLeft[nR] (-1 none), depth[nL] (-1 none), p3d[nL*3].
int computeStereoFishEyeMatches(const Kb8& c1, const Kb8& c2, const KeyPoint* kpL, const uint8_t* descL, int nL, int monoLeft,
                                const KeyPoint* kpR, const uint8_t* descR, int nR, int monoRight, const float R12[9],
                                const float t12[3], const float* sigma2, int32_t* leftToRight, int32_t* rightToLeft,
                                float* depth, float* p3d, float* margins);

// SearchForTriangulation / SearchForTriangulation_ with Kannala-Brandt cameras (ORBmatcher.cc:965-1432 with
// KannalaBrandt8::epipolarConstrain_, KannalaBrandt8.cpp:239-242: a pair passes iff TriangulateMatches_ > 1e-4).
// nleft* = KeyFrame::NLeft (-1: single camera; else features [0,nleft) come from mpCamera, the rest from
// mpCamera2).  cam*[0] = mpCamera, cam*[1] = mpCamera2.  R12 / t12 [4] = {ll, lr, rl, rr} (:1248-1262),
// indexed 2*bRight1 + bRight2; entry 0 alone is used unless both keyframes have a second camera (:1332).
// mvuRight is never set with a fisheye camera, so bStereo1/2 are false: onlyStereo yields no matches, the
// epipole test (:1321-1327) applies iff pKF1 has no second camera.  points12 (nullable, nA*3): the
// triangulated point (camera-1 frame) of each accepted match; margins (nullable, nA): smallest threshold
// margin over every candidate pair of feature i that reached the geometric test.
struct TriKb8Params {
    int checkOri, onlyStereo, coarse;
    Kb8 camA[2], camB[2];
    int nleftA, nleftB;
    float R12[4][9], t12[4][3];
    float ex, ey;
};
int searchForTriangulationKb8(const KeyPoint* kpA, const uint8_t* descA, int nA, const FeatureVectorCsr& fvA,
                              const uint8_t* hasMpA, const float* sigma2A, const KeyPoint* kpB, const uint8_t* descB, int nB,
                              const FeatureVectorCsr& fvB, const uint8_t* hasMpB, const float* sigma2B, const float* scaleB,
                              const TriKb8Params& p, int32_t* matches12, float* points12, float* margins);

}  // namespace orbref
