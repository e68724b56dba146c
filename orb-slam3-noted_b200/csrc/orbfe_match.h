// orbfe_match.h — device-side problem descriptors and launchers of the matcher kernels
// (k_match.cu, k_stereo.cu, k_bow.cu), shared with the matcher C ABI (orbfe_match_api.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "orbfe_internal.h"

// One BoW / triangulation problem: all pointers are device pointers into the matcher's staging arena.
struct MatchProblem {
    const uint8_t* descA;     // nA x 32
    const uint8_t* descB;     // nB x 32
    const int32_t* idxA;      // CSR indices of A (all nodes, concatenated)
    const int32_t* idxB;
    const uint8_t* maskA;     // SearchByBoW: has a good MapPoint; SearchForTriangulation: already has a MapPoint
    const uint8_t* maskB;     // may be null in SearchByBoW form 0
    const float4* kpA;        // (x, y, angle, octave as int bits)
    const float4* kpB;
    const float* uRightA;     // triangulation only; may be null (all mono)
    const float* uRightB;
    int32_t* a2b;             // nA
    int32_t* b2a;             // nB
    int nA, nB;
};

// One vocabulary node present in both FeatureVectors of problem `prob`.
struct MatchItem {
    int prob;
    int a0, a1;               // range in idxA
    int b0, b1;               // range in idxB
};

struct TriGeo {
    float F[9];
    float ex, ey;
    float sigma2B[ORBFE_MAX_LEVELS], scaleB[ORBFE_MAX_LEVELS];
};

struct UndistortGeo { double fx, fy, cx, cy, k1, k2, p1, p2, k3; };
// brute-force 2-NN: the POPC kernel (k_match.cu) and the tensor-core kernel (k_knn2_tc.cu), same arguments and results
cudaError_t orbfe_launch_knn2_merge(cudaStream_t st, const uint2* dPartial, const int* dNq, int nqMax, int nsplit, int batch, int32_t* dIdx2,
                                    int32_t* dDist2, long long* launches);
void orbfe_knn2_tc_plan(int nqMax, int ntMax, int batch, int smCount, int* nsplit, int* splitLen);
cudaError_t orbfe_launch_knn2_tc(cudaStream_t st, const uint8_t* dq, long long qStride, const int* dNq, int nqMax, const uint8_t* dt,
                                 long long tStride, const int* dNt, int ntMax, int batch, int nsplit, int splitLen, uint2* dPartial,
                                 int32_t* dIdx2, int32_t* dDist2, long long* launches);
cudaError_t orbfe_launch_undistort(cudaStream_t st, orbfe_keypoint* dKps, int n, const UndistortGeo& g, long long* launches);
cudaError_t orbfe_launch_distinctive(cudaStream_t st, const uint8_t* dDesc, const int32_t* dOffsets, int npoints, int32_t* dBest,
                                     long long* launches);

// ---- projection-guided window searches (k_project.cu) ----
struct ProjQueryDev {          // = orbfe_proj_query (include/orbfe.h), 64 bytes
    float u, v, r, ur, angle;
    int16_t minLevel, maxLevel;
    uint8_t right, obs, checkUr, valid, skipIfPrevRatioRejected, chi2, skipIfPrevEmpty, pad;
    uint8_t desc[32];
};
static_assert(sizeof(ProjQueryDev) == 64, "orbfe_proj_query layout");

struct ProjProblem {
    const float4* kp;          // (x, y, angle, octave as int bits) per Frame feature
    const uint8_t* desc;
    const uint8_t* blocked;    // nullable
    const float* uRight;       // nullable
    const int32_t* l2r;        // nullable
    const int32_t* r2l;        // nullable
    int n, nleft;
    float minX, minY, gridWInv, gridHInv;
    const float* invSigma2;    // mvInvLevelSigma2 (windows with chi2; nullable)
    const ProjQueryDev* q;
    int nq;
    int32_t* assign;           // n
    int32_t* choice;           // nq
    int* nmatches;
    // device scratch
    uint32_t* sorted;          // n grid keys (side, cell) << 16 | index, ascending
    int* cellStart;            // 2 * 3072
    int* cellEnd;              // 2 * 3072
    unsigned long long* cand;  // nq * PROJ_K smallest candidate keys per window
    unsigned long long* limit; // nq: keys below it are a correct prefix of the window's sorted candidates
    int* scratch;              // n + 5 * nq ints
};
#define PROJ_K 8
cudaError_t orbfe_launch_search_by_projection(cudaStream_t st, const ProjProblem* dProbs, int nprob, int nMax, int nqMax, int mode,
                                              int th, float nnRatio, int linkStereo, int rotHist, long long* launches);

cudaError_t orbfe_launch_search_for_initialization(cudaStream_t st, const ProjProblem* dProbs, int nprob, int nMax, int nqMax, int th,
                                                   float nnRatio, int rotHist, long long* launches);

// SearchForTriangulation with Kannala-Brandt cameras: cameras [0] = mpCamera, [1] = mpCamera2 of each
// keyframe, relative poses {ll, lr, rl, rr} (x1 = R12 x2 + t12), NLeft of both keyframes (-1: one camera).
struct TriKb8Geo {
    float camA[2][8], camB[2][8];
    float R12[4][9], t12[4][3];
    float sigma2A[ORBFE_MAX_LEVELS], sigma2B[ORBFE_MAX_LEVELS], scaleB[ORBFE_MAX_LEVELS];
    float ex, ey;
    int nleftA, nleftB;
};

// Kannala-Brandt stereo geometry (cameras 1 = left, 2 = right; x1 = R12 x2 + t12).
struct Kb8Geo {
    float cam1[8], cam2[8];     // fx, fy, cx, cy, k0..k3 (KannalaBrandt8::mvParameters)
    float R12[9], t12[3];
    float sigma2[ORBFE_MAX_LEVELS];
};

cudaError_t orbfe_launch_fisheye_triangulate(cudaStream_t st, const Kb8Geo& G, const orbfe_keypoint* kpL, const orbfe_keypoint* kpR,
                                             int monoLeft, int monoRight, int nq, const int32_t* dMatch, int32_t* l2r, int32_t* r2l,
                                             float* depth, float* p3d, int* count, long long* launches);
cudaError_t orbfe_launch_tri_kb8(cudaStream_t st, const MatchProblem* dProbs, const MatchItem* dItems, int nitems,
                                 const TriKb8Geo& geo, int coarse, int checkOri, float* dPoints, int* dCounts, long long* launches);
void orbfe_knn2_plan(int nqMax, int ntMax, int batch, int smCount, int* nsplit, int* splitLen);
cudaError_t orbfe_launch_knn2(cudaStream_t st, const uint8_t* dq, long long qStride, const int* dNq, int nqMax, const uint8_t* dt,
                              long long tStride, const int* dNt, int ntMax, int batch, int nsplit, int splitLen, uint2* dPartial,
                              int32_t* dIdx2, int32_t* dDist2, long long* launches);
cudaError_t orbfe_launch_ratio_filter(cudaStream_t st, const int32_t* dIdx2, const int32_t* dDist2, int nq, double ratio,
                                      int32_t* dMatch, int* dCount, long long* launches);
cudaError_t orbfe_launch_stereo(cudaStream_t st, const orbfe_handle* hl, int frameL0, int frameLStep, const orbfe_handle* hr,
                                int frameR0, int frameRStep, int npairs, float bf, float b, float* dURight, float* dDepth,
                                int* dSad, int* dCount, int outStride, long long* launches);
cudaError_t orbfe_launch_bow(cudaStream_t st, const MatchProblem* dProbs, int nprob, const MatchItem* dItems, int nitems,
                             float nnRatio, int th, int strictTh, int nleftB, int checkOri, int* dCounts, long long* launches);
cudaError_t orbfe_launch_tri(cudaStream_t st, const MatchProblem* dProbs, const MatchItem* dItems, int nitems, const TriGeo& geo,
                             int onlyStereo, int coarse, int checkOri, int* dCounts, long long* launches);
