// k_kb8.cu — the geometric half of Frame::ComputeStereoFishEyeMatches (Frame.cc:1141-1153):
// KannalaBrandt8::TriangulateMatches (src/CameraModels/KannalaBrandt8.cpp:338-407) for every lapping-area
// pair that survived the 2-NN + Lowe-ratio test (k_match.cu).  One thread per left feature:
// unproject both keypoints (Newton iteration of :96-123), parallax test, linear triangulation,
// cheirality and the two reprojection tests.  Triangulate (:514-526) takes the right singular vector of
// the smallest singular value of a 4x4 system with cv::SVD; here it is the eigenvector of A^T A by cyclic
// Jacobi rotations in double precision (B200 keeps full-rate FP64) — float-tolerance parity, as the oracle.
// Several left features may pick the same right feature; the reference overwrites mvRightToLeftMatch in
// query order, i.e. the largest left index wins -> atomicMax.
#include "orbfe_internal.h"
#include "orbfe_match.h"
#include "kb8_device.cuh"

// match[q]: lapping-area train index of query q (or -1).  One thread per query.
__global__ void __launch_bounds__(128) k_fisheye_triangulate(Kb8Geo G, const orbfe_keypoint* __restrict__ kpL,
                                                             const orbfe_keypoint* __restrict__ kpR, int monoLeft, int monoRight,
                                                             int nq, const int32_t* __restrict__ match, int32_t* __restrict__ l2r,
                                                             int32_t* __restrict__ r2l, float* __restrict__ depth,
                                                             float* __restrict__ p3d, int* __restrict__ count) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    const int tr = match[q];
    if (tr < 0) return;
    const int iL = q + monoLeft, iR = tr + monoRight;
    const orbfe_keypoint a = kpL[iL], b = kpR[iR];
    float p[3];
    const float d = kb8_triangulate_matches(G.cam1, G.cam2, G.R12, G.t12, a.x, a.y, b.x, b.y, G.sigma2[a.octave], G.sigma2[b.octave], p);
    if (d > 0.0001f) {
        l2r[iL] = iR;
        atomicMax(&r2l[iR], iL);
        depth[iL] = d;
        p3d[3 * iL] = p[0]; p3d[3 * iL + 1] = p[1]; p3d[3 * iL + 2] = p[2];
        atomicAdd(count, 1);
    }
}

cudaError_t orbfe_launch_fisheye_triangulate(cudaStream_t st, const Kb8Geo& G, const orbfe_keypoint* kpL, const orbfe_keypoint* kpR,
                                             int monoLeft, int monoRight, int nq, const int32_t* dMatch, int32_t* l2r, int32_t* r2l,
                                             float* depth, float* p3d, int* count, long long* launches) {
    if (nq <= 0) return cudaSuccess;
    k_fisheye_triangulate<<<(nq + 127) / 128, 128, 0, st>>>(G, kpL, kpR, monoLeft, monoRight, nq, dMatch, l2r, r2l, depth, p3d, count);
    ++*launches;
    return cudaGetLastError();
}
