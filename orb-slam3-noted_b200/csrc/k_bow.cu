// k_bow.cu — vocabulary-node-restricted matchers: ORBmatcher::SearchByBoW (ORBmatcher.cc:269-471 and
// :823-963), ORBmatcher::SearchForTriangulation_ (:1208-1432) and the rotation-consistency filter
// they share (:384-399, :450-468, ComputeThreeMaxima :2515-2556).
//
// The host flattens each DBoW2::FeatureVector to CSR and merge-joins the two node lists (the
// while / lower_bound walk of :289-429) into work items (problem, A-range, B-range).  A feature
// belongs to exactly one node, so items are independent; one warp takes one item.
//   SearchByBoW: inside a node the reference is sequential over the A features because a B
//   feature claimed by an earlier A is skipped (:324,:345 / :884,:911).  The warp keeps that order:
//   for each A in turn the 32 lanes scan the node's unclaimed B features, each lane holding its
//   two smallest keys  dist << 16 | position ; a 5-step butterfly merges them into the warp's best
//   and second best (strict '<' in list order = smallest position among equal distances); lane 0
//   applies TH_LOW and the ratio test and publishes the claim before the next A.
//   SearchForTriangulation: no claims are ever made (vbMatched2 is never set), every A is
//   independent, '<=' with in-order replacement means the LAST minimal candidate that passes the
//   epipolar test wins -> key  dist << 16 | (0xFFFF - position) , warp min.
//   k_match_finish: one CTA per problem builds the 30-bin rotation histogram from the surviving
//   matches (a match's bin is a function of the two angles, so no per-bin lists are needed),
//   picks the three maxima like the reference and clears everything else.
#include "orbfe_internal.h"
#include "orbfe_match.h"
#include "kb8_device.cuh"

#define BOW_WARPS 4

__device__ __forceinline__ int rot_bin(float a1, float a2) {
    const float factor = 1.0f / ORBFE_HISTO_LENGTH;
    float rot = __fsub_rn(a1, a2);
    if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
    int bin = (int)roundf(__fmul_rn(rot, factor));
    return bin == ORBFE_HISTO_LENGTH ? 0 : bin;
}

__device__ __forceinline__ void merge2(uint32_t& k1, uint32_t& k2, uint32_t o1, uint32_t o2) {
    const uint32_t lo = min(k1, o1), hi = max(k1, o1);
    k2 = min(hi, min(k2, o2));
    k1 = lo;
}

__device__ __forceinline__ void warp_min2(uint32_t& k1, uint32_t& k2) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const uint32_t o1 = __shfl_xor_sync(0xffffffffu, k1, o), o2 = __shfl_xor_sync(0xffffffffu, k2, o);
        merge2(k1, k2, o1, o2);
    }
}

__device__ __forceinline__ int ham(const uint4 a0, const uint4 a1, const uint8_t* __restrict__ d) {
    const uint4 b0 = __ldg(reinterpret_cast<const uint4*>(d)), b1 = __ldg(reinterpret_cast<const uint4*>(d) + 1);
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x) +
           __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__global__ void __launch_bounds__(BOW_WARPS * 32) k_bow_nodes(const MatchProblem* __restrict__ probs,
                                                              const MatchItem* __restrict__ items, int nitems, float nnRatio,
                                                              int th, int strictTh, int nleftB) {
    const int item = blockIdx.x * BOW_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (item >= nitems) return;
    const MatchItem it = items[item];
    const MatchProblem P = probs[it.prob];
    volatile int32_t* b2a = P.b2a;
    const uint32_t NONE = (256u << 16) | 0xFFFFu;
    for (int ia = it.a0; ia < it.a1; ++ia) {
        const int idxA = P.idxA[ia];
        if (!P.maskA[idxA]) continue;
        const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(P.descA + 32ll * idxA));
        const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(P.descA + 32ll * idxA) + 1);
        uint32_t k1 = NONE, k2 = NONE, r1 = NONE, r2 = NONE;
        for (int jb = it.b0 + lane; jb < it.b1; jb += 32) {
            const int idxB = P.idxB[jb];
            if (b2a[idxB] >= 0) continue;
            if (P.maskB && !P.maskB[idxB]) continue;
            const uint32_t key = (uint32_t)ham(a0, a1, P.descB + 32ll * idxB) << 16 | (uint32_t)(jb - it.b0);
            if (nleftB < 0 || idxB < nleftB) merge2(k1, k2, key, NONE);
            else merge2(r1, r2, key, NONE);
        }
        warp_min2(k1, k2);
        if (nleftB >= 0) warp_min2(r1, r2);
        if (lane == 0) {
            const int d1 = (int)(k1 >> 16), d2 = (int)(k2 >> 16);
            if (strictTh ? d1 < th : d1 <= th) {
                if ((float)d1 < __fmul_rn(nnRatio, (float)d2)) {
                    const int bi = P.idxB[it.b0 + (int)(k1 & 0xFFFF)];
                    b2a[bi] = idxA;
                    P.a2b[idxA] = bi;
                }
                if (nleftB >= 0 && (int)(r1 >> 16) <= th) {
                    const int bi = P.idxB[it.b0 + (int)(r1 & 0xFFFF)];
                    b2a[bi] = idxA;
                    if (P.a2b[idxA] < 0) P.a2b[idxA] = bi;
                }
            }
        }
        __syncwarp();
    }
}

__global__ void __launch_bounds__(BOW_WARPS * 32) k_tri_nodes(const MatchProblem* __restrict__ probs,
                                                              const MatchItem* __restrict__ items, int nitems, TriGeo geo,
                                                              int onlyStereo, int coarse) {
    const int item = blockIdx.x * BOW_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (item >= nitems) return;
    const MatchItem it = items[item];
    const MatchProblem P = probs[it.prob];
    for (int ia = it.a0; ia < it.a1; ++ia) {
        const int idx1 = P.idxA[ia];
        if (P.maskA[idx1]) continue;                           // already has a MapPoint (:1279-1282)
        const bool st1 = P.uRightA && P.uRightA[idx1] >= 0.f;
        if (onlyStereo && !st1) continue;
        const float4 k1p = P.kpA[idx1];                        // x, y, angle, octave
        const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(P.descA + 32ll * idx1));
        const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(P.descA + 32ll * idx1) + 1);
        // epipolar line of kp1 in image 2 (Pinhole.cpp:167-169), unfused float like the scalar reference
        const float la = __fadd_rn(__fadd_rn(__fmul_rn(k1p.x, geo.F[0]), __fmul_rn(k1p.y, geo.F[3])), geo.F[6]);
        const float lb = __fadd_rn(__fadd_rn(__fmul_rn(k1p.x, geo.F[1]), __fmul_rn(k1p.y, geo.F[4])), geo.F[7]);
        const float lc = __fadd_rn(__fadd_rn(__fmul_rn(k1p.x, geo.F[2]), __fmul_rn(k1p.y, geo.F[5])), geo.F[8]);
        const float den = __fadd_rn(__fmul_rn(la, la), __fmul_rn(lb, lb));
        uint32_t best = 0xFFFFFFFFu;
        for (int jb = it.b0 + lane; jb < it.b1; jb += 32) {
            const int idx2 = P.idxB[jb];
            if (P.maskB[idx2]) continue;
            const bool st2 = P.uRightB && P.uRightB[idx2] >= 0.f;
            if (onlyStereo && !st2) continue;
            const int dist = ham(a0, a1, P.descB + 32ll * idx2);
            if (dist > ORBFE_TH_LOW) continue;
            const float4 k2p = P.kpB[idx2];
            const int oct2 = __float_as_int(k2p.w);
            if (!st1 && !st2) {
                const float dx = __fsub_rn(geo.ex, k2p.x), dy = __fsub_rn(geo.ey, k2p.y);
                if (__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < __fmul_rn(100.f, geo.scaleB[oct2])) continue;
            }
            bool ok = coarse != 0;
            if (!ok && den != 0.f) {
                const float num = __fadd_rn(__fadd_rn(__fmul_rn(la, k2p.x), __fmul_rn(lb, k2p.y)), lc);
                const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
                ok = (double)dsqr < 3.84 * (double)geo.sigma2B[oct2];
            }
            if (ok) best = min(best, (uint32_t)dist << 16 | (uint32_t)(0xFFFF - (jb - it.b0)));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
        if (lane == 0 && best != 0xFFFFFFFFu) P.a2b[idx1] = P.idxB[it.b0 + (0xFFFF - (int)(best & 0xFFFF))];
    }
}

// The same search with Kannala-Brandt cameras (ORBmatcher.cc:1208-1432 + KannalaBrandt8::epipolarConstrain_,
// KannalaBrandt8.cpp:239-242): a candidate passes iff TriangulateMatches_ returns a depth > 1e-4 for the
// camera pairing (left/right of each keyframe) the two features belong to.  Fisheye keyframes carry no
// mvuRight, so bStereo1/2 are false; the epipole test applies iff keyframe 1 has a single camera.
__global__ void __launch_bounds__(BOW_WARPS * 32) k_tri_nodes_kb8(const MatchProblem* __restrict__ probs,
                                                                  const MatchItem* __restrict__ items, int nitems,
                                                                  const __grid_constant__ TriKb8Geo geo, int coarse,
                                                                  float* __restrict__ points) {
    const int item = blockIdx.x * BOW_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (item >= nitems) return;
    const MatchItem it = items[item];
    const MatchProblem P = probs[it.prob];
    const bool dual = geo.nleftA >= 0 && geo.nleftB >= 0;
    for (int ia = it.a0; ia < it.a1; ++ia) {
        const int idx1 = P.idxA[ia];
        if (P.maskA[idx1]) continue;
        const float4 k1p = P.kpA[idx1];
        const int oct1 = __float_as_int(k1p.w);
        const int right1 = !(geo.nleftA == -1 || idx1 < geo.nleftA);
        const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(P.descA + 32ll * idx1));
        const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(P.descA + 32ll * idx1) + 1);
        uint32_t best = 0xFFFFFFFFu, mine = 0xFFFFFFFFu;
        float bp[3] = {0.f, 0.f, 0.f};
        for (int jb = it.b0 + lane; jb < it.b1; jb += 32) {
            const int idx2 = P.idxB[jb];
            if (P.maskB[idx2]) continue;
            const int dist = ham(a0, a1, P.descB + 32ll * idx2);
            if (dist > ORBFE_TH_LOW) continue;
            const uint32_t key = (uint32_t)dist << 16 | (uint32_t)(0xFFFF - (jb - it.b0));
            if (key > mine) continue;                           // cannot beat this lane's best: skip the geometry
            const float4 k2p = P.kpB[idx2];
            const int oct2 = __float_as_int(k2p.w);
            const int right2 = !(geo.nleftB == -1 || idx2 < geo.nleftB);
            if (geo.nleftA < 0) {
                const float dx = __fsub_rn(geo.ex, k2p.x), dy = __fsub_rn(geo.ey, k2p.y);
                if (__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < __fmul_rn(100.f, geo.scaleB[oct2])) continue;
            }
            const int c = dual ? 2 * right1 + right2 : 0;
            float p[3] = {0.f, 0.f, 0.f};
            bool ok = coarse != 0;
            if (!ok)
                ok = kb8_triangulate_matches(geo.camA[dual ? right1 : 0], geo.camB[dual ? right2 : 0], geo.R12[c], geo.t12[c], k1p.x,
                                             k1p.y, k2p.x, k2p.y, geo.sigma2A[oct1], geo.sigma2B[oct2], p) > 0.0001f;
            if (ok) { mine = key; bp[0] = p[0]; bp[1] = p[1]; bp[2] = p[2]; }
        }
        best = mine;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
        if (best != 0xFFFFFFFFu && mine == best) {              // keys are unique (they carry the position)
            P.a2b[idx1] = P.idxB[it.b0 + (0xFFFF - (int)(best & 0xFFFF))];
            if (points) { points[3 * idx1] = bp[0]; points[3 * idx1 + 1] = bp[1]; points[3 * idx1 + 2] = bp[2]; }
        }
    }
}

// One CTA per problem.  overB != 0: matches are the entries of b2a (SearchByBoW); else of a2b
// (SearchForTriangulation's vMatches12).  rot = angle(A) - angle(B) in both cases.
__global__ void __launch_bounds__(256) k_match_finish(const MatchProblem* __restrict__ probs, int overB, int checkOri,
                                                      int* __restrict__ nmatches) {
    __shared__ int hist[ORBFE_HISTO_LENGTH];
    __shared__ int keep[3];
    __shared__ int total;
    const MatchProblem P = probs[blockIdx.x];
    const int tid = threadIdx.x;
    const int n = overB ? P.nB : P.nA;
    int32_t* m = overB ? P.b2a : P.a2b;
    if (tid < ORBFE_HISTO_LENGTH) hist[tid] = 0;
    if (tid == 0) total = 0;
    __syncthreads();
    if (!checkOri) {
        int c = 0;
        for (int i = tid; i < n; i += 256) c += m[i] >= 0;
        atomicAdd(&total, c);
        __syncthreads();
        if (tid == 0) nmatches[blockIdx.x] = total;
        return;
    }
    for (int i = tid; i < n; i += 256) {
        const int o = m[i];
        if (o < 0) continue;
        const float angA = overB ? P.kpA[o].z : P.kpA[i].z, angB = overB ? P.kpB[i].z : P.kpB[o].z;
        atomicAdd(&hist[rot_bin(angA, angB)], 1);
    }
    __syncthreads();
    if (tid == 0) {
        int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < ORBFE_HISTO_LENGTH; i++) {
            const int s = hist[i];
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
            else if (s > max3) { max3 = s; ind3 = i; }
        }
        if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
        else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { ind3 = -1; }
        keep[0] = ind1; keep[1] = ind2; keep[2] = ind3;
    }
    __syncthreads();
    int c = 0;
    for (int i = tid; i < n; i += 256) {
        const int o = m[i];
        if (o < 0) continue;
        const float angA = overB ? P.kpA[o].z : P.kpA[i].z, angB = overB ? P.kpB[i].z : P.kpB[o].z;
        const int bin = rot_bin(angA, angB);
        if (bin == keep[0] || bin == keep[1] || bin == keep[2]) { ++c; continue; }
        m[i] = -1;
        if (overB && P.a2b[o] == i) P.a2b[o] = -1;
    }
    atomicAdd(&total, c);
    __syncthreads();
    if (tid == 0) nmatches[blockIdx.x] = total;
}

cudaError_t orbfe_launch_bow(cudaStream_t st, const MatchProblem* dProbs, int nprob, const MatchItem* dItems, int nitems,
                             float nnRatio, int th, int strictTh, int nleftB, int checkOri, int* dCounts, long long* launches) {
    if (nitems > 0) {
        k_bow_nodes<<<(nitems + BOW_WARPS - 1) / BOW_WARPS, BOW_WARPS * 32, 0, st>>>(dProbs, dItems, nitems, nnRatio, th, strictTh,
                                                                                     nleftB);
        ++*launches;
    }
    k_match_finish<<<nprob, 256, 0, st>>>(dProbs, 1, checkOri, dCounts);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t orbfe_launch_tri_kb8(cudaStream_t st, const MatchProblem* dProbs, const MatchItem* dItems, int nitems,
                                 const TriKb8Geo& geo, int coarse, int checkOri, float* dPoints, int* dCounts, long long* launches) {
    if (nitems > 0) {
        k_tri_nodes_kb8<<<(nitems + BOW_WARPS - 1) / BOW_WARPS, BOW_WARPS * 32, 0, st>>>(dProbs, dItems, nitems, geo, coarse, dPoints);
        ++*launches;
    }
    k_match_finish<<<1, 256, 0, st>>>(dProbs, 0, checkOri, dCounts);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t orbfe_launch_tri(cudaStream_t st, const MatchProblem* dProbs, const MatchItem* dItems, int nitems, const TriGeo& geo,
                             int onlyStereo, int coarse, int checkOri, int* dCounts, long long* launches) {
    if (nitems > 0) {
        k_tri_nodes<<<(nitems + BOW_WARPS - 1) / BOW_WARPS, BOW_WARPS * 32, 0, st>>>(dProbs, dItems, nitems, geo, onlyStereo, coarse);
        ++*launches;
    }
    k_match_finish<<<1, 256, 0, st>>>(dProbs, 0, checkOri, dCounts);
    ++*launches;
    return cudaGetLastError();
}
