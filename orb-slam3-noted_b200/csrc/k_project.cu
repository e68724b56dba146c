// k_project.cu — the projection-guided window searches: Frame::GetFeaturesInArea (src/Frame.cc:640-706) over the
// 64 x 48 feature grid of Frame::AssignFeaturesToGrid (:380-411) + the matching loop shared by
// ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th, ...) (src/ORBmatcher.cc:44-236) and the best-only
// overloads (Frame&, const Frame& LastFrame, ...) :2173-2389, (Frame&, KeyFrame*, ...) :2391-2510.
//
// One CTA per problem (one Frame and its ordered list of windows):
//   1. grid: every feature's cell by PosInGrid's round() rule; a shared-memory bitonic sort of
//      (side, cell, index) keys gives each cell's features in ascending index = the reference's push order.
//   2. search: one warp per window.  Lanes take the window's cells in GetFeaturesInArea's order (x outer, y
//      inner); a candidate's key is dist << 40 | position of its cell << 16 | index, so the two smallest keys
//      are exactly the reference's best / second best under its strict '<' updates.
//   3. order: the reference assigns MapPoints one after another and a feature holding a MapPoint with
//      observations is skipped by every later window.  That recurrence (window i depends on the choices of
//      windows < i) is solved by Jacobi sweeps: all windows are re-evaluated in parallel against the claims of
//      the previous sweep, blockedAt[f] = smallest claiming window index, until nothing changes.  Sweep k makes
//      windows 0..k-1 final, so the fixed point is the sequential result; conflicts are rare and 2-4 sweeps do.
//   4. finish: last writer per feature (atomicMax), stereo links, rotation histogram, count.
#include "orbfe_internal.h"
#include "orbfe_match.h"

#define PROJ_T 512
#define GRID_COLS 64
#define GRID_ROWS 48
#define GRID_CELLS (GRID_COLS * GRID_ROWS)

__device__ __forceinline__ int rot_bin_p(float a1, float a2) {
    const float factor = 1.0f / ORBFE_HISTO_LENGTH;
    float rot = __fsub_rn(a1, a2);
    if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
    int bin = (int)roundf(__fmul_rn(rot, factor));
    return bin == ORBFE_HISTO_LENGTH ? 0 : bin;
}

__device__ __forceinline__ int ham_q(const uint4& a0, const uint4& a1, const uint8_t* __restrict__ b) {
    const uint4 b0 = __ldg(reinterpret_cast<const uint4*>(b)), b1 = __ldg(reinterpret_cast<const uint4*>(b) + 1);
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x) +
           __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__device__ __forceinline__ void top2_insert(unsigned long long& k1, unsigned long long& k2, unsigned long long k) {
    if (k < k1) { k2 = k1; k1 = k; }
    else if (k < k2) k2 = k;
}

// Result of one window against the current claims: chosen feature (or -1) and whether the ratio rule rejected it.
__device__ int2 search_window(const ProjProblem& P, int i, int lane, const uint32_t* __restrict__ sorted,
                              const int* __restrict__ cellStart, const int* __restrict__ cellEnd,
                              const int* __restrict__ blockedAt, int mode, int th, float nnRatio) {
    const ProjQueryDev& Q = P.q[i];
    const float x = Q.u, y = Q.v, r = Q.r;
    const int nMinCellX = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, P.minX), r), P.gridWInv)));
    const int nMaxCellX = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, P.minX), r), P.gridWInv)));
    const int nMinCellY = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, P.minY), r), P.gridHInv)));
    const int nMaxCellY = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, P.minY), r), P.gridHInv)));
    if (nMinCellX >= GRID_COLS || nMaxCellX < 0 || nMinCellY >= GRID_ROWS || nMaxCellY < 0 || nMinCellX > nMaxCellX ||
        nMinCellY > nMaxCellY)
        return make_int2(-1, 0);
    const int minLevel = Q.minLevel, maxLevel = Q.maxLevel;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    const bool right = Q.right != 0;
    const bool checkUr = !right && Q.checkUr && P.nleft == -1 && P.uRight != nullptr;
    const uint4 a0 = *reinterpret_cast<const uint4*>(Q.desc), a1 = *(reinterpret_cast<const uint4*>(Q.desc) + 1);
    const int ny = nMaxCellY - nMinCellY + 1, ncell = (nMaxCellX - nMinCellX + 1) * ny;
    const int side = right ? GRID_CELLS : 0;
    unsigned long long k1 = ~0ull, k2 = ~0ull;
    for (int e = lane; e < ncell; e += 32) {
        const int ix = nMinCellX + e / ny, iy = nMinCellY + e % ny;
        const int c = side + ix * GRID_ROWS + iy;
        for (int j = cellStart[c]; j < cellEnd[c]; ++j) {
            const int idx = (int)(sorted[j] & 0xFFFFu);
            const float4 kp = P.kp[idx];
            const int oct = __float_as_int(kp.w);
            if (bCheckLevels) {
                if (oct < minLevel) continue;
                if (maxLevel >= 0 && oct > maxLevel) continue;
            }
            if (!(fabsf(__fsub_rn(kp.x, x)) < r && fabsf(__fsub_rn(kp.y, y)) < r)) continue;
            if (blockedAt[idx] < i) continue;
            if (checkUr) {
                const float ur = P.uRight[idx];
                if (ur > 0.f && fabsf(__fsub_rn(Q.ur, ur)) > r) continue;
            }
            const int dist = ham_q(a0, a1, P.desc + 32ll * idx);
            top2_insert(k1, k2, (unsigned long long)dist << 40 | (unsigned long long)e << 16 | (unsigned)idx);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long o1 = __shfl_xor_sync(0xffffffffu, k1, o), o2 = __shfl_xor_sync(0xffffffffu, k2, o);
        const unsigned long long lo = min(k1, o1), hi = max(k1, o1);
        k2 = min(hi, min(k2, o2));
        k1 = lo;
    }
    if (k1 == ~0ull) return make_int2(-1, 0);
    const int bestDist = (int)(k1 >> 40), bestIdx = (int)(k1 & 0xFFFFu);
    if (bestDist > th) return make_int2(-1, 0);
    if (mode == 1 && k2 != ~0ull) {
        const int bestDist2 = (int)(k2 >> 40), idx2 = (int)(k2 & 0xFFFFu);
        const int l1 = __float_as_int(P.kp[bestIdx].w), l2 = __float_as_int(P.kp[idx2].w);
        if (l1 == l2 && (float)bestDist > __fmul_rn(nnRatio, (float)bestDist2)) return make_int2(-1, 1);
    }
    return make_int2(bestIdx, 0);
}

__device__ __forceinline__ int stereo_link(const ProjProblem& P, int idx) {
    if (P.nleft == -1) return -1;
    if (idx < P.nleft) {
        if (!P.l2r) return -1;
        const int r = P.l2r[idx];
        return r == -1 ? -1 : r + P.nleft;
    }
    if (!P.r2l) return -1;
    return P.r2l[idx - P.nleft];          // -1 or a left index
}

__global__ void __launch_bounds__(PROJ_T) k_search_by_projection(const ProjProblem* __restrict__ probs, int mode, int th,
                                                                 float nnRatio, int linkStereo, int rotHist) {
    extern __shared__ __align__(16) uint32_t sKeys[];
    __shared__ int hist[ORBFE_HISTO_LENGTH];
    __shared__ int keep[3];
    __shared__ int sChanged, sCount;
    const ProjProblem P = probs[blockIdx.x];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = P.n, nq = P.nq;
    int np2 = 1;
    while (np2 < n) np2 <<= 1;
    int* cellStart = P.scratch;                       // 2 * GRID_CELLS
    int* cellEnd = cellStart + 2 * GRID_CELLS;        // 2 * GRID_CELLS
    int* blockedAt = cellEnd + 2 * GRID_CELLS;        // n
    int* chA = blockedAt + n;                         // nq   (choice, previous sweep)
    int* chB = chA + nq;                              // nq
    int* rjA = chB + nq;                              // nq   (ratio-rejected flag, previous sweep)
    int* rjB = rjA + nq;                              // nq
    // ---- 1. grid
    for (int t = tid; t < 2 * GRID_CELLS; t += PROJ_T) { cellStart[t] = 0; cellEnd[t] = 0; }
    for (int t = tid; t < np2; t += PROJ_T) {
        uint32_t key = 0xFFFFFFFFu;
        if (t < n) {
            const float4 kp = P.kp[t];
            const int px = (int)roundf(__fmul_rn(__fsub_rn(kp.x, P.minX), P.gridWInv));
            const int py = (int)roundf(__fmul_rn(__fsub_rn(kp.y, P.minY), P.gridHInv));
            if (px >= 0 && px < GRID_COLS && py >= 0 && py < GRID_ROWS) {
                const int side = (P.nleft == -1 || t < P.nleft) ? 0 : GRID_CELLS;
                key = (uint32_t)(side + px * GRID_ROWS + py) << 16 | (uint32_t)t;
            }
        }
        sKeys[t] = key;
    }
    __syncthreads();
    for (int k = 2; k <= np2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = tid; t < np2; t += PROJ_T) {
                const int p = t ^ j;
                if (p > t) {
                    const uint32_t a = sKeys[t], b = sKeys[p];
                    if ((a > b) == ((t & k) == 0)) { sKeys[t] = b; sKeys[p] = a; }
                }
            }
            __syncthreads();
        }
    for (int t = tid; t < n; t += PROJ_T) {
        const uint32_t key = sKeys[t];
        if (key == 0xFFFFFFFFu) continue;
        const int c = (int)(key >> 16);
        if (t == 0 || (int)(sKeys[t - 1] >> 16) != c) cellStart[c] = t;
        if (t == n - 1 || sKeys[t + 1] == 0xFFFFFFFFu || (int)(sKeys[t + 1] >> 16) != c) cellEnd[c] = t + 1;
    }
    for (int i = tid; i < nq; i += PROJ_T) { chA[i] = -1; rjA[i] = 0; }
    __syncthreads();
    // ---- 2./3. Jacobi sweeps
    int* chOld = chA; int* chNew = chB; int* rjOld = rjA; int* rjNew = rjB;
    for (int sweep = 0; sweep <= nq; ++sweep) {
        for (int t = tid; t < n; t += PROJ_T) blockedAt[t] = (P.blocked && P.blocked[t]) ? -1 : 0x7fffffff;
        if (tid == 0) sChanged = 0;
        __syncthreads();
        for (int i = tid; i < nq; i += PROJ_T) {
            const int c = chOld[i];
            if (c >= 0 && P.q[i].obs) {
                atomicMin(&blockedAt[c], i);
                if (linkStereo) { const int e = stereo_link(P, c); if (e >= 0) atomicMin(&blockedAt[e], i); }
            }
        }
        __syncthreads();
        int changed = 0;
        for (int i = warp; i < nq; i += PROJ_T / 32) {
            int2 res = make_int2(-1, 0);
            const ProjQueryDev& Q = P.q[i];
            const bool skip = !Q.valid || (Q.skipIfPrevRatioRejected && i > 0 && rjOld[i - 1]);
            if (!skip) res = search_window(P, i, lane, sKeys, cellStart, cellEnd, blockedAt, mode, th, nnRatio);
            if (lane == 0) {
                if (res.x != chOld[i] || res.y != rjOld[i]) changed = 1;
                chNew[i] = res.x; rjNew[i] = res.y;
            }
        }
        if (changed) sChanged = 1;
        __syncthreads();
        int* t1 = chOld; chOld = chNew; chNew = t1;
        int* t2 = rjOld; rjOld = rjNew; rjNew = t2;
        const int any = sChanged;
        __syncthreads();
        if (!any) break;
    }
    // ---- 4. finish (chOld holds the fixed point)
    for (int t = tid; t < n; t += PROJ_T) P.assign[t] = -1;
    if (tid < ORBFE_HISTO_LENGTH) hist[tid] = 0;
    if (tid == 0) sCount = 0;
    __syncthreads();
    int cnt = 0;
    for (int i = tid; i < nq; i += PROJ_T) {
        const int c = chOld[i];
        P.choice[i] = c;
        if (c < 0) continue;
        atomicMax(&P.assign[c], i);
        ++cnt;
        if (linkStereo) { const int e = stereo_link(P, c); if (e >= 0) { atomicMax(&P.assign[e], i); ++cnt; } }
        if (rotHist) atomicAdd(&hist[rot_bin_p(P.q[i].angle, P.kp[c].z)], 1);
    }
    __syncthreads();
    if (rotHist) {
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
        for (int i = tid; i < nq; i += PROJ_T) {
            const int c = chOld[i];
            if (c < 0) continue;
            const int bin = rot_bin_p(P.q[i].angle, P.kp[c].z);
            if (bin == keep[0] || bin == keep[1] || bin == keep[2]) continue;
            P.assign[c] = -2;                 // mvpMapPoints[c] = NULL (:2381), whoever wrote it last
            --cnt;
        }
    }
    atomicAdd(&sCount, cnt);
    __syncthreads();
    if (tid == 0) *P.nmatches = sCount;
}

cudaError_t orbfe_launch_search_by_projection(cudaStream_t st, const ProjProblem* dProbs, int nprob, int nMax, int mode, int th,
                                              float nnRatio, int linkStereo, int rotHist, long long* launches) {
    if (nprob <= 0) return cudaSuccess;
    int np2 = 1;
    while (np2 < nMax) np2 <<= 1;
    const size_t smem = (size_t)np2 * 4;
    static size_t configured = 0;
    if (smem > 48 * 1024 && smem > configured) {
        cudaError_t e = cudaFuncSetAttribute(k_search_by_projection, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured = smem;
    }
    k_search_by_projection<<<nprob, PROJ_T, smem, st>>>(dProbs, mode, th, nnRatio, linkStereo, rotHist);
    ++*launches;
    return cudaGetLastError();
}
