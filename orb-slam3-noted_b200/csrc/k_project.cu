// k_project.cu — the projection-guided window searches: Frame::GetFeaturesInArea (src/Frame.cc:640-706) over the
// 64 x 48 feature grid of Frame::AssignFeaturesToGrid (:380-411) + the matching loop shared by
// ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th, ...) (src/ORBmatcher.cc:44-236) and the best-only
// overloads (Frame&, const Frame& LastFrame, ...) :2173-2389, (Frame&, KeyFrame*, ...) :2391-2510.
//
// Three kernels per call (any number of problems = Frames with their ordered windows per launch):
//   k_proj_grid (one CTA per problem): every feature's cell by PosInGrid's round() rule; a shared-memory counting
//      sort by (side, cell) with every cell's segment in ascending index = the reference's push order.
//   k_proj_candidates (one warp per window, CTAs over all SMs): lanes take the window's cells in
//      GetFeaturesInArea's order (x outer, y inner); a candidate's key is dist << 40 | position of its cell << 16 |
//      index, so ascending keys are exactly the order in which the reference's strict '<' updates prefer
//      candidates.  The PROJ_K smallest keys of the window are stored together with `limit`, the smallest key
//      that was dropped anywhere (lane-local third best or list overflow): entries below it are a correct prefix
//      of the window's full sorted candidate list.
//   k_proj_resolve (one CTA per problem): the reference assigns MapPoints one after another and a feature
//      holding a MapPoint with observations is skipped by every later window.  That recurrence (window i depends
//      on the choices of windows < i) is solved by Jacobi sweeps: every window re-reads its short list against the
//      claims of the previous sweep (blockedAt[f] = smallest claiming window index) until nothing changes.  Sweep
//      k makes windows 0..k-1 final, so the fixed point is the sequential result; conflicts are rare and 2-4
//      sweeps do.  A window that runs past its list's valid prefix redoes the full search (rare).  Then: last
//      writer per feature (atomicMax), stereo links, rotation histogram, count.
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

struct Window {
    int minCX, maxCX, minCY, maxCY;   // cell range, empty if minCX > maxCX
};

__device__ __forceinline__ Window window_cells(const ProjProblem& P, const ProjQueryDev& Q) {
    const float x = Q.u, y = Q.v, r = Q.r;
    Window w;
    w.minCX = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, P.minX), r), P.gridWInv)));
    w.maxCX = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, P.minX), r), P.gridWInv)));
    w.minCY = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, P.minY), r), P.gridHInv)));
    w.maxCY = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, P.minY), r), P.gridHInv)));
    if (w.minCX >= GRID_COLS || w.maxCX < 0 || w.minCY >= GRID_ROWS || w.maxCY < 0 || w.minCY > w.maxCY) { w.minCX = 1; w.maxCX = 0; }
    return w;
}

// Visits the candidates of window i that lane `lane` of `nlanes` owns, in cell order; f(key, idx).
// blockedAt == nullptr: only the caller's static mask applies.
struct AlwaysVisible { __device__ __forceinline__ bool operator()(int, int) const { return true; } };

// vis(idx, dist): evaluated after the distance, false = the candidate is skipped (SearchForInitialization :744).
template <class F, class V = AlwaysVisible>
__device__ __forceinline__ void scan_window(const ProjProblem& P, int i, int lane, int nlanes, const int* __restrict__ blockedAt, F f,
                                            V vis = V(), int* areaHits = nullptr) {
    const ProjQueryDev& Q = P.q[i];
    const Window w = window_cells(P, Q);
    if (w.minCX > w.maxCX) return;
    const float x = Q.u, y = Q.v, r = Q.r;
    const int minLevel = Q.minLevel, maxLevel = Q.maxLevel;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    const bool right = Q.right != 0;
    const bool checkUr = !right && Q.checkUr && P.nleft == -1 && P.uRight != nullptr;
    const bool chi2 = Q.chi2 != 0 && P.invSigma2 != nullptr;
    const uint4 a0 = *reinterpret_cast<const uint4*>(Q.desc), a1 = *(reinterpret_cast<const uint4*>(Q.desc) + 1);
    const int ny = w.maxCY - w.minCY + 1, ncell = (w.maxCX - w.minCX + 1) * ny;
    const int side = right ? GRID_CELLS : 0;
    for (int e = lane; e < ncell; e += nlanes) {
        const int ix = w.minCX + e / ny, iy = w.minCY + e % ny;
        const int c = side + ix * GRID_ROWS + iy;
        for (int j = P.cellStart[c]; j < P.cellEnd[c]; ++j) {
            const int idx = (int)(P.sorted[j] & 0xFFFFu);
            const float4 kp = P.kp[idx];
            const int oct = __float_as_int(kp.w);
            if (bCheckLevels) {
                if (oct < minLevel) continue;
                if (maxLevel >= 0 && oct > maxLevel) continue;
            }
            if (!(fabsf(__fsub_rn(kp.x, x)) < r && fabsf(__fsub_rn(kp.y, y)) < r)) continue;
            if (areaHits) ++*areaHits;          // the feature is in GetFeaturesInArea's result, whatever the loop does with it
            if (blockedAt ? blockedAt[idx] < i : (P.blocked && P.blocked[idx])) continue;
            if (checkUr) {
                const float ur = P.uRight[idx];
                if (ur > 0.f && fabsf(__fsub_rn(Q.ur, ur)) > r) continue;
            }
            if (chi2) {             // Fuse :1744-1770 (mvuRight is read with the camera-local index, as the reference does)
                const float ex = __fsub_rn(x, kp.x), ey = __fsub_rn(y, kp.y);
                const float kpr = P.uRight ? P.uRight[right ? idx - P.nleft : idx] : -1.f;
                float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
                double gate = 5.99;
                if (kpr >= 0.f) {
                    const float er = __fsub_rn(Q.ur, kpr);
                    e2 = __fadd_rn(e2, __fmul_rn(er, er));
                    gate = 7.8;
                }
                if ((double)__fmul_rn(e2, P.invSigma2[oct]) > gate) continue;
            }
            const int dist = ham_q(a0, a1, P.desc + 32ll * idx);
            if (!vis(idx, dist)) continue;
            f((unsigned long long)dist << 40 | (unsigned long long)e << 16 | (unsigned)idx);
        }
    }
}

// Accept rule on the best / second-best keys: chosen feature (or -1) and whether the ratio rule rejected it.
__device__ __forceinline__ int2 decide(const ProjProblem& P, unsigned long long k1, unsigned long long k2, int mode, int th, float nnRatio) {
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

// ---- 1. grid
// Counting sort (round 2; the bitonic sort of 2048 keys took 27 us of a 100 us call): a histogram of the 2 x 3072 cells in shared
// memory, an exclusive scan, a scatter with per-cell fill counters, and - because the scatter's order within a cell is whatever
// the atomics made it - an insertion sort of every cell's short segment by feature index (= the reference's push order).
#define GRID_TOTAL (2 * GRID_CELLS)
#define GRID_PER_THREAD (GRID_TOTAL / PROJ_T)      // 12 cells per thread
__global__ void __launch_bounds__(PROJ_T) k_proj_grid(const ProjProblem* __restrict__ probs) {
    extern __shared__ __align__(16) uint32_t sGrid[];
    int* sCount = reinterpret_cast<int*>(sGrid);                    // GRID_TOTAL: histogram, then fill cursor
    uint32_t* sKeys = sGrid + GRID_TOTAL;                           // n sorted keys
    __shared__ int sWarp[PROJ_T / 32];
    const ProjProblem P = probs[blockIdx.x];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, n = P.n;
    for (int t = tid; t < GRID_TOTAL; t += PROJ_T) sCount[t] = 0;
    __syncthreads();
    // the cell of every feature (PosInGrid's round() rule), -1 outside the grid
    auto cell_of = [&](int t) -> int {
        const float4 kp = P.kp[t];
        const int px = (int)roundf(__fmul_rn(__fsub_rn(kp.x, P.minX), P.gridWInv));
        const int py = (int)roundf(__fmul_rn(__fsub_rn(kp.y, P.minY), P.gridHInv));
        if (px < 0 || px >= GRID_COLS || py < 0 || py >= GRID_ROWS) return -1;
        return ((P.nleft == -1 || t < P.nleft) ? 0 : GRID_CELLS) + px * GRID_ROWS + py;
    };
    for (int t = tid; t < n; t += PROJ_T) {
        const int c = cell_of(t);
        if (c >= 0) atomicAdd(&sCount[c], 1);
    }
    __syncthreads();
    // exclusive scan over the cells: thread = 12 consecutive cells
    int cnt[GRID_PER_THREAD], sum = 0;
#pragma unroll
    for (int k = 0; k < GRID_PER_THREAD; ++k) { cnt[k] = sCount[tid * GRID_PER_THREAD + k]; sum += cnt[k]; }
    int inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
    }
    if (lane == 31) sWarp[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int w = lane < PROJ_T / 32 ? sWarp[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += v;
        }
        if (lane < PROJ_T / 32) sWarp[lane] = w;          // inclusive over warps
    }
    __syncthreads();
    int run = (wid > 0 ? sWarp[wid - 1] : 0) + inc - sum;   // first slot of this thread's first cell
    int start[GRID_PER_THREAD];
#pragma unroll
    for (int k = 0; k < GRID_PER_THREAD; ++k) {
        start[k] = run;
        const int c = tid * GRID_PER_THREAD + k;
        P.cellStart[c] = cnt[k] ? run : 0;                  // (empty cells read 0 / 0, as before)
        P.cellEnd[c] = cnt[k] ? run + cnt[k] : 0;
        sCount[c] = run;                                    // fill cursor
        run += cnt[k];
    }
    __syncthreads();
    for (int t = tid; t < n; t += PROJ_T) {
        const int c = cell_of(t);
        if (c >= 0) sKeys[atomicAdd(&sCount[c], 1)] = (uint32_t)c << 16 | (uint32_t)t;
    }
    __syncthreads();
    // ascending feature index inside every cell (segments are a handful of entries: 3072 cells for a few thousand features; a
    // frame with hundreds of features in ONE 12 x 10 px cell would make this quadratic for that thread - slow, still correct)
#pragma unroll
    for (int k = 0; k < GRID_PER_THREAD; ++k) {
        const int b0 = start[k], e0 = b0 + cnt[k];
        for (int i = b0 + 1; i < e0; ++i) {
            const uint32_t key = sKeys[i];
            int j = i - 1;
            while (j >= b0 && sKeys[j] > key) { sKeys[j + 1] = sKeys[j]; --j; }
            sKeys[j + 1] = key;
        }
    }
    __syncthreads();
    const int valid = sWarp[PROJ_T / 32 - 1];
    for (int t = tid; t < n; t += PROJ_T) P.sorted[t] = t < valid ? sKeys[t] : 0xFFFFFFFFu;
}

// ---- 2. candidate lists: blockIdx.y = problem, one warp per window
__global__ void __launch_bounds__(256) k_proj_candidates(const ProjProblem* __restrict__ probs) {
    const ProjProblem P = probs[blockIdx.y];
    const int i = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= P.nq) return;
    unsigned long long* list = P.cand + (long long)i * PROJ_K;
    int* areaEmpty = P.scratch + P.n + 4 * P.nq;                    // nq flags: GetFeaturesInArea returned nothing
    if (!P.q[i].valid) {
        if (lane < PROJ_K) list[lane] = ~0ull;
        if (lane == 0) { P.limit[i] = ~0ull; areaEmpty[i] = 1; }
        return;
    }
    unsigned long long k1 = ~0ull, k2 = ~0ull, dropped = ~0ull;     // lane-local best two + smallest key let go
    int hits = 0;
    scan_window(P, i, lane, 32, nullptr, [&](unsigned long long k) {
        if (k < k1) { dropped = min(dropped, k2); k2 = k1; k1 = k; }
        else if (k < k2) { dropped = min(dropped, k2); k2 = k; }
        else dropped = min(dropped, k);
    }, AlwaysVisible(), &hits);
    {
        const unsigned any = __ballot_sync(0xffffffffu, hits > 0);
        if (lane == 0) areaEmpty[i] = any == 0u;
    }
    // the PROJ_K smallest of the 64 kept keys: repeated warp-min extraction
    unsigned long long lim = dropped;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) lim = min(lim, __shfl_xor_sync(0xffffffffu, lim, o));
    for (int s = 0; s <= PROJ_K; ++s) {
        unsigned long long m = k1;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = min(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (s == PROJ_K) { lim = min(lim, m); break; }          // first key that did not fit
        if (lane == 0) list[s] = m;
        if (m == ~0ull) {                                         // exhausted: pad
            if (lane > s && lane < PROJ_K) list[lane] = ~0ull;
            break;
        }
        if (k1 == m) { k1 = k2; k2 = ~0ull; }                     // keys are unique (they carry the feature index)
    }
    if (lane == 0) P.limit[i] = lim;
}

// ---- 3. ordered assignment
__global__ void __launch_bounds__(PROJ_T) k_proj_resolve(const ProjProblem* __restrict__ probs, int mode, int th, float nnRatio,
                                                         int linkStereo, int rotHist) {
    __shared__ int hist[ORBFE_HISTO_LENGTH];
    __shared__ int keep[3];
    __shared__ int sChanged, sCount;
    const ProjProblem P = probs[blockIdx.x];
    const int tid = threadIdx.x;
    const int n = P.n, nq = P.nq;
    int* blockedAt = P.scratch;                       // n
    int* chA = blockedAt + n;                         // nq   (choice, previous sweep)
    int* chB = chA + nq;
    int* rjA = chB + nq;                              // nq   (ratio-rejected flag, previous sweep)
    int* rjB = rjA + nq;
    const int* areaEmpty = rjB + nq;                  // nq   (written by k_proj_candidates)
    for (int i = tid; i < nq; i += PROJ_T) { chA[i] = -1; rjA[i] = 0; }
    __syncthreads();
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
        for (int i = tid; i < nq; i += PROJ_T) {
            int2 res = make_int2(-1, 0);
            const ProjQueryDev& Q = P.q[i];
            const bool skip = !Q.valid || (Q.skipIfPrevRatioRejected && i > 0 && rjOld[i - 1]) ||
                              (Q.skipIfPrevEmpty && i > 0 && areaEmpty[i - 1]);
            if (!skip) {
                const unsigned long long* list = P.cand + (long long)i * PROJ_K;
                const unsigned long long lim = P.limit[i];
                unsigned long long k1 = ~0ull, k2 = ~0ull;
                bool complete = false;                  // enough unblocked entries seen inside the valid prefix
                for (int s = 0; s < PROJ_K; ++s) {
                    const unsigned long long k = list[s];
                    if (k == ~0ull) { complete = lim == ~0ull; break; }
                    if (k >= lim) break;
                    if (blockedAt[(int)(k & 0xFFFFu)] < i) continue;
                    if (k1 == ~0ull) { k1 = k; if (mode == 0) { complete = true; break; } }
                    else { k2 = k; complete = true; break; }
                }
                if (!complete && lim == ~0ull) complete = true;     // list held every candidate
                if (!complete) {                                     // past the valid prefix: full scan (rare)
                    k1 = ~0ull; k2 = ~0ull;
                    scan_window(P, i, 0, 1, blockedAt, [&](unsigned long long k) { top2_insert(k1, k2, k); });
                }
                res = decide(P, k1, k2, mode, th, nnRatio);
            }
            if (res.x != chOld[i] || res.y != rjOld[i]) changed = 1;
            chNew[i] = res.x; rjNew[i] = res.y;
        }
        if (changed) sChanged = 1;
        __syncthreads();
        int* t1 = chOld; chOld = chNew; chNew = t1;
        int* t2 = rjOld; rjOld = rjNew; rjNew = t2;
        const int any = sChanged;
        __syncthreads();
        if (!any) break;
    }
    // finish (chOld holds the fixed point)
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

// ---- 3b. SearchForInitialization (ORBmatcher.cc:706-821): one window per F1 feature, windows in index order.
// The reference keeps, per F2 feature, the distance of its current match (vMatchedDistance) and hides the feature
// from any later window whose distance is not strictly smaller (:744); a later, closer window steals it (:757-768).
// Same Jacobi scheme as k_proj_resolve: in every sweep the accepted windows of the previous sweep are chained per
// feature (head / next), window i treats a candidate as visible iff no window j < i on that chain has
// dist_j <= dist, and takes the two smallest visible keys of its list.  Sweep k makes windows < k final.
// Afterwards: vnMatches12[i] survives iff i is the LAST window that chose its feature; the rotation histogram
// counts every accepted window, stolen or not (rotHist[bin].push_back(i1) is never undone, :772-781).
__global__ void __launch_bounds__(PROJ_T) k_init_resolve(const ProjProblem* __restrict__ probs, int th, float nnRatio, int rotHist) {
    __shared__ int hist[ORBFE_HISTO_LENGTH];
    __shared__ int keep[3];
    __shared__ int sChanged, sCount;
    const ProjProblem P = probs[blockIdx.x];
    const int tid = threadIdx.x;
    const int n = P.n, nq = P.nq;
    int* head = P.scratch;                            // n: newest claimant of each F2 feature (previous sweep)
    int* chA = head + n;                              // nq: choice
    int* chB = chA + nq;
    int* dA = chB + nq;                               // nq: distance of the choice
    int* dB = dA + nq;
    int* next = P.assign;                             // nq <= capacity of assign (the API sizes it max(n, nq))
    for (int i = tid; i < nq; i += PROJ_T) { chA[i] = -1; dA[i] = 0; }
    __syncthreads();
    int* chOld = chA; int* chNew = chB; int* dOld = dA; int* dNew = dB;
    for (int sweep = 0; sweep <= nq; ++sweep) {
        for (int t = tid; t < n; t += PROJ_T) head[t] = -1;
        if (tid == 0) sChanged = 0;
        __syncthreads();
        for (int i = tid; i < nq; i += PROJ_T) {
            const int c = chOld[i];
            if (c >= 0) next[i] = atomicExch(&head[c], i);
        }
        __syncthreads();
        auto visible = [&](int i, int idx, int dist) {
            for (int j = head[idx]; j >= 0; j = next[j])
                if (j < i && dOld[j] <= dist) return false;
            return true;
        };
        int changed = 0;
        for (int i = tid; i < nq; i += PROJ_T) {
            int res = -1, resDist = 0;
            if (P.q[i].valid) {
                const unsigned long long* list = P.cand + (long long)i * PROJ_K;
                const unsigned long long lim = P.limit[i];
                unsigned long long k1 = ~0ull, k2 = ~0ull;
                bool complete = false;
                for (int s = 0; s < PROJ_K; ++s) {
                    const unsigned long long k = list[s];
                    if (k == ~0ull) { complete = lim == ~0ull; break; }
                    if (k >= lim) break;
                    if (!visible(i, (int)(k & 0xFFFFu), (int)(k >> 40))) continue;
                    if (k1 == ~0ull) k1 = k;
                    else { k2 = k; complete = true; break; }
                }
                if (!complete && lim == ~0ull) complete = true;
                if (!complete) {
                    k1 = ~0ull; k2 = ~0ull;
                    scan_window(P, i, 0, 1, nullptr, [&](unsigned long long k) { top2_insert(k1, k2, k); },
                                [&](int idx, int dist) { return visible(i, idx, dist); });
                }
                if (k1 != ~0ull) {
                    const int bestDist = (int)(k1 >> 40);
                    const float second = k2 == ~0ull ? 2147483648.0f : (float)(int)(k2 >> 40);      // (float)INT_MAX
                    if (bestDist <= th && (float)bestDist < __fmul_rn(second, nnRatio)) { res = (int)(k1 & 0xFFFFu); resDist = bestDist; }
                }
            }
            if (res != chOld[i] || resDist != dOld[i]) changed = 1;
            chNew[i] = res; dNew[i] = resDist;
        }
        if (changed) sChanged = 1;
        __syncthreads();
        int* t1 = chOld; chOld = chNew; chNew = t1;
        int* t2 = dOld; dOld = dNew; dNew = t2;
        const int any = sChanged;
        __syncthreads();
        if (!any) break;
    }
    // finish: head[f] <- last window that chose f (vnMatches21)
    for (int t = tid; t < n; t += PROJ_T) head[t] = -1;
    if (tid < ORBFE_HISTO_LENGTH) hist[tid] = 0;
    if (tid == 0) sCount = 0;
    __syncthreads();
    for (int i = tid; i < nq; i += PROJ_T) {
        const int c = chOld[i];
        if (c < 0) continue;
        atomicMax(&head[c], i);
        if (rotHist) atomicAdd(&hist[rot_bin_p(P.q[i].angle, P.kp[c].z)], 1);
    }
    __syncthreads();
    if (rotHist && tid == 0) {
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
    int cnt = 0;
    for (int i = tid; i < nq; i += PROJ_T) {
        int c = chOld[i];
        if (c >= 0 && head[c] != i) c = -1;                        // stolen by a later window (:759-763)
        if (c >= 0 && rotHist) {
            const int bin = rot_bin_p(P.q[i].angle, P.kp[c].z);
            if (!(bin == keep[0] || bin == keep[1] || bin == keep[2])) c = -1;
        }
        P.choice[i] = c;
        if (c >= 0) ++cnt;
    }
    atomicAdd(&sCount, cnt);
    __syncthreads();
    if (tid == 0) *P.nmatches = sCount;
}

cudaError_t orbfe_launch_search_for_initialization(cudaStream_t st, const ProjProblem* dProbs, int nprob, int nMax, int nqMax, int th,
                                                   float nnRatio, int rotHist, long long* launches) {
    if (nprob <= 0) return cudaSuccess;
    const size_t smem = ((size_t)2 * 64 * 48 + (size_t)nMax) * 4;      // cell counters + one key per feature
    {
        const cudaError_t ea = orbfe_ensure_dyn_smem(reinterpret_cast<const void*>(k_proj_grid), smem);
        if (ea != cudaSuccess) return ea;
    }
    k_proj_grid<<<nprob, PROJ_T, smem, st>>>(dProbs);
    dim3 grid((nqMax + 7) / 8, nprob);
    k_proj_candidates<<<grid, 256, 0, st>>>(dProbs);
    k_init_resolve<<<nprob, PROJ_T, 0, st>>>(dProbs, th, nnRatio, rotHist);
    *launches += 3;
    return cudaGetLastError();
}

cudaError_t orbfe_launch_search_by_projection(cudaStream_t st, const ProjProblem* dProbs, int nprob, int nMax, int nqMax, int mode,
                                              int th, float nnRatio, int linkStereo, int rotHist, long long* launches) {
    if (nprob <= 0) return cudaSuccess;
    const size_t smem = ((size_t)2 * 64 * 48 + (size_t)nMax) * 4;      // cell counters + one key per feature
    {   // the attribute is per device: one high-water mark per (device, kernel), see orbfe_ensure_dyn_smem
        const cudaError_t ea = orbfe_ensure_dyn_smem(reinterpret_cast<const void*>(k_proj_grid), smem);
        if (ea != cudaSuccess) return ea;
    }
    k_proj_grid<<<nprob, PROJ_T, smem, st>>>(dProbs);
    if (nqMax > 0) {
        dim3 grid((nqMax + 7) / 8, nprob);
        k_proj_candidates<<<grid, 256, 0, st>>>(dProbs);
        ++*launches;
    }
    k_proj_resolve<<<nprob, PROJ_T, 0, st>>>(dProbs, mode, th, nnRatio, linkStereo, rotHist);
    *launches += 2;
    return cudaGetLastError();
}
