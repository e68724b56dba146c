// k_match.cu — brute-force Hamming 2-NN: cv::BFMatcher(NORM_HAMMING).knnMatch(k=2) as used by
// Frame::ComputeStereoFishEyeMatches (Frame.cc:43,1134), with ORBmatcher::DescriptorDistance
// (ORBmatcher.cc:2561-2577) as the metric.
//
// One thread owns one query descriptor (8 x u32 in registers) and scans a slice of the train set,
// which the CTA streams through shared memory in 128-row tiles (4 KB, uint4 loads; every thread
// reads the same train row at the same time, so the shared reads are broadcasts).  The distance
// is 8 x (xor, popc); best and second best are kept as packed keys  dist << 20 | trainIndex , so
// "two smallest keys" is exactly "two smallest distances, ties to the lowest train index"
// (OpenCV's order) and the update is three min/max ops without a branch.  The train set is split
// over blockIdx.y to fill the GPU when there are few queries; partial results are merged by a
// second tiny kernel.  Integer-pipe (popc) bound by construction: HBM traffic is nq*32 + nt*32
// bytes per tile pass, re-used 128 times from registers / shared memory.
#include "orbfe_internal.h"
#include "orbfe_match.h"

#define KNN_THREADS 128
#define KNN_TILE 128
#define KNN_IDX_BITS 20
#define KNN_IDX_MASK ((1u << KNN_IDX_BITS) - 1u)
#define KNN_NONE 0xFFFFFFFFu

__device__ __forceinline__ void knn_update(uint32_t& k0, uint32_t& k1, uint32_t k) {
    const uint32_t hi = max(k0, k);
    k0 = min(k0, k);
    k1 = min(k1, hi);
}

__device__ __forceinline__ void knn_store(uint32_t k0, uint32_t k1, int32_t* idx, int32_t* dist) {
    idx[0] = k0 == KNN_NONE ? -1 : (int)(k0 & KNN_IDX_MASK);
    dist[0] = k0 == KNN_NONE ? -1 : (int)(k0 >> KNN_IDX_BITS);
    idx[1] = k1 == KNN_NONE ? -1 : (int)(k1 & KNN_IDX_MASK);
    dist[1] = k1 == KNN_NONE ? -1 : (int)(k1 >> KNN_IDX_BITS);
}

// grid = (query tiles, train splits, batch)
__global__ void __launch_bounds__(KNN_THREADS) k_knn2(const uint8_t* __restrict__ q, long long qStride, const int* __restrict__ nqArr,
                                                      int nqMax, const uint8_t* __restrict__ t, long long tStride,
                                                      const int* __restrict__ ntArr, int ntMax, int splitLen,
                                                      uint2* __restrict__ partial, int32_t* __restrict__ idx2,
                                                      int32_t* __restrict__ dist2) {
    __shared__ uint4 tile[KNN_TILE * 2];
    const int b = blockIdx.z, split = blockIdx.y, nsplit = gridDim.y;
    const int nq = nqArr ? min(nqArr[b], nqMax) : nqMax;
    const int nt = ntArr ? min(ntArr[b], ntMax) : ntMax;
    const int qi = blockIdx.x * KNN_THREADS + threadIdx.x;
    if (blockIdx.x * KNN_THREADS >= nq) return;
    const bool live = qi < nq;
    uint32_t qa[8];
    {
        const uint4* qp = reinterpret_cast<const uint4*>(q + (long long)b * qStride) + 2 * (long long)(live ? qi : 0);
        const uint4 a = __ldg(qp), c = __ldg(qp + 1);
        qa[0] = a.x; qa[1] = a.y; qa[2] = a.z; qa[3] = a.w; qa[4] = c.x; qa[5] = c.y; qa[6] = c.z; qa[7] = c.w;
    }
    const uint4* tp = reinterpret_cast<const uint4*>(t + (long long)b * tStride);
    const int t0 = split * splitLen, t1 = min(t0 + splitLen, nt);
    uint32_t k0 = KNN_NONE, k1 = KNN_NONE;
    for (int base = t0; base < t1; base += KNN_TILE) {
        const int n = min(KNN_TILE, t1 - base);
        __syncthreads();
        for (int i = threadIdx.x; i < 2 * n; i += KNN_THREADS) tile[i] = __ldg(tp + 2 * (long long)base + i);
        __syncthreads();
#pragma unroll 4
        for (int j = 0; j < n; ++j) {
            const uint4 a = tile[2 * j], c = tile[2 * j + 1];
            const int d = __popc(qa[0] ^ a.x) + __popc(qa[1] ^ a.y) + __popc(qa[2] ^ a.z) + __popc(qa[3] ^ a.w) +
                          __popc(qa[4] ^ c.x) + __popc(qa[5] ^ c.y) + __popc(qa[6] ^ c.z) + __popc(qa[7] ^ c.w);
            knn_update(k0, k1, ((uint32_t)d << KNN_IDX_BITS) | (uint32_t)(base + j));
        }
    }
    if (!live) return;
    if (nsplit == 1) {
        const long long o = ((long long)b * nqMax + qi) * 2;
        knn_store(k0, k1, idx2 + o, dist2 + o);
    } else {
        partial[((long long)b * nsplit + split) * nqMax + qi] = make_uint2(k0, k1);
    }
}

__global__ void __launch_bounds__(KNN_THREADS) k_knn2_merge(const uint2* __restrict__ partial, const int* __restrict__ nqArr,
                                                            int nqMax, int nsplit, int32_t* __restrict__ idx2,
                                                            int32_t* __restrict__ dist2) {
    const int b = blockIdx.y, qi = blockIdx.x * KNN_THREADS + threadIdx.x;
    const int nq = nqArr ? min(nqArr[b], nqMax) : nqMax;
    if (qi >= nq) return;
    uint32_t k0 = KNN_NONE, k1 = KNN_NONE;
    for (int s = 0; s < nsplit; ++s) {
        const uint2 p = partial[((long long)b * nsplit + s) * nqMax + qi];
        knn_update(k0, k1, p.x);
        knn_update(k0, k1, p.y);   // KNN_NONE entries never displace a real key
    }
    const long long o = ((long long)b * nqMax + qi) * 2;
    knn_store(k0, k1, idx2 + o, dist2 + o);
}

// Lowe-ratio filter of Frame.cc:1141 on 2-NN results: match = idx0 if both exist and d0 < d1 * ratio.
// DMatch::distance is float and the literal 0.7 is double in the reference, hence the double compare.
__global__ void k_ratio_filter(const int32_t* __restrict__ idx2, const int32_t* __restrict__ dist2, int nq, double ratio,
                               int32_t* __restrict__ match, int* __restrict__ count) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    const bool ok = idx2[2 * i + 1] >= 0 && (double)(float)dist2[2 * i] < (double)(float)dist2[2 * i + 1] * ratio;
    match[i] = ok ? idx2[2 * i] : -1;
    if (ok) atomicAdd(count, 1);
}

// Number of train splits / rows per split for a problem set (host).
void orbfe_knn2_plan(int nqMax, int ntMax, int batch, int smCount, int* nsplit, int* splitLen) {
    const int qTiles = (nqMax + KNN_THREADS - 1) / KNN_THREADS;
    const long long ctas = (long long)qTiles * batch;
    const long long want = 4LL * smCount;                       // ~4 resident CTAs' worth of queries per SM
    int s = (int)((want + ctas - 1) / (ctas > 0 ? ctas : 1));
    const int maxSplit = (ntMax + KNN_TILE - 1) / KNN_TILE;
    s = s < 1 ? 1 : s > maxSplit ? maxSplit : s;
    if (s > 64) s = 64;
    int len = (ntMax + s - 1) / (s > 0 ? s : 1);
    len = (len + KNN_TILE - 1) / KNN_TILE * KNN_TILE;
    if (len < KNN_TILE) len = KNN_TILE;
    s = (ntMax + len - 1) / len;
    *nsplit = s < 1 ? 1 : s;
    *splitLen = len;
}

cudaError_t orbfe_launch_knn2(cudaStream_t st, const uint8_t* dq, long long qStride, const int* dNq, int nqMax, const uint8_t* dt,
                              long long tStride, const int* dNt, int ntMax, int batch, int nsplit, int splitLen, uint2* dPartial,
                              int32_t* dIdx2, int32_t* dDist2, long long* launches) {
    if (nqMax <= 0 || batch <= 0) return cudaSuccess;
    if (ntMax > (int)KNN_IDX_MASK) return cudaErrorInvalidValue;
    dim3 grid((nqMax + KNN_THREADS - 1) / KNN_THREADS, nsplit, batch);
    k_knn2<<<grid, KNN_THREADS, 0, st>>>(dq, qStride, dNq, nqMax, dt, tStride, dNt, ntMax, splitLen, dPartial, dIdx2, dDist2);
    ++*launches;
    if (nsplit > 1) {
        dim3 g2((nqMax + KNN_THREADS - 1) / KNN_THREADS, batch);
        k_knn2_merge<<<g2, KNN_THREADS, 0, st>>>(dPartial, dNq, nqMax, nsplit, dIdx2, dDist2);
        ++*launches;
    }
    return cudaGetLastError();
}

cudaError_t orbfe_launch_knn2_merge(cudaStream_t st, const uint2* dPartial, const int* dNq, int nqMax, int nsplit, int batch, int32_t* dIdx2,
                                    int32_t* dDist2, long long* launches) {
    dim3 g2((nqMax + KNN_THREADS - 1) / KNN_THREADS, batch);
    k_knn2_merge<<<g2, KNN_THREADS, 0, st>>>(dPartial, dNq, nqMax, nsplit, dIdx2, dDist2);
    ++*launches;
    return cudaGetLastError();
}

// ---- allocation-free 2-NN (no counterpart in the reference) -----------------
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it):
// the same k_knn2 / k_knn2_merge pair, with the per-split partial keys in a file-scope device array instead of the matcher's
// growable scratch.  Every GPU that loads the library holds the array (1 MiB) until exit.  All calls share it and it keeps the
// last call's contents; nothing in it has to start at zero (k_knn2 writes every (split, query) slot k_knn2_merge reads), so it
// is never cleared — but two calls that overlap on the GPU (different streams, no ordering) race on it.
#include <mutex>

#include "../../include/orbfe.h"
int orbfe_set_cuda_error(cudaError_t e, const char* what);   // orbfe_api.cu

#define NOALLOC_MAX_SPLIT 16
static __device__ __align__(256) uint2 g_knnScratch[(size_t)NOALLOC_MAX_SPLIT * ORBFE_NOALLOC_MAX_QUERIES];
static_assert(sizeof(uint2) * NOALLOC_MAX_SPLIT * ORBFE_NOALLOC_MAX_QUERIES == ORBFE_NOALLOC_SCRATCH_BYTES, "declared scratch size");

#define NOALLOC_MAX_DEVICES 64
static std::mutex g_noallocMu;
static uint2* g_knnScratchPtr[NOALLOC_MAX_DEVICES];    // device address of g_knnScratch on each device (the name itself is a host stand-in)
static int g_noallocSms[NOALLOC_MAX_DEVICES];

extern "C" int orbfe_hamming_knn2_noalloc(int device, void* stream, const uint8_t* d_q, int nq, const uint8_t* d_t, int nt,
                                          int32_t* d_idx2, int32_t* d_dist2) {
    if (device < 0 || device >= NOALLOC_MAX_DEVICES || nq < 0 || nt < 0 || (nq > 0 && (!d_q || !d_idx2 || !d_dist2)) || (nt > 0 && !d_t))
        return ORBFE_ERR_BAD_ARGUMENT;
    if (nq > ORBFE_NOALLOC_MAX_QUERIES || nt > ORBFE_NOALLOC_MAX_TRAIN) {
        orbfe_set_cuda_error(cudaErrorInvalidValue, nq > ORBFE_NOALLOC_MAX_QUERIES
                                                        ? "orbfe_hamming_knn2_noalloc: nq exceeds ORBFE_NOALLOC_MAX_QUERIES (8192), the static scratch is sized for that"
                                                        : "orbfe_hamming_knn2_noalloc: nt exceeds ORBFE_NOALLOC_MAX_TRAIN (65536)");
        return ORBFE_ERR_TOO_LARGE;
    }
    if (nq == 0) return ORBFE_OK;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return orbfe_set_cuda_error(e, "orbfe_hamming_knn2_noalloc: cudaSetDevice");
    uint2* scratch;
    int sms;
    {   // once per device, with that device current: the device address of the static array
        std::lock_guard<std::mutex> lock(g_noallocMu);
        if (!g_knnScratchPtr[device]) {
            uint2* p = nullptr;
            e = cudaGetSymbolAddress((void**)&p, g_knnScratch);
            if (e != cudaSuccess || !p) return orbfe_set_cuda_error(e != cudaSuccess ? e : cudaErrorUnknown, "orbfe_hamming_knn2_noalloc: cudaGetSymbolAddress(g_knnScratch)");
            int n = 0;
            e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device);
            if (e != cudaSuccess) return orbfe_set_cuda_error(e, "orbfe_hamming_knn2_noalloc: cudaDeviceGetAttribute");
            g_noallocSms[device] = n;
            g_knnScratchPtr[device] = p;
        }
        scratch = g_knnScratchPtr[device];
        sms = g_noallocSms[device];
    }
    int nsplit = 1, splitLen = nt;
    orbfe_knn2_plan(nq, nt, 1, sms, &nsplit, &splitLen);
    if (nsplit > NOALLOC_MAX_SPLIT) {      // fewer, longer splits: the scratch holds NOALLOC_MAX_SPLIT partial results per query
        splitLen = ((nt + NOALLOC_MAX_SPLIT - 1) / NOALLOC_MAX_SPLIT + KNN_TILE - 1) / KNN_TILE * KNN_TILE;
        nsplit = (nt + splitLen - 1) / splitLen;
    }
    long long launches = 0;
    e = orbfe_launch_knn2((cudaStream_t)stream, d_q, 0, nullptr, nq, d_t, 0, nullptr, nt, 1, nsplit, splitLen, scratch, d_idx2, d_dist2, &launches);
    if (e != cudaSuccess) return orbfe_set_cuda_error(e, "orbfe_hamming_knn2_noalloc: launch");
    return ORBFE_OK;
}

// Frame::UndistortKeyPoints (src/Frame.cc:730-763) = cv::undistortPoints with P = K: thread per keypoint, OpenCV's five
// fixed-point iterations in double, every product and sum rounded separately (no FMA contraction) so the float
// result is bit-identical to the library's.
__global__ void __launch_bounds__(128) k_undistort(orbfe_keypoint* __restrict__ kps, int n, const UndistortGeo g) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double u = (double)kps[i].x, v = (double)kps[i].y;
    const double ifx = __ddiv_rn(1.0, g.fx), ify = __ddiv_rn(1.0, g.fy);
    double x = __dmul_rn(__dsub_rn(u, g.cx), ifx), y = __dmul_rn(__dsub_rn(v, g.cy), ify);
    const double x0 = x, y0 = y;
    for (int j = 0; j < 5; ++j) {
        const double xx = __dmul_rn(x, x), yy = __dmul_rn(y, y), r2 = __dadd_rn(xx, yy);
        // (1 + ((k4*r2 + k5)*r2 + k6)*r2) with k4 = k5 = k6 = 0 is exactly 1
        const double den = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(g.k3, r2), g.k2), r2), g.k1), r2));
        const double icdist = __ddiv_rn(1.0, den);
        if (icdist < 0) { x = x0; y = y0; break; }
        // deltaX = 2*p1*x*y + p2*(r2 + 2*x*x);  deltaY = p1*(r2 + 2*y*y) + 2*p2*x*y   (left-to-right products)
        const double dX = __dadd_rn(__dmul_rn(__dmul_rn(__dmul_rn(2.0, g.p1), x), y), __dmul_rn(g.p2, __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, x), x))));
        const double dY = __dadd_rn(__dmul_rn(g.p1, __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, y), y))), __dmul_rn(__dmul_rn(__dmul_rn(2.0, g.p2), x), y));
        x = __dmul_rn(__dsub_rn(x0, dX), icdist);
        y = __dmul_rn(__dsub_rn(y0, dY), icdist);
    }
    kps[i].x = (float)__dadd_rn(__dmul_rn(g.fx, x), g.cx);
    kps[i].y = (float)__dadd_rn(__dmul_rn(g.fy, y), g.cy);
}

cudaError_t orbfe_launch_undistort(cudaStream_t st, orbfe_keypoint* dKps, int n, const UndistortGeo& g, long long* launches) {
    if (n <= 0) return cudaSuccess;
    k_undistort<<<(n + 127) / 128, 128, 0, st>>>(dKps, n, g);
    ++*launches;
    return cudaGetLastError();
}

// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:388-417): one warp per map point.  For every row i the
// lanes evaluate the Hamming distances to all rows into a 257-bin histogram (distances are 0..256), the median is
// the bin where the running count passes k = (n - 1) / 2; the row with the smallest median wins, first row on ties.
#define DIST_WARPS 4
__global__ void __launch_bounds__(DIST_WARPS * 32) k_distinctive(const uint8_t* __restrict__ desc, const int32_t* __restrict__ offsets,
                                                                 int npoints, int32_t* __restrict__ best) {
    __shared__ int hist[DIST_WARPS][288];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31, p = blockIdx.x * DIST_WARPS + w;
    if (p >= npoints) return;
    const int o = offsets[p], n = offsets[p + 1] - o;
    if (n <= 0) { if (lane == 0) best[p] = -1; return; }
    const uint4* D = reinterpret_cast<const uint4*>(desc + 32ll * o);
    const int k = (int)(0.5 * (n - 1));
    int bestMedian = 0x7fffffff, bestIdx = 0;
    int* h = hist[w];
    for (int i = 0; i < n; ++i) {
        for (int b = lane; b < 288; b += 32) h[b] = 0;
        __syncwarp();
        const uint4 a0 = __ldg(D + 2 * i), a1 = __ldg(D + 2 * i + 1);
        for (int j = lane; j < n; j += 32) {
            const uint4 b0 = __ldg(D + 2 * j), b1 = __ldg(D + 2 * j + 1);
            const int d = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                          __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
            atomicAdd(&h[d], 1);
        }
        __syncwarp();
        // lane owns bins 9*lane .. 9*lane+8; exclusive prefix of the lane totals, then the bin where count passes k
        int c[9], tot = 0;
#pragma unroll
        for (int b = 0; b < 9; ++b) { c[b] = h[9 * lane + b]; tot += c[b]; }
        int incl = tot;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, incl, s);
            if (lane >= s) incl += y;
        }
        int run = incl - tot, med = 0x7fffffff;
#pragma unroll
        for (int b = 0; b < 9; ++b) {
            if (med == 0x7fffffff && run <= k && k < run + c[b]) med = 9 * lane + b;
            run += c[b];
        }
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) med = min(med, __shfl_xor_sync(0xffffffffu, med, s));
        if (med < bestMedian) { bestMedian = med; bestIdx = i; }
        __syncwarp();
    }
    if (lane == 0) best[p] = bestIdx;
}

cudaError_t orbfe_launch_distinctive(cudaStream_t st, const uint8_t* dDesc, const int32_t* dOffsets, int npoints, int32_t* dBest,
                                     long long* launches) {
    if (npoints <= 0) return cudaSuccess;
    k_distinctive<<<(npoints + DIST_WARPS - 1) / DIST_WARPS, DIST_WARPS * 32, 0, st>>>(dDesc, dOffsets, npoints, dBest);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t orbfe_launch_ratio_filter(cudaStream_t st, const int32_t* dIdx2, const int32_t* dDist2, int nq, double ratio,
                                      int32_t* dMatch, int* dCount, long long* launches) {
    if (nq <= 0) return cudaSuccess;
    k_ratio_filter<<<(nq + 127) / 128, 128, 0, st>>>(dIdx2, dDist2, nq, ratio, dMatch, dCount);
    ++*launches;
    return cudaGetLastError();
}
