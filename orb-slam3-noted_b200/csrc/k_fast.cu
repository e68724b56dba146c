// k_fast.cu — grid FAST: the cell loop of ORBextractor::ComputeKeyPointsOctTree
// (ORBextractor.cc:787-853) with cv::FAST(cell, keys, th, nonmaxSuppression=true), type 9_16.
//
// One CTA per (cell, frame).  The cell ROI (core + 6 px, clamped at the level's max border) is staged
// in shared memory; the reference never lets FAST read outside that ROI, and neither do we:
//   A. every tested pixel (3-px margin) gets the 16-bit brighter/darker ring masks; pixels with a
//      9-contiguous arc go on a shared corner list (keeps the expensive score off divergent lanes);
//   B. list entries get the exact OpenCV score  max(max_k min(d_k..d_k+8), max_k min(-d_k..-d_k+8)) - 1;
//   C. 3x3 strict-greater non-max suppression against the cell's own score map (zero outside the
//      tested area, exactly like OpenCV's per-call row buffers);
//   D. if nothing survives at iniThFAST the same CTA redoes A-C at minThFAST (ORBextractor.cc:825-828).
// Survivors are appended to the level's candidate array as packed (x, y, response) words.  Their
// memory order is arbitrary; the reference's push order (cell-row-major, then FAST row-major) is a
// pure function of (x, y) and is re-derived where it matters (quadtree tie-break, debug tap).
#include "orbfe_internal.h"

#define FAST_THREADS 128

__device__ __forceinline__ bool has9(uint32_t m) {
    m |= m << 16;
    uint32_t r = m & (m >> 1);
    r &= r >> 2;
    r &= r >> 4;
    r &= m >> 8;
    return (r & 0xFFFFu) != 0;
}

// Exact FAST score from the 16 centre-minus-ring differences.
__device__ __forceinline__ int fast_score(const int (&d)[16]) {
    int mn[16], mx[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) { mn[k] = min(d[k], d[(k + 1) & 15]); mx[k] = max(d[k], d[(k + 1) & 15]); }
    int mn4[16], mx4[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) { mn4[k] = min(mn[k], mn[(k + 2) & 15]); mx4[k] = max(mx[k], mx[(k + 2) & 15]); }
    int a = -256, b = 256;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int m8 = min(mn4[k], mn4[(k + 4) & 15]);
        const int x8 = max(mx4[k], mx4[(k + 4) & 15]);
        a = max(a, min(m8, d[(k + 8) & 15]));
        b = min(b, max(x8, d[(k + 8) & 15]));
    }
    return max(a, -b) - 1;
}

__global__ void __launch_bounds__(FAST_THREADS) k_fast_cells(const uint8_t* __restrict__ pyr,
                                                             const LevelGeom* __restrict__ geom,
                                                             const CellDesc* __restrict__ cells, uint32_t* __restrict__ cand,
                                                             int* __restrict__ candCount, long long candPerFrame,
                                                             int nlevels, int iniTh, int minTh) {
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ int sCorners, sKept, sBase;
    const CellDesc cd = cells[blockIdx.x];
    const int frame = blockIdx.y;
    const LevelGeom& g = geom[cd.level];
    const int cw = cd.cw, ch = cd.ch;
    const int tp = (cw + 3) & ~3;            // tile pitch
    const int tw = cw - 6, th = ch - 6;      // tested area
    if (tw <= 0 || th <= 0) return;
    uint8_t* tile = smem;
    uint8_t* score = tile + tp * ch;
    uint16_t* list = reinterpret_cast<uint16_t*>(score + tp * ch);
    uint32_t* kept = reinterpret_cast<uint32_t*>(list + ((tw * th + 1) & ~1));
    const int tid = threadIdx.x;

    const uint8_t* src = pyr + g.base + (long long)frame * g.plane + (long long)(cd.y0 + ORBFE_EDGE) * g.pitch + ORBFE_XO + cd.x0;
    for (int r = tid / 64; r < ch; r += FAST_THREADS / 64)
        for (int c = tid % 64; c < cw; c += 64) tile[r * tp + c] = __ldg(src + (long long)r * g.pitch + c);

    int off[16];
    {
        const int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
        const int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
#pragma unroll
        for (int k = 0; k < 16; ++k) off[k] = dy[k] * tp + dx[k];
    }

    int t = iniTh;
    for (int pass = 0; pass < 2; ++pass) {
        if (tid == 0) { sCorners = 0; sKept = 0; }
        for (int i = tid; i < (tp * ch) / 4; i += FAST_THREADS) reinterpret_cast<uint32_t*>(score)[i] = 0;
        __syncthreads();
        // A: ring masks
        for (int p = tid; p < tw * th; p += FAST_THREADS) {
            const int py = p / tw, px = p - py * tw;
            const int idx = (py + 3) * tp + px + 3;
            const int c = tile[idx];
            const int lo = c - t, hi = c + t;
            const int v0 = tile[idx + off[0]], v8 = tile[idx + off[8]];
            bool dark = (v0 < lo) | (v8 < lo), bright = (v0 > hi) | (v8 > hi);
            if (!(dark | bright)) continue;
            const int v4 = tile[idx + off[4]], v12 = tile[idx + off[12]];
            dark = dark & ((v4 < lo) | (v12 < lo));
            bright = bright & ((v4 > hi) | (v12 > hi));
            if (!(dark | bright)) continue;
            uint32_t md = 0, mb = 0;
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                const int v = tile[idx + off[k]];
                md |= (uint32_t)(v < lo) << k;
                mb |= (uint32_t)(v > hi) << k;
            }
            if (has9(md) || has9(mb)) list[atomicAdd(&sCorners, 1)] = (uint16_t)idx;
        }
        __syncthreads();
        // B: exact scores
        const int nc = sCorners;
        for (int i = tid; i < nc; i += FAST_THREADS) {
            const int idx = list[i];
            const int c = tile[idx];
            int d[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) d[k] = c - (int)tile[idx + off[k]];
            score[idx] = (uint8_t)fast_score(d);
        }
        __syncthreads();
        // C: non-max suppression
        for (int i = tid; i < nc; i += FAST_THREADS) {
            const int idx = list[i];
            const int s = score[idx];
            const bool keep = s > score[idx - 1] && s > score[idx + 1] && s > score[idx - tp - 1] && s > score[idx - tp] &&
                              s > score[idx - tp + 1] && s > score[idx + tp - 1] && s > score[idx + tp] &&
                              s > score[idx + tp + 1];
            if (keep) {
                const int y = idx / tp, x = idx - y * tp;
                kept[atomicAdd(&sKept, 1)] = (uint32_t)(x + cd.ox) | (uint32_t)(y + cd.oy) << 11 | (uint32_t)s << 22;
            }
        }
        __syncthreads();
        if (sKept > 0 || pass == 1) break;
        t = minTh;   // D: retry at the lower threshold
        __syncthreads();
    }
    const int nk = sKept;
    if (nk == 0) return;
    if (tid == 0) sBase = atomicAdd(&candCount[frame * nlevels + cd.level], nk);
    __syncthreads();
    uint32_t* out = cand + (long long)frame * candPerFrame + g.candBase + sBase;
    for (int i = tid; i < nk; i += FAST_THREADS) out[i] = kept[i];
}

cudaError_t orbfe_launch_fast(orbfe_handle* h, int batch) {
    cudaError_t e = cudaMemsetAsync(h->dCandCount, 0, sizeof(int) * (size_t)batch * h->p.nlevels, h->stream);
    if (e != cudaSuccess) return e;
    dim3 grid(h->nCellsTotal, batch);
    k_fast_cells<<<grid, FAST_THREADS, h->maxCellBytes, h->stream>>>(h->dPyr, h->dGeom, h->dCells, h->dCand, h->dCandCount,
                                                                     h->candPerFrame, h->p.nlevels, h->p.ini_th_fast,
                                                                     h->p.min_th_fast);
    h->launches++;
    return cudaGetLastError();
}
