// k_fast.cu — grid FAST: the cell loop of ORBextractor::ComputeKeyPointsOctTree
// (ORBextractor.cc:787-853) with cv::FAST(cell, keys, th, nonmaxSuppression=true), type 9_16.
//
// One CTA per (cell, frame).  The cell ROI (core + 6 px, clamped at the level's max border) is staged
// in shared memory; the reference never lets FAST read outside that ROI, and neither do we:
//   A. A1: a packed 4-pixels-per-thread necessary test on two opposite ring pairs (VABSDIFF4 + a
//      carry-trick byte compare) compacts the tested area (3-px margin) to a survivor list; A2: the
//      survivors get the 16-bit brighter/darker ring masks, pixels with a 9-contiguous arc go on the
//      corner list.  Each stage works on a dense list, so the expensive code never runs on mostly-idle warps;
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

// Per-byte unsigned "a > b" of two packed u8x4 words; the answer is bit 7 of each byte (other bits are
// garbage, mask with 0x80808080).  With a = a7:al, b = b7:bl:  a > b  <=>  (a7 & ~b7) | (a7 == b7 & al > bl),
// and bit 7 of (b | 0x80) - (a & 0x7f) is "bl >= al" (128 + bl - al never borrows across bytes).
__device__ __forceinline__ uint32_t gtu4_bit7(uint32_t a, uint32_t b) {
    const uint32_t ge_ba = (b | 0x80808080u) - (a & 0x7F7F7F7Fu);
    return (a & ~b) | (~(a ^ b) & ~ge_ba);
}

__global__ void __launch_bounds__(FAST_THREADS) k_fast_cells(const uint8_t* __restrict__ pyr,
                                                             const LevelGeom* __restrict__ geom,
                                                             const CellDesc* __restrict__ cells, uint32_t* __restrict__ cand,
                                                             int* __restrict__ candCount, long long candPerFrame,
                                                             int nlevels, int iniTh, int minTh) {
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ int sCand, sCorners, sKept, sBase;
    const CellDesc cd = cells[blockIdx.x];
    const int frame = blockIdx.y;
    const LevelGeom& g = geom[cd.level];
    const int cw = cd.cw, ch = cd.ch;
    const int sh = cd.x0 & 3;                // the tile keeps the global word alignment: tile byte = sh + column
    const int tp = (sh + cw + 3) & ~3;       // tile pitch (bytes), whole words
    const int tw = cw - 6, th = ch - 6;      // tested area
    if (tw <= 0 || th <= 0) return;
    uint8_t* tile = smem;
    uint8_t* score = tile + tp * ch;
    uint16_t* list1 = reinterpret_cast<uint16_t*>(score + tp * ch);          // survivors of the quick test
    uint16_t* list2 = list1 + ((tw * th + 1) & ~1);                          // corners
    uint32_t* kept = reinterpret_cast<uint32_t*>(list2 + ((tw * th + 1) & ~1));
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int nw = tp >> 2;

    {
        const uint32_t* src = reinterpret_cast<const uint32_t*>(pyr + g.base + (long long)frame * g.plane +
                                                                (long long)(cd.y0 + ORBFE_EDGE) * g.pitch + ORBFE_XO + cd.x0 - sh);
        const int pw = g.pitch >> 2;
        uint32_t* tw32 = reinterpret_cast<uint32_t*>(tile);
        for (int r = ty; r < ch; r += FAST_THREADS / 16)
            for (int c = tx; c < nw; c += 16) tw32[r * nw + c] = __ldg(src + (long long)r * pw + c);
    }

    int off[16];
    {
        const int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
        const int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
#pragma unroll
        for (int k = 0; k < 16; ++k) off[k] = dy[k] * tp + dx[k];
    }
    const int wLo = (sh + 3) >> 2, wHi = (sh + cw - 4) >> 2;   // words holding tested columns [3, cw-3)

    int t = iniTh;
    for (int pass = 0; pass < 2; ++pass) {
        if (tid == 0) { sCand = 0; sCorners = 0; sKept = 0; }
        for (int i = tid; i < (tp * ch) / 4; i += FAST_THREADS) reinterpret_cast<uint32_t*>(score)[i] = 0;
        __syncthreads();
        // A1: necessary condition on two opposite ring pairs, 4 pixels per step.  A 9-arc covers at least
        // one pixel of every opposite pair, so a corner has |c - v| > t for v0 or v8 AND for v4 or v12.
        {
            const uint32_t t4 = (uint32_t)t * 0x01010101u;
            const uint32_t* tw32 = reinterpret_cast<const uint32_t*>(tile);
            for (int r = 3 + ty; r < ch - 3; r += FAST_THREADS / 16)
                for (int wi = wLo + tx; wi <= wHi; wi += 16) {
                    const uint32_t c = tw32[r * nw + wi];
                    const uint32_t up = tw32[(r - 3) * nw + wi], dn = tw32[(r + 3) * nw + wi];
                    uint32_t m = (gtu4_bit7(__vabsdiffu4(c, up), t4) | gtu4_bit7(__vabsdiffu4(c, dn), t4)) & 0x80808080u;
                    if (m == 0) continue;
                    const uint32_t lf = tw32[r * nw + wi - 1], rt = tw32[r * nw + wi + 1];
                    const uint32_t vl = __byte_perm(lf, c, 0x4321);      // bytes at column - 3
                    const uint32_t vr = __byte_perm(c, rt, 0x6543);      // bytes at column + 3
                    m &= gtu4_bit7(__vabsdiffu4(c, vl), t4) | gtu4_bit7(__vabsdiffu4(c, vr), t4);
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int col = 4 * wi + k - sh;
                        if ((m >> (8 * k + 7) & 1u) && col >= 3 && col < cw - 3)
                            list1[atomicAdd(&sCand, 1)] = (uint16_t)(r * tp + 4 * wi + k);
                    }
                }
        }
        __syncthreads();
        // A2: full ring masks for the survivors
        const int n1 = sCand;
        for (int i = tid; i < n1; i += FAST_THREADS) {
            const int idx = list1[i];
            const int c = tile[idx];
            const int lo = c - t, hi = c + t;
            uint32_t md = 0, mb = 0;
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                const int v = tile[idx + off[k]];
                md |= (uint32_t)(v < lo) << k;
                mb |= (uint32_t)(v > hi) << k;
            }
            if (has9(md) || has9(mb)) list2[atomicAdd(&sCorners, 1)] = (uint16_t)idx;
        }
        __syncthreads();
        // B: exact scores
        const int nc = sCorners;
        for (int i = tid; i < nc; i += FAST_THREADS) {
            const int idx = list2[i];
            const int c = tile[idx];
            int d[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) d[k] = c - (int)tile[idx + off[k]];
            score[idx] = (uint8_t)fast_score(d);
        }
        __syncthreads();
        // C: non-max suppression
        for (int i = tid; i < nc; i += FAST_THREADS) {
            const int idx = list2[i];
            const int s = score[idx];
            const bool keep = s > score[idx - 1] && s > score[idx + 1] && s > score[idx - tp - 1] && s > score[idx - tp] &&
                              s > score[idx - tp + 1] && s > score[idx + tp - 1] && s > score[idx + tp] &&
                              s > score[idx + tp + 1];
            if (keep) {
                const int y = idx / tp, x = idx - y * tp - sh;
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
