// k_fast.cu — grid FAST: the cell loop of ORBextractor::ComputeKeyPointsOctTree
// (ORBextractor.cc:787-853) with cv::FAST(cell, keys, th, nonmaxSuppression=true), type 9_16.
//
// One CTA per (cell, frame).  The cell ROI (core + 6 px, clamped at the level's max border) is brought
// into shared memory by ONE TMA tiled load (cp.async.bulk.tensor.3d over a per-level tensor map of the
// pyramid arena, completion on an mbarrier) issued by thread 0 while the other threads clear the score
// map; the reference never lets FAST read outside the cell ROI, and neither do we (the TMA box may be
// larger than the cell, the extra bytes are never tested or used as ring pixels):
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

// ---- TMA + mbarrier helpers (sm_100a PTX) -----------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");   // make the init visible to the async proxy
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(phase)
        : "memory");
}
// 3-D tiled TMA load: box (boxW x boxH x 1) of the level's tensor map at (x, y, frame) -> dense smem tile.
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, int x, int y, int z, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar))
        : "memory");
}

// Warp-aggregated append: lanes with pred get consecutive slots of a shared counter (one atomic per warp).
__device__ __forceinline__ int warp_push(bool pred, int* counter) {
    const uint32_t m = __ballot_sync(0xffffffffu, pred);
    if (m == 0) return -1;
    const int lane = threadIdx.x & 31;
    int base = 0;
    if (lane == 0) base = atomicAdd(counter, __popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    return pred ? base + __popc(m & ((1u << lane) - 1u)) : -1;
}

__global__ void __launch_bounds__(FAST_THREADS) k_fast_cells(const __grid_constant__ FastMaps maps,
                                                             const LevelGeom* __restrict__ geom,
                                                             const CellDesc* __restrict__ cells, uint32_t* __restrict__ cand,
                                                             int* __restrict__ candCount, long long candPerFrame,
                                                             int nlevels, int iniTh, int minTh, int boxW, int boxH, int frame0) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t sBar;
    __shared__ int sCand, sCorners, sKept, sBase;
    const CellDesc cd = cells[blockIdx.x];
    const int frame = frame0 + blockIdx.y;
    const LevelGeom& g = geom[cd.level];
    const int cw = cd.cw, ch = cd.ch;
    const int tp = boxW;                     // dense tile as delivered by TMA
    // TMA wants the innermost start coordinate on a 16-byte boundary: load from the aligned column below
    // the cell; tile byte = sh + column.
    const int sh = (ORBFE_XO + cd.x0) & 15;
    const int tw = cw - 6, th = ch - 6;      // tested area
    if (tw <= 0 || th <= 0) return;
    const int tileBytes = boxW * boxH;
    uint8_t* tile = smem;
    uint8_t* score = tile + tileBytes;
    uint16_t* list1 = reinterpret_cast<uint16_t*>(score + tileBytes);        // survivors of the quick test
    const int listCap = ((boxW - 6) * (boxH - 6) + 1) & ~1;
    uint16_t* list2 = list1 + listCap;                                       // corners
    uint32_t* kept = reinterpret_cast<uint32_t*>(list2 + listCap);
    const int tid = threadIdx.x;
    const int nw = tp >> 2;

    if (tid == 0) {
        mbar_init(&sBar, 1);
        mbar_expect_tx(&sBar, (uint32_t)tileBytes);
        tma_load_3d(tile, &maps.m[cd.level], ORBFE_XO + cd.x0 - sh, ORBFE_EDGE + cd.y0, frame, &sBar);
    }

    int off[16];
    {
        const int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
        const int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
#pragma unroll
        for (int k = 0; k < 16; ++k) off[k] = dy[k] * tp + dx[k];
    }
    // A1 work units: (row r in [3, ch-3)) x (word wi in [wLo, wHi]); word wi holds columns 4wi-sh .. 4wi-sh+3
    const int wLo = (sh + 3) >> 2, wHi = (sh + cw - 4) >> 2, nWordsA1 = wHi - wLo + 1;
    const int units = th * nWordsA1;
    const uint32_t inv = ((1u << 20) + nWordsA1 - 1) / nWordsA1;   // u / nWordsA1 == (u * inv) >> 20 for u < 2^20 / nWordsA1

    int t = iniTh;
    for (int pass = 0; pass < 2; ++pass) {
        if (tid == 0) { sCand = 0; sCorners = 0; sKept = 0; }
        for (int i = tid; i < tileBytes / 4; i += FAST_THREADS) reinterpret_cast<uint32_t*>(score)[i] = 0;
        __syncthreads();                       // also orders the mbarrier init before everyone's wait
        if (pass == 0) mbar_wait(&sBar, 0);
        // A1: necessary condition on two opposite ring pairs, 4 pixels per step.  A 9-arc covers at least
        // one pixel of every opposite pair, so a corner has |c - v| > t for v0 or v8 AND for v4 or v12.
        {
            const uint32_t t4 = (uint32_t)t * 0x01010101u;
            const uint32_t* tw32 = reinterpret_cast<const uint32_t*>(tile);
            for (int u0 = 0; u0 < units; u0 += FAST_THREADS) {
                const int u = u0 + tid;
                uint32_t vb = 0;
                int idx = 0;
                if (u < units) {
                    const int rr = (int)(((uint32_t)u * inv) >> 20), wi = wLo + u - rr * nWordsA1, r = rr + 3;
                    idx = r * tp + 4 * wi;
                    const uint32_t c = tw32[r * nw + wi];
                    const uint32_t up = tw32[(r - 3) * nw + wi], dn = tw32[(r + 3) * nw + wi];
                    uint32_t m = (gtu4_bit7(__vabsdiffu4(c, up), t4) | gtu4_bit7(__vabsdiffu4(c, dn), t4)) & 0x80808080u;
                    if (m != 0) {
                        const uint32_t lf = tw32[r * nw + wi - 1], rt = tw32[r * nw + wi + 1];
                        const uint32_t vl = __byte_perm(lf, c, 0x4321);      // bytes at column - 3
                        const uint32_t vr = __byte_perm(c, rt, 0x6543);      // bytes at column + 3
                        m &= gtu4_bit7(__vabsdiffu4(c, vl), t4) | gtu4_bit7(__vabsdiffu4(c, vr), t4);
                        // nibble of survivors, restricted to tested columns [3, cw-3)
                        vb = (m >> 7 & 1u) | (m >> 14 & 2u) | (m >> 21 & 4u) | (m >> 28 & 8u);
                        const int c0 = 4 * wi - sh;                      // column of byte 0 (may be < 0)
                        const int kLo = max(3 - c0, 0), kHi = min(cw - 3 - c0, 4);   // valid bytes [kLo, kHi)
                        vb &= ((1u << kHi) - 1u) & ~((1u << kLo) - 1u);
                    }
                }
                // warp-aggregated append of up to 4 survivors per lane
                const uint32_t b0 = __ballot_sync(0xffffffffu, vb & 1u), b1 = __ballot_sync(0xffffffffu, vb & 2u);
                const uint32_t b2 = __ballot_sync(0xffffffffu, vb & 4u), b3 = __ballot_sync(0xffffffffu, vb & 8u);
                const int total = __popc(b0) + __popc(b1) + __popc(b2) + __popc(b3);
                if (total == 0) continue;
                const uint32_t lt = (1u << (tid & 31)) - 1u;
                int base = 0;
                if ((tid & 31) == 0) base = atomicAdd(&sCand, total);
                base = __shfl_sync(0xffffffffu, base, 0) + __popc(b0 & lt) + __popc(b1 & lt) + __popc(b2 & lt) + __popc(b3 & lt);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (vb >> k & 1u) list1[base++] = (uint16_t)(idx + k);
            }
        }
        __syncthreads();
        // A2: full ring masks for the survivors
        const int n1 = sCand;
        for (int i0 = 0; i0 < n1; i0 += FAST_THREADS) {
            const int i = i0 + tid;
            bool corner = false;
            int idx = 0;
            if (i < n1) {
                idx = list1[i];
                const int c = tile[idx];
                const int lo = c - t, hi = c + t;
                uint32_t md = 0, mb = 0;
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    const int v = tile[idx + off[k]];
                    md |= (uint32_t)(v < lo) << k;
                    mb |= (uint32_t)(v > hi) << k;
                }
                corner = has9(md) || has9(mb);
            }
            const int slot = warp_push(corner, &sCorners);
            if (slot >= 0) list2[slot] = (uint16_t)idx;
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
        for (int i0 = 0; i0 < nc; i0 += FAST_THREADS) {
            const int i = i0 + tid;
            bool keep = false;
            int idx = 0, s = 0;
            if (i < nc) {
                idx = list2[i];
                s = score[idx];
                keep = s > score[idx - 1] && s > score[idx + 1] && s > score[idx - tp - 1] && s > score[idx - tp] &&
                       s > score[idx - tp + 1] && s > score[idx + tp - 1] && s > score[idx + tp] && s > score[idx + tp + 1];
            }
            const int slot = warp_push(keep, &sKept);
            if (slot >= 0) {
                const int y = idx / tp, x = idx - y * tp - sh;
                kept[slot] = (uint32_t)(x + cd.ox) | (uint32_t)(y + cd.oy) << 11 | (uint32_t)s << 22;
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

cudaError_t orbfe_launch_fast(orbfe_handle* h, int frame0, int batch) {
    cudaError_t e = cudaMemsetAsync(h->dCandCount + (size_t)frame0 * h->p.nlevels, 0, sizeof(int) * (size_t)batch * h->p.nlevels, h->stream);
    if (e != cudaSuccess) return e;
    static size_t configured = 0;
    if ((size_t)h->maxCellBytes > 48 * 1024 && (size_t)h->maxCellBytes > configured) {
        e = cudaFuncSetAttribute(k_fast_cells, cudaFuncAttributeMaxDynamicSharedMemorySize, h->maxCellBytes);
        if (e != cudaSuccess) return e;
        configured = h->maxCellBytes;
    }
    dim3 grid(h->nCellsTotal, batch);
    k_fast_cells<<<grid, FAST_THREADS, h->maxCellBytes, h->stream>>>(h->fastMaps, h->dGeom, h->dCells, h->dCand, h->dCandCount,
                                                                     h->candPerFrame, h->p.nlevels, h->p.ini_th_fast,
                                                                     h->p.min_th_fast, h->fastBoxW, h->fastBoxH, frame0);
    h->launches++;
    return cudaGetLastError();
}
