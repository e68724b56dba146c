// k_fast.cu — grid FAST: the cell loop of ORBextractor::ComputeKeyPointsOctTree
// (ORBextractor.cc:787-853) with cv::FAST(cell, keys, th, nonmaxSuppression=true), type 9_16.
//
// One CTA per (cell, frame) (two to four cells per CTA were measured and lost: 1.56 -> 1.67 / 1.77 / 1.91 ms per 512 frames).  The cell ROI (core + 6 px, clamped at the level's max border) is brought
// into shared memory by ONE TMA tiled load (cp.async.bulk.tensor.3d over a per-level tensor map of the
// pyramid arena, completion on an mbarrier) issued by thread 0 while the other threads clear the score
// map; the reference never lets FAST read outside the cell ROI, and neither do we (the TMA box may be
// larger than the cell, the extra bytes are never tested or used as ring pixels):
//   A. a packed 4-pixels-per-thread necessary test on two opposite ring pairs (VABSDIFF4 + a
//      carry-trick byte compare) compacts the tested area (3-px margin) to a survivor list, so the
//      expensive code never runs on mostly-idle warps;
//   B. survivors get the exact OpenCV score  max(max_k min(d_k..d_k+8), max_k min(-d_k..-d_k+8)) - 1
//      (packed 16-bit halves, VIMNMX3.U16x2); score >= t <=> a 9-arc exists, those are the corners;
//   C. 3x3 strict-greater non-max suppression against the cell's own score map (zero outside the
//      tested area, exactly like OpenCV's per-call row buffers);
//   D. if nothing survives at iniThFAST the same CTA redoes A-C at minThFAST (ORBextractor.cc:825-828).
// Survivors are appended to the level's candidate array as packed (x, y, response) words.  Their
// memory order is arbitrary; the reference's push order (cell-row-major, then FAST row-major) is a
// pure function of (x, y) and is re-derived where it matters (quadtree tie-break, debug tap).
#include <stdlib.h>

#include <algorithm>

#include "orbfe_internal.h"

#ifndef FAST_THREADS
#define FAST_THREADS 128
#endif

// Exact FAST score of OpenCV, max(max_k min(d_k..d_k+8), max_k min(-d_k..-d_k+8)) - 1 with d_k = centre -
// ring pixel k, computed on packed halves (DPX VIMNMX3.U16x2): ring pixel k becomes P_k = (256 + c - v_k) |
// (256 - c + v_k) << 16 (one IMAD: v_k * 0xFFFF + C0, both halves stay in 1..511), the minimum over every
// 9-arc is two levels of min3, the maximum over the 16 arcs a max3 tree; low half = 256 + max_k min9(d),
// high half = 256 + max_k min9(-d), so score = max(low, high) - 257.  A pixel is a corner at threshold t
// iff score >= t (score + 1 is the largest threshold margin some 9-arc clears), which is how the kernel
// decides corner-ness: no separate ring-mask pass.
__device__ __forceinline__ int fast_score_packed(const uint8_t* __restrict__ tile, int idx, const int (&off)[16]) {
    const uint32_t c = tile[idx];
    const uint32_t c0 = (256u + c) | (256u - c) << 16;
    uint32_t P[16], m3[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) P[k] = (uint32_t)tile[idx + off[k]] * 0xFFFFu + c0;
#pragma unroll
    for (int k = 0; k < 16; ++k) m3[k] = __vimin3_u16x2(P[k], P[(k + 1) & 15], P[(k + 2) & 15]);
    uint32_t best = 0;
#pragma unroll
    for (int k = 0; k < 16; k += 2) {
        const uint32_t a = __vimin3_u16x2(m3[k], m3[(k + 3) & 15], m3[(k + 6) & 15]);
        const uint32_t b = __vimin3_u16x2(m3[k + 1], m3[(k + 4) & 15], m3[(k + 7) & 15]);
        best = __vimax3_u16x2(best, a, b);
    }
    return (int)max(best & 0xFFFFu, best >> 16) - 257;
}

// Per-byte unsigned "a > t" of a packed u8x4 word against ONE threshold t (the same in all four bytes); the answer is
// bit 7 of each byte (other bits are garbage, mask with 0x80808080).  With a = a7:al and t = t7:tl, bit 7 of
// x = al + (127 - tl) is "al > tl" (the sum never carries across bytes), and  a > t  <=>  t7 ? (a7 & x7) : (a7 | x7).
// k127 = (127 - tl) in every byte, lowT = t < 128 ? ~0 : 0.  Three instructions: LOP3, IADD, LOP3.
__device__ __forceinline__ uint32_t gtu4_bit7(uint32_t a, uint32_t k127, uint32_t lowT) {
    const uint32_t x = (a & 0x7F7F7F7Fu) + k127;
    return (x & a) | ((x | a) & lowT);
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

#ifndef FAST_MINB
#define FAST_MINB 8
#endif
#ifndef FAST_PAD
#define FAST_PAD 0             // measured on B200 (profiles/r02g_variant_sweep.txt): 1.287 ms with the re-pitch vs 1.243 without - the extra pass and
#endif                         // barrier cost more than the bank conflicts they remove; kept selectable
// TPC = tile pitch known at compile time (64: every BASELINE shape, cells up to 49 px wide) or 0 = runtime boxW.  With a constant
// pitch the 16 ring offsets, the +-3-row word offsets of stage A and the index -> (x, y) split are immediates / shifts.
// Everything that depends only on the cell (tested-word range, edge masks, TMA coordinates) or only on the launch (list
// capacities, reciprocal of the pitch) comes precomputed from the host (CellDesc, kernel parameters): the prologue of a
// CTA that lives for ~3000 warp instructions used to spend 14 % of them on this arithmetic (profiles/r01o_*).
template <int TPC>
__global__ void __launch_bounds__(FAST_THREADS, FAST_MINB) k_fast_cells(const __grid_constant__ FastMaps maps,
                                                             const LevelGeom* __restrict__ geom,
                                                             const CellDesc* __restrict__ cells, uint32_t* __restrict__ cand,
                                                             int* __restrict__ candCount, long long candPerFrame,
                                                             int nlevels, int iniTh, int minTh, int boxW, int tileBytes, int listCap,
                                                             uint32_t invTp, int frame0) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t sBar;
    __shared__ int sCand, sCorners, sKept, sBase;
    const int frame = frame0 + blockIdx.y;
    // FAST_PAD (TPC = 64 only): TMA delivers a dense 64-byte-pitch tile, which puts rows r and r + 2 on the same banks - a
    // vertical edge (survivors at one column of consecutive rows) then makes every ring load a 16-way bank conflict, and stage
    // A's three-row word loads conflict 2-way.  The tile is re-pitched once to 68 bytes (17 words, odd) right after it lands;
    // the dense TMA destination aliases the two lists (written only after the re-pitch).
    constexpr bool kPad = FAST_PAD && TPC == 64;
    const int tp = kPad ? 68 : TPC ? TPC : boxW;         // pitch of the tile and of the score map the stages work on
    const int mapBytes = kPad ? (tileBytes >> 6) * 68 : tileBytes;
    uint8_t* tile = smem;
    uint8_t* score = tile + ((mapBytes + 127) & ~127);   // TMA destinations must be 128-byte aligned
    uint16_t* list1 = reinterpret_cast<uint16_t*>(score + ((mapBytes + 127) & ~127));      // survivors of the quick test
    uint16_t* list2 = list1 + listCap;                                       // corners
    uint32_t* kept = reinterpret_cast<uint32_t*>(list2 + listCap);
    uint8_t* landing = kPad ? reinterpret_cast<uint8_t*>(list1) : tile;      // where TMA writes
    const int tid = threadIdx.x;
    const int nw = tp >> 2;
    const CellDesc cd = cells[blockIdx.x];

    const int myTileBytes = (TPC ? TPC : boxW) * geom[cd.level].fastBoxH;     // this level's TMA box
    if (tid == 0) {
        mbar_init(&sBar, 1);
        mbar_expect_tx(&sBar, (uint32_t)myTileBytes);
        // TMA wants the innermost start coordinate on a 16-byte boundary: the box starts at the aligned column below the
        // cell (cd.tx); tile byte = cd.sh + column.
        tma_load_3d(landing, &maps.m[cd.level], cd.tx, cd.ty, frame, &sBar);
    }

    int off[16];
    {
        const int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
        const int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
#pragma unroll
        for (int k = 0; k < 16; ++k) off[k] = dy[k] * tp + dx[k];
    }
    const int nWords = cd.nWords, units = cd.units;
    const uint32_t inv = cd.inv;             // host: ceil(2^20 / nWords); u / nWords == (u * inv) >> 20 for u < 2^20 / nWords
    const int w0 = 3 * nw + cd.wLo;          // first tested word: row 3, word wLo
    const uint32_t* tw32 = reinterpret_cast<const uint32_t*>(tile);

    int t = iniTh;
    for (int pass = 0; pass < 2; ++pass) {
        if (tid == 0) { sCand = 0; sCorners = 0; sKept = 0; }
        for (int i = tid; i < (mapBytes + 15) / 16; i += FAST_THREADS) reinterpret_cast<uint4*>(score)[i] = make_uint4(0u, 0u, 0u, 0u);
        __syncthreads();                       // also orders the mbarrier init before everyone's wait
        if (pass == 0) {
            mbar_wait(&sBar, 0);
            if (kPad) {                        // dense 16-word rows -> 17-word rows; both sides conflict-free
                const uint32_t* src = reinterpret_cast<const uint32_t*>(landing);
                uint32_t* dst = reinterpret_cast<uint32_t*>(tile);
                for (int i = tid; i < myTileBytes / 4; i += FAST_THREADS) dst[i + (i >> 4)] = src[i];
                __syncthreads();               // the lists reuse the landing area
            }
        }
        // A: necessary condition on two opposite ring pairs, 4 pixels per step.  A 9-arc covers at least one pixel of every
        // opposite pair, so a corner has |c - v| > t for v0 or v8 AND for v4 or v12.  Bit 7 of every byte of m = "may be a corner".
        {
            const uint32_t k127 = (uint32_t)(127 - (t & 127)) * 0x01010101u, lowT = t < 128 ? 0xFFFFFFFFu : 0u;
            for (int u = tid; u < units; u += FAST_THREADS) {
                const int rr = (int)(((uint32_t)u * inv) >> 20), wj = u - rr * nWords;
                const int w = w0 + rr * nw + wj;
                const uint32_t c = tw32[w];
                const uint32_t up = tw32[w - 3 * nw], dn = tw32[w + 3 * nw];
                uint32_t m = (gtu4_bit7(__vabsdiffu4(c, up), k127, lowT) | gtu4_bit7(__vabsdiffu4(c, dn), k127, lowT)) & 0x80808080u;
                if (wj == 0) m &= cd.firstMask;                  // tested columns are [3, cw - 3): edge words carry fewer than 4
                if (wj == nWords - 1) m &= cd.lastMask;
                if (m == 0) continue;
                const uint32_t lf = tw32[w - 1], rt = tw32[w + 1];
                const uint32_t vl = __byte_perm(lf, c, 0x4321);      // bytes at column - 3
                const uint32_t vr = __byte_perm(c, rt, 0x6543);      // bytes at column + 3
                m &= gtu4_bit7(__vabsdiffu4(c, vl), k127, lowT) | gtu4_bit7(__vabsdiffu4(c, vr), k127, lowT);
                if (m == 0u) continue;
                // survivors are sparse (a few lanes per warp): each lane reserves its own slots with one shared-memory atomic.
                // Measured against the warp-aggregated form (5 ballots + prefix popcounts + one atomic per warp) on B200:
                // FAST 1.56 -> 1.42 ms per 512 frames; the same change in stages B and C: -> 1.31 ms.
                int base = atomicAdd(&sCand, __popc(m));
                const int idx = 4 * w;
                if (m & 0x00000080u) list1[base++] = (uint16_t)idx;
                if (m & 0x00008000u) list1[base++] = (uint16_t)(idx + 1);
                if (m & 0x00800000u) list1[base++] = (uint16_t)(idx + 2);
                if (m & 0x80000000u) list1[base] = (uint16_t)(idx + 3);
            }
        }
        __syncthreads();
        // B: exact score of every survivor; the corners (score >= t) go on list2 and into the score map
        const int n1 = sCand;
        for (int i0 = 0; i0 < n1; i0 += FAST_THREADS) {
            const int i = i0 + tid;
            bool corner = false;
            int idx = 0, s = 0;
            if (i < n1) {
                idx = list1[i];
                s = fast_score_packed(tile, idx, off);
                corner = s >= t;
            }
            const int slot = corner ? atomicAdd(&sCorners, 1) : -1;
            if (slot >= 0) { list2[slot] = (uint16_t)idx; score[idx] = (uint8_t)s; }
        }
        __syncthreads();
        const int nc = sCorners;
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
            const int slot = keep ? atomicAdd(&sKept, 1) : -1;
            if (slot >= 0) {
                const int y = kPad ? (int)(((uint32_t)idx * 15421u) >> 20) : TPC ? idx / TPC : (int)(((uint32_t)idx * invTp) >> 20);
                const int x = idx - y * tp - cd.sh;
                kept[slot] = (uint32_t)(x + cd.ox) | (uint32_t)(y + cd.oy) << 11 | (uint32_t)s << 22;
            }
        }
        __syncthreads();
        if (sKept > 0 || pass == 1) break;
        t = minTh;   // D: retry at the lower threshold
        __syncthreads();
    }
    const int nk = sKept;
    if (nk > 0) {
        if (tid == 0) sBase = atomicAdd(&candCount[frame * nlevels + cd.level], nk);
        __syncthreads();
        uint32_t* out = cand + (long long)frame * candPerFrame + geom[cd.level].candBase + sBase;
        for (int i = tid; i < nk; i += FAST_THREADS) out[i] = kept[i];
    }
}

// ---- warp-per-cell form ---------------------------------------------------------------------------------
// The same stages with ONE WARP per cell and no CTA-wide barrier at all: each warp owns a slice of the dynamic shared
// memory (tile, score map, corner list, a small survivor queue, its own mbarrier) and issues its own TMA load.  What this
// buys over the CTA-per-cell form (ncu, profiles/r02k_*): a cell has ~300 four-pixel units, ~170 survivors and ~40 corners,
// so 128 threads ran stage B at 66 % and stage C at 30 % lane occupancy, repeated the cell prologue in four warps and
// drained at six __syncthreads per cell.  Here stage A feeds a queue and the exact score runs only on FULL warps of 32
// survivors (plus one flush), stage C on 32-wide slices, the prologue once, and kept candidates go straight to the level's
// array (one global atomic per 32-wide NMS slice that kept something; the retry rule only needs the running count).
#ifndef FAST_WARP
#define FAST_WARP 1
#endif
#ifndef FAST_WARPS
#define FAST_WARPS 1       // measured (profiles/r02l_*, r02m_*): 1.22 / 1.36 / 1.58 / 1.72 ms per 512 frames with 1 / 2 / 4 / 8 warps per CTA - a CTA's
#endif                     // shared memory is only released when its slowest cell is done
#ifndef FAST_WMINB
#define FAST_WMINB 32
#endif
#ifndef FAST_CORNER_BALLOT
#define FAST_CORNER_BALLOT 1
#endif
#define FAST_QCAP 192      // survivor queue: < 32 left over + at most 4 x 32 new entries per step
#ifndef FAST_L2CAP
#define FAST_L2CAP 512     // corner list; a cell with more corners takes the map-scan form of stage C.  Measured (profiles/r02p_*, r02q_*):
#endif                     // 128 / 192 / 256 / 384 / 512 entries -> 1.19 / 1.15 / 1.10 / 1.08 / 1.07 ms per 512 frames (the scan is slow, textured cells overflow)
// Slots from ballots + prefix popcounts instead of the two shared-memory atomics were measured and lost (1.138 vs 1.122 ms,
// profiles/r02o_variants.txt), as did a tile pitch of 80 bytes for the bank spread (1.30 vs 1.24, profiles/r02m_variants.txt)
// and two units / two survivors per lane and step for more independent work per warp (1.047 vs 1.025, profiles/r02s_variants.txt).
//
// Shared memory per cell decides how many cells an SM works on at once (the kernel is latency-bound: 1.22 ms at 24 cells per
// SM, 1.10 at 29), so the score map is NOT tile-shaped: it covers the tested area plus a one-pixel zero frame, at its own
// pitch sp: score index = (y - 2) * sp + (x - 2) for cell-relative (x, y).
template <int TPC>
__global__ void __launch_bounds__(32 * FAST_WARPS, FAST_WMINB) k_fast_warp(const __grid_constant__ FastMaps maps,
                                                             const LevelGeom* __restrict__ geom,
                                                             const CellDesc* __restrict__ cells, uint32_t* __restrict__ cand,
                                                             int* __restrict__ candCount, long long candPerFrame,
                                                             int nlevels, int iniTh, int minTh, int boxW, int tileBytes, int listCap,
                                                             uint32_t invTp, int frame0, int cell0, int cellEnd, int warpBytes,
                                                             int sp, uint32_t invSp, int scoreStride) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int cell = cell0 + blockIdx.x * FAST_WARPS + wid;
    if (cell >= cellEnd) return;                       // whole warps leave; nothing below synchronises the CTA
    const int frame = frame0 + blockIdx.y;
    const int tp = TPC ? TPC : boxW;
    const int tileStride = (tileBytes + 127) & ~127;
    uint8_t* tile = smem + (size_t)wid * warpBytes;   // TMA destination, 128-byte aligned (warpBytes is a multiple of 128)
    uint8_t* score = tile + tileStride;
    uint16_t* list2 = reinterpret_cast<uint16_t*>(score + scoreStride);  // corners (score indices)
    uint16_t* queue = list2 + listCap;                                     // survivors of the quick test waiting for a full warp
    uint64_t* bar = reinterpret_cast<uint64_t*>(queue + FAST_QCAP);
    int* cnt = reinterpret_cast<int*>(bar + 1);                            // [0] queue fill, [1] corners
    const int nw = tp >> 2;
    const CellDesc cd = cells[cell];
    const int myBoxH = geom[cd.level].fastBoxH;                // this level's TMA box (tileBytes = the largest of the launch)
    const int myScoreWords = (sp * (myBoxH - 4) + 3) >> 2;

    if (lane == 0) {
        mbar_init(bar, 1);
        mbar_expect_tx(bar, (uint32_t)(tp * myBoxH));
        tma_load_3d(tile, &maps.m[cd.level], cd.tx, cd.ty, frame, bar);
    }
    int off[16];
    {
        const int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
        const int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
#pragma unroll
        for (int k = 0; k < 16; ++k) off[k] = dy[k] * tp + dx[k];
    }
    const int nWords = cd.nWords, units = cd.units;
    const uint32_t inv = cd.inv;
    const int w0 = 3 * nw + cd.wLo;
    const int sOff = 2 * sp + cd.sh + 2;           // score index = y * sp + (tile column) - sOff
    const uint32_t* tw32 = reinterpret_cast<const uint32_t*>(tile);
    uint32_t* out = cand + (long long)frame * candPerFrame + geom[cd.level].candBase;
    int* outCount = candCount + frame * nlevels + cd.level;
    const uint32_t ltMask = (1u << lane) - 1u;
    const uint32_t outBase = (uint32_t)(cd.ox + 2) | (uint32_t)(cd.oy + 2) << 11;

    // exact score of one queued survivor; a corner goes into the score map and onto the corner list
#if FAST_CORNER_BALLOT
    // corner slots from one ballot (the count lives in a register, warp-uniform): the compiler turns atomicAdd(&cnt[1], 1)
    // into a match / elect / popc / shfl sequence of a dozen instructions around the shared-memory atomic
#define FAST_SCORE_ONE(idxExpr, active)                                                                      \
    {                                                                                                        \
        int idx = 0, s = -1;                                                                                 \
        if (active) {                                                                                        \
            idx = (idxExpr);                                                                                 \
            s = fast_score_packed(tile, idx, off);                                                           \
        }                                                                                                    \
        const bool corner = s >= t;                                                                          \
        const uint32_t cb = __ballot_sync(0xffffffffu, corner);                                              \
        if (corner) {                                                                                        \
            const int y = TPC ? idx / TPC : (int)(((uint32_t)idx * invTp) >> 20);                            \
            const int sidx = y * sp + (idx - y * tp) - sOff;                                                 \
            const int slot = ncr + __popc(cb & ltMask);                                                      \
            if (slot < listCap) list2[slot] = (uint16_t)sidx;                                                \
            score[sidx] = (uint8_t)s;                                                                        \
        }                                                                                                    \
        ncr += __popc(cb);                                                                                   \
    }
#else
#define FAST_SCORE_ONE(idxExpr, active)                                                                      \
    if (active) {                                                                                            \
        const int idx = (idxExpr);                                                                           \
        const int s = fast_score_packed(tile, idx, off);                                                     \
        if (s >= t) {                                                                                        \
            const int y = TPC ? idx / TPC : (int)(((uint32_t)idx * invTp) >> 20);                            \
            const int sidx = y * sp + (idx - y * tp) - sOff;                                                 \
            const int slot = atomicAdd(&cnt[1], 1);                                                          \
            if (slot < listCap) list2[slot] = (uint16_t)sidx;                                                \
            score[sidx] = (uint8_t)s;                                                                        \
        }                                                                                                    \
    }
#endif
#define FAST_NMS(sidx, s)                                                                                    \
    ((s) > score[(sidx) - 1] && (s) > score[(sidx) + 1] && (s) > score[(sidx) - sp - 1] && (s) > score[(sidx) - sp] && \
     (s) > score[(sidx) - sp + 1] && (s) > score[(sidx) + sp - 1] && (s) > score[(sidx) + sp] && (s) > score[(sidx) + sp + 1])

    int t = iniTh;
    for (int pass = 0; pass < 2; ++pass) {
        if (lane == 0) { cnt[0] = 0; cnt[1] = 0; }
        int ncr = 0;
        (void)ncr;
        for (int i = lane; i < (myScoreWords + 3) >> 2; i += 32) reinterpret_cast<uint4*>(score)[i] = make_uint4(0u, 0u, 0u, 0u);
        __syncwarp();                          // also orders lane 0's mbarrier init before the other lanes' wait
        if (pass == 0) mbar_wait(bar, 0);
        const uint32_t k127 = (uint32_t)(127 - (t & 127)) * 0x01010101u, lowT = t < 128 ? 0xFFFFFFFFu : 0u;
        // A + B: quick test over the cell, exact score whenever 32 survivors are waiting
        for (int u0 = 0; u0 < units; u0 += 32) {
            const int u = u0 + lane;
            if (u < units) {
                const int rr = (int)(((uint32_t)u * inv) >> 20), wj = u - rr * nWords;
                const int w = w0 + rr * nw + wj;
                const uint32_t c = tw32[w];
                const uint32_t up = tw32[w - 3 * nw], dn = tw32[w + 3 * nw];
                uint32_t m = (gtu4_bit7(__vabsdiffu4(c, up), k127, lowT) | gtu4_bit7(__vabsdiffu4(c, dn), k127, lowT)) & 0x80808080u;
                if (wj == 0) m &= cd.firstMask;
                if (wj == nWords - 1) m &= cd.lastMask;
                if (m != 0u) {
                    const uint32_t lf = tw32[w - 1], rt = tw32[w + 1];
                    const uint32_t vl = __byte_perm(lf, c, 0x4321);
                    const uint32_t vr = __byte_perm(c, rt, 0x6543);
                    m &= gtu4_bit7(__vabsdiffu4(c, vl), k127, lowT) | gtu4_bit7(__vabsdiffu4(c, vr), k127, lowT);
                    if (m != 0u) {
                        int base = atomicAdd(&cnt[0], __popc(m));
                        const int idx = 4 * w;
                        if (m & 0x00000080u) queue[base++] = (uint16_t)idx;
                        if (m & 0x00008000u) queue[base++] = (uint16_t)(idx + 1);
                        if (m & 0x00800000u) queue[base++] = (uint16_t)(idx + 2);
                        if (m & 0x80000000u) queue[base] = (uint16_t)(idx + 3);
                    }
                }
            }
            __syncwarp();
            int qn = *reinterpret_cast<volatile int*>(&cnt[0]);
            if (qn >= 32) {
                do {
                    qn -= 32;
                    FAST_SCORE_ONE(queue[qn + lane], true)
                } while (qn >= 32);
                __syncwarp();                  // every lane has read its queue entry before the fill count drops
                if (lane == 0) cnt[0] = qn;
                __syncwarp();
            }
        }
        {
            const bool act = lane < *reinterpret_cast<volatile int*>(&cnt[0]);     // flush: fewer than 32 left
            FAST_SCORE_ONE(queue[lane], act)
        }
        __syncwarp();
        // C: non-max suppression; kept candidates go to the level's array in 32-wide slices
#if FAST_CORNER_BALLOT
        const int nc = ncr;
#else
        const int nc = *reinterpret_cast<volatile int*>(&cnt[1]);
#endif
        int keptTotal = 0;
        if (nc > listCap) {                    // rare: the list overflowed, walk the score map itself
            int mine = 0;
            for (int wi = lane; wi < myScoreWords; wi += 32) {
                const uint32_t sw = reinterpret_cast<const uint32_t*>(score)[wi];
                if (sw == 0u) continue;
                for (int k = 0; k < 4; ++k) {
                    const int s = (int)((sw >> (8 * k)) & 0xFFu), sidx = 4 * wi + k;
                    if (s == 0 || !FAST_NMS(sidx, s)) continue;
                    const int ys = (int)(((uint32_t)sidx * invSp) >> 20), xs = sidx - ys * sp;
                    out[atomicAdd(outCount, 1)] = (outBase + (uint32_t)xs + ((uint32_t)ys << 11)) | (uint32_t)s << 22;
                    ++mine;
                }
            }
            keptTotal = __reduce_add_sync(0xffffffffu, mine);
        } else {
            for (int i0 = 0; i0 < nc; i0 += 32) {
                const int i = i0 + lane;
                bool keep = false;
                int sidx = 0, s = 0;
                if (i < nc) {
                    sidx = list2[i];
                    s = score[sidx];
                    keep = FAST_NMS(sidx, s);
                }
                const uint32_t bal = __ballot_sync(0xffffffffu, keep);
                if (bal != 0u) {
                    int gbase = 0;
                    if (lane == 0) gbase = atomicAdd(outCount, __popc(bal));
                    gbase = __shfl_sync(0xffffffffu, gbase, 0);
                    if (keep) {
                        const int ys = (int)(((uint32_t)sidx * invSp) >> 20), xs = sidx - ys * sp;
                        out[gbase + __popc(bal & ltMask)] = (outBase + (uint32_t)xs + ((uint32_t)ys << 11)) | (uint32_t)s << 22;
                    }
                    keptTotal += __popc(bal);
                }
            }
        }
        if (keptTotal > 0 || pass == 1) break;
        t = minTh;   // D: retry at the lower threshold
        __syncwarp();
    }
#undef FAST_SCORE_ONE
#undef FAST_NMS
}

#ifndef FAST_SIDE
#define FAST_SIDE 2       // measured (profiles/r05c_fast_side.txt, r05d): 0 / 1 / 2 = 200.8 / 201.6 / 201.8 k frames/s, e2e 125.2 / 126.3 / 126.8 k
#endif
#ifndef FAST_SMALL_ROWS
#define FAST_SMALL_ROWS 48     // levels whose cells are at most this tall share a launch with a small shared-memory footprint
#endif
static cudaError_t launch_fast_warp(orbfe_handle* h, int frame0, int batch, cudaStream_t sideStream) {
    // ORBFE_FAST_L2CAP (environment, tests only) shrinks the corner list so that ordinary frames take the overflow path
    static const int capEnv = [] { const char* e = getenv("ORBFE_FAST_L2CAP"); const int v = e ? atoi(e) : 0; return v >= 4 ? (v & ~3) : FAST_L2CAP; }();
    const uint32_t invTp = ((1u << 20) + h->fastBoxW - 1) / h->fastBoxW;
    const int nl = h->p.nlevels;
    // Cells are stored level by level.  Almost all of them are ~44 rows tall, but the coarsest levels have a few of up to ~60
    // rows; sizing every CTA's shared memory for those costs a quarter of the occupancy, so consecutive levels of the same
    // class (small / tall) go out as one launch with the footprint of their own tallest box.
    // ... unless the whole launch fits the GPU at once anyway (single frames, stereo pairs): then occupancy is not the issue,
    // the second launch's ~15 us of cell latency is, and everything goes out together with the largest footprint.
    const bool oneLaunch = (long long)h->nCellsTotal * batch <= 24LL * h->smCount;
    int nLaunch = 0;
    for (int l = 0; l < nl;) {
        const bool small = oneLaunch || h->geom[l].fastBoxH <= FAST_SMALL_ROWS;
        int l1 = l, maxH = 0;
        while (l1 < nl && (oneLaunch || (h->geom[l1].fastBoxH <= FAST_SMALL_ROWS) == small)) { maxH = std::max(maxH, h->geom[l1].fastBoxH); ++l1; }
        const int cell0 = h->geom[l].cellBase, cellEnd = h->geom[l1 - 1].cellBase + h->geom[l1 - 1].nCells;
        l = l1;
        if (cellEnd <= cell0) continue;
        const int tileBytes = h->fastBoxW * maxH;
        const int tileStride = (tileBytes + 127) & ~127;
        const int listCap = std::min(((h->fastMaxCw - 6) * (maxH - 6) + 3) & ~3, capEnv);     // at most: every tested pixel a corner
        const int sp = (h->fastMaxCw - 4 + 3) & ~3;                       // score map: tested area + a one-pixel frame
        const int scoreStride = (sp * (maxH - 4) + 15) & ~15;
        const uint32_t invSp = ((1u << 20) + sp - 1) / sp;               // sidx / sp == (sidx * invSp) >> 20 for sidx < 2^14
        const int warpBytes = (tileStride + scoreStride + 2 * listCap + 2 * FAST_QCAP + 16 + 127) & ~127;
        const size_t smem = (size_t)warpBytes * FAST_WARPS;
        dim3 grid((cellEnd - cell0 + FAST_WARPS - 1) / FAST_WARPS, batch);
        cudaStream_t st = (sideStream && nLaunch > 0) ? sideStream : h->cur;
        ++nLaunch;
#define ORBFE_FAST_WLAUNCH(TPC)                                                                                                  \
        do {                                                                                                                     \
            const cudaError_t ea = orbfe_ensure_dyn_smem(reinterpret_cast<const void*>(k_fast_warp<TPC>), smem);                 \
            if (ea != cudaSuccess) return ea;                                                                                    \
            k_fast_warp<TPC><<<grid, 32 * FAST_WARPS, smem, st>>>(h->fastMaps, h->dGeom, h->dCells, h->dCand, h->dCandCount,     \
                                                                      h->candPerFrame, nl, h->p.ini_th_fast, h->p.min_th_fast,   \
                                                                      h->fastBoxW, tileBytes, listCap, invTp, frame0, cell0, cellEnd, warpBytes, \
                                                                      sp, invSp, scoreStride);                                     \
        } while (0)
        if (h->fastBoxW == 64) ORBFE_FAST_WLAUNCH(64);
        else if (h->fastBoxW == 80) ORBFE_FAST_WLAUNCH(80);
        else ORBFE_FAST_WLAUNCH(0);
#undef ORBFE_FAST_WLAUNCH
        h->launches++;
    }
    return cudaGetLastError();
}

// ORBFE_FAST_SIDE (read once): 0 = every launch in line on the compute stream; 1 = FAST's launches after the first (the tall-cell
// class) on the side stream, under the main launch; 2 = the pyramid levels only those cells read go there too.
int orbfe_fast_side_mode() {
    static const int mode = [] { const char* e = getenv("ORBFE_FAST_SIDE"); return e ? atoi(e) : FAST_SIDE; }();
    return FAST_WARP ? mode : 0;
}

// First level of the trailing run of tall-cell levels that k_fast_warp launches separately; nlevels when there is no such run
// (one launch for everything, or no tall class at the end).
int orbfe_fast_tall_from(const orbfe_handle* h, int batch) {
    const int nl = h->p.nlevels;
    if (!FAST_WARP || (long long)h->nCellsTotal * batch <= 24LL * h->smCount) return nl;
    int l = nl;
    while (l > 0 && h->geom[l - 1].fastBoxH > FAST_SMALL_ROWS) --l;
    return l > 0 ? l : nl;
}

cudaError_t orbfe_clear_candidates(orbfe_handle* h, int frame0, int batch) {
    return cudaMemsetAsync(h->dCandCount + (size_t)frame0 * h->p.nlevels, 0, sizeof(int) * (size_t)batch * h->p.nlevels, h->cur);
}

// The candidate counters must have been cleared (orbfe_clear_candidates) on a stream both h->cur and sideStream are ordered after.
cudaError_t orbfe_launch_fast(orbfe_handle* h, int frame0, int batch, cudaStream_t sideStream) {
    if (FAST_WARP) return launch_fast_warp(h, frame0, batch, sideStream);
    const int tileBytes = h->fastBoxW * h->fastBoxH;
    const size_t tileStride = ((size_t)tileBytes + 127) & ~(size_t)127;
    // maxCellBytes counts two unpadded maps; FAST_PAD re-pitches both to 68 bytes and 128-byte-aligns them
    const size_t smem = (size_t)h->maxCellBytes + 2 * (tileStride - (size_t)tileBytes) + (FAST_PAD && h->fastBoxW == 64 ? 2 * ((size_t)h->fastBoxH * 4 + 128) : 0);
    const int listCap = ((h->fastBoxW - 6) * (h->fastBoxH - 6) + 1) & ~1;
    const uint32_t invTp = ((1u << 20) + h->fastBoxW - 1) / h->fastBoxW;   // idx / tp == (idx * invTp) >> 20 (a tile has < 2^14 bytes)
    dim3 grid(h->nCellsTotal, batch);
#define ORBFE_FAST_LAUNCH(TPC)                                                                                                   \
    do {                                                                                                                         \
        const cudaError_t ea = orbfe_ensure_dyn_smem(reinterpret_cast<const void*>(k_fast_cells<TPC>), smem);                    \
        if (ea != cudaSuccess) return ea;                                                                                        \
        k_fast_cells<TPC><<<grid, FAST_THREADS, smem, h->cur>>>(h->fastMaps, h->dGeom, h->dCells, h->dCand, h->dCandCount,       \
                                                                h->candPerFrame, h->p.nlevels, h->p.ini_th_fast, h->p.min_th_fast, \
                                                                h->fastBoxW, tileBytes, listCap, invTp, frame0);                 \
    } while (0)
    if (h->fastBoxW == 64) ORBFE_FAST_LAUNCH(64);
    else if (h->fastBoxW == 80) ORBFE_FAST_LAUNCH(80);
    else ORBFE_FAST_LAUNCH(0);
#undef ORBFE_FAST_LAUNCH
    h->launches++;
    return cudaGetLastError();
}
