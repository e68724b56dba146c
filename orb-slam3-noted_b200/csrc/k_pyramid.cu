// k_pyramid.cu — image pyramid: ORBextractor::ComputePyramid (ORBextractor.cc:1152-1177).
//
// Level 0 is the input with a 19-px reflect-101 border; level l > 0 is the OpenCV 8-bit bilinear
// resize (11-bit fixed-point taps, see ResizeTap) of level l-1's ROI, again with the border.  A
// border pixel equals the ROI pixel at the reflected coordinate, so every thread simply evaluates
// the level function at reflect101(x), reflect101(y): one kernel writes ROI and border together
// and the level never has to be re-read to make its border.
//
// Layout: each level lives in a bordered buffer whose ROI column 0 sits at byte ORBFE_XO (32) of a
// 128-byte-multiple pitch, so ROI words are 4-byte aligned for the uchar4 stores below.  One thread
// produces one 32-bit word (4 pixels) of a bordered row; blockIdx.y = row, blockIdx.z = frame.
#include "orbfe_internal.h"

__device__ __forceinline__ int reflect101(int i, int n) {
    // valid for -n < i < 2n-1, which the 19-px border guarantees for every level with n >= 20
    i = i < 0 ? -i : i;
    return i >= n ? 2 * n - 2 - i : i;
}

// One thread = one 16-byte chunk of a bordered row; a warp = 8 chunks (128 bytes) x 4 rows, so only the
// warps over the left / right border columns diverge into the per-byte reflected path.  Chunks inside the
// ROI (ROI column 0 sits on a 32-byte boundary, so they start at a multiple of 16 source bytes) are one
// 128-bit load + store when the source row is 16-byte aligned, four 32-bit loads when 4-byte aligned.
#ifndef PYR_L0_ROWS
#define PYR_L0_ROWS 4          // rows per thread (32 apart): the chunk classification and the column arithmetic are shared
#endif
__global__ void __launch_bounds__(256) k_pyr_level0(const uint8_t* __restrict__ in, long long stride,
                                                    long long frameStride, uint8_t* __restrict__ out, LevelGeom g, int frame0) {
    const int chunk = blockIdx.x * 8 + (threadIdx.x & 7), rowA = blockIdx.z * (32 * PYR_L0_ROWS) + (threadIdx.x >> 3);
    if (chunk >= (g.pitch >> 4)) return;
    const int x0 = chunk * 16 - ORBFE_XO;
    const uint8_t* frame = in + (long long)blockIdx.y * frameStride;
    uint8_t* dst = out + g.base + (long long)(frame0 + blockIdx.y) * g.plane;
    const bool inside = x0 >= 0 && x0 + 16 <= g.w;
    const int al = (int)((reinterpret_cast<uintptr_t>(frame) | (uintptr_t)stride) & 15);     // every row start shares this alignment
    if (inside && al == 0) {      // the common case: every load of the thread issued before its first store
        uint4 v[PYR_L0_ROWS];
#pragma unroll
        for (int k = 0; k < PYR_L0_ROWS; ++k) {
            const int y = reflect101(min(rowA + 32 * k, g.rows - 1) - ORBFE_EDGE, g.h);
            v[k] = __ldg(reinterpret_cast<const uint4*>(frame + (long long)y * stride + x0));
        }
#pragma unroll
        for (int k = 0; k < PYR_L0_ROWS; ++k) {
            const int row = rowA + 32 * k;
            if (row < g.rows) reinterpret_cast<uint4*>(dst + (long long)row * g.pitch)[chunk] = v[k];
        }
        return;
    }
#pragma unroll
    for (int k = 0; k < PYR_L0_ROWS; ++k) {
        const int row = rowA + 32 * k;
        if (row >= g.rows) break;
        const int y = reflect101(row - ORBFE_EDGE, g.h);
        const uint8_t* src = frame + (long long)y * stride;
        uint4 v;
        if (inside && al == 0) {
            v = __ldg(reinterpret_cast<const uint4*>(src + x0));
        } else if (inside && (al & 3) == 0) {
            const uint32_t* s4 = reinterpret_cast<const uint32_t*>(src + x0);
            v = make_uint4(__ldg(s4), __ldg(s4 + 1), __ldg(s4 + 2), __ldg(s4 + 3));
        } else {
            uint32_t w[4] = {0u, 0u, 0u, 0u};
#pragma unroll
            for (int b = 0; b < 16; ++b) {
                const int x = x0 + b;
                if (x >= -ORBFE_EDGE && x < g.w + ORBFE_EDGE) w[b >> 2] |= (uint32_t)__ldg(src + reflect101(x, g.w)) << (8 * (b & 3));
            }
            v = make_uint4(w[0], w[1], w[2], w[3]);
        }
        reinterpret_cast<uint4*>(dst + (long long)row * g.pitch)[chunk] = v;
    }
}

// Colour input: cv::cvtColor(..., RGB2GRAY / BGR2GRAY / RGBA2GRAY / BGRA2GRAY) of OpenCV 4.x for 8-bit images,
// (R * 9798 + G * 19235 + B * 3735 + (1 << 14)) >> 15, fused with the border copy.  One thread = one 32-bit word of
// a bordered row; BPP = 3 or 4, rFirst = the first channel is red.
template <int BPP>
__global__ void __launch_bounds__(128) k_pyr_level0_color(const uint8_t* __restrict__ in, long long stride, long long frameStride,
                                                          uint8_t* __restrict__ out, LevelGeom g, int frame0, int rFirst) {
    const int word = blockIdx.x * blockDim.x + threadIdx.x;
    const int row = blockIdx.y;
    if (word * 4 >= g.pitch) return;
    const int y = reflect101(row - ORBFE_EDGE, g.h);
    const uint8_t* src = in + (long long)blockIdx.z * frameStride + (long long)y * stride;
    const int x0 = word * 4 - ORBFE_XO;
    const int cr = rFirst ? 9798 : 3735, cb = rFirst ? 3735 : 9798;
    uint32_t v = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int x = x0 + k;
        if (x >= -ORBFE_EDGE && x < g.w + ORBFE_EDGE) {
            const uint8_t* p = src + (long long)reflect101(x, g.w) * BPP;
            const int grey = ((int)__ldg(p) * cr + (int)__ldg(p + 1) * 19235 + (int)__ldg(p + 2) * cb + (1 << 14)) >> 15;
            v |= (uint32_t)grey << (8 * k);
        }
    }
    uint8_t* dst = out + g.base + (long long)(frame0 + blockIdx.z) * g.plane + (long long)row * g.pitch;
    reinterpret_cast<uint32_t*>(dst)[word] = v;
}

// gs = source level geometry, gd = destination level geometry.
__global__ void __launch_bounds__(128) k_pyr_resize(uint8_t* __restrict__ pyr, LevelGeom gs, LevelGeom gd,
                                                    const ResizeTap* __restrict__ taps, int frame0) {
    const int word = blockIdx.x * blockDim.x + threadIdx.x;
    const int row = blockIdx.y;
    if (word * 4 >= gd.pitch) return;
    const int y = reflect101(row - ORBFE_EDGE, gd.h);
    const ResizeTap ty = taps[gd.tabY + y];
    const uint8_t* sbase = pyr + gs.base + (long long)(frame0 + blockIdx.z) * gs.plane + (long long)ORBFE_EDGE * gs.pitch + ORBFE_XO;
    const uint8_t* S0 = sbase + (long long)ty.s0 * gs.pitch;
    const uint8_t* S1 = sbase + (long long)ty.s1 * gs.pitch;
    const int x0 = word * 4 - ORBFE_XO;
    uint32_t v = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int x = x0 + k;
        if (x >= -ORBFE_EDGE && x < gd.w + ORBFE_EDGE) {
            const ResizeTap tx = taps[gd.tabX + reflect101(x, gd.w)];
            const int r0 = (int)S0[tx.s0] * tx.a0 + (int)S0[tx.s1] * tx.a1;
            const int r1 = (int)S1[tx.s0] * tx.a0 + (int)S1[tx.s1] * tx.a1;
            const int o = ((((int)ty.a0 * (r0 >> 4)) >> 16) + (((int)ty.a1 * (r1 >> 4)) >> 16) + 2) >> 2;
            v |= (uint32_t)(o & 0xFF) << (8 * k);
        }
    }
    uint8_t* dst = pyr + gd.base + (long long)(frame0 + blockIdx.z) * gd.plane + (long long)row * gd.pitch;
    reinterpret_cast<uint32_t*>(dst)[word] = v;
}

// ---- fast path: column plans ---------------------------------------------------------------------------
// The x side of the resize does not depend on the row, so the host precomputes, per 32-bit word of a
// bordered destination row, a ColPlan: where the (at most 8-byte) source windows of its column pairs
// start, the PRMT selectors that gather (S[s0], S[s1]) of two columns into one word, the packed int16
// weights (a0 | a1 << 16) of the four columns and a byte mask for columns outside the 19-px border.
// A thread owns one destination word and walks PYR_TH rows with the plan in registers: per row
// 8 word loads (L1-resident), 4 PRMT, 8 DP2A (the 11-bit horizontal blend), the vertical blend, 1 store.
// One warp = a 128-byte x PYR_TH-row tile, 4 tiles per CTA.
#ifndef PYR_TH
#define PYR_TH 16
#endif

#ifndef PYR_MINB
#define PYR_MINB 1
#endif
__global__ void __launch_bounds__(128, PYR_MINB) k_pyr_resize_plan(uint8_t* __restrict__ pyr, LevelGeom gs, LevelGeom gd,
                                                         const ResizeTap* __restrict__ taps, const ColPlan* __restrict__ plans,
                                                         int tilesX, int totalTiles, int frame0) {
    const int tile = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31, frame = frame0 + blockIdx.y;
    if (tile >= totalTiles) return;
    const int ty = tile / tilesX, tx = tile - ty * tilesX;
    const int word = tx * 32 + lane;
    const ColPlan cp = plans[gd.planBase + word];
    const int row0 = ty * PYR_TH, row1 = min(row0 + PYR_TH, gd.rows);
    const uint32_t* sbase = reinterpret_cast<const uint32_t*>(pyr + gs.base + (long long)frame * gs.plane +
                                                              (long long)ORBFE_EDGE * gs.pitch + ORBFE_XO);
    const int spw = gs.pitch >> 2;
    uint32_t* dst = reinterpret_cast<uint32_t*>(pyr + gd.base + (long long)frame * gd.plane);     // uniform; the word goes into the offset
    const int dpw = gd.pitch >> 2;
    // lane j prepares the vertical tap of row0 + j once (reflection, table read, word offsets of the two
    // source rows, weights pre-shifted so that (a * x) >> 16 is one IMAD.HI); the row loop broadcasts it
    // All addresses below are a uniform 64-bit base plus an UNSIGNED 32-bit word offset (one IMAD.WIDE.U32 each): with signed
    // ints the compiler sign-extended and rebuilt every pointer in four instructions, a quarter of the row loop (r02 SASS).
    unsigned myOff0 = 0, myOff1 = 0;
    uint32_t myA0 = 0, myA1 = 0;
    if (lane < PYR_TH && row0 + lane < row1) {
        const ResizeTap t = taps[gd.tabY + reflect101(row0 + lane - ORBFE_EDGE, gd.h)];
        myOff0 = (unsigned)t.s0 * (unsigned)spw; myOff1 = (unsigned)t.s1 * (unsigned)spw;
        myA0 = (uint32_t)t.a0 << 16; myA1 = (uint32_t)t.a1 << 16;
    }
    const unsigned b01 = (unsigned)cp.base01, b23 = (unsigned)cp.base23;
    unsigned doff = (unsigned)row0 * (unsigned)dpw + (unsigned)word;
#pragma unroll 4
    for (int row = row0; row < row1; ++row) {
        const int j = row - row0;
        const unsigned o0 = __shfl_sync(0xffffffffu, myOff0, j), o1 = __shfl_sync(0xffffffffu, myOff1, j);
        const uint32_t a0s = __shfl_sync(0xffffffffu, myA0, j), a1s = __shfl_sync(0xffffffffu, myA1, j);
        const uint32_t* q0a = sbase + (o0 + b01);
        const uint32_t* q0b = sbase + (o0 + b23);
        const uint32_t* q1a = sbase + (o1 + b01);
        const uint32_t* q1b = sbase + (o1 + b23);
        const uint32_t p0a = __byte_perm(__ldg(q0a), __ldg(q0a + 1), cp.sel & 0xFFFF);
        const uint32_t p0b = __byte_perm(__ldg(q0b), __ldg(q0b + 1), cp.sel >> 16);
        const uint32_t p1a = __byte_perm(__ldg(q1a), __ldg(q1a + 1), cp.sel & 0xFFFF);
        const uint32_t p1b = __byte_perm(__ldg(q1b), __ldg(q1b + 1), cp.sel >> 16);
        int r0[4], r1[4];
        r0[0] = __dp2a_lo(cp.a[0], p0a, 0u); r0[1] = __dp2a_hi(cp.a[1], p0a, 0u);
        r0[2] = __dp2a_lo(cp.a[2], p0b, 0u); r0[3] = __dp2a_hi(cp.a[3], p0b, 0u);
        r1[0] = __dp2a_lo(cp.a[0], p1a, 0u); r1[1] = __dp2a_hi(cp.a[1], p1a, 0u);
        r1[2] = __dp2a_lo(cp.a[2], p1b, 0u); r1[3] = __dp2a_hi(cp.a[3], p1b, 0u);
        uint32_t o[4];     // ((a0 * (r0 >> 4)) >> 16) + ((a1 * (r1 >> 4)) >> 16) + 2) >> 2, always <= 255
#pragma unroll
        for (int k = 0; k < 4; ++k) o[k] = (__umulhi(a0s, (uint32_t)r0[k] >> 4) + __umulhi(a1s, (uint32_t)r1[k] >> 4) + 2u) >> 2;
        const uint32_t v = __byte_perm(__byte_perm(o[0], o[1], 0x0040), __byte_perm(o[2], o[3], 0x0040), 0x5410);
        dst[doff] = v & cp.keep;
        doff += (unsigned)dpw;
    }
}

// ---- the same arithmetic with the source staged by TMA (the default since round 2) ---------------------------------------
// north_star prescribes "staged through shared memory with TMA tile loads" for the pyramid.  A CTA of four warps covers 128
// destination bytes x 64 destination rows; every source row / column its taps touch lies in one box of the parent level
// (PyrTmaLevel: per tile column the first source word, per row group the first source row), which thread 0 fetches with one
// cp.async.bulk.tensor.3d; the row loop then reads the two source rows from shared memory instead of L1.  Stores are the same
// full-line 32-bit stores (a warp writes 128 contiguous bytes per row).  Measured against k_pyr_resize_plan: see
// profiles/r03m_pyr_tma.txt - 0.659 -> 0.581 ms per 512 frames, so this is the default; ORBFE_PYR_TMA=0 selects the plain kernel.
namespace {
__device__ __forceinline__ uint32_t pyr_s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
}
#define PYRTMA_THREADS (ORBFE_PYRTMA_ROWS / PYR_TH * 32)
#ifndef PYRTMA_TAPS_SMEM
#define PYRTMA_TAPS_SMEM 1   // measured (profiles/r06b_variants.txt): pyramid 0.550 -> 0.504 ms per 512 frames, 201.7 -> 205.3 k frames/s
#endif
#ifndef PYRTMA_MINB
#define PYRTMA_MINB 1
#endif
__global__ void __launch_bounds__(PYRTMA_THREADS, PYRTMA_MINB) k_pyr_resize_tma(const __grid_constant__ FastMaps maps, uint8_t* __restrict__ pyr, LevelGeom gd,
                                                        const ResizeTap* __restrict__ taps, const ColPlan* __restrict__ plans,
                                                        const PyrTmaPlan* __restrict__ plan, int level, int frame0) {
    extern __shared__ __align__(128) uint8_t tileBytes[];
    __shared__ __align__(8) uint64_t bar;
    const PyrTmaLevel& L = plan->lv[level];
    const int tx = blockIdx.x, ty = blockIdx.y, frame = frame0 + blockIdx.z;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int boxWords = L.boxWords, colLo = L.colLo[tx], rowLo = L.rowLo[ty];
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(pyr_s32(&bar)), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(pyr_s32(&bar)), "r"((uint32_t)(boxWords * 4 * L.boxH)) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
                         pyr_s32(tileBytes)),
                     "l"(&maps.m[level]), "r"(ORBFE_XO + colLo * 4), "r"(ORBFE_EDGE + rowLo), "r"(frame), "r"(pyr_s32(&bar))
                     : "memory");
    }
    const int word = tx * 32 + lane;
    const ColPlan cp = plans[gd.planBase + word];
    const int row0 = ty * ORBFE_PYRTMA_ROWS + warp * PYR_TH, row1 = min(row0 + PYR_TH, gd.rows);
#if PYRTMA_TAPS_SMEM
    // The row's two source offsets and two weights come from shared memory as ONE broadcast 128-bit load instead of four shuffles
    // (the kernel sits at 80 % of the shared-memory / L1 wavefront rate: 8 loads + 4 shuffles + 1 store per row), and the
    // offsets are in bytes so that a load address is one add.
    __shared__ __align__(16) uint4 rowTaps[PYRTMA_THREADS / 32][PYR_TH];
    if (lane < PYR_TH) {
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (row0 + lane < row1) {
            const ResizeTap t = taps[gd.tabY + reflect101(row0 + lane - ORBFE_EDGE, gd.h)];
            v.x = (unsigned)(t.s0 - rowLo) * (unsigned)boxWords * 4u; v.y = (unsigned)(t.s1 - rowLo) * (unsigned)boxWords * 4u;
            v.z = (uint32_t)t.a0 << 16; v.w = (uint32_t)t.a1 << 16;
        }
        rowTaps[warp][lane] = v;
    }
    const unsigned b01 = (unsigned)(cp.base01 - colLo) * 4u, b23 = (unsigned)(cp.base23 - colLo) * 4u;
    const uint8_t* S = tileBytes;
#define PYRTMA_LD(off) (*reinterpret_cast<const uint32_t*>(S + (off)))
#define PYRTMA_LD1(off) (*reinterpret_cast<const uint32_t*>(S + (off) + 4))
#else
    unsigned myOff0 = 0, myOff1 = 0;
    uint32_t myA0 = 0, myA1 = 0;
    if (lane < PYR_TH && row0 + lane < row1) {
        const ResizeTap t = taps[gd.tabY + reflect101(row0 + lane - ORBFE_EDGE, gd.h)];
        myOff0 = (unsigned)(t.s0 - rowLo) * (unsigned)boxWords; myOff1 = (unsigned)(t.s1 - rowLo) * (unsigned)boxWords;
        myA0 = (uint32_t)t.a0 << 16; myA1 = (uint32_t)t.a1 << 16;
    }
    const unsigned b01 = (unsigned)(cp.base01 - colLo), b23 = (unsigned)(cp.base23 - colLo);
    const uint32_t* S = reinterpret_cast<const uint32_t*>(tileBytes);
#define PYRTMA_LD(off) S[off]
#define PYRTMA_LD1(off) S[(off) + 1]
#endif
    uint32_t* dst = reinterpret_cast<uint32_t*>(pyr + gd.base + (long long)frame * gd.plane);
    const unsigned dpw = (unsigned)(gd.pitch >> 2);
    unsigned doff = (unsigned)row0 * dpw + (unsigned)word;
    __syncthreads();                   // the barrier is initialised (and the row taps are written) before anyone goes on
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "PW_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra PD_%=;\n"
        "bra PW_%=;\n"
        "PD_%=:\n"
        "}\n" ::"r"(pyr_s32(&bar)),
        "r"(0)
        : "memory");
    if (row0 >= gd.rows) return;
#pragma unroll 4
    for (int row = row0; row < row1; ++row) {
        const int j = row - row0;
#if PYRTMA_TAPS_SMEM
        const uint4 rt = rowTaps[warp][j];
        const unsigned o0 = rt.x, o1 = rt.y;
        const uint32_t a0s = rt.z, a1s = rt.w;
#else
        const unsigned o0 = __shfl_sync(0xffffffffu, myOff0, j), o1 = __shfl_sync(0xffffffffu, myOff1, j);
        const uint32_t a0s = __shfl_sync(0xffffffffu, myA0, j), a1s = __shfl_sync(0xffffffffu, myA1, j);
#endif
        const uint32_t p0a = __byte_perm(PYRTMA_LD(o0 + b01), PYRTMA_LD1(o0 + b01), cp.sel & 0xFFFF);
        const uint32_t p0b = __byte_perm(PYRTMA_LD(o0 + b23), PYRTMA_LD1(o0 + b23), cp.sel >> 16);
        const uint32_t p1a = __byte_perm(PYRTMA_LD(o1 + b01), PYRTMA_LD1(o1 + b01), cp.sel & 0xFFFF);
        const uint32_t p1b = __byte_perm(PYRTMA_LD(o1 + b23), PYRTMA_LD1(o1 + b23), cp.sel >> 16);
        int r0[4], r1[4];
        r0[0] = __dp2a_lo(cp.a[0], p0a, 0u); r0[1] = __dp2a_hi(cp.a[1], p0a, 0u);
        r0[2] = __dp2a_lo(cp.a[2], p0b, 0u); r0[3] = __dp2a_hi(cp.a[3], p0b, 0u);
        r1[0] = __dp2a_lo(cp.a[0], p1a, 0u); r1[1] = __dp2a_hi(cp.a[1], p1a, 0u);
        r1[2] = __dp2a_lo(cp.a[2], p1b, 0u); r1[3] = __dp2a_hi(cp.a[3], p1b, 0u);
        uint32_t o[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) o[k] = (__umulhi(a0s, (uint32_t)r0[k] >> 4) + __umulhi(a1s, (uint32_t)r1[k] >> 4) + 2u) >> 2;
        const uint32_t v = __byte_perm(__byte_perm(o[0], o[1], 0x0040), __byte_perm(o[2], o[3], 0x0040), 0x5410);
        dst[doff] = v & cp.keep;
        doff += dpw;
    }
#undef PYRTMA_LD
#undef PYRTMA_LD1
}

// Levels [lFirst, lEnd) on stream st (level l needs level l - 1: the caller orders the streams).
cudaError_t orbfe_launch_pyramid_levels(orbfe_handle* h, const uint8_t* dImages, int frame0, int batch, long long stride,
                                        long long frameStride, int lFirst, int lEnd, cudaStream_t st) {
    if (lFirst == 0 && lEnd > 0) {
        const LevelGeom& g = h->geom[0];
        if (h->inFormat == ORBFE_FORMAT_GRAY8) {
            dim3 grid(((g.pitch >> 4) + 7) / 8, batch, (g.rows + 32 * PYR_L0_ROWS - 1) / (32 * PYR_L0_ROWS));
            k_pyr_level0<<<grid, 256, 0, st>>>(dImages, stride, frameStride, h->dPyr, g, frame0);
        } else {
            dim3 grid((g.pitch / 4 + 127) / 128, g.rows, batch);
            const int rFirst = h->inFormat == ORBFE_FORMAT_RGB8 || h->inFormat == ORBFE_FORMAT_RGBA8;
            if (h->inFormat <= ORBFE_FORMAT_BGR8)
                k_pyr_level0_color<3><<<grid, 128, 0, st>>>(dImages, stride, frameStride, h->dPyr, g, frame0, rFirst);
            else
                k_pyr_level0_color<4><<<grid, 128, 0, st>>>(dImages, stride, frameStride, h->dPyr, g, frame0, rFirst);
        }
        h->launches++;
    }
    for (int l = lFirst > 1 ? lFirst : 1; l < lEnd; ++l) {
        const LevelGeom& gd = h->geom[l];
        if (gd.planBase >= 0 && h->pyrUseTma && h->pyrTma.lv[l].ok) {
            const PyrTmaLevel& L = h->pyrTma.lv[l];
            const size_t smem = (size_t)L.boxWords * 4 * L.boxH;
            dim3 grid(L.tilesX, L.groupsY, batch);
            k_pyr_resize_tma<<<grid, PYRTMA_THREADS, smem, st>>>(h->pyrMaps, h->dPyr, gd, h->dTaps, h->dPlans, h->dPyrTma, l, frame0);
        } else if (gd.planBase >= 0) {
            const int tilesX = gd.pitch / 128, total = tilesX * ((gd.rows + PYR_TH - 1) / PYR_TH);
            dim3 grid((total + 3) / 4, batch);
            k_pyr_resize_plan<<<grid, 128, 0, st>>>(h->dPyr, h->geom[l - 1], gd, h->dTaps, h->dPlans, tilesX, total, frame0);
        } else {   // scale factors whose 2-column source window exceeds 8 bytes: generic kernel
            dim3 grid((gd.pitch / 4 + 127) / 128, gd.rows, batch);
            k_pyr_resize<<<grid, 128, 0, st>>>(h->dPyr, h->geom[l - 1], gd, h->dTaps, frame0);
        }
        h->launches++;
    }
    return cudaGetLastError();
}

cudaError_t orbfe_launch_pyramid(orbfe_handle* h, const uint8_t* dImages, int frame0, int batch, int width, int height,
                                 long long stride, long long frameStride) {
    (void)width; (void)height;
    return orbfe_launch_pyramid_levels(h, dImages, frame0, batch, stride, frameStride, 0, h->p.nlevels, h->cur);
}
