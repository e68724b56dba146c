// k_describe.cu — output assembly, orientation and descriptors.
//
//   k_assemble  ORBextractor.cc:862-872 (add border, octave, size) and :1100-1149 (scale to level-0
//               units, lapping-area two-ended placement, monoIndex).  One CTA per frame; the
//               sequential monoIndex++/stereoIndex-- walk is a prefix sum over "in lapping area" flags.
//   k_blur      GaussianBlur 7x7 sigma 2, BORDER_REFLECT_101 (:1114-1115) of every level, OpenCV's 8.8
//               fixed point {18,34,48,56,48,34,18}.  One warp owns a 128 x 32 tile, one thread a 4-pixel-wide
//               column strip: per row it forms the 4 horizontal sums with 6 PRMT + 8 DP4A from three aligned
//               words and keeps the last 7 rows in registers for the vertical pass, so there is no shared
//               memory, no barrier, and every level is read once (L1 absorbs the 3x word overlap) and the
//               blurred copy written once.  The stored 19-px reflect-101 border IS the border GaussianBlur
//               synthesises, so edges need no special case.  Only ROI pixels are written: rotated pattern
//               samples stay >= 1 px inside the ROI (keypoints sit >= 19 px from its edge, reach <= 18 px).
//   k_describe  IC_Angle (:75-102) + computeOrbDescriptor (:106-145), one WARP per keypoint, no barriers:
//               the 31 disc rows are read straight from the (L1/L2-resident) level, lanes = columns,
//               butterfly reduction of m10/m01; then lane i evaluates tests 8i..8i+7 on the blurred level
//               and owns byte i of the descriptor, exactly the reference's inner loop.
#include <algorithm>

#include "orbfe_internal.h"

#ifndef ASM_THREADS
#define ASM_THREADS 1024       // 128 / 256 / 512 / 1024 threads: 0.034 / 0.029 / 0.025 / 0.021 ms per 512 frames (profiles/r04s_variants.txt, r04t)
#endif

// bit_pattern_31_ (ORBextractor.cc:147-406) as 256 tests x (x0, y0, x1, y1) int8; staged per CTA into
// shared memory in [test & 7][test >> 3] order so that lane i's k-th test is a conflict-free word read.
__device__ const int8_t d_pattern[1024] = {
#include "orb_pattern.inc"
};
// umax of ORBextractor.cc:452-467 (radius-15 disc)
__constant__ int c_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

__global__ void __launch_bounds__(ASM_THREADS) k_assemble(const LevelGeom* __restrict__ geom, const uint32_t* __restrict__ lvlKp,
                                                          const int* __restrict__ lvlCount, int kpPerFrame, int nlevels,
                                                          int lap0, int lap1, orbfe_keypoint* __restrict__ kps,
                                                          uint32_t* __restrict__ slot, int* __restrict__ nOut,
                                                          int* __restrict__ monoOut, int frame0) {
    __shared__ int lvlStart[ORBFE_MAX_LEVELS + 1];
    __shared__ int warpSums[ASM_THREADS / 32 + 1];
    const int frame = frame0 + blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) {
        int s = 0;
        for (int l = 0; l < nlevels; ++l) { lvlStart[l] = s; s += lvlCount[frame * nlevels + l]; }
        lvlStart[nlevels] = s;
    }
    __syncthreads();
    const int n = lvlStart[nlevels];
    const float flap0 = (float)lap0, flap1 = (float)lap1;
    const int per = (n + ASM_THREADS - 1) / ASM_THREADS;
    const int b = min(tid * per, n), e = min(b + per, n);
    const uint32_t* src = lvlKp + (long long)frame * kpPerFrame;
    // pass 1: count lapping keypoints of this thread's chunk
    int cnt = 0, level = 0;
    for (int i = b; i < e; ++i) {
        while (i >= lvlStart[level + 1]) ++level;
        const LevelGeom& g = geom[level];
        const uint32_t c = src[g.kpBase + i - lvlStart[level]];
        float x = (float)(ORBFE_CAND_X(c) + ORBFE_MIN_BORDER);
        if (level != 0) x = __fmul_rn(x, g.scale);
        cnt += (x >= flap0 && x <= flap1);
    }
    int inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
    }
    if (lane == 31) warpSums[wid] = inc;
    __syncthreads();
    if (tid == 0) {
        int s = 0;
        for (int w = 0; w < ASM_THREADS / 32; ++w) { const int v = warpSums[w]; warpSums[w] = s; s += v; }
        warpSums[ASM_THREADS / 32] = s;
    }
    __syncthreads();
    int before = warpSums[wid] + inc - cnt;   // lapping keypoints before this chunk
    const int totalLap = warpSums[ASM_THREADS / 32];
    // pass 2: place
    level = 0;
    orbfe_keypoint* kout = kps + (long long)frame * kpPerFrame;
    uint32_t* sout = slot + (long long)frame * kpPerFrame;
    for (int i = b; i < e; ++i) {
        while (i >= lvlStart[level + 1]) ++level;
        const LevelGeom& g = geom[level];
        const uint32_t c = src[g.kpBase + i - lvlStart[level]];
        const int xi = ORBFE_CAND_X(c) + ORBFE_MIN_BORDER, yi = ORBFE_CAND_Y(c) + ORBFE_MIN_BORDER;
        float x = (float)xi, y = (float)yi;
        if (level != 0) { x = __fmul_rn(x, g.scale); y = __fmul_rn(y, g.scale); }
        const bool lap = (x >= flap0 && x <= flap1);
        const int dst = lap ? n - 1 - before : i - before;
        before += lap;
        orbfe_keypoint k;
        k.x = x; k.y = y; k.size = g.size; k.angle = -1.f; k.response = (float)ORBFE_CAND_R(c);
        k.octave = level; k.class_id = -1;
        kout[dst] = k;
        sout[dst] = (uint32_t)level << 28 | (uint32_t)yi << 14 | (uint32_t)xi;
    }
    if (tid == 0) { nOut[frame] = n; monoOut[frame] = n - totalLap; }
}

// OpenCV fastAtan2 (degrees), unfused float32.
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    const float s = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * s, p3 = -0.3258083974640975f * s, p5 = 0.1555786518463281f * s,
                p7 = -0.04432655554792128f * s;
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, 2.2204460492503131e-16f));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, 2.2204460492503131e-16f));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

#define BLUR_TW 128            // tile width  (one warp, 4 px per lane)
#define BLUR_TH ORBFE_BLUR_TH   // tile height (rows walked by a thread); the host counts tiles with the same constant

// grid.x = ceil(total tiles / 4) (4 warps per CTA, each decodes its own tile), grid.y = frame
#ifndef BLUR_MINB
#define BLUR_MINB 4             // with the unconditional row loop the compiler hoists every load (198 registers, 2 CTAs / SM) unless capped:
#endif                         // describe stage 1.114 ms uncapped, 1.001 at 3, 0.951 at 4 (128 registers) - profiles/r02g_variant_sweep.txt
#ifndef BLUR_PTR
#define BLUR_PTR 1
#endif
#ifndef BLUR_DP2A
#define BLUR_DP2A 1
#endif
__device__ __forceinline__ void blur_tile(const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blur, const LevelGeom* __restrict__ geom,
                                          int nlevels, int tile, int frame, int lane) {
    int level = 0;
    while (level + 1 < nlevels && tile >= geom[level + 1].blurTileBase) ++level;
    const LevelGeom& g = geom[level];
    const int t = tile - g.blurTileBase;
    const int ty = t / g.blurTilesX, tx = t - ty * g.blurTilesX;
    const int x0 = tx * BLUR_TW + lane * 4, y0 = ty * BLUR_TH;
    if (x0 >= g.w) return;
    const int rows = min(BLUR_TH, g.h - y0);
    const long long lvl = g.base + (long long)frame * g.plane;
    const int pw = g.pitch >> 2;
    // word pointer to (x0, y0 - 3) of the bordered level
    const uint32_t* src = reinterpret_cast<const uint32_t*>(pyr + lvl + (long long)(y0 - 3 + ORBFE_EDGE) * g.pitch + ORBFE_XO + x0);
    uint32_t* dst = reinterpret_cast<uint32_t*>(blur + lvl + (long long)(y0 + ORBFE_EDGE) * g.pitch + ORBFE_XO + x0);
    const uint32_t qa = 18u | 34u << 8 | 48u << 16 | 56u << 24, qb = 48u | 34u << 8 | 18u << 16;
    // The row loop is fully unrolled and its body unconditional, so the 7-row window rotates by register renaming.  Rows
    // past a short last tile are simply read (the level's own bottom border and, beyond it, the arena's slack: the results
    // of those rows are never stored) - only the store is predicated, and source / destination advance by one pitch per
    // row (a clamped row index cost a 64-bit multiply-add per row: 12.6 % of the kernel's instructions, profiles/r01o_*).
#if BLUR_DP2A
    // Vertical pass on PAIRS of rows: the horizontal sums are 16-bit (<= 255 * 256), two rows share a register as packed
    // halves and one IDP.2A applies two taps: per pixel column and row one PRMT (new pair), three IDP.2A and one IMAD instead
    // of three adds and four multiply-adds.  When row r + 6 arrives pr[j] = (h(r + j), h(r + j + 1)), j = 0..4: output row r
    // uses pr[0], pr[2], pr[4] and the new row on its own.
    uint32_t pr[5][4], hp[4] = {0u, 0u, 0u, 0u};
    const uint32_t w01 = 18u | 34u << 8, w23 = 48u | 56u << 8, w45 = 48u | 34u << 8;
#pragma unroll
    for (int r = 0; r < BLUR_TH + 6; ++r) {
        const uint32_t* p = src;
        src += pw;
        const uint32_t wm = __ldg(p - 1), w0 = __ldg(p), wp = __ldg(p + 1);
        uint32_t hn[4];
        hn[0] = __dp4a(__byte_perm(wm, w0, 0x4321), qa, __dp4a(__byte_perm(w0, wp, 0x4321), qb, 0u));
        hn[1] = __dp4a(__byte_perm(wm, w0, 0x5432), qa, __dp4a(__byte_perm(w0, wp, 0x5432), qb, 0u));
        hn[2] = __dp4a(__byte_perm(wm, w0, 0x6543), qa, __dp4a(__byte_perm(w0, wp, 0x6543), qb, 0u));
        hn[3] = __dp4a(w0, qa, __dp4a(wp, qb, 0u));
        if (r >= 6) {
            uint32_t v[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)     // < 2^24: the rounded result is byte 2
                v[k] = __dp2a_lo(pr[0][k], w01, __dp2a_lo(pr[2][k], w23, __dp2a_lo(pr[4][k], w45, 18u * hn[k] + 32768u)));
            const uint32_t o = __byte_perm(__byte_perm(v[0], v[1], 0x0062), __byte_perm(v[2], v[3], 0x0062), 0x5410);
            if (r - 6 < rows) *dst = o;
            dst += pw;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
#pragma unroll
            for (int j = 0; j < 4; ++j) pr[j][k] = pr[j + 1][k];
            pr[4][k] = __byte_perm(hp[k], hn[k], 0x5410);
            hp[k] = hn[k];
        }
    }
}
#else
    uint32_t h[7][4];
#pragma unroll
    for (int r = 0; r < BLUR_TH + 6; ++r) {
#if BLUR_PTR
        const uint32_t* p = src;
        src += pw;
#else
        const uint32_t* p = src + (long long)min(r, rows + 5) * pw;
#endif
        const uint32_t wm = __ldg(p - 1), w0 = __ldg(p), wp = __ldg(p + 1);
#pragma unroll
        for (int j = 0; j < 6; ++j)
#pragma unroll
            for (int k = 0; k < 4; ++k) h[j][k] = h[j + 1][k];
        h[6][0] = __dp4a(__byte_perm(wm, w0, 0x4321), qa, __dp4a(__byte_perm(w0, wp, 0x4321), qb, 0u));
        h[6][1] = __dp4a(__byte_perm(wm, w0, 0x5432), qa, __dp4a(__byte_perm(w0, wp, 0x5432), qb, 0u));
        h[6][2] = __dp4a(__byte_perm(wm, w0, 0x6543), qa, __dp4a(__byte_perm(w0, wp, 0x6543), qb, 0u));
        h[6][3] = __dp4a(w0, qa, __dp4a(wp, qb, 0u));
        if (r >= 6) {
            uint32_t v[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)     // < 2^24: the rounded result is byte 2
                v[k] = 18u * (h[0][k] + h[6][k]) + 34u * (h[1][k] + h[5][k]) + 48u * (h[2][k] + h[4][k]) + 56u * h[3][k] + 32768u;
            const uint32_t o = __byte_perm(__byte_perm(v[0], v[1], 0x0062), __byte_perm(v[2], v[3], 0x0062), 0x5410);
#if BLUR_PTR
            if (r - 6 < rows) *dst = o;
            dst += pw;
#else
            if (r - 6 < rows) dst[(long long)(r - 6) * pw] = o;
#endif
        }
    }
}
#endif

__global__ void __launch_bounds__(128, BLUR_MINB) k_blur(const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blur,
                                              const LevelGeom* __restrict__ geom, int nlevels, int totalTiles, int frame0) {
    const int tile = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (tile < totalTiles) blur_tile(pyr, blur, geom, nlevels, tile, frame0 + blockIdx.y, threadIdx.x & 31);
}
// The same tiles walked by a fixed number of resident CTAs (grid = SMs x CTAs per SM): the form that runs beside FAST.
__global__ void __launch_bounds__(128, BLUR_MINB) k_blur_resident(const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blur,
                                                       const LevelGeom* __restrict__ geom, int nlevels, int totalTiles, int frame0,
                                                       int batch) {
    const int lane = threadIdx.x & 31;
    const long long all = (long long)totalTiles * batch;
    for (long long g = (long long)blockIdx.x * 4 + (threadIdx.x >> 5); g < all; g += (long long)gridDim.x * 4) {
        const int f = (int)(g / totalTiles);
        blur_tile(pyr, blur, geom, nlevels, (int)(g - (long long)f * totalTiles), frame0 + f, lane);
    }
}

#define DESC_WARPS 8
#ifndef DESC_STAGE
#define DESC_STAGE 0       // measured (profiles/r02m_variants.txt): 0.962 ms with the staged window vs 0.955 without
#endif

#ifndef DESC_KPW
#define DESC_KPW 1         // key points per warp.  > 1 keeps the lane's part of the pattern in registers across key points: measured
#endif                     // (profiles/r03q_variants.txt) 0.921 ms at 8 (85 or 51 registers), 1.09 at 4-8 with 64 registers, vs 0.914 at 1
#ifndef DESC_MINB
#define DESC_MINB (DESC_KPW > 1 ? 4 : 8)      // 8 CTAs of 8 warps = every warp slot of the SM at 32 registers: describe stage 0.916 -> 0.881 ms (profiles/r04p_variants.txt)
#endif
__global__ void __launch_bounds__(DESC_WARPS * 32, DESC_MINB) k_describe(const uint8_t* __restrict__ pyr, const uint8_t* __restrict__ blur,
                                                              const LevelGeom* __restrict__ geom,
                                                              const uint32_t* __restrict__ slot, const int* __restrict__ nOut,
                                                              int kpPerFrame, orbfe_keypoint* __restrict__ kps,
                                                              uint8_t* __restrict__ desc, int frame0) {
#if DESC_KPW > 1
    // A warp describes DESC_KPW key points one after the other and keeps ITS part of the pattern - the eight tests of
    // descriptor byte `lane` - in registers for all of them: the per-key-point reads of the pattern from shared memory were
    // a quarter of the wavefronts of the data pipe this kernel is bound by (l1tex__data_pipe_lsu_wavefronts 91 %, shared 25 %).
    const int frame = frame0 + blockIdx.y, lane = threadIdx.x & 31;
    float4 pat[8];
    {
        const uint4* pp = reinterpret_cast<const uint4*>(d_pattern) + 2 * lane;      // tests 8 * lane .. 8 * lane + 7, four bytes each
        const uint4 p0 = __ldg(pp), p1 = __ldg(pp + 1);
        const uint32_t w[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
#pragma unroll
        for (int k = 0; k < 8; ++k)
            pat[k] = make_float4((float)(int8_t)(w[k] & 0xFF), (float)(int8_t)(w[k] >> 8 & 0xFF), (float)(int8_t)(w[k] >> 16 & 0xFF),
                                 (float)(int8_t)(w[k] >> 24));
    }
    const int nKp = nOut[frame];
    const int iFirst = (blockIdx.x * DESC_WARPS + (threadIdx.x >> 5)) * DESC_KPW;
    for (int i = iFirst; i < min(iFirst + DESC_KPW, nKp); ++i) {
#else
    __shared__ float4 sPat[256];     // (x0, y0, x1, y1) of test T at [(T & 7) * 32 + (T >> 3)]
    {
        const int T = threadIdx.x;   // DESC_WARPS * 32 == 256 tests
        const uint32_t pw = __ldg(reinterpret_cast<const uint32_t*>(d_pattern) + T);
        sPat[(T & 7) * 32 + (T >> 3)] = make_float4((float)(int8_t)(pw & 0xFF), (float)(int8_t)(pw >> 8 & 0xFF),
                                                    (float)(int8_t)(pw >> 16 & 0xFF), (float)(int8_t)(pw >> 24));
    }
    __syncthreads();
    const int frame = frame0 + blockIdx.y, i = blockIdx.x * DESC_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= nOut[frame]) return;
    {
#endif
    const uint32_t s = slot[(long long)frame * kpPerFrame + i];
    const int level = s >> 28, cy = (s >> 14) & 0x3FFF, cx = s & 0x3FFF;
    const LevelGeom& g = geom[level];
    const long long centre = g.base + (long long)frame * g.plane + (long long)(cy + ORBFE_EDGE) * g.pitch + ORBFE_XO + cx;
#if DESC_STAGE
    // The 512 samples of a key point fall anywhere in a 37 x 37 window of the blurred level (pattern radius 18.4): read
    // straight from global memory, every one of the 16 gathers of a warp touches ~20 different lines and the kernel sat on
    // the L1 tag stage (l1tex 91 %, profiles/r02k_*).  The window is staged once with coalesced aligned words (37 rows x 10
    // words; the pitch is a multiple of 128, so all rows share the byte shift sh) and the gathers hit shared memory.
    __shared__ uint32_t sWin[DESC_WARPS][37 * 10];
    const int sh = (int)((centre - 18) & 3);
    {
        const uint32_t* wbase = reinterpret_cast<const uint32_t*>(blur + (centre - 18 - sh)) - 18 * (g.pitch >> 2);
        const int pw = g.pitch >> 2;
        uint32_t* dstw = sWin[threadIdx.x >> 5];
#pragma unroll 4
        for (int u = lane; u < 370; u += 32) {
            const int r = (u * 6554) >> 16, j = u - 10 * r;      // u / 10 for u < 16384
            dstw[u] = __ldg(wbase + r * pw + j);
        }
    }
#endif
    // ---- intensity centroid (IC_Angle): lane = column u = lane - 15, rows v = -15..15 clipped to the disc
    int m10 = 0, m01 = 0;
    {
        // the disc is symmetric: |u| <= umax[|v|]  <=>  |v| <= umax[|u|], one limit per lane
        const int u = lane - 15;
        const int lim = lane < 31 ? c_umax[u < 0 ? -u : u] : -1;
        // addresses = one per-lane 64-bit pointer to the TOP row of the window + an unsigned 32-bit offset (one IMAD.WIDE.U32
        // per load); signed row offsets from the centre cost a multiply, a sign extension and a three-input 64-bit add each
        const unsigned upitch = (unsigned)g.pitch;
        const uint8_t* cTop = pyr + (centre + u - 15ll * g.pitch);
        asm volatile("" : "+l"(cTop));      // keep it ONE register pair (else the uniform arena base is re-added at every load)
        // rows v and -v together (m01 += v * (I(v) - I(-v)), as IC_Angle itself does); the loads are unconditional - every
        // row of the 31 x 31 window lies inside the 19-px border and touches the same lines whichever lanes are inside the
        // disc - so only the three accumulations are predicated and the unrolled loop has no branches
        int sum = lane < 31 ? (int)__ldg(cTop + 15ull * upitch) : 0;
#pragma unroll
        for (int v = 1; v <= 15; ++v) {
            const int Ip = __ldg(cTop + (unsigned long long)upitch * (unsigned)(15 + v)), Im = __ldg(cTop + (unsigned long long)upitch * (unsigned)(15 - v));
            if (v <= lim) {
                sum += Ip + Im;
                m01 += v * (Ip - Im);
            }
        }
        m10 = u * sum;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m10 += __shfl_xor_sync(0xffffffffu, m10, o);
        m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    }
    const float angle = fast_atan2_deg((float)m01, (float)m10);
    if (lane == 0) kps[(long long)frame * kpPerFrame + i].angle = angle;
    // a = cosf(angle * factorPI), b = sinf(...): float argument, result rounded from double precision
    const float rad = __fmul_rn(angle, (float)(3.14159265358979323846 / 180.f));
    double sd, cd;                                  // one shared argument reduction for both
    sincos((double)rad, &sd, &cd);
    const float a = (float)cd, b = (float)sd;
    // ---- lane computes descriptor byte `lane`: tests 8*lane .. 8*lane+7
#if DESC_STAGE
    __syncwarp();
    const uint8_t* Bc = reinterpret_cast<const uint8_t*>(sWin[threadIdx.x >> 5]) + sh;
    const int pitch = 40;
    const int kOff = 18 * 40 + 18;
#define DESC_LD(p) (*(p))
#else
    // one per-lane 64-bit pointer to the centre + a signed 32-bit offset
    const uint8_t* Bc = blur + centre;
    asm volatile("" : "+l"(Bc));
    const int pitch = g.pitch;
    const int kOff = 0;
#define DESC_LD(p) __ldg(p)
#endif
    uint32_t byte = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
#if DESC_KPW > 1
        const float4 pt = pat[k];                // test 8*lane + k
#else
        const float4 pt = sPat[k * 32 + lane];   // test 8*lane + k
#endif
        const float x0 = pt.x, y0 = pt.y, x1 = pt.z, y1 = pt.w;
        const int t0 = DESC_LD(Bc + (long long)(__float2int_rn(__fadd_rn(__fmul_rn(x0, b), __fmul_rn(y0, a))) * pitch +
                                                __float2int_rn(__fsub_rn(__fmul_rn(x0, a), __fmul_rn(y0, b))) + kOff));
        const int t1 = DESC_LD(Bc + (long long)(__float2int_rn(__fadd_rn(__fmul_rn(x1, b), __fmul_rn(y1, a))) * pitch +
                                                __float2int_rn(__fsub_rn(__fmul_rn(x1, a), __fmul_rn(y1, b))) + kOff));
        byte |= (uint32_t)(t0 < t1) << k;
    }
    // gather 4 bytes per word: lanes 4j..4j+3 -> word j, written by lane 4j
    uint32_t w = byte << (8 * (lane & 3));
    w |= __shfl_xor_sync(0xffffffffu, w, 1);
    w |= __shfl_xor_sync(0xffffffffu, w, 2);
    if ((lane & 3) == 0) reinterpret_cast<uint32_t*>(desc + ((long long)frame * kpPerFrame + i) * 32)[lane >> 2] = w;
    }
}

cudaError_t orbfe_launch_assemble(orbfe_handle* h, int frame0, int batch, int lap0, int lap1) {
    k_assemble<<<batch, ASM_THREADS, 0, h->cur>>>(h->dGeom, h->dLvlKp, h->dLvlCount, h->kpPerFrame, h->p.nlevels, lap0, lap1,
                                                     h->dKps, h->dSlot, h->dN, h->dMono, frame0);
    h->launches++;
    return cudaGetLastError();
}

cudaError_t orbfe_launch_blur(orbfe_handle* h, int frame0, int batch, cudaStream_t st, int persistPerSm) {
    if (h->blurBackend == ORBFE_BLUR_TENSOR) return orbfe_launch_blur_tc(h, frame0, batch, st);
    if (persistPerSm > 0) {
        const long long groups = ((long long)h->blurTiles * batch + 3) / 4;
        const int grid = (int)std::min<long long>(groups, (long long)h->smCount * persistPerSm);
        k_blur_resident<<<grid, 128, 0, st>>>(h->dPyr, h->dBlur, h->dGeom, h->p.nlevels, h->blurTiles, frame0, batch);
    } else {
        dim3 grid((h->blurTiles + 3) / 4, batch);
        k_blur<<<grid, 128, 0, st>>>(h->dPyr, h->dBlur, h->dGeom, h->p.nlevels, h->blurTiles, frame0);
    }
    h->launches++;
    return cudaGetLastError();
}

cudaError_t orbfe_launch_describe(orbfe_handle* h, int frame0, int batch) {
    dim3 grid((h->kpPerFrame + DESC_WARPS * DESC_KPW - 1) / (DESC_WARPS * DESC_KPW), batch);
    k_describe<<<grid, DESC_WARPS * 32, 0, h->cur>>>(h->dPyr, h->dBlur, h->dGeom, h->dSlot, h->dN, h->kpPerFrame, h->dKps,
                                                       h->dDesc, frame0);
    h->launches++;
    return cudaGetLastError();
}
