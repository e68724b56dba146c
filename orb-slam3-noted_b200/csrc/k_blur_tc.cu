// k_blur_tc.cu — the 7x7 Gaussian of ORBextractor::operator() (GaussianBlur(workingMat, workingMat, Size(7, 7), 2, 2,
// BORDER_REFLECT_101), ORBextractor.cc:1114-1115) on the 5th-generation tensor cores.
//
// OpenCV's 8-bit path is exact integer arithmetic: h = sum_i k_i p(x + i - 3) with the 8.8 fixed-point taps
// k = (18, 34, 48, 56, 48, 34, 18) (no rounding, h <= 255 * 256), then out = (sum_j k_j h(y + j - 3) + 2^15) >> 16.  Both
// passes are products with the same banded Toeplitz matrix T[m][k] = k_{k - m - 13} (0 <= k - m - 13 <= 6), i.e. GEMMs:
//     G1:  D1[x][r] = sum_c T[x][c] * P[r][c]        M = 128 output columns, N = 160 window rows, K = 160 window columns
//     G2:  D2[y][x] = sum_r T[y][r] * H[x][r]        M = 128 output rows, N = 2 x 128 (byte planes of h), K = 160 window rows
// on tcgen05.mma kind::i8 with UNSIGNED 8-bit operands and 32-bit accumulators in TMEM.  The 16-bit h does not fit an 8-bit
// operand: it is split into its two bytes, H = 256 Hhi + Hlo, both planes go through G2 as one N = 256 operand, and
//     out = (256 Dhi + Dlo + 2^15) >> 16 = (Dhi + (Dlo >> 8) + 128) >> 8        (exact: Dlo = 256 q + r, r < 256)
// with Dhi + (Dlo >> 8) + 128 <= 65408, so the whole epilogue runs on packed 16-bit halves.  G1 is taken TRANSPOSED (x on
// the TMEM lanes) so that a thread of the first epilogue holds 160 consecutive window rows of one column: exactly the K-major
// rows of G2's second operand, written with 16-byte stores.
//
// One tile = 128 x 128 output pixels of one level of one frame; its window is 160 x 160 (16 to the left / above, so that it
// starts on a 16-byte boundary; 3 would do for the filter).  Persistent CTAs (one per SM: 416 TMEM columns), four roles:
//   warps 8-11  producers: window -> shared memory in the K-major no-swizzle UMMA layout ([16-byte column chunk][row][16]);
//   warp 12     one thread issues G1(i) and G2(i - 1) alternately (5 + 5 MMAs of K = 32) and commits to mbarriers;
//   warps 0-3   first epilogue: D1 (tcgen05.ld pack::16b) -> byte planes of h in shared memory (G2's operand), two stages;
//   warps 4-7   second epilogue: D2 -> blurred pixels, 32 bytes per thread and step.
// ~2.4 CUDA-core instructions per pixel against 11 for k_blur (k_describe.cu), which stays as the selectable ALU form.
#include "orbfe_internal.h"

#ifndef BT_TIMING
#define BT_TIMING 0
#endif
#ifndef BT_NOLOAD
#define BT_NOLOAD 0            // timing experiments only: the producers skip their copies (wrong pixels)
#endif
#ifndef BT_NOSTORE
#define BT_NOSTORE 0           // timing experiments only: the second epilogue skips its stores
#endif
#define BT_TILE 128
#define BT_WIN 160
#define BT_CHUNKS 10
#define BT_T_BYTES (BT_TILE * BT_WIN)            // 20480: T, [chunk][128 rows][16]
#define BT_P_BYTES (BT_WIN * BT_WIN)             // 25600 per stage: the window, [chunk][160 rows][16]
#define BT_H_BYTES (BT_CHUNKS * 256 * 16)        // 40960 per stage: byte planes of h, [chunk][256 rows][16], rows 0-127 hi, 128-255 lo
#ifndef BT_STAGES
#define BT_STAGES 4                                // window stages (cp.async groups in flight: BT_STAGES - 1)
#endif
#define BT_SMEM (BT_T_BYTES + BT_STAGES * BT_P_BYTES + 2 * BT_H_BYTES + 1024)
#define BT_THREADS 416
#define BT_D1_COL 0
#define BT_D2_COL 256

struct BlurTcPlan {
    int nlevels, tilesPerFrame;
    int tileBase[ORBFE_MAX_LEVELS + 1], tilesX[ORBFE_MAX_LEVELS];
    int w[ORBFE_MAX_LEVELS], h[ORBFE_MAX_LEVELS], pitch[ORBFE_MAX_LEVELS], rows[ORBFE_MAX_LEVELS];
    long long base[ORBFE_MAX_LEVELS], plane[ORBFE_MAX_LEVELS];
};

namespace {

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(bar)), "r"(count));
}
__device__ __forceinline__ void bar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(bar)) : "memory");
}
__device__ __forceinline__ void bar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "W_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra D_%=;\n"
        "bra W_%=;\n"
        "D_%=:\n"
        "}\n" ::"r"(s32(bar)),
        "r"(parity)
        : "memory");
}
#if BT_TIMING
// debug build: every wait adds its duration to a per-role, per-barrier counter (g_btTime, read back by tools/probe)
__device__ long long g_btTime[16];
#define BAR_WAIT_T(bar, parity, slot)                                           \
    do {                                                                        \
        const long long t0_ = clock64();                                        \
        bar_wait(bar, parity);                                                  \
        if ((threadIdx.x & 31) == 0 && blockIdx.x == 0) atomicAdd((unsigned long long*)&g_btTime[slot], (unsigned long long)(clock64() - t0_)); \
    } while (0)
#else
#define BAR_WAIT_T(bar, parity, slot) bar_wait(bar, parity)
#endif
__device__ __forceinline__ void fence_async_proxy() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// K-major, no swizzle (see k_knn2_tc.cu): LBO = distance between the two 16-byte K chunks of one MMA, SBO = between 8-row groups.
__device__ __forceinline__ uint64_t umma_desc(uint32_t smemAddr, uint32_t lbo16, uint32_t sbo16) {
    return (uint64_t)((smemAddr >> 4) & 0x3FFFu) | (uint64_t)(lbo16 & 0x3FFFu) << 16 | (uint64_t)(sbo16 & 0x3FFFu) << 32 | 1ull << 46;
}
// kind::i8, D = S32, A = B = UNSIGNED 8-bit (format 0), both K-major, M = 128.
__device__ constexpr uint32_t idesc_u8(int n) { return 2u << 4 | (uint32_t)(n >> 3) << 17 | (uint32_t)(BT_TILE >> 4) << 24; }

__device__ __forceinline__ void umma_u8(uint32_t tmemD, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n"
        "}\n" ::"r"(tmemD),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(0u)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(bar)) : "memory");
}
// 32 accumulator columns of this thread's lane as 16 registers: the low 16 bits of two adjacent columns per register.
__device__ __forceinline__ void tmem_ld32p(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.pack::16b.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
          "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait(uint32_t (&v)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]), "+r"(v[9]), "+r"(v[10]),
                   "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15])
                 :
                 : "memory");
}

struct TileAt {
    int level, x0, y0, frame;
};
__device__ __forceinline__ TileAt decode_tile(const BlurTcPlan& pl, long long g, int frame0) {
    TileAt t;
    const int f = (int)(g / pl.tilesPerFrame);
    int r = (int)(g - (long long)f * pl.tilesPerFrame);
    int level = 0;
    while (level + 1 < pl.nlevels && r >= pl.tileBase[level + 1]) ++level;
    r -= pl.tileBase[level];
    const int ty = r / pl.tilesX[level];
    t.level = level; t.frame = frame0 + f; t.y0 = ty * BT_TILE; t.x0 = (r - ty * pl.tilesX[level]) * BT_TILE;
    return t;
}

}  // namespace

__global__ void __launch_bounds__(BT_THREADS, 1) k_blur_tc(const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blur,
                                                          const BlurTcPlan pl, int frame0, int batch) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t pFull[BT_STAGES], pEmpty[BT_STAGES], hFull[2], hEmpty[2], d1Full, d1Empty, d2Full, d2Empty;
    __shared__ uint32_t sTmem;
    uint8_t* sT = smem;
    uint8_t* sP = smem + BT_T_BYTES;
    uint8_t* sH = sP + BT_STAGES * BT_P_BYTES;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long total = (long long)pl.tilesPerFrame * batch;

    if (tid == 0) {
        for (int s = 0; s < BT_STAGES; ++s) { bar_init(&pFull[s], 128); bar_init(&pEmpty[s], 1); }
        for (int s = 0; s < 2; ++s) { bar_init(&hFull[s], 128); bar_init(&hEmpty[s], 1); }
        bar_init(&d1Full, 1); bar_init(&d1Empty, 128); bar_init(&d2Full, 1); bar_init(&d2Empty, 128);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 12) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&sTmem)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // the Toeplitz operand, once per CTA: byte (m, k) at chunk (k >> 4), row m
    for (int u = tid; u < BT_CHUNKS * BT_TILE; u += BT_THREADS) {
        const int chunk = u >> 7, m = u & 127;
        uint32_t wds[4] = {0u, 0u, 0u, 0u};
#pragma unroll
        for (int b = 0; b < 16; ++b) {
            const int j = chunk * 16 + b - m - 13;     // tap index
            const uint32_t wgt = j == 0 || j == 6 ? 18u : j == 1 || j == 5 ? 34u : j == 2 || j == 4 ? 48u : j == 3 ? 56u : 0u;
            wds[b >> 2] |= wgt << (8 * (b & 3));
        }
        *reinterpret_cast<uint4*>(sT + chunk * (BT_TILE * 16) + (m >> 3) * 128 + (m & 7) * 16) = make_uint4(wds[0], wds[1], wds[2], wds[3]);
    }
    fence_async_proxy();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = sTmem;
#if BT_TIMING
    const long long tStart = clock64();
#endif

    if (warp >= 8 && warp < 12) {
        // ---- producers: 1600 16-byte pieces per window; lanes 0-15 rows r..r+15 of chunk c, lanes 16-31 the same rows of
        // chunk c + 1 (32 contiguous bytes per row), 8 consecutive rows = one 128-byte core matrix in shared memory
        // Copies are cp.async (global -> shared without registers), one commit group per tile and BT_STAGES - 1 groups in
        // flight: a tile's window costs one HBM round trip (~1.5 us under load, longer than everything else a tile needs),
        // so the loads of the next tiles must be on their way while this one is consumed.
        const int p = tid - 256;                       // 0..127
        const int rsub = p & 15, csub = (p >> 4) & 1, grp = p >> 5;      // 4 warps
        int n = 0;
        for (long long g = blockIdx.x; g < total; g += gridDim.x) ++n;
        for (int it = 0; it < n + BT_STAGES - 1; ++it) {
            if (it < n) {      // issue the copies of tile `it`
                const int s = it % BT_STAGES;
                const TileAt t = decode_tile(pl, (long long)blockIdx.x + (long long)it * gridDim.x, frame0);
                const int pitch = pl.pitch[t.level], rows = pl.rows[t.level];
                const uint8_t* lvl = pyr + pl.base[t.level] + (long long)t.frame * pl.plane[t.level];
                BAR_WAIT_T(&pEmpty[s], ((it / BT_STAGES) & 1) ^ 1, 0);
                const uint32_t dstP = s32(sP + s * BT_P_BYTES);
#pragma unroll
                for (int i = 0; i < 13; ++i) {
                    const int u = grp + 4 * i;
                    if (u < 50 && !BT_NOLOAD) {
                        const int rg = u / 5, cp = u - rg * 5;
                        const int r = rg * 16 + rsub, c = cp * 2 + csub;
                        const int br = t.y0 + 3 + r, bx = t.x0 + 16 + 16 * c;       // bordered row / byte column
                        const bool in = br < rows && bx + 16 <= pitch;              // outside the level: zero fill (src-size 0)
                        const uint8_t* src = lvl + (in ? (long long)br * pitch + bx : 0);
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dstP + c * (BT_WIN * 16) + (r >> 3) * 128 + (r & 7) * 16),
                                     "l"(src), "r"(in ? 16 : 0)
                                     : "memory");
                    }
                }
            }
            asm volatile("cp.async.commit_group;" ::: "memory");       // (an empty group past the last tile keeps the count uniform)
            const int done = it - (BT_STAGES - 1);                      // this tile's group is the oldest but BT_STAGES - 1
            if (done >= 0) {
                asm volatile("cp.async.wait_group %0;" ::"n"(BT_STAGES - 1) : "memory");
                fence_async_proxy();
                bar_arrive(&pFull[done % BT_STAGES]);
            }
        }
    } else if (warp == 12) {
        // ---- MMA issuer
        if (lane == 0) {
            const uint32_t tBase = s32(sT), pBase = s32(sP), hBase = s32(sH);
            int n = 0;
            for (long long g = blockIdx.x; g < total; g += gridDim.x) ++n;
            for (int it = 0; it <= n; ++it) {
                if (it < n) {      // G1(it): D1 = T x window^T
                    const int s = it % BT_STAGES;
                    BAR_WAIT_T(&pFull[s], (it / BT_STAGES) & 1, 1);
                    BAR_WAIT_T(&d1Empty, (it & 1) ^ 1, 2);
                    tc_fence_after();
#pragma unroll
                    for (int k = 0; k < BT_CHUNKS / 2; ++k) {
                        const uint64_t ad = umma_desc(tBase + 2 * k * (BT_TILE * 16), BT_TILE, 8);
                        const uint64_t bd = umma_desc(pBase + s * BT_P_BYTES + 2 * k * (BT_WIN * 16), BT_WIN, 8);
                        umma_u8(tmem + BT_D1_COL, ad, bd, idesc_u8(BT_WIN), k > 0);
                    }
                    umma_commit(&pEmpty[s]);
                    umma_commit(&d1Full);
                }
                if (it > 0) {      // G2(it - 1): D2 = T x (byte planes of h)^T
                    const int j = it - 1, s = j & 1;
                    BAR_WAIT_T(&hFull[s], (j >> 1) & 1, 3);
                    BAR_WAIT_T(&d2Empty, (j & 1) ^ 1, 4);
                    tc_fence_after();
#pragma unroll
                    for (int k = 0; k < BT_CHUNKS / 2; ++k) {
                        const uint64_t ad = umma_desc(tBase + 2 * k * (BT_TILE * 16), BT_TILE, 8);
                        const uint64_t bd = umma_desc(hBase + s * BT_H_BYTES + 2 * k * (256 * 16), 256, 8);
                        umma_u8(tmem + BT_D2_COL, ad, bd, idesc_u8(256), k > 0);
                    }
                    umma_commit(&hEmpty[s]);
                    umma_commit(&d2Full);
                }
            }
        }
    } else if (warp < 4) {
        // ---- first epilogue: thread = output column m = TMEM lane; D1[m][r] -> hi / lo byte planes, K-major rows of G2's operand
        const int m = tid;
        const uint32_t laneBase = tmem + ((uint32_t)(warp * 32) << 16) + BT_D1_COL;
        const int rowOff = (m >> 3) * 128 + (m & 7) * 16;
        int it = 0;
        for (long long g = blockIdx.x; g < total; g += gridDim.x, ++it) {
            const int s = it & 1;
            BAR_WAIT_T(&d1Full, it & 1, 5);
            tc_fence_after();
            BAR_WAIT_T(&hEmpty[s], ((it >> 1) & 1) ^ 1, 6);
            uint8_t* dstH = sH + s * BT_H_BYTES;
            uint32_t v[2][16];
            tmem_ld32p(laneBase, v[0]);
#pragma unroll
            for (int cb = 0; cb < 5; ++cb) {      // 32 window rows = two K chunks per step
                uint32_t (&x)[16] = v[cb & 1];
                tmem_ld_wait(x);
                if (cb < 4) tmem_ld32p(laneBase + (cb + 1) * 32, v[(cb + 1) & 1]);
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    const uint32_t* y = &x[8 * half];
                    const uint4 lo = make_uint4(__byte_perm(y[0], y[1], 0x6420), __byte_perm(y[2], y[3], 0x6420),
                                                __byte_perm(y[4], y[5], 0x6420), __byte_perm(y[6], y[7], 0x6420));
                    const uint4 hi = make_uint4(__byte_perm(y[0], y[1], 0x7531), __byte_perm(y[2], y[3], 0x7531),
                                                __byte_perm(y[4], y[5], 0x7531), __byte_perm(y[6], y[7], 0x7531));
                    uint8_t* at = dstH + (2 * cb + half) * (256 * 16) + rowOff;
                    *reinterpret_cast<uint4*>(at) = hi;
                    *reinterpret_cast<uint4*>(at + 128 * 16) = lo;
                }
            }
            tc_fence_before();
            bar_arrive(&d1Empty);
            fence_async_proxy();
            bar_arrive(&hFull[s]);
        }
    } else {
        // ---- second epilogue: thread = output row m = TMEM lane; columns 0-127 = sums over the hi plane, 128-255 over the lo plane
        const int m = tid - 128;
        const uint32_t laneBase = tmem + ((uint32_t)((warp & 3) * 32) << 16) + BT_D2_COL;
        int it = 0;
        for (long long g = blockIdx.x; g < total; g += gridDim.x, ++it) {
            const TileAt t = decode_tile(pl, g, frame0);
            const int w = pl.w[t.level], pitch = pl.pitch[t.level];
            const bool rowOk = t.y0 + m < pl.h[t.level];
            uint8_t* dstRow = blur + pl.base[t.level] + (long long)t.frame * pl.plane[t.level] +
                              (long long)(ORBFE_EDGE + t.y0 + m) * pitch + ORBFE_XO + t.x0;
            BAR_WAIT_T(&d2Full, it & 1, 7);
            tc_fence_after();
#pragma unroll
            for (int gx = 0; gx < 4; ++gx) {      // 32 pixels per step
                uint32_t vh[16], vl[16];
                tmem_ld32p(laneBase + gx * 32, vh);
                tmem_ld32p(laneBase + 128 + gx * 32, vl);
                tmem_ld_wait(vh);
                tmem_ld_wait(vl);
                uint32_t sum[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) sum[j] = vh[j] + __byte_perm(vl[j], 0u, 0x4341) + 0x00800080u;   // both halves < 2^16
                uint32_t o[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) o[q] = __byte_perm(sum[2 * q], sum[2 * q + 1], 0x7531);
                if (rowOk && !BT_NOSTORE) {
                    if (t.x0 + gx * 32 < w) *reinterpret_cast<uint4*>(dstRow + gx * 32) = make_uint4(o[0], o[1], o[2], o[3]);
                    if (t.x0 + gx * 32 + 16 < w) *reinterpret_cast<uint4*>(dstRow + gx * 32 + 16) = make_uint4(o[4], o[5], o[6], o[7]);
                }
            }
            tc_fence_before();
            bar_arrive(&d2Empty);
        }
    }
#if BT_TIMING
    if (tid == 0 && blockIdx.x == 0) {
        atomicAdd((unsigned long long*)&g_btTime[8], (unsigned long long)(clock64() - tStart));
        long long nt = 0;
        for (long long g = blockIdx.x; g < total; g += gridDim.x) ++nt;
        atomicAdd((unsigned long long*)&g_btTime[9], (unsigned long long)nt);
    }
#endif
    tc_fence_before();
    __syncthreads();
    if (warp == 12) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

cudaError_t orbfe_launch_blur_tc(orbfe_handle* h, int frame0, int batch, cudaStream_t st) {
    BlurTcPlan pl;
    memset(&pl, 0, sizeof(pl));
    pl.nlevels = h->p.nlevels;
    int tiles = 0;
    for (int l = 0; l < pl.nlevels; ++l) {
        const LevelGeom& g = h->geom[l];
        pl.tileBase[l] = tiles;
        pl.tilesX[l] = (g.w + BT_TILE - 1) / BT_TILE;
        tiles += pl.tilesX[l] * ((g.h + BT_TILE - 1) / BT_TILE);
        pl.w[l] = g.w; pl.h[l] = g.h; pl.pitch[l] = g.pitch; pl.rows[l] = g.rows; pl.base[l] = g.base; pl.plane[l] = g.plane;
    }
    pl.tileBase[pl.nlevels] = tiles;
    pl.tilesPerFrame = tiles;
    const cudaError_t ea = orbfe_ensure_dyn_smem(reinterpret_cast<const void*>(k_blur_tc), BT_SMEM);
    if (ea != cudaSuccess) return ea;
    const long long total = (long long)tiles * batch;
    const int grid = (int)(total < h->smCount ? total : h->smCount);
    k_blur_tc<<<grid, BT_THREADS, BT_SMEM, st>>>(h->dPyr, h->dBlur, pl, frame0, batch);
    h->launches++;
    return cudaGetLastError();
}

#if BT_TIMING
extern "C" int orbfe_debug_blur_tc_times(long long* out16, int reset) {
    if (out16 && cudaMemcpyFromSymbol(out16, g_btTime, sizeof(long long) * 16) != cudaSuccess) return -1;
    if (reset) { long long z[16] = {0}; if (cudaMemcpyToSymbol(g_btTime, z, sizeof(z)) != cudaSuccess) return -1; }
    return 0;
}
#endif
