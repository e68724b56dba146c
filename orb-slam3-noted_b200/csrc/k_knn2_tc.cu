// k_knn2_tc.cu — brute-force Hamming 2-NN on the 5th-generation tensor cores (tcgen05.mma kind::i8, TMEM accumulators).
//
// Same contract as k_knn2 (k_match.cu): cv::BFMatcher(NORM_HAMMING).knnMatch(k = 2) as Frame::ComputeStereoFishEyeMatches
// uses it (Frame.cc:43,1134), distances of ORBmatcher::DescriptorDistance (ORBmatcher.cc:2561-2577), ties to the lowest
// train index.  Integer arithmetic end to end, so indices and distances are bit-identical to the POPC kernel.
//
// All-pairs Hamming IS a dense contraction.  Expand the descriptor bits to signed bytes, a_k = +1 / -1 for a set / clear
// query bit and b_k = -128 / 0 for a set / clear train bit: a_k * b_k is -128 where both bits are set, +128 where only the
// train bit is, 0 elsewhere, so
//     sum_k a_k b_k = 128 * (n01 - n11) = 128 * (hamming - popcount(query)).
// One more 32-byte K slice carries (2, 1, u0, u1, u2, u3, 0, ...) on the query side and (j >> 1, j & 1, 64, 64, 64, 64, 0, ...)
// on the train side, with j = the train row's position inside its 128-row half tile and u0 + u1 + u2 + u3 =
// 2 * (popcount(query) - 128): it adds j + 128 * (popcount(query) - 128).  The accumulator therefore holds
//     D = 128 * (hamming - 128) + j,     -16384 <= D <= 16511,
// a 16-bit quantity whose signed minimum over a half tile IS "smallest distance, ties to the lowest index".  The epilogue
// reads the accumulators two columns per register (tcgen05.ld ... pack::16b) and keeps the two smallest per 16-bit lane
// with packed DPX min / max: 1.25 integer instructions per descriptor pair instead of the 27 of the XOR + POPC form;
// distance and index are unpacked from the survivors of a tile only.
//
// One CTA = 128 queries (the 128 TMEM lanes) x one split of the train set, streamed in tiles of 256 rows:
//   warps 8-11 producers: expand the tile's descriptors (32 B each) to 288-byte K-major rows (shift + mask per 4 bytes: a bit
//              moved to bit 7 of its byte IS the byte -128 / 0) in the canonical no-swizzle UMMA layout (8-row x 16-byte
//              core matrices), two stages;
//   warp 12    one elected thread issues 9 x tcgen05.mma (M 128, N 256, K 32) per tile into one of two 256-column TMEM
//              accumulators and commits to mbarriers (frees the smem stage, publishes the accumulator);
//   warps 0-7  epilogue: thread = query row; warps 0-3 take columns 0-127, warps 4-7 columns 128-255 of the same lanes;
//              tcgen05.ld 32 columns at a time, two in flight, four independent two-smallest chains per thread.
// Roles meet only through mbarriers; generic-proxy writes are fenced (fence.proxy.async) before the tensor core reads them.
#include <limits.h>

#include "orbfe_internal.h"
#include "orbfe_match.h"

#define TC_M 128                 // queries per CTA = TMEM lanes
#define TC_N 256                 // train rows per tile = accumulator columns
#define TC_CHUNKS 18             // 16-byte K chunks per row: 16 of descriptor bits + 2 carrying the row index
#define TC_THREADS 416
#define TC_PRODUCER_WARP0 8
#define TC_MMA_WARP 12
#define TC_A_BYTES (TC_M * 16 * TC_CHUNKS)        // 36864
#define TC_B_BYTES (TC_N * 16 * TC_CHUNKS)        // 73728
#define TC_SMEM (TC_A_BYTES + 2 * TC_B_BYTES + TC_M * 8 + 256)
#define KNN_IDX_BITS 20
#define KNN_IDX_MASK ((1u << KNN_IDX_BITS) - 1u)
#define KNN_NONE 0xFFFFFFFFu

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
__device__ __forceinline__ void fence_async_proxy() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// Shared-memory matrix descriptor, K-major, no swizzle: 8-row x 16-byte core matrices of 128 contiguous bytes; LBO = distance
// between the two 16-byte K chunks of one MMA, SBO = distance between 8-row groups (both in 16-byte units); version 1 (sm_100).
__device__ __forceinline__ uint64_t umma_desc(uint32_t smemAddr, uint32_t lbo16, uint32_t sbo16) {
    return (uint64_t)((smemAddr >> 4) & 0x3FFFu) | (uint64_t)(lbo16 & 0x3FFFu) << 16 | (uint64_t)(sbo16 & 0x3FFFu) << 32 | 1ull << 46;
}
// Instruction descriptor of kind::i8: D = S32, A = B = signed 8-bit, both K-major, N = 256, M = 128.
__device__ constexpr uint32_t kIdesc = 2u << 4 | 1u << 7 | 1u << 10 | (uint32_t)(TC_N >> 3) << 17 | (uint32_t)(TC_M >> 4) << 24;

__device__ __forceinline__ void umma_i8(uint32_t tmemD, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n"
        "}\n" ::"r"(tmemD),
        "l"(adesc), "l"(bdesc), "r"(kIdesc), "r"(accumulate), "r"(0u)
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
// The loaded registers are valid only after tcgen05.wait::ld; naming them as in-out operands keeps every use behind the wait.
__device__ __forceinline__ void tmem_ld_wait(uint32_t (&v)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]), "+r"(v[9]), "+r"(v[10]),
                   "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15])
                 :
                 : "memory");
}

// Per 16-bit lane: the two smallest of {m0 <= m1} and two new values (5 DPX instructions per four accumulators).
__device__ __forceinline__ void two_min2(uint32_t& m0, uint32_t& m1, uint32_t a, uint32_t b) {
    // (the two-input packed min / max intrinsics are emulated on sm_100a; the three-input DPX forms are one instruction)
    const uint32_t lo = __vimin3_s16x2(a, b, b), hi = __vimax3_s16x2(a, b, b);
    m1 = __vimin3_s16x2(__vimax3_s16x2(m0, lo, lo), m1, hi);
    m0 = __vimin3_s16x2(m0, lo, lo);
}
__device__ __forceinline__ void two_min_scalar(int& m0, int& m1, int a) {
    const int hi = max(m0, a);
    m0 = min(m0, a);
    m1 = min(m1, hi);
}
__device__ __forceinline__ void key_update(uint32_t& k0, uint32_t& k1, uint32_t k) {
    const uint32_t hi = max(k0, k);
    k0 = min(k0, k);
    k1 = min(k1, hi);
}

// One descriptor row (32 bytes) -> 18 K chunks of signed bytes at row `r` of a tile of R rows.  Output word (i, p) holds bit p
// of the four bytes of descriptor word i: the word shifted left by 7 - p has those bits at bit 7 of each byte.  kTrain: the
// byte is -128 / 0 (mask only); query: +1 / -1 (a few more operations, once per CTA).  Any fixed order of the K elements
// serves, as long as both operands use the same one.  extra = the two bytes of the index-carrying chunk.
template <bool kTrain>
__device__ __forceinline__ void expand_row(uint8_t* tile, int R, int r, const uint4& w0, const uint4& w1, uint32_t extra0, uint32_t extra1) {
    uint8_t* row = tile + (r >> 3) * 128 + (r & 7) * 16;
    const uint32_t w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
    const int cstride = R * 16;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        uint32_t o[8];
#pragma unroll
        for (int p = 0; p < 8; ++p) {
            const uint32_t sh = w[i] * (1u << (7 - p));      // a multiply: the FMA pipe is idle here, the integer ALU is what this kernel runs out of
            const uint32_t m = sh & 0x80808080u;                                   // 0x80 = -128 where the bit is set
            o[p] = kTrain ? m : ((((m >> 7) ^ 0x01010101u) * 0xFFu) | (m >> 7));    // query: 0x01 = +1 where set, 0xFF = -1 where clear
        }
        *reinterpret_cast<uint4*>(row + (2 * i) * cstride) = make_uint4(o[0], o[1], o[2], o[3]);
        *reinterpret_cast<uint4*>(row + (2 * i + 1) * cstride) = make_uint4(o[4], o[5], o[6], o[7]);
    }
    *reinterpret_cast<uint4*>(row + 16 * cstride) = make_uint4(extra0, extra1, 0u, 0u);
    *reinterpret_cast<uint4*>(row + 17 * cstride) = make_uint4(0u, 0u, 0u, 0u);
}

}  // namespace

// grid = (query tiles of 128, train splits, batch)
__global__ void __launch_bounds__(TC_THREADS, 1) k_knn2_tc(const uint8_t* __restrict__ q, long long qStride, const int* __restrict__ nqArr,
                                                          int nqMax, const uint8_t* __restrict__ t, long long tStride,
                                                          const int* __restrict__ ntArr, int ntMax, int splitLen,
                                                          uint2* __restrict__ partial, int32_t* __restrict__ idx2, int32_t* __restrict__ dist2) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t aFull, bFull[2], bEmpty[2], accFull[2], accEmpty[2];
    __shared__ uint32_t sTmem;
    const int b = blockIdx.z, split = blockIdx.y, nsplit = gridDim.y;
    const int nq = nqArr ? min(nqArr[b], nqMax) : nqMax;
    const int nt = ntArr ? min(ntArr[b], ntMax) : ntMax;
    const int q0 = blockIdx.x * TC_M;
    if (q0 >= nq) return;
    const int t0 = split * splitLen, t1 = min(t0 + splitLen, nt);
    const int ntiles = t1 > t0 ? (t1 - t0 + TC_N - 1) / TC_N : 0;
    uint8_t* sA = smem;
    uint8_t* sB = smem + TC_A_BYTES;
    uint2* sKeys = reinterpret_cast<uint2*>(smem + TC_A_BYTES + 2 * TC_B_BYTES);     // the upper column half's result per query
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (tid == 0) {
        bar_init(&aFull, 128);
        for (int s = 0; s < 2; ++s) { bar_init(&bFull[s], 128); bar_init(&bEmpty[s], 1); bar_init(&accFull[s], 1); bar_init(&accEmpty[s], 256); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == TC_MMA_WARP) {  // the whole TMEM of the SM: two accumulators of 256 columns (one CTA per SM by shared memory)
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&sTmem)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = sTmem;

    if (warp >= TC_PRODUCER_WARP0 && warp < TC_MMA_WARP) {
        // ---- producers
        const int p = tid - TC_PRODUCER_WARP0 * 32;
        {
            uint4 w0 = make_uint4(0u, 0u, 0u, 0u), w1 = w0;
            if (q0 + p < nq) {
                const uint4* src = reinterpret_cast<const uint4*>(q + (long long)b * qStride) + 2 * (long long)(q0 + p);
                w0 = __ldg(src); w1 = __ldg(src + 1);
            }
            // bytes 2, 1 (they weigh the train side's index bytes) and four bytes that sum to 2 * (popcount - 128)
            const int pop = __popc(w0.x) + __popc(w0.y) + __popc(w0.z) + __popc(w0.w) + __popc(w1.x) + __popc(w1.y) + __popc(w1.z) + __popc(w1.w);
            const int two = 2 * pop - 256, q4 = two >> 2, rem = two & 3;       // floor division: two == 4 * q4 + rem
            const uint32_t u = (uint32_t)((q4 + (rem > 0)) & 0xFF) | (uint32_t)((q4 + (rem > 1)) & 0xFF) << 8 |
                               (uint32_t)((q4 + (rem > 2)) & 0xFF) << 16 | (uint32_t)(q4 & 0xFF) << 24;
            expand_row<false>(sA, TC_M, p, w0, w1, 0x00000102u | u << 16, u >> 16);
            fence_async_proxy();
            bar_arrive(&aFull);
        }
        const uint4* tp = reinterpret_cast<const uint4*>(t + (long long)b * tStride);
        for (int i = 0; i < ntiles; ++i) {
            const int s = i & 1;
            uint4 w[4];
#pragma unroll
            for (int h = 0; h < 2; ++h) {      // the loads do not depend on the stage: issue them before waiting for it
                const int g = t0 + i * TC_N + p + 128 * h;
                w[2 * h] = w[2 * h + 1] = make_uint4(0u, 0u, 0u, 0u);
                if (g < t1) { w[2 * h] = __ldg(tp + 2 * (long long)g); w[2 * h + 1] = __ldg(tp + 2 * (long long)g + 1); }
            }
            bar_wait(&bEmpty[s], ((i >> 1) & 1) ^ 1);
            uint8_t* tile = sB + s * TC_B_BYTES;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int r = p + 128 * h;
                const uint32_t j = (uint32_t)(r & 127);         // position inside its 128-row half tile
                expand_row<true>(tile, TC_N, r, w[2 * h], w[2 * h + 1], (j >> 1) | (j & 1u) << 8 | 0x40400000u, 0x00004040u);
            }
            fence_async_proxy();
            bar_arrive(&bFull[s]);
        }
    } else if (warp == TC_MMA_WARP) {
        // ---- MMA issuer: one thread
        if (lane == 0 && ntiles > 0) {
            bar_wait(&aFull, 0);
            const uint32_t aBase = s32(sA), bBase = s32(sB);
            for (int i = 0; i < ntiles; ++i) {
                const int s = i & 1;
                bar_wait(&bFull[s], (i >> 1) & 1);
                bar_wait(&accEmpty[s], ((i >> 1) & 1) ^ 1);
                tc_fence_after();
#pragma unroll
                for (int k = 0; k < TC_CHUNKS / 2; ++k) {
                    const uint64_t ad = umma_desc(aBase + 2 * k * TC_M * 16, TC_M, 8);
                    const uint64_t bd = umma_desc(bBase + s * TC_B_BYTES + 2 * k * TC_N * 16, TC_N, 8);
                    umma_i8(tmem + s * TC_N, ad, bd, k > 0);
                }
                umma_commit(&bEmpty[s]);       // the stage may be overwritten once these MMAs have read it
                umma_commit(&accFull[s]);      // ... and the accumulator is complete
            }
        }
    } else {
        // ---- epilogue: thread = query row = TMEM lane; `half` = which 128 columns of every tile
        const int half = warp >> 2, row = (warp & 3) * 32 + lane;
        const int qi = q0 + row;
        uint32_t k0 = KNN_NONE, k1 = KNN_NONE;
        const uint32_t laneBase = tmem + ((uint32_t)((warp & 3) * 32) << 16) + half * 128;
        const uint32_t kBig = 0x7FFF7FFFu;     // above every real accumulator in both lanes
        for (int i = 0; i < ntiles; ++i) {
            const int s = i & 1;
            bar_wait(&accFull[s], (i >> 1) & 1);
            tc_fence_after();
            const int base = t0 + i * TC_N + half * 128, valid = min(128, t1 - base);      // this half's rows: [base, base + valid)
            uint32_t m0[4] = {kBig, kBig, kBig, kBig}, m1[4] = {kBig, kBig, kBig, kBig};
            uint32_t v[2][16];
            tmem_ld32p(laneBase + s * TC_N, v[0]);
#pragma unroll
            for (int cb = 0; cb < 4; ++cb) {
                uint32_t (&x)[16] = v[cb & 1];
                tmem_ld_wait(x);
                if (cb < 3) tmem_ld32p(laneBase + s * TC_N + (cb + 1) * 32, v[(cb + 1) & 1]);      // in flight while this block is reduced
                if (cb * 32 + 32 > valid) {    // the last tile of a split: rows past the end do not exist
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const int c = cb * 32 + 2 * j;
                        if (c >= valid) x[j] = kBig;
                        else if (c + 1 >= valid) x[j] = (x[j] & 0xFFFFu) | 0x7FFF0000u;
                    }
                }
#pragma unroll
                for (int j = 0; j < 16; j += 8) {
                    two_min2(m0[0], m1[0], x[j], x[j + 1]);
                    two_min2(m0[1], m1[1], x[j + 2], x[j + 3]);
                    two_min2(m0[2], m1[2], x[j + 4], x[j + 5]);
                    two_min2(m0[3], m1[3], x[j + 6], x[j + 7]);
                }
            }
            tc_fence_before();
            bar_arrive(&accEmpty[s]);
            two_min2(m0[0], m1[0], m0[1], m1[1]);
            two_min2(m0[2], m1[2], m0[3], m1[3]);
            two_min2(m0[0], m1[0], m0[2], m1[2]);
            // the two lanes hold disjoint columns: the tile's two smallest are among their four survivors
            int a0 = INT_MAX, a1 = INT_MAX;
            two_min_scalar(a0, a1, (int)(short)(m0[0] & 0xFFFFu));
            two_min_scalar(a0, a1, (int)(short)(m0[0] >> 16));
            two_min_scalar(a0, a1, (int)(short)(m1[0] & 0xFFFFu));
            two_min_scalar(a0, a1, (int)(short)(m1[0] >> 16));
            // D = 128 * (hamming - 128) + j  ->  packed key  hamming << 20 | train index
            if (a0 != 0x7FFF) key_update(k0, k1, (uint32_t)((a0 >> 7) + 128) << KNN_IDX_BITS | (uint32_t)(base + (a0 & 127)));
            if (a1 != 0x7FFF) key_update(k0, k1, (uint32_t)((a1 >> 7) + 128) << KNN_IDX_BITS | (uint32_t)(base + (a1 & 127)));
        }
        if (half == 1) sKeys[row] = make_uint2(k0, k1);
        tc_fence_before();
        __syncthreads();
        if (half == 0 && qi < nq) {
            const uint2 o = sKeys[row];
            key_update(k0, k1, o.x);
            key_update(k0, k1, o.y);           // KNN_NONE entries never displace a real key
            if (nsplit == 1) {
                const long long at = ((long long)b * nqMax + qi) * 2;
                idx2[at] = k0 == KNN_NONE ? -1 : (int)(k0 & KNN_IDX_MASK);
                dist2[at] = k0 == KNN_NONE ? -1 : (int)(k0 >> KNN_IDX_BITS);
                idx2[at + 1] = k1 == KNN_NONE ? -1 : (int)(k1 & KNN_IDX_MASK);
                dist2[at + 1] = k1 == KNN_NONE ? -1 : (int)(k1 >> KNN_IDX_BITS);
            } else {
                partial[((long long)b * nsplit + split) * nqMax + qi] = make_uint2(k0, k1);
            }
        }
        return;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == TC_MMA_WARP) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

// Train splits for the tensor-core kernel: tiles of 256 rows, about two CTAs' worth of work per SM.
void orbfe_knn2_tc_plan(int nqMax, int ntMax, int batch, int smCount, int* nsplit, int* splitLen) {
    const long long ctas = (long long)((nqMax + TC_M - 1) / TC_M) * batch;
    const int tiles = (ntMax + TC_N - 1) / TC_N;
    long long s = ctas > 0 ? (2LL * smCount + ctas - 1) / ctas : 1;
    s = s < 1 ? 1 : s > tiles ? tiles : s;
    if (s > 64) s = 64;
    const int per = (int)((tiles + s - 1) / (s > 0 ? s : 1));
    *splitLen = (per < 1 ? 1 : per) * TC_N;
    const int n = (ntMax + *splitLen - 1) / *splitLen;
    *nsplit = n < 1 ? 1 : n;
}

cudaError_t orbfe_launch_knn2_tc(cudaStream_t st, const uint8_t* dq, long long qStride, const int* dNq, int nqMax, const uint8_t* dt,
                                 long long tStride, const int* dNt, int ntMax, int batch, int nsplit, int splitLen, uint2* dPartial,
                                 int32_t* dIdx2, int32_t* dDist2, long long* launches) {
    if (nqMax <= 0 || batch <= 0) return cudaSuccess;
    if (ntMax > (int)KNN_IDX_MASK || splitLen % TC_N != 0) return cudaErrorInvalidValue;
    const cudaError_t ea = orbfe_ensure_dyn_smem(reinterpret_cast<const void*>(k_knn2_tc), TC_SMEM);
    if (ea != cudaSuccess) return ea;
    dim3 grid((nqMax + TC_M - 1) / TC_M, nsplit, batch);
    k_knn2_tc<<<grid, TC_THREADS, TC_SMEM, st>>>(dq, qStride, dNq, nqMax, dt, tStride, dNt, ntMax, splitLen, dPartial, dIdx2, dDist2);
    ++*launches;
    if (nsplit > 1) return orbfe_launch_knn2_merge(st, dPartial, dNq, nqMax, nsplit, batch, dIdx2, dDist2, launches);
    return cudaGetLastError();
}
