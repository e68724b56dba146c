// k_describe.cu — output assembly, orientation and descriptors.
//
//   k_assemble  ORBextractor.cc:862-872 (add border, octave, size) and :1100-1149 (scale to level-0
//               units, lapping-area two-ended placement, monoIndex).  One CTA per frame; the
//               sequential monoIndex++/stereoIndex-- walk is a prefix sum over "in lapping area" flags.
//   k_describe  IC_Angle (:75-102) + GaussianBlur 7x7 sigma 2 (:1114-1115) + computeOrbDescriptor
//               (:106-145), fused per keypoint: one CTA stages the 43x43 unblurred patch around the
//               keypoint in shared memory, takes the intensity-centroid moments from it, runs the
//               separable 8.8 fixed-point Gaussian on the 37x37 window the rotated pattern can reach
//               (|coord| <= 18), and evaluates the 256 tests; __ballot_sync packs 32 tests per word.
//               Blurring only what the descriptors read replaces two full passes over every level.
//               All taps fall inside the stored 19-px reflect-101 border (keypoints are >= 19 px from
//               the ROI edge, taps reach 21 px), which equals what GaussianBlur synthesises.
#include "orbfe_internal.h"

#define ASM_THREADS 256
#define DESC_THREADS 128

__constant__ int8_t c_pattern[1024] = {
#include "orb_pattern.inc"
};
// umax of ORBextractor.cc:452-467 (radius-15 disc)
__constant__ int c_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

__global__ void __launch_bounds__(ASM_THREADS) k_assemble(const LevelGeom* __restrict__ geom, const uint32_t* __restrict__ lvlKp,
                                                          const int* __restrict__ lvlCount, int kpPerFrame, int nlevels,
                                                          int lap0, int lap1, orbfe_keypoint* __restrict__ kps,
                                                          uint32_t* __restrict__ slot, int* __restrict__ nOut,
                                                          int* __restrict__ monoOut) {
    __shared__ int lvlStart[ORBFE_MAX_LEVELS + 1];
    __shared__ int warpSums[ASM_THREADS / 32 + 1];
    const int frame = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
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

#define PR 21                 // patch radius: 18 (pattern reach) + 3 (blur taps)
#define PW 43                 // patch width
#define PP 48                 // patch pitch in shared memory (12 words)
#define BW 37                 // blurred window width
#define HP 38                 // pitch of the horizontal-pass buffer (uint16)
#define BP 40                 // pitch of the blurred window

__global__ void __launch_bounds__(DESC_THREADS) k_describe(const uint8_t* __restrict__ pyr, const LevelGeom* __restrict__ geom,
                                                           const uint32_t* __restrict__ slot, const int* __restrict__ nOut,
                                                           int kpPerFrame, orbfe_keypoint* __restrict__ kps,
                                                           uint8_t* __restrict__ desc) {
    __shared__ __align__(16) uint8_t P[PW * PP + 16];
    __shared__ uint16_t H[PW * HP];
    __shared__ uint8_t B[BW * BP];
    __shared__ int red[2 * (DESC_THREADS / 32)];
    __shared__ float rot[2];
    const int frame = blockIdx.y, i = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (i >= nOut[frame]) return;
    const uint32_t s = slot[(long long)frame * kpPerFrame + i];
    const int level = s >> 28, cy = (s >> 14) & 0x3FFF, cx = s & 0x3FFF;
    const LevelGeom& g = geom[level];
    // ---- stage the 43x43 patch with aligned 32-bit loads
    const long long col0 = ORBFE_XO + cx - PR;               // byte column of patch column 0 in the bordered row
    const int sh = (int)(col0 & 3);
    const uint8_t* base = pyr + g.base + (long long)frame * g.plane + (long long)(cy + ORBFE_EDGE - PR) * g.pitch + (col0 - sh);
    for (int w = tid; w < PW * (PP / 4); w += DESC_THREADS) {
        const int r = w / (PP / 4), c = w - r * (PP / 4);
        reinterpret_cast<uint32_t*>(P)[r * (PP / 4) + c] = __ldg(reinterpret_cast<const uint32_t*>(base + (long long)r * g.pitch) + c);
    }
    __syncthreads();
    const uint8_t* Pc = P + sh;   // Pc[r * PP + c] = pixel (cx - 21 + c, cy - 21 + r)
    // ---- intensity centroid (IC_Angle): 31 rows x 4 segments
    int m10 = 0, m01 = 0;
    if (tid < 124) {
        const int v = (tid >> 2) - 15, d = c_umax[v < 0 ? -v : v];
        const uint8_t* row = Pc + (v + PR) * PP + PR;
        for (int u = -d + (tid & 3); u <= d; u += 4) {
            const int I = row[u];
            m10 += u * I;
            m01 += v * I;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m10 += __shfl_xor_sync(0xffffffffu, m10, o);
        m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    }
    if (lane == 0) { red[wid] = m10; red[DESC_THREADS / 32 + wid] = m01; }
    // ---- horizontal pass of the 7x7 Gaussian, 8.8 fixed point {18,34,48,56,48,34,18}
    for (int o = tid; o < PW * BW; o += DESC_THREADS) {
        const int r = o / BW, c = o - r * BW;
        const uint8_t* p = Pc + r * PP + c;
        H[r * HP + c] = (uint16_t)(18 * (p[0] + p[6]) + 34 * (p[1] + p[5]) + 48 * (p[2] + p[4]) + 56 * p[3]);
    }
    __syncthreads();
    if (tid == 0) {
        int a10 = 0, a01 = 0;
#pragma unroll
        for (int w = 0; w < DESC_THREADS / 32; ++w) { a10 += red[w]; a01 += red[DESC_THREADS / 32 + w]; }
        const float angle = fast_atan2_deg((float)a01, (float)a10);
        kps[(long long)frame * kpPerFrame + i].angle = angle;
        // a = cosf(angle * factorPI), b = sinf(...): float argument, result rounded from double precision
        const float rad = __fmul_rn(angle, (float)(3.14159265358979323846 / 180.f));
        rot[0] = (float)cos((double)rad);
        rot[1] = (float)sin((double)rad);
    }
    // ---- vertical pass
    for (int o = tid; o < BW * BW; o += DESC_THREADS) {
        const int r = o / BW, c = o - r * BW;
        const uint16_t* hcol = H + r * HP + c;
        const uint32_t v = 18u * (hcol[0] + hcol[6 * HP]) + 34u * (hcol[HP] + hcol[5 * HP]) + 48u * (hcol[2 * HP] + hcol[4 * HP]) +
                           56u * hcol[3 * HP];
        B[r * BP + c] = (uint8_t)((v + 32768u) >> 16);
    }
    __syncthreads();
    // ---- 256 rotated-pattern tests
    const float a = rot[0], b = rot[1];
    const uint8_t* Bc = B + 18 * BP + 18;
    uint32_t* dout = reinterpret_cast<uint32_t*>(desc + ((long long)frame * kpPerFrame + i) * 32);
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        const int T = half * DESC_THREADS + tid;
        const float x0 = (float)c_pattern[4 * T], y0 = (float)c_pattern[4 * T + 1];
        const float x1 = (float)c_pattern[4 * T + 2], y1 = (float)c_pattern[4 * T + 3];
        const int t0 = Bc[__float2int_rn(__fadd_rn(__fmul_rn(x0, b), __fmul_rn(y0, a))) * BP +
                          __float2int_rn(__fsub_rn(__fmul_rn(x0, a), __fmul_rn(y0, b)))];
        const int t1 = Bc[__float2int_rn(__fadd_rn(__fmul_rn(x1, b), __fmul_rn(y1, a))) * BP +
                          __float2int_rn(__fsub_rn(__fmul_rn(x1, a), __fmul_rn(y1, b)))];
        const uint32_t bits = __ballot_sync(0xffffffffu, t0 < t1);
        if (lane == 0) dout[half * (DESC_THREADS / 32) + wid] = bits;
    }
}

cudaError_t orbfe_launch_assemble(orbfe_handle* h, int batch, int lap0, int lap1) {
    k_assemble<<<batch, ASM_THREADS, 0, h->stream>>>(h->dGeom, h->dLvlKp, h->dLvlCount, h->kpPerFrame, h->p.nlevels, lap0, lap1,
                                                     h->dKps, h->dSlot, h->dN, h->dMono);
    h->launches++;
    return cudaGetLastError();
}

cudaError_t orbfe_launch_describe(orbfe_handle* h, int batch) {
    dim3 grid(h->kpPerFrame, batch);
    k_describe<<<grid, DESC_THREADS, 0, h->stream>>>(h->dPyr, h->dGeom, h->dSlot, h->dN, h->kpPerFrame, h->dKps, h->dDesc);
    h->launches++;
    return cudaGetLastError();
}
