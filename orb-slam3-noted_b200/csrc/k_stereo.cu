// k_stereo.cu — Frame::ComputeStereoMatches (Frame.cc:794-964), the rectified-stereo "SearchStereo".
//
//   k_stereo_search  grid (left-keypoint chunks of 128, pairs).  The CTA first builds the right
//       image's row-band table in shared memory: per right keypoint its x and the packed
//       [minr, maxr] row interval (floor/ceil of y -+ 2*scale[octave], :812-821) and octave.  The
//       reference materialises vRowIndices[row] lists; testing  minr <= int(vL) <= maxr  per right
//       keypoint is the same candidate set and needs no list.  Phase 1: one thread per left
//       keypoint scans the table (octave within +-1, uR in [uL - bf/b, uL], :862-870), evaluates
//       the 256-bit Hamming distance for survivors and keeps min(dist << 16 | iR) — strict '<' in
//       ascending iR order in the reference = lowest iR among equal distances (:874-878).
//       Phase 2: one warp per accepted keypoint (dist < 75) does the 11x11 SAD over the 11 shifts
//       on the pyramid level of the left keypoint (:883-915), the parabola fit and the disparity
//       tests (:917-946), all float ops unfused like the reference's scalar code.
//   k_stereo_filter  one CTA per pair: the median of the accepted SADs (rank n/2 of the sorted
//       list, :950-951) by a two-pass histogram select, then drops matches with SAD >= 2.1*median.
// The pyramids, keypoints and descriptors are read where the extractor left them in HBM.
#include "orbfe_internal.h"

#define ST_THREADS 128

struct StereoSide {
    const orbfe_keypoint* kps;   // [frame][kpPerFrame]
    const uint8_t* desc;         // [frame][kpPerFrame][32]
    const int* n;                // [frame]
    const uint8_t* pyr;
    const LevelGeom* geom;
    int kpPerFrame;
    int frame0, frameStep;       // pair p uses frame0 + p * frameStep
};

struct StereoTables {
    float scale[ORBFE_MAX_LEVELS], invScale[ORBFE_MAX_LEVELS];
    int nlevels;
};

__device__ __forceinline__ int hamming256(const uint4 a0, const uint4 a1, const uint4* __restrict__ b) {
    const uint4 b0 = __ldg(b), b1 = __ldg(b + 1);
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x) +
           __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__global__ void __launch_bounds__(ST_THREADS) k_stereo_search(StereoSide L, StereoSide R, StereoTables tb, float mbf, float mb,
                                                              float* __restrict__ uRightOut, float* __restrict__ depthOut,
                                                              int* __restrict__ sadOut, int outStride) {
    extern __shared__ __align__(8) uint8_t smem[];
    __shared__ int sBestR[ST_THREADS];
    const int pair = blockIdx.y, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int fL = L.frame0 + pair * L.frameStep, fR = R.frame0 + pair * R.frameStep;
    const int nL = min(L.n[fL], L.kpPerFrame), nR = min(R.n[fR], R.kpPerFrame);
    const int i0 = blockIdx.x * ST_THREADS;
    if (i0 >= nL) return;
    float* tabX = reinterpret_cast<float*>(smem);
    uint32_t* tabB = reinterpret_cast<uint32_t*>(tabX + R.kpPerFrame);
    const orbfe_keypoint* kR = R.kps + (long long)fR * R.kpPerFrame;
    const int nRows = L.geom[0].h;
    for (int i = tid; i < nR; i += ST_THREADS) {
        const orbfe_keypoint k = kR[i];
        const float r = __fmul_rn(2.0f, tb.scale[k.octave]);
        int maxr = (int)ceilf(__fadd_rn(k.y, r)), minr = (int)floorf(__fsub_rn(k.y, r));
        minr = max(minr, 0);
        maxr = min(maxr, nRows - 1);
        tabX[i] = k.x;
        // an empty interval (minr > maxr after clamping) can never contain a row
        tabB[i] = minr > maxr ? 0xFFFu : ((uint32_t)minr | (uint32_t)maxr << 12 | (uint32_t)k.octave << 24);
    }
    __syncthreads();
    const float minD = 0.f, maxD = __fdiv_rn(mbf, mb);
    const int iL = i0 + tid;
    const orbfe_keypoint* kLp = L.kps + (long long)fL * L.kpPerFrame;
    orbfe_keypoint kp;
    int bestR = -1;
    if (iL < nL) {
        kp = kLp[iL];
        const float uL = kp.x;
        const float minU = __fsub_rn(uL, maxD), maxU = __fsub_rn(uL, minD);
        const int yL = (int)kp.y;
        if (!(maxU < 0) && yL >= 0 && yL < nRows) {
            const uint4* dl = reinterpret_cast<const uint4*>(L.desc + ((long long)fL * L.kpPerFrame + iL) * 32);
            const uint4 a0 = __ldg(dl), a1 = __ldg(dl + 1);
            const uint4* dR = reinterpret_cast<const uint4*>(R.desc + (long long)fR * R.kpPerFrame * 32);
            uint32_t best = (uint32_t)ORBFE_TH_HIGH << 16;
            for (int iR = 0; iR < nR; ++iR) {
                const uint32_t pb = tabB[iR];
                const int mn = pb & 0xFFF, mx = (pb >> 12) & 0xFFF, oc = pb >> 24;
                if (yL < mn || yL > mx || oc < kp.octave - 1 || oc > kp.octave + 1) continue;
                const float uR = tabX[iR];
                if (uR >= minU && uR <= maxU) {
                    const uint32_t key = (uint32_t)hamming256(a0, a1, dR + 2 * (long long)iR) << 16 | (uint32_t)iR;
                    best = min(best, key);
                }
            }
            if ((int)(best >> 16) < (ORBFE_TH_HIGH + ORBFE_TH_LOW) / 2) bestR = (int)(best & 0xFFFF);
        }
    }
    sBestR[tid] = bestR;
    if (iL < nL) {
        const long long o = (long long)pair * outStride + iL;
        uRightOut[o] = -1.f;
        depthOut[o] = -1.f;
        sadOut[o] = -1;
    }
    __syncthreads();
    // ---- phase 2: SAD refinement, one warp per accepted keypoint
    for (int k = wid; k < ST_THREADS; k += ST_THREADS / 32) {
        const int bR = sBestR[k];
        if (bR < 0) continue;
        const int i = i0 + k;
        const orbfe_keypoint kl = kLp[i];
        const float uR0 = tabX[bR];
        const float sf = tb.invScale[kl.octave];
        const float scaleduL = roundf(__fmul_rn(kl.x, sf)), scaledvL = roundf(__fmul_rn(kl.y, sf));
        const float scaleduR0 = roundf(__fmul_rn(uR0, sf));
        const LevelGeom gl = L.geom[kl.octave], gr = R.geom[kl.octave];
        const float iniu = scaleduR0, endu = scaleduR0 + 11.f;   // scaleduR0 + L - w, scaleduR0 + L + w + 1 with L = w = 5
        if (iniu < 0 || endu >= (float)gr.w) continue;
        const int y0 = (int)(scaledvL - 5.f), xl0 = (int)(scaleduL - 5.f), xr0 = (int)(scaleduR0 - 10.f);
        // windows stay inside the stored 19-px border for every keypoint the extractor can emit; clamp regardless
        if (y0 < -ORBFE_EDGE || y0 + 11 > gl.h + ORBFE_EDGE || xl0 < -ORBFE_EDGE || xl0 + 11 > gl.w + ORBFE_EDGE ||
            xr0 < -ORBFE_EDGE || xr0 + 21 > gr.w + ORBFE_EDGE)
            continue;
        const uint8_t* pl = L.pyr + gl.base + (long long)fL * gl.plane + (long long)(y0 + ORBFE_EDGE) * gl.pitch + ORBFE_XO + xl0;
        const uint8_t* pr = R.pyr + gr.base + (long long)fR * gr.plane + (long long)(y0 + ORBFE_EDGE) * gr.pitch + ORBFE_XO + xr0;
        int s[11];
#pragma unroll
        for (int d = 0; d < 11; ++d) s[d] = 0;
        for (int p = lane; p < 121; p += 32) {
            const int r = p / 11, c = p - r * 11;
            const int a = __ldg(pl + (long long)r * gl.pitch + c);
            const uint8_t* rr = pr + (long long)r * gr.pitch + c;
#pragma unroll
            for (int d = 0; d < 11; ++d) s[d] += abs(a - (int)__ldg(rr + d));
        }
#pragma unroll
        for (int d = 0; d < 11; ++d)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s[d] += __shfl_xor_sync(0xffffffffu, s[d], o);
        if (lane == 0) {
            int bestS = 0x7FFFFFFF, bestInc = 0;
#pragma unroll
            for (int d = 0; d < 11; ++d)
                if (s[d] < bestS) { bestS = s[d]; bestInc = d - 5; }
            if (bestInc != -5 && bestInc != 5) {
                float d1 = 0.f, d2 = 0.f, d3 = 0.f;
#pragma unroll
                for (int d = 1; d < 10; ++d)
                    if (d - 5 == bestInc) { d1 = (float)s[d - 1]; d2 = (float)s[d]; d3 = (float)s[d + 1]; }
                const float deltaR = __fdiv_rn(__fsub_rn(d1, d3), __fmul_rn(2.0f, __fsub_rn(__fadd_rn(d1, d3), __fmul_rn(2.0f, d2))));
                if (!(deltaR < -1.f || deltaR > 1.f)) {
                    float bestuR = __fmul_rn(tb.scale[kl.octave], __fadd_rn(__fadd_rn(scaleduR0, (float)bestInc), deltaR));
                    float disparity = __fsub_rn(kl.x, bestuR);
                    if (disparity >= minD && disparity < maxD) {
                        if (disparity <= 0.f) { disparity = 0.01f; bestuR = (float)((double)kl.x - 0.01); }
                        const long long o = (long long)pair * outStride + i;
                        depthOut[o] = __fdiv_rn(mbf, disparity);
                        uRightOut[o] = bestuR;
                        sadOut[o] = bestS;
                    }
                }
            }
        }
    }
}

// One CTA per pair: median SAD (rank n/2) by histogram select, then the 2.1 x median cut (Frame.cc:950-963).
__global__ void __launch_bounds__(256) k_stereo_filter(StereoSide L, float* __restrict__ uRight, float* __restrict__ depth,
                                                       const int* __restrict__ sad, int outStride, int* __restrict__ nMatched) {
    __shared__ int hist[256];
    __shared__ int sSel, sBefore, sTotal;
    const int pair = blockIdx.x, tid = threadIdx.x;
    const int fL = L.frame0 + pair * L.frameStep;
    const int nL = min(L.n[fL], L.kpPerFrame);
    const int* s = sad + (long long)pair * outStride;
    hist[tid] = 0;
    __syncthreads();
    for (int i = tid; i < nL; i += 256)
        if (s[i] >= 0) atomicAdd(&hist[min(s[i] >> 7, 255)], 1);   // SAD <= 121*255 = 30855 -> bins 0..241
    __syncthreads();
    if (tid == 0) {
        int tot = 0;
        for (int b = 0; b < 256; ++b) tot += hist[b];
        sTotal = tot;
        int rank = tot / 2, acc = 0, b = 0;
        for (; b < 256; ++b) { if (acc + hist[b] > rank) break; acc += hist[b]; }
        sSel = b;
        sBefore = acc;
    }
    __syncthreads();
    const int total = sTotal;
    if (total == 0) { if (tid == 0) nMatched[pair] = 0; return; }   // the reference reads an empty vector here (:951)
    const int hiBin = sSel, before = sBefore;
    __syncthreads();
    hist[tid] = 0;
    __syncthreads();
    for (int i = tid; i < nL; i += 256)
        if (s[i] >= 0 && min(s[i] >> 7, 255) == hiBin) atomicAdd(&hist[s[i] & 127], 1);
    __syncthreads();
    if (tid == 0) {
        int rank = total / 2 - before, acc = 0, b = 0;
        for (; b < 128; ++b) { if (acc + hist[b] > rank) break; acc += hist[b]; }
        sSel = (hiBin << 7) | b;
        sBefore = 0;
    }
    __syncthreads();
    const float median = (float)sSel;
    const float thDist = __fmul_rn(__fmul_rn(1.5f, 1.4f), median);
    int kept = 0;
    for (int i = tid; i < nL; i += 256) {
        if (s[i] < 0) continue;
        if ((float)s[i] < thDist) ++kept;
        else {
            uRight[(long long)pair * outStride + i] = -1.f;
            depth[(long long)pair * outStride + i] = -1.f;
        }
    }
    atomicAdd(&sBefore, kept);
    __syncthreads();
    if (tid == 0) nMatched[pair] = sBefore;
}

cudaError_t orbfe_launch_stereo(cudaStream_t st, const orbfe_handle* hl, int frameL0, int frameLStep, const orbfe_handle* hr,
                                int frameR0, int frameRStep, int npairs, float bf, float b, float* dURight, float* dDepth,
                                int* dSad, int* dCount, int outStride, long long* launches) {
    StereoSide L{hl->dKps, hl->dDesc, hl->dN, hl->dPyr, hl->dGeom, hl->kpPerFrame, frameL0, frameLStep};
    StereoSide R{hr->dKps, hr->dDesc, hr->dN, hr->dPyr, hr->dGeom, hr->kpPerFrame, frameR0, frameRStep};
    StereoTables tb;
    tb.nlevels = hl->p.nlevels;
    for (int l = 0; l < ORBFE_MAX_LEVELS; ++l) {
        tb.scale[l] = l < hl->p.nlevels ? hl->scale[l] : 1.f;
        tb.invScale[l] = l < hl->p.nlevels ? hl->invScale[l] : 1.f;
    }
    const size_t smem = (size_t)hr->kpPerFrame * 8;
    {   // the attribute is per device: one high-water mark per (device, kernel), see orbfe_ensure_dyn_smem
        const cudaError_t ea = orbfe_ensure_dyn_smem(reinterpret_cast<const void*>(k_stereo_search), smem);
        if (ea != cudaSuccess) return ea;
    }
    dim3 grid((hl->kpPerFrame + ST_THREADS - 1) / ST_THREADS, npairs);
    k_stereo_search<<<grid, ST_THREADS, smem, st>>>(L, R, tb, bf, b, dURight, dDepth, dSad, outStride);
    ++*launches;
    k_stereo_filter<<<npairs, 256, 0, st>>>(L, dURight, dDepth, dSad, outStride, dCount);
    ++*launches;
    return cudaGetLastError();
}
