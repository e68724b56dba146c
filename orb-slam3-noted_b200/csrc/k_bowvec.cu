// k_bowvec.cu — DBoW2 TemplatedVocabulary::transform(features, BowVector&, FeatureVector&, levelsup)
// (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1137-1206, :1230-1271), the step Frame::ComputeBoW
// (src/Frame.cc:721-728) and KeyFrame::ComputeBoW (src/KeyFrame.cc:105-114) run between extraction and the
// BoW-restricted searches.
//
// k_bow_descend: one thread per feature walks the tree: at every level the child with the smallest Hamming
// distance (strict '<': the first child wins ties, :1256), remembering the node at level L - levelsup.  The
// tree (35 MB for k = 10, L = 6) stays resident in the 126 MB L2; a descent is L x k 32-byte reads.
// k_bow_assemble: one CTA per frame turns the per-feature (word, node) pairs into the two std::map's the
// reference builds, flattened in key order: a bitonic sort of (key << 32 | feature index) in shared memory,
// head flags + a block scan for the segments.  Word values are accumulated in feature order and the L1 / L2
// norm in ascending word order, sequentially in double like the map iteration of BowVector::normalize
// (BowVector.cpp:58-79), so every value is bit-identical to the reference's.
#include "orbfe_internal.h"
#include "orbfe_voc.h"

__device__ __forceinline__ int ham256(const uint4& a0, const uint4& a1, const uint4* __restrict__ b) {
    const uint4 b0 = __ldg(b), b1 = __ldg(b + 1);
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x) +
           __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__global__ void __launch_bounds__(128) k_bow_descend(const __grid_constant__ VocDev v, const uint8_t* __restrict__ desc,
                                                     long long frameStride, const int* __restrict__ nPerFrame, int nHost,
                                                     int nMax, int levelsup, uint32_t* __restrict__ word,
                                                     uint32_t* __restrict__ node, double* __restrict__ wOut) {
    const int frame = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = nPerFrame ? min(nPerFrame[frame], nMax) : nHost;
    if (i >= n) return;
    const uint4* f = reinterpret_cast<const uint4*>(desc + frame * frameStride + 32ll * i);
    const uint4 a0 = __ldg(f), a1 = __ldg(f + 1);
    const int nidLevel = v.L - levelsup;
    int finalId = 0, level = 0;
    uint32_t nid = 0;
    int2 ch = __ldg(v.child);
    do {
        ++level;
        int best = 0x7fffffff, bestId = 0;
        const int first = v.contiguous ? __ldg(v.childIdx + ch.x) : 0;
        for (int j = 0; j < ch.y; ++j) {
            const int id = v.contiguous ? first + j : __ldg(v.childIdx + ch.x + j);
            const int d = ham256(a0, a1, v.desc + 2ll * id);
            if (d < best) { best = d; bestId = id; }
        }
        finalId = bestId;
        if (level == nidLevel) nid = (uint32_t)finalId;
        ch = __ldg(v.child + finalId);
    } while (ch.y != 0);
    if (nidLevel > 0 && level < nidLevel) nid = (uint32_t)finalId;    // early leaf: see oracle/cpu_ref/bow_ref.h
    const double w = __ldg(v.weight + finalId);
    const long long o = (long long)frame * nMax + i;
    word[o] = w > 0 ? (uint32_t)__ldg(v.wordId + finalId) : 0xFFFFFFFFu;
    node[o] = nid;
    wOut[o] = w;
}

#define ASM_T 256

__device__ void bitonic_sort_u64(unsigned long long* key, int np2) {
    for (int k = 2; k <= np2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < np2; t += ASM_T) {
                const int p = t ^ j;
                if (p > t) {
                    const unsigned long long a = key[t], b = key[p];
                    const bool up = (t & k) == 0;
                    if ((a > b) == up) { key[t] = b; key[p] = a; }
                }
            }
            __syncthreads();
        }
}

// exclusive scan of one int per slot of flag[0..m) into pos[], returns the total; ASM_T threads, chunked
__device__ int block_scan_flags(const unsigned long long* key, int m, int shift, int* pos, int* warpSums, int* carry) {
    // flag(j) = j == 0 || key[j] >> shift != key[j-1] >> shift
    if (threadIdx.x == 0) *carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int base = 0; base < m; base += ASM_T) {
        const int j = base + threadIdx.x;
        int f = 0;
        if (j < m) f = (j == 0) || ((key[j] >> shift) != (key[j - 1] >> shift));
        int x = f;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warpSums[wid] = x;
        __syncthreads();
        if (wid == 0) {
            int s = lane < ASM_T / 32 ? warpSums[lane] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int y = __shfl_up_sync(0xffffffffu, s, o);
                if (lane >= o) s += y;
            }
            if (lane < ASM_T / 32) warpSums[lane] = s;
        }
        __syncthreads();
        const int c = *carry;
        const int incl = c + x + (wid > 0 ? warpSums[wid - 1] : 0);
        if (j < m) pos[j] = f ? incl - 1 : -1;
        __syncthreads();
        if (threadIdx.x == ASM_T - 1) *carry = incl;
        __syncthreads();
    }
    return *carry;
}

__global__ void __launch_bounds__(ASM_T) k_bow_assemble(const __grid_constant__ VocDev v, const int* __restrict__ nPerFrame,
                                                        int nHost, int nMax, const uint32_t* __restrict__ word,
                                                        const uint32_t* __restrict__ node, const double* __restrict__ wIn,
                                                        uint32_t* __restrict__ words, double* __restrict__ values,
                                                        int* __restrict__ nWords, uint32_t* __restrict__ nodeIds,
                                                        int32_t* __restrict__ offsets, int32_t* __restrict__ indices,
                                                        int* __restrict__ nNodes) {
    extern __shared__ __align__(16) unsigned char smemRaw[];
    __shared__ int warpSums[ASM_T / 32];
    __shared__ int carry, sM;
    __shared__ double sNorm;
    const int frame = blockIdx.x;
    const int n = nPerFrame ? min(nPerFrame[frame], nMax) : nHost;
    int np2 = 1;
    while (np2 < n) np2 <<= 1;
    unsigned long long* key = reinterpret_cast<unsigned long long*>(smemRaw);
    int* pos = reinterpret_cast<int*>(key + np2);
    const long long fo = (long long)frame * nMax;
    const uint32_t* fw = word + fo;
    const uint32_t* fn = node + fo;
    const double* fwt = wIn + fo;
    words += fo; values += fo; nodeIds += fo; indices += fo;
    offsets += (long long)frame * (nMax + 1);
    if (n == 0) {
        if (threadIdx.x == 0) { nWords[frame] = 0; nNodes[frame] = 0; offsets[0] = 0; }
        return;
    }
    // ---- FeatureVector: sort by (node, feature index); stop words sort to the end
    if (threadIdx.x == 0) sM = 0;
    for (int t = threadIdx.x; t < np2; t += ASM_T)
        key[t] = (t < n && fw[t] != 0xFFFFFFFFu) ? ((unsigned long long)fn[t] << 32 | (unsigned)t) : ~0ull;
    __syncthreads();
    {
        int c = 0;
        for (int t = threadIdx.x; t < n; t += ASM_T) c += fw[t] != 0xFFFFFFFFu;
        atomicAdd(&sM, c);
    }
    __syncthreads();
    bitonic_sort_u64(key, np2);
    const int m = sM;                                  // kept features
    const int nn = m ? block_scan_flags(key, m, 32, pos, warpSums, &carry) : 0;
    for (int j = threadIdx.x; j < m; j += ASM_T) {
        indices[j] = (int32_t)(key[j] & 0xFFFFFFFFu);
        if (pos[j] >= 0) { nodeIds[pos[j]] = (uint32_t)(key[j] >> 32); offsets[pos[j]] = j; }
    }
    if (threadIdx.x == 0) { offsets[nn] = m; nNodes[frame] = nn; }
    __syncthreads();
    // ---- BowVector: sort by (word, feature index)
    for (int t = threadIdx.x; t < np2; t += ASM_T)
        key[t] = (t < n && fw[t] != 0xFFFFFFFFu) ? ((unsigned long long)fw[t] << 32 | (unsigned)t) : ~0ull;
    __syncthreads();
    bitonic_sort_u64(key, np2);
    const int nw = m ? block_scan_flags(key, m, 32, pos, warpSums, &carry) : 0;
    const bool tf = v.weighting == 0 || v.weighting == 1;
    for (int j = threadIdx.x; j < m; j += ASM_T) {
        if (pos[j] < 0) continue;
        const uint32_t wd = (uint32_t)(key[j] >> 32);
        double s = fwt[(uint32_t)(key[j] & 0xFFFFFFFFu)];          // first occurrence: insert(id, w)
        if (tf)
            for (int e = j + 1; e < m && (uint32_t)(key[e] >> 32) == wd; ++e) s += fwt[(uint32_t)(key[e] & 0xFFFFFFFFu)];   // addWeight
        words[pos[j]] = wd;
        values[pos[j]] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        nWords[frame] = nw;
        double nrm = 0.0;
        if (v.norm == 1) { for (int i = 0; i < nw; ++i) nrm += fabs(values[i]); }
        else if (v.norm == 2) { for (int i = 0; i < nw; ++i) nrm = __dadd_rn(nrm, __dmul_rn(values[i], values[i])); nrm = sqrt(nrm); }
        else if (tf) nrm = (double)nw;                                // :1180-1186: divided by the number of words
        else nrm = 0.0;
        sNorm = nrm;
    }
    __syncthreads();
    const double nrm = sNorm;
    if (nrm > 0.0)
        for (int i = threadIdx.x; i < nw; i += ASM_T) values[i] = __ddiv_rn(values[i], nrm);
}

size_t orbfe_bow_assemble_smem(int nMax) {
    int np2 = 1;
    while (np2 < nMax) np2 <<= 1;
    return (size_t)np2 * 12;
}

cudaError_t orbfe_launch_bow_descend(cudaStream_t st, const VocDev& v, const uint8_t* dDesc, long long frameStride, const int* dN,
                                     int nHost, int nMax, int batch, int levelsup, uint32_t* dWord, uint32_t* dNode, double* dW,
                                     long long* launches) {
    if (nMax <= 0 || batch <= 0) return cudaSuccess;
    dim3 grid((nMax + 127) / 128, batch);
    k_bow_descend<<<grid, 128, 0, st>>>(v, dDesc, frameStride, dN, nHost, nMax, levelsup, dWord, dNode, dW);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t orbfe_launch_bow_assemble(cudaStream_t st, const VocDev& v, const int* dN, int nHost, int nMax, int batch,
                                      const uint32_t* dWord, const uint32_t* dNode, const double* dW, uint32_t* dWords,
                                      double* dValues, int* dNWords, uint32_t* dNodeIds, int32_t* dOffsets, int32_t* dIndices,
                                      int* dNNodes, long long* launches) {
    if (batch <= 0) return cudaSuccess;
    const size_t smem = orbfe_bow_assemble_smem(nMax > 0 ? nMax : 1);
    {   // the attribute is per device: one high-water mark per (device, kernel), see orbfe_ensure_dyn_smem
        const cudaError_t ea = orbfe_ensure_dyn_smem(reinterpret_cast<const void*>(k_bow_assemble), smem);
        if (ea != cudaSuccess) return ea;
    }
    k_bow_assemble<<<batch, ASM_T, smem, st>>>(v, dN, nHost, nMax, dWord, dNode, dW, dWords, dValues, dNWords, dNodeIds, dOffsets,
                                              dIndices, dNNodes);
    ++*launches;
    return cudaGetLastError();
}
