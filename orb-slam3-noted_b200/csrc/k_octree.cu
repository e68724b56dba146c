// k_octree.cu — deterministic GPU DistributeOctTree (ORBextractor.cc:479-761).
//
// One CTA per (frame, level).  The reference walks a std::list of nodes, each owning a vector of
// keypoints, splitting nodes into quadrants until the list holds >= N nodes.  Restated level-
// synchronously: the list is an ordered array of (box, count) in shared memory, every candidate
// carries the list position of the node that owns it (nodeOf[]), and one "pass" is
//   1. all candidates of expandable nodes vote their quadrant into cnt4[node][q] (shared atomics);
//   2. the new list order is computed with scans:
//        coarse pass (:593-660): every node with > 1 key is split; push_front while iterating means
//        new list = reverse(children in creation order) ++ untouched single-key nodes in old order;
//        fine pass (:671-735): nodes are split in descending (size, pointer) order until the list
//        reaches N.  Pointer order is a malloc artefact in the reference; the documented rule
//        (shared with the oracle) is "later-created node first", i.e. ascending list position;
//   3. candidates remap nodeOf[] through the child / survivor position tables.
// Finally each node keeps its max-response key, first in reference push order on ties (:748-755);
// push order is a function of (x, y) (cell-row-major, then row-major inside the cell), so it is
// folded into the 64-bit atomicMax key instead of being materialised.
#include "orbfe_internal.h"

// Threads per CTA are a template parameter chosen per launch: a level of a 1000-feature frame keeps <= ~250 nodes and 128 threads are
// fastest (barriers dominate: 0.160 -> 0.147 ms per 512 frames of 752x480), the 4000-feature levels of 1920x1080 frames want 256
// (27.5 k vs 30.7 k frames/s).  profiles/r04i_variants.txt, r04j_variants.txt.
#define OCT_MAX_THREADS 256

struct NodeS {
    short x0, x1, y0, y1;
    int cnt;
};

// In-place exclusive scan of data[0..n) by the whole CTA; returns the total to every thread.
template <int OCT_THREADS>
__device__ int block_exclusive_scan(int* data, int n, int* warpSums) {
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int per = (n + OCT_THREADS - 1) / OCT_THREADS;
    const int b = min(tid * per, n), e = min(b + per, n);
    int s = 0;
    for (int i = b; i < e; ++i) s += data[i];
    int inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
    }
    if (lane == 31) warpSums[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int w = lane < OCT_THREADS / 32 ? warpSums[lane] : 0;
        int winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += v;
        }
        if (lane < OCT_THREADS / 32) warpSums[lane] = winc - w;
        if (lane == OCT_THREADS / 32 - 1) warpSums[OCT_THREADS / 32] = winc;
    }
    __syncthreads();
    int run = warpSums[wid] + inc - s;
    for (int i = b; i < e; ++i) {
        const int v = data[i];
        data[i] = run;
        run += v;
    }
    const int total = warpSums[OCT_THREADS / 32];
    __syncthreads();
    return total;
}

// Ascending bitonic sort of n2 (power of two) 64-bit keys in shared memory.
template <int OCT_THREADS>
__device__ void block_bitonic_sort(unsigned long long* k, int n2) {
    for (int size = 2; size <= n2; size <<= 1)
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = threadIdx.x; i < n2 / 2; i += OCT_THREADS) {
                const int lo = 2 * i - (i & (stride - 1));
                const int hi = lo + stride;
                const bool up = (lo & size) == 0;
                const unsigned long long a = k[lo], b = k[hi];
                if ((a > b) == up) { k[lo] = b; k[hi] = a; }
            }
            __syncthreads();
        }
}

__device__ __forceinline__ int quadrant(const NodeS& nd, int x, int y) {
    const int mx = nd.x0 + ((nd.x1 - nd.x0 + 1) >> 1);   // UL.x + ceil((UR.x-UL.x)/2)
    const int my = nd.y0 + ((nd.y1 - nd.y0 + 1) >> 1);
    return (x >= mx ? 1 : 0) + (y >= my ? 2 : 0);
}

__device__ __forceinline__ NodeS child_box(const NodeS& nd, int q, int cnt) {
    const int mx = nd.x0 + ((nd.x1 - nd.x0 + 1) >> 1);
    const int my = nd.y0 + ((nd.y1 - nd.y0 + 1) >> 1);
    NodeS c;
    c.x0 = (q & 1) ? (short)mx : nd.x0;
    c.x1 = (q & 1) ? nd.x1 : (short)mx;
    c.y0 = (q & 2) ? (short)my : nd.y0;
    c.y1 = (q & 2) ? nd.y1 : (short)my;
    c.cnt = cnt;
    return c;
}

template <int OCT_THREADS>
__global__ void __launch_bounds__(OCT_THREADS) k_octree(const LevelGeom* __restrict__ geom, const uint32_t* __restrict__ cand,
                                                        uint32_t* __restrict__ nodeOfAll, const int* __restrict__ candCount,
                                                        long long candPerFrame, uint32_t* __restrict__ lvlKp,
                                                        int* __restrict__ lvlCount, int kpPerFrame, int nlevels, int Lcap,
                                                        int sortCap, int frame0) {
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ int warpSums[OCT_THREADS / 32 + 1];
    const int level = blockIdx.x, frame = frame0 + blockIdx.y, tid = threadIdx.x;
    const LevelGeom& g = geom[level];
    const int K = candCount[frame * nlevels + level];
    const uint32_t* keys = cand + (long long)frame * candPerFrame + g.candBase;
    uint32_t* nodeOf = nodeOfAll + (long long)frame * candPerFrame + g.candBase;
    const int N = g.nfeat;

    unsigned long long* skey = reinterpret_cast<unsigned long long*>(smem);       // [sortCap]
    NodeS* listA = reinterpret_cast<NodeS*>(skey + sortCap);                       // [Lcap]
    NodeS* listB = listA + Lcap;                                                   // [Lcap]
    int* cnt4 = reinterpret_cast<int*>(listB + Lcap);                              // [Lcap*4]
    int* scanA = cnt4 + Lcap * 4;                                                  // [Lcap]
    int* scanB = scanA + Lcap;                                                     // [Lcap]
    short* cpos = reinterpret_cast<short*>(scanB + Lcap);                          // [Lcap*4]
    short* upos = cpos + Lcap * 4;                                                 // [Lcap]

    NodeS* cur = listA;
    NodeS* nxt = listB;

    // ---- initial nodes (:541-586)
    const int nIni = g.nIni;
    const float hX = g.hX;
    for (int i = tid; i < nIni; i += OCT_THREADS) {
        NodeS nd;
        nd.x0 = (short)(int)__fmul_rn(hX, (float)i);
        nd.x1 = (short)(int)__fmul_rn(hX, (float)(i + 1));
        nd.y0 = 0;
        nd.y1 = (short)(g.maxBY - ORBFE_MIN_BORDER);
        nd.cnt = 0;
        cur[i] = nd;
    }
    __syncthreads();
    for (int k = tid; k < K; k += OCT_THREADS) {
        const int x = ORBFE_CAND_X(keys[k]);
        const int p = (int)__fdiv_rn((float)x, hX);
        nodeOf[k] = (uint32_t)p;
        atomicAdd(&cur[p].cnt, 1);
    }
    __syncthreads();
    // drop empty initial nodes, keep order (:570-586)
    for (int i = tid; i < nIni; i += OCT_THREADS) scanA[i] = cur[i].cnt > 0;
    __syncthreads();
    int L = block_exclusive_scan<OCT_THREADS>(scanA, nIni, warpSums);
    for (int i = tid; i < nIni; i += OCT_THREADS) {
        upos[i] = (short)scanA[i];
        if (cur[i].cnt > 0) nxt[scanA[i]] = cur[i];
    }
    __syncthreads();
    for (int k = tid; k < K; k += OCT_THREADS) nodeOf[k] = (uint32_t)upos[nodeOf[k]];
    { NodeS* t = cur; cur = nxt; nxt = t; }
    __syncthreads();

    bool fine = false;
    while (true) {
        const int prevSize = L;
        // 1. quadrant votes of every key that sits in an expandable node
        for (int i = tid; i < L * 4; i += OCT_THREADS) cnt4[i] = 0;
        __syncthreads();
        for (int k = tid; k < K; k += OCT_THREADS) {
            const int p = (int)nodeOf[k];
            const NodeS nd = cur[p];
            if (nd.cnt > 1) {
                const uint32_t c = keys[k];
                atomicAdd(&cnt4[p * 4 + quadrant(nd, ORBFE_CAND_X(c), ORBFE_CAND_Y(c))], 1);
            }
        }
        __syncthreads();
        int TC, newL, nToExpand = 0;
        if (!fine) {
            // 2a. coarse pass: split every expandable node, in list order
            for (int p = tid; p < L; p += OCT_THREADS) {
                int nch = 0, nex = 0;
                if (cur[p].cnt > 1) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) { nch += cnt4[p * 4 + q] > 0; nex += cnt4[p * 4 + q] > 1; }
                }
                scanA[p] = nch | ((cur[p].cnt == 1) << 16);
                scanB[p] = nex;
            }
            __syncthreads();
            const int tot = block_exclusive_scan<OCT_THREADS>(scanA, L, warpSums);
            nToExpand = block_exclusive_scan<OCT_THREADS>(scanB, L, warpSums);
            TC = tot & 0xFFFF;
            newL = TC + (tot >> 16);
            for (int p = tid; p < L; p += OCT_THREADS) {
                const NodeS nd = cur[p];
                if (nd.cnt > 1) {
                    int ci = scanA[p] & 0xFFFF;
                    upos[p] = -1;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int c = cnt4[p * 4 + q];
                        if (c > 0) {
                            const int np = TC - 1 - ci;
                            cpos[p * 4 + q] = (short)np;
                            nxt[np] = child_box(nd, q, c);
                            ++ci;
                        }
                    }
                } else {
                    const int np = TC + (scanA[p] >> 16);
                    upos[p] = (short)np;
                    nxt[np] = nd;
                }
            }
        } else {
            // 2b. fine pass: split in descending (size, later-created-first) order until the list holds N
            int n2 = 1;
            while (n2 < L) n2 <<= 1;
            for (int p = tid; p < n2; p += OCT_THREADS) {
                unsigned long long key = ~0ull;
                if (p < L && cur[p].cnt > 1) key = ((unsigned long long)(0xFFFFFFFFu - (uint32_t)cur[p].cnt) << 32) | (uint32_t)p;
                skey[p] = key;
            }
            __syncthreads();
            block_bitonic_sort<OCT_THREADS>(skey, n2);
            // growth of the list if sorted entry j is split: (#non-empty children - 1)
            for (int j = tid; j < L; j += OCT_THREADS) {
                int grow = 0;
                if (skey[j] != ~0ull) {
                    const int p = (int)(skey[j] & 0xFFFFFFFFu);
                    int nch = 0;
#pragma unroll
                    for (int q = 0; q < 4; ++q) nch += cnt4[p * 4 + q] > 0;
                    grow = (nch - 1) | (1 << 16);   // high half counts expandable entries
                }
                scanA[j] = grow;
            }
            __syncthreads();
            const int tot = block_exclusive_scan<OCT_THREADS>(scanA, L, warpSums);
            const int E = tot >> 16;
            // entry j is split iff the list was still short before it: L + growth of entries < j  < N
            // (the reference breaks right after the split that reaches N, :728-729)
            for (int p = tid; p < L; p += OCT_THREADS) { upos[p] = 0; scanB[p] = 0; }
            __syncthreads();
            for (int j = tid; j < E; j += OCT_THREADS) {
                const int before = L + (scanA[j] & 0xFFFF);
                const int p = (int)(skey[j] & 0xFFFFFFFFu);
                if (before < N) {
                    upos[p] = -1;
                    int nch = 0;
#pragma unroll
                    for (int q = 0; q < 4; ++q) nch += cnt4[p * 4 + q] > 0;
                    scanB[j] = nch;
                }
            }
            __syncthreads();
            TC = block_exclusive_scan<OCT_THREADS>(scanB, L, warpSums);            // creation index over sorted order
            for (int p = tid; p < L; p += OCT_THREADS) scanA[p] = upos[p] != -1;
            __syncthreads();
            const int nKeep = block_exclusive_scan<OCT_THREADS>(scanA, L, warpSums);   // survivors keep list order
            newL = TC + nKeep;
            for (int j = tid; j < E; j += OCT_THREADS) {
                const int p = (int)(skey[j] & 0xFFFFFFFFu);
                if (upos[p] == -1) {
                    const NodeS nd = cur[p];
                    int ci = scanB[j];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int c = cnt4[p * 4 + q];
                        if (c > 0) {
                            const int np = TC - 1 - ci;
                            cpos[p * 4 + q] = (short)np;
                            nxt[np] = child_box(nd, q, c);
                            ++ci;
                        }
                    }
                }
            }
            for (int p = tid; p < L; p += OCT_THREADS)
                if (upos[p] != -1) {
                    const int np = TC + scanA[p];
                    upos[p] = (short)np;
                    nxt[np] = cur[p];
                }
        }
        __syncthreads();
        // 3. keys follow their node
        for (int k = tid; k < K; k += OCT_THREADS) {
            const int p = (int)nodeOf[k];
            int np = upos[p];
            if (np == -1) {
                const uint32_t c = keys[k];
                np = cpos[p * 4 + quadrant(cur[p], ORBFE_CAND_X(c), ORBFE_CAND_Y(c))];
            }
            nodeOf[k] = (uint32_t)np;
        }
        __syncthreads();
        { NodeS* t = cur; cur = nxt; nxt = t; }
        L = newL;
        // finish tests (:667 and :732)
        if (L >= N || L == prevSize) break;
        if (!fine && L + nToExpand * 3 > N) fine = true;
    }

    // ---- best key per node (:742-758): max response, earliest push order on ties
    unsigned long long* best = skey;
    for (int i = tid; i < L; i += OCT_THREADS) best[i] = 0ull;
    __syncthreads();
    for (int k = tid; k < K; k += OCT_THREADS) {
        const uint32_t c = keys[k];
        const int x = ORBFE_CAND_X(c), y = ORBFE_CAND_Y(c);
        const int cj = (x - 3) / g.wCell, ci = (y - 3) / g.hCell;
        const unsigned long long order =
            ((unsigned long long)(ci * g.nCols + cj) << 14) | (unsigned long long)((y - ci * g.hCell) << 7) | (unsigned long long)(x - cj * g.wCell);
        const unsigned long long v = ((unsigned long long)ORBFE_CAND_R(c) << 56) | ((0x3FFFFFFFFull - order) << 22) |
                                     ((unsigned long long)y << 11) | (unsigned long long)x;
        atomicMax(&best[nodeOf[k]], v);
    }
    __syncthreads();
    uint32_t* out = lvlKp + (long long)frame * kpPerFrame + g.kpBase;
    for (int i = tid; i < L; i += OCT_THREADS) {
        const unsigned long long v = best[i];
        out[i] = (uint32_t)(v & 0x3FFFFFu) | ((uint32_t)(v >> 56) << 22);
    }
    if (tid == 0) lvlCount[frame * nlevels + level] = L;
}

static size_t octree_smem_bytes(int Lcap, int sortCap) {
    return (size_t)sortCap * 8 + (size_t)Lcap * (2 * sizeof(NodeS) + 4 * 4 + 4 + 4 + 4 * 2 + 2) + 64;
}

cudaError_t orbfe_launch_octree(orbfe_handle* h, int frame0, int batch) {
    int Lcap = 0;
    for (int l = 0; l < h->p.nlevels; ++l) Lcap = h->geom[l].kpCap > Lcap ? h->geom[l].kpCap : Lcap;
    Lcap = (Lcap + 7) & ~7;
    int sortCap = 1;
    while (sortCap < Lcap) sortCap <<= 1;
    const size_t smem = octree_smem_bytes(Lcap, sortCap);
    dim3 grid(h->p.nlevels, batch);
#define ORBFE_OCT_LAUNCH(NT)                                                                                                     \
    do {    /* the attribute is per device: one high-water mark per (device, kernel), see orbfe_ensure_dyn_smem */              \
        const cudaError_t ea = orbfe_ensure_dyn_smem(reinterpret_cast<const void*>(k_octree<NT>), smem);                         \
        if (ea != cudaSuccess) return ea;                                                                                        \
        k_octree<NT><<<grid, NT, smem, h->cur>>>(h->dGeom, h->dCand, h->dNodeOf, h->dCandCount, h->candPerFrame, h->dLvlKp,      \
                                                 h->dLvlCount, h->kpPerFrame, h->p.nlevels, Lcap, sortCap, frame0);             \
    } while (0)
    // 128 threads win on throughput (fewer lanes per barrier, more CTAs per SM); a launch that does not fill the GPU anyway
    // (single frames: 8 CTAs) is a matter of one CTA's latency, and that is shorter with 256 (28 vs 35 us)
    const bool fewCtas = (long long)batch * h->p.nlevels <= 2LL * h->smCount;
    if (Lcap <= 320 && !fewCtas) ORBFE_OCT_LAUNCH(128);     // up to ~1400 features per frame
    else ORBFE_OCT_LAUNCH(256);
#undef ORBFE_OCT_LAUNCH
    h->launches++;
    return cudaGetLastError();
}
