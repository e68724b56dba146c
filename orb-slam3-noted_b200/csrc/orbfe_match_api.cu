// orbfe_match_api.cu — C ABI of the matchers (include/orbfe.h, "matchers" section): the
// orbfe_matcher object (one stream, a pinned host / device staging arena pair and device scratch,
// all grown on demand), host-side flattening of the reference's inputs and the launches.
// No CPU fallback: every entry point needs the CUDA device the matcher was created on.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "orbfe_internal.h"
#include "orbfe_match.h"

int orbfe_set_cuda_error(cudaError_t e, const char* what);   // orbfe_api.cu

#define CK(call)                                                         \
    do {                                                                 \
        cudaError_t e_ = (call);                                         \
        if (e_ != cudaSuccess) return orbfe_set_cuda_error(e_, #call);   \
    } while (0)

struct orbfe_matcher {
    float nnRatio;
    int checkOri;
    int device;
    int smCount;
    int knnBackend;              // ORBFE_KNN2_POPC / ORBFE_KNN2_TENSOR
    cudaStream_t stream;
    cudaEvent_t ev;
    long long launches;
    // staging arena: host (pinned) and device mirrors with identical offsets
    uint8_t* hStage;
    uint8_t* dStage;
    size_t stageCap, stageUsed;
    // device-only scratch (knn2 partials, stereo SAD, counters)
    uint8_t* dScratch;
    size_t scratchCap;
};

static int ensure_stage(orbfe_matcher* m, size_t bytes) {
    if (bytes <= m->stageCap) return ORBFE_OK;
    CK(cudaStreamSynchronize(m->stream));
    size_t cap = std::max(bytes + bytes / 4, (size_t)1 << 20);
    if (m->hStage) cudaFreeHost(m->hStage);
    if (m->dStage) cudaFree(m->dStage);
    m->hStage = nullptr; m->dStage = nullptr; m->stageCap = 0;
    CK(cudaMallocHost(&m->hStage, cap));
    CK(cudaMalloc(&m->dStage, cap));
    m->stageCap = cap;
    return ORBFE_OK;
}

static int ensure_scratch(orbfe_matcher* m, size_t bytes) {
    if (bytes <= m->scratchCap) return ORBFE_OK;
    CK(cudaStreamSynchronize(m->stream));
    size_t cap = std::max(bytes + bytes / 4, (size_t)1 << 20);
    if (m->dScratch) cudaFree(m->dScratch);
    m->dScratch = nullptr; m->scratchCap = 0;
    CK(cudaMalloc(&m->dScratch, cap));
    m->scratchCap = cap;
    return ORBFE_OK;
}

// Bump allocation inside the staging arena (256-byte aligned so uint4 / float4 loads are legal).
static inline size_t stage_take(size_t& used, size_t bytes) {
    const size_t off = (used + 255) & ~(size_t)255;
    used = off + bytes;
    return off;
}

extern "C" int orbfe_matcher_create(float nnratio, int check_orientation, int device, orbfe_matcher** out) {
    if (!out) return ORBFE_ERR_BAD_ARGUMENT;
    *out = nullptr;
    if (!(nnratio > 0.f)) return ORBFE_ERR_BAD_ARGUMENT;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) { cudaGetLastError(); return ORBFE_ERR_NO_DEVICE; }
    if (device < 0 || device >= ndev) return ORBFE_ERR_BAD_ARGUMENT;
    orbfe_matcher* m = new orbfe_matcher();
    memset(m, 0, sizeof(*m));
    m->nnRatio = nnratio;
    m->checkOri = check_orientation ? 1 : 0;
    m->device = device;
    {
        const char* env = getenv("ORBFE_KNN2");
        m->knnBackend = env && !strcmp(env, "popc") ? ORBFE_KNN2_POPC : env && !strcmp(env, "tensor") ? ORBFE_KNN2_TENSOR : ORBFE_KNN2_DEFAULT;
    }
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&m->ev, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&m->smCount, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess) { orbfe_set_cuda_error(e, "orbfe_matcher_create"); orbfe_matcher_destroy(m); return ORBFE_ERR_CUDA; }
    *out = m;
    return ORBFE_OK;
}

extern "C" void orbfe_matcher_destroy(orbfe_matcher* m) {
    if (!m) return;
    cudaSetDevice(m->device);
    if (m->stream) cudaStreamSynchronize(m->stream);
    if (m->hStage) cudaFreeHost(m->hStage);
    if (m->dStage) cudaFree(m->dStage);
    if (m->dScratch) cudaFree(m->dScratch);
    if (m->ev) cudaEventDestroy(m->ev);
    if (m->stream) cudaStreamDestroy(m->stream);
    delete m;
}

extern "C" void* orbfe_matcher_stream(orbfe_matcher* m) { return m ? (void*)m->stream : nullptr; }
extern "C" int orbfe_matcher_synchronize(orbfe_matcher* m) {
    if (!m) return ORBFE_ERR_BAD_ARGUMENT;
    CK(cudaSetDevice(m->device));
    CK(cudaStreamSynchronize(m->stream));
    return ORBFE_OK;
}
extern "C" long long orbfe_matcher_launch_count(const orbfe_matcher* m) { return m ? m->launches : 0; }

// ---- brute-force 2-NN ----------------------------------------------------------------------------------
extern "C" int orbfe_hamming_knn2_batch_device(orbfe_matcher* m, const uint8_t* d_q, ptrdiff_t q_stride, const int* d_nq,
                                               int nq_max, const uint8_t* d_t, ptrdiff_t t_stride, const int* d_nt, int nt_max,
                                               int batch, int32_t* d_idx2, int32_t* d_dist2) {
    if (!m || !d_q || !d_t || !d_idx2 || !d_dist2 || nq_max < 0 || nt_max < 0 || batch < 0) return ORBFE_ERR_BAD_ARGUMENT;
    if (nt_max >= (1 << 20)) return ORBFE_ERR_TOO_LARGE;
    if (nq_max == 0 || batch == 0) return ORBFE_OK;
    CK(cudaSetDevice(m->device));
    const bool tensor = m->knnBackend == ORBFE_KNN2_TENSOR;
    int nsplit = 1, splitLen = nt_max;
    if (tensor) orbfe_knn2_tc_plan(nq_max, nt_max, batch, m->smCount, &nsplit, &splitLen);
    else orbfe_knn2_plan(nq_max, nt_max, batch, m->smCount, &nsplit, &splitLen);
    uint2* partial = nullptr;
    if (nsplit > 1) {
        int st = ensure_scratch(m, sizeof(uint2) * (size_t)batch * nsplit * nq_max);
        if (st != ORBFE_OK) return st;
        partial = reinterpret_cast<uint2*>(m->dScratch);
    }
    if (tensor)
        CK(orbfe_launch_knn2_tc(m->stream, d_q, q_stride, d_nq, nq_max, d_t, t_stride, d_nt, nt_max, batch, nsplit, splitLen, partial, d_idx2,
                                d_dist2, &m->launches));
    else
        CK(orbfe_launch_knn2(m->stream, d_q, q_stride, d_nq, nq_max, d_t, t_stride, d_nt, nt_max, batch, nsplit, splitLen, partial, d_idx2,
                             d_dist2, &m->launches));
    return ORBFE_OK;
}

extern "C" int orbfe_matcher_set_knn2_backend(orbfe_matcher* m, int backend) {
    if (!m || (backend != ORBFE_KNN2_POPC && backend != ORBFE_KNN2_TENSOR)) return ORBFE_ERR_BAD_ARGUMENT;
    m->knnBackend = backend;
    return ORBFE_OK;
}

// Stages q and t, runs the 2-NN; leaves idx2 / dist2 (nq x 2 each) in the device staging arena at the returned offsets.
static int knn2_staged(orbfe_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt, size_t* offIdx, size_t* offDist,
                       size_t* used) {
    size_t u = 0;
    const size_t oq = stage_take(u, (size_t)nq * 32), ot = stage_take(u, (size_t)std::max(nt, 1) * 32);
    const size_t inBytes = u;
    const size_t oi = stage_take(u, sizeof(int32_t) * 2 * nq), od = stage_take(u, sizeof(int32_t) * 2 * nq);
    const size_t extra = stage_take(u, 256 + sizeof(int32_t) * (size_t)nq);   // room for the lapping filter's outputs
    (void)extra;
    int st = ensure_stage(m, u);
    if (st != ORBFE_OK) return st;
    memcpy(m->hStage + oq, q, (size_t)nq * 32);
    if (nt > 0) memcpy(m->hStage + ot, t, (size_t)nt * 32);
    CK(cudaMemcpyAsync(m->dStage, m->hStage, inBytes, cudaMemcpyHostToDevice, m->stream));
    st = orbfe_hamming_knn2_batch_device(m, m->dStage + oq, 0, nullptr, nq, m->dStage + ot, 0, nullptr, nt, 1,
                                         reinterpret_cast<int32_t*>(m->dStage + oi), reinterpret_cast<int32_t*>(m->dStage + od));
    if (st != ORBFE_OK) return st;
    *offIdx = oi; *offDist = od; *used = u;
    return ORBFE_OK;
}

extern "C" int orbfe_hamming_knn2(orbfe_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx2,
                                  int32_t* dist2) {
    if (!m || nq < 0 || nt < 0 || (nq > 0 && (!q || !idx2 || !dist2)) || (nt > 0 && !t)) return ORBFE_ERR_BAD_ARGUMENT;
    if (nq == 0) return ORBFE_OK;
    CK(cudaSetDevice(m->device));
    size_t oi, od, used;
    int st = knn2_staged(m, q, nq, t, nt, &oi, &od, &used);
    if (st != ORBFE_OK) return st;
    // idx and dist are adjacent allocations: one copy back covers both
    CK(cudaMemcpyAsync(m->hStage + oi, m->dStage + oi, od + sizeof(int32_t) * 2 * nq - oi, cudaMemcpyDeviceToHost, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    memcpy(idx2, m->hStage + oi, sizeof(int32_t) * 2 * nq);
    memcpy(dist2, m->hStage + od, sizeof(int32_t) * 2 * nq);
    return ORBFE_OK;
}

extern "C" int orbfe_match_lapping(orbfe_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt, double ratio,
                                   int32_t* match) {
    if (!m || nq < 0 || nt < 0 || (nq > 0 && (!q || !match)) || (nt > 0 && !t)) return ORBFE_ERR_BAD_ARGUMENT;
    if (nq == 0) return 0;
    CK(cudaSetDevice(m->device));
    size_t oi, od, used;
    int st = knn2_staged(m, q, nq, t, nt, &oi, &od, &used);
    if (st != ORBFE_OK) return st;
    const size_t oc = (od + sizeof(int32_t) * 2 * nq + 255) & ~(size_t)255, om = oc + 256;
    CK(cudaMemsetAsync(m->dStage + oc, 0, sizeof(int), m->stream));
    const double r = ratio;   // the reference multiplies DMatch::distance (float) by the double literal 0.7 (Frame.cc:1141)
    CK(orbfe_launch_ratio_filter(m->stream, reinterpret_cast<int32_t*>(m->dStage + oi), reinterpret_cast<int32_t*>(m->dStage + od), nq,
                                 r, reinterpret_cast<int32_t*>(m->dStage + om), reinterpret_cast<int*>(m->dStage + oc),
                                 &m->launches));
    CK(cudaMemcpyAsync(m->hStage + oc, m->dStage + oc, 256 + sizeof(int32_t) * (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    memcpy(match, m->hStage + om, sizeof(int32_t) * (size_t)nq);
    return *reinterpret_cast<int*>(m->hStage + oc);
}

// ---- rectified stereo ----------------------------------------------------------------------------------
// Orders the matcher's stream after everything already enqueued on an extractor's stream.
static int wait_for(orbfe_matcher* m, orbfe_handle* h) {
    CK(cudaEventRecord(m->ev, h->stream));
    CK(cudaStreamWaitEvent(m->stream, m->ev, 0));
    return ORBFE_OK;
}

// ... and the reverse: the extractor's next call (which overwrites its results and pyramid) waits for what the matcher enqueued.
static int release_to(orbfe_matcher* m, orbfe_handle* h) {
    CK(cudaEventRecord(m->ev, m->stream));
    CK(cudaStreamWaitEvent(h->stream, m->ev, 0));
    return ORBFE_OK;
}

extern "C" int orbfe_stereo_match_batch_device(orbfe_matcher* m, orbfe_handle* h, int npairs, float bf, float b, float* d_u_right,
                                               float* d_depth) {
    if (!m || !h || !d_u_right || !d_depth || npairs <= 0 || !(b > 0.f)) return ORBFE_ERR_BAD_ARGUMENT;
    if (h->device != m->device) return ORBFE_ERR_BAD_ARGUMENT;
    if (h->lastBatch < 2 * npairs) return ORBFE_ERR_NOT_READY;
    CK(cudaSetDevice(m->device));
    int st = ensure_scratch(m, sizeof(int) * ((size_t)npairs * h->kpCapPerFrame + npairs));
    if (st != ORBFE_OK) return st;
    if ((st = wait_for(m, h)) != ORBFE_OK) return st;
    int* dSad = reinterpret_cast<int*>(m->dScratch);
    int* dCount = dSad + (size_t)npairs * h->kpCapPerFrame;
    CK(orbfe_launch_stereo(m->stream, h, 0, 2, h, 1, 2, npairs, bf, b, d_u_right, d_depth, dSad, dCount, h->kpCapPerFrame, &m->launches));
    return release_to(m, h);
}

extern "C" int orbfe_stereo_match(orbfe_matcher* m, orbfe_handle* left, int frame_l, orbfe_handle* right, int frame_r, float bf,
                                  float b, float* u_right, float* depth, int cap) {
    if (!m || !left || !right || !u_right || !depth || !(b > 0.f)) return ORBFE_ERR_BAD_ARGUMENT;
    if (left->device != m->device || right->device != m->device) return ORBFE_ERR_BAD_ARGUMENT;
    if (frame_l < 0 || frame_l >= left->lastBatch || frame_r < 0 || frame_r >= right->lastBatch) return ORBFE_ERR_NOT_READY;
    if (left->curW != right->curW || left->curH != right->curH || left->p.nlevels != right->p.nlevels) return ORBFE_ERR_BAD_ARGUMENT;
    CK(cudaSetDevice(m->device));
    const int kpf = left->kpPerFrame;
    size_t u = 0;
    const size_t oU = stage_take(u, sizeof(float) * kpf), oD = stage_take(u, sizeof(float) * kpf), oN = stage_take(u, 256);
    int st = ensure_stage(m, u);
    if (st != ORBFE_OK) return st;
    if ((st = ensure_scratch(m, sizeof(int) * ((size_t)kpf + 64))) != ORBFE_OK) return st;
    if ((st = wait_for(m, left)) != ORBFE_OK) return st;
    if (right != left && (st = wait_for(m, right)) != ORBFE_OK) return st;
    int* dSad = reinterpret_cast<int*>(m->dScratch);
    int* dCount = dSad + kpf;
    float* dU = reinterpret_cast<float*>(m->dStage + oU);
    float* dD = reinterpret_cast<float*>(m->dStage + oD);
    CK(orbfe_launch_stereo(m->stream, left, frame_l, 0, right, frame_r, 0, 1, bf, b, dU, dD, dSad, dCount, kpf, &m->launches));
    CK(cudaMemcpyAsync(m->dStage + oN, left->dN + frame_l, sizeof(int), cudaMemcpyDeviceToDevice, m->stream));
    CK(cudaMemcpyAsync(m->dStage + oN + 4, dCount, sizeof(int), cudaMemcpyDeviceToDevice, m->stream));
    CK(cudaMemcpyAsync(m->hStage, m->dStage, u, cudaMemcpyDeviceToHost, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    const int nL = reinterpret_cast<int*>(m->hStage + oN)[0], cnt = reinterpret_cast<int*>(m->hStage + oN)[1];
    if (nL > cap) return ORBFE_ERR_CAPACITY;
    memcpy(u_right, m->hStage + oU, sizeof(float) * nL);
    memcpy(depth, m->hStage + oD, sizeof(float) * nL);
    return cnt;
}

// ---- BoW-guided searches --------------------------------------------------------------------------------
struct SideLayout {
    size_t desc, idx, mask, kp, ur, out;   // offsets in the staging arena
};

static bool valid_side(const orbfe_bow_side& s) {
    if (s.n < 0 || (s.n > 0 && !s.desc)) return false;
    if (s.fv.nnodes < 0 || (s.fv.nnodes > 0 && (!s.fv.node_ids || !s.fv.offsets || !s.fv.indices))) return false;
    for (int k = 0; k < s.fv.nnodes; ++k) {
        if (s.fv.offsets[k] > s.fv.offsets[k + 1] || s.fv.offsets[k] < 0) return false;
        if (k > 0 && s.fv.node_ids[k] <= s.fv.node_ids[k - 1]) return false;
    }
    const int tot = s.fv.nnodes ? s.fv.offsets[s.fv.nnodes] : 0;
    for (int i = 0; i < tot; ++i)
        if (s.fv.indices[i] < 0 || s.fv.indices[i] >= s.n) return false;
    return true;
}

static void reserve_side(const orbfe_bow_side& s, bool needUr, size_t& u, SideLayout& L) {
    const int tot = s.fv.nnodes ? s.fv.offsets[s.fv.nnodes] : 0;
    L.desc = stage_take(u, (size_t)std::max(s.n, 1) * 32);
    L.idx = stage_take(u, sizeof(int32_t) * std::max(tot, 1));
    L.mask = stage_take(u, (size_t)std::max(s.n, 1));
    L.kp = stage_take(u, sizeof(float4) * std::max(s.n, 1));
    L.ur = needUr ? stage_take(u, sizeof(float) * std::max(s.n, 1)) : 0;
}

// maskDefault: value used when the caller passes no mask.
static void fill_side(uint8_t* base, const orbfe_bow_side& s, const SideLayout& L, uint8_t maskDefault, const float* uRight) {
    const int tot = s.fv.nnodes ? s.fv.offsets[s.fv.nnodes] : 0;
    if (s.n > 0) memcpy(base + L.desc, s.desc, (size_t)s.n * 32);
    if (tot > 0) memcpy(base + L.idx, s.fv.indices, sizeof(int32_t) * tot);
    if (s.mask) memcpy(base + L.mask, s.mask, s.n);
    else memset(base + L.mask, maskDefault, std::max(s.n, 1));
    float4* kp = reinterpret_cast<float4*>(base + L.kp);
    for (int i = 0; i < s.n; ++i) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (s.kps) {
            v.x = s.kps[i].x; v.y = s.kps[i].y; v.z = s.kps[i].angle;
            memcpy(&v.w, &s.kps[i].octave, 4);
        }
        kp[i] = v;
    }
    if (L.ur) {
        float* ur = reinterpret_cast<float*>(base + L.ur);
        for (int i = 0; i < s.n; ++i) ur[i] = uRight ? uRight[i] : -1.f;
    }
}

// The merge-join of ORBmatcher.cc:289-429 over two CSR FeatureVectors.
static void join_nodes(int prob, const orbfe_feature_vector& a, const orbfe_feature_vector& b, std::vector<MatchItem>& items) {
    int i = 0, j = 0;
    while (i < a.nnodes && j < b.nnodes) {
        if (a.node_ids[i] == b.node_ids[j]) {
            MatchItem it{prob, a.offsets[i], a.offsets[i + 1], b.offsets[j], b.offsets[j + 1]};
            if (it.a1 > it.a0 && it.b1 > it.b0) items.push_back(it);
            ++i; ++j;
        } else if (a.node_ids[i] < b.node_ids[j]) ++i;
        else ++j;
    }
}

extern "C" int orbfe_search_bow(orbfe_matcher* m, int form, const orbfe_bow_side* A, const orbfe_bow_side* B, int nprob,
                                int b_shared, int nleft_b, int32_t* const* a2b, int32_t* const* b2a, int* nmatches) {
    if (!m || !A || !B || nprob <= 0 || (form != 0 && form != 1)) return ORBFE_ERR_BAD_ARGUMENT;
    const int nB = b_shared ? 1 : nprob;
    for (int p = 0; p < nprob; ++p)
        if (!valid_side(A[p]) || !A[p].mask || (m->checkOri && !A[p].kps)) return ORBFE_ERR_BAD_ARGUMENT;
    for (int p = 0; p < nB; ++p)
        if (!valid_side(B[p]) || (form == 1 && !B[p].mask) || (m->checkOri && !B[p].kps) || B[p].n >= 65536)
            return ORBFE_ERR_BAD_ARGUMENT;
    CK(cudaSetDevice(m->device));
    // ---- layout: inputs first (one H2D), then outputs (one D2H)
    size_t u = 0;
    std::vector<SideLayout> LA(nprob), LB(nB);
    for (int p = 0; p < nprob; ++p) reserve_side(A[p], false, u, LA[p]);
    for (int p = 0; p < nB; ++p) reserve_side(B[p], false, u, LB[p]);
    std::vector<MatchItem> items;
    for (int p = 0; p < nprob; ++p) join_nodes(p, A[p].fv, B[b_shared ? 0 : p].fv, items);
    const size_t oProb = stage_take(u, sizeof(MatchProblem) * nprob);
    const size_t oItems = stage_take(u, sizeof(MatchItem) * std::max<size_t>(items.size(), 1));
    const size_t inBytes = u;
    const size_t oOut = stage_take(u, 0);
    std::vector<size_t> oA2B(nprob), oB2A(nprob);
    for (int p = 0; p < nprob; ++p) {
        oA2B[p] = stage_take(u, sizeof(int32_t) * std::max(A[p].n, 1));
        oB2A[p] = stage_take(u, sizeof(int32_t) * std::max(B[b_shared ? 0 : p].n, 1));
    }
    const size_t oCnt = stage_take(u, sizeof(int) * nprob);
    int st = ensure_stage(m, u);
    if (st != ORBFE_OK) return st;
    CK(cudaStreamSynchronize(m->stream));   // the host arena may still be the source / target of an earlier copy
    for (int p = 0; p < nprob; ++p) fill_side(m->hStage, A[p], LA[p], 1, nullptr);
    for (int p = 0; p < nB; ++p) fill_side(m->hStage, B[p], LB[p], 1, nullptr);
    MatchProblem* probs = reinterpret_cast<MatchProblem*>(m->hStage + oProb);
    for (int p = 0; p < nprob; ++p) {
        const int q = b_shared ? 0 : p;
        MatchProblem P;
        memset(&P, 0, sizeof(P));
        P.descA = m->dStage + LA[p].desc; P.descB = m->dStage + LB[q].desc;
        P.idxA = reinterpret_cast<const int32_t*>(m->dStage + LA[p].idx);
        P.idxB = reinterpret_cast<const int32_t*>(m->dStage + LB[q].idx);
        P.maskA = m->dStage + LA[p].mask;
        P.maskB = (form == 1) ? m->dStage + LB[q].mask : nullptr;
        P.kpA = reinterpret_cast<const float4*>(m->dStage + LA[p].kp);
        P.kpB = reinterpret_cast<const float4*>(m->dStage + LB[q].kp);
        P.a2b = reinterpret_cast<int32_t*>(m->dStage + oA2B[p]);
        P.b2a = reinterpret_cast<int32_t*>(m->dStage + oB2A[p]);
        P.nA = A[p].n; P.nB = B[q].n;
        probs[p] = P;
    }
    if (!items.empty()) memcpy(m->hStage + oItems, items.data(), sizeof(MatchItem) * items.size());
    CK(cudaMemcpyAsync(m->dStage, m->hStage, inBytes, cudaMemcpyHostToDevice, m->stream));
    CK(cudaMemsetAsync(m->dStage + oOut, 0xFF, u - oOut, m->stream));   // every match slot = -1
    CK(orbfe_launch_bow(m->stream, reinterpret_cast<const MatchProblem*>(m->dStage + oProb), nprob,
                        reinterpret_cast<const MatchItem*>(m->dStage + oItems), (int)items.size(), m->nnRatio, ORBFE_TH_LOW,
                        form == 1, form == 0 ? nleft_b : -1, m->checkOri, reinterpret_cast<int*>(m->dStage + oCnt), &m->launches));
    CK(cudaMemcpyAsync(m->hStage + oOut, m->dStage + oOut, u - oOut, cudaMemcpyDeviceToHost, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    for (int p = 0; p < nprob; ++p) {
        const int q = b_shared ? 0 : p;
        if (a2b && a2b[p] && A[p].n > 0) memcpy(a2b[p], m->hStage + oA2B[p], sizeof(int32_t) * A[p].n);
        if (b2a && b2a[p] && B[q].n > 0) memcpy(b2a[p], m->hStage + oB2A[p], sizeof(int32_t) * B[q].n);
        if (nmatches) nmatches[p] = reinterpret_cast<int*>(m->hStage + oCnt)[p];
    }
    return ORBFE_OK;
}

extern "C" int orbfe_search_triangulation(orbfe_matcher* m, const orbfe_bow_side* A, const float* u_right_a,
                                          const orbfe_bow_side* B, const float* u_right_b, const orbfe_epipolar* geo,
                                          int only_stereo, int coarse, int32_t* matches12) {
    if (!m || !A || !B || !geo || !matches12 || !valid_side(*A) || !valid_side(*B) || !A->kps || !B->kps) return ORBFE_ERR_BAD_ARGUMENT;
    if (!geo->sigma2_b || !geo->scale_b || geo->nlevels < 1 || geo->nlevels > ORBFE_MAX_LEVELS || B->n >= 65536)
        return ORBFE_ERR_BAD_ARGUMENT;
    for (int i = 0; i < B->n; ++i)
        if (B->kps[i].octave < 0 || B->kps[i].octave >= geo->nlevels) return ORBFE_ERR_BAD_ARGUMENT;
    CK(cudaSetDevice(m->device));
    size_t u = 0;
    SideLayout LA, LB;
    reserve_side(*A, true, u, LA);
    reserve_side(*B, true, u, LB);
    std::vector<MatchItem> items;
    join_nodes(0, A->fv, B->fv, items);
    const size_t oProb = stage_take(u, sizeof(MatchProblem));
    const size_t oItems = stage_take(u, sizeof(MatchItem) * std::max<size_t>(items.size(), 1));
    const size_t inBytes = u;
    const size_t oOut = stage_take(u, 0);
    const size_t oM = stage_take(u, sizeof(int32_t) * std::max(A->n, 1));
    const size_t oCnt = stage_take(u, sizeof(int));
    int st = ensure_stage(m, u);
    if (st != ORBFE_OK) return st;
    CK(cudaStreamSynchronize(m->stream));
    fill_side(m->hStage, *A, LA, 0, u_right_a);   // no mask = no feature has a MapPoint yet
    fill_side(m->hStage, *B, LB, 0, u_right_b);
    MatchProblem P;
    memset(&P, 0, sizeof(P));
    P.descA = m->dStage + LA.desc; P.descB = m->dStage + LB.desc;
    P.idxA = reinterpret_cast<const int32_t*>(m->dStage + LA.idx);
    P.idxB = reinterpret_cast<const int32_t*>(m->dStage + LB.idx);
    P.maskA = m->dStage + LA.mask; P.maskB = m->dStage + LB.mask;
    P.kpA = reinterpret_cast<const float4*>(m->dStage + LA.kp);
    P.kpB = reinterpret_cast<const float4*>(m->dStage + LB.kp);
    P.uRightA = reinterpret_cast<const float*>(m->dStage + LA.ur);
    P.uRightB = reinterpret_cast<const float*>(m->dStage + LB.ur);
    P.a2b = reinterpret_cast<int32_t*>(m->dStage + oM);
    P.b2a = nullptr;
    P.nA = A->n; P.nB = B->n;
    memcpy(m->hStage + oProb, &P, sizeof(P));
    if (!items.empty()) memcpy(m->hStage + oItems, items.data(), sizeof(MatchItem) * items.size());
    TriGeo g;
    memset(&g, 0, sizeof(g));
    memcpy(g.F, geo->F12, sizeof(g.F));
    g.ex = geo->ex; g.ey = geo->ey;
    for (int l = 0; l < geo->nlevels; ++l) { g.sigma2B[l] = geo->sigma2_b[l]; g.scaleB[l] = geo->scale_b[l]; }
    CK(cudaMemcpyAsync(m->dStage, m->hStage, inBytes, cudaMemcpyHostToDevice, m->stream));
    CK(cudaMemsetAsync(m->dStage + oOut, 0xFF, u - oOut, m->stream));
    CK(orbfe_launch_tri(m->stream, reinterpret_cast<const MatchProblem*>(m->dStage + oProb),
                        reinterpret_cast<const MatchItem*>(m->dStage + oItems), (int)items.size(), g, only_stereo, coarse,
                        m->checkOri, reinterpret_cast<int*>(m->dStage + oCnt), &m->launches));
    CK(cudaMemcpyAsync(m->hStage + oOut, m->dStage + oOut, u - oOut, cudaMemcpyDeviceToHost, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    if (A->n > 0) memcpy(matches12, m->hStage + oM, sizeof(int32_t) * A->n);
    return *reinterpret_cast<int*>(m->hStage + oCnt);
}

extern "C" int orbfe_search_triangulation_kb8(orbfe_matcher* m, const orbfe_bow_side* A, const orbfe_bow_side* B,
                                              const orbfe_kb8_epipolar* geo, int only_stereo, int coarse, int32_t* matches12,
                                              float* points12) {
    if (!m || !A || !B || !geo || !matches12 || !valid_side(*A) || !valid_side(*B) || !A->kps || !B->kps) return ORBFE_ERR_BAD_ARGUMENT;
    if (!geo->sigma2_a || !geo->sigma2_b || !geo->scale_b || geo->nlevels < 1 || geo->nlevels > ORBFE_MAX_LEVELS || B->n >= 65536)
        return ORBFE_ERR_BAD_ARGUMENT;
    if (geo->nleft_a < -1 || geo->nleft_a > A->n || geo->nleft_b < -1 || geo->nleft_b > B->n) return ORBFE_ERR_BAD_ARGUMENT;
    for (int i = 0; i < A->n; ++i)
        if (A->kps[i].octave < 0 || A->kps[i].octave >= geo->nlevels) return ORBFE_ERR_BAD_ARGUMENT;
    for (int i = 0; i < B->n; ++i)
        if (B->kps[i].octave < 0 || B->kps[i].octave >= geo->nlevels) return ORBFE_ERR_BAD_ARGUMENT;
    for (int i = 0; i < A->n; ++i) matches12[i] = -1;
    if (points12) memset(points12, 0, sizeof(float) * 3 * (size_t)A->n);
    if (only_stereo || A->n == 0) return 0;                       // bStereo1 is false for every fisheye feature
    CK(cudaSetDevice(m->device));
    size_t u = 0;
    SideLayout LA, LB;
    reserve_side(*A, false, u, LA);
    reserve_side(*B, false, u, LB);
    std::vector<MatchItem> items;
    join_nodes(0, A->fv, B->fv, items);
    const size_t oProb = stage_take(u, sizeof(MatchProblem));
    const size_t oItems = stage_take(u, sizeof(MatchItem) * std::max<size_t>(items.size(), 1));
    const size_t inBytes = u;
    const size_t oOut = stage_take(u, 0);
    const size_t oM = stage_take(u, sizeof(int32_t) * A->n);
    const size_t oCnt = stage_take(u, sizeof(int));
    const size_t oNeg = u;
    const size_t oPts = stage_take(u, sizeof(float) * 3 * A->n);
    int st = ensure_stage(m, u);
    if (st != ORBFE_OK) return st;
    CK(cudaStreamSynchronize(m->stream));
    fill_side(m->hStage, *A, LA, 0, nullptr);
    fill_side(m->hStage, *B, LB, 0, nullptr);
    MatchProblem P;
    memset(&P, 0, sizeof(P));
    P.descA = m->dStage + LA.desc; P.descB = m->dStage + LB.desc;
    P.idxA = reinterpret_cast<const int32_t*>(m->dStage + LA.idx);
    P.idxB = reinterpret_cast<const int32_t*>(m->dStage + LB.idx);
    P.maskA = m->dStage + LA.mask; P.maskB = m->dStage + LB.mask;
    P.kpA = reinterpret_cast<const float4*>(m->dStage + LA.kp);
    P.kpB = reinterpret_cast<const float4*>(m->dStage + LB.kp);
    P.a2b = reinterpret_cast<int32_t*>(m->dStage + oM);
    P.nA = A->n; P.nB = B->n;
    memcpy(m->hStage + oProb, &P, sizeof(P));
    if (!items.empty()) memcpy(m->hStage + oItems, items.data(), sizeof(MatchItem) * items.size());
    TriKb8Geo g;
    memset(&g, 0, sizeof(g));
    memcpy(g.camA, geo->cam_a, sizeof(g.camA));
    memcpy(g.camB, geo->cam_b, sizeof(g.camB));
    memcpy(g.R12, geo->R12, sizeof(g.R12));
    memcpy(g.t12, geo->t12, sizeof(g.t12));
    g.ex = geo->ex; g.ey = geo->ey; g.nleftA = geo->nleft_a; g.nleftB = geo->nleft_b;
    for (int l = 0; l < geo->nlevels; ++l) { g.sigma2A[l] = geo->sigma2_a[l]; g.sigma2B[l] = geo->sigma2_b[l]; g.scaleB[l] = geo->scale_b[l]; }
    CK(cudaMemcpyAsync(m->dStage, m->hStage, inBytes, cudaMemcpyHostToDevice, m->stream));
    CK(cudaMemsetAsync(m->dStage + oOut, 0xFF, oNeg - oOut, m->stream));
    CK(cudaMemsetAsync(m->dStage + oPts, 0, u - oPts, m->stream));
    CK(orbfe_launch_tri_kb8(m->stream, reinterpret_cast<const MatchProblem*>(m->dStage + oProb),
                            reinterpret_cast<const MatchItem*>(m->dStage + oItems), (int)items.size(), g, coarse, m->checkOri,
                            reinterpret_cast<float*>(m->dStage + oPts), reinterpret_cast<int*>(m->dStage + oCnt), &m->launches));
    CK(cudaMemcpyAsync(m->hStage + oOut, m->dStage + oOut, u - oOut, cudaMemcpyDeviceToHost, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    memcpy(matches12, m->hStage + oM, sizeof(int32_t) * A->n);
    if (points12) memcpy(points12, m->hStage + oPts, sizeof(float) * 3 * (size_t)A->n);
    return *reinterpret_cast<int*>(m->hStage + oCnt);
}

// ---- Frame::UndistortKeyPoints ---------------------------------------------------------------------------
extern "C" int orbfe_undistort_keypoints(orbfe_matcher* m, const orbfe_keypoint* kps, int n, const float* K4, const float* dist,
                                         int ndist, orbfe_keypoint* kps_un) {
    if (!m || n < 0 || !K4 || !dist || ndist < 4 || ndist > 5 || (n > 0 && (!kps || !kps_un))) return ORBFE_ERR_BAD_ARGUMENT;
    if (n == 0) return ORBFE_OK;
    if (dist[0] == 0.0f) {                                   // Frame.cc:732-736: no distortion, plain copy
        if (kps_un != kps) memmove(kps_un, kps, sizeof(orbfe_keypoint) * (size_t)n);
        return ORBFE_OK;
    }
    CK(cudaSetDevice(m->device));
    size_t u = 0;
    const size_t oK = stage_take(u, sizeof(orbfe_keypoint) * (size_t)n);
    int st = ensure_stage(m, u);
    if (st != ORBFE_OK) return st;
    CK(cudaStreamSynchronize(m->stream));
    memcpy(m->hStage + oK, kps, sizeof(orbfe_keypoint) * (size_t)n);
    CK(cudaMemcpyAsync(m->dStage + oK, m->hStage + oK, sizeof(orbfe_keypoint) * (size_t)n, cudaMemcpyHostToDevice, m->stream));
    UndistortGeo g;
    g.fx = K4[0]; g.fy = K4[1]; g.cx = K4[2]; g.cy = K4[3];
    g.k1 = dist[0]; g.k2 = dist[1]; g.p1 = dist[2]; g.p2 = dist[3]; g.k3 = ndist > 4 ? dist[4] : 0.0;
    CK(orbfe_launch_undistort(m->stream, reinterpret_cast<orbfe_keypoint*>(m->dStage + oK), n, g, &m->launches));
    CK(cudaMemcpyAsync(m->hStage + oK, m->dStage + oK, sizeof(orbfe_keypoint) * (size_t)n, cudaMemcpyDeviceToHost, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    memcpy(kps_un, m->hStage + oK, sizeof(orbfe_keypoint) * (size_t)n);
    return ORBFE_OK;
}

// ---- distinctive descriptors ----------------------------------------------------------------------------
extern "C" int orbfe_distinctive_descriptors(orbfe_matcher* m, const uint8_t* desc, const int32_t* offsets, int npoints,
                                             int32_t* best) {
    if (!m || npoints < 0 || (npoints > 0 && (!offsets || !best))) return ORBFE_ERR_BAD_ARGUMENT;
    if (npoints == 0) return ORBFE_OK;
    if (offsets[0] != 0) return ORBFE_ERR_BAD_ARGUMENT;
    for (int p = 0; p < npoints; ++p)
        if (offsets[p + 1] < offsets[p]) return ORBFE_ERR_BAD_ARGUMENT;
    const int total = offsets[npoints];
    if (total > 0 && !desc) return ORBFE_ERR_BAD_ARGUMENT;
    CK(cudaSetDevice(m->device));
    size_t u = 0;
    const size_t oD = stage_take(u, 32 * (size_t)std::max(total, 1)), oOff = stage_take(u, sizeof(int32_t) * ((size_t)npoints + 1));
    const size_t inBytes = u;
    const size_t oBest = stage_take(u, sizeof(int32_t) * (size_t)npoints);
    int st = ensure_stage(m, u);
    if (st != ORBFE_OK) return st;
    CK(cudaStreamSynchronize(m->stream));
    if (total > 0) memcpy(m->hStage + oD, desc, 32 * (size_t)total);
    memcpy(m->hStage + oOff, offsets, sizeof(int32_t) * ((size_t)npoints + 1));
    CK(cudaMemcpyAsync(m->dStage, m->hStage, inBytes, cudaMemcpyHostToDevice, m->stream));
    CK(orbfe_launch_distinctive(m->stream, m->dStage + oD, reinterpret_cast<const int32_t*>(m->dStage + oOff), npoints,
                                reinterpret_cast<int32_t*>(m->dStage + oBest), &m->launches));
    CK(cudaMemcpyAsync(m->hStage + oBest, m->dStage + oBest, sizeof(int32_t) * (size_t)npoints, cudaMemcpyDeviceToHost, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    memcpy(best, m->hStage + oBest, sizeof(int32_t) * (size_t)npoints);
    return ORBFE_OK;
}

// ---- projection-guided window searches ------------------------------------------------------------------
extern "C" int orbfe_search_by_projection(orbfe_matcher* m, const orbfe_grid_frame* F, const orbfe_proj_query* queries, int nq,
                                          int mode, int th, int link_stereo, int rot_hist, int32_t* assign, int32_t* choice) {
    static_assert(sizeof(orbfe_proj_query) == sizeof(ProjQueryDev), "orbfe_proj_query layout");
    if (!m || !F || nq < 0 || (nq > 0 && !queries) || !assign || mode < 0 || mode > 1 || F->n < 0) return ORBFE_ERR_BAD_ARGUMENT;
    if (F->n > 0 && (!F->kps || !F->desc)) return ORBFE_ERR_BAD_ARGUMENT;
    if (F->nleft < -1 || F->nleft > F->n) return ORBFE_ERR_BAD_ARGUMENT;
    if (F->n > 16384 || nq > (1 << 24)) return ORBFE_ERR_TOO_LARGE;      // grid keys: 16-bit index, sort in shared memory
    const int n = F->n;
    for (int i = 0; i < n; ++i) assign[i] = -1;
    if (choice) for (int i = 0; i < nq; ++i) choice[i] = -1;
    if (n == 0 || nq == 0) return 0;
    bool anyChi2 = false;
    for (int i = 0; i < nq && !anyChi2; ++i) anyChi2 = queries[i].chi2 != 0;
    if (anyChi2) {          // Fuse's gate indexes mvInvLevelSigma2 by the feature's octave
        if (!F->inv_level_sigma2 || F->n_levels < 1 || F->n_levels > ORBFE_MAX_LEVELS) return ORBFE_ERR_BAD_ARGUMENT;
        for (int i = 0; i < n; ++i) if (F->kps[i].octave < 0 || F->kps[i].octave >= F->n_levels) return ORBFE_ERR_BAD_ARGUMENT;
    }
    if (link_stereo && F->nleft != -1) {
        if (F->left_to_right) for (int i = 0; i < F->nleft; ++i) if (F->left_to_right[i] < -1 || F->left_to_right[i] >= n - F->nleft) return ORBFE_ERR_BAD_ARGUMENT;
        if (F->right_to_left) for (int i = 0; i < n - F->nleft; ++i) if (F->right_to_left[i] < -1 || F->right_to_left[i] >= F->nleft) return ORBFE_ERR_BAD_ARGUMENT;
    }
    CK(cudaSetDevice(m->device));
    const int nr = F->nleft == -1 ? 0 : n - F->nleft, nl = F->nleft == -1 ? 0 : F->nleft;
    size_t u = 0;
    const size_t oKp = stage_take(u, sizeof(float4) * n), oDesc = stage_take(u, 32 * (size_t)n), oBlk = stage_take(u, n);
    const size_t oUr = stage_take(u, sizeof(float) * n), oL2R = stage_take(u, sizeof(int32_t) * std::max(nl, 1));
    const size_t oR2L = stage_take(u, sizeof(int32_t) * std::max(nr, 1)), oQ = stage_take(u, sizeof(ProjQueryDev) * (size_t)nq);
    const size_t oSig = stage_take(u, sizeof(float) * ORBFE_MAX_LEVELS);
    const size_t oProb = stage_take(u, sizeof(ProjProblem));
    const size_t inBytes = u;
    const size_t oSorted = stage_take(u, sizeof(uint32_t) * n), oCell = stage_take(u, sizeof(int) * 4 * 3072);
    const size_t oCand = stage_take(u, sizeof(unsigned long long) * (size_t)nq * PROJ_K), oLim = stage_take(u, sizeof(unsigned long long) * (size_t)nq);
    const size_t oScr = stage_take(u, sizeof(int) * ((size_t)n + 5 * (size_t)nq));
    const size_t oOut = stage_take(u, 0);
    const size_t oAsg = stage_take(u, sizeof(int32_t) * n), oCh = stage_take(u, sizeof(int32_t) * (size_t)nq), oCnt = stage_take(u, 4);
    int st = ensure_stage(m, u);
    if (st != ORBFE_OK) return st;
    CK(cudaStreamSynchronize(m->stream));
    uint8_t* hs = m->hStage;
    uint8_t* ds = m->dStage;
    float4* kp = reinterpret_cast<float4*>(hs + oKp);
    for (int i = 0; i < n; ++i) {
        float4 v = make_float4(F->kps[i].x, F->kps[i].y, F->kps[i].angle, 0.f);
        memcpy(&v.w, &F->kps[i].octave, 4);
        kp[i] = v;
    }
    memcpy(hs + oDesc, F->desc, 32 * (size_t)n);
    if (F->blocked) memcpy(hs + oBlk, F->blocked, n);
    if (F->u_right) memcpy(hs + oUr, F->u_right, sizeof(float) * n);
    if (F->left_to_right && nl) memcpy(hs + oL2R, F->left_to_right, sizeof(int32_t) * nl);
    if (F->right_to_left && nr) memcpy(hs + oR2L, F->right_to_left, sizeof(int32_t) * nr);
    memcpy(hs + oQ, queries, sizeof(ProjQueryDev) * (size_t)nq);
    if (anyChi2) memcpy(hs + oSig, F->inv_level_sigma2, sizeof(float) * F->n_levels);
    ProjProblem P;
    memset(&P, 0, sizeof(P));
    P.kp = reinterpret_cast<const float4*>(ds + oKp);
    P.desc = ds + oDesc;
    P.blocked = F->blocked ? ds + oBlk : nullptr;
    P.uRight = F->u_right ? reinterpret_cast<const float*>(ds + oUr) : nullptr;
    P.l2r = (F->left_to_right && nl) ? reinterpret_cast<const int32_t*>(ds + oL2R) : nullptr;
    P.r2l = (F->right_to_left && nr) ? reinterpret_cast<const int32_t*>(ds + oR2L) : nullptr;
    P.invSigma2 = anyChi2 ? reinterpret_cast<const float*>(ds + oSig) : nullptr;
    P.n = n; P.nleft = F->nleft;
    P.minX = F->min_x; P.minY = F->min_y; P.gridWInv = F->grid_w_inv; P.gridHInv = F->grid_h_inv;
    P.q = reinterpret_cast<const ProjQueryDev*>(ds + oQ);
    P.nq = nq;
    P.assign = reinterpret_cast<int32_t*>(ds + oAsg);
    P.choice = reinterpret_cast<int32_t*>(ds + oCh);
    P.nmatches = reinterpret_cast<int*>(ds + oCnt);
    P.sorted = reinterpret_cast<uint32_t*>(ds + oSorted);
    P.cellStart = reinterpret_cast<int*>(ds + oCell);
    P.cellEnd = P.cellStart + 2 * 3072;
    P.cand = reinterpret_cast<unsigned long long*>(ds + oCand);
    P.limit = reinterpret_cast<unsigned long long*>(ds + oLim);
    P.scratch = reinterpret_cast<int*>(ds + oScr);
    memcpy(hs + oProb, &P, sizeof(P));
    CK(cudaMemcpyAsync(ds, hs, inBytes, cudaMemcpyHostToDevice, m->stream));
    CK(orbfe_launch_search_by_projection(m->stream, reinterpret_cast<const ProjProblem*>(ds + oProb), 1, n, nq, mode, th, m->nnRatio,
                                         link_stereo ? 1 : 0, (rot_hist && m->checkOri) ? 1 : 0, &m->launches));
    CK(cudaMemcpyAsync(hs + oOut, ds + oOut, u - oOut, cudaMemcpyDeviceToHost, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    memcpy(assign, hs + oAsg, sizeof(int32_t) * n);
    if (choice) memcpy(choice, hs + oCh, sizeof(int32_t) * (size_t)nq);
    return *reinterpret_cast<int*>(hs + oCnt);
}

// ORBmatcher::SearchForInitialization (ORBmatcher.cc:706-821): one level-0 window per F1 feature around vbPrevMatched.
extern "C" int orbfe_search_for_initialization(orbfe_matcher* m, const orbfe_keypoint* kps1, const uint8_t* desc1, int n1,
                                               const orbfe_grid_frame* F2, float* prev_matched, int window_size, int32_t* matches12) {
    if (!m || !F2 || n1 < 0 || F2->n < 0 || window_size < 0) return ORBFE_ERR_BAD_ARGUMENT;
    if (n1 > 0 && (!kps1 || !desc1 || !prev_matched || !matches12)) return ORBFE_ERR_BAD_ARGUMENT;
    if (F2->n > 0 && (!F2->kps || !F2->desc)) return ORBFE_ERR_BAD_ARGUMENT;
    if (F2->n > 16384 || n1 > (1 << 24)) return ORBFE_ERR_TOO_LARGE;
    const int n = F2->n, nq = n1;
    for (int i = 0; i < n1; ++i) matches12[i] = -1;
    if (n == 0 || nq == 0) return 0;
    CK(cudaSetDevice(m->device));
    size_t u = 0;
    const size_t oKp = stage_take(u, sizeof(float4) * n), oDesc = stage_take(u, 32 * (size_t)n);
    const size_t oQ = stage_take(u, sizeof(ProjQueryDev) * (size_t)nq), oProb = stage_take(u, sizeof(ProjProblem));
    const size_t inBytes = u;
    const size_t oSorted = stage_take(u, sizeof(uint32_t) * n), oCell = stage_take(u, sizeof(int) * 4 * 3072);
    const size_t oCand = stage_take(u, sizeof(unsigned long long) * (size_t)nq * PROJ_K), oLim = stage_take(u, sizeof(unsigned long long) * (size_t)nq);
    const size_t oScr = stage_take(u, sizeof(int) * ((size_t)n + 5 * (size_t)nq));
    const size_t oAsg = stage_take(u, sizeof(int32_t) * (size_t)std::max(n, nq));      // claim chains of k_init_resolve
    const size_t oOut = stage_take(u, 0);
    const size_t oCh = stage_take(u, sizeof(int32_t) * (size_t)nq), oCnt = stage_take(u, 4);
    int st = ensure_stage(m, u);
    if (st != ORBFE_OK) return st;
    CK(cudaStreamSynchronize(m->stream));
    uint8_t* hs = m->hStage;
    uint8_t* ds = m->dStage;
    float4* kp = reinterpret_cast<float4*>(hs + oKp);
    for (int i = 0; i < n; ++i) {
        float4 v = make_float4(F2->kps[i].x, F2->kps[i].y, F2->kps[i].angle, 0.f);
        memcpy(&v.w, &F2->kps[i].octave, 4);
        kp[i] = v;
    }
    memcpy(hs + oDesc, F2->desc, 32 * (size_t)n);
    ProjQueryDev* q = reinterpret_cast<ProjQueryDev*>(hs + oQ);
    memset(q, 0, sizeof(ProjQueryDev) * (size_t)nq);
    for (int i = 0; i < nq; ++i) {
        ProjQueryDev& e = q[i];
        e.u = prev_matched[2 * i];
        e.v = prev_matched[2 * i + 1];
        e.r = (float)window_size;
        e.angle = kps1[i].angle;
        e.minLevel = 0;                               // GetFeaturesInArea(x, y, windowSize, level1, level1), level1 == 0 (:721-727)
        e.maxLevel = 0;
        e.valid = kps1[i].octave > 0 ? 0 : 1;
        memcpy(e.desc, desc1 + 32 * (size_t)i, 32);
    }
    ProjProblem P;
    memset(&P, 0, sizeof(P));
    P.kp = reinterpret_cast<const float4*>(ds + oKp);
    P.desc = ds + oDesc;
    P.n = n; P.nleft = -1;
    P.minX = F2->min_x; P.minY = F2->min_y; P.gridWInv = F2->grid_w_inv; P.gridHInv = F2->grid_h_inv;
    P.q = reinterpret_cast<const ProjQueryDev*>(ds + oQ);
    P.nq = nq;
    P.assign = reinterpret_cast<int32_t*>(ds + oAsg);
    P.choice = reinterpret_cast<int32_t*>(ds + oCh);
    P.nmatches = reinterpret_cast<int*>(ds + oCnt);
    P.sorted = reinterpret_cast<uint32_t*>(ds + oSorted);
    P.cellStart = reinterpret_cast<int*>(ds + oCell);
    P.cellEnd = P.cellStart + 2 * 3072;
    P.cand = reinterpret_cast<unsigned long long*>(ds + oCand);
    P.limit = reinterpret_cast<unsigned long long*>(ds + oLim);
    P.scratch = reinterpret_cast<int*>(ds + oScr);
    memcpy(hs + oProb, &P, sizeof(P));
    CK(cudaMemcpyAsync(ds, hs, inBytes, cudaMemcpyHostToDevice, m->stream));
    CK(orbfe_launch_search_for_initialization(m->stream, reinterpret_cast<const ProjProblem*>(ds + oProb), 1, n, nq, ORBFE_TH_LOW,
                                              m->nnRatio, m->checkOri ? 1 : 0, &m->launches));
    CK(cudaMemcpyAsync(hs + oOut, ds + oOut, u - oOut, cudaMemcpyDeviceToHost, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    memcpy(matches12, hs + oCh, sizeof(int32_t) * (size_t)nq);
    for (int i = 0; i < nq; ++i)                      // :815-817
        if (matches12[i] >= 0) { prev_matched[2 * i] = F2->kps[matches12[i]].x; prev_matched[2 * i + 1] = F2->kps[matches12[i]].y; }
    return *reinterpret_cast<int*>(hs + oCnt);
}

// ---- Kannala-Brandt fisheye stereo ----------------------------------------------------------------------
extern "C" int orbfe_fisheye_stereo_match(orbfe_matcher* m, const orbfe_keypoint* kps_left, const uint8_t* desc_left, int n_left,
                                          int mono_left, const orbfe_keypoint* kps_right, const uint8_t* desc_right, int n_right,
                                          int mono_right, const float* cam1, const float* cam2, const float* R12, const float* t12,
                                          const float* sigma2, int nlevels, int32_t* left_to_right, int32_t* right_to_left,
                                          float* depth, float* p3d) {
    if (!m || n_left < 0 || n_right < 0 || mono_left < 0 || mono_right < 0 || mono_left > n_left || mono_right > n_right)
        return ORBFE_ERR_BAD_ARGUMENT;
    if (!cam1 || !cam2 || !R12 || !t12 || !sigma2 || nlevels < 1 || nlevels > ORBFE_MAX_LEVELS) return ORBFE_ERR_BAD_ARGUMENT;
    if ((n_left > 0 && (!kps_left || !desc_left || !left_to_right || !depth || !p3d)) || (n_right > 0 && (!kps_right || !desc_right || !right_to_left)))
        return ORBFE_ERR_BAD_ARGUMENT;
    for (int i = 0; i < n_left; ++i)
        if (kps_left[i].octave < 0 || kps_left[i].octave >= nlevels) return ORBFE_ERR_BAD_ARGUMENT;
    for (int i = 0; i < n_right; ++i)
        if (kps_right[i].octave < 0 || kps_right[i].octave >= nlevels) return ORBFE_ERR_BAD_ARGUMENT;
    const int nq = n_left - mono_left, nt = n_right - mono_right;
    for (int i = 0; i < n_left; ++i) { left_to_right[i] = -1; depth[i] = -1.f; p3d[3 * i] = p3d[3 * i + 1] = p3d[3 * i + 2] = 0.f; }
    for (int i = 0; i < n_right; ++i) right_to_left[i] = -1;
    if (nq <= 0) return 0;
    CK(cudaSetDevice(m->device));
    // inputs
    size_t u = 0;
    const size_t oKL = stage_take(u, sizeof(orbfe_keypoint) * n_left), oKR = stage_take(u, sizeof(orbfe_keypoint) * std::max(n_right, 1));
    const size_t oQ = stage_take(u, (size_t)nq * 32), oT = stage_take(u, (size_t)std::max(nt, 1) * 32);
    const size_t inBytes = u;
    // device-only intermediates
    const size_t oIdx = stage_take(u, sizeof(int32_t) * 2 * nq), oDist = stage_take(u, sizeof(int32_t) * 2 * nq);
    const size_t oMatch = stage_take(u, sizeof(int32_t) * nq), oTmp = stage_take(u, 256);
    // outputs (one D2H)
    const size_t oOut = stage_take(u, 0);
    const size_t oL2R = stage_take(u, sizeof(int32_t) * n_left), oR2L = stage_take(u, sizeof(int32_t) * std::max(n_right, 1));
    const size_t oCnt = stage_take(u, 256);
    const size_t oNeg = u;                                   // everything from oOut to here is initialised to -1 / 0xFF
    const size_t oDepth = stage_take(u, sizeof(float) * n_left), oP3d = stage_take(u, sizeof(float) * 3 * n_left);
    int st = ensure_stage(m, u);
    if (st != ORBFE_OK) return st;
    CK(cudaStreamSynchronize(m->stream));
    memcpy(m->hStage + oKL, kps_left, sizeof(orbfe_keypoint) * n_left);
    if (n_right > 0) memcpy(m->hStage + oKR, kps_right, sizeof(orbfe_keypoint) * n_right);
    memcpy(m->hStage + oQ, desc_left + (size_t)mono_left * 32, (size_t)nq * 32);
    if (nt > 0) memcpy(m->hStage + oT, desc_right + (size_t)mono_right * 32, (size_t)nt * 32);
    CK(cudaMemcpyAsync(m->dStage, m->hStage, inBytes, cudaMemcpyHostToDevice, m->stream));
    int32_t* dIdx = reinterpret_cast<int32_t*>(m->dStage + oIdx);
    int32_t* dDist = reinterpret_cast<int32_t*>(m->dStage + oDist);
    st = orbfe_hamming_knn2_batch_device(m, m->dStage + oQ, 0, nullptr, nq, m->dStage + oT, 0, nullptr, nt, 1, dIdx, dDist);
    if (st != ORBFE_OK) return st;
    CK(cudaMemsetAsync(m->dStage + oTmp, 0, sizeof(int), m->stream));
    CK(orbfe_launch_ratio_filter(m->stream, dIdx, dDist, nq, 0.7, reinterpret_cast<int32_t*>(m->dStage + oMatch),
                                 reinterpret_cast<int*>(m->dStage + oTmp), &m->launches));
    CK(cudaMemsetAsync(m->dStage + oOut, 0xFF, oNeg - oOut, m->stream));
    CK(cudaMemsetAsync(m->dStage + oCnt, 0, sizeof(int), m->stream));
    CK(cudaMemsetAsync(m->dStage + oDepth, 0, u - oDepth, m->stream));
    Kb8Geo G;
    memset(&G, 0, sizeof(G));
    memcpy(G.cam1, cam1, sizeof(G.cam1));
    memcpy(G.cam2, cam2, sizeof(G.cam2));
    memcpy(G.R12, R12, sizeof(G.R12));
    memcpy(G.t12, t12, sizeof(G.t12));
    for (int l = 0; l < nlevels; ++l) G.sigma2[l] = sigma2[l];
    CK(orbfe_launch_fisheye_triangulate(m->stream, G, reinterpret_cast<const orbfe_keypoint*>(m->dStage + oKL),
                                        reinterpret_cast<const orbfe_keypoint*>(m->dStage + oKR), mono_left, mono_right, nq,
                                        reinterpret_cast<const int32_t*>(m->dStage + oMatch), reinterpret_cast<int32_t*>(m->dStage + oL2R),
                                        reinterpret_cast<int32_t*>(m->dStage + oR2L), reinterpret_cast<float*>(m->dStage + oDepth),
                                        reinterpret_cast<float*>(m->dStage + oP3d), reinterpret_cast<int*>(m->dStage + oCnt),
                                        &m->launches));
    CK(cudaMemcpyAsync(m->hStage + oOut, m->dStage + oOut, u - oOut, cudaMemcpyDeviceToHost, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    const int32_t* hl = reinterpret_cast<const int32_t*>(m->hStage + oL2R);
    const float* hd = reinterpret_cast<const float*>(m->hStage + oDepth);
    const float* hp = reinterpret_cast<const float*>(m->hStage + oP3d);
    for (int i = 0; i < n_left; ++i) {
        left_to_right[i] = hl[i];
        if (hl[i] >= 0) { depth[i] = hd[i]; p3d[3 * i] = hp[3 * i]; p3d[3 * i + 1] = hp[3 * i + 1]; p3d[3 * i + 2] = hp[3 * i + 2]; }
    }
    if (n_right > 0) memcpy(right_to_left, m->hStage + oR2L, sizeof(int32_t) * n_right);
    return *reinterpret_cast<int*>(m->hStage + oCnt);
}
