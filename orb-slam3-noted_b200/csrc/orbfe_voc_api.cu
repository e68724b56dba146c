// orbfe_voc_api.cu — C ABI of the vocabulary tree (include/orbfe.h, "BoW transform" section): upload of a
// flattened DBoW2 TemplatedVocabulary, the ORBvoc text-format loader (loadFromTextFile,
// TemplatedVocabulary.h:1350-1428) and the transform entry points.  No CPU fallback.
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "orbfe_internal.h"
#include "orbfe_voc.h"

int orbfe_set_cuda_error(cudaError_t e, const char* what);   // orbfe_api.cu

#define CK(call)                                                         \
    do {                                                                 \
        cudaError_t e_ = (call);                                         \
        if (e_ != cudaSuccess) return orbfe_set_cuda_error(e_, #call);   \
    } while (0)

struct orbfe_vocabulary {
    int k, L, weighting, norm, nnodes, nwords, device;
    VocDev dev;
    void* dArena;
    cudaStream_t stream;
    long long launches;
    // staging for the host-buffer entry point
    uint8_t* hStage;
    uint8_t* dStage;
    size_t stageCap;
    // per-feature intermediates of the batch entry point
    uint8_t* dTmp;
    size_t tmpCap;
};

static inline size_t take(size_t& used, size_t bytes) {
    const size_t off = (used + 255) & ~(size_t)255;
    used = off + bytes;
    return off;
}

extern "C" void orbfe_vocabulary_destroy(orbfe_vocabulary* v) {
    if (!v) return;
    cudaSetDevice(v->device);
    if (v->stream) cudaStreamSynchronize(v->stream);
    if (v->dArena) cudaFree(v->dArena);
    if (v->dStage) cudaFree(v->dStage);
    if (v->hStage) cudaFreeHost(v->hStage);
    if (v->dTmp) cudaFree(v->dTmp);
    if (v->stream) cudaStreamDestroy(v->stream);
    delete v;
}

extern "C" int orbfe_vocabulary_create(int k, int L, int weighting, int norm, int nnodes, const int32_t* parent,
                                       const uint8_t* is_leaf, const uint8_t* descriptors, const double* weights, int device,
                                       orbfe_vocabulary** out) {
    if (!out) return ORBFE_ERR_BAD_ARGUMENT;
    *out = nullptr;
    if (k < 1 || L < 1 || nnodes < 2 || !parent || !is_leaf || !descriptors || !weights || weighting < 0 || weighting > 3 ||
        norm < 0 || norm > 2)
        return ORBFE_ERR_BAD_ARGUMENT;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) { cudaGetLastError(); return ORBFE_ERR_NO_DEVICE; }
    if (device < 0 || device >= ndev) return ORBFE_ERR_BAD_ARGUMENT;
    // children lists in file order, word ids in file order of the leaves
    std::vector<int2> child(nnodes, make_int2(0, 0));
    for (int i = 1; i < nnodes; ++i) {
        if (parent[i] < 0 || parent[i] >= i) return ORBFE_ERR_BAD_ARGUMENT;
        child[parent[i]].y++;
    }
    int tot = 0;
    for (int i = 0; i < nnodes; ++i) { child[i].x = tot; tot += child[i].y; }
    std::vector<int32_t> childIdx(std::max(tot, 1)), fill(nnodes), wordId(nnodes, -1);
    for (int i = 0; i < nnodes; ++i) fill[i] = child[i].x;
    for (int i = 1; i < nnodes; ++i) childIdx[fill[parent[i]]++] = i;
    int nwords = 0, contiguous = 1;
    for (int i = 1; i < nnodes; ++i) {
        if ((is_leaf[i] != 0) != (child[i].y == 0)) return ORBFE_ERR_BAD_ARGUMENT;
        if (is_leaf[i]) wordId[i] = nwords++;
    }
    if (child[0].y == 0) return ORBFE_ERR_BAD_ARGUMENT;
    for (int i = 0; i < nnodes && contiguous; ++i)
        for (int j = 1; j < child[i].y; ++j)
            if (childIdx[child[i].x + j] != childIdx[child[i].x] + j) { contiguous = 0; break; }

    orbfe_vocabulary* v = new orbfe_vocabulary();
    memset(v, 0, sizeof(*v));
    v->k = k; v->L = L; v->weighting = weighting; v->norm = norm; v->nnodes = nnodes; v->nwords = nwords; v->device = device;
    size_t u = 0;
    const size_t oDesc = take(u, 32 * (size_t)nnodes), oChild = take(u, sizeof(int2) * (size_t)nnodes);
    const size_t oIdx = take(u, sizeof(int32_t) * childIdx.size()), oWord = take(u, sizeof(int32_t) * (size_t)nnodes);
    const size_t oW = take(u, sizeof(double) * (size_t)nnodes);
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&v->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc(&v->dArena, u);
    uint8_t* a = static_cast<uint8_t*>(v->dArena);
    if (e == cudaSuccess) e = cudaMemcpy(a + oDesc, descriptors, 32 * (size_t)nnodes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(a + oChild, child.data(), sizeof(int2) * (size_t)nnodes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(a + oIdx, childIdx.data(), sizeof(int32_t) * childIdx.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(a + oWord, wordId.data(), sizeof(int32_t) * (size_t)nnodes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(a + oW, weights, sizeof(double) * (size_t)nnodes, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { orbfe_set_cuda_error(e, "orbfe_vocabulary_create"); orbfe_vocabulary_destroy(v); return ORBFE_ERR_CUDA; }
    v->dev.desc = reinterpret_cast<const uint4*>(a + oDesc);
    v->dev.child = reinterpret_cast<const int2*>(a + oChild);
    v->dev.childIdx = reinterpret_cast<const int32_t*>(a + oIdx);
    v->dev.wordId = reinterpret_cast<const int32_t*>(a + oWord);
    v->dev.weight = reinterpret_cast<const double*>(a + oW);
    v->dev.L = L; v->dev.weighting = weighting; v->dev.norm = norm; v->dev.contiguous = contiguous;
    *out = v;
    return ORBFE_OK;
}

// loadFromTextFile (TemplatedVocabulary.h:1350-1428): "k L scoring weighting", then one node per line:
// "parent isLeaf d0 .. d31 weight".  Empty lines are skipped (the reference's eof loop would turn a trailing
// one into a spurious node).  The caller names the normalisation its scoring type implies (L1 for ORBvoc).
extern "C" int orbfe_vocabulary_load_text(const char* path, int norm, int device, orbfe_vocabulary** out) {
    if (!path || !out) return ORBFE_ERR_BAD_ARGUMENT;
    *out = nullptr;
    std::ifstream f(path);
    if (!f.is_open()) return ORBFE_ERR_BAD_ARGUMENT;
    std::string line;
    if (!std::getline(f, line)) return ORBFE_ERR_BAD_ARGUMENT;
    int k = 0, L = 0, n1 = 0, n2 = 0;
    {
        std::stringstream ss(line);
        ss >> k >> L >> n1 >> n2;
        if (ss.fail() || k < 0 || k > 20 || L < 1 || L > 10 || n1 < 0 || n1 > 5 || n2 < 0 || n2 > 3) return ORBFE_ERR_BAD_ARGUMENT;
    }
    std::vector<int32_t> parent(1, 0);
    std::vector<uint8_t> leaf(1, 0), desc(32, 0);
    std::vector<double> weight(1, 0.0);
    while (std::getline(f, line)) {
        std::stringstream ss(line);
        int pid, isLeaf;
        ss >> pid;
        if (ss.fail()) continue;
        ss >> isLeaf;
        uint8_t d[32];
        for (int i = 0; i < 32; ++i) { int x = 0; ss >> x; d[i] = (uint8_t)x; }
        double w = 0.0;
        ss >> w;
        if (ss.fail()) return ORBFE_ERR_BAD_ARGUMENT;
        parent.push_back(pid); leaf.push_back(isLeaf > 0); weight.push_back(w);
        desc.insert(desc.end(), d, d + 32);
    }
    return orbfe_vocabulary_create(k, L, n2, norm, (int)parent.size(), parent.data(), leaf.data(), desc.data(), weight.data(), device, out);
}

extern "C" int orbfe_vocabulary_words(const orbfe_vocabulary* v) { return v ? v->nwords : 0; }
extern "C" int orbfe_vocabulary_nodes(const orbfe_vocabulary* v) { return v ? v->nnodes : 0; }
extern "C" void* orbfe_vocabulary_stream(orbfe_vocabulary* v) { return v ? (void*)v->stream : nullptr; }
extern "C" long long orbfe_vocabulary_launch_count(const orbfe_vocabulary* v) { return v ? v->launches : 0; }
extern "C" int orbfe_vocabulary_synchronize(orbfe_vocabulary* v) {
    if (!v) return ORBFE_ERR_BAD_ARGUMENT;
    CK(cudaSetDevice(v->device));
    CK(cudaStreamSynchronize(v->stream));
    return ORBFE_OK;
}

static int ensure_tmp(orbfe_vocabulary* v, size_t bytes) {
    if (bytes <= v->tmpCap) return ORBFE_OK;
    CK(cudaStreamSynchronize(v->stream));
    if (v->dTmp) cudaFree(v->dTmp);
    v->dTmp = nullptr; v->tmpCap = 0;
    CK(cudaMalloc(&v->dTmp, bytes + bytes / 4));
    v->tmpCap = bytes + bytes / 4;
    return ORBFE_OK;
}

extern "C" int orbfe_bow_transform_batch_device(orbfe_vocabulary* v, const uint8_t* d_desc, ptrdiff_t frame_stride, const int* d_n,
                                                int n_max, int batch, int levelsup, uint32_t* d_words, double* d_values,
                                                int* d_nwords, uint32_t* d_node_ids, int32_t* d_offsets, int32_t* d_indices,
                                                int* d_nnodes) {
    if (!v || !d_desc || !d_n || n_max < 0 || batch < 0 || !d_words || !d_values || !d_nwords || !d_node_ids || !d_offsets ||
        !d_indices || !d_nnodes)
        return ORBFE_ERR_BAD_ARGUMENT;
    if (n_max > 65536) return ORBFE_ERR_TOO_LARGE;
    if (batch == 0) return ORBFE_OK;
    CK(cudaSetDevice(v->device));
    const size_t per = (size_t)batch * std::max(n_max, 1);
    size_t u = 0;
    const size_t oWord = take(u, 4 * per), oNode = take(u, 4 * per), oW = take(u, 8 * per);
    int st = ensure_tmp(v, u);
    if (st != ORBFE_OK) return st;
    uint32_t* dWord = reinterpret_cast<uint32_t*>(v->dTmp + oWord);
    uint32_t* dNode = reinterpret_cast<uint32_t*>(v->dTmp + oNode);
    double* dW = reinterpret_cast<double*>(v->dTmp + oW);
    CK(orbfe_launch_bow_descend(v->stream, v->dev, d_desc, (long long)frame_stride, d_n, 0, n_max, batch, levelsup, dWord, dNode, dW,
                                &v->launches));
    CK(orbfe_launch_bow_assemble(v->stream, v->dev, d_n, 0, n_max, batch, dWord, dNode, dW, d_words, d_values, d_nwords, d_node_ids,
                                 d_offsets, d_indices, d_nnodes, &v->launches));
    return ORBFE_OK;
}

extern "C" int orbfe_bow_transform(orbfe_vocabulary* v, const uint8_t* desc, int n, int levelsup, uint32_t* words, double* values,
                                   int* nwords, uint32_t* node_ids, int32_t* offsets, int32_t* indices, int* nnodes) {
    if (!v || n < 0 || !nwords || !nnodes || !offsets || (n > 0 && (!desc || !words || !values || !node_ids || !indices)))
        return ORBFE_ERR_BAD_ARGUMENT;
    if (n > 65536) return ORBFE_ERR_TOO_LARGE;
    *nwords = 0; *nnodes = 0; offsets[0] = 0;
    if (n == 0) return ORBFE_OK;
    CK(cudaSetDevice(v->device));
    size_t u = 0;
    const size_t oDesc = take(u, 32 * (size_t)n), oN = take(u, 4);
    const size_t inBytes = u;
    const size_t oOut = take(u, 0);
    const size_t oWords = take(u, 4 * (size_t)n), oVal = take(u, 8 * (size_t)n), oNodes = take(u, 4 * (size_t)n);
    const size_t oOff = take(u, 4 * ((size_t)n + 1)), oIdx = take(u, 4 * (size_t)n), oCnt = take(u, 8);
    if (u > v->stageCap) {
        CK(cudaStreamSynchronize(v->stream));
        if (v->hStage) cudaFreeHost(v->hStage);
        if (v->dStage) cudaFree(v->dStage);
        v->hStage = nullptr; v->dStage = nullptr; v->stageCap = 0;
        const size_t cap = std::max(u + u / 4, (size_t)1 << 20);
        CK(cudaMallocHost(&v->hStage, cap));
        CK(cudaMalloc(&v->dStage, cap));
        v->stageCap = cap;
    }
    CK(cudaStreamSynchronize(v->stream));
    memcpy(v->hStage + oDesc, desc, 32 * (size_t)n);
    memcpy(v->hStage + oN, &n, 4);
    CK(cudaMemcpyAsync(v->dStage, v->hStage, inBytes, cudaMemcpyHostToDevice, v->stream));
    uint8_t* d = v->dStage;
    int st = orbfe_bow_transform_batch_device(v, d + oDesc, 0, reinterpret_cast<const int*>(d + oN), n, 1, levelsup,
                                              reinterpret_cast<uint32_t*>(d + oWords), reinterpret_cast<double*>(d + oVal),
                                              reinterpret_cast<int*>(d + oCnt), reinterpret_cast<uint32_t*>(d + oNodes),
                                              reinterpret_cast<int32_t*>(d + oOff), reinterpret_cast<int32_t*>(d + oIdx),
                                              reinterpret_cast<int*>(d + oCnt) + 1);
    if (st != ORBFE_OK) return st;
    CK(cudaMemcpyAsync(v->hStage + oOut, d + oOut, u - oOut, cudaMemcpyDeviceToHost, v->stream));
    CK(cudaStreamSynchronize(v->stream));
    const int* cnt = reinterpret_cast<const int*>(v->hStage + oCnt);
    const int nw = cnt[0], nn = cnt[1];
    *nwords = nw; *nnodes = nn;
    memcpy(words, v->hStage + oWords, 4 * (size_t)nw);
    memcpy(values, v->hStage + oVal, 8 * (size_t)nw);
    memcpy(node_ids, v->hStage + oNodes, 4 * (size_t)nn);
    memcpy(offsets, v->hStage + oOff, 4 * ((size_t)nn + 1));
    const int kept = reinterpret_cast<const int32_t*>(v->hStage + oOff)[nn];
    memcpy(indices, v->hStage + oIdx, 4 * (size_t)kept);
    return ORBFE_OK;
}
