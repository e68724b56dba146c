// orbfe_voc.h — device view of a DBoW2 vocabulary tree and the launchers of the BoW transform kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

struct VocDev {
    const uint4* desc;         // 2 x uint4 per node (256-bit descriptor), node 0 = root
    const int2* child;         // (first slot in childIdx, number of children) per node
    const int32_t* childIdx;   // children in file order (TemplatedVocabulary.h:1405 push_back order)
    const int32_t* wordId;     // per node, -1 for inner nodes
    const double* weight;      // per node (idf of the word for leaves)
    int L, weighting, norm;
    int contiguous;            // every node's children are consecutive node ids: child j = childIdx[first] + j
};

// K1: one thread per (frame, feature): word / node / weight of the leaf the descriptor descends to.
// word = 0xFFFFFFFF marks a stop word (weight <= 0).  Per-frame stride of the outputs = nMax.
cudaError_t orbfe_launch_bow_descend(cudaStream_t st, const VocDev& v, const uint8_t* dDesc, long long frameStride, const int* dN,
                                     int nHost, int nMax, int batch, int levelsup, uint32_t* dWord, uint32_t* dNode, double* dW,
                                     long long* launches);
// K2: one CTA per frame: BowVector (ascending words, summed + normalised values) and FeatureVector CSR.
// Outputs per frame at strides nMax (words, values, nodeIds, indices) and nMax + 1 (offsets).
cudaError_t orbfe_launch_bow_assemble(cudaStream_t st, const VocDev& v, const int* dN, int nHost, int nMax, int batch,
                                      const uint32_t* dWord, const uint32_t* dNode, const double* dW, uint32_t* dWords,
                                      double* dValues, int* dNWords, uint32_t* dNodeIds, int32_t* dOffsets, int32_t* dIndices,
                                      int* dNNodes, long long* launches);
size_t orbfe_bow_assemble_smem(int nMax);
