// k_peaks.cu — measures the integer-pipe ceiling the Hamming matchers are bound by: sustained
// 32-bit POPC issue rate of the whole GPU (SURVEY.md §8d asks for the measured figure, not the
// nominal 16/clk/SM, as the denominator of the matching roofline).  Pure register arithmetic, no
// memory traffic: 8 independent xor+popc chains per thread so the pipe, not latency, is the limit.
#include "orbfe_internal.h"

__global__ void __launch_bounds__(256) k_popc_peak(uint32_t* __restrict__ sink, int iters, uint32_t seed) {
    uint32_t a0 = threadIdx.x ^ seed, a1 = a0 * 3u + 1u, a2 = a0 * 5u + 2u, a3 = a0 * 7u + 3u, a4 = a0 * 11u + 4u, a5 = a0 * 13u + 5u,
             a6 = a0 * 17u + 6u, a7 = a0 * 19u + 7u;
    const uint32_t k = 0x9E3779B9u * (blockIdx.x + 1u);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            a0 = __popc(a0 ^ k) + a1; a1 = __popc(a1 ^ k) + a2; a2 = __popc(a2 ^ k) + a3; a3 = __popc(a3 ^ k) + a4;
            a4 = __popc(a4 ^ k) + a5; a5 = __popc(a5 ^ k) + a6; a6 = __popc(a6 ^ k) + a7; a7 = __popc(a7 ^ k) + a0;
        }
    }
    sink[blockIdx.x * blockDim.x + threadIdx.x] = a0 ^ a1 ^ a2 ^ a3 ^ a4 ^ a5 ^ a6 ^ a7;
}

// Returns 0 and the sustained rate in POPC instructions (thread-level) per second, or a negative status.
extern "C" int orbfe_measure_popc_peak(int device, double* popc_per_s) {
    if (!popc_per_s) return ORBFE_ERR_BAD_ARGUMENT;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) { cudaGetLastError(); return ORBFE_ERR_NO_DEVICE; }
    if (device < 0 || device >= ndev) return ORBFE_ERR_BAD_ARGUMENT;
    if (cudaSetDevice(device) != cudaSuccess) return ORBFE_ERR_CUDA;
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const int blocks = sms * 8, threads = 256, iters = 4096;
    uint32_t* sink = nullptr;
    if (cudaMalloc(&sink, sizeof(uint32_t) * blocks * threads) != cudaSuccess) return ORBFE_ERR_CUDA;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0, 0);
        k_popc_peak<<<blocks, threads>>>(sink, iters, 1234u + rep);
        cudaEventRecord(e1, 0);
        if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(sink); return ORBFE_ERR_CUDA; }
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        const double rate = (double)blocks * threads * iters * 32.0 / (ms * 1e-3);
        if (rep > 0 && rate > best) best = rate;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    *popc_per_s = best;
    return ORBFE_OK;
}
