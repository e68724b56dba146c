// tma_probe.cu — stand-alone check of the TMA tiled load used by k_fast_cells (debug tool).
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
struct Maps { CUtensorMap m[16]; };
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void probe(const __grid_constant__ Maps maps, const int* lvl, uint8_t* out, int boxW, int boxH, int x, int y, int z, int mode) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    const int level = mode ? lvl[blockIdx.x] : 0;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(boxW * boxH) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(smem_u32(smem)), "l"(&maps.m[level]), "r"(x), "r"(y), "r"(z), "r"(smem_u32(&bar)) : "memory");
    }
    __syncthreads();
    asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(smem_u32(&bar)), "r"(0) : "memory");
    for (int i = threadIdx.x; i < boxW * boxH; i += blockDim.x) out[i] = smem[i];
}
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char** argv) {
    const int ax = argc > 1 ? atoi(argv[1]) : 85, abw = argc > 2 ? atoi(argv[2]) : 48;
    const int pitch = 896, rows = 518, B = 3, boxW = abw, boxH = 44;
    std::vector<uint8_t> h((size_t)pitch * rows * B);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (uint8_t)(i * 2654435761u >> 13);
    uint8_t *d, *o; int* dl;
    cudaMalloc(&d, h.size()); cudaMalloc(&o, boxW * boxH); cudaMalloc(&dl, 4);
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    int zero = 0; cudaMemcpy(dl, &zero, 4, cudaMemcpyHostToDevice);
    void* fn = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    Maps maps; memset(&maps, 0, sizeof(maps));
    cuuint64_t dims[3] = {(cuuint64_t)pitch, (cuuint64_t)rows, (cuuint64_t)B}, strides[2] = {(cuuint64_t)pitch, (cuuint64_t)pitch * rows};
    cuuint32_t box[3] = {(cuuint32_t)boxW, (cuuint32_t)boxH, 1}, es[3] = {1, 1, 1};
    CUresult r = ((EncodeTiledFn)fn)(&maps.m[0], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode: %d (qres %d)\n", (int)r, (int)q);
    for (int mode = 0; mode < 2; ++mode) {
        const int x = ax, y = 19 + 16, z = 1;
        probe<<<1, 128, boxW * boxH>>>(maps, dl, o, boxW, boxH, x, y, z, mode);
        cudaError_t e = cudaDeviceSynchronize();
        printf("mode %d: %s\n", mode, cudaGetErrorString(e));
        if (e != cudaSuccess) return 1;
        std::vector<uint8_t> got(boxW * boxH);
        cudaMemcpy(got.data(), o, got.size(), cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int rr = 0; rr < boxH; ++rr)
            for (int c = 0; c < boxW; ++c) bad += got[rr * boxW + c] != h[(size_t)z * pitch * rows + (size_t)(y + rr) * pitch + x + c];
        printf("mode %d: %d mismatches\n", mode, bad);
    }
    return 0;
}
