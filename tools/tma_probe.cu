// Diagnostic: one 3-D tensor-map box copy (u8 plane -> shared) with the map in global memory or in kernel parameters.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -o tools/tma_probe tools/tma_probe.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <vector>
#include <cstdlib>

__global__ void probe(const __grid_constant__ CUtensorMap pm, const CUtensorMap* gm, int useParam, int x, int y, int z,
                      int boxBytes, uint8_t* out, int mode)
{
    __shared__ __align__(128) uint8_t buf[4096];
    __shared__ __align__(8) unsigned long long bar;
    const uint32_t b = (uint32_t)__cvta_generic_to_shared(&bar), d = (uint32_t)__cvta_generic_to_shared(buf);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"((uint32_t)boxBytes) : "memory");
        const CUtensorMap* m = useParam ? &pm : gm;
        if (mode == 0)
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"(d), "l"(m), "r"(x), "r"(y), "r"(z), "r"(b) : "memory");
        else if (mode == 1)
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         ::"r"(d), "l"(m), "r"(x), "r"(y + z * 480), "r"(b) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.3d.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"(d), "l"(m), "r"(x), "r"(y), "r"(z), "r"(b) : "memory");
    }
    uint32_t done = 0;
    while (!done)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done) : "r"(b) : "memory");
    __syncwarp();
    for (int i = threadIdx.x; i < boxBytes; i += 32) out[i] = buf[i];
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv)
{
    const int pitch = 640, h = 480, frames = 4;
    const int bw = argc > 1 ? atoi(argv[1]) : 48, bh = argc > 2 ? atoi(argv[2]) : 37, up = argc > 3 ? atoi(argv[3]) : 1;
    const int l2 = argc > 4 ? atoi(argv[4]) : 1, mode = argc > 5 ? atoi(argv[5]) : 0;
    std::vector<uint8_t> img((size_t)pitch * h * frames);
    for (size_t i = 0; i < img.size(); ++i) img[i] = (uint8_t)(i * 2654435761u >> 24);
    uint8_t *d, *o; CUtensorMap* dm;
    cudaMalloc(&d, img.size()); cudaMalloc(&o, 4096); cudaMalloc(&dm, sizeof(CUtensorMap));
    cudaMemcpy(d, img.data(), img.size(), cudaMemcpyHostToDevice);
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    printf("entry point: %s q=%d p=%p\n", cudaGetErrorString(e), (int)q, p);
    CUtensorMap m; memset(&m, 0, sizeof(m));
    const cuuint64_t dims[3] = {(cuuint64_t)pitch, (cuuint64_t)h, (cuuint64_t)frames};
    const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)pitch * h};
    const cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, 1u}, es[3] = {1u, 1u, 1u};
    const cuuint64_t dims2[2] = {(cuuint64_t)pitch, (cuuint64_t)h * frames};
    CUresult r = mode == 1 ? ((EncodeTiledFn)p)(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, dims2, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) :
                 ((EncodeTiledFn)p)(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                    CU_TENSOR_MAP_SWIZZLE_NONE, l2 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B : CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode: %d\n", (int)r);
    cudaMemcpy(dm, &m, sizeof(m), cudaMemcpyHostToDevice);
    for (int useParam = up; useParam == up; ++useParam) {
        const int x = 101, y = 57, z = 2;
        probe<<<1, 32>>>(m, dm, useParam, x, y, z, bw * bh, o, mode);
        for (int i = 0; i < 16; ++i) printf("%016llx ", (unsigned long long)m.opaque[i]); printf("\n");
        e = cudaDeviceSynchronize();
        printf("useParam=%d: %s\n", useParam, cudaGetErrorString(e));
        if (e != cudaSuccess) return 1;
        std::vector<uint8_t> got(bw * bh);
        cudaMemcpy(got.data(), o, got.size(), cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int r2 = 0; r2 < bh; ++r2) for (int c = 0; c < bw; ++c)
            if (got[r2 * bw + c] != img[((size_t)z * h + y + r2) * pitch + x + c]) ++bad;
        printf("  mismatches: %d\n", bad);
    }
    return 0;
}
