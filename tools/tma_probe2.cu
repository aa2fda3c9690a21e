// Diagnostic: the CUDA programming guide's 2-D TMA example (libcu++ wrappers), to compare against hand-written PTX.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda/barrier>
#include <cstdio>
#include <cstring>
#include <vector>
using barrier = cuda::barrier<cuda::thread_scope_block>;
namespace cde = cuda::device::experimental;

__global__ void probe(const __grid_constant__ CUtensorMap tm, int x, int y, uint8_t* out)
{
    __shared__ alignas(128) uint8_t buf[64 * 32];
    #pragma nv_diag_suppress static_var_with_dynamic_init
    __shared__ barrier bar;
    if (threadIdx.x == 0) { init(&bar, blockDim.x); cde::fence_proxy_async_shared_cta(); }
    __syncthreads();
    barrier::arrival_token token;
    if (threadIdx.x == 0) {
        cde::cp_async_bulk_tensor_2d_global_to_shared(&buf, &tm, x, y, bar);
        token = cuda::device::barrier_arrive_tx(bar, 1, sizeof(buf));
    } else token = bar.arrive();
    bar.wait(std::move(token));
    for (int i = threadIdx.x; i < 64 * 32; i += blockDim.x) out[i] = buf[i];
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main()
{
    const int pitch = 640, h = 480;
    std::vector<uint8_t> img((size_t)pitch * h);
    for (size_t i = 0; i < img.size(); ++i) img[i] = (uint8_t)(i * 2654435761u >> 24);
    uint8_t *d, *o; cudaMalloc(&d, img.size()); cudaMalloc(&o, 4096);
    cudaMemcpy(d, img.data(), img.size(), cudaMemcpyHostToDevice);
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    CUtensorMap m; memset(&m, 0, sizeof(m));
    const cuuint64_t dims[2] = {(cuuint64_t)pitch, (cuuint64_t)h}, strides[1] = {(cuuint64_t)pitch};
    const cuuint32_t box[2] = {64u, 32u}, es[2] = {1u, 1u};
    CUresult r = ((EncodeTiledFn)p)(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode: %d\n", (int)r);
    probe<<<1, 128>>>(m, 101, 57, o);
    cudaError_t e = cudaDeviceSynchronize();
    printf("libcu++ 2d: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    std::vector<uint8_t> got(64 * 32);
    cudaMemcpy(got.data(), o, got.size(), cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int r2 = 0; r2 < 32; ++r2) for (int c = 0; c < 64; ++c) if (got[r2 * 64 + c] != img[(size_t)(57 + r2) * pitch + 101 + c]) ++bad;
    printf("  mismatches: %d\n", bad);
    return 0;
}
