// K4b -- GaussianBlur(Size(9,9), sigma 2, BORDER_REFLECT_101) on every level
// (src/ORBextractor.cc:1617-1625), with OpenCV's 8-bit fixed-point arithmetic: integer kernel
// [7,17,32,46,52,46,32,17,7] (sum 256), horizontal sums kept in 16 bits, out = (V + 32768) >> 16.
// One launch for all levels and frames; 64x32 output tile per CTA, separable through shared memory.
#include "orbx_internal.h"

namespace orbx {

constexpr int kBlurTW = 64, kBlurTH = 32, kBlurR = 4, kBlurThreads = 256;

__device__ __forceinline__ int reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = i < 0 ? -i : 2 * (n - 1) - i;
    return i;
}

__global__ void __launch_bounds__(kBlurThreads)
gauss9Kernel(const GeomSet* __restrict__ G, const uint8_t* __restrict__ lvl0, size_t lvl0Pitch, size_t lvl0Frame,
             const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blur)
{
    __shared__ uint8_t raw[kBlurTH + 2 * kBlurR][kBlurTW + 2 * kBlurR + 8];
    __shared__ uint16_t hsum[kBlurTH + 2 * kBlurR][kBlurTW];
    const int frame = blockIdx.y;
    int level = 0;
    while (level + 1 < G->nlevels && (int)blockIdx.x >= G->lv[level + 1].tileBase) ++level;
    const LevelGeom& g = G->lv[level];
    const int t = blockIdx.x - g.tileBase;
    const int ty = t / g.tilesX, tx = t - ty * g.tilesX;
    const int x0 = tx * kBlurTW, y0 = ty * kBlurTH;
    const uint8_t* img; size_t pitch;
    if (level == 0) { img = lvl0 + (size_t)frame * lvl0Frame; pitch = lvl0Pitch; }
    else { img = pyr + g.planeOff + (size_t)frame * g.pitch * g.h; pitch = (size_t)g.pitch; }

    constexpr int RW = kBlurTW + 2 * kBlurR, RH = kBlurTH + 2 * kBlurR;
    for (int i = threadIdx.x; i < RW * RH; i += kBlurThreads) {
        const int y = i / RW, x = i - y * RW;
        const int sx = reflect101(x0 + x - kBlurR, g.w), sy = reflect101(y0 + y - kBlurR, g.h);
        raw[y][x] = img[(size_t)sy * pitch + sx];
    }
    __syncthreads();
    for (int i = threadIdx.x; i < RH * kBlurTW; i += kBlurThreads) {
        const int y = i / kBlurTW, x = i - y * kBlurTW;
        const uint8_t* r = &raw[y][x];
        const int s = 7 * (r[0] + r[8]) + 17 * (r[1] + r[7]) + 32 * (r[2] + r[6]) + 46 * (r[3] + r[5]) + 52 * r[4];
        hsum[y][x] = (uint16_t)s;
    }
    __syncthreads();
    uint8_t* out = blur + g.blurOff + (size_t)frame * g.pitch * g.h;
    for (int i = threadIdx.x; i < kBlurTH * kBlurTW; i += kBlurThreads) {
        const int y = i / kBlurTW, x = i - y * kBlurTW;
        if (x0 + x < g.w && y0 + y < g.h) {
            const uint32_t v = 7u * (hsum[y][x] + hsum[y + 8][x]) + 17u * (hsum[y + 1][x] + hsum[y + 7][x]) +
                               32u * (hsum[y + 2][x] + hsum[y + 6][x]) + 46u * (hsum[y + 3][x] + hsum[y + 5][x]) +
                               52u * hsum[y + 4][x];
            out[(size_t)(y0 + y) * g.pitch + (x0 + x)] = (uint8_t)((v + 32768u) >> 16);
        }
    }
}

int launchBlur(Ctx* c, int batch)
{
    const GeomSet& g = c->geom;
    stageBegin(c, ST_BLUR);
    dim3 grid(g.blurTilesPerFrame, batch), block(kBlurThreads);
    gauss9Kernel<<<grid, block, 0, c->stream>>>(c->dGeom, c->lvl0, c->lvl0Pitch, c->lvl0FrameStride, c->dPyr, c->dBlur);
    stageEnd(c, ST_BLUR, 1);
    ORBX_CUDA(c, cudaGetLastError());
    return ORBX_OK;
}

} // namespace orbx
