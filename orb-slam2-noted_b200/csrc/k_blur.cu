// K4b -- GaussianBlur(Size(9,9), sigma 2, BORDER_REFLECT_101) on every level
// (src/ORBextractor.cc:1617-1625), with OpenCV's 8-bit fixed-point arithmetic: integer kernel
// [7,17,32,46,52,46,32,17,7] (sum 256), exact 16-bit horizontal sums, out = (V + 32768) >> 16.
// One launch for all levels and frames; 128x64 output tile per CTA, separable through shared memory.
// Integer dot-product instructions do the taps: the horizontal pass is 3 x IDP.4A per pixel on byte windows
// (funnel-shifted from three aligned words), the vertical pass is 5 x IDP.2A per pixel on row pairs of the 16-bit
// sums; 4 pixels per thread step, one 32-bit store per step.  HBM traffic = one read + one write of every level.
#include "orbx_internal.h"

namespace orbx {

constexpr int kBlurTW = 128, kBlurTH = 64, kBlurR = 4, kBlurThreads = 256;
constexpr int kRawW = kBlurTW + 2 * kBlurR + 8;     // bytes per raw row (144: 36 words, 3-word windows never leave the row)
constexpr int kRawH = kBlurTH + 2 * kBlurR;

__device__ __forceinline__ int reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = i < 0 ? -i : 2 * (n - 1) - i;
    return i;
}

__global__ void __launch_bounds__(kBlurThreads)
gauss9Kernel(const __grid_constant__ GeomSet G, const uint8_t* __restrict__ lvl0, size_t lvl0Pitch, size_t lvl0Frame,
             const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blur)
{
    __shared__ __align__(16) uint32_t raw[kRawH][kRawW / 4];
    __shared__ __align__(16) uint2 hsum[kRawH][kBlurTW / 4];   // 4 x u16 per entry
    const int frame = blockIdx.y + G.frame0;
    int level = 0;
    while (level + 1 < G.nlevels && (int)blockIdx.x >= G.lv[level + 1].tileBase) ++level;
    const LevelGeom& g = G.lv[level];
    const int t = blockIdx.x - g.tileBase;
    const int ty = t / g.tilesX, tx = t - ty * g.tilesX;
    const int x0 = tx * kBlurTW, y0 = ty * kBlurTH;
    const uint8_t* img; size_t pitch;
    if (level == 0) { img = lvl0 + (size_t)frame * lvl0Frame; pitch = lvl0Pitch; }
    else { img = pyr + g.planeOff + (size_t)frame * g.pitch * g.h; pitch = (size_t)g.pitch; }
    const int vw = min(kBlurTW, g.w - x0), vh = min(kBlurTH, g.h - y0);   // valid outputs of this tile
    const int ng = (vw + 3) >> 2;
    const int nrr = vh + 2 * kBlurR;
    const bool wordOk = ((((size_t)img) | pitch) & 3) == 0;

    // stage: raw word (r, k) holds image columns x0-4+4k .. +3 of row reflect(y0-4+r); warp = row phase, lane = word
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    for (int r = wrp; r < nrr; r += kBlurThreads / 32) {
        const uint8_t* row = img + (size_t)reflect101(y0 - kBlurR + r, g.h) * pitch;
        for (int k = lane; k < ng + 2; k += 32) {
            const int x = x0 - kBlurR + 4 * k;
            uint32_t v;
            if (wordOk && x >= 0 && x + 3 < g.w) v = __ldg(reinterpret_cast<const uint32_t*>(row + x));
            else {
                v = 0;
                #pragma unroll
                for (int b = 0; b < 4; ++b) v |= (uint32_t)row[reflect101(x + b, g.w)] << (8 * b);
            }
            raw[r][k] = v;
        }
    }
    __syncthreads();
    // horizontal: H(x) = sum_i k_i * src(x - 4 + i), 9 consecutive bytes starting at raw column lx
    const uint32_t c0 = 7u | (17u << 8) | (32u << 16) | (46u << 24), c1 = 52u | (46u << 8) | (32u << 16) | (17u << 24);
    if (lane < ng) {
        for (int r = wrp; r < nrr; r += kBlurThreads / 32) {
            const uint32_t w0 = raw[r][lane], w1 = raw[r][lane + 1], w2 = raw[r][lane + 2];
            uint32_t h[4];
            #pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t a = j ? __funnelshift_r(w0, w1, 8 * j) : w0;
                const uint32_t b = j ? __funnelshift_r(w1, w2, 8 * j) : w1;
                h[j] = __dp4a(w2, 7u << (8 * j), __dp4a(b, c1, __dp4a(a, c0, 0u)));
            }
            hsum[r][lane] = make_uint2(h[0] | (h[1] << 16), h[2] | (h[3] << 16));
        }
    }
    __syncthreads();
    // vertical: V = sum_j k_j * H(row y + j), row pairs through IDP.2A; out = (V + 32768) >> 16
    uint8_t* out = blur + g.blurOff + (size_t)frame * g.pitch * g.h + (size_t)y0 * g.pitch + x0 + 4 * lane;
    const uint32_t k01 = 7u | (17u << 8), k23 = 32u | (46u << 8), k45 = 52u | (46u << 8), k67 = 32u | (17u << 8);
    if (lane < ng) {
        for (int y = wrp; y < vh; y += kBlurThreads / 32) {
            uint2 r[9];
            #pragma unroll
            for (int j = 0; j < 9; ++j) r[j] = hsum[y + j][lane];
            uint32_t v[4];
            #pragma unroll
            for (int p = 0; p < 4; ++p) {
                const uint32_t sel = (p & 1) ? 0x7632u : 0x5410u;
                #define ROWW(j) ((p & 2) ? r[j].y : r[j].x)
                uint32_t acc = 32768u;
                acc = __dp2a_lo(__byte_perm(ROWW(0), ROWW(1), sel), k01, acc);
                acc = __dp2a_lo(__byte_perm(ROWW(2), ROWW(3), sel), k23, acc);
                acc = __dp2a_lo(__byte_perm(ROWW(4), ROWW(5), sel), k45, acc);
                acc = __dp2a_lo(__byte_perm(ROWW(6), ROWW(7), sel), k67, acc);
                acc = __dp2a_lo(ROWW(8), (p & 1) ? (7u << 8) : 7u, acc);
                #undef ROWW
                v[p] = acc;
            }
            // byte 2 of each accumulator
            const uint32_t lo = __byte_perm(v[0], v[1], 0x0062), hi = __byte_perm(v[2], v[3], 0x0062);
            *reinterpret_cast<uint32_t*>(out + (size_t)y * g.pitch) = __byte_perm(lo, hi, 0x5410);
        }
    }
}

int launchBlur(Ctx* c, int frame0, int batch)
{
    GeomSet g = c->geom;
    g.frame0 = frame0;
    stageBegin(c, ST_BLUR);
    dim3 grid(g.blurTilesPerFrame, batch), block(kBlurThreads);
    gauss9Kernel<<<grid, block, 0, c->stream>>>(g, c->lvl0, c->lvl0Pitch, c->lvl0FrameStride, c->dPyr, c->dBlur);
    stageEnd(c, ST_BLUR, 1);
    ORBX_CUDA(c, cudaGetLastError());
    return ORBX_OK;
}

} // namespace orbx
