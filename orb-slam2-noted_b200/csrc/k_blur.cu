// K4b -- GaussianBlur(Size(9,9), sigma 2, BORDER_REFLECT_101) on every level
// (src/ORBextractor.cc:1617-1625), with OpenCV's 8-bit fixed-point arithmetic: integer kernel
// [7,17,32,46,52,46,32,17,7] (sum 256), exact 16-bit horizontal sums, out = (V + 32768) >> 16.
// One launch for all levels and frames; one CTA per tile of 128 columns x up to 256 rows.
//   stage   the tile's rows (+4 reflected rows above/below, 16 columns either side) come into shared memory as 1-D bulk
//           async copies (cp.async.bulk, completion on an mbarrier); the <= 4 reflected columns outside the image are
//           then filled in shared memory.  Plain loads when the source is not 16-byte aligned.
//   filter  each warp owns a band of rows, each lane 4 columns.  Walking down its band a lane computes the horizontal
//           sums of a row with 3 x IDP.4A per pixel (three aligned words against byte-shifted weight constants) and keeps
//           the last 8 rows as (row, row+1) pairs of 16-bit sums in registers, so the vertical pass is 5 x IDP.2A per
//           pixel with no second trip through shared memory; one 32-bit store per lane per row.
// HBM traffic = one read + one write of every level.
#include "orbx_internal.h"

namespace orbx {

constexpr int kBlurR = 4, kBlurThreads = 128, kBlurWarps = kBlurThreads / 32;
constexpr int kRawPad = 16;                              // columns staged left of the tile (keeps bulk copies 16-byte aligned)
constexpr int kRawW = kBlurTW + 2 * kRawPad;             // bytes per staged row (160: 40 words)
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
    __shared__ __align__(128) uint32_t raw[kRawH][kRawW / 4];
    __shared__ __align__(8) unsigned long long sBar;
    const int frame = blockIdx.y + G.frame0;
    int level = 0;
    while (level + 1 < G.nlevels && (int)blockIdx.x >= G.lv[level + 1].tileBase) ++level;
    const LevelGeom& g = G.lv[level];
    const int t = blockIdx.x - g.tileBase;
    const int ty = t / g.tilesX, tx = t - ty * g.tilesX;
    const int tileH = (g.h + g.tilesY - 1) / g.tilesY;                    // balanced row split, <= kBlurTH
    const int x0 = tx * kBlurTW, y0 = ty * tileH;
    const uint8_t* img; size_t pitch;
    if (level == 0) { img = lvl0 + (size_t)frame * lvl0Frame; pitch = lvl0Pitch; }
    else { img = pyr + g.planeOff + (size_t)frame * g.pitch * g.h; pitch = (size_t)g.pitch; }
    const int vw = min(kBlurTW, g.w - x0), vh = min(tileH, g.h - y0);     // valid outputs of this tile
    const int ng = (vw + 3) >> 2;
    const int nrr = vh + 2 * kBlurR;
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    uint8_t* rawB = reinterpret_cast<uint8_t*>(&raw[0][0]);               // raw byte (r, c) = image column x0 - 16 + c

    const bool bulk = ((((size_t)img) | pitch) & 15) == 0 && g.w >= 8 && g.h >= 8;
    if (bulk) {
        const int xs = max(x0 - kRawPad, 0), xe = min(x0 + kBlurTW + kRawPad, (int)pitch);
        const uint32_t rowBytes = (uint32_t)(xe - xs);
        const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&sBar);
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
        if (threadIdx.x == 0)
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(rowBytes * (uint32_t)nrr) : "memory");
        for (int r = threadIdx.x; r < nrr; r += kBlurThreads) {
            int yy = y0 - kBlurR + r;
            yy = yy < 0 ? -yy : (yy >= g.h ? 2 * (g.h - 1) - yy : yy);
            const uint8_t* src = img + (size_t)yy * pitch + xs;
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(rawB + r * kRawW + (xs - (x0 - kRawPad)));
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(dst), "l"(src), "r"(rowBytes), "r"(bar) : "memory");
        }
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(bar) : "memory");
        if (x0 == 0 || x0 + vw + kBlurR > g.w) {     // block-uniform: tiles touching the left / right image border
            __syncthreads();
            for (int i = threadIdx.x; i < nrr * 8; i += kBlurThreads) {
                const int r = i >> 3, j = i & 7;
                const int x = j < 4 ? j - 4 : g.w + (j - 4);
                if (x < x0 - kBlurR || x >= x0 + vw + kBlurR) continue;
                const int src = x < 0 ? -x : 2 * (g.w - 1) - x;
                rawB[r * kRawW + (x - (x0 - kRawPad))] = rawB[r * kRawW + (src - (x0 - kRawPad))];
            }
            __syncthreads();
        }
    } else {
        const int nb = 4 * ng + 2 * kBlurR;          // columns x0-4 .. x0+4ng+3
        for (int i = threadIdx.x; i < nrr * nb; i += kBlurThreads) {
            const int r = i / nb, cidx = i - r * nb;
            rawB[r * kRawW + (kRawPad - kBlurR) + cidx] =
                img[(size_t)reflect101(y0 - kBlurR + r, g.h) * pitch + reflect101(x0 - kBlurR + cidx, g.w)];
        }
        __syncthreads();
    }

    // rows of this warp: a balanced band of the tile
    const int rowsW = (vh + kBlurWarps - 1) / kBlurWarps;
    const int yb = wrp * rowsW;                                   // first output row of the band (tile-relative)
    const int nOut = min(rowsW, vh - yb);
    if (nOut <= 0 || lane >= ng) return;
    const int nH = nOut + 2 * kBlurR;                             // horizontal rows the band needs
    const uint32_t k01 = 7u | (17u << 8), k23 = 32u | (46u << 8), k45 = 52u | (46u << 8), k67 = 32u | (17u << 8);
    const uint32_t* rp = &raw[yb][(kRawPad - kBlurR) / 4 + lane];  // words holding columns x0-4+4*lane .. +11 of the band's row 0
    uint8_t* out = blur + g.blurOff + (size_t)frame * g.pitch * g.h + (size_t)(y0 + yb) * g.pitch + x0 + 4 * lane;
    const int opitch = g.pitch;

    uint32_t P[8][4];       // P[r & 7][p] = H_r(p) | H_{r+1}(p) << 16 for the last 8 rows
    uint32_t hp01 = 0, hp23 = 0;
    #pragma unroll
    for (int s = 0; s < 8; ++s) { P[s][0] = P[s][1] = P[s][2] = P[s][3] = 0; }

    // one row: horizontal sums, the (r-1, r) pairs, and -- once 9 rows are in -- one output row.  j = r mod 8.
    auto step = [&](const int j, const bool emit) {
        // horizontal: H(x) = sum_i k_i * src(x - 4 + i), 9 consecutive bytes
        const uint32_t w0 = rp[0], w1 = rp[1], w2 = rp[2];
        rp += kRawW / 4;
        // pixel q needs bytes q .. q+8 of (w0, w1, w2): the nine weights are laid over the three aligned words, shifted by q
        // bytes (compile-time constants), so no byte window has to be cut out of the data
        uint32_t h[4];
        #pragma unroll
        for (int q = 0; q < 4; ++q) {
            constexpr unsigned long long kLo = 7ull | (17ull << 8) | (32ull << 16) | (46ull << 24) | (52ull << 32) | (46ull << 40) | (32ull << 48) | (17ull << 56);
            const uint32_t k0 = (uint32_t)(kLo << (8 * q));                                  // weights on w0
            const uint32_t k1 = (uint32_t)(q ? (kLo >> (32 - 8 * q)) : (kLo >> 32));          // weights on w1
            const uint32_t k2 = (uint32_t)((q ? (kLo >> (64 - 8 * q)) : 0ull) | (7ull << (8 * q)));   // weights on w2 (tap 8 = 7)
            h[q] = __dp4a(w2, k2, __dp4a(w1, k1, __dp4a(w0, k0, 0u)));
        }
        const uint32_t h01 = h[1] * 65536u + h[0], h23 = h[3] * 65536u + h[2];   // one IMAD each (sums < 65536)
        // pair (r-1, r) completes slot (r-1) & 7
        const int sNew = (j + 7) & 7;
        P[sNew][0] = __byte_perm(hp01, h01, 0x5410); P[sNew][1] = __byte_perm(hp01, h01, 0x7632);
        P[sNew][2] = __byte_perm(hp23, h23, 0x5410); P[sNew][3] = __byte_perm(hp23, h23, 0x7632);
        hp01 = h01; hp23 = h23;
        if (emit) {
            // vertical for output row r-8: pairs (r-8,r-7) (r-6,r-5) (r-4,r-3) (r-2,r-1) and row r
            uint32_t v[4];
            #pragma unroll
            for (int p = 0; p < 4; ++p) {
                uint32_t acc = 32768u;
                acc = __dp2a_lo(P[j & 7][p], k01, acc);
                acc = __dp2a_lo(P[(j + 2) & 7][p], k23, acc);
                acc = __dp2a_lo(P[(j + 4) & 7][p], k45, acc);
                acc = __dp2a_lo(P[(j + 6) & 7][p], k67, acc);
                acc = __dp2a_lo((p & 2) ? h23 : h01, (p & 1) ? (7u << 8) : 7u, acc);
                v[p] = acc;
            }
            // byte 2 of each accumulator
            const uint32_t lo = __byte_perm(v[0], v[1], 0x0062), hi = __byte_perm(v[2], v[3], 0x0062);
            *reinterpret_cast<uint32_t*>(out) = __byte_perm(lo, hi, 0x5410);
            out += opitch;
        }
    };
    // warm-up: the 8 rows above the band's first output row
    #pragma unroll
    for (int j = 0; j < 8; ++j) step(j, false);
    // steady state, 8 rows per trip without any per-row test, then the remainder
    int r = 2 * kBlurR;
    for (; r + 8 <= nH; r += 8) {
        #pragma unroll
        for (int j = 0; j < 8; ++j) step(j, true);
    }
    #pragma unroll
    for (int j = 0; j < 7; ++j)
        if (r + j < nH) step(j, true);                             // warp-uniform
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
