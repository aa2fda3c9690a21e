// K1 -- ComputePyramid (src/ORBextractor.cc:1664-1733): chained 8-bit bilinear resize with the
// exact fixed-point arithmetic of cv::resize(INTER_LINEAR) (11-bit weights, (b*(T>>4))>>16, +2>>2),
// plus the reflect-101 border copy used only when a caller asks for mvImagePyramid with its frame.
// HBM-streaming integer kernel: 4 output pixels per thread, one 32-bit store, taps from a tiny
// per-level table (L1/L2 resident).  Level l is produced from level l-1, so levels are separate
// launches; the batch dimension provides the parallelism (grid.z = frames).
#include "orbx_internal.h"
#include <algorithm>

namespace orbx {

constexpr int kResizeRows = 8;   // destination rows per thread: taps stay in registers, the horizontal pass of a source
                                 // row is reused by the next destination row (scale 1.2: ~1.2 row passes per row, not 2)

__device__ __forceinline__ void hpass(const uint8_t* __restrict__ S, const int (&sx)[4], const int (&sx1)[4],
                                      const int (&c0)[4], const int (&c1)[4], int (&t)[4])
{
    #pragma unroll
    for (int i = 0; i < 4; ++i) t[i] = ((int)__ldg(S + sx[i]) * c0[i] + (int)__ldg(S + sx1[i]) * c1[i]) >> 4;
}

__global__ void __launch_bounds__(128)
resizeLinearKernel(const uint8_t* __restrict__ src, size_t sPitch, size_t sFrame, int sw, int sh,
                   uint8_t* __restrict__ dst, size_t dPitch, size_t dFrame, int dw, int dh,
                   const ResizeTap* __restrict__ tapX, const ResizeTap* __restrict__ tapY, int frame0)
{
    const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (x0 >= dw) return;
    int sx[4], sx1[4], c0[4], c1[4];
    #pragma unroll
    for (int i = 0; i < 4; ++i) {
        const ResizeTap tx = tapX[min(x0 + i, dw - 1)];
        sx[i] = tx.s; sx1[i] = min(tx.s + 1, sw - 1); c0[i] = tx.c0; c1[i] = tx.c1;
    }
    const uint8_t* S = src + (size_t)(blockIdx.z + frame0) * sFrame;
    uint8_t* D = dst + (size_t)(blockIdx.z + frame0) * dFrame;
    const int yb = blockIdx.y * kResizeRows, ye = min(yb + kResizeRows, dh);
    int rowA = -1, rowB = -1, tA[4] = {0, 0, 0, 0}, tB[4] = {0, 0, 0, 0};
    for (int y = yb; y < ye; ++y) {
        const ResizeTap ty = tapY[y];
        const int r0 = ty.s, r1 = min(ty.s + 1, sh - 1);   // rows are clipped, weights kept (OpenCV VResize)
        if (r0 == rowB) {
            #pragma unroll
            for (int i = 0; i < 4; ++i) tA[i] = tB[i];
            rowA = rowB;
        } else if (r0 != rowA) { hpass(S + (size_t)r0 * sPitch, sx, sx1, c0, c1, tA); rowA = r0; }
        if (r1 == rowA) {
            #pragma unroll
            for (int i = 0; i < 4; ++i) tB[i] = tA[i];
            rowB = rowA;
        } else { hpass(S + (size_t)r1 * sPitch, sx, sx1, c0, c1, tB); rowB = r1; }
        const int b0 = ty.c0, b1 = ty.c1;
        uint32_t packed = 0;
        #pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int v = (((b0 * tA[i]) >> 16) + ((b1 * tB[i]) >> 16) + 2) >> 2;
            packed |= (uint32_t)(v & 0xff) << (8 * i);
        }
        // rows are padded to a multiple of 16 bytes, so the 4-byte store never leaves the row
        *reinterpret_cast<uint32_t*>(D + (size_t)y * dPitch + x0) = packed;
    }
}


// Fast path (scale <= ~1.33, word-aligned source): the 4 destination pixels of a thread read at most 8 consecutive
// source bytes counted from an aligned word, so one horizontal pass is 3 aligned 32-bit loads, and per pixel one
// funnel shift (bytes s, s+1 into the low half) and one IDP.2A against the (c0, c1) weight pair.
// kRows destination rows per thread: 16 for batches that fill the GPU (a source row is reduced once for ~1.2 destination rows,
// taps and pointers amortised), 2 for a single frame or a stereo pair, where the chain of dependent loads per thread is the latency
template <int kRows>
__global__ void __launch_bounds__(128)
resizeLinearWordKernel(const uint8_t* __restrict__ src, size_t sPitch, size_t sFrame, int sw, int sh,
                       uint8_t* __restrict__ dst, size_t dPitch, size_t dFrame, int dw, int dh,
                       const ResizeTap* __restrict__ tapX, const ResizeTap* __restrict__ tapY, int frame0)
{
    // work items = (row band, 4-pixel group) flattened, so warps stay full whatever the level width
    const int groups = (dw + 3) >> 2;
    const int item = blockIdx.x * blockDim.x + threadIdx.x;
    const int band = item / groups;
    const int x0 = (item - band * groups) * 4;
    if (band * kRows >= dh) return;
    uint32_t cpair[4], shift[4], pick[4];       // pick = all ones when the pixel's byte pair starts in the second word
    int base = 0;
    #pragma unroll
    for (int i = 0; i < 4; ++i) {
        const ResizeTap tx = tapX[min(x0 + i, dw - 1)];
        if (i == 0) base = tx.s >> 2;
        const int off = tx.s - 4 * base;                    // 0..7 (checked on the host)
        pick[i] = off >= 4 ? 0xffffffffu : 0u; shift[i] = 8u * (off & 3);
        cpair[i] = (uint32_t)(uint16_t)tx.c0 | ((uint32_t)(uint16_t)tx.c1 << 16);
    }
    const int pw = (int)(sPitch >> 2);
    const bool has1 = base + 1 < pw, has2 = base + 2 < pw;   // words past the row end only ever meet zero weights
    const int yb = band * kRows, ye = min(yb + kRows, dh);
    int y = yb;
    ResizeTap ty = tapY[y];
    int r = ty.s;                                             // source row the pointer is on
    const uint32_t* p = reinterpret_cast<const uint32_t*>(src + (size_t)(blockIdx.z + frame0) * sFrame + (size_t)r * sPitch) + base;
    uint8_t* D = dst + (size_t)(blockIdx.z + frame0) * dFrame + (size_t)y * dPitch + x0;

    // Source rows are walked once, four at a time (12 independent loads in flight per thread); a destination row is
    // written as soon as its second source row has been reduced horizontally.
    auto hcompute = [&](const uint32_t (&w)[3], uint32_t (&t)[4]) {
        const uint32_t x01 = w[0] ^ w[1], x12 = w[1] ^ w[2];
        #pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t lo = w[0] ^ (x01 & pick[i]), hi = w[1] ^ (x12 & pick[i]);
            t[i] = __dp2a_lo(cpair[i], __funnelshift_r(lo, hi, shift[i]), 0u) >> 4;
        }
    };
    auto loadRow = [&](uint32_t (&w)[3]) {                    // loads the row under the pointer, then steps down
        w[0] = __ldg(p);
        w[1] = has1 ? __ldg(p + 1) : w[0];
        w[2] = has2 ? __ldg(p + 2) : w[1];
        if (r < sh - 1) p += pw;                              // rows are clipped, weights kept (OpenCV VResize)
        ++r;
    };
    uint32_t tp[4];
    {
        uint32_t w[3];
        loadRow(w);
        hcompute(w, tp);
    }
    int rb = ty.s;                                            // tp holds row rb
    while (y < ye) {
        uint32_t w[4][3];
        #pragma unroll
        for (int k = 0; k < 4; ++k) loadRow(w[k]);
        #pragma unroll
        for (int k = 0; k < 4; ++k) {
            uint32_t tc[4];
            hcompute(w[k], tc);                               // row rb + 1
            if (y < ye && ty.s == rb) {                       // rows (rb, rb+1) are this destination row's pair: at most one
                                                              // destination row per source row (the host checks that the taps rise)
                // (b * t) >> 16 as the high word of (b << 16) * t: weights and sums are non-negative
                const uint32_t b0 = (uint32_t)ty.c0 << 16, b1 = (uint32_t)ty.c1 << 16;
                uint32_t v[4];
                #pragma unroll
                for (int i = 0; i < 4; ++i) v[i] = (__umulhi(b0, tp[i]) + __umulhi(b1, tc[i]) + 2u) >> 2;   // <= 255
                *reinterpret_cast<uint32_t*>(D) = (v[0] + (v[1] << 8)) + ((v[2] << 16) + (v[3] << 24));
                D += dPitch;
                if (++y < ye) ty = tapY[y];
            }
            ++rb;
            #pragma unroll
            for (int i = 0; i < 4; ++i) tp[i] = tc[i];
        }
    }
}

// Staged path (batches that fill the GPU).  ncu on the register-only kernels shows the LSU data pipe, not DRAM, as the
// limiter (l1tex data-pipe wavefronts 71 %, DRAM 31 %): a warp's lanes sit ~1.2 x their pixel count apart in the source,
// so every warp-wide global load spans 3-4 cache lines (one wavefront each), twice that where a warp straddles two row
// bands.  Here a CTA takes kBands bands of 16 destination rows; the source rows they need are CONTIGUOUS in memory (whole
// rows incl. padding), so ONE bulk async copy (cp.async.bulk -> UBLKCP, completion on an mbarrier) brings them into shared
// memory without touching the LSU, and the strided per-thread reads become 32-bit LDS with at most 2-way bank conflicts.
// Thread = 8 destination pixels x 16 rows: three funnel shifts align the 12-byte window to the first pixel's source column
// once per source row, then pixel i's byte pair (s_i, s_i + 1) is one PRMT with a per-thread selector (pixels 0-3 start 0..6
// bytes into words v0:v1, pixels 4-7 start 4..10 bytes in = words v1:v2; checked per group on the host: resize8Ok), one
// IDP.2A against the (c0, c1) pair and one shift; a destination row is 2 IMAD.HI + add + shift per pixel, 3 PRMT per packed
// word and one 64-bit store.  Reads past a row's end stay inside the staged block (+16 bytes of slack) and meet zero weights.
constexpr int kStageRows = 16;
__global__ void __launch_bounds__(512)
resizeLinearStagedKernel(const uint8_t* __restrict__ src, size_t sPitch, size_t sFrame, int sw, int sh,
                         uint8_t* __restrict__ dst, size_t dPitch, size_t dFrame, int dw, int dh,
                         const uint32_t* __restrict__ tapG, const ResizeTap* __restrict__ tapY, int frame0, int nBands)
{
    extern __shared__ __align__(128) uint8_t sRows[];
    __shared__ __align__(8) unsigned long long sBar;
    const int groups = (dw + 7) >> 3;
    const int yCta = blockIdx.x * nBands * kStageRows;
    const int yCtaEnd = min(yCta + nBands * kStageRows, dh);
    const int rFirst = tapY[yCta].s, rLast = min((int)tapY[yCtaEnd - 1].s + 1, sh - 1);
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&sBar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        const uint32_t bytes = (uint32_t)((size_t)(rLast - rFirst + 1) * sPitch);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        const uint8_t* g = src + (size_t)(blockIdx.z + frame0) * sFrame + (size_t)rFirst * sPitch;
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"((uint32_t)__cvta_generic_to_shared(sRows)), "l"(g), "r"(bytes), "r"(bar) : "memory");
    }
    // per-thread constants while the copy is in flight
    const int band = threadIdx.x / groups;
    const int x0 = (threadIdx.x - band * groups) * 8;
    (void)dw;
    const int yb = yCta + band * kStageRows, ye = min(yb + kStageRows, yCtaEnd);
    const bool active = band < nBands && yb < yCtaEnd;
    // Per-group constants from the host-built table (buildResizeGroups), structure of arrays: a warp's 11 loads are 11
    // lane-contiguous lines.  (Reading the eight 8-byte ResizeTap records per thread instead puts the lanes 64 bytes apart:
    // 16 cache lines per request, as many LSU wavefronts as the whole band's row work -- measured with ncu.)
    uint32_t cpair[8], sel[8];
    const int gi = min(threadIdx.x - band * groups, groups - 1);
    const int s0 = (int)__ldg(tapG + gi);
    {
        const uint32_t sa = __ldg(tapG + groups + gi), sb = __ldg(tapG + 2 * groups + gi);
        // PRMT reads the low four nibbles of its selector; bytes 2-3 of the result are never used, so a shifted copy will do
        #pragma unroll
        for (int i = 0; i < 4; ++i) { sel[i] = sa >> (8 * i); sel[4 + i] = sb >> (8 * i); }
        #pragma unroll
        for (int i = 0; i < 8; ++i) cpair[i] = __ldg(tapG + (3 + i) * groups + gi);
    }
    const uint32_t ash = 8u * (uint32_t)(s0 & 3);
    const int pw = (int)(sPitch >> 2);
    int y = active ? yb : 0;
    ResizeTap ty = tapY[y];
    __syncthreads();                                          // the mbarrier is initialised before anybody polls it
    uint32_t done = 0;
    while (!done)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done) : "r"(bar) : "memory");
    if (!active) return;
    int r = ty.s;                                             // source row the pointer is on
    const uint32_t* p = reinterpret_cast<const uint32_t*>(sRows) + (size_t)(r - rFirst) * pw + (s0 >> 2);
    uint8_t* D = dst + (size_t)(blockIdx.z + frame0) * dFrame + (size_t)y * dPitch + x0;

    auto hrow = [&](uint32_t (&t)[8]) {                       // reduces the row under the pointer, then steps down
        const uint32_t w0 = p[0], w1 = p[1], w2 = p[2], w3 = p[3];
        if (r < rLast) p += pw;                               // rows are clipped, weights kept (OpenCV VResize)
        ++r;
        const uint32_t v0 = __funnelshift_r(w0, w1, ash), v1 = __funnelshift_r(w1, w2, ash), v2 = __funnelshift_r(w2, w3, ash);
        #pragma unroll
        for (int i = 0; i < 4; ++i) t[i] = __dp2a_lo(cpair[i], __byte_perm(v0, v1, sel[i]), 0u) >> 4;
        #pragma unroll
        for (int i = 4; i < 8; ++i) t[i] = __dp2a_lo(cpair[i], __byte_perm(v1, v2, sel[i]), 0u) >> 4;
    };
    uint32_t tp[8];
    hrow(tp);
    int rb = ty.s;                                            // tp holds row rb
    while (y < ye) {
        #pragma unroll
        for (int k = 0; k < 4; ++k) {
            uint32_t tc[8];
            hrow(tc);                                         // row rb + 1
            if (y < ye && ty.s == rb) {                       // at most one destination row per source row (host-checked)
                const uint32_t b0 = (uint32_t)ty.c0 << 16, b1 = (uint32_t)ty.c1 << 16;
                uint32_t v[8];
                #pragma unroll
                for (int i = 0; i < 8; ++i) v[i] = (__umulhi(b0, tp[i]) + __umulhi(b1, tc[i]) + 2u) >> 2;   // <= 255
                uint2 o;
                o.x = __byte_perm(__byte_perm(v[0], v[1], 0x0040), __byte_perm(v[2], v[3], 0x0040), 0x5410);
                o.y = __byte_perm(__byte_perm(v[4], v[5], 0x0040), __byte_perm(v[6], v[7], 0x0040), 0x5410);
                *reinterpret_cast<uint2*>(D) = o;             // x0 is a multiple of 8 and rows are padded to 16 bytes
                D += dPitch;
                if (++y < ye) ty = tapY[y];
            }
            ++rb;
            #pragma unroll
            for (int i = 0; i < 8; ++i) tp[i] = tc[i];
        }
    }
}

// cvtColor(RGB / BGR / RGBA / BGRA -> GRAY), the step in front of the extractor (src/Tracking.cc:252-278, 313-326, 377-390),
// with OpenCV's 8-bit fixed point: (R * 9798 + G * 19235 + B * 3735 + 16384) >> 15 (cv2 4.13, the pinned dependency).
// 4 pixels per thread: 12 or 16 source bytes in, one 32-bit store out.
template <int kCh>
__global__ void __launch_bounds__(256)
cvtGrayKernel(const uint8_t* __restrict__ src, size_t sPitch, int w, int h, int bgr, uint8_t* __restrict__ dst, size_t dPitch)
{
    const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y;
    if (x0 >= w) return;
    const uint8_t* s = src + (size_t)y * sPitch + (size_t)x0 * kCh;
    const uint32_t cr = bgr ? 3735u : 9798u, cb = bgr ? 9798u : 3735u;   // weights of channel 0 and channel 2
    uint32_t px[4];
    if (x0 + 4 <= w && ((((size_t)s) & 3) == 0)) {
        const uint32_t* sw = reinterpret_cast<const uint32_t*>(s);
        if (kCh == 4) {
            #pragma unroll
            for (int i = 0; i < 4; ++i) px[i] = __ldg(sw + i);
        } else {
            const uint32_t a = __ldg(sw), b = __ldg(sw + 1), c = __ldg(sw + 2);
            px[0] = a; px[1] = __funnelshift_r(a, b, 24); px[2] = __funnelshift_r(b, c, 16); px[3] = c >> 8;
        }
    } else {
        #pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint8_t* q = s + (size_t)min(i, w - 1 - x0) * kCh;
            px[i] = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16);
        }
    }
    uint32_t out = 0;
    #pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t c0 = px[i] & 0xffu, c1 = (px[i] >> 8) & 0xffu, c2 = (px[i] >> 16) & 0xffu;
        out |= ((c0 * cr + c1 * 19235u + c2 * cb + 16384u) >> 15) << (8 * i);
    }
    uint8_t* d = dst + (size_t)y * dPitch + x0;
    if (x0 + 4 <= w && ((((size_t)d) & 3) == 0)) *reinterpret_cast<uint32_t*>(d) = out;
    else for (int i = 0; i < min(4, w - x0); ++i) d[i] = (uint8_t)(out >> (8 * i));
}

int launchCvtGray(Ctx* c, const uint8_t* src, size_t sPitch, int w, int h, int format, uint8_t* dst, size_t dPitch)
{
    dim3 grid((w + 1023) / 1024, h), block(256);
    const int bgr = format & 1;
    if (format & 2) cvtGrayKernel<4><<<grid, block, 0, c->stream>>>(src, sPitch, w, h, bgr, dst, dPitch);
    else cvtGrayKernel<3><<<grid, block, 0, c->stream>>>(src, sPitch, w, h, bgr, dst, dPitch);
    c->kernelLaunches += 1;
    ORBX_CUDA(c, cudaGetLastError());
    return ORBX_OK;
}

int launchPyramid(Ctx* c, int frame0, int batch)
{
    const GeomSet& g = c->geom;
    stageBegin(c, ST_PYRAMID);
    int launches = 0;
    for (int l = 1; l < g.nlevels; ++l) {
        const LevelGeom& d = g.lv[l];
        const LevelGeom& s = g.lv[l - 1];
        const uint8_t* src; size_t sp, sf;
        if (l == 1) { src = c->lvl0; sp = c->lvl0Pitch; sf = c->lvl0FrameStride; }
        else { src = c->dPyr + s.planeOff; sp = (size_t)s.pitch; sf = (size_t)s.pitch * s.h; }
        const bool wordOk = c->resizeWordOk[l] && ((((size_t)src) | sp | sf) & 3) == 0;
        // staged path: 16-byte aligned rows for the bulk copy, 8-byte aligned destination rows, enough CTAs to fill the GPU
        const int groups8 = (d.w + 7) / 8;
        int nBands = 0; size_t smem = 0;
        if (wordOk && c->resize8Ok[l] && c->dTapG[l] && c->resizeWide && ((((size_t)src) | sp | sf) & 15) == 0 && (d.pitch & 7) == 0 &&
            ((d.planeOff | (size_t)d.pitch * d.h) & 7) == 0 && groups8 <= 512) {
            const int allBands = (d.h + kStageRows - 1) / kStageRows;
            for (int nb = std::min(4, 512 / groups8); nb >= 1; --nb) {     // most bands per CTA (fewer idle lanes) that fit 64 KB
                const size_t rows = (size_t)((long long)nb * kStageRows * s.h / d.h) + 4;
                const size_t need = rows * sp + 16;
                if (need <= 64 * 1024 && (long long)batch * ((allBands + nb - 1) / nb) >= 148LL * 3) { nBands = nb; smem = need; break; }
            }
        }
        if (nBands > 0) {
            if (smem > 48 * 1024 && smem > c->smemCfg[SMEM_RESIZE]) {
                ORBX_CUDA(c, cudaFuncSetAttribute(resizeLinearStagedKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
                c->smemCfg[SMEM_RESIZE] = 64 * 1024;
            }
            const int allBands = (d.h + kStageRows - 1) / kStageRows;
            dim3 grid((allBands + nBands - 1) / nBands, 1, batch), block(((groups8 * nBands + 31) / 32) * 32);
            resizeLinearStagedKernel<<<grid, block, smem, c->stream>>>(src, sp, sf, s.w, s.h, c->dPyr + d.planeOff, (size_t)d.pitch,
                                                                      (size_t)d.pitch * d.h, d.w, d.h, c->dTapG[l], c->dTapY[l], frame0, nBands);
        } else if (wordOk) {
            const int groups = (d.w + 3) / 4;
            const bool small = (long long)batch * groups * ((d.h + 15) / 16) < 148LL * 640;     // too few threads to fill the GPU
            const int rows = small ? 2 : 16;
            const int items = groups * ((d.h + rows - 1) / rows);
            dim3 grid((items + 127) / 128, 1, batch), block(128);
            if (small)
                resizeLinearWordKernel<2><<<grid, block, 0, c->stream>>>(src, sp, sf, s.w, s.h, c->dPyr + d.planeOff, (size_t)d.pitch,
                                                                        (size_t)d.pitch * d.h, d.w, d.h, c->dTapX[l], c->dTapY[l], frame0);
            else
                resizeLinearWordKernel<16><<<grid, block, 0, c->stream>>>(src, sp, sf, s.w, s.h, c->dPyr + d.planeOff, (size_t)d.pitch,
                                                                         (size_t)d.pitch * d.h, d.w, d.h, c->dTapX[l], c->dTapY[l], frame0);
        } else {
            dim3 grid((d.w + 511) / 512, (d.h + kResizeRows - 1) / kResizeRows, batch), block(128);
            resizeLinearKernel<<<grid, block, 0, c->stream>>>(src, sp, sf, s.w, s.h, c->dPyr + d.planeOff, (size_t)d.pitch,
                                                             (size_t)d.pitch * d.h, d.w, d.h, c->dTapX[l], c->dTapY[l], frame0);
        }
        ++launches;
    }
    stageEnd(c, ST_PYRAMID, launches);
    ORBX_CUDA(c, cudaGetLastError());
    return ORBX_OK;
}

// reflect-101 frame (cv::copyMakeBorder, src/ORBextractor.cc:1704-1727): on demand only, the
// extraction kernels never read the border (FAST/IC/BRIEF stay >= 1 px inside the level).
__global__ void borderCopyKernel(const uint8_t* __restrict__ src, int w, int h, size_t pitch,
                                 uint8_t* __restrict__ dst, size_t dpitch, int border)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w + 2 * border) return;
    int sx = x - border, sy = y - border;
    if (w == 1) sx = 0; else { while (sx < 0 || sx >= w) sx = sx < 0 ? -sx : 2 * (w - 1) - sx; }
    if (h == 1) sy = 0; else { while (sy < 0 || sy >= h) sy = sy < 0 ? -sy : 2 * (h - 1) - sy; }
    dst[(size_t)y * dpitch + x] = src[(size_t)sy * pitch + sx];
}

// All levels of one frame with their reflect-101 frame in ONE launch, packed back to back (level l at off[l], row pitch
// w_l + 2 * border): what ORBextractor::mvImagePyramid holds after operator() (src/ORBextractor.cc:1674-1727).  4 output bytes
// per thread; grid.z = level.
struct BorderedLevels { const uint8_t* src[kMaxLevels]; unsigned long long pitch[kMaxLevels]; int w[kMaxLevels], h[kMaxLevels];
                        unsigned long long off[kMaxLevels]; };
__global__ void __launch_bounds__(128)
borderedPyramidKernel(const __grid_constant__ BorderedLevels L, uint8_t* __restrict__ dst, int border)
{
    const int l = blockIdx.z, w = L.w[l], h = L.h[l];
    const int ow = w + 2 * border, oh = h + 2 * border;
    const int y = blockIdx.y, x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (y >= oh || x0 >= ow) return;
    int sy = y - border;
    if (h == 1) sy = 0; else { while (sy < 0 || sy >= h) sy = sy < 0 ? -sy : 2 * (h - 1) - sy; }
    const uint8_t* S = L.src[l] + (size_t)sy * L.pitch[l];
    uint8_t* D = dst + L.off[l] + (size_t)y * ow + x0;
    #pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (x0 + i >= ow) break;
        int sx = x0 + i - border;
        if (w == 1) sx = 0; else { while (sx < 0 || sx >= w) sx = sx < 0 ? -sx : 2 * (w - 1) - sx; }
        D[i] = S[sx];
    }
}

int launchBorderedPyramid(Ctx* c, int frame, int border, uint8_t* dst, size_t* offsets, size_t* total)
{
    const GeomSet& g = c->geom;
    BorderedLevels L;
    size_t off = 0; int maxW = 0, maxH = 0;
    for (int l = 0; l < g.nlevels; ++l) {
        const LevelGeom& q = g.lv[l];
        if (l == 0) { L.src[l] = c->lvl0 + (size_t)frame * c->lvl0FrameStride; L.pitch[l] = c->lvl0Pitch; }
        else { L.src[l] = c->dPyr + q.planeOff + (size_t)frame * q.pitch * q.h; L.pitch[l] = (unsigned long long)q.pitch; }
        L.w[l] = q.w; L.h[l] = q.h; L.off[l] = off;
        if (offsets) offsets[l] = off;
        off += (size_t)(q.w + 2 * border) * (q.h + 2 * border);
        maxW = std::max(maxW, q.w + 2 * border); maxH = std::max(maxH, q.h + 2 * border);
    }
    if (total) *total = off;
    if (!dst) return ORBX_OK;
    dim3 grid((maxW + 511) / 512, maxH, g.nlevels), block(128);
    borderedPyramidKernel<<<grid, block, 0, c->stream>>>(L, dst, border);
    c->kernelLaunches++;
    ORBX_CUDA(c, cudaGetLastError());
    return ORBX_OK;
}

int launchBorderCopy(Ctx* c, const uint8_t* src, int w, int h, size_t pitch, uint8_t* dst, size_t dpitch, int border)
{
    dim3 grid((w + 2 * border + 127) / 128, h + 2 * border), block(128);
    borderCopyKernel<<<grid, block, 0, c->stream>>>(src, w, h, pitch, dst, dpitch, border);
    c->kernelLaunches++;
    ORBX_CUDA(c, cudaGetLastError());
    return ORBX_OK;
}

} // namespace orbx
