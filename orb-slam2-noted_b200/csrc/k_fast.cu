// K2 -- per-cell FAST-9/16 with the iniThFAST -> minThFAST retry
// (ComputeKeyPointsOctTree cell loop, src/ORBextractor.cc:1058-1153; cv::FAST(roi, th, nms=true)).
//
// One CTA per STRIP = one cell-row of one level of one frame, split into segments of whole cells so the tile fits in
// shared memory; all levels and frames in ONE launch; the strip table is built on the host (StripDesc, 16 bytes).
// The candidate regions of the reference's cells tile the detection window without overlap, so a strip is detected as
// one image with per-cell NMS boundaries.
//   stage   rows of the strip (+3 px apron) are brought into shared memory with 1-D bulk async copies
//           (cp.async.bulk -> UBLKCP, completion on an mbarrier); plain loads when the source is not 16-byte aligned
//   sweep 0 detects the whole strip at iniTh; sweep 1 (only if some cell kept nothing, :1128) detects those cells again
//           at minTh.  Inside a sweep every WARP owns the rows y = warp, warp + nWarps, ... and runs the detection
//           phases on them without CTA barriers (lists and their counters are private to the warp):
//     pass 0  quick rejection, 4 pixels per lane step, byte-parallel in 32-bit registers; surviving groups -> warp list
//     pass 1  full test of the listed groups: brighter/darker flags of the 16 ring pixels by SWAR unsigned-byte compares,
//             9-contiguity as bit-sliced AND/OR trees (no per-pixel branches); corners -> warp list
//     score   listed corners: exact OpenCV cornerScore (max over the 16 arcs of the arc's min margin, -1) -> u8 score map
//   one CTA barrier (the score map is complete), then
//     nms     3x3 strict-> NMS inside the pixel's own cell (neighbours outside the cell count 0, exactly like the per-ROI
//             cv::FAST calls) over the warp's own corners; survivors go straight to the global candidate records
//             (one global atomic per warp step); a cell that kept something is flagged for the retry decision
// Output records are order-independent: key = (cell, ly, lx) reproduces the reference's push_back order
// (cell-row-major, then pixel-row-major :1084-1149) whenever an ordered view is needed.
#include "orbx_internal.h"
#include <algorithm>
#include <vector>

namespace orbx {

#ifndef ORBX_P0_UNROLL
#define ORBX_P0_UNROLL 1
#endif
#ifndef ORBX_FAST_MINCTAS
#define ORBX_FAST_MINCTAS 8
#endif
constexpr int kFastMinCtas128 = ORBX_FAST_MINCTAS;   // __launch_bounds__ of the 128-thread form: 8 -> 64 registers, 6 -> 80 (shared memory allows 6 CTAs per SM)
constexpr int kP0Unroll = ORBX_P0_UNROLL;        // rows per trip of the pass-0 loop (A/B builds: tools/build_variant.sh)
constexpr int kCornerCapW = 384;                 // per-warp corner list (scored and dropped when the next step could overflow it)
constexpr int kMaxCellsPerStrip = 32;

__device__ __forceinline__ uint32_t smemAddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// per-byte unsigned a > b, result in bit 7 of each byte (other bits undefined); nb7 = ~b & 0x7f7f7f7f
__device__ __forceinline__ uint32_t gtu4(uint32_t a, uint32_t b, uint32_t nb7)
{
    const uint32_t x = (a & 0x7f7f7f7fu) + nb7;
    return (a & ~b) | (~(a ^ b) & x);
}

__device__ __forceinline__ uint32_t nineOfSixteen(const uint32_t (&f)[16])
{
    uint32_t a3[16];
    #pragma unroll
    for (int k = 0; k < 16; ++k) a3[k] = f[k] & f[(k + 1) & 15] & f[(k + 2) & 15];
    uint32_t a9[16];
    #pragma unroll
    for (int k = 0; k < 16; ++k) a9[k] = a3[k] & a3[(k + 3) & 15] & a3[(k + 6) & 15];
    // OR tree in 3-input steps
    const uint32_t t0 = a9[0] | a9[1] | a9[2], t1 = a9[3] | a9[4] | a9[5], t2 = a9[6] | a9[7] | a9[8];
    const uint32_t t3 = a9[9] | a9[10] | a9[11], t4 = a9[12] | a9[13] | a9[14];
    return (t0 | t1 | t2) | (t3 | t4 | a9[15]);
}

template <int kFastThreads, int kMinCtas>
__global__ void __launch_bounds__(kFastThreads, kMinCtas)
fastStripKernel(const __grid_constant__ GeomSet G, const uint8_t* __restrict__ lvl0, size_t lvl0Pitch, size_t lvl0Frame,
                const uint8_t* __restrict__ pyr, const StripDesc* __restrict__ strips, unsigned long long* __restrict__ cand,
                int* __restrict__ candCount, int tilePitchMax, int tileRowsMax, int scorePitchMax, int groupCapW)
{
    constexpr int nW = kFastThreads / 32;
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) unsigned long long sBar;
    __shared__ int sCellKept[kMaxCellsPerStrip];
    __shared__ int sRetryN;

    const uint4 sdw = __ldg(reinterpret_cast<const uint4*>(strips) + blockIdx.x);
    const int level = sdw.x & 0xff, nc = (sdw.x >> 8) & 0xff, ci = sdw.x >> 16;
    const int j0 = sdw.y & 0xffff, cx0 = sdw.y >> 16;
    const int cy0 = sdw.z & 0xffff, rw = sdw.z >> 16;
    const int rh = sdw.w & 0xffff;
    const uint32_t rcpW = sdw.w >> 16;                               // (lx * rcpW) >> 16 == lx / wCell for lx < 1024
    const int frame = blockIdx.y + G.frame0;
    const LevelGeom& g = G.lv[level];

    const uint8_t* img; size_t pitch;
    if (level == 0) { img = lvl0 + (size_t)frame * lvl0Frame; pitch = lvl0Pitch; }
    else { img = pyr + g.planeOff + (size_t)frame * g.pitch * g.h; pitch = (size_t)g.pitch; }

    const int TP = tilePitchMax, SP = scorePitchMax;               // bytes per tile / score row
    const int maxRh = tileRowsMax - 6;
    const int wp = TP >> 2;
    uint8_t* tile = smem;                                           // (rh + 6) rows
    uint8_t* score = smem + (size_t)TP * tileRowsMax;               // (rh + 2) x SP, zero apron
    uint32_t* colValid = reinterpret_cast<uint32_t*>(score + (size_t)SP * (maxRh + 2)); // per tile word column: bit 7 of the bytes to detect
    uint16_t* cornerLists = reinterpret_cast<uint16_t*>(colValid + wp);                 // nW x kCornerCapW
    uint16_t* groupLists = cornerLists + nW * kCornerCapW;          // nW x groupCapW
    uint8_t* colCell = reinterpret_cast<uint8_t*>(groupLists + nW * groupCapW);          // cell column of candidate column lx
    uint16_t* retryCols = reinterpret_cast<uint16_t*>(colCell + SP);                    // sweep 1: word columns with pixels to detect
    const int TH = rh + 6;
    const bool aligned = ((((size_t)img) | pitch) & 15) == 0;
    const int xa = aligned ? ((cx0 - 4) & ~15) : (cx0 - 4) & ~3;    // image column of tile column 0 (>= 4 px left of cx0)
    const int off0 = cx0 - xa;                                      // tile column of the first candidate pixel (>= 4)
    const int cx1 = cx0 + rw;
    const int wCell = g.wCell;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t ltMask = (1u << lane) - 1u;
    asm volatile("" : "+r"(ltMask));                          // keep it in a register (ptxas would recompute it per use)

    if (threadIdx.x < kMaxCellsPerStrip) sCellKept[threadIdx.x] = 0;
    if (aligned) {
        const int xb = min((cx1 + 3 + 15) & ~15, (int)pitch);
        const uint32_t rowBytes = (uint32_t)(xb - xa);
        const uint32_t bar = smemAddr(&sBar);
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
        if (threadIdx.x == 0)
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(rowBytes * (uint32_t)TH) : "memory");
        if ((int)threadIdx.x < TH) {
            const uint8_t* src = img + (size_t)(cy0 - 3 + threadIdx.x) * pitch + xa;
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(smemAddr(tile + (size_t)threadIdx.x * TP)), "l"(src), "r"(rowBytes), "r"(bar) : "memory");
        }
    } else {
        const int tw = min(cx1 + 3, (int)g.w) - xa;
        for (int i = threadIdx.x; i < TH * tw; i += kFastThreads) {
            const int y = i / tw, x = i - y * tw;
            tile[y * TP + x] = img[(size_t)(cy0 - 3 + y) * pitch + xa + x];
        }
    }
    // zero the score map, build the column tables while the copies are in flight
    {
        const int nbytes = SP * (rh + 2);                           // SP is a multiple of 4, the map starts 16-byte aligned
        uint4* z = reinterpret_cast<uint4*>(score);
        for (int i = threadIdx.x; i < (nbytes >> 4); i += kFastThreads) z[i] = make_uint4(0, 0, 0, 0);
        if ((int)threadIdx.x < ((nbytes & 15) >> 2)) reinterpret_cast<uint32_t*>(score)[(nbytes >> 4) * 4 + threadIdx.x] = 0;
        for (int i = threadIdx.x; i < rw; i += kFastThreads) colCell[i] = (uint8_t)(((uint32_t)i * rcpW) >> 16);
        for (int i = threadIdx.x; i < wp; i += kFastThreads) {
            const int lx0 = i * 4 - off0;
            uint32_t m = 0;
            #pragma unroll
            for (int b = 0; b < 4; ++b) if (lx0 + b >= 0 && lx0 + b < rw) m |= 0x80u << (8 * b);
            colValid[i] = m;
        }
    }
    if (aligned) {
        const uint32_t bar = smemAddr(&sBar);
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(bar) : "memory");
    }
    __syncthreads();

    const int tmin = G.minTh, tini = G.iniTh;
    const int g0 = off0 >> 2, g1 = (off0 + rw + 3) >> 2;            // word columns holding candidate pixels
    const uint32_t* W = reinterpret_cast<const uint32_t*>(tile);
    const int o[16] = {3 * TP, 3 * TP + 1, 2 * TP + 2, TP + 3, 3, -TP + 3, -2 * TP + 2, -3 * TP + 1,
                       -3 * TP, -3 * TP - 1, -2 * TP - 2, -TP - 3, -3, TP - 3, 2 * TP - 2, 3 * TP - 1};
    uint32_t myGoff = (uint32_t)(warp * groupCapW), myCoff = (uint32_t)(warp * kCornerCapW);
    asm volatile("" : "+r"(myGoff), "+r"(myCoff));
    uint16_t* myG = groupLists + myGoff;
    uint16_t* myC = cornerLists + myCoff;

    // exact score of the warp's listed corners
    auto scoreWarp = [&](int n) {
        for (int i = lane; i < n; i += 32) {
            const int code = myC[i], ly = code >> 10, lx = code & 1023;
            const uint8_t* c = tile + (ly + 3) * TP + (lx + off0);
            const int v = c[0];
            int d[16];
            #pragma unroll
            for (int k = 0; k < 16; ++k) d[k] = v - (int)c[o[k]];
            int mn2[16], mx2[16];
            #pragma unroll
            for (int k = 0; k < 16; ++k) { mn2[k] = min(d[k], d[(k + 1) & 15]); mx2[k] = max(d[k], d[(k + 1) & 15]); }
            int mn4[16], mx4[16];
            #pragma unroll
            for (int k = 0; k < 16; ++k) { mn4[k] = min(mn2[k], mn2[(k + 2) & 15]); mx4[k] = max(mx2[k], mx2[(k + 2) & 15]); }
            int a0 = -255, b0 = 255;
            #pragma unroll
            for (int k = 0; k < 16; ++k) {
                const int mn9 = min(min(mn4[k], mn4[(k + 4) & 15]), d[(k + 8) & 15]);
                const int mx9 = max(max(mx4[k], mx4[(k + 4) & 15]), d[(k + 8) & 15]);
                a0 = max(a0, mn9);
                b0 = min(b0, mx9);
            }
            score[(ly + 1) * SP + (lx + 1)] = (uint8_t)(max(a0, -b0) - 1); // > th-1 >= 0 for a detected corner
        }
    };

    unsigned long long* out = cand + (size_t)frame * G.candPerFrame + g.candBase;
    int* outCount = candCount + frame * G.nlevels + level;
    const uint32_t cellKey0 = (uint32_t)(ci * g.nCols + j0);

    for (int sweep = 0; sweep < 2; ++sweep) {
        const int th = sweep ? tmin : tini;
        uint32_t retry = 0xffffffffu;
        if (sweep) {
            // the retry set (:1128): cells of this strip that kept nothing at iniTh
            __syncthreads();                                        // every warp's survivors of sweep 0 are flagged
            retry = __ballot_sync(0xffffffffu, lane < nc && sCellKept[lane] == 0);
            if (!retry || tmin == tini) break;
            for (int i = threadIdx.x; i < wp; i += kFastThreads) {
                const int lx0 = i * 4 - off0;
                uint32_t m = 0;
                #pragma unroll
                for (int b = 0; b < 4; ++b) {
                    const int lx = lx0 + b;
                    if (lx >= 0 && lx < rw && ((retry >> colCell[lx]) & 1u)) m |= 0x80u << (8 * b);
                }
                colValid[i] = m;
            }
            __syncthreads();
            if (warp == 0) {                                        // ordered list of the word columns left to detect
                int n = 0;
                for (int w0 = g0; w0 < g1; w0 += 32) {
                    const int w = w0 + lane;
                    const bool on = w < g1 && colValid[w] != 0;
                    const uint32_t m = __ballot_sync(0xffffffffu, on);
                    if (on) retryCols[n + __popc(m & ltMask)] = (uint16_t)w;
                    n += __popc(m);
                }
                if (lane == 0) sRetryN = n;
            }
            if (tmin > tini) {   // scores in [iniTh, minTh) are not corners of this sweep
                uint32_t* z = reinterpret_cast<uint32_t*>(score);
                const int nz = (SP * (rh + 2) + 3) / 4;
                for (int i = threadIdx.x; i < nz; i += kFastThreads) z[i] = 0;
            }
            __syncthreads();
        }
        const uint32_t T4 = (uint32_t)th * 0x01010101u, nT7 = ~T4 & 0x7f7f7f7fu;

        // ---- pass 0: quick rejection, 4 pixels per lane step.  A 9-arc contains two adjacent compass pixels, i.e. one
        //      of {0, 8} and one of {4, 12}, both differing from the centre by more than the threshold. ----
        int gcnt = 0;
        // one lane step: the 4 pixels of tile word *c (row y, word column in the low bits of code)
        auto quick = [&](const uint32_t* c, const uint32_t vm, const uint32_t code) {
            const uint32_t v = c[0], up = c[-3 * wp], dn = c[3 * wp];
            const uint32_t p4 = __funnelshift_r(v, c[1], 24), p12 = __funnelshift_r(c[-1], v, 8);
            // |p - v| per byte (VABSDIFF4); OR-ing two difference words can only over-estimate their maximum, and
            // the sign is ignored: a superset of the exact quick test (pass 1 is exact)
            const uint32_t dv = __vabsdiffu4(dn, v) | __vabsdiffu4(up, v), dh = __vabsdiffu4(p4, v) | __vabsdiffu4(p12, v);
            const bool pass = (gtu4(dv, T4, nT7) & gtu4(dh, T4, nT7) & vm) != 0;
            const uint32_t m = __ballot_sync(0xffffffffu, pass);
            if (pass) myG[gcnt + __popc(m & ltMask)] = (uint16_t)code;
            gcnt += __popc(m);
        };
        if (!sweep) {
            // column chunks of 32 words, rows inside: a lane keeps its column (and that column's mask) for the whole chunk.
            // A last chunk of <= 16 columns is laid over the warp q = 32 / width times, each copy taking every q-th row.
            for (int k0 = g0; k0 < g1; k0 += 32) {
                const int wk = min(32, g1 - k0);
                const int q = min(32 / wk, 8), sub = lane / wk;    // q >= 1; lanes with sub >= q idle
                const uint32_t col = (uint32_t)(k0 + lane - sub * wk);
                const uint32_t vm = sub < q ? colValid[col] : 0u;
                const int yStep = nW * q;
                int y = warp + nW * sub;
                const uint32_t* const cSafe = W + 3 * wp + col;     // lanes past the last row read row 0 (masked by vm = 0)
                const uint32_t* c = W + (y + 3) * wp + col;
                uint32_t code = (uint32_t)(y << 10) | col;
                #pragma unroll kP0Unroll
                for (int yl = warp; yl < rh; yl += yStep, y += yStep, c += yStep * wp, code += (uint32_t)(yStep << 10))
                    quick(y < rh ? c : cSafe, y < rh ? vm : 0u, code);
            }
        } else {
            // retry sweep: only the word columns of the cells that kept nothing (about a third of a strip's columns)
            const int nR = sRetryN;
            for (int k0 = 0; k0 < nR; k0 += 32) {
                const bool on = k0 + lane < nR;
                const uint32_t col = on ? retryCols[k0 + lane] : (uint32_t)g0;
                const uint32_t vm = on ? colValid[col] : 0u;
                const uint32_t* c = W + 3 * wp + col;
                #pragma unroll 1
                for (int y = warp; y < rh; y += nW) quick(c + y * wp, vm, (uint32_t)(y << 10) | col);
            }
        }
        __syncwarp();

        // ---- pass 1: full test of the warp's groups; corners go to the warp's list, scored in bulk ----
        int ccnt = 0, flushes = 0;
        for (int ib = 0; ib < gcnt; ib += 32) {
            const int i = ib + lane;
            uint32_t nib = 0; int y = 0, lx0 = 0;
            if (i < gcnt) {
                const int code = myG[i], gw = code & 1023;
                y = code >> 10; lx0 = gw * 4 - off0;
                const uint32_t* c = W + (y + 3) * wp + gw;
                uint32_t a[7][3];
                #pragma unroll
                for (int r = 0; r < 7; ++r) { a[r][0] = c[(r - 3) * wp - 1]; a[r][1] = c[(r - 3) * wp]; a[r][2] = c[(r - 3) * wp + 1]; }
                const uint32_t v = a[3][1];
                const uint32_t hi = __vaddus4(v, T4), lo = __vsubus4(v, T4);    // min(v + t, 255), max(v - t, 0) per byte
                const uint32_t nh7 = ~hi & 0x7f7f7f7fu, cl = (lo & 0x7f7f7f7fu) + 0x7f7f7f7fu;
                uint32_t p[16];
                p[0] = a[6][1];                                   // ( 0, 3)
                p[1] = __funnelshift_r(a[6][1], a[6][2], 8);      // ( 1, 3)
                p[2] = __funnelshift_r(a[5][1], a[5][2], 16);     // ( 2, 2)
                p[3] = __funnelshift_r(a[4][1], a[4][2], 24);     // ( 3, 1)
                p[4] = __funnelshift_r(a[3][1], a[3][2], 24);     // ( 3, 0)
                p[5] = __funnelshift_r(a[2][1], a[2][2], 24);     // ( 3,-1)
                p[6] = __funnelshift_r(a[1][1], a[1][2], 16);     // ( 2,-2)
                p[7] = __funnelshift_r(a[0][1], a[0][2], 8);      // ( 1,-3)
                p[8] = a[0][1];                                   // ( 0,-3)
                p[9] = __funnelshift_r(a[0][0], a[0][1], 24);     // (-1,-3)
                p[10] = __funnelshift_r(a[1][0], a[1][1], 16);    // (-2,-2)
                p[11] = __funnelshift_r(a[2][0], a[2][1], 8);     // (-3,-1)
                p[12] = __funnelshift_r(a[3][0], a[3][1], 8);     // (-3, 0)
                p[13] = __funnelshift_r(a[4][0], a[4][1], 8);     // (-3, 1)
                p[14] = __funnelshift_r(a[5][0], a[5][1], 16);    // (-2, 2)
                p[15] = __funnelshift_r(a[6][0], a[6][1], 24);    // (-1, 3)
                // brighter: p > hi, darker: p < lo, per byte in bit 7; the two compares share p & 0x7f7f7f7f:
                //   p > hi : carry of (p & 7f) + (~hi & 7f) decides when the top bits agree
                //   p < lo : carry of (lo & 7f) + 7f - (p & 7f) likewise (no borrow between bytes)
                uint32_t fb[16], fd[16];
                #pragma unroll
                for (int k = 0; k < 16; ++k) {
                    const uint32_t p7 = p[k] & 0x7f7f7f7fu;
                    const uint32_t xb = p7 + nh7, xd = cl - p7;
                    fb[k] = (p[k] & ~hi) | (~(p[k] ^ hi) & xb);
                    fd[k] = (lo & ~p[k]) | (~(p[k] ^ lo) & xd);
                }
                const uint32_t corner = (nineOfSixteen(fb) | nineOfSixteen(fd)) & colValid[gw];
                nib = ((corner >> 7) & 1u) | ((corner >> 14) & 2u) | ((corner >> 21) & 4u) | ((corner >> 28) & 8u);
            }
            if (__any_sync(0xffffffffu, nib != 0)) {
                #pragma unroll
                for (int b = 0; b < 4; ++b) {                       // list order is free: bit plane by bit plane
                    const bool on = (nib >> b) & 1u;
                    const uint32_t m = __ballot_sync(0xffffffffu, on);
                    if (on) myC[ccnt + __popc(m & ltMask)] = (uint16_t)((y << 10) | (lx0 + b));
                    ccnt += __popc(m);
                }
                if (ccnt > kCornerCapW - 128) {                     // the next step could overflow the list
                    __syncwarp();
                    scoreWarp(ccnt);
                    __syncwarp();
                    ccnt = 0; ++flushes;
                }
            }
        }
        __syncwarp();
        scoreWarp(ccnt);
        __syncthreads();                                            // the score map of this sweep is complete

        // ---- nms (strict >, 8 neighbours, zero outside the pixel's own cell) + emit ----
        auto nmsEmit = [&](bool have, int ly, int lx) {
            bool keep = false; int s = 0, cj = 0, cx = 0;
            if (have) {
                const uint8_t* r = score + (ly + 1) * SP + (lx + 1);
                s = r[0];
                cj = colCell[lx]; cx = lx - cj * wCell;
                keep = s > r[-SP] && s > r[SP];
                if (cx > 0) keep = keep && s > r[-1] && s > r[-SP - 1] && s > r[SP - 1];
                if (cx < wCell - 1) keep = keep && s > r[1] && s > r[-SP + 1] && s > r[SP + 1];
            }
            const uint32_t m = __ballot_sync(0xffffffffu, keep);
            if (!m) return;
            int base = 0;
            if (lane == 0) base = atomicAdd(outCount, __popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (keep) {
                if (!sweep) sCellKept[cj] = 1;
                const uint32_t key = ((cellKey0 + (uint32_t)cj) << 12) | (uint32_t)(ly << 6) | (uint32_t)cx;
                out[base + __popc(m & ltMask)] = ((unsigned long long)((uint32_t)s << 24) << 32) | key; // hi = score<<24 | node(0), lo = key
            }
        };
        if (!flushes) {
            for (int ib = 0; ib < ccnt; ib += 32) {
                const int i = ib + lane;
                const int code = i < ccnt ? myC[i] : 0;
                nmsEmit(i < ccnt, code >> 10, code & 1023);
            }
        } else {
            // the list was dropped on the way: walk the score map of the warp's rows (dense frames only)
            const int nWords = (rw + 1 + 3) >> 2;
            for (int y = warp; y < rh; y += nW) {
                const uint32_t* srow = reinterpret_cast<const uint32_t*>(score + (y + 1) * SP);
                for (int wb = 0; wb < nWords; wb += 32) {
                    const int wi = wb + lane;
                    const uint32_t word = wi < nWords ? srow[wi] : 0;
                    if (!__any_sync(0xffffffffu, word != 0)) continue;
                    #pragma unroll
                    for (int b = 0; b < 4; ++b) {
                        const int lx = wi * 4 + b - 1;
                        bool have = ((word >> (8 * b)) & 0xffu) >= (uint32_t)th && lx >= 0 && lx < rw;
                        if (have && sweep) have = (retry >> colCell[lx]) & 1u;
                        nmsEmit(have, y, lx);
                    }
                }
            }
        }
    }
}

// Host side of the strip table: one entry per non-empty strip of a frame (skipped cells :1094,:1112 have an empty
// candidate region and get no CTA).
void buildFastStrips(const GeomSet& g, std::vector<StripDesc>& out)
{
    out.clear();
    for (int l = 0; l < g.nlevels; ++l) {
        const LevelGeom& q = g.lv[l];
        for (int ci = 0; ci < q.nRows; ++ci)
            for (int seg = 0; seg < q.segsPerRow; ++seg) {
                const int j0 = seg * q.cellsPerSeg, j1 = std::min(j0 + q.cellsPerSeg, q.nCols);
                const int cx0 = kEdge + j0 * q.wCell, cy0 = kEdge + ci * q.hCell;
                const int cx1 = std::min(kEdge + j1 * q.wCell, q.w - kEdge), cy1 = std::min(cy0 + q.hCell, q.h - kEdge);
                const int rw = cx1 - cx0, rh = cy1 - cy0;
                if (rw <= 0 || rh <= 0 || j1 <= j0) continue;
                const int nc = (rw + q.wCell - 1) / q.wCell;         // cells that own at least one candidate column
                StripDesc d;
                d.a = (uint32_t)l | ((uint32_t)nc << 8) | ((uint32_t)ci << 16);
                d.b = (uint32_t)j0 | ((uint32_t)cx0 << 16);
                d.c = (uint32_t)cy0 | ((uint32_t)rw << 16);
                d.d = (uint32_t)rh | ((65536u / (uint32_t)q.wCell + 1u) << 16);
                out.push_back(d);
            }
    }
}

int launchFast(Ctx* c, int frame0, int batch)
{
    GeomSet g = c->geom;
    g.frame0 = frame0;
    stageBegin(c, ST_FAST);
    ORBX_CUDA(c, cudaMemsetAsync(c->dCandCount + (size_t)frame0 * g.nlevels, 0, sizeof(int) * (size_t)batch * g.nlevels, c->stream));
    int launches = 0;
    if (c->nStrips > 0) {
        const int TP = g.fastTilePitch, TR = g.fastTileRows, SP = g.fastScorePitch;
        const int maxRh = TR - 6;
        const int fastThreads = (long long)batch * c->nStrips < 148 * 8 ? 256 : c->fastThreads;   // few CTAs: more warps per strip
        const int nW = fastThreads / 32;
        const int groupCapW = (((maxRh + nW - 1) / nW) * ((SP >> 2) + 2) + 1) & ~1;   // rows of a warp x group columns of a row
        size_t smem = (size_t)TP * TR + (size_t)SP * (maxRh + 2) + sizeof(uint32_t) * (size_t)(TP >> 2);
        smem += sizeof(uint16_t) * (size_t)nW * (kCornerCapW + groupCapW) + (size_t)SP + sizeof(uint16_t) * (size_t)(TP >> 2);
        smem = (smem + 127) & ~(size_t)127;
        dim3 grid(c->nStrips, batch);
        auto go = [&](auto kern, int threads, size_t* configured) -> int {
            if (smem > 48 * 1024 && smem > *configured) {
                ORBX_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                *configured = smem;
            }
            kern<<<grid, dim3(threads), smem, c->stream>>>(g, c->lvl0, c->lvl0Pitch, c->lvl0FrameStride, c->dPyr, c->dStrips,
                                                           c->dCand, c->dCandCount, TP, TR, SP, groupCapW);
            return ORBX_OK;
        };
        const int rc = fastThreads == 128 ? go(fastStripKernel<128, kFastMinCtas128>, 128, &c->smemCfg[SMEM_FAST128])
                                             : go(fastStripKernel<256, 4>, 256, &c->smemCfg[SMEM_FAST256]);
        if (rc != ORBX_OK) return rc;
        launches = 1;
    }
    stageEnd(c, ST_FAST, launches);
    ORBX_CUDA(c, cudaGetLastError());
    return ORBX_OK;
}

} // namespace orbx
