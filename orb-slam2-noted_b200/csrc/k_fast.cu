// K2 -- per-cell FAST-9/16 with the iniThFAST -> minThFAST retry
// (ComputeKeyPointsOctTree cell loop, src/ORBextractor.cc:1058-1153; cv::FAST(roi, th, nms=true)).
//
// One CTA per STRIP = one cell-row of one level of one frame, split into segments of whole cells so the tile fits in
// shared memory; all levels and frames in ONE launch.  The candidate regions of the reference's cells tile the
// detection window without overlap, so a strip is detected as one image with per-cell NMS boundaries.
//   stage   rows of the strip (+3 px apron) are brought into shared memory with 1-D bulk async copies
//           (cp.async.bulk -> UBLKCP, completion on an mbarrier); plain loads when the source is not 16-byte aligned
//   sweep 0 detects the whole strip at iniTh; sweep 1 (only if some cell kept nothing, :1128) detects those cells again
//           at minTh.  Each sweep:
//     pass 0  quick rejection, 4 pixels per thread step, byte-parallel in 32-bit registers; surviving groups -> list
//     pass 1  full test of the listed groups: brighter/darker flags of the 16 ring pixels by SWAR unsigned-byte compares
//             (3 LOP3/IADD per flag word), 9-contiguity as bit-sliced AND/OR trees (no per-pixel branches); corners go
//             to a list and a bitmap
//     score   list entries: exact OpenCV cornerScore (max over the 16 arcs of the arc's min margin, -1) -> u8 score map
//     nms     3x3 strict-> NMS inside the pixel's own cell (neighbours outside the cell count 0, exactly like the per-ROI
//             cv::FAST calls); survivors -> compact list, per-cell survivor counts (the retry decision)
//     emit    survivors -> global candidate records
// Output records are order-independent: key = (cell, ly, lx) reproduces the reference's push_back order
// (cell-row-major, then pixel-row-major :1084-1149) whenever an ordered view is needed.
#include "orbx_internal.h"
#include <algorithm>

namespace orbx {

constexpr int kCornerCap = 4096;                 // corner list capacity (scored in bulk whenever the next batch could overflow it)
constexpr int kGroupBatch = 512;                 // 4-pixel groups per full-test batch (<= 2048 new corners)
constexpr int kMaxCellsPerStrip = 32;

__device__ __forceinline__ uint32_t smemAddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// per-byte unsigned a > b, result in bit 7 of each byte (other bits undefined); nb7 = ~b & 0x7f7f7f7f
__device__ __forceinline__ uint32_t gtu4(uint32_t a, uint32_t b, uint32_t nb7)
{
    const uint32_t x = (a & 0x7f7f7f7fu) + nb7;
    return (a & ~b) | (~(a ^ b) & x);
}
// per-byte unsigned a < b; b7 = b & 0x7f7f7f7f
__device__ __forceinline__ uint32_t ltu4(uint32_t a, uint32_t b, uint32_t b7)
{
    const uint32_t x = b7 + (~a & 0x7f7f7f7fu);
    return (b & ~a) | (~(a ^ b) & x);
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

// warp-aggregated append of `cnt` items per lane; returns the lane's first slot
__device__ __forceinline__ int warpAppend(int* counter, int cnt)
{
    const int lane = threadIdx.x & 31;
    int incl = cnt;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    int base = 0;
    if (lane == 31) base = atomicAdd(counter, total);
    base = __shfl_sync(0xffffffffu, base, 31);
    return base + incl - cnt;
}

// upper bound of NMS survivors in a strip: no two 8-neighbours inside one cell both survive
__host__ __device__ inline int fastSurvivorCap(int maxRw, int maxRh) { return ((maxRw + kMaxCellsPerStrip) / 2 + 1) * ((maxRh + 1) / 2); }

template <int kFastThreads, int kMinCtas>
__global__ void __launch_bounds__(kFastThreads, kMinCtas)
fastStripKernel(const __grid_constant__ GeomSet G, const uint8_t* __restrict__ lvl0, size_t lvl0Pitch, size_t lvl0Frame,
                const uint8_t* __restrict__ pyr, unsigned long long* __restrict__ cand, int* __restrict__ candCount,
                int tilePitchMax, int tileRowsMax, int scorePitchMax, int listBytes, uint32_t rcpBP4)
{
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) unsigned long long sBar;
    __shared__ int sGroupN, sCornerN, sSurvN, sBase;
    __shared__ uint32_t sRetry;
    __shared__ int sRegG0[kMaxCellsPerStrip], sRegNg[kMaxCellsPerStrip], sNgMax;   // pass-0 regions: first group column, group columns
    __shared__ int sCellCnt[kMaxCellsPerStrip];

    const int frame = blockIdx.y + G.frame0;
    int level = 0;
    while (level + 1 < G.nlevels && (int)blockIdx.x >= G.lv[level + 1].stripBase) ++level;
    const LevelGeom& g = G.lv[level];
    const int sIdx = blockIdx.x - g.stripBase;
    const int ci = sIdx / g.segsPerRow, seg = sIdx - ci * g.segsPerRow;
    const int j0 = seg * g.cellsPerSeg, j1 = min(j0 + g.cellsPerSeg, g.nCols);
    const int cx0 = kEdge + j0 * g.wCell, cy0 = kEdge + ci * g.hCell;
    const int cx1 = min(kEdge + j1 * g.wCell, g.w - kEdge), cy1 = min(cy0 + g.hCell, g.h - kEdge);
    const int rw = cx1 - cx0, rh = cy1 - cy0;
    if (rw <= 0 || rh <= 0) return; // skipped cells (:1094,:1112) have an empty candidate region

    const uint8_t* img; size_t pitch;
    if (level == 0) { img = lvl0 + (size_t)frame * lvl0Frame; pitch = lvl0Pitch; }
    else { img = pyr + g.planeOff + (size_t)frame * g.pitch * g.h; pitch = (size_t)g.pitch; }

    const int TP = tilePitchMax, SP = scorePitchMax;               // bytes per tile / score row
    const int BP = (SP + 31) >> 5;                                  // bitmap words per row
    const int maxRh = tileRowsMax - 6;
    const int wp = TP >> 2;
    uint8_t* tile = smem;                                           // (rh + 6) rows
    uint8_t* score = smem + (size_t)TP * tileRowsMax;               // (rh + 2) x SP, zero apron
    uint32_t* cornerBits = reinterpret_cast<uint32_t*>(score + (size_t)SP * (maxRh + 2)); // rh x BP
    uint32_t* colValid = cornerBits + maxRh * BP;                   // per tile word column: bit 7 of the bytes to detect
    uint16_t* cornerList = reinterpret_cast<uint16_t*>(colValid + wp);                    // kCornerCap
    uint16_t* groupList = cornerList + kCornerCap;                  // up to maxRh * (SP/4 + 2) groups
    uint32_t* survList = reinterpret_cast<uint32_t*>(groupList);    // reuses the group list once pass 1 is over
    uint8_t* colCell = reinterpret_cast<uint8_t*>(cornerList) + listBytes;                // cell column of candidate column lx
    const int TH = rh + 6;
    const bool aligned = ((((size_t)img) | pitch) & 15) == 0;
    const int xa = aligned ? ((cx0 - 4) & ~15) : (cx0 - 4) & ~3;    // image column of tile column 0 (>= 4 px left of cx0)
    const int off0 = cx0 - xa;                                      // tile column of the first candidate pixel (>= 4)
    const int nc = j1 - j0;

    if (threadIdx.x == 0) {
        sGroupN = 0; sCornerN = 0; sSurvN = 0;
    }
    if (threadIdx.x < kMaxCellsPerStrip) sCellCnt[threadIdx.x] = 0;
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
    // zero the score map and the bitmap, build the column tables while the copies are in flight
    {
        uint32_t* z = reinterpret_cast<uint32_t*>(score);
        const int nz = (SP * (rh + 2) + 3) / 4;
        for (int i = threadIdx.x; i < nz; i += kFastThreads) z[i] = 0;
        for (int i = threadIdx.x; i < maxRh * BP; i += kFastThreads) cornerBits[i] = 0;
        for (int i = threadIdx.x; i < rw; i += kFastThreads) colCell[i] = (uint8_t)(i / g.wCell);
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
    const int lane = threadIdx.x & 31;
    const int g0 = off0 >> 2, g1 = (off0 + rw + 3) >> 2;            // word columns holding candidate pixels
    const int ng = g1 - g0;
    const int wCell = g.wCell;
    const uint32_t* W = reinterpret_cast<const uint32_t*>(tile);
    const int o[16] = {3 * TP, 3 * TP + 1, 2 * TP + 2, TP + 3, 3, -TP + 3, -2 * TP + 2, -3 * TP + 1,
                       -3 * TP, -3 * TP - 1, -2 * TP - 2, -TP - 3, -3, TP - 3, 2 * TP - 2, 3 * TP - 1};

    // exact score of the listed corners
    auto scorePass = [&](int n) {
        for (int i = threadIdx.x; i < n; i += kFastThreads) {
            const int code = cornerList[i], ly = code >> 10, lx = code & 1023;
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
    const uint8_t* cornerBytes = reinterpret_cast<const uint8_t*>(cornerBits);

    for (int sweep = 0; sweep < 2; ++sweep) {
        const int th = sweep ? tmin : tini;
        int nReg = 1, ngMax = ng;                                   // sweep 0: one region, the whole strip
        if (sweep) {
            // the retry set (:1128): cells of this strip that kept nothing at iniTh; each becomes a pass-0 region
            if (threadIdx.x < 32) {
                const int j = threadIdx.x;
                const bool mine = j < nc && sCellCnt[j] == 0;
                const uint32_t m = __ballot_sync(0xffffffffu, mine);
                int gFirst = (off0 + j * wCell) >> 2;
                const int gLast = (off0 + min((j + 1) * wCell, rw) - 1) >> 2;
                // a group column shared with a retry cell on the left is enumerated by that cell
                if (j > 0 && ((m >> (j - 1)) & 1u) && ((off0 + j * wCell - 1) >> 2) == gFirst) ++gFirst;
                const int n = mine ? gLast - gFirst + 1 : 0;
                const int mx = __reduce_max_sync(0xffffffffu, n);
                if (mine) { const int slot = __popc(m & ((1u << j) - 1u)); sRegG0[slot] = gFirst; sRegNg[slot] = n; }
                if (j == 0) { sRetry = m; sNgMax = mx; sGroupN = 0; sCornerN = 0; sSurvN = 0; }
            }
            __syncthreads();
            const uint32_t retry = sRetry;
            if (!retry || tmin == tini) break;
            nReg = __popc(retry); ngMax = max(sNgMax, 1);
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
            for (int i = threadIdx.x; i < rh * BP; i += kFastThreads) cornerBits[i] = 0;
            if (tmin > tini) {   // scores in [iniTh, minTh) are not corners of this sweep
                uint32_t* z = reinterpret_cast<uint32_t*>(score);
                const int nz = (SP * (rh + 2) + 3) / 4;
                for (int i = threadIdx.x; i < nz; i += kFastThreads) z[i] = 0;
            }
            __syncthreads();
        }
        const uint32_t T4 = (uint32_t)th * 0x01010101u, nT7 = ~T4 & 0x7f7f7f7fu;

        // ---- pass 0: quick rejection, 4 pixels per step.  A 9-arc contains two adjacent compass pixels, i.e. one of
        //      {0, 8} and one of {4, 12}, both differing from the centre by more than the threshold. ----
        auto quick = [&](const uint32_t* c, uint32_t vm, int y, int gw) {
            const uint32_t v = c[0], up = c[-3 * wp], dn = c[3 * wp];
            const uint32_t p4 = __funnelshift_r(v, c[1], 24), p12 = __funnelshift_r(c[-1], v, 8);
            // |p - v| per byte (VABSDIFF4); OR-ing two difference words can only over-estimate their maximum, and the
            // sign is ignored: a superset of the exact quick test at 13 instructions (pass 1 is exact)
            const uint32_t dv = __vabsdiffu4(dn, v) | __vabsdiffu4(up, v), dh = __vabsdiffu4(p4, v) | __vabsdiffu4(p12, v);
            if (gtu4(dv, T4, nT7) & gtu4(dh, T4, nT7) & vm) groupList[atomicAdd(&sGroupN, 1)] = (uint16_t)((y << 10) | gw);   // aggregated per warp by the compiler
        };
        if (!sweep) {
            // whole strip: a warp per row, a lane per group column (no index arithmetic)
            for (int y = threadIdx.x >> 5; y < rh; y += kFastThreads / 32) {
                const uint32_t* crow = W + (y + 3) * wp;
                for (int gw = g0 + lane; gw < g1; gw += 32) quick(crow + gw, colValid[gw], y, gw);
            }
        } else {
            // retry cells only: flat enumeration of (region, row, group column)
            const int perReg = rh * ngMax;
            const uint32_t rcpPer = 0xffffffffu / (uint32_t)perReg + 1u, rcpNgMax = 0xffffffffu / (uint32_t)ngMax + 1u;   // exact below 65536
            for (int i = threadIdx.x; i < nReg * perReg; i += kFastThreads) {
                const int reg = (int)__umulhi((uint32_t)i, rcpPer), rem = i - reg * perReg;
                const int y = (int)__umulhi((uint32_t)rem, rcpNgMax), k = rem - y * ngMax;
                if (k >= sRegNg[reg]) continue;
                const int gw = sRegG0[reg] + k;
                const uint32_t vm = colValid[gw];
                if (vm) quick(W + (y + 3) * wp + gw, vm, y, gw);
            }
        }
        __syncthreads();
        const int nG = sGroupN;

        // ---- pass 1: full test of the surviving groups in batches; corners go to the list and the corner bitmap ----
        int flushes = 0, lastN = 0;
        for (int gb = 0; gb < nG; gb += kGroupBatch) {
            const int gEnd = min(nG, gb + kGroupBatch);
            for (int ib = gb; ib < gEnd; ib += kFastThreads) {
                const int i = ib + threadIdx.x;
                uint32_t nib = 0; int y = 0, lx0 = 0;
                if (i < gEnd) {
                    const int code = groupList[i], gw = code & 1023;
                    y = code >> 10; lx0 = gw * 4 - off0;
                    const uint32_t* c = W + (y + 3) * wp + gw;
                    uint32_t a[7][3];
                    #pragma unroll
                    for (int r = 0; r < 7; ++r) { a[r][0] = c[(r - 3) * wp - 1]; a[r][1] = c[(r - 3) * wp]; a[r][2] = c[(r - 3) * wp + 1]; }
                    const uint32_t v = a[3][1];
                    const uint32_t hi = __vaddus4(v, T4), lo = __vsubus4(v, T4);    // min(v + t, 255), max(v - t, 0) per byte
                    const uint32_t nh7 = ~hi & 0x7f7f7f7fu, lo7 = lo & 0x7f7f7f7fu;
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
                    uint32_t f[16];
                    #pragma unroll
                    for (int k = 0; k < 16; ++k) f[k] = gtu4(p[k], hi, nh7);
                    uint32_t corner = nineOfSixteen(f);
                    #pragma unroll
                    for (int k = 0; k < 16; ++k) f[k] = ltu4(p[k], lo, lo7);
                    corner = (corner | nineOfSixteen(f)) & colValid[gw];
                    nib = ((corner >> 7) & 1u) | ((corner >> 14) & 2u) | ((corner >> 21) & 4u) | ((corner >> 28) & 8u);
                }
                const uint32_t m = __ballot_sync(0xffffffffu, nib != 0);
                if (m) {
                    // few lanes hold corners in scene-like frames: direct shared atomics; dense warps use the scan
                    int slot = 0;
                    if (__popc(m) <= 6) { if (nib) slot = atomicAdd(&sCornerN, __popc(nib)); }
                    else slot = warpAppend(&sCornerN, __popc(nib));
                    if (nib) {
                        #pragma unroll
                        for (int b = 0; b < 4; ++b)
                            if ((nib >> b) & 1u) cornerList[slot++] = (uint16_t)((y << 10) | (lx0 + b));
                        // the 4 pixels may straddle two bitmap words
                        const int lxa = max(lx0, 0);
                        const uint32_t bitsA = (lx0 >= 0 ? nib : nib >> (-lx0));
                        const int wA = y * BP + (lxa >> 5), sh = lxa & 31;
                        atomicOr(&cornerBits[wA], bitsA << sh);
                        if (sh > 28 && (bitsA >> (32 - sh))) atomicOr(&cornerBits[wA + 1], bitsA >> (32 - sh));
                    }
                }
            }
            __syncthreads();
            if (sCornerN > kCornerCap - 4 * kGroupBatch || gEnd == nG) {   // uniform: score in bulk before the list could overflow
                const int n = sCornerN;
                lastN = n; ++flushes;
                scorePass(n);
                __syncthreads();
                if (threadIdx.x == 0) sCornerN = 0;
                __syncthreads();
            }
        }

        // ---- nms (strict >, 8 neighbours, zero outside the pixel's own cell): one list entry per thread step when the
        //      list still holds every corner of the sweep, else 8 bitmap columns per step ----
        auto nmsOne = [&](int ly, int lx) {
            const uint8_t* r = score + (ly + 1) * SP + (lx + 1);
            const int s = r[0];
            const int cj = colCell[lx], cx = lx - cj * wCell;
            bool keep = s > r[-SP] && s > r[SP];
            if (cx > 0) keep = keep && s > r[-1] && s > r[-SP - 1] && s > r[SP - 1];
            if (cx < wCell - 1) keep = keep && s > r[1] && s > r[-SP + 1] && s > r[SP + 1];
            if (keep) {
                survList[atomicAdd(&sSurvN, 1)] = ((uint32_t)s << 16) | (uint32_t)(ly << 10) | (uint32_t)lx;
                if (!sweep) atomicAdd(&sCellCnt[cj], 1);
            }
        };
        if (flushes == 1) {
            for (int i = threadIdx.x; i < lastN; i += kFastThreads) { const int code = cornerList[i]; nmsOne(code >> 10, code & 1023); }
        } else if (flushes > 1) {
            for (int it = threadIdx.x; it < rh * BP * 4; it += kFastThreads) {
                uint32_t bits = cornerBytes[it];
                if (!bits) continue;
                const int ly = (int)__umulhi((uint32_t)it, rcpBP4), lxb = (it - ly * BP * 4) << 3;
                while (bits) {
                    const int b = __ffs(bits) - 1;
                    bits &= bits - 1;
                    nmsOne(ly, lxb + b);
                }
            }
        }
        __syncthreads();
        // ---- emit ----
        const int nS = sSurvN;
        if (threadIdx.x == 0) sBase = nS ? atomicAdd(&candCount[frame * G.nlevels + level], nS) : 0;
        __syncthreads();
        for (int i = threadIdx.x; i < nS; i += kFastThreads) {
            const uint32_t r = survList[i];
            const int s = r >> 16, ly = (r >> 10) & 63, lx = r & 1023;
            const int cj = colCell[lx], cx = lx - cj * wCell;
            const uint32_t key = ((uint32_t)(ci * g.nCols + j0 + cj) << 12) | (uint32_t)(ly << 6) | (uint32_t)cx;
            out[sBase + i] = ((unsigned long long)((uint32_t)s << 24) << 32) | key; // hi = score<<24 | node(0), lo = key
        }
        __syncthreads();
    }
}

int launchFast(Ctx* c, int frame0, int batch)
{
    GeomSet g = c->geom;
    g.frame0 = frame0;
    stageBegin(c, ST_FAST);
    ORBX_CUDA(c, cudaMemsetAsync(c->dCandCount + (size_t)frame0 * g.nlevels, 0, sizeof(int) * (size_t)batch * g.nlevels, c->stream));
    int launches = 0;
    if (g.stripsPerFrame > 0) {
        const int TP = g.fastTilePitch, TR = g.fastTileRows, SP = g.fastScorePitch;
        const int maxRh = TR - 6, BP = (SP + 31) / 32;
        // corner list, then the group list whose space is reused as the survivor list (u32 records)
        size_t listBytes = sizeof(uint16_t) * (size_t)kCornerCap +
                           std::max(sizeof(uint16_t) * (size_t)maxRh * ((SP >> 2) + 2), sizeof(uint32_t) * (size_t)fastSurvivorCap(SP, maxRh));
        listBytes = (listBytes + 3) & ~(size_t)3;
        size_t smem = (size_t)TP * TR + (size_t)SP * (maxRh + 2);
        smem += sizeof(uint32_t) * ((size_t)maxRh * BP + (size_t)(TP >> 2));
        smem += listBytes + (size_t)SP;
        smem = (smem + 127) & ~(size_t)127;
        const uint32_t rcpBP4 = 0xffffffffu / (uint32_t)(BP * 4) + 1u;
        dim3 grid(g.stripsPerFrame, batch);
        auto go = [&](auto kern, int threads, size_t* configured) -> int {
            if (smem > 48 * 1024 && smem > *configured) {
                ORBX_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                *configured = smem;
            }
            kern<<<grid, dim3(threads), smem, c->stream>>>(g, c->lvl0, c->lvl0Pitch, c->lvl0FrameStride, c->dPyr,
                                                           c->dCand, c->dCandCount, TP, TR, SP, (int)listBytes, rcpBP4);
            return ORBX_OK;
        };
        static size_t cfg256 = 0, cfg128 = 0;
        const int rc = c->fastThreads == 128 ? go(fastStripKernel<128, 8>, 128, &cfg128) : go(fastStripKernel<256, 4>, 256, &cfg256);
        if (rc != ORBX_OK) return rc;
        launches = 1;
    }
    stageEnd(c, ST_FAST, launches);
    ORBX_CUDA(c, cudaGetLastError());
    return ORBX_OK;
}

} // namespace orbx
