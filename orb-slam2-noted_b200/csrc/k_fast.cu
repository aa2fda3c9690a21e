// K2 -- per-cell FAST-9/16 with the iniThFAST -> minThFAST retry
// (ComputeKeyPointsOctTree cell loop, src/ORBextractor.cc:1058-1153; cv::FAST(roi, th, nms=true)).
//
// One CTA per (cell, frame); all levels of all frames in ONE launch (cell table in GeomSet).
//   phase 0  stage the (wCell+6) x (hCell+6) tile in shared memory (coalesced row loads)
//   phase 1  every candidate pixel: 16-bit brighter/darker ring masks at minTh, 9-contiguous test;
//            corners are compacted into a shared list (no divergent heavy work)
//   phase 2  list entries: exact OpenCV cornerScore (max over the 16 arcs of the arc's min margin, -1),
//            which does not depend on the threshold for a detected corner -> u8 score map with a zero apron
//   phase 3  3x3 strict-> NMS inside the cell only (neighbours outside the cell's candidate region count 0,
//            exactly like the per-ROI cv::FAST calls), count survivors with score >= iniTh;
//            the cell keeps {survivors >= iniTh} if that set is non-empty, else all survivors (the retry).
// Output records are order-independent: key = (cell, ly, lx) reproduces the reference's push_back order
// (cell-row-major, then pixel-row-major :1084-1149) whenever an ordered view is needed.
#include "orbx_internal.h"

namespace orbx {

__device__ __forceinline__ bool nineContiguous(uint32_t m16)
{
    uint32_t m = m16 | (m16 << 16);
    uint32_t a = m & (m >> 1);
    a &= a >> 2;
    a &= a >> 4;      // runs of 8
    a &= m >> 8;      // runs of 9
    return (a & 0xffffu) != 0;
}

constexpr int kFastThreads = 256;

__global__ void __launch_bounds__(kFastThreads)
fastCellsKernel(const GeomSet* __restrict__ G, const uint8_t* __restrict__ lvl0, size_t lvl0Pitch, size_t lvl0Frame,
                const uint8_t* __restrict__ pyr, unsigned long long* __restrict__ cand, int* __restrict__ candCount,
                int tilePitchMax)
{
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ int sListN, sNms, sIni, sBase;
    __shared__ int sLevel;

    const int frame = blockIdx.y;
    // locate level of this cell (<= 16 levels: linear scan by one thread)
    if (threadIdx.x == 0) {
        int l = 0;
        const int cell = blockIdx.x;
        while (l + 1 < G->nlevels && cell >= G->lv[l + 1].cellBase) ++l;
        sLevel = l; sListN = 0; sNms = 0; sIni = 0;
    }
    __syncthreads();
    const int level = sLevel;
    const LevelGeom& g = G->lv[level];
    const int cellIdx = blockIdx.x - g.cellBase;
    const int ci = cellIdx / g.nCols, cj = cellIdx - ci * g.nCols;
    // candidate region in level coordinates
    const int cx0 = kEdge + cj * g.wCell, cy0 = kEdge + ci * g.hCell;
    const int cx1 = min(cx0 + g.wCell, g.w - kEdge), cy1 = min(cy0 + g.hCell, g.h - kEdge);
    const int rw = cx1 - cx0, rh = cy1 - cy0;
    if (rw <= 0 || rh <= 0) return; // skipped cells (:1094,:1112) have an empty candidate region

    const uint8_t* img; size_t pitch;
    if (level == 0) { img = lvl0 + (size_t)frame * lvl0Frame; pitch = lvl0Pitch; }
    else { img = pyr + g.planeOff + (size_t)frame * g.pitch * g.h; pitch = (size_t)g.pitch; }

    const int tw = rw + 6, th = rh + 6;
    const int tp = tilePitchMax;                       // shared tile pitch (bytes)
    uint8_t* tile = smem;                              // th x tp
    uint8_t* score = smem + (size_t)tp * (g.hCell + 6); // (rh+2) x sp, zero apron
    const int sp = g.wCell + 2;
    uint16_t* list = reinterpret_cast<uint16_t*>(score + (((size_t)sp * (g.hCell + 2) + 15) & ~(size_t)15));

    // phase 0: tile + zeroed score map
    for (int i = threadIdx.x; i < tw * th; i += kFastThreads) {
        const int y = i / tw, x = i - y * tw;
        tile[y * tp + x] = img[(size_t)(cy0 - 3 + y) * pitch + (cx0 - 3 + x)];
    }
    for (int i = threadIdx.x; i < sp * (rh + 2); i += kFastThreads) score[i] = 0;
    __syncthreads();

    const int o[16] = {3 * tp, 3 * tp + 1, 2 * tp + 2, tp + 3, 3, -tp + 3, -2 * tp + 2, -3 * tp + 1,
                       -3 * tp, -3 * tp - 1, -2 * tp - 2, -tp - 3, -3, tp - 3, 2 * tp - 2, 3 * tp - 1};
    const int tmin = G->minTh, tini = G->iniTh;

    // phase 1: corner test at minTh
    for (int i = threadIdx.x; i < rw * rh; i += kFastThreads) {
        const int ly = i / rw, lx = i - ly * rw;
        const uint8_t* c = tile + (ly + 3) * tp + (lx + 3);
        const int v = c[0];
        const int lo = v - tmin, hi = v + tmin;
        // any 9-arc contains ring pixel k or k+8: two cheap rejections first
        int p0 = c[o[0]], p8 = c[o[8]];
        if (!((p0 > hi) | (p0 < lo) | (p8 > hi) | (p8 < lo))) continue;
        int p4 = c[o[4]], p12 = c[o[12]];
        if (!((p4 > hi) | (p4 < lo) | (p12 > hi) | (p12 < lo))) continue;
        uint32_t br = 0, dk = 0;
        #pragma unroll
        for (int k = 0; k < 16; ++k) {
            const int p = c[o[k]];
            br |= (uint32_t)(p > hi) << k;
            dk |= (uint32_t)(p < lo) << k;
        }
        if (nineContiguous(br) || nineContiguous(dk)) {
            const int slot = atomicAdd(&sListN, 1);
            list[slot] = (uint16_t)(ly * 64 + lx);
        }
    }
    __syncthreads();
    const int nList = sListN;

    // phase 2: exact score for the listed corners
    for (int i = threadIdx.x; i < nList; i += kFastThreads) {
        const int code = list[i], ly = code >> 6, lx = code & 63;
        const uint8_t* c = tile + (ly + 3) * tp + (lx + 3);
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
        const int s = max(a0, -b0) - 1;  // > minTh-1 >= 0 for a detected corner
        score[(ly + 1) * sp + (lx + 1)] = (uint8_t)s;
    }
    __syncthreads();

    // phase 3: NMS (strict >, 8 neighbours, zero outside the cell)
    for (int i = threadIdx.x; i < nList; i += kFastThreads) {
        const int code = list[i], ly = code >> 6, lx = code & 63;
        const uint8_t* r = score + (ly + 1) * sp + (lx + 1);
        const int s = r[0];
        const bool keep = s > r[-1] && s > r[1] && s > r[-sp - 1] && s > r[-sp] && s > r[-sp + 1] &&
                          s > r[sp - 1] && s > r[sp] && s > r[sp + 1];
        if (keep) { atomicAdd(&sNms, 1); if (s >= tini) atomicAdd(&sIni, 1); }
        else list[i] = 0xffff;
    }
    __syncthreads();
    const int nIni = sIni;
    const int nOut = nIni > 0 ? nIni : sNms;
    if (nOut == 0) return;
    if (threadIdx.x == 0) { sBase = atomicAdd(&candCount[frame * G->nlevels + level], nOut); sNms = 0; }
    __syncthreads();
    unsigned long long* out = cand + (size_t)frame * G->candPerFrame + g.candBase + sBase;
    for (int i = threadIdx.x; i < nList; i += kFastThreads) {
        const int code = list[i];
        if (code == 0xffff) continue;
        const int ly = code >> 6, lx = code & 63;
        const int s = score[(ly + 1) * sp + (lx + 1)];
        if (nIni > 0 && s < tini) continue;
        const int slot = atomicAdd(&sNms, 1);
        const uint32_t key = ((uint32_t)cellIdx << 12) | (uint32_t)code;
        out[slot] = ((unsigned long long)((uint32_t)s << 24) << 32) | key; // hi = score<<24 | node(0), lo = key
    }
}

int launchFast(Ctx* c, int batch)
{
    const GeomSet& g = c->geom;
    int maxW = 0, maxH = 0;
    for (int l = 0; l < g.nlevels; ++l) if (g.lv[l].nCols > 0) { maxW = max(maxW, g.lv[l].wCell); maxH = max(maxH, g.lv[l].hCell); }
    stageBegin(c, ST_FAST);
    ORBX_CUDA(c, cudaMemsetAsync(c->dCandCount, 0, sizeof(int) * (size_t)batch * g.nlevels, c->stream));
    int launches = 0;
    if (g.cellsPerFrame > 0) {
        const int tp = (maxW + 6 + 3) & ~3;
        size_t smem = (size_t)tp * (maxH + 6);
        smem += (((size_t)(maxW + 2) * (maxH + 2) + 15) & ~(size_t)15);
        smem += sizeof(uint16_t) * (size_t)maxW * maxH; // worst case list (every pixel); NMS-free bound
        smem = (smem + 15) & ~(size_t)15;
        static size_t configured = 0;
        if (smem > 48 * 1024 && smem > configured) {
            ORBX_CUDA(c, cudaFuncSetAttribute(fastCellsKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            configured = smem;
        }
        dim3 grid(g.cellsPerFrame, batch), block(kFastThreads);
        fastCellsKernel<<<grid, block, smem, c->stream>>>(c->dGeom, c->lvl0, c->lvl0Pitch, c->lvl0FrameStride, c->dPyr,
                                                         c->dCand, c->dCandCount, tp);
        launches = 1;
    }
    stageEnd(c, ST_FAST, launches);
    ORBX_CUDA(c, cudaGetLastError());
    return ORBX_OK;
}

} // namespace orbx
