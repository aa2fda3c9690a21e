// Host side of the C ABI (include/orbx.h): context, constructor tables, per-size geometry, the
// extraction pipeline (H2D -> K1 pyramid -> K2 FAST cells -> K3 quadtree -> K4b blur -> K4a/K5 describe
// -> D2H), instrumentation.  No CPU fallback anywhere: every stage is a CUDA kernel on ctx->stream.
#include "orbx_internal.h"
#include <math.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>
#include <new>
#include <algorithm>

namespace orbx {

static const char* kStageNames[ST_COUNT] = {"pyramid", "fast", "octree", "blur", "describe", "h2d", "d2h",
                                            "hamming", "match", "stereo", "bow"};

int cudaFail(Ctx* c, cudaError_t e, const char* what)
{
    if (c) { c->lastCudaError = std::string(what) + ": " + cudaGetErrorString(e); }
    return ORBX_E_CUDA;
}

static cudaEvent_t takeEvent(Ctx* c)
{
    if (!c->eventPool.empty()) { cudaEvent_t e = c->eventPool.back(); c->eventPool.pop_back(); return e; }
    cudaEvent_t e = nullptr; cudaEventCreate(&e); return e;
}

void stageBegin(Ctx* c, int stage)
{
    if (!c->profiling) return;
    cudaEvent_t a = takeEvent(c), b = takeEvent(c);
    cudaEventRecord(a, c->stream);
    c->pendingEvents.push_back({stage, {a, b}});
}

void stageEnd(Ctx* c, int stage, int launches)
{
    c->kernelLaunches += launches;
    c->stageLaunches[stage] += launches;
    if (!c->profiling) return;
    cudaEventRecord(c->pendingEvents.back().second.second, c->stream);
}

static void drainEvents(Ctx* c)
{
    for (auto& pe : c->pendingEvents) {
        float ms = 0.f;
        cudaEventSynchronize(pe.second.second);
        if (cudaEventElapsedTime(&ms, pe.second.first, pe.second.second) == cudaSuccess) c->stageMs[pe.first] += ms;
        c->eventPool.push_back(pe.second.first); c->eventPool.push_back(pe.second.second);
    }
    c->pendingEvents.clear();
}

int ensureScratch(Ctx* c, size_t bytes)
{
    if (bytes <= c->scratchBytes) return ORBX_OK;
    if (c->dScratch) { cudaStreamSynchronize(c->stream); cudaFree(c->dScratch); c->dScratch = nullptr; c->scratchBytes = 0; }
    bytes = (bytes + (1u << 20)) & ~(size_t)((1u << 20) - 1);
    ORBX_CUDA(c, cudaMalloc(&c->dScratch, bytes));
    c->scratchBytes = bytes;
    return ORBX_OK;
}

int ensureMatchPinned(Ctx* c, size_t bytes)
{
    if (bytes <= c->hMatchBytes) return ORBX_OK;
    if (c->hMatch) { cudaStreamSynchronize(c->stream); cudaFreeHost(c->hMatch); c->hMatch = nullptr; c->hMatchBytes = 0; }
    bytes = (bytes + (1u << 18)) & ~(size_t)((1u << 18) - 1);
    ORBX_CUDA(c, cudaMallocHost((void**)&c->hMatch, bytes));
    c->hMatchBytes = bytes;
    return ORBX_OK;
}

// cvRound(float): round half to even under the default rounding mode
static inline int cvRoundF(float v) { return (int)lrintf(v); }
static inline int cvRoundD(double v) { return (int)lrint(v); }

// Constructor tables -- same formulas, same float/double mix as src/ORBextractor.cc:471-579.
static void initTables(Ctx* c)
{
    const int L = c->nlevels;
    c->scale[0] = 1.0f; c->sigma2[0] = 1.0f;
    for (int i = 1; i < L; ++i) { c->scale[i] = c->scale[i - 1] * c->scaleFactor; c->sigma2[i] = c->scale[i] * c->scale[i]; }
    for (int i = 0; i < L; ++i) { c->invScale[i] = 1.0f / c->scale[i]; c->invSigma2[i] = 1.0f / c->sigma2[i]; }
    const float factor = 1.0f / c->scaleFactor;
    float nd = (float)c->nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)L));
    int sum = 0;
    for (int l = 0; l < L - 1; ++l) { c->nfeatLevel[l] = cvRoundF(nd); sum += c->nfeatLevel[l]; nd *= factor; }
    c->nfeatLevel[L - 1] = c->nfeatures - sum > 0 ? c->nfeatures - sum : 0;
    int v, v0;
    const int vmax = (int)floor(kHalfPatch * sqrt(2.f) / 2 + 1), vmin = (int)ceil(kHalfPatch * sqrt(2.f) / 2);
    const double hp2 = kHalfPatch * kHalfPatch;
    for (v = 0; v <= vmax; ++v) c->umax[v] = cvRoundD(sqrt(hp2 - v * v));
    for (v = kHalfPatch, v0 = 0; v >= vmin; --v) { while (c->umax[v0] == c->umax[v0 + 1]) ++v0; c->umax[v] = v0; ++v0; }
}

static void levelSize(const Ctx* c, int l, int w0, int h0, int* w, int* h)
{   // src/ORBextractor.cc:1669-1672
    *w = cvRoundF((float)w0 * c->invScale[l]);
    *h = cvRoundF((float)h0 * c->invScale[l]);
}

// cv::resize INTER_LINEAR coefficient tables (OpenCV resize.cpp), computed once per geometry on the host
static void buildTaps(int dn, int sn, bool clampF, std::vector<ResizeTap>& t)
{
    t.resize(dn);
    const double scale = 1.0 / ((double)dn / (double)sn);
    for (int d = 0; d < dn; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)floorf(f);
        f -= (float)s;
        if (clampF) { if (s < 0) { f = 0; s = 0; } if (s >= sn - 1) { f = 0; s = sn - 1; } }
        else { if (s < 0) s = 0; if (s > sn - 1) s = sn - 1; } // rows: index clipped, weights kept (downscale: s>=0 always)
        const int a = cvRoundF((1.f - f) * 2048.f), b = cvRoundF(f * 2048.f);
        t[d].s = (uint16_t)s; t[d].c0 = (int16_t)a; t[d].c1 = (int16_t)b; t[d].pad = 0;
    }
}

static void freeGeometry(Ctx* c)
{
    for (int l = 0; l < kMaxLevels; ++l) {
        if (c->dTapX[l]) cudaFree(c->dTapX[l]);
        if (c->dTapY[l]) cudaFree(c->dTapY[l]);
        if (c->dTapG[l]) cudaFree(c->dTapG[l]);
        c->dTapX[l] = c->dTapY[l] = nullptr; c->dTapG[l] = nullptr;
    }
}

// (Re)compute everything that depends on the image size.  Arenas are sized for maxW x maxH x maxBatch once.
static int configure(Ctx* c, int w, int h)
{
    if (w == c->curW && h == c->curH) return ORBX_OK;
    GeomSet& g = c->geom;
    memset(&g, 0, sizeof(g));
    g.nlevels = c->nlevels; g.iniTh = c->iniTh; g.minTh = c->minTh;
    for (int i = 0; i < 16; ++i) g.umax[i] = c->umax[i];
    size_t planeOff = 0, blurOff = 0;
    int cellBase = 0, candBase = 0, outBase = 0, tileBase = 0, stripBase = 0, maxRw = 0, maxHc = 0;
    for (int l = 0; l < c->nlevels; ++l) {
        LevelGeom& q = g.lv[l];
        levelSize(c, l, w, h, &q.w, &q.h);
        if (q.w < 1 || q.h < 1 || q.w > 65535 || q.h > 32767) return ORBX_E_UNSUPPORTED;
        q.pitch = (q.w + 15) & ~15;
        q.scale = c->scale[l];
        q.patchSize = (float)(int)((float)kPatch * c->scale[l]);
        q.quota = c->nfeatLevel[l];
        const int minB = kWinOff, maxBX = q.w - kEdge + 3, maxBY = q.h - kEdge + 3;
        const float width = (float)(maxBX - minB), height = (float)(maxBY - minB);
        const int nCols = (int)(width / 30.f), nRows = (int)(height / 30.f); // :1076-1077
        if (nCols >= 1 && nRows >= 1) {
            q.nCols = nCols; q.nRows = nRows;
            q.wCell = (int)ceilf(width / nCols); q.hCell = (int)ceilf(height / nRows); // :1080-1081
            const int nIni = (int)roundf((float)(maxBX - minB) / (float)(maxBY - minB)); // :703
            if (nIni < 1 || nIni > 8) return ORBX_E_UNSUPPORTED; // nIni == 0: the reference divides by zero
            q.nIni = nIni; q.hX = (float)(maxBX - minB) / (float)nIni;                   // :706
            if (q.wCell >= 64 || q.hCell >= 64) return ORBX_E_UNSUPPORTED;
        } else { q.nCols = q.nRows = 0; q.wCell = q.hCell = 1; q.nIni = 1; q.hX = 1.f; } // reference: UB; here: no keypoints
        q.cellBase = cellBase; cellBase += q.nCols * q.nRows;
        // NMS survivors are never 8-adjacent: at most ceil(w/2)*ceil(h/2) per cell
        q.candCap = q.nCols * q.nRows * ((q.wCell + 1) / 2) * ((q.hCell + 1) / 2);
        q.candBase = candBase; candBase += (q.candCap + 3) & ~3;
        q.outCap = q.quota + 4 * q.nIni + 4;
        q.outBase = outBase; outBase += q.outCap;
        q.planeOff = planeOff; planeOff += (size_t)q.pitch * q.h * c->maxBatch; planeOff = (planeOff + 255) & ~(size_t)255;
        q.blurOff = blurOff; blurOff += (size_t)q.pitch * q.h * c->maxBatch; blurOff = (blurOff + 255) & ~(size_t)255;
        q.tilesX = (q.w + kBlurTW - 1) / kBlurTW; q.tilesY = (q.h + kBlurTH - 1) / kBlurTH; // blur tiles (k_blur.cu)
        q.tileBase = tileBase; tileBase += q.tilesX * q.tilesY;
        // FAST strips: segments of whole cells, at most 320 candidate columns and 32 cells each (k_fast.cu)
        q.stripBase = stripBase; q.segsPerRow = 0; q.cellsPerSeg = 1;
        if (q.nCols > 0) {
            const int maxCells = std::max(1, std::min(32, c->fastSegW / q.wCell));
            q.segsPerRow = (q.nCols + maxCells - 1) / maxCells;
            q.cellsPerSeg = (q.nCols + q.segsPerRow - 1) / q.segsPerRow;
            stripBase += q.segsPerRow * q.nRows;
            maxRw = std::max(maxRw, q.cellsPerSeg * q.wCell); maxHc = std::max(maxHc, q.hCell);
        }
    }
    g.stripsPerFrame = stripBase;
    g.fastTilePitch = ((maxRw + 37 + 15) & ~15) + 16; g.fastTileRows = maxHc + 6; g.fastScorePitch = (maxRw + 2 + 3) & ~3;
    g.cellsPerFrame = cellBase; g.candPerFrame = candBase; g.selPerFrame = outBase; g.blurTilesPerFrame = tileBase;
    ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    // arenas only ever grow (sized for max_w x max_h x max_batch at creation)
    if (planeOff > c->pyrBytes) { if (c->dPyr) cudaFree(c->dPyr); c->dPyr = nullptr; ORBX_CUDA(c, cudaMalloc(&c->dPyr, planeOff)); c->pyrBytes = planeOff; }
    if (blurOff > c->blurBytes) { if (c->dBlur) cudaFree(c->dBlur); c->dBlur = nullptr; ORBX_CUDA(c, cudaMalloc(&c->dBlur, blurOff)); c->blurBytes = blurOff; }
    const size_t candNeed = (size_t)candBase * c->maxBatch;
    if (candNeed > c->candRecs) { if (c->dCand) cudaFree(c->dCand); c->dCand = nullptr; ORBX_CUDA(c, cudaMalloc(&c->dCand, candNeed * sizeof(unsigned long long))); c->candRecs = candNeed; }
    const size_t selNeed = (size_t)outBase * c->maxBatch;
    if (selNeed > c->selRecs) { if (c->dSel) cudaFree(c->dSel); c->dSel = nullptr; ORBX_CUDA(c, cudaMalloc(&c->dSel, selNeed * sizeof(orbx_keypoint))); c->selRecs = selNeed; }
    freeGeometry(c);
    for (int l = 1; l < c->nlevels; ++l) {
        std::vector<ResizeTap> tx, ty;
        buildTaps(g.lv[l].w, g.lv[l - 1].w, true, tx);
        buildTaps(g.lv[l].h, g.lv[l - 1].h, false, ty);
        bool wordOk = true;
        for (size_t x = 0; x < tx.size(); x += 4) {
            const int last = tx[std::min(x + 3, tx.size() - 1)].s, first = tx[x].s;
            if (last < first || last - (first & ~3) > 7) wordOk = false;
        }
        for (const ResizeTap& t : tx) if (t.c0 < 0 || t.c1 < 0) wordOk = false;
        for (const ResizeTap& t : ty) if (t.c0 < 0 || t.c1 < 0) wordOk = false;
        for (size_t y = 1; y < ty.size(); ++y) if (ty[y].s <= ty[y - 1].s) wordOk = false;   // one destination row per source row at most
        c->resizeWordOk[l] = wordOk;
        bool ok8 = wordOk;
        for (size_t x = 0; x < tx.size(); x += 8)
            for (size_t i = 1; i < 8 && x + i < tx.size(); ++i) {
                const int d = tx[x + i].s - tx[x].s;
                if (i < 4 ? (d < 0 || d > 6) : (d < 4 || d > 10)) ok8 = false;
            }
        c->resize8Ok[l] = ok8;
        if (ok8) {   // constants of the staged kernel, one column per 8-pixel group (k_pyramid.cu: resizeLinearStagedKernel)
            const int groups = ((int)tx.size() + 7) / 8;
            std::vector<uint32_t> tg((size_t)11 * groups, 0u);
            for (int gI = 0; gI < groups; ++gI) {
                const int first = tx[(size_t)gI * 8].s;
                tg[gI] = (uint32_t)first;
                for (int i = 0; i < 8; ++i) {
                    const ResizeTap& t = tx[std::min((size_t)gI * 8 + i, tx.size() - 1)];
                    // pixels past the row end reuse the last tap (they land in the row's padding): keep the selector in range
                    const int dlt = std::min(std::max(t.s - first - (i >= 4 ? 4 : 0), 0), 6);
                    const uint32_t sel = (uint32_t)dlt | ((uint32_t)(dlt + 1) << 4);
                    tg[(size_t)(1 + i / 4) * groups + gI] |= sel << (8 * (i & 3));
                    tg[(size_t)(3 + i) * groups + gI] = (uint32_t)(uint16_t)t.c0 | ((uint32_t)(uint16_t)t.c1 << 16);
                }
            }
            ORBX_CUDA(c, cudaMalloc(&c->dTapG[l], sizeof(uint32_t) * tg.size()));
            ORBX_CUDA(c, cudaMemcpy(c->dTapG[l], tg.data(), sizeof(uint32_t) * tg.size(), cudaMemcpyHostToDevice));
        }
        ORBX_CUDA(c, cudaMalloc(&c->dTapX[l], sizeof(ResizeTap) * tx.size()));
        ORBX_CUDA(c, cudaMalloc(&c->dTapY[l], sizeof(ResizeTap) * ty.size()));
        ORBX_CUDA(c, cudaMemcpy(c->dTapX[l], tx.data(), sizeof(ResizeTap) * tx.size(), cudaMemcpyHostToDevice));
        ORBX_CUDA(c, cudaMemcpy(c->dTapY[l], ty.data(), sizeof(ResizeTap) * ty.size(), cudaMemcpyHostToDevice));
    }
    ORBX_CUDA(c, cudaMemcpy(c->dGeom, &g, sizeof(g), cudaMemcpyHostToDevice));
    {
        std::vector<StripDesc> strips;
        buildFastStrips(g, strips);
        if ((int)strips.size() > c->stripCap) {
            if (c->dStrips) cudaFree(c->dStrips);
            c->dStrips = nullptr; c->stripCap = 0;
            ORBX_CUDA(c, cudaMalloc(&c->dStrips, sizeof(StripDesc) * strips.size()));
            c->stripCap = (int)strips.size();
        }
        if (!strips.empty())
            ORBX_CUDA(c, cudaMemcpy(c->dStrips, strips.data(), sizeof(StripDesc) * strips.size(), cudaMemcpyHostToDevice));
        c->nStrips = (int)strips.size();
    }
    c->curW = w; c->curH = h; c->geomStamp++;
    return ORBX_OK;
}

} // namespace orbx
using namespace orbx;

extern "C" {

const char* orbx_strerror(int s)
{
    switch (s) {
    case ORBX_OK: return "ok";
    case ORBX_E_INVALID: return "invalid argument";
    case ORBX_E_CUDA: return "CUDA error";
    case ORBX_E_NOMEM: return "out of memory";
    case ORBX_E_CAPACITY: return "capacity exceeded";
    case ORBX_E_UNSUPPORTED: return "unsupported geometry";
    default: return "unknown status";
    }
}

const char* orbx_last_cuda_error(const orbx_ctx* ctx) { return ctx ? ctx->c.lastCudaError.c_str() : ""; }

int orbx_create(orbx_ctx** out, int device, int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th,
                int max_w, int max_h, int max_batch)
{
    if (!out) return ORBX_E_INVALID;
    *out = nullptr;
    if (nfeatures < 0 || nlevels < 1 || nlevels > kMaxLevels || !(scale_factor > 1.0f) || ini_th < 1 || min_th < 1 ||
        min_th > ini_th || ini_th > 254 || max_w < 1 || max_h < 1 || max_batch < 1)
        return ORBX_E_INVALID;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return ORBX_E_CUDA; // no CPU fallback
    orbx_ctx* h = new (std::nothrow) orbx_ctx();
    if (!h) return ORBX_E_NOMEM;
    Ctx* c = &h->c;
    c->device = device; c->nfeatures = nfeatures; c->scaleFactor = scale_factor; c->nlevels = nlevels;
    c->iniTh = ini_th; c->minTh = min_th; c->maxW = max_w; c->maxH = max_h; c->maxBatch = max_batch;
    initTables(c);
    int rc = ORBX_OK;
    do {
        if (cudaSetDevice(device) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        if (cudaStreamCreateWithFlags(&c->sIn, cudaStreamNonBlocking) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        if (cudaStreamCreateWithFlags(&c->sOut, cudaStreamNonBlocking) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        if (cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        if (cudaEventCreateWithFlags(&c->evJoin, cudaEventDisableTiming) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        if (const char* e = getenv("ORBX_CHUNK")) c->chunk = atoi(e);
        if (const char* e = getenv("ORBX_FRAME_GRAPH")) c->frameGraphOff = atoi(e) == 0;
        if (const char* e = getenv("ORBX_OVERLAP_BLUR")) { c->overlapBlur = atoi(e) != 0; c->blurAfterFast = atoi(e) == 2; }
        if (const char* e = getenv("ORBX_RESIZE_WIDE")) c->resizeWide = atoi(e) != 0;
        if (const char* e = getenv("ORBX_FAST_THREADS")) c->fastThreads = atoi(e) == 128 ? 128 : 256;
        if (const char* e = getenv("ORBX_FAST_SEGW")) c->fastSegW = std::max(64, std::min(640, atoi(e)));
        if (cudaMalloc(&c->dGeom, sizeof(GeomSet)) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        if (cudaMalloc(&c->dCandCount, sizeof(int) * (size_t)max_batch * kMaxLevels) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        if (cudaMalloc(&c->dSelCount, sizeof(int) * (size_t)max_batch * kMaxLevels) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        if (cudaMalloc(&c->dErr, sizeof(int)) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        cudaMemset(c->dErr, 0, sizeof(int));
        rc = configure(c, max_w, max_h);
    } while (0);
    if (rc != ORBX_OK) { orbx_destroy(h); return rc; }
    *out = h;
    return ORBX_OK;
}

// A context for the matcher entry points only: a device, a stream and (grown on demand) the matcher scratch; no pyramid,
// candidate or keypoint arenas, no geometry.  Extraction calls on it fail with ORBX_E_CAPACITY (max_w = max_h = 0).
int orbx_create_matcher(orbx_ctx** out, int device)
{
    if (!out) return ORBX_E_INVALID;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return ORBX_E_CUDA; // no CPU fallback
    orbx_ctx* h = new (std::nothrow) orbx_ctx();
    if (!h) return ORBX_E_NOMEM;
    Ctx* c = &h->c;
    c->device = device; c->nlevels = 1; c->scaleFactor = 1.2f; c->scale[0] = c->invScale[0] = c->sigma2[0] = c->invSigma2[0] = 1.0f;
    if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
        orbx_destroy(h);
        return ORBX_E_CUDA;
    }
    *out = h;
    return ORBX_OK;
}

void orbx_destroy(orbx_ctx* h)
{
    if (!h) return;
    Ctx* c = &h->c;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    drainEvents(c);
    for (cudaEvent_t e : c->eventPool) cudaEventDestroy(e);
    freeGeometry(c);
    void* ptrs[] = {c->dGeom, c->dPyr, c->dBlur, c->dCand, c->dCandCount, c->dSel, c->dSelCount, c->dErr, c->dFrameOff,
                    c->dOutKp, c->dOutDesc, c->dOutN, c->dIn, c->dScratch, c->dTensorMaps, c->dStrips, c->dStereoN};
    for (void* p : ptrs) if (p) cudaFree(p);
    if (c->hPinned) cudaFreeHost(c->hPinned);
    if (c->hMatch) cudaFreeHost(c->hMatch);
    if (c->evFence) cudaEventDestroy(c->evFence);
    for (cudaEvent_t e : c->evIn) cudaEventDestroy(e);
    for (cudaEvent_t e : c->evDone) cudaEventDestroy(e);
    if (c->sIn) cudaStreamDestroy(c->sIn);
    if (c->sOut) cudaStreamDestroy(c->sOut);
    if (c->frameGraph) cudaGraphExecDestroy(c->frameGraph);
    if (c->stream2) cudaStreamDestroy(c->stream2);
    if (c->evJoin) cudaEventDestroy(c->evJoin);
    if (c->evFork) cudaEventDestroy(c->evFork);
    if (c->evBlur) cudaEventDestroy(c->evBlur);
    if (c->sBlur) cudaStreamDestroy(c->sBlur);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete h;
}

int orbx_get_levels(const orbx_ctx* h) { return h ? h->c.nlevels : 0; }
float orbx_get_scale_factor(const orbx_ctx* h) { return h ? h->c.scaleFactor : 0.f; }

int orbx_get_tables(const orbx_ctx* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2, int* fpl)
{
    if (!h) return ORBX_E_INVALID;
    const Ctx* c = &h->c;
    for (int i = 0; i < c->nlevels; ++i) {
        if (scale) scale[i] = c->scale[i];
        if (inv_scale) inv_scale[i] = c->invScale[i];
        if (sigma2) sigma2[i] = c->sigma2[i];
        if (inv_sigma2) inv_sigma2[i] = c->invSigma2[i];
        if (fpl) fpl[i] = c->nfeatLevel[i];
    }
    return ORBX_OK;
}

int orbx_max_keypoints(const orbx_ctx* h)
{
    if (!h) return 0;
    int s = 0;
    for (int l = 0; l < h->c.nlevels; ++l) s += h->c.nfeatLevel[l] + 36; // quota + 4*nIni(<=8) + 4 per level
    return (s + 31) & ~31;
}

int orbx_level_size(const orbx_ctx* h, int level, int w0, int h0, int* w, int* hh)
{
    if (!h || level < 0 || level >= h->c.nlevels || !w || !hh) return ORBX_E_INVALID;
    levelSize(&h->c, level, w0, h0, w, hh);
    return ORBX_OK;
}

int orbx_sync(orbx_ctx* h)
{
    if (!h) return ORBX_E_INVALID;
    Ctx* c = &h->c;
    ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    drainEvents(c);
    return ORBX_OK;
}

void* orbx_stream(orbx_ctx* h) { return h ? (void*)h->c.stream : nullptr; }
int orbx_set_profiling(orbx_ctx* h, int on) { if (!h) return ORBX_E_INVALID; orbx_sync(h); h->c.profiling = on != 0; return ORBX_OK; }
int orbx_stage_count(void) { return ST_COUNT; }
const char* orbx_stage_name(int i) { return (i >= 0 && i < ST_COUNT) ? kStageNames[i] : ""; }

int orbx_stage_stats(orbx_ctx* h, int i, double* total_ms, long long* launches)
{
    if (!h || i < 0 || i >= ST_COUNT) return ORBX_E_INVALID;
    int rc = orbx_sync(h);
    if (rc) return rc;
    if (total_ms) *total_ms = h->c.stageMs[i];
    if (launches) *launches = h->c.stageLaunches[i];
    return ORBX_OK;
}

int orbx_reset_stats(orbx_ctx* h)
{
    if (!h) return ORBX_E_INVALID;
    int rc = orbx_sync(h);
    if (rc) return rc;
    for (int i = 0; i < ST_COUNT; ++i) { h->c.stageMs[i] = 0; h->c.stageLaunches[i] = 0; }
    return ORBX_OK;
}

long long orbx_kernel_launches(const orbx_ctx* h) { return h ? h->c.kernelLaunches : 0; }

} // extern "C"

namespace orbx {

static int checkDeviceError(Ctx* c)
{
    int e = 0;
    ORBX_CUDA(c, cudaMemcpyAsync(&e, c->dErr, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    if (e != 0) {
        cudaMemsetAsync(c->dErr, 0, sizeof(int), c->stream);
        c->lastCudaError = "device-side error flag " + std::to_string(e);
        return e == 5 ? ORBX_E_CAPACITY : ORBX_E_CUDA;
    }
    return ORBX_OK;
}

static int ensureOutStaging(Ctx* c, int cap)
{
    if (cap <= c->outCap && c->dOutKp) return ORBX_OK;
    ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    if (c->dOutKp) cudaFree(c->dOutKp);
    if (c->dOutDesc) cudaFree(c->dOutDesc);
    if (c->dOutN) cudaFree(c->dOutN);
    c->dOutKp = nullptr; c->dOutDesc = nullptr; c->dOutN = nullptr; c->outCap = 0;
    ORBX_CUDA(c, cudaMalloc(&c->dOutKp, sizeof(orbx_keypoint) * (size_t)cap * c->maxBatch));
    ORBX_CUDA(c, cudaMalloc(&c->dOutDesc, (size_t)32 * cap * c->maxBatch));
    ORBX_CUDA(c, cudaMalloc(&c->dOutN, sizeof(int) * (size_t)c->maxBatch));
    c->outCap = cap;
    return ORBX_OK;
}

} // namespace orbx

extern "C" {

int orbx_extract_batch(orbx_ctx* h, const uint8_t* imgs, int batch, int w, int hgt, size_t stride, size_t frame_stride,
                       int flags, orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out)
{
    if (!h || batch < 0 || !n_out) return ORBX_E_INVALID;
    Ctx* c = &h->c;
    if (!imgs || w <= 0 || hgt <= 0 || batch == 0) { // empty image: outputs untouched (src/ORBextractor.cc:1547)
        if (!(flags & ORBX_OUT_DEVICE)) for (int b = 0; b < batch; ++b) n_out[b] = 0;
        return ORBX_OK;
    }
    if (!kps || !desc || cap < 1 || stride < (size_t)w) return ORBX_E_INVALID;
    if (batch > 1 && frame_stride < stride * (size_t)hgt) return ORBX_E_INVALID;
    if (batch > c->maxBatch) return ORBX_E_CAPACITY;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    int rc = configure(c, w, hgt);
    if (rc) return rc;
    const GeomSet& g = c->geom;

    const bool inDev = flags & ORBX_IN_DEVICE, outDev = flags & ORBX_OUT_DEVICE;
    const bool dense = (flags & ORBX_OUT_DENSE) != 0;
    if (dense && !(inDev && outDev)) return ORBX_E_INVALID;      // dense rows need the whole batch in one pass, results on the device
    static const int devChunkEnv = getenv("ORBX_DEVICE_CHUNK") ? atoi(getenv("ORBX_DEVICE_CHUNK")) : 0;
    if (dense && devChunkEnv > 0 && devChunkEnv < batch) return ORBX_E_UNSUPPORTED;
    uint8_t* d0 = c->dPyr + g.lv[0].planeOff;
    const size_t dp = (size_t)g.lv[0].pitch;
    // a device image whose rows are not 16-byte aligned (e.g. a dense 1241-wide frame) is repacked into the arena once:
    // the bulk-copy staging of every stage needs aligned rows, and one extra pass over level 0 is cheaper than byte loads
    const bool repack = inDev && ((((size_t)imgs) | stride | (batch > 1 ? frame_stride : 0)) & 15) != 0;
    if (inDev && !repack) { c->lvl0 = imgs; c->lvl0Pitch = stride; c->lvl0FrameStride = frame_stride; }
    else { c->lvl0 = d0; c->lvl0Pitch = dp; c->lvl0FrameStride = dp * hgt; }
    if (repack) {
        if (frame_stride == stride * (size_t)hgt || batch == 1) {
            ORBX_CUDA(c, cudaMemcpy2DAsync(d0, dp, imgs, stride, (size_t)w, (size_t)hgt * batch, cudaMemcpyDeviceToDevice, c->stream));
        } else {
            for (int b = 0; b < batch; ++b)
                ORBX_CUDA(c, cudaMemcpy2DAsync(d0 + (size_t)b * dp * hgt, dp, imgs + (size_t)b * frame_stride, stride, (size_t)w,
                                               (size_t)hgt, cudaMemcpyDeviceToDevice, c->stream));
        }
    }
    c->lastBatch = batch;

    orbx_keypoint* dK = kps; uint8_t* dD = desc; int* dN = n_out;
    if (!outDev) {
        rc = ensureOutStaging(c, cap);
        if (rc) return rc;
        dK = c->dOutKp; dD = c->dOutDesc; dN = c->dOutN;
    }
    const int sc = outDev ? cap : c->outCap; // capacity per frame of the buffers the kernels write (>= cap)

    // Chunked pipeline: the stages of one chunk run back to back (its pyramid/blur planes stay in the 126 MB L2), and
    // with host buffers the H2D of chunk i+1 and the D2H of chunk i-1 overlap the kernels of chunk i.
    // device-resident input: one pass over the whole batch (larger launches are more efficient and there is no copy to hide)
    // (env ORBX_DEVICE_CHUNK = frames: device-resident batches are cut too and alternate between the two compute streams, so the
    // latency-bound stages of one chunk -- quadtree, describe -- run beside the ALU-bound FAST of the next)
    static const int devChunk = getenv("ORBX_DEVICE_CHUNK") ? atoi(getenv("ORBX_DEVICE_CHUNK")) : 0;
    const int chunk = (!inDev && c->chunk > 0) ? c->chunk : (inDev && devChunk > 0 ? devChunk : batch);
    if (!outDev && c->hPinnedBytes < sizeof(int) * (size_t)c->maxBatch) {   // pinned landing zone for n_out: keeps the D2H asynchronous
        if (c->hPinned) cudaFreeHost(c->hPinned);
        c->hPinned = nullptr; c->hPinnedBytes = 0;
        ORBX_CUDA(c, cudaMallocHost((void**)&c->hPinned, sizeof(int) * (size_t)c->maxBatch));
        c->hPinnedBytes = sizeof(int) * (size_t)c->maxBatch;
    }
    int* hN = reinterpret_cast<int*>(c->hPinned);
    std::vector<int> cuts(1, 0);                                   // chunk boundaries
    for (int f = chunk; f < batch; f += chunk) cuts.push_back(f);
    cuts.push_back(batch);
    const int nChunks = (int)cuts.size() - 1;
    while ((int)c->evIn.size() < nChunks) { cudaEvent_t e; ORBX_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); c->evIn.push_back(e); }
    while ((int)c->evDone.size() < nChunks) { cudaEvent_t e; ORBX_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); c->evDone.push_back(e); }
    if (!inDev || nChunks > 1) {
        if (!c->evFence) ORBX_CUDA(c, cudaEventCreateWithFlags(&c->evFence, cudaEventDisableTiming));
        ORBX_CUDA(c, cudaEventRecord(c->evFence, c->stream));
        if (!inDev) ORBX_CUDA(c, cudaStreamWaitEvent(c->sIn, c->evFence, 0));
        ORBX_CUDA(c, cudaStreamWaitEvent(c->stream2, c->evFence, 0));
    }
    cudaStream_t const mainStream = c->stream;
    struct StreamRestore { Ctx* c; cudaStream_t s; ~StreamRestore() { c->stream = s; } } restoreStream{c, mainStream};   // also on error returns
    const bool twoStreams = nChunks > 1;
    static const bool dbgTl = getenv("ORBX_DBG_TIMELINE") != nullptr;
    std::vector<cudaEvent_t> tl;
    auto mark = [&](cudaStream_t st) { if (!dbgTl) return; cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, st); tl.push_back(e); };
    mark(c->sIn);
    for (int i = 0; i < nChunks; ++i) {
        const int f0 = cuts[i], n = cuts[i + 1] - f0;
        c->stream = (twoStreams && (i & 1)) ? c->stream2 : mainStream;   // launchers enqueue on c->stream
        if (!inDev) {
            if ((frame_stride == stride * (size_t)hgt || n == 1) && stride == dp) {
                // both sides are one contiguous block: a plain 1-D copy.  (cudaMemcpy2DAsync with equal pitches still runs as a
                // strided DMA, one descriptor per row: 6 GB/s for 752-byte rows against 49 GB/s for the flat copy.)
                ORBX_CUDA(c, cudaMemcpyAsync(d0 + (size_t)f0 * dp * hgt, imgs + (size_t)f0 * frame_stride, dp * (size_t)hgt * n,
                                             cudaMemcpyHostToDevice, c->sIn));
            } else if (frame_stride == stride * (size_t)hgt || n == 1) {
                // rows of the caller's images and of the arena differ in pitch (e.g. 1241 -> 1248): the chunk crosses PCIe as one
                // flat block into a staging buffer and is repacked on the device, on the same copy stream (so the buffer is free
                // again before the next chunk's copy starts)
                const size_t bytes = stride * (size_t)hgt * n;
                if (bytes > c->dInBytes) {
                    ORBX_CUDA(c, cudaStreamSynchronize(c->sIn));
                    if (c->dIn) cudaFree(c->dIn);
                    c->dIn = nullptr; c->dInBytes = 0;
                    ORBX_CUDA(c, cudaMalloc(&c->dIn, bytes));
                    c->dInBytes = bytes;
                }
                ORBX_CUDA(c, cudaMemcpyAsync(c->dIn, imgs + (size_t)f0 * frame_stride, bytes, cudaMemcpyHostToDevice, c->sIn));
                ORBX_CUDA(c, cudaMemcpy2DAsync(d0 + (size_t)f0 * dp * hgt, dp, c->dIn, stride, (size_t)w, (size_t)hgt * n,
                                               cudaMemcpyDeviceToDevice, c->sIn));
            } else {
                for (int b = f0; b < f0 + n; ++b)
                    ORBX_CUDA(c, cudaMemcpy2DAsync(d0 + (size_t)b * dp * hgt, dp, imgs + (size_t)b * frame_stride, stride, (size_t)w,
                                                   (size_t)hgt, cudaMemcpyHostToDevice, c->sIn));
            }
            ORBX_CUDA(c, cudaEventRecord(c->evIn[i], c->sIn));
            mark(c->sIn);
            ORBX_CUDA(c, cudaStreamWaitEvent(c->stream, c->evIn[i], 0));
        }
        // The five stages of this chunk.  For ONE frame from and to host buffers (the per-frame call of a tracking thread) the
        // dozen launches are the latency, so their sequence is captured once per geometry into a CUDA graph and replayed
        // (the first call with a geometry runs eagerly: it sets function attributes and sizes arenas; the second captures).
        auto runStages = [&]() -> int {
        int rc = launchPyramid(c, f0, n);
        const bool fewCtas = (long long)n * c->nStrips < 148 * 8;   // a frame or two: kernels leave most SMs idle, so the blur runs beside FAST
        if (!rc && (c->overlapBlur || fewCtas) && !twoStreams) {
            // the blur only needs the pyramid: it runs on a side stream next to FAST + quadtree (FAST is limited to 5 CTAs per
            // SM by shared memory and leaves registers, shared memory and issue slots for a blur CTA) and joins before describe
            cudaStream_t const s0 = c->stream;
            if (!c->sBlur) {
                ORBX_CUDA(c, cudaStreamCreateWithFlags(&c->sBlur, cudaStreamNonBlocking));
                ORBX_CUDA(c, cudaEventCreateWithFlags(&c->evFork, cudaEventDisableTiming));
                ORBX_CUDA(c, cudaEventCreateWithFlags(&c->evBlur, cudaEventDisableTiming));
            }
            // blurAfterFast: fork AFTER FAST, so the blur shares the GPU with the quadtree only (FAST fills the shared memory of
            // every SM by itself; the quadtree waits on block scans most of the time and leaves issue slots)
            if (c->blurAfterFast && !fewCtas) rc = launchFast(c, f0, n);
            ORBX_CUDA(c, cudaEventRecord(c->evFork, s0));
            ORBX_CUDA(c, cudaStreamWaitEvent(c->sBlur, c->evFork, 0));
            if (!(c->blurAfterFast && !fewCtas)) rc = launchFast(c, f0, n);
            c->stream = c->sBlur;
            if (!rc) rc = launchBlur(c, f0, n);
            c->stream = s0;
            if (!rc) rc = launchOctree(c, f0, n);
            ORBX_CUDA(c, cudaEventRecord(c->evBlur, c->sBlur));
            ORBX_CUDA(c, cudaStreamWaitEvent(s0, c->evBlur, 0));
        } else {
            if (!rc) rc = launchFast(c, f0, n);
            if (!rc) rc = launchOctree(c, f0, n);
            if (!rc) rc = launchBlur(c, f0, n);
        }
        if (!rc) rc = launchDescribe(c, f0, n, dK, dD, sc, dN, dense);
        return rc;
        };
        const bool graphable = batch == 1 && !inDev && !outDev && !c->profiling && !c->frameGraphOff && !dbgTl;
        if (graphable) {
            const unsigned long long key[6] = {(unsigned long long)c->geomStamp, (unsigned long long)(uintptr_t)dK, (unsigned long long)(uintptr_t)dD,
                                               (unsigned long long)(uintptr_t)dN, (unsigned long long)sc,
                                               (unsigned long long)(uintptr_t)c->lvl0 ^ ((unsigned long long)c->lvl0Pitch << 48)};
            if (c->frameGraph && !memcmp(key, c->frameGraphKey, sizeof(key))) {
                if (cudaGraphLaunch(c->frameGraph, c->stream) != cudaSuccess) { cudaGetLastError(); c->frameGraphOff = true; rc = runStages(); }
                else c->kernelLaunches += c->frameGraphLaunches;
            } else if (!memcmp(key, c->frameGraphSeen, sizeof(key))) {          // second call with these buffers: capture
                if (c->frameGraph) { cudaGraphExecDestroy(c->frameGraph); c->frameGraph = nullptr; }
                const long long l0 = c->kernelLaunches;
                cudaGraph_t g = nullptr;
                bool ok = cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
                int crc = ok ? runStages() : ORBX_OK;
                if (ok) ok = cudaStreamEndCapture(c->stream, &g) == cudaSuccess && g && crc == ORBX_OK;
                const int captured = (int)(c->kernelLaunches - l0);
                c->kernelLaunches = l0;
                if (ok) ok = cudaGraphInstantiate(&c->frameGraph, g, 0) == cudaSuccess;
                if (g) cudaGraphDestroy(g);
                if (ok) ok = cudaGraphLaunch(c->frameGraph, c->stream) == cudaSuccess;
                if (ok) { memcpy(c->frameGraphKey, key, sizeof(key)); c->frameGraphLaunches = captured; c->kernelLaunches += captured; }
                else {                                                          // whatever went wrong: no graphs on this context, run as always
                    cudaGetLastError();
                    if (c->frameGraph) { cudaGraphExecDestroy(c->frameGraph); c->frameGraph = nullptr; }
                    c->frameGraphOff = true;
                    rc = runStages();
                }
            } else { memcpy(c->frameGraphSeen, key, sizeof(key)); rc = runStages(); }
        } else rc = runStages();
        if (rc) { c->stream = mainStream; return rc; }
        if (!outDev) {
            ORBX_CUDA(c, cudaEventRecord(c->evDone[i], c->stream));
            mark(c->stream);
            ORBX_CUDA(c, cudaStreamWaitEvent(c->sOut, c->evDone[i], 0));
            ORBX_CUDA(c, cudaMemcpyAsync(hN + f0, dN + f0, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, c->sOut));
            if (sc == cap) {
                ORBX_CUDA(c, cudaMemcpyAsync(kps + (size_t)f0 * cap, dK + (size_t)f0 * cap, sizeof(orbx_keypoint) * (size_t)cap * n,
                                             cudaMemcpyDeviceToHost, c->sOut));
                ORBX_CUDA(c, cudaMemcpyAsync(desc + (size_t)32 * f0 * cap, dD + (size_t)32 * f0 * cap, (size_t)32 * cap * n,
                                             cudaMemcpyDeviceToHost, c->sOut));
            } else {
                ORBX_CUDA(c, cudaMemcpy2DAsync(kps + (size_t)f0 * cap, sizeof(orbx_keypoint) * (size_t)cap, dK + (size_t)f0 * sc,
                                               sizeof(orbx_keypoint) * (size_t)sc, sizeof(orbx_keypoint) * (size_t)cap, (size_t)n,
                                               cudaMemcpyDeviceToHost, c->sOut));
                ORBX_CUDA(c, cudaMemcpy2DAsync(desc + (size_t)32 * f0 * cap, (size_t)32 * cap, dD + (size_t)32 * f0 * sc, (size_t)32 * sc,
                                               (size_t)32 * cap, (size_t)n, cudaMemcpyDeviceToHost, c->sOut));
            }
            mark(c->sOut);
        }
    }
    c->stream = mainStream;
    if (twoStreams) {   // later work on the context's stream sees both halves
        ORBX_CUDA(c, cudaEventRecord(c->evJoin, c->stream2));
        ORBX_CUDA(c, cudaStreamWaitEvent(c->stream, c->evJoin, 0));
    }
    if (outDev) return ORBX_OK; // asynchronous; orbx_sync() reports device-side errors
    ORBX_CUDA(c, cudaStreamSynchronize(c->sOut));
    if (dbgTl && !tl.empty()) {
        fprintf(stderr, "timeline (ms since start; per chunk: h2d done, kernels done, d2h done)\n");
        for (size_t k = 1; k < tl.size(); ++k) { float ms = 0; cudaEventElapsedTime(&ms, tl[0], tl[k]); fprintf(stderr, "%s%.3f", (k - 1) % 3 ? " " : "\n  ", ms); }
        fprintf(stderr, "\n");
        for (cudaEvent_t e : tl) cudaEventDestroy(e);
    }
    memcpy(n_out, hN, sizeof(int) * (size_t)batch);
    rc = checkDeviceError(c);
    if (rc) return rc;
    for (int b = 0; b < batch; ++b) if (n_out[b] > cap) return ORBX_E_CAPACITY;
    return ORBX_OK;
}

int orbx_cvt_gray(orbx_ctx* h, const uint8_t* src, int w, int hgt, size_t stride, int format, int flags, uint8_t* dst, size_t dst_stride)
{
    if (!h || !src || !dst || w < 1 || hgt < 1 || format < 0 || format > 3) return ORBX_E_INVALID;
    const int ch = (format & 2) ? 4 : 3;
    if (stride < (size_t)w * ch || dst_stride < (size_t)w) return ORBX_E_INVALID;
    Ctx* c = &h->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    const bool inDev = flags & ORBX_IN_DEVICE, outDev = flags & ORBX_OUT_DEVICE;
    const size_t inPitch = ((size_t)w * ch + 15) & ~(size_t)15, outPitch = ((size_t)w + 15) & ~(size_t)15;
    const size_t inBytes = inDev ? 0 : inPitch * hgt, outBytes = outDev ? 0 : outPitch * hgt;
    int rc = ensureScratch(c, inBytes + outBytes + 256);
    if (rc) return rc;
    uint8_t* S = (uint8_t*)c->dScratch;
    const uint8_t* dS = src; size_t sP = stride;
    if (!inDev) {
        ORBX_CUDA(c, cudaMemcpy2DAsync(S, inPitch, src, stride, (size_t)w * ch, (size_t)hgt, cudaMemcpyHostToDevice, c->stream));
        dS = S; sP = inPitch;
    }
    uint8_t* dD = outDev ? dst : S + inBytes; const size_t dP = outDev ? dst_stride : outPitch;
    rc = launchCvtGray(c, dS, sP, w, hgt, format, dD, dP);
    if (rc) return rc;
    if (!outDev) {
        ORBX_CUDA(c, cudaMemcpy2DAsync(dst, dst_stride, dD, dP, (size_t)w, (size_t)hgt, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return ORBX_OK;
}

int orbx_set_chunk(orbx_ctx* h, int frames)
{
    if (!h || frames < 0) return ORBX_E_INVALID;
    h->c.chunk = frames;
    return ORBX_OK;
}

int orbx_extract(orbx_ctx* h, const uint8_t* img, int w, int hgt, size_t stride, int flags, orbx_keypoint* kps,
                 uint8_t* desc, int cap, int* n_out)
{
    return orbx_extract_batch(h, img, 1, w, hgt, stride, stride * (size_t)(hgt > 0 ? hgt : 0), flags, kps, desc, cap, n_out);
}

// Every level of `frame` with its border, packed: level l starts at offsets[l] (bytes), rows of w_l + 2 * border bytes.
// dst == NULL: only offsets[] and *total_bytes are filled (sizing call).  One kernel, one copy, one synchronisation.
int orbx_pyramid_bordered(orbx_ctx* h, int frame, int border, int flags, uint8_t* dst, size_t dst_bytes, size_t* offsets, size_t* total_bytes)
{
    if (!h) return ORBX_E_INVALID;
    Ctx* c = &h->c;
    if (c->curW <= 0 || border < 0 || border > 64) return ORBX_E_INVALID;
    size_t total = 0;
    int rc = launchBorderedPyramid(c, 0, border, nullptr, offsets, &total);
    if (rc) return rc;
    if (total_bytes) *total_bytes = total;
    if (!dst) return ORBX_OK;
    if (frame < 0 || frame >= c->lastBatch) return ORBX_E_INVALID;
    if (dst_bytes < total) return ORBX_E_CAPACITY;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    if (flags & ORBX_OUT_DEVICE) return launchBorderedPyramid(c, frame, border, dst, nullptr, nullptr);
    if ((rc = ensureScratch(c, total))) return rc;
    if ((rc = ensureMatchPinned(c, total))) return rc;
    if ((rc = launchBorderedPyramid(c, frame, border, (uint8_t*)c->dScratch, nullptr, nullptr))) return rc;
    ORBX_CUDA(c, cudaMemcpyAsync(c->hMatch, c->dScratch, total, cudaMemcpyDeviceToHost, c->stream));   // pinned landing zone
    ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    memcpy(dst, c->hMatch, total);
    return ORBX_OK;
}

int orbx_pyramid_level(orbx_ctx* h, int frame, int level, int border, int flags, uint8_t* dst, size_t dst_stride, int* w, int* hh)
{
    if (!h || !dst) return ORBX_E_INVALID;
    Ctx* c = &h->c;
    if (level < 0 || level >= c->nlevels || frame < 0 || frame >= c->lastBatch || border < 0 || border > 64) return ORBX_E_INVALID;
    const LevelGeom& q = c->geom.lv[level];
    const int ow = q.w + 2 * border, oh = q.h + 2 * border;
    if (dst_stride < (size_t)ow) return ORBX_E_INVALID;
    if (w) *w = ow;
    if (hh) *hh = oh;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    const uint8_t* src; size_t pitch;
    if (level == 0) { src = c->lvl0 + (size_t)frame * c->lvl0FrameStride; pitch = c->lvl0Pitch; }
    else { src = c->dPyr + q.planeOff + (size_t)frame * q.pitch * q.h; pitch = (size_t)q.pitch; }
    if (border == 0) {
        ORBX_CUDA(c, cudaMemcpy2DAsync(dst, dst_stride, src, pitch, (size_t)q.w, (size_t)q.h,
                                       (flags & ORBX_OUT_DEVICE) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
        return ORBX_OK;
    }
    if (flags & ORBX_OUT_DEVICE) {
        int rc = launchBorderCopy(c, src, q.w, q.h, pitch, dst, dst_stride, border);
        if (rc) return rc;
        ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
        return ORBX_OK;
    }
    int rc = ensureScratch(c, (size_t)ow * oh);
    if (rc) return rc;
    if ((rc = launchBorderCopy(c, src, q.w, q.h, pitch, (uint8_t*)c->dScratch, (size_t)ow, border))) return rc;
    ORBX_CUDA(c, cudaMemcpy2DAsync(dst, dst_stride, c->dScratch, (size_t)ow, (size_t)ow, (size_t)oh, cudaMemcpyDeviceToHost, c->stream));
    ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    return ORBX_OK;
}

} // extern "C"
