// Internal declarations shared by the orbx CUDA translation units (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <string>
#include <vector>
#include "../../include/orbx.h"

namespace orbx {

constexpr int kEdge = 19;        // EDGE_THRESHOLD, src/ORBextractor.cc:81
constexpr int kHalfPatch = 15;   // HALF_PATCH_SIZE :80
constexpr int kPatch = 31;       // PATCH_SIZE :79
constexpr int kWinOff = 16;      // minBorderX = EDGE_THRESHOLD-3 :1061
constexpr int kMaxLevels = ORBX_MAX_LEVELS;
constexpr int kBlurTW = 128, kBlurTH = 160;   // blur tile: columns per CTA, max rows per CTA (k_blur.cu)

enum Stage { ST_PYRAMID = 0, ST_FAST, ST_OCTREE, ST_BLUR, ST_DESCRIBE, ST_H2D, ST_D2H, ST_HAMMING, ST_MATCH, ST_STEREO, ST_BOW, ST_COUNT };

// Per-level geometry, identical for every frame of a batch (device copy lives in __constant__).
struct LevelGeom {
    int w, h;            // level size (:1672)
    int pitch;           // bytes per row of the internal plane (multiple of 16)
    int nCols, nRows;    // FAST cell grid (:1076-1077); 0 when the window is narrower than a cell
    int wCell, hCell;    // :1080-1081
    int cellBase;        // first cell index of this level in the per-frame cell table
    int candCap;         // candidate capacity per frame (can never overflow, see host)
    int candBase;        // offset (in records) of this level inside a frame's candidate block
    int quota;           // mnFeaturesPerLevel[level] :521-531
    int nIni;            // quadtree roots :703
    float hX;            // root width :706
    int outCap;          // quota + slack: max nodes the quadtree can return
    int outBase;         // offset of this level inside a frame's selected-keypoint block
    float scale;         // mvScaleFactor[level]
    float patchSize;     // (int)(31*scale) as float :1168
    size_t planeOff;     // byte offset of this level's [B][h][pitch] block inside the pyramid arena
    size_t blurOff;      // same for the blurred arena
    int tileBase, tilesX, tilesY; // blur tile table
    int stripBase, segsPerRow, cellsPerSeg; // FAST strips: one cell-row, split into segments of whole cells
};

// One FAST strip (k_fast.cu): a = level | cells<<8 | cellRow<<16, b = firstCell | x0<<16, c = y0 | width<<16,
// d = height | (65536 / wCell + 1)<<16; x0/y0 = first candidate pixel in level coordinates.
struct StripDesc { uint32_t a, b, c, d; };

struct ResizeTap { uint16_t s; int16_t c0, c1; uint16_t pad; }; // source index + 11-bit fixed-point weights

struct GeomSet {            // what kernels need, uploaded to constant memory on (re)configuration
    LevelGeom lv[kMaxLevels];
    int nlevels;
    int cellsPerFrame;
    int candPerFrame;
    int selPerFrame;
    int blurTilesPerFrame;
    int stripsPerFrame, fastTilePitch, fastTileRows, fastScorePitch;
    int frame0;                 // first frame of this launch (chunked pipeline); frame = blockIdx.y + frame0
    int iniTh, minTh;
    int umax[16];
};

// kernels with more than 48 KB of dynamic shared memory: cudaFuncSetAttribute is per device, so what has been configured is
// remembered per context (a context belongs to one device and one caller), not in a process-wide static
enum { SMEM_FAST128, SMEM_FAST256, SMEM_DESCRIBE, SMEM_OCTREE, SMEM_BOW_RESOLVE, SMEM_BOW_VECTOR, SMEM_RESIZE, SMEM_SLOTS };

struct Ctx {
    int device = 0;
    size_t smemCfg[SMEM_SLOTS] = {};
    cudaStream_t stream = nullptr;
    cudaStream_t sIn = nullptr, sOut = nullptr;   // copy streams of the chunked pipeline (H2D / D2H overlap compute)
    cudaStream_t stream2 = nullptr;               // odd chunks of the pipeline run here, so kernels of neighbouring chunks fill each other's tails
    cudaEvent_t evJoin = nullptr;
    cudaStream_t sBlur = nullptr; cudaEvent_t evFork = nullptr, evBlur = nullptr;   // blur next to FAST + quadtree (device-resident batches)
    bool blurAfterFast = false;                    // ORBX_OVERLAP_BLUR=2: the blur forks after FAST and runs beside the quadtree only
    bool overlapBlur = false;                      // env ORBX_OVERLAP_BLUR=1: no gain at 640x480 (the kernels are throughput-bound), ~4 % at 1241x376
    std::vector<cudaEvent_t> evIn, evDone;
    cudaEvent_t evFence = nullptr;                // orders a new call's H2D after the previous call's kernels
    int fastThreads = 128, fastSegW = 320;         // FAST strip shape: threads per CTA, max candidate columns per strip
    int chunk = 64;                                // frames per pipeline chunk (stages of a chunk stay L2-resident)
    int nfeatures = 0, nlevels = 0, iniTh = 0, minTh = 0;
    float scaleFactor = 0;
    float scale[kMaxLevels], invScale[kMaxLevels], sigma2[kMaxLevels], invSigma2[kMaxLevels];
    int nfeatLevel[kMaxLevels];
    int umax[16];
    int maxW = 0, maxH = 0, maxBatch = 0;
    // current geometry
    int curW = 0, curH = 0;
    GeomSet geom;
    GeomSet* dGeom = nullptr;            // device copy (global memory; kernels read via __ldg / constant cache)
    ResizeTap* dTapX[kMaxLevels] = {};   // per-level tables (level l = taps from level l-1)
    ResizeTap* dTapY[kMaxLevels] = {};
    uint32_t* dTapG[kMaxLevels] = {};    // staged resize: per 8-pixel group [s0 | 4 selectors | 4 selectors | 8 x (c0, c1)], SoA over groups
    StripDesc* dStrips = nullptr; int nStrips = 0, stripCap = 0;   // FAST strip table of the current geometry
    int geomStamp = 0;                   // bumped whenever the geometry / arenas change
    void* dTensorMaps = nullptr; int mapsStamp = -1;   // CUtensorMap[2][kMaxLevels] of the blurred / pyramid planes (k_describe.cu)
    bool resizeWordOk[kMaxLevels] = {};  // level l: every 4-pixel group reads <= 8 source bytes from an aligned word (k_pyramid.cu fast path)
    bool resize8Ok[kMaxLevels] = {};     // level l: in every 8-pixel group pixels 0-3 start 0..6 and pixels 4-7 start 4..10 bytes after pixel 0 (staged path)
    bool resizeWide = true;              // env ORBX_RESIZE_WIDE=0 keeps the register-only 4-pixel kernel (A/B measurements)
    // arenas
    uint8_t* dPyr = nullptr; size_t pyrBytes = 0;      // levels (level 0 too unless used in place)
    uint8_t* dBlur = nullptr; size_t blurBytes = 0;
    unsigned long long* dCand = nullptr; size_t candRecs = 0; // [B][candPerFrame]
    int* dCandCount = nullptr;           // [B][nlevels]
    orbx_keypoint* dSel = nullptr; size_t selRecs = 0; // [B][selPerFrame] level-coord keypoints after the quadtree
    int* dSelCount = nullptr;            // [B][nlevels]
    int* dErr = nullptr;                 // device error flag
    int* dFrameOff = nullptr;            // ORBX_OUT_DENSE: first output row of every frame (+ total), maxBatch + 1 entries
    orbx_keypoint* dOutKp = nullptr; uint8_t* dOutDesc = nullptr; int* dOutN = nullptr; int outCap = 0; // staging for host outputs
    uint8_t* hPinned = nullptr; size_t hPinnedBytes = 0;
    uint8_t* dIn = nullptr; size_t dInBytes = 0;       // H2D staging for level 0 when input is on host
    const uint8_t* lvl0 = nullptr; size_t lvl0Pitch = 0, lvl0FrameStride = 0; // where level 0 of the last batch lives
    int lastBatch = 0;
    // the stages of a single host frame as a CUDA graph (orbx_api.cu); env ORBX_FRAME_GRAPH=0 turns it off
    cudaGraphExec_t frameGraph = nullptr;
    unsigned long long frameGraphKey[6] = {}, frameGraphSeen[6] = {};
    int frameGraphLaunches = 0;
    bool frameGraphOff = false;
    // instrumentation
    bool profiling = false;
    double stageMs[ST_COUNT] = {};
    long long stageLaunches[ST_COUNT] = {};
    long long kernelLaunches = 0;
    std::vector<std::pair<int, std::pair<cudaEvent_t, cudaEvent_t>>> pendingEvents;
    std::vector<cudaEvent_t> eventPool;
    std::string lastCudaError;
    // matcher scratch
    void* dScratch = nullptr; size_t scratchBytes = 0;
    int* dStereoN = nullptr;                       // the two counts of a single-pair orbx_stereo_match with device arrays
    uint8_t* hMatch = nullptr; size_t hMatchBytes = 0; // pinned staging for the small matcher inputs / results (one copy each way)
};

#ifdef __CUDACC__
// 256-bit Hamming distance (DescriptorDistance, src/ORBmatcher.cc:2123-2143).  POPC is a quarter-rate pipe, LOP3 a
// half-rate one: the eight XOR words go through a carry-save adder tree (each CSA = 2 LOP3) down to one word of weight 1,
// two of weight 2 and one of weight 4, so 4 POPCs are issued per pair instead of 8, and the weighted sum is IMADs (third pipe).
__device__ __forceinline__ int hamming256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1)
{
    const uint32_t x0 = a0.x ^ b0.x, x1 = a0.y ^ b0.y, x2 = a0.z ^ b0.z, x3 = a0.w ^ b0.w;
    const uint32_t x4 = a1.x ^ b1.x, x5 = a1.y ^ b1.y, x6 = a1.z ^ b1.z, x7 = a1.w ^ b1.w;
    const uint32_t s1 = x0 ^ x1 ^ x2, c1 = (x0 & x1) | (x2 & (x0 | x1));
    const uint32_t s2 = x3 ^ x4 ^ x5, c2 = (x3 & x4) | (x5 & (x3 | x4));
    const uint32_t s3 = s1 ^ s2 ^ x6, c3 = (s1 & s2) | (x6 & (s1 | s2));
    const uint32_t ones = s3 ^ x7, c4 = s3 & x7;
    const uint32_t t1 = c1 ^ c2 ^ c3, d1 = (c1 & c2) | (c3 & (c1 | c2));
    return __popc(ones) + 2 * (__popc(t1) + __popc(c4)) + 4 * __popc(d1);
}
#endif

// RAII-free helpers -----------------------------------------------------------------------------
int cudaFail(Ctx* c, cudaError_t e, const char* what);
#define ORBX_CUDA(c, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return orbx::cudaFail((c), e_, #call); } while (0)
void stageBegin(Ctx* c, int stage);
void stageEnd(Ctx* c, int stage, int launches);
int ensureScratch(Ctx* c, size_t bytes);
int ensureMatchPinned(Ctx* c, size_t bytes);

// kernels (one launcher per stage; all asynchronous on c->stream) -------------------------------
// frames [frame0, frame0 + batch) of the current batch
int launchPyramid(Ctx* c, int frame0, int batch);
int launchCvtGray(Ctx* c, const uint8_t* src, size_t sPitch, int w, int h, int format, uint8_t* dst, size_t dPitch);
int launchFast(Ctx* c, int frame0, int batch);
void buildFastStrips(const GeomSet& g, std::vector<StripDesc>& out);
int launchOctree(Ctx* c, int frame0, int batch);
int launchBlur(Ctx* c, int frame0, int batch);
int launchDescribe(Ctx* c, int frame0, int batch, orbx_keypoint* dKps, uint8_t* dDesc, int capPerFrame, int* dN, bool dense = false);
int launchBorderedPyramid(Ctx* c, int frame, int border, uint8_t* dst, size_t* offsets, size_t* total);   // dst == NULL: sizes only
int launchBorderCopy(Ctx* c, const uint8_t* src, int w, int h, size_t pitch, uint8_t* dst, size_t dpitch, int border);

} // namespace orbx

struct orbx_ctx { orbx::Ctx c; };
