// Frame::ComputeStereoMatches (src/Frame.cc:831-1128) on the GPU.
//
// Left keypoints are independent (the reference never masks right keypoints), so: one warp per left keypoint.
//   1. candidates = right keypoints whose row band [floor(y-r), ceil(y+r)], r = 2*scale[octave] (:869-888) contains
//      row (int)vL, with |octave difference| <= 1 and uR in [uL - maxD, uL]; best Hamming distance below TH_HIGH,
//      first candidate (lowest right index: vRowIndices is filled in index order) wins ties (:937-978).
//      The row table is replaced by testing every right keypoint (N_r tests per warp; cheaper than building it).
//   2. if best < 75: 11 SAD windows of 11x11 centre-subtracted pixels on the two extractors' device-resident
//      pyramid levels (:981-1058; integer arithmetic, the reference's float values are exact integers),
//      parabola refinement and disparity gates in the reference's float order (:1061-1105).
//   3. stereoMedianKernel: sort-free median of the SAD scores (element nvd/2 of the (SAD, index) order) and the
//      1.5*1.4*median rejection (:1112-1127).
#include "orbx_internal.h"
#include <algorithm>

namespace orbx {

struct PyrView { const uint8_t* base[kMaxLevels]; size_t frameStride[kMaxLevels]; int pitch[kMaxLevels], w[kMaxLevels], h[kMaxLevels];
                 float scale[kMaxLevels], inv[kMaxLevels]; int nlevels; };

// grid = (ceil(cap / 4), pairs): pair p = blockIdx.y uses rows [p * cap, p * cap + n[p]) of the keypoint / descriptor arrays
// (the layout orbx_extract_batch writes) and frame frame0 + p of both pyramids.
__global__ void __launch_bounds__(128)
stereoMatchKernel(const orbx_keypoint* __restrict__ kLall, const uint4* __restrict__ dLall, const int* __restrict__ nL,
                  const orbx_keypoint* __restrict__ kRall, const uint4* __restrict__ dRall, const int* __restrict__ nR, int cap, int frame0,
                  const __grid_constant__ PyrView L, const __grid_constant__ PyrView R, float mbf, float mb, float* __restrict__ uRightAll,
                  float* __restrict__ depthAll, int* __restrict__ sadListAll, int* __restrict__ sadCountAll)
{
    const int lane = threadIdx.x & 31, pair = blockIdx.y;
    const int iL = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int N = min(nL[pair], cap), Nr = min(nR[pair], cap);
    if (iL >= N) return;
    const orbx_keypoint* kL = kLall + (size_t)pair * cap; const uint4* dL = dLall + 2 * (size_t)pair * cap;
    const orbx_keypoint* kR = kRall + (size_t)pair * cap; const uint4* dR = dRall + 2 * (size_t)pair * cap;
    float* uRight = uRightAll + (size_t)pair * cap; float* depth = depthAll + (size_t)pair * cap;
    int* sadList = sadListAll + 2 * (size_t)pair * cap; int* sadCount = sadCountAll + 2 * pair;
    const int frame = frame0 + pair;
    const orbx_keypoint kp = kL[iL];
    if (lane == 0) { uRight[iL] = -1.0f; depth[iL] = -1.0f; }
    const int levelL = kp.octave;
    if (levelL < 0 || levelL >= L.nlevels) return;            // a keypoint that did not come from this extractor: no match
    const float vL = kp.y, uL = kp.x;
    const float maxD = __fdiv_rn(mbf, mb), minD = 0.f;
    const float minU = __fsub_rn(uL, maxD), maxU = __fsub_rn(uL, minD);
    if (maxU < 0) return;
    const int row = (int)vL;
    const uint4 a0 = dL[2 * iL], a1 = dL[2 * iL + 1];
    uint32_t best = 0xffffffffu;
    for (int iR = lane; iR < Nr; iR += 32) {
        const orbx_keypoint r = kR[iR];
        if (r.octave < 0 || r.octave >= R.nlevels) continue;
        const float rr = __fmul_rn(2.0f, R.scale[r.octave]);
        const int maxr = (int)ceilf(__fadd_rn(r.y, rr)), minr = (int)floorf(__fsub_rn(r.y, rr));
        if (row < minr || row > maxr) continue;
        if (r.octave < levelL - 1 || r.octave > levelL + 1) continue;
        if (!(r.x >= minU && r.x <= maxU)) continue;
        const uint4 b0 = dR[2 * iR], b1 = dR[2 * iR + 1];
        const int d = hamming256(a0, a1, b0, b1);
        best = min(best, ((uint32_t)d << 20) | (uint32_t)iR);
    }
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    const int bestDist = (int)(best >> 20);
    if (best == 0xffffffffu || bestDist >= 100) return;       // bestDist starts at TH_HIGH, strict '<' (:941,:966)
    if (!(bestDist < 75)) return;                             // thOrbDist = (TH_HIGH+TH_LOW)/2 (:906,:981)
    const int bestIdxR = (int)(best & 0xfffffu);
    const float uR0 = kR[bestIdxR].x;
    const float sf = L.inv[levelL];
    const float scaleduL = roundf(__fmul_rn(kp.x, sf)), scaledvL = roundf(__fmul_rn(kp.y, sf)), scaleduR0 = roundf(__fmul_rn(uR0, sf));
    const int w = 5, Lw = 5;
    const float iniu = __fsub_rn(__fsub_rn(scaleduR0, (float)Lw), (float)w), endu = __fadd_rn(__fadd_rn(__fadd_rn(scaleduR0, (float)Lw), (float)w), 1.f);
    if (iniu < 0 || endu >= (float)R.w[levelL]) return;
    const int y0 = (int)__fsub_rn(scaledvL, (float)w), xl0 = (int)__fsub_rn(scaleduL, (float)w);
    // the reference's rowRange / colRange would assert here; keypoints of this extractor are always >= 19 px inside
    if (y0 < 0 || y0 + 2 * w >= L.h[levelL] || xl0 < 0 || xl0 + 2 * w >= L.w[levelL]) return;
    const uint8_t* IL = L.base[levelL] + (size_t)frame * L.frameStride[levelL]; const int pl = L.pitch[levelL];
    const uint8_t* IR = R.base[levelL] + (size_t)frame * R.frameStride[levelL]; const int pr = R.pitch[levelL];
    const int cL = IL[(size_t)(y0 + w) * pl + xl0 + w];
    int sums[11];
    #pragma unroll
    for (int k = 0; k < 11; ++k) sums[k] = 0;
    int xr0[11], cR[11];
    #pragma unroll
    for (int k = 0; k < 11; ++k) {
        xr0[k] = (int)__fsub_rn(__fadd_rn(scaleduR0, (float)(k - Lw)), (float)w);
        cR[k] = IR[(size_t)(y0 + w) * pr + xr0[k] + w];
    }
    for (int p = lane; p < 121; p += 32) {
        const int yy = p / 11, xx = p - yy * 11;
        const int a = (int)IL[(size_t)(y0 + yy) * pl + xl0 + xx] - cL;
        const uint8_t* rrow = IR + (size_t)(y0 + yy) * pr + xx;
        #pragma unroll
        for (int k = 0; k < 11; ++k) sums[k] += abs(a - ((int)rrow[xr0[k]] - cR[k]));
    }
    #pragma unroll
    for (int k = 0; k < 11; ++k) {
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) sums[k] += __shfl_xor_sync(0xffffffffu, sums[k], o);
    }
    if (lane != 0) return;
    int bestSad = 0x7fffffff, bestinc = 0;
    #pragma unroll
    for (int k = 0; k < 11; ++k) if ((float)sums[k] < (float)bestSad) { bestSad = sums[k]; bestinc = k - Lw; }
    if (bestinc == -Lw || bestinc == Lw) return;
    float d1 = 0.f, d2 = 0.f, d3 = 0.f;
    #pragma unroll
    for (int k = 1; k < 10; ++k) if (k - Lw == bestinc) { d1 = (float)sums[k - 1]; d2 = (float)sums[k]; d3 = (float)sums[k + 1]; }
    const float deltaR = __fdiv_rn(__fsub_rn(d1, d3), __fmul_rn(2.0f, __fsub_rn(__fadd_rn(d1, d3), __fmul_rn(2.0f, d2))));
    if (deltaR < -1 || deltaR > 1) return;
    float bestuR = __fmul_rn(L.scale[levelL], __fadd_rn(__fadd_rn(scaleduR0, (float)bestinc), deltaR));
    float disparity = __fsub_rn(uL, bestuR);
    if (disparity >= minD && disparity < maxD) {
        if (disparity <= 0) { disparity = 0.01f; bestuR = (float)__dsub_rn((double)uL, 0.01); }
        depth[iL] = __fdiv_rn(mbf, disparity);
        uRight[iL] = bestuR;
        const int slot = atomicAdd(sadCount, 1);
        sadList[2 * slot] = bestSad; sadList[2 * slot + 1] = iL;
    }
}

// one CTA per pair.  sadCount[2 * pair] = matches before the rejection, sadCount[2 * pair + 1] = matches that keep their depth
__global__ void __launch_bounds__(1024)
stereoMedianKernel(const int* __restrict__ sadListAll, int* __restrict__ sadCountAll, int cap, float* __restrict__ uRightAll,
                   float* __restrict__ depthAll)
{
    __shared__ float thDist;
    __shared__ int kept;
    const int pair = blockIdx.x;
    const int* sadList = sadListAll + 2 * (size_t)pair * cap;
    float* uRight = uRightAll + (size_t)pair * cap; float* depth = depthAll + (size_t)pair * cap;
    const int n = sadCountAll[2 * pair];
    if (threadIdx.x == 0) { kept = 0; sadCountAll[2 * pair + 1] = 0; }
    if (n == 0) return;
    __syncthreads();
    const int k = n / 2;
    // element k of the ascending (SAD, index) order; only its SAD is used
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int v = sadList[2 * i];
        int less = 0, leq = 0;
        for (int j = 0; j < n; ++j) { const int u = sadList[2 * j]; less += u < v; leq += u <= v; }
        if (less <= k && k < leq) thDist = __fmul_rn(1.5f * 1.4f, (float)v);
    }
    __syncthreads();
    const float th = thDist;
    int mine = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        if (!((float)sadList[2 * i] < th)) { const int iL = sadList[2 * i + 1]; uRight[iL] = -1.f; depth[iL] = -1.f; }
        else ++mine;
    }
    if (mine) atomicAdd(&kept, mine);
    __syncthreads();
    if (threadIdx.x == 0) sadCountAll[2 * pair + 1] = kept;
}

static void fillView(const Ctx* c, PyrView& v)
{
    memset(&v, 0, sizeof(v));
    v.nlevels = c->nlevels;
    for (int l = 0; l < c->nlevels; ++l) {
        const LevelGeom& q = c->geom.lv[l];
        if (l == 0) { v.base[0] = c->lvl0; v.frameStride[0] = c->lvl0FrameStride; v.pitch[0] = (int)c->lvl0Pitch; }
        else { v.base[l] = c->dPyr + q.planeOff; v.frameStride[l] = (size_t)q.pitch * q.h; v.pitch[l] = q.pitch; }
        v.w[l] = q.w; v.h[l] = q.h; v.scale[l] = c->scale[l]; v.inv[l] = c->invScale[l];
    }
}

} // namespace orbx
using namespace orbx;

// Frame::ComputeStereoMatches for `pairs` stereo pairs: pair p = frames frame0 + p of the two contexts' LAST extract calls and
// rows [p * cap, p * cap + n[p]) of the keypoint / descriptor arrays, i.e. exactly what orbx_extract_batch wrote for the left
// and the right batch.  With ORBX_IN_DEVICE every input (the counts too) is a device pointer, so the call chains behind two
// orbx_extract_batch(..., ORBX_OUT_DEVICE) calls without a host round trip.
// hostNL: the left counts when the host knows them (host arrays, or the single-pair call): host outputs then receive exactly
// n_l[p] values per pair; otherwise the padded pairs x cap block is copied.
static int stereoBatchImpl(orbx_ctx* hl, orbx_ctx* hr, int frame0, int pairs, const orbx_keypoint* kps_l, const uint8_t* desc_l,
                           const int32_t* n_l, const orbx_keypoint* kps_r, const uint8_t* desc_r, const int32_t* n_r, int cap,
                           float bf, float b, int flags, float* u_right, float* depth, int32_t* n_matched, const int32_t* hostNL)
{
    if (!hl || !hr || pairs < 0 || cap < 1 || frame0 < 0) return ORBX_E_INVALID;
    if (pairs == 0) return ORBX_OK;
    if (!kps_l || !desc_l || !n_l || !kps_r || !desc_r || !n_r || !u_right || !depth) return ORBX_E_INVALID;
    if (cap >= (1 << 20)) return ORBX_E_CAPACITY;             // the candidate key keeps the right index in 20 bits
    Ctx* c = &hl->c; Ctx* cr = &hr->c;
    if (c->device != cr->device || c->nlevels != cr->nlevels || frame0 + pairs > c->lastBatch || frame0 + pairs > cr->lastBatch ||
        c->curW != cr->curW || c->curH != cr->curH)
        return ORBX_E_INVALID;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    ORBX_CUDA(c, cudaStreamSynchronize(cr->stream)); // the right pyramid is produced on the other ctx's stream
    const bool inDev = flags & ORBX_IN_DEVICE, outDev = flags & ORBX_OUT_DEVICE;
    const size_t rows = (size_t)pairs * cap;
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t oList = take(sizeof(int) * 2 * rows), oCnt = take(sizeof(int) * 2 * (size_t)pairs);
    size_t oKL = 0, oDL = 0, oKR = 0, oDR = 0, oNL = 0, oNR = 0, oU = 0, oZ = 0;
    if (!inDev) {
        oKL = take(sizeof(orbx_keypoint) * rows); oDL = take(32 * rows); oKR = take(sizeof(orbx_keypoint) * rows); oDR = take(32 * rows);
        oNL = take(sizeof(int) * (size_t)pairs); oNR = take(sizeof(int) * (size_t)pairs);
    }
    if (!outDev) { oU = take(sizeof(float) * rows); oZ = take(sizeof(float) * rows); }
    int rc = ensureScratch(c, off);
    if (rc) return rc;
    uint8_t* S = (uint8_t*)c->dScratch;
    const orbx_keypoint *dKL = kps_l, *dKR = kps_r; const uint4 *dDL = (const uint4*)desc_l, *dDR = (const uint4*)desc_r;
    const int *dNL = n_l, *dNR = n_r;
    stageBegin(c, ST_H2D);
    if (!inDev) {
        for (int p = 0; p < pairs; ++p) if (n_l[p] < 0 || n_r[p] < 0 || n_l[p] > cap || n_r[p] > cap) return ORBX_E_INVALID;
        // only the rows in use cross the bus: one copy per pair and array for few pairs, the padded block for many
        auto up = [&](size_t o, const void* src, size_t rowBytes, const int32_t* n) -> int {
            if (pairs > 16) { ORBX_CUDA(c, cudaMemcpyAsync(S + o, src, rowBytes * rows, cudaMemcpyHostToDevice, c->stream)); return ORBX_OK; }
            for (int p = 0; p < pairs; ++p)
                if (n[p] > 0) ORBX_CUDA(c, cudaMemcpyAsync(S + o + rowBytes * (size_t)p * cap, (const uint8_t*)src + rowBytes * (size_t)p * cap,
                                                          rowBytes * (size_t)n[p], cudaMemcpyHostToDevice, c->stream));
            return ORBX_OK;
        };
        if ((rc = up(oKL, kps_l, sizeof(orbx_keypoint), n_l)) || (rc = up(oDL, desc_l, 32, n_l)) ||
            (rc = up(oKR, kps_r, sizeof(orbx_keypoint), n_r)) || (rc = up(oDR, desc_r, 32, n_r))) return rc;
        ORBX_CUDA(c, cudaMemcpyAsync(S + oNL, n_l, sizeof(int) * (size_t)pairs, cudaMemcpyHostToDevice, c->stream));
        ORBX_CUDA(c, cudaMemcpyAsync(S + oNR, n_r, sizeof(int) * (size_t)pairs, cudaMemcpyHostToDevice, c->stream));
        dKL = (const orbx_keypoint*)(S + oKL); dDL = (const uint4*)(S + oDL); dKR = (const orbx_keypoint*)(S + oKR); dDR = (const uint4*)(S + oDR);
        dNL = (const int*)(S + oNL); dNR = (const int*)(S + oNR);
    }
    stageEnd(c, ST_H2D, 0);
    float* dU = outDev ? u_right : (float*)(S + oU);
    float* dZ = outDev ? depth : (float*)(S + oZ);
    int* dCnt = (int*)(S + oCnt);
    PyrView L, R;
    fillView(c, L); fillView(cr, R);
    ORBX_CUDA(c, cudaMemsetAsync(dCnt, 0, sizeof(int) * 2 * (size_t)pairs, c->stream));
    stageBegin(c, ST_STEREO);
    stereoMatchKernel<<<dim3((cap + 3) / 4, pairs), 128, 0, c->stream>>>(dKL, dDL, dNL, dKR, dDR, dNR, cap, frame0, L, R, bf, b, dU, dZ,
                                                                      (int*)(S + oList), dCnt);
    stereoMedianKernel<<<pairs, 1024, 0, c->stream>>>((int*)(S + oList), dCnt, cap, dU, dZ);
    stageEnd(c, ST_STEREO, 2);
    ORBX_CUDA(c, cudaGetLastError());
    if (!outDev) {
        if (hostNL && pairs <= 16) {
            for (int p = 0; p < pairs; ++p) {
                const size_t n = (size_t)std::min(std::max(hostNL[p], 0), cap);
                if (!n) continue;
                ORBX_CUDA(c, cudaMemcpyAsync(u_right + (size_t)p * cap, dU + (size_t)p * cap, sizeof(float) * n, cudaMemcpyDeviceToHost, c->stream));
                ORBX_CUDA(c, cudaMemcpyAsync(depth + (size_t)p * cap, dZ + (size_t)p * cap, sizeof(float) * n, cudaMemcpyDeviceToHost, c->stream));
            }
        } else {
            ORBX_CUDA(c, cudaMemcpyAsync(u_right, dU, sizeof(float) * rows, cudaMemcpyDeviceToHost, c->stream));
            ORBX_CUDA(c, cudaMemcpyAsync(depth, dZ, sizeof(float) * rows, cudaMemcpyDeviceToHost, c->stream));
        }
    }
    if (n_matched) {
        std::vector<int> cnt(2 * (size_t)pairs);
        ORBX_CUDA(c, cudaMemcpyAsync(cnt.data(), dCnt, sizeof(int) * 2 * (size_t)pairs, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
        for (int p = 0; p < pairs; ++p) n_matched[p] = cnt[2 * p + 1];
    } else if (!outDev) ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    return ORBX_OK;
}

extern "C" int orbx_stereo_match_batch(orbx_ctx* hl, orbx_ctx* hr, int frame0, int pairs, const orbx_keypoint* kps_l, const uint8_t* desc_l,
                                       const int32_t* n_l, const orbx_keypoint* kps_r, const uint8_t* desc_r, const int32_t* n_r, int cap,
                                       float bf, float b, int flags, float* u_right, float* depth, int32_t* n_matched)
{
    return stereoBatchImpl(hl, hr, frame0, pairs, kps_l, desc_l, n_l, kps_r, desc_r, n_r, cap, bf, b, flags, u_right, depth, n_matched,
                           (flags & ORBX_IN_DEVICE) ? nullptr : n_l);
}

// One pair (Frame::Frame(imLeft, imRight, ...), src/Frame.cc:119-217): the batch call with a single pair whose arrays hold n rows.
extern "C" int orbx_stereo_match(orbx_ctx* hl, orbx_ctx* hr, int frame, const orbx_keypoint* kps_l, const uint8_t* desc_l, int n_l,
                                 const orbx_keypoint* kps_r, const uint8_t* desc_r, int n_r, float bf, float b, int flags,
                                 float* u_right, float* depth, int* n_matched)
{
    if (!hl || !hr || n_l < 0 || n_r < 0 || (n_l > 0 && (!kps_l || !desc_l || !u_right || !depth)) || (n_r > 0 && (!kps_r || !desc_r)))
        return ORBX_E_INVALID;
    if (n_matched) *n_matched = 0;
    if (n_l == 0) return ORBX_OK;
    if (n_r >= (1 << 20) || n_l >= (1 << 20)) return ORBX_E_CAPACITY;
    Ctx* c = &hl->c;
    // the counts of a single pair are host values whatever the flags say: hand them to the kernel through the scratch tail
    const int cap = std::max(std::max(n_l, n_r), 1);
    static const orbx_keypoint kNone = {0, 0, 0, 0, 0, 0, 0};
    static const uint8_t dNone[32] = {0};
    if (n_r == 0) { kps_r = (flags & ORBX_IN_DEVICE) ? kps_l : &kNone; desc_r = (flags & ORBX_IN_DEVICE) ? desc_l : dNone; }
    if (!(flags & ORBX_IN_DEVICE)) {
        // host arrays may be shorter than cap rows: per-pair copies (pairs == 1) move exactly n rows
        return stereoBatchImpl(hl, hr, frame, 1, kps_l, desc_l, &n_l, kps_r, desc_r, &n_r, cap, bf, b, flags, u_right, depth, n_matched, &n_l);
    }
    ORBX_CUDA(c, cudaSetDevice(c->device));
    int rc = ensureMatchPinned(c, 64);
    if (rc) return rc;
    int* dN = nullptr;
    if (!c->dStereoN) ORBX_CUDA(c, cudaMalloc(&c->dStereoN, 2 * sizeof(int)));
    dN = c->dStereoN;
    const int hn[2] = {n_l, n_r};
    ORBX_CUDA(c, cudaMemcpyAsync(dN, hn, sizeof(hn), cudaMemcpyHostToDevice, c->stream));
    ORBX_CUDA(c, cudaStreamSynchronize(c->stream));            // hn lives on this stack frame
    return stereoBatchImpl(hl, hr, frame, 1, kps_l, desc_l, dN, kps_r, desc_r, dN + 1, cap, bf, b, flags, u_right, depth, n_matched, &n_l);
}
