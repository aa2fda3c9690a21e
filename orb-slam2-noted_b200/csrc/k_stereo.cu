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

struct PyrView { const uint8_t* base[kMaxLevels]; int pitch[kMaxLevels], w[kMaxLevels], h[kMaxLevels]; float scale[kMaxLevels], inv[kMaxLevels]; };

__global__ void __launch_bounds__(128)
stereoMatchKernel(const orbx_keypoint* __restrict__ kL, const uint4* __restrict__ dL, int N, const orbx_keypoint* __restrict__ kR,
                  const uint4* __restrict__ dR, int Nr, PyrView L, PyrView R, float mbf, float mb, float* __restrict__ uRight,
                  float* __restrict__ depth, int* __restrict__ sadList, int* __restrict__ sadCount)
{
    const int lane = threadIdx.x & 31;
    const int iL = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (iL >= N) return;
    const orbx_keypoint kp = kL[iL];
    if (lane == 0) { uRight[iL] = -1.0f; depth[iL] = -1.0f; }
    const int levelL = kp.octave;
    const float vL = kp.y, uL = kp.x;
    const float maxD = __fdiv_rn(mbf, mb), minD = 0.f;
    const float minU = __fsub_rn(uL, maxD), maxU = __fsub_rn(uL, minD);
    if (maxU < 0) return;
    const int row = (int)vL;
    const uint4 a0 = dL[2 * iL], a1 = dL[2 * iL + 1];
    uint32_t best = 0xffffffffu;
    for (int iR = lane; iR < Nr; iR += 32) {
        const orbx_keypoint r = kR[iR];
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
    const uint8_t* IL = L.base[levelL]; const int pl = L.pitch[levelL];
    const uint8_t* IR = R.base[levelL]; const int pr = R.pitch[levelL];
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

__global__ void __launch_bounds__(1024)
stereoMedianKernel(const int* __restrict__ sadList, const int* __restrict__ sadCount, float* __restrict__ uRight, float* __restrict__ depth)
{
    __shared__ float thDist;
    const int n = *sadCount;
    if (n == 0) return;
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
    for (int i = threadIdx.x; i < n; i += blockDim.x)
        if (!((float)sadList[2 * i] < th)) { const int iL = sadList[2 * i + 1]; uRight[iL] = -1.f; depth[iL] = -1.f; }
}

static void fillView(const Ctx* c, int frame, PyrView& v)
{
    for (int l = 0; l < c->nlevels; ++l) {
        const LevelGeom& q = c->geom.lv[l];
        if (l == 0) { v.base[0] = c->lvl0 + (size_t)frame * c->lvl0FrameStride; v.pitch[0] = (int)c->lvl0Pitch; }
        else { v.base[l] = c->dPyr + q.planeOff + (size_t)frame * q.pitch * q.h; v.pitch[l] = q.pitch; }
        v.w[l] = q.w; v.h[l] = q.h; v.scale[l] = c->scale[l]; v.inv[l] = c->invScale[l];
    }
}

} // namespace orbx
using namespace orbx;

extern "C" int orbx_stereo_match(orbx_ctx* hl, orbx_ctx* hr, int frame, const orbx_keypoint* kps_l, const uint8_t* desc_l, int n_l,
                                 const orbx_keypoint* kps_r, const uint8_t* desc_r, int n_r, float bf, float b, int flags,
                                 float* u_right, float* depth, int* n_matched)
{
    if (!hl || !hr || n_l < 0 || n_r < 0 || (n_l > 0 && (!kps_l || !desc_l || !u_right || !depth)) || (n_r > 0 && (!kps_r || !desc_r)))
        return ORBX_E_INVALID;
    Ctx* c = &hl->c; Ctx* cr = &hr->c;
    if (c->device != cr->device || c->nlevels != cr->nlevels || frame < 0 || frame >= c->lastBatch || frame >= cr->lastBatch ||
        c->curW != cr->curW || c->curH != cr->curH)
        return ORBX_E_INVALID;
    if (n_matched) *n_matched = 0;
    if (n_l == 0) return ORBX_OK;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    ORBX_CUDA(c, cudaStreamSynchronize(cr->stream)); // the right pyramid is produced on the other ctx's stream
    const bool inDev = flags & ORBX_IN_DEVICE, outDev = flags & ORBX_OUT_DEVICE;
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t oKL = take(sizeof(orbx_keypoint) * (size_t)n_l), oDL = take((size_t)32 * n_l),
                 oKR = take(sizeof(orbx_keypoint) * (size_t)std::max(n_r, 1)), oDR = take((size_t)32 * std::max(n_r, 1)),
                 oU = take(sizeof(float) * (size_t)n_l), oZ = take(sizeof(float) * (size_t)n_l),
                 oList = take(sizeof(int) * 2 * (size_t)n_l), oCnt = take(sizeof(int));
    int rc = ensureScratch(c, off);
    if (rc) return rc;
    uint8_t* S = (uint8_t*)c->dScratch;
    const orbx_keypoint *dKL = kps_l, *dKR = kps_r; const uint4 *dDL = (const uint4*)desc_l, *dDR = (const uint4*)desc_r;
    stageBegin(c, ST_H2D);
    if (!inDev) {
        ORBX_CUDA(c, cudaMemcpyAsync(S + oKL, kps_l, sizeof(orbx_keypoint) * (size_t)n_l, cudaMemcpyHostToDevice, c->stream));
        ORBX_CUDA(c, cudaMemcpyAsync(S + oDL, desc_l, (size_t)32 * n_l, cudaMemcpyHostToDevice, c->stream));
        if (n_r) {
            ORBX_CUDA(c, cudaMemcpyAsync(S + oKR, kps_r, sizeof(orbx_keypoint) * (size_t)n_r, cudaMemcpyHostToDevice, c->stream));
            ORBX_CUDA(c, cudaMemcpyAsync(S + oDR, desc_r, (size_t)32 * n_r, cudaMemcpyHostToDevice, c->stream));
        }
        dKL = (const orbx_keypoint*)(S + oKL); dDL = (const uint4*)(S + oDL); dKR = (const orbx_keypoint*)(S + oKR); dDR = (const uint4*)(S + oDR);
    }
    stageEnd(c, ST_H2D, 0);
    float* dU = outDev ? u_right : (float*)(S + oU);
    float* dZ = outDev ? depth : (float*)(S + oZ);
    int* dCnt = (int*)(S + oCnt);
    PyrView L, R;
    memset(&L, 0, sizeof(L)); memset(&R, 0, sizeof(R));
    fillView(c, frame, L); fillView(cr, frame, R);
    ORBX_CUDA(c, cudaMemsetAsync(dCnt, 0, sizeof(int), c->stream));
    stageBegin(c, ST_STEREO);
    stereoMatchKernel<<<(n_l + 3) / 4, 128, 0, c->stream>>>(dKL, dDL, n_l, dKR, dDR, n_r, L, R, bf, b, dU, dZ, (int*)(S + oList), dCnt);
    stereoMedianKernel<<<1, 1024, 0, c->stream>>>((int*)(S + oList), dCnt, dU, dZ);
    stageEnd(c, ST_STEREO, 2);
    ORBX_CUDA(c, cudaGetLastError());
    int cnt = 0;
    if (!outDev) {
        ORBX_CUDA(c, cudaMemcpyAsync(u_right, dU, sizeof(float) * (size_t)n_l, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaMemcpyAsync(depth, dZ, sizeof(float) * (size_t)n_l, cudaMemcpyDeviceToHost, c->stream));
    }
    ORBX_CUDA(c, cudaMemcpyAsync(&cnt, dCnt, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    if (n_matched) *n_matched = cnt;
    return ORBX_OK;
}
