// Window (projection) matchers and the frame grid:
//   Frame::AssignFeaturesToGrid / PosInGrid        src/Frame.cc:398-422, 682-698
//   Frame::GetFeaturesInArea                       src/Frame.cc:590-671  (KeyFrame twin src/KeyFrame.cc:794-838)
//   ORBmatcher::SearchByProjection x4 inner loops  src/ORBmatcher.cc:78-191, 431-560, 1741-1910, 1922-2065
//
// The reference handles projected points one after another and lets every accepted match hide its keypoint
// from the following points.  On the GPU the parallel part (window enumeration in the reference's cell order,
// Hamming distances) runs first; one warp then replays the greedy decisions over the precomputed candidate
// lists, so results are identical whatever the order of distances.
//   P1 projCountKernel   thread per query: number of window candidates
//   P2 (scan)            exclusive prefix over queries (single CTA)
//   P3 projFillKernel    thread per query: candidate indices in visit order (ix outer, iy inner, cell order)
//   P4 projDistKernel    thread per candidate: uright gate + 256-bit distance
//   P5 projResolveKernel one warp: greedy best / second-best, thresholds, claims, rotation votes
//   P6 projRotationKernel ComputeThreeMaxima + removal (only with check_orientation)
#include "orbx_internal.h"
#include <string.h>
#include <algorithm>

namespace orbx {

constexpr int kHistoP = 30;

struct GridDev { float minX, minY, invW, invH; int cols, rows; };

// ---------------- grid ----------------
__global__ void gridCellKernel(const orbx_keypoint* __restrict__ keys, int n, GridDev g, int* __restrict__ cell)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int px = (int)roundf(__fmul_rn(__fsub_rn(keys[i].x, g.minX), g.invW));
    const int py = (int)roundf(__fmul_rn(__fsub_rn(keys[i].y, g.minY), g.invH));
    cell[i] = (px < 0 || px >= g.cols || py < 0 || py >= g.rows) ? -1 : px * g.rows + py;
}

// one thread per cell: stable gather (ascending feature index = push_back order)
__global__ void gridCountKernel(const int* __restrict__ cell, int n, int ncell, int* __restrict__ cnt)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncell) return;
    int k = 0;
    for (int i = 0; i < n; ++i) k += cell[i] == c;
    cnt[c] = k;
}

__global__ void __launch_bounds__(1024) exclusiveScanKernel(const int* __restrict__ in, int n, int* __restrict__ out)
{   // out[0..n] (n+1 entries); single CTA, chunked
    __shared__ int warpSum[32];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + threadIdx.x;
        const int v = i < n ? in[i] : 0;
        int x = v;
        const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
        #pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += t; }
        if (lane == 31) warpSum[w] = x;
        __syncthreads();
        if (w == 0) {
            int s = warpSum[lane];
            #pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, s, o); if (lane >= o) s += t; }
            warpSum[lane] = s;
        }
        __syncthreads();
        const int excl = carry + (w ? warpSum[w - 1] : 0) + x - v;
        if (i < n) out[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) out[n] = carry;
}

__global__ void gridFillKernel(const int* __restrict__ cell, int n, int ncell, const int* __restrict__ off, int* __restrict__ idx)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncell) return;
    int o = off[c];
    for (int i = 0; i < n; ++i) if (cell[i] == c) idx[o++] = i;
}

// ---------------- window enumeration ----------------
struct Window { int c0, c1, r0, r1; bool empty; };
__device__ __forceinline__ Window windowOf(const GridDev& g, float x, float y, float r)
{   // src/Frame.cc:599-622
    Window w; w.empty = true;
    w.c0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, g.minX), r), g.invW)));
    if (w.c0 >= g.cols) return w;
    w.c1 = min(g.cols - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, g.minX), r), g.invW)));
    if (w.c1 < 0) return w;
    w.r0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, g.minY), r), g.invH)));
    if (w.r0 >= g.rows) return w;
    w.r1 = min(g.rows - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, g.minY), r), g.invH)));
    if (w.r1 < 0) return w;
    w.empty = false;
    return w;
}

template <bool FILL>
__global__ void projEnumKernel(const orbx_proj_query* __restrict__ q, int nq, const orbx_keypoint* __restrict__ keys, GridDev g,
                               const int* __restrict__ cellOff, const int* __restrict__ cellIdx, int* __restrict__ count,
                               const int* __restrict__ off, int* __restrict__ cand)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    const float x = q[i].u, y = q[i].v, r = q[i].radius;
    const int minL = q[i].min_level, maxL = q[i].max_level;
    const Window w = windowOf(g, x, y, r);
    int k = 0;
    int o = FILL ? off[i] : 0;
    if (!w.empty) {
        const bool check = (minL > 0) || (maxL >= 0);
        for (int ix = w.c0; ix <= w.c1; ++ix)
            for (int iy = w.r0; iy <= w.r1; ++iy) {
                const int c = ix * g.rows + iy;
                for (int j = cellOff[c]; j < cellOff[c + 1]; ++j) {
                    const int f = cellIdx[j];
                    const orbx_keypoint kp = keys[f];
                    if (check) {
                        if (kp.octave < minL) continue;
                        if (maxL >= 0 && kp.octave > maxL) continue;
                    }
                    const float dx = __fsub_rn(kp.x, x), dy = __fsub_rn(kp.y, y);
                    if (fabsf(dx) < r && fabsf(dy) < r) { if (FILL) cand[o + k] = f; ++k; }
                }
            }
    }
    if (!FILL) count[i] = k;
}

// candidate record: distance (9 bits) | octave << 9 | skip << 15 ; feature index kept in cand[]
__global__ void projDistKernel(const orbx_proj_query* __restrict__ q, int nq, const int* __restrict__ off, const int* __restrict__ cand,
                               const orbx_keypoint* __restrict__ keys, const uint4* __restrict__ desc, const float* __restrict__ uright,
                               int checkUr, int* __restrict__ rec)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= off[nq]) return;
    int lo = 0, hi = nq; // query of this entry
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (off[mid] <= e) lo = mid; else hi = mid; }
    const int f = cand[e];
    int skip = 0;
    if (checkUr && uright && uright[f] > 0) {
        const float er = fabsf(__fsub_rn(q[lo].ur, uright[f]));
        if (er > q[lo].radius) skip = 1;
    }
    const uint4* qd = reinterpret_cast<const uint4*>(q[lo].desc);
    const uint4 a0 = qd[0], a1 = qd[1], b0 = desc[2 * f], b1 = desc[2 * f + 1];
    const int d = hamming256(a0, a1, b0, b1);
    rec[e] = d | ((keys[f].octave & 63) << 9) | (skip << 15);
}

__device__ __forceinline__ int rotBinP(float a1, float a2)
{
    float rot = __fsub_rn(a1, a2);
    if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
    int bin = (int)roundf(__fmul_rn(rot, 30 / 360.0f));
    if (bin == kHistoP) bin = 0;
    return bin;
}

// One warp, queries in order.  key = dist << 22 | position-in-list: min = best with first-wins ties.
__global__ void __launch_bounds__(32)
projResolveKernel(const orbx_proj_query* __restrict__ q, int nq, const int* __restrict__ off, const int* __restrict__ cand,
                  const int* __restrict__ rec, const orbx_keypoint* __restrict__ keys, const uint8_t* __restrict__ occupied, int n,
                  int modeFlags, int thDist, float nnratio, int checkOri, int* __restrict__ assign, int* __restrict__ hist)
{
    extern __shared__ uint32_t occ[];
    const int lane = threadIdx.x;
    for (int i = lane; i < (n + 31) / 32; i += 32) {
        uint32_t m = 0;
        if (occupied) for (int b = 0; b < 32; ++b) { const int j = i * 32 + b; if (j < n && occupied[j]) m |= 1u << b; }
        occ[i] = m;
    }
    __syncwarp();
    int nm = 0;
    for (int i = 0; i < nq; ++i) {
        const int b = off[i], e = off[i + 1];
        if (e == b) continue;
        uint32_t k1 = 0xffffffffu, k2 = 0xffffffffu;
        for (int p = b + lane; p < e; p += 32) {
            const int f = cand[p], r = rec[p];
            if ((r >> 15) & 1) continue;
            if ((occ[f >> 5] >> (f & 31)) & 1u) continue;
            const uint32_t key = ((uint32_t)(r & 511) << 22) | (uint32_t)(p - b);
            if (key < k1) { k2 = k1; k1 = key; } else if (key < k2) k2 = key;
        }
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const uint32_t b1 = __shfl_xor_sync(0xffffffffu, k1, o), b2 = __shfl_xor_sync(0xffffffffu, k2, o);
            const uint32_t lo = min(k1, b1), hi = max(k1, b1);
            k2 = min(hi, min(k2, b2)); k1 = lo;
        }
        if (k1 == 0xffffffffu) continue;
        const int best = (int)(k1 >> 22);
        if (best > thDist) continue;
        const int p1 = b + (int)(k1 & 0x3fffffu);
        if (modeFlags & ORBX_PROJ_RATIO_SAME_LEVEL) {            // src/ORBmatcher.cc:177-186
            if (k2 != 0xffffffffu) {
                const int p2 = b + (int)(k2 & 0x3fffffu);
                const int l1 = (rec[p1] >> 9) & 63, l2 = (rec[p2] >> 9) & 63;
                if (l1 == l2 && (float)best > __fmul_rn(nnratio, (float)(int)(k2 >> 22))) continue;
            }   // bestLevel2 == -1 never equals a real level
        }
        const int f = cand[p1];
        __syncwarp();
        if (lane == 0) {
            occ[f >> 5] |= 1u << (f & 31);
            assign[f] = i;
            if (checkOri) hist[rotBinP(q[i].angle, keys[f].angle)]++;
        }
        ++nm;
        __syncwarp();
    }
    if (lane == 0) hist[kHistoP] = nm;
}

__global__ void __launch_bounds__(256)
projRotationKernel(const orbx_proj_query* __restrict__ q, const orbx_keypoint* __restrict__ keys, int n, int* __restrict__ assign,
                   int* __restrict__ hist)
{
    __shared__ int keep[3];
    __shared__ int total;
    if (threadIdx.x == 0) {
        total = 0;
        int max1 = 0, max2 = 0, max3 = 0, i1 = -1, i2 = -1, i3 = -1;
        for (int i = 0; i < kHistoP; ++i) {
            const int s = hist[i];
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; i3 = i2; i2 = i1; i1 = i; }
            else if (s > max2) { max3 = max2; max2 = s; i3 = i2; i2 = i; }
            else if (s > max3) { max3 = s; i3 = i; }
        }
        if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { i2 = -1; i3 = -1; }
        else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) i3 = -1;
        keep[0] = i1; keep[1] = i2; keep[2] = i3;
    }
    __syncthreads();
    int cnt = 0;
    for (int f = threadIdx.x; f < n; f += blockDim.x) {
        const int i = assign[f];
        if (i < 0) continue;
        const int bin = rotBinP(q[i].angle, keys[f].angle);
        if (bin != keep[0] && bin != keep[1] && bin != keep[2]) { assign[f] = -1; continue; }
        ++cnt;
    }
    atomicAdd(&total, cnt);
    __syncthreads();
    if (threadIdx.x == 0) hist[kHistoP] = total;
}

struct BumpP { size_t off = 0; size_t take(size_t b) { const size_t o = off; off += (b + 255) & ~(size_t)255; return o; } };

static int buildGrid(Ctx* c, const orbx_keypoint* dKeys, int n, const GridDev& g, int* dCell, int* dCnt, int* dOff, int* dIdx)
{
    const int ncell = g.cols * g.rows;
    if (n > 0) gridCellKernel<<<(n + 255) / 256, 256, 0, c->stream>>>(dKeys, n, g, dCell);
    gridCountKernel<<<(ncell + 127) / 128, 128, 0, c->stream>>>(dCell, n, ncell, dCnt);
    exclusiveScanKernel<<<1, 1024, 0, c->stream>>>(dCnt, ncell, dOff);
    gridFillKernel<<<(ncell + 127) / 128, 128, 0, c->stream>>>(dCell, n, ncell, dOff, dIdx);
    c->kernelLaunches += 4; c->stageLaunches[ST_MATCH] += 4;
    ORBX_CUDA(c, cudaGetLastError());
    return ORBX_OK;
}

} // namespace orbx
using namespace orbx;

extern "C" int orbx_assign_grid(orbx_ctx* h, const orbx_keypoint* keys, int n, const orbx_grid* g, int flags,
                                int32_t* cell_off, int32_t* cell_idx)
{
    if (!h || !g || n < 0 || g->cols < 1 || g->rows < 1 || !cell_off || (n > 0 && (!keys || !cell_idx))) return ORBX_E_INVALID;
    Ctx* c = &h->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    const bool inDev = flags & ORBX_IN_DEVICE, outDev = flags & ORBX_OUT_DEVICE;
    const int ncell = g->cols * g->rows;
    BumpP B;
    const size_t oK = B.take(sizeof(orbx_keypoint) * (size_t)std::max(n, 1)), oCell = B.take(sizeof(int) * (size_t)std::max(n, 1)),
                 oCnt = B.take(sizeof(int) * (size_t)ncell), oOff = B.take(sizeof(int) * (size_t)(ncell + 1)),
                 oIdx = B.take(sizeof(int) * (size_t)std::max(n, 1));
    int rc = ensureScratch(c, B.off);
    if (rc) return rc;
    uint8_t* S = (uint8_t*)c->dScratch;
    const orbx_keypoint* dK = keys;
    if (!inDev && n > 0) { ORBX_CUDA(c, cudaMemcpyAsync(S + oK, keys, sizeof(orbx_keypoint) * (size_t)n, cudaMemcpyHostToDevice, c->stream)); dK = (const orbx_keypoint*)(S + oK); }
    const GridDev gd = {g->min_x, g->min_y, g->inv_w, g->inv_h, g->cols, g->rows};
    int* dOff = outDev ? cell_off : (int*)(S + oOff);
    int* dIdx = outDev ? cell_idx : (int*)(S + oIdx);
    stageBegin(c, ST_MATCH);
    rc = buildGrid(c, dK, n, gd, (int*)(S + oCell), (int*)(S + oCnt), dOff, dIdx);
    stageEnd(c, ST_MATCH, 0);
    if (rc) return rc;
    if (!outDev) {
        ORBX_CUDA(c, cudaMemcpyAsync(cell_off, dOff, sizeof(int) * (size_t)(ncell + 1), cudaMemcpyDeviceToHost, c->stream));
        if (n > 0) ORBX_CUDA(c, cudaMemcpyAsync(cell_idx, dIdx, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return ORBX_OK;
}

extern "C" int orbx_search_by_projection(orbx_ctx* h, const orbx_proj_query* q, int nq, const orbx_keypoint* keys,
                                         const uint8_t* desc, const float* uright, int n, const orbx_grid* g,
                                         const uint8_t* occupied, int mode_flags, int th_dist, float nnratio,
                                         int check_ori, int flags, int32_t* assign, int* nmatches)
{
    if (!h || !g || nq < 0 || n < 0 || g->cols < 1 || g->rows < 1 || (n > 0 && (!keys || !desc || !assign)) || (nq > 0 && !q))
        return ORBX_E_INVALID;
    Ctx* c = &h->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    if (nmatches) *nmatches = 0;
    if (n == 0) return ORBX_OK;
    const bool inDev = flags & ORBX_IN_DEVICE, outDev = flags & ORBX_OUT_DEVICE;
    const int ncell = g->cols * g->rows;
    const size_t occBytes = (size_t)((n + 31) / 32 + 1) * 4;
    if (occBytes > 200 * 1024) return ORBX_E_CAPACITY;
    // pass 1 layout (fixed part)
    BumpP B;
    const size_t oQ = B.take(sizeof(orbx_proj_query) * (size_t)std::max(nq, 1)), oK = B.take(sizeof(orbx_keypoint) * (size_t)n),
                 oD = B.take((size_t)32 * n), oU = B.take(sizeof(float) * (size_t)n), oOcc = B.take((size_t)n),
                 oCell = B.take(sizeof(int) * (size_t)n), oCnt = B.take(sizeof(int) * (size_t)ncell),
                 oOff = B.take(sizeof(int) * (size_t)(ncell + 1)), oIdx = B.take(sizeof(int) * (size_t)n),
                 oQCnt = B.take(sizeof(int) * (size_t)std::max(nq, 1)), oQOff = B.take(sizeof(int) * (size_t)(nq + 1)),
                 oHist = B.take(sizeof(int) * (kHistoP + 1)), oAssign = B.take(sizeof(int) * (size_t)n);
    const size_t fixedBytes = B.off;
    // worst case candidate volume is nq * n; allocate on demand after the count pass
    int rc = ensureScratch(c, fixedBytes);
    if (rc) return rc;
    uint8_t* S = (uint8_t*)c->dScratch;
    const orbx_proj_query* dQ = q; const orbx_keypoint* dK = keys; const uint4* dD = (const uint4*)desc;
    const float* dU = uright; const uint8_t* dOcc = occupied;
    stageBegin(c, ST_H2D);
    if (!inDev) {
        if (nq) ORBX_CUDA(c, cudaMemcpyAsync(S + oQ, q, sizeof(orbx_proj_query) * (size_t)nq, cudaMemcpyHostToDevice, c->stream));
        ORBX_CUDA(c, cudaMemcpyAsync(S + oK, keys, sizeof(orbx_keypoint) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
        ORBX_CUDA(c, cudaMemcpyAsync(S + oD, desc, (size_t)32 * n, cudaMemcpyHostToDevice, c->stream));
        dQ = (const orbx_proj_query*)(S + oQ); dK = (const orbx_keypoint*)(S + oK); dD = (const uint4*)(S + oD);
        if (uright) { ORBX_CUDA(c, cudaMemcpyAsync(S + oU, uright, sizeof(float) * (size_t)n, cudaMemcpyHostToDevice, c->stream)); dU = (const float*)(S + oU); }
        if (occupied) { ORBX_CUDA(c, cudaMemcpyAsync(S + oOcc, occupied, (size_t)n, cudaMemcpyHostToDevice, c->stream)); dOcc = S + oOcc; }
    }
    stageEnd(c, ST_H2D, 0);
    const GridDev gd = {g->min_x, g->min_y, g->inv_w, g->inv_h, g->cols, g->rows};
    int* dAssign = outDev ? assign : (int*)(S + oAssign);
    int* dHist = (int*)(S + oHist);
    stageBegin(c, ST_MATCH);
    rc = buildGrid(c, dK, n, gd, (int*)(S + oCell), (int*)(S + oCnt), (int*)(S + oOff), (int*)(S + oIdx));
    if (rc) return rc;
    ORBX_CUDA(c, cudaMemsetAsync(dHist, 0, sizeof(int) * (kHistoP + 1), c->stream));
    ORBX_CUDA(c, cudaMemsetAsync(dAssign, 0xff, sizeof(int) * (size_t)n, c->stream));
    int total = 0, launches = 0;
    if (nq > 0) {
        projEnumKernel<false><<<(nq + 127) / 128, 128, 0, c->stream>>>(dQ, nq, dK, gd, (int*)(S + oOff), (int*)(S + oIdx),
                                                                    (int*)(S + oQCnt), nullptr, nullptr);
        exclusiveScanKernel<<<1, 1024, 0, c->stream>>>((int*)(S + oQCnt), nq, (int*)(S + oQOff));
        launches += 2;
        ORBX_CUDA(c, cudaMemcpyAsync(&total, (int*)(S + oQOff) + nq, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    // candidate arrays live after the fixed part; growing the scratch would move it, so use a second buffer
    int *dCand = nullptr, *dRec = nullptr;
    if (total > 0) {
        ORBX_CUDA(c, cudaMallocAsync((void**)&dCand, sizeof(int) * (size_t)total * 2, c->stream));
        dRec = dCand + total;
        projEnumKernel<true><<<(nq + 127) / 128, 128, 0, c->stream>>>(dQ, nq, dK, gd, (int*)(S + oOff), (int*)(S + oIdx), nullptr,
                                                                   (int*)(S + oQOff), dCand);
        projDistKernel<<<(total + 255) / 256, 256, 0, c->stream>>>(dQ, nq, (int*)(S + oQOff), dCand, dK, dD, dU,
                                                                  (mode_flags & ORBX_PROJ_CHECK_URIGHT) ? 1 : 0, dRec);
        launches += 2;
    }
    if (nq > 0) {
        projResolveKernel<<<1, 32, occBytes, c->stream>>>(dQ, nq, (int*)(S + oQOff), dCand, dRec, dK, dOcc, n, mode_flags, th_dist,
                                                        nnratio, check_ori ? 1 : 0, dAssign, dHist);
        ++launches;
        if (check_ori) { projRotationKernel<<<1, 256, 0, c->stream>>>(dQ, dK, n, dAssign, dHist); ++launches; }
    }
    stageEnd(c, ST_MATCH, launches);
    ORBX_CUDA(c, cudaGetLastError());
    if (dCand) ORBX_CUDA(c, cudaFreeAsync(dCand, c->stream));
    int nm = 0;
    if (!outDev) ORBX_CUDA(c, cudaMemcpyAsync(assign, dAssign, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
    ORBX_CUDA(c, cudaMemcpyAsync(&nm, dHist + kHistoP, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    if (nmatches) *nmatches = nm;
    return ORBX_OK;
}
