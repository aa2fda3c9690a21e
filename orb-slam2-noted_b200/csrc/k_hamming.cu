// K6 -- 256-bit Hamming core of ORBmatcher (DescriptorDistance, src/ORBmatcher.cc:2123-2143) as a
// segmented many-vs-many XOR+POPC kernel with running best / second-best (the inner loops at
// :325-345, :828-843, :1013-1052; Frame::ComputeStereoMatches src/Frame.cc:937-978).
//
// Integer/POPC-pipe bound by design (8 x POPC32 per descriptor pair); no tensor cores.
// One CTA = up to 256 query rows of one segment (kHamQPT = 1 per thread, its descriptor held in 8 registers);
// the valid train descriptors stream through shared memory in compacted chunks of up to kHamChunk = 256
// (one staged row per thread) and are read as warp-broadcast 128-bit loads: two LDS.128 feed one 256-bit
// distance (18 LOP3 of the carry-save tree + 4 POPC, see hamming256 in orbx_internal.h).
#include "orbx_internal.h"
#include <algorithm>

namespace orbx {

constexpr int kHamThreads = 256, kHamQPT = 1, kHamChunk = 256;   // kHamChunk == kHamThreads: one staged train row per thread

__device__ __forceinline__ int dist256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1)
{
    return hamming256(a0, a1, b0, b1);
}

__global__ void pairDistanceKernel(const uint4* __restrict__ a, const uint4* __restrict__ b, int n, int32_t* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = dist256(a[2 * i], a[2 * i + 1], b[2 * i], b[2 * i + 1]);
}

// key of a candidate: distance << 20 | scan position inside the segment; the smallest key is the reference's best
// (lowest distance, first in list order on ties), so best / second-best updates are three integer min/max.
constexpr int kHamPosBits = 20;
static_assert(kHamChunk == kHamThreads, "one staged train row per thread");

__global__ void __launch_bounds__(kHamThreads)
hammingBest2Kernel(const uint4* __restrict__ qDesc, const uint4* __restrict__ tDesc,
                   const int32_t* __restrict__ qSeg, const int32_t* __restrict__ qIdx,
                   const int32_t* __restrict__ tSeg, const int32_t* __restrict__ tIdx,
                   const int32_t* __restrict__ blockPrefix, int nseg, const uint8_t* __restrict__ tValid,
                   int32_t* __restrict__ bestIdx, int32_t* __restrict__ bestDist, int32_t* __restrict__ secondDist)
{
    __shared__ uint4 sT[kHamChunk * 2];
    __shared__ int sPos[kHamChunk];      // scan position of the staged (valid) train descriptors
    __shared__ int sCnt[kHamThreads / 32];
    // segment of this CTA: largest s with blockPrefix[s] <= blockIdx.x
    int lo = 0, hi = nseg;
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (blockPrefix[mid] <= (int)blockIdx.x) lo = mid; else hi = mid; }
    const int seg = lo;
    const int qb = (blockIdx.x - blockPrefix[seg]) * (kHamThreads * kHamQPT);
    const int q0 = qSeg[seg], q1 = qSeg[seg + 1], t0 = tSeg[seg], t1 = tSeg[seg + 1];
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;

    uint4 qa[kHamQPT], qbv[kHamQPT];
    int row[kHamQPT];
    uint32_t k1[kHamQPT], k2[kHamQPT];                      // best and second-best keys
    const uint32_t kNone = 256u << kHamPosBits;             // distance 256, position 0: the reference's initial state
    #pragma unroll
    for (int k = 0; k < kHamQPT; ++k) {
        row[k] = q0 + qb + k * kHamThreads + threadIdx.x;
        k1[k] = kNone; k2[k] = kNone;
        if (row[k] < q1) {
            const int qi = qIdx ? qIdx[row[k]] : row[k];
            qa[k] = qDesc[2 * qi]; qbv[k] = qDesc[2 * qi + 1];
        } else { qa[k] = make_uint4(0, 0, 0, 0); qbv[k] = qa[k]; }
    }
    for (int c = t0; c < t1; c += kHamChunk) {
        // stage the valid train descriptors of this chunk, compacted, in list order
        const int i = c + (int)threadIdx.x;
        int ti = -1;
        if (i < t1) { ti = tIdx ? tIdx[i] : i; if (tValid && !tValid[ti]) ti = -1; }
        const uint32_t mask = __ballot_sync(0xffffffffu, ti >= 0);
        __syncthreads();                                    // previous chunk fully consumed
        if (lane == 0) sCnt[wrp] = __popc(mask);
        __syncthreads();
        int base = 0, m = 0;
        #pragma unroll
        for (int w = 0; w < kHamThreads / 32; ++w) { const int n = sCnt[w]; if (w < wrp) base += n; m += n; }
        if (ti >= 0) {
            const int p = base + __popc(mask & ((1u << lane) - 1u));
            sPos[p] = i - t0;
            sT[2 * p] = tDesc[2 * ti]; sT[2 * p + 1] = tDesc[2 * ti + 1];
        }
        __syncthreads();
        #pragma unroll 4
        for (int j = 0; j < m; ++j) {
            const uint4 ta = sT[2 * j], tb = sT[2 * j + 1];
            const uint32_t pos = (uint32_t)sPos[j];
            #pragma unroll
            for (int k = 0; k < kHamQPT; ++k) {
                const uint32_t key = ((uint32_t)hamming256(qa[k], qbv[k], ta, tb) << kHamPosBits) + pos;
                k2[k] = min(k2[k], max(k1[k], key));
                k1[k] = min(k1[k], key);
            }
        }
    }
    #pragma unroll
    for (int k = 0; k < kHamQPT; ++k)
        if (row[k] < q1) {
            const int d1 = (int)(k1[k] >> kHamPosBits), d2 = (int)(k2[k] >> kHamPosBits);
            int bi = -1;
            if (k1[k] != kNone) { const int p = t0 + (int)(k1[k] & ((1u << kHamPosBits) - 1u)); bi = tIdx ? tIdx[p] : p; }
            bestIdx[row[k]] = bi; bestDist[row[k]] = d1; secondDist[row[k]] = d2;
        }
}

// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:366-448) for many map points in one launch: one warp per map
// point.  Row i of the point's N x N distance matrix is written to shared memory as it is computed; its median
// (element (N-1)/2 of the sorted row, self-distance 0 included) is found by bisection on the value (distances are 0..256)
// with warp-wide counts, so nothing is sorted.  The first row with the smallest median wins (strict <, :437).
constexpr int kDistinctWarps = 4, kDistinctMaxObs = 1024;
__global__ void __launch_bounds__(32 * kDistinctWarps)
distinctiveKernel(const uint4* __restrict__ desc, const int* __restrict__ seg, int npoints, int* __restrict__ best)
{
    __shared__ uint16_t sd[kDistinctWarps][kDistinctMaxObs];
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    const int p = blockIdx.x * kDistinctWarps + wrp;
    if (p >= npoints) return;
    const int s0 = seg[p], N = seg[p + 1] - s0;
    if (N <= 0 || N > kDistinctMaxObs) { if (lane == 0) best[p] = N <= 0 ? -1 : -2; return; }
    const uint4* D = desc + 2 * (size_t)s0;
    uint16_t* row = sd[wrp];
    const int k1 = ((N - 1) >> 1) + 1;                               // rank of the median, 1-based
    int bestMedian = 0x7fffffff, bestIdx = 0;
    for (int i = 0; i < N; ++i) {
        const uint4 a0 = D[2 * i], a1 = D[2 * i + 1];
        for (int j = lane; j < N; j += 32) row[j] = (uint16_t)(i == j ? 0 : hamming256(a0, a1, D[2 * j], D[2 * j + 1]));
        __syncwarp();
        int lo = 0, hi = 256;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            int cnt = 0;
            for (int j = lane; j < N; j += 32) cnt += row[j] <= mid;
            cnt = __reduce_add_sync(0xffffffffu, cnt);
            if (cnt >= k1) hi = mid; else lo = mid + 1;
        }
        if (lo < bestMedian) { bestMedian = lo; bestIdx = i; }
        __syncwarp();
    }
    if (lane == 0) best[p] = bestIdx;
}

} // namespace orbx
using namespace orbx;

extern "C" int orbx_descriptor_distance(orbx_ctx* h, const uint8_t* a, const uint8_t* b, int n, int flags, int32_t* dist)
{
    if (!h || n < 0 || (n > 0 && (!a || !b || !dist))) return ORBX_E_INVALID;
    if (n == 0) return ORBX_OK;
    Ctx* c = &h->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    const uint4 *da = (const uint4*)a, *db = (const uint4*)b; int32_t* dd = dist;
    const bool inDev = flags & ORBX_IN_DEVICE, outDev = flags & ORBX_OUT_DEVICE;
    if (!inDev || !outDev) {
        int rc = ensureScratch(c, (size_t)n * 68);
        if (rc) return rc;
        uint8_t* s = (uint8_t*)c->dScratch;
        if (!inDev) {
            ORBX_CUDA(c, cudaMemcpyAsync(s, a, (size_t)n * 32, cudaMemcpyHostToDevice, c->stream));
            ORBX_CUDA(c, cudaMemcpyAsync(s + (size_t)n * 32, b, (size_t)n * 32, cudaMemcpyHostToDevice, c->stream));
            da = (const uint4*)s; db = (const uint4*)(s + (size_t)n * 32);
        }
        if (!outDev) dd = (int32_t*)(s + (size_t)n * 64);
    }
    stageBegin(c, ST_HAMMING);
    pairDistanceKernel<<<(n + 255) / 256, 256, 0, c->stream>>>(da, db, n, dd);
    stageEnd(c, ST_HAMMING, 1);
    ORBX_CUDA(c, cudaGetLastError());
    if (!outDev) {
        ORBX_CUDA(c, cudaMemcpyAsync(dist, dd, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return ORBX_OK;
}

// seg is a HOST array of npoints + 1 offsets (rows of desc); desc / best follow `flags`.
extern "C" int orbx_distinctive_descriptors(orbx_ctx* h, const uint8_t* desc, const int32_t* seg, int npoints, int flags, int32_t* best)
{
    if (!h || npoints < 0 || (npoints > 0 && (!seg || !best))) return ORBX_E_INVALID;
    if (npoints == 0) return ORBX_OK;
    const int total = seg[npoints];
    if (total < 0 || (total > 0 && !desc)) return ORBX_E_INVALID;
    for (int p = 0; p < npoints; ++p) {
        if (seg[p + 1] < seg[p]) return ORBX_E_INVALID;
        if (seg[p + 1] - seg[p] > kDistinctMaxObs) return ORBX_E_CAPACITY;
    }
    Ctx* c = &h->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    const bool inDev = flags & ORBX_IN_DEVICE, outDev = flags & ORBX_OUT_DEVICE;
    const size_t dBytes = ((size_t)std::max(total, 1) * 32 + 255) & ~(size_t)255, sBytes = (sizeof(int) * (size_t)(npoints + 1) + 255) & ~(size_t)255;
    int rc = ensureScratch(c, dBytes + sBytes + sizeof(int) * (size_t)npoints);
    if (rc) return rc;
    uint8_t* S = (uint8_t*)c->dScratch;
    const uint4* dD = (const uint4*)desc;
    if (!inDev && total > 0) { ORBX_CUDA(c, cudaMemcpyAsync(S, desc, (size_t)total * 32, cudaMemcpyHostToDevice, c->stream)); dD = (const uint4*)S; }
    ORBX_CUDA(c, cudaMemcpyAsync(S + dBytes, seg, sizeof(int) * (size_t)(npoints + 1), cudaMemcpyHostToDevice, c->stream));
    int* dBest = outDev ? best : (int*)(S + dBytes + sBytes);
    stageBegin(c, ST_HAMMING);
    distinctiveKernel<<<(npoints + kDistinctWarps - 1) / kDistinctWarps, 32 * kDistinctWarps, 0, c->stream>>>(dD, (const int*)(S + dBytes), npoints, dBest);
    stageEnd(c, ST_HAMMING, 1);
    ORBX_CUDA(c, cudaGetLastError());
    if (!outDev) ORBX_CUDA(c, cudaMemcpyAsync(best, dBest, sizeof(int) * (size_t)npoints, cudaMemcpyDeviceToHost, c->stream));
    ORBX_CUDA(c, cudaStreamSynchronize(c->stream));   // seg is read from pageable host memory
    return ORBX_OK;
}

// q_seg / t_seg are always HOST arrays (nseg+1 CSR offsets); everything else follows `flags`.
extern "C" int orbx_hamming_best2(orbx_ctx* h, const uint8_t* q_desc, int nq, const uint8_t* t_desc, int nt,
                                  const int32_t* q_seg, const int32_t* q_idx, const int32_t* t_seg, const int32_t* t_idx,
                                  int nseg, const uint8_t* t_valid, int flags,
                                  int32_t* best_idx, int32_t* best_dist, int32_t* second_dist)
{
    if (!h || nq < 0 || nt < 0 || nseg < 1 || !q_seg || !t_seg || !best_idx || !best_dist || !second_dist) return ORBX_E_INVALID;
    if ((nq > 0 && !q_desc) || (nt > 0 && !t_desc)) return ORBX_E_INVALID;
    Ctx* c = &h->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    const int nrowQ = q_seg[nseg], nrowT = t_seg[nseg];
    if (nrowQ == 0) return ORBX_OK;
    std::vector<int32_t> prefix((size_t)nseg + 1);
    const int per = kHamThreads * kHamQPT;
    prefix[0] = 0;
    for (int s = 0; s < nseg; ++s) {
        if (q_seg[s + 1] < q_seg[s] || t_seg[s + 1] < t_seg[s]) return ORBX_E_INVALID;
        if (t_seg[s + 1] - t_seg[s] > (1 << kHamPosBits)) return ORBX_E_CAPACITY;   // scan position field of the candidate key
        prefix[s + 1] = prefix[s] + (q_seg[s + 1] - q_seg[s] + per - 1) / per;
    }
    const int nblocks = prefix[nseg];
    const bool inDev = flags & ORBX_IN_DEVICE, outDev = flags & ORBX_OUT_DEVICE;
    // scratch layout: [prefix | qSeg | tSeg | (host inputs) | (host outputs)]
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t oPrefix = take(sizeof(int32_t) * (nseg + 1)), oQs = take(sizeof(int32_t) * (nseg + 1)), oTs = take(sizeof(int32_t) * (nseg + 1));
    size_t oQ = 0, oT = 0, oQi = 0, oTi = 0, oTv = 0, oOut = 0;
    if (!inDev) {
        oQ = take((size_t)nq * 32); oT = take((size_t)nt * 32);
        if (q_idx) oQi = take(sizeof(int32_t) * (size_t)nrowQ);
        if (t_idx) oTi = take(sizeof(int32_t) * (size_t)nrowT);
        if (t_valid) oTv = take((size_t)nt);
    }
    if (!outDev) oOut = take(sizeof(int32_t) * 3 * (size_t)nrowQ);
    int rc = ensureScratch(c, off);
    if (rc) return rc;
    uint8_t* S = (uint8_t*)c->dScratch;
    stageBegin(c, ST_H2D);
    ORBX_CUDA(c, cudaMemcpyAsync(S + oPrefix, prefix.data(), sizeof(int32_t) * (nseg + 1), cudaMemcpyHostToDevice, c->stream));
    ORBX_CUDA(c, cudaMemcpyAsync(S + oQs, q_seg, sizeof(int32_t) * (nseg + 1), cudaMemcpyHostToDevice, c->stream));
    ORBX_CUDA(c, cudaMemcpyAsync(S + oTs, t_seg, sizeof(int32_t) * (nseg + 1), cudaMemcpyHostToDevice, c->stream));
    const uint4 *dq = (const uint4*)q_desc, *dt = (const uint4*)t_desc;
    const int32_t *dqi = q_idx, *dti = t_idx; const uint8_t* dtv = t_valid;
    if (!inDev) {
        ORBX_CUDA(c, cudaMemcpyAsync(S + oQ, q_desc, (size_t)nq * 32, cudaMemcpyHostToDevice, c->stream)); dq = (const uint4*)(S + oQ);
        ORBX_CUDA(c, cudaMemcpyAsync(S + oT, t_desc, (size_t)nt * 32, cudaMemcpyHostToDevice, c->stream)); dt = (const uint4*)(S + oT);
        if (q_idx) { ORBX_CUDA(c, cudaMemcpyAsync(S + oQi, q_idx, sizeof(int32_t) * (size_t)nrowQ, cudaMemcpyHostToDevice, c->stream)); dqi = (const int32_t*)(S + oQi); }
        if (t_idx) { ORBX_CUDA(c, cudaMemcpyAsync(S + oTi, t_idx, sizeof(int32_t) * (size_t)nrowT, cudaMemcpyHostToDevice, c->stream)); dti = (const int32_t*)(S + oTi); }
        if (t_valid) { ORBX_CUDA(c, cudaMemcpyAsync(S + oTv, t_valid, (size_t)nt, cudaMemcpyHostToDevice, c->stream)); dtv = S + oTv; }
    }
    stageEnd(c, ST_H2D, 0);
    int32_t *dbi = best_idx, *dbd = best_dist, *dsd = second_dist;
    if (!outDev) { dbi = (int32_t*)(S + oOut); dbd = dbi + nrowQ; dsd = dbd + nrowQ; }
    stageBegin(c, ST_HAMMING);
    hammingBest2Kernel<<<nblocks, kHamThreads, 0, c->stream>>>(dq, dt, (const int32_t*)(S + oQs), dqi, (const int32_t*)(S + oTs), dti,
                                                               (const int32_t*)(S + oPrefix), nseg, dtv, dbi, dbd, dsd);
    stageEnd(c, ST_HAMMING, 1);
    ORBX_CUDA(c, cudaGetLastError());
    if (!outDev) {
        stageBegin(c, ST_D2H);
        ORBX_CUDA(c, cudaMemcpyAsync(best_idx, dbi, sizeof(int32_t) * (size_t)nrowQ, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaMemcpyAsync(best_dist, dbd, sizeof(int32_t) * (size_t)nrowQ, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaMemcpyAsync(second_dist, dsd, sizeof(int32_t) * (size_t)nrowQ, cudaMemcpyDeviceToHost, c->stream));
        stageEnd(c, ST_D2H, 0);
        ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return ORBX_OK;
}

// ---- POPC-pipe peak micro-benchmark (roofline denominator for K6; MEASURED_PEAKS.json has no integer peak) ----
namespace orbx {
__global__ void __launch_bounds__(256) popcPeakKernel(uint32_t* out, int iters)
{
    uint32_t a[8], s[8];
    #pragma unroll
    for (int j = 0; j < 8; ++j) { a[j] = (threadIdx.x + 1u) * 2654435761u + j * 0x9E3779B9u + blockIdx.x; s[j] = 0; }
    for (int i = 0; i < iters; ++i) {
        #pragma unroll
        for (int j = 0; j < 8; ++j) { s[j] += __popc(a[j]); a[j] ^= s[j]; } // 8 independent POPC chains per thread
    }
    uint32_t r = 0;
    #pragma unroll
    for (int j = 0; j < 8; ++j) r += s[j];
    if (r == 0xffffffffu) out[0] = r; // never true in practice; keeps the loop alive
}
}

extern "C" int orbx_test_popc_peak(int device, int iters, double* popc_per_s, double* ms_out)
{
    if (cudaSetDevice(device) != cudaSuccess) return ORBX_E_CUDA;
    uint32_t* d = nullptr;
    if (cudaMalloc(&d, 64) != cudaSuccess) return ORBX_E_CUDA;
    cudaDeviceProp p; cudaGetDeviceProperties(&p, device);
    const int blocks = p.multiProcessorCount * 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    orbx::popcPeakKernel<<<blocks, 256>>>(d, iters / 8); // warm-up
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        orbx::popcPeakKernel<<<blocks, 256>>>(d, iters);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(d); return ORBX_E_CUDA; }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
    if (popc_per_s) *popc_per_s = (double)blocks * 256.0 * 8.0 * (double)iters / (best * 1e-3);
    if (ms_out) *ms_out = best;
    return ORBX_OK;
}

// ---- orbx_hamming_best2_static: the same kernel with NO allocation on any path ---------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  orbx_hamming_best2 keeps its CSR
// arrays and host staging in the context's scratch buffer, which ensureScratch() grows with cudaFree + cudaMalloc; here
// the scratch is a file-scope device array of ORBX_HAMMING_STATIC_BYTES that every GPU loading this library holds until
// the process exits.
//
// gHamStaticScratch is, in HOST code, only a stand-in for the symbol: it is NOT a device address and must never be handed
// to a kernel or a memcpy.  The device address is fetched per device with cudaGetSymbolAddress (staticScratch below).
// All calls on a device share the array and it keeps the previous call's contents: every byte the kernel reads is
// rewritten by this call's copies first (nothing has to start at zero), and two calls that overlap on the GPU -- two
// contexts, two streams, two host threads -- race on it.  Serialise them.
#include <mutex>
namespace orbx {
static __device__ __align__(256) unsigned char gHamStaticScratch[ORBX_HAMMING_STATIC_BYTES];

constexpr int kStaticMaxDevices = 64;
static unsigned char* gHamStaticPtr[kStaticMaxDevices] = {};   // device address of the array on each GPU (0 = not asked yet)
static std::mutex gHamStaticMutex;

// Set-up, once per device, with that device current (the caller has just done cudaSetDevice(c->device)).
static int staticScratch(Ctx* c, unsigned char** out)
{
    if (c->device < 0 || c->device >= kStaticMaxDevices) return ORBX_E_INVALID;
    std::lock_guard<std::mutex> lock(gHamStaticMutex);
    if (!gHamStaticPtr[c->device]) {
        unsigned char* p = nullptr;
        ORBX_CUDA(c, cudaGetSymbolAddress((void**)&p, gHamStaticScratch));
        if (!p) return ORBX_E_CUDA;
        gHamStaticPtr[c->device] = p;
    }
    *out = gHamStaticPtr[c->device];
    return ORBX_OK;
}
} // namespace orbx

extern "C" int orbx_hamming_best2_static(orbx_ctx* h, const uint8_t* q_desc, int nq, const uint8_t* t_desc, int nt,
                                         const int32_t* q_seg, const int32_t* q_idx, const int32_t* t_seg, const int32_t* t_idx,
                                         int nseg, const uint8_t* t_valid, int flags,
                                         int32_t* best_idx, int32_t* best_dist, int32_t* second_dist)
{
    if (!h || nq < 0 || nt < 0 || nseg < 1 || !q_seg || !t_seg || !best_idx || !best_dist || !second_dist) return ORBX_E_INVALID;
    if ((nq > 0 && !q_desc) || (nt > 0 && !t_desc)) return ORBX_E_INVALID;
    Ctx* c = &h->c;
    if (nseg > ORBX_HAMMING_STATIC_MAX_SEG) {
        c->lastCudaError = "orbx_hamming_best2_static: nseg exceeds ORBX_HAMMING_STATIC_MAX_SEG";
        return ORBX_E_CAPACITY;
    }
    ORBX_CUDA(c, cudaSetDevice(c->device));
    const int nrowQ = q_seg[nseg], nrowT = t_seg[nseg];
    if (nrowQ == 0) return ORBX_OK;
    int32_t prefix[ORBX_HAMMING_STATIC_MAX_SEG + 1];          // on the stack: no host allocation either
    const int per = kHamThreads * kHamQPT;
    prefix[0] = 0;
    for (int s = 0; s < nseg; ++s) {
        if (q_seg[s + 1] < q_seg[s] || t_seg[s + 1] < t_seg[s]) return ORBX_E_INVALID;
        if (t_seg[s + 1] - t_seg[s] > (1 << kHamPosBits)) return ORBX_E_CAPACITY;
        prefix[s + 1] = prefix[s] + (q_seg[s + 1] - q_seg[s] + per - 1) / per;
    }
    const int nblocks = prefix[nseg];
    const bool inDev = flags & ORBX_IN_DEVICE, outDev = flags & ORBX_OUT_DEVICE;
    // layout inside the static array: [prefix | qSeg | tSeg | (host inputs) | (host outputs)], as in orbx_hamming_best2
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t oPrefix = take(sizeof(int32_t) * (nseg + 1)), oQs = take(sizeof(int32_t) * (nseg + 1)), oTs = take(sizeof(int32_t) * (nseg + 1));
    size_t oQ = 0, oT = 0, oQi = 0, oTi = 0, oTv = 0, oOut = 0;
    if (!inDev) {
        oQ = take((size_t)nq * 32); oT = take((size_t)nt * 32);
        if (q_idx) oQi = take(sizeof(int32_t) * (size_t)nrowQ);
        if (t_idx) oTi = take(sizeof(int32_t) * (size_t)nrowT);
        if (t_valid) oTv = take((size_t)nt);
    }
    if (!outDev) oOut = take(sizeof(int32_t) * 3 * (size_t)nrowQ);
    if (off > (size_t)ORBX_HAMMING_STATIC_BYTES) {
        c->lastCudaError = "orbx_hamming_best2_static: this call needs more scratch than ORBX_HAMMING_STATIC_BYTES "
                           "(pass device buffers, or use orbx_hamming_best2)";
        return ORBX_E_CAPACITY;
    }
    unsigned char* S = nullptr;                               // the array's DEVICE address on this GPU
    int rc = staticScratch(c, &S);
    if (rc) return rc;
    ORBX_CUDA(c, cudaMemcpyAsync(S + oPrefix, prefix, sizeof(int32_t) * (nseg + 1), cudaMemcpyHostToDevice, c->stream));
    ORBX_CUDA(c, cudaMemcpyAsync(S + oQs, q_seg, sizeof(int32_t) * (nseg + 1), cudaMemcpyHostToDevice, c->stream));
    ORBX_CUDA(c, cudaMemcpyAsync(S + oTs, t_seg, sizeof(int32_t) * (nseg + 1), cudaMemcpyHostToDevice, c->stream));
    const uint4 *dq = (const uint4*)q_desc, *dt = (const uint4*)t_desc;
    const int32_t *dqi = q_idx, *dti = t_idx; const uint8_t* dtv = t_valid;
    if (!inDev) {
        ORBX_CUDA(c, cudaMemcpyAsync(S + oQ, q_desc, (size_t)nq * 32, cudaMemcpyHostToDevice, c->stream)); dq = (const uint4*)(S + oQ);
        ORBX_CUDA(c, cudaMemcpyAsync(S + oT, t_desc, (size_t)nt * 32, cudaMemcpyHostToDevice, c->stream)); dt = (const uint4*)(S + oT);
        if (q_idx) { ORBX_CUDA(c, cudaMemcpyAsync(S + oQi, q_idx, sizeof(int32_t) * (size_t)nrowQ, cudaMemcpyHostToDevice, c->stream)); dqi = (const int32_t*)(S + oQi); }
        if (t_idx) { ORBX_CUDA(c, cudaMemcpyAsync(S + oTi, t_idx, sizeof(int32_t) * (size_t)nrowT, cudaMemcpyHostToDevice, c->stream)); dti = (const int32_t*)(S + oTi); }
        if (t_valid) { ORBX_CUDA(c, cudaMemcpyAsync(S + oTv, t_valid, (size_t)nt, cudaMemcpyHostToDevice, c->stream)); dtv = S + oTv; }
    }
    int32_t *dbi = best_idx, *dbd = best_dist, *dsd = second_dist;
    if (!outDev) { dbi = (int32_t*)(S + oOut); dbd = dbi + nrowQ; dsd = dbd + nrowQ; }
    // launch counters only: stageBegin() would take events from a pool that can grow while profiling is on
    hammingBest2Kernel<<<nblocks, kHamThreads, 0, c->stream>>>(dq, dt, (const int32_t*)(S + oQs), dqi, (const int32_t*)(S + oTs), dti,
                                                               (const int32_t*)(S + oPrefix), nseg, dtv, dbi, dbd, dsd);
    c->kernelLaunches += 1; c->stageLaunches[ST_HAMMING] += 1;
    ORBX_CUDA(c, cudaGetLastError());
    if (!outDev) {
        ORBX_CUDA(c, cudaMemcpyAsync(best_idx, dbi, sizeof(int32_t) * (size_t)nrowQ, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaMemcpyAsync(best_dist, dbd, sizeof(int32_t) * (size_t)nrowQ, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaMemcpyAsync(second_dist, dsd, sizeof(int32_t) * (size_t)nrowQ, cudaMemcpyDeviceToHost, c->stream));
    }
    // the CSR arrays (and `prefix` on this stack frame) are read by the copies above: wait before returning
    ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    return ORBX_OK;
}
