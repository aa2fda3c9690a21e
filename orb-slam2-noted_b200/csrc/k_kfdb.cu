// KeyFrameDatabase on the device: add / erase / clear and the two candidate searches of
// src/KeyFrameDatabase.cc -- DetectLoopCandidates :113-270 and DetectRelocalizationCandidates :275-399 -- with the
// reference's results in the reference's ORDER (the output is an ordered vector<KeyFrame*>).
//
// The reference walks an inverted file (word -> std::list<KeyFrame*>): pointer chasing whose only products are, per keyframe,
// the NUMBER of words shared with the query and the list order of first encounter.  On the GPU both come out of a forward
// scan: the query's words are a bitmap over the vocabulary (1 M words = 128 KB, L1/L2-resident), every keyframe's BowVector
// lies back to back in one arena in insertion order, and one warp per keyframe tests its words against the bitmap with
// coalesced loads.  The order of first encounter is (smallest shared word, position in that word's list), and the position
// in a list is the insertion order of the keyframes that carry the word -- so the sort key is (first shared word, slot),
// with slots handed out in insertion order and never reused.  No inverted file is kept at all; add is an append, erase a flag.
//   K1 kfdbBeginKernel   counters, the query's bitmap and word -> query-position table, exclusion marks of the connected keyframes
//   K2 kfdbSharedKernel  one warp per keyframe: shared-word count, first shared word, the mnLoopQuery / mnLoopWords state
//                        machine of :120-146 (connected keyframes, repeated query ids), max common words
//   K3 kfdbScoreKernel   one CTA per keyframe above 0.8 * max: L1Scoring::score in double; all threads compute the terms of the
//                        common words, one warp sums them in ascending word order through shuffles -> float score, candidate list
//   K4 kfdbFinalKernel   one CTA: group accumulation over the <= 10 best covisible keyframes :193-224 / :337-364, the
//                        0.75 * best gate, first-occurrence de-duplication and output position by rank among the candidates'
//                        keys (O(C^2) over a handful to a few thousand candidates), then the query's bits leave the bitmap
// Four launches per query and no memset: 0.07 ms per query in a batch, 0.11 ms for a single call (4000 keyframes x 860 words).
// The per-keyframe state the reference keeps in KeyFrame (mnLoopQuery, mnLoopWords, mLoopScore, mnRelocQuery, mnRelocWords,
// mRelocScore; include/KeyFrame.h:322-333) lives in slot arrays here and persists between queries exactly like there: a
// relocalisation query adds a neighbour's mRelocScore even when that score stems from an earlier query (:351).  The two
// scores have no initialiser in the reference (src/KeyFrame.cc:43-60); they start at 0 here.
#include "orbx_internal.h"
#include <climits>
#include <unordered_map>
#include <vector>

struct orbx_kfdb {
    orbx_ctx* ctx = nullptr;
    int nwords = 0;
    // BowVector arena (insertion order)
    int32_t* dIds = nullptr; double* dVals = nullptr; size_t entries = 0, entriesCap = 0;
    // slots
    int nslots = 0, slotsCap = 0;
    int *dSeg = nullptr, *dLen = nullptr, *dKfId = nullptr, *dAlive = nullptr, *dNeigh = nullptr;     // dNeigh: 10 ids per slot, -1 padded
    long long *dLoopQuery = nullptr, *dRelocQuery = nullptr;
    int *dLoopWords = nullptr, *dRelocWords = nullptr;
    float *dLoopScore = nullptr, *dRelocScore = nullptr;
    int *dExclStamp = nullptr, *dInList = nullptr, *dFirstWord = nullptr, *dListed = nullptr;   // dListed: slots that share words with the query
    // candidates of the current query
    unsigned long long* dKey = nullptr; int *dCandSlot = nullptr, *dBest = nullptr, *dEmit = nullptr; float* dAcc = nullptr;
    // id -> slot
    int* dSlotOfId = nullptr; int idCap = 0;
    // query
    uint32_t* dBitmap = nullptr;
    int32_t* dQPos = nullptr;               // word -> index in the query's BowVector; valid where the bitmap bit is set (never cleared)
    int32_t* dQIds = nullptr; double* dQVals = nullptr; int qCap = 0;
    int* dConn = nullptr; int connCap = 0;
    int* dCounters = nullptr;               // [0] max common words, [1] candidates, [2] best accumulated score (float bits), [3] outputs
    int32_t* dOut = nullptr; int* dNOut = nullptr; int outCap = 0, outQueries = 0;
    int stamp = 0;
    std::unordered_map<int, int> slotOf;    // host mirror: keyframe id -> live slot
    std::vector<int> idOfSlot;
};

namespace orbx {

__global__ void kfdbInitSlotKernel(int slot, int seg, int len, int id, int* dSeg, int* dLen, int* dKfId, int* dAlive, int* dNeigh,
                                   long long* lq, long long* rq, int* lw, int* rw, float* ls, float* rs, int* excl, int* slotOfId)
{
    if (threadIdx.x == 0) {
        dSeg[slot] = seg; dLen[slot] = len; dKfId[slot] = id; dAlive[slot] = 1;
        lq[slot] = 0; rq[slot] = 0; lw[slot] = 0; rw[slot] = 0; ls[slot] = 0.f; rs[slot] = 0.f; excl[slot] = 0;   // src/KeyFrame.cc:47
        slotOfId[id] = slot;
    }
    if (threadIdx.x < 10) dNeigh[slot * 10 + threadIdx.x] = -1;
}

struct Neigh10 { int id[10]; };
__global__ void kfdbSetNeighKernel(int slot, Neigh10 nb, int* dNeigh) { if (threadIdx.x < 10) dNeigh[slot * 10 + threadIdx.x] = nb.id[threadIdx.x]; }
// neigh: n x 10 ids (-1 padded) of the keyframes ids[0..n)
__global__ void kfdbSetNeighBulkKernel(const int* __restrict__ ids, const int* __restrict__ neigh, int n, const int* __restrict__ slotOfId, int idCap,
                                       int* __restrict__ dNeigh)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * 10) return;
    const int id = ids[i / 10];
    if (id < 0 || id >= idCap) return;
    const int s = slotOfId[id];
    if (s >= 0) dNeigh[s * 10 + i % 10] = neigh[i];
}
__global__ void kfdbEraseKernel(int slot, int id, int* dAlive, int* slotOfId) { dAlive[slot] = 0; slotOfId[id] = -1; }
__global__ void kfdbFillKernel(int* p, int n, int v) { const int i = blockIdx.x * blockDim.x + threadIdx.x; if (i < n) p[i] = v; }

// K1: counters, exclusion stamps of the connected keyframes, and the query's bitmap (cleared by a memset before)
__global__ void kfdbBeginKernel(const int32_t* __restrict__ qIds, int nq, int nwords, uint32_t* __restrict__ bitmap, int32_t* __restrict__ qpos,
                                const int* __restrict__ conn,
                                int nconn, const int* __restrict__ slotOfId, int idCap, int* __restrict__ exclStamp, int stamp,
                                int* __restrict__ counters, float bestAcc0)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { counters[0] = 0; counters[1] = 0; counters[2] = __float_as_int(bestAcc0); counters[3] = 0; }
    if (i < nq) { const int w = qIds[i]; if (w >= 0 && w < nwords) { atomicOr(&bitmap[w >> 5], 1u << (w & 31)); qpos[w] = i; } }
    if (i < nconn) { const int id = conn[i]; if (id >= 0 && id < idCap) { const int s = slotOfId[id]; if (s >= 0) exclStamp[s] = stamp; } }
}

// K2: one warp per slot
__global__ void __launch_bounds__(256)
kfdbSharedKernel(int nslots, const int* __restrict__ dSeg, const int* __restrict__ dLen, const int* __restrict__ dAlive,
                 const int32_t* __restrict__ ids, const uint32_t* __restrict__ bitmap, int loop, long long qid,
                 const int* __restrict__ exclStamp, int stamp, long long* __restrict__ query, int* __restrict__ words,
                 int* __restrict__ inList, int* __restrict__ firstWord, int* __restrict__ counters, int* __restrict__ listed)
{
    const int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (s >= nslots) return;
    if (!dAlive[s]) { if (lane == 0) inList[s] = 0; return; }
    const int seg = dSeg[s], len = dLen[s];
    int cnt = 0, first = INT_MAX;
    #pragma unroll 8
    for (int j = lane; j < len; j += 32) {                       // independent iterations: eight pairs of loads in flight
        const int w = ids[seg + j];
        if ((bitmap[w >> 5] >> (w & 31)) & 1u) { ++cnt; first = min(first, w); }
    }
    #pragma unroll
    for (int o = 16; o; o >>= 1) { cnt += __shfl_xor_sync(0xffffffffu, cnt, o); first = min(first, __shfl_xor_sync(0xffffffffu, first, o)); }
    if (lane) return;
    int in = 0;
    if (cnt) {                                                   // the keyframe is met `cnt` times while the query's words are walked
        if (query[s] != qid) {
            if (loop && exclStamp[s] == stamp) words[s] = 1;     // connected: mnLoopWords reset at every visit, then ++ (:131-141)
            else {
                query[s] = qid; words[s] = cnt; in = 1; firstWord[s] = first; atomicMax(&counters[0], cnt);
                listed[atomicAdd(&counters[3], 1)] = s;          // lKFsSharingWords, unordered (the order comes from the keys)
            }
        } else words[s] += cnt;                                  // a query id seen before: counted on, not listed again
    }
    inList[s] = in;
}

__device__ __forceinline__ void atomicMaxFloat(int* addr, float v)
{
    int old = *addr;
    while (v > __int_as_float(old)) { const int assumed = old; old = atomicCAS(addr, assumed, __float_as_int(v)); if (old == assumed) break; }
}

// K3: the CTAs walk the list of keyframes that share words (K2) and take those with more than minCommon of them (a handful).
// All 256 threads compute the terms of the common words (independent loads, so the keyframe's ~1000 entries cost two or three
// round trips), then warp 0 adds them in ascending word order -- the order of L1Scoring::score's merge -- from shared memory.
constexpr int kScoreTile = 2048;
__global__ void __launch_bounds__(256)
kfdbScoreKernel(int nslots, const int* __restrict__ dSeg, const int* __restrict__ dLen, const int32_t* __restrict__ ids,
                const double* __restrict__ vals, const uint32_t* __restrict__ bitmap, const int32_t* __restrict__ qpos,
                const double* __restrict__ qVals, int nq, int loop, float minScore, const int* __restrict__ inList,
                const int* __restrict__ words, const int* __restrict__ firstWord, float* __restrict__ score,
                int* __restrict__ counters, unsigned long long* __restrict__ key, int* __restrict__ candSlot,
                const int* __restrict__ listed)
{
    __shared__ double sTerm[kScoreTile];
    __shared__ uint32_t sHit[kScoreTile / 32];
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    const int minCommon = (int)((float)counters[0] * 0.8f);      // int * float -> float -> int, :161 / :312
    const int nListed = counters[3];
  for (int li = blockIdx.x; li < nListed; li += gridDim.x) {
    const int s = listed[li];
    if (!(words[s] > minCommon)) continue;                        // uniform over the CTA
    const int seg = dSeg[s], len = dLen[s];
    double sum = 0.0;                                             // lives in warp 0
    for (int t0 = 0; t0 < len; t0 += kScoreTile) {
        const int tn = min(kScoreTile, len - t0);
        for (int j = threadIdx.x; j < ((tn + 31) & ~31); j += 256) {
            double term = 0.0;
            bool hit = false;
            if (j < tn) {
                const int w = ids[seg + t0 + j];
                if ((bitmap[w >> 5] >> (w & 31)) & 1u) {
                    const double vi = qVals[qpos[w]], wi = vals[seg + t0 + j];   // the word is in the query: qpos holds its place there
                    term = fabs(vi - wi) - fabs(vi) - fabs(wi);  // ScoringObject.cpp:41
                    hit = true;
                }
            }
            sTerm[j] = term;
            const uint32_t m = __ballot_sync(0xffffffffu, hit);
            if (lane == 0) sHit[j >> 5] = m;
        }
        __syncthreads();
        if (wrp == 0) {
            for (int base = 0; base < tn; base += 32) {
                const double term = sTerm[base + lane];
                uint32_t m = sHit[base >> 5];
                while (m) {                                       // ascending word order = ascending lane
                    const int l = __ffs(m) - 1;
                    m &= m - 1;
                    sum += __shfl_sync(0xffffffffu, term, l);
                }
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const float si = (float)(-sum / 2.0);
        score[s] = si;
        if (!loop || si >= minScore) {
            const int c = atomicAdd(&counters[1], 1);
            key[c] = ((unsigned long long)(unsigned)firstWord[s] << 32) | (unsigned)s;
            candSlot[c] = s;
        }
    }
  }
}

// K4 + K5 in one CTA: group accumulation :193-224 / :337-364 (one candidate per thread step), the 0.75 * best gate,
// first-occurrence de-duplication and output position by rank among the candidates' keys, then the query's bits are wiped
// from the bitmap so that the next query starts from zeros without a memset.  Candidates are a handful to a few hundred; the
// O(C^2) rank loops of a single CTA stay below 40 us up to 8 k candidates.
__global__ void __launch_bounds__(1024)
kfdbFinalKernel(const int* __restrict__ counters, const int* __restrict__ candSlot, const unsigned long long* __restrict__ key,
                const int* __restrict__ dNeigh, const int* __restrict__ slotOfId, int idCap, const int* __restrict__ dKfId, int loop,
                long long qid, const long long* __restrict__ query, const int* __restrict__ words, const float* __restrict__ score,
                float bestAcc0, float* __restrict__ acc, int* __restrict__ best, int* __restrict__ emit, int32_t* __restrict__ out, int cap,
                int* __restrict__ nOut, const int32_t* __restrict__ qIds, int nq, int nwords, uint32_t* __restrict__ bitmap)
{
    __shared__ int sBestAcc, sOut;
    const int n = counters[1];
    if (threadIdx.x == 0) { sBestAcc = __float_as_int(bestAcc0); sOut = 0; }
    __syncthreads();
    const int minCommon = (int)((float)counters[0] * 0.8f);      // int * float -> float -> int, :161 / :312
    {   // one warp per candidate: lanes 0-9 fetch its covisible keyframes' state side by side, then the sum runs in list order
        const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5, nWarps = blockDim.x >> 5;
        for (int c = wrp; c < n; c += nWarps) {
            const int s = candSlot[c];
            const int id = lane < 10 ? dNeigh[s * 10 + lane] : -1;
            float sc = 0.f;
            bool use = false;
            if (id >= 0 && id < idCap) {
                const int s2 = slotOfId[id];                     // < 0: not (or no longer) in the database, never met by this query
                if (s2 >= 0 && query[s2] == qid && !(loop && !(words[s2] > minCommon))) { sc = score[s2]; use = true; }   // :206 / :346-348
            }
            // the list ends at the first -1 (set_covisible pads with -1), so entries behind a gap do not exist
            const uint32_t gap = __ballot_sync(0xffffffffu, lane < 10 && id < 0);
            const uint32_t valid = __ballot_sync(0xffffffffu, use) & (gap ? ((1u << (__ffs(gap) - 1)) - 1u) : 0x3ffu);
            float bestScore = score[s], a = bestScore;
            int bestId = dKfId[s];
            #pragma unroll
            for (int j = 0; j < 10; ++j) {
                const float scj = __shfl_sync(0xffffffffu, sc, j);
                const int idj = __shfl_sync(0xffffffffu, id, j);
                if ((valid >> j) & 1u) { a += scj; if (scj > bestScore) { bestId = idj; bestScore = scj; } }
            }
            if (lane == 0) { acc[c] = a; best[c] = bestId; atomicMaxFloat(&sBestAcc, a); }
        }
    }
    __syncthreads();
    const float retain = 0.75f * __int_as_float(sBestAcc);       // :227 / :368
    for (int c = threadIdx.x; c < n; c += blockDim.x) {          // kept, and not preceded (in list order) by a kept candidate with the same best keyframe
        int e = acc[c] > retain;
        if (e) {
            const unsigned long long k = key[c];
            const int b = best[c];
            for (int j = 0; j < n && e; ++j) if (key[j] < k && best[j] == b && acc[j] > retain) e = 0;
        }
        emit[c] = e;
    }
    __syncthreads();
    for (int c = threadIdx.x; c < n; c += blockDim.x) {          // output position = emitting candidates in front
        if (!emit[c]) continue;
        const unsigned long long k = key[c];
        int pos = 0;
        for (int j = 0; j < n; ++j) pos += (emit[j] && key[j] < k);
        if (pos < cap) out[pos] = best[c];
        atomicAdd(&sOut, 1);
    }
    for (int i = threadIdx.x; i < nq; i += blockDim.x) { const int w = qIds[i]; if (w >= 0 && w < nwords) bitmap[w >> 5] = 0u; }
    __syncthreads();
    if (threadIdx.x == 0) *nOut = sOut;
}

template <typename T> static int growArray(Ctx* c, T*& p, size_t oldN, size_t newN)
{
    T* q = nullptr;
    ORBX_CUDA(c, cudaMalloc(&q, sizeof(T) * newN));
    if (p && oldN) ORBX_CUDA(c, cudaMemcpyAsync(q, p, sizeof(T) * oldN, cudaMemcpyDeviceToDevice, c->stream));
    ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    if (p) cudaFree(p);
    p = q;
    return ORBX_OK;
}

} // namespace orbx
using namespace orbx;

static int kfdbGrowSlots(orbx_kfdb* d, int need)
{
    if (need <= d->slotsCap) return ORBX_OK;
    Ctx* c = &d->ctx->c;
    size_t n = d->slotsCap ? (size_t)d->slotsCap : 1024;
    while (n < (size_t)need) n *= 2;
    const size_t o = (size_t)d->nslots;
    int rc;
#define G(p) if ((rc = growArray(c, d->p, o, n))) return rc
    G(dSeg); G(dLen); G(dKfId); G(dAlive); G(dLoopQuery); G(dRelocQuery); G(dLoopWords); G(dRelocWords); G(dLoopScore); G(dRelocScore);
    G(dExclStamp); G(dInList); G(dFirstWord); G(dListed); G(dKey); G(dCandSlot); G(dBest); G(dEmit); G(dAcc);
#undef G
    if ((rc = growArray(c, d->dNeigh, o * 10, n * 10))) return rc;
    d->slotsCap = (int)n;
    return ORBX_OK;
}

static int kfdbGrowIds(orbx_kfdb* d, int id)
{
    if (id < d->idCap) return ORBX_OK;
    Ctx* c = &d->ctx->c;
    size_t n = d->idCap ? (size_t)d->idCap : 1024;
    while (n <= (size_t)id) n *= 2;
    int rc = growArray(c, d->dSlotOfId, (size_t)d->idCap, n);
    if (rc) return rc;
    const int fresh = (int)n - d->idCap;
    kfdbFillKernel<<<(fresh + 255) / 256, 256, 0, c->stream>>>(d->dSlotOfId + d->idCap, fresh, -1);
    ORBX_CUDA(c, cudaGetLastError());
    d->idCap = (int)n;
    return ORBX_OK;
}

// KeyFrameDatabase::KeyFrameDatabase, src/KeyFrameDatabase.cc:38-42
extern "C" int orbx_kfdb_create(orbx_ctx* h, int nwords, orbx_kfdb** out)
{
    if (!h || !out || nwords < 1) return ORBX_E_INVALID;
    *out = nullptr;
    Ctx* c = &h->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    orbx_kfdb* d = new (std::nothrow) orbx_kfdb();
    if (!d) return ORBX_E_NOMEM;
    d->ctx = h; d->nwords = nwords;
    int rc = ORBX_OK;
    if (cudaMalloc(&d->dBitmap, sizeof(uint32_t) * ((size_t)nwords / 32 + 1)) != cudaSuccess ||
        cudaMalloc(&d->dQPos, sizeof(int32_t) * (size_t)nwords) != cudaSuccess ||
        cudaMalloc(&d->dCounters, sizeof(int) * 8) != cudaSuccess) rc = ORBX_E_CUDA;
    if (!rc && cudaMemsetAsync(d->dBitmap, 0, sizeof(uint32_t) * ((size_t)nwords / 32 + 1), c->stream) != cudaSuccess) rc = ORBX_E_CUDA;
    if (!rc) rc = kfdbGrowSlots(d, 1);
    if (!rc) rc = kfdbGrowIds(d, 0);
    if (rc) { orbx_kfdb_destroy(d); return rc; }
    *out = d;
    return ORBX_OK;
}

extern "C" void orbx_kfdb_destroy(orbx_kfdb* d)
{
    if (!d) return;
    Ctx* c = &d->ctx->c;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    void* all[] = {d->dIds, d->dVals, d->dSeg, d->dLen, d->dKfId, d->dAlive, d->dNeigh, d->dLoopQuery, d->dRelocQuery, d->dLoopWords,
                   d->dRelocWords, d->dLoopScore, d->dRelocScore, d->dExclStamp, d->dInList, d->dFirstWord, d->dListed, d->dKey, d->dCandSlot, d->dBest,
                   d->dEmit, d->dAcc, d->dSlotOfId, d->dBitmap, d->dQPos, d->dQIds, d->dQVals, d->dConn, d->dCounters, d->dOut, d->dNOut};
    for (void* p : all) if (p) cudaFree(p);
    delete d;
}

extern "C" int orbx_kfdb_size(const orbx_kfdb* d) { return d ? (int)d->slotOf.size() : ORBX_E_INVALID; }

// KeyFrameDatabase::add, :50-56.  words: ascending word ids of pKF->mBowVec, vals their weights (HOST arrays).
extern "C" int orbx_kfdb_add(orbx_kfdb* d, int kf_id, const int32_t* words, const double* vals, int n)
{
    if (!d || kf_id < 0 || kf_id >= (1 << 28) || n < 0 || (n > 0 && (!words || !vals))) return ORBX_E_INVALID;
    if (d->slotOf.count(kf_id)) return ORBX_E_INVALID;                        // the reference would list it twice; nothing does that
    for (int i = 0; i < n; ++i) if (words[i] < 0 || words[i] >= d->nwords || (i && words[i] <= words[i - 1])) return ORBX_E_INVALID;
    Ctx* c = &d->ctx->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    int rc;
    if ((rc = kfdbGrowSlots(d, d->nslots + 1)) || (rc = kfdbGrowIds(d, kf_id))) return rc;
    if (d->entries + (size_t)n > d->entriesCap) {
        size_t cap = d->entriesCap ? d->entriesCap : ((size_t)1 << 20);
        while (cap < d->entries + (size_t)n) cap *= 2;
        if ((rc = growArray(c, d->dIds, d->entries, cap)) || (rc = growArray(c, d->dVals, d->entries, cap))) return rc;
        d->entriesCap = cap;
    }
    if (n) {
        ORBX_CUDA(c, cudaMemcpyAsync(d->dIds + d->entries, words, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
        ORBX_CUDA(c, cudaMemcpyAsync(d->dVals + d->entries, vals, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    }
    const int slot = d->nslots;
    kfdbInitSlotKernel<<<1, 32, 0, c->stream>>>(slot, (int)d->entries, n, kf_id, d->dSeg, d->dLen, d->dKfId, d->dAlive, d->dNeigh, d->dLoopQuery,
                                                d->dRelocQuery, d->dLoopWords, d->dRelocWords, d->dLoopScore, d->dRelocScore, d->dExclStamp, d->dSlotOfId);
    ORBX_CUDA(c, cudaGetLastError());
    c->kernelLaunches++;
    d->entries += (size_t)n; d->nslots++;
    d->slotOf[kf_id] = slot; d->idOfSlot.push_back(kf_id);
    return ORBX_OK;
}

// KeyFrameDatabase::erase, :64-83
extern "C" int orbx_kfdb_erase(orbx_kfdb* d, int kf_id)
{
    if (!d) return ORBX_E_INVALID;
    auto it = d->slotOf.find(kf_id);
    if (it == d->slotOf.end()) return ORBX_OK;                                // like the reference: nothing to remove
    Ctx* c = &d->ctx->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    kfdbEraseKernel<<<1, 1, 0, c->stream>>>(it->second, kf_id, d->dAlive, d->dSlotOfId);
    ORBX_CUDA(c, cudaGetLastError());
    c->kernelLaunches++;
    d->slotOf.erase(it);
    return ORBX_OK;
}

// KeyFrameDatabase::clear, :85-89
extern "C" int orbx_kfdb_clear(orbx_kfdb* d)
{
    if (!d) return ORBX_E_INVALID;
    Ctx* c = &d->ctx->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    if (d->idCap) { kfdbFillKernel<<<(d->idCap + 255) / 256, 256, 0, c->stream>>>(d->dSlotOfId, d->idCap, -1); ORBX_CUDA(c, cudaGetLastError()); }
    d->nslots = 0; d->entries = 0; d->slotOf.clear(); d->idOfSlot.clear();
    return ORBX_OK;
}

// The list pKF->GetBestCovisibilityKeyFrames(10) returns (src/KeyFrame.cc:248-258): keyframe ids, strongest first; the
// database keeps the first 10.  Ids that are not in the database (yet, or any more) are fine: such a keyframe is never met
// by a query.  Call it whenever the keyframe's ordered connections change (KeyFrame::UpdateBestCovisibles).
extern "C" int orbx_kfdb_set_covisible(orbx_kfdb* d, int kf_id, const int32_t* ids, int n)
{
    if (!d || n < 0 || (n > 0 && !ids)) return ORBX_E_INVALID;
    auto it = d->slotOf.find(kf_id);
    if (it == d->slotOf.end()) return ORBX_E_INVALID;
    Ctx* c = &d->ctx->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    Neigh10 nb;
    int m = 0;
    for (int i = 0; i < n && m < 10; ++i) if (ids[i] >= 0) nb.id[m++] = ids[i];   // the first 10 of the list
    for (; m < 10; ++m) nb.id[m] = -1;
    kfdbSetNeighKernel<<<1, 32, 0, c->stream>>>(it->second, nb, d->dNeigh);
    ORBX_CUDA(c, cudaGetLastError());
    c->kernelLaunches++;
    return ORBX_OK;
}

// The same for n keyframes in one call (one copy, one launch): kf_ids[i]'s list is neigh[10 * i ... 10 * i + 10), -1 padded.
// What a caller without a hook in KeyFrame::UpdateBestCovisibles does in front of a query: refresh every keyframe's list.
extern "C" int orbx_kfdb_set_covisible_bulk(orbx_kfdb* d, int n, const int32_t* kf_ids, const int32_t* neigh)
{
    if (!d || n < 0 || (n > 0 && (!kf_ids || !neigh))) return ORBX_E_INVALID;
    if (n == 0) return ORBX_OK;
    Ctx* c = &d->ctx->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    int rc = ensureScratch(c, sizeof(int) * (size_t)n * 11 + 256);
    if (rc) return rc;
    int* dI = (int*)c->dScratch; int* dN = dI + n;
    ORBX_CUDA(c, cudaMemcpyAsync(dI, kf_ids, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    ORBX_CUDA(c, cudaMemcpyAsync(dN, neigh, sizeof(int) * (size_t)n * 10, cudaMemcpyHostToDevice, c->stream));
    kfdbSetNeighBulkKernel<<<(n * 10 + 255) / 256, 256, 0, c->stream>>>(dI, dN, n, d->dSlotOfId, d->idCap, d->dNeigh);
    ORBX_CUDA(c, cudaGetLastError());
    c->kernelLaunches++;
    return ORBX_OK;
}

namespace orbx {
// enqueue one query; dQ* = the query's BowVector on the device; results to dOut[0, cap) / *dNOut
static int kfdbEnqueue(orbx_kfdb* d, int loop, long long qid, const int32_t* dQIds, const double* dQVals, int nq, int nconn, float minScore,
                       int32_t* dOut, int cap, int* dNOut)
{
    Ctx* c = &d->ctx->c;
    const int N = d->nslots;
    if (!N || !nq) { ORBX_CUDA(c, cudaMemsetAsync(dNOut, 0, sizeof(int), c->stream)); return ORBX_OK; }
    d->stamp++;
    const int nb = std::max(std::max(nq, nconn), 1);
    kfdbBeginKernel<<<(nb + 255) / 256, 256, 0, c->stream>>>(dQIds, nq, d->nwords, d->dBitmap, d->dQPos, d->dConn, nconn, d->dSlotOfId, d->idCap,
                                                            d->dExclStamp, d->stamp, d->dCounters, loop ? minScore : 0.f);
    long long* query = loop ? d->dLoopQuery : d->dRelocQuery;
    int* words = loop ? d->dLoopWords : d->dRelocWords;
    float* score = loop ? d->dLoopScore : d->dRelocScore;
    const int warpGrid = (N + 7) / 8;
    kfdbSharedKernel<<<warpGrid, 256, 0, c->stream>>>(N, d->dSeg, d->dLen, d->dAlive, d->dIds, d->dBitmap, loop, qid, d->dExclStamp, d->stamp,
                                                     query, words, d->dInList, d->dFirstWord, d->dCounters, d->dListed);
    kfdbScoreKernel<<<std::min(N, 296), 256, 0, c->stream>>>(N, d->dSeg, d->dLen, d->dIds, d->dVals, d->dBitmap, d->dQPos, dQVals, nq, loop, minScore,
                                                            d->dInList, words, d->dFirstWord, score, d->dCounters, d->dKey, d->dCandSlot, d->dListed);
    // the candidate count stays on the device: one CTA walks however many there are
    kfdbFinalKernel<<<1, 1024, 0, c->stream>>>(d->dCounters, d->dCandSlot, d->dKey, d->dNeigh, d->dSlotOfId, d->idCap, d->dKfId, loop, qid, query,
                                              words, score, loop ? minScore : 0.f, d->dAcc, d->dBest, d->dEmit, dOut, cap, dNOut, dQIds, nq,
                                              d->nwords, d->dBitmap);
    ORBX_CUDA(c, cudaGetLastError());
    return ORBX_OK;
}

static int kfdbEnsureQuery(orbx_kfdb* d, int nq, int nconn, int outInts, int queries)
{
    Ctx* c = &d->ctx->c;
    if (nq > d->qCap) {
        if (d->dQIds) { cudaStreamSynchronize(c->stream); cudaFree(d->dQIds); cudaFree(d->dQVals); d->dQIds = nullptr; d->dQVals = nullptr; d->qCap = 0; }
        const int cap = std::max(4096, nq + nq / 2);
        ORBX_CUDA(c, cudaMalloc(&d->dQIds, sizeof(int32_t) * (size_t)cap));
        ORBX_CUDA(c, cudaMalloc(&d->dQVals, sizeof(double) * (size_t)cap));
        d->qCap = cap;
    }
    if (nconn > d->connCap) {
        if (d->dConn) { cudaStreamSynchronize(c->stream); cudaFree(d->dConn); d->dConn = nullptr; d->connCap = 0; }
        const int cap = std::max(1024, 2 * nconn);
        ORBX_CUDA(c, cudaMalloc(&d->dConn, sizeof(int) * (size_t)cap));
        d->connCap = cap;
    }
    if (outInts > d->outCap || queries > d->outQueries) {
        if (d->dOut) { cudaStreamSynchronize(c->stream); cudaFree(d->dOut); cudaFree(d->dNOut); d->dOut = nullptr; d->dNOut = nullptr; }
        d->outCap = std::max(outInts, 4096); d->outQueries = std::max(queries, 64);
        ORBX_CUDA(c, cudaMalloc(&d->dOut, sizeof(int32_t) * (size_t)d->outCap));
        ORBX_CUDA(c, cudaMalloc(&d->dNOut, sizeof(int) * (size_t)d->outQueries));
    }
    return ORBX_OK;
}

static bool ascending(const int32_t* w, int n) { for (int i = 1; i < n; ++i) if (w[i] <= w[i - 1]) return false; return true; }
} // namespace orbx

// KeyFrameDatabase::DetectLoopCandidates(pKF, minScore), :113-270.  query_id = pKF->mnId, words / vals = pKF->mBowVec (ascending
// word ids), connected = ids of pKF->GetConnectedKeyFrames().  out_ids receives the candidates in the reference's order,
// *n_out their number (which may exceed cap: the first cap are written).
extern "C" int orbx_kfdb_detect_loop(orbx_kfdb* d, long long query_id, const int32_t* words, const double* vals, int nq, const int32_t* connected,
                                     int nconnected, float min_score, int32_t* out_ids, int cap, int* n_out)
{
    if (!d || nq < 0 || nconnected < 0 || cap < 0 || !n_out || (nq > 0 && (!words || !vals)) || (nconnected > 0 && !connected) || (cap > 0 && !out_ids))
        return ORBX_E_INVALID;
    if (!ascending(words, nq)) return ORBX_E_INVALID;
    Ctx* c = &d->ctx->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    int rc = kfdbEnsureQuery(d, nq, nconnected, std::max(cap, 1), 1);
    if (rc) return rc;
    if (nq) { ORBX_CUDA(c, cudaMemcpyAsync(d->dQIds, words, sizeof(int32_t) * (size_t)nq, cudaMemcpyHostToDevice, c->stream));
              ORBX_CUDA(c, cudaMemcpyAsync(d->dQVals, vals, sizeof(double) * (size_t)nq, cudaMemcpyHostToDevice, c->stream)); }
    if (nconnected) ORBX_CUDA(c, cudaMemcpyAsync(d->dConn, connected, sizeof(int) * (size_t)nconnected, cudaMemcpyHostToDevice, c->stream));
    stageBegin(c, ST_BOW);
    rc = kfdbEnqueue(d, 1, query_id, d->dQIds, d->dQVals, nq, nconnected, min_score, d->dOut, cap, d->dNOut);
    stageEnd(c, ST_BOW, (d->nslots && nq) ? 4 : 0);
    if (rc) return rc;
    int n = 0;
    ORBX_CUDA(c, cudaMemcpyAsync(&n, d->dNOut, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    if (cap) ORBX_CUDA(c, cudaMemcpyAsync(out_ids, d->dOut, sizeof(int32_t) * (size_t)cap, cudaMemcpyDeviceToHost, c->stream));
    ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    *n_out = n;
    return ORBX_OK;
}

// KeyFrameDatabase::DetectRelocalizationCandidates(F), :275-399, for nqueries frames per call, in order (a later query sees the
// mRelocScore values the earlier ones left, as it would in the reference).  query_ids[q] = F->mnId; the BowVector of query q
// is words / vals [seg[q], seg[q + 1]) (HOST arrays, seg has nqueries + 1 entries).  Candidates of query q go to
// out_ids[q * cap ...], their number to n_out[q].  One synchronisation for the whole batch.
extern "C" int orbx_kfdb_detect_reloc(orbx_kfdb* d, int nqueries, const long long* query_ids, const int32_t* words, const double* vals,
                                      const int32_t* seg, int32_t* out_ids, int cap, int* n_out)
{
    if (!d || nqueries < 0 || cap < 0 || (nqueries > 0 && (!query_ids || !seg || !n_out)) || (cap > 0 && nqueries > 0 && !out_ids)) return ORBX_E_INVALID;
    if (nqueries == 0) return ORBX_OK;
    const int total = seg[nqueries];
    if (seg[0] != 0 || total < 0 || (total > 0 && (!words || !vals))) return ORBX_E_INVALID;
    for (int q = 0; q < nqueries; ++q) if (seg[q + 1] < seg[q] || !ascending(words + seg[q], seg[q + 1] - seg[q])) return ORBX_E_INVALID;
    if ((size_t)nqueries * (size_t)std::max(cap, 1) > ((size_t)1 << 30)) return ORBX_E_CAPACITY;
    Ctx* c = &d->ctx->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    int rc = kfdbEnsureQuery(d, total, 0, nqueries * std::max(cap, 1), nqueries);
    if (rc) return rc;
    if (total) { ORBX_CUDA(c, cudaMemcpyAsync(d->dQIds, words, sizeof(int32_t) * (size_t)total, cudaMemcpyHostToDevice, c->stream));
                 ORBX_CUDA(c, cudaMemcpyAsync(d->dQVals, vals, sizeof(double) * (size_t)total, cudaMemcpyHostToDevice, c->stream)); }
    stageBegin(c, ST_BOW);
    int launches = 0;
    for (int q = 0; q < nqueries && !rc; ++q) {
        const int nq = seg[q + 1] - seg[q];
        rc = kfdbEnqueue(d, 0, query_ids[q], d->dQIds + seg[q], d->dQVals + seg[q], nq, 0, 0.f, d->dOut + (size_t)q * cap, cap, d->dNOut + q);
        if (d->nslots && nq) launches += 4;
    }
    stageEnd(c, ST_BOW, launches);
    if (rc) return rc;
    ORBX_CUDA(c, cudaMemcpyAsync(n_out, d->dNOut, sizeof(int) * (size_t)nqueries, cudaMemcpyDeviceToHost, c->stream));
    if (cap) ORBX_CUDA(c, cudaMemcpyAsync(out_ids, d->dOut, sizeof(int32_t) * (size_t)nqueries * cap, cudaMemcpyDeviceToHost, c->stream));
    ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    return ORBX_OK;
}
