// Greedy BoW matchers on the GPU, bit-equal to the reference's order-dependent loops:
//   ORBmatcher::SearchByBoW(KeyFrame*, Frame&, ...)      src/ORBmatcher.cc:255-419   (mode KF_FRAME)
//   ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, ...)   src/ORBmatcher.cc:760-903   (mode KF_KF)
//   ORBmatcher::SearchForTriangulation                   src/ORBmatcher.cc:915-1127  (mode TRIANGULATE)
//
// The reference walks the features of a shared vocabulary node in order; each query row scans the node's
// train rows, SKIPPING rows claimed by earlier queries, keeps best / second-best, and claims its best on
// acceptance.  Nodes are independent, rows inside a node are not.  Three launches:
//   A  bowTopKKernel   (the POPC work, fully parallel): every query row gets its K best statically-valid
//      train rows of the node, ordered by (distance, scan position) -- first-wins for SearchByBoW,
//      last-wins for SearchForTriangulation (its `dist > bestDist` rule, :1026).
//   B  bowResolveKernel (one warp per node, rows in reference order): drops claimed entries from the
//      K-list; the first two survivors ARE the reference's best / second-best whenever the list is not
//      exhausted; otherwise the warp rescans the node for that row (rare, exact).  Acceptance test,
//      claim, rotation-histogram vote.
//   C  bowRotationKernel (one CTA per job): ComputeThreeMaxima (:2076-2118) and removal of the other bins.
#include "orbx_internal.h"
#include <string.h>
#include <algorithm>

namespace orbx {

constexpr int kTopK = 4;
constexpr int kThLow = 50, kHisto = 30;      // TH_LOW, HISTO_LENGTH  src/ORBmatcher.cc:57-58
constexpr int kATh = 128, kAQpt = 1, kAChunk = 128, kARows = kATh * kAQpt;   // one query row per thread: more CTAs per node, more warps per SM
constexpr uint32_t kNoKey = 0xffffffffu;
constexpr int kPosBits = 20;
constexpr uint32_t kPosMask = (1u << kPosBits) - 1;

struct BowJob {
    const uint4 *desc1, *desc2;
    const orbx_keypoint *keys1, *keys2;
    const uint8_t *flag1, *flag2;
    const float *ur1, *ur2;
    const int32_t *idx1, *idx2;
    const int4* seg;              // {qb, qe, tb, te} positions into idx1 / idx2
    const int32_t* tilePrefix;    // nseg + 1
    uint32_t* topk;               // [len(idx1)][kTopK]
    int32_t* match;
    int* hist;                    // kHisto bins + [kHisto] = nmatches
    int nseg, n1, n2, mode, checkOri, onlyStereo;
    float nnratio, ex, ey;
    float F[9], scale2[kMaxLevels], sigma2[kMaxLevels];
};

__device__ __forceinline__ int dist256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1)
{
    return hamming256(a0, a1, b0, b1);
}

__device__ __forceinline__ bool queryEligible(const BowJob& J, int i1)
{
    const bool has = J.flag1 && J.flag1[i1];
    if (J.mode != ORBX_BOW_TRIANGULATE) return has;                 // :307-313, :802-808
    if (has) return false;                                           // :981
    if (J.onlyStereo && !(J.ur1 && J.ur1[i1] >= 0)) return false;    // :986-990
    return true;
}

__device__ __forceinline__ bool trainEligible(const BowJob& J, int i2)
{
    if (J.mode == ORBX_BOW_KF_FRAME) return true;
    const bool has = J.flag2 && J.flag2[i2];
    if (J.mode == ORBX_BOW_KF_KF) return has;                        // :822-826
    if (has) return false;                                           // :1011
    if (J.onlyStereo && !(J.ur2 && J.ur2[i2] >= 0)) return false;
    return true;
}

// Epipolar-line coefficients of a KF1 keypoint (CheckDistEpipolarLine :217-221), plain rounded float ops.
struct Epi { float a, b, c, den; bool stereo1; };
__device__ __forceinline__ Epi makeEpi(const BowJob& J, int i1)
{
    Epi e;
    const float x = J.keys1[i1].x, y = J.keys1[i1].y;
    e.a = __fadd_rn(__fadd_rn(__fmul_rn(x, J.F[0]), __fmul_rn(y, J.F[3])), J.F[6]);
    e.b = __fadd_rn(__fadd_rn(__fmul_rn(x, J.F[1]), __fmul_rn(y, J.F[4])), J.F[7]);
    e.c = __fadd_rn(__fadd_rn(__fmul_rn(x, J.F[2]), __fmul_rn(y, J.F[5])), J.F[8]);
    e.den = __fadd_rn(__fmul_rn(e.a, e.a), __fmul_rn(e.b, e.b));
    e.stereo1 = J.ur1 && J.ur1[i1] >= 0;
    return e;
}

// geometric gates of SearchForTriangulation for one (kp1, kp2) pair, :1034-1052
__device__ __forceinline__ bool triGeometryOk(const BowJob& J, const Epi& e, float x2, float y2, int oct2, bool stereo2)
{
    if (!e.stereo1 && !stereo2) {
        const float dx = __fsub_rn(J.ex, x2), dy = __fsub_rn(J.ey, y2);
        if (__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < __fmul_rn(100.f, J.scale2[oct2])) return false;
    }
    const float num = __fadd_rn(__fadd_rn(__fmul_rn(e.a, x2), __fmul_rn(e.b, y2)), e.c);
    if (e.den == 0.f) return false;
    const float dsqr = __fdiv_rn(__fmul_rn(num, num), e.den);
    return (double)dsqr < __dmul_rn(3.84, (double)J.sigma2[oct2]);
}

__device__ __forceinline__ void insertKey(uint32_t (&t)[kTopK], uint32_t key)
{
    if (key < t[kTopK - 1]) {
        t[kTopK - 1] = key;
        #pragma unroll
        for (int j = kTopK - 1; j > 0; --j) {
            const uint32_t lo = min(t[j - 1], t[j]), hi = max(t[j - 1], t[j]);
            t[j - 1] = lo; t[j] = hi;
        }
    }
}

__global__ void __launch_bounds__(kATh)
bowTopKKernel(const BowJob* __restrict__ jobs)
{
    const BowJob& J = jobs[blockIdx.y];
    __shared__ uint4 sT[kAChunk * 2];
    __shared__ int sOk[kAChunk];        // train feature index, or -1 when statically invalid
    __shared__ float sX[kAChunk], sY[kAChunk];
    __shared__ int sOct[kAChunk];       // octave | stereo2 << 8
    int lo = 0, hi = J.nseg;
    if ((int)blockIdx.x >= J.tilePrefix[J.nseg]) return;
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (J.tilePrefix[mid] <= (int)blockIdx.x) lo = mid; else hi = mid; }
    const int4 sg = J.seg[lo];
    const int qBase = sg.x + ((int)blockIdx.x - J.tilePrefix[lo]) * kARows;
    const bool tri = J.mode == ORBX_BOW_TRIANGULATE;

    uint4 qa[kAQpt], qb[kAQpt];
    uint32_t top[kAQpt][kTopK];
    Epi epi[kAQpt];
    int row[kAQpt]; bool act[kAQpt];
    #pragma unroll
    for (int k = 0; k < kAQpt; ++k) {
        row[k] = qBase + k * kATh + threadIdx.x;
        act[k] = false;
        #pragma unroll
        for (int j = 0; j < kTopK; ++j) top[k][j] = kNoKey;
        qa[k] = make_uint4(0, 0, 0, 0); qb[k] = qa[k];
        epi[k].a = epi[k].b = epi[k].c = epi[k].den = 0.f; epi[k].stereo1 = false;
        if (row[k] < sg.y) {
            const int i1 = J.idx1[row[k]];
            if (queryEligible(J, i1)) {
                act[k] = true;
                qa[k] = J.desc1[2 * i1]; qb[k] = J.desc1[2 * i1 + 1];
                if (tri) epi[k] = makeEpi(J, i1);
            }
        }
    }
    for (int c = sg.z; c < sg.w; c += kAChunk) {
        const int m = min(kAChunk, sg.w - c);
        __syncthreads();
        for (int i = threadIdx.x; i < m; i += kATh) {
            const int i2 = J.idx2[c + i];
            const bool ok = trainEligible(J, i2);
            sOk[i] = ok ? i2 : -1;
            sT[2 * i] = J.desc2[2 * i2]; sT[2 * i + 1] = J.desc2[2 * i2 + 1];
            if (tri) {
                const orbx_keypoint kp = J.keys2[i2];
                sX[i] = kp.x; sY[i] = kp.y;
                sOct[i] = kp.octave | ((J.ur2 && J.ur2[i2] >= 0) ? 256 : 0);
            }
        }
        __syncthreads();
        for (int i = 0; i < m; ++i) {
            if (sOk[i] < 0) continue; // uniform
            const uint4 ta = sT[2 * i], tb = sT[2 * i + 1];
            const uint32_t pos = (uint32_t)(c + i - sg.z);
            #pragma unroll
            for (int k = 0; k < kAQpt; ++k) {
                const int d = dist256(qa[k], qb[k], ta, tb);
                if (!tri) {
                    const uint32_t key = ((uint32_t)d << kPosBits) | pos;
                    if (key < top[k][kTopK - 1]) insertKey(top[k], key);   // rare once the list has warmed up
                } else if (d <= kThLow) {
                    const uint32_t key = ((uint32_t)d << kPosBits) | (kPosMask - pos); // last wins ties
                    if (act[k] && key < top[k][kTopK - 1] && triGeometryOk(J, epi[k], sX[i], sY[i], sOct[i] & 255, (sOct[i] & 256) != 0))
                        insertKey(top[k], key);
                }
            }
        }
    }
    #pragma unroll
    for (int k = 0; k < kAQpt; ++k)
        if (act[k]) *reinterpret_cast<uint4*>(J.topk + (size_t)row[k] * kTopK) = make_uint4(top[k][0], top[k][1], top[k][2], top[k][3]);
}

__device__ __forceinline__ int rotBin(float a1, float a2)
{   // :364-371 with factor = HISTO_LENGTH/360.0f (:275)
    float rot = __fsub_rn(a1, a2);
    if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
    int bin = (int)roundf(__fmul_rn(rot, 30 / 360.0f));
    if (bin == kHisto) bin = 0;
    return bin;
}

// merge two ascending key pairs, keep the two smallest
__device__ __forceinline__ void merge2(uint32_t& a1, uint32_t& a2, uint32_t b1, uint32_t b2)
{
    const uint32_t lo = min(a1, b1), hi = max(a1, b1);
    a2 = min(hi, min(a2, b2));
    a1 = lo;
}

constexpr int kBWarps = 4;

// Greedy replay of one vocabulary node by one warp, in the reference's row order.  Rows are taken 32 at a time: lane l
// loads everything row r0+l needs (global memory, all lanes in parallel), then the lanes take their turns in order, each
// turn touching only its own registers and the node's claim mask in shared memory; accepted matches are written back
// in parallel after the batch.  A row whose top-K list cannot decide (all listed rows claimed, or the second-best
// unknown) is rescanned exactly by the whole warp.
__global__ void __launch_bounds__(kBWarps * 32)
bowResolveKernel(const BowJob* __restrict__ jobs, int maskWords)
{
    extern __shared__ uint32_t sMaskAll[];
    const BowJob& J = jobs[blockIdx.y];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int s = blockIdx.x * kBWarps + w;
    if (s >= J.nseg) return;
    uint32_t* mask = sMaskAll + (size_t)w * maskWords;
    const int4 sg = J.seg[s];
    const int nt = sg.w - sg.z;
    for (int i = lane; i < (nt + 31) / 32; i += 32) mask[i] = 0;
    __syncwarp();
    const bool tri = J.mode == ORBX_BOW_TRIANGULATE;
    const int thAccept = J.mode == ORBX_BOW_KF_KF ? kThLow - 1 : kThLow; // '<' vs '<=' (:845 vs :348)
    const float nnratio = J.nnratio;

    // accept test on (best, second) keys; returns the claimed scan position or -1
    auto settle = [&](uint32_t u0, uint32_t u1) -> int {
        if (u0 == kNoKey) return -1;
        const int best = (int)(u0 >> kPosBits);
        bool accept;
        if (tri) accept = true; // d <= TH_LOW and geometry were applied when the key was built
        else {
            const int second = u1 == kNoKey ? 256 : (int)(u1 >> kPosBits);
            accept = best <= thAccept && (float)best < __fmul_rn(nnratio, (float)second);
        }
        if (!accept) return -1;
        const uint32_t p = tri ? kPosMask - (u0 & kPosMask) : (u0 & kPosMask);
        mask[p >> 5] |= 1u << (p & 31);
        return (int)p;
    };

    for (int r0 = sg.x; r0 < sg.y; r0 += 32) {
        const int r = r0 + lane;
        int i1 = -1; bool elig = false;
        uint32_t e[kTopK] = {kNoKey, kNoKey, kNoKey, kNoKey};
        if (r < sg.y) {
            i1 = J.idx1[r];
            elig = queryEligible(J, i1);
            if (elig) {
                const uint4 e4 = *reinterpret_cast<const uint4*>(J.topk + (size_t)r * kTopK);
                e[0] = e4.x; e[1] = e4.y; e[2] = e4.z; e[3] = e4.w;
            }
        }
        int accP = -1;
        // a row whose best listed distance is already beyond the acceptance threshold can never claim anything
        // (claims only remove candidates): it needs no turn
        const bool canAccept = elig && e[0] != kNoKey && (tri || (int)(e[0] >> kPosBits) <= thAccept);
        uint32_t turns = __ballot_sync(0xffffffffu, canAccept);
        while (turns) {
            const int t = __ffs(turns) - 1;
            turns &= turns - 1;
            bool need = false;
            if (lane == t) {
                uint32_t u0 = kNoKey, u1 = kNoKey;
                const bool full = e[kTopK - 1] != kNoKey;
                #pragma unroll
                for (int j = 0; j < kTopK; ++j) {
                    if (e[j] == kNoKey) continue;
                    const uint32_t p = tri ? kPosMask - (e[j] & kPosMask) : (e[j] & kPosMask);
                    const bool taken = (mask[p >> 5] >> (p & 31)) & 1u;
                    if (!taken) { if (u0 == kNoKey) u0 = e[j]; else if (u1 == kNoKey) u1 = e[j]; }
                }
                // every unlisted row is at least as far as the last listed one: when that is already beyond the acceptance
                // threshold nothing can be accepted and the rescan is pointless (SearchForTriangulation lists only hold
                // acceptable candidates, so there it must be done)
                if (u0 == kNoKey) need = full && (tri || (int)(e[kTopK - 1] >> kPosBits) <= thAccept);
                else if (!tri && (int)(u0 >> kPosBits) <= thAccept && u1 == kNoKey && full) {
                    // second-best unknown: every unlisted row is >= the last listed distance
                    const float lb = (float)(e[kTopK - 1] >> kPosBits);
                    if (!((float)(u0 >> kPosBits) < __fmul_rn(nnratio, lb))) need = true;
                    else u1 = e[kTopK - 1]; // accepted whatever the true second is
                }
                if (!need) accP = settle(u0, u1);
            }
            __syncwarp();
            if (__shfl_sync(0xffffffffu, (int)need, t)) { // exact rescan of the node for this row under the current claims
                const int i1t = __shfl_sync(0xffffffffu, i1, t);
                const uint4 qa = J.desc1[2 * i1t], qb = J.desc1[2 * i1t + 1];
                Epi epi; if (tri) epi = makeEpi(J, i1t);
                uint32_t k1 = kNoKey, k2 = kNoKey;
                for (int p = lane; p < nt; p += 32) {
                    if ((mask[p >> 5] >> (p & 31)) & 1u) continue;
                    const int i2 = J.idx2[sg.z + p];
                    if (!trainEligible(J, i2)) continue;
                    const int d = dist256(qa, qb, J.desc2[2 * i2], J.desc2[2 * i2 + 1]);
                    uint32_t key;
                    if (!tri) key = ((uint32_t)d << kPosBits) | (uint32_t)p;
                    else {
                        if (d > kThLow) continue;
                        const orbx_keypoint kp = J.keys2[i2];
                        if (!triGeometryOk(J, epi, kp.x, kp.y, kp.octave, J.ur2 && J.ur2[i2] >= 0)) continue;
                        key = ((uint32_t)d << kPosBits) | (kPosMask - (uint32_t)p);
                    }
                    if (key < k1) { k2 = k1; k1 = key; } else if (key < k2) k2 = key;
                }
                #pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const uint32_t b1 = __shfl_xor_sync(0xffffffffu, k1, o), b2 = __shfl_xor_sync(0xffffffffu, k2, o);
                    merge2(k1, k2, b1, b2);
                }
                __syncwarp();
                if (lane == t) accP = settle(k1, k2);
                __syncwarp();
            }
        }
        if (accP >= 0) {                                   // results of the batch, all lanes in parallel
            const int i2 = J.idx2[sg.z + accP];
            if (J.mode == ORBX_BOW_KF_FRAME) J.match[i2] = i1; else J.match[i1] = i2;
            if (J.checkOri) atomicAdd(&J.hist[rotBin(J.keys1[i1].angle, J.keys2[i2].angle)], 1);
        }
    }
}

__global__ void __launch_bounds__(256)
bowRotationKernel(const BowJob* __restrict__ jobs)
{
    const BowJob& J = jobs[blockIdx.x];
    __shared__ int keep[3];
    __shared__ int total;
    if (threadIdx.x == 0) {
        total = 0;
        int max1 = 0, max2 = 0, max3 = 0, i1 = -1, i2 = -1, i3 = -1;
        if (J.checkOri) {
            for (int i = 0; i < kHisto; ++i) {
                const int s = J.hist[i];
                if (s > max1) { max3 = max2; max2 = max1; max1 = s; i3 = i2; i2 = i1; i1 = i; }
                else if (s > max2) { max3 = max2; max2 = s; i3 = i2; i2 = i; }
                else if (s > max3) { max3 = s; i3 = i; }
            }
            if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { i2 = -1; i3 = -1; }
            else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) i3 = -1;
        }
        keep[0] = i1; keep[1] = i2; keep[2] = i3;
    }
    __syncthreads();
    const int n = J.mode == ORBX_BOW_KF_FRAME ? J.n2 : J.n1;
    int cnt = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int m = J.match[i];
        if (m < 0) continue;
        if (J.checkOri) {
            const int bin = J.mode == ORBX_BOW_KF_FRAME ? rotBin(J.keys1[m].angle, J.keys2[i].angle)
                                                        : rotBin(J.keys1[i].angle, J.keys2[m].angle);
            if (bin != keep[0] && bin != keep[1] && bin != keep[2]) { J.match[i] = -1; continue; }
        }
        ++cnt;
    }
    atomicAdd(&total, cnt);
    __syncthreads();
    if (threadIdx.x == 0) J.hist[kHisto] = total;
}

// bump allocator over the ctx scratch buffer
struct Bump {
    size_t off = 0;
    size_t take(size_t bytes) { const size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; }
};

int launchBowJobs(Ctx* c, const BowJob* dJobs, int njobs, int maxTiles, int maxSeg, int maxN2)
{
    if (njobs == 0) return ORBX_OK;
    stageBegin(c, ST_MATCH);
    int launches = 0;
    if (maxTiles > 0) {
        bowTopKKernel<<<dim3(maxTiles, njobs), kATh, 0, c->stream>>>(dJobs);
        ++launches;
        const int maskWords = (maxN2 + 31) / 32 + 1;
        const size_t smem = (size_t)kBWarps * maskWords * sizeof(uint32_t);
        if (smem > 200 * 1024) return ORBX_E_CAPACITY;
        if (smem > 48 * 1024 && smem > c->smemCfg[SMEM_BOW_RESOLVE]) {
            ORBX_CUDA(c, cudaFuncSetAttribute(bowResolveKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            c->smemCfg[SMEM_BOW_RESOLVE] = smem;
        }
        bowResolveKernel<<<dim3((maxSeg + kBWarps - 1) / kBWarps, njobs), kBWarps * 32, smem, c->stream>>>(dJobs, maskWords);
        ++launches;
    }
    bowRotationKernel<<<njobs, 256, 0, c->stream>>>(dJobs);
    ++launches;
    stageEnd(c, ST_MATCH, launches);
    ORBX_CUDA(c, cudaGetLastError());
    return ORBX_OK;
}

} // namespace orbx
using namespace orbx;

// One or many keyframe-pair jobs in ONE set of launches (grid.y = job): the top-K kernels of all pairs fill the GPU while
// the per-node greedy replays (one warp per node) of different pairs run side by side.
static int matchBowPairs(Ctx* c, int mode, int npairs, const orbx_bow_side* const* s1v, const orbx_bow_side* const* s2v,
                         const orbx_triangulation* const* triv, float nnratio, int check_ori, int flags,
                         int32_t* const* matchv, int* nmatches)
{
    const bool inDev = flags & ORBX_IN_DEVICE, outDev = flags & ORBX_OUT_DEVICE;
    struct Plan {
        std::vector<int4> segs; std::vector<int32_t> tilePrefix;
        int nseg = 0, len1 = 0, len2 = 0, nOut = 0, maxT = 0;
        size_t oSeg = 0, oTp = 0, oTop = 0, oHist = 0, oIdx1 = 0, oIdx2 = 0, oD1 = 0, oD2 = 0, oK1 = 0, oK2 = 0, oF1 = 0, oF2 = 0, oU1 = 0, oU2 = 0, oMatch = 0;
    };
    std::vector<Plan> plans((size_t)npairs);
    Bump B;
    const size_t oJobs = B.take(sizeof(BowJob) * (size_t)npairs);
    int maxTiles = 0, maxSeg = 0, maxN2 = 0;
    for (int j = 0; j < npairs; ++j) {
        const orbx_bow_side *s1 = s1v[j], *s2 = s2v[j];
        if (!s1 || !s2 || !matchv[j] || (mode == ORBX_BOW_TRIANGULATE && !(triv && triv[j]))) return ORBX_E_INVALID;
        if (s1->n < 0 || s2->n < 0 || s1->nnodes < 0 || s2->nnodes < 0) return ORBX_E_INVALID;
        if ((s1->n > 0 && (!s1->desc || !s1->keys)) || (s2->n > 0 && (!s2->desc || !s2->keys))) return ORBX_E_INVALID;
        Plan& P = plans[j];
        P.nOut = mode == ORBX_BOW_KF_FRAME ? s2->n : s1->n;
        P.tilePrefix.assign(1, 0);
        // merge-join of the two FeatureVectors (the while/lower_bound loops, e.g. :285-390)
        for (int a = 0, b = 0; a < s1->nnodes && b < s2->nnodes;) {
            if (s1->node_id[a] < s2->node_id[b]) ++a;
            else if (s1->node_id[a] > s2->node_id[b]) ++b;
            else {
                const int4 sg = make_int4(s1->node_off[a], s1->node_off[a + 1], s2->node_off[b], s2->node_off[b + 1]);
                if (sg.y < sg.x || sg.w < sg.z || sg.w - sg.z > (int)kPosMask) return ORBX_E_INVALID;
                if (sg.y > sg.x && sg.w > sg.z) {
                    P.segs.push_back(sg);
                    P.tilePrefix.push_back(P.tilePrefix.back() + (sg.y - sg.x + kARows - 1) / kARows);
                    P.maxT = std::max(P.maxT, sg.w - sg.z);
                }
                ++a; ++b;
            }
        }
        P.nseg = (int)P.segs.size();
        P.len1 = s1->nnodes ? s1->node_off[s1->nnodes] : 0; P.len2 = s2->nnodes ? s2->node_off[s2->nnodes] : 0;
        if ((P.len1 > 0 && !s1->feat) || (P.len2 > 0 && !s2->feat) || P.len1 < 0 || P.len2 < 0) return ORBX_E_INVALID;
        for (int i = 0; i < P.len1; ++i) if ((unsigned)s1->feat[i] >= (unsigned)s1->n) return ORBX_E_INVALID;   // the kernels index desc / keys with these
        for (int i = 0; i < P.len2; ++i) if ((unsigned)s2->feat[i] >= (unsigned)s2->n) return ORBX_E_INVALID;
        P.oSeg = B.take(sizeof(int4) * (size_t)std::max(P.nseg, 1)); P.oTp = B.take(sizeof(int32_t) * (size_t)(P.nseg + 1));
        P.oTop = B.take(sizeof(uint32_t) * kTopK * (size_t)std::max(P.len1, 1)); P.oHist = B.take(sizeof(int) * (kHisto + 1));
        P.oIdx1 = B.take(sizeof(int32_t) * (size_t)std::max(P.len1, 1)); P.oIdx2 = B.take(sizeof(int32_t) * (size_t)std::max(P.len2, 1));
        if (!inDev) {
            P.oD1 = B.take((size_t)32 * std::max(s1->n, 1)); P.oD2 = B.take((size_t)32 * std::max(s2->n, 1));
            P.oK1 = B.take(sizeof(orbx_keypoint) * (size_t)std::max(s1->n, 1)); P.oK2 = B.take(sizeof(orbx_keypoint) * (size_t)std::max(s2->n, 1));
            P.oF1 = B.take((size_t)std::max(s1->n, 1)); P.oF2 = B.take((size_t)std::max(s2->n, 1));
            P.oU1 = B.take(sizeof(float) * (size_t)std::max(s1->n, 1)); P.oU2 = B.take(sizeof(float) * (size_t)std::max(s2->n, 1));
        }
        if (!outDev) P.oMatch = B.take(sizeof(int32_t) * (size_t)std::max(P.nOut, 1));
        maxTiles = std::max(maxTiles, P.tilePrefix.back()); maxSeg = std::max(maxSeg, P.nseg); maxN2 = std::max(maxN2, P.maxT);
    }
    int rc = ensureScratch(c, B.off);
    if (rc) return rc;
    uint8_t* S = (uint8_t*)c->dScratch;
    auto h2d = [&](size_t off, const void* src, size_t bytes) -> cudaError_t {
        return bytes ? cudaMemcpyAsync(S + off, src, bytes, cudaMemcpyHostToDevice, c->stream) : cudaSuccess;
    };
    std::vector<BowJob> jobs((size_t)npairs);
    stageBegin(c, ST_H2D);
    for (int j = 0; j < npairs; ++j) {
        const orbx_bow_side *s1 = s1v[j], *s2 = s2v[j];
        const Plan& P = plans[j];
        BowJob& J = jobs[j];
        memset(&J, 0, sizeof(J));
        if (!inDev) {
            ORBX_CUDA(c, h2d(P.oD1, s1->desc, (size_t)32 * s1->n)); ORBX_CUDA(c, h2d(P.oD2, s2->desc, (size_t)32 * s2->n));
            ORBX_CUDA(c, h2d(P.oK1, s1->keys, sizeof(orbx_keypoint) * (size_t)s1->n)); ORBX_CUDA(c, h2d(P.oK2, s2->keys, sizeof(orbx_keypoint) * (size_t)s2->n));
            J.desc1 = (const uint4*)(S + P.oD1); J.desc2 = (const uint4*)(S + P.oD2);
            J.keys1 = (const orbx_keypoint*)(S + P.oK1); J.keys2 = (const orbx_keypoint*)(S + P.oK2);
            if (s1->has_point) { ORBX_CUDA(c, h2d(P.oF1, s1->has_point, (size_t)s1->n)); J.flag1 = S + P.oF1; }
            if (s2->has_point) { ORBX_CUDA(c, h2d(P.oF2, s2->has_point, (size_t)s2->n)); J.flag2 = S + P.oF2; }
            if (s1->uright) { ORBX_CUDA(c, h2d(P.oU1, s1->uright, sizeof(float) * (size_t)s1->n)); J.ur1 = (const float*)(S + P.oU1); }
            if (s2->uright) { ORBX_CUDA(c, h2d(P.oU2, s2->uright, sizeof(float) * (size_t)s2->n)); J.ur2 = (const float*)(S + P.oU2); }
        } else {
            J.desc1 = (const uint4*)s1->desc; J.desc2 = (const uint4*)s2->desc; J.keys1 = s1->keys; J.keys2 = s2->keys;
            J.flag1 = s1->has_point; J.flag2 = s2->has_point; J.ur1 = s1->uright; J.ur2 = s2->uright;
        }
        ORBX_CUDA(c, h2d(P.oIdx1, s1->feat, sizeof(int32_t) * (size_t)P.len1));
        ORBX_CUDA(c, h2d(P.oIdx2, s2->feat, sizeof(int32_t) * (size_t)P.len2));
        ORBX_CUDA(c, h2d(P.oSeg, P.segs.data(), sizeof(int4) * (size_t)P.nseg));
        ORBX_CUDA(c, h2d(P.oTp, P.tilePrefix.data(), sizeof(int32_t) * (size_t)(P.nseg + 1)));
        J.idx1 = (const int32_t*)(S + P.oIdx1); J.idx2 = (const int32_t*)(S + P.oIdx2);
        J.seg = (const int4*)(S + P.oSeg); J.tilePrefix = (const int32_t*)(S + P.oTp);
        J.topk = (uint32_t*)(S + P.oTop); J.hist = (int*)(S + P.oHist);
        J.match = outDev ? matchv[j] : (int32_t*)(S + P.oMatch);
        J.nseg = P.nseg; J.n1 = s1->n; J.n2 = s2->n; J.mode = mode; J.checkOri = check_ori ? 1 : 0; J.nnratio = nnratio;
        if (mode == ORBX_BOW_TRIANGULATE) {
            const orbx_triangulation* tri = triv[j];
            memcpy(J.F, tri->F12, sizeof(J.F)); J.ex = tri->ex; J.ey = tri->ey; J.onlyStereo = tri->only_stereo;
            memcpy(J.scale2, tri->scale2, sizeof(J.scale2)); memcpy(J.sigma2, tri->sigma2_2, sizeof(J.sigma2));
        }
        ORBX_CUDA(c, cudaMemsetAsync(J.hist, 0, sizeof(int) * (kHisto + 1), c->stream));
        if (P.nOut > 0) ORBX_CUDA(c, cudaMemsetAsync(J.match, 0xff, sizeof(int32_t) * (size_t)P.nOut, c->stream));
    }
    ORBX_CUDA(c, h2d(oJobs, jobs.data(), sizeof(BowJob) * (size_t)npairs));
    stageEnd(c, ST_H2D, 0);
    rc = launchBowJobs(c, (const BowJob*)(S + oJobs), npairs, maxTiles, maxSeg, maxN2);
    if (rc) return rc;
    std::vector<int> nm((size_t)npairs, 0);
    stageBegin(c, ST_D2H);
    for (int j = 0; j < npairs; ++j) {
        if (!outDev && plans[j].nOut > 0)
            ORBX_CUDA(c, cudaMemcpyAsync(matchv[j], jobs[j].match, sizeof(int32_t) * (size_t)plans[j].nOut, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaMemcpyAsync(&nm[j], jobs[j].hist + kHisto, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    }
    stageEnd(c, ST_D2H, 0);
    ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    if (nmatches) for (int j = 0; j < npairs; ++j) nmatches[j] = nm[j];
    return ORBX_OK;
}

extern "C" int orbx_match_bow(orbx_ctx* h, int mode, const orbx_bow_side* s1, const orbx_bow_side* s2,
                              const orbx_triangulation* tri, float nnratio, int check_ori, int flags,
                              int32_t* match, int* nmatches)
{
    if (!h || !s1 || !s2 || !match || mode < 0 || mode > 2 || (mode == ORBX_BOW_TRIANGULATE && !tri)) return ORBX_E_INVALID;
    Ctx* c = &h->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    return matchBowPairs(c, mode, 1, &s1, &s2, &tri, nnratio, check_ori, flags, &match, nmatches);
}

extern "C" int orbx_match_bow_batch(orbx_ctx* h, int mode, int npairs, const orbx_bow_side* const* s1, const orbx_bow_side* const* s2,
                                    const orbx_triangulation* const* tri, float nnratio, int check_ori, int flags,
                                    int32_t* const* match, int* nmatches)
{
    if (!h || npairs < 0 || mode < 0 || mode > 2 || (npairs > 0 && (!s1 || !s2 || !match))) return ORBX_E_INVALID;
    if (npairs == 0) return ORBX_OK;
    if (npairs > 65535) return ORBX_E_CAPACITY;   // grid.y
    Ctx* c = &h->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    return matchBowPairs(c, mode, npairs, s1, s2, tri, nnratio, check_ori, flags, match, nmatches);
}
