// Window (projection) matchers and the frame grid:
//   Frame::AssignFeaturesToGrid / PosInGrid        src/Frame.cc:398-422, 682-698
//   Frame::GetFeaturesInArea                       src/Frame.cc:590-671  (KeyFrame twin src/KeyFrame.cc:794-838)
//   ORBmatcher::SearchByProjection x4 inner loops  src/ORBmatcher.cc:78-191, 431-560, 1741-1910, 1922-2065
//
// The reference handles projected points one after another and lets every accepted match hide its keypoint
// from the following points.  GetFeaturesInArea visits cells with ix outer / iy inner, i.e. in ascending cell id
// c = ix * rows + iy, and each cell's list in ascending feature index -- which is exactly the order of the grid's
// flat index array.  So the position j of a feature in that array is its visit rank for EVERY query, a query's
// candidates are one contiguous j-range per window column, and "first visited wins a tie" is "smaller j wins".
//   P1 gridBuildKernel    one CTA: cell of every keypoint, counts, prefix, stable fill; for the matcher also the
//                         keypoints / descriptors gathered into cell order (so P2 reads them with unit stride)
//   P2 projScanKernel     thread per query: window columns -> j-ranges -> filters + 256-bit distance; keeps the four
//                         smallest keys (dist << 22 | j) and the number of candidates; nothing else is materialised
//   P3 projResolveKernel  one CTA.  Warp 0 replays the greedy order 32 queries at a time: every lane decides for
//                         its query from its four keys and the claim bitmask, lanes whose lists do not contain an
//                         earlier lane's claim commit together, an exact warp-wide rescan runs only when the four keys
//                         cannot decide.  All warps then write assign[], vote the rotation histogram and apply
//                         ComputeThreeMaxima (check_orientation).
#include "orbx_internal.h"
#include <string.h>
#include <algorithm>

namespace orbx {

constexpr int kHistoP = 30;
constexpr int kJBits = 22;                       // position bits of a key: n < 2^22
constexpr uint32_t kJMask = (1u << kJBits) - 1;
constexpr uint32_t kNoKeyP = 0xffffffffu;
constexpr int kGridThreads = 1024;
constexpr int kResolveThreads = 256;

struct GridDev { float minX, minY, invW, invH; int cols, rows; };

// ---------------- grid ----------------
__device__ __forceinline__ int cellOfKey(float x, float y, const GridDev& g)
{   // Frame::PosInGrid, src/Frame.cc:682-698
    const int px = (int)roundf(__fmul_rn(__fsub_rn(x, g.minX), g.invW));
    const int py = (int)roundf(__fmul_rn(__fsub_rn(y, g.minY), g.invH));
    return (px < 0 || px >= g.cols || py < 0 || py >= g.rows) ? -1 : px * g.rows + py;
}

// cinfo[j] = (x, y, uright or -1, octave | occupied << 8) of the feature at position j of the flat grid array.
__global__ void __launch_bounds__(kGridThreads)
gridBuildKernel(const orbx_keypoint* __restrict__ keys, int n, GridDev g, int* __restrict__ cellTmp, int* __restrict__ tmpIdx,
                int* __restrict__ off, int* __restrict__ idx, const uint4* __restrict__ desc, const float* __restrict__ uright,
                const uint8_t* __restrict__ occupied, float4* __restrict__ cinfo, uint4* __restrict__ dsort,
                int* __restrict__ assign, int nAssign)
{
    extern __shared__ int sGrid[];
    __shared__ int warpSum[32];
    __shared__ int carry;
    const int ncell = g.cols * g.rows;
    int* cnt = sGrid;                 // [ncell]    counts, later the fill cursors
    int* start = sGrid + ncell;       // [ncell+1]  exclusive prefix
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    for (int c = tid; c < ncell; c += kGridThreads) cnt[c] = 0;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int i = tid; i < n; i += kGridThreads) {
        const int cell = cellOfKey(keys[i].x, keys[i].y, g);
        cellTmp[i] = cell;
        if (cell >= 0) atomicAdd(&cnt[cell], 1);
    }
    if (assign) for (int i = tid; i < nAssign; i += kGridThreads) assign[i] = -1;
    __syncthreads();
    for (int base = 0; base < ncell; base += kGridThreads) {
        const int i = base + tid;
        const int v = i < ncell ? cnt[i] : 0;
        int x = v;
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
        if (i < ncell) start[i] = excl;
        __syncthreads();
        if (tid == kGridThreads - 1) carry = excl + v;
        __syncthreads();
    }
    if (tid == 0) start[ncell] = carry;
    __syncthreads();
    for (int c = tid; c <= ncell; c += kGridThreads) { off[c] = start[c]; if (c < ncell) cnt[c] = start[c]; }
    __syncthreads();
    for (int i = tid; i < n; i += kGridThreads) {     // unordered fill ...
        const int cell = cellTmp[i];
        if (cell >= 0) tmpIdx[atomicAdd(&cnt[cell], 1)] = i;
    }
    __syncthreads();
    const int placed = start[ncell];
    for (int p = tid; p < placed; p += kGridThreads) { // ... then every entry finds its rank inside its cell (push_back order)
        const int i = tmpIdx[p], cell = cellTmp[i];
        const int b = start[cell], e = start[cell + 1];
        int rank = 0;
        for (int k = b; k < e; ++k) rank += tmpIdx[k] < i;
        const int j = b + rank;
        idx[j] = i;
        if (cinfo) {
            const orbx_keypoint kp = keys[i];
            const int fl = (kp.octave & 0xff) | ((occupied && occupied[i]) ? 0x100 : 0);
            cinfo[j] = make_float4(kp.x, kp.y, uright ? uright[i] : -1.0f, __int_as_float(fl));
            dsort[2 * j] = desc[2 * i]; dsort[2 * j + 1] = desc[2 * i + 1];
        }
    }
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

struct QueryGeom { float x, y, r, ur; int minL, maxL; bool checkLevel; };
__device__ __forceinline__ QueryGeom geomOf(const orbx_proj_query& q)
{
    QueryGeom s;
    s.x = q.u; s.y = q.v; s.r = q.radius; s.ur = q.ur; s.minL = q.min_level; s.maxL = q.max_level;
    s.checkLevel = (s.minL > 0) || (s.maxL >= 0);
    return s;
}

// GetFeaturesInArea's level / distance filter (src/Frame.cc:640-665), the caller's "already matched" skip and the
// uright gate (src/ORBmatcher.cc:141-151, :1846-1852).  Pure filters: their order does not matter.
__device__ __forceinline__ bool candPass(const float4& ci, const QueryGeom& s, int checkUr, int& octave)
{
    const int fl = __float_as_int(ci.w);
    octave = fl & 0xff;
    if (fl & 0x100) return false;
    if (s.checkLevel) {
        if (octave < s.minL) return false;
        if (s.maxL >= 0 && octave > s.maxL) return false;
    }
    const float dx = __fsub_rn(ci.x, s.x), dy = __fsub_rn(ci.y, s.y);
    if (!(fabsf(dx) < s.r && fabsf(dy) < s.r)) return false;
    if (checkUr && ci.z > 0) {
        const float er = fabsf(__fsub_rn(s.ur, ci.z));
        if (er > s.r) return false;
    }
    return true;
}

// P2: four smallest keys of every query, the octaves that go with them, and the number of candidates.
// One warp per query: the window's columns (one contiguous j-range each) are flattened over the lanes.
constexpr int kScanWarps = 4;
// kGate: the reprojection chi-square gate of Fuse (src/ORBmatcher.cc:1254-1285): e2 * invSigma2[level] > 5.99 (7.8 with a
// right coordinate) rejects the candidate; plain float arithmetic in the reference's order, compared as double.
struct ReprojGate { float invSigma2[ORBX_MAX_LEVELS]; };
template <bool kGate>
__global__ void __launch_bounds__(32 * kScanWarps)
projScanKernel(const orbx_proj_query* __restrict__ q, int nq, GridDev g, const int* __restrict__ off,
               const float4* __restrict__ cinfo, const uint4* __restrict__ dsort, int checkUr, uint4* __restrict__ topKey,
               int2* __restrict__ meta, const ReprojGate gate)
{
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * kScanWarps + (threadIdx.x >> 5);
    if (i >= nq) return;
    const QueryGeom s = geomOf(q[i]);
    const Window w = windowOf(g, s.x, s.y, s.r);
    uint32_t e0 = kNoKeyP, e1 = kNoKeyP, e2 = kNoKeyP, e3 = kNoKeyP;
    int cnt = 0;
    if (!w.empty) {
        const uint4* qd = reinterpret_cast<const uint4*>(q[i].desc);
        const uint4 a0 = qd[0], a1 = qd[1];
        const int ncol = w.c1 - w.c0 + 1;
        for (int cb = 0; cb < ncol; cb += 32) {
            int jb = 0, len = 0;
            if (cb + lane < ncol) {
                const int c = (w.c0 + cb + lane) * g.rows;
                jb = off[c + w.r0]; len = off[c + w.r1 + 1] - jb;
            }
            int incl = len;
            #pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
            const int excl = incl - len;
            const int T = __shfl_sync(0xffffffffu, incl, 31);
            for (int tb = 0; tb < T; tb += 32) {
                const int t = tb + lane;
                int k = 0;                      // column of entry t: the last one whose first entry is <= t
                #pragma unroll
                for (int b = 16; b > 0; b >>= 1) { const int ex = __shfl_sync(0xffffffffu, excl, k + b); if (ex <= t) k += b; }
                const int jbk = __shfl_sync(0xffffffffu, jb, k), exk = __shfl_sync(0xffffffffu, excl, k);
                if (t >= T) continue;
                const int j = jbk + (t - exk);
                int oct;
                const float4 ci = cinfo[j];
                if (!candPass(ci, s, checkUr, oct)) continue;
                if (kGate) {
                    const float ex = __fsub_rn(s.x, ci.x), ey = __fsub_rn(s.y, ci.y);
                    float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
                    double lim = 5.99;
                    if (ci.z >= 0.0f) { const float er = __fsub_rn(s.ur, ci.z); e2 = __fadd_rn(e2, __fmul_rn(er, er)); lim = 7.8; }
                    if ((double)__fmul_rn(e2, gate.invSigma2[min(oct, ORBX_MAX_LEVELS - 1)]) > lim) continue;
                }
                const uint32_t key = ((uint32_t)hamming256(a0, a1, dsort[2 * j], dsort[2 * j + 1]) << kJBits) | (uint32_t)j;
                ++cnt;
                if (key < e3) {
                    e3 = key;
                    if (e3 < e2) { const uint32_t u = e2; e2 = e3; e3 = u; }
                    if (e2 < e1) { const uint32_t u = e1; e1 = e2; e2 = u; }
                    if (e1 < e0) { const uint32_t u = e0; e0 = e1; e1 = u; }
                }
            }
        }
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    uint32_t mine = kNoKeyP;                    // lane k < 4 ends up with the k-th smallest key of the warp
    #pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t m = __reduce_min_sync(0xffffffffu, e0);
        if (lane == k) mine = m;
        if (e0 == m && m != kNoKeyP) { e0 = e1; e1 = e2; e2 = e3; e3 = kNoKeyP; }   // keys are unique: one owner pops
    }
    int lev = 0;
    if (lane < 4 && mine != kNoKeyP) lev = (__float_as_int(cinfo[mine & kJMask].w) & 0xff) << (8 * lane);
    lev = __reduce_or_sync(0xffffffffu, lev);
    const uint32_t k1 = __shfl_sync(0xffffffffu, mine, 1), k2 = __shfl_sync(0xffffffffu, mine, 2), k3 = __shfl_sync(0xffffffffu, mine, 3);
    if (lane == 0) {
        topKey[i] = make_uint4(mine, k1, k2, k3);
        meta[i] = make_int2(cnt, lev);
    }
}

__device__ __forceinline__ int rotBinP(float a1, float a2)
{
    float rot = __fsub_rn(a1, a2);
    if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
    int bin = (int)roundf(__fmul_rn(rot, 30 / 360.0f));
    if (bin == kHistoP) bin = 0;
    return bin;
}

// exact best / second-best key of query i among the candidates nobody has claimed yet (all lanes get the result)
__device__ void rescanQuery(const orbx_proj_query* __restrict__ q, int i, const GridDev& g, const int* __restrict__ off,
                            const float4* __restrict__ cinfo, const uint4* __restrict__ dsort, const uint32_t* occ,
                            int checkUr, int lane, uint32_t& k1, uint32_t& k2)
{
    const QueryGeom s = geomOf(q[i]);
    const Window w = windowOf(g, s.x, s.y, s.r);
    k1 = kNoKeyP; k2 = kNoKeyP;
    if (!w.empty) {
        const uint4* qd = reinterpret_cast<const uint4*>(q[i].desc);
        const uint4 a0 = qd[0], a1 = qd[1];
        for (int ix = w.c0; ix <= w.c1; ++ix) {
            const int jb = off[ix * g.rows + w.r0], je = off[ix * g.rows + w.r1 + 1];
            for (int j = jb + lane; j < je; j += 32) {
                int oct;
                if (!candPass(cinfo[j], s, checkUr, oct)) continue;
                if ((occ[j >> 5] >> (j & 31)) & 1u) continue;
                const uint32_t key = ((uint32_t)hamming256(a0, a1, dsort[2 * j], dsort[2 * j + 1]) << kJBits) | (uint32_t)j;
                k2 = min(k2, max(k1, key)); k1 = min(k1, key);
            }
        }
    }
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const uint32_t b1 = __shfl_xor_sync(0xffffffffu, k1, o), b2 = __shfl_xor_sync(0xffffffffu, k2, o);
        const uint32_t lo = min(k1, b1), hi = max(k1, b1);
        k2 = min(hi, min(k2, b2)); k1 = lo;
    }
}

// P3.  claims are bits over j (position in the flat grid array); matchJ[i] = claimed j of query i or -1.
__global__ void __launch_bounds__(kResolveThreads)
projResolveKernel(const orbx_proj_query* __restrict__ q, int nq, GridDev g, const int* __restrict__ off, const int* __restrict__ idx,
                  const float4* __restrict__ cinfo, const uint4* __restrict__ dsort, const uint4* __restrict__ topKey,
                  const int2* __restrict__ meta, const orbx_keypoint* __restrict__ keys, int n, int modeFlags, int thDist,
                  float nnratio, int checkOri, int* __restrict__ matchJ, int* __restrict__ assign, int* __restrict__ nmOut)
{
    extern __shared__ uint32_t occ[];
    __shared__ int sHist[kHistoP];
    __shared__ int keep[3];
    __shared__ int total;
    const int tid = threadIdx.x, lane = tid & 31;
    for (int i = tid; i < (n + 31) / 32; i += kResolveThreads) occ[i] = 0;
    if (tid < kHistoP) sHist[tid] = 0;
    if (tid == 0) total = 0;
    __syncthreads();
    if (tid < 32) {
        const bool ratio = modeFlags & ORBX_PROJ_RATIO_SAME_LEVEL;
        const int checkUr = (modeFlags & ORBX_PROJ_CHECK_URIGHT) ? 1 : 0;
        const uint4 none = make_uint4(kNoKeyP, kNoKeyP, kNoKeyP, kNoKeyP);
        uint4 eN = none; int2 mN = make_int2(0, 0);
        if (lane < nq) { eN = topKey[lane]; mN = meta[lane]; }
        for (int i0 = 0; i0 < nq; i0 += 32) {
            const int i = i0 + lane;
            const uint4 e = eN; const int2 m = mN;
            if (i + 32 < nq) { eN = topKey[i + 32]; mN = meta[i + 32]; } else eN = none;
            const uint32_t j0 = e.x & kJMask, j1 = e.y & kJMask, j2 = e.z & kJMask, j3 = e.w & kJMask;
            const bool complete = m.x <= 4;
            // A transient query (ORBX_PROJ_QUERY_TRANSIENT: its map point has Observations() == 0) takes a keypoint without
            // hiding it: the reference skips a keypoint only if its CURRENT holder is observed (src/ORBmatcher.cc:136-138,
            // :1840-1842), so a later query may take the same keypoint again (assign[] keeps the last one, every accept counts).
            const bool occupies = !(i < nq && (q[i].flags & ORBX_PROJ_QUERY_TRANSIENT));
            // a query whose overall best is beyond the threshold can never be accepted
            uint32_t pending = __ballot_sync(0xffffffffu, i < nq && e.x != kNoKeyP && (int)(e.x >> kJBits) <= thDist);
            int myJ = -1;
            while (pending) {
                const int t0 = __ffs(pending) - 1;
                int claim = -1; bool need = false;
                uint32_t dep0 = kJMask, dep1 = kJMask;   // unclaimed entries (besides the claim) the decision rests on
                if ((pending >> lane) & 1u) {
                    // first and second unclaimed entry of the list
                    uint32_t u0 = kNoKeyP, u1 = kNoKeyP; int lv0 = 0, lv1 = 0;
                    const uint32_t ks[4] = {e.x, e.y, e.z, e.w}, js[4] = {j0, j1, j2, j3};
                    #pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const bool free_ = ks[k] != kNoKeyP && !((occ[js[k] >> 5] >> (js[k] & 31)) & 1u);
                        if (free_) {
                            if (u0 == kNoKeyP) { u0 = ks[k]; lv0 = (m.y >> (8 * k)) & 255; }
                            else if (u1 == kNoKeyP) { u1 = ks[k]; lv1 = (m.y >> (8 * k)) & 255; }
                        }
                    }
                    if (u0 == kNoKeyP) {
                        // everything listed is taken: unlisted candidates are no closer than the last listed one
                        need = !complete && (int)(e.w >> kJBits) <= thDist;
                    } else {
                        const int best = (int)(u0 >> kJBits);
                        if (best <= thDist) {          // else: later claims can only push the best further away
                            bool ok = true;
                            if (ratio) {                             // src/ORBmatcher.cc:177-186
                                if (u1 != kNoKeyP) {
                                    dep1 = u1 & kJMask;
                                    if (lv0 == lv1 && (float)best > __fmul_rn(nnratio, (float)(int)(u1 >> kJBits))) { ok = false; dep0 = u0 & kJMask; }
                                } else if (!complete) {
                                    // the runner-up is unlisted, its distance >= the last listed one: if the ratio rule
                                    // passes against that bound it passes against the real one, else look it up
                                    if ((float)best > __fmul_rn(nnratio, (float)(int)(e.w >> kJBits))) { ok = false; need = true; }
                                }   // no runner-up at all: bestLevel2 == -1 never equals a real level
                            }
                            if (ok) claim = (int)(u0 & kJMask);
                        }
                    }
                }
                const uint32_t needMask = __ballot_sync(0xffffffffu, need);
                if ((needMask >> t0) & 1u) {
                    uint32_t k1, k2;
                    rescanQuery(q, i0 + t0, g, off, cinfo, dsort, occ, checkUr, lane, k1, k2);
                    if (lane == t0 && k1 != kNoKeyP && (int)(k1 >> kJBits) <= thDist) {
                        bool ok = true;
                        if (ratio && k2 != kNoKeyP) {
                            const int lv1 = __float_as_int(cinfo[k1 & kJMask].w) & 0xff, lv2 = __float_as_int(cinfo[k2 & kJMask].w) & 0xff;
                            if (lv1 == lv2 && (float)(int)(k1 >> kJBits) > __fmul_rn(nnratio, (float)(int)(k2 >> kJBits))) ok = false;
                        }
                        if (ok) { myJ = (int)(k1 & kJMask); if (occupies) occ[myJ >> 5] |= 1u << (myJ & 31); }
                    }
                    pending &= ~(1u << t0);
                    __syncwarp();
                    continue;
                }
                // Speculative commit.  Lanes claiming the same entry: the lowest one (the leader) owns it.  Leaders mark
                // their claim; a lane whose decision rests on an entry that somebody marked in this round (whoever it
                // is: conservative) waits for the next round, as does everything behind it.  The lowest pending lane
                // has nobody before it and always commits.
                // (match.any costs a pass per distinct value, so it only runs when an atomicOr has seen a duplicate)
                // A transient lane marks nothing and contends with nobody; its own claim is one more entry its decision
                // rests on (an earlier occupying lane of this round may take it first).
                const bool marks = claim >= 0 && occupies;
                bool leader = marks;
                uint32_t seen = 0;
                if (leader) seen = atomicOr(&occ[claim >> 5], 1u << (claim & 31)) & (1u << (claim & 31));
                if (__ballot_sync(0xffffffffu, seen != 0)) {
                    const uint32_t peers = __match_any_sync(0xffffffffu, marks ? (uint32_t)claim : (0x80000000u | (uint32_t)lane));
                    leader = marks && (peers & ((1u << lane) - 1u)) == 0;
                }
                bool unsafe = need || (marks && !leader);
                __syncwarp();
                if (lane != t0) {
                    if (dep0 != kJMask && ((occ[dep0 >> 5] >> (dep0 & 31)) & 1u)) unsafe = true;
                    if (dep1 != kJMask && ((occ[dep1 >> 5] >> (dep1 & 31)) & 1u)) unsafe = true;
                    if (claim >= 0 && !occupies && ((occ[claim >> 5] >> (claim & 31)) & 1u)) unsafe = true;
                }
                const uint32_t unsafeMask = __ballot_sync(0xffffffffu, unsafe) & pending;
                const uint32_t commit = unsafeMask ? (pending & ((1u << (__ffs(unsafeMask) - 1)) - 1u)) : pending;
                if (leader) {
                    if ((commit >> lane) & 1u) myJ = claim;
                    else atomicAnd(&occ[claim >> 5], ~(1u << (claim & 31)));
                } else if (claim >= 0 && !occupies && ((commit >> lane) & 1u)) myJ = claim;
                pending &= ~commit;
                __syncwarp();
            }
            if (i < nq) matchJ[i] = myJ;
        }
    }
    __syncthreads();
    int cnt = 0;
    for (int i = tid; i < nq; i += kResolveThreads) {
        const int j = matchJ[i];
        if (j < 0) continue;
        const int f = idx[j];
        atomicMax(&assign[f], i); ++cnt;        // assign[] starts at -1; with transient queries the last accept stays (:185, :1858)
        if (checkOri) atomicAdd(&sHist[rotBinP(q[i].angle, keys[f].angle)], 1);
    }
    if (checkOri) {     // ComputeThreeMaxima + removal, src/ORBmatcher.cc:1884-1903, :2076-2120
        __syncthreads();
        if (tid == 0) {
            int max1 = 0, max2 = 0, max3 = 0, i1 = -1, i2 = -1, i3 = -1;
            for (int i = 0; i < kHistoP; ++i) {
                const int s = sHist[i];
                if (s > max1) { max3 = max2; max2 = max1; max1 = s; i3 = i2; i2 = i1; i1 = i; }
                else if (s > max2) { max3 = max2; max2 = s; i3 = i2; i2 = i; }
                else if (s > max3) { max3 = s; i3 = i; }
            }
            if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { i2 = -1; i3 = -1; }
            else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) i3 = -1;
            keep[0] = i1; keep[1] = i2; keep[2] = i3;
        }
        __syncthreads();
        cnt = 0;
        for (int i = tid; i < nq; i += kResolveThreads) {
            const int j = matchJ[i];
            if (j < 0) continue;
            const int f = idx[j];
            const int bin = rotBinP(q[i].angle, keys[f].angle);
            // every accept of a dropped bin clears its keypoint, whoever holds it by now, and un-counts itself (:1893-1903)
            if (bin != keep[0] && bin != keep[1] && bin != keep[2]) assign[f] = -1; else ++cnt;
        }
    }
    atomicAdd(&total, cnt);
    __syncthreads();
    if (tid == 0) *nmOut = total;
}

// ---------------- SearchForInitialization (src/ORBmatcher.cc:580-748) ----------------
// F1 keypoint i1 (level 0 only, :609) searches the square window of half-size windowSize around vbPrevMatched[i1] among
// the level-0 keypoints of F2 (:615).  The reference's loop is sequential with a DISTANCE-dependent mask (a candidate
// whose current match is at least as close is skipped, :643) and later queries steal matches (:667-671):
//   I1 initQueryKernel     every F1 keypoint becomes a window query (radius -1 = no candidates for levels > 0)
//   P2 projScanKernel      (shared with SearchByProjection) the four smallest (distance, visit rank) keys per query
//   I2 initResolveKernel   one warp replays the queries in order: a query is decided from its four keys when two of
//                          them survive the mask (or the list is complete), otherwise the warp rescans its window with
//                          the mask applied; then the rotation histogram (every accepted i1 votes, stolen or not, :692)
__global__ void initQueryKernel(const orbx_keypoint* __restrict__ keys1, const uint4* __restrict__ desc1,
                                const float2* __restrict__ prev, int n1, float window, orbx_proj_query* __restrict__ q)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n1) return;
    const orbx_keypoint k = keys1[i];
    const float2 p = prev[i];
    uint4* o = reinterpret_cast<uint4*>(q + i);
    o[0] = make_uint4(__float_as_uint(p.x), __float_as_uint(p.y), __float_as_uint(k.octave > 0 ? -1.0f : window), 0u);
    o[1] = make_uint4(__float_as_uint(k.angle), 0u, 0u, 0u);     // min_level = max_level = 0: level-0 candidates only
    o[2] = desc1[2 * i]; o[3] = desc1[2 * i + 1];
}

__global__ void __launch_bounds__(32)
initResolveKernel(const orbx_proj_query* __restrict__ q, int n1, GridDev g, const int* __restrict__ off, const int* __restrict__ idx,
                  const float4* __restrict__ cinfo, const uint4* __restrict__ dsort, const uint4* __restrict__ topKey,
                  const int2* __restrict__ meta, const orbx_keypoint* __restrict__ keys2, int n2, float nnratio, int checkOri,
                  int* __restrict__ binOf, int* __restrict__ matches12, float2* __restrict__ prev, int* __restrict__ nmOut)
{
    extern __shared__ uint32_t smemI[];
    int* m21 = reinterpret_cast<int*>(smemI);                      // vnMatches21 over visit ranks j
    uint16_t* md = reinterpret_cast<uint16_t*>(m21 + n2);           // vMatchedDistance (0xffff = INT_MAX)
    __shared__ int sHist[kHistoP];
    __shared__ int keep[3];
    const int lane = threadIdx.x;
    for (int j = lane; j < n2; j += 32) { m21[j] = -1; md[j] = 0xffffu; }
    for (int i = lane; i < n1; i += 32) { matches12[i] = -1; binOf[i] = -1; }
    if (lane < kHistoP) sHist[lane] = 0;
    __syncwarp();
    const uint4 none = make_uint4(kNoKeyP, kNoKeyP, kNoKeyP, kNoKeyP);
    for (int i0 = 0; i0 < n1; i0 += 32) {
        uint4 eL = none; int cntL = 0;
        if (i0 + lane < n1) { eL = topKey[i0 + lane]; cntL = meta[i0 + lane].x; }
        const int kEnd = min(32, n1 - i0);
        for (int k = 0; k < kEnd; ++k) {
            const int i = i0 + k;
            uint32_t e[4];
            e[0] = __shfl_sync(0xffffffffu, eL.x, k); e[1] = __shfl_sync(0xffffffffu, eL.y, k);
            e[2] = __shfl_sync(0xffffffffu, eL.z, k); e[3] = __shfl_sync(0xffffffffu, eL.w, k);
            const int cnt = __shfl_sync(0xffffffffu, cntL, k);
            // the unmasked best of the whole window is a lower bound of the masked best (:660)
            if (e[0] == kNoKeyP || (int)(e[0] >> kJBits) > 50) continue;
            uint32_t b1 = kNoKeyP, b2 = kNoKeyP;
            #pragma unroll
            for (int t = 0; t < 4; ++t) {
                if (e[t] == kNoKeyP) continue;
                const uint32_t d = e[t] >> kJBits, j = e[t] & kJMask;
                if ((uint32_t)md[j] <= d) continue;
                if (b1 == kNoKeyP) b1 = e[t]; else if (b2 == kNoKeyP) b2 = e[t];
            }
            if (b2 == kNoKeyP && cnt > 4) {
                // fewer than two listed keys survive the mask and the window holds more candidates: exact rescan
                const QueryGeom s = geomOf(q[i]);
                const Window w = windowOf(g, s.x, s.y, s.r);
                uint32_t k1 = kNoKeyP, k2 = kNoKeyP;
                if (!w.empty) {
                    const uint4* qd = reinterpret_cast<const uint4*>(q[i].desc);
                    const uint4 a0 = qd[0], a1 = qd[1];
                    for (int ix = w.c0; ix <= w.c1; ++ix) {
                        const int jb = off[ix * g.rows + w.r0], je = off[ix * g.rows + w.r1 + 1];
                        for (int j = jb + lane; j < je; j += 32) {
                            int oct;
                            if (!candPass(cinfo[j], s, 0, oct)) continue;
                            const uint32_t d = (uint32_t)hamming256(a0, a1, dsort[2 * j], dsort[2 * j + 1]);
                            if ((uint32_t)md[j] <= d) continue;
                            const uint32_t key = (d << kJBits) | (uint32_t)j;
                            k2 = min(k2, max(k1, key)); k1 = min(k1, key);
                        }
                    }
                }
                #pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const uint32_t c1 = __shfl_xor_sync(0xffffffffu, k1, o), c2 = __shfl_xor_sync(0xffffffffu, k2, o);
                    const uint32_t lo = min(k1, c1), hi = max(k1, c1);
                    k2 = min(hi, min(k2, c2)); k1 = lo;
                }
                b1 = k1; b2 = k2;
            }
            if (b1 == kNoKeyP) continue;
            const int best = (int)(b1 >> kJBits);
            const float best2 = b2 == kNoKeyP ? 2147483648.0f : (float)(int)(b2 >> kJBits);   // (float)INT_MAX
            if (best <= 50 && (float)best < __fmul_rn(best2, nnratio)) {            // :660, :664
                const int j = (int)(b1 & kJMask);
                if (lane == 0) {
                    const int old = m21[j];
                    if (old >= 0) matches12[old] = -1;                              // :667-671
                    matches12[i] = j; m21[j] = i; md[j] = (uint16_t)best;
                    if (checkOri) {
                        const int bin = rotBinP(q[i].angle, keys2[idx[j]].angle);   // :683-692
                        binOf[i] = bin; sHist[bin] += 1;
                    }
                }
                __syncwarp();
            }
        }
    }
    __syncwarp();
    if (checkOri && lane == 0) {    // ComputeThreeMaxima :2076-2118
        int max1 = 0, max2 = 0, max3 = 0, i1 = -1, i2 = -1, i3 = -1;
        for (int i = 0; i < kHistoP; ++i) {
            const int s = sHist[i];
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; i3 = i2; i2 = i1; i1 = i; }
            else if (s > max2) { max3 = max2; max2 = s; i3 = i2; i2 = i; }
            else if (s > max3) { max3 = s; i3 = i; }
        }
        if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { i2 = -1; i3 = -1; }
        else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) i3 = -1;
        keep[0] = i1; keep[1] = i2; keep[2] = i3;
    }
    __syncwarp();
    int cnt = 0;
    for (int i = lane; i < n1; i += 32) {
        int j = matches12[i];
        if (j >= 0 && checkOri) {
            const int bin = binOf[i];
            if (bin != keep[0] && bin != keep[1] && bin != keep[2]) j = -1;         // :706-722
        }
        int f = -1;
        if (j >= 0) { f = idx[j]; prev[i] = make_float2(keys2[f].x, keys2[f].y); ++cnt; }   // :741-744
        matches12[i] = f;
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if (lane == 0) *nmOut = cnt;
}

// ---------------- independent window searches (Fuse x2, SearchBySim3) ----------------
// No claims between queries: the smallest key of the scan IS the reference's answer (strict <, first visited wins).
__global__ void windowBestKernel(const uint4* __restrict__ topKey, const int* __restrict__ idx, int nq, int thDist,
                                 int* __restrict__ bestIdx, int* __restrict__ bestDist)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    const uint32_t k = topKey[i].x;
    const int d = k == kNoKeyP ? -1 : (int)(k >> kJBits);
    bestDist[i] = d;
    bestIdx[i] = (d >= 0 && d <= thDist) ? idx[k & kJMask] : -1;
}

struct BumpP { size_t off = 0; size_t take(size_t b) { const size_t o = off; off += (b + 255) & ~(size_t)255; return o; } };

static int launchGridBuild(Ctx* c, const orbx_keypoint* dKeys, int n, const GridDev& g, int* dCellTmp, int* dTmpIdx, int* dOff, int* dIdx,
                           const uint4* dDesc, const float* dUright, const uint8_t* dOcc, float4* dCinfo, uint4* dDsort,
                           int* dAssign, int nAssign)
{
    const size_t smem = sizeof(int) * (size_t)(2 * g.cols * g.rows + 1);
    if (smem > 200 * 1024) return ORBX_E_CAPACITY;
    if (smem > 48 * 1024) ORBX_CUDA(c, cudaFuncSetAttribute(gridBuildKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    gridBuildKernel<<<1, kGridThreads, smem, c->stream>>>(dKeys, n, g, dCellTmp, dTmpIdx, dOff, dIdx, dDesc, dUright, dOcc, dCinfo,
                                                          dDsort, dAssign, nAssign);
    c->kernelLaunches += 1; c->stageLaunches[ST_MATCH] += 1;
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
    const size_t n1 = (size_t)std::max(n, 1);
    BumpP B;
    const size_t oK = B.take(sizeof(orbx_keypoint) * n1), oCell = B.take(sizeof(int) * n1), oTmp = B.take(sizeof(int) * n1),
                 oOff = B.take(sizeof(int) * (size_t)(ncell + 1)), oIdx = B.take(sizeof(int) * n1);
    int rc = ensureScratch(c, B.off);
    if (rc) return rc;
    uint8_t* S = (uint8_t*)c->dScratch;
    const orbx_keypoint* dK = keys;
    if (!inDev && n > 0) { ORBX_CUDA(c, cudaMemcpyAsync(S + oK, keys, sizeof(orbx_keypoint) * (size_t)n, cudaMemcpyHostToDevice, c->stream)); dK = (const orbx_keypoint*)(S + oK); }
    const GridDev gd = {g->min_x, g->min_y, g->inv_w, g->inv_h, g->cols, g->rows};
    int* dOff = outDev ? cell_off : (int*)(S + oOff);
    int* dIdx = outDev ? cell_idx : (int*)(S + oIdx);
    stageBegin(c, ST_MATCH);
    rc = launchGridBuild(c, dK, n, gd, (int*)(S + oCell), (int*)(S + oTmp), dOff, dIdx, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0);
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
    if (occBytes > 200 * 1024 || n >= (1 << kJBits)) return ORBX_E_CAPACITY;
    const size_t nq1 = (size_t)std::max(nq, 1);
    // scratch: [inputs, one block so that host inputs arrive with a single copy][work][results]
    BumpP B;
    const size_t oQ = B.take(sizeof(orbx_proj_query) * nq1), oK = B.take(sizeof(orbx_keypoint) * (size_t)n),
                 oD = B.take((size_t)32 * n), oU = B.take(sizeof(float) * (size_t)n), oOcc = B.take((size_t)n);
    const size_t inBytes = B.off;
    const size_t oCell = B.take(sizeof(int) * (size_t)n), oTmp = B.take(sizeof(int) * (size_t)n),
                 oOff = B.take(sizeof(int) * (size_t)(ncell + 1)), oIdx = B.take(sizeof(int) * (size_t)n),
                 oInfo = B.take(sizeof(float4) * (size_t)n), oSort = B.take((size_t)32 * n),
                 oTop = B.take(sizeof(uint4) * nq1), oMeta = B.take(sizeof(int2) * nq1), oMJ = B.take(sizeof(int) * nq1),
                 oOut = B.take(sizeof(int) * (size_t)(n + 1));
    int rc = ensureScratch(c, B.off);
    if (rc) return rc;
    const size_t outBytes = sizeof(int) * (size_t)(n + 1);
    rc = ensureMatchPinned(c, inBytes + outBytes);
    if (rc) return rc;
    uint8_t* S = (uint8_t*)c->dScratch;
    const orbx_proj_query* dQ = q; const orbx_keypoint* dK = keys; const uint4* dD = (const uint4*)desc;
    const float* dU = uright; const uint8_t* dOcc = occupied;
    stageBegin(c, ST_H2D);
    if (!inDev) {
        uint8_t* H = c->hMatch;
        if (nq) memcpy(H + oQ, q, sizeof(orbx_proj_query) * (size_t)nq);
        memcpy(H + oK, keys, sizeof(orbx_keypoint) * (size_t)n);
        memcpy(H + oD, desc, (size_t)32 * n);
        if (uright) memcpy(H + oU, uright, sizeof(float) * (size_t)n);
        if (occupied) memcpy(H + oOcc, occupied, (size_t)n);
        ORBX_CUDA(c, cudaMemcpyAsync(S, H, inBytes, cudaMemcpyHostToDevice, c->stream));
        dQ = (const orbx_proj_query*)(S + oQ); dK = (const orbx_keypoint*)(S + oK); dD = (const uint4*)(S + oD);
        if (uright) dU = (const float*)(S + oU);
        if (occupied) dOcc = S + oOcc;
    }
    stageEnd(c, ST_H2D, 0);
    const GridDev gd = {g->min_x, g->min_y, g->inv_w, g->inv_h, g->cols, g->rows};
    int* dOut = (int*)(S + oOut);                    // assign[n] then the match count
    int* dAssign = outDev ? assign : dOut;
    stageBegin(c, ST_MATCH);
    rc = launchGridBuild(c, dK, n, gd, (int*)(S + oCell), (int*)(S + oTmp), (int*)(S + oOff), (int*)(S + oIdx), dD, dU, dOcc,
                         (float4*)(S + oInfo), (uint4*)(S + oSort), dAssign, n);
    if (rc) return rc;
    int launches = 0;
    if (nq > 0) {
        projScanKernel<false><<<(nq + kScanWarps - 1) / kScanWarps, 32 * kScanWarps, 0, c->stream>>>(dQ, nq, gd, (int*)(S + oOff), (float4*)(S + oInfo), (uint4*)(S + oSort),
                                                           (mode_flags & ORBX_PROJ_CHECK_URIGHT) ? 1 : 0, (uint4*)(S + oTop),
                                                           (int2*)(S + oMeta), ReprojGate());
        if (occBytes > 40 * 1024)
            ORBX_CUDA(c, cudaFuncSetAttribute(projResolveKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)occBytes));
        projResolveKernel<<<1, kResolveThreads, occBytes, c->stream>>>(dQ, nq, gd, (int*)(S + oOff), (int*)(S + oIdx),
                                                                       (float4*)(S + oInfo), (uint4*)(S + oSort), (uint4*)(S + oTop),
                                                                       (int2*)(S + oMeta), dK, n, mode_flags, th_dist, nnratio,
                                                                       check_ori ? 1 : 0, (int*)(S + oMJ), dAssign, dOut + n);
        launches += 2;
    } else {
        ORBX_CUDA(c, cudaMemsetAsync(dOut + n, 0, sizeof(int), c->stream));
    }
    stageEnd(c, ST_MATCH, launches);
    ORBX_CUDA(c, cudaGetLastError());
    int nm = 0;
    if (!outDev) {
        int* HO = (int*)(c->hMatch + inBytes);
        ORBX_CUDA(c, cudaMemcpyAsync(HO, dOut, outBytes, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
        memcpy(assign, HO, sizeof(int) * (size_t)n);
        nm = HO[n];
    } else {
        int* HO = (int*)(c->hMatch + inBytes);
        ORBX_CUDA(c, cudaMemcpyAsync(HO, dOut + n, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
        nm = HO[0];
    }
    if (nmatches) *nmatches = nm;
    return ORBX_OK;
}

extern "C" int orbx_search_for_initialization(orbx_ctx* h, const orbx_keypoint* keys1, const uint8_t* desc1, int n1, float* prev_xy,
                                              const orbx_keypoint* keys2, const uint8_t* desc2, int n2, const orbx_grid* g,
                                              int window, float nnratio, int check_ori, int flags, int32_t* matches12,
                                              int* nmatches)
{
    if (!h || !g || n1 < 0 || n2 < 0 || g->cols < 1 || g->rows < 1 || (n1 > 0 && (!keys1 || !desc1 || !prev_xy || !matches12)) ||
        (n2 > 0 && (!keys2 || !desc2)))
        return ORBX_E_INVALID;
    Ctx* c = &h->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    if (nmatches) *nmatches = 0;
    if (n1 == 0) return ORBX_OK;
    const bool inDev = flags & ORBX_IN_DEVICE, outDev = flags & ORBX_OUT_DEVICE;
    if (n2 == 0) {
        if (outDev) ORBX_CUDA(c, cudaMemsetAsync(matches12, 0xff, sizeof(int) * (size_t)n1, c->stream));
        else for (int i = 0; i < n1; ++i) matches12[i] = -1;
        return ORBX_OK;
    }
    const int ncell = g->cols * g->rows;
    const size_t resolveSmem = (size_t)n2 * 6 + 16;
    if (resolveSmem > 200 * 1024 || n2 >= (1 << kJBits)) return ORBX_E_CAPACITY;
    BumpP B;
    const size_t oK1 = B.take(sizeof(orbx_keypoint) * (size_t)n1), oD1 = B.take((size_t)32 * n1), oP = B.take(sizeof(float2) * (size_t)n1),
                 oK2 = B.take(sizeof(orbx_keypoint) * (size_t)n2), oD2 = B.take((size_t)32 * n2);
    const size_t inBytes = B.off;
    const size_t oQ = B.take(sizeof(orbx_proj_query) * (size_t)n1), oCell = B.take(sizeof(int) * (size_t)n2),
                 oTmp = B.take(sizeof(int) * (size_t)n2), oOff = B.take(sizeof(int) * (size_t)(ncell + 1)),
                 oIdx = B.take(sizeof(int) * (size_t)n2), oInfo = B.take(sizeof(float4) * (size_t)n2), oSort = B.take((size_t)32 * n2),
                 oTop = B.take(sizeof(uint4) * (size_t)n1), oMeta = B.take(sizeof(int2) * (size_t)n1), oBin = B.take(sizeof(int) * (size_t)n1),
                 oOut = B.take(sizeof(int) * (size_t)(n1 + 1));
    int rc = ensureScratch(c, B.off);
    if (rc) return rc;
    const size_t outBytes = sizeof(int) * (size_t)(n1 + 1);
    rc = ensureMatchPinned(c, inBytes + outBytes + sizeof(float2) * (size_t)n1);
    if (rc) return rc;
    uint8_t* S = (uint8_t*)c->dScratch;
    const orbx_keypoint *dK1 = keys1, *dK2 = keys2;
    const uint4 *dD1 = (const uint4*)desc1, *dD2 = (const uint4*)desc2;
    float2* dPrev = (float2*)prev_xy;
    stageBegin(c, ST_H2D);
    if (!inDev) {
        uint8_t* H = c->hMatch;
        memcpy(H + oK1, keys1, sizeof(orbx_keypoint) * (size_t)n1); memcpy(H + oD1, desc1, (size_t)32 * n1);
        memcpy(H + oP, prev_xy, sizeof(float2) * (size_t)n1);
        memcpy(H + oK2, keys2, sizeof(orbx_keypoint) * (size_t)n2); memcpy(H + oD2, desc2, (size_t)32 * n2);
        ORBX_CUDA(c, cudaMemcpyAsync(S, H, inBytes, cudaMemcpyHostToDevice, c->stream));
        dK1 = (const orbx_keypoint*)(S + oK1); dD1 = (const uint4*)(S + oD1); dPrev = (float2*)(S + oP);
        dK2 = (const orbx_keypoint*)(S + oK2); dD2 = (const uint4*)(S + oD2);
    }
    stageEnd(c, ST_H2D, 0);
    const GridDev gd = {g->min_x, g->min_y, g->inv_w, g->inv_h, g->cols, g->rows};
    int* dOut = (int*)(S + oOut);
    int* dM12 = outDev ? matches12 : dOut;
    stageBegin(c, ST_MATCH);
    rc = launchGridBuild(c, dK2, n2, gd, (int*)(S + oCell), (int*)(S + oTmp), (int*)(S + oOff), (int*)(S + oIdx), dD2, nullptr, nullptr,
                         (float4*)(S + oInfo), (uint4*)(S + oSort), nullptr, 0);
    if (rc) return rc;
    initQueryKernel<<<(n1 + 255) / 256, 256, 0, c->stream>>>(dK1, dD1, dPrev, n1, (float)window, (orbx_proj_query*)(S + oQ));
    projScanKernel<false><<<(n1 + kScanWarps - 1) / kScanWarps, 32 * kScanWarps, 0, c->stream>>>((orbx_proj_query*)(S + oQ), n1, gd, (int*)(S + oOff),
                                                       (float4*)(S + oInfo), (uint4*)(S + oSort), 0, (uint4*)(S + oTop), (int2*)(S + oMeta), ReprojGate());
    if (resolveSmem > 40 * 1024)
        ORBX_CUDA(c, cudaFuncSetAttribute(initResolveKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)resolveSmem));
    initResolveKernel<<<1, 32, resolveSmem, c->stream>>>((orbx_proj_query*)(S + oQ), n1, gd, (int*)(S + oOff), (int*)(S + oIdx),
                                                         (float4*)(S + oInfo), (uint4*)(S + oSort), (uint4*)(S + oTop), (int2*)(S + oMeta),
                                                         dK2, n2, nnratio, check_ori ? 1 : 0, (int*)(S + oBin), dM12, dPrev, dOut + n1);
    stageEnd(c, ST_MATCH, 3);
    ORBX_CUDA(c, cudaGetLastError());
    int* HO = (int*)(c->hMatch + inBytes);
    if (!outDev) {
        ORBX_CUDA(c, cudaMemcpyAsync(HO, dOut, outBytes, cudaMemcpyDeviceToHost, c->stream));
        float2* HP = (float2*)(c->hMatch + inBytes + outBytes);
        if (!inDev) ORBX_CUDA(c, cudaMemcpyAsync(HP, dPrev, sizeof(float2) * (size_t)n1, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
        memcpy(matches12, HO, sizeof(int) * (size_t)n1);
        if (!inDev) memcpy(prev_xy, HP, sizeof(float2) * (size_t)n1);
        if (nmatches) *nmatches = HO[n1];
    } else {
        ORBX_CUDA(c, cudaMemcpyAsync(HO, dOut + n1, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
        if (nmatches) *nmatches = HO[0];
    }
    return ORBX_OK;
}

extern "C" int orbx_window_best(orbx_ctx* h, const orbx_proj_query* q, int nq, const orbx_keypoint* keys, const uint8_t* desc,
                                const float* uright, int n, const orbx_grid* g, const float* inv_sigma2, int th_dist, int flags,
                                int32_t* best_idx, int32_t* best_dist)
{
    if (!h || !g || nq < 0 || n < 0 || g->cols < 1 || g->rows < 1 || (n > 0 && (!keys || !desc)) || (nq > 0 && (!q || !best_idx || !best_dist)))
        return ORBX_E_INVALID;
    Ctx* c = &h->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    if (nq == 0) return ORBX_OK;
    const bool inDev = flags & ORBX_IN_DEVICE, outDev = flags & ORBX_OUT_DEVICE;
    if (n == 0) {
        if (outDev) { ORBX_CUDA(c, cudaMemsetAsync(best_idx, 0xff, sizeof(int) * (size_t)nq, c->stream)); ORBX_CUDA(c, cudaMemsetAsync(best_dist, 0xff, sizeof(int) * (size_t)nq, c->stream)); }
        else for (int i = 0; i < nq; ++i) { best_idx[i] = -1; best_dist[i] = -1; }
        return ORBX_OK;
    }
    if (n >= (1 << kJBits)) return ORBX_E_CAPACITY;
    const int ncell = g->cols * g->rows;
    BumpP B;
    const size_t oQ = B.take(sizeof(orbx_proj_query) * (size_t)nq), oK = B.take(sizeof(orbx_keypoint) * (size_t)n),
                 oD = B.take((size_t)32 * n), oU = B.take(sizeof(float) * (size_t)n);
    const size_t inBytes = B.off;
    const size_t oCell = B.take(sizeof(int) * (size_t)n), oTmp = B.take(sizeof(int) * (size_t)n),
                 oOff = B.take(sizeof(int) * (size_t)(ncell + 1)), oIdx = B.take(sizeof(int) * (size_t)n),
                 oInfo = B.take(sizeof(float4) * (size_t)n), oSort = B.take((size_t)32 * n),
                 oTop = B.take(sizeof(uint4) * (size_t)nq), oMeta = B.take(sizeof(int2) * (size_t)nq),
                 oOut = B.take(sizeof(int) * 2 * (size_t)nq);
    int rc = ensureScratch(c, B.off);
    if (rc) return rc;
    const size_t outBytes = sizeof(int) * 2 * (size_t)nq;
    rc = ensureMatchPinned(c, inBytes + outBytes);
    if (rc) return rc;
    uint8_t* S = (uint8_t*)c->dScratch;
    const orbx_proj_query* dQ = q; const orbx_keypoint* dK = keys; const uint4* dD = (const uint4*)desc; const float* dU = uright;
    stageBegin(c, ST_H2D);
    if (!inDev) {
        uint8_t* H = c->hMatch;
        memcpy(H + oQ, q, sizeof(orbx_proj_query) * (size_t)nq);
        memcpy(H + oK, keys, sizeof(orbx_keypoint) * (size_t)n);
        memcpy(H + oD, desc, (size_t)32 * n);
        if (uright) memcpy(H + oU, uright, sizeof(float) * (size_t)n);
        ORBX_CUDA(c, cudaMemcpyAsync(S, H, inBytes, cudaMemcpyHostToDevice, c->stream));
        dQ = (const orbx_proj_query*)(S + oQ); dK = (const orbx_keypoint*)(S + oK); dD = (const uint4*)(S + oD);
        if (uright) dU = (const float*)(S + oU);
    }
    stageEnd(c, ST_H2D, 0);
    const GridDev gd = {g->min_x, g->min_y, g->inv_w, g->inv_h, g->cols, g->rows};
    ReprojGate gate;
    for (int l = 0; l < ORBX_MAX_LEVELS; ++l) gate.invSigma2[l] = inv_sigma2 ? inv_sigma2[l] : 0.0f;   // inv_sigma2 is a HOST array
    int* dOut = (int*)(S + oOut);
    int* dBI = outDev ? best_idx : dOut;
    int* dBD = outDev ? best_dist : dOut + nq;
    stageBegin(c, ST_MATCH);
    rc = launchGridBuild(c, dK, n, gd, (int*)(S + oCell), (int*)(S + oTmp), (int*)(S + oOff), (int*)(S + oIdx), dD, dU, nullptr,
                         (float4*)(S + oInfo), (uint4*)(S + oSort), nullptr, 0);
    if (rc) return rc;
    const int blocks = (nq + kScanWarps - 1) / kScanWarps;
    if (inv_sigma2)
        projScanKernel<true><<<blocks, 32 * kScanWarps, 0, c->stream>>>(dQ, nq, gd, (int*)(S + oOff), (float4*)(S + oInfo), (uint4*)(S + oSort), 0,
                                                                        (uint4*)(S + oTop), (int2*)(S + oMeta), gate);
    else
        projScanKernel<false><<<blocks, 32 * kScanWarps, 0, c->stream>>>(dQ, nq, gd, (int*)(S + oOff), (float4*)(S + oInfo), (uint4*)(S + oSort), 0,
                                                                         (uint4*)(S + oTop), (int2*)(S + oMeta), gate);
    windowBestKernel<<<(nq + 255) / 256, 256, 0, c->stream>>>((uint4*)(S + oTop), (int*)(S + oIdx), nq, th_dist, dBI, dBD);
    stageEnd(c, ST_MATCH, 2);
    ORBX_CUDA(c, cudaGetLastError());
    if (!outDev) {
        int* HO = (int*)(c->hMatch + inBytes);
        ORBX_CUDA(c, cudaMemcpyAsync(HO, dOut, outBytes, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
        memcpy(best_idx, HO, sizeof(int) * (size_t)nq);
        memcpy(best_dist, HO + nq, sizeof(int) * (size_t)nq);
    }
    return ORBX_OK;
}
