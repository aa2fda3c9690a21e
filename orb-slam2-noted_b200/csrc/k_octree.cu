// K3 -- DistributeOctTree (src/ORBextractor.cc:696-1042, ExtractorNode::DivideNode :610-682) as a
// deterministic, level-parallel quadtree: one CTA per (level, frame), node list in shared memory.
//
// The reference keeps a std::list of nodes, each holding a vector of keypoints, and splits nodes in
// passes.  What it computes depends only on (a) how many keypoints fall in each child box and (b) the
// ORDER of the node list (children are push_front'ed in order n1..n4, parents erased).  So here:
//   * a keypoint record only carries the list position of its node (24 bits) - no per-node vectors;
//   * a pass = count children (shared atomics) -> choose which nodes split, in which order ->
//     prefix sums give every child / surviving node its position in the new list -> relabel keypoints;
//   * "breadth" passes (:808-891) split every expandable node in list order;
//   * "fine" passes (:917-999) split in order (size desc, most-recently-created first) and stop at the
//     first split that reaches N nodes.  The reference sorts pair<size, node*> (:935), i.e. ties are
//     broken by heap address; canonical rule here (and in the oracle): later-created node first, which
//     is exactly "smaller list position first" because every candidate was push_front'ed in the
//     previous pass.  The early `break` (:991) becomes a prefix-sum search;
//   * final pick per node (:1013-1038): highest response, first in detection order on ties ->
//     atomicMax of (score, ~orderKey); output in list order.
#include "orbx_internal.h"

namespace orbx {

constexpr int kOctThreadsMax = 512;   // CTA size is chosen per launch (128 for VGA-class frames, more when a level holds many candidates)

// inclusive block scan of data[0..n) in place; returns the total. All threads must call.
__device__ int blockInclusiveScan(int* data, int n, int* warpSums)
{
    const int T = blockDim.x;
    const int per = (n + T - 1) / T;
    const int b = threadIdx.x * per, e = min(b + per, n);
    int sum = 0;
    for (int i = b; i < e; ++i) { sum += data[i]; data[i] = sum; }
    // scan of thread sums
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int v = sum;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += t; }
    if (lane == 31) warpSums[wid] = v;
    __syncthreads();
    if (wid == 0) {
        int w = lane < T / 32 ? warpSums[lane] : 0;
        #pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += t; }
        if (lane < T / 32) warpSums[lane] = w;
    }
    __syncthreads();
    const int offset = (wid ? warpSums[wid - 1] : 0) + v - sum; // exclusive prefix of this thread's chunk
    for (int i = b; i < e; ++i) data[i] += offset;
    const int total = warpSums[T / 32 - 1];
    __syncthreads();
    return total;
}

struct Box { short ulx, uly, brx, bry; };

__device__ __forceinline__ void decodeKey(const LevelGeom& g, uint32_t key, int& x, int& y)
{
    const int cell = key >> 12, ly = (key >> 6) & 63, lx = key & 63;
    const int ci = cell / g.nCols, cj = cell - ci * g.nCols;
    x = 3 + cj * g.wCell + lx;   // window-relative, as pushed at :1146-1147
    y = 3 + ci * g.hCell + ly;
}

__device__ __forceinline__ int quadrant(const Box& b, int x, int y)
{
    const int mx = b.ulx + ((b.brx - b.ulx + 1) >> 1); // UL.x + ceil((UR.x-UL.x)/2)  :616
    const int my = b.uly + ((b.bry - b.uly + 1) >> 1); // :617
    return (x < mx ? 0 : 1) + (y < my ? 0 : 2);        // 0..3 = n1,n2,n3,n4  :659-669
}

__global__ void __launch_bounds__(kOctThreadsMax)
octreeKernel(const __grid_constant__ GeomSet G, unsigned long long* __restrict__ candAll, const int* __restrict__ candCount,
             orbx_keypoint* __restrict__ selAll, int* __restrict__ selCount, int M, int* __restrict__ err)
{
    extern __shared__ __align__(16) uint8_t smemRaw[];
    __shared__ int warpSums[kOctThreadsMax / 32];
    __shared__ int sL, sNSplit, sC, sFinish, sMode;

    const int level = blockIdx.x, frame = blockIdx.y + G.frame0;
    const LevelGeom& g = G.lv[level];
    const int n = candCount[frame * G.nlevels + level];
    const int N = g.quota;
    if (n == 0) { if (threadIdx.x == 0) selCount[frame * G.nlevels + level] = 0; return; }
    unsigned long long* rec = candAll + (size_t)frame * G.candPerFrame + g.candBase;
    uint32_t* recHi = reinterpret_cast<uint32_t*>(rec) + 1; // little endian: hi word = score<<24 | pos
    const uint32_t* recLo = reinterpret_cast<const uint32_t*>(rec);

    // shared arrays (M entries each)
    Box* boxA = reinterpret_cast<Box*>(smemRaw);
    Box* boxB = boxA + M;
    int* cntA = reinterpret_cast<int*>(boxB + M);
    int* cntB = cntA + M;
    int* childCnt = cntB + M;          // 4*M
    int* ord = childCnt + 4 * M;       // processing order -> position          (M)
    int* scanA = ord + M;              // scratch scans                         (M)
    int* newPos = scanA + M;           // position of a non-split node in the new list, or -1-rank if split (M)
    unsigned short* childPos = reinterpret_cast<unsigned short*>(newPos + M); // 4*M
    unsigned long long* best = reinterpret_cast<unsigned long long*>(childCnt); // aliases childCnt (final phase)

    const int T = blockDim.x, tid = threadIdx.x;
    const int winH = (g.h - kEdge + 3) - kWinOff; // maxY - minY

    // ---- roots (:719-749) ----
    for (int i = tid; i < M; i += T) { cntA[i] = 0; }
    __syncthreads();
    const int nIni = g.nIni;
    const float hX = g.hX;
    for (int i = tid; i < nIni; i += T) {
        Box b; b.ulx = (short)(int)__fmul_rn(hX, (float)i); b.brx = (short)(int)__fmul_rn(hX, (float)(i + 1));
        b.uly = 0; b.bry = (short)winH;
        boxA[i] = b;
    }
    for (int k = tid; k < n; k += T) {
        int x, y; decodeKey(g, recLo[2 * k], x, y);
        const int r = (int)__fdiv_rn((float)x, hX);
        atomicAdd(&cntA[r], 1);
        recHi[2 * k] = (recHi[2 * k] & 0xff000000u) | (uint32_t)r;
    }
    __syncthreads();
    // erase empty roots (:756-774): serial compaction, nIni is tiny
    if (tid == 0) {
        int L = 0;
        for (int i = 0; i < nIni; ++i) {
            if (cntA[i] > 0) { newPos[i] = L; boxB[L] = boxA[i]; cntB[L] = cntA[i]; ++L; } else newPos[i] = -1;
        }
        sL = L; sFinish = 0; sMode = 0;
    }
    __syncthreads();
    for (int k = tid; k < n; k += T) {
        const uint32_t h = recHi[2 * k];
        recHi[2 * k] = (h & 0xff000000u) | (uint32_t)newPos[h & 0xffffffu];
    }
    __syncthreads();
    Box* box = boxB; int* cnt = cntB; Box* boxN = boxA; int* cntN = cntA;

    // ---- passes ----
    for (int guard = 0; guard < 64; ++guard) {
        const int L = sL;
        const int mode = sMode; // 0 = breadth pass, 1 = fine pass
        // (a) child counts of every expandable node
        for (int i = tid; i < 4 * L; i += T) childCnt[i] = 0;
        __syncthreads();
        for (int k = tid; k < n; k += T) {
            const int p = recHi[2 * k] & 0xffffffu;
            if (cnt[p] > 1) {
                int x, y; decodeKey(g, recLo[2 * k], x, y);
                atomicAdd(&childCnt[4 * p + quadrant(box[p], x, y)], 1);
            }
        }
        // (b) processing order of expandable nodes
        for (int p = tid; p < L; p += T) scanA[p] = cnt[p] > 1 ? 1 : 0;
        __syncthreads();
        const int E = blockInclusiveScan(scanA, L, warpSums);
        if (E == 0) break; // nothing to split: lNodes.size()==prevSize (:897 / :996)
        if (mode == 0) {
            for (int p = tid; p < L; p += T) if (cnt[p] > 1) ord[scanA[p] - 1] = p;
        } else {
            // rank among expandable: size desc, then list position asc (canonical tie-break).  The E expandable nodes are
            // packed in list order with key = count : (4095 - position) (M <= 3400, counts < 2^20), rank = larger keys.
            uint32_t* key = reinterpret_cast<uint32_t*>(newPos);                 // free until step (d)
            for (int p = tid; p < L; p += T) if (cnt[p] > 1) key[scanA[p] - 1] = ((uint32_t)cnt[p] << 12) | (uint32_t)(4095 - p);
            __syncthreads();
            for (int e = tid; e < E; e += T) {
                const uint32_t k = key[e];
                int r = 0;
                for (int q = 0; q < E; ++q) r += key[q] > k;
                ord[r] = 4095 - (int)(k & 4095u);
            }
        }
        __syncthreads();
        // (c) children per processed node, prefix of list growth
        for (int r = tid; r < E; r += T) {
            const int* cc = childCnt + 4 * ord[r];
            scanA[r] = (cc[0] > 0) + (cc[1] > 0) + (cc[2] > 0) + (cc[3] > 0);
        }
        __syncthreads();
        const int Call = blockInclusiveScan(scanA, E, warpSums); // scanA[r] = children of ord[0..r]
        if (tid == 0) { sNSplit = E; sC = Call; }
        __syncthreads();
        if (mode == 1) {
            // first r with L + (scanA[r] - (r+1)) >= N  (:991 break)
            for (int r = tid; r < E; r += T) {
                const bool hit = L + scanA[r] - (r + 1) >= N;
                const bool prevHit = r > 0 && (L + scanA[r - 1] - r >= N);
                if (hit && !prevHit) { sNSplit = r + 1; sC = scanA[r]; }
            }
            __syncthreads();
        }
        const int nSplit = sNSplit, C = sC;
        // (e) mark split nodes, (f) positions of the others
        for (int p = tid; p < L; p += T) newPos[p] = 0;
        __syncthreads();
        for (int r = tid; r < nSplit; r += T) newPos[ord[r]] = -1 - r;
        __syncthreads();
        // offsets of child blocks: block of ord[r] starts at C - scanA[r]
        for (int r = tid; r < nSplit; r += T) {
            const int p = ord[r];
            const int* cc = childCnt + 4 * p;
            int pos = C - scanA[r];
            const Box b = box[p];
            const short mx = (short)(b.ulx + ((b.brx - b.ulx + 1) >> 1)), my = (short)(b.uly + ((b.bry - b.uly + 1) >> 1));
            #pragma unroll
            for (int q = 3; q >= 0; --q) { // front of the block is n4, then n3, n2, n1
                if (cc[q] > 0) {
                    Box c;
                    c.ulx = (q & 1) ? mx : b.ulx; c.brx = (q & 1) ? b.brx : mx;
                    c.uly = (q & 2) ? my : b.uly; c.bry = (q & 2) ? b.bry : my;
                    boxN[pos] = c; cntN[pos] = cc[q];
                    childPos[4 * p + q] = (unsigned short)pos;
                    ++pos;
                }
            }
        }
        __syncthreads();
        // stationary nodes keep their relative order behind the new children
        for (int p = tid; p < L; p += T) scanA[p] = newPos[p] >= 0 ? 1 : 0;
        __syncthreads();
        blockInclusiveScan(scanA, L, warpSums);
        for (int p = tid; p < L; p += T) {
            if (newPos[p] >= 0) { const int np = C + scanA[p] - 1; newPos[p] = np; boxN[np] = box[p]; cntN[np] = cnt[p]; }
        }
        __syncthreads();
        // (h) relabel keypoints
        for (int k = tid; k < n; k += T) {
            const uint32_t h = recHi[2 * k];
            const int p = h & 0xffffffu;
            int np = newPos[p];
            if (np < 0) { int x, y; decodeKey(g, recLo[2 * k], x, y); np = childPos[4 * p + quadrant(box[p], x, y)]; }
            recHi[2 * k] = (h & 0xff000000u) | (uint32_t)np;
        }
        // nToExpand (:833...) = new children with more than one keypoint
        const int newL = C + (L - nSplit);
        for (int p = tid; p < C; p += T) scanA[p] = cntN[p] > 1 ? 1 : 0;
        __syncthreads();
        const int toExpand = blockInclusiveScan(scanA, C, warpSums);
        if (tid == 0) {
            sL = newL;
            if (newL >= N || newL == L) sFinish = 1;                       // :897-903 / :996
            else if (mode == 0 && newL + 3 * toExpand > N) sMode = 1;      // :917
        }
        { Box* tb = box; box = boxN; boxN = tb; int* tc = cnt; cnt = cntN; cntN = tc; }
        __syncthreads();
        if (sFinish) break;
        if (guard == 63 && tid == 0) atomicExch(err, 3);
    }
    __syncthreads();

    // ---- best keypoint per node, list order output (:1013-1038, :1171-1186) ----
    const int L = sL;
    for (int p = tid; p < L; p += T) best[p] = 0ull;
    __syncthreads();
    for (int k = tid; k < n; k += T) {
        const uint32_t h = recHi[2 * k];
        const unsigned long long v = ((unsigned long long)(h >> 24) << 32) | (uint32_t)(~recLo[2 * k]);
        atomicMax(&best[h & 0xffffffu], v);
    }
    __syncthreads();
    orbx_keypoint* sel = selAll + (size_t)frame * G.selPerFrame + g.outBase;
    for (int p = tid; p < L; p += T) {
        const unsigned long long v = best[p];
        int x, y; decodeKey(g, ~(uint32_t)v, x, y);
        orbx_keypoint kp;
        kp.x = (float)(x + kWinOff); kp.y = (float)(y + kWinOff);
        kp.size = g.patchSize; kp.angle = -1.f; kp.response = (float)(uint32_t)(v >> 32);
        kp.octave = level; kp.class_id = -1;
        if (p < g.outCap) sel[p] = kp; else atomicExch(err, 4);
    }
    if (tid == 0) selCount[frame * G.nlevels + level] = min(L, g.outCap);
}

static size_t octreeSmem(int M) { return (size_t)M * (8 + 8 + 4 + 4 + 16 + 4 + 4 + 4 + 8) + 64; }

int launchOctree(Ctx* c, int frame0, int batch)
{
    GeomSet g = c->geom;
    g.frame0 = frame0;
    int M = 0;
    for (int l = 0; l < g.nlevels; ++l) M = max(M, g.lv[l].outCap);
    M = (M + 7) & ~7;
    const size_t smem = octreeSmem(M);
    if (smem > 200 * 1024) return ORBX_E_UNSUPPORTED;
    if (smem > 48 * 1024 && smem > c->smemCfg[SMEM_OCTREE]) {
        ORBX_CUDA(c, cudaFuncSetAttribute(octreeKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        c->smemCfg[SMEM_OCTREE] = smem;
    }
    stageBegin(c, ST_OCTREE);
    // the rounds are barrier-bound: small CTAs for VGA-class levels, larger ones when level 0 can hold tens of thousands of candidates
    static const int envThreads = getenv("ORBX_OCT_THREADS") ? atoi(getenv("ORBX_OCT_THREADS")) : 0;
    const int area = g.lv[0].w * g.lv[0].h;
    int threads = area > 1000 * 1000 ? 512 : area > 400 * 1000 ? 256 : 128;   // measured: VGA 128, KITTI 1241x376 256, 1080p 512
    if (batch * g.nlevels < 2 * 148) threads = 512;                        // a frame or two: latency of the rounds, not throughput
    if (envThreads >= 32 && envThreads <= kOctThreadsMax && envThreads % 32 == 0) threads = envThreads;
    dim3 grid(g.nlevels, batch), block(threads);
    octreeKernel<<<grid, block, smem, c->stream>>>(g, c->dCand, c->dCandCount, c->dSel, c->dSelCount, M, c->dErr);
    stageEnd(c, ST_OCTREE, 1);
    ORBX_CUDA(c, cudaGetLastError());
    return ORBX_OK;
}

} // namespace orbx
