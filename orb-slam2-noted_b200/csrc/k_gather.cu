// Multi-GPU collection of results: frames and keyframe-pair jobs are sharded over one process per GPU with no
// exchange during the computation; the only collective is a FINAL GATHER of what each rank found, to one root rank.
//   1. compaction on the device: an exclusive scan of the per-frame counts and a dense copy of the records in use
//      (28-byte keypoints, 32-byte descriptors; 8-byte (idx1, idx2) records for match lists) -- no padding crosses NVLink;
//   2. the per-rank totals are all-gathered (one 8-byte value per rank) and read by the host: the only host synchronisation;
//   3. the dense blocks go to the root, on the communicator's own HIGH-PRIORITY stream so that they overlap the next step's
//      extraction (the priority matters: without it the collectives' few CTAs queue behind the extraction kernels' grids and
//      the chain scan -> totals -> host -> compaction -> transfer outlasts a step):
//      a. PEER WINDOW (orbx_comm_window): when the root's buffers lie in memory the other ranks have mapped (CUDA IPC over
//         NVLink / NVSwitch), nothing goes through NCCL's staged protocol: every rank compacts locally and then writes its
//         blocks STRAIGHT INTO THE ROOT'S HBM at their final place, by default with one peer copy per block on the DMA
//         engines (no SM involved), or with the compaction kernel's own stores (ORBX_GATHER_COPY=0: compaction and transfer
//         in one kernel).  The per-rank counts still travel by ncclSend/ncclRecv: 4 KB that double as the "my writes have
//         landed" handshake, enqueued behind the copies / the kernel whose completion makes the peer writes visible.
//         Measured on 8 B200s (profiles/r2_gather_sweep_8gpu.log, 433 MB into rank 0 per step, 4.37 ms per step on 1 GPU):
//         DMA copies 4.52 ms per step; SM stores 4.64 (12 CTAs) .. 5.0 (one CTA per frame): seven writers into one NVLink
//         ingress is an in-cast, the stores back up into the writers' memory systems and their extraction kernels slow down
//         by 0.1 - 0.6 ms per step, more the more CTAs store; few CTAs avoid that but stretch the chain (4 CTAs: 4.85);
//      b. otherwise grouped ncclSend / ncclRecv of staged blocks with few CTAs (ncclConfig_t::maxCTAs): 4.60 ms per step.
//   Records that are dense already (orbx_extract_batch with ORBX_OUT_DENSE: the describe stage writes frame b's records right
//   behind frame b - 1's) skip step 1's copy: orbx_gather_frames_dense scans the counts for the total and sends the caller's
//   blocks as they lie (4.49 ms per step on 8 B200s = 0.971 of 8 x one GPU).
// NCCL is loaded at run time (dlopen "libnccl.so.2"): a single-GPU user of liborbx.so needs no NCCL, and inside a process that
// already carries NCCL (torch) the same library is used.  The reference has no counterpart (it is single-process CPU code);
// the contract is north_star's: "only a final NCCL gather of keypoint and match lists over NVLink".
#include "orbx_internal.h"
#include <dlfcn.h>
#include <nccl.h>
#include <mutex>
#include <vector>

namespace orbx {

struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRankConfig)(ncclComm_t*, int, ncclUniqueId, int, ncclConfig_t*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};

static NcclApi* ncclApi()
{
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, []() {
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) return;
        auto sym = [&](const char* n) { return dlsym(h, n); };
        api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
        api.CommInitRankConfig = (decltype(api.CommInitRankConfig))sym("ncclCommInitRankConfig");
        api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
        api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
        api.Send = (decltype(api.Send))sym("ncclSend");
        api.Recv = (decltype(api.Recv))sym("ncclRecv");
        api.AllGather = (decltype(api.AllGather))sym("ncclAllGather");
        api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
        api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
        api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
        api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.Send && api.Recv && api.AllGather && api.GroupStart && api.GroupEnd;
    });
    return &api;
}

constexpr int kMarks = 8;
constexpr int kOwn = 4;                         // int64 values each rank contributes per gather: total, capacity, window offset of block 0 / 1

} // namespace orbx

struct orbx_comm {
    ncclComm_t comm = nullptr;
    int nranks = 1, rank = 0, device = 0;
    cudaStream_t stream = nullptr;              // the collection's own stream
    cudaEvent_t marks[orbx::kMarks] = {};       // "results ready" events recorded on a context's stream (orbx_comm_mark)
    int nmarks = 0;
    cudaEvent_t done = nullptr;                 // after the last gather
    uint8_t* dStage = nullptr; size_t stageBytes = 0;      // dense records + offsets of this rank
    long long* dTotals = nullptr;               // [nranks][kOwn] all-gathered (total, output capacity, window offsets) + this rank's own entry
    long long* hTotals = nullptr;               // pinned mirror
    // peer window: winBase[r] is rank r's window as mapped HERE (own rank: the allocation itself), winBytes[r] its size
    std::vector<uint8_t*> winBase; std::vector<size_t> winBytes;
    bool remoteByCopyEngine = true;             // ORBX_GATHER_COPY=0: the compaction kernel stores into the root's window itself (no staging, no DMA engine)
    int directGathers = 0, stagedGathers = 0;   // how the gathers of this communicator travelled (orbx_comm_stats)
    std::string lastError;
};

namespace orbx {

static int ncclFail(orbx_comm* c, ncclResult_t r, const char* what)
{
    NcclApi* a = ncclApi();
    c->lastError = std::string(what) + ": " + (a->GetErrorString ? a->GetErrorString(r) : "NCCL error");
    return ORBX_E_CUDA;
}
#define ORBX_NCCL(c, call) do { ncclResult_t r_ = (call); if (r_ != ncclSuccess) return orbx::ncclFail((c), r_, #call); } while (0)
#define ORBX_CU(c, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { (c)->lastError = std::string(#call) + ": " + cudaGetErrorString(e_); return ORBX_E_CUDA; } } while (0)

static int ensureStage(orbx_comm* c, size_t bytes)
{
    if (bytes <= c->stageBytes) return ORBX_OK;
    if (c->dStage) { cudaStreamSynchronize(c->stream); cudaFree(c->dStage); c->dStage = nullptr; c->stageBytes = 0; }
    bytes = (bytes + (1u << 20)) & ~(size_t)((1u << 20) - 1);
    ORBX_CU(c, cudaMalloc(&c->dStage, bytes));
    c->stageBytes = bytes;
    return ORBX_OK;
}

// off[i] = sum of min(n[j], cap) for j < i; off[count] = total; *total64 = total.  One CTA (count <= a few thousand).
__global__ void __launch_bounds__(1024) scanCountsKernel(const int* __restrict__ n, int count, int cap, int* __restrict__ off, long long* __restrict__ total64)
{
    __shared__ int warpSum[32];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int base = 0; base < count; base += 1024) {
        const int i = base + threadIdx.x;
        const int v = i < count ? min(max(n[i], 0), cap) : 0;
        int s = v;
        #pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, s, o); if (lane >= o) s += t; }
        if (lane == 31) warpSum[w] = s;
        __syncthreads();
        if (w == 0) {
            int t = warpSum[lane];
            #pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(0xffffffffu, t, o); if (lane >= o) t += u; }
            warpSum[lane] = t;
        }
        __syncthreads();
        const int before = carry + (w ? warpSum[w - 1] : 0) + s - v;
        if (i < count) off[i] = before;
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) { off[count] = carry; *total64 = carry; }
}

// frames strided over the CTAs: rows [0, n) of each padded block -> rows [off, off + n) of the dense blocks.  A launch that writes
// into local memory uses one CTA per frame; one that stores into the root's peer window (ORBX_GATHER_COPY=0) uses a few CTAs
// only (kRemoteCtas, env ORBX_GATHER_CTAS): see 3a of the file comment.
constexpr int kRemoteCtas = 12;
__global__ void __launch_bounds__(256)
compactFramesKernel(const uint32_t* __restrict__ kps, const uint4* __restrict__ desc, const int* __restrict__ n, const int* __restrict__ off,
                    int frames, int cap, uint32_t* __restrict__ dk, uint4* __restrict__ dd)
{
    for (int f = blockIdx.x; f < frames; f += gridDim.x) {
        const int cnt = min(max(n[f], 0), cap), o = off[f];
        const uint32_t* sk = kps + (size_t)f * cap * 7;
        const uint4* sd = desc + (size_t)f * cap * 2;
        #pragma unroll 4
        for (int i = threadIdx.x; i < cnt * 7; i += blockDim.x) dk[(size_t)o * 7 + i] = sk[i];
        #pragma unroll 4
        for (int i = threadIdx.x; i < cnt * 2; i += blockDim.x) dd[(size_t)o * 2 + i] = sd[i];
    }
}

// one CTA per keyframe pair: count (pass 0) or write (pass 1) the (row, match[row]) records with match >= 0, in row order
__global__ void __launch_bounds__(256)
matchListKernel(const int* __restrict__ match, const int* __restrict__ rows, int cap, int* __restrict__ cnt, const int* __restrict__ off,
                int2* __restrict__ out, int pass)
{
    __shared__ int warpSum[8];
    __shared__ int carry;
    const int p = blockIdx.x, nr = min(max(rows[p], 0), cap);
    const int* m = match + (size_t)p * cap;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int base = 0; base < nr; base += 256) {
        const int i = base + threadIdx.x;
        const int v = i < nr ? m[i] : -1;
        const uint32_t bal = __ballot_sync(0xffffffffu, v >= 0);
        if (lane == 0) warpSum[w] = __popc(bal);
        __syncthreads();
        int before = carry;
        for (int k = 0; k < w; ++k) before += warpSum[k];
        if (pass && v >= 0) out[off[p] + before + __popc(bal & ((1u << lane) - 1u))] = make_int2(i, v);
        __syncthreads();
        if (threadIdx.x == 0) { int t = carry; for (int k = 0; k < 8; ++k) t += warpSum[k]; carry = t; }
        __syncthreads();
    }
    if (!pass && threadIdx.x == 0) cnt[p] = carry;
}

// steps 2 and 3 of the file comment; `bytesPerRecord[k]` blocks of `total` records each are staged at staging[k] / land at atRoot[k].
// `compact(dst[])` launches the kernels that write this rank's dense blocks to dst[k]; it runs AFTER the totals are known, so
// the root writes its own records straight to their final place, and so does every other rank when the root's buffers lie in
// its peer window (3a) -- otherwise the blocks are staged locally and sent (3b).
template <typename Compact>
static int gatherDense(orbx_comm* c, int root, const int* dCounts, int countsPerRank, int* allCounts, int nblocks, uint8_t* const* staging,
                       uint8_t* const* atRoot, const size_t* bytesPerRecord, size_t capRecords, long long* totalOut, Compact compact,
                       const uint8_t* const* denseSrc = nullptr)
{   // capRecords: this rank's output capacity (only the root's counts)
    // denseSrc: the caller's records are dense already (block k at denseSrc[k]): nothing is compacted or staged, the blocks are
    // copied / sent from where they lie
    // Every rank contributes (its total, the capacity of ITS output buffers, where ITS output buffers lie in its window): the
    // total was written to dTotals[kOwn * nranks] by the scan kernel; the rest travels along so that all ranks judge the ROOT's
    // buffers and take the same decisions (a rank that gave up alone would leave the others waiting in their receives).
    NcclApi* a = ncclApi();
    long long* own = c->dTotals + kOwn * (size_t)c->nranks;
    long long mine[kOwn - 1] = {(long long)capRecords, -1, -1};
    if (c->rank == root && !c->winBase.empty() && c->winBase[root] && !getenv("ORBX_GATHER_STAGED"))
        for (int k = 0; k < nblocks && k < 2; ++k) {
            const uint8_t* lo = c->winBase[root]; const uint8_t* hi = lo + c->winBytes[root];
            if (atRoot[k] >= lo && atRoot[k] + capRecords * bytesPerRecord[k] <= hi) mine[1 + k] = (long long)(atRoot[k] - lo);
        }
    ORBX_CU(c, cudaMemcpyAsync(own + 1, mine, sizeof(mine), cudaMemcpyHostToDevice, c->stream));
    if (c->nranks > 1) ORBX_NCCL(c, a->AllGather(own, c->dTotals, kOwn, ncclInt64, c->comm, c->stream));
    else ORBX_CU(c, cudaMemcpyAsync(c->dTotals, own, kOwn * sizeof(long long), cudaMemcpyDeviceToDevice, c->stream));
    ORBX_CU(c, cudaMemcpyAsync(c->hTotals, c->dTotals, kOwn * sizeof(long long) * c->nranks, cudaMemcpyDeviceToHost, c->stream));
    ORBX_CU(c, cudaStreamSynchronize(c->stream));
    const long long rootCap = c->hTotals[kOwn * root + 1];
    long long winOff[2] = {c->hTotals[kOwn * root + 2], c->hTotals[kOwn * root + 3]};
    bool direct = c->nranks > 1 && !c->winBase.empty() && c->winBase[root];
    for (int k = 0; k < nblocks; ++k) direct = direct && k < 2 && winOff[k] >= 0;
    long long sum = 0;
    for (int r = 0; r < c->nranks; ++r) { c->hTotals[r] = c->hTotals[kOwn * r]; sum += c->hTotals[r]; }   // totals packed to the front
    const long long total = c->hTotals[c->rank];
    if (totalOut) *totalOut = sum;
    if (sum > rootCap) { c->lastError = "orbx_gather: the root's buffers are smaller than the records found"; return ORBX_E_CAPACITY; }   // on every rank alike
    {   // this rank's dense blocks: behind the lower ranks' records in the root's buffers (own or mapped), else into the staging buffer
        uint8_t* dst[4];
        long long mineBefore = 0;
        for (int r = 0; r < c->rank; ++r) mineBefore += c->hTotals[r];
        const bool remote = direct && c->rank != root;
        const bool byCopyEngine = remote && (c->remoteByCopyEngine || denseSrc);
        uint8_t* far[4];
        for (int k = 0; k < nblocks; ++k) {
            far[k] = remote ? c->winBase[root] + winOff[k] + (size_t)mineBefore * bytesPerRecord[k] : nullptr;
            dst[k] = c->rank == root ? atRoot[k] + (size_t)mineBefore * bytesPerRecord[k] : (remote && !byCopyEngine) ? far[k] : staging[k];
        }
        if (denseSrc) {
            for (int k = 0; k < nblocks && total && c->rank == root; ++k)
                ORBX_CU(c, cudaMemcpyAsync(dst[k], denseSrc[k], (size_t)total * bytesPerRecord[k], cudaMemcpyDeviceToDevice, c->stream));
        } else {
            if (total) compact(dst, remote && !byCopyEngine);
            ORBX_CU(c, cudaGetLastError());
        }
        for (int k = 0; k < nblocks && byCopyEngine && total; ++k)
            ORBX_CU(c, cudaMemcpyAsync(far[k], denseSrc ? denseSrc[k] : staging[k], (size_t)total * bytesPerRecord[k], cudaMemcpyDefault, c->stream));
    }
    (direct ? c->directGathers : c->stagedGathers)++;
    const uint8_t* const* local = denseSrc ? denseSrc : staging;
    if (c->nranks > 1) ORBX_NCCL(c, a->GroupStart());
    if (c->rank == root) {
        long long before = 0;
        for (int r = 0; r < c->nranks; ++r) {
            const long long t = c->hTotals[r];
            if (r == c->rank) {
                ORBX_CU(c, cudaMemcpyAsync(allCounts + (size_t)r * countsPerRank, dCounts, sizeof(int) * (size_t)countsPerRank, cudaMemcpyDeviceToDevice, c->stream));
            } else {
                // in the direct form this receive is also the handshake: rank r enqueued its send behind its compaction kernel
                ORBX_NCCL(c, a->Recv(allCounts + (size_t)r * countsPerRank, (size_t)countsPerRank, ncclInt32, r, c->comm, c->stream));
                for (int k = 0; k < nblocks && !direct; ++k)
                    if (t) ORBX_NCCL(c, a->Recv(atRoot[k] + (size_t)before * bytesPerRecord[k], (size_t)t * bytesPerRecord[k], ncclUint8, r, c->comm, c->stream));
            }
            before += t;
        }
    } else {
        ORBX_NCCL(c, a->Send(dCounts, (size_t)countsPerRank, ncclInt32, root, c->comm, c->stream));
        for (int k = 0; k < nblocks && !direct; ++k)
            if (total) ORBX_NCCL(c, a->Send(local[k], (size_t)total * bytesPerRecord[k], ncclUint8, root, c->comm, c->stream));
    }
    if (c->nranks > 1) ORBX_NCCL(c, a->GroupEnd());
    ORBX_CU(c, cudaEventRecord(c->done, c->stream));
    return ORBX_OK;
}

static int waitTicket(orbx_comm* c, orbx_ctx* ctx, int ticket)
{
    if (ticket >= 0) { ORBX_CU(c, cudaStreamWaitEvent(c->stream, c->marks[ticket % kMarks], 0)); return ORBX_OK; }
    if (!ctx) return ORBX_OK;
    ORBX_CU(c, cudaEventRecord(c->marks[kMarks - 1], ctx->c.stream));        // everything enqueued on the context so far
    ORBX_CU(c, cudaStreamWaitEvent(c->stream, c->marks[kMarks - 1], 0));
    return ORBX_OK;
}

} // namespace orbx
using namespace orbx;

extern "C" int orbx_comm_unique_id(uint8_t* id)
{
    if (!id) return ORBX_E_INVALID;
    NcclApi* a = ncclApi();
    if (!a->ok) return ORBX_E_UNSUPPORTED;
    static_assert(sizeof(ncclUniqueId) == ORBX_COMM_ID_BYTES, "ncclUniqueId size");
    ncclUniqueId u;
    if (a->GetUniqueId(&u) != ncclSuccess) return ORBX_E_CUDA;
    memcpy(id, &u, sizeof(u));
    return ORBX_OK;
}

extern "C" int orbx_comm_create(orbx_comm** out, orbx_ctx* ctx, int nranks, int rank, const uint8_t* id)
{
    if (!out || !ctx || nranks < 1 || rank < 0 || rank >= nranks || (nranks > 1 && !id)) return ORBX_E_INVALID;
    *out = nullptr;
    NcclApi* a = ncclApi();
    if (nranks > 1 && !a->ok) return ORBX_E_UNSUPPORTED;          // no NCCL library in this process' reach
    orbx_comm* c = new (std::nothrow) orbx_comm();
    if (!c) return ORBX_E_NOMEM;
    c->nranks = nranks; c->rank = rank; c->device = ctx->c.device;
    if (const char* e = getenv("ORBX_GATHER_COPY")) c->remoteByCopyEngine = atoi(e) != 0;
    int rc = ORBX_OK;
    do {
        if (cudaSetDevice(c->device) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        // highest priority: the scan, the collectives' few CTAs and the compaction get the next free SM slots ahead of the queued
        // CTAs of the extraction kernels they overlap, so the chain (scan -> totals -> host -> compaction -> handshake) stays short
        int prLeast = 0, prGreatest = 0;
        cudaDeviceGetStreamPriorityRange(&prLeast, &prGreatest);
        if (cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, prGreatest) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        for (int i = 0; i < kMarks && rc == ORBX_OK; ++i) if (cudaEventCreateWithFlags(&c->marks[i], cudaEventDisableTiming) != cudaSuccess) rc = ORBX_E_CUDA;
        if (rc) break;
        if (cudaEventCreateWithFlags(&c->done, cudaEventDisableTiming) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        if (cudaMalloc(&c->dTotals, kOwn * sizeof(long long) * (size_t)(nranks + 1)) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        if (cudaMallocHost((void**)&c->hTotals, kOwn * sizeof(long long) * (size_t)nranks) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        if (cudaEventRecord(c->done, c->stream) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        if (nranks > 1) {
            ncclUniqueId u;
            memcpy(&u, id, sizeof(u));
            ncclResult_t r;
            if (a->CommInitRankConfig) {
                ncclConfig_t cfg = NCCL_CONFIG_INITIALIZER;
                cfg.maxCTAs = 4;                                  // a gather of a few hundred MB per step: keep the SMs for the extraction
                if (const char* e = getenv("ORBX_NCCL_MAX_CTAS")) cfg.maxCTAs = std::max(1, atoi(e));
                r = a->CommInitRankConfig(&c->comm, nranks, u, rank, &cfg);
            } else r = a->CommInitRank(&c->comm, nranks, u, rank);
            if (r != ncclSuccess) { rc = ORBX_E_CUDA; break; }
        }
    } while (0);
    if (rc) { orbx_comm_destroy(c); return rc; }
    *out = c;
    return ORBX_OK;
}

extern "C" void orbx_comm_destroy(orbx_comm* c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    for (size_t r = 0; r < c->winBase.size(); ++r)             // the others' windows first, then this rank's own (all gathers are done: see orbx.h)
        if (c->winBase[r]) { if ((int)r == c->rank) cudaFree(c->winBase[r]); else cudaIpcCloseMemHandle(c->winBase[r]); }
    if (c->comm) ncclApi()->CommDestroy(c->comm);
    for (int i = 0; i < kMarks; ++i) if (c->marks[i]) cudaEventDestroy(c->marks[i]);
    if (c->done) cudaEventDestroy(c->done);
    if (c->dStage) cudaFree(c->dStage);
    if (c->dTotals) cudaFree(c->dTotals);
    if (c->hTotals) cudaFreeHost(c->hTotals);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

extern "C" const char* orbx_comm_last_error(const orbx_comm* c) { return c ? c->lastError.c_str() : ""; }

extern "C" int orbx_comm_mark(orbx_comm* c, orbx_ctx* ctx)
{
    if (!c || !ctx) return ORBX_E_INVALID;
    const int t = c->nmarks % (kMarks - 1);                       // the last slot belongs to ticket < 0
    if (cudaEventRecord(c->marks[t], ctx->c.stream) != cudaSuccess) return ORBX_E_CUDA;
    c->nmarks++;
    return t;
}

extern "C" int orbx_comm_wait(orbx_comm* c, orbx_ctx* ctx)
{
    if (!c || !ctx) return ORBX_E_INVALID;
    return cudaStreamWaitEvent(ctx->c.stream, c->done, 0) == cudaSuccess ? ORBX_OK : ORBX_E_CUDA;
}

extern "C" int orbx_comm_sync(orbx_comm* c)
{
    if (!c) return ORBX_E_INVALID;
    return cudaStreamSynchronize(c->stream) == cudaSuccess ? ORBX_OK : ORBX_E_CUDA;
}

// The peer window (3a of the file comment).  Collective: every rank calls it once, with the bytes IT wants to be able to receive
// (0 on ranks that are never a root).  The allocation is a plain cudaMalloc (IPC handles cover whole allocations), its handle
// and size are all-gathered, and every rank maps every non-empty window of the others.  All ranks end with the same answer:
// ORBX_OK and the windows in place, or ORBX_E_UNSUPPORTED (some rank could not export / map: no peer access between the
// devices, IPC not permitted in this container) with nothing left mapped -- the gathers then use ncclSend/ncclRecv.
extern "C" int orbx_comm_window(orbx_comm* c, size_t bytes, void** base)
{
    if (!c || !base) return ORBX_E_INVALID;
    *base = nullptr;
    if (!c->winBase.empty()) { c->lastError = "orbx_comm_window: this communicator has its window already"; return ORBX_E_INVALID; }
    ORBX_CU(c, cudaSetDevice(c->device));
    struct Entry { cudaIpcMemHandle_t h; unsigned long long bytes; long long ok; };
    static_assert(sizeof(Entry) == 80, "window entry");
    NcclApi* a = ncclApi();
    const int R = c->nranks;
    uint8_t* mineDev = nullptr;
    Entry me; memset(&me, 0, sizeof(me)); me.ok = 1;
    bytes = (bytes + 255) & ~(size_t)255;
    if (bytes) {
        if (cudaMalloc(&mineDev, bytes) != cudaSuccess) { cudaGetLastError(); me.ok = 0; mineDev = nullptr; }
        else if (R > 1 && cudaIpcGetMemHandle(&me.h, mineDev) != cudaSuccess) { cudaGetLastError(); me.ok = 0; }
        me.bytes = me.ok ? bytes : 0;
    }
    std::vector<Entry> all((size_t)R);
    Entry* dAll = nullptr;
    auto exchange = [&](void) -> int {                     // all-gather `me` into `all` (through device memory: NCCL is the only transport here)
        if (R == 1) { all[0] = me; return ORBX_OK; }
        if (!dAll) ORBX_CU(c, cudaMalloc(&dAll, sizeof(Entry) * (size_t)(R + 1)));
        ORBX_CU(c, cudaMemcpyAsync(dAll + R, &me, sizeof(Entry), cudaMemcpyHostToDevice, c->stream));
        ORBX_NCCL(c, a->AllGather(dAll + R, dAll, sizeof(Entry), ncclUint8, c->comm, c->stream));
        ORBX_CU(c, cudaMemcpyAsync(all.data(), dAll, sizeof(Entry) * (size_t)R, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CU(c, cudaStreamSynchronize(c->stream));
        return ORBX_OK;
    };
    int rc = exchange();
    std::vector<uint8_t*> mapped((size_t)R, nullptr);
    if (rc == ORBX_OK) {
        bool ok = true;
        for (int r = 0; r < R; ++r) ok = ok && all[r].ok;
        for (int r = 0; r < R && ok; ++r) {
            if (r == c->rank) { mapped[r] = mineDev; continue; }
            if (!all[r].bytes) continue;
            void* p = nullptr;
            if (cudaIpcOpenMemHandle(&p, all[r].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                c->lastError = std::string("orbx_comm_window: cudaIpcOpenMemHandle: ") + cudaGetErrorString(cudaGetLastError());
                ok = false;
            } else mapped[r] = (uint8_t*)p;
        }
        me.ok = ok ? 1 : 0;                                 // second round: did EVERY rank map everything?
        rc = exchange();
        if (rc == ORBX_OK) for (int r = 0; r < R; ++r) if (!all[r].ok) rc = ORBX_E_UNSUPPORTED;
    }
    if (dAll) cudaFree(dAll);
    if (rc != ORBX_OK) {
        for (int r = 0; r < R; ++r) if (r != c->rank && mapped[r]) cudaIpcCloseMemHandle(mapped[r]);
        if (mineDev) cudaFree(mineDev);
        if (rc == ORBX_E_UNSUPPORTED && c->lastError.empty()) c->lastError = "orbx_comm_window: another rank could not export or map its window";
        return rc;
    }
    c->winBase = mapped;
    c->winBytes.assign((size_t)R, 0);
    for (int r = 0; r < R; ++r) c->winBytes[r] = (size_t)all[r].bytes;
    c->winBytes[c->rank] = mineDev ? bytes : 0;
    *base = mineDev;
    return ORBX_OK;
}

extern "C" int orbx_comm_stats(const orbx_comm* c, int* direct_gathers, int* staged_gathers)
{
    if (!c) return ORBX_E_INVALID;
    if (direct_gathers) *direct_gathers = c->directGathers;
    if (staged_gathers) *staged_gathers = c->stagedGathers;
    return ORBX_OK;
}

extern "C" int orbx_gather_frames(orbx_comm* c, orbx_ctx* ctx, int ticket, const orbx_keypoint* kps, const uint8_t* desc, const int32_t* n,
                                  int frames, int cap, int root, int32_t* all_n, orbx_keypoint* all_kps, uint8_t* all_desc,
                                  size_t all_records_cap, long long* total_records)
{
    if (!c || frames < 1 || cap < 1 || !kps || !desc || !n || root < 0 || root >= c->nranks) return ORBX_E_INVALID;
    if (c->rank == root && (!all_n || !all_kps || !all_desc)) return ORBX_E_INVALID;
    ORBX_CU(c, cudaSetDevice(c->device));
    int rc = waitTicket(c, ctx, ticket);
    if (rc) return rc;
    const size_t rows = (size_t)frames * cap;
    const size_t oOff = 0, oK = ((sizeof(int) * (size_t)(frames + 1)) + 255) & ~(size_t)255, oD = (oK + 28 * rows + 255) & ~(size_t)255;
    if ((rc = ensureStage(c, oD + 32 * rows + 256))) return rc;
    int* dOff = (int*)(c->dStage + oOff);
    scanCountsKernel<<<1, 1024, 0, c->stream>>>(n, frames, cap, dOff, c->dTotals + kOwn * (size_t)c->nranks);
    ORBX_CU(c, cudaGetLastError());
    uint8_t* staging[2] = {c->dStage + oK, c->dStage + oD};
    uint8_t* atRoot[2] = {(uint8_t*)all_kps, all_desc};
    const size_t bpr[2] = {28, 32};
    cudaStream_t st = c->stream;
    int remoteCtas = kRemoteCtas;
    if (const char* e = getenv("ORBX_GATHER_CTAS")) remoteCtas = std::max(1, atoi(e));
    return gatherDense(c, root, n, frames, all_n, 2, staging, atRoot, bpr, all_records_cap, total_records, [=](uint8_t* const* dst, bool remote) {
        const int grid = remote ? std::min(frames, remoteCtas) : frames;
        compactFramesKernel<<<grid, 256, 0, st>>>((const uint32_t*)kps, (const uint4*)desc, n, dOff, frames, cap, (uint32_t*)dst[0], (uint4*)dst[1]);
    });
}

extern "C" int orbx_gather_frames_dense(orbx_comm* c, orbx_ctx* ctx, int ticket, const orbx_keypoint* kps, const uint8_t* desc, const int32_t* n,
                                        int frames, int cap, int root, int32_t* all_n, orbx_keypoint* all_kps, uint8_t* all_desc,
                                        size_t all_records_cap, long long* total_records)
{
    if (!c || frames < 1 || cap < 1 || !kps || !desc || !n || root < 0 || root >= c->nranks) return ORBX_E_INVALID;
    if (c->rank == root && (!all_n || !all_kps || !all_desc)) return ORBX_E_INVALID;
    ORBX_CU(c, cudaSetDevice(c->device));
    int rc = waitTicket(c, ctx, ticket);
    if (rc) return rc;
    if ((rc = ensureStage(c, sizeof(int) * (size_t)(frames + 1) + 256))) return rc;
    scanCountsKernel<<<1, 1024, 0, c->stream>>>(n, frames, cap, (int*)c->dStage, c->dTotals + kOwn * (size_t)c->nranks);   // only the total is used
    ORBX_CU(c, cudaGetLastError());
    uint8_t* staging[2] = {nullptr, nullptr};
    uint8_t* atRoot[2] = {(uint8_t*)all_kps, all_desc};
    const uint8_t* dense[2] = {(const uint8_t*)kps, desc};
    const size_t bpr[2] = {28, 32};
    return gatherDense(c, root, n, frames, all_n, 2, staging, atRoot, bpr, all_records_cap, total_records, [](uint8_t* const*, bool) {}, dense);
}

extern "C" int orbx_gather_matches(orbx_comm* c, orbx_ctx* ctx, int ticket, const int32_t* match, const int32_t* rows, int pairs, int cap,
                                   int root, int32_t* all_n, int32_t* all_pairs, size_t all_records_cap, long long* total_records)
{
    if (!c || pairs < 1 || cap < 1 || !match || !rows || root < 0 || root >= c->nranks) return ORBX_E_INVALID;
    if (c->rank == root && (!all_n || !all_pairs)) return ORBX_E_INVALID;
    ORBX_CU(c, cudaSetDevice(c->device));
    int rc = waitTicket(c, ctx, ticket);
    if (rc) return rc;
    const size_t oCnt = 0, oOff = (sizeof(int) * (size_t)pairs + 255) & ~(size_t)255, oRec = (oOff + sizeof(int) * (size_t)(pairs + 1) + 255) & ~(size_t)255;
    if ((rc = ensureStage(c, oRec + 8 * (size_t)pairs * cap + 256))) return rc;
    int* dCnt = (int*)(c->dStage + oCnt); int* dOff = (int*)(c->dStage + oOff);
    matchListKernel<<<pairs, 256, 0, c->stream>>>(match, rows, cap, dCnt, nullptr, nullptr, 0);
    scanCountsKernel<<<1, 1024, 0, c->stream>>>(dCnt, pairs, cap, dOff, c->dTotals + kOwn * (size_t)c->nranks);
    ORBX_CU(c, cudaGetLastError());
    uint8_t* staging[1] = {c->dStage + oRec};
    uint8_t* atRoot[1] = {(uint8_t*)all_pairs};
    const size_t bpr[1] = {8};
    cudaStream_t st = c->stream;
    return gatherDense(c, root, dCnt, pairs, all_n, 1, staging, atRoot, bpr, all_records_cap, total_records, [=](uint8_t* const* dst, bool) {
        matchListKernel<<<pairs, 256, 0, st>>>(match, rows, cap, dCnt, dOff, (int2*)dst[0], 1);
    });
}
