// Multi-GPU collection of results: frames and keyframe-pair jobs are sharded over one process per GPU with no
// exchange during the computation; the only collective is a FINAL GATHER of what each rank found, to one root rank.
//   1. compaction on the device: an exclusive scan of the per-frame counts and a dense copy of the records in use
//      (28-byte keypoints, 32-byte descriptors; 8-byte (idx1, idx2) records for match lists) -- no padding crosses NVLink;
//   2. the per-rank totals are all-gathered (one 8-byte value per rank) and read by the host: the only host synchronisation;
//   3. grouped ncclSend / ncclRecv of the dense blocks to the root, on the communicator's own stream with few CTAs
//      (ncclConfig_t::maxCTAs), so it overlaps the next step's extraction without taking its SMs.
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

} // namespace orbx

struct orbx_comm {
    ncclComm_t comm = nullptr;
    int nranks = 1, rank = 0, device = 0;
    cudaStream_t stream = nullptr;              // the collection's own stream
    cudaEvent_t marks[orbx::kMarks] = {};       // "results ready" events recorded on a context's stream (orbx_comm_mark)
    int nmarks = 0;
    cudaEvent_t done = nullptr;                 // after the last gather
    uint8_t* dStage = nullptr; size_t stageBytes = 0;      // dense records + offsets of this rank
    long long* dTotals = nullptr;               // [nranks][2] all-gathered (total, output capacity) pairs + this rank's own pair
    long long* hTotals = nullptr;               // pinned mirror
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

// one CTA per frame: rows [0, n) of the padded block -> rows [off, off + n) of the dense blocks
__global__ void __launch_bounds__(256)
compactFramesKernel(const uint32_t* __restrict__ kps, const uint4* __restrict__ desc, const int* __restrict__ n, const int* __restrict__ off,
                    int cap, uint32_t* __restrict__ dk, uint4* __restrict__ dd)
{
    const int f = blockIdx.x, cnt = min(max(n[f], 0), cap), o = off[f];
    const uint32_t* sk = kps + (size_t)f * cap * 7;
    const uint4* sd = desc + (size_t)f * cap * 2;
    for (int i = threadIdx.x; i < cnt * 7; i += blockDim.x) dk[(size_t)o * 7 + i] = sk[i];
    for (int i = threadIdx.x; i < cnt * 2; i += blockDim.x) dd[(size_t)o * 2 + i] = sd[i];
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

// steps 2 and 3 of the file comment; `bytesPerRecord[k]` blocks of `total` records each start at local[k] / land at root[k]
// `compact(dst[])` launches the kernels that write this rank's dense blocks to dst[k]; it runs AFTER the totals are known, so
// the root writes its own records straight to their final place (no staging copy on the rank that is busiest anyway).
template <typename Compact>
static int gatherDense(orbx_comm* c, int root, const int* dCounts, int countsPerRank, int* allCounts, int nblocks, uint8_t* const* staging,
                       uint8_t* const* atRoot, const size_t* bytesPerRecord, size_t capRecords, long long* totalOut, Compact compact)
{   // capRecords: this rank's output capacity (only the root's counts)
    // Every rank contributes (its total, the capacity of ITS output buffers): the total was written to dTotals[2 * nranks] by the
    // scan kernel; the capacity travels along so that all ranks judge the ROOT's buffers and take the same decision (a rank
    // that gave up alone would leave the others waiting in their receives).
    NcclApi* a = ncclApi();
    long long* own = c->dTotals + 2 * (size_t)c->nranks;
    const long long capLL = (long long)capRecords;
    ORBX_CU(c, cudaMemcpyAsync(own + 1, &capLL, sizeof(long long), cudaMemcpyHostToDevice, c->stream));
    if (c->nranks > 1) ORBX_NCCL(c, a->AllGather(own, c->dTotals, 2, ncclInt64, c->comm, c->stream));
    else ORBX_CU(c, cudaMemcpyAsync(c->dTotals, own, 2 * sizeof(long long), cudaMemcpyDeviceToDevice, c->stream));
    ORBX_CU(c, cudaMemcpyAsync(c->hTotals, c->dTotals, 2 * sizeof(long long) * c->nranks, cudaMemcpyDeviceToHost, c->stream));
    ORBX_CU(c, cudaStreamSynchronize(c->stream));
    const long long rootCap = c->hTotals[2 * root + 1];
    long long sum = 0;
    for (int r = 0; r < c->nranks; ++r) { c->hTotals[r] = c->hTotals[2 * r]; sum += c->hTotals[r]; }   // totals packed to the front
    const long long total = c->hTotals[c->rank];
    if (totalOut) *totalOut = sum;
    if (sum > rootCap) { c->lastError = "orbx_gather: the root's buffers are smaller than the records found"; return ORBX_E_CAPACITY; }   // on every rank alike
    {   // this rank's dense blocks: at the root directly behind the lower ranks' records, elsewhere into the staging buffer
        uint8_t* dst[4];
        long long mineBefore = 0;
        for (int r = 0; r < c->rank; ++r) mineBefore += c->hTotals[r];
        for (int k = 0; k < nblocks; ++k) dst[k] = c->rank == root ? atRoot[k] + (size_t)mineBefore * bytesPerRecord[k] : staging[k];
        if (total) compact(dst);
        ORBX_CU(c, cudaGetLastError());
    }
    const uint8_t* const* local = staging;
    if (c->nranks > 1) ORBX_NCCL(c, a->GroupStart());
    if (c->rank == root) {
        long long before = 0;
        for (int r = 0; r < c->nranks; ++r) {
            const long long t = c->hTotals[r];
            if (r == c->rank) {
                ORBX_CU(c, cudaMemcpyAsync(allCounts + (size_t)r * countsPerRank, dCounts, sizeof(int) * (size_t)countsPerRank, cudaMemcpyDeviceToDevice, c->stream));
            } else {
                ORBX_NCCL(c, a->Recv(allCounts + (size_t)r * countsPerRank, (size_t)countsPerRank, ncclInt32, r, c->comm, c->stream));
                for (int k = 0; k < nblocks; ++k)
                    if (t) ORBX_NCCL(c, a->Recv(atRoot[k] + (size_t)before * bytesPerRecord[k], (size_t)t * bytesPerRecord[k], ncclUint8, r, c->comm, c->stream));
            }
            before += t;
        }
    } else {
        ORBX_NCCL(c, a->Send(dCounts, (size_t)countsPerRank, ncclInt32, root, c->comm, c->stream));
        for (int k = 0; k < nblocks; ++k)
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
    int rc = ORBX_OK;
    do {
        if (cudaSetDevice(c->device) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        for (int i = 0; i < kMarks && rc == ORBX_OK; ++i) if (cudaEventCreateWithFlags(&c->marks[i], cudaEventDisableTiming) != cudaSuccess) rc = ORBX_E_CUDA;
        if (rc) break;
        if (cudaEventCreateWithFlags(&c->done, cudaEventDisableTiming) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        if (cudaMalloc(&c->dTotals, 2 * sizeof(long long) * (size_t)(nranks + 1)) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
        if (cudaMallocHost((void**)&c->hTotals, 2 * sizeof(long long) * (size_t)nranks) != cudaSuccess) { rc = ORBX_E_CUDA; break; }
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
    scanCountsKernel<<<1, 1024, 0, c->stream>>>(n, frames, cap, dOff, c->dTotals + 2 * (size_t)c->nranks);
    ORBX_CU(c, cudaGetLastError());
    uint8_t* staging[2] = {c->dStage + oK, c->dStage + oD};
    uint8_t* atRoot[2] = {(uint8_t*)all_kps, all_desc};
    const size_t bpr[2] = {28, 32};
    cudaStream_t st = c->stream;
    return gatherDense(c, root, n, frames, all_n, 2, staging, atRoot, bpr, all_records_cap, total_records, [=](uint8_t* const* dst) {
        compactFramesKernel<<<frames, 256, 0, st>>>((const uint32_t*)kps, (const uint4*)desc, n, dOff, cap, (uint32_t*)dst[0], (uint4*)dst[1]);
    });
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
    scanCountsKernel<<<1, 1024, 0, c->stream>>>(dCnt, pairs, cap, dOff, c->dTotals + 2 * (size_t)c->nranks);
    ORBX_CU(c, cudaGetLastError());
    uint8_t* staging[1] = {c->dStage + oRec};
    uint8_t* atRoot[1] = {(uint8_t*)all_pairs};
    const size_t bpr[1] = {8};
    cudaStream_t st = c->stream;
    return gatherDense(c, root, dCnt, pairs, all_n, 1, staging, atRoot, bpr, all_records_cap, total_records, [=](uint8_t* const* dst) {
        matchListKernel<<<pairs, 256, 0, st>>>(match, rows, cap, dCnt, dOff, (int2*)dst[0], 1);
    });
}
