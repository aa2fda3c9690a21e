// K7 -- DBoW2 tree descent (TemplatedVocabulary::transform, Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1255-1306)
// with FORB::distance (FORB.cpp:81-101) = the same 256-bit XOR+POPC as K6.
//
// One warp per feature: lane c evaluates child c of the current node (k <= 32), the warp takes the
// (distance, child position) minimum -- strict '<' in the reference means the FIRST child wins ties -- and
// descends; L sequential steps.  The node table (k=10, L=6: 1.11 M nodes x 32 B = 35 MB) is L2-resident on
// B200 (126 MB), so the 2 x 16-byte gathers per lane are L2 hits after the first frames.
#include "orbx_internal.h"
#include <cstdio>
#include <cstdlib>
#include <stdio.h>
#include <string.h>
#include <algorithm>
#include <fstream>
#include <sstream>
#include <new>
#include <vector>

struct orbx_vocab {
    int device = 0, k = 0, L = 0, nnodes = 0, nwords = 0;
    int32_t* dChildOff = nullptr; int32_t* dChild = nullptr; uint4* dDesc = nullptr; int32_t* dWord = nullptr; float* dWeight = nullptr;
    double* dWordWeight = nullptr;      // weight of word id w as a double (WordValue), for the BowVector
    int wordCap = 0;                    // 1 + largest word id
    int scoring = 0, weighting = 0;     // DBoW2 ScoringType / WeightingType of the text header (0, 0 = L1_NORM, TF_IDF: ORBvoc)
};

namespace orbx {

__global__ void __launch_bounds__(256)
bowTransformKernel(const int32_t* __restrict__ childOff, const int32_t* __restrict__ child, const uint4* __restrict__ nodeDesc,
                   const int32_t* __restrict__ wordId, const float* __restrict__ weight, const uint4* __restrict__ desc, int n,
                   int nidLevel, int32_t* __restrict__ outWord, float* __restrict__ outWeight, int32_t* __restrict__ outNode)
{
    const int lane = threadIdx.x & 31;
    const int f = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (f >= n) return;
    const uint4 a0 = desc[2 * f], a1 = desc[2 * f + 1];
    int node = 0, nid = 0, level = 0;
    for (;;) {
        const int c0 = childOff[node], c1 = childOff[node + 1];
        if (c1 <= c0) break; // leaf
        ++level;
        uint32_t best = 0xffffffffu;
        for (int base = c0; base < c1; base += 32) { // k <= 32 in practice: one pass
            uint32_t key = 0xffffffffu;
            if (base + lane < c1) {
                const int id = child[base + lane];
                const uint4 b0 = nodeDesc[2 * id], b1 = nodeDesc[2 * id + 1];
                const int d = hamming256(a0, a1, b0, b1);
                key = ((uint32_t)d << 20) | (uint32_t)(base + lane - c0); // first child wins ties
            }
            #pragma unroll
            for (int o = 16; o > 0; o >>= 1) key = min(key, __shfl_xor_sync(0xffffffffu, key, o));
            best = min(best, key);
        }
        node = child[c0 + (int)(best & 0xfffffu)];
        if (level == nidLevel) nid = node;
    }
    if (lane == 0) { outWord[f] = wordId[node]; outWeight[f] = weight[node]; outNode[f] = nid; }
}

// ---------------- BowVector + L1 score (the step after transform) ----------------
// TemplatedVocabulary::transform's BowVector half (TemplatedVocabulary.h:1166-1221: addWeight per feature in feature
// order, then BowVector::normalize(L1), BowVector.cpp:37-49,59-82) and L1Scoring::score (ScoringObject.cpp:28-61).
// WordValue is a double and the reference adds in a fixed order (features of a word in feature order; the norm and the
// score in ascending word id), so the sums here are sequential in exactly that order; everything around them is parallel.
constexpr int kBowThreads = 1024;

__global__ void __launch_bounds__(kBowThreads)
bowVectorKernel(const int32_t* __restrict__ word, int n, int npow2, const double* __restrict__ wordWeight, int wordCap,
                int32_t* __restrict__ ids, double* __restrict__ values, int* __restrict__ nOut)
{
    extern __shared__ unsigned long long keys[];        // (word << 32 | feature), sorted; later reused for the sums
    __shared__ int sWarpTot[32];
    __shared__ int sTotal;
    __shared__ double sNorm;
    const int tid = threadIdx.x;
    for (int i = tid; i < npow2; i += kBowThreads) {
        unsigned long long k = ~0ull;
        if (i < n) {
            const int w = word[i];
            if (w >= 0 && w < wordCap && wordWeight[w] > 0.0) k = ((unsigned long long)(uint32_t)w << 32) | (uint32_t)i;   // w > 0: not stopped (:1181)
        }
        keys[i] = k;
    }
    __syncthreads();
    for (int size = 2; size <= npow2; size <<= 1)
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = tid; t < (npow2 >> 1); t += kBowThreads) {
                const int lo = 2 * t - (t & (stride - 1)), hi = lo + stride;
                const bool up = (lo & size) == 0;
                const unsigned long long a = keys[lo], b = keys[hi];
                if ((a > b) == up) { keys[lo] = b; keys[hi] = a; }
            }
            __syncthreads();
        }
    // heads of the runs of equal word id, compacted in order
    const int per = npow2 / kBowThreads > 0 ? npow2 / kBowThreads : 1;
    const int i0 = tid * per;
    int cnt = 0;
    for (int i = i0; i < i0 + per && i < npow2; ++i) {
        const unsigned long long k = keys[i];
        if (k != ~0ull && (i == 0 || (uint32_t)(keys[i - 1] >> 32) != (uint32_t)(k >> 32))) ++cnt;
    }
    int incl = cnt;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if ((tid & 31) >= o) incl += t; }
    if ((tid & 31) == 31) sWarpTot[tid >> 5] = incl;
    __syncthreads();
    if (tid < 32) {
        int v = sWarpTot[tid], iv = v;
        #pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, iv, o); if (tid >= o) iv += t; }
        sWarpTot[tid] = iv - v;
        if (tid == 31) sTotal = iv;
    }
    __syncthreads();
    int pos = sWarpTot[tid >> 5] + incl - cnt;
    for (int i = i0; i < i0 + per && i < npow2; ++i) {
        const unsigned long long k = keys[i];
        if (k == ~0ull) break;
        const uint32_t w = (uint32_t)(k >> 32);
        if (i == 0 || (uint32_t)(keys[i - 1] >> 32) != w) {
            const double ww = wordWeight[w];
            double sum = ww;                                 // first addWeight inserts, the others add (feature order)
            for (int j = i + 1; j < npow2 && (uint32_t)(keys[j] >> 32) == w && keys[j] != ~0ull; ++j) sum += ww;
            ids[pos] = (int32_t)w; values[pos] = sum;
            ++pos;
        }
    }
    __syncthreads();
    const int nw = sTotal;
    double* sums = reinterpret_cast<double*>(keys);          // the sort buffer is free now
    for (int p = tid; p < nw; p += kBowThreads) sums[p] = values[p];
    __syncthreads();
    if (tid == 0) {
        double norm = 0.0;
        for (int p = 0; p < nw; ++p) norm += fabs(sums[p]);  // BowVector::normalize(L1): ascending word id
        sNorm = norm;
        *nOut = nw;
    }
    __syncthreads();
    const double norm = sNorm;
    if (norm > 0.0) for (int p = tid; p < nw; p += kBowThreads) values[p] = sums[p] / norm;
}

// One thread per database vector: L1Scoring::score(query, vector).  The reference's lower_bound jumps visit the common
// word ids in ascending order, which is what the linear merge does.
__global__ void __launch_bounds__(128)
bowScoreL1Kernel(const int32_t* __restrict__ qIds, const double* __restrict__ qVal, int nq, const int32_t* __restrict__ ids,
                 const double* __restrict__ vals, const int32_t* __restrict__ seg, int nvec, double* __restrict__ score)
{
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nvec) return;
    int a = 0, b = seg[v];
    const int bEnd = seg[v + 1];
    double s = 0.0;
    while (a < nq && b < bEnd) {
        const int ia = qIds[a], ib = ids[b];
        if (ia == ib) {
            const double vi = qVal[a], wi = vals[b];
            s += fabs(vi - wi) - fabs(vi) - fabs(wi);
            ++a; ++b;
        } else if (ia < ib) ++a; else ++b;
    }
    score[v] = -s / 2.0;
}

} // namespace orbx
using namespace orbx;

extern "C" int orbx_vocab_create(orbx_ctx* h, int k, int L, int nnodes, const int32_t* child_off, const int32_t* child,
                                 const uint8_t* node_desc, const int32_t* word_id, const float* weight, orbx_vocab** out)
{
    if (!h || !out || k < 1 || L < 1 || nnodes < 1 || !child_off || !child || !node_desc || !word_id || !weight) return ORBX_E_INVALID;
    *out = nullptr;
    Ctx* c = &h->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    orbx_vocab* v = new (std::nothrow) orbx_vocab();
    if (!v) return ORBX_E_NOMEM;
    v->device = c->device; v->k = k; v->L = L; v->nnodes = nnodes;
    int nw = 0;
    for (int i = 0; i < nnodes; ++i) if (child_off[i + 1] == child_off[i]) ++nw;
    v->nwords = nw;
    const int nchild = child_off[nnodes];
    cudaError_t e = cudaSuccess;
    auto up = [&](void** d, const void* s, size_t bytes) {
        if (e != cudaSuccess) return;
        e = cudaMalloc(d, bytes ? bytes : 16);
        if (e == cudaSuccess && bytes) e = cudaMemcpy(*d, s, bytes, cudaMemcpyHostToDevice);
    };
    up((void**)&v->dChildOff, child_off, sizeof(int32_t) * ((size_t)nnodes + 1));
    up((void**)&v->dChild, child, sizeof(int32_t) * (size_t)nchild);
    up((void**)&v->dDesc, node_desc, (size_t)32 * nnodes);
    up((void**)&v->dWord, word_id, sizeof(int32_t) * (size_t)nnodes);
    up((void**)&v->dWeight, weight, sizeof(float) * (size_t)nnodes);
    {
        int cap = 0;
        for (int i = 0; i < nnodes; ++i) if (child_off[i + 1] == child_off[i] && word_id[i] >= cap) cap = word_id[i] + 1;
        std::vector<double> ww((size_t)std::max(cap, 1), 0.0);
        for (int i = 0; i < nnodes; ++i) if (child_off[i + 1] == child_off[i] && word_id[i] >= 0) ww[word_id[i]] = (double)weight[i];
        v->wordCap = cap;
        up((void**)&v->dWordWeight, ww.data(), sizeof(double) * ww.size());
    }
    if (e != cudaSuccess) { orbx_vocab_destroy(v); return cudaFail(c, e, "orbx_vocab_create"); }
    *out = v;
    return ORBX_OK;
}

extern "C" void orbx_vocab_destroy(orbx_vocab* v)
{
    if (!v) return;
    cudaSetDevice(v->device);
    void* p[] = {v->dChildOff, v->dChild, v->dDesc, v->dWord, v->dWeight, v->dWordWeight};
    for (void* q : p) if (q) cudaFree(q);
    delete v;
}

extern "C" int orbx_vocab_info(const orbx_vocab* v, int* k, int* L, int* nnodes, int* nwords)
{
    if (!v) return ORBX_E_INVALID;
    if (k) *k = v->k;
    if (L) *L = v->L;
    if (nnodes) *nnodes = v->nnodes;
    if (nwords) *nwords = v->nwords;
    return ORBX_OK;
}

// ORBvoc text format: header "k L scoring weighting", then one line per node (creation order, ids 1..):
// "parent isLeaf b0 ... b31 weight" (TemplatedVocabulary.h:1385-1487; FORB::fromString FORB.cpp:120-135).
// One pass over the file buffer (strtol-style scanning; a 170 MB ORBvoc-sized file parses in about a second, the reference's
// per-line stringstreams take ten).  Format, TemplatedVocabulary.h:1385-1487: "k L scoring weighting", then one line per
// node in id order: "parent isLeaf d0 ... d31 weight".
namespace {
inline const char* skipBlank(const char* p, const char* e) { while (p < e && (*p == ' ' || *p == '\t' || *p == '\r')) ++p; return p; }
inline bool scanInt(const char*& p, const char* e, long& v)
{
    p = skipBlank(p, e);
    if (p >= e) return false;
    bool neg = false;
    if (*p == '-' || *p == '+') { neg = *p == '-'; ++p; }
    if (p >= e || *p < '0' || *p > '9') return false;
    long r = 0;
    while (p < e && *p >= '0' && *p <= '9') { r = r * 10 + (*p - '0'); ++p; }
    v = neg ? -r : r;
    return true;
}
}

extern "C" int orbx_vocab_load_text(orbx_ctx* h, const char* path, orbx_vocab** out)
{
    if (!h || !path || !out) return ORBX_E_INVALID;
    std::string buf;
    {
        FILE* f = fopen(path, "rb");
        if (!f) return ORBX_E_INVALID;
        fseek(f, 0, SEEK_END);
        const long sz = ftell(f);
        fseek(f, 0, SEEK_SET);
        if (sz <= 0) { fclose(f); return ORBX_E_INVALID; }
        buf.resize((size_t)sz);
        const size_t got = fread(&buf[0], 1, (size_t)sz, f);
        fclose(f);
        if (got != (size_t)sz) return ORBX_E_INVALID;
    }
    const char* p = buf.data(); const char* const e = p + buf.size();
    long k = 0, L = 0, n1 = 0, n2 = 0;
    if (!scanInt(p, e, k) || !scanInt(p, e, L) || !scanInt(p, e, n1) || !scanInt(p, e, n2)) return ORBX_E_INVALID;
    if (k < 1 || k > 20 || L < 1 || L > 10 || n1 < 0 || n1 > 5 || n2 < 0 || n2 > 3) return ORBX_E_INVALID; // same gate as :1409
    while (p < e && *p != '\n') ++p;
    size_t expect = 1, pw = 1;
    for (long l = 0; l < L; ++l) { pw *= (size_t)k; expect += pw; }
    std::vector<int32_t> parent(1, -1), leaf(1, 0);
    std::vector<uint8_t> desc(32, 0);
    std::vector<float> weight(1, 0.f);
    std::vector<double> weight64(1, 0.0);          // WordValue is a double in DBoW2: the BowVector uses the parsed value
    parent.reserve(expect); leaf.reserve(expect); desc.reserve(32 * expect); weight.reserve(expect); weight64.reserve(expect);
    while (p < e) {
        if (*p == '\n') { ++p; continue; }
        p = skipBlank(p, e);
        if (p >= e) break;
        if (*p == '\n') continue;                  // an empty line: the reference's trailing phantom node is never visited
        long pid = 0, isLeaf = 0;
        if (!scanInt(p, e, pid) || !scanInt(p, e, isLeaf)) return ORBX_E_INVALID;
        if (pid < 0 || pid >= (long)parent.size()) return ORBX_E_INVALID;
        parent.push_back((int32_t)pid); leaf.push_back((int32_t)isLeaf);
        for (int i = 0; i < 32; ++i) { long b = 0; if (!scanInt(p, e, b)) return ORBX_E_INVALID; desc.push_back((uint8_t)b); }
        p = skipBlank(p, e);
        char* endw = nullptr;
        const double w = (p < e && *p != '\n') ? strtod(p, &endw) : 0.0;   // buf is NUL-terminated (std::string)
        if (endw) p = endw;
        weight.push_back((float)w); weight64.push_back(w);
        while (p < e && *p != '\n') ++p;
    }
    const int nn = (int)parent.size();
    std::vector<int32_t> off((size_t)nn + 1, 0), child((size_t)(nn > 1 ? nn - 1 : 0)), word((size_t)nn, 0);
    for (int i = 1; i < nn; ++i) off[parent[i] + 1]++;
    for (int i = 0; i < nn; ++i) off[i + 1] += off[i];
    std::vector<int32_t> fill(off.begin(), off.end() - 1);
    for (int i = 1; i < nn; ++i) child[fill[parent[i]]++] = i; // ascending id = push_back order (:1448)
    int wid = 0;
    for (int i = 1; i < nn; ++i) if (leaf[i] > 0) word[i] = wid++; // :1467-1473
    int rc = orbx_vocab_create(h, (int)k, (int)L, nn, off.data(), child.data(), desc.data(), word.data(), weight.data(), out);
    if (rc) return rc;
    (*out)->scoring = (int)n1; (*out)->weighting = (int)n2;
    std::vector<double> ww((size_t)std::max(wid, 1), 0.0);
    for (int i = 1; i < nn; ++i) if (leaf[i] > 0) ww[word[i]] = weight64[i];
    return orbx_vocab_set_word_weights(*out, ww.data(), wid);
}

// TemplatedVocabulary::saveToTextFile (TemplatedVocabulary.h:1492-1517): header "k L  scoring weighting" (the reference writes
// two blanks there), then one line per node 1..N-1 in id order: parent, isLeaf (= no children), the 32 descriptor bytes each
// followed by a blank, one more blank and the weight as ostream prints a double (6 significant digits = "%g").
extern "C" int orbx_vocab_save_text(const orbx_vocab* v, const char* path)
{
    if (!v || !path) return ORBX_E_INVALID;
    if (cudaSetDevice(v->device) != cudaSuccess) return ORBX_E_CUDA;
    const int nn = v->nnodes;
    std::vector<int32_t> off((size_t)nn + 1), child((size_t)std::max(nn - 1, 1)), word((size_t)nn);
    std::vector<uint8_t> desc((size_t)32 * nn);
    std::vector<float> weight((size_t)nn);
    std::vector<double> ww((size_t)std::max(v->wordCap, 1), 0.0);
    if (cudaMemcpy(off.data(), v->dChildOff, sizeof(int32_t) * (size_t)(nn + 1), cudaMemcpyDeviceToHost) != cudaSuccess ||
        (nn > 1 && cudaMemcpy(child.data(), v->dChild, sizeof(int32_t) * (size_t)(nn - 1), cudaMemcpyDeviceToHost) != cudaSuccess) ||
        cudaMemcpy(word.data(), v->dWord, sizeof(int32_t) * (size_t)nn, cudaMemcpyDeviceToHost) != cudaSuccess ||
        cudaMemcpy(desc.data(), v->dDesc, (size_t)32 * nn, cudaMemcpyDeviceToHost) != cudaSuccess ||
        cudaMemcpy(weight.data(), v->dWeight, sizeof(float) * (size_t)nn, cudaMemcpyDeviceToHost) != cudaSuccess ||
        (v->dWordWeight && v->wordCap > 0 && cudaMemcpy(ww.data(), v->dWordWeight, sizeof(double) * (size_t)v->wordCap, cudaMemcpyDeviceToHost) != cudaSuccess))
        return ORBX_E_CUDA;
    std::vector<int32_t> parent((size_t)nn, 0);
    for (int i = 0; i < nn; ++i) for (int j = off[i]; j < off[i + 1]; ++j) parent[child[j]] = i;
    FILE* f = fopen(path, "w");
    if (!f) return ORBX_E_INVALID;
    fprintf(f, "%d %d  %d %d\n", v->k, v->L, v->scoring, v->weighting);
    std::string line;
    char num[64];
    for (int i = 1; i < nn; ++i) {
        const bool leaf = off[i + 1] == off[i];
        line.clear();
        snprintf(num, sizeof(num), "%d %d ", parent[i], leaf ? 1 : 0); line += num;
        for (int b = 0; b < 32; ++b) { snprintf(num, sizeof(num), "%d ", (int)desc[(size_t)32 * i + b]); line += num; }
        const double w = (leaf && v->dWordWeight && word[i] >= 0 && word[i] < v->wordCap) ? ww[word[i]] : (double)weight[i];
        snprintf(num, sizeof(num), " %g\n", w); line += num;
        fputs(line.c_str(), f);
    }
    return fclose(f) == 0 ? ORBX_OK : ORBX_E_INVALID;
}

extern "C" int orbx_vocab_modes(const orbx_vocab* v, int* scoring, int* weighting)
{
    if (!v) return ORBX_E_INVALID;
    if (scoring) *scoring = v->scoring;
    if (weighting) *weighting = v->weighting;
    return ORBX_OK;
}

extern "C" int orbx_vocab_set_word_weights(orbx_vocab* v, const double* weights, int nwords)
{
    if (!v || !weights || nwords != v->wordCap) return ORBX_E_INVALID;
    if (cudaSetDevice(v->device) != cudaSuccess) return ORBX_E_CUDA;
    if (nwords > 0 && cudaMemcpy(v->dWordWeight, weights, sizeof(double) * (size_t)nwords, cudaMemcpyHostToDevice) != cudaSuccess) return ORBX_E_CUDA;
    return ORBX_OK;
}

// word: the word id of every feature (orbx_bow_transform); ids / values receive the BowVector (ascending word id).
extern "C" int orbx_bow_vector(orbx_ctx* h, const orbx_vocab* v, const int32_t* word, int n, int flags, int32_t* ids, double* values,
                               int* nwords)
{
    if (!h || !v || n < 0 || !nwords || (n > 0 && (!word || !ids || !values))) return ORBX_E_INVALID;
    *nwords = 0;
    // TemplatedVocabulary::transform accumulates with addWeight only for TF_IDF (0) / TF (1) weighting (IDF / BINARY use
    // addIfNotExist, :1166-1221) and this entry point normalises with the L1 norm of L1_NORM scoring (0): anything else
    // would be silently wrong, so it is refused.
    if (v->scoring != 0 || v->weighting > 1) return ORBX_E_UNSUPPORTED;
    if (n == 0) return ORBX_OK;
    Ctx* c = &h->c;
    if (c->device != v->device) return ORBX_E_INVALID;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    int npow2 = 32;
    while (npow2 < n) npow2 <<= 1;
    const size_t smem = sizeof(unsigned long long) * (size_t)npow2;
    if (smem > 200 * 1024) return ORBX_E_CAPACITY;              // at most 16 384 features per call (power-of-two sort buffer)
    const bool inDev = flags & ORBX_IN_DEVICE, outDev = flags & ORBX_OUT_DEVICE;
    const size_t oW = 0, oI = (sizeof(int32_t) * (size_t)n + 255) & ~(size_t)255, oV = oI + ((sizeof(int32_t) * (size_t)n + 255) & ~(size_t)255),
                 oN = oV + sizeof(double) * (size_t)n;
    int rc = ensureScratch(c, oN + 256);
    if (rc) return rc;
    rc = ensureMatchPinned(c, sizeof(int));
    if (rc) return rc;
    uint8_t* S = (uint8_t*)c->dScratch;
    const int32_t* dW = word;
    if (!inDev) { ORBX_CUDA(c, cudaMemcpyAsync(S + oW, word, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, c->stream)); dW = (const int32_t*)(S + oW); }
    int32_t* dI = outDev ? ids : (int32_t*)(S + oI);
    double* dV = outDev ? values : (double*)(S + oV);
    if (smem > 48 * 1024 && smem > c->smemCfg[SMEM_BOW_VECTOR]) {
        ORBX_CUDA(c, cudaFuncSetAttribute(bowVectorKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        c->smemCfg[SMEM_BOW_VECTOR] = smem;
    }
    stageBegin(c, ST_BOW);
    bowVectorKernel<<<1, kBowThreads, smem, c->stream>>>(dW, n, npow2, v->dWordWeight, v->wordCap, dI, dV, (int*)(S + oN));
    stageEnd(c, ST_BOW, 1);
    ORBX_CUDA(c, cudaGetLastError());
    int* HN = (int*)c->hMatch;
    ORBX_CUDA(c, cudaMemcpyAsync(HN, S + oN, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    *nwords = *HN;
    if (!outDev && *nwords > 0) {
        ORBX_CUDA(c, cudaMemcpyAsync(ids, dI, sizeof(int32_t) * (size_t)*nwords, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaMemcpyAsync(values, dV, sizeof(double) * (size_t)*nwords, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return ORBX_OK;
}

// L1Scoring::score of one query BowVector against nvec database vectors (ids / vals back to back, seg = HOST offsets).
extern "C" int orbx_bow_score_l1(orbx_ctx* h, const int32_t* q_ids, const double* q_vals, int nq, const int32_t* ids, const double* vals,
                                 const int32_t* seg, int nvec, int flags, double* score)
{
    if (!h || nq < 0 || nvec < 0 || (nq > 0 && (!q_ids || !q_vals)) || (nvec > 0 && (!seg || !score))) return ORBX_E_INVALID;
    if (nvec == 0) return ORBX_OK;
    const int total = seg[nvec];
    if (total < 0 || (total > 0 && (!ids || !vals))) return ORBX_E_INVALID;
    Ctx* c = &h->c;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    const bool inDev = flags & ORBX_IN_DEVICE, outDev = flags & ORBX_OUT_DEVICE;
    auto al = [](size_t b) { return (b + 255) & ~(size_t)255; };
    const size_t oQV = 0, oV = oQV + al(sizeof(double) * (size_t)std::max(nq, 1)), oS = oV + al(sizeof(double) * (size_t)std::max(total, 1)),
                 oQI = oS + al(sizeof(double) * (size_t)nvec), oI = oQI + al(sizeof(int32_t) * (size_t)std::max(nq, 1)),
                 oSeg = oI + al(sizeof(int32_t) * (size_t)std::max(total, 1)), end = oSeg + al(sizeof(int32_t) * (size_t)(nvec + 1));
    int rc = ensureScratch(c, end);
    if (rc) return rc;
    uint8_t* S = (uint8_t*)c->dScratch;
    const int32_t *dQI = q_ids, *dI = ids; const double *dQV = q_vals, *dV = vals;
    if (!inDev) {
        if (nq) { ORBX_CUDA(c, cudaMemcpyAsync(S + oQI, q_ids, sizeof(int32_t) * (size_t)nq, cudaMemcpyHostToDevice, c->stream));
                  ORBX_CUDA(c, cudaMemcpyAsync(S + oQV, q_vals, sizeof(double) * (size_t)nq, cudaMemcpyHostToDevice, c->stream)); }
        if (total) { ORBX_CUDA(c, cudaMemcpyAsync(S + oI, ids, sizeof(int32_t) * (size_t)total, cudaMemcpyHostToDevice, c->stream));
                     ORBX_CUDA(c, cudaMemcpyAsync(S + oV, vals, sizeof(double) * (size_t)total, cudaMemcpyHostToDevice, c->stream)); }
        dQI = (const int32_t*)(S + oQI); dQV = (const double*)(S + oQV); dI = (const int32_t*)(S + oI); dV = (const double*)(S + oV);
    }
    ORBX_CUDA(c, cudaMemcpyAsync(S + oSeg, seg, sizeof(int32_t) * (size_t)(nvec + 1), cudaMemcpyHostToDevice, c->stream));
    double* dS = outDev ? score : (double*)(S + oS);
    stageBegin(c, ST_BOW);
    bowScoreL1Kernel<<<(nvec + 127) / 128, 128, 0, c->stream>>>(dQI, dQV, nq, dI, dV, (const int32_t*)(S + oSeg), nvec, dS);
    stageEnd(c, ST_BOW, 1);
    ORBX_CUDA(c, cudaGetLastError());
    if (!outDev) ORBX_CUDA(c, cudaMemcpyAsync(score, dS, sizeof(double) * (size_t)nvec, cudaMemcpyDeviceToHost, c->stream));
    ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    return ORBX_OK;
}

extern "C" int orbx_bow_transform(orbx_ctx* h, const orbx_vocab* v, const uint8_t* desc, int n, int levelsup, int flags,
                                  int32_t* word, float* weight, int32_t* node)
{
    if (!h || !v || n < 0 || (n > 0 && (!desc || !word || !weight || !node))) return ORBX_E_INVALID;
    if (n == 0) return ORBX_OK;
    Ctx* c = &h->c;
    if (c->device != v->device) return ORBX_E_INVALID;
    ORBX_CUDA(c, cudaSetDevice(c->device));
    const bool inDev = flags & ORBX_IN_DEVICE, outDev = flags & ORBX_OUT_DEVICE;
    const uint4* dD = (const uint4*)desc; int32_t* dW = word; float* dWt = weight; int32_t* dN = node;
    if (!inDev || !outDev) {
        int rc = ensureScratch(c, (size_t)n * (32 + 12) + 1024);
        if (rc) return rc;
        uint8_t* S = (uint8_t*)c->dScratch;
        if (!inDev) { ORBX_CUDA(c, cudaMemcpyAsync(S, desc, (size_t)n * 32, cudaMemcpyHostToDevice, c->stream)); dD = (const uint4*)S; }
        if (!outDev) { dW = (int32_t*)(S + (size_t)n * 32); dWt = (float*)(dW + n); dN = (int32_t*)(dWt + n); }
    }
    stageBegin(c, ST_BOW);
    bowTransformKernel<<<(n + 7) / 8, 256, 0, c->stream>>>(v->dChildOff, v->dChild, v->dDesc, v->dWord, v->dWeight, dD, n,
                                                          v->L - levelsup, dW, dWt, dN);
    stageEnd(c, ST_BOW, 1);
    ORBX_CUDA(c, cudaGetLastError());
    if (!outDev) {
        ORBX_CUDA(c, cudaMemcpyAsync(word, dW, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaMemcpyAsync(weight, dWt, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaMemcpyAsync(node, dN, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
        ORBX_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return ORBX_OK;
}

// FeatureVector::addFeature for all kept features: std::map<NodeId, vector<unsigned>> flattened to CSR.
extern "C" int orbx_feature_vector(const int32_t* node, const float* weight, int n, int32_t* node_id, int32_t* node_off,
                                   int32_t* feat, int* nnodes)
{
    if (n < 0 || !nnodes || (n > 0 && (!node || !node_id || !node_off || !feat))) return ORBX_E_INVALID;
    std::vector<int32_t> order;
    order.reserve((size_t)n);
    for (int i = 0; i < n; ++i) if (!weight || weight[i] > 0) order.push_back(i);
    std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return node[a] < node[b]; });
    int nn = 0;
    for (size_t i = 0; i < order.size(); ++i) {
        if (i == 0 || node[order[i]] != node[order[i - 1]]) { node_id[nn] = node[order[i]]; node_off[nn] = (int32_t)i; ++nn; }
        feat[i] = order[i];
    }
    if (node_off) node_off[nn] = (int32_t)order.size();
    *nnodes = nn;
    return ORBX_OK;
}
