// K7 -- DBoW2 tree descent (TemplatedVocabulary::transform, Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1255-1306)
// with FORB::distance (FORB.cpp:81-101) = the same 256-bit XOR+POPC as K6.
//
// One warp per feature: lane c evaluates child c of the current node (k <= 32), the warp takes the
// (distance, child position) minimum -- strict '<' in the reference means the FIRST child wins ties -- and
// descends; L sequential steps.  The node table (k=10, L=6: 1.11 M nodes x 32 B = 35 MB) is L2-resident on
// B200 (126 MB), so the 2 x 16-byte gathers per lane are L2 hits after the first frames.
#include "orbx_internal.h"
#include <stdio.h>
#include <string.h>
#include <algorithm>
#include <fstream>
#include <sstream>
#include <new>

struct orbx_vocab {
    int device = 0, k = 0, L = 0, nnodes = 0, nwords = 0;
    int32_t* dChildOff = nullptr; int32_t* dChild = nullptr; uint4* dDesc = nullptr; int32_t* dWord = nullptr; float* dWeight = nullptr;
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
    if (e != cudaSuccess) { orbx_vocab_destroy(v); return cudaFail(c, e, "orbx_vocab_create"); }
    *out = v;
    return ORBX_OK;
}

extern "C" void orbx_vocab_destroy(orbx_vocab* v)
{
    if (!v) return;
    cudaSetDevice(v->device);
    void* p[] = {v->dChildOff, v->dChild, v->dDesc, v->dWord, v->dWeight};
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
extern "C" int orbx_vocab_load_text(orbx_ctx* h, const char* path, orbx_vocab** out)
{
    if (!h || !path || !out) return ORBX_E_INVALID;
    std::ifstream f(path);
    if (!f.good()) return ORBX_E_INVALID;
    std::string line;
    if (!std::getline(f, line)) return ORBX_E_INVALID;
    int k = 0, L = 0, n1 = 0, n2 = 0;
    { std::stringstream ss(line); ss >> k >> L >> n1 >> n2; }
    if (k < 1 || k > 20 || L < 1 || L > 10 || n1 < 0 || n1 > 5 || n2 < 0 || n2 > 3) return ORBX_E_INVALID; // same gate as :1409
    std::vector<int32_t> parent(1, -1), leaf(1, 0);
    std::vector<uint8_t> desc(32, 0);
    std::vector<float> weight(1, 0.f);
    while (std::getline(f, line)) {
        if (line.find_first_not_of(" \t\r\n") == std::string::npos) continue; // the reference's trailing phantom node is never visited
        std::stringstream ss(line);
        int pid = 0, isLeaf = 0;
        ss >> pid >> isLeaf;
        if (pid < 0 || pid >= (int)parent.size()) return ORBX_E_INVALID;
        parent.push_back(pid); leaf.push_back(isLeaf);
        for (int i = 0; i < 32; ++i) { int b = 0; ss >> b; desc.push_back((uint8_t)b); }
        double w = 0; ss >> w; weight.push_back((float)w);
    }
    const int nn = (int)parent.size();
    std::vector<int32_t> off((size_t)nn + 1, 0), child((size_t)(nn > 1 ? nn - 1 : 0)), word((size_t)nn, 0);
    for (int i = 1; i < nn; ++i) off[parent[i] + 1]++;
    for (int i = 0; i < nn; ++i) off[i + 1] += off[i];
    std::vector<int32_t> fill(off.begin(), off.end() - 1);
    for (int i = 1; i < nn; ++i) child[fill[parent[i]]++] = i; // ascending id = push_back order (:1448)
    int wid = 0;
    for (int i = 1; i < nn; ++i) if (leaf[i] > 0) word[i] = wid++; // :1467-1473
    return orbx_vocab_create(h, k, L, nn, off.data(), child.data(), desc.data(), word.data(), weight.data(), out);
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
