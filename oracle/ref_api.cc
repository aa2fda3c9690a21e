// TEST INFRASTRUCTURE ONLY (oracle/).  Flat C entry points over the UNMODIFIED reference classes, compiled by
// `make -C oracle ref` from the sources where they lie under /root/reference into oracle/_ref/liborb_ref.so:
//   src/ORBextractor.cc, src/ORBmatcher.cc, src/Frame.cc, src/KeyFrame.cc, src/MapPoint.cc, src/Map.cc,
//   src/KeyFrameDatabase.cc, Thirdparty/DBoW2/DBoW2/{FORB,FeatureVector,BowVector,ScoringObject}.cpp, DUtils/Random.cpp
// against oracle/cvshim (a cv:: namespace whose image primitives are the cv2-pinned orc_* restatements).
// This file only builds the reference's own objects (Frame, KeyFrame, MapPoint, Map, ORBVocabulary) from flat arrays,
// calls the reference's own methods and copies the results out; it contains no restatement of any algorithm.
// Used by tests/ (oracle == reference, GPU == reference) and by bench.py's cpu_baseline / --impl reference legs.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <list>
#include <map>
#include <mutex>
#include <set>
#include <sstream>
#include <string>
#include <thread>
#include <vector>
#include <opencv2/opencv.hpp>

// The reference keeps the members this file has to fill (MapPoint::mDescriptor, KeyFrame::mvpMapPoints,
// Frame::AssignFeaturesToGrid, ORBextractor::DistributeOctTree ...) private or protected; the access keywords are
// widened for THIS translation unit only (the reference's own .cc files are compiled untouched).
#define private public
#define protected public
#include "ORBextractor.h"
#include "ORBmatcher.h"
#include "Frame.h"
#include "KeyFrame.h"
#include "MapPoint.h"
#include "Map.h"
#include "KeyFrameDatabase.h"
#include "ORBVocabulary.h"
#undef private
#undef protected

using namespace ORB_SLAM2;

// ------------------------------------------------------------------------------------------------ allocator model
// ORBextractor::DistributeOctTree sorts (size, ExtractorNode*) pairs (src/ORBextractor.cc:935), so among equal-size nodes
// the split order follows the ADDRESSES of the std::list nodes: the reference's output depends on the allocator and on the
// heap's history.  Two models are selectable per call (ref_set_allocator):
//   0  the process's real heap (glibc malloc: tcache/fastbin LIFO reuse of freed nodes) -- what a real run does;
//   1  monotonic: every small allocation (<= 256 B) gets the next higher address and nothing is reused, so address
//      order == creation order.  This is the canonical rule oracle/orb_extract.c and the GPU kernel implement
//      ("among equal sizes the most recently created node splits first").
// Nothing in the reference sources changes; only where operator new finds its memory does.
#include <atomic>
#include <new>
#include <sys/mman.h>
namespace {
const size_t kArenaBytes = (size_t)4 << 30;    // address space only (MAP_NORESERVE); reset between calls
char* g_arena = NULL;
std::atomic<size_t> g_arena_off(0);
std::atomic<int> g_alloc_mode(0);     // what operator new does right now
std::atomic<int> g_alloc_model(0);    // the model selected for extraction calls
inline bool inArena(const void* p) { return g_arena && (const char*)p >= g_arena && (const char*)p < g_arena + kArenaBytes; }
inline void* refAlloc(size_t n)
{
    if (g_alloc_mode.load(std::memory_order_relaxed) == 1 && n <= 256 && g_arena) {
        const size_t sz = (n + 15) & ~(size_t)15;
        const size_t o = g_arena_off.fetch_add(sz, std::memory_order_relaxed);
        if (o + sz <= kArenaBytes) return g_arena + o;
    }
    void* p = std::malloc(n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
}
inline void refFree(void* p) { if (p && !inArena(p)) std::free(p); }
}
void* operator new(size_t n) { return refAlloc(n); }
void* operator new[](size_t n) { return refAlloc(n); }
void operator delete(void* p) noexcept { refFree(p); }
void operator delete[](void* p) noexcept { refFree(p); }
void operator delete(void* p, size_t) noexcept { refFree(p); }
void operator delete[](void* p, size_t) noexcept { refFree(p); }

extern "C" int ref_set_allocator(int mode)
{   // call between API calls only (no reference object alive): switches the model and rewinds the arena
    if (!g_arena) {
        void* m = mmap(NULL, kArenaBytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        if (m == MAP_FAILED) return -1;
        g_arena = (char*)m;
    }
    if (g_arena_off.load() > ((size_t)1 << 30)) madvise(g_arena, kArenaBytes, MADV_DONTNEED);   // give the pages back
    g_arena_off.store(0);
    g_alloc_model.store(mode);
    return 0;
}
namespace {
// The model is in force only inside the extraction entry points: there every small block is allocated and freed by code
// instantiated in this library.  Elsewhere (vocabulary loading, matcher scenes) libstdc++.so itself frees blocks that
// this library allocated (std::getline growing a string), so those calls always use the real heap.
struct ExtractionScope {
    ExtractionScope() { g_alloc_mode.store(g_alloc_model.load()); }
    ~ExtractionScope() { g_alloc_mode.store(0); }
};
}

extern "C" {

// ------------------------------------------------------------------------------------------------ extraction
// ORBextractor::operator() (src/ORBextractor.cc:1543-1658).  pyr (optional): the un-bordered levels back to back.
// pyr_bordered (optional): the bordered levels (w+38)x(h+38) back to back, i.e. the parents of mvImagePyramid[l].
int ref_extract(int nfeatures, float scale, int nlevels, int ini_th, int min_th, const uint8_t* img, int w, int h, size_t stride,
                orc_keypoint* kps, uint8_t* desc, int cap, uint8_t* pyr, uint8_t* pyr_bordered)
{
    ExtractionScope scope;
    ORBextractor ex(nfeatures, scale, nlevels, ini_th, min_th);
    cv::Mat image(h, w, CV_8UC1, (void*)img, stride);
    std::vector<cv::KeyPoint> keys;
    cv::Mat descriptors;
    ex(image, cv::Mat(), keys, descriptors);
    const int n = (int)keys.size();
    if (n > cap) return -1;
    static_assert(sizeof(cv::KeyPoint) == sizeof(orc_keypoint), "keypoint layout");
    if (n) {
        std::memcpy(kps, keys.data(), sizeof(orc_keypoint) * (size_t)n);
        for (int i = 0; i < n; ++i) std::memcpy(desc + 32 * (size_t)i, descriptors.ptr(i), 32);
    }
    if (pyr || pyr_bordered) {
        size_t o = 0, ob = 0;
        for (int l = 0; l < nlevels; ++l) {
            const cv::Mat& m = ex.mvImagePyramid[l];
            if (pyr) for (int y = 0; y < m.rows; ++y) { std::memcpy(pyr + o, m.ptr(y), (size_t)m.cols); o += (size_t)m.cols; }
            if (pyr_bordered) {
                const int E = 19;
                for (int y = -E; y < m.rows + E; ++y) { std::memcpy(pyr_bordered + ob, m.data + (ptrdiff_t)y * (ptrdiff_t)m.step.p[0] - E, (size_t)m.cols + 2 * E); ob += (size_t)m.cols + 2 * E; }
            }
        }
    }
    return n;
}

// The six getters Frame reads (src/Frame.cc:128-140) + mnFeaturesPerLevel + umax.
void ref_extractor_tables(int nfeatures, float scale, int nlevels, int ini_th, int min_th, float* sf, float* isf, float* s2, float* is2,
                          int* nfeat, int* umax)
{
    ORBextractor ex(nfeatures, scale, nlevels, ini_th, min_th);
    std::vector<float> a = ex.GetScaleFactors(), b = ex.GetInverseScaleFactors(), c = ex.GetScaleSigmaSquares(), d = ex.GetInverseScaleSigmaSquares();
    for (int i = 0; i < nlevels; ++i) { sf[i] = a[i]; isf[i] = b[i]; s2[i] = c[i]; is2[i] = d[i]; nfeat[i] = ex.mnFeaturesPerLevel[i]; }
    for (int i = 0; i < 16; ++i) umax[i] = ex.umax[i];
}

// ORBextractor::DistributeOctTree (src/ORBextractor.cc:696-1042) with the process's real heap: the sort at :935 orders
// equal-size nodes by the ADDRESS of their std::list node.
int ref_distribute_octree(const orc_keypoint* in, int n, int minX, int maxX, int minY, int maxY, int N, int level, orc_keypoint* out, int cap)
{
    ExtractionScope scope;
    ORBextractor ex(1000, 1.2f, 8, 20, 7);
    std::vector<cv::KeyPoint> v((const cv::KeyPoint*)in, (const cv::KeyPoint*)in + n);
    std::vector<cv::KeyPoint> r = ex.DistributeOctTree(v, minX, maxX, minY, maxY, N, level);
    if ((int)r.size() > cap) return -1;
    if (!r.empty()) std::memcpy(out, r.data(), sizeof(orc_keypoint) * r.size());
    return (int)r.size();
}

// The reference's extraction timed on `threads` host threads, one frame per call per thread, each thread with its own
// ORBextractor (as Tracking owns one per camera): frames/s over `frames` images of w x h taken round-robin from `nimg`.
double ref_extract_throughput(int nfeatures, float scale, int nlevels, int ini_th, int min_th, const uint8_t* imgs, int nimg, int w, int h,
                              int frames, int threads, long long* total_keypoints)
{
    ExtractionScope scope;
    std::vector<std::thread> pool;
    std::vector<long long> cnt(threads, 0);
    const auto t0 = std::chrono::steady_clock::now();
    for (int t = 0; t < threads; ++t)
        pool.emplace_back([&, t]() {
            ORBextractor ex(nfeatures, scale, nlevels, ini_th, min_th);
            std::vector<cv::KeyPoint> keys; cv::Mat d;
            for (int f = t; f < frames; f += threads) {
                cv::Mat image(h, w, CV_8UC1, (void*)(imgs + (size_t)(f % nimg) * w * h), (size_t)w);
                ex(image, cv::Mat(), keys, d);
                cnt[t] += (long long)keys.size();
            }
        });
    for (auto& th : pool) th.join();
    const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    long long tot = 0; for (long long c : cnt) tot += c;
    if (total_keypoints) *total_keypoints = tot;
    return frames / s;
}

// ------------------------------------------------------------------------------------------------ stereo frame
// Frame::Frame(imLeft, imRight, ...) (src/Frame.cc:119-217): both extractions, UndistortKeyPoints, ComputeStereoMatches
// (:831-1128) and AssignFeaturesToGrid, exactly as Tracking::GrabImageStereo builds it.  dist = (k1,k2,p1,p2[,k3]) or NULL.
int ref_stereo_frame(int nfeatures, float scale, int nlevels, int ini_th, int min_th, const uint8_t* imL, const uint8_t* imR, int w, int h,
                     float fx, float fy, float cx, float cy, const float* dist, int ndist, float bf, float th_depth,
                     orc_keypoint* keysL, uint8_t* descL, int* nL, orc_keypoint* keysR, uint8_t* descR, int* nR, int cap,
                     float* uRight, float* depth, orc_keypoint* keysUn)
{
    ExtractionScope scope;
    ORBextractor exL(nfeatures, scale, nlevels, ini_th, min_th), exR(nfeatures, scale, nlevels, ini_th, min_th);
    cv::Mat L(h, w, CV_8UC1, (void*)imL, (size_t)w), R(h, w, CV_8UC1, (void*)imR, (size_t)w);
    cv::Mat K = cv::Mat::eye(3, 3, CV_32F);
    K.at<float>(0, 0) = fx; K.at<float>(1, 1) = fy; K.at<float>(0, 2) = cx; K.at<float>(1, 2) = cy;
    cv::Mat D = cv::Mat::zeros(ndist > 4 ? ndist : 4, 1, CV_32F);
    for (int i = 0; i < ndist; ++i) D.at<float>(i) = dist[i];
    Frame::mbInitialComputations = true;
    Frame F(L, R, 0.0, &exL, &exR, (ORBVocabulary*)NULL, K, D, bf, th_depth);
    *nL = F.N; *nR = (int)F.mvKeysRight.size();
    if (F.N > cap || *nR > cap) return -1;
    for (int i = 0; i < F.N; ++i) {
        std::memcpy(&keysL[i], &F.mvKeys[i], 28); std::memcpy(descL + 32 * (size_t)i, F.mDescriptors.ptr(i), 32);
        if (keysUn) std::memcpy(&keysUn[i], &F.mvKeysUn[i], 28);
        uRight[i] = F.mvuRight[i]; depth[i] = F.mvDepth[i];
    }
    for (int i = 0; i < *nR; ++i) { std::memcpy(&keysR[i], &F.mvKeysRight[i], 28); std::memcpy(descR + 32 * (size_t)i, F.mDescriptorsRight.ptr(i), 32); }
    int n = 0;
    for (int i = 0; i < F.N; ++i) if (F.mvDepth[i] > 0) ++n;
    return n;
}

// ------------------------------------------------------------------------------------------------ vocabulary
// ORBVocabulary::loadFromTextFile (TemplatedVocabulary.h:1385-1487) + transform (:1145-1221, :1255-1306) on the real DBoW2.
void* ref_vocab_load_text(const char* path)
{
    ORBVocabulary* v = new ORBVocabulary();
    if (!v->loadFromTextFile(path)) { delete v; return NULL; }
    return v;
}
void ref_vocab_save_text(void* v, const char* path) { ((ORBVocabulary*)v)->saveToTextFile(path); }   // TemplatedVocabulary.h:1492-1517
void ref_vocab_free(void* v) { delete (ORBVocabulary*)v; }
void ref_vocab_info(void* vv, int* k, int* L, int* nwords, int* nnodes, int* scoring, int* weighting)
{
    ORBVocabulary* v = (ORBVocabulary*)vv;
    *k = v->getBranchingFactor(); *L = v->getDepthLevels(); *nwords = (int)v->size(); *nnodes = (int)v->m_nodes.size();
    *scoring = (int)v->getScoringType(); *weighting = (int)v->getWeightingType();
}
// per feature: word id, word weight, node id at levelsup (the values FeatureVector::addFeature / BowVector::addWeight receive)
void ref_vocab_transform_each(void* vv, const uint8_t* desc, int n, int levelsup, int32_t* word, double* weight, int32_t* node)
{
    ORBVocabulary* v = (ORBVocabulary*)vv;
    for (int i = 0; i < n; ++i) {
        cv::Mat d(1, 32, CV_8U, (void*)(desc + 32 * (size_t)i));
        DBoW2::WordId id; DBoW2::WordValue w; DBoW2::NodeId nid;
        v->transform(d, id, w, &nid, levelsup);
        word[i] = (int32_t)id; weight[i] = w; node[i] = (int32_t)nid;
    }
}
// the full transform Frame::ComputeBoW calls (src/Frame.cc:704-719): BowVector (ids ascending, normalised values) and
// FeatureVector as CSR.  Returns the number of BowVector entries; *nnodes the FeatureVector size.
int ref_vocab_transform(void* vv, const uint8_t* desc, int n, int levelsup, int32_t* bow_ids, double* bow_vals,
                        int32_t* node_id, int32_t* node_off, int32_t* feat, int* nnodes)
{
    ORBVocabulary* v = (ORBVocabulary*)vv;
    std::vector<cv::Mat> f; f.reserve(n);
    for (int i = 0; i < n; ++i) f.push_back(cv::Mat(1, 32, CV_8U, (void*)(desc + 32 * (size_t)i)));
    DBoW2::BowVector bv; DBoW2::FeatureVector fv;
    v->transform(f, bv, fv, levelsup);
    int nb = 0;
    for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it, ++nb) { bow_ids[nb] = (int32_t)it->first; bow_vals[nb] = it->second; }
    int nn = 0, o = 0;
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++nn) {
        node_id[nn] = (int32_t)it->first; node_off[nn] = o;
        for (size_t j = 0; j < it->second.size(); ++j) feat[o++] = (int32_t)it->second[j];
    }
    node_off[nn] = o; *nnodes = nn;
    return nb;
}
double ref_vocab_score(void* vv, const int32_t* ids1, const double* v1, int n1, const int32_t* ids2, const double* v2, int n2)
{
    ORBVocabulary* v = (ORBVocabulary*)vv;
    DBoW2::BowVector a, b;
    for (int i = 0; i < n1; ++i) a.insert(a.end(), std::make_pair((DBoW2::WordId)ids1[i], v1[i]));
    for (int i = 0; i < n2; ++i) b.insert(b.end(), std::make_pair((DBoW2::WordId)ids2[i], v2[i]));
    return v->score(a, b);
}

// The shim's float matrix algebra, exposed so that tests/test_shim_algebra.py can pin it to the executable OpenCV of this
// image (cv2.gemm / cv2.norm, 4.13).  op: 0 = A*B, 1 = A*B + C, 2 = A.t()*B, 3 = -A.t()*B, 4 = A / s (s = C[0]),
// 5 = -A*B (unary minus on a materialised Mat, KeyFrame::SetPose's Ow).  Returns the element count written to out.
int ref_shim_algebra(int op, const float* A, int ar, int ac, const float* B, int br, int bc, const float* C, float* out)
{
    cv::Mat a(ar, ac, CV_32F, (void*)A), b, c, r;
    if (B) b = cv::Mat(br, bc, CV_32F, (void*)B);
    switch (op) {
    case 0: r = a * b; break;
    case 1: c = cv::Mat(ar, bc, CV_32F, (void*)C); r = a * b + c; break;
    case 2: r = a.t() * b; break;
    case 3: r = -a.t() * b; break;
    case 4: r = a / C[0]; break;
    case 5: r = -a * b; break;
    default: return -1;
    }
    for (int y = 0; y < r.rows; ++y) for (int x = 0; x < r.cols; ++x) out[y * r.cols + x] = r.at<float>(y, x);
    return r.rows * r.cols;
}
double ref_shim_norm(const float* a, int n) { return cv::norm(cv::Mat(n, 1, CV_32F, (void*)a)); }
double ref_shim_dot(const float* a, const float* b, int n) { return cv::Mat(n, 1, CV_32F, (void*)a).dot(cv::Mat(n, 1, CV_32F, (void*)b)); }

// CPU baseline of the Hamming core: the reference's inner matching loop (src/ORBmatcher.cc:325-345: best / second-best
// ORBmatcher::DescriptorDistance over the candidates of one vocabulary node; with one node holding everything this is the
// brute-force mode of configs[3]) on `threads` host threads, queries dealt round-robin.  The rows are taken the way the
// reference takes them (`const cv::Mat& d = descriptors.row(idx)` per candidate) and the distance function is the
// reference's own.  Returns descriptor-pair evaluations per second; best_idx / best_dist / second_dist as the loop leaves them.
double ref_hamming_throughput(const uint8_t* q, int nq, const uint8_t* t, int nt, int threads, int32_t* best_idx, int32_t* best_dist,
                              int32_t* second_dist)
{
    cv::Mat Q0(nq, 32, CV_8U, (void*)q), T0(nt, 32, CV_8U, (void*)t);
    std::vector<std::thread> pool;
    const auto t0 = std::chrono::steady_clock::now();
    for (int th = 0; th < threads; ++th)
        pool.emplace_back([&, th]() {
            // every thread owns its descriptor matrices, as every thread of the reference works on its own keyframe pair:
            // row() bumps the matrix's reference count, which threads sharing one cv::Mat would fight over
            const cv::Mat Q = Q0.clone(), T = T0.clone();
            for (int i = th; i < nq; i += threads) {
                const cv::Mat& dKF = Q.row(i);
                int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
                for (int j = 0; j < nt; ++j) {
                    const cv::Mat& dF = T.row(j);
                    const int dist = ORBmatcher::DescriptorDistance(dKF, dF);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = j; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
                if (best_idx) { best_idx[i] = bestIdxF; best_dist[i] = bestDist1; second_dist[i] = bestDist2; }
            }
        });
    for (auto& th : pool) th.join();
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return (double)nq * (double)nt / sec;
}

int ref_descriptor_distance(const uint8_t* a, const uint8_t* b)
{
    cv::Mat A(1, 32, CV_8U, (void*)a), B(1, 32, CV_8U, (void*)b);
    return ORBmatcher::DescriptorDistance(A, B);
}

} // extern "C"

// ------------------------------------------------------------------------------------------------ matcher scenes
// The scene file of host/orbx_cpp_matcher_driver.cc (same arrays, same order, same output records), run through the
// reference's own Frame / KeyFrame / MapPoint / Map objects and ORBmatcher methods.
namespace {

typedef std::vector<std::vector<char> > Arrays;
Arrays readAll(const char* path)
{
    Arrays out;
    FILE* f = fopen(path, "rb");
    if (!f) return out;
    int n;
    while (fread(&n, 4, 1, f) == 1) { std::vector<char> b(n); if (n && fread(b.data(), 1, n, f) != (size_t)n) break; out.push_back(b); }
    fclose(f);
    return out;
}
template <typename T> const T* as(const std::vector<char>& b) { return reinterpret_cast<const T*>(b.data()); }
void put(FILE* f, const void* p, int nbytes) { fwrite(&nbytes, 4, 1, f); fwrite(p, 1, nbytes, f); }
cv::Mat vec3(const float* p) { cv::Mat m(3, 1, CV_32F); std::memcpy(m.data, p, 12); return m; }
// optional 4x4 row-major pose at in[k]; absent or empty -> the given default
cv::Mat pose4(const Arrays& in, size_t k, const cv::Mat& dflt)
{
    if (in.size() <= k || in[k].size() < 64) return dflt;
    cv::Mat m(4, 4, CV_32F); std::memcpy(m.data, in[k].data(), 64); return m;
}
cv::Mat rowDesc(const char* p) { cv::Mat m(1, 32, CV_8U); std::memcpy(m.data, p, 32); return m; }

struct Scene {
    const float* P;
    Map map;
    KeyFrameDatabase* db;
    Frame base[2];
    KeyFrame* kf[2];
    std::vector<MapPoint*> pool[2];
    Scene() : P(NULL), db(NULL) { kf[0] = kf[1] = NULL; }
};

// a Frame holding one side of the scene (keys, descriptors, stereo coordinates, FeatureVector, scale tables, grid)
void fillFrame(Frame& F, const Arrays& in, int s, const float* P, float logsf, float mbf)
{
    const std::vector<char>&d = in[1 + 5 * s], &k = in[2 + 5 * s], &u = in[4 + 5 * s], &nd = in[5 + 5 * s];
    F.N = (int)d.size() / 32;
    F.mDescriptors = cv::Mat(F.N, 32, CV_8U);
    std::memcpy(F.mDescriptors.data, d.data(), d.size());
    F.mvKeys.assign(as<cv::KeyPoint>(k), as<cv::KeyPoint>(k) + F.N);
    F.mvKeysUn = F.mvKeys;
    F.mvuRight.assign(as<float>(u), as<float>(u) + F.N);
    F.mvDepth.assign(F.N, -1.f);
    F.mvpMapPoints.assign(F.N, (MapPoint*)NULL);
    F.mvbOutlier.assign(F.N, false);
    const int* node = as<int>(nd);
    for (int i = 0; i < F.N; ++i) if (node[i] >= 0) F.mFeatVec.addFeature((DBoW2::NodeId)node[i], (unsigned)i);
    F.mnScaleLevels = 8; F.mfScaleFactor = 1.2f; F.mfLogScaleFactor = logsf;
    F.mvScaleFactors.assign(as<float>(in[11]), as<float>(in[11]) + 8);
    F.mvLevelSigma2.assign(as<float>(in[12]), as<float>(in[12]) + 8);
    F.mvInvScaleFactors.resize(8); F.mvInvLevelSigma2.resize(8);
    for (int l = 0; l < 8; ++l) { F.mvInvScaleFactors[l] = 1.0f / F.mvScaleFactors[l]; F.mvInvLevelSigma2[l] = 1.0f / F.mvLevelSigma2[l]; }
    F.mK = cv::Mat::eye(3, 3, CV_32F);
    F.mK.at<float>(0, 0) = P[3]; F.mK.at<float>(1, 1) = P[4]; F.mK.at<float>(0, 2) = P[5]; F.mK.at<float>(1, 2) = P[6];
    F.mDistCoef = cv::Mat::zeros(4, 1, CV_32F);
    F.mbf = mbf; F.mb = mbf / P[3]; F.mThDepth = 0;
    F.mpORBvocabulary = NULL; F.mpORBextractorLeft = F.mpORBextractorRight = NULL; F.mpReferenceKF = NULL;
    F.mTimeStamp = 0; F.mnId = Frame::nNextId++;
    F.SetPose(cv::Mat::eye(4, 4, CV_32F));
    F.AssignFeaturesToGrid();
}

} // namespace

extern "C" int ref_run_matcher_scene(const char* in_path, const char* out_path)
{
    Arrays in = readAll(in_path);
    if (in.size() < 18) return 2;
    const float* P = as<float>(in[0]);
    const float logsf = in.size() > 20 ? as<float>(in[19])[4] : std::log(1.2f);
    const float mbf = in.size() > 24 ? as<float>(in[24])[1] : 40.f;
    Frame::fx = P[3]; Frame::fy = P[4]; Frame::cx = P[5]; Frame::cy = P[6]; Frame::invfx = 1.0f / P[3]; Frame::invfy = 1.0f / P[4];
    Frame::mnMinX = P[10]; Frame::mnMinY = P[11]; Frame::mnMaxX = P[12]; Frame::mnMaxY = P[13];
    Frame::mfGridElementWidthInv = P[8]; Frame::mfGridElementHeightInv = P[9];
    Frame::mbInitialComputations = false;

    Map map;
    KeyFrameDatabase* db = NULL;       // KeyFrame only stores the pointer
    Frame base[2];
    KeyFrame* ks[2];
    std::vector<MapPoint*> pool[2];
    for (int s = 0; s < 2; ++s) {
        fillFrame(base[s], in, s, P, logsf, mbf);
        ks[s] = new KeyFrame(base[s], &map, db);
    }
    KeyFrame &k1 = *ks[0], &k2 = *ks[1];
    const bool full = in.size() > 20, fuse = in.size() > 24;
    const float* W[2] = {full ? as<float>(in[18]) : NULL, fuse ? as<float>(in[21]) : NULL};
    const char* MD[2] = {full ? in[20].data() : NULL, fuse ? in[22].data() : NULL};
    const int* obsN = fuse ? as<int>(in[23]) : NULL;
    // one MapPoint per feature of each side (whether the keyframe holds it is `has`)
    auto makePoint = [&](int s, int i) {
        static const float zero[8] = {0, 0, 1, 0, 0, 0, 0, 1};
        const float* w = W[s] ? W[s] + 8 * i : zero;
        MapPoint* mp = new MapPoint(vec3(w), ks[s], &map);
        mp->mfMaxDistance = w[3]; mp->mfMinDistance = w[4];
        mp->mNormalVector = vec3(w + 5);
        mp->mDescriptor = MD[s] ? rowDesc(MD[s] + 32 * i) : ks[s]->mDescriptors.row(i).clone();
        mp->nObs = 1;
        return mp;
    };
    auto reset = [&]() {     // fresh MapPoints, keyframe <-> point links as in the scene file
        for (int s = 0; s < 2; ++s) {
            const std::vector<char>& h = in[3 + 5 * s];
            pool[s].assign(ks[s]->N, (MapPoint*)NULL);
            for (int i = 0; i < ks[s]->N; ++i) {
                pool[s][i] = makePoint(s, i);
                ks[s]->mvpMapPoints[i] = h[i] ? pool[s][i] : NULL;
                if (h[i]) { pool[s][i]->mObservations[ks[s]] = i; if (obsN) pool[s][i]->nObs = obsN[(s ? k1.N : 0) + i]; }
            }
        }
    };
    reset();
    auto code1 = [&](MapPoint* p, int s) { for (int i = 0; i < ks[s]->N; ++i) if (pool[s][i] == p) return i; return -2; };
    std::map<MapPoint*, int> codeMap;
    auto rebuildCodes = [&]() { codeMap.clear(); for (int i = 0; i < k1.N; ++i) codeMap[pool[0][i]] = i; for (int i = 0; i < k2.N; ++i) codeMap[pool[1][i]] = 100000 + i; };
    auto code = [&](MapPoint* p) { if (!p) return -1; std::map<MapPoint*, int>::iterator it = codeMap.find(p); return it == codeMap.end() ? -2 : it->second; };
    (void)code1;

    // F12 / epipole inputs of SearchForTriangulation: C1 = k1's camera centre, pose of k2 = identity
    cv::Mat F12(3, 3, CV_32F); std::memcpy(F12.data, in[13].data(), 36);
    {   // KeyFrame 1 sits at Ow (rotation identity): Tcw = [I | -Ow]
        cv::Mat T = cv::Mat::eye(4, 4, CV_32F);
        for (int i = 0; i < 3; ++i) T.at<float>(i, 3) = -as<float>(in[14])[i];
        k1.SetPose(T);
    }
    FILE* o = fopen(out_path, "wb");
    if (!o) return 2;
    ORBmatcher m(P[0], P[1] != 0);
    std::vector<int> res;
    // (a) SearchByBoW(KF, F): the Frame is side 2
    Frame F(base[1]);
    std::vector<MapPoint*> mm;
    int n = m.SearchByBoW(&k1, F, mm);
    rebuildCodes();
    res.assign(F.N, -1);
    for (int j = 0; j < F.N; ++j) if (mm[j]) res[j] = code(mm[j]);
    put(o, &n, 4); put(o, res.data(), 4 * F.N);
    // (b) SearchByBoW(KF, KF)
    n = m.SearchByBoW(&k1, &k2, mm);
    res.assign(k1.N, -1);
    for (int i = 0; i < k1.N; ++i) if (mm[i]) res[i] = code(mm[i]) - 100000;
    put(o, &n, 4); put(o, res.data(), 4 * k1.N);
    // (c) SearchForTriangulation
    std::vector<std::pair<size_t, size_t> > pairs;
    n = m.SearchForTriangulation(&k1, &k2, F12, pairs, P[2] != 0);
    res.assign(k1.N, -1);
    for (size_t i = 0; i < pairs.size(); ++i) res[pairs[i].first] = (int)pairs[i].second;
    put(o, &n, 4); put(o, res.data(), 4 * k1.N);
    k1.SetPose(cv::Mat::eye(4, 4, CV_32F));
    // (d) SearchByProjection(Frame, MapPoints, th)
    const int np = (int)in[15].size() / 20;
    std::vector<MapPoint*> vp(np);
    for (int i = 0; i < np; ++i) {
        const float* q = as<float>(in[15]) + 5 * i;
        static const float w0[3] = {0, 0, 1};
        MapPoint* mp = new MapPoint(vec3(w0), &k1, &map);
        mp->nObs = in.size() > 25 ? as<int>(in[25])[i] : 1;      // 0: a temporal point (Tracking::UpdateLastFrame)
        mp->mbTrackInView = true; mp->mTrackProjX = q[0]; mp->mTrackProjY = q[1]; mp->mTrackProjXR = q[2];
        mp->mTrackViewCos = q[3]; mp->mnTrackScaleLevel = (int)q[4];
        mp->mDescriptor = rowDesc(in[16].data() + 32 * i);
        vp[i] = mp;
    }
    static const float w0[3] = {0, 0, 1};
    MapPoint* occupiedMP = new MapPoint(vec3(w0), &k1, &map);
    occupiedMP->nObs = 1;
    F.mvpMapPoints.assign(F.N, (MapPoint*)NULL);
    for (int i = 0; i < F.N; ++i) if (in[17][i]) F.mvpMapPoints[i] = occupiedMP;
    n = ORBmatcher(P[0], true).SearchByProjection(F, vp, P[7]);
    res.assign(F.N, -1);
    for (int i = 0; i < F.N; ++i)
        if (F.mvpMapPoints[i] && F.mvpMapPoints[i] != occupiedMP) res[i] = (int)(std::find(vp.begin(), vp.end(), F.mvpMapPoints[i]) - vp.begin());
    put(o, &n, 4); put(o, res.data(), 4 * F.N);
    int dd = ORBmatcher::DescriptorDistance(k1.mDescriptors.row(0), k2.mDescriptors.row(0));
    put(o, &dd, 4);
    if (full) {
        const float* Q = as<float>(in[19]);
        // (e) SearchByProjection(Frame, KeyFrame, sAlreadyFound, th, ORBdist): relocalisation
        F.SetPose(pose4(in, 31, cv::Mat::eye(4, 4, CV_32F)));
        F.mvpMapPoints.assign(F.N, (MapPoint*)NULL);
        for (int i = 0; i < F.N; ++i) if (in[17][i]) F.mvpMapPoints[i] = occupiedMP;
        std::set<MapPoint*> found;
        for (int i = 0; i < k1.N; i += 17) if (k1.mvpMapPoints[i]) found.insert(k1.mvpMapPoints[i]);
        n = ORBmatcher(P[0], true).SearchByProjection(F, &k1, found, Q[0], (int)Q[1]);
        res.assign(F.N, -1);
        for (int i = 0; i < F.N; ++i) if (F.mvpMapPoints[i] && F.mvpMapPoints[i] != occupiedMP) res[i] = code(F.mvpMapPoints[i]);
        put(o, &n, 4); put(o, res.data(), 4 * F.N);
        // (f) SearchByProjection(KeyFrame, Scw, vpPoints, vpMatched, th): loop closing, KF = side 2
        cv::Mat Scw = cv::Mat::zeros(4, 4, CV_32F);
        for (int i = 0; i < 3; ++i) Scw.at<float>(i, i) = Q[5];
        Scw.at<float>(3, 3) = 1.f;
        Scw = pose4(in, 29, Scw);
        std::vector<MapPoint*> vpPoints, vpMatched(k2.N, (MapPoint*)NULL);
        for (int i = 0; i < k1.N; ++i) if (k1.mvpMapPoints[i]) vpPoints.push_back(k1.mvpMapPoints[i]);
        for (int i = 0; i < k2.N; ++i) if (in[17][i]) vpMatched[i] = occupiedMP;
        n = ORBmatcher(P[0], true).SearchByProjection(&k2, Scw, vpPoints, vpMatched, (int)Q[2]);
        res.assign(k2.N, -1);
        for (int i = 0; i < k2.N; ++i) if (vpMatched[i] && vpMatched[i] != occupiedMP) res[i] = code(vpMatched[i]);
        put(o, &n, 4); put(o, res.data(), 4 * k2.N);
        // (g) SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)
        Frame F1(base[0]);
        std::vector<cv::Point2f> prev(F1.N);
        for (int i = 0; i < F1.N; ++i) prev[i] = F1.mvKeysUn[i].pt;
        std::vector<int> m12;
        n = ORBmatcher(P[0], true).SearchForInitialization(F1, F, prev, m12, (int)Q[3]);
        put(o, &n, 4); put(o, m12.data(), 4 * F1.N); put(o, prev.data(), 8 * F1.N);
    }
    if (fuse) {
        const float* Q2 = as<float>(in[24]);
        // (h) Fuse(KeyFrame, vpMapPoints, th): KF1's points (with NULL gaps and a few duplicates) into KF2
        reset(); rebuildCodes();
        std::vector<MapPoint*> vpf = k1.GetMapPointMatches();
        for (int i = 0; i + 50 < k1.N; i += 50) vpf.push_back(vpf[i]);
        n = ORBmatcher(P[0], true).Fuse(&k2, vpf, Q2[0]);
        res.assign(k2.N, -1);
        { std::vector<MapPoint*> cur = k2.GetMapPointMatches(); for (int i = 0; i < k2.N; ++i) res[i] = code(cur[i]); }
        put(o, &n, 4); put(o, res.data(), 4 * k2.N);
        std::vector<int> st(k1.N + k2.N);
        for (int i = 0; i < k1.N; ++i) st[i] = (pool[0][i]->isBad() ? 1 : 0) | (pool[0][i]->Observations() << 1);
        for (int i = 0; i < k2.N; ++i) st[k1.N + i] = (pool[1][i]->isBad() ? 1 : 0) | (pool[1][i]->Observations() << 1);
        put(o, st.data(), 4 * (int)st.size());
        // (i) Fuse(KeyFrame, Scw, vpPoints, th, vpReplacePoint)
        reset(); rebuildCodes();
        cv::Mat Scw = cv::Mat::zeros(4, 4, CV_32F);
        for (int i = 0; i < 3; ++i) Scw.at<float>(i, i) = 2.f;
        Scw.at<float>(3, 3) = 1.f;
        Scw = pose4(in, 30, Scw);
        std::vector<MapPoint*> vpts;
        { std::vector<MapPoint*> cur = k1.GetMapPointMatches(); for (int i = 0; i < k1.N; ++i) if (cur[i]) vpts.push_back(cur[i]); }
        for (size_t i = 0, nv = vpts.size(); i + 40 < nv; i += 40) vpts.push_back(vpts[i]);
        std::vector<MapPoint*> repl(vpts.size(), (MapPoint*)NULL);
        n = ORBmatcher(P[0], true).Fuse(&k2, Scw, vpts, Q2[2], repl);
        { std::vector<MapPoint*> cur = k2.GetMapPointMatches(); for (int i = 0; i < k2.N; ++i) res[i] = code(cur[i]); }
        put(o, &n, 4); put(o, res.data(), 4 * k2.N);
        std::vector<int> rc(repl.size());
        for (size_t i = 0; i < repl.size(); ++i) rc[i] = code(repl[i]);
        put(o, rc.data(), 4 * (int)rc.size());
        // (j) SearchBySim3(KF1, KF2, vpMatches12, s12 = 1, R12 = I, t12 = 0, th)
        reset(); rebuildCodes();
        std::vector<MapPoint*> m12(k1.N, (MapPoint*)NULL);
        { std::vector<MapPoint*> c1 = k1.GetMapPointMatches(), c2 = k2.GetMapPointMatches();
          for (int i = 0; i < k1.N; i += 23) if (c1[i] && i < k2.N && c2[i]) m12[i] = c2[i]; }
        cv::Mat R12 = cv::Mat::eye(3, 3, CV_32F), t12 = cv::Mat::zeros(3, 1, CV_32F);
        const float s12 = 1.0f;
        n = ORBmatcher(P[0], true).SearchBySim3(&k1, &k2, m12, s12, R12, t12, Q2[3]);
        res.assign(k1.N, -1);
        for (int i = 0; i < k1.N; ++i) res[i] = code(m12[i]);
        put(o, &n, 4); put(o, res.data(), 4 * k1.N);
    }
    if (in.size() > 28) {
        // (k) SearchByProjection(CurrentFrame, LastFrame, th, bMono) (:1741-1910): the last frame is side 1 with its map points
        //     (observation counts in[26]: 0 = temporal point; every 13th keypoint an outlier), the current frame is side 2
        //     at pose in[27], holding the occupied pattern plus temporal points (nObs 0) that must not hide their keypoint
        reset(); rebuildCodes();
        const int* lastObs = as<int>(in[26]);
        const float thK = as<float>(in[28])[0];
        MapPoint* transientMP = new MapPoint(vec3(w0), &k1, &map);
        transientMP->nObs = 0;
        Frame Fl(base[0]);
        Fl.SetPose(cv::Mat::eye(4, 4, CV_32F));
        for (int i = 0; i < Fl.N; ++i) {
            Fl.mvpMapPoints[i] = in[3][i] ? pool[0][i] : (MapPoint*)NULL;
            Fl.mvbOutlier[i] = (i % 13) == 5;
            pool[0][i]->nObs = lastObs[i];
        }
        for (int mono = 0; mono < 2; ++mono) {
            Frame Fc(base[1]);
            Fc.SetPose(pose4(in, 27, cv::Mat::eye(4, 4, CV_32F)));
            for (int i = 0; i < Fc.N; ++i) Fc.mvpMapPoints[i] = in[17][i] ? occupiedMP : ((i % 7) == 3 ? transientMP : (MapPoint*)NULL);
            n = ORBmatcher(P[0], true).SearchByProjection(Fc, Fl, thK, mono != 0);
            res.assign(Fc.N, -1);
            for (int i = 0; i < Fc.N; ++i)
                if (Fc.mvpMapPoints[i] && Fc.mvpMapPoints[i] != occupiedMP && Fc.mvpMapPoints[i] != transientMP) res[i] = code(Fc.mvpMapPoints[i]);
            put(o, &n, 4); put(o, res.data(), 4 * Fc.N);
        }
    }
    fclose(o);
    return 0;   // the scene's objects are left to the process (the Map owns raw pointers exactly as in the reference)
}

// ------------------------------------------------------------------------------------------------ KeyFrameDatabase
// The reference's own KeyFrameDatabase (src/KeyFrameDatabase.cc) over keyframes that carry nothing but an id, a BowVector
// and their covisibility lists.  The vocabulary is a default ORBVocabulary (L1_NORM scoring) whose word table is sized to
// `nwords` -- the database only asks it for size() and score().  mLoopScore / mRelocScore have no initialiser in the
// reference (src/KeyFrame.cc:43-60); they are set to 0 here, the value orb_kfdb.c and the device code start from.
namespace {
struct RefKfdb {
    ORBVocabulary voc;
    KeyFrameDatabase* db;
    Map map;
    Frame base;
    std::map<int, KeyFrame*> kfs;
};
DBoW2::BowVector bowOf(const int32_t* w, const double* v, int n)
{
    DBoW2::BowVector b;
    for (int i = 0; i < n; ++i) b.insert(b.end(), std::make_pair((DBoW2::WordId)w[i], v[i]));
    return b;
}
}

extern "C" {
void* ref_kfdb_create(int nwords)
{
    RefKfdb* r = new RefKfdb();
    r->voc.m_words.resize((size_t)nwords, NULL);
    r->db = new KeyFrameDatabase(r->voc);
    Frame& F = r->base;
    F.N = 0; F.mnScaleLevels = 1; F.mfScaleFactor = 1.2f; F.mfLogScaleFactor = std::log(1.2f);
    F.mK = cv::Mat::eye(3, 3, CV_32F); F.mDistCoef = cv::Mat::zeros(4, 1, CV_32F);
    F.mbf = 40.f; F.mb = 0.1f; F.mThDepth = 0; F.mTimeStamp = 0; F.mnId = 0;
    F.mpORBvocabulary = &r->voc; F.mpORBextractorLeft = F.mpORBextractorRight = NULL; F.mpReferenceKF = NULL;
    F.SetPose(cv::Mat::eye(4, 4, CV_32F));
    return r;
}
void ref_kfdb_destroy(void* h) { RefKfdb* r = (RefKfdb*)h; delete r->db; delete r; }   // keyframes stay with the process, as in the reference
void ref_kfdb_clear(void* h) { ((RefKfdb*)h)->db->clear(); }
int ref_kfdb_new_keyframe(void* h, int id, const int32_t* w, const double* v, int n)
{
    RefKfdb* r = (RefKfdb*)h;
    if (r->kfs.count(id)) return -1;
    KeyFrame* k = new KeyFrame(r->base, &r->map, r->db);
    k->mnId = (long unsigned)id;
    k->mBowVec = bowOf(w, v, n);
    k->mLoopScore = 0.f; k->mRelocScore = 0.f;
    r->kfs[id] = k;
    return 0;
}
int ref_kfdb_add(void* h, int id) { RefKfdb* r = (RefKfdb*)h; if (!r->kfs.count(id)) return -1; r->db->add(r->kfs[id]); return 0; }
int ref_kfdb_erase(void* h, int id) { RefKfdb* r = (RefKfdb*)h; if (!r->kfs.count(id)) return -1; r->db->erase(r->kfs[id]); return 0; }
int ref_kfdb_set_covisible(void* h, int id, const int32_t* ids, int n)
{
    RefKfdb* r = (RefKfdb*)h;
    if (!r->kfs.count(id)) return -1;
    std::vector<KeyFrame*> v;
    for (int i = 0; i < n; ++i) if (r->kfs.count(ids[i])) v.push_back(r->kfs[ids[i]]);
    r->kfs[id]->mvpOrderedConnectedKeyFrames = v;          // what GetBestCovisibilityKeyFrames(10) reads (src/KeyFrame.cc:248-258)
    return 0;
}
int ref_kfdb_detect_loop(void* h, int query, const int32_t* conn, int nconn, float minScore, int32_t* out, int cap)
{
    RefKfdb* r = (RefKfdb*)h;
    if (!r->kfs.count(query)) return -1;
    KeyFrame* q = r->kfs[query];
    q->mConnectedKeyFrameWeights.clear();                  // GetConnectedKeyFrames returns its keys (src/KeyFrame.cc:225-233)
    for (int i = 0; i < nconn; ++i) if (r->kfs.count(conn[i])) q->mConnectedKeyFrameWeights[r->kfs[conn[i]]] = 15;
    std::vector<KeyFrame*> c = r->db->DetectLoopCandidates(q, minScore);
    for (size_t i = 0; i < c.size() && (int)i < cap; ++i) out[i] = (int32_t)c[i]->mnId;
    return (int)c.size();
}
int ref_kfdb_detect_reloc(void* h, long long frame_id, const int32_t* w, const double* v, int n, int32_t* out, int cap)
{
    RefKfdb* r = (RefKfdb*)h;
    Frame F(r->base);
    F.mnId = (long unsigned)frame_id;
    F.mBowVec = bowOf(w, v, n);
    std::vector<KeyFrame*> c = r->db->DetectRelocalizationCandidates(&F);
    for (size_t i = 0; i < c.size() && (int)i < cap; ++i) out[i] = (int32_t)c[i]->mnId;
    return (int)c.size();
}
}
