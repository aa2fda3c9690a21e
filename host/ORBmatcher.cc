// Host glue of the drop-in ORB_SLAM2::ORBmatcher (include/ORBmatcher.h) over the C ABI (include/orbx.h).
// Replaces the bodies of the reference's src/ORBmatcher.cc named below; every Hamming distance, best/second-best
// decision, claim, ratio test and rotation-histogram vote runs on the GPU (csrc/k_match.cu, csrc/k_project.cu).
#include "ORBmatcher.h"
#include "orbx.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace ORB_SLAM2
{

const int ORBmatcher::TH_HIGH = 100;     // src/ORBmatcher.cc:56
const int ORBmatcher::TH_LOW = 50;       // :57
const int ORBmatcher::HISTO_LENGTH = 30; // :58

namespace {
struct CtxHolder {
    orbx_ctx* ctx = nullptr;
    ~CtxHolder() { if (ctx) orbx_destroy(ctx); }
};
// The reference's matchers neither throw nor return codes: a library error (no device, a CUDA failure) is recorded per
// thread, reported once on stderr and turned into "no matches" (every result array is reset to its no-match state).
thread_local int tlsLastStatus = ORBX_OK;
thread_local int tlsDevice = -1;             // -1: env ORBX_DEVICE or 0
bool ok(int rc, const char* what)
{
    if (rc == ORBX_OK) { tlsLastStatus = ORBX_OK; return true; }
    if (tlsLastStatus == ORBX_OK) fprintf(stderr, "ORBmatcher: %s failed: %s\n", what, orbx_strerror(rc));
    tlsLastStatus = rc;
    return false;
}
template <typename V> void noMatch(V& v, int& n) { std::fill(v.begin(), v.end(), -1); n = 0; }
static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint must be the 28-byte POD the kernels read");

// DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned>>) -> CSR in map order (ascending node id)
struct FlatFeatureVector {
    std::vector<int32_t> node_id, node_off, feat;
    explicit FlatFeatureVector(const DBoW2::FeatureVector& fv)
    {
        node_off.push_back(0);
        for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it) {
            node_id.push_back((int32_t)it->first);
            for (size_t k = 0; k < it->second.size(); ++k) feat.push_back((int32_t)it->second[k]);
            node_off.push_back((int32_t)feat.size());
        }
    }
};

// descriptors of a Frame / KeyFrame are rows of an N x 32 CV_8U Mat; the kernels want them contiguous
const uint8_t* contiguous(const cv::Mat& d, std::vector<uint8_t>& tmp)
{
    if (d.isContinuous()) return d.data;
    tmp.resize((size_t)d.rows * 32);
    for (int i = 0; i < d.rows; ++i) memcpy(&tmp[(size_t)i * 32], d.ptr<uint8_t>(i), 32);
    return tmp.data();
}

orbx_bow_side makeSide(const cv::Mat& desc, std::vector<uint8_t>& tmp, const std::vector<cv::KeyPoint>& keys,
                       const std::vector<uint8_t>* has, const std::vector<float>* uright, int N, const FlatFeatureVector& fv)
{
    orbx_bow_side s;
    s.desc = contiguous(desc, tmp);
    s.keys = reinterpret_cast<const orbx_keypoint*>(keys.data());
    s.has_point = has ? has->data() : nullptr;
    s.uright = (uright && !uright->empty()) ? uright->data() : nullptr;
    s.n = N;
    s.nnodes = (int)fv.node_id.size();
    s.node_id = fv.node_id.data(); s.node_off = fv.node_off.data(); s.feat = fv.feat.data();
    return s;
}
} // namespace

// One light context (a device and a stream, no extraction arenas: orbx_create_matcher) per calling thread: the reference
// builds an ORBmatcher on the stack at every call site, from the tracking, mapping and loop-closing threads concurrently.
orbx_ctx* ORBmatcher::Context()
{
    static thread_local CtxHolder holder;
    static thread_local int holderDevice = -1;
    int dev = tlsDevice;
    if (dev < 0) { const char* e = getenv("ORBX_DEVICE"); dev = e ? atoi(e) : 0; }
    if (holder.ctx && holderDevice != dev) { orbx_destroy(holder.ctx); holder.ctx = nullptr; }
    if (!holder.ctx) {
        if (!ok(orbx_create_matcher(&holder.ctx, dev), "orbx_create_matcher")) holder.ctx = nullptr;
        holderDevice = dev;
    }
    return holder.ctx;                       // NULL after a failure: every orbx_* call then returns ORBX_E_INVALID
}

void ORBmatcher::SetDevice(int device) { tlsDevice = device; }
int ORBmatcher::LastStatus() { return tlsLastStatus; }

ORBmatcher::ORBmatcher(float nnratio, bool checkOri) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

int ORBmatcher::DescriptorDistance(const cv::Mat& a, const cv::Mat& b)
{
    const uint32_t* pa = a.ptr<uint32_t>();
    const uint32_t* pb = b.ptr<uint32_t>();
    int dist = 0;
    for (int i = 0; i < 8; ++i) dist += __builtin_popcount(pa[i] ^ pb[i]);
    return dist;
}

float ORBmatcher::RadiusByViewingCos(const float& viewCos) { return viewCos > 0.998 ? 2.5f : 4.0f; } // :194-201

int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches)
{
    const std::vector<MapPoint*> vpMapPointsKF = pKF->GetMapPointMatches();
    vpMapPointMatches = std::vector<MapPoint*>(F.N, static_cast<MapPoint*>(NULL));
    std::vector<uint8_t> has(pKF->N, 0), t1, t2;
    for (int i = 0; i < pKF->N; ++i) has[i] = vpMapPointsKF[i] && !vpMapPointsKF[i]->isBad();     // :307-313
    const FlatFeatureVector fvKF(pKF->mFeatVec), fvF(F.mFeatVec);
    const orbx_bow_side s1 = makeSide(pKF->mDescriptors, t1, pKF->mvKeysUn, &has, nullptr, pKF->N, fvKF);
    const orbx_bow_side s2 = makeSide(F.mDescriptors, t2, F.mvKeys, nullptr, nullptr, F.N, fvF);   // F.mvKeys[..].angle, :362
    std::vector<int32_t> match(F.N > 0 ? F.N : 1);
    int n = 0;
    if (!ok(orbx_match_bow(Context(), ORBX_BOW_KF_FRAME, &s1, &s2, nullptr, mfNNratio, mbCheckOrientation, 0, match.data(), &n),
            "orbx_match_bow")) noMatch(match, n);
    for (int j = 0; j < F.N; ++j) if (match[j] >= 0) vpMapPointMatches[j] = vpMapPointsKF[match[j]];
    return n;
}

int ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12)
{
    const std::vector<MapPoint*> vpMapPoints1 = pKF1->GetMapPointMatches();
    const std::vector<MapPoint*> vpMapPoints2 = pKF2->GetMapPointMatches();
    vpMatches12 = std::vector<MapPoint*>(vpMapPoints1.size(), static_cast<MapPoint*>(NULL));
    std::vector<uint8_t> has1(pKF1->N, 0), has2(pKF2->N, 0), t1, t2;
    for (int i = 0; i < pKF1->N; ++i) has1[i] = vpMapPoints1[i] && !vpMapPoints1[i]->isBad();
    for (int i = 0; i < pKF2->N; ++i) has2[i] = vpMapPoints2[i] && !vpMapPoints2[i]->isBad();
    const FlatFeatureVector fv1(pKF1->mFeatVec), fv2(pKF2->mFeatVec);
    const orbx_bow_side s1 = makeSide(pKF1->mDescriptors, t1, pKF1->mvKeysUn, &has1, nullptr, pKF1->N, fv1);
    const orbx_bow_side s2 = makeSide(pKF2->mDescriptors, t2, pKF2->mvKeysUn, &has2, nullptr, pKF2->N, fv2);
    std::vector<int32_t> match(pKF1->N > 0 ? pKF1->N : 1);
    int n = 0;
    if (!ok(orbx_match_bow(Context(), ORBX_BOW_KF_KF, &s1, &s2, nullptr, mfNNratio, mbCheckOrientation, 0, match.data(), &n),
            "orbx_match_bow")) noMatch(match, n);
    for (int i = 0; i < pKF1->N; ++i) if (match[i] >= 0) vpMatches12[i] = vpMapPoints2[match[i]];
    return n;
}

int ORBmatcher::SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12,
                                       std::vector<std::pair<size_t, size_t> >& vMatchedPairs, const bool bOnlyStereo)
{
    // epipole of KF1's centre in KF2, src/ORBmatcher.cc:920-929 (plain float arithmetic on the 3x3 / 3x1 matrices)
    cv::Mat Cw = pKF1->GetCameraCenter(), R2w = pKF2->GetRotation(), t2w = pKF2->GetTranslation();
    float C2[3];
    for (int r = 0; r < 3; ++r)
        C2[r] = R2w.at<float>(r, 0) * Cw.at<float>(0) + R2w.at<float>(r, 1) * Cw.at<float>(1) + R2w.at<float>(r, 2) * Cw.at<float>(2) + t2w.at<float>(r);
    const float invz = 1.0f / C2[2];
    orbx_triangulation tri;
    memset(&tri, 0, sizeof(tri));
    tri.ex = pKF2->fx * C2[0] * invz + pKF2->cx;
    tri.ey = pKF2->fy * C2[1] * invz + pKF2->cy;
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) tri.F12[3 * r + c] = F12.at<float>(r, c);
    for (size_t l = 0; l < pKF2->mvScaleFactors.size() && l < ORBX_MAX_LEVELS; ++l) {
        tri.scale2[l] = pKF2->mvScaleFactors[l]; tri.sigma2_2[l] = pKF2->mvLevelSigma2[l];
    }
    tri.only_stereo = bOnlyStereo ? 1 : 0;
    std::vector<uint8_t> has1(pKF1->N, 0), has2(pKF2->N, 0), t1, t2;
    for (int i = 0; i < pKF1->N; ++i) has1[i] = pKF1->GetMapPoint(i) != NULL;      // :981 (any MapPoint, bad or not)
    for (int i = 0; i < pKF2->N; ++i) has2[i] = pKF2->GetMapPoint(i) != NULL;      // :1011
    const FlatFeatureVector fv1(pKF1->mFeatVec), fv2(pKF2->mFeatVec);
    const orbx_bow_side s1 = makeSide(pKF1->mDescriptors, t1, pKF1->mvKeysUn, &has1, &pKF1->mvuRight, pKF1->N, fv1);
    const orbx_bow_side s2 = makeSide(pKF2->mDescriptors, t2, pKF2->mvKeysUn, &has2, &pKF2->mvuRight, pKF2->N, fv2);
    std::vector<int32_t> match(pKF1->N > 0 ? pKF1->N : 1);
    int n = 0;
    if (!ok(orbx_match_bow(Context(), ORBX_BOW_TRIANGULATE, &s1, &s2, &tri, mfNNratio, mbCheckOrientation, 0, match.data(), &n),
            "orbx_match_bow")) noMatch(match, n);
    vMatchedPairs.clear();
    vMatchedPairs.reserve(n);
    for (int i = 0; i < pKF1->N; ++i) if (match[i] >= 0) vMatchedPairs.push_back(std::make_pair((size_t)i, (size_t)match[i])); // :1116-1124
    return n;
}

namespace {
// -R^T t the way OpenCV evaluates `-Rcw.t() * tcw` (src/ORBmatcher.cc:446, :1335, :1760, :1932): a product with a transposed
// operand leaves gemm's small-matrix float path and accumulates in DOUBLE, one rounding to float at the end (alpha = -1).
// oracle/cvshim states the same rule; tests/test_shim_algebra.py pins it to cv2 4.13 (a float sum differs in 4 of 5 cases).
void centreOf(const float R[9], const float t[3], float Ow[3])
{
    for (int r = 0; r < 3; ++r)
        Ow[r] = (float)(-((double)R[r] * (double)t[0] + (double)R[3 + r] * (double)t[1] + (double)R[6 + r] * (double)t[2]));
}

int projectionSearch(orbx_ctx* ctx, Frame& F, const std::vector<orbx_proj_query>& q, const std::vector<MapPoint*>& src, int modeFlags,
                     int thDist, float nnratio, bool checkOri, bool anyPointOccupies = false)
{
    std::vector<uint8_t> occupied(F.N, 0), tmp;
    for (int i = 0; i < F.N; ++i)                                                                              // :136-138, :2000
        occupied[i] = F.mvpMapPoints[i] && (anyPointOccupies || F.mvpMapPoints[i]->Observations() > 0);
    const orbx_grid g = {Frame::mnMinX, Frame::mnMinY, Frame::mfGridElementWidthInv, Frame::mfGridElementHeightInv, FRAME_GRID_COLS, FRAME_GRID_ROWS};
    std::vector<int32_t> assign(F.N > 0 ? F.N : 1);
    int n = 0;
    const bool stereo = !F.mvuRight.empty() && !anyPointOccupies;   // the relocalisation overload has no uright gate
    if (!ok(orbx_search_by_projection(ctx, q.data(), (int)q.size(), reinterpret_cast<const orbx_keypoint*>(F.mvKeysUn.data()),
                                      contiguous(F.mDescriptors, tmp), stereo ? F.mvuRight.data() : nullptr, F.N, &g, occupied.data(),
                                      modeFlags | (stereo ? ORBX_PROJ_CHECK_URIGHT : 0), thDist, nnratio, checkOri ? 1 : 0, 0,
                                      assign.data(), &n), "orbx_search_by_projection")) noMatch(assign, n);
    for (int i = 0; i < F.N; ++i) if (assign[i] >= 0) F.mvpMapPoints[i] = src[assign[i]];
    return n;
}
} // namespace

int ORBmatcher::SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th)
{
    const bool bFactor = th != 1.0;
    std::vector<orbx_proj_query> q;
    std::vector<MapPoint*> src;
    q.reserve(vpMapPoints.size());
    for (size_t iMP = 0; iMP < vpMapPoints.size(); iMP++) {
        MapPoint* pMP = vpMapPoints[iMP];
        if (!pMP->mbTrackInView) continue;
        if (pMP->isBad()) continue;
        const int nPredictedLevel = pMP->mnTrackScaleLevel;
        float r = RadiusByViewingCos(pMP->mTrackViewCos);
        if (bFactor) r *= th;
        orbx_proj_query e;
        memset(&e, 0, sizeof(e));
        e.u = pMP->mTrackProjX; e.v = pMP->mTrackProjY; e.ur = pMP->mTrackProjXR;
        e.radius = r * F.mvScaleFactors[nPredictedLevel];
        e.min_level = nPredictedLevel - 1; e.max_level = nPredictedLevel;
        e.flags = pMP->Observations() > 0 ? 0 : ORBX_PROJ_QUERY_TRANSIENT;        // :136-138: only an observed holder hides a keypoint
        const cv::Mat d = pMP->GetDescriptor();
        memcpy(e.desc, d.ptr<uint8_t>(), 32);
        q.push_back(e); src.push_back(pMP);
    }
    return projectionSearch(Context(), F, q, src, ORBX_PROJ_RATIO_SAME_LEVEL, TH_HIGH, mfNNratio, false);
}

int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono)
{
    // projection step of src/ORBmatcher.cc:1756-1808 on the host (plain float arithmetic, row . vector + t)
    const cv::Mat& Tc = CurrentFrame.mTcw; const cv::Mat& Tl = LastFrame.mTcw;
    float Rcw[9], tcw[3], Rlw[9], tlw[3], twc[3], tlc[3];
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) { Rcw[3 * r + c] = Tc.at<float>(r, c); Rlw[3 * r + c] = Tl.at<float>(r, c); }
        tcw[r] = Tc.at<float>(r, 3); tlw[r] = Tl.at<float>(r, 3);
    }
    centreOf(Rcw, tcw, twc);                                                                                     // -Rcw^T tcw, :1760
    for (int r = 0; r < 3; ++r) tlc[r] = Rlw[3 * r] * twc[0] + Rlw[3 * r + 1] * twc[1] + Rlw[3 * r + 2] * twc[2] + tlw[r];
    const bool bForward = tlc[2] > CurrentFrame.mb && !bMono;
    const bool bBackward = -tlc[2] > CurrentFrame.mb && !bMono;
    std::vector<orbx_proj_query> q;
    std::vector<MapPoint*> src;
    for (int i = 0; i < LastFrame.N; i++) {
        MapPoint* pMP = LastFrame.mvpMapPoints[i];
        if (!pMP || LastFrame.mvbOutlier[i]) continue;
        cv::Mat X = pMP->GetWorldPos();
        const float x = X.at<float>(0), y = X.at<float>(1), z = X.at<float>(2);
        const float xc = Rcw[0] * x + Rcw[1] * y + Rcw[2] * z + tcw[0];
        const float yc = Rcw[3] * x + Rcw[4] * y + Rcw[5] * z + tcw[1];
        const float invzc = 1.0 / (Rcw[6] * x + Rcw[7] * y + Rcw[8] * z + tcw[2]);
        if (invzc < 0) continue;
        const float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;
        const float v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
        if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
        if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
        const int nLastOctave = LastFrame.mvKeys[i].octave;
        orbx_proj_query e;
        memset(&e, 0, sizeof(e));
        e.u = u; e.v = v; e.radius = th * CurrentFrame.mvScaleFactors[nLastOctave];
        e.ur = u - CurrentFrame.mbf * invzc;
        e.angle = LastFrame.mvKeysUn[i].angle;
        if (bForward) { e.min_level = nLastOctave; e.max_level = -1; }
        else if (bBackward) { e.min_level = 0; e.max_level = nLastOctave; }
        else { e.min_level = nLastOctave - 1; e.max_level = nLastOctave + 1; }
        e.flags = pMP->Observations() > 0 ? 0 : ORBX_PROJ_QUERY_TRANSIENT;        // :1840-1842 (UpdateLastFrame's temporal points)
        const cv::Mat d = pMP->GetDescriptor();
        memcpy(e.desc, d.ptr<uint8_t>(), 32);
        q.push_back(e); src.push_back(pMP);
    }
    return projectionSearch(Context(), CurrentFrame, q, src, 0, TH_HIGH, mfNNratio, mbCheckOrientation);
}

namespace {
// 3x3 (row-major) and 3x1 float blocks of a pose matrix, and the camera centre -R^T t
void poseOf(const cv::Mat& T, float R[9], float t[3], float Ow[3], float invScale = 1.0f)
{
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) R[3 * r + c] = T.at<float>(r, c) * invScale;
        t[r] = T.at<float>(r, 3) * invScale;
    }
    centreOf(R, t, Ow);
}
// Sim3 pose Scw -> (Rcw, tcw, Ow) as src/ORBmatcher.cc:441-446 / :1330-1335 evaluate it: scw = sqrt(row0 . row0) with
// cv::Mat::dot and sqrt in double, rounded to float; Mat / scw = Mat * (float)(1.0 / scw) element by element.
void sim3PoseOf(const cv::Mat& Scw, float R[9], float t[3], float Ow[3])
{
    const double r00 = Scw.at<float>(0, 0), r01 = Scw.at<float>(0, 1), r02 = Scw.at<float>(0, 2);
    const float scw = (float)std::sqrt(r00 * r00 + r01 * r01 + r02 * r02);
    poseOf(Scw, R, t, Ow, (float)(1.0 / scw));
}
float norm3(float x, float y, float z) { return (float)std::sqrt((double)x * x + (double)y * y + (double)z * z); }   // cv::norm(NORM_L2)
} // namespace

// Relocalisation: project the KeyFrame's MapPoints into the current frame (:1949-1978) and search around them.
int ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound, const float th,
                                   const int ORBdist)
{
    float Rcw[9], tcw[3], Ow[3];
    poseOf(CurrentFrame.mTcw, Rcw, tcw, Ow);
    const std::vector<MapPoint*> vpMPs = pKF->GetMapPointMatches();
    std::vector<orbx_proj_query> q;
    std::vector<MapPoint*> src;
    for (size_t i = 0, iend = vpMPs.size(); i < iend; i++) {
        MapPoint* pMP = vpMPs[i];
        if (!pMP || pMP->isBad() || sAlreadyFound.count(pMP)) continue;
        cv::Mat X = pMP->GetWorldPos();
        const float x = X.at<float>(0), y = X.at<float>(1), z = X.at<float>(2);
        const float xc = Rcw[0] * x + Rcw[1] * y + Rcw[2] * z + tcw[0];
        const float yc = Rcw[3] * x + Rcw[4] * y + Rcw[5] * z + tcw[1];
        const float invzc = 1.0 / (Rcw[6] * x + Rcw[7] * y + Rcw[8] * z + tcw[2]);
        const float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;
        const float v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
        if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
        if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
        const float dist3D = norm3(x - Ow[0], y - Ow[1], z - Ow[2]);
        const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        const int nPredictedLevel = pMP->PredictScale(dist3D, &CurrentFrame);
        orbx_proj_query e;
        memset(&e, 0, sizeof(e));
        e.u = u; e.v = v; e.radius = th * CurrentFrame.mvScaleFactors[nPredictedLevel];
        e.angle = pKF->mvKeysUn[i].angle;
        e.min_level = nPredictedLevel - 1; e.max_level = nPredictedLevel + 1;
        const cv::Mat d = pMP->GetDescriptor();
        memcpy(e.desc, d.ptr<uint8_t>(), 32);
        q.push_back(e); src.push_back(pMP);
    }
    return projectionSearch(Context(), CurrentFrame, q, src, 0, ORBdist, mfNNratio, mbCheckOrientation, true);
}

// Loop closing: project candidate MapPoints with the Sim3 Scw into the KeyFrame (:437-512) and search around them.
int ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, std::vector<MapPoint*>& vpMatched, int th)
{
    const float fx = pKF->fx, fy = pKF->fy, cx = pKF->cx, cy = pKF->cy;
    float Rcw[9], tcw[3], Ow[3];
    sim3PoseOf(Scw, Rcw, tcw, Ow);
    std::set<MapPoint*> spAlreadyFound(vpMatched.begin(), vpMatched.end());
    spAlreadyFound.erase(static_cast<MapPoint*>(NULL));
    std::vector<orbx_proj_query> q;
    std::vector<MapPoint*> src;
    for (int iMP = 0, iendMP = (int)vpPoints.size(); iMP < iendMP; iMP++) {
        MapPoint* pMP = vpPoints[iMP];
        if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
        cv::Mat X = pMP->GetWorldPos();
        const float x = X.at<float>(0), y = X.at<float>(1), z = X.at<float>(2);
        const float xc = Rcw[0] * x + Rcw[1] * y + Rcw[2] * z + tcw[0];
        const float yc = Rcw[3] * x + Rcw[4] * y + Rcw[5] * z + tcw[1];
        const float zc = Rcw[6] * x + Rcw[7] * y + Rcw[8] * z + tcw[2];
        if (zc < 0.0) continue;
        const float invz = 1 / zc;
        const float u = fx * (xc * invz) + cx, v = fy * (yc * invz) + cy;
        if (!pKF->IsInImage(u, v)) continue;
        const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
        const float POx = x - Ow[0], POy = y - Ow[1], POz = z - Ow[2];
        const float dist = norm3(POx, POy, POz);
        if (dist < minDistance || dist > maxDistance) continue;
        cv::Mat Pn = pMP->GetNormal();
        if ((double)POx * Pn.at<float>(0) + (double)POy * Pn.at<float>(1) + (double)POz * Pn.at<float>(2) < 0.5 * dist) continue;   // cv::Mat::dot sums in double
        const int nPredictedLevel = pMP->PredictScale(dist, pKF);
        orbx_proj_query e;
        memset(&e, 0, sizeof(e));
        e.u = u; e.v = v; e.radius = th * pKF->mvScaleFactors[nPredictedLevel];
        e.min_level = nPredictedLevel - 1; e.max_level = nPredictedLevel;      // :536 (KeyFrame::GetFeaturesInArea has no level filter)
        const cv::Mat d = pMP->GetDescriptor();
        memcpy(e.desc, d.ptr<uint8_t>(), 32);
        q.push_back(e); src.push_back(pMP);
    }
    std::vector<uint8_t> occupied(pKF->N, 0), tmp;
    for (int i = 0; i < pKF->N; ++i) occupied[i] = vpMatched[i] != NULL;                                     // :531
    const orbx_grid g = {(float)pKF->mnMinX, (float)pKF->mnMinY, pKF->mfGridElementWidthInv, pKF->mfGridElementHeightInv,
                         pKF->mnGridCols, pKF->mnGridRows};
    std::vector<int32_t> assign(pKF->N > 0 ? pKF->N : 1);
    int n = 0;
    if (!ok(orbx_search_by_projection(Context(), q.data(), (int)q.size(), reinterpret_cast<const orbx_keypoint*>(pKF->mvKeysUn.data()),
                                      contiguous(pKF->mDescriptors, tmp), nullptr, pKF->N, &g, occupied.data(), 0, TH_LOW, mfNNratio, 0, 0,
                                      assign.data(), &n), "orbx_search_by_projection")) noMatch(assign, n);
    for (int i = 0; i < pKF->N; ++i) if (assign[i] >= 0) vpMatched[i] = src[assign[i]];                      // :554
    return n;
}

int ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12,
                                        int windowSize)
{
    static_assert(sizeof(cv::Point2f) == 2 * sizeof(float), "vbPrevMatched is read as (x, y) float pairs");
    const int n1 = (int)F1.mvKeysUn.size(), n2 = (int)F2.mvKeysUn.size();
    vnMatches12 = std::vector<int>(n1, -1);                                                                  // :590
    if (n1 == 0) return 0;
    std::vector<uint8_t> t1, t2;
    const orbx_grid g = {Frame::mnMinX, Frame::mnMinY, Frame::mfGridElementWidthInv, Frame::mfGridElementHeightInv, FRAME_GRID_COLS, FRAME_GRID_ROWS};
    int n = 0;
    if (!ok(orbx_search_for_initialization(Context(), reinterpret_cast<const orbx_keypoint*>(F1.mvKeysUn.data()), contiguous(F1.mDescriptors, t1),
                                           n1, reinterpret_cast<float*>(vbPrevMatched.data()),
                                           reinterpret_cast<const orbx_keypoint*>(F2.mvKeysUn.data()), contiguous(F2.mDescriptors, t2), n2, &g,
                                           windowSize, mfNNratio, mbCheckOrientation ? 1 : 0, 0, vnMatches12.data(), &n),
            "orbx_search_for_initialization")) noMatch(vnMatches12, n);
    return n;
}

namespace {
// the independent window search of Fuse x2 / SearchBySim3 on a KeyFrame's keypoints (orbx_window_best)
void keyFrameWindowBest(orbx_ctx* ctx, KeyFrame* pKF, const std::vector<orbx_proj_query>& q, bool reprojGate, int thDist,
                        std::vector<int32_t>& bestIdx)
{
    bestIdx.assign(q.size() > 0 ? q.size() : 1, -1);
    if (q.empty() || pKF->N == 0) return;
    std::vector<int32_t> bestDist(q.size());
    std::vector<uint8_t> tmp;
    float invSigma2[ORBX_MAX_LEVELS] = {0};
    for (size_t l = 0; l < pKF->mvInvLevelSigma2.size() && l < (size_t)ORBX_MAX_LEVELS; ++l) invSigma2[l] = pKF->mvInvLevelSigma2[l];
    const orbx_grid g = {(float)pKF->mnMinX, (float)pKF->mnMinY, pKF->mfGridElementWidthInv, pKF->mfGridElementHeightInv,
                         pKF->mnGridCols, pKF->mnGridRows};
    if (!ok(orbx_window_best(ctx, q.data(), (int)q.size(), reinterpret_cast<const orbx_keypoint*>(pKF->mvKeysUn.data()),
                             contiguous(pKF->mDescriptors, tmp), pKF->mvuRight.empty() ? nullptr : pKF->mvuRight.data(), pKF->N, &g,
                             reprojGate ? invSigma2 : nullptr, thDist, 0, bestIdx.data(), bestDist.data()), "orbx_window_best"))
        std::fill(bestIdx.begin(), bestIdx.end(), -1);          // nothing found: Fuse / SearchBySim3 then change nothing
}

// project a world point with (R, t) and the KeyFrame's intrinsics; the gates shared by Fuse x2 (:1166-1222, :1345-1395)
bool projectForFuse(KeyFrame* pKF, MapPoint* pMP, const float R[9], const float t[3], const float Ow[3], float th, float bf,
                    orbx_proj_query& e)
{
    cv::Mat X = pMP->GetWorldPos();
    const float x = X.at<float>(0), y = X.at<float>(1), z = X.at<float>(2);
    const float xc = R[0] * x + R[1] * y + R[2] * z + t[0];
    const float yc = R[3] * x + R[4] * y + R[5] * z + t[1];
    const float zc = R[6] * x + R[7] * y + R[8] * z + t[2];
    if (zc < 0.0f) return false;
    const float invz = 1 / zc;
    const float u = pKF->fx * (xc * invz) + pKF->cx, v = pKF->fy * (yc * invz) + pKF->cy;
    if (!pKF->IsInImage(u, v)) return false;
    const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
    const float POx = x - Ow[0], POy = y - Ow[1], POz = z - Ow[2];
    const float dist3D = norm3(POx, POy, POz);
    if (dist3D < minDistance || dist3D > maxDistance) return false;
    cv::Mat Pn = pMP->GetNormal();
    if ((double)POx * Pn.at<float>(0) + (double)POy * Pn.at<float>(1) + (double)POz * Pn.at<float>(2) < 0.5 * dist3D) return false;   // cv::Mat::dot sums in double
    const int nPredictedLevel = pMP->PredictScale(dist3D, pKF);
    memset(&e, 0, sizeof(e));
    e.u = u; e.v = v; e.ur = u - bf * invz;
    e.radius = th * pKF->mvScaleFactors[nPredictedLevel];
    e.min_level = nPredictedLevel - 1; e.max_level = nPredictedLevel;
    const cv::Mat d = pMP->GetDescriptor();
    memcpy(e.desc, d.ptr<uint8_t>(), 32);
    return true;
}
} // namespace

// LocalMapping: fuse the neighbours' MapPoints into pKF.  The candidate search of every point is independent (GPU); the map
// is then updated point by point in the reference's order, re-checking the skip conditions the earlier updates can change.
int ORBmatcher::Fuse(KeyFrame* pKF, const std::vector<MapPoint*>& vpMapPoints, const float th)
{
    float Rcw[9], tcw[3], Ow[3];
    const cv::Mat R = pKF->GetRotation(), t = pKF->GetTranslation(), O = pKF->GetCameraCenter();
    for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) Rcw[3 * r + c] = R.at<float>(r, c); tcw[r] = t.at<float>(r); Ow[r] = O.at<float>(r); }
    std::vector<orbx_proj_query> q;
    std::vector<int> src;
    for (int i = 0, n = (int)vpMapPoints.size(); i < n; i++) {
        MapPoint* pMP = vpMapPoints[i];
        if (!pMP) continue;
        if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
        orbx_proj_query e;
        if (!projectForFuse(pKF, pMP, Rcw, tcw, Ow, th, pKF->mbf, e)) continue;
        q.push_back(e); src.push_back(i);
    }
    std::vector<int32_t> bestIdx;
    keyFrameWindowBest(Context(), pKF, q, true, TH_LOW, bestIdx);
    int nFused = 0;
    for (size_t k = 0; k < q.size(); ++k) {
        MapPoint* pMP = vpMapPoints[src[k]];
        if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;          // changed by an earlier Replace / AddObservation (:1157)
        if (bestIdx[k] < 0) continue;
        MapPoint* pMPinKF = pKF->GetMapPoint(bestIdx[k]);               // :1288-1300
        if (pMPinKF) {
            if (!pMPinKF->isBad()) {
                if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
                else pMPinKF->Replace(pMP);
            }
        } else {
            pMP->AddObservation(pKF, bestIdx[k]);
            pKF->AddMapPoint(pMP, bestIdx[k]);
        }
        nFused++;
    }
    return nFused;
}

// LoopClosing: fuse the loop's MapPoints into pKF with the corrected Sim3 pose (:1310-1458).
int ORBmatcher::Fuse(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, float th, std::vector<MapPoint*>& vpReplacePoint)
{
    float Rcw[9], tcw[3], Ow[3];
    sim3PoseOf(Scw, Rcw, tcw, Ow);
    const std::set<MapPoint*> spAlreadyFound = pKF->GetMapPoints();
    std::vector<orbx_proj_query> q;
    std::vector<int> src;
    for (int iMP = 0, n = (int)vpPoints.size(); iMP < n; iMP++) {
        MapPoint* pMP = vpPoints[iMP];
        if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
        orbx_proj_query e;
        if (!projectForFuse(pKF, pMP, Rcw, tcw, Ow, th, 0.0f, e)) continue;
        q.push_back(e); src.push_back(iMP);
    }
    std::vector<int32_t> bestIdx;
    keyFrameWindowBest(Context(), pKF, q, false, TH_LOW, bestIdx);
    int nFused = 0;
    for (size_t k = 0; k < q.size(); ++k) {
        if (bestIdx[k] < 0) continue;
        MapPoint* pMP = vpPoints[src[k]];
        MapPoint* pMPinKF = pKF->GetMapPoint(bestIdx[k]);               // :1442-1456
        if (pMPinKF) {
            if (!pMPinKF->isBad()) vpReplacePoint[src[k]] = pMPinKF;
        } else {
            pMP->AddObservation(pKF, bestIdx[k]);
            pKF->AddMapPoint(pMP, bestIdx[k]);
        }
        nFused++;
    }
    return nFused;
}

// LoopClosing: more matches between two keyframes once a Sim3 is known; both directions must agree (:1472-1723).
int ORBmatcher::SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const float& s12, const cv::Mat& R12,
                             const cv::Mat& t12, const float th)
{
    float R1w[9], t1w[3], R2w[9], t2w[3], sR12[9], sR21[9], t21[3], T12[3];
    const cv::Mat r1 = pKF1->GetRotation(), tt1 = pKF1->GetTranslation(), r2 = pKF2->GetRotation(), tt2 = pKF2->GetTranslation();
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) {
            R1w[3 * r + c] = r1.at<float>(r, c); R2w[3 * r + c] = r2.at<float>(r, c);
            sR12[3 * r + c] = s12 * R12.at<float>(r, c);
            sR21[3 * r + c] = (float)(1.0 / s12) * R12.at<float>(c, r);     // cv::Mat scaling of CV_32F works in float
        }
        t1w[r] = tt1.at<float>(r); t2w[r] = tt2.at<float>(r); T12[r] = t12.at<float>(r);
    }
    for (int r = 0; r < 3; ++r) t21[r] = -(sR21[3 * r] * T12[0] + sR21[3 * r + 1] * T12[1] + sR21[3 * r + 2] * T12[2]);
    const std::vector<MapPoint*> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
    const int N1 = (int)vpMapPoints1.size(), N2 = (int)vpMapPoints2.size();
    std::vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
    for (int i = 0; i < N1; i++) {
        MapPoint* pMP = vpMatches12[i];
        if (pMP) {
            vbAlreadyMatched1[i] = true;
            const int idx2 = pMP->GetIndexInKeyFrame(pKF2);
            if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
        }
    }
    // one direction: points of `from` (pose Raw, taw) into `to` through (sRba, tba)
    auto direction = [&](const std::vector<MapPoint*>& pts, const std::vector<bool>& done, const float* Raw, const float* taw,
                         const float* sRba, const float* tba, KeyFrame* to, std::vector<int>& match) {
        std::vector<orbx_proj_query> q;
        std::vector<int> src;
        for (int i = 0; i < (int)pts.size(); i++) {
            MapPoint* pMP = pts[i];
            if (!pMP || done[i]) continue;
            if (pMP->isBad()) continue;
            cv::Mat X = pMP->GetWorldPos();
            const float x = X.at<float>(0), y = X.at<float>(1), z = X.at<float>(2);
            const float ax = Raw[0] * x + Raw[1] * y + Raw[2] * z + taw[0], ay = Raw[3] * x + Raw[4] * y + Raw[5] * z + taw[1],
                        az = Raw[6] * x + Raw[7] * y + Raw[8] * z + taw[2];
            const float bx = sRba[0] * ax + sRba[1] * ay + sRba[2] * az + tba[0], by = sRba[3] * ax + sRba[4] * ay + sRba[5] * az + tba[1],
                        bz = sRba[6] * ax + sRba[7] * ay + sRba[8] * az + tba[2];
            if (bz < 0.0) continue;
            const float invz = 1.0 / bz;
            const float u = to->fx * (bx * invz) + to->cx, v = to->fy * (by * invz) + to->cy;
            if (!to->IsInImage(u, v)) continue;
            const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
            const float dist3D = norm3(bx, by, bz);
            if (dist3D < minDistance || dist3D > maxDistance) continue;
            const int nPredictedLevel = pMP->PredictScale(dist3D, to);
            orbx_proj_query e;
            memset(&e, 0, sizeof(e));
            e.u = u; e.v = v; e.radius = th * to->mvScaleFactors[nPredictedLevel];
            e.min_level = nPredictedLevel - 1; e.max_level = nPredictedLevel;
            const cv::Mat d = pMP->GetDescriptor();
            memcpy(e.desc, d.ptr<uint8_t>(), 32);
            q.push_back(e); src.push_back(i);
        }
        std::vector<int32_t> bestIdx;
        keyFrameWindowBest(Context(), to, q, false, TH_HIGH, bestIdx);
        for (size_t k = 0; k < q.size(); ++k) if (bestIdx[k] >= 0) match[src[k]] = bestIdx[k];
    };
    std::vector<int> vnMatch1(N1, -1), vnMatch2(N2, -1);
    direction(vpMapPoints1, vbAlreadyMatched1, R1w, t1w, sR21, t21, pKF2, vnMatch1);
    direction(vpMapPoints2, vbAlreadyMatched2, R2w, t2w, sR12, T12, pKF1, vnMatch2);
    int nFound = 0;
    for (int i1 = 0; i1 < N1; i1++) {                                                                        // :1700-1720
        const int idx2 = vnMatch1[i1];
        if (idx2 >= 0 && vnMatch2[idx2] == i1) { vpMatches12[i1] = vpMapPoints2[idx2]; nFound++; }
    }
    return nFound;
}

// static members of the compat Frame stand-in (the reference defines them in src/Frame.cc:38-42)
#ifdef ORBX_COMPAT_SLAM_FRAME_H
float Frame::fx, Frame::fy, Frame::cx, Frame::cy;
float Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY;
float Frame::mfGridElementWidthInv, Frame::mfGridElementHeightInv;
#endif

} // namespace ORB_SLAM2
