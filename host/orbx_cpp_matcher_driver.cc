// Test driver for the C++ drop-in ORB_SLAM2::ORBmatcher (host/ORBmatcher.cc) built against the compat stand-ins.
//   orbx_cpp_matcher_driver <in.bin> <out.bin>
// in.bin: a sequence of arrays, each "int32 nbytes, bytes" (see tests/test_gpu_cpp.py); out.bin likewise.
#include "ORBmatcher.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <set>
#include <vector>
using namespace ORB_SLAM2;

static std::vector<std::vector<char> > readAll(const char* path)
{
    std::vector<std::vector<char> > out;
    FILE* f = fopen(path, "rb");
    if (!f) { fprintf(stderr, "cannot open %s\n", path); exit(2); }
    int n;
    while (fread(&n, 4, 1, f) == 1) { std::vector<char> b(n); if (n && fread(b.data(), 1, n, f) != (size_t)n) exit(2); out.push_back(b); }
    fclose(f);
    return out;
}
template <typename T> static const T* as(const std::vector<char>& b) { return reinterpret_cast<const T*>(b.data()); }
static void put(FILE* f, const void* p, int nbytes) { fwrite(&nbytes, 4, 1, f); fwrite(p, 1, nbytes, f); }

// optional 4x4 row-major pose at in[k]; absent or empty -> the given default
static cv::Mat pose4(const std::vector<std::vector<char> >& in, size_t k, const cv::Mat& dflt)
{
    if (in.size() <= k || in[k].size() < 64) return dflt;
    cv::Mat m(4, 4, CV_32F); memcpy(m.data, in[k].data(), 64); return m;
}
static cv::Mat eye4() { cv::Mat m = cv::Mat::zeros(4, 4, CV_32F); for (int i = 0; i < 4; ++i) m.at<float>(i, i) = 1.f; return m; }

static void fillFeatVec(DBoW2::FeatureVector& fv, const int* node, int n)
{
    for (int i = 0; i < n; ++i) if (node[i] >= 0) fv[(unsigned)node[i]].push_back((unsigned)i);
}

int main(int argc, char** argv)
{
    if (argc < 3) return 2;
    std::vector<std::vector<char> > in = readAll(argv[1]);
    // 0: params float[16]: nnratio, checkOri, onlyStereo, fx, fy, cx, cy, th, gridInvW, gridInvH, minX, minY, maxX, maxY
    const float* P = as<float>(in[0]);
    std::vector<MapPoint> pool1, pool2;
    KeyFrame k1, k2;
    KeyFrame* ks[2] = {&k1, &k2};
    std::vector<MapPoint>* pools[2] = {&pool1, &pool2};
    for (int s = 0; s < 2; ++s) {            // 1..5 / 6..10: desc, keys, has, uright, node
        const std::vector<char>&d = in[1 + 5 * s], &k = in[2 + 5 * s], &h = in[3 + 5 * s], &u = in[4 + 5 * s], &nd = in[5 + 5 * s];
        KeyFrame& K = *ks[s];
        K.N = (int)d.size() / 32;
        K.mDescriptors = cv::Mat(K.N, 32, CV_8U);
        memcpy(K.mDescriptors.data, d.data(), d.size());
        K.mvKeysUn.assign(as<cv::KeyPoint>(k), as<cv::KeyPoint>(k) + K.N);
        K.mvuRight.assign(as<float>(u), as<float>(u) + K.N);
        pools[s]->resize(K.N);
        K.mvpMapPoints.assign(K.N, (MapPoint*)NULL);
        for (int i = 0; i < K.N; ++i) if (h[i]) K.mvpMapPoints[i] = &(*pools[s])[i];
        fillFeatVec(K.mFeatVec, as<int>(nd), K.N);
        K.fx = P[3]; K.fy = P[4]; K.cx = P[5]; K.cy = P[6];
        K.mvScaleFactors.assign(as<float>(in[11]), as<float>(in[11]) + 8);
        K.mvLevelSigma2.assign(as<float>(in[12]), as<float>(in[12]) + 8);
    }
    // 13: F12 (9 floats), 14: Ow (3), R2w = I, t2w = 0
    cv::Mat F12(3, 3, CV_32F); memcpy(F12.data, in[13].data(), 36);
    k1.Ow = cv::Mat(3, 1, CV_32F); memcpy(k1.Ow.data, in[14].data(), 12);
    k2.Rcw = cv::Mat::zeros(3, 3, CV_32F); for (int i = 0; i < 3; ++i) k2.Rcw.at<float>(i, i) = 1.f;
    k2.tcw = cv::Mat::zeros(3, 1, CV_32F);
    FILE* o = fopen(argv[2], "wb");
    try {
        ORBmatcher m(P[0], P[1] != 0);
        // (a) SearchByBoW(KF, F): the Frame is side 2
        Frame F;
        F.N = k2.N; F.mDescriptors = k2.mDescriptors; F.mvKeys = k2.mvKeysUn; F.mvKeysUn = k2.mvKeysUn; F.mFeatVec = k2.mFeatVec;
        std::vector<MapPoint*> mm;
        int n = m.SearchByBoW(&k1, F, mm);
        std::vector<int> res(F.N, -1);
        for (int j = 0; j < F.N; ++j) if (mm[j]) res[j] = (int)(mm[j] - pool1.data());
        put(o, &n, 4); put(o, res.data(), 4 * F.N);
        // (b) SearchByBoW(KF, KF)
        n = m.SearchByBoW(&k1, &k2, mm);
        res.assign(k1.N, -1);
        for (int i = 0; i < k1.N; ++i) if (mm[i]) res[i] = (int)(mm[i] - pool2.data());
        put(o, &n, 4); put(o, res.data(), 4 * k1.N);
        // (c) SearchForTriangulation
        std::vector<std::pair<size_t, size_t> > pairs;
        n = m.SearchForTriangulation(&k1, &k2, F12, pairs, P[2] != 0);
        res.assign(k1.N, -1);
        for (size_t i = 0; i < pairs.size(); ++i) res[pairs[i].first] = (int)pairs[i].second;
        put(o, &n, 4); put(o, res.data(), 4 * k1.N);
        // (d) SearchByProjection(Frame, MapPoints, th): 15: points (u, v, ur, viewCos, level as float), 16: their descriptors,
        //     17: occupied flags of the frame keypoints
        const int np = (int)in[15].size() / 20;
        std::vector<MapPoint> pts(np);
        std::vector<MapPoint*> vp(np);
        for (int i = 0; i < np; ++i) {
            const float* q = as<float>(in[15]) + 5 * i;
            pts[i].nObs = in.size() > 25 ? as<int>(in[25])[i] : 1;       // 0: a temporal point (Tracking::UpdateLastFrame)
            pts[i].mbTrackInView = true; pts[i].mTrackProjX = q[0]; pts[i].mTrackProjY = q[1]; pts[i].mTrackProjXR = q[2];
            pts[i].mTrackViewCos = q[3]; pts[i].mnTrackScaleLevel = (int)q[4];
            pts[i].mDescriptor = cv::Mat(1, 32, CV_8U); memcpy(pts[i].mDescriptor.data, in[16].data() + 32 * i, 32);
            vp[i] = &pts[i];
        }
        Frame::mnMinX = P[10]; Frame::mnMinY = P[11]; Frame::mnMaxX = P[12]; Frame::mnMaxY = P[13];
        Frame::mfGridElementWidthInv = P[8]; Frame::mfGridElementHeightInv = P[9];
        F.mvScaleFactors = k2.mvScaleFactors; F.mvuRight = k2.mvuRight;
        MapPoint occupiedMP;
        F.mvpMapPoints.assign(F.N, (MapPoint*)NULL);
        for (int i = 0; i < F.N; ++i) if (in[17][i]) F.mvpMapPoints[i] = &occupiedMP;
        n = ORBmatcher(P[0], true).SearchByProjection(F, vp, P[7]);
        res.assign(F.N, -1);
        for (int i = 0; i < F.N; ++i) if (F.mvpMapPoints[i] && F.mvpMapPoints[i] != &occupiedMP) res[i] = (int)(F.mvpMapPoints[i] - pts.data());
        put(o, &n, 4); put(o, res.data(), 4 * F.N);
        int dd = ORBmatcher::DescriptorDistance(k1.mDescriptors.row(0), k2.mDescriptors.row(0));
        put(o, &dd, 4);
        if (in.size() > 20) {
            // 18: per KF1 feature 8 floats (world x, y, z, mfMaxDistance, mfMinDistance, normal x, y, z)
            // 19: float[8]: th_reloc, ORBdist, th_sim3, window, logScaleFactor, sim3 scale
            const float* Wd = as<float>(in[18]);
            const float* Q = as<float>(in[19]);
            for (int i = 0; i < k1.N; ++i) {
                MapPoint& mp = pool1[i];
                mp.mWorldPos = cv::Mat(3, 1, CV_32F); memcpy(mp.mWorldPos.data, Wd + 8 * i, 12);
                mp.mfMaxDistance = Wd[8 * i + 3]; mp.mfMinDistance = Wd[8 * i + 4];
                mp.mNormalVector = cv::Mat(3, 1, CV_32F); memcpy(mp.mNormalVector.data, Wd + 8 * i + 5, 12);
                mp.mDescriptor = cv::Mat(1, 32, CV_8U); memcpy(mp.mDescriptor.data, in[20].data() + 32 * i, 32);   // 20: MapPoint descriptors
            }
            // (e) SearchByProjection(Frame, KeyFrame, sAlreadyFound, th, ORBdist): relocalisation
            F.mTcw = pose4(in, 31, eye4());
            Frame::fx = P[3]; Frame::fy = P[4]; Frame::cx = P[5]; Frame::cy = P[6];
            F.mnScaleLevels = 8; F.mfLogScaleFactor = Q[4];
            F.mvpMapPoints.assign(F.N, (MapPoint*)NULL);
            for (int i = 0; i < F.N; ++i) if (in[17][i]) F.mvpMapPoints[i] = &occupiedMP;
            std::set<MapPoint*> found;
            for (int i = 0; i < k1.N; i += 17) if (k1.mvpMapPoints[i]) found.insert(k1.mvpMapPoints[i]);
            n = ORBmatcher(P[0], true).SearchByProjection(F, &k1, found, Q[0], (int)Q[1]);
            res.assign(F.N, -1);
            for (int i = 0; i < F.N; ++i) if (F.mvpMapPoints[i] && F.mvpMapPoints[i] != &occupiedMP) res[i] = (int)(F.mvpMapPoints[i] - pool1.data());
            put(o, &n, 4); put(o, res.data(), 4 * F.N);
            // (f) SearchByProjection(KeyFrame, Scw, vpPoints, vpMatched, th): loop closing, KF = side 2
            k2.mnScaleLevels = 8; k2.mfLogScaleFactor = Q[4];
            k2.mnGridCols = FRAME_GRID_COLS; k2.mnGridRows = FRAME_GRID_ROWS;
            k2.mfGridElementWidthInv = P[8]; k2.mfGridElementHeightInv = P[9];
            k2.mnMinX = (int)P[10]; k2.mnMinY = (int)P[11]; k2.mnMaxX = (int)P[12]; k2.mnMaxY = (int)P[13];
            cv::Mat Scw = cv::Mat::zeros(4, 4, CV_32F);
            for (int i = 0; i < 3; ++i) Scw.at<float>(i, i) = Q[5];
            Scw.at<float>(3, 3) = 1.f;
            Scw = pose4(in, 29, Scw);
            std::vector<MapPoint*> vpPoints, vpMatched(k2.N, (MapPoint*)NULL);
            for (int i = 0; i < k1.N; ++i) if (k1.mvpMapPoints[i]) vpPoints.push_back(k1.mvpMapPoints[i]);
            for (int i = 0; i < k2.N; ++i) if (in[17][i]) vpMatched[i] = &occupiedMP;
            n = ORBmatcher(P[0], true).SearchByProjection(&k2, Scw, vpPoints, vpMatched, (int)Q[2]);
            res.assign(k2.N, -1);
            for (int i = 0; i < k2.N; ++i) if (vpMatched[i] && vpMatched[i] != &occupiedMP) res[i] = (int)(vpMatched[i] - pool1.data());
            put(o, &n, 4); put(o, res.data(), 4 * k2.N);
            // (g) SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)
            Frame F1;
            F1.N = k1.N; F1.mDescriptors = k1.mDescriptors; F1.mvKeys = k1.mvKeysUn; F1.mvKeysUn = k1.mvKeysUn;
            std::vector<cv::Point2f> prev(F1.N);
            for (int i = 0; i < F1.N; ++i) prev[i] = F1.mvKeysUn[i].pt;
            std::vector<int> m12;
            n = ORBmatcher(P[0], true).SearchForInitialization(F1, F, prev, m12, (int)Q[3]);
            put(o, &n, 4); put(o, m12.data(), 4 * F1.N); put(o, prev.data(), 8 * F1.N);
        }
        if (in.size() > 24) {
            // 21: world2 (per KF2 feature: x, y, z, maxDist, minDist, normal), 22: KF2 MapPoint descriptors,
            // 23: int32 observation counts of pool1 then pool2, 24: float[8]: th_fuse, bf, th_fuse_sim3, th_sim3
            const float* W1 = as<float>(in[18]); const float* W2 = as<float>(in[21]);
            const int* obsN = as<int>(in[23]);
            const float* Q2 = as<float>(in[24]);
            const std::vector<MapPoint*> has1 = k1.mvpMapPoints, has2 = k2.mvpMapPoints;
            auto setPoint = [&](MapPoint& mp, const float* w, const char* d) {
                mp = MapPoint();
                mp.mWorldPos = cv::Mat(3, 1, CV_32F); memcpy(mp.mWorldPos.data, w, 12);
                mp.mfMaxDistance = w[3]; mp.mfMinDistance = w[4];
                mp.mNormalVector = cv::Mat(3, 1, CV_32F); memcpy(mp.mNormalVector.data, w + 5, 12);
                mp.mDescriptor = cv::Mat(1, 32, CV_8U); memcpy(mp.mDescriptor.data, d, 32);
            };
            auto reset = [&]() {
                for (int i = 0; i < k1.N; ++i) { setPoint(pool1[i], W1 + 8 * i, in[20].data() + 32 * i); k1.mvpMapPoints[i] = has1[i]; }
                for (int i = 0; i < k2.N; ++i) { setPoint(pool2[i], W2 + 8 * i, in[22].data() + 32 * i); k2.mvpMapPoints[i] = has2[i]; }
                for (int i = 0; i < k1.N; ++i) if (has1[i]) { pool1[i].mObs[&k1] = i; pool1[i].nObs = obsN[i]; }
                for (int i = 0; i < k2.N; ++i) if (has2[i]) { pool2[i].mObs[&k2] = i; pool2[i].nObs = obsN[k1.N + i]; }
            };
            auto code = [&](MapPoint* p) { return !p ? -1 : (p >= pool1.data() && p < pool1.data() + k1.N) ? (int)(p - pool1.data()) : 100000 + (int)(p - pool2.data()); };
            for (int s = 0; s < 2; ++s) {
                KeyFrame& K = *ks[s];
                K.Rcw = cv::Mat::zeros(3, 3, CV_32F); for (int i = 0; i < 3; ++i) K.Rcw.at<float>(i, i) = 1.f;
                K.tcw = cv::Mat::zeros(3, 1, CV_32F); K.Ow = cv::Mat::zeros(3, 1, CV_32F);
                K.mbf = Q2[1];
                K.mvInvLevelSigma2.resize(8); for (int l = 0; l < 8; ++l) K.mvInvLevelSigma2[l] = 1.0f / K.mvLevelSigma2[l];
                K.mnScaleLevels = 8; K.mfLogScaleFactor = as<float>(in[19])[4];
                K.mnGridCols = FRAME_GRID_COLS; K.mnGridRows = FRAME_GRID_ROWS;
                K.mfGridElementWidthInv = P[8]; K.mfGridElementHeightInv = P[9];
                K.mnMinX = (int)P[10]; K.mnMinY = (int)P[11]; K.mnMaxX = (int)P[12]; K.mnMaxY = (int)P[13];
            }
            // (h) Fuse(KeyFrame, vpMapPoints, th): KF1's points (with NULL gaps and a few duplicates) into KF2
            reset();
            std::vector<MapPoint*> vp = k1.mvpMapPoints;
            for (int i = 0; i + 50 < k1.N; i += 50) vp.push_back(vp[i]);
            n = ORBmatcher(P[0], true).Fuse(&k2, vp, Q2[0]);
            res.assign(k2.N, -1);
            for (int i = 0; i < k2.N; ++i) res[i] = code(k2.mvpMapPoints[i]);
            put(o, &n, 4); put(o, res.data(), 4 * k2.N);
            std::vector<int> st(k1.N + k2.N);
            for (int i = 0; i < k1.N; ++i) st[i] = (pool1[i].mbBad ? 1 : 0) | (pool1[i].nObs << 1);
            for (int i = 0; i < k2.N; ++i) st[k1.N + i] = (pool2[i].mbBad ? 1 : 0) | (pool2[i].nObs << 1);
            put(o, st.data(), 4 * (int)st.size());
            // (i) Fuse(KeyFrame, Scw, vpPoints, th, vpReplacePoint)
            reset();
            cv::Mat Scw = cv::Mat::zeros(4, 4, CV_32F);
            for (int i = 0; i < 3; ++i) Scw.at<float>(i, i) = 2.f;
            Scw.at<float>(3, 3) = 1.f;
            Scw = pose4(in, 30, Scw);
            std::vector<MapPoint*> vpts;
            for (int i = 0; i < k1.N; ++i) if (k1.mvpMapPoints[i]) vpts.push_back(k1.mvpMapPoints[i]);
            for (size_t i = 0, nv = vpts.size(); i + 40 < nv; i += 40) vpts.push_back(vpts[i]);
            std::vector<MapPoint*> repl(vpts.size(), (MapPoint*)NULL);
            n = ORBmatcher(P[0], true).Fuse(&k2, Scw, vpts, Q2[2], repl);
            res.assign(k2.N, -1);
            for (int i = 0; i < k2.N; ++i) res[i] = code(k2.mvpMapPoints[i]);
            put(o, &n, 4); put(o, res.data(), 4 * k2.N);
            std::vector<int> rc(repl.size());
            for (size_t i = 0; i < repl.size(); ++i) rc[i] = code(repl[i]);
            put(o, rc.data(), 4 * (int)rc.size());
            // (j) SearchBySim3(KF1, KF2, vpMatches12, s12 = 1, R12 = I, t12 = 0, th)
            reset();
            std::vector<MapPoint*> m12(k1.N, (MapPoint*)NULL);
            for (int i = 0; i < k1.N; i += 23) if (k1.mvpMapPoints[i] && i < k2.N && k2.mvpMapPoints[i]) m12[i] = k2.mvpMapPoints[i];
            cv::Mat R12 = cv::Mat::zeros(3, 3, CV_32F); for (int i = 0; i < 3; ++i) R12.at<float>(i, i) = 1.f;
            cv::Mat t12 = cv::Mat::zeros(3, 1, CV_32F);
            const float s12 = 1.0f;
            n = ORBmatcher(P[0], true).SearchBySim3(&k1, &k2, m12, s12, R12, t12, Q2[3]);
            res.assign(k1.N, -1);
            for (int i = 0; i < k1.N; ++i) res[i] = code(m12[i]);
            put(o, &n, 4); put(o, res.data(), 4 * k1.N);
            if (in.size() > 28) {
                // (k) SearchByProjection(CurrentFrame, LastFrame, th, bMono): see oracle/ref_api.cc for the scene
                reset();
                const int* lastObs = as<int>(in[26]);
                const float thK = as<float>(in[28])[0];
                MapPoint transientMP; transientMP.nObs = 0;
                MapPoint occMP;
                Frame Fl;
                Fl.N = k1.N; Fl.mDescriptors = k1.mDescriptors; Fl.mvKeys = k1.mvKeysUn; Fl.mvKeysUn = k1.mvKeysUn;
                Fl.mTcw = eye4();
                Fl.mvpMapPoints.assign(Fl.N, (MapPoint*)NULL); Fl.mvbOutlier.assign(Fl.N, false);
                for (int i = 0; i < Fl.N; ++i) {
                    Fl.mvpMapPoints[i] = has1[i] ? &pool1[i] : (MapPoint*)NULL;
                    Fl.mvbOutlier[i] = (i % 13) == 5;
                    pool1[i].nObs = lastObs[i];
                }
                for (int mono = 0; mono < 2; ++mono) {
                    Frame Fc;
                    Fc.N = k2.N; Fc.mDescriptors = k2.mDescriptors; Fc.mvKeys = k2.mvKeysUn; Fc.mvKeysUn = k2.mvKeysUn;
                    Fc.mvuRight = k2.mvuRight; Fc.mvScaleFactors = k2.mvScaleFactors;
                    Fc.mbf = Q2[1]; Fc.mb = Q2[1] / P[3];
                    Fc.mTcw = pose4(in, 27, eye4());
                    Fc.mvpMapPoints.assign(Fc.N, (MapPoint*)NULL);
                    for (int i = 0; i < Fc.N; ++i) Fc.mvpMapPoints[i] = in[17][i] ? &occMP : ((i % 7) == 3 ? &transientMP : (MapPoint*)NULL);
                    n = ORBmatcher(P[0], true).SearchByProjection(Fc, Fl, thK, mono != 0);
                    res.assign(Fc.N, -1);
                    for (int i = 0; i < Fc.N; ++i)
                        if (Fc.mvpMapPoints[i] && Fc.mvpMapPoints[i] != &occMP && Fc.mvpMapPoints[i] != &transientMP) res[i] = code(Fc.mvpMapPoints[i]);
                    put(o, &n, 4); put(o, res.data(), 4 * Fc.N);
                }
            }
        }
    } catch (const std::exception& e) { fprintf(stderr, "error: %s\n", e.what()); return 1; }
    fclose(o);
    if (ORBmatcher::LastStatus() != 0) {  // the class itself never throws; the driver reports what it recorded
        fprintf(stderr, "ORBmatcher::LastStatus() = %d\n", ORBmatcher::LastStatus());
        return 3;
    }
    return 0;
}
