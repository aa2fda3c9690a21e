// Latency of the C++ drop-in classes as the reference's Tracking thread would see it (src/Tracking.cc:1356-1363,
// src/Frame.cc:430-443): ORBextractor::operator() on one host image, with the host mirror of mvImagePyramid on and off,
// and one ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, mono) with the last frame's map points.
//   orbx_cpp_latency <in.raw> <w> <h> <nfeatures> [iterations]      -> one JSON line on stdout
// Built against the compat stand-ins (this image has no OpenCV C++); bench.py runs it and copies the line.
#include "ORBextractor.h"
#include "ORBmatcher.h"
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
using namespace ORB_SLAM2;

static double median(std::vector<double> v) { std::sort(v.begin(), v.end()); return v.empty() ? 0.0 : v[v.size() / 2]; }
static double nowMs() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main(int argc, char** argv)
{
    if (argc < 5) { fprintf(stderr, "usage: orbx_cpp_latency in.raw w h nfeatures [iterations]\n"); return 2; }
    const int w = atoi(argv[2]), h = atoi(argv[3]), nf = atoi(argv[4]), iters = argc > 5 ? atoi(argv[5]) : 200;
    std::vector<unsigned char> buf((size_t)w * h);
    FILE* f = fopen(argv[1], "rb");
    if (!f || fread(buf.data(), 1, buf.size(), f) != buf.size()) { fprintf(stderr, "read failed\n"); return 2; }
    fclose(f);
    ORBextractor ex(nf, 1.2f, 8, 20, 7);
    cv::Mat im(h, w, CV_8UC1, buf.data());
    std::vector<cv::KeyPoint> keys;
    cv::Mat desc;
    double msMirror = 0, msPlain = 0;
    for (int mirror = 1; mirror >= 0; --mirror) {
        ex.SetHostPyramidMirror(mirror != 0);
        for (int i = 0; i < 20; ++i) ex(im, cv::Mat(), keys, desc);
        std::vector<double> t;
        for (int i = 0; i < iters; ++i) { const double t0 = nowMs(); ex(im, cv::Mat(), keys, desc); t.push_back(nowMs() - t0); }
        (mirror ? msMirror : msPlain) = median(t);
    }
    if (ex.LastStatus() != 0 || keys.empty()) { fprintf(stderr, "extractor status %d\n", ex.LastStatus()); return 3; }

    // last frame = this frame's keypoints, each with a map point 2..6 m in front of the camera; the current frame sees the same
    // scene after a small forward motion, so the projections land a few pixels from the keypoints they came from
    const int N = (int)keys.size();
    const float fx = 520.9f, fy = 521.0f, cx = w * 0.5f, cy = h * 0.5f, bf = 40.f;
    Frame::fx = fx; Frame::fy = fy; Frame::cx = cx; Frame::cy = cy;
    Frame::mnMinX = 0; Frame::mnMinY = 0; Frame::mnMaxX = (float)w; Frame::mnMaxY = (float)h;
    Frame::mfGridElementWidthInv = (float)FRAME_GRID_COLS / (float)w; Frame::mfGridElementHeightInv = (float)FRAME_GRID_ROWS / (float)h;
    std::vector<MapPoint> pts(N);
    Frame last, cur;
    for (Frame* F : {&last, &cur}) {
        F->N = N; F->mDescriptors = desc; F->mvKeys = keys; F->mvKeysUn = keys;
        F->mvuRight.assign(N, -1.f); F->mvScaleFactors = ex.GetScaleFactors();
        F->mvpMapPoints.assign(N, (MapPoint*)NULL); F->mvbOutlier.assign(N, false);
        F->mbf = bf; F->mb = bf / fx;
        F->mTcw = cv::Mat::zeros(4, 4, CV_32F);
        for (int i = 0; i < 4; ++i) F->mTcw.at<float>(i, i) = 1.f;
    }
    cur.mTcw.at<float>(2, 3) = -0.02f;
    unsigned rng = 12345u;
    for (int i = 0; i < N; ++i) {
        rng = rng * 1664525u + 1013904223u;
        const float z = 2.f + 4.f * (float)(rng >> 8) / 16777216.f;
        pts[i].mWorldPos = cv::Mat(3, 1, CV_32F);
        pts[i].mWorldPos.at<float>(0) = (keys[i].pt.x - cx) * z / fx;
        pts[i].mWorldPos.at<float>(1) = (keys[i].pt.y - cy) * z / fy;
        pts[i].mWorldPos.at<float>(2) = z;
        pts[i].mDescriptor = cv::Mat(1, 32, CV_8U);
        memcpy(pts[i].mDescriptor.data, desc.ptr<unsigned char>(i), 32);
        pts[i].nObs = (i % 3) ? 1 : 0;                       // a third are temporal points (stereo / RGB-D tracking)
        last.mvpMapPoints[i] = &pts[i];
    }
    int nm = 0;
    std::vector<double> t;
    for (int i = 0; i < 20 + iters; ++i) {
        std::fill(cur.mvpMapPoints.begin(), cur.mvpMapPoints.end(), (MapPoint*)NULL);     // Tracking does this per frame
        const double t0 = nowMs();
        nm = ORBmatcher(0.9f, true).SearchByProjection(cur, last, 15.f, true);
        if (i >= 20) t.push_back(nowMs() - t0);
    }
    if (ORBmatcher::LastStatus() != 0) { fprintf(stderr, "matcher status %d\n", ORBmatcher::LastStatus()); return 3; }
    printf("{\"extract_ms_with_pyramid_mirror\": %.4f, \"extract_ms_no_mirror\": %.4f, \"keypoints\": %d, "
           "\"search_by_projection_ms\": %.4f, \"projected_points\": %d, \"matches\": %d, \"iterations\": %d, \"image\": \"%dx%d\"}\n",
           msMirror, msPlain, N, median(t), N, nm, iters, w, h);
    return 0;
}
