// Drop-in replacement for the reference's include/ORBmatcher.h (QiuYue-bit/ORB-SLAM2-noted, :48-264): same namespace,
// class, constructor, static members and method signatures, so Tracking / LocalMapping / LoopClosing call it unchanged.
// The Hamming work runs in hand-written sm_100a CUDA kernels behind the C ABI (include/orbx.h); the bodies in
// host/ORBmatcher.cc only flatten the SLAM objects.  Implemented here: DescriptorDistance, both SearchByBoW overloads,
// SearchForTriangulation, all four SearchByProjection overloads, SearchForInitialization, SearchBySim3 and Fuse x2: the
// complete public surface of the reference class.
#ifndef ORBMATCHER_H
#define ORBMATCHER_H

#include <set>
#include <vector>
#include <opencv2/core/core.hpp>
#include <opencv2/features2d/features2d.hpp>

#include "MapPoint.h"
#include "KeyFrame.h"
#include "Frame.h"

struct orbx_ctx;

namespace ORB_SLAM2
{

class ORBmatcher
{
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true);

    // reference src/ORBmatcher.cc:2123-2143 (computed on the host: one pair is not worth a launch)
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);

    int SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th = 3);                // :78-191
    int SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono);          // :1741-1910
    int SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound, const float th,
                           const int ORBdist);                                                                      // :1922-2065
    int SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints,
                           std::vector<MapPoint*>& vpMatched, int th);                                              // :431-560
    int SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12,
                                int windowSize = 10);                                                               // :580-748
    int SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const float& s12, const cv::Mat& R12,
                     const cv::Mat& t12, const float th);                                                           // :1472-1723
    int Fuse(KeyFrame* pKF, const std::vector<MapPoint*>& vpMapPoints, const float th = 3.0);                       // :1139-1300
    int Fuse(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, float th,
             std::vector<MapPoint*>& vpReplacePoint);                                                               // :1310-1458
    int SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches);                            // :255-419
    int SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12);                           // :760-903
    int SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12,
                               std::vector<std::pair<size_t, size_t> >& vMatchedPairs, const bool bOnlyStereo);     // :915-1127

public:
    static const int TH_LOW;
    static const int TH_HIGH;
    static const int HISTO_LENGTH;

    // the per-thread CUDA context the matcher calls run on (matchers are constructed on the stack at every call site and
    // used from the tracking, mapping and loop-closing threads concurrently)
    static orbx_ctx* Context();
    // ---- additions (not in the reference) ----
    static void SetDevice(int device);   // CUDA device of the calling thread's context (default: env ORBX_DEVICE, else 0)
    // Like the reference, no method throws or returns a code.  A library error (no device, a CUDA failure) makes the call
    // report no matches, prints one line to stderr, and is kept here per thread (0 = ORBX_OK) until a call succeeds.
    static int LastStatus();

protected:
    float RadiusByViewingCos(const float& viewCos);
    float mfNNratio;
    bool mbCheckOrientation;
};

} // namespace ORB_SLAM2

#endif // ORBMATCHER_H
