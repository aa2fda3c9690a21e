// Minimal stand-ins for the members of ORB_SLAM2::Frame / KeyFrame / MapPoint that ORBmatcher reads
// (reference include/Frame.h, include/KeyFrame.h, include/MapPoint.h).  Used ONLY to compile and test
// host/ORBmatcher.cc in this image, which has neither OpenCV C++, Eigen, g2o nor DBoW2 installed.  In the reference
// tree the real headers are used and this directory is not on the include path.
#ifndef ORBX_COMPAT_SLAM_FRAME_H
#define ORBX_COMPAT_SLAM_FRAME_H
#include <map>
#include <set>
#include <vector>
#include <opencv2/core.hpp>

namespace DBoW2 { typedef std::map<unsigned int, std::vector<unsigned int> > FeatureVector; }

namespace ORB_SLAM2
{
#define FRAME_GRID_ROWS 48
#define FRAME_GRID_COLS 64

class MapPoint
{
public:
    MapPoint() : mTrackProjX(0), mTrackProjY(0), mTrackProjXR(0), mbTrackInView(false), mnTrackScaleLevel(0), mTrackViewCos(1), mbBad(false), nObs(1) {}
    bool isBad() { return mbBad; }
    int Observations() { return nObs; }
    cv::Mat GetDescriptor() { return mDescriptor; }
    cv::Mat GetWorldPos() { return mWorldPos; }
    float mTrackProjX, mTrackProjY, mTrackProjXR;
    bool mbTrackInView;
    int mnTrackScaleLevel;
    float mTrackViewCos;
    // test-side state
    bool mbBad; int nObs; cv::Mat mDescriptor, mWorldPos;
};

class Frame
{
public:
    int N;
    cv::Mat mDescriptors;
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
    std::vector<float> mvuRight;
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    DBoW2::FeatureVector mFeatVec;
    std::vector<float> mvScaleFactors;
    cv::Mat mTcw;
    float mb, mbf;
    static float fx, fy, cx, cy;
    static float mnMinX, mnMaxX, mnMinY, mnMaxY;
    static float mfGridElementWidthInv, mfGridElementHeightInv;
};

class KeyFrame
{
public:
    int N;
    cv::Mat mDescriptors;
    std::vector<cv::KeyPoint> mvKeysUn;
    std::vector<float> mvuRight;
    DBoW2::FeatureVector mFeatVec;
    std::vector<float> mvScaleFactors, mvLevelSigma2;
    float fx, fy, cx, cy;
    std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
    MapPoint* GetMapPoint(const size_t& idx) { return mvpMapPoints[idx]; }
    cv::Mat GetCameraCenter() { return Ow; }
    cv::Mat GetRotation() { return Rcw; }
    cv::Mat GetTranslation() { return tcw; }
    // test-side state
    std::vector<MapPoint*> mvpMapPoints;
    cv::Mat Ow, Rcw, tcw;
};

} // namespace ORB_SLAM2
#endif
