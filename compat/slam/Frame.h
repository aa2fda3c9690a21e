// Minimal stand-ins for the members of ORB_SLAM2::Frame / KeyFrame / MapPoint that ORBmatcher reads
// (reference include/Frame.h, include/KeyFrame.h, include/MapPoint.h).  Used ONLY to compile and test
// host/ORBmatcher.cc in this image, which has neither OpenCV C++, Eigen, g2o nor DBoW2 installed.  In the reference
// tree the real headers are used and this directory is not on the include path.
#ifndef ORBX_COMPAT_SLAM_FRAME_H
#define ORBX_COMPAT_SLAM_FRAME_H
#include <cmath>
#include <map>
#include <set>
#include <vector>
#include <opencv2/core.hpp>

namespace DBoW2 { typedef std::map<unsigned int, std::vector<unsigned int> > FeatureVector; typedef std::map<unsigned int, double> BowVector; }

namespace ORB_SLAM2
{
#define FRAME_GRID_ROWS 48
#define FRAME_GRID_COLS 64

class Frame;
class KeyFrame;

class MapPoint
{
public:
    MapPoint() : mTrackProjX(0), mTrackProjY(0), mTrackProjXR(0), mbTrackInView(false), mnTrackScaleLevel(0), mTrackViewCos(1), mbBad(false), nObs(1),
                 mfMaxDistance(0), mfMinDistance(0) {}
    float GetMinDistanceInvariance() { return 0.8f * mfMinDistance; }   // src/MapPoint.cc:549-559
    float GetMaxDistanceInvariance() { return 1.2f * mfMaxDistance; }
    cv::Mat GetNormal() { return mNormalVector; }
    // observation bookkeeping, reduced to what ORBmatcher::Fuse / SearchBySim3 touch (src/MapPoint.cc:139-330)
    bool IsInKeyFrame(KeyFrame* pKF) { return mObs.count(pKF) != 0; }
    int GetIndexInKeyFrame(KeyFrame* pKF) { return mObs.count(pKF) ? (int)mObs[pKF] : -1; }
    void AddObservation(KeyFrame* pKF, size_t idx);                      // :139-150 (+2 for a stereo keypoint)
    void Replace(MapPoint* pMP);
    int PredictScale(const float& currentDist, KeyFrame* pKF);          // src/MapPoint.cc:584-604
    int PredictScale(const float& currentDist, Frame* pF);              // :613-631
    bool isBad() { return mbBad; }
    int Observations() { return nObs; }
    cv::Mat GetDescriptor() { return mDescriptor; }
    cv::Mat GetWorldPos() { return mWorldPos; }
    float mTrackProjX, mTrackProjY, mTrackProjXR;
    bool mbTrackInView;
    int mnTrackScaleLevel;
    float mTrackViewCos;
    // test-side state
    bool mbBad; int nObs; cv::Mat mDescriptor, mWorldPos, mNormalVector; float mfMaxDistance, mfMinDistance;
    std::map<KeyFrame*, size_t> mObs; MapPoint* mpReplaced = 0;
};

class Frame
{
public:
    long unsigned int mnId = 0;
    DBoW2::BowVector mBowVec;
    int N;
    cv::Mat mDescriptors;
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
    std::vector<float> mvuRight;
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    DBoW2::FeatureVector mFeatVec;
    std::vector<float> mvScaleFactors;
    int mnScaleLevels; float mfLogScaleFactor;
    cv::Mat mTcw;
    float mb, mbf;
    static float fx, fy, cx, cy;
    static float mnMinX, mnMaxX, mnMinY, mnMaxY;
    static float mfGridElementWidthInv, mfGridElementHeightInv;
};

class KeyFrame
{
public:
    long unsigned int mnId = 0;
    DBoW2::BowVector mBowVec;
    // covisibility graph, reduced to what KeyFrameDatabase reads (src/KeyFrame.cc:225-233, :248-258)
    std::set<KeyFrame*> GetConnectedKeyFrames() { std::set<KeyFrame*> s; for (std::map<KeyFrame*, int>::iterator it = mConnectedKeyFrameWeights.begin(); it != mConnectedKeyFrameWeights.end(); ++it) s.insert(it->first); return s; }
    std::vector<KeyFrame*> GetBestCovisibilityKeyFrames(const int& n) { return (int)mvpOrderedConnectedKeyFrames.size() < n ? mvpOrderedConnectedKeyFrames : std::vector<KeyFrame*>(mvpOrderedConnectedKeyFrames.begin(), mvpOrderedConnectedKeyFrames.begin() + n); }
    std::map<KeyFrame*, int> mConnectedKeyFrameWeights;
    std::vector<KeyFrame*> mvpOrderedConnectedKeyFrames;
    int N;
    cv::Mat mDescriptors;
    std::vector<cv::KeyPoint> mvKeysUn;
    std::vector<float> mvuRight;
    DBoW2::FeatureVector mFeatVec;
    std::vector<float> mvScaleFactors, mvLevelSigma2;
    float fx, fy, cx, cy;
    float mbf; std::vector<float> mvInvLevelSigma2;
    void AddMapPoint(MapPoint* pMP, const size_t& idx) { mvpMapPoints[idx] = pMP; }
    std::set<MapPoint*> GetMapPoints() { std::set<MapPoint*> s; for (size_t i = 0; i < mvpMapPoints.size(); ++i) if (mvpMapPoints[i] && !mvpMapPoints[i]->isBad()) s.insert(mvpMapPoints[i]); return s; }
    int mnScaleLevels; float mfLogScaleFactor;
    int mnGridCols, mnGridRows; float mfGridElementWidthInv, mfGridElementHeightInv;
    int mnMinX, mnMinY, mnMaxX, mnMaxY;
    bool IsInImage(const float& x, const float& y) const { return (x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY); } // src/KeyFrame.cc:841
    std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
    MapPoint* GetMapPoint(const size_t& idx) { return mvpMapPoints[idx]; }
    cv::Mat GetCameraCenter() { return Ow; }
    cv::Mat GetRotation() { return Rcw; }
    cv::Mat GetTranslation() { return tcw; }
    // test-side state
    std::vector<MapPoint*> mvpMapPoints;
    cv::Mat Ow, Rcw, tcw;
};

inline void MapPoint::AddObservation(KeyFrame* pKF, size_t idx)
{
    if (mObs.count(pKF)) return;
    mObs[pKF] = idx;
    nObs += (idx < pKF->mvuRight.size() && pKF->mvuRight[idx] >= 0) ? 2 : 1;
}

// MapPoint::Replace, src/MapPoint.cc:257-311, reduced: this point goes bad, its observations move to pMP
inline void MapPoint::Replace(MapPoint* pMP)
{
    if (pMP == this) return;
    mbBad = true; mpReplaced = pMP;
    std::map<KeyFrame*, size_t> obs = mObs;
    mObs.clear();
    for (std::map<KeyFrame*, size_t>::iterator it = obs.begin(); it != obs.end(); ++it) {
        KeyFrame* kf = it->first;
        if (!pMP->IsInKeyFrame(kf)) { kf->mvpMapPoints[it->second] = pMP; pMP->AddObservation(kf, it->second); }
        else kf->mvpMapPoints[it->second] = 0;
    }
}

inline int predictScaleImpl(float maxDistance, float currentDist, float logScaleFactor, int nLevels)
{
    const float ratio = maxDistance / currentDist;
    int nScale = (int)ceil(log(ratio) / logScaleFactor);
    if (nScale < 0) nScale = 0; else if (nScale >= nLevels) nScale = nLevels - 1;
    return nScale;
}
inline int MapPoint::PredictScale(const float& d, KeyFrame* pKF) { return predictScaleImpl(mfMaxDistance, d, pKF->mfLogScaleFactor, pKF->mnScaleLevels); }
inline int MapPoint::PredictScale(const float& d, Frame* pF) { return predictScaleImpl(mfMaxDistance, d, pF->mfLogScaleFactor, pF->mnScaleLevels); }

} // namespace ORB_SLAM2
#endif
