// TEST INFRASTRUCTURE ONLY (oracle/): cv::FAST of the cv:: shim (see ../core.hpp).
#ifndef ORB_REF_CVSHIM_FEATURES2D_HPP
#define ORB_REF_CVSHIM_FEATURES2D_HPP
#include "../core.hpp"
namespace cv {
// cv::FAST(image, keypoints, threshold, nonmaxSuppression = true), TYPE_9_16 (src/ORBextractor.cc:1122, :1131):
// KeyPoint((float)x, (float)y, 7.f, -1, (float)score), row-major.  orc_fast_roi is pinned to cv2.FastFeatureDetector.
static inline void FAST(InputArray img_, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression = true)
{
    Mat img = img_.getMat();
    assert(img.type() == CV_8UC1 && nonmaxSuppression);
    (void)nonmaxSuppression;
    keypoints.clear();
    const int cap = img.rows * img.cols;
    if (cap <= 0) return;
    std::vector<int> xys(3 * (size_t)cap);
    const int n = orc_fast_roi(img.data, img.cols, img.rows, img.step, threshold, xys.data(), cap);
    keypoints.reserve(n);
    for (int i = 0; i < n; ++i) keypoints.push_back(KeyPoint((float)xys[3 * i], (float)xys[3 * i + 1], 7.f, -1, (float)xys[3 * i + 2]));
}
// cv::KeyPointsFilter::retainBest: only used by ComputeKeyPointsOld (src/ORBextractor.cc:1461, :1495), which is not on the path.
struct KeyPointsFilter {
    static void retainBest(std::vector<KeyPoint>& kps, int n)
    {
        if (n < 0 || (int)kps.size() <= n) return;
        if (n == 0) { kps.clear(); return; }
        std::nth_element(kps.begin(), kps.begin() + n - 1, kps.end(), [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; });
        const float ambiguous = kps[n - 1].response;
        std::vector<KeyPoint>::iterator e = std::partition(kps.begin() + n, kps.end(), [ambiguous](const KeyPoint& k) { return k.response >= ambiguous; });
        kps.resize(e - kps.begin());
    }
};
} // namespace cv
#endif
