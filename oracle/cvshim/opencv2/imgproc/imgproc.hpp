// TEST INFRASTRUCTURE ONLY (oracle/): image primitives of the cv:: shim, bound to the cv2-pinned oracle restatements
// (see ../core.hpp for the contract).
#ifndef ORB_REF_CVSHIM_IMGPROC_HPP
#define ORB_REF_CVSHIM_IMGPROC_HPP
#include "../core.hpp"
namespace cv {

// cv::resize, INTER_LINEAR, CV_8UC1 (called at src/ORBextractor.cc:1686).  dst keeps its buffer when it already has the
// requested size (it is a ROI of the bordered level there), as cv::Mat::create does.
static inline void resize(InputArray src_, OutputArray dst_, Size dsize, double = 0, double = 0, int interpolation = INTER_LINEAR)
{
    Mat src = src_.getMat();
    assert(interpolation == INTER_LINEAR && src.type() == CV_8UC1);
    (void)interpolation;
    dst_.create(dsize, src.type());
    Mat dst = dst_.getMat();
    orc_resize_linear(src.data, src.cols, src.rows, src.step, dst.data, dst.cols, dst.rows, dst.step);
}

// cv::copyMakeBorder, BORDER_REFLECT_101 (+ BORDER_ISOLATED), CV_8UC1 (:1704, :1725).  src may be the interior ROI of dst.
// Without BORDER_ISOLATED a ROI source would read its parent's pixels; the reference passes the caller's whole image
// at :1725, which is not a ROI in any test here (oracle decision 5 in DESIGN.md).
static inline void copyMakeBorder(InputArray src_, OutputArray dst_, int top, int bottom, int left, int right, int borderType, const Scalar& = Scalar())
{
    Mat src = src_.getMat();
    assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101 && src.type() == CV_8UC1);
    assert(top == bottom && left == right && top == left);
    assert((borderType & BORDER_ISOLATED) || !src.isSubmatrix());
    (void)borderType; (void)bottom; (void)left; (void)right;
    dst_.create(src.rows + 2 * top, src.cols + 2 * top, src.type());
    Mat dst = dst_.getMat();
    orc_border_reflect101(src.data, src.cols, src.rows, src.step, dst.data, dst.step, top);
}

// cv::GaussianBlur(Size(k,k), 2, 2, BORDER_REFLECT_101), CV_8UC1, k = 7 or 9 (:1620).  In place in the reference.
static inline void GaussianBlur(InputArray src_, OutputArray dst_, Size ksize, double sigmaX, double sigmaY = 0, int borderType = BORDER_DEFAULT)
{
    Mat src = src_.getMat().clone();
    assert(src.type() == CV_8UC1 && ksize.width == ksize.height && (ksize.width == 7 || ksize.width == 9));
    assert(sigmaX == 2 && sigmaY == 2 && borderType == BORDER_REFLECT_101);
    (void)sigmaX; (void)sigmaY; (void)borderType;
    dst_.create(src.size(), src.type());
    Mat dst = dst_.getMat();
    orc_gauss_blur(src.data, src.cols, src.rows, src.step, dst.data, dst.step, ksize.width);
}

// cv::undistortPoints(src Nx1 32FC2, dst, K, dist, R = empty, P = K) (src/Frame.cc:751, :802): 5 fixed-point iterations
// in doubles, restated as orc_undistort_keypoints and pinned to cv2.undistortPoints goldens.
static inline void undistortPoints(InputArray src_, OutputArray dst_, InputArray K_, InputArray dist_, InputArray R_ = noArray(), InputArray P_ = noArray())
{
    Mat src = src_.getMat(), K = K_.getMat(), D = dist_.getMat(), P = P_.getMat();
    assert(src.type() == CV_32FC2 && src.cols == 1 && R_.empty() && K.depth() == CV_32F && D.depth() == CV_32F);
    assert(!P.empty() && P.data == K.data);
    (void)R_;
    const int n = src.rows, nd = (int)D.total();
    std::vector<orc_keypoint> in(n > 0 ? n : 1), out(n > 0 ? n : 1);
    for (int i = 0; i < n; ++i) { std::memset(&in[i], 0, sizeof(orc_keypoint)); in[i].x = src.ptr<float>(i)[0]; in[i].y = src.ptr<float>(i)[1]; }
    float d[8] = {0};
    for (int i = 0; i < nd && i < 8; ++i) d[i] = D.rows == 1 ? D.at<float>(0, i) : D.at<float>(i, 0);
    orc_undistort_keypoints(in.data(), n, K.at<float>(0, 0), K.at<float>(1, 1), K.at<float>(0, 2), K.at<float>(1, 2), d, nd, out.data());
    dst_.create(n, 1, CV_32FC2);
    Mat dst = dst_.getMat();
    for (int i = 0; i < n; ++i) { dst.ptr<float>(i)[0] = out[i].x; dst.ptr<float>(i)[1] = out[i].y; }
}

} // namespace cv
#endif
