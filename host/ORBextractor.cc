// Host glue of the drop-in ORB_SLAM2::ORBextractor (include/ORBextractor.h) over the C ABI (include/orbx.h).
// Replaces the reference's src/ORBextractor.cc; compile this file instead of it and link liborbx.so.
#include "ORBextractor.h"
#include "orbx.h"
#include <cassert>
#include <stdexcept>
#include <string>

namespace ORB_SLAM2
{

static const int EDGE_THRESHOLD = 19; // src/ORBextractor.cc:81

static void check(int rc, orbx_ctx* ctx, const char* what)
{
    if (rc != ORBX_OK)
        throw std::runtime_error(std::string(what) + ": " + orbx_strerror(rc) + " " + (ctx ? orbx_last_cuda_error(ctx) : ""));
}

ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST), minThFAST(_minThFAST),
      mpCtx(nullptr), mnDevice(0), mnMaxW(0), mnMaxH(0), mbMirrorPyramid(true)
{
    // The tables come from the library (same formulas as src/ORBextractor.cc:484-531) so that the getters,
    // which Frame's constructors copy (src/Frame.cc:128-140), agree bit for bit with what the kernels use.
    EnsureContext(64, 64);
}

ORBextractor::~ORBextractor() { if (mpCtx) orbx_destroy(mpCtx); }

void ORBextractor::SetDevice(int device)
{
    if (device == mnDevice) return;
    if (mpCtx) { orbx_destroy(mpCtx); mpCtx = nullptr; }
    mnDevice = device; mnMaxW = mnMaxH = 0;
    EnsureContext(64, 64);
}

void ORBextractor::EnsureContext(int w, int h)
{
    if (mpCtx && w <= mnMaxW && h <= mnMaxH) return;
    if (mpCtx) { orbx_destroy(mpCtx); mpCtx = nullptr; }
    mnMaxW = w > mnMaxW ? w : mnMaxW; mnMaxH = h > mnMaxH ? h : mnMaxH;
    check(orbx_create(&mpCtx, mnDevice, nfeatures, (float)scaleFactor, nlevels, iniThFAST, minThFAST, mnMaxW, mnMaxH, 1), nullptr, "orbx_create");
    mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels); mvLevelSigma2.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
    mnFeaturesPerLevel.resize(nlevels);
    check(orbx_get_tables(mpCtx, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(), mvInvLevelSigma2.data(),
                          mnFeaturesPerLevel.data()), mpCtx, "orbx_get_tables");
    mvImagePyramid.resize(nlevels);
    mvBordered.resize(nlevels);
}

void ORBextractor::operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints, cv::OutputArray _descriptors)
{
    if (_image.empty()) return;                       // src/ORBextractor.cc:1547-1548
    cv::Mat image = _image.getMat();
    assert(image.type() == CV_8UC1);                  // :1553
    EnsureContext(image.cols, image.rows);

    const int cap = orbx_max_keypoints(mpCtx);
    _keypoints.resize(cap);
    cv::Mat desc(cap, 32, CV_8U);
    int n = 0;
    static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint must be the 28-byte POD the kernels write");
    check(orbx_extract(mpCtx, image.data, image.cols, image.rows, image.step, 0, reinterpret_cast<orbx_keypoint*>(_keypoints.data()),
                       desc.data, cap, &n), mpCtx, "orbx_extract");
    _keypoints.resize(n);
    if (n == 0) _descriptors.release();               // :1582
    else {
        _descriptors.create(n, 32, CV_8U);            // :1586
        cv::Mat out = _descriptors.getMat();
        for (int i = 0; i < n; ++i) memcpy(out.ptr<unsigned char>(i), desc.ptr<unsigned char>(i), 32);
    }
    if (mbMirrorPyramid) {
        for (int level = 0; level < nlevels; ++level) {
            int w = 0, h = 0;
            check(orbx_level_size(mpCtx, level, image.cols, image.rows, &w, &h), mpCtx, "orbx_level_size");
            cv::Mat& temp = mvBordered[level];
            temp.create(h + 2 * EDGE_THRESHOLD, w + 2 * EDGE_THRESHOLD, CV_8UC1);
            check(orbx_pyramid_level(mpCtx, 0, level, EDGE_THRESHOLD, 0, temp.data, temp.step, nullptr, nullptr), mpCtx, "orbx_pyramid_level");
            mvImagePyramid[level] = temp(cv::Rect(EDGE_THRESHOLD, EDGE_THRESHOLD, w, h)); // :1677,1704-1727
        }
    }
}

} // namespace ORB_SLAM2
