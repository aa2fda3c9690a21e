// Host glue of the drop-in ORB_SLAM2::ORBextractor (include/ORBextractor.h) over the C ABI (include/orbx.h).
// Replaces the reference's src/ORBextractor.cc; compile this file instead of it and link liborbx.so.
#include "ORBextractor.h"
#include "orbx.h"
#include <cassert>
#include <cstdio>
#include <cstring>

namespace ORB_SLAM2
{

static const int EDGE_THRESHOLD = 19; // src/ORBextractor.cc:81

// The reference's contract is "no exceptions, no return codes" (Tracking calls operator() from its own thread and checks
// nothing): an error from the library is recorded, reported once on stderr and turned into "no keypoints".
bool ORBextractor::Ok(int rc, const char* what)
{
    if (rc == ORBX_OK) return true;
    if (mnLastStatus == ORBX_OK)
        fprintf(stderr, "ORBextractor: %s failed: %s %s\n", what, orbx_strerror(rc), mpCtx ? orbx_last_cuda_error(mpCtx) : "");
    mnLastStatus = rc;
    return false;
}

ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST), minThFAST(_minThFAST),
      mpCtx(nullptr), mnDevice(0), mnMaxW(0), mnMaxH(0), mbMirrorPyramid(true), mnLastStatus(0)
{
    // The tables come from the library (same formulas as src/ORBextractor.cc:484-531) so that the getters,
    // which Frame's constructors copy (src/Frame.cc:128-140), agree bit for bit with what the kernels use.
    EnsureContext(64, 64);          // a failure (no CUDA device) is remembered in LastStatus(); operator() then returns nothing
}

ORBextractor::~ORBextractor() { if (mpCtx) orbx_destroy(mpCtx); }

void ORBextractor::SetDevice(int device)
{
    if (device == mnDevice) return;
    if (mpCtx) { orbx_destroy(mpCtx); mpCtx = nullptr; }
    mnDevice = device; mnMaxW = mnMaxH = 0;
    EnsureContext(64, 64);
}

bool ORBextractor::EnsureContext(int w, int h)
{
    if (mpCtx && w <= mnMaxW && h <= mnMaxH) return true;
    if (mpCtx) { orbx_destroy(mpCtx); mpCtx = nullptr; }
    mnMaxW = w > mnMaxW ? w : mnMaxW; mnMaxH = h > mnMaxH ? h : mnMaxH;
    mvScaleFactor.assign(nlevels, 1.f); mvInvScaleFactor.assign(nlevels, 1.f); mvLevelSigma2.assign(nlevels, 1.f); mvInvLevelSigma2.assign(nlevels, 1.f);
    mnFeaturesPerLevel.assign(nlevels, 0);
    mvImagePyramid.resize(nlevels);
    mvBordered.resize(nlevels);
    if (!Ok(orbx_create(&mpCtx, mnDevice, nfeatures, (float)scaleFactor, nlevels, iniThFAST, minThFAST, mnMaxW, mnMaxH, 1), "orbx_create")) {
        mpCtx = nullptr; mnMaxW = mnMaxH = 0;
        return false;
    }
    return Ok(orbx_get_tables(mpCtx, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(), mvInvLevelSigma2.data(),
                              mnFeaturesPerLevel.data()), "orbx_get_tables");
}

void ORBextractor::operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints, cv::OutputArray _descriptors)
{
    if (_image.empty()) return;                       // src/ORBextractor.cc:1547-1548
    cv::Mat image = _image.getMat();
    assert(image.type() == CV_8UC1);                  // :1553
    if (!EnsureContext(image.cols, image.rows)) { _keypoints.clear(); _descriptors.release(); return; }

    const int cap = orbx_max_keypoints(mpCtx);
    _keypoints.resize(cap);
    cv::Mat desc(cap, 32, CV_8U);
    int n = 0;
    static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint must be the 28-byte POD the kernels write");
    if (!Ok(orbx_extract(mpCtx, image.data, image.cols, image.rows, image.step, 0, reinterpret_cast<orbx_keypoint*>(_keypoints.data()),
                         desc.data, cap, &n), "orbx_extract")) { _keypoints.clear(); _descriptors.release(); return; }
    mnLastStatus = ORBX_OK;
    _keypoints.resize(n);
    if (n == 0) _descriptors.release();               // :1582
    else {
        _descriptors.create(n, 32, CV_8U);            // :1586
        cv::Mat out = _descriptors.getMat();
        for (int i = 0; i < n; ++i) memcpy(out.ptr<unsigned char>(i), desc.ptr<unsigned char>(i), 32);
    }
    if (mbMirrorPyramid) {
        // every level with its 19-px frame in one device pass and one copy; the Mats are views into one host block
        std::vector<size_t> off(nlevels);
        size_t total = 0;
        if (!Ok(orbx_pyramid_bordered(mpCtx, 0, EDGE_THRESHOLD, 0, nullptr, 0, off.data(), &total), "orbx_pyramid_bordered")) return;
        if (mvPyramidBlock.size() != total) mvPyramidBlock.resize(total);
        if (!Ok(orbx_pyramid_bordered(mpCtx, 0, EDGE_THRESHOLD, 0, mvPyramidBlock.data(), total, off.data(), &total), "orbx_pyramid_bordered")) return;
        for (int level = 0; level < nlevels; ++level) {
            int w = 0, h = 0;
            if (!Ok(orbx_level_size(mpCtx, level, image.cols, image.rows, &w, &h), "orbx_level_size")) return;
            mvBordered[level] = cv::Mat(h + 2 * EDGE_THRESHOLD, w + 2 * EDGE_THRESHOLD, CV_8UC1, mvPyramidBlock.data() + off[level],
                                        (size_t)(w + 2 * EDGE_THRESHOLD));
            mvImagePyramid[level] = mvBordered[level](cv::Rect(EDGE_THRESHOLD, EDGE_THRESHOLD, w, h)); // :1677,1704-1727
        }
    }
}

} // namespace ORB_SLAM2
