// Drop-in replacement for the reference's include/ORBextractor.h (QiuYue-bit/ORB-SLAM2-noted):
// same namespace, class name, constructor, operator(), getters and public mvImagePyramid
// (reference include/ORBextractor.h:75-158), so src/Frame.cc:128-153,430-443 and src/Tracking.cc:166-190
// compile and behave unchanged.  The work is done by hand-written sm_100a CUDA kernels behind the C ABI in
// include/orbx.h (liborbx.so); this class is host glue only.  There is no CPU fallback and, like the reference, no
// exception: without a CUDA device every operator() call returns no keypoints and LastStatus() says why.
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <vector>
#include <opencv2/core.hpp>   // real OpenCV, or compat/opencv2/core.hpp when OpenCV C++ is not installed

struct orbx_ctx;

namespace ORB_SLAM2
{

class ORBextractor
{
public:
    // reference: src/ORBextractor.cc:471-579
    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
    ~ORBextractor();
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    // reference: src/ORBextractor.cc:1543-1658.  `mask` is ignored, as in the reference.
    void operator()(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint>& keypoints, cv::OutputArray descriptors);

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return scaleFactor; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    // Level images of the last frame, each a ROI inside its 19-px reflect-101 frame exactly like
    // src/ORBextractor.cc:1704-1727 (Frame::ComputeStereoMatches reads them, src/Frame.cc:853,1000-1040).
    std::vector<cv::Mat> mvImagePyramid;

    // ---- additions (not in the reference) ----
    // The host mirror of the pyramid costs one device-to-host copy per level; callers that use
    // orbx_stereo_match (device pyramids) can switch it off.
    void SetHostPyramidMirror(bool on) { mbMirrorPyramid = on; }
    void SetDevice(int device);            // must be called before the first operator(); default device 0
    orbx_ctx* Context() { return mpCtx; }  // for orbx_stereo_match / orbx_match_* calls
    // Like the reference, nothing here throws or returns a code (src/ORBextractor.cc:1547-1553 returns silently).  When the
    // library reports an error (no device, a CUDA failure) the call leaves `keypoints` empty and `descriptors` released,
    // prints one line to stderr the first time, and keeps the status here (0 = ORBX_OK) until the next successful call.
    int LastStatus() const { return mnLastStatus; }

protected:
    bool EnsureContext(int w, int h);
    bool Ok(int rc, const char* what);

    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;
    std::vector<int> mnFeaturesPerLevel;
    std::vector<float> mvScaleFactor;
    std::vector<float> mvInvScaleFactor;
    std::vector<float> mvLevelSigma2;
    std::vector<float> mvInvLevelSigma2;

    orbx_ctx* mpCtx;
    int mnDevice, mnMaxW, mnMaxH;
    bool mbMirrorPyramid;
    std::vector<cv::Mat> mvBordered; // the bordered level images: views into mvPyramidBlock
    std::vector<unsigned char> mvPyramidBlock; // all bordered levels of the last frame, back to back (one copy per frame)
    int mnLastStatus;
};

} // namespace ORB_SLAM2

#endif
