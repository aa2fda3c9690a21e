// Minimal stand-in for the parts of <opencv2/core.hpp> that the ORB front-end boundary touches
// (cv::Mat as an 8-bit / 32-bit row-major matrix with ROI views, cv::KeyPoint, Point, Size, Rect,
// InputArray / OutputArray).  It exists ONLY so that include/ORBextractor.h and include/ORBmatcher.h compile
// and can be tested on machines without OpenCV C++ (this build image has none).  With a real OpenCV
// installed, do not add compat/ to the include path: the same headers then bind to the real cv:: types.
#ifndef ORBX_COMPAT_OPENCV_CORE_HPP
#define ORBX_COMPAT_OPENCV_CORE_HPP
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#define CV_8U 0
#define CV_32F 5
#define CV_8UC1 0
#define CV_32FC1 5
#define CV_MAT_DEPTH(t) ((t) & 7)

namespace cv {

typedef unsigned char uchar;

template <typename T> struct Point_ { T x, y; Point_() : x(0), y(0) {} Point_(T a, T b) : x(a), y(b) {} };
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;
struct Size { int width, height; Size() : width(0), height(0) {} Size(int w, int h) : width(w), height(h) {} };
struct Rect { int x, y, width, height; Rect() : x(0), y(0), width(0), height(0) {} Rect(int a, int b, int w, int h) : x(a), y(b), width(w), height(h) {} };

struct KeyPoint { // same 28-byte layout as OpenCV's
    Point2f pt; float size; float angle; float response; int octave; int class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float s, float a = -1, float r = 0, int o = 0, int c = -1) : pt(x, y), size(s), angle(a), response(r), octave(o), class_id(c) {}
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

class Mat {
public:
    int rows, cols; uchar* data; size_t step;
    Mat() : rows(0), cols(0), data(nullptr), step(0), type_(0) {}
    Mat(int r, int c, int type) : Mat() { create(r, c, type); }
    Mat(int r, int c, int type, void* ext, size_t stp = 0) : rows(r), cols(c), data((uchar*)ext), step(stp ? stp : (size_t)c * esz(type)), type_(type) {}
    void create(int r, int c, int type)
    {
        if (r == rows && c == cols && type == type_ && data && isContinuous()) return;
        rows = r; cols = c; type_ = type; step = (size_t)c * esz(type);
        buf_ = std::shared_ptr<uchar>(new uchar[(size_t)r * step + 1], std::default_delete<uchar[]>());
        data = buf_.get();
    }
    void release() { buf_.reset(); data = nullptr; rows = cols = 0; step = 0; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return type_; }
    size_t elemSize() const { return esz(type_); }
    bool isContinuous() const { return step == (size_t)cols * esz(type_); }
    Size size() const { return Size(cols, rows); }
    Mat clone() const { Mat m(rows, cols, type_); for (int y = 0; y < rows; ++y) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, (size_t)cols * esz(type_)); return m; }
    Mat operator()(const Rect& r) const { Mat m(*this); m.data = data + (size_t)r.y * step + (size_t)r.x * esz(type_); m.rows = r.height; m.cols = r.width; return m; }
    Mat row(int y) const { return (*this)(Rect(0, y, cols, 1)); }
    Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
    Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }
    template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
    template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step); }
    template <typename T> T& at(int y, int x) { return ((T*)(data + (size_t)y * step))[x]; }
    template <typename T> const T& at(int y, int x) const { return ((const T*)(data + (size_t)y * step))[x]; }
    template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    template <typename T> const T& at(int i) const { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    static Mat zeros(int r, int c, int type) { Mat m(r, c, type); std::memset(m.data, 0, (size_t)r * m.step); return m; }
private:
    static size_t esz(int type) { return CV_MAT_DEPTH(type) == CV_32F ? 4 : 1; }
    int type_;
    std::shared_ptr<uchar> buf_;
};

class _InputArray {
public:
    _InputArray() : m_(nullptr) {}
    _InputArray(const Mat& m) : m_(&m) {}
    Mat getMat() const { return m_ ? *m_ : Mat(); }
    bool empty() const { return !m_ || m_->empty(); }
private:
    const Mat* m_;
};
typedef const _InputArray& InputArray;

class _OutputArray {
public:
    _OutputArray(Mat& m) : m_(&m) {}
    void create(int r, int c, int type) const { m_->create(r, c, type); }
    void release() const { m_->release(); }
    Mat getMat() const { return *m_; }
private:
    Mat* m_;
};
typedef const _OutputArray& OutputArray;

} // namespace cv
#endif
