// TEST INFRASTRUCTURE ONLY (oracle/).  A small cv:: namespace that lets the UNMODIFIED reference sources
// (/root/reference/src/ORBextractor.cc, ORBmatcher.cc, Frame.cc, KeyFrame.cc, MapPoint.cc, Map.cc,
// KeyFrameDatabase.cc and Thirdparty/DBoW2) compile in an image without OpenCV C++.  The recipe is oracle/Makefile
// (target `ref`), the output oracle/_ref/liborb_ref.so.
//
// Arithmetic contract.  The image primitives the reference calls are bound to the oracle's restatements, which
// tests/test_oracle_golden.py pins bit for bit to the executable cv2 4.13 of this image:
//   cv::resize -> orc_resize_linear, cv::copyMakeBorder -> orc_border_reflect101, cv::GaussianBlur -> orc_gauss_blur,
//   cv::FAST -> orc_fast_roi, cv::fastAtan2 -> orc_fast_atan2, cv::undistortPoints -> orc_undistort_keypoints.
// cosf/sinf/sqrt/log come from the host libm, as in a real build.  The float matrix algebra follows OpenCV 4.x:
//   * A*B without transposes and inner length 2..4: float products and sums left to right (gemm's small-matrix path);
//   * every other product (a transposed operand, as in -Rcw.t()*tcw): double accumulation, one rounding to float;
//   * Mat::dot, cv::norm: double accumulation;  Mat*s, Mat/s: element * (float)s resp. (float)(1.0/s) (cvtScale, wtype float);
//   * cvRound = round half to even (cvtsd2si).
// No FMA anywhere (compile with -ffp-contract=off, like the reference's -O3 x86-64 baseline).
#ifndef ORB_REF_CVSHIM_CORE_HPP
#define ORB_REF_CVSHIM_CORE_HPP
#include <algorithm>
#include <cassert>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>
#include "orb_oracle.h"

#define CV_8U 0
#define CV_8S 1
#define CV_16U 2
#define CV_16S 3
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_CN_SHIFT 3
#define CV_MAT_DEPTH(t) ((t) & 7)
#define CV_MAT_CN(t) ((((t) >> CV_CN_SHIFT) & 511) + 1)
#define CV_MAKETYPE(d, cn) (CV_MAT_DEPTH(d) + (((cn) - 1) << CV_CN_SHIFT))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_32FC2 CV_MAKETYPE(CV_32F, 2)
#define CV_64FC1 CV_MAKETYPE(CV_64F, 1)
#define CV_PI 3.1415926535897932384626433832795

typedef unsigned char uchar;

static inline int cvRound(double v) { return (int)lrint(v); }
static inline int cvRound(float v) { return (int)lrintf(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { return (int)std::floor(v); }
static inline int cvCeil(double v) { return (int)std::ceil(v); }

namespace cv {

typedef ::uchar uchar;
using std::string;
typedef std::string String;

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T a, T b) : x(a), y(b) {}
    template <typename U> Point_(const Point_<U>& o) : x((T)o.x), y((T)o.y) {}
    Point_& operator+=(const Point_& o) { x += o.x; y += o.y; return *this; }
    Point_& operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }
};
template <typename T> static inline Point_<T> operator+(const Point_<T>& a, const Point_<T>& b) { return Point_<T>(a.x + b.x, a.y + b.y); }
template <typename T> static inline Point_<T> operator-(const Point_<T>& a, const Point_<T>& b) { return Point_<T>(a.x - b.x, a.y - b.y); }
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;
typedef Point_<double> Point2d;
template <typename T> struct Point3_ { T x, y, z; Point3_() : x(0), y(0), z(0) {} Point3_(T a, T b, T c) : x(a), y(b), z(c) {} };
typedef Point3_<float> Point3f;
typedef Point3_<double> Point3d;
template <typename T> struct Size_ { T width, height; Size_() : width(0), height(0) {} Size_(T w, T h) : width(w), height(h) {} };
typedef Size_<int> Size;
template <typename T> struct Rect_ { T x, y, width, height; Rect_() : x(0), y(0), width(0), height(0) {} Rect_(T a, T b, T w, T h) : x(a), y(b), width(w), height(h) {} };
typedef Rect_<int> Rect;
struct Range { int start, end; Range() : start(0), end(0) {} Range(int a, int b) : start(a), end(b) {} };
template <typename T> struct Scalar_ { T val[4]; Scalar_() { val[0] = val[1] = val[2] = val[3] = 0; } Scalar_(T a) { val[0] = a; val[1] = val[2] = val[3] = 0; } };
typedef Scalar_<double> Scalar;

class KeyPoint { // 28-byte layout of OpenCV's
public:
    Point2f pt; float size; float angle; float response; int octave; int class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(Point2f p, float s, float a = -1, float r = 0, int o = 0, int c = -1) : pt(p), size(s), angle(a), response(r), octave(o), class_id(c) {}
    KeyPoint(float x, float y, float s, float a = -1, float r = 0, int o = 0, int c = -1) : pt(x, y), size(s), angle(a), response(r), octave(o), class_id(c) {}
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

enum { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3, BORDER_REFLECT_101 = 4, BORDER_REFLECT101 = 4,
       BORDER_DEFAULT = 4, BORDER_ISOLATED = 16 };
enum { INTER_NEAREST = 0, INTER_LINEAR = 1, INTER_CUBIC = 2, INTER_AREA = 3 };
enum { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4 };
enum { DECOMP_LU = 0 };

struct MatStep {
    size_t p[2];
    MatStep() { p[0] = p[1] = 0; }
    operator size_t() const { return p[0]; }
    size_t operator[](int i) const { return p[i]; }
    MatStep& operator=(size_t s) { p[0] = s; return *this; }
};

class Mat;
class MatR;   // the value of a matrix expression (see below)
class MatT;   // A.t() (possibly scaled) before it is materialised: a product with it takes gemm's general path
class _OutputArray;

class Mat {
public:
    int flags, dims, rows, cols;
    uchar* data;
    MatStep step;

    Mat() : flags(0), dims(0), rows(0), cols(0), data(nullptr), sub_(false) {}
    Mat(int r, int c, int type) : Mat() { create(r, c, type); }
    Mat(Size s, int type) : Mat() { create(s.height, s.width, type); }
    Mat(int r, int c, int type, const Scalar& s) : Mat() { create(r, c, type); setTo(s.val[0]); }
    Mat(int r, int c, int type, void* ext, size_t stp = 0) : flags(type), dims(2), rows(r), cols(c), data((uchar*)ext), sub_(false)
    {
        step.p[1] = esz(type);
        step.p[0] = stp ? stp : (size_t)c * esz(type);
    }
    Mat(const Mat& m, const Rect& r) : Mat(m) { data = m.data + (size_t)r.y * m.step.p[0] + (size_t)r.x * m.elemSize(); rows = r.height; cols = r.width; sub_ = true; }
    Mat(const MatT& t);
    Mat& operator=(const MatT& t);
    Mat& operator=(const MatR& r);   // cv::Mat = cv::MatExpr: create() (a matching buffer, also a ROI view, is kept) and write in place

    void create(int r, int c, int type)
    {
        if (data && r == rows && c == cols && type == this->type()) return;  // cv::Mat::create keeps a matching buffer (also a ROI view)
        flags = type; dims = 2; rows = r; cols = c; sub_ = false;
        step.p[1] = esz(type);
        step.p[0] = (size_t)c * esz(type);
        buf_ = std::shared_ptr<uchar>(new uchar[(size_t)r * step.p[0] + 64], std::default_delete<uchar[]>());
        data = buf_.get();
    }
    void create(Size s, int type) { create(s.height, s.width, type); }
    void release() { buf_.reset(); data = nullptr; rows = cols = 0; step.p[0] = 0; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return flags & 0xfff; }
    int depth() const { return CV_MAT_DEPTH(flags); }
    int channels() const { return CV_MAT_CN(flags); }
    size_t elemSize() const { return esz(type()); }
    size_t elemSize1() const { return esz(depth()); }
    size_t step1(int i = 0) const { return step.p[i] / elemSize1(); }
    size_t total() const { return (size_t)rows * cols; }
    bool isContinuous() const { return rows <= 1 || step.p[0] == (size_t)cols * elemSize(); }
    bool isSubmatrix() const { return sub_; }
    Size size() const { return Size(cols, rows); }

    Mat clone() const { Mat m; copyData(m); return m; }
    void copyTo(const _OutputArray& dst) const;
    void convertTo(const _OutputArray& dst, int rtype, double alpha = 1, double beta = 0) const;
    Mat reshape(int cn, int newRows = 0) const
    {
        assert(newRows == 0 && isContinuous());
        Mat m(*this);
        const int totalCh = cols * channels();
        assert(totalCh % cn == 0);
        m.flags = CV_MAKETYPE(depth(), cn); m.cols = totalCh / cn; m.step.p[1] = esz(m.flags);
        return m;
    }
    Mat operator()(const Rect& r) const { return Mat(*this, r); }
    Mat operator()(const Range& rr, const Range& cr) const { return Mat(*this, Rect(cr.start, rr.start, cr.end - cr.start, rr.end - rr.start)); }
    Mat row(int y) const { return Mat(*this, Rect(0, y, cols, 1)); }
    Mat col(int x) const { return Mat(*this, Rect(x, 0, 1, rows)); }
    Mat rowRange(int a, int b) const { return Mat(*this, Rect(0, a, cols, b - a)); }
    Mat colRange(int a, int b) const { return Mat(*this, Rect(a, 0, b - a, rows)); }
    uchar* ptr(int y = 0) { return data + (size_t)y * step.p[0]; }
    const uchar* ptr(int y = 0) const { return data + (size_t)y * step.p[0]; }
    template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step.p[0]); }
    template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step.p[0]); }
    template <typename T> T& at(int y, int x) { return ((T*)(data + (size_t)y * step.p[0]))[x]; }
    template <typename T> const T& at(int y, int x) const { return ((const T*)(data + (size_t)y * step.p[0]))[x]; }
    template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    template <typename T> const T& at(int i) const { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    template <typename T> T& at(Point p) { return at<T>(p.y, p.x); }

    Mat& setTo(double v)
    {
        for (int y = 0; y < rows; ++y)
            for (int x = 0; x < cols * channels(); ++x) {
                if (depth() == CV_32F) ptr<float>(y)[x] = (float)v;
                else if (depth() == CV_64F) ptr<double>(y)[x] = v;
                else if (depth() == CV_32S) ptr<int>(y)[x] = (int)v;
                else ptr<uchar>(y)[x] = (uchar)v;
            }
        return *this;
    }
    Mat& operator=(const Scalar& s) { return setTo(s.val[0]); }
    static MatR zeros(int r, int c, int type);
    static MatR zeros(Size s, int type);
    static MatR ones(int r, int c, int type);
    static MatR eye(int r, int c, int type);

    MatT t() const;
    MatR inv(int = DECOMP_LU) const;
    double dot(const Mat& m) const
    {   // dotProd_32f for short vectors: double products, double sum, in order
        assert(depth() == CV_32F && m.depth() == CV_32F && total() == m.total());
        double r = 0;
        for (int y = 0; y < rows; ++y)
            for (int x = 0; x < cols; ++x) {
                const int i = y * cols + x;
                r += (double)at<float>(y, x) * (double)m.at<float>(i / m.cols, i % m.cols);
            }
        return r;
    }
    void push_back(const Mat& m);

    static size_t esz(int type)
    {
        static const size_t d[8] = {1, 1, 2, 2, 4, 4, 8, 2};
        return d[CV_MAT_DEPTH(type)] * CV_MAT_CN(type);
    }
protected:
    void copyData(Mat& m) const
    {
        m.create(rows, cols, type());
        const size_t n = (size_t)cols * elemSize();
        for (int y = 0; y < rows; ++y) std::memmove(m.data + (size_t)y * m.step.p[0], data + (size_t)y * step.p[0], n);
    }
    std::shared_ptr<uchar> buf_;
    bool sub_;
};

// What a cv::MatExpr evaluates to.  It is a Mat; the only difference is assignment: `m = expr` writes into m's existing
// buffer when size and type match (descriptors.rowRange(..) = Mat::zeros(..) at src/ORBextractor.cc:1524 relies on it).
class MatR : public Mat {
public:
    MatR(const Mat& m) : Mat(m) {}
};
inline Mat& Mat::operator=(const MatR& r)
{
    if (data && rows == r.rows && cols == r.cols && type() == r.type()) {
        const size_t n = (size_t)cols * elemSize();
        for (int y = 0; y < rows; ++y) std::memmove(data + (size_t)y * step.p[0], r.data + (size_t)y * r.step.p[0], n);
    } else *this = static_cast<const Mat&>(r);
    return *this;
}
inline MatR Mat::zeros(int r, int c, int type) { Mat m(r, c, type); m.setTo(0); return m; }
inline MatR Mat::zeros(Size s, int type) { return zeros(s.height, s.width, type); }
inline MatR Mat::ones(int r, int c, int type) { Mat m(r, c, type); m.setTo(1); return m; }
inline MatR Mat::eye(int r, int c, int type)
{
    Mat m(r, c, type); m.setTo(0);
    for (int i = 0; i < (r < c ? r : c); ++i) { if (CV_MAT_DEPTH(type) == CV_32F) m.at<float>(i, i) = 1.f; else m.at<double>(i, i) = 1.0; }
    return m;
}

template <typename T> struct DepthOf;
template <> struct DepthOf<float> { enum { value = CV_32F }; };
template <> struct DepthOf<double> { enum { value = CV_64F }; };
template <> struct DepthOf<uchar> { enum { value = CV_8U }; };
template <> struct DepthOf<int> { enum { value = CV_32S }; };

template <typename T> class Mat_ : public Mat {
public:
    Mat_() : Mat() {}
    Mat_(int r, int c) : Mat(r, c, DepthOf<T>::value) {}
    Mat_(const Mat& m) : Mat(m) {}
    T& operator()(int y, int x) { return this->template at<T>(y, x); }
    const T& operator()(int y, int x) const { return this->template at<T>(y, x); }
};
template <typename T> class MatCommaInitializer_ {
public:
    MatCommaInitializer_(const Mat_<T>& m) : m_(m), i_(0) {}
    template <typename U> MatCommaInitializer_& operator,(U v) { put((T)v); return *this; }
    template <typename U> void put(U v) { m_.template at<T>(i_ / m_.cols, i_ % m_.cols) = (T)v; ++i_; }
    operator Mat() const { return m_; }
    operator Mat_<T>() const { return m_; }
private:
    Mat_<T> m_; int i_;
};
template <typename T, typename U> static inline MatCommaInitializer_<T> operator<<(const Mat_<T>& m, U v) { MatCommaInitializer_<T> c(m); c.put((T)v); return c; }

class MatT {
public:
    MatT(const Mat& m, double a) : m_(m), alpha_(a) {}
    Mat materialise() const
    {
        Mat d(m_.cols, m_.rows, m_.type());
        assert(m_.depth() == CV_32F);
        for (int y = 0; y < m_.rows; ++y) for (int x = 0; x < m_.cols; ++x) d.at<float>(x, y) = alpha_ == 1 ? m_.at<float>(y, x) : (float)(m_.at<float>(y, x) * (float)alpha_);
        return d;
    }
    MatT operator-() const { return MatT(m_, -alpha_); }
    const Mat& src() const { return m_; }
    double alpha() const { return alpha_; }
    double dot(const Mat& o) const { return materialise().dot(o); }
private:
    Mat m_; double alpha_;
};
inline Mat::Mat(const MatT& t) : Mat() { *this = t.materialise(); }
inline Mat& Mat::operator=(const MatT& t) { *this = MatR(t.materialise()); return *this; }
inline MatT Mat::t() const { return MatT(*this, 1.0); }

// ---- arrays as function arguments
class _InputArray {
public:
    _InputArray() {}
    _InputArray(const Mat& m) : m_(m) {}
    _InputArray(const MatT& m) : m_(m.materialise()) {}
    template <typename T> _InputArray(const Mat_<T>& m) : m_(m) {}
    _InputArray(const std::vector<Point2f>& v) : m_((int)v.size(), 1, CV_32FC2, (void*)v.data()) {}
    Mat getMat(int = -1) const { return m_; }
    bool empty() const { return m_.empty(); }
private:
    Mat m_;
};
typedef const _InputArray& InputArray;

class _OutputArray {
public:
    _OutputArray(Mat& m) : m_(&m) {}
    _OutputArray(const Mat& m) : m_(const_cast<Mat*>(&m)) {}   // ROI temporaries: Rwc.copyTo(Twc.rowRange(0,3).colRange(0,3))
    void create(int r, int c, int type) const { m_->create(r, c, type); }
    void create(Size s, int type) const { m_->create(s, type); }
    void release() const { m_->release(); }
    Mat getMat(int = -1) const { return *m_; }
    Mat& getMatRef() const { return *m_; }
    bool needed() const { return true; }
private:
    Mat* m_;
};
typedef const _OutputArray& OutputArray;
typedef const _OutputArray& InputOutputArray;
static inline _InputArray noArray() { return _InputArray(); }

inline void Mat::copyTo(const _OutputArray& o) const
{
    if (empty()) { o.release(); return; }
    o.create(rows, cols, type());
    Mat& d = o.getMatRef();
    const size_t n = (size_t)cols * elemSize();
    for (int y = 0; y < rows; ++y) std::memmove(d.data + (size_t)y * d.step.p[0], data + (size_t)y * step.p[0], n);
}
inline void Mat::convertTo(const _OutputArray& o, int rtype, double alpha, double beta) const
{
    const int dd = rtype < 0 ? depth() : CV_MAT_DEPTH(rtype);
    Mat src = (o.getMatRef().data == data) ? clone() : *this;   // in-place conversion (IL.convertTo(IL, CV_32F))
    Mat d(rows, cols, CV_MAKETYPE(dd, channels()));
    const int n = cols * channels();
    for (int y = 0; y < rows; ++y)
        for (int x = 0; x < n; ++x) {
            double v;
            switch (src.depth()) {
            case CV_8U: v = src.ptr<uchar>(y)[x]; break;
            case CV_32F: v = src.ptr<float>(y)[x]; break;
            case CV_64F: v = src.ptr<double>(y)[x]; break;
            case CV_32S: v = src.ptr<int>(y)[x]; break;
            default: assert(!"convertTo: depth"); v = 0;
            }
            if (alpha != 1 || beta != 0) v = (dd == CV_64F) ? v * alpha + beta : (double)((float)v * (float)alpha + (float)beta);
            switch (dd) {
            case CV_8U: { int iv = cvRound(v); d.ptr<uchar>(y)[x] = (uchar)(iv < 0 ? 0 : iv > 255 ? 255 : iv); break; }
            case CV_32F: d.ptr<float>(y)[x] = (float)v; break;
            case CV_64F: d.ptr<double>(y)[x] = v; break;
            case CV_32S: d.ptr<int>(y)[x] = cvRound(v); break;
            default: assert(!"convertTo: depth");
            }
        }
    o.getMatRef() = d;
}
inline void Mat::push_back(const Mat& m)
{
    if (empty()) { *this = m.clone(); return; }
    assert(m.cols == cols && m.type() == type());
    Mat d(rows + m.rows, cols, type());
    rowRange(0, rows).copyTo(d.rowRange(0, rows));
    m.copyTo(d.rowRange(rows, rows + m.rows));
    *this = d;
}

// ---- float matrix algebra (eager; see the arithmetic contract at the top)
static inline MatR gemmGeneral(const Mat& A, bool tA, const Mat& B, double alpha)
{   // GEMMSingleMul<float,double>: double accumulation over k in order, d = float(s*alpha)
    assert(A.depth() == CV_32F && B.depth() == CV_32F);
    const int M = tA ? A.cols : A.rows, K = tA ? A.rows : A.cols, N = B.cols;
    assert(K == B.rows);
    Mat D(M, N, CV_32F);
    for (int i = 0; i < M; ++i)
        for (int j = 0; j < N; ++j) {
            double s = 0;
            for (int k = 0; k < K; ++k) s += (double)(tA ? A.at<float>(k, i) : A.at<float>(i, k)) * (double)B.at<float>(k, j);
            D.at<float>(i, j) = (float)(s * alpha);
        }
    return D;
}
static inline MatR operator*(const Mat& A, const Mat& B)
{
    assert(A.depth() == CV_32F && B.depth() == CV_32F && A.cols == B.rows);
    const int len = A.cols, M = A.rows, N = B.cols;
    if (len >= 2 && len <= 4 && (len == N || len == M)) {   // gemm small-matrix path: float, left to right, then *alpha(=1) in double
        Mat D(M, N, CV_32F);
        for (int i = 0; i < M; ++i)
            for (int j = 0; j < N; ++j) {
                float t = A.at<float>(i, 0) * B.at<float>(0, j);
                for (int k = 1; k < len; ++k) t = t + A.at<float>(i, k) * B.at<float>(k, j);
                D.at<float>(i, j) = (float)((double)t * 1.0);
            }
        return D;
    }
    return gemmGeneral(A, false, B, 1.0);
}
static inline MatR operator*(const MatT& A, const Mat& B) { return gemmGeneral(A.src(), true, B, A.alpha()); }
static inline MatR operator*(const MatT& A, const MatT& B) { return gemmGeneral(A.src(), true, B.materialise(), A.alpha()); }
static inline MatR operator*(const Mat& A, const MatT& B) { return A * B.materialise(); }
template <typename F> static inline MatR elementwise(const Mat& a, const Mat& b, F f)
{
    assert(a.depth() == CV_32F && b.depth() == CV_32F && a.rows == b.rows && a.cols == b.cols);
    Mat d(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) d.at<float>(y, x) = f(a.at<float>(y, x), b.at<float>(y, x));
    return d;
}
static inline MatR operator+(const Mat& a, const Mat& b) { return elementwise(a, b, [](float x, float y) { return x + y; }); }
static inline MatR operator-(const Mat& a, const Mat& b) { return elementwise(a, b, [](float x, float y) { return x - y; }); }
static inline MatR operator+(const Mat& a, const MatT& b) { return a + b.materialise(); }
static inline MatR operator-(const Mat& a, const MatT& b) { return a - b.materialise(); }
static inline MatR scaled(const Mat& a, float s)
{
    assert(a.depth() == CV_32F);
    Mat d(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) d.at<float>(y, x) = a.at<float>(y, x) * s;
    return d;
}
static inline MatR operator-(const Mat& a) { return scaled(a, -1.f); }
static inline MatR operator*(const Mat& a, double s) { return scaled(a, (float)s); }
static inline MatR operator*(double s, const Mat& a) { return scaled(a, (float)s); }
static inline MatR operator/(const Mat& a, double s) { return scaled(a, (float)(1.0 / s)); }
static inline Mat& operator+=(Mat& a, const Mat& b) { a = a + b; return a; }
static inline Mat& operator-=(Mat& a, const Mat& b) { a = a - b; return a; }
inline MatR Mat::inv(int) const
{   // only 3x3 / 4x4 float, by Gauss-Jordan in double (not on the pinned path)
    assert(rows == cols && depth() == CV_32F && rows <= 4);
    const int n = rows; double a[4][8];
    for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) { a[i][j] = at<float>(i, j); a[i][n + j] = i == j; }
    for (int c = 0; c < n; ++c) {
        int p = c; for (int r = c + 1; r < n; ++r) if (std::fabs(a[r][c]) > std::fabs(a[p][c])) p = r;
        for (int j = 0; j < 2 * n; ++j) std::swap(a[c][j], a[p][j]);
        const double d = a[c][c];
        for (int j = 0; j < 2 * n; ++j) a[c][j] /= d;
        for (int r = 0; r < n; ++r) if (r != c) { const double f = a[r][c]; for (int j = 0; j < 2 * n; ++j) a[r][j] -= f * a[c][j]; }
    }
    Mat d(n, n, CV_32F);
    for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) d.at<float>(i, j) = (float)a[i][n + j];
    return d;
}

static inline double norm(const Mat& a, int type = NORM_L2)
{   // normL2_32f / normL1_32f: double accumulation in order
    assert(a.depth() == CV_32F);
    double s = 0;
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols * a.channels(); ++x) { const double v = a.ptr<float>(y)[x]; s += type == NORM_L1 ? std::fabs(v) : v * v; }
    return type == NORM_L1 ? s : std::sqrt(s);
}
static inline double norm(const MatT& a, int type = NORM_L2) { return norm(a.materialise(), type); }
static inline double norm(const Mat& a, const Mat& b, int type = NORM_L2)
{
    assert(a.depth() == CV_32F && b.depth() == CV_32F && a.rows == b.rows && a.cols == b.cols);
    double s = 0;
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) { const double v = (double)(a.at<float>(y, x) - b.at<float>(y, x)); s += type == NORM_L1 ? std::fabs(v) : v * v; }
    return type == NORM_L1 ? s : std::sqrt(s);
}

static inline float fastAtan2(float y, float x) { return orc_fast_atan2(y, x); }

// ---- persistence: declared so that DBoW2's cv::FileStorage save/load (never called here) compile
class FileNode {
public:
    FileNode operator[](const char*) const { return FileNode(); }
    FileNode operator[](const std::string&) const { return FileNode(); }
    FileNode operator[](int) const { return FileNode(); }
    size_t size() const { return 0; }
    bool empty() const { return true; }
    operator int() const { return 0; }
    operator float() const { return 0; }
    operator double() const { return 0; }
    operator std::string() const { return std::string(); }
};
class FileStorage {
public:
    enum { READ = 0, WRITE = 1 };
    FileStorage() {}
    FileStorage(const std::string&, int) {}
    bool isOpened() const { return false; }
    void release() {}
    FileNode operator[](const char*) const { return FileNode(); }
    FileNode operator[](const std::string&) const { return FileNode(); }
};
template <typename T> static inline FileStorage& operator<<(FileStorage& f, const T&) { return f; }
template <typename T> static inline void operator>>(const FileNode&, T&) {}

} // namespace cv
#endif
