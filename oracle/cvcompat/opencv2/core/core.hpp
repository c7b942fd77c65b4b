/*
 * oracle/cvcompat -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A minimal stand-in for the slice of the OpenCV C++ API that the reference's
 * src/ORBextractor.cc and include/ORBextractor.h use, so that the reference translation unit can
 * be compiled where it lies (this image has no OpenCV C++ headers or libraries; only the cv2
 * Python wheel).  Types follow OpenCV's public semantics: cv::Mat is a ref-counted header over a
 * u8 plane with ROI views; assignment from Mat::zeros() fills an already-allocated destination in
 * place (as cv::MatExpr does), which the reference relies on at src/ORBextractor.cc:1313.
 * The five image primitives forward to oracle/cvprim.c, pinned bit-exact to OpenCV 4.13.0.
 * For the reference's src/Frame.cc (oracle/_ref/libreframe.so) cv::Mat also carries an element type
 * (8-bit or float, 1-2 channels) with the small float algebra that file uses (t(), *, +, -, norm,
 * Mat_<float> comma initialiser, reshape); cv::undistortPoints forwards to oracle/frame_oracle.c,
 * pinned bit-exact to OpenCV 4.13.0.
 *
 * The same header doubles as the "no OpenCV on this box" build of the product's C++ host shim
 * (orb_slam2_test_b200/shim), which needs nothing beyond Mat/KeyPoint/InputArray/OutputArray.
 */
#ifndef ORACLE_CVCOMPAT_CORE_HPP
#define ORACLE_CVCOMPAT_CORE_HPP

#include <algorithm>
#include <cassert>
#include <climits>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <iterator>
#include <memory>
#include <sstream> /* the real core.hpp pulls these in; DBoW2 relies on it */
#include <iostream>
#include <string>
#include <vector>

#if defined(__SSE2__)
#include <emmintrin.h>
#endif

#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0
#define CV_8UC3 16
#define CV_8UC4 24
#define CV_32F 5
#define CV_32FC1 5
#define CV_32FC2 13
#define CVCOMPAT_SHIM 1

typedef unsigned char uchar;

static inline int cvRound(double v)
{
#if defined(__SSE2__)
    return _mm_cvtsd_si32(_mm_set_sd(v));
#else
    return (int)lrint(v);
#endif
}
static inline int cvRound(float v)
{
#if defined(__SSE2__)
    return _mm_cvtss_si32(_mm_set_ss(v));
#else
    return (int)lrintf(v);
#endif
}
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v)
{
    int i = (int)v;
    return i - (v < i);
}
static inline int cvCeil(double v)
{
    int i = (int)v;
    return i + (v > i);
}

namespace cv {

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
    template <typename U> Point_(const Point_<U> &o) : x((T)o.x), y((T)o.y) {}
    Point_ &operator*=(float s)
    {
        x = (T)(x * s);
        y = (T)(y * s);
        return *this;
    }
};
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};

struct Rect {
    int x, y, width, height;
    Rect() : x(0), y(0), width(0), height(0) {}
    Rect(int _x, int _y, int w, int h) : x(_x), y(_y), width(w), height(h) {}
};

struct KeyPoint {
    Point2f pt;
    float size;
    float angle;
    float response;
    int octave;
    int class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0,
             int _octave = 0, int _class_id = -1)
        : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave),
          class_id(_class_id) {}
};

struct MatZeros { int rows, cols, type; };

class Mat {
public:
    int rows, cols;
    uchar *data;
    size_t step;

    Mat() : rows(0), cols(0), data(nullptr), step(0), type_(CV_8UC1) {}
    Mat(Size sz, int type) : rows(0), cols(0), data(nullptr), step(0), type_(type) { create(sz.height, sz.width, type); }
    Mat(int r, int c, int type) : rows(0), cols(0), data(nullptr), step(0), type_(type) { create(r, c, type); }
    /* header over caller-owned memory (no ownership) */
    Mat(int r, int c, int type, void *ext, size_t ext_step = 0)
        : rows(r), cols(c), data((uchar *)ext), step(0), type_(type)
    {
        step = ext_step ? ext_step : (size_t)c * elemSize();
    }

    void create(int r, int c, int type)
    {
        if (data && r == rows && c == cols && type == type_) return;
        type_ = type;
        rows = r;
        cols = c;
        step = (size_t)c * elemSize();
        const size_t n = (size_t)r * step;
        store_.reset(n ? new uchar[n] : nullptr, std::default_delete<uchar[]>());
        data = store_.get();
    }
    void create(Size sz, int type) { create(sz.height, sz.width, type); }
    void release()
    {
        store_.reset();
        data = nullptr;
        rows = cols = 0;
        step = 0;
    }

    int type() const { return type_; }
    int channels() const { return (type_ >> 3) + 1; }
    size_t elemSize() const { return (size_t)channels() * ((type_ & 7) == CV_32F ? 4 : 1); }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    size_t step1() const { return step; }
    bool isContinuous() const { return step == (size_t)cols * elemSize() || rows <= 1; }
    size_t total() const { return (size_t)rows * (size_t)cols; }

    Mat operator()(const Rect &r) const
    {
        Mat m(*this);
        m.data = data + (size_t)r.y * step + (size_t)r.x * elemSize();
        m.rows = r.height;
        m.cols = r.width;
        return m;
    }
    Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
    Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }
    Mat row(int y) const { return rowRange(y, y + 1); }
    Mat col(int x) const { return colRange(x, x + 1); }

    Mat clone() const
    {
        Mat m(rows, cols, type_);
        for (int y = 0; y < rows; ++y)
            std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, (size_t)cols * elemSize());
        return m;
    }
    void copyTo(Mat &dst) const
    {
        dst.create(rows, cols, type_);
        for (int y = 0; y < rows; ++y)
            std::memcpy(dst.data + (size_t)y * dst.step, data + (size_t)y * step, (size_t)cols * elemSize());
    }
    /* same data, `cn` channels per element (rows unchanged), as cv::Mat::reshape(cn) on a continuous matrix */
    Mat reshape(int cn) const
    {
        Mat m(*this);
        const int scalars = cols * channels();
        m.type_ = (type_ & 7) | ((cn - 1) << 3);
        m.cols = scalars / cn;
        return m;
    }

    template <typename T> T &at(int y, int x) { return *(T *)(data + (size_t)y * step + (size_t)x * sizeof(T)); }
    template <typename T> const T &at(int y, int x) const { return *(const T *)(data + (size_t)y * step + (size_t)x * sizeof(T)); }
    /* single index: element i of a row or column vector */
    template <typename T> T &at(int i) { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    template <typename T> const T &at(int i) const { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    uchar *ptr(int y = 0) { return data + (size_t)y * step; }
    const uchar *ptr(int y = 0) const { return data + (size_t)y * step; }
    template <typename T> T *ptr(int y = 0) { return (T *)(data + (size_t)y * step); }
    template <typename T> const T *ptr(int y = 0) const { return (const T *)(data + (size_t)y * step); }

    static MatZeros zeros(int r, int c, int type) { return MatZeros{r, c, type}; }
    /* cv::MatExpr semantics: create() is a no-op on a matching destination, then fill */
    Mat &operator=(const MatZeros &z)
    {
        create(z.rows, z.cols, z.type);
        for (int y = 0; y < rows; ++y) std::memset(data + (size_t)y * step, 0, (size_t)cols * elemSize());
        return *this;
    }

    /* 8-bit -> float conversion (the SAD windows of Frame::ComputeStereoMatches); may alias */
    void convertTo(Mat &dst, int type) const
    {
        assert(type == CV_32F);
        Mat m(rows, cols, CV_32F);
        for (int y = 0; y < rows; ++y)
            for (int x = 0; x < cols; ++x)
                m.at<float>(y, x) = (type_ & 7) == CV_32F ? at<float>(y, x) : (float)at<uchar>(y, x);
        dst = m;
    }
    static Mat ones(int r, int c, int type)
    {
        Mat m(r, c, type);
        for (int y = 0; y < r; ++y)
            for (int x = 0; x < c; ++x) m.at<float>(y, x) = 1.f;
        return m;
    }
    double dot(const Mat &o) const
    {
        double s = 0;
        for (int y = 0; y < rows; ++y)
            for (int x = 0; x < cols; ++x) s += (double)at<float>(y, x) * o.at<float>(y, x);
        return s;
    }

    /* float (CV_32F, one channel) algebra used by the reference's Frame.cc pose code.  t() is an EXPRESSION, as
     * cv::MatExpr is: -A.t() * B reaches cv::gemm with GEMM_1_T and alpha = -1, a different arithmetic path from A * B
     * (MatT below). */
    inline class MatT t() const;

private:
    std::shared_ptr<uchar> store_;
    int type_;
};

/* OpenCV's own gemm / arithmetic live in a library built without FMA contraction (cv2 4.13: the small-
 * matrix float product is a left-to-right sum of individually rounded products, tests/test_oracle_match.py);
 * a reference translation unit compiled with -march=x86-64-v3 must not contract them when it inlines
 * these stand-ins. */
#if defined(__GNUC__) && !defined(__clang__) && !defined(__CUDACC__)
#define CVCOMPAT_NOFMA __attribute__((optimize("fp-contract=off"), noinline))
#else
#define CVCOMPAT_NOFMA
#endif
CVCOMPAT_NOFMA inline Mat operator*(const Mat &a, const Mat &b)
{
    Mat m(a.rows, b.cols, CV_32F);
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < b.cols; ++x) {
            float s = 0.f;
            for (int k = 0; k < a.cols; ++k) s += a.at<float>(y, k) * b.at<float>(k, x);
            m.at<float>(y, x) = s;
        }
    return m;
}
CVCOMPAT_NOFMA inline Mat mat_lin_(const Mat &a, float sa, const Mat *b, float sb)
{
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) m.at<float>(y, x) = sa * a.at<float>(y, x) + (b ? sb * b->at<float>(y, x) : 0.f);
    return m;
}
inline Mat operator+(const Mat &a, const Mat &b) { return mat_lin_(a, 1.f, &b, 1.f); }
inline Mat operator-(const Mat &a, const Mat &b) { return mat_lin_(a, 1.f, &b, -1.f); }
inline Mat operator-(const Mat &a) { return mat_lin_(a, -1.f, nullptr, 0.f); }
inline Mat operator*(float s, const Mat &a) { return mat_lin_(a, s, nullptr, 0.f); }
inline Mat operator/(const Mat &a, float s) { return mat_lin_(a, 1.f / s, nullptr, 0.f); } /* only named by keyframe matchers nobody runs here */
/* A.t() [scaled]: what cv::MatExpr carries for a transpose.  Used as a matrix it is the plain transpose
 * (cv::transpose, src/Frame.cc:332); multiplied, it is cv::gemm(A, B, alpha, noArray(), 0, GEMM_1_T), which -- unlike
 * the plain small product above -- takes OpenCV's generic path: the sum over k in DOUBLE, left to right, times alpha,
 * one rounding to float (cv2 4.13: tests/test_oracle_match.py::test_transposed_gemm_form_is_cv2s).  This is
 * Ow = -Rcw.t() * tcw of src/Frame.cc:334 and src/ORBmatcher.cc:366, 1146, 1516, 1676. */
CVCOMPAT_NOFMA inline float mul_nofma_(float a, float b) { return a * b; }
class MatT {
public:
    MatT(const Mat &a, double alpha) : a_(a), alpha_(alpha) {}
    operator Mat() const
    {
        /* MatOp_T::assign (OpenCV modules/core/src/matop.cpp): cv::transpose, then convertTo(m, -1, alpha) when alpha != 1,
         * which for CV_32F is a float multiply by (float)alpha (src/ORBmatcher.cc:1280: (1.0 / s12) * R12.t()) */
        Mat m(a_.cols, a_.rows, CV_32F);
        const float fa = (float)alpha_;
        for (int y = 0; y < a_.rows; ++y)
            for (int x = 0; x < a_.cols; ++x) m.at<float>(x, y) = alpha_ != 1.0 ? mul_nofma_(a_.at<float>(y, x), fa) : a_.at<float>(y, x);
        return m;
    }
    Mat a_;
    double alpha_;
};
inline MatT Mat::t() const { return MatT(*this, 1.0); }
inline MatT operator-(const MatT &e) { return MatT(e.a_, -e.alpha_); }
inline MatT operator*(double s, const MatT &e) { return MatT(e.a_, s * e.alpha_); }
CVCOMPAT_NOFMA inline Mat operator*(const MatT &e, const Mat &b)
{
    const Mat &a = e.a_; /* result = alpha * a^T * b */
    Mat m(a.cols, b.cols, CV_32F);
    for (int y = 0; y < a.cols; ++y)
        for (int x = 0; x < b.cols; ++x) {
            double s = 0.0;
            for (int k = 0; k < a.rows; ++k) s += (double)a.at<float>(k, y) * (double)b.at<float>(k, x);
            m.at<float>(y, x) = (float)(s * e.alpha_);
        }
    return m;
}
enum { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4 };
/* L1 norm of a - b, accumulated in double as OpenCV does for CV_32F */
inline double norm(const Mat &a, const Mat &b, int type)
{
    assert(type == NORM_L1);
    (void)type;
    double s = 0;
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) s += std::fabs((double)a.at<float>(y, x) - (double)b.at<float>(y, x));
    return s;
}
inline double norm(const Mat &a)
{
    double s = 0;
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) s += (double)a.at<float>(y, x) * a.at<float>(y, x);
    return std::sqrt(s);
}

/* (cv::Mat_<float>(r, c) << a, b, c) */
template <typename T> class Mat_;
template <typename T> class MatCommaInit_ {
public:
    MatCommaInit_(Mat m) : m_(m), i_(0) {}
    MatCommaInit_ &operator,(T v)
    {
        m_.template at<T>(i_ / m_.cols, i_ % m_.cols) = v;
        ++i_;
        return *this;
    }
    operator Mat() const { return m_; }
    Mat m_;
    int i_;
};
template <typename T> class Mat_ : public Mat {
public:
    Mat_(int r, int c) : Mat(r, c, CV_32F) {}
    MatCommaInit_<T> operator<<(T v)
    {
        MatCommaInit_<T> ci(*this);
        ci, v;
        return ci;
    }
};

class _InputArray {
public:
    _InputArray() : m_(nullptr) {}
    _InputArray(const Mat &m) : m_(&m) {}
    bool empty() const { return !m_ || m_->empty(); }
    Mat getMat() const { return m_ ? *m_ : Mat(); }

protected:
    const Mat *m_;
};

class _OutputArray : public _InputArray {
public:
    _OutputArray() : o_(nullptr) {}
    _OutputArray(Mat &m) : _InputArray(m), o_(&m) {}
    void create(int r, int c, int type) const { o_->create(r, c, type); }
    void create(Size sz, int type) const { o_->create(sz.height, sz.width, type); }
    void release() const { if (o_) o_->release(); }
    Mat getMat() const { return o_ ? *o_ : Mat(); }
    Mat &getMatRef() const { return *o_; }

private:
    Mat *o_;
};

typedef const _InputArray &InputArray;
typedef const _OutputArray &OutputArray;

enum { INTER_NEAREST = 0, INTER_LINEAR = 1 };
enum { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3,
       BORDER_REFLECT_101 = 4, BORDER_REFLECT101 = 4, BORDER_DEFAULT = 4, BORDER_ISOLATED = 16 };

void resize(InputArray src, OutputArray dst, Size dsize, double fx = 0, double fy = 0,
            int interpolation = INTER_LINEAR);
void copyMakeBorder(InputArray src, OutputArray dst, int top, int bottom, int left, int right,
                    int borderType);
void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sigmaX, double sigmaY = 0,
                  int borderType = BORDER_DEFAULT);
void FAST(InputArray image, std::vector<KeyPoint> &keypoints, int threshold,
          bool nonmaxSuppression = true);
float fastAtan2(float y, float x);
/* float points [N x 1, 2 channels] (or N x 2), K and P 3x3 CV_32F, dist 4/5 x 1 CV_32F, R empty */
void undistortPoints(InputArray src, OutputArray dst, InputArray K, InputArray dist, InputArray R, InputArray P);

struct KeyPointsFilter {
    static void retainBest(std::vector<KeyPoint> &keypoints, int npoints);
};

/* cv::FileStorage / cv::FileNode: named by the YAML save()/load() members of DBoW2's TemplatedVocabulary
 * (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1452-1632), which are virtual and therefore instantiated with the
 * class.  Nothing here reads or writes YAML (the vocabulary comes from loadFromTextFile :1338): every member aborts. */
class FileNode {
public:
    FileNode operator[](const std::string &) const { std::abort(); }
    FileNode operator[](const char *) const { std::abort(); }
    FileNode operator[](int) const { std::abort(); }
    size_t size() const { std::abort(); }
    operator int() const { std::abort(); }
    operator double() const { std::abort(); }
    operator std::string() const { std::abort(); }
};
class FileStorage {
public:
    enum { READ = 0, WRITE = 1 };
    FileStorage(const char *, int) {}
    bool isOpened() const { return false; }
    FileNode operator[](const std::string &) const { std::abort(); }
};
template <typename T> inline FileStorage &operator<<(FileStorage &, const T &) { std::abort(); }

} // namespace cv

#endif
