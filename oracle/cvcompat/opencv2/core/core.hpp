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
 *
 * The same header doubles as the "no OpenCV on this box" build of the product's C++ host shim
 * (orb_slam2_test_b200/shim), which needs nothing beyond Mat/KeyPoint/InputArray/OutputArray.
 */
#ifndef ORACLE_CVCOMPAT_CORE_HPP
#define ORACLE_CVCOMPAT_CORE_HPP

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <iterator>
#include <memory>
#include <vector>

#if defined(__SSE2__)
#include <emmintrin.h>
#endif

#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0
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

    Mat() : rows(0), cols(0), data(nullptr), step(0) {}
    Mat(Size sz, int type) : rows(0), cols(0), data(nullptr), step(0) { create(sz.height, sz.width, type); }
    Mat(int r, int c, int type) : rows(0), cols(0), data(nullptr), step(0) { create(r, c, type); }
    /* header over caller-owned memory (no ownership) */
    Mat(int r, int c, int /*type*/, void *ext, size_t ext_step = 0)
        : rows(r), cols(c), data((uchar *)ext), step(ext_step ? ext_step : (size_t)c) {}

    void create(int r, int c, int /*type*/)
    {
        if (data && r == rows && c == cols) return;
        rows = r;
        cols = c;
        step = (size_t)c;
        const size_t n = (size_t)r * (size_t)c;
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

    int type() const { return CV_8UC1; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    size_t step1() const { return step; }
    bool isContinuous() const { return step == (size_t)cols || rows <= 1; }
    size_t total() const { return (size_t)rows * (size_t)cols; }

    Mat operator()(const Rect &r) const
    {
        Mat m(*this);
        m.data = data + (size_t)r.y * step + r.x;
        m.rows = r.height;
        m.cols = r.width;
        return m;
    }
    Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
    Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }

    Mat clone() const
    {
        Mat m(rows, cols, CV_8UC1);
        for (int y = 0; y < rows; ++y) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, (size_t)cols);
        return m;
    }
    void copyTo(Mat &dst) const
    {
        dst.create(rows, cols, CV_8UC1);
        for (int y = 0; y < rows; ++y) std::memcpy(dst.data + (size_t)y * dst.step, data + (size_t)y * step, (size_t)cols);
    }

    template <typename T> T &at(int y, int x) { return *(T *)(data + (size_t)y * step + (size_t)x * sizeof(T)); }
    template <typename T> const T &at(int y, int x) const { return *(const T *)(data + (size_t)y * step + (size_t)x * sizeof(T)); }
    uchar *ptr(int y = 0) { return data + (size_t)y * step; }
    const uchar *ptr(int y = 0) const { return data + (size_t)y * step; }
    template <typename T> T *ptr(int y = 0) { return (T *)(data + (size_t)y * step); }
    template <typename T> const T *ptr(int y = 0) const { return (const T *)(data + (size_t)y * step); }

    static MatZeros zeros(int r, int c, int type) { return MatZeros{r, c, type}; }
    /* cv::MatExpr semantics: create() is a no-op on a matching destination, then fill */
    Mat &operator=(const MatZeros &z)
    {
        create(z.rows, z.cols, z.type);
        for (int y = 0; y < rows; ++y) std::memset(data + (size_t)y * step, 0, (size_t)cols);
        return *this;
    }

private:
    std::shared_ptr<uchar> store_;
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

struct KeyPointsFilter {
    static void retainBest(std::vector<KeyPoint> &keypoints, int npoints);
};

} // namespace cv

#endif
