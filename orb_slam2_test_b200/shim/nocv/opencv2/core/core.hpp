/*
 * nocv/opencv2/core/core.hpp -- build-only stand-in for <opencv2/core/core.hpp>.
 *
 * This image (and the GPU box) has no OpenCV C++ headers or libraries.  The host shim
 * (../../ORBextractor.{h,cc}) touches only four OpenCV types -- cv::Mat (8-bit 1/3/4 channels, float maps),
 * cv::KeyPoint, cv::InputArray, cv::OutputArray -- and a caller written like the reference's
 * Frame.cc additionally uses Mat::rowRange/colRange/at/ptr.  This header declares exactly that
 * slice with OpenCV's public semantics so the shim and its link test can be compiled here.  With a
 * real OpenCV, pass its include flags instead (shim/Makefile: OPENCV_CFLAGS) and this directory is
 * unused.  No image processing lives here: all pixel work is in the CUDA library.
 */
#ifndef BORB_NOCV_CORE_HPP
#define BORB_NOCV_CORE_HPP

#include <cstddef>
#include <cstring>
#include <memory>
#include <vector>

#define CV_8U 0
#define CV_8UC1 0
#define CV_8UC3 16
#define CV_8UC4 24
#define CV_32F 5
#define CV_32FC1 5
#define BORB_NOCV 1

typedef unsigned char uchar;

namespace cv
{

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
};
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;

struct KeyPoint { /* 28 bytes, same field order as OpenCV's */
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float size_, float angle_ = -1, float response_ = 0, int octave_ = 0, int class_id_ = -1)
        : pt(x, y), size(size_), angle(angle_), response(response_), octave(octave_), class_id(class_id_) {}
};

class Mat
{
public:
    int rows, cols;
    uchar *data;
    size_t step;

    Mat() : rows(0), cols(0), data(NULL), step(0), type_(CV_8UC1) {}
    Mat(int r, int c, int type) : rows(0), cols(0), data(NULL), step(0), type_(type) { create(r, c, type); }
    Mat(int r, int c, int type, void *ext, size_t ext_step = 0)
        : rows(r), cols(c), data((uchar *)ext), step(0), type_(type) { step = ext_step ? ext_step : (size_t)c * elemSize(); }

    void create(int r, int c, int type)
    {
        if (data && own_ && r == rows && c == cols && type == type_) return;
        type_ = type;
        rows = r; cols = c; step = (size_t)c * elemSize();
        own_.reset(new uchar[(size_t)r * step + 1], std::default_delete<uchar[]>());
        data = own_.get();
    }
    void release() { own_.reset(); data = NULL; rows = cols = 0; step = 0; }
    bool empty() const { return !data || rows == 0 || cols == 0; }
    int type() const { return type_; }
    int channels() const { return (type_ >> 3) + 1; }
    size_t elemSize() const { return (size_t)channels() * ((type_ & 7) == CV_32F ? 4 : 1); }
    bool isContinuous() const { return step == (size_t)cols * elemSize() || rows <= 1; }

    Mat rowRange(int a, int b) const { Mat m(*this); m.data = data + (size_t)a * step; m.rows = b - a; return m; }
    Mat colRange(int a, int b) const { Mat m(*this); m.data = data + (size_t)a * elemSize(); m.cols = b - a; return m; }
    Mat clone() const
    {
        Mat m(rows, cols, type_);
        for (int y = 0; y < rows; ++y) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, (size_t)cols * elemSize());
        return m;
    }
    uchar *ptr(int y = 0) { return data + (size_t)y * step; }
    const uchar *ptr(int y = 0) const { return data + (size_t)y * step; }
    template <typename T> T &at(int y, int x) { return *(T *)(data + (size_t)y * step + (size_t)x * sizeof(T)); }
    template <typename T> const T &at(int y, int x) const { return *(const T *)(data + (size_t)y * step + (size_t)x * sizeof(T)); }

private:
    std::shared_ptr<uchar> own_;
    int type_;
};

class _InputArray
{
public:
    _InputArray() : in_(NULL) {}
    _InputArray(const Mat &m) : in_(&m) {}
    bool empty() const { return !in_ || in_->empty(); }
    Mat getMat() const { return in_ ? *in_ : Mat(); }
protected:
    const Mat *in_;
};

class _OutputArray : public _InputArray
{
public:
    _OutputArray() : out_(NULL) {}
    _OutputArray(Mat &m) : _InputArray(m), out_(&m) {}
    void create(int r, int c, int type) const { out_->create(r, c, type); }
    void release() const { if (out_) out_->release(); }
    Mat getMat() const { return out_ ? *out_ : Mat(); }
private:
    Mat *out_;
};

typedef const _InputArray &InputArray;
typedef const _OutputArray &OutputArray;

} // namespace cv

#endif
