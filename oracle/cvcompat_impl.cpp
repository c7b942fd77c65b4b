/*
 * oracle/cvcompat_impl.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Out-of-line half of the cvcompat shim: the five cv:: image functions the reference calls,
 * forwarding to the pinned plain-C restatements in cvprim.c.
 *   cv::resize          <- /root/reference/src/ORBextractor.cc:1429
 *   cv::copyMakeBorder  <- :1432, :1438
 *   cv::GaussianBlur    <- :1377
 *   cv::FAST            <- :1069, :1074 (and the dead legacy path :1210, :1214)
 *   cv::fastAtan2       <- :110
 * and, for the reference's src/Frame.cc (oracle/_ref/libreframe.so),
 *   cv::undistortPoints <- /root/reference/src/Frame.cc:560, :594  -> oracle/frame_oracle.c
 */
#include "opencv2/core/core.hpp"

#include <cstdlib>

#include "cvprim.h"

namespace cv {

void resize(InputArray _src, OutputArray _dst, Size dsize, double, double, int interpolation)
{
    assert(interpolation == INTER_LINEAR);
    (void)interpolation;
    Mat src = _src.getMat();
    _dst.create(dsize, CV_8UC1); /* no-op when dst already is a dsize view: ROI stays bound */
    Mat dst = _dst.getMat();
    cvp_resize_linear_u8(src.data, src.step, src.cols, src.rows, dst.data, dst.step, dst.cols,
                         dst.rows);
}

void copyMakeBorder(InputArray _src, OutputArray _dst, int top, int bottom, int left, int right,
                    int borderType)
{
    /* Only the reference's use is supported: equal margins, reflect-101.  Without
     * BORDER_ISOLATED OpenCV would read real pixels around a sub-matrix source; every in-repo
     * caller passes a whole image at that site (src/ORBextractor.cc:1438), so the source is
     * treated as isolated (SURVEY.md hazard H12). */
    assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101);
    assert(top == bottom && left == right && top == left);
    (void)borderType; (void)bottom; (void)right;
    Mat src = _src.getMat();
    _dst.create(src.rows + 2 * top, src.cols + 2 * left, CV_8UC1);
    Mat dst = _dst.getMat();
    uchar *interior = dst.data + (size_t)top * dst.step + left;
    if (interior != src.data)
        for (int y = 0; y < src.rows; ++y)
            std::memcpy(interior + (size_t)y * dst.step, src.data + (size_t)y * src.step, (size_t)src.cols);
    cvp_border_reflect101(dst.data, dst.step, src.cols, src.rows, top);
}

void GaussianBlur(InputArray _src, OutputArray _dst, Size ksize, double sigmaX, double sigmaY,
                  int borderType)
{
    assert(ksize.width == 7 && ksize.height == 7 && sigmaX == 2 && sigmaY == 2);
    assert(borderType == BORDER_REFLECT_101);
    (void)ksize; (void)sigmaX; (void)sigmaY; (void)borderType;
    Mat src = _src.getMat();
    _dst.create(src.rows, src.cols, CV_8UC1);
    Mat dst = _dst.getMat();
    cvp_gaussian7_sigma2(src.data, src.step, dst.data, dst.step, src.cols, src.rows);
}

void FAST(InputArray _img, std::vector<KeyPoint> &keypoints, int threshold, bool nms)
{
    assert(nms);
    (void)nms;
    Mat img = _img.getMat();
    keypoints.clear();
    if (img.cols < 7 || img.rows < 7) return;
    /* strict 3x3 NMS leaves at most one survivor per 2x2 block */
    const int cap = ((img.cols + 1) / 2) * ((img.rows + 1) / 2);
    std::vector<cvp_kp> buf((size_t)cap);
    const int n = cvp_fast9_16(img.data, img.step, img.cols, img.rows, threshold, buf.data(), cap);
    keypoints.reserve((size_t)n);
    for (int i = 0; i < n; ++i)
        keypoints.push_back(KeyPoint((float)buf[i].x, (float)buf[i].y, 7.f, -1.f, (float)buf[i].score));
}

float fastAtan2(float y, float x) { return cvp_fast_atan2(y, x); }

extern "C" void fo_undistort_points(const float *xy, int n, const float *cam, float *out);

/* the one call shape the reference uses: N float points, P == K, R empty */
void undistortPoints(InputArray _src, OutputArray _dst, InputArray _K, InputArray _dist, InputArray, InputArray)
{
    const Mat src = _src.getMat(), K = _K.getMat(), D = _dist.getMat();
    const int n = src.rows;
    float cam[9] = {K.at<float>(0, 0), K.at<float>(1, 1), K.at<float>(0, 2), K.at<float>(1, 2), 0, 0, 0, 0, 0};
    const int nd = D.rows * D.cols;
    for (int i = 0; i < nd && i < 5; ++i) cam[4 + i] = D.at<float>(i);
    std::vector<float> in((size_t)2 * n), out((size_t)2 * n);
    for (int i = 0; i < n; ++i) {
        in[2 * i] = src.ptr<float>(i)[0];
        in[2 * i + 1] = src.ptr<float>(i)[1];
    }
    fo_undistort_points(in.data(), n, cam, out.data());
    Mat &dst = _dst.getMatRef(); /* in place in the reference: same geometry */
    for (int i = 0; i < n; ++i) {
        dst.ptr<float>(i)[0] = out[2 * i];
        dst.ptr<float>(i)[1] = out[2 * i + 1];
    }
}

void KeyPointsFilter::retainBest(std::vector<KeyPoint> &, int)
{
    /* only reachable from ComputeKeyPointsOld (src/ORBextractor.cc:1130-1307), whose single call
     * is commented out at :1344 */
    std::abort();
}

} // namespace cv
