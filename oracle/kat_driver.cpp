/*
 * oracle/kat_driver.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Known-answer access to the reference's file-static descriptor routine: the reference's OWN, UNMODIFIED
 * src/ORBextractor.cc is included here as text, so that computeOrbDescriptor (:117-157) and bit_pattern_31_
 * (:160-418), both `static`, can be called on a chosen image and angle.  Compiled with the reference's flags
 * (REFFLAGS in oracle/Makefile: -O3 -march=x86-64-v3, contraction as GCC decides), like _ref/ORBextractor.o, so the
 * float form of `pattern[idx].x*b + pattern[idx].y*a` is the one the reference object has.
 */
#include <cstdint>
#include <cstring>

#include "../src/ORBextractor.cc"

extern "C" void refkat_describe(const uint8_t *img, int cols, int rows, size_t step, int x, int y, float angle_deg,
                                uint8_t *desc32)
{
    cv::Mat m(rows, cols, CV_8UC1);
    for (int r = 0; r < rows; ++r) std::memcpy(m.ptr(r), img + (size_t)r * step, (size_t)cols);
    cv::KeyPoint kp;
    kp.pt.x = (float)x;
    kp.pt.y = (float)y;
    kp.angle = angle_deg;
    ORB_SLAM2::computeOrbDescriptor(kp, m, (const cv::Point *)ORB_SLAM2::bit_pattern_31_, desc32);
}
