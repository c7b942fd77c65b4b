/* oracle/framecompat -- TEST INFRASTRUCTURE.  Shadow of include/ORBmatcher.h with the three
 * members src/Frame.cc uses in ComputeStereoMatches (:624, :689).  The reference's
 * src/ORBmatcher.cc cannot be compiled here (it needs the whole map model), so these are
 * RESTATED, not compiled: TH_LOW = 50, TH_HIGH = 100 (src/ORBmatcher.cc:34-35) and
 * DescriptorDistance = the 256-bit Hamming distance (src/ORBmatcher.cc:1846-1862). */
#ifndef ORBMATCHER_H
#define ORBMATCHER_H
#include <cstdint>
#include <cstring>
#include <opencv2/core/core.hpp>
namespace ORB_SLAM2 {
class ORBmatcher {
public:
    static const int TH_LOW = 50;
    static const int TH_HIGH = 100;
    static int DescriptorDistance(const cv::Mat &a, const cv::Mat &b)
    {
        int dist = 0;
        for (int i = 0; i < 8; ++i) {
            uint32_t x, y;
            std::memcpy(&x, a.ptr() + 4 * i, 4);
            std::memcpy(&y, b.ptr() + 4 * i, 4);
            dist += __builtin_popcount(x ^ y);
        }
        return dist;
    }
};
} // namespace ORB_SLAM2
#endif
