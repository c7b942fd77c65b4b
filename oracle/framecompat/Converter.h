/* oracle/framecompat -- TEST INFRASTRUCTURE.  Shadow of include/Converter.h (g2o/Eigen
 * conversions): src/Frame.cc uses only toDescriptorVector, in Frame::ComputeBoW (:536). */
#ifndef CONVERTER_H
#define CONVERTER_H
#include <vector>
#include <opencv2/core/core.hpp>
namespace ORB_SLAM2 {
class Converter {
public:
    static std::vector<cv::Mat> toDescriptorVector(const cv::Mat &d)
    {
        std::vector<cv::Mat> v;
        for (int i = 0; i < d.rows; ++i) v.push_back(d.row(i));
        return v;
    }
};
} // namespace ORB_SLAM2
#endif
