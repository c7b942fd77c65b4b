/* oracle/framecompat -- TEST INFRASTRUCTURE.  Shadow of include/ORBVocabulary.h (a DBoW2 template
 * instance; the vocabulary blob is missing from the reference tree): src/Frame.cc only calls
 * transform() from Frame::ComputeBoW (:532-539), which the oracle driver never runs. */
#ifndef ORBVOCABULARY_H
#define ORBVOCABULARY_H
#include <vector>
#include <opencv2/core/core.hpp>
#include "Thirdparty/DBoW2/DBoW2/BowVector.h"
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"
/* the real header pulls in Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h, whose file-scope
 * `using namespace std;` (:36) include/Frame.h and src/Frame.cc silently rely on */
using namespace std;
namespace ORB_SLAM2 {
class ORBVocabulary {
public:
    void transform(const std::vector<cv::Mat> &, DBoW2::BowVector &, DBoW2::FeatureVector &, int) {}
};
} // namespace ORB_SLAM2
#endif
