/* oracle/matchcompat -- TEST INFRASTRUCTURE.  Shadow of include/KeyFrame.h: the members
 * src/ORBmatcher.cc names in its keyframe matchers.  A keyframe here is a plain record of what
 * SearchByBoW(KeyFrame*, Frame&, ...) (:195-351) and the relocalisation SearchByProjection(Frame&, KeyFrame*, ...)
 * (:1670-1798) read -- map-point matches, feature vector, descriptors, undistorted keypoints -- which
 * the drivers (oracle/bow_driver.cpp) fill; the methods of the matchers nobody runs here abort. */
#ifndef KEYFRAME_H
#define KEYFRAME_H
#include <cstdlib>
#include <set>
#include <vector>
#include <opencv2/core/core.hpp>
#include "Thirdparty/DBoW2/DBoW2/BowVector.h"
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"
namespace ORB_SLAM2 {
class MapPoint;
class KeyFrame {
public:
    std::vector<MapPoint *> mvpMapPoints;
    std::vector<MapPoint *> GetMapPointMatches() { return mvpMapPoints; }
    std::set<MapPoint *> GetMapPoints() { std::abort(); }
    MapPoint *GetMapPoint(const size_t &idx) { return mvpMapPoints[idx]; }
    /* pose accessors as src/KeyFrame.cc has them (clones of Tcw's blocks and of Ow, which KeyFrame::SetPose computes as
     * -Rcw.t() * tcw): the drivers fill Tcw_ and Ow_ */
    cv::Mat Tcw_, Ow_;
    cv::Mat GetRotation() { return Tcw_.rowRange(0, 3).colRange(0, 3).clone(); }
    cv::Mat GetTranslation() { return Tcw_.rowRange(0, 3).col(3).clone(); }
    cv::Mat GetCameraCenter() { return Ow_.clone(); }
    cv::Mat GetPose() { return Tcw_.clone(); }
    std::vector<size_t> GetFeaturesInArea(const float &, const float &, const float &) const { std::abort(); }
    bool IsInImage(const float &, const float &) const { std::abort(); }
    void AddMapPoint(MapPoint *, const size_t &) { std::abort(); }
    DBoW2::FeatureVector mFeatVec;
    std::vector<cv::KeyPoint> mvKeysUn, mvKeys;
    std::vector<float> mvuRight, mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
    cv::Mat mDescriptors;
    float fx, fy, cx, cy, invfx, invfy, mbf, mb, mThDepth, mfLogScaleFactor;
    int N, mnScaleLevels;
    long unsigned int mnId;
    int mnMinX, mnMinY, mnMaxX, mnMaxY;
};
} // namespace ORB_SLAM2
#endif
