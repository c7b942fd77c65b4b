/* oracle/framecompat -- TEST INFRASTRUCTURE.  Shadow of the reference's include/MapPoint.h with
 * only what src/Frame.cc touches (Frame::isInFrustum, src/Frame.cc:348-407), so that Frame.cc can
 * be compiled where it lies without the map, g2o and Eigen.  Never called by the oracle driver. */
#ifndef MAPPOINT_H
#define MAPPOINT_H
#include <opencv2/core/core.hpp>
namespace ORB_SLAM2 {
class Frame;
class MapPoint {
public:
    cv::Mat GetWorldPos() { return cv::Mat(); }
    cv::Mat GetNormal() { return cv::Mat(); }
    float GetMaxDistanceInvariance() { return 0.f; }
    float GetMinDistanceInvariance() { return 0.f; }
    int PredictScale(const float &, Frame *) { return 0; }
    bool isBad() { return true; }
    float mTrackProjX, mTrackProjY, mTrackProjXR, mTrackViewCos;
    bool mbTrackInView;
    int mnTrackScaleLevel;
};
} // namespace ORB_SLAM2
#endif
