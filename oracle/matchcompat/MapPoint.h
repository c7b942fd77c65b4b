/* oracle/matchcompat -- TEST INFRASTRUCTURE.  Shadow of the reference's include/MapPoint.h for
 * compiling src/ORBmatcher.cc (and src/Frame.cc beside it) where they lie, without the map, g2o
 * and Eigen.  A map point here is a plain record holding what the matchers read through the
 * reference's accessors (GetWorldPos, GetDescriptor, Observations, isBad); the driver
 * (oracle/matcher_driver.cpp) fills it.  Members the exercised functions never reach abort. */
#ifndef MAPPOINT_H
#define MAPPOINT_H
#include <cstdlib>
#include <map>
#include <opencv2/core/core.hpp>
namespace ORB_SLAM2 {
class Frame;
class KeyFrame;
class MapPoint {
public:
    cv::Mat mWorldPos, mDescriptor; /* 3x1 CV_32F, 1x32 CV_8U */
    int nObs = 0;
    bool mbBad = false;

    cv::Mat GetWorldPos() { return mWorldPos.clone(); }
    cv::Mat GetDescriptor() { return mDescriptor.clone(); }
    int Observations() { return nObs; }
    bool isBad() { return mbBad; }

    cv::Mat mNormalVector; /* 3x1 CV_32F */
    cv::Mat GetNormal() { return mNormalVector.clone(); }
    /* src/MapPoint.cc:520-530 and :575-590, restated (the reference's MapPoint.cc needs the whole map model); the body of
     * PredictScale is in MapPoint_impl.h, which the drivers include where Frame is complete */
    float mfMinDistance = 0, mfMaxDistance = 0;
    float GetMaxDistanceInvariance() { return 1.2f * mfMaxDistance; }
    float GetMinDistanceInvariance() { return 0.8f * mfMinDistance; }
    int PredictScale(const float &currentDist, Frame *pF);
    int PredictScale(const float &currentDist, KeyFrame *pKF); /* src/MapPoint.cc:557-572, body in MapPoint_impl.h */
    /* the observation graph reduced to the one keyframe a matcher call works on (src/MapPoint.cc: IsInKeyFrame,
     * AddObservation, Replace, GetIndexInKeyFrame); what the keyframe matchers do with the map AFTER choosing a keypoint is
     * recorded, not modelled: the drivers read the choice out of KeyFrame::GetMapPoint's log */
    bool mbInKF = false;
    int mnIdxInKF = -1;
    bool IsInKeyFrame(KeyFrame *) { return mbInKF; }
    void Replace(MapPoint *) { mbBad = true; }
    void AddObservation(KeyFrame *, size_t idx) { mbInKF = true; mnIdxInKF = (int)idx; ++nObs; }
    int GetIndexInKeyFrame(KeyFrame *) { return mbInKF ? mnIdxInKF : -1; }

    float mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0, mTrackViewCos = 0;
    bool mbTrackInView = false;
    int mnTrackScaleLevel = 0;
    long unsigned int mnId = 0, mnLastFrameSeen = 0, mnFuseCandidateForKF = 0;
};
} // namespace ORB_SLAM2
#endif
