/* oracle/framecompat -- TEST INFRASTRUCTURE.  Shadow of include/KeyFrame.h: src/Frame.cc only
 * holds a KeyFrame* (mpReferenceKF). */
#ifndef KEYFRAME_H
#define KEYFRAME_H
namespace ORB_SLAM2 {
class KeyFrame {};
} // namespace ORB_SLAM2
#endif
