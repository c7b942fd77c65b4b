/* oracle/matchcompat -- TEST INFRASTRUCTURE.  MapPoint::PredictScale(const float &currentDist, Frame *pF) of
 * /root/reference/src/MapPoint.cc:575-590, restated for the record type of matchcompat/MapPoint.h (the reference's
 * MapPoint.cc needs KeyFrame, Map and the observation graph).  The reference file sees `using namespace std` (through
 * ORBmatcher.h -> ... -> Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:36) and <cmath>, so log(float) and ceil(float)
 * resolve to the FLOAT overloads: logf of the host libm, a float division by mfLogScaleFactor, ceilf.  The same
 * context is set up here.  Included by the drivers, once per library, after Frame.h. */
#ifndef MATCHCOMPAT_MAPPOINT_IMPL_H
#define MATCHCOMPAT_MAPPOINT_IMPL_H
#include <cmath>
#include "Frame.h"
namespace ORB_SLAM2 {
using namespace std;
int MapPoint::PredictScale(const float &currentDist, Frame *pF)
{
    float ratio;
    ratio = mfMaxDistance / currentDist;
    int nScale = ceil(log(ratio) / pF->mfLogScaleFactor);
    if (nScale < 0)
        nScale = 0;
    else if (nScale >= pF->mnScaleLevels)
        nScale = pF->mnScaleLevels - 1;
    return nScale;
}
int MapPoint::PredictScale(const float &currentDist, KeyFrame *pKF) /* src/MapPoint.cc:557-572 */
{
    float ratio;
    ratio = mfMaxDistance / currentDist;
    int nScale = ceil(log(ratio) / pKF->mfLogScaleFactor);
    if (nScale < 0)
        nScale = 0;
    else if (nScale >= pKF->mnScaleLevels)
        nScale = pKF->mnScaleLevels - 1;
    return nScale;
}
bool KeyFrame::mp_is_bad_(MapPoint *p) { return p->isBad(); }
} // namespace ORB_SLAM2
#endif
