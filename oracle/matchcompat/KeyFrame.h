/* oracle/matchcompat -- TEST INFRASTRUCTURE.  Shadow of include/KeyFrame.h: the members
 * src/ORBmatcher.cc names in its keyframe matchers.  A keyframe here is a plain record of what
 * SearchByBoW(KeyFrame*, Frame&, ...) (:195-351) and the relocalisation SearchByProjection(Frame&, KeyFrame*, ...)
 * (:1670-1798) read -- map-point matches, feature vector, descriptors, undistorted keypoints -- which
 * the drivers (oracle/bow_driver.cpp) fill; the methods of the matchers nobody runs here abort. */
#ifndef KEYFRAME_H
#define KEYFRAME_H
#include <algorithm>
#include <cmath>
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
    std::set<MapPoint *> GetMapPoints() /* src/KeyFrame.cc: the non-NULL entries of mvpMapPoints that are not bad */
    {
        std::set<MapPoint *> s;
        for (size_t i = 0; i < mvpMapPoints.size(); ++i)
            if (mvpMapPoints[i] && !mp_is_bad_(mvpMapPoints[i])) s.insert(mvpMapPoints[i]);
        return s;
    }
    static bool mp_is_bad_(MapPoint *p); /* MapPoint is incomplete here; defined in MapPoint_impl.h */
    mutable long mnLastGetMapPoint = -1; /* the drivers read a matcher's choice out of this */
    MapPoint *GetMapPoint(const size_t &idx) { mnLastGetMapPoint = (long)idx; return mvpMapPoints[idx]; }
    /* pose accessors as src/KeyFrame.cc has them (clones of Tcw's blocks and of Ow, which KeyFrame::SetPose computes as
     * -Rcw.t() * tcw): the drivers fill Tcw_ and Ow_ */
    cv::Mat Tcw_, Ow_;
    cv::Mat GetRotation() { return Tcw_.rowRange(0, 3).colRange(0, 3).clone(); }
    cv::Mat GetTranslation() { return Tcw_.rowRange(0, 3).col(3).clone(); }
    cv::Mat GetCameraCenter() { return Ow_.clone(); }
    cv::Mat GetPose() { return Tcw_.clone(); }
    /* src/KeyFrame.cc:750-789, restated on the members below (src/KeyFrame.cc needs the map, the keyframe database and
     * g2o): note the INT bounds -- KeyFrame truncates the Frame's float mnMinX .. mnMaxY (include/KeyFrame.h) -- beside the
     * Frame's float grid factors */
    std::vector<size_t> GetFeaturesInArea(const float &x, const float &y, const float &r) const
    {
        std::vector<size_t> vIndices;
        vIndices.reserve(N);
        const int nMinCellX = std::max(0, (int)std::floor((x - mnMinX - r) * mfGridElementWidthInv));
        if (nMinCellX >= mnGridCols) return vIndices;
        const int nMaxCellX = std::min((int)mnGridCols - 1, (int)std::ceil((x - mnMinX + r) * mfGridElementWidthInv));
        if (nMaxCellX < 0) return vIndices;
        const int nMinCellY = std::max(0, (int)std::floor((y - mnMinY - r) * mfGridElementHeightInv));
        if (nMinCellY >= mnGridRows) return vIndices;
        const int nMaxCellY = std::min((int)mnGridRows - 1, (int)std::ceil((y - mnMinY + r) * mfGridElementHeightInv));
        if (nMaxCellY < 0) return vIndices;
        for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
            for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
                const std::vector<size_t> &vCell = mGrid[ix][iy];
                for (size_t j = 0, jend = vCell.size(); j < jend; j++) {
                    const cv::KeyPoint &kpUn = mvKeysUn[vCell[j]];
                    const float distx = kpUn.pt.x - x;
                    const float disty = kpUn.pt.y - y;
                    if (std::fabs(distx) < r && std::fabs(disty) < r) vIndices.push_back(vCell[j]);
                }
            }
        return vIndices;
    }
    bool IsInImage(const float &x, const float &y) const { return (x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY); } /* :792-795 */
    void AddMapPoint(MapPoint *pMP, const size_t &idx) { mvpMapPoints[idx] = pMP; }
    int mnGridCols = 64, mnGridRows = 48;
    float mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;
    std::vector<std::vector<std::vector<size_t> > > mGrid;
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
