/*
 * oracle/matcher_driver.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * C entry point over the reference's OWN src/ORBmatcher.cc and src/Frame.cc (compiled where they
 * lie, see oracle/Makefile: _ref/librefmatch.so) for the descriptor consumer built on the device:
 *   ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, th, bMono)
 *       src/ORBmatcher.cc:1503-1668, with Frame::GetFeaturesInArea src/Frame.cc:421-508,
 *       ORBmatcher::ComputeThreeMaxima :1800-1843 and DescriptorDistance :1846-1862.
 * The two reference Frame objects are default-constructed (src/Frame.cc:43) and their public members
 * filled from flat arrays -- keypoints, descriptors, grid and stereo coordinates are the outputs the
 * other rows already pin (tests/test_oracle_frame_vs_reference.py); map points are the records of
 * oracle/matchcompat/MapPoint.h.  The call itself, including every float expression in it, is the
 * reference's object code (-O3 -march=x86-64-v3, contraction as GCC decides).
 */
#include <cstdint>
#include <cstring>
#include <set>
#include <vector>

#include "Frame.h"
#include "ORBmatcher.h"
#include "matchcompat/MapPoint_impl.h"

using namespace ORB_SLAM2;

namespace {

cv::Mat pose_mat(const float *T12)
{
    cv::Mat T(4, 4, CV_32F);
    for (int c = 0; c < 4; ++c) T.at<float>(3, c) = c == 3 ? 1.f : 0.f;
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 4; ++c) T.at<float>(r, c) = T12[4 * r + c];
    return T;
}

} // namespace

extern "C" {

struct refmatch_frame { /* one side of the call, flat */
    int n;
    const uint8_t *kps;      /* n x 28 B cv::KeyPoint (mvKeys) */
    const float *un_xy;      /* n x 2 (mvKeysUn positions; octave / angle as in kps) */
    const uint8_t *desc;     /* n x 32 */
    const float *u_right;    /* n (mvuRight), may be null: all -1 */
    const int32_t *grid_count, *grid_idx; /* 64*48 cell-major counts + concatenated indices, may be null */
    const float *Tcw;        /* 3 x 4 row-major */
};

/* cam = fx, fy, cx, cy, mbf, mb; bounds = mnMinX, mnMaxX, mnMinY, mnMaxY; scale = mvScaleFactors.
 * Map points of the last frame: mp_flags[i] bit 0 = LastFrame.mvpMapPoints[i] != NULL, bit 1 =
 * Observations() > 0, bit 2 = LastFrame.mvbOutlier[i]; world position mp_xyz[3i..], descriptor
 * mp_desc[32i..].  out_match[i2] = index i of the last-frame map point CurrentFrame.mvpMapPoints[i2]
 * ends up with, or -1.  Returns nmatches. */
int refmatch_search_by_projection(const refmatch_frame *cur, const refmatch_frame *last, const float *cam,
                                  const float *bounds, const float *scale, int nlevels, const uint8_t *mp_flags,
                                  const float *mp_xyz, const uint8_t *mp_desc, float th, int mono, int check_ori,
                                  int32_t *out_match)
{
    Frame::fx = cam[0]; Frame::fy = cam[1]; Frame::cx = cam[2]; Frame::cy = cam[3];
    Frame::invfx = 1.0f / cam[0]; Frame::invfy = 1.0f / cam[1];
    Frame::mnMinX = bounds[0]; Frame::mnMaxX = bounds[1]; Frame::mnMinY = bounds[2]; Frame::mnMaxY = bounds[3];
    /* as the Frame constructors compute them, src/Frame.cc:135-136 */
    Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(Frame::mnMaxX - Frame::mnMinX);
    Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(Frame::mnMaxY - Frame::mnMinY);
    Frame::mbInitialComputations = false;

    Frame F[2];
    const refmatch_frame *src[2] = {cur, last};
    for (int s = 0; s < 2; ++s) {
        Frame &f = F[s];
        const refmatch_frame &r = *src[s];
        f.N = r.n;
        f.mbf = cam[4];
        f.mb = cam[5];
        f.mvKeys.resize(r.n);
        if (r.n) std::memcpy(f.mvKeys.data(), r.kps, (size_t)r.n * 28);
        f.mvKeysUn = f.mvKeys;
        for (int i = 0; i < r.n; ++i) { f.mvKeysUn[i].pt.x = r.un_xy[2 * i]; f.mvKeysUn[i].pt.y = r.un_xy[2 * i + 1]; }
        f.mDescriptors = cv::Mat(r.n, 32, CV_8UC1);
        for (int i = 0; i < r.n; ++i) std::memcpy(f.mDescriptors.ptr(i), r.desc + (size_t)i * 32, 32);
        f.mvuRight.assign(r.n, -1.f);
        if (r.u_right) for (int i = 0; i < r.n; ++i) f.mvuRight[i] = r.u_right[i];
        f.mvDepth.assign(r.n, -1.f);
        f.mvpMapPoints.assign(r.n, static_cast<MapPoint *>(NULL));
        f.mvbOutlier.assign(r.n, false);
        f.mnScaleLevels = nlevels;
        f.mvScaleFactors.assign(scale, scale + nlevels);
        if (r.grid_count) {
            int k = 0;
            for (int i = 0; i < FRAME_GRID_COLS; ++i)
                for (int j = 0; j < FRAME_GRID_ROWS; ++j) {
                    const int c = r.grid_count[i * FRAME_GRID_ROWS + j];
                    f.mGrid[i][j].assign(r.grid_idx + k, r.grid_idx + k + c);
                    k += c;
                }
        }
        f.SetPose(pose_mat(r.Tcw));
    }
    Frame &C = F[0], &L = F[1];
    std::vector<MapPoint> mps(last->n);
    for (int i = 0; i < last->n; ++i) {
        if (!(mp_flags[i] & 1)) continue;
        MapPoint &m = mps[i];
        m.mWorldPos = cv::Mat(3, 1, CV_32F);
        for (int k = 0; k < 3; ++k) m.mWorldPos.at<float>(k) = mp_xyz[3 * i + k];
        m.mDescriptor = cv::Mat(1, 32, CV_8UC1);
        std::memcpy(m.mDescriptor.ptr(0), mp_desc + (size_t)i * 32, 32);
        m.nObs = (mp_flags[i] & 2) ? 1 : 0;
        L.mvpMapPoints[i] = &m;
        L.mvbOutlier[i] = (mp_flags[i] & 4) != 0;
    }
    ORBmatcher matcher(0.9f, check_ori != 0);
    const int nm = matcher.SearchByProjection(C, L, th, mono != 0);
    for (int i2 = 0; i2 < cur->n; ++i2)
        out_match[i2] = C.mvpMapPoints[i2] ? (int32_t)(C.mvpMapPoints[i2] - mps.data()) : -1;
    return nm;
}

/* ORBmatcher::SearchByProjection(Frame &F, const vector<MapPoint*> &vpMapPoints, th) (src/ORBmatcher.cc:59-154).
 * proj[4i..] = mTrackProjX, mTrackProjY, mTrackProjXR, mTrackViewCos; level[i] = mnTrackScaleLevel; mp_flags bit 0 =
 * mbTrackInView && !isBad(), bit 1 = Observations() > 0.  occupied[i2]: 0 = F.mvpMapPoints[i2] is NULL on entry, 1 =
 * it holds a point without observations, 2 = with.  out_match[i2] = index of the map point this call put there, -1
 * where the slot still holds what it held. */
int refmatch_search_local_points(const refmatch_frame *cur, const float *bounds, const float *scale, int nlevels, int n_mp,
                                 const float *proj, const uint8_t *mp_desc, const uint8_t *mp_flags, const int32_t *level,
                                 const uint8_t *occupied, float th, float nnratio, int32_t *out_match)
{
    Frame::mnMinX = bounds[0]; Frame::mnMaxX = bounds[1]; Frame::mnMinY = bounds[2]; Frame::mnMaxY = bounds[3];
    Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(Frame::mnMaxX - Frame::mnMinX);
    Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(Frame::mnMaxY - Frame::mnMinY);
    Frame::mbInitialComputations = false;
    Frame f;
    const refmatch_frame &r = *cur;
    f.N = r.n;
    f.mvKeys.resize(r.n);
    if (r.n) std::memcpy(f.mvKeys.data(), r.kps, (size_t)r.n * 28);
    f.mvKeysUn = f.mvKeys;
    for (int i = 0; i < r.n; ++i) { f.mvKeysUn[i].pt.x = r.un_xy[2 * i]; f.mvKeysUn[i].pt.y = r.un_xy[2 * i + 1]; }
    f.mDescriptors = cv::Mat(r.n, 32, CV_8UC1);
    for (int i = 0; i < r.n; ++i) std::memcpy(f.mDescriptors.ptr(i), r.desc + (size_t)i * 32, 32);
    f.mvuRight.assign(r.n, -1.f);
    if (r.u_right) for (int i = 0; i < r.n; ++i) f.mvuRight[i] = r.u_right[i];
    f.mnScaleLevels = nlevels;
    f.mvScaleFactors.assign(scale, scale + nlevels);
    int k = 0;
    for (int i = 0; i < FRAME_GRID_COLS; ++i)
        for (int j = 0; j < FRAME_GRID_ROWS; ++j) {
            const int c = r.grid_count[i * FRAME_GRID_ROWS + j];
            f.mGrid[i][j].assign(r.grid_idx + k, r.grid_idx + k + c);
            k += c;
        }
    MapPoint held[2]; /* what occupied slots point at on entry */
    held[0].nObs = 0;
    held[1].nObs = 1;
    f.mvpMapPoints.assign(r.n, static_cast<MapPoint *>(NULL));
    for (int i = 0; i < r.n; ++i)
        if (occupied[i]) f.mvpMapPoints[i] = &held[occupied[i] == 2];
    std::vector<MapPoint> mps(n_mp);
    std::vector<MapPoint *> vp(n_mp);
    for (int i = 0; i < n_mp; ++i) {
        MapPoint &m = mps[i];
        m.mbTrackInView = (mp_flags[i] & 1) != 0;
        m.mbBad = false;
        m.nObs = (mp_flags[i] & 2) ? 1 : 0;
        m.mTrackProjX = proj[4 * i]; m.mTrackProjY = proj[4 * i + 1]; m.mTrackProjXR = proj[4 * i + 2];
        m.mTrackViewCos = proj[4 * i + 3];
        m.mnTrackScaleLevel = level[i];
        m.mDescriptor = cv::Mat(1, 32, CV_8UC1);
        std::memcpy(m.mDescriptor.ptr(0), mp_desc + (size_t)i * 32, 32);
        vp[i] = &m;
    }
    ORBmatcher matcher(nnratio, true);
    const int nm = matcher.SearchByProjection(f, vp, th);
    for (int i2 = 0; i2 < r.n; ++i2) {
        MapPoint *p = f.mvpMapPoints[i2];
        out_match[i2] = (p && p != &held[0] && p != &held[1]) ? (int32_t)(p - mps.data()) : -1;
    }
    return nm;
}

/* ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) (src/ORBmatcher.cc:487-632).
 * f1 / f2 as refmatch_frame (f1 needs no grid; Tcw unused); prev[2 n1] = vbPrevMatched in and out. */
int refmatch_search_for_initialization(const refmatch_frame *f1, const refmatch_frame *f2, const float *bounds, float *prev,
                                       int window_size, float nnratio, int check_ori, int32_t *matches12)
{
    Frame::mnMinX = bounds[0]; Frame::mnMaxX = bounds[1]; Frame::mnMinY = bounds[2]; Frame::mnMaxY = bounds[3];
    Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(Frame::mnMaxX - Frame::mnMinX);
    Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(Frame::mnMaxY - Frame::mnMinY);
    Frame::mbInitialComputations = false;
    Frame F[2];
    const refmatch_frame *src[2] = {f1, f2};
    for (int s = 0; s < 2; ++s) {
        Frame &f = F[s];
        const refmatch_frame &r = *src[s];
        f.N = r.n;
        f.mvKeys.resize(r.n);
        if (r.n) std::memcpy(f.mvKeys.data(), r.kps, (size_t)r.n * 28);
        f.mvKeysUn = f.mvKeys;
        for (int i = 0; i < r.n; ++i) { f.mvKeysUn[i].pt.x = r.un_xy[2 * i]; f.mvKeysUn[i].pt.y = r.un_xy[2 * i + 1]; }
        f.mDescriptors = cv::Mat(r.n, 32, CV_8UC1);
        for (int i = 0; i < r.n; ++i) std::memcpy(f.mDescriptors.ptr(i), r.desc + (size_t)i * 32, 32);
        if (r.grid_count) {
            int k = 0;
            for (int i = 0; i < FRAME_GRID_COLS; ++i)
                for (int j = 0; j < FRAME_GRID_ROWS; ++j) {
                    const int c = r.grid_count[i * FRAME_GRID_ROWS + j];
                    f.mGrid[i][j].assign(r.grid_idx + k, r.grid_idx + k + c);
                    k += c;
                }
        }
    }
    std::vector<cv::Point2f> vbPrev(f1->n);
    for (int i = 0; i < f1->n; ++i) vbPrev[i] = cv::Point2f(prev[2 * i], prev[2 * i + 1]);
    std::vector<int> m12;
    ORBmatcher matcher(nnratio, check_ori != 0);
    const int nm = matcher.SearchForInitialization(F[0], F[1], vbPrev, m12, window_size);
    for (int i = 0; i < f1->n; ++i) { matches12[i] = m12[i]; prev[2 * i] = vbPrev[i].x; prev[2 * i + 1] = vbPrev[i].y; }
    return nm;
}

/* ORBmatcher::SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, const set<MapPoint*> &sAlreadyFound, th, ORBdist)
 * (src/ORBmatcher.cc:1670-1798), the relocalisation matcher.  kf_flags[i] bit 0 = vpMPs[i] != NULL, bit 1 = isBad(),
 * bit 2 = member of sAlreadyFound; kf_dist[2i], [2i + 1] = mfMinDistance, mfMaxDistance; kf_angle[i] =
 * pKF->mvKeysUn[i].angle; occupied[i2] != 0: CurrentFrame.mvpMapPoints[i2] is set on entry.  out_match[i2] = index of the
 * keyframe keypoint whose map point this call put into slot i2, -1 elsewhere. */
int refmatch_search_by_projection_kf(const refmatch_frame *cur, const float *cam, const float *bounds, const float *scale,
                                     int nlevels, float log_scale, int n_kf, const uint8_t *kf_flags, const float *kf_xyz,
                                     const uint8_t *kf_desc, const float *kf_dist, const float *kf_angle,
                                     const uint8_t *occupied, float th, int orb_dist, int check_ori, int32_t *out_match)
{
    Frame::fx = cam[0]; Frame::fy = cam[1]; Frame::cx = cam[2]; Frame::cy = cam[3];
    Frame::invfx = 1.0f / cam[0]; Frame::invfy = 1.0f / cam[1];
    Frame::mnMinX = bounds[0]; Frame::mnMaxX = bounds[1]; Frame::mnMinY = bounds[2]; Frame::mnMaxY = bounds[3];
    Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(Frame::mnMaxX - Frame::mnMinX);
    Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(Frame::mnMaxY - Frame::mnMinY);
    Frame::mbInitialComputations = false;
    Frame f;
    const refmatch_frame &r = *cur;
    f.N = r.n;
    f.mvKeys.resize(r.n);
    if (r.n) std::memcpy(f.mvKeys.data(), r.kps, (size_t)r.n * 28);
    f.mvKeysUn = f.mvKeys;
    for (int i = 0; i < r.n; ++i) { f.mvKeysUn[i].pt.x = r.un_xy[2 * i]; f.mvKeysUn[i].pt.y = r.un_xy[2 * i + 1]; }
    f.mDescriptors = cv::Mat(r.n, 32, CV_8UC1);
    for (int i = 0; i < r.n; ++i) std::memcpy(f.mDescriptors.ptr(i), r.desc + (size_t)i * 32, 32);
    f.mvuRight.assign(r.n, -1.f);
    f.mnScaleLevels = nlevels;
    f.mfLogScaleFactor = log_scale;
    f.mvScaleFactors.assign(scale, scale + nlevels);
    int k = 0;
    for (int i = 0; i < FRAME_GRID_COLS; ++i)
        for (int j = 0; j < FRAME_GRID_ROWS; ++j) {
            const int c = r.grid_count[i * FRAME_GRID_ROWS + j];
            f.mGrid[i][j].assign(r.grid_idx + k, r.grid_idx + k + c);
            k += c;
        }
    f.SetPose(pose_mat(r.Tcw));
    MapPoint held;
    f.mvpMapPoints.assign(r.n, static_cast<MapPoint *>(NULL));
    if (occupied)
        for (int i = 0; i < r.n; ++i)
            if (occupied[i]) f.mvpMapPoints[i] = &held;
    std::vector<MapPoint> mps(n_kf);
    KeyFrame kf;
    kf.mvpMapPoints.assign(n_kf, static_cast<MapPoint *>(NULL));
    kf.mvKeysUn.resize(n_kf);
    std::set<MapPoint *> found;
    for (int i = 0; i < n_kf; ++i) {
        kf.mvKeysUn[i].angle = kf_angle[i];
        if (!(kf_flags[i] & 1)) continue;
        MapPoint &m = mps[i];
        m.mWorldPos = cv::Mat(3, 1, CV_32F);
        for (int c = 0; c < 3; ++c) m.mWorldPos.at<float>(c) = kf_xyz[3 * i + c];
        m.mDescriptor = cv::Mat(1, 32, CV_8UC1);
        std::memcpy(m.mDescriptor.ptr(0), kf_desc + (size_t)i * 32, 32);
        m.mbBad = (kf_flags[i] & 2) != 0;
        m.mfMinDistance = kf_dist[2 * i];
        m.mfMaxDistance = kf_dist[2 * i + 1];
        kf.mvpMapPoints[i] = &m;
        if (kf_flags[i] & 4) found.insert(&m);
    }
    ORBmatcher matcher(0.9f, check_ori != 0);
    const int nm = matcher.SearchByProjection(f, &kf, found, th, orb_dist);
    for (int i2 = 0; i2 < r.n; ++i2) {
        MapPoint *p = f.mvpMapPoints[i2];
        out_match[i2] = (p && p != &held) ? (int32_t)(p - mps.data()) : -1;
    }
    return nm;
}

namespace {
/* a keyframe stand-in (oracle/matchcompat/KeyFrame.h) filled the way KeyFrame::KeyFrame(Frame &F, ...) copies a Frame
 * (src/KeyFrame.cc): keypoints, descriptors, grid, the Frame's float grid factors and its bounds TRUNCATED to int */
void fill_keyframe(KeyFrame &kf, const refmatch_frame &r, const float *cam, const float *bounds, const float *scale,
                   const float *inv_sigma2, int nlevels, float log_scale)
{
    kf.N = r.n;
    kf.mvKeys.resize(r.n);
    if (r.n) std::memcpy(kf.mvKeys.data(), r.kps, (size_t)r.n * 28);
    kf.mvKeysUn = kf.mvKeys;
    for (int i = 0; i < r.n; ++i) { kf.mvKeysUn[i].pt.x = r.un_xy[2 * i]; kf.mvKeysUn[i].pt.y = r.un_xy[2 * i + 1]; }
    kf.mDescriptors = cv::Mat(r.n, 32, CV_8UC1);
    for (int i = 0; i < r.n; ++i) std::memcpy(kf.mDescriptors.ptr(i), r.desc + (size_t)i * 32, 32);
    kf.mvuRight.assign(r.n, -1.f);
    if (r.u_right) kf.mvuRight.assign(r.u_right, r.u_right + r.n);
    kf.mvpMapPoints.assign(r.n, static_cast<MapPoint *>(NULL));
    kf.fx = cam[0]; kf.fy = cam[1]; kf.cx = cam[2]; kf.cy = cam[3]; kf.mbf = cam[4]; kf.mb = cam[5];
    kf.invfx = 1.0f / cam[0]; kf.invfy = 1.0f / cam[1];
    kf.mnMinX = bounds[0]; kf.mnMaxX = bounds[1]; kf.mnMinY = bounds[2]; kf.mnMaxY = bounds[3]; /* float -> int, as the const int members */
    kf.mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(bounds[1] - bounds[0]);  /* the Frame's, src/Frame.cc:135-136 */
    kf.mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(bounds[3] - bounds[2]);
    kf.mnScaleLevels = nlevels;
    kf.mfLogScaleFactor = log_scale;
    kf.mvScaleFactors.assign(scale, scale + nlevels);
    kf.mvInvLevelSigma2.assign(inv_sigma2, inv_sigma2 + nlevels);
    kf.mGrid.assign(FRAME_GRID_COLS, std::vector<std::vector<size_t> >(FRAME_GRID_ROWS));
    int k = 0;
    for (int i = 0; i < FRAME_GRID_COLS; ++i)
        for (int j = 0; j < FRAME_GRID_ROWS; ++j) {
            const int c = r.grid_count[i * FRAME_GRID_ROWS + j];
            kf.mGrid[i][j].assign(r.grid_idx + k, r.grid_idx + k + c);
            k += c;
        }
    kf.Tcw_ = pose_mat(r.Tcw);
    cv::Mat Rcw = kf.Tcw_.rowRange(0, 3).colRange(0, 3);
    cv::Mat tcw = kf.Tcw_.rowRange(0, 3).col(3);
    kf.Ow_ = -Rcw.t() * tcw; /* KeyFrame::SetPose */
}
} // namespace

/* ORBmatcher::Fuse(KeyFrame *pKF, const vector<MapPoint*> &vpMapPoints, const float th) (src/ORBmatcher.cc:968-1131), the
 * matcher of LocalMapping::SearchInNeighbors.  What it does to the map after choosing a keypoint (Replace / AddObservation /
 * AddMapPoint, :1103-1127) is the map model's business; what is pinned here is the CHOICE: the reference's function is run
 * on one map point at a time -- its loop carries no state from one point to the next besides the map it mutates, and the
 * choice reads none of that -- and the keypoint it hands to pKF->GetMapPoint (:1105) is recorded.  A second run over the
 * whole vector must fuse the same number.  mp_flags[i] bit 0 = vpMapPoints[i] != NULL, bit 1 = isBad(), bit 2 =
 * IsInKeyFrame(pKF); mp_dist[2i], [2i + 1] = mfMinDistance, mfMaxDistance; mp_normal[3i..] = GetNormal().
 * out_best[i] = the chosen keypoint of pKF or -1.  Returns nFused of the whole-vector run. */
int refmatch_fuse(const refmatch_frame *kfr, const float *cam, const float *bounds, const float *scale, const float *inv_sigma2,
                  int nlevels, float log_scale, int n_mp, const uint8_t *mp_flags, const float *mp_xyz, const uint8_t *mp_desc,
                  const float *mp_dist, const float *mp_normal, float th, int32_t *out_best)
{
    int total = 0;
    for (int pass = 0; pass < 2; ++pass) {
        KeyFrame kf;
        fill_keyframe(kf, *kfr, cam, bounds, scale, inv_sigma2, nlevels, log_scale);
        std::vector<MapPoint> mps(n_mp);
        std::vector<MapPoint *> vp(n_mp, static_cast<MapPoint *>(NULL));
        for (int i = 0; i < n_mp; ++i) {
            if (!(mp_flags[i] & 1)) continue;
            MapPoint &m = mps[i];
            m.mWorldPos = cv::Mat(3, 1, CV_32F);
            m.mNormalVector = cv::Mat(3, 1, CV_32F);
            for (int c = 0; c < 3; ++c) { m.mWorldPos.at<float>(c) = mp_xyz[3 * i + c]; m.mNormalVector.at<float>(c) = mp_normal[3 * i + c]; }
            m.mDescriptor = cv::Mat(1, 32, CV_8UC1);
            std::memcpy(m.mDescriptor.ptr(0), mp_desc + (size_t)i * 32, 32);
            m.mbBad = (mp_flags[i] & 2) != 0;
            m.mbInKF = (mp_flags[i] & 4) != 0;
            m.mfMinDistance = mp_dist[2 * i];
            m.mfMaxDistance = mp_dist[2 * i + 1];
            m.nObs = 1;
            vp[i] = &m;
        }
        ORBmatcher matcher(0.6f, true);
        if (pass == 0) {
            for (int i = 0; i < n_mp; ++i) {
                kf.mnLastGetMapPoint = -1;
                std::vector<MapPoint *> one(1, vp[i]);
                const int nf = matcher.Fuse(&kf, one, th);
                out_best[i] = nf ? (int32_t)kf.mnLastGetMapPoint : -1;
                total += nf;
            }
        } else if (matcher.Fuse(&kf, vp, th) != total)
            std::abort();
    }
    return total;
}

namespace {
cv::Mat mat44(const float *m16)
{
    cv::Mat T(4, 4, CV_32F);
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c) T.at<float>(r, c) = m16[4 * r + c];
    return T;
}
void fill_points(std::vector<MapPoint> &mps, int n, const uint8_t *flags, const float *xyz, const uint8_t *desc, const float *dist,
                 const float *normal)
{
    mps.assign((size_t)std::max(n, 1), MapPoint());
    for (int i = 0; i < n; ++i) {
        MapPoint &m = mps[i];
        m.mWorldPos = cv::Mat(3, 1, CV_32F);
        m.mNormalVector = cv::Mat(3, 1, CV_32F);
        for (int c = 0; c < 3; ++c) { m.mWorldPos.at<float>(c) = xyz[3 * i + c]; m.mNormalVector.at<float>(c) = normal ? normal[3 * i + c] : 0.f; }
        m.mDescriptor = cv::Mat(1, 32, CV_8UC1);
        std::memcpy(m.mDescriptor.ptr(0), desc + (size_t)i * 32, 32);
        m.mbBad = (flags[i] & 2) != 0;
        m.mfMinDistance = dist[2 * i];
        m.mfMaxDistance = dist[2 * i + 1];
        m.nObs = 1;
    }
}
} // namespace

/* ORBmatcher::Fuse(KeyFrame *pKF, cv::Mat Scw, const vector<MapPoint*> &vpPoints, float th, vector<MapPoint*>
 * &vpReplacePoint) (src/ORBmatcher.cc:1133-1260, LoopClosing::SearchAndFuse): the CHOICE per point, pinned as in
 * refmatch_fuse (one point per call of the reference's function, the keypoint handed to pKF->GetMapPoint recorded; a run
 * over the whole vector must fuse as many).  Scw = the 4x4 similarity, row-major.  mp_flags[i] bit 1 = isBad(), bit 2 = the
 * point is one of pKF->GetMapPoints() (spAlreadyFound, :1147). */
int refmatch_fuse_sim3(const refmatch_frame *kfr, const float *cam, const float *bounds, const float *scale, const float *inv_sigma2,
                       int nlevels, float log_scale, const float *Scw, int n_mp, const uint8_t *mp_flags, const float *mp_xyz,
                       const uint8_t *mp_desc, const float *mp_dist, const float *mp_normal, float th, int32_t *out_best)
{
    int total = 0;
    for (int pass = 0; pass < 2; ++pass) {
        KeyFrame kf;
        fill_keyframe(kf, *kfr, cam, bounds, scale, inv_sigma2, nlevels, log_scale);
        std::vector<MapPoint> mps;
        fill_points(mps, n_mp, mp_flags, mp_xyz, mp_desc, mp_dist, mp_normal);
        std::vector<MapPoint *> vp(n_mp);
        for (int i = 0; i < n_mp; ++i) {
            vp[i] = &mps[i];
            if (mp_flags[i] & 4) kf.mvpMapPoints.push_back(&mps[i]); /* beyond N: only GetMapPoints() sees them */
        }
        ORBmatcher matcher(0.6f, true);
        cv::Mat S = mat44(Scw);
        if (pass == 0) {
            for (int i = 0; i < n_mp; ++i) {
                kf.mnLastGetMapPoint = -1;
                std::vector<MapPoint *> one(1, vp[i]), rep(1, static_cast<MapPoint *>(NULL));
                const int nf = matcher.Fuse(&kf, S, one, th, rep);
                out_best[i] = nf ? (int32_t)kf.mnLastGetMapPoint : -1;
                total += nf;
            }
        } else {
            std::vector<MapPoint *> rep(n_mp, static_cast<MapPoint *>(NULL));
            if (matcher.Fuse(&kf, S, vp, th, rep) != total) std::abort();
        }
    }
    return total;
}

/* ORBmatcher::SearchByProjection(KeyFrame *pKF, cv::Mat Scw, const vector<MapPoint*> &vpPoints, vector<MapPoint*>
 * &vpMatched, int th) (src/ORBmatcher.cc:353-485, LoopClosing::ComputeSim3).  occupied[idx] != 0: vpMatched[idx] is set on
 * entry; mp_flags[i] bit 1 = isBad(), bit 2 = the point is in vpMatched already (spAlreadyFound, :367-368) -- such points
 * are what the occupied slots hold here.  out_match[idx] = the point this call put into vpMatched[idx], -1 elsewhere. */
int refmatch_search_by_projection_kf_sim3(const refmatch_frame *kfr, const float *cam, const float *bounds, const float *scale,
                                          const float *inv_sigma2, int nlevels, float log_scale, const float *Scw, int n_mp,
                                          const uint8_t *mp_flags, const float *mp_xyz, const uint8_t *mp_desc, const float *mp_dist,
                                          const float *mp_normal, const uint8_t *occupied, int th, int32_t *out_match)
{
    KeyFrame kf;
    fill_keyframe(kf, *kfr, cam, bounds, scale, inv_sigma2, nlevels, log_scale);
    std::vector<MapPoint> mps;
    fill_points(mps, n_mp, mp_flags, mp_xyz, mp_desc, mp_dist, mp_normal);
    std::vector<MapPoint *> vp(n_mp), found;
    for (int i = 0; i < n_mp; ++i) {
        vp[i] = &mps[i];
        if (mp_flags[i] & 4) found.push_back(&mps[i]);
    }
    MapPoint held;
    std::vector<MapPoint *> matched(kfr->n, static_cast<MapPoint *>(NULL));
    size_t k = 0;
    for (int idx = 0; idx < kfr->n; ++idx)
        if (occupied && occupied[idx]) matched[idx] = k < found.size() ? found[k++] : &held;
    if (k < found.size()) std::abort(); /* the scenario must offer a slot per already-found point */
    std::vector<MapPoint *> before = matched;
    ORBmatcher matcher(0.75f, true);
    const int nm = matcher.SearchByProjection(&kf, mat44(Scw), vp, matched, th);
    for (int idx = 0; idx < kfr->n; ++idx)
        out_match[idx] = (matched[idx] && matched[idx] != before[idx]) ? (int32_t)(matched[idx] - mps.data()) : -1;
    return nm;
}

/* ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th) (src/ORBmatcher.cc:1262-1501, LoopClosing::ComputeSim3).
 * Each keyframe: frame record with its pose in Tcw, map points per keypoint (flags bit 0 = GetMapPointMatches()[i] != NULL,
 * bit 1 = isBad()).  pre12[i1] = the pKF2 keypoint whose map point vpMatches12[i1] holds on entry, or -1.
 * out12[i1] = the pKF2 keypoint whose map point vpMatches12[i1] holds on return, -1 = NULL.  Returns nFound. */
int refmatch_search_by_sim3(const refmatch_frame *k1, const refmatch_frame *k2, const float *cam, const float *bounds,
                            const float *scale, const float *inv_sigma2, int nlevels, float log_scale, const uint8_t *flags1,
                            const float *xyz1, const uint8_t *desc1, const float *dist1, const uint8_t *flags2, const float *xyz2,
                            const uint8_t *desc2, const float *dist2, const int32_t *pre12, float s12, const float *R12,
                            const float *t12, float th, int32_t *out12)
{
    KeyFrame kf1, kf2;
    fill_keyframe(kf1, *k1, cam, bounds, scale, inv_sigma2, nlevels, log_scale);
    fill_keyframe(kf2, *k2, cam, bounds, scale, inv_sigma2, nlevels, log_scale);
    std::vector<MapPoint> m1, m2;
    fill_points(m1, k1->n, flags1, xyz1, desc1, dist1, NULL);
    fill_points(m2, k2->n, flags2, xyz2, desc2, dist2, NULL);
    for (int i = 0; i < k1->n; ++i) if (flags1[i] & 1) kf1.mvpMapPoints[i] = &m1[i];
    for (int i = 0; i < k2->n; ++i) {
        if (flags2[i] & 1) kf2.mvpMapPoints[i] = &m2[i];
        m2[i].mbInKF = true; /* GetIndexInKeyFrame(pKF2) of a pKF2 point: its own index (:1297) */
        m2[i].mnIdxInKF = i;
    }
    std::vector<MapPoint *> v12(k1->n, static_cast<MapPoint *>(NULL));
    for (int i = 0; i < k1->n; ++i) if (pre12[i] >= 0) v12[i] = &m2[pre12[i]];
    cv::Mat R(3, 3, CV_32F), t(3, 1, CV_32F);
    for (int r = 0; r < 3; ++r) {
        t.at<float>(r) = t12[r];
        for (int c = 0; c < 3; ++c) R.at<float>(r, c) = R12[3 * r + c];
    }
    ORBmatcher matcher(0.75f, true);
    const int nf = matcher.SearchBySim3(&kf1, &kf2, v12, s12, R, t, th);
    for (int i = 0; i < k1->n; ++i) out12[i] = v12[i] ? (int32_t)(v12[i] - m2.data()) : -1;
    return nf;
}

} // extern "C"



