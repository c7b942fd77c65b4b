/*
 * oracle/frame_driver.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * C entry point over the reference's OWN src/Frame.cc (compiled where it lies, see oracle/Makefile:
 * _ref/libreframe.so) so that the rows built around the extractor are pinned to reference code:
 *   Frame::Frame(imLeft, imRight, ...)   src/Frame.cc:79-151  runs, on the reference's own objects,
 *     ExtractORB x2 on two threads (:110-113) -> UndistortKeyPoints (:116, :542-572)
 *     -> ComputeStereoMatches (:124, :619-834) -> ComputeImageBounds (:132, :574-609)
 *     -> AssignFeaturesToGrid (:150, :292-307)
 * and this driver copies mvKeys, mDescriptors, mvKeysUn, mvuRight, mvDepth, mGrid and the bounds out.
 *
 * The reference reads its member `mb` in ComputeStereoMatches (:661) BEFORE the constructor assigns
 * it (:148): the value is whatever the Frame's storage held.  The driver makes that defined by
 * constructing the Frame with placement new in storage where `mb` already holds the value the
 * caller wants (the intended one is mbf / fx).
 *
 * ORBmatcher::DescriptorDistance / TH_LOW / TH_HIGH come from oracle/framecompat/ORBmatcher.h
 * (restated; src/ORBmatcher.cc is not compilable here), cv::undistortPoints from
 * oracle/frame_oracle.c (pinned to cv2 4.13.0).
 */
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <new>

#include "Frame.h"

using namespace ORB_SLAM2;

extern "C" {

#ifdef FRAME_DRIVER_NO_MONOTONE
/* built against the B200 drop-in ORBextractor (tests/shim/Makefile: libshimframe.so): there is no host quadtree
 * and so no allocator to switch -- the device quadtree IS the canonical tie-break */
static void reforb_force_monotone(int) {}
#else
void reforb_force_monotone(int on); /* oracle/ref_driver.cpp, monotone-allocator build */
#endif

struct refframe_out {
    int n_left, n_right;
    float bounds[4];
};

/* Buffers sized by the caller for `cap` keypoints: kps 28 B each, desc 32 B each, un_xy 2 floats,
 * u_right / depth 1 float, grid_count 64*48 ints, grid_idx cap ints.  Returns 0, or -1 if cap is
 * too small.  canonical != 0: the extractor threads use the monotone node allocator. */
int refframe_stereo(const uint8_t *left, const uint8_t *right, int cols, int rows, int nfeatures,
                    float scale, int nlevels, int ini_th, int min_th, const float *K4, const float *dist,
                    int n_dist, float bf, float mb, int canonical, int cap, uint8_t *kps_l, uint8_t *desc_l,
                    uint8_t *kps_r, uint8_t *desc_r, float *un_xy, float *u_right, float *depth,
                    int32_t *grid_count, int32_t *grid_idx, refframe_out *out)
{
    static_assert(sizeof(cv::KeyPoint) == 28, "cv::KeyPoint layout");
    cv::Mat imL(rows, cols, CV_8UC1, (void *)left), imR(rows, cols, CV_8UC1, (void *)right);
    cv::Mat K(3, 3, CV_32F), D(n_dist, 1, CV_32F);
    std::memset(K.data, 0, 36);
    K.at<float>(0, 0) = K4[0]; K.at<float>(1, 1) = K4[1]; K.at<float>(0, 2) = K4[2]; K.at<float>(1, 2) = K4[3];
    K.at<float>(2, 2) = 1.f;
    for (int i = 0; i < n_dist; ++i) D.at<float>(i) = dist[i];
    ORBextractor extL(nfeatures, scale, nlevels, ini_th, min_th), extR(nfeatures, scale, nlevels, ini_th, min_th);
    ORBVocabulary voc;

    Frame::mbInitialComputations = true; /* bounds and grid constants are recomputed for this camera */
    void *mem = std::calloc(1, sizeof(Frame) + 64);
    Frame *probe = reinterpret_cast<Frame *>(mem);
    *reinterpret_cast<float *>(reinterpret_cast<char *>(mem) +
                               (reinterpret_cast<char *>(&probe->mb) - reinterpret_cast<char *>(probe))) = mb;
    /* canonical quadtree tie-break inside the two extraction threads */
    reforb_force_monotone(canonical);
    Frame *F = new (mem) Frame(imL, imR, 0.0, &extL, &extR, &voc, K, D, bf, 35.f * bf / K4[0]);
    reforb_force_monotone(0);

    int rc = 0;
    const int nl = (int)F->mvKeys.size(), nr = (int)F->mvKeysRight.size();
    out->n_left = nl;
    out->n_right = nr;
    out->bounds[0] = Frame::mnMinX; out->bounds[1] = Frame::mnMaxX;
    out->bounds[2] = Frame::mnMinY; out->bounds[3] = Frame::mnMaxY;
    if (nl > cap || nr > cap) rc = -1;
    else {
        if (nl) std::memcpy(kps_l, F->mvKeys.data(), (size_t)nl * 28);
        if (nr) std::memcpy(kps_r, F->mvKeysRight.data(), (size_t)nr * 28);
        for (int i = 0; i < nl; ++i) std::memcpy(desc_l + (size_t)i * 32, F->mDescriptors.ptr(i), 32);
        for (int i = 0; i < nr; ++i) std::memcpy(desc_r + (size_t)i * 32, F->mDescriptorsRight.ptr(i), 32);
        for (int i = 0; i < nl; ++i) {
            un_xy[2 * i] = F->mvKeysUn[i].pt.x;
            un_xy[2 * i + 1] = F->mvKeysUn[i].pt.y;
            u_right[i] = F->mvuRight[i];
            depth[i] = F->mvDepth[i];
        }
        int k = 0;
        for (int i = 0; i < FRAME_GRID_COLS; ++i)
            for (int j = 0; j < FRAME_GRID_ROWS; ++j) {
                grid_count[i * FRAME_GRID_ROWS + j] = (int)F->mGrid[i][j].size();
                for (size_t q = 0; q < F->mGrid[i][j].size(); ++q) grid_idx[k++] = (int)F->mGrid[i][j][q];
            }
    }
    F->~Frame();
    std::free(mem);
    return rc;
}

} // extern "C"
