/*
 * include/borb.h -- C ABI of the B200 ORB extraction library (liborb_b200.so).
 *
 * This is the boundary a maintainer of the reference binds instead of the CPU body of
 * ORB_SLAM2::ORBextractor (reference: include/ORBextractor.h:46-144, src/ORBextractor.cc).
 * Plain pointers and sizes only; no OpenCV, torch or CUDA types.  The C++ host shim
 * (orb_slam2_test_b200/shim/ORBextractor.{h,cc}) keeps the reference's class interface on top
 * of these calls; INTEGRATION.md shows the binding.
 *
 * All functions return BORB_OK (0) or a negative borb_status; borb_last_error() gives the
 * message for the calling thread.  There is no CPU fallback: without a CUDA device every compute
 * entry point fails with BORB_ERR_CUDA.
 *
 * Threading (reference: src/Frame.cc:110-113 runs the left and right extractor on two threads):
 * one handle may be used by one thread at a time; different handles are fully independent (own
 * stream, graph, device and pinned buffers).
 */
#ifndef BORB_H
#define BORB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct borb_extractor borb_extractor;

typedef enum borb_status {
    BORB_OK = 0,
    BORB_ERR_ARG = -1,      /* bad argument / unsupported parameter range               */
    BORB_ERR_CUDA = -2,     /* CUDA runtime failure, or no device                        */
    BORB_ERR_GEOMETRY = -3, /* image too small for the pyramid / aspect ratio < 0.5 (the
                               reference divides by zero there, src/ORBextractor.cc:674-680) */
    BORB_ERR_CAPACITY = -4, /* batch larger than max_batch, or internal buffer overflow   */
    BORB_ERR_STATE = -5     /* result requested before any extraction                     */
} borb_status;

/* Binary image of cv::KeyPoint (7 x 4 bytes), the element type of the reference's output
 * vector (src/ORBextractor.cc:1361-1395). */
typedef struct borb_keypoint {
    float x, y;     /* pt, original-image coordinates (level coords * mvScaleFactor[octave]) */
    float size;     /* (float)(int)(31 * mvScaleFactor[octave])                              */
    float angle;    /* degrees [0,360)                                                       */
    float response; /* FAST score                                                            */
    int32_t octave;
    int32_t class_id; /* always -1 */
} borb_keypoint;

/* -- lifetime: replaces ORBextractor::ORBextractor (src/ORBextractor.cc:432-521) ------------- */

/* device < 0 selects the current CUDA device.  max_batch >= 1 is the largest number of frames
 * one borb_extract_* call may carry (frames of one call run through the kernels together). */
int borb_create(int nfeatures, float scale_factor, int nlevels, int ini_th_fast, int min_th_fast,
                int device, int max_batch, borb_extractor **out);
void borb_destroy(borb_extractor *h);
const char *borb_last_error(void);
const char *borb_version(void);

/* -- constructor-derived tables: replace the inline getters (include/ORBextractor.h:78-98) ---- */
int borb_levels(const borb_extractor *h);
float borb_scale_factor(const borb_extractor *h); /* GetScaleFactor(): the double member as float */
/* which: 0 mvScaleFactor, 1 mvInvScaleFactor, 2 mvLevelSigma2, 3 mvInvLevelSigma2; out[nlevels] */
int borb_table(const borb_extractor *h, int which, float *out);
int borb_features_per_level(const borb_extractor *h, int32_t *out); /* mnFeaturesPerLevel */
int borb_umax(const borb_extractor *h, int32_t *out);               /* 16 entries         */

/* -- the step before the extractor (optional): replaces cv::cvtColor in Tracking::GrabImage*
 *    (src/Tracking.cc:176-201, :213-227, :245-259) and the stereo drivers' cv::remap
 *    (Examples/Stereo/stereo_euroc.cc:136-137), in the order the reference applies them ------ */

/* channel layout of the 8-bit images handed to the extract calls from now on (default gray);
 * RGB/BGR = the reference's Camera.RGB 1/0.  Colour frames are converted on the device with
 * OpenCV's 8-bit formula, Y = (B*3735 + G*19235 + R*9798 + 2^14) >> 15. */
#define BORB_GRAY8 0
#define BORB_BGR8 1
#define BORB_RGB8 2
#define BORB_BGRA8 3
#define BORB_RGBA8 4
int borb_set_input_format(borb_extractor *h, int format);
/* Rectification maps as cv::initUndistortRectifyMap(..., CV_32F, M1, M2) gives them: map_x/map_y
 * [rows][map_step floats], (cols, rows) = size of the RECTIFIED image, which becomes the extractor's
 * geometry; the extract calls' (cols, rows, step) then describe the raw source frames.  Sampling is
 * cv::remap(INTER_LINEAR, BORDER_CONSTANT 0) bit for bit (1/32-pixel fixed point).  NULL, NULL
 * removes the map.  The maps are copied; the call synchronises the handle. */
int borb_set_rectify_map(borb_extractor *h, const float *map_x, const float *map_y, int cols, int rows,
                         size_t map_step);

/* -- extraction: replaces ORBextractor::operator() (src/ORBextractor.cc:1330-1397) ------------ */

/* flags */
#define BORB_WANT_PYRAMID 1u /* also bring the padded pyramid back to pinned host memory
                                (mvImagePyramid, include/ORBextractor.h:101)                  */

/* n_frames 8-bit single-channel images in HOST memory: frame f starts at img + f*frame_stride,
 * rows are `step` bytes apart.  Copies them to the device, runs the whole path, copies the
 * results back and returns when they are readable through borb_result(). */
int borb_extract_host(borb_extractor *h, const uint8_t *img, int n_frames, int cols, int rows,
                      size_t step, size_t frame_stride, unsigned flags);

/* The two halves of borb_extract_host: submit enqueues copy-in, kernels and copy-out on the
 * handle's stream and returns; wait blocks until the results are readable.  Several handles kept
 * in flight this way overlap their transfers and kernels. */
int borb_submit_host(borb_extractor *h, const uint8_t *img, int n_frames, int cols, int rows,
                     size_t step, size_t frame_stride, unsigned flags);
int borb_wait(borb_extractor *h);

/* Page-locked host memory for image batches: borb_submit_host copies from such memory straight
 * to the device (densely packed frames only) instead of staging through its own pinned buffer. */
void *borb_host_alloc(size_t bytes);
void borb_host_free(void *p);

/* Same, but the images already are in DEVICE memory and the results stay on the device until
 * borb_fetch() is called (kernel-only timing, pipelines that keep descriptors on the GPU). */
int borb_extract_device(borb_extractor *h, const uint8_t *dev_img, int n_frames, int cols,
                        int rows, size_t step, size_t frame_stride);
int borb_fetch(borb_extractor *h, unsigned flags);
/* Put a frame extracted earlier -- a KeyFrame, which carries its Frame's mvKeys and mDescriptors (src/KeyFrame.cc) -- back
 * into the handle as if an extract call had just produced it (frame 0 of a one-frame call): with a camera set, mvKeysUn,
 * mGrid and the bounds are rebuilt on the device from the records, and every matcher below takes the handle as that
 * frame.  cols, rows = the image size the frame came from.  No counterpart in the reference (its keyframes live in host
 * memory); this is how the keyframe matchers reach a keyframe on the device.
 * u_right = the keyframe's mvuRight (what BORB_MATCH_USE_STEREO then reads) or NULL for a monocular one. */
int borb_load_frame(borb_extractor *h, const borb_keypoint *kps, const uint8_t *desc, const float *u_right, int n, int cols,
                    int rows);
/* blocks until everything enqueued on the handle's stream has finished */
int borb_sync(borb_extractor *h);

/* Results of the last call, frame f: pointers into the handle's pinned staging memory, valid
 * until the next extract call on this handle.  *n may be 0 (desc/kps then still valid). */
int borb_result(const borb_extractor *h, int frame, const borb_keypoint **kps,
                const uint8_t **desc, int *n);
/* keypoint counts of all frames of the last call: out[n_frames] */
int borb_counts(const borb_extractor *h, int32_t *out);
/* Level `level` of frame f: *data points at pixel (0,0) of the level image inside its padded
 * (19-pixel reflect-101 frame) parent, rows *step bytes apart -- the same view
 * mvImagePyramid[level] is (src/ORBextractor.cc:1417).  Needs BORB_WANT_PYRAMID. */
int borb_pyramid(const borb_extractor *h, int frame, int level, const uint8_t **data, int *cols,
                 int *rows, size_t *step);
/* -- the per-keypoint step after the extractor: replaces Frame::UndistortKeyPoints
 *    (src/Frame.cc:542-572), Frame::ComputeImageBounds (:574-609) and
 *    Frame::AssignFeaturesToGrid / PosInGrid (:292-307, :510-520) -------------------------------- */

/* K = {fx, fy, cx, cy} and dist = {k1, k2, p1, p2[, k3]} as the reference holds them in mK and
 * mDistCoef (floats).  While a camera is set, every extract call also produces, per frame, the
 * undistorted keypoint positions (cv::undistortPoints(pts, pts, K, dist, Mat(), K), bit for bit;
 * identity when k1 == 0 exactly as the reference's shortcut) and the 64 x 48 keypoint grid.
 * K = NULL turns it off. */
int borb_set_camera(borb_extractor *h, const float *K, const float *dist, int n_dist);
/* frame f of the last call: xy_un[2n] = mvKeysUn[i].pt; cell (i, j), i < 64 columns, j < 48 rows, is
 * cell_idx[cell_start[i*48 + j] .. cell_start[i*48 + j + 1]) = mGrid[i][j] (keypoint indices in
 * ascending order); cell_start has 64*48 + 1 entries; bounds[4] = mnMinX, mnMaxX, mnMinY, mnMaxY. */
int borb_frame_grid(const borb_extractor *h, int frame, const float **xy_un, const int32_t **cell_start,
                    const int32_t **cell_idx, float *bounds);

/* -- stereo association: replaces Frame::ComputeStereoMatches (src/Frame.cc:619-834) ---------- */

#define BORB_STEREO_KEEP_ON_DEVICE 1u /* enqueue only; no download, borb_stereo_result unavailable */

/* For every keypoint of `left`'s last call, frame by frame against the same frame of `right`'s
 * last call (both handles on one device, same image size and pyramid parameters, same number of
 * frames): best right keypoint by descriptor distance inside the row band, 11x11 SAD refinement on
 * the two device-resident pyramids, parabola sub-pixel fit, the median-SAD outlier cut -- giving
 * mvuRight[i] and mvDepth[i] (-1 where unmatched).  mbf = baseline * fx (Frame::mbf), mb = the
 * baseline used for the search range minD = 0, maxD = mbf / mb (src/Frame.cc:661-663; the reference reads
 * its mb member there before its constructor assigns it, :124 vs :148 -- pass mbf/fx).  The work
 * runs on left's stream, ordered after right's; neither pyramid leaves the device.
 * The row-band bookkeeping uses (size_t)y of the LEFT keypoint exactly as the reference does. */
int borb_stereo_match(borb_extractor *left, borb_extractor *right, float mbf, float mb, unsigned flags);
/* frame f of the last borb_stereo_match: u_right[n], depth[n] parallel to borb_result()'s
 * keypoints of `left`; *n_matched = entries with depth > 0.  Pinned memory of `left`. */
int borb_stereo_result(const borb_extractor *left, int frame, const float **u_right, const float **depth,
                       int *n, int *n_matched);

/* -- frame-to-frame matcher: replaces ORBmatcher::SearchByProjection(Frame &CurrentFrame,
 *    const Frame &LastFrame, const float th, const bool bMono) (src/ORBmatcher.cc:1503-1668), the
 *    matcher of Tracking::TrackWithMotionModel (src/Tracking.cc:868-936) ------------------------- */

/* The last frame as the matcher reads it: entry i is LastFrame keypoint i.
 *   flags[i]  bit 0: LastFrame.mvpMapPoints[i] != NULL, bit 1: that map point's Observations() > 0,
 *             bit 2: LastFrame.mvbOutlier[i]
 *   xyz[3i..] pMP->GetWorldPos(), desc[32i..] pMP->GetDescriptor() (read only where bit 0 is set)
 *   octave[i] LastFrame.mvKeys[i].octave, angle[i] LastFrame.mvKeysUn[i].angle
 *   Tcw       LastFrame.mTcw, rows 0-2 (3 x 4, row-major) */
typedef struct borb_last_frame {
    int n;
    const float *xyz;
    const uint8_t *desc;
    const uint8_t *flags;
    const int32_t *octave;
    const float *angle;
    const float *Tcw;
} borb_last_frame;

#define BORB_MATCH_MONO 1u          /* bMono: no forward / backward octave rule */
#define BORB_MATCH_NO_ORIENTATION 2u /* ORBmatcher(nnratio, checkOri = false) */
#define BORB_MATCH_USE_STEREO 4u    /* CurrentFrame.mvuRight = the result of the last borb_stereo_match with
                                       `cur` as the left handle (else all -1, a monocular frame) */

/* CurrentFrame = frame `frame` of cur's last extract call: keypoints, descriptors, mvKeysUn, mGrid and the
 * image bounds stay on the device (borb_set_camera must be on: K gives fx, fy, cx, cy).  Tcw_cur =
 * CurrentFrame.mTcw rows 0-2; mbf, mb = CurrentFrame.mbf, mb; th as the caller passes it (15 mono, 7
 * stereo, doubled on the retry, src/Tracking.cc:893-906).  match[i2], i2 < *n_cur, = index into `last` of
 * the map point CurrentFrame.mvpMapPoints[i2] ends up with, or -1 (the reference expects the caller to
 * have cleared mvpMapPoints, :890); *nmatches = the function's return value.  match must hold
 * borb_result()'s n entries of that frame. */
int borb_search_by_projection(borb_extractor *cur, int frame, const borb_last_frame *last, const float *Tcw_cur,
                              float mbf, float mb, float th, unsigned flags, int32_t *match, int *n_cur,
                              int *nmatches);
/* -- local-map matcher: replaces ORBmatcher::SearchByProjection(Frame &F, const vector<MapPoint*>
 *    &vpMapPoints, const float th) (src/ORBmatcher.cc:59-154), the matcher of
 *    Tracking::SearchLocalPoints (src/Tracking.cc:1055-1103) -------------------------------------- */

/* The local map as Frame::isInFrustum (src/Frame.cc:348-407) left it: entry i is vpMapPoints[i].
 *   flags[i]   bit 0: pMP->mbTrackInView && !pMP->isBad(), bit 1: pMP->Observations() > 0
 *   proj[4i..] mTrackProjX, mTrackProjY, mTrackProjXR, mTrackViewCos;  level[i] mnTrackScaleLevel
 *   desc[32i..] pMP->GetDescriptor() */
typedef struct borb_map_points {
    int n;
    const float *proj;
    const uint8_t *desc;
    const uint8_t *flags;
    const int32_t *level;
} borb_map_points;

/* F = frame `frame` of cur's last extract call (as borb_search_by_projection).  occupied[i2], i2 < the
 * frame's keypoint count: 2 where F.mvpMapPoints[i2] holds a map point with Observations() > 0 on entry
 * (such keypoints are skipped, :92-94), anything else where the slot is free to take (NULL = all free).
 * th as the caller passes it (1, 3 for RGB-D, 5 after a relocalisation), nnratio = the matcher's
 * mfNNratio (0.8 in SearchLocalPoints).  flags: BORB_MATCH_USE_STEREO.  match[i2] = index into mps of the
 * map point this call puts into F.mvpMapPoints[i2], -1 where the slot keeps what it held. */
int borb_search_local_points(borb_extractor *cur, int frame, const borb_map_points *mps, const uint8_t *occupied,
                             float th, float nnratio, unsigned flags, int32_t *match, int *n_cur, int *nmatches);

/* -- relocalisation matcher: replaces ORBmatcher::SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF,
 *    const set<MapPoint*> &sAlreadyFound, const float th, const int ORBdist) (src/ORBmatcher.cc:1670-1798), the second and
 *    third search of Tracking::Relocalization (src/Tracking.cc: th 10 / ORBdist 100, then 3 / 64) ------------------- */

/* The keyframe's map points as the matcher reads them: entry i is keyframe keypoint i.
 *   flags[i]   bit 0: pKF->GetMapPointMatches()[i] != NULL, bit 1: that point isBad(), bit 2: it is in sAlreadyFound
 *   xyz[3i..]  pMP->GetWorldPos(), desc[32i..] pMP->GetDescriptor() (read only where bit 0 is set)
 *   dist[2i]   pMP->mfMinDistance, dist[2i + 1] pMP->mfMaxDistance -- the raw members: the 0.8f / 1.2f factors of
 *              GetMinDistanceInvariance / GetMaxDistanceInvariance and MapPoint::PredictScale(dist3D, &CurrentFrame)
 *              (src/MapPoint.cc:520-530, :575-590, with glibc's logf) are applied on the device
 *   angle[i]   pKF->mvKeysUn[i].angle */
typedef struct borb_keyframe_points {
    int n;
    const float *xyz;
    const uint8_t *desc;
    const uint8_t *flags;
    const float *dist;
    const float *angle;
} borb_keyframe_points;

/* CurrentFrame = frame `frame` of cur's last extract call (as borb_search_by_projection; mfLogScaleFactor and
 * mnScaleLevels are the handle's).  Tcw_cur = CurrentFrame.mTcw rows 0-2.  occupied[i2], i2 < the frame's keypoint
 * count: non-zero where CurrentFrame.mvpMapPoints[i2] is set on entry -- every set slot is skipped (:1735-1736), NULL =
 * none set.  orb_dist in 0..255.  flags: BORB_MATCH_NO_ORIENTATION.  match[i2] = index of the keyframe keypoint whose map
 * point this call puts into CurrentFrame.mvpMapPoints[i2], -1 where the slot keeps what it held; *nmatches = the return
 * value. */
int borb_search_by_projection_keyframe(borb_extractor *cur, int frame, const borb_keyframe_points *kf, const float *Tcw_cur,
                                       const uint8_t *occupied, float th, int orb_dist, unsigned flags, int32_t *match,
                                       int *n_cur, int *nmatches);

/* -- the fuse matcher: replaces the search inside ORBmatcher::Fuse(KeyFrame *pKF, const vector<MapPoint*> &vpMapPoints,
 *    const float th) (src/ORBmatcher.cc:968-1131, LocalMapping::SearchInNeighbors) ---------------------------------- */

/* The map points as the matcher reads them: entry i is vpMapPoints[i].
 *   flags[i]    bit 0: vpMapPoints[i] != NULL, bit 1: isBad(), bit 2: IsInKeyFrame(pKF)
 *   xyz, desc   GetWorldPos(), GetDescriptor();  dist[2i], [2i + 1] = mfMinDistance, mfMaxDistance (raw members, as
 *               borb_keyframe_points);  normal[3i..] = GetNormal() */
typedef struct borb_fuse_points {
    int n;
    const float *xyz;
    const uint8_t *desc;
    const uint8_t *flags;
    const float *dist;
    const float *normal;
} borb_fuse_points;

/* pKF = frame `frame` of kf's last call: a keyframe put into the handle by borb_load_frame (camera set; its mvScaleFactors,
 * mvInvLevelSigma2, mfLogScaleFactor are the handle's).  Tcw = pKF's pose rows 0-2 (GetRotation / GetTranslation), Ow =
 * pKF->GetCameraCenter(), bf = pKF->mbf, th as the caller passes it (3.0 by default).  flags: BORB_MATCH_USE_STEREO = the
 * keyframe's mvuRight was loaded (its stereo keypoints take the 3-dof chi-square test :1063-1070).
 * best_idx[i] = the keypoint bestIdx of pKF that map point i is fused into (bestDist <= TH_LOW, :1103) or -1; *nfused = the
 * function's return value.  What follows in the reference (:1103-1127: Replace / AddObservation / AddMapPoint, in the order
 * of i) mutates the map and stays with the caller: it reads nothing the choice depends on. */
int borb_fuse_search(borb_extractor *kf, int frame, const borb_fuse_points *mps, const float *Tcw, const float *Ow, float bf,
                     float th, unsigned flags, int32_t *best_idx, int *nfused);

/* -- the loop-closing matchers: the searches of ORBmatcher::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint)
 *    (src/ORBmatcher.cc:1133-1260, LoopClosing::SearchAndFuse), ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched,
 *    th) (:353-485) and ORBmatcher::SearchBySim3 (:1262-1501), both LoopClosing::ComputeSim3 ------------------------- */

/* How the points reach the keyframe's image.  T1 = the first transform, 3 x 4 row-major [R | t]: [Rcw | tcw] of the
 * similarity as :360-364 / :1140-1145 derive them (Rcw = sRcw / scw, tcw = Scw.col(3) / scw), or a keyframe's pose
 * (SearchBySim3: R1w, t1w).  T2 = a second transform applied to the camera-frame point ([sR21 | t21] resp. [sR12 | t12],
 * :1279-1281) or NULL.  Ow = the camera centre -Rcw.t() * tcw for dist3D = |p3Dw - Ow| and the viewing-angle test, or NULL:
 * dist3D of the camera-frame point itself (SearchBySim3, :1330 / :1418).  The matrices are the CALLER's cv::Mat results,
 * so their arithmetic is OpenCV's own. */
typedef struct borb_projection {
    const float *T1;
    const float *T2;
    const float *Ow;
} borb_projection;

/* pKF = frame `frame` of kf's last call (borb_load_frame).  pts as borb_fuse_points: bit 0 of a flag = the entry is a point,
 * bit 1 = isBad(), bit 2 = already found (in pKF->GetMapPoints(), :1147; in vpMatched, :367); normal = NULL switches the
 * viewing-angle test off (SearchBySim3 has none).  best_idx[i] = the keypoint with the first smallest distance among the
 * admitted ones when that distance is <= max_dist (TH_LOW = 50 for Fuse, :1240), else -1.  What Fuse then does with the
 * map (:1240-1256) stays with the caller. */
int borb_project_search(borb_extractor *kf, int frame, const borb_fuse_points *pts, const borb_projection *pr, float th,
                        int max_dist, int32_t *best_idx, int *nfound);
/* occupied[idx] != 0 where vpMatched[idx] is set on entry (NULL = none).  match[idx], idx < *n_kf, = the point this call puts
 * into vpMatched[idx] or -1; *nmatches = the return value. */
int borb_search_by_projection_sim3(borb_extractor *kf, int frame, const borb_fuse_points *pts, const borb_projection *pr,
                                   const uint8_t *occupied, int th, int32_t *match, int *n_kf, int *nmatches);
/* kf1 / kf2 = handles holding pKF1 / pKF2 (frame 0); pts1[i1] = pKF1->GetMapPointMatches()[i1] (bit 0 non-NULL, bit 1 bad),
 * pts2 likewise; T1w / T2w = the keyframes' poses; T12 = [s12 R12 | t12], T21 = [sR21 | t21] (:1279-1281); pre12[i1] = the
 * pKF2 keypoint whose map point vpMatches12[i1] holds on entry or -1 (NULL = none).  match12[i1] = the same on return;
 * *nfound = the return value. */
int borb_search_by_sim3(borb_extractor *kf1, borb_extractor *kf2, const borb_fuse_points *pts1, const borb_fuse_points *pts2,
                        const float *T1w, const float *T2w, const float *T12, const float *T21, float th, const int32_t *pre12,
                        int32_t *match12, int *nfound);

/* -- monocular-initialisation matcher: replaces ORBmatcher::SearchForInitialization(Frame &F1, Frame &F2,
 *    vector<cv::Point2f> &vbPrevMatched, vector<int> &vnMatches12, int windowSize)
 *    (src/ORBmatcher.cc:487-632), called by Tracking::MonocularInitialization (src/Tracking.cc:570-646) --- */

/* The initial frame F1 as the matcher reads it: octave[i] = F1.mvKeysUn[i].octave (only level 0 is
 * matched), angle[i] = F1.mvKeysUn[i].angle, desc[32i..] = F1.mDescriptors.row(i). */
typedef struct borb_init_frame {
    int n;
    const int32_t *octave;
    const float *angle;
    const uint8_t *desc;
} borb_init_frame;

/* F2 = frame `frame` of cur's last extract call (device-resident keypoints, descriptors, mvKeysUn, mGrid,
 * bounds; borb_set_camera on).  prev_matched[2 n] = vbPrevMatched on entry, updated for the matched entries
 * as the reference does (:621-623).  nnratio = mfNNratio (0.9 there), flags: BORB_MATCH_NO_ORIENTATION.
 * matches12[i] = vnMatches12[i] (index into F2's keypoints or -1); *nmatches = the return value. */
int borb_search_for_initialization(borb_extractor *cur, int frame, const borb_init_frame *f1, float *prev_matched,
                                   int window_size, float nnratio, unsigned flags, int32_t *matches12, int *nmatches);

/* -- pipelined sequence feeder: replaces the drivers' frame loop around the extractor
 *    (Examples/Stereo/stereo_kitti.cc:66-108 -> System::TrackStereo -> Frame::Frame, src/Frame.cc:79-151, whose two
 *    ExtractORB threads :110-113 and ComputeStereoMatches :124 are the device work) for callers that have many
 *    frames: SURVEY.md 8(e), one feeder per GPU --------------------------------------------------------------------
 *
 * A stream owns one extractor (two for stereo pairs) and a ring of `slots` batches.  Each slot has page-locked
 * input and result buffers on the host and their twins on the device; uploads, kernels and downloads run on three
 * CUDA streams chained by events, so batch k+1's upload and batch k-1's download overlap batch k's kernels on ONE
 * stream object.  The kernels of a batch are one cached CUDA graph per (slot, eye).  Nothing is exchanged between
 * streams (or GPUs): a sequence is sharded item-wise over one stream per GPU with no collective. */
typedef struct borb_stream borb_stream;

#define BORB_STREAM_STEREO 1u  /* an item is a stereo pair: left + right extractor, Frame::ComputeStereoMatches on
                                  the device (needs borb_stream_set_stereo), mvuRight / mvDepth come back with it */
#define BORB_STREAM_PYRAMID 2u /* also bring every item's padded pyramids back (mvImagePyramid for a host-side
                                  consumer; ~4x the download) */

/* batch = items (frames, or pairs) per push, slots >= 2 = ring depth.  Geometry is fixed at open. */
int borb_stream_open(int nfeatures, float scale_factor, int nlevels, int ini_th_fast, int min_th_fast, int device,
                     int cols, int rows, int batch, int slots, unsigned flags, borb_stream **out);
void borb_stream_close(borb_stream *s);
/* Frame::mbf and the baseline of the search range, as borb_stereo_match takes them */
int borb_stream_set_stereo(borb_stream *s, float mbf, float mb);
/* as borb_set_camera: undistorted positions + 64 x 48 grid come back with every item (both eyes: the right eye's
 * are computed and not downloaded, as the reference's Frame never undistorts mvKeysRight) */
int borb_stream_set_camera(borb_stream *s, const float *K, const float *dist, int n_dist);

/* The page-locked input buffer of the slot the next push will use (eye 0 = left / mono, 1 = right): a producer
 * that decodes or renders frames straight into it saves the staging copy.  capacity = batch * rows * cols bytes,
 * frames densely packed.  BORB_ERR_STATE when the ring is full (pop first). */
int borb_stream_input(borb_stream *s, int eye, uint8_t **buf, size_t *capacity);
/* Enqueue one batch of n <= batch items and return at once.  left (and right, for pairs): n images in host
 * memory, rows `step` bytes apart, images `frame_stride` bytes apart; page-locked dense memory is uploaded from
 * where it lies, anything else is packed into the slot's buffer first.  left == NULL: the frames already are in
 * the slot's input buffers (borb_stream_input).  tag travels with the batch.  BORB_ERR_STATE when the ring is full. */
int borb_stream_push(borb_stream *s, const uint8_t *left, const uint8_t *right, int n, size_t step,
                     size_t frame_stride, uint64_t tag);
/* Block until the OLDEST batch in flight has its results on the host; they stay readable (and its slot stays
 * reserved) until the next pop or borb_stream_release.  BORB_ERR_STATE when nothing is in flight. */
int borb_stream_pop(borb_stream *s, uint64_t *tag, int *n);
int borb_stream_release(borb_stream *s);
int borb_stream_in_flight(const borb_stream *s); /* pushed, not yet popped */
int borb_stream_slots_free(const borb_stream *s); /* pushes possible right now */
/* results of the popped batch, item i, as borb_result / borb_stereo_result / borb_frame_grid / borb_pyramid */
int borb_stream_result(const borb_stream *s, int eye, int item, const borb_keypoint **kps, const uint8_t **desc,
                       int *n);
int borb_stream_stereo_result(const borb_stream *s, int item, const float **u_right, const float **depth, int *n,
                              int *n_matched);
int borb_stream_frame_grid(const borb_stream *s, int item, const float **xy_un, const int32_t **cell_start,
                           const int32_t **cell_idx, float *bounds);
int borb_stream_pyramid(const borb_stream *s, int eye, int item, int level, const uint8_t **data, int *cols,
                        int *rows, size_t *step);
/* bytes one full batch moves over the host link: [0] host->device, [1] device->host */
int borb_stream_link_bytes(const borb_stream *s, uint64_t *bytes);
/* kernel launches one push enqueues (bench.py's gpu_launches) */
int borb_stream_launches_per_push(const borb_stream *s);

/* The feeder thread.  produce(user, index, left, right, max_items) writes up to max_items frames (pairs) of
 * batch `index` (0, 1, 2 ...) into the page-locked buffers it is handed and returns how many (0 = end of the
 * sequence, < 0 = abort); consume(user, s, tag, n) is called for every finished batch, in push order, tag = index,
 * and reads it through the borb_stream_* result calls.  Both run on the feeder thread.  start returns at once;
 * join waits for the end of the sequence and returns the first error (message through borb_last_error). */
typedef int (*borb_produce_fn)(void *user, uint64_t index, uint8_t *left, uint8_t *right, int max_items);
typedef void (*borb_consume_fn)(void *user, borb_stream *s, uint64_t tag, int n);
int borb_stream_start(borb_stream *s, borb_produce_fn produce, borb_consume_fn consume, void *user);
int borb_stream_join(borb_stream *s);

/* -- allocation-free variant of the per-keypoint step.  NO COUNTERPART IN THE REFERENCE: added
 *    for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 *    loops, harnesses that forbid allocation) --------------------------------------------------------------------
 *
 * The arithmetic of borb_set_camera + borb_frame_grid (cv::undistortPoints with P = K, the image bounds, the
 * 64 x 48 keypoint grid; bit for bit the same results), but on buffers the CALLER owns, all in DEVICE memory of the
 * current CUDA device, and with the kernel's scratch memory (one int per keypoint slot) in a static device array of
 * BORB_NOALLOC_MAX_SLOTS entries (1 MiB, held by every GPU that loads the library until the process exits) instead
 * of an allocation.  The call allocates and frees nothing (no cudaMalloc / cudaFree / cudaMallocHost, no library
 * routine that allocates) and does not synchronise: the kernel is enqueued on `stream` (a cudaStream_t cast to
 * void *; NULL = the default stream) and the outputs are valid when that stream reaches them.
 *   dev_kps [n_frames][kp_stride] keypoint records, dev_counts [n_frames] keypoints per frame (<= kp_stride)
 *   K = {fx, fy, cx, cy}, dist = {k1, k2, p1, p2[, k3]} (host memory, as borb_set_camera; k1 == 0: identity)
 *   dev_xy_un [n_frames][kp_stride][2], dev_cell_start [n_frames][64*48 + 1], dev_cell_idx [n_frames][kp_stride],
 *   dev_bounds [n_frames][4]: as borb_frame_grid describes them
 * n_frames * kp_stride > BORB_NOALLOC_MAX_SLOTS is refused with BORB_ERR_CAPACITY.  Every call on a device shares
 * the one scratch array (nothing in it needs clearing between calls: each entry a call reads it has written first),
 * so calls that OVERLAP on the GPU -- two streams, two threads -- race: serialise them. */
#define BORB_NOALLOC_MAX_SLOTS 262144
int borb_undistort_grid_noalloc(const borb_keypoint *dev_kps, const int32_t *dev_counts, int n_frames, int kp_stride,
                                int cols, int rows, const float *K, const float *dist, int n_dist, float *dev_xy_un,
                                int32_t *dev_cell_start, int32_t *dev_cell_idx, float *dev_bounds, void *stream);

/* -- bag of words: Frame::ComputeBoW (src/Frame.cc:532-539) and the matcher that consumes it ------------------
 *
 * The vocabulary is DBoW2's TemplatedVocabulary<FORB::TDescriptor, FORB> (include/ORBVocabulary.h): a k-ary tree
 * of 256-bit descriptors, words at the leaves.  borb_voc_load_text reads the text format of
 * TemplatedVocabulary::loadFromTextFile (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1338-1421) that
 * System::System loads (src/System.cc); borb_voc_create takes the same tree as arrays.  The tree is uploaded once and
 * used read-only: any number of extractor handles on that device may share it. */
typedef struct borb_vocabulary borb_vocabulary;
int borb_voc_load_text(const char *path, int device, borb_vocabulary **out);
/* nodes 1 .. n_nodes in file order (node 0 = root): parent id (< own id), leaf flag (> 0: a word; word ids count the
 * flagged nodes in order), 32-byte descriptor, weight.  scoring / weighting: DBoW2's ScoringType / WeightingType
 * (BowVector.h:36-53; ORBvoc.txt: L1_NORM = 0, TF_IDF = 0). */
int borb_voc_create(int k, int L, int scoring, int weighting, int n_nodes, const int32_t *parent, const uint8_t *is_leaf,
                    const uint8_t *desc, const double *weight, int device, borb_vocabulary **out);
void borb_voc_destroy(borb_vocabulary *v);
int borb_voc_info(const borb_vocabulary *v, int *k, int *L, int *n_nodes /* with the root */, int *n_words,
                  int *scoring, int *weighting);

/* TemplatedVocabulary::transform(features, BowVector&, FeatureVector&, levelsup) (:1127-1206) on frame `frame` of
 * h's last extract call, from the descriptors that call left on the device; the reference calls it with
 * levelsup = 4 (Frame::ComputeBoW, src/Frame.cc:536).  Results stay on the device for borb_search_by_bow and come
 * back in one copy.  Up to 16384 keypoint slots per frame (BORB_ERR_CAPACITY beyond). */
int borb_compute_bow(borb_extractor *h, int frame, const borb_vocabulary *voc, int levelsup);
/* mBowVec as (word id ascending, value) pairs -- std::map order -- and mFeatVec in CSR form: node ids ascending, the
 * features of node u are feat_idx[node_start[u] .. node_start[u + 1]), ascending (= push_back order).  Values are
 * the reference's doubles bit for bit (weights added in feature order, norm summed in word order).  Pointers into
 * page-locked memory owned by the handle, valid until its next borb_compute_bow. */
int borb_bow_result(borb_extractor *h, int frame, const uint32_t **word_id, const double **word_val, int *n_words,
                    const uint32_t **node_id, const int32_t **node_start, const uint32_t **feat_idx, int *n_nodes);
/* the same transform for descriptors in host memory (KeyFrame::ComputeBoW on a keyframe no handle holds).
 * Capacities: word_id / word_val / node_id / feat_idx [n], node_start [n + 1].  One call at a time per vocabulary. */
int borb_voc_transform(borb_vocabulary *v, const uint8_t *desc, int n, int levelsup, uint32_t *word_id, double *word_val,
                       int *n_words, uint32_t *node_id, int32_t *node_start, uint32_t *feat_idx, int *n_nodes);

/* ORBmatcher::SearchByBoW(KeyFrame *pKF, Frame &F, vector<MapPoint*> &vpMapPointMatches) (src/ORBmatcher.cc:195-351;
 * Tracking::TrackReferenceKeyFrame src/Tracking.cc, Tracking::Relocalization): F = frame `frame` of cur's last call
 * after borb_compute_bow; the keyframe arrives as flat arrays. */
typedef struct borb_keyframe {
    int n;                   /* pKF->N */
    const uint8_t *desc;     /* pKF->mDescriptors, n x 32 */
    const float *angle;      /* pKF->mvKeysUn[i].angle */
    const uint8_t *mp_flags; /* bit 0: GetMapPointMatches()[i] != NULL, bit 1: that map point isBad() */
    int n_nodes;             /* pKF->mFeatVec, CSR as borb_bow_result returns it */
    const uint32_t *node_id;
    const int32_t *node_start;
    const uint32_t *feat_idx;
} borb_keyframe;
/* match [borb_max_keypoints]: per keypoint of F the index of the keyframe keypoint whose map point it receives, or -1
 * (the caller maps it to vpMapPointsKF[index]); *nmatches = the reference's return value.  nnratio = mfNNratio;
 * flags: BORB_MATCH_NO_ORIENTATION = !mbCheckOrientation. */
int borb_search_by_bow(borb_extractor *cur, int frame, const borb_keyframe *kf, float nnratio, unsigned flags,
                       int32_t *match, int *n_cur, int *nmatches);

/* ORBmatcher::SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, vector<MapPoint*> &vpMatches12) (src/ORBmatcher.cc:634-777;
 * LoopClosing::ComputeSim3): both keyframes as flat arrays (their mFeatVec e.g. from borb_voc_transform), matched on the
 * vocabulary's device with its stream and scratch -- one call at a time per vocabulary.  match12 [kf1->n]: per pKF1
 * keypoint the index of the pKF2 keypoint whose map point vpMatches12 receives, or -1; *nmatches = the return value.
 * Up to 65535 keypoints per keyframe. */
int borb_voc_search_by_bow(borb_vocabulary *v, const borb_keyframe *kf1, const borb_keyframe *kf2, float nnratio,
                           unsigned flags, int32_t *match12, int *nmatches);

/* ORBmatcher::SearchForTriangulation(KeyFrame *pKF1, KeyFrame *pKF2, cv::Mat F12, vector<pair<size_t, size_t> >
 * &vMatchedPairs, const bool bOnlyStereo) (src/ORBmatcher.cc:779-966, with CheckDistEpipolarLine :165-182;
 * LocalMapping::CreateNewMapPoints): the keyframes as borb_keyframe records (only bit 0 of mp_flags is read here: the
 * keypoint has a map point) and what the epipolar test reads beside them. */
typedef struct borb_triangulation {
    const float *F12;            /* 3 x 3 row-major, as LocalMapping::ComputeF12 returns it */
    const float *Ow1;            /* pKF1->GetCameraCenter(), 3 floats */
    const float *Tcw2;           /* pKF2's pose rows 0-2 (GetRotation / GetTranslation), 3 x 4 row-major */
    float fx2, fy2, cx2, cy2;    /* pKF2->fx, fy, cx, cy */
    const float *xy1, *xy2;      /* mvKeysUn[i].pt of either keyframe, 2 floats per keypoint */
    const int32_t *octave2;      /* pKF2->mvKeysUn[i].octave */
    const float *u_right1, *u_right2; /* mvuRight; NULL = a monocular keyframe (all -1) */
    int nlevels;
    const float *scale_factors2, *level_sigma2_2; /* pKF2->mvScaleFactors, mvLevelSigma2 */
} borb_triangulation;
#define BORB_MATCH_ONLY_STEREO 8u /* bOnlyStereo */
/* match12 [kf1->n]: per pKF1 keypoint the pKF2 keypoint it is paired with, or -1 -- vMatchedPairs is the list of
 * (i1, match12[i1]) for match12[i1] >= 0 in ascending i1 (:950-956); *nmatches = the return value.  flags:
 * BORB_MATCH_NO_ORIENTATION, BORB_MATCH_ONLY_STEREO.  One call at a time per vocabulary. */
int borb_voc_search_for_triangulation(borb_vocabulary *v, const borb_keyframe *kf1, const borb_keyframe *kf2,
                                      const borb_triangulation *g, unsigned flags, int32_t *match12, int *nmatches);

/* capacity of a frame's result arrays (keypoints per frame can never exceed it; valid after any extract call) */
int borb_max_keypoints(const borb_extractor *h);
/* low 8 bits: rounds the last call's fixed-point iteration took (>= 1); the rest: cell rescans
 * (queries whose cached candidates were all held; see csrc/k_match.cuh) */
int borb_match_rounds(const borb_extractor *cur);

/* geometry only (valid after any extract call) */
int borb_level_size(const borb_extractor *h, int level, int *cols, int *rows);

/* number of kernel launches an extract call of the last call's shape enqueues (a frame or two: the level chain is
 * one launch; before the first call: the count of the batched shape); for bench.py's gpu_launches */
int borb_launches_per_call(const borb_extractor *h);
/* CUDA-event time of the last borb_extract_device call's kernels, milliseconds (<0: n/a) */
float borb_last_device_ms(const borb_extractor *h);

/* CUDA-event stopwatch on the handle's stream (bench.py: the library does not run on torch's
 * stream, so torch.cuda.Event cannot time it) */
int borb_timer_start(borb_extractor *h);
int borb_timer_stop(borb_extractor *h);
float borb_timer_ms(borb_extractor *h);
/* one un-graphed pass over n frames in device memory with an event between stages:
 * ms[0..4] = pyramid, FAST, quadtree, blur, orientation+descriptor */
int borb_stage_times(borb_extractor *h, const uint8_t *dev_img, int n_frames, int cols, int rows,
                     size_t step, size_t frame_stride, float *ms);

/* Per-stage device times of REAL calls: with stage timing on, every extract call records a CUDA
 * event between its stages (event-record nodes inside the captured graph), and
 * borb_last_stage_ms() returns those of the most recent call (ms[0..4] as above). */
int borb_set_stage_timing(borb_extractor *h, int on);
int borb_last_stage_ms(borb_extractor *h, float *ms);
/* time of the pre-step kernel (colour conversion / rectification) inside the last call's pyramid
 * stage, milliseconds; < 0 when stage timing is off or no pre-step is configured */
float borb_last_prestep_ms(borb_extractor *h);

/* -- stage taps for parity tests (device -> caller buffers; not part of the hot path) -------- */
/* FAST candidates of (frame, level) before distribution, coords relative to (16,16);
 * unordered.  Returns count via *n (may exceed cap; only cap entries are written). */
int borb_debug_candidates(borb_extractor *h, int frame, int level, int32_t *xs, int32_t *ys,
                          int32_t *scores, int cap, int *n);
int borb_debug_blurred(borb_extractor *h, int frame, int level, uint8_t *out /* cols*rows */);
/* computeOrbDescriptor (src/ORBextractor.cc:117-157) of the pixel (x, y) of a caller-supplied (already blurred) image
 * at a caller-supplied angle, by the device function the extract kernels use: the known-answer tap for the float
 * form of the pattern rotation.  18 <= x < cols - 18, 18 <= y < rows - 18. */
int borb_debug_describe(borb_extractor *h, const uint8_t *img, int cols, int rows, size_t step, int x, int y,
                        float angle_deg, uint8_t *desc32);
/* the frame-to-frame / local-map matchers run at most this many rounds of their fixed point before one thread finishes
 * the still-open suffix in the reference's order (default 16; 0 restores it): lets a test reach the sequential finish */
void borb_debug_set_match_rounds(int max_rounds);
/* host evaluation of the device sin/cos routine (same inline function, host build) */
void borb_debug_sincos(float radians, float *s, float *c);
/* glibc's logf as the relocalisation matcher's level prediction uses it (csrc/orb_logf.cuh): the host copy and the
 * device copy (one small kernel; allocates) */
void borb_debug_logf(const float *x, float *y, int n);
int borb_debug_logf_device(const float *x, float *y, int n);
/* host evaluation of the per-pixel FAST score used by the cell kernel: 0 = not a FAST-9 corner at
 * threshold t_lo, else the largest threshold at which p still is one */
int borb_debug_fast_score(const uint8_t *p, int pitch, int t_lo);

#ifdef __cplusplus
}
#endif
#endif
