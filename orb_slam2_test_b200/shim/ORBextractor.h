/*
 * ORBextractor.h -- drop-in replacement for the reference's include/ORBextractor.h.
 *
 * Same namespace, class name and public interface as ORB_SLAM2::ORBextractor
 * (reference include/ORBextractor.h:46-144): the five-argument constructor, operator()(image,
 * mask, keypoints, descriptors), the six scale getters and the public mvImagePyramid, so that the
 * reference's src/Frame.cc (:101-113, :310-316, :626-757) and src/Tracking.cc (:121-127) compile
 * and link against it unchanged.  The body is a thin host shim over the C ABI of liborb_b200.so
 * (include/borb.h): every pixel operation runs in hand-written sm_100a CUDA kernels.  There is no
 * CPU fallback: if the CUDA library cannot run, the constructor / operator() throw
 * std::runtime_error.
 *
 * Deviations from the reference header, all source-compatible for the in-tree callers:
 *   - includes <opencv2/core/core.hpp> instead of the legacy <opencv/cv.h> (absent in OpenCV 4);
 *   - the destructor releases the device handle (the reference's is empty), the class is not
 *     copyable (callers only ever hold ORBextractor* -- src/Tracking.cc:121-127);
 *   - ExtractorNode stays declared for source compatibility but its DivideNode is not provided:
 *     the quadtree lives on the device (csrc/k_quadtree.cuh), and no caller outside the
 *     reference's own ORBextractor.cc uses the type.  Likewise the protected ComputePyramid /
 *     ComputeKeyPointsOctTree / DistributeOctTree / ComputeKeyPointsOld members are gone.
 */
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <list>
#include <map>
#include <string>
#include <vector>

#include <opencv2/core/core.hpp>

struct borb_extractor;
struct borb_vocabulary;

namespace ORB_SLAM2
{

class ExtractorNode
{
public:
    ExtractorNode() : bNoMore(false) {}

    void DivideNode(ExtractorNode &n1, ExtractorNode &n2, ExtractorNode &n3, ExtractorNode &n4);

    std::vector<cv::KeyPoint> vKeys;
    cv::Point2i UL, UR, BL, BR;
    std::list<ExtractorNode>::iterator lit;
    bool bNoMore;
};

// Device copy of the vocabulary (extension).  The reference's ORBVocabulary (include/ORBVocabulary.h, DBoW2's
// TemplatedVocabulary<FORB::TDescriptor, FORB>) stays what System owns and KeyFrameDatabase scores with; this class holds
// the same tree on the GPU for the per-frame transform: construct it from the text file System::System loads with
// mpVocabulary->loadFromTextFile(strVocFile) (reference src/System.cc), once, next to that call.
class DeviceVocabulary
{
public:
    explicit DeviceVocabulary(const std::string &strVocFile, int device = -1);
    ~DeviceVocabulary();
    // transform(features, BowVector, FeatureVector, levelsup) for descriptors in host memory (KeyFrame::ComputeBoW on a
    // keyframe no extractor holds); DBoW2::BowVector IS-A std::map<unsigned int, double> and DBoW2::FeatureVector IS-A
    // std::map<unsigned int, std::vector<unsigned int> >, so mBowVec / mFeatVec bind to these references directly
    void transform(const cv::Mat &descriptors, std::map<unsigned int, double> &bowVec,
                   std::map<unsigned int, std::vector<unsigned int> > &featVec, int levelsup = 4);
    // ORBmatcher::SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, vector<MapPoint*> &vpMatches12) (reference
    // src/ORBmatcher.cc:634-777, LoopClosing::ComputeSim3) on the device.  Each keyframe arrives as its mDescriptors,
    // mvKeysUn, mFeatVec and one flag byte per keypoint (bit 0: GetMapPointMatches()[i] != NULL, bit 1: that point
    // isBad()).  vnMatches12[i1] = index of the pKF2 keypoint whose map point vpMatches12[i1] receives, or -1; returns nmatches.
    int SearchByBoW(const cv::Mat &descriptors1, const std::vector<cv::KeyPoint> &keysUn1, const unsigned char *mapPointFlags1,
                    const std::map<unsigned int, std::vector<unsigned int> > &featVec1, const cv::Mat &descriptors2,
                    const std::vector<cv::KeyPoint> &keysUn2, const unsigned char *mapPointFlags2,
                    const std::map<unsigned int, std::vector<unsigned int> > &featVec2, float nnRatio, bool checkOrientation,
                    std::vector<int> &vnMatches12);
    // ORBmatcher::SearchForTriangulation(KeyFrame *pKF1, KeyFrame *pKF2, cv::Mat F12, vector<pair<size_t, size_t> >
    // &vMatchedPairs, const bool bOnlyStereo) (reference src/ORBmatcher.cc:779-966, LocalMapping::CreateNewMapPoints) on the
    // device.  Keyframes as for SearchByBoW (bit 0 of the flag byte: the keypoint has a map point) plus mvuRight (empty =
    // monocular); Ow1 = pKF1->GetCameraCenter() (3x1 CV_32F), Tcw2 = pKF2->GetPose() (4x4 CV_32F), pKF2's intrinsics and
    // level tables.  Returns nmatches; vMatchedPairs as the reference fills it.
    int SearchForTriangulation(const cv::Mat &descriptors1, const std::vector<cv::KeyPoint> &keysUn1, const unsigned char *mapPointFlags1,
                               const std::map<unsigned int, std::vector<unsigned int> > &featVec1, const std::vector<float> &uRight1,
                               const cv::Mat &Ow1, const cv::Mat &descriptors2, const std::vector<cv::KeyPoint> &keysUn2,
                               const unsigned char *mapPointFlags2, const std::map<unsigned int, std::vector<unsigned int> > &featVec2,
                               const std::vector<float> &uRight2, const cv::Mat &Tcw2, float fx2, float fy2, float cx2, float cy2,
                               const std::vector<float> &scaleFactors2, const std::vector<float> &levelSigma2_2, const cv::Mat &F12,
                               std::vector<std::pair<size_t, size_t> > &vMatchedPairs, bool bOnlyStereo, bool checkOrientation = true);
    borb_vocabulary *get() const { return voc; }

private:
    DeviceVocabulary(const DeviceVocabulary &);
    DeviceVocabulary &operator=(const DeviceVocabulary &);
    borb_vocabulary *voc;
};

class ORBextractor
{
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);

    ~ORBextractor();

    // Compute the ORB features and descriptors on an image.
    // Mask is ignored, as in the reference.
    void operator()(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint> &keypoints,
                    cv::OutputArray descriptors);

    int inline GetLevels() { return nlevels; }

    float inline GetScaleFactor() { return scaleFactor; }

    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }

    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }

    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }

    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    // Level views into page-locked memory owned by the device library; valid until the next
    // operator() call on this instance (the reference re-creates them every call as well).
    std::vector<cv::Mat> mvImagePyramid;

    // -- extensions (not in the reference) --
    // CUDA device of this instance: environment variable BORB_DEVICE at construction, default 0.
    int GetDevice() const { return device; }
    // Mono / RGB-D callers never read mvImagePyramid; turning the download off saves the
    // device-to-host copy of the 8 levels.  Default: on (stereo matching needs it).
    void SetPyramidDownload(bool on) { downloadPyramid = on; }
    // The step before the extractor, on the device.  (1) Colour frames: with colour input enabled,
    // operator() also accepts CV_8UC3 / CV_8UC4 images and converts them exactly as
    // Tracking::GrabImage* does with cv::cvtColor (reference src/Tracking.cc:176-201); rgb = the
    // settings file's Camera.RGB.  (2) Stereo rectification: M1, M2 are the CV_32F maps of
    // cv::initUndistortRectifyMap; operator() then takes the RAW frame and rectifies it exactly as
    // cv::remap(im, rect, M1, M2, cv::INTER_LINEAR) does (Examples/Stereo/stereo_euroc.cc:136-137).
    // Empty maps remove the rectification.
    void SetColorInput(bool enabled, bool rgb) { colorInput = enabled; colorRGB = rgb; }
    void SetRectifyMap(const cv::Mat &M1, const cv::Mat &M2);
    // The per-keypoint step after the extractor, on the device: with a camera set (mK 3x3 CV_32F,
    // mDistCoef 4x1 / 5x1 CV_32F as Tracking builds them, src/Tracking.cc:56-80), every operator()
    // call also undistorts the keypoints and bins them into the 64 x 48 grid; FillFrame then
    // hands back what Frame::UndistortKeyPoints (reference src/Frame.cc:542-572),
    // Frame::ComputeImageBounds (:574-609) and Frame::AssignFeaturesToGrid (:292-307) compute:
    // mvKeysUn, mGrid[64][48] and {mnMinX, mnMaxX, mnMinY, mnMaxY}.
    void SetCamera(const cv::Mat &K, const cv::Mat &DistCoef);
    void FillFrame(const std::vector<cv::KeyPoint> &mvKeys, std::vector<cv::KeyPoint> &mvKeysUn,
                   std::vector<std::size_t> mGrid[64][48], float bounds[4]);
    // Frame::ComputeStereoMatches (reference src/Frame.cc:619-834) on the device, for *this as the
    // left extractor: fills mvuRight / mvDepth (size = left keypoint count, -1 where unmatched)
    // from the last operator() results of both instances.  Call it after both extraction threads
    // have joined (src/Frame.cc:110-113); neither pyramid has to be downloaded for it.
    void ComputeStereoMatches(ORBextractor &right, float mbf, float mb, std::vector<float> &mvuRight,
                              std::vector<float> &mvDepth);
    // ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, th, bMono) (reference
    // src/ORBmatcher.cc:1503-1668) on the device, for *this as CurrentFrame.mpORBextractorLeft after
    // its operator() call (SetCamera on; useStereo: ComputeStereoMatches has run).  The last frame
    // arrives as the flat arrays include/borb.h describes (entry i = LastFrame keypoint i: flags bit 0
    // map point present, bit 1 Observations() > 0, bit 2 outlier; GetWorldPos(), GetDescriptor(),
    // mvKeys[i].octave, mvKeysUn[i].angle); TcwCur / TcwLast are the 4x4 CV_32F mTcw.
    // vnMatches[i2] = i of the map point CurrentFrame.mvpMapPoints[i2] receives, or -1; returns nmatches.
    int SearchByProjection(const cv::Mat &TcwCur, const cv::Mat &TcwLast, int nLast, const float *worldPos,
                           const unsigned char *mpDescriptors, const unsigned char *flags, const int *octave,
                           const float *angle, float mbf, float mb, float th, bool bMono, bool checkOrientation,
                           bool useStereo, std::vector<int> &vnMatches);
    // ORBmatcher::SearchByProjection(Frame &F, const vector<MapPoint*> &vpMapPoints, th) (reference
    // src/ORBmatcher.cc:59-154, the matcher of Tracking::SearchLocalPoints) on the device, *this = F's left
    // extractor.  Entry i = vpMapPoints[i] as Frame::isInFrustum left it: proj[4i..] = mTrackProjX, mTrackProjY,
    // mTrackProjXR, mTrackViewCos; level[i] = mnTrackScaleLevel; flags bit 0 = mbTrackInView && !isBad(), bit 1 =
    // Observations() > 0; occupied[i2] = 2 where F.mvpMapPoints[i2] holds a point with observations (may be NULL).
    // vnMatches[i2] = i put into F.mvpMapPoints[i2] by this call, -1 = slot unchanged; returns nmatches.
    int SearchLocalPoints(int nPoints, const float *proj, const unsigned char *mpDescriptors, const unsigned char *flags,
                          const int *level, const unsigned char *occupied, float th, float nnRatio, bool useStereo,
                          std::vector<int> &vnMatches);
    // ORBmatcher::SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, const set<MapPoint*> &sAlreadyFound, th, ORBdist)
    // (reference src/ORBmatcher.cc:1670-1798, the second and third search of Tracking::Relocalization) on the device,
    // *this = CurrentFrame's left extractor after its operator() call (SetCamera on).  Entry i = keyframe keypoint i:
    // flags bit 0 = GetMapPointMatches()[i] != NULL, bit 1 = isBad(), bit 2 = member of sAlreadyFound; GetWorldPos(),
    // GetDescriptor(), minMaxDistance[2i] = mfMinDistance, [2i + 1] = mfMaxDistance, angle[i] = pKF->mvKeysUn[i].angle;
    // occupied[i2] != 0 where CurrentFrame.mvpMapPoints[i2] is set on entry (may be NULL).
    // vnMatches[i2] = i put into CurrentFrame.mvpMapPoints[i2] by this call, -1 = slot unchanged; returns nmatches.
    int SearchByProjectionKF(const cv::Mat &TcwCur, int nKF, const float *worldPos, const unsigned char *mpDescriptors,
                             const unsigned char *flags, const float *minMaxDistance, const float *angle,
                             const unsigned char *occupied, float th, int ORBdist, bool checkOrientation,
                             std::vector<int> &vnMatches);
    // A KeyFrame on the device: the keyframe's mvKeys, mDescriptors (and mvuRight, empty = monocular) put back into this
    // instance as if operator() had just produced them; mvKeysUn, mGrid and the bounds are rebuilt on the device (SetCamera
    // on).  The keyframe matchers below then take *this as that keyframe.  cols, rows = the image size.
    void LoadKeyFrame(const std::vector<cv::KeyPoint> &keys, const cv::Mat &descriptors, const std::vector<float> &uRight,
                      int cols, int rows);
    // The search inside ORBmatcher::Fuse(KeyFrame *pKF, const vector<MapPoint*> &vpMapPoints, const float th) (reference
    // src/ORBmatcher.cc:968-1131, LocalMapping::SearchInNeighbors), *this = pKF after LoadKeyFrame.  Entry i =
    // vpMapPoints[i]: flags bit 0 = non-NULL, bit 1 = isBad(), bit 2 = IsInKeyFrame(pKF); GetWorldPos(), GetDescriptor(),
    // minMaxDistance[2i] = mfMinDistance, [2i + 1] = mfMaxDistance, normals[3i..] = GetNormal().  Tcw = pKF->GetPose() (4x4
    // CV_32F), Ow = pKF->GetCameraCenter() (3x1).  vnBestIdx[i] = the keypoint of pKF the reference hands to
    // pKF->GetMapPoint(bestIdx) (:1105) or -1; returns nFused.  The caller then runs the reference's own :1103-1127 per i.
    int FuseSearch(const cv::Mat &Tcw, const cv::Mat &Ow, float bf, int nPoints, const float *worldPos,
                   const unsigned char *mpDescriptors, const unsigned char *flags, const float *minMaxDistance,
                   const float *normals, float th, bool useStereo, std::vector<int> &vnBestIdx);
    // The loop-closing matchers (reference src/ORBmatcher.cc; LoopClosing::ComputeSim3 / SearchAndFuse), *this = the keyframe
    // after LoadKeyFrame.  Points as FuseSearch's (flags bit 0 = an entry, bit 1 = isBad(), bit 2 = already found); the
    // matrices are the caller's cv::Mat results, exactly as the reference derives them: Rcw = sRcw / scw, tcw = Scw.col(3) /
    // scw, Ow = -Rcw.t() * tcw (:360-365, :1140-1146); sR12, sR21, t21 (:1279-1281).
    //   FuseSim3Search: the search of Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (:1133-1260): vnBestIdx[i] = the keypoint
    //   handed to pKF->GetMapPoint (:1242) or -1; the caller runs :1240-1256 per i.
    int FuseSim3Search(const cv::Mat &Rcw, const cv::Mat &tcw, const cv::Mat &Ow, int nPoints, const float *worldPos,
                       const unsigned char *mpDescriptors, const unsigned char *flags, const float *minMaxDistance,
                       const float *normals, float th, std::vector<int> &vnBestIdx);
    //   SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (:353-485): occupied[idx] != 0 where vpMatched[idx] is set on
    //   entry (may be NULL); vnMatches[idx] = the point put into vpMatched[idx] by this call or -1; returns nmatches.
    int SearchByProjectionSim3(const cv::Mat &Rcw, const cv::Mat &tcw, const cv::Mat &Ow, int nPoints, const float *worldPos,
                               const unsigned char *mpDescriptors, const unsigned char *flags, const float *minMaxDistance,
                               const float *normals, const unsigned char *occupied, int th, std::vector<int> &vnMatches);
    //   SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th) (:1262-1501): *this = pKF1, kf2 = pKF2 (both loaded);
    //   points i = GetMapPointMatches()[i] of either keyframe; T1w / T2w = the poses (4x4); sR12, t12, sR21, t21 as :1279-1281;
    //   vnMatches12 in/out = vpMatches12 as pKF2 keypoint indices (-1 = NULL); returns nFound.
    int SearchBySim3(ORBextractor &kf2, const cv::Mat &T1w, const cv::Mat &T2w, const cv::Mat &sR12, const cv::Mat &t12,
                     const cv::Mat &sR21, const cv::Mat &t21, int n1, const float *worldPos1, const unsigned char *desc1,
                     const unsigned char *flags1, const float *minMaxDistance1, int n2, const float *worldPos2,
                     const unsigned char *desc2, const unsigned char *flags2, const float *minMaxDistance2, float th,
                     std::vector<int> &vnMatches12);
    // ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) (reference
    // src/ORBmatcher.cc:487-632, Tracking::MonocularInitialization) on the device, *this = F2's extractor after its
    // operator() call (SetCamera on).  F1 arrives as its mvKeysUn (octave, angle) and descriptors (CV_8U, n x 32).
    // vbPrevMatched is updated as the reference does; returns nmatches.
    int SearchForInitialization(const std::vector<cv::KeyPoint> &F1KeysUn, const cv::Mat &F1Descriptors,
                                std::vector<cv::Point2f> &vbPrevMatched, std::vector<int> &vnMatches12, int windowSize,
                                float nnRatio, bool checkOrientation);

    // Frame::ComputeBoW (reference src/Frame.cc:532-539: mpORBvocabulary->transform(vCurrentDesc, mBowVec, mFeatVec, 4))
    // on the device, from the descriptors the last operator() call left there; fills the frame's mBowVec / mFeatVec
    // with the reference's values bit for bit.
    void ComputeBoW(DeviceVocabulary &voc, std::map<unsigned int, double> &mBowVec,
                    std::map<unsigned int, std::vector<unsigned int> > &mFeatVec, int levelsup = 4);
    // ORBmatcher::SearchByBoW(KeyFrame *pKF, Frame &F, vector<MapPoint*> &vpMapPointMatches) (reference
    // src/ORBmatcher.cc:195-351) on the device, *this = F's extractor after ComputeBoW.  The keyframe arrives as
    // pKF->mDescriptors, pKF->mvKeysUn, pKF->mFeatVec and one flag byte per keypoint (bit 0: GetMapPointMatches()[i] !=
    // NULL, bit 1: that point isBad()).  vnMatches[iF] = index of the keyframe keypoint whose map point F's keypoint iF
    // receives, or -1; returns nmatches.
    int SearchByBoW(const cv::Mat &kfDescriptors, const std::vector<cv::KeyPoint> &kfKeysUn, const unsigned char *kfMapPointFlags,
                    const std::map<unsigned int, std::vector<unsigned int> > &kfFeatVec, float nnRatio, bool checkOrientation,
                    std::vector<int> &vnMatches);

protected:
    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;

    std::vector<int> mnFeaturesPerLevel;

    std::vector<int> umax;

    std::vector<float> mvScaleFactor;
    std::vector<float> mvInvScaleFactor;
    std::vector<float> mvLevelSigma2;
    std::vector<float> mvInvLevelSigma2;

private:
    ORBextractor(const ORBextractor &);
    ORBextractor &operator=(const ORBextractor &);

    borb_extractor *handle;
    int device;
    bool downloadPyramid;
    bool colorInput, colorRGB;
    int inputFormat; // BORB_* format the handle is currently set to
};

} // namespace ORB_SLAM2

#endif
