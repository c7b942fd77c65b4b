/*
 * ORBextractor.cc -- host shim: the reference's ORB_SLAM2::ORBextractor interface over the C ABI
 * of liborb_b200.so (include/borb.h).  Replaces the reference's src/ORBextractor.cc.
 *
 * What stays on the host is only what the reference's callers observe: argument conventions,
 * the empty-image early return (reference src/ORBextractor.cc:1333-1334), the CV_8UC1 assert
 * (:1337), clear-and-refill of the caller's containers (:1351-1362, :1395) and cv::Mat headers for
 * mvImagePyramid (:1417).  Scale tables come from the library so that there is a single source.
 */
#include "ORBextractor.h"

#include <cassert>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>

#include "borb.h"

namespace ORB_SLAM2
{

namespace
{
void check(int rc, const char *what)
{
    if (rc != BORB_OK)
        throw std::runtime_error(std::string("ORBextractor (B200): ") + what + ": " + borb_last_error());
}
} // namespace

ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST),
      minThFAST(_minThFAST), handle(NULL), device(0), downloadPyramid(true), colorInput(false), colorRGB(false),
      inputFormat(BORB_GRAY8)
{
    if (const char *d = std::getenv("BORB_DEVICE")) device = std::atoi(d);
    check(borb_create(nfeatures, _scaleFactor, nlevels, iniThFAST, minThFAST, device, 1, &handle), "create");
    mvScaleFactor.resize(nlevels);
    mvInvScaleFactor.resize(nlevels);
    mvLevelSigma2.resize(nlevels);
    mvInvLevelSigma2.resize(nlevels);
    mnFeaturesPerLevel.resize(nlevels);
    umax.resize(16);
    borb_table(handle, 0, &mvScaleFactor[0]);
    borb_table(handle, 1, &mvInvScaleFactor[0]);
    borb_table(handle, 2, &mvLevelSigma2[0]);
    borb_table(handle, 3, &mvInvLevelSigma2[0]);
    borb_features_per_level(handle, &mnFeaturesPerLevel[0]);
    borb_umax(handle, &umax[0]);
    mvImagePyramid.resize(nlevels);
}

ORBextractor::~ORBextractor()
{
    mvImagePyramid.clear();
    borb_destroy(handle);
}

void ORBextractor::operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint> &_keypoints,
                              cv::OutputArray _descriptors)
{
    if (_image.empty())
        return;

    cv::Mat image = _image.getMat();
    assert(image.type() == CV_8UC1 || (colorInput && (image.type() == CV_8UC3 || image.type() == CV_8UC4)));

    // colour frames are converted on the device (extension, see SetColorInput)
    const int format = image.type() == CV_8UC1 ? BORB_GRAY8
                       : image.type() == CV_8UC3 ? (colorRGB ? BORB_RGB8 : BORB_BGR8)
                                                 : (colorRGB ? BORB_RGBA8 : BORB_BGRA8);
    if (format != inputFormat) {
        check(borb_set_input_format(handle, format), "input format");
        inputFormat = format;
    }

    // rows of a non-continuous cv::Mat (ROI) are image.step bytes apart; the library packs them
    check(borb_extract_host(handle, image.data, 1, image.cols, image.rows, (size_t)image.step, 0,
                            downloadPyramid ? BORB_WANT_PYRAMID : 0u),
          "extract");

    const borb_keypoint *kps = NULL;
    const uint8_t *desc = NULL;
    int n = 0;
    check(borb_result(handle, 0, &kps, &desc, &n), "result");

    if (n == 0)
        _descriptors.release();
    else {
        _descriptors.create(n, 32, CV_8U);
        cv::Mat descriptors = _descriptors.getMat();
        for (int i = 0; i < n; ++i) std::memcpy(descriptors.ptr(i), desc + (size_t)i * 32, 32);
    }

    _keypoints.clear();
    _keypoints.reserve(n);
    for (int i = 0; i < n; ++i) {
        const borb_keypoint &k = kps[i];
        _keypoints.push_back(cv::KeyPoint(k.x, k.y, k.size, k.angle, k.response, k.octave, k.class_id));
    }

    if (downloadPyramid) {
        mvImagePyramid.resize(nlevels);
        for (int level = 0; level < nlevels; ++level) {
            const uint8_t *data = NULL;
            int cols = 0, rows = 0;
            size_t step = 0;
            check(borb_pyramid(handle, 0, level, &data, &cols, &rows, &step), "pyramid");
            // header over the library's pinned staging memory, at pixel (0,0) of the level inside
            // its 19-px reflect-101 frame: the same view the reference keeps (:1417)
            mvImagePyramid[level] = cv::Mat(rows, cols, CV_8UC1, (void *)data, step);
        }
    }
}

void ORBextractor::SetRectifyMap(const cv::Mat &M1, const cv::Mat &M2)
{
    if (M1.empty() && M2.empty()) {
        check(borb_set_rectify_map(handle, NULL, NULL, 0, 0, 0), "rectify map");
        return;
    }
    assert(M1.type() == CV_32FC1 && M2.type() == CV_32FC1 && M1.rows == M2.rows && M1.cols == M2.cols &&
           M1.step == M2.step);
    check(borb_set_rectify_map(handle, (const float *)M1.data, (const float *)M2.data, M1.cols, M1.rows,
                               (size_t)M1.step / sizeof(float)),
          "rectify map");
}

void ORBextractor::SetCamera(const cv::Mat &K, const cv::Mat &DistCoef)
{
    if (K.empty()) {
        check(borb_set_camera(handle, NULL, NULL, 0), "camera");
        return;
    }
    assert(K.type() == CV_32F && K.rows == 3 && K.cols == 3 && DistCoef.type() == CV_32F);
    const float k4[4] = {K.at<float>(0, 0), K.at<float>(1, 1), K.at<float>(0, 2), K.at<float>(1, 2)};
    float d[5] = {0, 0, 0, 0, 0};
    const int nd = DistCoef.rows * DistCoef.cols;
    assert(nd == 4 || nd == 5);
    for (int i = 0; i < nd; ++i) d[i] = DistCoef.rows == 1 ? DistCoef.at<float>(0, i) : DistCoef.at<float>(i, 0);
    check(borb_set_camera(handle, k4, d, nd), "camera");
}

void ORBextractor::FillFrame(const std::vector<cv::KeyPoint> &mvKeys, std::vector<cv::KeyPoint> &mvKeysUn,
                             std::vector<std::size_t> mGrid[64][48], float bounds[4])
{
    const float *xy = NULL;
    const int32_t *start = NULL, *idx = NULL;
    check(borb_frame_grid(handle, 0, &xy, &start, &idx, bounds), "frame grid");
    mvKeysUn = mvKeys; // reference :563-571: a copy of each keypoint with pt replaced
    for (size_t i = 0; i < mvKeysUn.size(); ++i) {
        mvKeysUn[i].pt.x = xy[2 * i];
        mvKeysUn[i].pt.y = xy[2 * i + 1];
    }
    for (int i = 0; i < 64; ++i)
        for (int j = 0; j < 48; ++j) {
            const int c = i * 48 + j;
            mGrid[i][j].assign(idx + start[c], idx + start[c + 1]);
        }
}

void ORBextractor::ComputeStereoMatches(ORBextractor &right, float mbf, float mb, std::vector<float> &mvuRight,
                                        std::vector<float> &mvDepth)
{
    check(borb_stereo_match(handle, right.handle, mbf, mb, 0u), "stereo match");
    const float *u = NULL, *z = NULL;
    int n = 0;
    check(borb_stereo_result(handle, 0, &u, &z, &n, NULL), "stereo result");
    mvuRight.assign(u, u + n);
    mvDepth.assign(z, z + n);
}

int ORBextractor::SearchByProjection(const cv::Mat &TcwCur, const cv::Mat &TcwLast, int nLast, const float *worldPos,
                                     const unsigned char *mpDescriptors, const unsigned char *flags, const int *octave,
                                     const float *angle, float mbf, float mb, float th, bool bMono,
                                     bool checkOrientation, bool useStereo, std::vector<int> &vnMatches)
{
    assert(TcwCur.type() == CV_32F && TcwCur.rows >= 3 && TcwCur.cols == 4 && TcwLast.type() == CV_32F);
    float tc[12], tl[12];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 4; ++c) {
            tc[4 * r + c] = TcwCur.at<float>(r, c);
            tl[4 * r + c] = TcwLast.at<float>(r, c);
        }
    borb_last_frame last;
    last.n = nLast; last.xyz = worldPos; last.desc = mpDescriptors; last.flags = flags;
    last.octave = octave; last.angle = angle; last.Tcw = tl;
    vnMatches.assign((size_t)borb_max_keypoints(handle), -1);
    int n = 0, nmatches = 0;
    const unsigned f = (bMono ? BORB_MATCH_MONO : 0u) | (checkOrientation ? 0u : BORB_MATCH_NO_ORIENTATION) |
                       (useStereo ? BORB_MATCH_USE_STEREO : 0u);
    check(borb_search_by_projection(handle, 0, &last, tc, mbf, mb, th, f, vnMatches.data(), &n, &nmatches),
          "search by projection");
    vnMatches.resize((size_t)n);
    return nmatches;
}

int ORBextractor::SearchLocalPoints(int nPoints, const float *proj, const unsigned char *mpDescriptors,
                                    const unsigned char *flags, const int *level, const unsigned char *occupied, float th,
                                    float nnRatio, bool useStereo, std::vector<int> &vnMatches)
{
    borb_map_points mps;
    mps.n = nPoints; mps.proj = proj; mps.desc = mpDescriptors; mps.flags = flags; mps.level = level;
    vnMatches.assign((size_t)borb_max_keypoints(handle), -1);
    int n = 0, nmatches = 0;
    check(borb_search_local_points(handle, 0, &mps, occupied, th, nnRatio, useStereo ? BORB_MATCH_USE_STEREO : 0u,
                                   vnMatches.data(), &n, &nmatches),
          "search local points");
    vnMatches.resize((size_t)n);
    return nmatches;
}

int ORBextractor::SearchByProjectionKF(const cv::Mat &TcwCur, int nKF, const float *worldPos, const unsigned char *mpDescriptors,
                                       const unsigned char *flags, const float *minMaxDistance, const float *angle,
                                       const unsigned char *occupied, float th, int ORBdist, bool checkOrientation,
                                       std::vector<int> &vnMatches)
{
    assert(TcwCur.rows == 4 && TcwCur.cols == 4 && TcwCur.type() == CV_32F);
    float tc[12];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 4; ++c) tc[4 * r + c] = TcwCur.at<float>(r, c);
    borb_keyframe_points kf;
    kf.n = nKF; kf.xyz = worldPos; kf.desc = mpDescriptors; kf.flags = flags; kf.dist = minMaxDistance; kf.angle = angle;
    vnMatches.assign((size_t)borb_max_keypoints(handle), -1);
    int n = 0, nmatches = 0;
    check(borb_search_by_projection_keyframe(handle, 0, &kf, tc, occupied, th, ORBdist,
                                             checkOrientation ? 0u : BORB_MATCH_NO_ORIENTATION, vnMatches.data(), &n, &nmatches),
          "search by projection (keyframe)");
    vnMatches.resize((size_t)n);
    return nmatches;
}

void ORBextractor::LoadKeyFrame(const std::vector<cv::KeyPoint> &keys, const cv::Mat &descriptors, const std::vector<float> &uRight,
                                int cols, int rows)
{
    const int n = (int)keys.size();
    assert(descriptors.rows == n && (n == 0 || descriptors.cols == 32) && (uRight.empty() || (int)uRight.size() == n));
    static_assert(sizeof(cv::KeyPoint) == sizeof(borb_keypoint), "cv::KeyPoint is the 28-byte record of include/borb.h");
    std::vector<unsigned char> desc((size_t)32 * n);
    for (int i = 0; i < n; ++i) memcpy(&desc[(size_t)32 * i], descriptors.ptr(i), 32);
    check(borb_load_frame(handle, reinterpret_cast<const borb_keypoint *>(keys.data()), desc.data(),
                          uRight.empty() ? NULL : uRight.data(), n, cols, rows),
          "load keyframe");
}

int ORBextractor::FuseSearch(const cv::Mat &Tcw, const cv::Mat &Ow, float bf, int nPoints, const float *worldPos,
                             const unsigned char *mpDescriptors, const unsigned char *flags, const float *minMaxDistance,
                             const float *normals, float th, bool useStereo, std::vector<int> &vnBestIdx)
{
    assert(Tcw.rows == 4 && Tcw.cols == 4 && Tcw.type() == CV_32F && Ow.rows == 3 && Ow.cols == 1 && Ow.type() == CV_32F);
    float tc[12], ow[3];
    for (int r = 0; r < 3; ++r) {
        ow[r] = Ow.at<float>(r, 0);
        for (int c = 0; c < 4; ++c) tc[4 * r + c] = Tcw.at<float>(r, c);
    }
    borb_fuse_points mps;
    mps.n = nPoints; mps.xyz = worldPos; mps.desc = mpDescriptors; mps.flags = flags; mps.dist = minMaxDistance; mps.normal = normals;
    vnBestIdx.assign((size_t)std::max(nPoints, 0), -1);
    int nfused = 0;
    check(borb_fuse_search(handle, 0, &mps, tc, ow, bf, th, useStereo ? BORB_MATCH_USE_STEREO : 0u, vnBestIdx.data(), &nfused),
          "fuse search");
    return nfused;
}

namespace {
void rt34(const cv::Mat &R, const cv::Mat &t, float *T) /* [R | t], 3 x 4 row-major */
{
    assert(R.rows == 3 && R.cols == 3 && R.type() == CV_32F && t.rows == 3 && t.cols == 1 && t.type() == CV_32F);
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) T[4 * r + c] = R.at<float>(r, c);
        T[4 * r + 3] = t.at<float>(r, 0);
    }
}
void pose34(const cv::Mat &Tcw, float *T)
{
    assert(Tcw.rows == 4 && Tcw.cols == 4 && Tcw.type() == CV_32F);
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 4; ++c) T[4 * r + c] = Tcw.at<float>(r, c);
}
} // namespace

int ORBextractor::FuseSim3Search(const cv::Mat &Rcw, const cv::Mat &tcw, const cv::Mat &Ow, int nPoints, const float *worldPos,
                                 const unsigned char *mpDescriptors, const unsigned char *flags, const float *minMaxDistance,
                                 const float *normals, float th, std::vector<int> &vnBestIdx)
{
    float T[12], ow[3] = {Ow.at<float>(0, 0), Ow.at<float>(1, 0), Ow.at<float>(2, 0)};
    rt34(Rcw, tcw, T);
    borb_fuse_points pts;
    pts.n = nPoints; pts.xyz = worldPos; pts.desc = mpDescriptors; pts.flags = flags; pts.dist = minMaxDistance; pts.normal = normals;
    borb_projection pr;
    pr.T1 = T; pr.T2 = NULL; pr.Ow = ow;
    vnBestIdx.assign((size_t)std::max(nPoints, 0), -1);
    int nf = 0;
    check(borb_project_search(handle, 0, &pts, &pr, th, 50 /* TH_LOW */, vnBestIdx.data(), &nf), "fuse (Sim3) search");
    return nf;
}

int ORBextractor::SearchByProjectionSim3(const cv::Mat &Rcw, const cv::Mat &tcw, const cv::Mat &Ow, int nPoints, const float *worldPos,
                                         const unsigned char *mpDescriptors, const unsigned char *flags, const float *minMaxDistance,
                                         const float *normals, const unsigned char *occupied, int th, std::vector<int> &vnMatches)
{
    float T[12], ow[3] = {Ow.at<float>(0, 0), Ow.at<float>(1, 0), Ow.at<float>(2, 0)};
    rt34(Rcw, tcw, T);
    borb_fuse_points pts;
    pts.n = nPoints; pts.xyz = worldPos; pts.desc = mpDescriptors; pts.flags = flags; pts.dist = minMaxDistance; pts.normal = normals;
    borb_projection pr;
    pr.T1 = T; pr.T2 = NULL; pr.Ow = ow;
    vnMatches.assign((size_t)borb_max_keypoints(handle), -1);
    int n = 0, nm = 0;
    check(borb_search_by_projection_sim3(handle, 0, &pts, &pr, occupied, th, vnMatches.data(), &n, &nm), "search by projection (Sim3)");
    vnMatches.resize((size_t)n);
    return nm;
}

int ORBextractor::SearchBySim3(ORBextractor &kf2, const cv::Mat &T1w, const cv::Mat &T2w, const cv::Mat &sR12, const cv::Mat &t12,
                               const cv::Mat &sR21, const cv::Mat &t21, int n1, const float *worldPos1, const unsigned char *desc1,
                               const unsigned char *flags1, const float *minMaxDistance1, int n2, const float *worldPos2,
                               const unsigned char *desc2, const unsigned char *flags2, const float *minMaxDistance2, float th,
                               std::vector<int> &vnMatches12)
{
    float a[12], b[12], t12m[12], t21m[12];
    pose34(T1w, a);
    pose34(T2w, b);
    rt34(sR12, t12, t12m);
    rt34(sR21, t21, t21m);
    borb_fuse_points p1, p2;
    p1.n = n1; p1.xyz = worldPos1; p1.desc = desc1; p1.flags = flags1; p1.dist = minMaxDistance1; p1.normal = NULL;
    p2.n = n2; p2.xyz = worldPos2; p2.desc = desc2; p2.flags = flags2; p2.dist = minMaxDistance2; p2.normal = NULL;
    std::vector<int> pre(vnMatches12);
    pre.resize((size_t)std::max(n1, 0), -1);
    vnMatches12.assign((size_t)std::max(n1, 0), -1);
    int nf = 0;
    check(borb_search_by_sim3(handle, kf2.handle, &p1, &p2, a, b, t12m, t21m, th, pre.data(), vnMatches12.data(), &nf), "search by Sim3");
    return nf;
}

int ORBextractor::SearchForInitialization(const std::vector<cv::KeyPoint> &F1KeysUn, const cv::Mat &F1Descriptors,
                                          std::vector<cv::Point2f> &vbPrevMatched, std::vector<int> &vnMatches12,
                                          int windowSize, float nnRatio, bool checkOrientation)
{
    const int n = (int)F1KeysUn.size();
    assert(F1Descriptors.rows == n && (n == 0 || F1Descriptors.cols == 32) && (int)vbPrevMatched.size() == n);
    std::vector<int> octave((size_t)n);
    std::vector<float> angle((size_t)n), prev((size_t)2 * n);
    std::vector<unsigned char> desc((size_t)32 * n);
    for (int i = 0; i < n; ++i) {
        octave[i] = F1KeysUn[i].octave;
        angle[i] = F1KeysUn[i].angle;
        prev[2 * i] = vbPrevMatched[i].x;
        prev[2 * i + 1] = vbPrevMatched[i].y;
        memcpy(&desc[(size_t)32 * i], F1Descriptors.ptr(i), 32);
    }
    borb_init_frame f1;
    f1.n = n; f1.octave = octave.data(); f1.angle = angle.data(); f1.desc = desc.data();
    vnMatches12.assign((size_t)n, -1);
    int nmatches = 0;
    check(borb_search_for_initialization(handle, 0, &f1, prev.data(), windowSize, nnRatio,
                                         checkOrientation ? 0u : BORB_MATCH_NO_ORIENTATION, vnMatches12.data(), &nmatches),
          "search for initialization");
    for (int i = 0; i < n; ++i) vbPrevMatched[i] = cv::Point2f(prev[2 * i], prev[2 * i + 1]);
    return nmatches;
}

namespace {

void fill_maps(const uint32_t *wid, const double *wval, int nw, const uint32_t *nid, const int32_t *nstart, const uint32_t *fidx,
               int nn, std::map<unsigned int, double> &bowVec, std::map<unsigned int, std::vector<unsigned int> > &featVec)
{
    bowVec.clear();
    featVec.clear();
    for (int i = 0; i < nw; ++i) bowVec.insert(bowVec.end(), std::make_pair((unsigned int)wid[i], wval[i])); /* ascending ids */
    for (int u = 0; u < nn; ++u)
        featVec.insert(featVec.end(), std::make_pair((unsigned int)nid[u], std::vector<unsigned int>(fidx + nstart[u], fidx + nstart[u + 1])));
}

} // namespace

DeviceVocabulary::DeviceVocabulary(const std::string &strVocFile, int device) : voc(NULL)
{
    check(borb_voc_load_text(strVocFile.c_str(), device, &voc), "vocabulary");
}

DeviceVocabulary::~DeviceVocabulary() { borb_voc_destroy(voc); }

void DeviceVocabulary::transform(const cv::Mat &descriptors, std::map<unsigned int, double> &bowVec,
                                 std::map<unsigned int, std::vector<unsigned int> > &featVec, int levelsup)
{
    const int n = descriptors.rows;
    assert(n == 0 || descriptors.cols == 32);
    std::vector<unsigned char> d((size_t)32 * n);
    for (int i = 0; i < n; ++i) memcpy(&d[(size_t)32 * i], descriptors.ptr(i), 32);
    std::vector<uint32_t> wid((size_t)n + 1), nid((size_t)n + 1), fidx((size_t)n + 1);
    std::vector<double> wval((size_t)n + 1);
    std::vector<int32_t> nstart((size_t)n + 2);
    int nw = 0, nn = 0;
    check(borb_voc_transform(voc, d.data(), n, levelsup, wid.data(), wval.data(), &nw, nid.data(), nstart.data(), fidx.data(), &nn),
          "vocabulary transform");
    fill_maps(wid.data(), wval.data(), nw, nid.data(), nstart.data(), fidx.data(), nn, bowVec, featVec);
}

void ORBextractor::ComputeBoW(DeviceVocabulary &voc, std::map<unsigned int, double> &mBowVec,
                              std::map<unsigned int, std::vector<unsigned int> > &mFeatVec, int levelsup)
{
    check(borb_compute_bow(handle, 0, voc.get(), levelsup), "compute bow");
    const uint32_t *wid = NULL, *nid = NULL, *fidx = NULL;
    const double *wval = NULL;
    const int32_t *nstart = NULL;
    int nw = 0, nn = 0;
    check(borb_bow_result(handle, 0, &wid, &wval, &nw, &nid, &nstart, &fidx, &nn), "bow result");
    fill_maps(wid, wval, nw, nid, nstart, fidx, nn, mBowVec, mFeatVec);
}

namespace {
struct FlatKeyFrame { /* a keyframe as include/borb.h's borb_keyframe wants it */
    std::vector<unsigned char> desc;
    std::vector<float> angle;
    std::vector<uint32_t> nid, fidx;
    std::vector<int32_t> nstart;
    borb_keyframe rec;
    FlatKeyFrame(const cv::Mat &descriptors, const std::vector<cv::KeyPoint> &keysUn, const unsigned char *flags,
                 const std::map<unsigned int, std::vector<unsigned int> > &featVec)
        : nstart(1, 0)
    {
        const int n = (int)keysUn.size();
        assert(descriptors.rows == n && (n == 0 || descriptors.cols == 32));
        desc.resize((size_t)32 * n);
        angle.resize((size_t)n);
        for (int i = 0; i < n; ++i) {
            memcpy(&desc[(size_t)32 * i], descriptors.ptr(i), 32);
            angle[i] = keysUn[i].angle;
        }
        for (std::map<unsigned int, std::vector<unsigned int> >::const_iterator it = featVec.begin(); it != featVec.end(); ++it) {
            nid.push_back(it->first);
            fidx.insert(fidx.end(), it->second.begin(), it->second.end());
            nstart.push_back((int32_t)fidx.size());
        }
        rec.n = n; rec.desc = desc.data(); rec.angle = angle.data(); rec.mp_flags = flags;
        rec.n_nodes = (int)nid.size(); rec.node_id = nid.data(); rec.node_start = nstart.data(); rec.feat_idx = fidx.data();
    }
};
} // namespace

int DeviceVocabulary::SearchByBoW(const cv::Mat &descriptors1, const std::vector<cv::KeyPoint> &keysUn1,
                                  const unsigned char *mapPointFlags1, const std::map<unsigned int, std::vector<unsigned int> > &featVec1,
                                  const cv::Mat &descriptors2, const std::vector<cv::KeyPoint> &keysUn2,
                                  const unsigned char *mapPointFlags2, const std::map<unsigned int, std::vector<unsigned int> > &featVec2,
                                  float nnRatio, bool checkOrientation, std::vector<int> &vnMatches12)
{
    FlatKeyFrame a(descriptors1, keysUn1, mapPointFlags1, featVec1), b(descriptors2, keysUn2, mapPointFlags2, featVec2);
    vnMatches12.assign(keysUn1.size(), -1);
    int nmatches = 0;
    check(borb_voc_search_by_bow(voc, &a.rec, &b.rec, nnRatio, checkOrientation ? 0u : BORB_MATCH_NO_ORIENTATION,
                                 vnMatches12.data(), &nmatches),
          "search by bow (keyframes)");
    return nmatches;
}

int DeviceVocabulary::SearchForTriangulation(const cv::Mat &descriptors1, const std::vector<cv::KeyPoint> &keysUn1,
                                             const unsigned char *mapPointFlags1, const std::map<unsigned int, std::vector<unsigned int> > &featVec1,
                                             const std::vector<float> &uRight1, const cv::Mat &Ow1, const cv::Mat &descriptors2,
                                             const std::vector<cv::KeyPoint> &keysUn2, const unsigned char *mapPointFlags2,
                                             const std::map<unsigned int, std::vector<unsigned int> > &featVec2,
                                             const std::vector<float> &uRight2, const cv::Mat &Tcw2, float fx2, float fy2, float cx2,
                                             float cy2, const std::vector<float> &scaleFactors2, const std::vector<float> &levelSigma2_2,
                                             const cv::Mat &F12, std::vector<std::pair<size_t, size_t> > &vMatchedPairs,
                                             bool bOnlyStereo, bool checkOrientation)
{
    FlatKeyFrame a(descriptors1, keysUn1, mapPointFlags1, featVec1), b(descriptors2, keysUn2, mapPointFlags2, featVec2);
    assert(F12.rows == 3 && F12.cols == 3 && F12.type() == CV_32F && Tcw2.rows == 4 && Tcw2.cols == 4 && Tcw2.type() == CV_32F);
    assert(Ow1.rows == 3 && Ow1.cols == 1 && Ow1.type() == CV_32F && scaleFactors2.size() == levelSigma2_2.size());
    assert(uRight1.empty() || uRight1.size() == keysUn1.size());
    assert(uRight2.empty() || uRight2.size() == keysUn2.size());
    float F[9], ow[3], tc[12];
    for (int r = 0; r < 3; ++r) {
        ow[r] = Ow1.at<float>(r, 0);
        for (int c = 0; c < 3; ++c) F[3 * r + c] = F12.at<float>(r, c);
        for (int c = 0; c < 4; ++c) tc[4 * r + c] = Tcw2.at<float>(r, c);
    }
    std::vector<float> xy1(2 * keysUn1.size()), xy2(2 * keysUn2.size());
    std::vector<int32_t> oct2(keysUn2.size());
    for (size_t i = 0; i < keysUn1.size(); ++i) { xy1[2 * i] = keysUn1[i].pt.x; xy1[2 * i + 1] = keysUn1[i].pt.y; }
    for (size_t i = 0; i < keysUn2.size(); ++i) { xy2[2 * i] = keysUn2[i].pt.x; xy2[2 * i + 1] = keysUn2[i].pt.y; oct2[i] = keysUn2[i].octave; }
    borb_triangulation g;
    g.F12 = F; g.Ow1 = ow; g.Tcw2 = tc; g.fx2 = fx2; g.fy2 = fy2; g.cx2 = cx2; g.cy2 = cy2;
    g.xy1 = xy1.data(); g.xy2 = xy2.data(); g.octave2 = oct2.data();
    g.u_right1 = uRight1.empty() ? NULL : uRight1.data();
    g.u_right2 = uRight2.empty() ? NULL : uRight2.data();
    g.nlevels = (int)scaleFactors2.size(); g.scale_factors2 = scaleFactors2.data(); g.level_sigma2_2 = levelSigma2_2.data();
    std::vector<int> vMatches12(keysUn1.size(), -1);
    int nmatches = 0;
    check(borb_voc_search_for_triangulation(voc, &a.rec, &b.rec, &g,
                                            (checkOrientation ? 0u : BORB_MATCH_NO_ORIENTATION) | (bOnlyStereo ? BORB_MATCH_ONLY_STEREO : 0u),
                                            vMatches12.data(), &nmatches),
          "search for triangulation");
    vMatchedPairs.clear();
    vMatchedPairs.reserve((size_t)std::max(nmatches, 0));
    for (size_t i = 0; i < vMatches12.size(); ++i) /* :950-956 */
        if (vMatches12[i] >= 0) vMatchedPairs.push_back(std::make_pair(i, (size_t)vMatches12[i]));
    return nmatches;
}

int ORBextractor::SearchByBoW(const cv::Mat &kfDescriptors, const std::vector<cv::KeyPoint> &kfKeysUn,
                              const unsigned char *kfMapPointFlags, const std::map<unsigned int, std::vector<unsigned int> > &kfFeatVec,
                              float nnRatio, bool checkOrientation, std::vector<int> &vnMatches)
{
    const int n = (int)kfKeysUn.size();
    assert(kfDescriptors.rows == n && (n == 0 || kfDescriptors.cols == 32));
    std::vector<unsigned char> desc((size_t)32 * n);
    std::vector<float> angle((size_t)n);
    for (int i = 0; i < n; ++i) {
        memcpy(&desc[(size_t)32 * i], kfDescriptors.ptr(i), 32);
        angle[i] = kfKeysUn[i].angle;
    }
    std::vector<uint32_t> nid, fidx;
    std::vector<int32_t> nstart(1, 0);
    for (std::map<unsigned int, std::vector<unsigned int> >::const_iterator it = kfFeatVec.begin(); it != kfFeatVec.end(); ++it) {
        nid.push_back(it->first);
        fidx.insert(fidx.end(), it->second.begin(), it->second.end());
        nstart.push_back((int32_t)fidx.size());
    }
    borb_keyframe kf;
    kf.n = n; kf.desc = desc.data(); kf.angle = angle.data(); kf.mp_flags = kfMapPointFlags;
    kf.n_nodes = (int)nid.size(); kf.node_id = nid.data(); kf.node_start = nstart.data(); kf.feat_idx = fidx.data();
    vnMatches.assign((size_t)borb_max_keypoints(handle), -1);
    int nc = 0, nmatches = 0;
    check(borb_search_by_bow(handle, 0, &kf, nnRatio, checkOrientation ? 0u : BORB_MATCH_NO_ORIENTATION, vnMatches.data(), &nc, &nmatches),
          "search by bow");
    vnMatches.resize((size_t)nc);
    return nmatches;
}

} // namespace ORB_SLAM2
