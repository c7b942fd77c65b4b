/*
 * tests/shim/frame_like.cc -- TEST: drives the drop-in ORB_SLAM2::ORBextractor exactly the way
 * the reference's callers do, to show that their source compiles and links unchanged:
 *   src/Tracking.cc:121-127  `new ORBextractor(nFeatures, fScaleFactor, nLevels, fIniThFAST, fMinThFAST)`
 *                            for left, right and the 2*nFeatures monocular-init extractor
 *   src/Frame.cc:101-107     the seven getters
 *   src/Frame.cc:110-113     left and right operator() on two std::threads
 *   src/Frame.cc:310-316     (*mpORBextractorLeft)(im, cv::Mat(), mvKeys, mDescriptors)
 *   src/Frame.cc:626,738-757 mvImagePyramid[level].rows/.cols/.rowRange().colRange()
 *   src/Frame.cc:120         ComputeStereoMatches() -- here the one-line device version
 *
 *   src/Tracking.cc:176-201 + Examples/Stereo/stereo_euroc.cc:136-137 -- `pre` mode: a raw RGB
 *                            frame goes in, rectification and gray conversion run on the device
 *
 * usage: frame_like pre <rgb.raw> <cols> <rows> <mapx.f32> <mapy.f32> <nfeatures> <out.bin>
 * usage: frame_like <left.raw> <right.raw> <cols> <rows> <nfeatures> <out.bin>
 * Writes, for left then right: int32 n, n x 28-byte keypoints, n x 32 descriptor bytes, then for
 * each of the 8 levels int32 cols, rows and an 11x11 window checksum (sum of all level pixels via
 * rowRange/colRange walks); at the very end int32 n, n x float mvuRight, n x float mvDepth of the
 * stereo pair, then int32 n, n x 2 float mvKeysUn positions, the 64x48 grid as int32 counts and
 * indices, and the four image bounds (camera: TUM1.yaml).  The pytest compares everything with the
 * CPU oracle.
 */
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <chrono>
#include <thread>
#include <map>
#include <vector>

#include "ORBextractor.h"

using namespace std;
using namespace ORB_SLAM2;

struct FrameLike {
    ORBextractor *mpORBextractorLeft, *mpORBextractorRight;
    vector<cv::KeyPoint> mvKeys, mvKeysRight;
    cv::Mat mDescriptors, mDescriptorsRight;
    int mnScaleLevels;
    float mfScaleFactor, mfLogScaleFactor;
    vector<float> mvScaleFactors, mvInvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
    vector<float> mvuRight, mvDepth;
    float mbf, mb;
    vector<cv::KeyPoint> mvKeysUn;
    std::vector<std::size_t> mGrid[64][48];
    float mnBounds[4];

    FrameLike(const cv::Mat &imLeft, const cv::Mat &imRight, ORBextractor *l, ORBextractor *r)
        : mpORBextractorLeft(l), mpORBextractorRight(r)
    {
        mnScaleLevels = mpORBextractorLeft->GetLevels();
        mfScaleFactor = mpORBextractorLeft->GetScaleFactor();
        mfLogScaleFactor = log(mfScaleFactor);
        mvScaleFactors = mpORBextractorLeft->GetScaleFactors();
        mvInvScaleFactors = mpORBextractorLeft->GetInverseScaleFactors();
        mvLevelSigma2 = mpORBextractorLeft->GetScaleSigmaSquares();
        mvInvLevelSigma2 = mpORBextractorLeft->GetInverseScaleSigmaSquares();

        thread threadLeft(&FrameLike::ExtractORB, this, 0, imLeft);
        thread threadRight(&FrameLike::ExtractORB, this, 1, imRight);
        threadLeft.join();
        threadRight.join();

        /* src/Frame.cc:86-87 (mb is derived from mbf and fx), :120 ComputeStereoMatches() */
        mbf = 386.1448f;
        mb = mbf / 718.856f;
        ComputeStereoMatches();

        /* src/Frame.cc:116 UndistortKeyPoints(), :132 ComputeImageBounds(), :150 AssignFeaturesToGrid() */
        mpORBextractorLeft->FillFrame(mvKeys, mvKeysUn, mGrid, mnBounds);
    }

    void ComputeStereoMatches()
    {
        mpORBextractorLeft->ComputeStereoMatches(*mpORBextractorRight, mbf, mb, mvuRight, mvDepth);
    }

    void ExtractORB(int flag, const cv::Mat &im)
    {
        if (flag == 0)
            (*mpORBextractorLeft)(im, cv::Mat(), mvKeys, mDescriptors);
        else
            (*mpORBextractorRight)(im, cv::Mat(), mvKeysRight, mDescriptorsRight);
    }
};

static vector<unsigned char> slurp(const char *path, size_t n)
{
    vector<unsigned char> b(n);
    FILE *f = fopen(path, "rb");
    if (!f || fread(b.data(), 1, n, f) != n) { fprintf(stderr, "cannot read %s\n", path); exit(2); }
    fclose(f);
    return b;
}

static void dump(FILE *f, const vector<cv::KeyPoint> &k, const cv::Mat &d, ORBextractor *e)
{
    static_assert(sizeof(cv::KeyPoint) == 28, "cv::KeyPoint layout");
    int n = (int)k.size();
    fwrite(&n, 4, 1, f);
    fwrite(k.data(), 28, (size_t)n, f);
    for (int i = 0; i < n; ++i) fwrite(d.ptr(i), 1, 32, f);
    for (int level = 0; level < e->GetLevels(); ++level) {
        const cv::Mat &L = e->mvImagePyramid[level];
        int cols = L.cols, rows = L.rows;
        long long sum = 0;
        const int w = 5;
        for (int y = w; y + w + 1 <= rows; y += 2 * w + 1)
            for (int x = w; x + w + 1 <= cols; x += 2 * w + 1) {
                cv::Mat I = L.rowRange(y - w, y + w + 1).colRange(x - w, x + w + 1);
                for (int r = 0; r < I.rows; ++r)
                    for (int c = 0; c < I.cols; ++c) sum += I.at<uchar>(r, c) * (long long)(1 + ((r * 11 + c) % 7));
            }
        fwrite(&cols, 4, 1, f);
        fwrite(&rows, 4, 1, f);
        fwrite(&sum, 8, 1, f);
    }
}

static int pre_mode(char **argv)
{
    const int cols = atoi(argv[3]), rows = atoi(argv[4]), nFeatures = atoi(argv[7]);
    vector<unsigned char> rgb = slurp(argv[2], (size_t)cols * rows * 3);
    vector<unsigned char> mx = slurp(argv[5], (size_t)cols * rows * 4), my = slurp(argv[6], (size_t)cols * rows * 4);
    cv::Mat im(rows, cols, CV_8UC3, rgb.data());
    cv::Mat M1(rows, cols, CV_32FC1, mx.data()), M2(rows, cols, CV_32FC1, my.data());
    ORBextractor ext(nFeatures, 1.2f, 8, 20, 7);
    ext.SetColorInput(true, /*rgb=*/true); /* Camera.RGB: 1 */
    ext.SetRectifyMap(M1, M2);
    vector<cv::KeyPoint> k;
    cv::Mat d;
    ext(im, cv::Mat(), k, d);
    FILE *f = fopen(argv[8], "wb");
    if (!f) return 2;
    dump(f, k, d, &ext);
    fclose(f);
    printf("frame_like pre ok\n");
    return 0;
}

/* `match` mode: the tracking thread's frame-to-frame matcher through the shim.  argv: match left right cols
 * rows nfeatures last.bin out.bin; last.bin = int32 n, float K[4], float mbf, mb, th, int32 mono, float
 * TcwCur[16], TcwLast[16], then xyz[3n], desc[32n], flags[n], octave[n], angle[n] (the arrays of
 * include/borb.h: borb_last_frame).  The current frame is a stereo frame built the way Frame's constructor
 * does (extraction on two threads, device stereo association), then
 * ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (reference src/ORBmatcher.cc:1503-1668)
 * runs on the device.  out.bin = int32 nmatches, n_cur, then n_cur int32 matches. */
static int match_mode(char **argv)
{
    const int cols = atoi(argv[4]), rows = atoi(argv[5]), nFeatures = atoi(argv[6]);
    vector<unsigned char> l = slurp(argv[2], (size_t)cols * rows), r = slurp(argv[3], (size_t)cols * rows);
    cv::Mat imLeft(rows, cols, CV_8UC1, l.data()), imRight(rows, cols, CV_8UC1, r.data());
    FILE *in = fopen(argv[7], "rb");
    if (!in) return 2;
    int n = 0, mono = 0;
    float K4[4], mbf, mb, th, tc[16], tl[16];
    if (fread(&n, 4, 1, in) != 1 || fread(K4, 4, 4, in) != 4 || fread(&mbf, 4, 1, in) != 1 || fread(&mb, 4, 1, in) != 1 ||
        fread(&th, 4, 1, in) != 1 || fread(&mono, 4, 1, in) != 1 || fread(tc, 4, 16, in) != 16 || fread(tl, 4, 16, in) != 16)
        return 2;
    vector<float> xyz((size_t)3 * n), angle((size_t)n);
    vector<unsigned char> desc((size_t)32 * n), flags((size_t)n);
    vector<int> octave((size_t)n);
    if (fread(xyz.data(), 4, xyz.size(), in) != xyz.size() || fread(desc.data(), 1, desc.size(), in) != desc.size() ||
        fread(flags.data(), 1, flags.size(), in) != flags.size() || fread(octave.data(), 4, octave.size(), in) != octave.size() ||
        fread(angle.data(), 4, angle.size(), in) != angle.size())
        return 2;
    fclose(in);
    cv::Mat K(3, 3, CV_32F), DistCoef(4, 1, CV_32F), TcwCur(4, 4, CV_32F), TcwLast(4, 4, CV_32F);
    memset(K.data, 0, 36);
    memset(DistCoef.data, 0, 16);
    K.at<float>(0, 0) = K4[0]; K.at<float>(1, 1) = K4[1]; K.at<float>(0, 2) = K4[2]; K.at<float>(1, 2) = K4[3]; K.at<float>(2, 2) = 1.f;
    memcpy(TcwCur.data, tc, 64);
    memcpy(TcwLast.data, tl, 64);
    ORBextractor left(nFeatures, 1.2f, 8, 20, 7), right(nFeatures, 1.2f, 8, 20, 7);
    left.SetCamera(K, DistCoef);
    FrameLike fr(imLeft, imRight, &left, &right); /* KITTI mbf / mb inside, as the other mode */
    vector<int> vnMatches;
    const int nmatches = left.SearchByProjection(TcwCur, TcwLast, n, xyz.data(), desc.data(), flags.data(), octave.data(),
                                                 angle.data(), mbf, mb, th, mono != 0, true, true, vnMatches);
    if (getenv("FRAME_LIKE_TIME")) { /* C++ call latency of the matcher (bench note in DESIGN.md 4e) */
        vector<double> us;
        vector<int> tmp;
        for (int it = 0; it < 220; ++it) {
            const auto t0 = std::chrono::steady_clock::now();
            left.SearchByProjection(TcwCur, TcwLast, n, xyz.data(), desc.data(), flags.data(), octave.data(), angle.data(), mbf,
                                    mb, th, mono != 0, true, true, tmp);
            const auto t1 = std::chrono::steady_clock::now();
            if (it >= 20) us.push_back(std::chrono::duration<double, std::micro>(t1 - t0).count());
        }
        std::sort(us.begin(), us.end());
        printf("SearchByProjection through the shim: p50 %.1f us, p99 %.1f us\n", us[us.size() / 2], us[us.size() * 99 / 100]);
    }
    FILE *f = fopen(argv[8], "wb");
    if (!f) return 2;
    int nc = (int)vnMatches.size();
    fwrite(&nmatches, 4, 1, f);
    fwrite(&nc, 4, 1, f);
    fwrite(vnMatches.data(), 4, vnMatches.size(), f);
    /* then the local-map matcher on the same frame (reference src/ORBmatcher.cc:59-154): the map points are the last
     * frame's again, projected the way the motion-model matcher just did (what Frame::isInFrustum would leave), viewing
     * cosine 1, predicted level = the octave they were seen at; the slots the first matcher filled count as held */
    {
        vector<float> proj((size_t)4 * n);
        vector<unsigned char> lflags((size_t)n), occupied((size_t)nc);
        for (int i2 = 0; i2 < nc; ++i2) occupied[i2] = vnMatches[i2] >= 0 ? 2 : 0;
        for (int i = 0; i < n; ++i) {
            const float *X = &xyz[(size_t)3 * i];
            const double xc = (double)tc[0] * X[0] + (double)tc[1] * X[1] + (double)tc[2] * X[2] + tc[3];
            const double yc = (double)tc[4] * X[0] + (double)tc[5] * X[1] + (double)tc[6] * X[2] + tc[7];
            const double zc = (double)tc[8] * X[0] + (double)tc[9] * X[1] + (double)tc[10] * X[2] + tc[11];
            const bool front = (flags[i] & 1) && zc > 0.1;
            const float u = front ? (float)(K4[0] * xc / zc + K4[2]) : 0.f, v = front ? (float)(K4[1] * yc / zc + K4[3]) : 0.f;
            proj[4 * i] = u; proj[4 * i + 1] = v; proj[4 * i + 2] = front ? (float)(u - mbf / zc) : 0.f; proj[4 * i + 3] = 1.f;
            lflags[i] = (unsigned char)((front && u > 0.f && u < (float)cols && v > 0.f && v < (float)rows ? 1 : 0) | (flags[i] & 2));
        }
        vector<int> vnLocal;
        const int nlocal = left.SearchLocalPoints(n, proj.data(), desc.data(), lflags.data(), octave.data(), occupied.data(), 3.f,
                                                  0.8f, true, vnLocal);
        int nl = (int)vnLocal.size();
        fwrite(&nlocal, 4, 1, f);
        fwrite(&nl, 4, 1, f);
        fwrite(vnLocal.data(), 4, vnLocal.size(), f);
        fwrite(proj.data(), 4, proj.size(), f);
        fwrite(lflags.data(), 1, lflags.size(), f);
    }
    fclose(f);
    printf("frame_like match ok\n");
    return 0;
}

/* `reloc` mode: the relocalisation matcher through the shim (reference src/ORBmatcher.cc:1670-1798).  argv: reloc cur.raw
 * cols rows nfeatures kf.bin out.bin; kf.bin = int32 n, n_occ, float K[4], th, int32 ORBdist, float TcwCur[16], then
 * xyz[3n], desc[32n], flags[n], dist[2n], angle[n], occupied[n_occ] (the arrays of include/borb.h: borb_keyframe_points).
 * out.bin = int32 nmatches, n_cur, then n_cur int32 matches. */
static int reloc_mode(char **argv)
{
    const int cols = atoi(argv[3]), rows = atoi(argv[4]), nFeatures = atoi(argv[5]);
    vector<unsigned char> im = slurp(argv[2], (size_t)cols * rows);
    cv::Mat image(rows, cols, CV_8UC1, im.data());
    FILE *in = fopen(argv[6], "rb");
    if (!in) return 2;
    int n = 0, nocc = 0, orbDist = 0;
    float K4[4], th, tc[16];
    if (fread(&n, 4, 1, in) != 1 || fread(&nocc, 4, 1, in) != 1 || fread(K4, 4, 4, in) != 4 || fread(&th, 4, 1, in) != 1 ||
        fread(&orbDist, 4, 1, in) != 1 || fread(tc, 4, 16, in) != 16)
        return 2;
    vector<float> xyz((size_t)3 * n), dist((size_t)2 * n), angle((size_t)n);
    vector<unsigned char> desc((size_t)32 * n), flags((size_t)n), occupied((size_t)nocc);
    if (fread(xyz.data(), 4, xyz.size(), in) != xyz.size() || fread(desc.data(), 1, desc.size(), in) != desc.size() ||
        fread(flags.data(), 1, flags.size(), in) != flags.size() || fread(dist.data(), 4, dist.size(), in) != dist.size() ||
        fread(angle.data(), 4, angle.size(), in) != angle.size() || fread(occupied.data(), 1, occupied.size(), in) != occupied.size())
        return 2;
    fclose(in);
    cv::Mat K(3, 3, CV_32F), DistCoef(4, 1, CV_32F), TcwCur(4, 4, CV_32F);
    memset(K.data, 0, 36);
    memset(DistCoef.data, 0, 16);
    K.at<float>(0, 0) = K4[0]; K.at<float>(1, 1) = K4[1]; K.at<float>(0, 2) = K4[2]; K.at<float>(1, 2) = K4[3]; K.at<float>(2, 2) = 1.f;
    memcpy(TcwCur.data, tc, 64);
    ORBextractor ext(nFeatures, 1.2f, 8, 20, 7);
    ext.SetCamera(K, DistCoef);
    vector<cv::KeyPoint> keys;
    cv::Mat descriptors;
    ext(image, cv::Mat(), keys, descriptors);
    if ((int)keys.size() != nocc) return 3;
    vector<int> vnMatches;
    const int nmatches = ext.SearchByProjectionKF(TcwCur, n, xyz.data(), desc.data(), flags.data(), dist.data(), angle.data(),
                                                  occupied.data(), th, orbDist, true, vnMatches);
    FILE *f = fopen(argv[7], "wb");
    if (!f) return 2;
    int nc = (int)vnMatches.size();
    fwrite(&nmatches, 4, 1, f);
    fwrite(&nc, 4, 1, f);
    fwrite(vnMatches.data(), 4, vnMatches.size(), f);
    fclose(f);
    printf("frame_like reloc ok\n");
    return 0;
}

/* `init` mode: ORBmatcher::SearchForInitialization through the shim (reference src/ORBmatcher.cc:487-632).  argv: init
 * image cols rows nfeatures f1.bin out.bin; f1.bin = int32 n, window, float nnratio, then octave[n] (int32), angle[n],
 * desc[32n], prev[2n].  The image is the CURRENT frame (the mpIniORBextractor instance, TUM1 camera); out.bin = int32
 * nmatches, n, then n int32 vnMatches12 and 2n float vbPrevMatched. */
static int init_mode(char **argv)
{
    const int cols = atoi(argv[3]), rows = atoi(argv[4]), nFeatures = atoi(argv[5]);
    vector<unsigned char> im = slurp(argv[2], (size_t)cols * rows);
    cv::Mat image(rows, cols, CV_8UC1, im.data());
    FILE *in = fopen(argv[6], "rb");
    if (!in) return 2;
    int n = 0, window = 0;
    float ratio = 0.f;
    if (fread(&n, 4, 1, in) != 1 || fread(&window, 4, 1, in) != 1 || fread(&ratio, 4, 1, in) != 1) return 2;
    vector<int> octave((size_t)n);
    vector<float> angle((size_t)n), prev((size_t)2 * n);
    cv::Mat desc(n, 32, CV_8UC1);
    if (fread(octave.data(), 4, (size_t)n, in) != (size_t)n || fread(angle.data(), 4, (size_t)n, in) != (size_t)n) return 2;
    for (int i = 0; i < n; ++i) if (fread(desc.ptr(i), 1, 32, in) != 32) return 2;
    if (fread(prev.data(), 4, prev.size(), in) != prev.size()) return 2;
    fclose(in);
    cv::Mat K(3, 3, CV_32F), DistCoef(5, 1, CV_32F);
    memset(K.data, 0, 36);
    K.at<float>(0, 0) = 517.306408f; K.at<float>(1, 1) = 516.469215f;
    K.at<float>(0, 2) = 318.643040f; K.at<float>(1, 2) = 255.313989f; K.at<float>(2, 2) = 1.f;
    DistCoef.at<float>(0, 0) = 0.262383f; DistCoef.at<float>(1, 0) = -0.953104f; DistCoef.at<float>(2, 0) = -0.005358f;
    DistCoef.at<float>(3, 0) = 0.002628f; DistCoef.at<float>(4, 0) = 1.163314f;
    ORBextractor ini(nFeatures, 1.2f, 8, 20, 7);
    ini.SetCamera(K, DistCoef);
    vector<cv::KeyPoint> k2;
    cv::Mat d2;
    ini(image, cv::Mat(), k2, d2);
    vector<cv::KeyPoint> F1KeysUn((size_t)n);
    vector<cv::Point2f> vbPrevMatched((size_t)n);
    for (int i = 0; i < n; ++i) {
        F1KeysUn[i].octave = octave[i];
        F1KeysUn[i].angle = angle[i];
        vbPrevMatched[i] = cv::Point2f(prev[2 * i], prev[2 * i + 1]);
    }
    vector<int> vnMatches12;
    const int nmatches = ini.SearchForInitialization(F1KeysUn, desc, vbPrevMatched, vnMatches12, window, ratio, true);
    FILE *f = fopen(argv[7], "wb");
    if (!f) return 2;
    fwrite(&nmatches, 4, 1, f);
    fwrite(&n, 4, 1, f);
    fwrite(vnMatches12.data(), 4, vnMatches12.size(), f);
    for (int i = 0; i < n; ++i) { fwrite(&vbPrevMatched[i].x, 4, 1, f); fwrite(&vbPrevMatched[i].y, 4, 1, f); }
    fclose(f);
    printf("frame_like init ok\n");
    return 0;
}

/* `bow` mode: Frame::ComputeBoW and ORBmatcher::SearchByBoW(KeyFrame*, Frame&) through the shim (reference
 * src/Frame.cc:532-539, src/ORBmatcher.cc:195-351).  argv: bow image cols rows nfeatures vocabulary.txt kf.bin out.bin;
 * kf.bin = int32 n, float nnratio, then angle[n], desc[32n], flags[n] of the keyframe.  out.bin = int32 nmatches, nF,
 * words, nodes, entries; vnMatches[nF]; word ids [words] (u32), values [words] (f64); node ids [nodes], node sizes [nodes],
 * feature indices [entries] of the FRAME's mBowVec / mFeatVec; then the same five arrays of the keyframe's. */
static void dump_bow(FILE *f, const std::map<unsigned int, double> &bow, const std::map<unsigned int, std::vector<unsigned int> > &fv)
{
    int words = (int)bow.size(), nodes = (int)fv.size(), entries = 0;
    for (std::map<unsigned int, std::vector<unsigned int> >::const_iterator it = fv.begin(); it != fv.end(); ++it) entries += (int)it->second.size();
    fwrite(&words, 4, 1, f); fwrite(&nodes, 4, 1, f); fwrite(&entries, 4, 1, f);
    for (std::map<unsigned int, double>::const_iterator it = bow.begin(); it != bow.end(); ++it) fwrite(&it->first, 4, 1, f);
    for (std::map<unsigned int, double>::const_iterator it = bow.begin(); it != bow.end(); ++it) fwrite(&it->second, 8, 1, f);
    for (std::map<unsigned int, std::vector<unsigned int> >::const_iterator it = fv.begin(); it != fv.end(); ++it) fwrite(&it->first, 4, 1, f);
    for (std::map<unsigned int, std::vector<unsigned int> >::const_iterator it = fv.begin(); it != fv.end(); ++it) { int c = (int)it->second.size(); fwrite(&c, 4, 1, f); }
    for (std::map<unsigned int, std::vector<unsigned int> >::const_iterator it = fv.begin(); it != fv.end(); ++it) fwrite(it->second.data(), 4, it->second.size(), f);
}

static int bow_mode(char **argv)
{
    const int cols = atoi(argv[3]), rows = atoi(argv[4]), nFeatures = atoi(argv[5]);
    vector<unsigned char> im = slurp(argv[2], (size_t)cols * rows);
    cv::Mat image(rows, cols, CV_8UC1, im.data());
    FILE *in = fopen(argv[7], "rb");
    if (!in) return 2;
    int n = 0;
    float ratio = 0.f;
    if (fread(&n, 4, 1, in) != 1 || fread(&ratio, 4, 1, in) != 1) return 2;
    vector<cv::KeyPoint> kfKeysUn((size_t)n);
    vector<float> angle((size_t)n);
    vector<unsigned char> flags((size_t)n);
    cv::Mat kfDesc(n, 32, CV_8UC1);
    if (fread(angle.data(), 4, (size_t)n, in) != (size_t)n) return 2;
    for (int i = 0; i < n; ++i) { if (fread(kfDesc.ptr(i), 1, 32, in) != 32) return 2; kfKeysUn[i].angle = angle[i]; }
    if (fread(flags.data(), 1, (size_t)n, in) != (size_t)n) return 2;
    fclose(in);
    DeviceVocabulary voc(argv[6]);                       /* next to System's loadFromTextFile */
    std::map<unsigned int, double> kfBow, mBowVec;       /* DBoW2::BowVector / FeatureVector derive from these */
    std::map<unsigned int, std::vector<unsigned int> > kfFeatVec, mFeatVec;
    voc.transform(kfDesc, kfBow, kfFeatVec, 4);          /* KeyFrame::ComputeBoW */
    ORBextractor ext(nFeatures, 1.2f, 8, 20, 7);
    vector<cv::KeyPoint> keys;
    cv::Mat desc;
    ext(image, cv::Mat(), keys, desc);
    ext.ComputeBoW(voc, mBowVec, mFeatVec);              /* Frame::ComputeBoW */
    vector<int> vnMatches;
    const int nmatches = ext.SearchByBoW(kfDesc, kfKeysUn, flags.data(), kfFeatVec, ratio, true, vnMatches);
    FILE *f = fopen(argv[8], "wb");
    if (!f) return 2;
    const int nF = (int)vnMatches.size();
    fwrite(&nmatches, 4, 1, f);
    fwrite(&nF, 4, 1, f);
    fwrite(vnMatches.data(), 4, vnMatches.size(), f);
    dump_bow(f, mBowVec, mFeatVec);
    dump_bow(f, kfBow, kfFeatVec);
    /* ORBmatcher::SearchByBoW(pKF1, pKF2, vpMatches12) (reference src/ORBmatcher.cc:634-777): the keyframe against the
     * current frame promoted to a keyframe -- every keypoint but each fourth carries a map point, each 16th a bad one */
    {
        vector<unsigned char> flags2(keys.size());
        for (size_t i = 0; i < keys.size(); ++i) flags2[i] = (unsigned char)((i % 4 ? 1 : 0) | (i % 16 == 5 ? 2 : 0));
        vector<int> vnMatches12;
        const int nm12 = voc.SearchByBoW(kfDesc, kfKeysUn, flags.data(), kfFeatVec, desc, keys, flags2.data(), mFeatVec, 0.75f, true,
                                         vnMatches12);
        const int n1 = (int)vnMatches12.size();
        fwrite(&nm12, 4, 1, f);
        fwrite(&n1, 4, 1, f);
        fwrite(vnMatches12.data(), 4, vnMatches12.size(), f);
    }
    fclose(f);
    printf("frame_like bow ok\n");
    return 0;
}

int main(int argc, char **argv)
{
    if (argc == 9 && !strcmp(argv[1], "bow")) return bow_mode(argv);
    if (argc == 8 && !strcmp(argv[1], "init")) return init_mode(argv);
    if (argc == 8 && !strcmp(argv[1], "reloc")) return reloc_mode(argv);
    if (argc == 9 && !strcmp(argv[1], "match")) return match_mode(argv);
    if (argc == 9 && !strcmp(argv[1], "pre")) return pre_mode(argv);
    if (argc != 7) { fprintf(stderr, "usage: %s left right cols rows nfeatures out\n", argv[0]); return 2; }
    const int cols = atoi(argv[3]), rows = atoi(argv[4]), nFeatures = atoi(argv[5]);
    vector<unsigned char> l = slurp(argv[1], (size_t)cols * rows), r = slurp(argv[2], (size_t)cols * rows);
    cv::Mat imLeft(rows, cols, CV_8UC1, l.data()), imRight(rows, cols, CV_8UC1, r.data());

    const float fScaleFactor = 1.2f;
    const int nLevels = 8, fIniThFAST = 20, fMinThFAST = 7;
    ORBextractor *mpORBextractorLeft = new ORBextractor(nFeatures, fScaleFactor, nLevels, fIniThFAST, fMinThFAST);
    ORBextractor *mpORBextractorRight = new ORBextractor(nFeatures, fScaleFactor, nLevels, fIniThFAST, fMinThFAST);
    ORBextractor *mpIniORBextractor = new ORBextractor(2 * nFeatures, fScaleFactor, nLevels, fIniThFAST, fMinThFAST);

    /* mK and mDistCoef as src/Tracking.cc:56-80 builds them (TUM1.yaml values) */
    cv::Mat K(3, 3, CV_32F), DistCoef(5, 1, CV_32F);
    memset(K.data, 0, 36);
    K.at<float>(0, 0) = 517.306408f; K.at<float>(1, 1) = 516.469215f;
    K.at<float>(0, 2) = 318.643040f; K.at<float>(1, 2) = 255.313989f; K.at<float>(2, 2) = 1.f;
    DistCoef.at<float>(0, 0) = 0.262383f; DistCoef.at<float>(1, 0) = -0.953104f; DistCoef.at<float>(2, 0) = -0.005358f;
    DistCoef.at<float>(3, 0) = 0.002628f; DistCoef.at<float>(4, 0) = 1.163314f;
    mpORBextractorLeft->SetCamera(K, DistCoef);

    FILE *f = fopen(argv[6], "wb");
    if (!f) return 2;
    vector<float> uRight, depth, unxy, bnd;
    vector<int> gcount, gidx;
    for (int rep = 0; rep < 3; ++rep) { /* instances are reused frame after frame */
        FrameLike fr(imLeft, imRight, mpORBextractorLeft, mpORBextractorRight);
        if (rep == 2) {
            dump(f, fr.mvKeys, fr.mDescriptors, mpORBextractorLeft);
            dump(f, fr.mvKeysRight, fr.mDescriptorsRight, mpORBextractorRight);
            uRight = fr.mvuRight;
            depth = fr.mvDepth;
            for (size_t i = 0; i < fr.mvKeysUn.size(); ++i) { unxy.push_back(fr.mvKeysUn[i].pt.x); unxy.push_back(fr.mvKeysUn[i].pt.y); }
            for (int i = 0; i < 64; ++i)
                for (int j = 0; j < 48; ++j) {
                    gcount.push_back((int)fr.mGrid[i][j].size());
                    for (size_t q = 0; q < fr.mGrid[i][j].size(); ++q) gidx.push_back((int)fr.mGrid[i][j][q]);
                }
            bnd.assign(fr.mnBounds, fr.mnBounds + 4);
        }
    }
    /* monocular initialisation extractor + the empty-image convention */
    vector<cv::KeyPoint> k;
    cv::Mat d;
    (*mpIniORBextractor)(imLeft, cv::Mat(), k, d);
    dump(f, k, d, mpIniORBextractor);
    const size_t before = k.size();
    (*mpIniORBextractor)(cv::Mat(), cv::Mat(), k, d); /* returns without touching the outputs */
    int untouched = (k.size() == before && d.rows == (int)before);
    fwrite(&untouched, 4, 1, f);
    int ns = (int)uRight.size();
    fwrite(&ns, 4, 1, f);
    fwrite(uRight.data(), 4, (size_t)ns, f);
    fwrite(depth.data(), 4, (size_t)ns, f);
    int nu = (int)unxy.size() / 2, ng = (int)gidx.size();
    fwrite(&nu, 4, 1, f);
    fwrite(unxy.data(), 4, unxy.size(), f);
    fwrite(gcount.data(), 4, gcount.size(), f);
    fwrite(&ng, 4, 1, f);
    fwrite(gidx.data(), 4, gidx.size(), f);
    fwrite(bnd.data(), 4, 4, f);
    fclose(f);
    delete mpORBextractorLeft;
    delete mpORBextractorRight;
    delete mpIniORBextractor;
    printf("frame_like ok\n");
    return 0;
}
