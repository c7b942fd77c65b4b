/*
 * oracle/bow_driver.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * C entry points over the reference's OWN bag-of-words code, compiled where it lies (oracle/Makefile:
 * _ref/librefbow.so):
 *   DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB> (= ORB_SLAM2::ORBVocabulary, include/ORBVocabulary.h)
 *       loadFromTextFile   Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1338-1421
 *       transform          :1127-1206 (bow + feature vector) and :1218-1270 (descent of one descriptor)
 *       FORB::distance     Thirdparty/DBoW2/DBoW2/FORB.cpp:81-101
 *   Frame::ComputeBoW      src/Frame.cc:532-539
 *   ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&)   src/ORBmatcher.cc:195-351
 * Frames and keyframes are the reference's classes (keyframe / map point: the plain records of
 * oracle/matchcompat) filled from flat arrays; every call below is the reference's object code.
 * The vocabulary blob (Vocabulary/ORBvoc.txt) is NOT in the reference tree: vocabularies here are synthetic files
 * in its text format (tests/golden/make_vocabulary.py).
 */
#include <cstdint>
#include <cstring>
#include <map>
#include <vector>

#include "Frame.h"
#include "KeyFrame.h"
#include "MapPoint.h"
#include "ORBVocabulary.h"
#include "ORBmatcher.h"
#include "matchcompat/MapPoint_impl.h"

using namespace ORB_SLAM2;

namespace {

cv::Mat desc_mat(const uint8_t *desc, int n)
{
    cv::Mat m(n, 32, CV_8UC1);
    for (int i = 0; i < n; ++i) std::memcpy(m.ptr(i), desc + (size_t)i * 32, 32);
    return m;
}

void flatten(const DBoW2::BowVector &bv, const DBoW2::FeatureVector &fv, uint32_t *word_id, double *word_val,
             int *n_words, uint32_t *node_id, int32_t *node_start, uint32_t *feat_idx, int *n_nodes)
{
    int w = 0;
    for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it, ++w) {
        word_id[w] = it->first;
        word_val[w] = it->second;
    }
    *n_words = w;
    int nn = 0, pos = 0;
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++nn) {
        node_id[nn] = it->first;
        node_start[nn] = pos;
        for (size_t j = 0; j < it->second.size(); ++j) feat_idx[pos++] = it->second[j];
    }
    node_start[nn] = pos;
    *n_nodes = nn;
}

} // namespace

extern "C" {

void *refbow_load(const char *path)
{
    ORBVocabulary *voc = new ORBVocabulary();
    if (!voc->loadFromTextFile(path) || voc->empty()) {
        delete voc;
        return nullptr;
    }
    return voc;
}

void refbow_free(void *v) { delete static_cast<ORBVocabulary *>(v); }

void refbow_info(void *v, int *k, int *L, int *n_words, int *scoring, int *weighting)
{
    ORBVocabulary *voc = static_cast<ORBVocabulary *>(v);
    *k = voc->getBranchingFactor();
    *L = voc->getDepthLevels();
    *n_words = (int)voc->size();
    *scoring = (int)voc->getScoringType();
    *weighting = (int)voc->getWeightingType();
}

/* Frame::ComputeBoW (levelsup == 4, the reference's constant) on a default-constructed reference Frame, or the
 * same transform() call with another levelsup.  Capacities: word_* [n], node_id [n], node_start [n + 1], feat_idx [n]. */
void refbow_compute_bow(void *v, const uint8_t *desc, int n, int levelsup, uint32_t *word_id, double *word_val,
                        int *n_words, uint32_t *node_id, int32_t *node_start, uint32_t *feat_idx, int *n_nodes)
{
    ORBVocabulary *voc = static_cast<ORBVocabulary *>(v);
    Frame F;
    F.N = n;
    F.mpORBvocabulary = voc;
    F.mDescriptors = desc_mat(desc, n);
    if (levelsup == 4) F.ComputeBoW();
    else {
        std::vector<cv::Mat> vd;
        for (int i = 0; i < n; ++i) vd.push_back(F.mDescriptors.row(i));
        voc->transform(vd, F.mBowVec, F.mFeatVec, levelsup);
    }
    flatten(F.mBowVec, F.mFeatVec, word_id, word_val, n_words, node_id, node_start, feat_idx, n_nodes);
}

/* L1 score of two bow vectors through the vocabulary's scoring object (TemplatedVocabulary::score :1208-1213) */
double refbow_score(void *v, const uint32_t *id1, const double *val1, int n1, const uint32_t *id2, const double *val2, int n2)
{
    ORBVocabulary *voc = static_cast<ORBVocabulary *>(v);
    DBoW2::BowVector a, b;
    for (int i = 0; i < n1; ++i) a.insert(a.end(), DBoW2::BowVector::value_type(id1[i], val1[i]));
    for (int i = 0; i < n2; ++i) b.insert(b.end(), DBoW2::BowVector::value_type(id2[i], val2[i]));
    return voc->score(a, b);
}

/* ORBmatcher::SearchByBoW(pKF, F, vpMapPointMatches).  Keyframe: n_kf keypoints (28-B cv::KeyPoint records, used
 * as mvKeysUn: only the angle is read), descriptors, kf_mp[i] bit 0 = has a map point, bit 1 = that point isBad();
 * its feature vector is computed here with the reference transform (KeyFrame::ComputeBoW, src/KeyFrame.cc, is the
 * same call as Frame::ComputeBoW).  out_match[iF] = index of the keyframe keypoint whose map point F keypoint iF
 * ends up with, or -1.  Returns nmatches. */
int refbow_search_by_bow(void *v, int n_kf, const uint8_t *kf_kps, const uint8_t *kf_desc, const uint8_t *kf_mp,
                         int n_f, const uint8_t *f_kps, const uint8_t *f_desc, float nnratio, int check_ori,
                         int32_t *out_match)
{
    ORBVocabulary *voc = static_cast<ORBVocabulary *>(v);
    KeyFrame KF;
    std::vector<MapPoint> mps((size_t)std::max(n_kf, 1));
    KF.N = n_kf;
    KF.mvKeysUn.resize(n_kf);
    if (n_kf) std::memcpy(KF.mvKeysUn.data(), kf_kps, (size_t)n_kf * 28);
    KF.mvKeys = KF.mvKeysUn;
    KF.mDescriptors = desc_mat(kf_desc, n_kf);
    KF.mvpMapPoints.assign(n_kf, static_cast<MapPoint *>(nullptr));
    for (int i = 0; i < n_kf; ++i)
        if (kf_mp[i] & 1) {
            mps[i].mbBad = (kf_mp[i] & 2) != 0;
            KF.mvpMapPoints[i] = &mps[i];
        }
    {
        std::vector<cv::Mat> vd;
        for (int i = 0; i < n_kf; ++i) vd.push_back(KF.mDescriptors.row(i));
        DBoW2::BowVector bv;
        voc->transform(vd, bv, KF.mFeatVec, 4);
    }
    Frame F;
    F.N = n_f;
    F.mpORBvocabulary = voc;
    F.mvKeys.resize(n_f);
    if (n_f) std::memcpy(F.mvKeys.data(), f_kps, (size_t)n_f * 28);
    F.mvKeysUn = F.mvKeys;
    F.mDescriptors = desc_mat(f_desc, n_f);
    F.ComputeBoW();

    ORBmatcher matcher(nnratio, check_ori != 0);
    std::vector<MapPoint *> matches;
    const int nm = matcher.SearchByBoW(&KF, F, matches);
    for (int i = 0; i < n_f; ++i) out_match[i] = matches[i] ? (int32_t)(matches[i] - mps.data()) : -1;
    return nm;
}

/* ORBmatcher::SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, vector<MapPoint*> &vpMatches12) (src/ORBmatcher.cc:634-777, the
 * matcher of LoopClosing::ComputeSim3).  Both keyframes as refbow_search_by_bow's keyframe side; their feature vectors are
 * computed here with the reference transform.  out_match12[i1] = index of the pKF2 keypoint whose map point pKF1's keypoint
 * i1 is matched with, or -1.  Returns nmatches. */
int refbow_search_by_bow_kf(void *v, int n1, const uint8_t *kps1, const uint8_t *desc1, const uint8_t *mp1,
                            int n2, const uint8_t *kps2, const uint8_t *desc2, const uint8_t *mp2, float nnratio,
                            int check_ori, int32_t *out_match12)
{
    ORBVocabulary *voc = static_cast<ORBVocabulary *>(v);
    KeyFrame KF[2];
    const int n[2] = {n1, n2};
    const uint8_t *kps[2] = {kps1, kps2}, *desc[2] = {desc1, desc2}, *mp[2] = {mp1, mp2};
    std::vector<MapPoint> mps[2];
    for (int s = 0; s < 2; ++s) {
        KeyFrame &K = KF[s];
        mps[s].resize((size_t)std::max(n[s], 1));
        K.N = n[s];
        K.mvKeysUn.resize(n[s]);
        if (n[s]) std::memcpy(K.mvKeysUn.data(), kps[s], (size_t)n[s] * 28);
        K.mvKeys = K.mvKeysUn;
        K.mDescriptors = desc_mat(desc[s], n[s]);
        K.mvpMapPoints.assign(n[s], static_cast<MapPoint *>(nullptr));
        for (int i = 0; i < n[s]; ++i)
            if (mp[s][i] & 1) {
                mps[s][i].mbBad = (mp[s][i] & 2) != 0;
                K.mvpMapPoints[i] = &mps[s][i];
            }
        std::vector<cv::Mat> vd;
        for (int i = 0; i < n[s]; ++i) vd.push_back(K.mDescriptors.row(i));
        DBoW2::BowVector bv;
        voc->transform(vd, bv, K.mFeatVec, 4);
    }
    ORBmatcher matcher(nnratio, check_ori != 0);
    std::vector<MapPoint *> matches;
    const int nm = matcher.SearchByBoW(&KF[0], &KF[1], matches);
    for (int i = 0; i < n1; ++i) out_match12[i] = matches[i] ? (int32_t)(matches[i] - mps[1].data()) : -1;
    return nm;
}

/* ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) (src/ORBmatcher.cc:779-966, with
 * CheckDistEpipolarLine :165-182), the matcher of LocalMapping::CreateNewMapPoints.  Keyframes as in
 * refbow_search_by_bow_kf (only bit 0 of mp is read: the keypoint has a map point), plus mvKeysUn positions / octaves
 * inside the 28-B records, mvuRight (may be null: all -1), the poses (3 x 4 row-major), pKF2's intrinsics cam2 = fx, fy,
 * cx, cy and its level tables.  out_match12[i1] = pKF2 keypoint paired with pKF1 keypoint i1 or -1 (vMatchedPairs is that
 * list in i1 order).  Returns nmatches. */
int refbow_search_for_triangulation(void *v, int n1, const uint8_t *kps1, const uint8_t *desc1, const uint8_t *mp1,
                                    const float *ur1, const float *Tcw1, int n2, const uint8_t *kps2, const uint8_t *desc2,
                                    const uint8_t *mp2, const float *ur2, const float *Tcw2, const float *cam2, int nlevels,
                                    const float *scale2, const float *sigma2_2, const float *F12, int only_stereo,
                                    int check_ori, int32_t *out_match12)
{
    ORBVocabulary *voc = static_cast<ORBVocabulary *>(v);
    KeyFrame KF[2];
    const int n[2] = {n1, n2};
    const uint8_t *kps[2] = {kps1, kps2}, *desc[2] = {desc1, desc2}, *mp[2] = {mp1, mp2};
    const float *ur[2] = {ur1, ur2}, *Tcw[2] = {Tcw1, Tcw2};
    std::vector<MapPoint> mps[2];
    for (int s = 0; s < 2; ++s) {
        KeyFrame &K = KF[s];
        mps[s].resize((size_t)std::max(n[s], 1));
        K.N = n[s];
        K.mvKeysUn.resize(n[s]);
        if (n[s]) std::memcpy(K.mvKeysUn.data(), kps[s], (size_t)n[s] * 28);
        K.mvKeys = K.mvKeysUn;
        K.mDescriptors = desc_mat(desc[s], n[s]);
        K.mvpMapPoints.assign(n[s], static_cast<MapPoint *>(nullptr));
        K.mvuRight.assign(n[s], -1.f);
        if (ur[s]) K.mvuRight.assign(ur[s], ur[s] + n[s]);
        for (int i = 0; i < n[s]; ++i)
            if (mp[s][i] & 1) K.mvpMapPoints[i] = &mps[s][i];
        K.Tcw_ = cv::Mat(4, 4, CV_32F);
        for (int c = 0; c < 4; ++c) K.Tcw_.at<float>(3, c) = c == 3 ? 1.f : 0.f;
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 4; ++c) K.Tcw_.at<float>(r, c) = Tcw[s][4 * r + c];
        {   /* KeyFrame::SetPose (src/KeyFrame.cc): Ow = -Rcw.t() * tcw */
            cv::Mat Rcw = K.Tcw_.rowRange(0, 3).colRange(0, 3);
            cv::Mat tcw = K.Tcw_.rowRange(0, 3).col(3);
            K.Ow_ = -Rcw.t() * tcw;
        }
        K.fx = cam2[0]; K.fy = cam2[1]; K.cx = cam2[2]; K.cy = cam2[3];
        K.mnScaleLevels = nlevels;
        K.mvScaleFactors.assign(scale2, scale2 + nlevels);
        K.mvLevelSigma2.assign(sigma2_2, sigma2_2 + nlevels);
        std::vector<cv::Mat> vd;
        for (int i = 0; i < n[s]; ++i) vd.push_back(K.mDescriptors.row(i));
        DBoW2::BowVector bv;
        voc->transform(vd, bv, K.mFeatVec, 4);
    }
    cv::Mat F(3, 3, CV_32F);
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) F.at<float>(r, c) = F12[3 * r + c];
    ORBmatcher matcher(0.6f, check_ori != 0);
    std::vector<std::pair<size_t, size_t> > pairs;
    const int nm = matcher.SearchForTriangulation(&KF[0], &KF[1], F, pairs, only_stereo != 0);
    for (int i = 0; i < n1; ++i) out_match12[i] = -1;
    size_t prev = 0;
    for (size_t k = 0; k < pairs.size(); ++k) {
        if (k && pairs[k].first <= prev) std::abort(); /* the list is in i1 order */
        prev = pairs[k].first;
        out_match12[pairs[k].first] = (int32_t)pairs[k].second;
    }
    if ((int)pairs.size() != nm) std::abort();
    return nm;
}

} // extern "C"


