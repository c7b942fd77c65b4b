/*
 * oracle/match_oracle.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement, on flat arrays, of the tracking thread's frame-to-frame matcher
 *   ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, th, bMono)
 *       /root/reference/src/ORBmatcher.cc:1503-1668
 * with Frame::GetFeaturesInArea (src/Frame.cc:421-508), ORBmatcher::ComputeThreeMaxima
 * (src/ORBmatcher.cc:1800-1843), DescriptorDistance (:1846-1862), TH_HIGH = 100 (:35),
 * HISTO_LENGTH = 30 (:36).
 * PINNED against the reference's own object code: oracle/Makefile compiles src/ORBmatcher.cc and
 * src/Frame.cc where they lie into oracle/_ref/librefmatch.so (oracle/matcher_driver.cpp feeds it),
 * and tests/test_oracle_match_vs_reference.py requires identical CurrentFrame.mvpMapPoints and
 * nmatches; tests/golden/match_*.npz carries such outputs to machines without the reference.
 *
 * Float forms, read off the reference object (g++ -O3 -march=x86-64-v3, objdump of
 * _ref/ORBmatcher.o) and kept here with explicit fmaf under -ffp-contract=off:
 *   x3Dc   = Rcw * x3Dw + tcw   OpenCV gemm on 3x3 * 3x1 floats: left-to-right sum of individually
 *                               rounded products (cv2 4.13: tests/test_oracle_match.py), then + tcw
 *   invzc  = 1.0f / zc          (the double division, narrowed, is the float division)
 *   u      = fmaf(fx * xc, invzc, cx),  v likewise           (:1548-1549)
 *   radius = th * mvScaleFactors[octave]                     (:1561)
 *   ur     = fmaf(-mbf, invzc, u);  er = fabsf(ur - mvuRight[i2])   (:1593-1594)
 *   rot    = angleLast - angleCur; if (rot < 0) rot += 360.f; bin = round(rot * (1.f / 30))  (:1617-1622)
 *   twc    = -Rcw.t() * tcw     cv::gemm with GEMM_1_T, alpha = -1: OpenCV's generic path, the sum over k in DOUBLE,
 *                               times alpha, one rounding (cv2 4.13: test_transposed_gemm_form_is_cv2s)   (:1516)
 *   tlc    = Rlw * twc + tlw, the small-product form of x3Dc   (:1521)
 *
 * The reference loop is sequential: a current keypoint that already holds a map point with
 * Observations() > 0 is skipped by every later query (:1583-1585), one holding a map point without
 * observations can be overwritten (:1610).  nmatches counts every accepted query, and every entry
 * of a discarded orientation bin subtracts one and clears its slot, whoever holds it (:1638-1647).
 */
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

extern "C" {

struct mo_kp { float x, y, size, angle, response; int octave, class_id; };

struct mo_frame { /* same layout as refmatch_frame in oracle/matcher_driver.cpp */
    int n;
    const uint8_t *kps;
    const float *un_xy;
    const uint8_t *desc;
    const float *u_right;
    const int32_t *grid_count, *grid_idx;
    const float *Tcw;
};

static int mo_hamming256(const uint8_t *a, const uint8_t *b)
{
    int d = 0;
    for (int i = 0; i < 8; ++i) {
        uint32_t x, y;
        std::memcpy(&x, a + 4 * i, 4);
        std::memcpy(&y, b + 4 * i, 4);
        d += __builtin_popcount(x ^ y);
    }
    return d;
}

/* row r of a 3x4 pose times a 3-vector: OpenCV's small-matrix float product */
static float mo_dot3(const float *row, int stride, const float *x)
{
    float s = 0.f;
    for (int k = 0; k < 3; ++k) s += row[k * stride] * x[k];
    return s;
}

enum { MO_COLS = 64, MO_ROWS = 48, MO_HISTO = 30, MO_TH_HIGH = 100 };

/* ORBmatcher::ComputeThreeMaxima on the bin sizes */
void mo_three_maxima(const int *sz, int L, int *ind1, int *ind2, int *ind3)
{
    int max1 = 0, max2 = 0, max3 = 0;
    *ind1 = *ind2 = *ind3 = -1;
    for (int i = 0; i < L; ++i) {
        const int s = sz[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; *ind3 = *ind2; *ind2 = *ind1; *ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; *ind3 = *ind2; *ind2 = i; }
        else if (s > max3) { max3 = s; *ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { *ind2 = -1; *ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { *ind3 = -1; }
}

int mo_search_by_projection(const mo_frame *cur, const mo_frame *last, const float *cam, const float *bounds,
                            const float *scale, int nlevels, const uint8_t *mp_flags, const float *mp_xyz,
                            const uint8_t *mp_desc, float th, int mono, int check_ori, int32_t *out_match)
{
    (void)nlevels;
    const float fx = cam[0], fy = cam[1], cx = cam[2], cy = cam[3], mbf = cam[4], mb = cam[5];
    const float min_x = bounds[0], max_x = bounds[1], min_y = bounds[2], max_y = bounds[3];
    const float winv = (float)MO_COLS / (max_x - min_x), hinv = (float)MO_ROWS / (max_y - min_y);
    const mo_kp *ck = reinterpret_cast<const mo_kp *>(cur->kps), *lk = reinterpret_cast<const mo_kp *>(last->kps);
    const float *Tc = cur->Tcw, *Tl = last->Tcw;
    /* twc = -Rcw^T * tcw; tlc = Rlw * twc + tlw */
    float twc[3], tcw[3] = {Tc[3], Tc[7], Tc[11]};
    for (int y = 0; y < 3; ++y) {
        double s = 0.0;
        for (int k = 0; k < 3; ++k) s += (double)Tc[4 * k + y] * (double)tcw[k];
        twc[y] = (float)(s * -1.0);
    }
    const float tlc_z = mo_dot3(Tl + 8, 1, twc) + Tl[11];
    const bool forward = tlc_z > mb && !mono, backward = -tlc_z > mb && !mono;

    /* cell-major prefix of the current frame's grid */
    std::vector<int> cell_start(MO_COLS * MO_ROWS + 1, 0);
    if (cur->grid_count)
        for (int c = 0; c < MO_COLS * MO_ROWS; ++c) cell_start[c + 1] = cell_start[c] + cur->grid_count[c];

    for (int i = 0; i < cur->n; ++i) out_match[i] = -1;
    std::vector<int> hist[MO_HISTO];
    const float factor = 1.0f / MO_HISTO;
    int nmatches = 0;

    for (int i = 0; i < last->n; ++i) {
        if (!(mp_flags[i] & 1) || (mp_flags[i] & 4)) continue;
        const float *X = mp_xyz + 3 * i;
        const float xc = mo_dot3(Tc, 1, X) + Tc[3], yc = mo_dot3(Tc + 4, 1, X) + Tc[7], zc = mo_dot3(Tc + 8, 1, X) + Tc[11];
        const float invzc = 1.0f / zc;
        if (invzc < 0) continue;
        const float u = std::fmaf(fx * xc, invzc, cx), v = std::fmaf(fy * yc, invzc, cy);
        if (u < min_x || u > max_x) continue;
        if (v < min_y || v > max_y) continue;
        const int oct = lk[i].octave;
        const float radius = th * scale[oct];
        int min_level, max_level;
        if (forward) { min_level = oct; max_level = -1; }
        else if (backward) { min_level = 0; max_level = oct; }
        else { min_level = oct - 1; max_level = oct + 1; }

        /* Frame::GetFeaturesInArea(u, v, radius, min_level, max_level) fused with the candidate loop */
        const int cx0 = std::max(0, (int)std::floor((u - min_x - radius) * winv));
        if (cx0 >= MO_COLS) continue;
        const int cx1 = std::min((int)MO_COLS - 1, (int)std::ceil((u - min_x + radius) * winv));
        if (cx1 < 0) continue;
        const int cy0 = std::max(0, (int)std::floor((v - min_y - radius) * hinv));
        if (cy0 >= MO_ROWS) continue;
        const int cy1 = std::min((int)MO_ROWS - 1, (int)std::ceil((v - min_y + radius) * hinv));
        if (cy1 < 0) continue;
        const bool check_levels = (min_level > 0) || (max_level >= 0);

        int best = 256, best_idx = -1;
        for (int ix = cx0; ix <= cx1; ++ix)
            for (int iy = cy0; iy <= cy1; ++iy) {
                const int c = ix * MO_ROWS + iy;
                for (int q = cell_start[c]; q < cell_start[c + 1]; ++q) {
                    const int i2 = cur->grid_idx[q];
                    if (check_levels) {
                        if (ck[i2].octave < min_level) continue;
                        if (max_level >= 0 && ck[i2].octave > max_level) continue;
                    }
                    const float dx = cur->un_xy[2 * i2] - u, dy = cur->un_xy[2 * i2 + 1] - v;
                    if (!(std::fabs(dx) < radius && std::fabs(dy) < radius)) continue;
                    /* the candidate loop proper (:1578-1607) */
                    if (out_match[i2] >= 0 && (mp_flags[out_match[i2]] & 2)) continue;
                    const float ur2 = cur->u_right ? cur->u_right[i2] : -1.f;
                    if (ur2 > 0) {
                        const float ur = std::fmaf(-mbf, invzc, u);
                        const float er = std::fabs(ur - ur2);
                        if (er > radius) continue;
                    }
                    const int d = mo_hamming256(mp_desc + (size_t)i * 32, cur->desc + (size_t)i2 * 32);
                    if (d < best) { best = d; best_idx = i2; }
                }
            }
        if (best <= MO_TH_HIGH) {
            out_match[best_idx] = i;
            ++nmatches;
            if (check_ori) {
                float rot = lk[i].angle - ck[best_idx].angle;
                if (rot < 0.0f) rot += 360.0f;
                int bin = (int)std::round(rot * factor);
                if (bin == MO_HISTO) bin = 0;
                hist[bin].push_back(best_idx);
            }
        }
    }
    if (check_ori) {
        int sz[MO_HISTO], i1, i2, i3;
        for (int b = 0; b < MO_HISTO; ++b) sz[b] = (int)hist[b].size();
        mo_three_maxima(sz, MO_HISTO, &i1, &i2, &i3);
        for (int b = 0; b < MO_HISTO; ++b) {
            if (b == i1 || b == i2 || b == i3) continue;
            for (size_t j = 0; j < hist[b].size(); ++j) {
                out_match[hist[b][j]] = -1;
                --nmatches;
            }
        }
    }
    return nmatches;
}

/* ---------------------------------------------------------------------------------------------
 * ORBmatcher::SearchByProjection(Frame &F, const vector<MapPoint*> &vpMapPoints, const float th)
 *     /root/reference/src/ORBmatcher.cc:59-154 -- the matcher of Tracking::SearchLocalPoints
 *     (src/Tracking.cc:1055-1103), with RadiusByViewingCos (:156-162).
 * Map point iMP arrives as what Frame::isInFrustum left in it: proj[4 iMP ..] = mTrackProjX,
 * mTrackProjY, mTrackProjXR, mTrackViewCos; level = mnTrackScaleLevel; flags bit 0 = mbTrackInView
 * && !isBad(), bit 1 = Observations() > 0.  occupied[i2] = 2 where F.mvpMapPoints[i2] holds a point
 * with observations on entry (such slots are never touched, :92-94), anything else is free.
 * out_match[i2] = iMP assigned by this call, -1 where the slot is left as it was.
 * Float forms: r = 2.5 / 4.0 by (double)viewCos > 0.998, r *= th if th != 1.0, radius = r * scale[level]
 * (two single multiplications); the ratio test compares (float)bestDist with mfNNratio * (float)bestDist2.
 */
int mo_search_local_points(const mo_frame *cur, const float *bounds, const float *scale, int nlevels, int n_mp,
                           const float *proj, const uint8_t *mp_desc, const uint8_t *mp_flags, const int32_t *level,
                           const uint8_t *occupied, float th, float nnratio, int32_t *out_match)
{
    (void)nlevels;
    const float min_x = bounds[0], max_x = bounds[1], min_y = bounds[2], max_y = bounds[3];
    const float winv = (float)MO_COLS / (max_x - min_x), hinv = (float)MO_ROWS / (max_y - min_y);
    const mo_kp *ck = reinterpret_cast<const mo_kp *>(cur->kps);
    std::vector<int> cell_start(MO_COLS * MO_ROWS + 1, 0);
    if (cur->grid_count)
        for (int c = 0; c < MO_COLS * MO_ROWS; ++c) cell_start[c + 1] = cell_start[c] + cur->grid_count[c];
    for (int i = 0; i < cur->n; ++i) out_match[i] = -1;
    const bool b_factor = th != 1.0;
    int nmatches = 0;
    for (int i = 0; i < n_mp; ++i) {
        if (!(mp_flags[i] & 1)) continue;
        const int lvl = level[i];
        float r = (double)proj[4 * i + 3] > 0.998 ? 2.5f : 4.0f;
        if (b_factor) r *= th;
        const float radius = r * scale[lvl];
        const float u = proj[4 * i], v = proj[4 * i + 1], xr = proj[4 * i + 2];
        const int min_level = lvl - 1, max_level = lvl;
        const int cx0 = std::max(0, (int)std::floor((u - min_x - radius) * winv));
        if (cx0 >= MO_COLS) continue;
        const int cx1 = std::min((int)MO_COLS - 1, (int)std::ceil((u - min_x + radius) * winv));
        if (cx1 < 0) continue;
        const int cy0 = std::max(0, (int)std::floor((v - min_y - radius) * hinv));
        if (cy0 >= MO_ROWS) continue;
        const int cy1 = std::min((int)MO_ROWS - 1, (int)std::ceil((v - min_y + radius) * hinv));
        if (cy1 < 0) continue;
        const bool check_levels = (min_level > 0) || (max_level >= 0);
        int best = 256, best_level = -1, best2 = 256, best_level2 = -1, best_idx = -1;
        for (int ix = cx0; ix <= cx1; ++ix)
            for (int iy = cy0; iy <= cy1; ++iy) {
                const int c = ix * MO_ROWS + iy;
                for (int q = cell_start[c]; q < cell_start[c + 1]; ++q) {
                    const int i2 = cur->grid_idx[q];
                    if (check_levels) {
                        if (ck[i2].octave < min_level) continue;
                        if (max_level >= 0 && ck[i2].octave > max_level) continue;
                    }
                    const float dx = cur->un_xy[2 * i2] - u, dy = cur->un_xy[2 * i2 + 1] - v;
                    if (!(std::fabs(dx) < radius && std::fabs(dy) < radius)) continue;
                    if (occupied[i2] == 2) continue;
                    if (out_match[i2] >= 0 && (mp_flags[out_match[i2]] & 2)) continue;
                    const float ur2 = cur->u_right ? cur->u_right[i2] : -1.f;
                    if (ur2 > 0) {
                        const float er = std::fabs(xr - ur2);
                        if (er > radius) continue;
                    }
                    const int d = mo_hamming256(mp_desc + (size_t)i * 32, cur->desc + (size_t)i2 * 32);
                    if (d < best) { best2 = best; best = d; best_level2 = best_level; best_level = ck[i2].octave; best_idx = i2; }
                    else if (d < best2) { best_level2 = ck[i2].octave; best2 = d; }
                }
            }
        if (best <= MO_TH_HIGH) {
            if (best_level == best_level2 && (float)best > nnratio * (float)best2) continue;
            out_match[best_idx] = i;
            ++nmatches;
        }
    }
    return nmatches;
}

/* ---------------------------------------------------------------------------------------------
 * ORBmatcher::SearchForInitialization(Frame &F1, Frame &F2, vector<cv::Point2f> &vbPrevMatched,
 *     vector<int> &vnMatches12, int windowSize)      /root/reference/src/ORBmatcher.cc:487-632
 * -- the matcher of Tracking::MonocularInitialization (src/Tracking.cc:570-646): level-0 keypoints
 * of the initial frame F1 against the level-0 keypoints of F2 in a window around where each was
 * matched last; TH_LOW = 50 (:34), ratio test against (float)bestDist2 * mfNNratio, a match on an
 * already matched F2 keypoint needs a strictly smaller distance and then takes it over (:526-527,
 * :552-558), orientation histogram over ALL accepted matches incl. the ones taken over later (:562-577).
 * f1: kps (octave, angle), desc; f2: kps, un_xy, desc, grid.  prev[2 n1] is vbPrevMatched, updated in
 * place for the matched entries (:621-623).  matches12[n1] = vnMatches12.  Returns nmatches.
 */
int mo_search_for_initialization(const mo_frame *f1, const mo_frame *f2, const float *bounds, float *prev,
                                 int window_size, float nnratio, int check_ori, int32_t *matches12)
{
    const float min_x = bounds[0], max_x = bounds[1], min_y = bounds[2], max_y = bounds[3];
    const float winv = (float)MO_COLS / (max_x - min_x), hinv = (float)MO_ROWS / (max_y - min_y);
    const mo_kp *k1 = reinterpret_cast<const mo_kp *>(f1->kps), *k2 = reinterpret_cast<const mo_kp *>(f2->kps);
    std::vector<int> cell_start(MO_COLS * MO_ROWS + 1, 0);
    for (int c = 0; c < MO_COLS * MO_ROWS; ++c) cell_start[c + 1] = cell_start[c] + f2->grid_count[c];
    int nmatches = 0;
    for (int i = 0; i < f1->n; ++i) matches12[i] = -1;
    std::vector<int> hist[MO_HISTO];
    const float factor = 1.0f / MO_HISTO;
    std::vector<int> matched_distance(f2->n, INT_MAX), matches21(f2->n, -1);
    const float r = (float)window_size;
    for (int i1 = 0; i1 < f1->n; ++i1) {
        if (k1[i1].octave > 0) continue;
        const float x = prev[2 * i1], y = prev[2 * i1 + 1];
        const int cx0 = std::max(0, (int)std::floor((x - min_x - r) * winv));
        if (cx0 >= MO_COLS) continue;
        const int cx1 = std::min((int)MO_COLS - 1, (int)std::ceil((x - min_x + r) * winv));
        if (cx1 < 0) continue;
        const int cy0 = std::max(0, (int)std::floor((y - min_y - r) * hinv));
        if (cy0 >= MO_ROWS) continue;
        const int cy1 = std::min((int)MO_ROWS - 1, (int)std::ceil((y - min_y + r) * hinv));
        if (cy1 < 0) continue;
        int best = INT_MAX, best2 = INT_MAX, best_idx = -1;
        for (int ix = cx0; ix <= cx1; ++ix)
            for (int iy = cy0; iy <= cy1; ++iy) {
                const int c = ix * MO_ROWS + iy;
                for (int q = cell_start[c]; q < cell_start[c + 1]; ++q) {
                    const int i2 = f2->grid_idx[q];
                    if (k2[i2].octave < 0 || k2[i2].octave > 0) continue; /* minLevel = maxLevel = 0 */
                    const float dx = f2->un_xy[2 * i2] - x, dy = f2->un_xy[2 * i2 + 1] - y;
                    if (!(std::fabs(dx) < r && std::fabs(dy) < r)) continue;
                    const int d = mo_hamming256(f1->desc + (size_t)i1 * 32, f2->desc + (size_t)i2 * 32);
                    if (matched_distance[i2] <= d) continue;
                    if (d < best) { best2 = best; best = d; best_idx = i2; }
                    else if (d < best2) best2 = d;
                }
            }
        if (best <= 50) {
            if ((float)best < (float)best2 * nnratio) {
                if (matches21[best_idx] >= 0) { matches12[matches21[best_idx]] = -1; --nmatches; }
                matches12[i1] = best_idx;
                matches21[best_idx] = i1;
                matched_distance[best_idx] = best;
                ++nmatches;
                if (check_ori) {
                    float rot = k1[i1].angle - k2[best_idx].angle;
                    if (rot < 0.0f) rot += 360.0f;
                    int bin = (int)std::round(rot * factor);
                    if (bin == MO_HISTO) bin = 0;
                    hist[bin].push_back(i1);
                }
            }
        }
    }
    if (check_ori) {
        int sz[MO_HISTO], i1, i2, i3;
        for (int b = 0; b < MO_HISTO; ++b) sz[b] = (int)hist[b].size();
        mo_three_maxima(sz, MO_HISTO, &i1, &i2, &i3);
        for (int b = 0; b < MO_HISTO; ++b) {
            if (b == i1 || b == i2 || b == i3) continue;
            for (size_t j = 0; j < hist[b].size(); ++j) {
                const int idx1 = hist[b][j];
                if (matches12[idx1] >= 0) { matches12[idx1] = -1; --nmatches; }
            }
        }
    }
    for (int i1 = 0; i1 < f1->n; ++i1)
        if (matches12[i1] >= 0) { prev[2 * i1] = f2->un_xy[2 * matches12[i1]]; prev[2 * i1 + 1] = f2->un_xy[2 * matches12[i1] + 1]; }
    return nmatches;
}

/* ---------------------------------------------------------------------------------------------
 * ORBmatcher::SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, const set<MapPoint*> &sAlreadyFound,
 *                                const float th, const int ORBdist)
 *     /root/reference/src/ORBmatcher.cc:1670-1798 -- the matcher of Tracking::Relocalization's second and third
 *     search (src/Tracking.cc: "matcher2.SearchByProjection(mCurrentFrame, vpCandidateKFs[i], sFound, 10, 100)",
 *     then "(…, 3, 64)").
 * Entry i is keyframe keypoint i: kf_flags[i] bit 0 = vpMPs[i] != NULL, bit 1 = isBad(), bit 2 = the point is in
 * sAlreadyFound; kf_xyz = GetWorldPos(), kf_desc = GetDescriptor(), kf_dist[2i] = mfMinDistance, [2i + 1] =
 * mfMaxDistance (src/MapPoint.cc:520-530, :575-590), kf_angle[i] = pKF->mvKeysUn[i].angle.  occupied[i2] != 0 where
 * CurrentFrame.mvpMapPoints[i2] is set on entry (every set slot is skipped, :1735-1736 -- also the ones this call sets).
 * log_scale = CurrentFrame.mfLogScaleFactor.  Float forms as the reference object has them:
 *   Ow      = -Rcw.t() * tcw          cv::gemm GEMM_1_T, double accumulation (see the header)            (:1676)
 *   x3Dc, u, v                        as the frame-to-frame matcher; NO invzc < 0 test here             (:1693-1705)
 *   dist3D  = (float)cv::norm(x3Dw - Ow)   squares summed in double left to right, double sqrt          (:1708-1709)
 *   0.8f * mfMinDistance, 1.2f * mfMaxDistance                                                          (:1711-1716)
 *   PredictScale: ratio = mfMaxDistance / dist3D; nScale = ceilf(logf(ratio) / mfLogScaleFactor), clamped to
 *                 [0, mnScaleLevels - 1] (with `using namespace std` in effect both calls are the float overloads;
 *                 logf is glibc's, which the device restates: csrc/orb_logf.cuh)
 *   radius  = th * mvScaleFactors[nPredictedLevel], octaves nPredictedLevel - 1 .. nPredictedLevel + 1   (:1721-1723)
 */
int mo_search_by_projection_kf(const mo_frame *cur, const float *cam, const float *bounds, const float *scale, int nlevels,
                               float log_scale, int n_kf, const uint8_t *kf_flags, const float *kf_xyz, const uint8_t *kf_desc,
                               const float *kf_dist, const float *kf_angle, const uint8_t *occupied, float th, int orb_dist,
                               int check_ori, int32_t *out_match)
{
    const float fx = cam[0], fy = cam[1], cx = cam[2], cy = cam[3];
    const float min_x = bounds[0], max_x = bounds[1], min_y = bounds[2], max_y = bounds[3];
    const float winv = (float)MO_COLS / (max_x - min_x), hinv = (float)MO_ROWS / (max_y - min_y);
    const mo_kp *ck = reinterpret_cast<const mo_kp *>(cur->kps);
    const float *Tc = cur->Tcw;
    float Ow[3];
    for (int y = 0; y < 3; ++y) {
        double s = 0.0;
        for (int k = 0; k < 3; ++k) s += (double)Tc[4 * k + y] * (double)Tc[4 * k + 3];
        Ow[y] = (float)(s * -1.0);
    }
    std::vector<int> cell_start(MO_COLS * MO_ROWS + 1, 0);
    if (cur->grid_count)
        for (int c = 0; c < MO_COLS * MO_ROWS; ++c) cell_start[c + 1] = cell_start[c] + cur->grid_count[c];
    for (int i = 0; i < cur->n; ++i) out_match[i] = -1;
    std::vector<int> hist[MO_HISTO];
    const float factor = 1.0f / MO_HISTO;
    int nmatches = 0;

    for (int i = 0; i < n_kf; ++i) {
        if (!(kf_flags[i] & 1) || (kf_flags[i] & 6)) continue;
        const float *X = kf_xyz + 3 * i;
        const float xc = mo_dot3(Tc, 1, X) + Tc[3], yc = mo_dot3(Tc + 4, 1, X) + Tc[7], zc = mo_dot3(Tc + 8, 1, X) + Tc[11];
        const float invzc = 1.0f / zc;
        const float u = std::fmaf(fx * xc, invzc, cx), v = std::fmaf(fy * yc, invzc, cy);
        if (u < min_x || u > max_x) continue;
        if (v < min_y || v > max_y) continue;
        double s2 = 0.0;
        for (int k = 0; k < 3; ++k) {
            const float po = X[k] - Ow[k];
            s2 += (double)po * (double)po;
        }
        const float dist3D = (float)std::sqrt(s2);
        const float max_d = 1.2f * kf_dist[2 * i + 1], min_d = 0.8f * kf_dist[2 * i];
        if (dist3D < min_d || dist3D > max_d) continue;
        const float ratio = kf_dist[2 * i + 1] / dist3D;
        int lvl = (int)std::ceil(std::log(ratio) / log_scale); /* float overloads */
        if (lvl < 0) lvl = 0;
        else if (lvl >= nlevels) lvl = nlevels - 1;
        const float radius = th * scale[lvl];
        const int min_level = lvl - 1, max_level = lvl + 1;

        const int cx0 = std::max(0, (int)std::floor((u - min_x - radius) * winv));
        if (cx0 >= MO_COLS) continue;
        const int cx1 = std::min((int)MO_COLS - 1, (int)std::ceil((u - min_x + radius) * winv));
        if (cx1 < 0) continue;
        const int cy0 = std::max(0, (int)std::floor((v - min_y - radius) * hinv));
        if (cy0 >= MO_ROWS) continue;
        const int cy1 = std::min((int)MO_ROWS - 1, (int)std::ceil((v - min_y + radius) * hinv));
        if (cy1 < 0) continue;
        const bool check_levels = (min_level > 0) || (max_level >= 0);

        int best = 256, best_idx = -1;
        for (int ix = cx0; ix <= cx1; ++ix)
            for (int iy = cy0; iy <= cy1; ++iy) {
                const int c = ix * MO_ROWS + iy;
                for (int q = cell_start[c]; q < cell_start[c + 1]; ++q) {
                    const int i2 = cur->grid_idx[q];
                    if (check_levels) {
                        if (ck[i2].octave < min_level) continue;
                        if (max_level >= 0 && ck[i2].octave > max_level) continue;
                    }
                    const float dx = cur->un_xy[2 * i2] - u, dy = cur->un_xy[2 * i2 + 1] - v;
                    if (!(std::fabs(dx) < radius && std::fabs(dy) < radius)) continue;
                    if ((occupied && occupied[i2]) || out_match[i2] >= 0) continue; /* :1735-1736 */
                    const int d = mo_hamming256(kf_desc + (size_t)i * 32, cur->desc + (size_t)i2 * 32);
                    if (d < best) { best = d; best_idx = i2; }
                }
            }
        if (best <= orb_dist && best_idx >= 0) {
            out_match[best_idx] = i;
            ++nmatches;
            if (check_ori) {
                float rot = kf_angle[i] - ck[best_idx].angle;
                if (rot < 0.0f) rot += 360.0f;
                int bin = (int)std::round(rot * factor);
                if (bin == MO_HISTO) bin = 0;
                hist[bin].push_back(best_idx);
            }
        }
    }
    if (check_ori) {
        int sz[MO_HISTO], i1, i2, i3;
        for (int b = 0; b < MO_HISTO; ++b) sz[b] = (int)hist[b].size();
        mo_three_maxima(sz, MO_HISTO, &i1, &i2, &i3);
        for (int b = 0; b < MO_HISTO; ++b) {
            if (b == i1 || b == i2 || b == i3) continue;
            for (size_t j = 0; j < hist[b].size(); ++j) {
                out_match[hist[b][j]] = -1;
                --nmatches;
            }
        }
    }
    return nmatches;
}

/* ---------------------------------------------------------------------------------------------
 * ORBmatcher::Fuse(KeyFrame *pKF, const vector<MapPoint*> &vpMapPoints, const float th)
 *     /root/reference/src/ORBmatcher.cc:968-1131 -- LocalMapping::SearchInNeighbors.  The CHOICE of a keypoint per map
 *     point (what the function does to the map afterwards, :1103-1127, stays with the caller).  kf = the keyframe as a frame
 *     record (kps with octaves, un_xy, desc, u_right, grid, Tcw); bounds = the Frame's FLOAT mnMinX, mnMaxX, mnMinY, mnMaxY:
 *     the keyframe's own are these truncated to int (include/KeyFrame.h: const int), its grid factors the Frame's floats.
 * Float forms as the reference object has them:
 *   p3Dc = Rcw * p3Dw + tcw (small float product); skipped when z < 0; invz = 1.0f / z; x = xc * invz;
 *   u = fmaf(x, fx, cx), v likewise; IsInImage on the int bounds; ur = fmaf(-bf, invz, u)                       (:993-1011)
 *   dist3D = (float)cv::norm(p3Dw - Ow) with Ow = pKF->GetCameraCenter(); 0.8f * min, 1.2f * max                (:1013-1021)
 *   PO.dot(Pn) summed in double, < 0.5 * (double)dist3D -> skipped                                              (:1024-1027)
 *   PredictScale(dist3D, pKF); radius = th * mvScaleFactors[level]; keypoint levels level-1 .. level             (:1029-1054)
 *   t = fmaf(ex, ex, ey * ey); stereo keypoint: (double)(fmaf(er, er, t) * invSigma2) > 7.8, else
 *   (double)(invSigma2 * t) > 5.99 -> skipped                                                                    (:1056-1083)
 *   first smallest Hamming distance in GetFeaturesInArea order; chosen when <= TH_LOW                            (:1085-1101)
 */
int mo_fuse(const mo_frame *kf, const float *cam, const float *bounds, const float *scale, const float *inv_sigma2, int nlevels,
            float log_scale, int n_mp, const uint8_t *mp_flags, const float *mp_xyz, const uint8_t *mp_desc, const float *mp_dist,
            const float *mp_normal, float th, int32_t *out_best)
{
    const float fx = cam[0], fy = cam[1], cx = cam[2], cy = cam[3], bf = cam[4];
    const int min_x = (int)bounds[0], max_x = (int)bounds[1], min_y = (int)bounds[2], max_y = (int)bounds[3];
    const float winv = (float)MO_COLS / (bounds[1] - bounds[0]), hinv = (float)MO_ROWS / (bounds[3] - bounds[2]);
    const mo_kp *kk = reinterpret_cast<const mo_kp *>(kf->kps);
    const float *Tc = kf->Tcw;
    float Ow[3];
    for (int y = 0; y < 3; ++y) {
        double s = 0.0;
        for (int k = 0; k < 3; ++k) s += (double)Tc[4 * k + y] * (double)Tc[4 * k + 3];
        Ow[y] = (float)(s * -1.0);
    }
    std::vector<int> cell_start(MO_COLS * MO_ROWS + 1, 0);
    for (int c = 0; c < MO_COLS * MO_ROWS; ++c) cell_start[c + 1] = cell_start[c] + kf->grid_count[c];
    int nfused = 0;
    for (int i = 0; i < n_mp; ++i) {
        out_best[i] = -1;
        if (!(mp_flags[i] & 1) || (mp_flags[i] & 6)) continue;
        const float *X = mp_xyz + 3 * i;
        const float xc = mo_dot3(Tc, 1, X) + Tc[3], yc = mo_dot3(Tc + 4, 1, X) + Tc[7], zc = mo_dot3(Tc + 8, 1, X) + Tc[11];
        if (zc < 0.0f) continue;
        const float invz = 1.0f / zc;
        const float x = xc * invz, y = yc * invz;
        const float u = std::fmaf(x, fx, cx), v = std::fmaf(y, fy, cy);
        if (!(u >= min_x && u < max_x && v >= min_y && v < max_y)) continue;
        const float ur = std::fmaf(-bf, invz, u);
        double s2 = 0.0, dotn = 0.0;
        float PO[3];
        for (int k = 0; k < 3; ++k) {
            PO[k] = X[k] - Ow[k];
            s2 += (double)PO[k] * (double)PO[k];
        }
        const float dist3D = (float)std::sqrt(s2);
        if (dist3D < 0.8f * mp_dist[2 * i] || dist3D > 1.2f * mp_dist[2 * i + 1]) continue;
        for (int k = 0; k < 3; ++k) dotn += (double)PO[k] * (double)mp_normal[3 * i + k];
        if (dotn < 0.5 * (double)dist3D) continue;
        const float ratio = mp_dist[2 * i + 1] / dist3D;
        int lvl = (int)std::ceil(std::log(ratio) / log_scale);
        if (lvl < 0) lvl = 0;
        else if (lvl >= nlevels) lvl = nlevels - 1;
        const float radius = th * scale[lvl];
        /* KeyFrame::GetFeaturesInArea (src/KeyFrame.cc:750-789) fused with the candidate loop */
        const int cx0 = std::max(0, (int)std::floor((u - min_x - radius) * winv));
        if (cx0 >= MO_COLS) continue;
        const int cx1 = std::min((int)MO_COLS - 1, (int)std::ceil((u - min_x + radius) * winv));
        if (cx1 < 0) continue;
        const int cy0 = std::max(0, (int)std::floor((v - min_y - radius) * hinv));
        if (cy0 >= MO_ROWS) continue;
        const int cy1 = std::min((int)MO_ROWS - 1, (int)std::ceil((v - min_y + radius) * hinv));
        if (cy1 < 0) continue;
        int best = 256, best_idx = -1;
        for (int ix = cx0; ix <= cx1; ++ix)
            for (int iy = cy0; iy <= cy1; ++iy) {
                const int c = ix * MO_ROWS + iy;
                for (int q = cell_start[c]; q < cell_start[c + 1]; ++q) {
                    const int idx = kf->grid_idx[q];
                    const float kx = kf->un_xy[2 * idx], ky = kf->un_xy[2 * idx + 1];
                    if (!(std::fabs(kx - u) < radius && std::fabs(ky - v) < radius)) continue;
                    const int lv = kk[idx].octave;
                    if (lv < lvl - 1 || lv > lvl) continue;
                    const float ex = u - kx, ey = v - ky;
                    const float t = std::fmaf(ex, ex, ey * ey);
                    const float kr = kf->u_right ? kf->u_right[idx] : -1.f;
                    if (kr >= 0) {
                        const float er = ur - kr;
                        const float e2 = std::fmaf(er, er, t);
                        if ((double)(e2 * inv_sigma2[lv]) > 7.8) continue;
                    } else if ((double)(inv_sigma2[lv] * t) > 5.99) continue;
                    const int d = mo_hamming256(mp_desc + (size_t)i * 32, kf->desc + (size_t)idx * 32);
                    if (d < best) { best = d; best_idx = idx; }
                }
            }
        if (best <= 50) { out_best[i] = best_idx; ++nfused; }
    }
    return nfused;
}

/* ---------------------------------------------------------------------------------------------
 * The keyframe matchers that project points with a given transform and take, per point, the first smallest Hamming
 * distance among the keypoints of the predicted level and the one below inside a window of th * scale[level]:
 *   ORBmatcher::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint)                 src/ORBmatcher.cc:1133-1260
 *   ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th)        :353-485
 *   ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th)     :1262-1501  (both directions)
 * Shared float forms (the reference object): p3Dc = R * p3Dw + t (small float product; SearchBySim3 applies a second one,
 * p3Dc2 = sR21 * p3Dc1 + t21); skipped when z < 0; invz = 1.0f / z; x = xc * invz; u = fmaf(x, fx, cx); KeyFrame::IsInImage
 * on the int bounds; dist3D = (float)cv::norm(p3Dw - Ow) -- or, SearchBySim3, of the camera-frame point -- in double;
 * 0.8f * mfMinDistance <= dist3D <= 1.2f * mfMaxDistance; the two Scw functions also PO.dot(Pn) < 0.5 * dist3D in double;
 * MapPoint::PredictScale; radius = th * mvScaleFactors[level].
 * T1 / T2: 3 x 4 row-major [R | t]; T2 = NULL for one step; Ow = NULL: dist3D of the camera-frame point; normal = NULL: no
 * viewing-angle test.  A point is a query when bit 0 of its flag is set and bits 1, 2 are clear.
 * out_best[i] = keypoint or -1 (distance <= max_dist).  Returns the number chosen.
 */
int mo_project_search(const mo_frame *kf, const float *cam, const float *bounds, const float *scale, int nlevels, float log_scale,
                      const float *T1, const float *T2, const float *Ow, int n, const uint8_t *flags, const float *xyz,
                      const uint8_t *desc, const float *dist, const float *normal, float th, int max_dist, int32_t *out_best)
{
    const float fx = cam[0], fy = cam[1], cx = cam[2], cy = cam[3];
    const int min_x = (int)bounds[0], max_x = (int)bounds[1], min_y = (int)bounds[2], max_y = (int)bounds[3];
    const float winv = (float)MO_COLS / (bounds[1] - bounds[0]), hinv = (float)MO_ROWS / (bounds[3] - bounds[2]);
    const mo_kp *kk = reinterpret_cast<const mo_kp *>(kf->kps);
    std::vector<int> cell_start(MO_COLS * MO_ROWS + 1, 0);
    for (int c = 0; c < MO_COLS * MO_ROWS; ++c) cell_start[c + 1] = cell_start[c] + kf->grid_count[c];
    int nfound = 0;
    for (int i = 0; i < n; ++i) {
        out_best[i] = -1;
        if (!(flags[i] & 1) || (flags[i] & 6)) continue;
        const float *X = xyz + 3 * i;
        float P[3] = {mo_dot3(T1, 1, X) + T1[3], mo_dot3(T1 + 4, 1, X) + T1[7], mo_dot3(T1 + 8, 1, X) + T1[11]};
        if (T2) {
            const float Q[3] = {mo_dot3(T2, 1, P) + T2[3], mo_dot3(T2 + 4, 1, P) + T2[7], mo_dot3(T2 + 8, 1, P) + T2[11]};
            P[0] = Q[0]; P[1] = Q[1]; P[2] = Q[2];
        }
        if (P[2] < 0.0f) continue;
        const float invz = 1.0f / P[2];
        const float x = P[0] * invz, y = P[1] * invz;
        const float u = std::fmaf(x, fx, cx), v = std::fmaf(y, fy, cy);
        if (!(u >= min_x && u < max_x && v >= min_y && v < max_y)) continue;
        double s2 = 0.0, dotn = 0.0;
        float PO[3];
        for (int k = 0; k < 3; ++k) {
            PO[k] = Ow ? X[k] - Ow[k] : P[k];
            s2 += (double)PO[k] * (double)PO[k];
        }
        const float dist3D = (float)std::sqrt(s2);
        if (dist3D < 0.8f * dist[2 * i] || dist3D > 1.2f * dist[2 * i + 1]) continue;
        if (normal) {
            for (int k = 0; k < 3; ++k) dotn += (double)PO[k] * (double)normal[3 * i + k];
            if (dotn < 0.5 * (double)dist3D) continue;
        }
        const float ratio = dist[2 * i + 1] / dist3D;
        int lvl = (int)std::ceil(std::log(ratio) / log_scale);
        if (lvl < 0) lvl = 0;
        else if (lvl >= nlevels) lvl = nlevels - 1;
        const float radius = th * scale[lvl];
        const int cx0 = std::max(0, (int)std::floor((u - min_x - radius) * winv));
        if (cx0 >= MO_COLS) continue;
        const int cx1 = std::min((int)MO_COLS - 1, (int)std::ceil((u - min_x + radius) * winv));
        if (cx1 < 0) continue;
        const int cy0 = std::max(0, (int)std::floor((v - min_y - radius) * hinv));
        if (cy0 >= MO_ROWS) continue;
        const int cy1 = std::min((int)MO_ROWS - 1, (int)std::ceil((v - min_y + radius) * hinv));
        if (cy1 < 0) continue;
        int best = INT_MAX, best_idx = -1;
        for (int ix = cx0; ix <= cx1; ++ix)
            for (int iy = cy0; iy <= cy1; ++iy) {
                const int c = ix * MO_ROWS + iy;
                for (int q = cell_start[c]; q < cell_start[c + 1]; ++q) {
                    const int idx = kf->grid_idx[q];
                    if (!(std::fabs(kf->un_xy[2 * idx] - u) < radius && std::fabs(kf->un_xy[2 * idx + 1] - v) < radius)) continue;
                    const int lv = kk[idx].octave;
                    if (lv < lvl - 1 || lv > lvl) continue;
                    const int d = mo_hamming256(desc + (size_t)i * 32, kf->desc + (size_t)idx * 32);
                    if (d < best) { best = d; best_idx = idx; }
                }
            }
        if (best <= max_dist) { out_best[i] = best_idx; ++nfound; }
    }
    return nfound;
}

/* Scw -> Rcw, tcw, Ow as :360-365 / :1140-1146 compute them: scw = (float)sqrt(row0 . row0) with the dot product in double;
 * A / scw is A * (float)(1.0 / scw) (cv::MatExpr: convertTo with alpha = 1. / s); Ow = -Rcw.t() * tcw through cv::gemm's
 * GEMM_1_T path.  T = [Rcw | tcw], 3 x 4. */
void mo_sim3_decompose(const float *Scw /* 4 x 4 row-major */, float *T, float *Ow)
{
    double d = 0.0;
    for (int k = 0; k < 3; ++k) d += (double)Scw[k] * (double)Scw[k];
    const float scw = (float)std::sqrt(d);
    const float inv = 1.0f / scw;
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 4; ++c) T[4 * r + c] = inv * Scw[4 * r + c];
    for (int y = 0; y < 3; ++y) {
        double s = 0.0;
        for (int k = 0; k < 3; ++k) s += (double)T[4 * k + y] * (double)T[4 * k + 3];
        Ow[y] = (float)(s * -1.0);
    }
}

/* sR12 = s12 * R12; sR21 = (1.0 / s12) * R12.t(); t21 = -sR21 * t12 (:1279-1281): T12 = [sR12 | t12], T21 = [sR21 | t21] */
void mo_sim3_relative(float s12, const float *R12, const float *t12, float *T12, float *T21)
{
    const float is = (float)(1.0 / (double)s12);
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) {
            T12[4 * r + c] = s12 * R12[3 * r + c];
            T21[4 * r + c] = R12[3 * c + r] * is;
        }
        T12[4 * r + 3] = t12[r];
    }
    for (int r = 0; r < 3; ++r) {
        float s = 0.f;
        for (int k = 0; k < 3; ++k) s += (-T21[4 * r + k]) * t12[k];
        T21[4 * r + 3] = s;
    }
}

/* ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) :353-485: sequential in the points -- a keypoint whose
 * vpMatched entry is set (on entry, or by an earlier point of this call) is skipped (:437-438).  T = [Rcw | tcw], Ow from
 * mo_sim3_decompose.  out_match[idx] = the point put into vpMatched[idx] by this call or -1.  Returns nmatches. */
int mo_search_by_projection_kf_sim3(const mo_frame *kf, const float *cam, const float *bounds, const float *scale, int nlevels,
                                    float log_scale, const float *T, const float *Ow, int n, const uint8_t *flags, const float *xyz,
                                    const uint8_t *desc, const float *dist, const float *normal, const uint8_t *occupied, int th,
                                    int32_t *out_match)
{
    const float fx = cam[0], fy = cam[1], cx = cam[2], cy = cam[3];
    const int min_x = (int)bounds[0], max_x = (int)bounds[1], min_y = (int)bounds[2], max_y = (int)bounds[3];
    const float winv = (float)MO_COLS / (bounds[1] - bounds[0]), hinv = (float)MO_ROWS / (bounds[3] - bounds[2]);
    const mo_kp *kk = reinterpret_cast<const mo_kp *>(kf->kps);
    std::vector<int> cell_start(MO_COLS * MO_ROWS + 1, 0);
    for (int c = 0; c < MO_COLS * MO_ROWS; ++c) cell_start[c + 1] = cell_start[c] + kf->grid_count[c];
    for (int i = 0; i < kf->n; ++i) out_match[i] = -1;
    int nmatches = 0;
    for (int i = 0; i < n; ++i) {
        if (!(flags[i] & 1) || (flags[i] & 6)) continue;
        const float *X = xyz + 3 * i;
        const float P[3] = {mo_dot3(T, 1, X) + T[3], mo_dot3(T + 4, 1, X) + T[7], mo_dot3(T + 8, 1, X) + T[11]};
        if (P[2] < 0.0f) continue;
        const float invz = 1.0f / P[2];
        const float x = P[0] * invz, y = P[1] * invz;
        const float u = std::fmaf(x, fx, cx), v = std::fmaf(y, fy, cy);
        if (!(u >= min_x && u < max_x && v >= min_y && v < max_y)) continue;
        double s2 = 0.0, dotn = 0.0;
        float PO[3];
        for (int k = 0; k < 3; ++k) {
            PO[k] = X[k] - Ow[k];
            s2 += (double)PO[k] * (double)PO[k];
        }
        const float dist3D = (float)std::sqrt(s2);
        if (dist3D < 0.8f * dist[2 * i] || dist3D > 1.2f * dist[2 * i + 1]) continue;
        for (int k = 0; k < 3; ++k) dotn += (double)PO[k] * (double)normal[3 * i + k];
        if (dotn < 0.5 * (double)dist3D) continue;
        const float ratio = dist[2 * i + 1] / dist3D;
        int lvl = (int)std::ceil(std::log(ratio) / log_scale);
        if (lvl < 0) lvl = 0;
        else if (lvl >= nlevels) lvl = nlevels - 1;
        const float radius = (float)th * scale[lvl];
        const int cx0 = std::max(0, (int)std::floor((u - min_x - radius) * winv));
        if (cx0 >= MO_COLS) continue;
        const int cx1 = std::min((int)MO_COLS - 1, (int)std::ceil((u - min_x + radius) * winv));
        if (cx1 < 0) continue;
        const int cy0 = std::max(0, (int)std::floor((v - min_y - radius) * hinv));
        if (cy0 >= MO_ROWS) continue;
        const int cy1 = std::min((int)MO_ROWS - 1, (int)std::ceil((v - min_y + radius) * hinv));
        if (cy1 < 0) continue;
        int best = 256, best_idx = -1;
        for (int ix = cx0; ix <= cx1; ++ix)
            for (int iy = cy0; iy <= cy1; ++iy) {
                const int c = ix * MO_ROWS + iy;
                for (int q = cell_start[c]; q < cell_start[c + 1]; ++q) {
                    const int idx = kf->grid_idx[q];
                    if (!(std::fabs(kf->un_xy[2 * idx] - u) < radius && std::fabs(kf->un_xy[2 * idx + 1] - v) < radius)) continue;
                    if ((occupied && occupied[idx]) || out_match[idx] >= 0) continue;
                    const int lv = kk[idx].octave;
                    if (lv < lvl - 1 || lv > lvl) continue;
                    const int d = mo_hamming256(desc + (size_t)i * 32, kf->desc + (size_t)idx * 32);
                    if (d < best) { best = d; best_idx = idx; }
                }
            }
        if (best <= 50) { out_match[best_idx] = i; ++nmatches; }
    }
    return nmatches;
}

/* ORBmatcher::SearchBySim3 :1262-1501: both directions through mo_project_search, then the agreement check (:1484-1498).
 * T1w / T2w = the keyframes' poses; flags bit 0 = the keypoint has a map point, bit 1 = bad; pre12 as the driver's. */
int mo_search_by_sim3(const mo_frame *k1, const mo_frame *k2, const float *cam, const float *bounds, const float *scale, int nlevels,
                      float log_scale, const uint8_t *flags1, const float *xyz1, const uint8_t *desc1, const float *dist1,
                      const uint8_t *flags2, const float *xyz2, const uint8_t *desc2, const float *dist2, const int32_t *pre12,
                      float s12, const float *R12, const float *t12, float th, int32_t *out12)
{
    float T12[12], T21[12];
    mo_sim3_relative(s12, R12, t12, T12, T21);
    const int n1 = k1->n, n2 = k2->n;
    std::vector<uint8_t> f1(flags1, flags1 + n1), f2(flags2, flags2 + n2);
    for (int i = 0; i < n1; ++i)
        if (pre12[i] >= 0) { /* vbAlreadyMatched1 / 2 (:1291-1302) */
            f1[i] |= 4;
            if (pre12[i] < n2) f2[pre12[i]] |= 4;
        }
    std::vector<int32_t> m1((size_t)std::max(n1, 1)), m2((size_t)std::max(n2, 1));
    mo_project_search(k2, cam, bounds, scale, nlevels, log_scale, k1->Tcw, T21, nullptr, n1, f1.data(), xyz1, desc1, dist1, nullptr,
                      th, MO_TH_HIGH, m1.data());
    mo_project_search(k1, cam, bounds, scale, nlevels, log_scale, k2->Tcw, T12, nullptr, n2, f2.data(), xyz2, desc2, dist2, nullptr,
                      th, MO_TH_HIGH, m2.data());
    int nfound = 0;
    for (int i = 0; i < n1; ++i) {
        out12[i] = pre12[i];
        const int idx2 = m1[i];
        if (idx2 >= 0 && m2[idx2] == i) { out12[i] = idx2; ++nfound; }
    }
    return nfound;
}

} // extern "C"



