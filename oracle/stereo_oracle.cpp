/*
 * oracle/stereo_oracle.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of Frame::ComputeStereoMatches (/root/reference/src/Frame.cc:619-834) with
 * ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1846-1862, TH_HIGH/TH_LOW :37-38), on flat
 * arrays.  PINNED against the reference's own src/Frame.cc: oracle/Makefile compiles that file where
 * it lies into oracle/_ref/libreframe.so (stand-ins for the five headers it cannot have here are in
 * oracle/framecompat/), and tests/test_oracle_frame_vs_reference.py requires this restatement, fed
 * the reference Frame's own keypoints and descriptors, to reproduce its mvuRight / mvDepth bit for
 * bit; tests/golden/frame_expected.npz carries such outputs to machines without the reference.
 * (DescriptorDistance and TH_LOW/TH_HIGH are restated in framecompat/ORBmatcher.h: src/ORBmatcher.cc
 * is not compilable here.)
 *
 * Notes on the reference's behaviour that are kept:
 *  - minZ = mb is read BEFORE the constructor assigns mb (src/Frame.cc:117 vs :140), i.e. it is
 *    whatever the caller's member holds; here it is an explicit argument.
 *  - candidates of a left keypoint are the right keypoints whose row band
 *    [floor(y - 2*scale), ceil(y + 2*scale)] contains (size_t)yL, visited in right-index order, the
 *    first smallest Hamming distance wins (strict <);
 *  - SAD over an 11x11 window of the keypoint's pyramid level, 11 shifts, parabola fit; all sums
 *    are small integers, exact in float;
 *  - outlier cut at 1.5*1.4*median of the SAD distances (element size/2 of the sorted list).
 * Build with -ffp-contract=off: every float operation individually rounded.
 */
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <utility>
#include <vector>

extern "C" {

struct so_kp { float x, y, size, angle, response; int octave, class_id; };
struct so_level { const uint8_t *data; int cols, rows; size_t step; };

static int hamming256(const uint8_t *a, const uint8_t *b)
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

/* returns the number of matches kept; u_right / depth get -1 where there is none;
 * sad_out (optional) gets the SAD distance of every left keypoint that reached the final list
 * (before the outlier cut), -1 otherwise */
int so_stereo_match(const so_kp *kl, const uint8_t *dl, int nl, const so_kp *kr, const uint8_t *dr, int nr,
                    const so_level *pyr_l, const so_level *pyr_r, int nlevels, const float *scale,
                    const float *inv_scale, float mbf, float mb, float *u_right, float *depth, int *sad_out)
{
    (void)nlevels;
    for (int i = 0; i < nl; ++i) { u_right[i] = -1.0f; depth[i] = -1.0f; if (sad_out) sad_out[i] = -1; }
    if (nl == 0) return 0;
    const int th_orb = (100 + 50) / 2;
    const int n_rows = pyr_l[0].rows;
    std::vector<std::vector<int>> rows((size_t)n_rows);
    for (int i = 0; i < nr; ++i) {
        const float y = kr[i].y, r = 2.0f * scale[kr[i].octave];
        const int maxr = (int)std::ceil(y + r), minr = (int)std::floor(y - r);
        for (int yi = minr; yi <= maxr; ++yi)
            if (yi >= 0 && yi < n_rows) rows[(size_t)yi].push_back(i);
    }
    const float min_z = mb, min_d = 0.f, max_d = mbf / min_z;
    std::vector<std::pair<int, int>> dist_idx;
    for (int il = 0; il < nl; ++il) {
        const so_kp &k = kl[il];
        const int lv = k.octave;
        const float vl = k.y, ul = k.x;
        const size_t row = (size_t)vl;
        if (row >= (size_t)n_rows) continue;
        const std::vector<int> &cand = rows[row];
        if (cand.empty()) continue;
        const float min_u = ul - max_d, max_u = ul - min_d;
        if (max_u < 0) continue;
        int best = 100, best_r = 0;
        for (int ir : cand) {
            const so_kp &q = kr[ir];
            if (q.octave < lv - 1 || q.octave > lv + 1) continue;
            if (q.x >= min_u && q.x <= max_u) {
                const int d = hamming256(dl + (size_t)il * 32, dr + (size_t)ir * 32);
                if (d < best) { best = d; best_r = ir; }
            }
        }
        if (best >= th_orb) continue;
        const float ur0 = kr[best_r].x, sf = inv_scale[lv];
        const float su = std::round(k.x * sf), sv = std::round(k.y * sf), sr = std::round(ur0 * sf);
        const int w = 5, L = 5;
        const so_level &PL = pyr_l[lv], &PR = pyr_r[lv];
        const float iniu = sr + L - w, endu = sr + L + w + 1;
        if (iniu < 0 || endu >= PR.cols) continue;
        const int cu = (int)su, cv = (int)sv, cr = (int)sr;
        int best_sad = INT_MAX, best_inc = 0;
        float dists[2 * 5 + 1];
        const int cl = PL.data[(size_t)cv * PL.step + cu];
        for (int inc = -L; inc <= L; ++inc) {
            const int cc = PR.data[(size_t)cv * PR.step + cr + inc];
            int sad = 0;
            for (int dy = -w; dy <= w; ++dy)
                for (int dx = -w; dx <= w; ++dx) {
                    const int a = PL.data[(size_t)(cv + dy) * PL.step + cu + dx] - cl;
                    const int b = PR.data[(size_t)(cv + dy) * PR.step + cr + inc + dx] - cc;
                    sad += std::abs(a - b);
                }
            const float d = (float)sad;
            if (d < (float)best_sad) { best_sad = (int)d; best_inc = inc; }
            dists[L + inc] = d;
        }
        if (best_inc == -L || best_inc == L) continue;
        const float d1 = dists[L + best_inc - 1], d2 = dists[L + best_inc], d3 = dists[L + best_inc + 1];
        const float delta = (d1 - d3) / (2.0f * (d1 + d3 - 2.0f * d2));
        if (delta < -1 || delta > 1) continue;
        float best_ur = scale[lv] * ((float)sr + (float)best_inc + delta);
        float disparity = ul - best_ur;
        if (disparity >= min_d && disparity < max_d) {
            if (disparity <= 0) {
                disparity = 0.01;
                best_ur = ul - 0.01; /* double arithmetic, as written in the reference */
            }
            depth[il] = mbf / disparity;
            u_right[il] = best_ur;
            if (sad_out) sad_out[il] = best_sad;
            dist_idx.push_back(std::make_pair(best_sad, il));
        }
    }
    if (dist_idx.empty()) return 0;
    std::sort(dist_idx.begin(), dist_idx.end());
    const float median = (float)dist_idx[dist_idx.size() / 2].first;
    const float th = 1.5f * 1.4f * median;
    int kept = (int)dist_idx.size();
    for (int i = (int)dist_idx.size() - 1; i >= 0; --i) {
        if ((float)dist_idx[i].first < th) break;
        u_right[dist_idx[i].second] = -1;
        depth[dist_idx[i].second] = -1;
        --kept;
    }
    return kept;
}

} // extern "C"
