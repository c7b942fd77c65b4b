/*
 * k_match.cuh -- M1: the tracking thread's frame-to-frame matcher on the device,
 *   ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, th, bMono)
 *       /root/reference/src/ORBmatcher.cc:1503-1668
 * with Frame::GetFeaturesInArea (src/Frame.cc:421-508), ComputeThreeMaxima (:1800-1843) and
 * DescriptorDistance (:1846-1862), over the CURRENT frame's device-resident results: cv::KeyPoint
 * records and descriptors (K7), undistorted positions, the 64 x 48 grid and the image bounds (K8),
 * mvuRight (S1/S2).  The last frame's map points arrive as flat arrays (world position, representative
 * descriptor, octave and angle of the keypoint they were seen at, flags).
 *
 * The reference loop is sequential in the last frame's index i: a current keypoint that holds a map
 * point with Observations() > 0 is skipped by every later query (:1583-1585).  Restated as a fixed
 * point: with taker[c] = the smallest i with observations whose best candidate is c, query i may use
 * candidate c iff taker[c] >= i.  Rounds of (all queries choose given the previous takers -> new
 * takers) reach the unique fixed point of that recurrence -- the sequential result -- after as many
 * rounds as the longest chain of displaced queries (2-4 on real frames); a round in which no choice
 * changes ends the loop.  The rest is order-free: nmatches counts the accepted queries, a slot ends
 * up with the LARGEST query index that chose it (later assignments overwrite, :1610), every entry of
 * an orientation bin outside the three maxima clears its slot and subtracts one (:1638-1647).
 *
 * Two kernels per call:
 *   M1a k_match_candidates  warp per query, grid over the whole GPU: projection, the grid cells of
 *       GetFeaturesInArea walked 32 candidates at a time (in a column of cells the cell-major lists are
 *       consecutive, and the list position q grows in the reference's scan order ix, iy, j), octave /
 *       window / mvuRight filters, 256-bit Hamming by popc; the MT_K best (distance << 16 | q) keys --
 *       i.e. best first, ties in scan order as the reference's strict < keeps the first -- are cached
 *       per query together with "there were more".
 *   M1b k_match_resolve     one CTA: the rounds of the fixed point on the cached keys (a query's
 *       choice is its first cached candidate nobody earlier holds; only a query whose MT_K cached
 *       candidates are all held AND that had more rescans its cells), then assignment, orientation
 *       histogram, ComputeThreeMaxima, removal.
 * Float forms as the reference object has them (oracle/match_oracle.cpp): products summed left to
 * right for the 3x3 * 3x1 gemm, u = fma(fx*xc, 1/zc, cx), ur = fma(-mbf, 1/zc, u); the library is
 * built with --fmad=false, so only the fmaf() written here fuses.
 */
#ifndef K_MATCH_CUH
#define K_MATCH_CUH

#include "borb_internal.h"
#include "k_describe.cuh" /* BorbKpRec */
#include "orb_logf.cuh"

#define MT_NT 1024
#define MT_COLS 64
#define MT_ROWS 48
#define MT_HISTO 30
#define MT_TH_HIGH 100
#define MT_TH_LOW 50

struct MatchParams {
    float Tc[12], Tl[12]; /* Tcw of the current and the last frame, 3 x 4 row-major */
    float fx, fy, cx, cy, mbf, mb, th;
    float nnratio; /* M2: mfNNratio of the local-map matcher */
    int mono, check_ori, n_last;
    int dir; /* 1 forward, 2 backward, 0 neither: from tlc.z against mb (:1515-1526), evaluated on the host */
    int max_rounds; /* fixed-point rounds of M1b before the open suffix is finished sequentially (MT_MAX_ROUNDS) */
    float scale[BORB_MAXL];
    /* MODE 4 (relocalisation matcher): camera centre Ow = -Rcw.t() * tcw, CurrentFrame.mfLogScaleFactor, mnScaleLevels, ORBdist */
    float Ow[3], log_scale;
    int nlevels, orb_dist;
    float inv_sigma2[BORB_MAXL]; /* MODE 5: pKF->mvInvLevelSigma2 for Fuse's chi-square tests */
    /* MODE 6 (the loop-closing matchers): a second transform applied to the camera-frame point (SearchBySim3), dist3D taken
     * of the camera-frame point instead of p3Dw - Ow (SearchBySim3), the viewing-angle test on / off */
    int two_step, cam_norm, use_normal;
};

__device__ __forceinline__ float mt_dot3(const float *row, const float *x)
{
    float s = 0.f;
    s += row[0] * x[0];
    s += row[1] * x[1];
    s += row[2] * x[2];
    return s;
}

__device__ __forceinline__ int mt_hamming(const uint4 a0, const uint4 a1, const uint4 *b)
{
    const uint4 b0 = b[0], b1 = b[1];
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

/* half away from zero, exactly (std::round) */
__device__ __forceinline__ int mt_round(float x)
{
    float t = truncf(x);
    if (fabsf(x - t) >= 0.5f) t += copysignf(1.f, x);
    return (int)t;
}

__device__ __forceinline__ int mt_bin(float a_last, float a_cur)
{
    float rot = a_last - a_cur;
    if (rot < 0.0f) rot += 360.0f;
    int bin = mt_round(rot * (1.0f / MT_HISTO));
    if (bin == MT_HISTO) bin = 0;
    return bin;
}

#define MT_K 8            /* cached candidates per query (M1b loads them as two uint4) */
#ifndef MT_MAX_ROUNDS
#define MT_MAX_ROUNDS 16  /* fixed-point rounds of M1b before the open suffix is finished sequentially */
#endif
#define MT_CAND_NT 256    /* M1a: 8 queries per CTA */
#define MT_NOKEY 0xffffffffu

struct MtQuery { /* what the candidate scan of one query needs (:1537-1572) */
    bool ok, check_levels;
    float u, v, radius, ur;
    int min_level, max_level, cx0, cx1, cy0, cy1;
    float inv_s2[2]; /* kind 1: mvInvLevelSigma2 of the two admitted levels */
    int kind; /* the per-candidate test beside level and window: 0 = mvuRight within the radius (:1587-1595, :96-102),
                 1 = Fuse's chi-square test on the pixel error (:1056-1083), 2 = none */
};

/* MODE 1: the frame-to-frame matcher (geo = world position; src/ORBmatcher.cc:1537-1572).
 * MODE 3: the monocular-initialisation matcher (k_match_resolve_init below): geo = vbPrevMatched[i1].
 * MODE 2: the local-map matcher, ORBmatcher::SearchByProjection(F, vpMapPoints, th) (src/ORBmatcher.cc:59-154) of
 *         Tracking::SearchLocalPoints: geo = mTrackProjX, mTrackProjY, mTrackProjXR, mTrackViewCos as
 *         Frame::isInFrustum left them, lvl = mnTrackScaleLevel; r = RadiusByViewingCos (:156-162), r *= th unless
 *         th == 1, window r * scale[lvl], octaves lvl-1 .. lvl (:72-79).
 * MODE 4: the relocalisation matcher, ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist)
 *         (src/ORBmatcher.cc:1670-1798): geo = world position and mfMaxDistance, `oct` = the BIT PATTERN of mfMinDistance
 *         (this mode predicts the level, so the octave column carries the second distance).  Projection as MODE 1 but with
 *         no invzc < 0 test (:1697-1705); dist3D = (float)cv::norm(x3Dw - Ow): squares summed in double, left to right,
 *         double square root; 0.8f * mfMinDistance <= dist3D <= 1.2f * mfMaxDistance (:1711-1716); MapPoint::PredictScale
 *         (src/MapPoint.cc:575-590): ceilf(logf(mfMaxDistance / dist3D) / mfLogScaleFactor) clamped to the levels, logf =
 *         glibc's (orb_logf.cuh); window th * scale[level], octaves level-1 .. level+1. */
template <int MODE>
__device__ __forceinline__ MtQuery mt_query(const MatchParams &mp, const float *__restrict__ l_geo, int oct, int i,
                                            float min_x, float max_x, float min_y, float max_y, float winv, float hinv,
                                            const float4 *__restrict__ l_extra = nullptr)
{
    MtQuery Q;
    Q.kind = MODE >= 5 ? 2 : 0;
    const float4 g = reinterpret_cast<const float4 *>(l_geo)[i];
    if (MODE >= 5) { /* a KeyFrame's bounds are its Frame's float bounds truncated to int (include/KeyFrame.h: const int mnMinX ...) */
        min_x = truncf(min_x); max_x = truncf(max_x); min_y = truncf(min_y); max_y = truncf(max_y);
    }
    if (MODE == 1) {
        const float X[3] = {g.x, g.y, g.z};
        const float xc = mt_dot3(mp.Tc, X) + mp.Tc[3], yc = mt_dot3(mp.Tc + 4, X) + mp.Tc[7], zc = mt_dot3(mp.Tc + 8, X) + mp.Tc[11];
        const float invzc = 1.0f / zc;
        Q.u = fmaf(mp.fx * xc, invzc, mp.cx);
        Q.v = fmaf(mp.fy * yc, invzc, mp.cy);
        Q.ok = !(invzc < 0) && !(Q.u < min_x || Q.u > max_x) && !(Q.v < min_y || Q.v > max_y);
        Q.radius = mp.th * mp.scale[oct];
        if (mp.dir == 1) { Q.min_level = oct; Q.max_level = -1; }
        else if (mp.dir == 2) { Q.min_level = 0; Q.max_level = oct; }
        else { Q.min_level = oct - 1; Q.max_level = oct + 1; }
        Q.ur = fmaf(-mp.mbf, invzc, Q.u);
    } else if (MODE == 4) {
        const float X[3] = {g.x, g.y, g.z};
        const float xc = mt_dot3(mp.Tc, X) + mp.Tc[3], yc = mt_dot3(mp.Tc + 4, X) + mp.Tc[7], zc = mt_dot3(mp.Tc + 8, X) + mp.Tc[11];
        const float invzc = 1.0f / zc;
        Q.u = fmaf(mp.fx * xc, invzc, mp.cx);
        Q.v = fmaf(mp.fy * yc, invzc, mp.cy);
        Q.ok = !(Q.u < min_x || Q.u > max_x) && !(Q.v < min_y || Q.v > max_y);
        double s2 = 0.0;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float po = X[k] - mp.Ow[k];
            s2 = __dadd_rn(s2, __dmul_rn((double)po, (double)po));
        }
        const float dist3D = (float)__dsqrt_rn(s2);
        const float max_d = 1.2f * g.w, min_d = 0.8f * __int_as_float(oct);
        Q.ok = Q.ok && !(dist3D < min_d || dist3D > max_d);
        const float ratio = g.w / dist3D;
        int lvl = orb_cvtt_x86(ceilf(orb_logf_exact(ratio) / mp.log_scale));
        if (lvl < 0) lvl = 0;
        else if (lvl >= mp.nlevels) lvl = mp.nlevels - 1;
        Q.radius = mp.th * mp.scale[lvl];
        Q.min_level = lvl - 1;
        Q.max_level = lvl + 1;
        Q.ur = 0.f;
    } else if (MODE == 5) {
        /* ORBmatcher::Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:968-1131): geo = world position and mfMaxDistance, `oct`
         * = the bit pattern of mfMinDistance, extra = GetNormal().  Float forms as the reference object has them
         * (oracle/match_oracle.cpp: mo_fuse). */
        const float X[3] = {g.x, g.y, g.z};
        const float xc = mt_dot3(mp.Tc, X) + mp.Tc[3], yc = mt_dot3(mp.Tc + 4, X) + mp.Tc[7], zc = mt_dot3(mp.Tc + 8, X) + mp.Tc[11];
        const float invz = 1.0f / zc;
        const float x = xc * invz, y = yc * invz;
        Q.u = fmaf(x, mp.fx, mp.cx);
        Q.v = fmaf(y, mp.fy, mp.cy);
        Q.ok = !(zc < 0.0f) && Q.u >= min_x && Q.u < max_x && Q.v >= min_y && Q.v < max_y; /* KeyFrame::IsInImage */
        Q.ur = fmaf(-mp.mbf, invz, Q.u);
        const float4 nv = l_extra[i];
        const float N[3] = {nv.x, nv.y, nv.z};
        double s2 = 0.0, dotn = 0.0;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float po = X[k] - mp.Ow[k];
            s2 = __dadd_rn(s2, __dmul_rn((double)po, (double)po));
            dotn = __dadd_rn(dotn, __dmul_rn((double)po, (double)N[k]));
        }
        const float dist3D = (float)__dsqrt_rn(s2);
        Q.ok = Q.ok && !(dist3D < 0.8f * __int_as_float(oct) || dist3D > 1.2f * g.w);
        Q.ok = Q.ok && !(dotn < __dmul_rn(0.5, (double)dist3D));
        const float ratio = g.w / dist3D;
        int lvl = orb_cvtt_x86(ceilf(orb_logf_exact(ratio) / mp.log_scale));
        if (lvl < 0) lvl = 0;
        else if (lvl >= mp.nlevels) lvl = mp.nlevels - 1;
        Q.radius = mp.th * mp.scale[lvl];
        Q.min_level = lvl - 1;
        Q.max_level = lvl;
        Q.kind = 1;
        Q.inv_s2[0] = lvl > 0 ? mp.inv_sigma2[lvl - 1] : 0.f;
        Q.inv_s2[1] = mp.inv_sigma2[lvl];
    } else if (MODE == 6) {
        /* Fuse(pKF, Scw, ...) :1133-1260, SearchByProjection(pKF, Scw, ...) :353-485, one direction of SearchBySim3
         * :1262-1501 (oracle/match_oracle.cpp: mo_project_search).  Tc = the first transform, Tl the second one. */
        const float X[3] = {g.x, g.y, g.z};
        float P[3] = {mt_dot3(mp.Tc, X) + mp.Tc[3], mt_dot3(mp.Tc + 4, X) + mp.Tc[7], mt_dot3(mp.Tc + 8, X) + mp.Tc[11]};
        if (mp.two_step) {
            const float P2[3] = {mt_dot3(mp.Tl, P) + mp.Tl[3], mt_dot3(mp.Tl + 4, P) + mp.Tl[7], mt_dot3(mp.Tl + 8, P) + mp.Tl[11]};
            P[0] = P2[0]; P[1] = P2[1]; P[2] = P2[2];
        }
        const float invz = 1.0f / P[2];
        const float x = P[0] * invz, y = P[1] * invz;
        Q.u = fmaf(x, mp.fx, mp.cx);
        Q.v = fmaf(y, mp.fy, mp.cy);
        Q.ok = !(P[2] < 0.0f) && Q.u >= min_x && Q.u < max_x && Q.v >= min_y && Q.v < max_y;
        Q.ur = 0.f;
        double s2 = 0.0, dotn = 0.0;
        float N[3] = {0.f, 0.f, 0.f};
        if (mp.use_normal) { const float4 nv = l_extra[i]; N[0] = nv.x; N[1] = nv.y; N[2] = nv.z; }
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float po = mp.cam_norm ? P[k] : X[k] - mp.Ow[k];
            s2 = __dadd_rn(s2, __dmul_rn((double)po, (double)po));
            dotn = __dadd_rn(dotn, __dmul_rn((double)po, (double)N[k]));
        }
        const float dist3D = (float)__dsqrt_rn(s2);
        Q.ok = Q.ok && !(dist3D < 0.8f * __int_as_float(oct) || dist3D > 1.2f * g.w);
        if (mp.use_normal) Q.ok = Q.ok && !(dotn < __dmul_rn(0.5, (double)dist3D));
        const float ratio = g.w / dist3D;
        int lvl = orb_cvtt_x86(ceilf(orb_logf_exact(ratio) / mp.log_scale));
        if (lvl < 0) lvl = 0;
        else if (lvl >= mp.nlevels) lvl = mp.nlevels - 1;
        Q.radius = mp.th * mp.scale[lvl];
        Q.min_level = lvl - 1;
        Q.max_level = lvl;
        Q.kind = 2;
    } else if (MODE == 3) { /* SearchForInitialization: window of windowSize (= mp.th) around vbPrevMatched, level 0 only (:510) */
        Q.radius = mp.th;
        Q.u = g.x;
        Q.v = g.y;
        Q.ur = 0.f;
        Q.min_level = 0;
        Q.max_level = 0;
        Q.ok = true;
    } else {
        float r = (double)g.w > 0.998 ? 2.5f : 4.0f;
        if (mp.th != 1.0f) r *= mp.th;
        Q.radius = r * mp.scale[oct];
        Q.u = g.x;
        Q.v = g.y;
        Q.ur = g.z;
        Q.min_level = oct - 1;
        Q.max_level = oct;
        Q.ok = true;
    }
    /* Frame::GetFeaturesInArea (src/Frame.cc:446-462) */
    Q.cx0 = max(0, (int)floorf((Q.u - min_x - Q.radius) * winv));
    Q.cx1 = min(MT_COLS - 1, (int)ceilf((Q.u - min_x + Q.radius) * winv));
    Q.cy0 = max(0, (int)floorf((Q.v - min_y - Q.radius) * hinv));
    Q.cy1 = min(MT_ROWS - 1, (int)ceilf((Q.v - min_y + Q.radius) * hinv));
    Q.ok = Q.ok && Q.cx0 < MT_COLS && Q.cx1 >= 0 && Q.cy0 < MT_ROWS && Q.cy1 >= 0;
    Q.check_levels = (Q.min_level > 0) || (Q.max_level >= 0);
    return Q;
}

/* the filters of GetFeaturesInArea (:480-495) and of the candidate loop that do not depend on earlier
 * queries (:1587-1595), then the distance; -1 = rejected */
__device__ __forceinline__ int mt_candidate(const MtQuery &Q, int i2, const BorbKpRec *__restrict__ kp,
                                            const float2 *__restrict__ un_xy, const float *__restrict__ u_right,
                                            const uint8_t *__restrict__ occupied, const uint8_t *__restrict__ desc,
                                            const uint4 d0, const uint4 d1)
{
    if (occupied && occupied[i2] == 2) return -1; /* M2: holds a point with observations on entry (:92-94) */
    const int o2 = kp[i2].octave;
    if (Q.check_levels && (o2 < Q.min_level || (Q.max_level >= 0 && o2 > Q.max_level))) return -1;
    const float2 p = un_xy[i2];
    if (!(fabsf(p.x - Q.u) < Q.radius && fabsf(p.y - Q.v) < Q.radius)) return -1;
    if (Q.kind == 0) {
        if (u_right) {
            const float ur2 = u_right[i2];
            if (ur2 > 0 && fabsf(Q.ur - ur2) > Q.radius) return -1;
        }
    } else if (Q.kind == 1) { /* Fuse: chi-square test of the pixel error, 3 dof for a stereo keypoint (:1056-1083) */
        const float ex = Q.u - p.x, ey = Q.v - p.y;
        const float t = fmaf(ex, ex, ey * ey);
        const float kr = u_right ? u_right[i2] : -1.f;
        const float inv = Q.inv_s2[o2 - Q.min_level];
        if (kr >= 0.f) {
            const float er = Q.ur - kr;
            if ((double)(fmaf(er, er, t) * inv) > 7.8) return -1;
        } else if ((double)(inv * t) > 5.99) return -1;
    }
    return mt_hamming(d0, d1, reinterpret_cast<const uint4 *>(desc) + 2 * i2);
}

/* last frame / local map: geo [4n], octave [n], desc [n] x 32 B (16-byte aligned), flags [n] (bit 0 map point,
 * bit 1 Observations() > 0, bit 2 outlier).  Out: cache [n_last][MT_K] entries (distance << 16 | keypoint
 * index, in ascending (distance, list position q) order; MT_NOKEY = none), more [n_last] = the query had more than MT_K candidates. */
template <int MODE>
__global__ void __launch_bounds__(MT_CAND_NT)
k_match_candidates(const __grid_constant__ MatchParams mp, const BorbKpRec *__restrict__ kp,
                   const uint8_t *__restrict__ desc, const float2 *__restrict__ un_xy,
                   const int *__restrict__ cell_start, const int *__restrict__ cell_idx,
                   const float *__restrict__ bounds, const float *__restrict__ u_right,
                   const uint8_t *__restrict__ occupied, const float *__restrict__ l_geo, const int *__restrict__ l_octave,
                   const uint4 *__restrict__ l_desc, const uint8_t *__restrict__ l_flags,
                   uint32_t *__restrict__ cache, uint8_t *__restrict__ more, const float4 *__restrict__ l_extra = nullptr)
{
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * (MT_CAND_NT / 32) + (threadIdx.x >> 5);
    if (i >= mp.n_last) return; /* whole warp */
    uint32_t T[MT_K];
#pragma unroll
    for (int k = 0; k < MT_K; ++k) T[k] = MT_NOKEY;
    int cnt = 0;
    const uint32_t fl = l_flags[i];
    if ((fl & 1u) && !(fl & 4u)) {
        const float min_x = bounds[0], max_x = bounds[1], min_y = bounds[2], max_y = bounds[3];
        const float winv = (float)MT_COLS / (max_x - min_x), hinv = (float)MT_ROWS / (max_y - min_y);
        const MtQuery Q = mt_query<MODE>(mp, l_geo, l_octave[i], i, min_x, max_x, min_y, max_y, winv, hinv, l_extra);
        if (Q.ok) {
            const uint4 d0 = l_desc[2 * i], d1 = l_desc[2 * i + 1];
            for (int ix = Q.cx0; ix <= Q.cx1; ++ix) {
                const int q0 = cell_start[ix * MT_ROWS + Q.cy0], q1 = cell_start[ix * MT_ROWS + Q.cy1 + 1];
                for (int qb = q0; qb < q1; qb += 32) {
                    const int q = qb + lane;
                    uint32_t key = MT_NOKEY;
                    if (q < q1) {
                        const int d = mt_candidate(Q, cell_idx[q], kp, un_xy, u_right, occupied, desc, d0, d1);
                        if (d >= 0) key = ((uint32_t)d << 16) | (uint32_t)q;
                    }
                    cnt += __popc(__ballot_sync(0xffffffffu, key != MT_NOKEY));
                    while (true) { /* move this chunk's keys that beat the cached worst into the sorted cache */
                        const uint32_t m = __reduce_min_sync(0xffffffffu, key);
                        if (m >= T[MT_K - 1]) break;
                        if (key == m) key = MT_NOKEY;
                        uint32_t carry = m;
#pragma unroll
                        for (int k = 0; k < MT_K; ++k) {
                            const uint32_t lo = min(T[k], carry);
                            carry = max(T[k], carry);
                            T[k] = lo;
                        }
                    }
                }
            }
        }
    }
    if (lane < MT_K) {
        uint32_t v = T[0];
#pragma unroll
        for (int k = 1; k < MT_K; ++k) if (lane == k) v = T[k];
        /* the order is fixed now: keep (distance << 16 | keypoint index), which is what the rounds of M1b read */
        if (v != MT_NOKEY) v = (v & 0xffff0000u) | (uint32_t)cell_idx[v & 0xffffu];
        cache[(size_t)i * MT_K + lane] = v;
    }
    if (lane == 0) more[i] = cnt > MT_K;
}

/* Scratch: choice [n_last], taker [2][kp_total].  Out: match [kp_total] = index of the last-frame map point
 * per current keypoint or -1; nm[0] = nmatches, nm[1] = rounds taken, nm[2] = rescans. */
template <int MODE>
__global__ void __launch_bounds__(MT_NT, 1)
k_match_resolve(const __grid_constant__ MatchParams mp, const BorbKpRec *__restrict__ kp,
                const uint8_t *__restrict__ desc, const float2 *__restrict__ un_xy,
                const int *__restrict__ cell_start, const int *__restrict__ cell_idx,
                const float *__restrict__ bounds, const int *__restrict__ n_cur_p,
                const float *__restrict__ u_right, const uint8_t *__restrict__ occupied,
                const float *__restrict__ l_geo, const float *__restrict__ l_angle, const int *__restrict__ l_octave,
                const uint4 *__restrict__ l_desc, const uint8_t *__restrict__ l_flags,
                const uint32_t *__restrict__ cache, const uint8_t *__restrict__ more,
                int *__restrict__ choice, int *__restrict__ taker_g, int kp_total, int *__restrict__ match,
                int *__restrict__ nm, int taker_in_smem, const float4 *__restrict__ l_extra = nullptr)
{
    /* the two taker arrays live in shared memory when they fit (2 x kp_total ints): a round is then three
     * barriers apart instead of three global-memory round trips */
    extern __shared__ int mt_taker_s[];
    int *taker = taker_in_smem ? mt_taker_s : taker_g;
    __shared__ int s_changed, s_nm, s_hist[MT_HISTO], s_keep[MT_HISTO], s_rescans, s_first;
    const int tid = threadIdx.x;
    const int n_cur = min(*n_cur_p, kp_total), n_last = mp.n_last;
    const float min_x = bounds[0], max_x = bounds[1], min_y = bounds[2], max_y = bounds[3];
    const float winv = (float)MT_COLS / (max_x - min_x), hinv = (float)MT_ROWS / (max_y - min_y);

    if (tid == 0) { s_nm = 0; s_rescans = 0; }
    if (tid < MT_HISTO) s_hist[tid] = 0;
    for (int i = tid; i < n_last; i += MT_NT) choice[i] = -1;
    for (int c = tid; c < kp_total; c += MT_NT) { taker[c] = 0x7fffffff; match[c] = -1; }
    __syncthreads();

    /* what query i chooses when tk[c] = the earliest query (with observations) that holds keypoint c */
    auto choose = [&](int i, const int *tk) -> int {
            int ch = -1;
            /* the query's cached entries and who holds each of them, loaded side by side (two 16-byte loads, then
             * MT_K independent ones) */
            uint32_t keys[MT_K];
            int held_by[MT_K];
            {
                const uint4 k0 = reinterpret_cast<const uint4 *>(cache)[(size_t)i * (MT_K / 4)];
                const uint4 k1 = reinterpret_cast<const uint4 *>(cache)[(size_t)i * (MT_K / 4) + 1];
                keys[0] = k0.x; keys[1] = k0.y; keys[2] = k0.z; keys[3] = k0.w;
                keys[4] = k1.x; keys[5] = k1.y; keys[6] = k1.z; keys[7] = k1.w;
#pragma unroll
                for (int k = 0; k < MT_K; ++k) held_by[k] = keys[k] != MT_NOKEY ? tk[keys[k] & 0xffffu] : 0x7fffffff;
            }
            if (MODE == 1 || MODE == 4 || MODE == 6) {
                /* MODE 6: SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) -- the rule of MODE 4 with TH_LOW (:437-438, :454).
                 * MODE 4: every set slot is closed (:1735-1736; the host sets the "observations" bit of every query) and
                 * the acceptance threshold is the caller's ORBdist (:1750) */
                const uint32_t th_d = MODE >= 4 ? (uint32_t)mp.orb_dist : (uint32_t)MT_TH_HIGH;
                bool all_held = true; /* every cached candidate is held by an earlier query with observations (:1583-1585) */
#pragma unroll
                for (int k = 0; k < MT_K; ++k) {
                    if (!all_held) continue;
                    const uint32_t key = keys[k];
                    if (key == MT_NOKEY || (key >> 16) > th_d) { all_held = false; continue; }
                    if (held_by[k] >= i) { ch = (int)(key & 0xffffu); all_held = false; }
                }
                if (all_held && more[i]) { /* rare: walk the cells again, now with the holders known */
                    atomicAdd(&s_rescans, 1);
                    const MtQuery Q = mt_query<MODE == 2 ? 1 : MODE>(mp, l_geo, l_octave[i], i, min_x, max_x, min_y, max_y, winv, hinv, l_extra);
                    const uint4 d0 = l_desc[2 * i], d1 = l_desc[2 * i + 1];
                    int best = 256, best_idx = -1;
                    for (int ix = Q.cx0; ix <= Q.cx1; ++ix) {
                        const int q0 = cell_start[ix * MT_ROWS + Q.cy0], q1 = cell_start[ix * MT_ROWS + Q.cy1 + 1];
                        for (int q = q0; q < q1; ++q) {
                            const int i2 = cell_idx[q];
                            if (tk[i2] < i) continue;
                            const int d = mt_candidate(Q, i2, kp, un_xy, u_right, occupied, desc, d0, d1);
                            if (d >= 0 && d < best) { best = d; best_idx = i2; }
                        }
                    }
                    ch = best <= (int)th_d ? best_idx : -1;
                }
            } else {
                /* best and second best among the candidates nobody earlier holds (:84-115): the first two free
                 * entries of the cached (distance, scan order) list */
                int found = 0, bd = 256, bl = -1, bd2 = 256, bl2 = -1, bidx = -1;
                bool decided = false;
#pragma unroll
                for (int k = 0; k < MT_K; ++k) {
                    if (decided) continue;
                    const uint32_t key = keys[k];
                    if (key == MT_NOKEY) { decided = true; continue; }
                    const int c = (int)(key & 0xffffu);
                    if (held_by[k] < i) continue;
                    const int d = (int)(key >> 16);
                    if (found == 0) { bd = d; bl = kp[c].octave; bidx = c; found = 1; if (d > MT_TH_HIGH) decided = true; }
                    else { bd2 = d; bl2 = kp[c].octave; found = 2; decided = true; }
                }
                if (!decided && more[i]) { /* fewer than two free candidates among the cached ones and there were more */
                    atomicAdd(&s_rescans, 1);
                    const MtQuery Q = mt_query<2>(mp, l_geo, l_octave[i], i, min_x, max_x, min_y, max_y, winv, hinv);
                    const uint4 d0 = l_desc[2 * i], d1 = l_desc[2 * i + 1];
                    bd = 256; bl = -1; bd2 = 256; bl2 = -1; bidx = -1;
                    for (int ix = Q.cx0; ix <= Q.cx1; ++ix) {
                        const int q0 = cell_start[ix * MT_ROWS + Q.cy0], q1 = cell_start[ix * MT_ROWS + Q.cy1 + 1];
                        for (int q = q0; q < q1; ++q) {
                            const int i2 = cell_idx[q];
                            if (tk[i2] < i) continue;
                            const int d = mt_candidate(Q, i2, kp, un_xy, u_right, occupied, desc, d0, d1);
                            if (d < 0) continue;
                            if (d < bd) { bd2 = bd; bd = d; bl2 = bl; bl = kp[i2].octave; bidx = i2; }
                            else if (d < bd2) { bl2 = kp[i2].octave; bd2 = d; }
                        }
                    }
                }
                if (bd <= MT_TH_HIGH && !(bl == bl2 && (float)bd > mp.nnratio * (float)bd2)) ch = bidx;
            }
            return ch;
    };

    /* Rounds of the fixed point.  After round r every query below the smallest index whose choice changed in round r
     * is final (a choice depends on earlier queries only), so the rounds needed equal the longest chain of displaced
     * queries: 2-10 on real frames, but up to n_last on adversarial input (many map points projecting onto the same few
     * keypoints).  After MT_MAX_ROUNDS the still-open suffix is finished by ONE thread walking it in the reference's own
     * order -- exact by construction, slow only where the input is pathological. */
    int round = 0;
    bool settled = false;
    const int max_rounds = mp.max_rounds > 0 ? mp.max_rounds : MT_MAX_ROUNDS;
    for (; round < max_rounds; ++round) {
        const int *tk_prev = taker + (round & 1) * kp_total;
        int *tk_new = taker + ((round + 1) & 1) * kp_total;
        for (int c = tid; c < n_cur; c += MT_NT) tk_new[c] = 0x7fffffff;
        if (tid == 0) { s_changed = 0; s_first = 0x7fffffff; }
        __syncthreads();
        for (int i = tid; i < n_last; i += MT_NT) {
            const int ch = choose(i, tk_prev);
            if (ch != choice[i]) { choice[i] = ch; s_changed = 1; atomicMin(&s_first, i); }
            if (ch >= 0 && (l_flags[i] & 2u)) atomicMin(&tk_new[ch], i);
        }
        __syncthreads();
        const int changed = s_changed;
        __syncthreads();
        if (!changed) { settled = true; break; }
    }
    if (!settled) {
        /* queries below `first` are final; rebuild the takers from them alone, then walk the rest in order */
        const int first = s_first;
        int *tk = taker + (round & 1) * kp_total; /* the array the last round filled: overwritten here */
        __syncthreads();
        for (int c = tid; c < n_cur; c += MT_NT) tk[c] = 0x7fffffff;
        __syncthreads();
        for (int i = tid; i < first && i < n_last; i += MT_NT) {
            const int ch = choice[i];
            if (ch >= 0 && (l_flags[i] & 2u)) atomicMin(&tk[ch], i);
        }
        __syncthreads();
        if (tid == 0) {
            for (int i = max(first, 0); i < n_last; ++i) {
                const int ch = choose(i, tk);
                choice[i] = ch;
                if (ch >= 0 && (l_flags[i] & 2u) && tk[ch] > i) tk[ch] = i;
            }
        }
        __syncthreads();
    }

    /* ---- assignment, orientation histogram, three maxima ---- */
    for (int i = tid; i < n_last; i += MT_NT) {
        const int ch = choice[i];
        if (ch >= 0) {
            atomicAdd(&s_nm, 1);
            atomicMax(&match[ch], i);
            if ((MODE == 1 || MODE == 4) && mp.check_ori) atomicAdd(&s_hist[mt_bin(l_angle[i], kp[ch].angle)], 1);
        }
    }
    __syncthreads();
    if ((MODE == 1 || MODE == 4) && mp.check_ori) {
        if (tid == 0) {
            int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
            for (int b = 0; b < MT_HISTO; ++b) {
                const int s = s_hist[b];
                if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = b; }
                else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = b; }
                else if (s > max3) { max3 = s; ind3 = b; }
            }
            if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
            else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
            for (int b = 0; b < MT_HISTO; ++b) s_keep[b] = (b == ind1 || b == ind2 || b == ind3);
        }
        __syncthreads();
        for (int i = tid; i < n_last; i += MT_NT) {
            const int ch = choice[i];
            if (ch >= 0 && !s_keep[mt_bin(l_angle[i], kp[ch].angle)]) {
                match[ch] = -1;
                atomicSub(&s_nm, 1);
            }
        }
        __syncthreads();
    }
    if (tid == 0) { nm[0] = s_nm; nm[1] = settled ? round + 1 : max_rounds + 1; nm[2] = s_rescans; } /* max_rounds + 1: the sequential finish ran */
}

/* ---------------------------------------------------------------------------------------------
 * M3: ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)
 *     /root/reference/src/ORBmatcher.cc:487-632, the matcher of Tracking::MonocularInitialization.
 * Candidates by k_match_candidates<3>.  The sequential rule here: a match on an F2 keypoint that is
 * already matched needs a strictly smaller distance (vMatchedDistance[i2] <= dist skips it, :526-527) and
 * then takes the keypoint over (:552-558).  With a(i) = the accepted (keypoint, distance) of query i,
 * candidate (c, d) is open to query i iff no earlier query i' < i has a(i') = (c, d') with d' <= d; a(i)
 * depends on earlier queries only, so the rounds below (all queries choose given the previous round's
 * chooser lists) reach the unique sequential solution.  At the end keypoint c belongs to its LAST
 * accepted chooser, the orientation histogram counts every accepted query (also the ones taken over
 * later, :562-577), and entries of discarded bins are cleared only where they still hold (:606-613).
 * The choosers of a keypoint are a linked list threaded through the queries themselves: head [2][kp_total] = the
 * query pushed last (-1 = none), node [2][n] = (next query + 1) << 8 | accepted distance.  A query holds at most
 * one keypoint per round, so a list can have any length (on repetitive texture with a 100-px window dozens of
 * level-0 keypoints may pick the same one; the reference never fails there) and nothing can overflow.
 */
__device__ __forceinline__ bool mt_init_closed(const uint32_t *__restrict__ node, const int *__restrict__ head, int c, int i, int d)
{
    for (int k = head[c]; k >= 0;) {
        const uint32_t e = node[k];
        if (k < i && (int)(e & 0xffu) <= d) return true;
        k = (int)(e >> 8) - 1;
    }
    return false;
}

__global__ void __launch_bounds__(MT_NT, 1)
k_match_resolve_init(const __grid_constant__ MatchParams mp, const BorbKpRec *__restrict__ kp,
                     const uint8_t *__restrict__ desc, const float2 *__restrict__ un_xy,
                     const int *__restrict__ cell_start, const int *__restrict__ cell_idx,
                     const float *__restrict__ bounds, const float *__restrict__ l_geo,
                     const float *__restrict__ l_angle, const int *__restrict__ l_octave,
                     const uint4 *__restrict__ l_desc, const uint8_t *__restrict__ l_flags,
                     const uint32_t *__restrict__ cache, const uint8_t *__restrict__ more,
                     int *__restrict__ choice /* [n][2]: keypoint, distance */, uint32_t *__restrict__ node /* [2][node_cap] */,
                     int node_cap, int *__restrict__ head /* [2][kp_total] */, int kp_total, int *__restrict__ owner /* [kp_total] */,
                     int *__restrict__ m12 /* [n] */, int *__restrict__ nm)
{
    __shared__ int s_changed, s_nm, s_hist[MT_HISTO], s_keep[MT_HISTO], s_rescans;
    const int tid = threadIdx.x, n1 = mp.n_last;
    const float min_x = bounds[0], max_x = bounds[1], min_y = bounds[2], max_y = bounds[3];
    const float winv = (float)MT_COLS / (max_x - min_x), hinv = (float)MT_ROWS / (max_y - min_y);
    if (tid == 0) { s_nm = 0; s_rescans = 0; }
    if (tid < MT_HISTO) { s_hist[tid] = 0; s_keep[tid] = 1; }
    for (int i = tid; i < n1; i += MT_NT) { choice[2 * i] = -1; choice[2 * i + 1] = 0; }
    for (int c = tid; c < kp_total; c += MT_NT) { head[c] = -1; owner[c] = -1; }
    __syncthreads();

    int round = 0;
    for (;; ++round) {
        const uint32_t *node_prev = node + (size_t)(round & 1) * node_cap;
        uint32_t *node_new = node + (size_t)((round + 1) & 1) * node_cap;
        const int *head_prev = head + (round & 1) * kp_total;
        int *head_new = head + ((round + 1) & 1) * kp_total;
        for (int c = tid; c < kp_total; c += MT_NT) head_new[c] = -1;
        if (tid == 0) s_changed = 0;
        __syncthreads();
        for (int i = tid; i < n1; i += MT_NT) {
            int found = 0, bd = 0x7fffffff, bd2 = 0x7fffffff, bidx = -1;
            bool decided = !(l_flags[i] & 1u);
#pragma unroll
            for (int k = 0; k < MT_K; ++k) {
                if (decided) continue;
                const uint32_t key = cache[(size_t)i * MT_K + k];
                if (key == MT_NOKEY) { decided = true; continue; }
                const int c = (int)(key & 0xffffu), d = (int)(key >> 16);
                if (found == 0 && d > MT_TH_LOW) { decided = true; continue; } /* the best open one can only be worse */
                if (mt_init_closed(node_prev, head_prev, c, i, d)) continue;
                if (found == 0) { bd = d; bidx = c; found = 1; }
                else { bd2 = d; found = 2; decided = true; }
            }
            if (!decided && more[i]) { /* fewer than two open candidates among the cached ones and there were more */
                atomicAdd(&s_rescans, 1);
                const MtQuery Q = mt_query<3>(mp, l_geo, 0, i, min_x, max_x, min_y, max_y, winv, hinv);
                const uint4 d0 = l_desc[2 * i], d1 = l_desc[2 * i + 1];
                bd = 0x7fffffff; bd2 = 0x7fffffff; bidx = -1;
                if (Q.ok)
                    for (int ix = Q.cx0; ix <= Q.cx1; ++ix) {
                        const int q0 = cell_start[ix * MT_ROWS + Q.cy0], q1 = cell_start[ix * MT_ROWS + Q.cy1 + 1];
                        for (int q = q0; q < q1; ++q) {
                            const int i2 = cell_idx[q];
                            const int d = mt_candidate(Q, i2, kp, un_xy, nullptr, nullptr, desc, d0, d1);
                            if (d < 0 || mt_init_closed(node_prev, head_prev, i2, i, d)) continue;
                            if (d < bd) { bd2 = bd; bd = d; bidx = i2; }
                            else if (d < bd2) bd2 = d;
                        }
                    }
            }
            int ch = -1, chd = 0;
            if (bidx >= 0 && bd <= MT_TH_LOW && (float)bd < (float)bd2 * mp.nnratio) { ch = bidx; chd = bd; }
            if (ch != choice[2 * i] || chd != choice[2 * i + 1]) { choice[2 * i] = ch; choice[2 * i + 1] = chd; s_changed = 1; }
            if (ch >= 0) node_new[i] = ((uint32_t)(atomicExch(&head_new[ch], i) + 1) << 8) | (uint32_t)chd; /* push */
        }
        __syncthreads();
        const int changed = s_changed;
        __syncthreads();
        if (!changed) break;
    }

    /* ---- ownership (the last accepted chooser), orientation histogram over every accepted query ---- */
    for (int i = tid; i < n1; i += MT_NT) {
        const int ch = choice[2 * i];
        if (ch >= 0) {
            atomicMax(&owner[ch], i);
            if (mp.check_ori) atomicAdd(&s_hist[mt_bin(l_angle[i], kp[ch].angle)], 1);
        }
    }
    __syncthreads();
    if (mp.check_ori) {
        if (tid == 0) {
            int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
            for (int b = 0; b < MT_HISTO; ++b) {
                const int s = s_hist[b];
                if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = b; }
                else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = b; }
                else if (s > max3) { max3 = s; ind3 = b; }
            }
            if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
            else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
            for (int b = 0; b < MT_HISTO; ++b) s_keep[b] = (b == ind1 || b == ind2 || b == ind3);
        }
        __syncthreads();
    }
    for (int i = tid; i < n1; i += MT_NT) {
        const int ch = choice[2 * i];
        int out = -1;
        if (ch >= 0 && owner[ch] == i && (!mp.check_ori || s_keep[mt_bin(l_angle[i], kp[ch].angle)])) {
            out = ch;
            atomicAdd(&s_nm, 1);
        }
        m12[i] = out;
    }
    __syncthreads();
    if (tid == 0) { nm[0] = s_nm; nm[1] = round + 1; nm[2] = s_rescans; nm[3] = 0; }
}

/* tap for tests/test_logf.py: the device copy of orb_logf_exact, in place */
__global__ void k_debug_logf(float *x, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) x[i] = orb_logf_exact(x[i]);
}

#endif
