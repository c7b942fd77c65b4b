/*
 * oracle/orb_oracle.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.  See orb_oracle.h.
 *
 * A restatement, on flat arrays and index-linked lists, of what the reference's ORB extractor
 * computes.  Every stage cites the reference lines (relative to /root/reference) it follows.
 * It is deliberately NOT a transliteration: no cv::Mat, no std::list<ExtractorNode>, no
 * KeyPoint copies during quadtree splits -- nodes hold candidate indices.
 *
 * Build with -ffp-contract=off.  The one place where the reference binary contracts
 * floating-point (the descriptor rotation, src/ORBextractor.cc:128-130, built with
 * -O3 -march=native on an FMA host) is written with explicit fmaf() in the contracted form the
 * compiled reference object actually contains (checked by disassembly).
 */
#include "orb_oracle.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <utility>
#include <vector>

#include "cvprim.h"

namespace {

const int kPatch = 31;      /* PATCH_SIZE       src/ORBextractor.cc:72 */
const int kHalfPatch = 15;  /* HALF_PATCH_SIZE  :73 */
const int kEdge = 19;       /* EDGE_THRESHOLD   :74 */

const signed char kPattern[1024] = {
#include "../orb_slam2_test_b200/csrc/orb_pattern.inc"
};

struct Plane {
    int w = 0, h = 0;
    size_t step = 0;
    std::vector<uint8_t> buf;
    uint8_t *at0 = nullptr; /* first interior pixel */
};

struct LevelKp {
    int x, y;     /* level pixel coordinates */
    int score;
    float angle;
};

/* ---------------------------------------------------------------- quadtree (:537-614, :668-951) */

struct QNode {
    int x0, y0, x1, y1;
    std::vector<int> pts; /* candidate indices, in parent order */
    int prev, next;
};

struct QList {
    std::vector<QNode> pool; /* pool index == creation order */
    int head = -1, tail = -1, count = 0;

    int push_front(QNode &&n)
    {
        const int id = (int)pool.size();
        n.prev = -1;
        n.next = head;
        pool.push_back(std::move(n));
        if (head >= 0) pool[head].prev = id; else tail = id;
        head = id;
        ++count;
        return id;
    }
    int push_back(QNode &&n)
    {
        const int id = (int)pool.size();
        n.next = -1;
        n.prev = tail;
        pool.push_back(std::move(n));
        if (tail >= 0) pool[tail].next = id; else head = id;
        tail = id;
        ++count;
        return id;
    }
    void unlink(int id)
    {
        QNode &n = pool[id];
        if (n.prev >= 0) pool[n.prev].next = n.next; else head = n.next;
        if (n.next >= 0) pool[n.next].prev = n.prev; else tail = n.prev;
        std::vector<int>().swap(n.pts);
        --count;
    }
};

/* quarter a node (:537-614): the cut lines sit at ceil(extent/2) from the upper-left corner,
 * membership is decided by comparing the point with the cut lines only (points are not
 * guaranteed to lie inside their node's box, see root assignment below) */
void quarter(const QNode &n, const int *xs, const int *ys, QNode out[4])
{
    const int hx = (int)std::ceil((float)(n.x1 - n.x0) / 2);
    const int hy = (int)std::ceil((float)(n.y1 - n.y0) / 2);
    const int mx = n.x0 + hx, my = n.y0 + hy;
    out[0].x0 = n.x0; out[0].y0 = n.y0; out[0].x1 = mx;   out[0].y1 = my;   /* upper left  */
    out[1].x0 = mx;   out[1].y0 = n.y0; out[1].x1 = n.x1; out[1].y1 = my;   /* upper right */
    out[2].x0 = n.x0; out[2].y0 = my;   out[2].x1 = mx;   out[2].y1 = n.y1; /* lower left  */
    out[3].x0 = mx;   out[3].y0 = my;   out[3].x1 = n.x1; out[3].y1 = n.y1; /* lower right */
    for (int q = 0; q < 4; ++q) out[q].pts.clear();
    for (int id : n.pts) {
        const int right = !((float)xs[id] < (float)mx);
        const int low = !((float)ys[id] < (float)my);
        out[low * 2 + right].pts.push_back(id);
    }
}

/* returns survivors (candidate indices) in the reference's output order, or false if nIni==0 */
bool distribute(const int *xs, const int *ys, const int *resp, int n, int min_x, int max_x,
                int min_y, int max_y, int n_target, std::vector<int> &out)
{
    out.clear();
    /* roots (:674-739) */
    const int n_ini = (int)std::round((float)(max_x - min_x) / (float)(max_y - min_y));
    if (n_ini <= 0) return false;
    const float hX = (float)(max_x - min_x) / (float)n_ini;

    QList L;
    L.pool.reserve((size_t)std::max(64, 8 * n_target));
    std::vector<int> root_id((size_t)n_ini);
    for (int i = 0; i < n_ini; ++i) {
        QNode r;
        r.x0 = (int)(hX * (float)i);
        r.x1 = (int)(hX * (float)(i + 1));
        r.y0 = 0;
        r.y1 = max_y - min_y;
        root_id[i] = L.push_back(std::move(r));
    }
    for (int i = 0; i < n; ++i) {
        const int r = (int)((float)xs[i] / hX); /* float division, not the integer box */
        if (r < 0 || r >= n_ini) return false;  /* the reference would index out of bounds */
        L.pool[root_id[r]].pts.push_back(i);
    }
    for (int id = L.head; id >= 0;) {
        const int nx = L.pool[id].next;
        if (L.pool[id].pts.empty()) L.unlink(id);
        id = nx;
    }

    /* expandable = nodes created by the most recent pass that hold more than one point,
     * in creation order (:745-746, vSizeAndPointerToNode) */
    std::vector<std::pair<int, int>> expandable; /* (point count, pool id) */
    QNode kids[4];
    bool done = false;
    while (!done) {
        /* breadth pass (:751-843): every multi-point node is replaced by its non-empty
         * quarters, which go to the FRONT of the list and are therefore not revisited */
        const int before = L.count;
        expandable.clear();
        for (int id = L.head; id >= 0;) {
            const int nx = L.pool[id].next;
            if (L.pool[id].pts.size() > 1) {
                quarter(L.pool[id], xs, ys, kids);
                for (int q = 0; q < 4; ++q) {
                    if (kids[q].pts.empty()) continue;
                    const int sz = (int)kids[q].pts.size();
                    const int kid = L.push_front(std::move(kids[q]));
                    if (sz > 1) expandable.push_back(std::make_pair(sz, kid));
                }
                L.unlink(id);
            }
            id = nx;
        }
        const int n_expand = (int)expandable.size();
        if (L.count >= n_target || L.count == before) {
            done = true; /* :849-852 */
        } else if (L.count + 3 * n_expand > n_target) {
            /* ordered phase (:853-924): biggest nodes first, stop the moment the target is hit.
             * Ties in point count: the reference orders by node ADDRESS; canonical rule = by
             * creation order, later-created first (pool id is creation order). */
            while (!done) {
                const int before2 = L.count;
                std::vector<std::pair<int, int>> todo = expandable;
                expandable.clear();
                std::sort(todo.begin(), todo.end());
                for (int j = (int)todo.size() - 1; j >= 0; --j) {
                    const int id = todo[j].second;
                    quarter(L.pool[id], xs, ys, kids);
                    for (int q = 0; q < 4; ++q) {
                        if (kids[q].pts.empty()) continue;
                        const int sz = (int)kids[q].pts.size();
                        const int kid = L.push_front(std::move(kids[q]));
                        if (sz > 1) expandable.push_back(std::make_pair(sz, kid));
                    }
                    L.unlink(id);
                    if (L.count >= n_target) break;
                }
                if (L.count >= n_target || L.count == before2) done = true;
            }
        }
    }

    /* one survivor per node: highest response, first in node order on ties (:930-948) */
    out.reserve((size_t)L.count);
    for (int id = L.head; id >= 0; id = L.pool[id].next) {
        const std::vector<int> &p = L.pool[id].pts;
        int best = p[0];
        for (size_t k = 1; k < p.size(); ++k)
            if ((float)resp[p[k]] > (float)resp[best]) best = p[k];
        out.push_back(best);
    }
    return true;
}

/* ---------------------------------------------------------------- orientation (:83-111) */

const int kUmax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

float ic_angle(const uint8_t *c, size_t step, const int *umax)
{
    int m01 = 0, m10 = 0;
    for (int u = -kHalfPatch; u <= kHalfPatch; ++u) m10 += u * c[u];
    for (int v = 1; v <= kHalfPatch; ++v) {
        const uint8_t *lo = c + (ptrdiff_t)v * (ptrdiff_t)step;
        const uint8_t *up = c - (ptrdiff_t)v * (ptrdiff_t)step;
        int vsum = 0;
        const int d = umax[v];
        for (int u = -d; u <= d; ++u) {
            const int a = lo[u], b = up[u];
            vsum += a - b;
            m10 += u * (a + b);
        }
        m01 += v * vsum;
    }
    return cvp_fast_atan2((float)m01, (float)m10);
}

/* ---------------------------------------------------------------- descriptor (:117-157) */

void describe(const uint8_t *c, size_t step, float angle_deg, uint8_t *desc)
{
    const float factor_pi = (float)(3.1415926535897932384626433832795 / 180.f); /* :114 */
    const float ang = angle_deg * factor_pi;
    float a, b; /* a = cos, b = sin; the compiled reference makes one sincosf call */
    sincosf(ang, &b, &a);
    const int istep = (int)step;
    const signed char *p = kPattern;
    for (int i = 0; i < 32; ++i, p += 32) {
        int val = 0;
        for (int k = 0; k < 8; ++k) {
            int t[2];
            for (int e = 0; e < 2; ++e) {
                const float px = (float)p[4 * k + 2 * e], py = (float)p[4 * k + 2 * e + 1];
                /* contracted form found in the reference object: fma(x,b,rn(y*a)), fma(x,a,-rn(y*b)) */
                const int row = cvp_round_f(fmaf(px, b, py * a));
                const int col = cvp_round_f(fmaf(px, a, -(py * b)));
                t[e] = c[row * istep + col];
            }
            val |= (t[0] < t[1]) << k;
        }
        desc[i] = (uint8_t)val;
    }
}

/* ---------------------------------------------------------------- extractor state */

struct Oracle {
    int nfeatures, nlevels, ini_th, min_th;
    double scale_factor; /* the reference member is a double (include/ORBextractor.h:128) */
    std::vector<float> scale, inv_scale, sigma2, inv_sigma2;
    std::vector<int> quota;
    int umax[16];

    std::vector<Plane> pyr, blur;
    std::vector<std::vector<orbo_cand>> cands;
    std::vector<std::vector<LevelKp>> kps;

    Oracle(int nf, float sf, int nl, int it, int mt)
        : nfeatures(nf), nlevels(nl), ini_th(it), min_th(mt), scale_factor(sf)
    {
        /* scale tables (:437-457) */
        scale.assign(nl, 1.f);
        sigma2.assign(nl, 1.f);
        for (int i = 1; i < nl; ++i) {
            scale[i] = (float)(scale[i - 1] * scale_factor);
            sigma2[i] = scale[i] * scale[i];
        }
        inv_scale.resize(nl);
        inv_sigma2.resize(nl);
        for (int i = 0; i < nl; ++i) {
            inv_scale[i] = 1.0f / scale[i];
            inv_sigma2[i] = 1.0f / sigma2[i];
        }
        /* per-level quotas: geometric series, remainder to the last level (:465-476) */
        quota.assign(nl, 0);
        const float factor = (float)(1.0f / scale_factor);
        float want = nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nl));
        int sum = 0;
        for (int l = 0; l < nl - 1; ++l) {
            quota[l] = cvp_round_f(want);
            sum += quota[l];
            want *= factor;
        }
        quota[nl - 1] = std::max(nfeatures - sum, 0);
        /* row half-widths of the 31-px disc (:504-520) */
        const int vmax = (int)std::floor((double)(kHalfPatch * std::sqrt(2.f) / 2 + 1));
        const int vmin = (int)std::ceil((double)(kHalfPatch * std::sqrt(2.f) / 2));
        const double hp2 = kHalfPatch * kHalfPatch;
        for (int v = 0; v <= vmax; ++v) umax[v] = cvp_round_d(std::sqrt(hp2 - v * v));
        for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
            while (umax[v0] == umax[v0 + 1]) ++v0;
            umax[v] = v0;
            ++v0;
        }
    }

    /* :1400-1443 */
    void build_pyramid(const uint8_t *img, int cols, int rows, size_t step)
    {
        pyr.assign(nlevels, Plane());
        for (int l = 0; l < nlevels; ++l) {
            Plane &P = pyr[l];
            P.w = cvp_round_f((float)cols * inv_scale[l]);
            P.h = cvp_round_f((float)rows * inv_scale[l]);
            P.step = (size_t)(P.w + 2 * kEdge);
            P.buf.assign(P.step * (size_t)(P.h + 2 * kEdge), 0);
            P.at0 = P.buf.data() + (size_t)kEdge * P.step + kEdge;
            if (l == 0) {
                for (int y = 0; y < rows; ++y) std::memcpy(P.at0 + (size_t)y * P.step, img + (size_t)y * step, (size_t)cols);
            } else {
                const Plane &S = pyr[l - 1];
                cvp_resize_linear_u8(S.at0, S.step, S.w, S.h, P.at0, P.step, P.w, P.h);
            }
            cvp_border_reflect101(P.buf.data(), P.step, P.w, P.h, kEdge);
        }
    }

    /* per-cell FAST with the threshold fallback (:966-1094); false on degenerate grids */
    bool detect(int l)
    {
        const Plane &P = pyr[l];
        std::vector<orbo_cand> &out = cands[l];
        out.clear();
        const int min_bx = kEdge - 3, min_by = min_bx;
        const int max_bx = P.w - kEdge + 3, max_by = P.h - kEdge + 3;
        const float width = (float)(max_bx - min_bx), height = (float)(max_by - min_by);
        const int n_cols = (int)(width / 30.f), n_rows = (int)(height / 30.f);
        if (n_cols <= 0 || n_rows <= 0) return false;
        const int w_cell = (int)std::ceil(width / n_cols), h_cell = (int)std::ceil(height / n_rows);
        std::vector<cvp_kp> cell((size_t)(((w_cell + 7) / 2) * ((h_cell + 7) / 2)));
        for (int i = 0; i < n_rows; ++i) {
            const float ini_y = (float)(min_by + i * h_cell);
            float max_y = ini_y + h_cell + 6;
            if (ini_y >= max_by - 3) continue;
            if (max_y > max_by) max_y = (float)max_by;
            for (int j = 0; j < n_cols; ++j) {
                const float ini_x = (float)(min_bx + j * w_cell);
                float max_x = ini_x + w_cell + 6;
                if (ini_x >= max_bx - 6) continue;
                if (max_x > max_bx) max_x = (float)max_bx;
                const uint8_t *win = P.at0 + (size_t)(int)ini_y * P.step + (int)ini_x;
                const int ww = (int)max_x - (int)ini_x, wh = (int)max_y - (int)ini_y;
                int n = cvp_fast9_16(win, P.step, ww, wh, ini_th, cell.data(), (int)cell.size());
                if (n == 0) n = cvp_fast9_16(win, P.step, ww, wh, min_th, cell.data(), (int)cell.size());
                for (int k = 0; k < n; ++k) {
                    orbo_cand c;
                    c.x = cell[k].x + j * w_cell;
                    c.y = cell[k].y + i * h_cell;
                    c.score = cell[k].score;
                    out.push_back(c);
                }
            }
        }
        return true;
    }

    int run(const uint8_t *img, int cols, int rows, size_t step, orbo_kp *okp, uint8_t *odesc, int cap)
    {
        if (!img || cols <= 0 || rows <= 0) return -1; /* :1333-1334 */
        build_pyramid(img, cols, rows, step);
        cands.assign(nlevels, std::vector<orbo_cand>());
        kps.assign(nlevels, std::vector<LevelKp>());
        blur.assign(nlevels, Plane());
        std::vector<int> xs, ys, rs, keep;
        for (int l = 0; l < nlevels; ++l) {
            if (!detect(l)) return -2;
            const std::vector<orbo_cand> &c = cands[l];
            xs.resize(c.size()); ys.resize(c.size()); rs.resize(c.size());
            for (size_t i = 0; i < c.size(); ++i) { xs[i] = c[i].x; ys[i] = c[i].y; rs[i] = c[i].score; }
            const Plane &P = pyr[l];
            const int min_b = kEdge - 3;
            if (!distribute(xs.data(), ys.data(), rs.data(), (int)c.size(), min_b, P.w - kEdge + 3,
                            min_b, P.h - kEdge + 3, quota[l], keep))
                return -2;
            kps[l].reserve(keep.size());
            for (int id : keep) { /* :1112-1122 */
                LevelKp k;
                k.x = c[id].x + min_b;
                k.y = c[id].y + min_b;
                k.score = c[id].score;
                k.angle = -1.f;
                kps[l].push_back(k);
            }
        }
        for (int l = 0; l < nlevels; ++l) /* :1126-1127 */
            for (LevelKp &k : kps[l]) k.angle = ic_angle(pyr[l].at0 + (size_t)k.y * pyr[l].step + k.x, pyr[l].step, umax);

        int total = 0;
        for (int l = 0; l < nlevels; ++l) { /* :1366-1396 */
            if (kps[l].empty()) continue;
            const Plane &P = pyr[l];
            Plane &B = blur[l];
            B.w = P.w; B.h = P.h; B.step = (size_t)P.w;
            B.buf.resize((size_t)P.w * (size_t)P.h);
            B.at0 = B.buf.data();
            cvp_gaussian7_sigma2(P.at0, P.step, B.at0, B.step, P.w, P.h);
            const float s = scale[l];
            const float patch = (float)(int)(kPatch * s);
            for (const LevelKp &k : kps[l]) {
                if (total < cap) {
                    describe(B.at0 + (size_t)k.y * B.step + k.x, B.step, k.angle, odesc + (size_t)total * 32);
                    orbo_kp &o = okp[total];
                    o.x = (float)k.x; o.y = (float)k.y;
                    if (l != 0) { o.x = o.x * s; o.y = o.y * s; }
                    o.size = patch;
                    o.angle = k.angle;
                    o.response = (float)k.score;
                    o.octave = l;
                    o.class_id = -1;
                }
                ++total;
            }
        }
        return total;
    }
};

} // namespace

extern "C" {

void *orbo_create(int nf, float sf, int nl, int it, int mt) { return new Oracle(nf, sf, nl, it, mt); }
void orbo_destroy(void *h) { delete (Oracle *)h; }

int orbo_extract(void *h, const uint8_t *img, int cols, int rows, size_t step, orbo_kp *kps,
                 uint8_t *desc, int cap)
{
    return ((Oracle *)h)->run(img, cols, rows, step, kps, desc, cap);
}

int orbo_levels(void *h) { return ((Oracle *)h)->nlevels; }
void orbo_table(void *h, int which, float *out)
{
    Oracle *o = (Oracle *)h;
    const std::vector<float> &v = which == 0 ? o->scale : which == 1 ? o->inv_scale : which == 2 ? o->sigma2 : o->inv_sigma2;
    std::memcpy(out, v.data(), v.size() * sizeof(float));
}
void orbo_quotas(void *h, int *out)
{
    Oracle *o = (Oracle *)h;
    std::memcpy(out, o->quota.data(), o->quota.size() * sizeof(int));
}
void orbo_umax(void *h, int *out) { std::memcpy(out, ((Oracle *)h)->umax, sizeof(int) * 16); }

int orbo_level_size(void *h, int level, int *cols, int *rows)
{
    Oracle *o = (Oracle *)h;
    if (level < 0 || level >= (int)o->pyr.size()) return -1;
    *cols = o->pyr[level].w;
    *rows = o->pyr[level].h;
    return 0;
}
int orbo_pyramid_copy(void *h, int level, int border, uint8_t *out)
{
    Oracle *o = (Oracle *)h;
    if (level < 0 || level >= (int)o->pyr.size() || border > kEdge) return -1;
    const Plane &P = o->pyr[level];
    const int W = P.w + 2 * border, H = P.h + 2 * border;
    const uint8_t *base = P.at0 - (size_t)border * P.step - border;
    for (int y = 0; y < H; ++y) std::memcpy(out + (size_t)y * W, base + (size_t)y * P.step, (size_t)W);
    return 0;
}
int orbo_blurred_copy(void *h, int level, uint8_t *out)
{
    Oracle *o = (Oracle *)h;
    if (level < 0 || level >= (int)o->blur.size() || o->blur[level].buf.empty()) return -1;
    std::memcpy(out, o->blur[level].buf.data(), o->blur[level].buf.size());
    return 0;
}
int orbo_candidates(void *h, int level, orbo_cand *out, int cap)
{
    Oracle *o = (Oracle *)h;
    if (level < 0 || level >= (int)o->cands.size()) return -1;
    const std::vector<orbo_cand> &c = o->cands[level];
    for (size_t i = 0; i < c.size() && (int)i < cap; ++i) out[i] = c[i];
    return (int)c.size();
}
int orbo_level_keypoints(void *h, int level, orbo_kp *out, int cap)
{
    Oracle *o = (Oracle *)h;
    if (level < 0 || level >= (int)o->kps.size()) return -1;
    const std::vector<LevelKp> &k = o->kps[level];
    for (size_t i = 0; i < k.size() && (int)i < cap; ++i) {
        out[i].x = (float)k[i].x; out[i].y = (float)k[i].y;
        out[i].size = (float)(int)(kPatch * o->scale[level]);
        out[i].angle = k[i].angle;
        out[i].response = (float)k[i].score;
        out[i].octave = level;
        out[i].class_id = -1;
    }
    return (int)k.size();
}

int orbo_distribute(const int *xs, const int *ys, const int *resp, int n, int min_x, int max_x,
                    int min_y, int max_y, int n_target, int *out_idx, int cap)
{
    std::vector<int> keep;
    if (!distribute(xs, ys, resp, n, min_x, max_x, min_y, max_y, n_target, keep)) return -2;
    for (size_t i = 0; i < keep.size() && (int)i < cap; ++i) out_idx[i] = keep[i];
    return (int)keep.size();
}

void orbo_describe(const uint8_t *img, size_t step, int x, int y, float angle_deg, uint8_t *desc32)
{
    describe(img + (size_t)y * step + x, step, angle_deg, desc32);
}
float orbo_ic_angle(const uint8_t *img, size_t step, int x, int y)
{
    return ic_angle(img + (size_t)y * step + x, step, kUmax);
}
const signed char *orbo_pattern(void) { return kPattern; }

} // extern "C"
