/*
 * borb_api.cu -- host side of liborb_b200.so: the C ABI declared in include/borb.h.
 *
 * Per handle: one CUDA stream, one device "plan" per image geometry (level sizes, resize tap
 * tables, FAST cell list, blur tile list, quadtree parameters -- everything that the reference
 * recomputes per call from (cols, rows) in ComputePyramid / ComputeKeyPointsOctTree,
 * src/ORBextractor.cc:1400-1443, :966-996), device slabs for max_batch frames, pinned staging,
 * and a cache of instantiated CUDA graphs (one per call shape) so a call is one graph launch.
 *
 * No CPU fallback: if the device work cannot run the call fails with BORB_ERR_CUDA.
 */
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/borb.h"
#include "borb_internal.h"
#include "k_blur.cuh"
#include "k_describe.cuh"
#include "k_fast.cuh"
#include "k_frame.cuh"
#include "k_prestep.cuh"
#include "k_pyramid.cuh"
#include "k_quadtree.cuh"
#include "k_stereo.cuh"
#include "k_match.cuh"

static_assert(sizeof(borb_keypoint) == 28 && sizeof(BorbKpRec) == 28, "cv::KeyPoint layout");

namespace {

thread_local std::string g_err;

int fail(int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CU(call)                                                                          \
    do {                                                                                  \
        cudaError_t e_ = (call);                                                          \
        if (e_ != cudaSuccess)                                                            \
            return fail(BORB_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_),   \
                        __FILE__, __LINE__);                                              \
    } while (0)

const signed char kPattern[1024] = {
#include "orb_pattern.inc"
};

inline int round_half_even(float v) { return (int)lrintf(v); }
inline int round_half_even_d(double v) { return (int)lrint(v); }
inline int floor_f(float v)
{
    int i = (int)v;
    return i - (v < (float)i);
}
inline short sat16(int v) { return (short)(v < -32768 ? -32768 : v > 32767 ? 32767 : v); }
inline long long align_up(long long v, long long a) { return (v + a - 1) / a * a; }

struct GraphKey {
    int mode;            /* 0 host path, 1 device path */
    const void *src;
    int n;
    size_t step, stride;
    unsigned flags;
    int scols, srows; /* source image size (differs from the plan's when a rectification map is set) */
    bool operator==(const GraphKey &o) const
    {
        return mode == o.mode && src == o.src && n == o.n && step == o.step && stride == o.stride &&
               flags == o.flags && scols == o.scols && srows == o.srows;
    }
};
struct GraphEntry {
    GraphKey key;
    cudaGraphExec_t exec;
};

} // namespace

struct BowState; /* borb_bow.inl */
static void bow_release(BowState *&b);

struct borb_extractor {
    /* constructor state, reference src/ORBextractor.cc:432-521 */
    int nfeatures, nlevels, ini_th, min_th;
    double scale_factor;
    std::vector<float> scale, inv_scale, sigma2, inv_sigma2;
    std::vector<int> quota;
    int umax[16];

    int device, max_batch;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;       /* around the last device call */
    cudaEvent_t evt0 = nullptr, evt1 = nullptr;     /* user timer */
    cudaEvent_t evs[8] = {};                        /* stage boundaries */
    bool pending = false;                           /* submitted, not yet waited for */
    unsigned pending_flags = 0;
    bool use_graph = true;
    bool stage_timing = false;                      /* record stage events inside every call */

    /* geometry-dependent state */
    bool have_plan = false;
    BorbPlan plan;
    int qt_M = 0;
    int fast_smem4 = 0; /* K2 with four warps per cell (a frame or two per call) */
    int fast_smem = 0, n_sm = 148;
    int fast_wp = 64;                   /* staged window pitch of K2 = TMA box width */
    FastMaps fast_maps;                 /* one TMA descriptor per level, passed by value to K2 */
    BlurMaps blur_maps;                 /* the same planes with K5's box */
    DescMaps desc_maps;                 /* unblurred + blurred planes with K4/K6's boxes, and the plan fields they read */
    int chunk = 0; /* frames per pass of the kernel chain (0 = whole call) */
    /* side stream for the latency shape: K5 (blur) needs only the pyramid, so for a frame or two it runs beside
     * K2 + K3 (fork / join by events; inside a capture this becomes a parallel graph branch) */
    cudaStream_t stream2 = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    int pyr_sp[BORB_MAXL] = {}, pyr_srows[BORB_MAXL] = {}; /* staged source tile of K1: pitch, rows */
    int pyr_yinv_off[BORB_MAXL] = {};  /* k_pyr_stream: per source row the output row that ends there (uint2 pairs inside tab_i) */
    bool pyr_stream[BORB_MAXL] = {};   /* the level meets k_pyr_stream's preconditions */
    /* single-launch level chain of the latency shape (k_pyr_chain): per (tile, level) rectangles, shared-memory layout */
    PyrChainTile *d_chain = nullptr;
    uint2 *d_chain_tabs = nullptr;
    int chain_tstride = 0;
    int chain_tiles = 0, chain_off_b = 0, chain_off_t = 0, chain_smem = 0;
    std::vector<BorbCell> cells;
    std::vector<BorbTile> tiles;
    BorbPlan *d_plan = nullptr;
    BorbCell *d_cells = nullptr;
    BorbTile *d_tiles = nullptr;
    int *d_tab_i = nullptr;
    short *d_tab_s = nullptr;
    signed char *d_pattern = nullptr;
    uint8_t *d_in = nullptr, *d_pyr = nullptr, *d_blur = nullptr;
    uint32_t *d_cand_xy = nullptr, *d_cand_score = nullptr, *d_cand_node = nullptr;
    int *d_ncand = nullptr, *d_nlkp = nullptr, *d_nout = nullptr, *d_err = nullptr;
    uint32_t *d_lkp_xy = nullptr, *d_lkp_score = nullptr;
    /* one result slab on each side -- [records][descriptors][counts][error word] -- so that a call that fills the
     * handle (n == max_batch, the single-frame case) brings everything back with ONE copy; d_out_kp, d_out_desc,
     * d_nout, d_err and their h_ twins point into it */
    uint8_t *d_res = nullptr, *h_res = nullptr;
    uint8_t *h_res_dev = nullptr; /* device alias of h_res (k_results_to_host), null if the platform has none */
    size_t res_bytes = 0, res_o_desc = 0, res_o_nout = 0, res_o_err = 0;
    BorbKpRec *d_out_kp = nullptr;
    uint8_t *d_out_desc = nullptr;
    /* pinned staging */
    uint8_t *h_in = nullptr, *h_pyr = nullptr, *h_out_desc = nullptr;
    BorbKpRec *h_out_kp = nullptr;
    int *h_nout = nullptr, *h_err = nullptr;
    /* the step before the extractor (K0): channel layout of the images handed in, rectification map */
    int in_format = 0, in_ch = 1;
    int map_cols = 0, map_rows = 0, map_pitch = 0;
    uint32_t *d_map_xy = nullptr;
    uint16_t *d_map_fr = nullptr;
    uint8_t *d_gray = nullptr; /* [B][rows][gray_pitch]: K0's output, K1's input */
    int gray_pitch = 0;
    size_t in_cap = 0;         /* bytes per frame d_in / h_in can hold */
    /* the per-keypoint step after the extractor (K8): undistortion + 64x48 grid, when a camera is set */
    bool have_cam = false;
    BorbCamera cam;
    uint8_t *d_fr = nullptr, *h_fr = nullptr; /* one slab for K8's outputs, like d_res / h_res: [un_xy][cell_start][cell_idx][bounds] */
    size_t fr_bytes = 0, fr_o_cs = 0, fr_o_ci = 0, fr_o_bd = 0;
    float2 *d_un_xy = nullptr, *h_un_xy = nullptr;
    int *d_cell_start = nullptr, *d_cell_idx = nullptr, *d_cell_of = nullptr, *h_cell_start = nullptr, *h_cell_idx = nullptr;
    float *d_bounds = nullptr, *h_bounds = nullptr;
    bool frame_on_host = false;
    /* stereo association (left handle only), allocated on first use */
    float *d_st_u = nullptr, *d_st_z = nullptr, *h_st_u = nullptr, *h_st_z = nullptr;
    int *d_st_sad = nullptr, *d_st_kept = nullptr, *h_st_kept = nullptr;
    int *d_st_rows = nullptr;      /* [B][rows+1] CSR starts of the right keypoints listed per image row */
    uint16_t *d_st_ent = nullptr;  /* [B][kp_total * st_span] */
    int st_span = 0;
    cudaEvent_t ev_peer = nullptr; /* recorded on this handle's stream for a peer handle to wait on */
    bool stereo_on_host = false;
    unsigned long long extract_gen = 0; /* bumped by every extract call */
    unsigned long long stereo_gen = 0;  /* extract_gen the last borb_stereo_match saw (0 = none) */
    /* frame-to-frame matcher (M1), allocated on first use: the last frame's packed arrays
     * [desc 32c][xyz 12c][angle 4c][octave 4c][flags c] for a capacity of c = mt_cap queries */
    unsigned char *d_mt_last = nullptr, *h_mt_last = nullptr;
    size_t mt_cap = 0;
    int *d_mt_choice = nullptr, *d_mt_taker = nullptr, *d_mt_match = nullptr, *h_mt_match = nullptr; /* match[K] + nm[4] */
    uint32_t *d_mt_cache = nullptr; /* [mt_cap][MT_K] candidate keys */
    uint8_t *d_mt_more = nullptr, *d_mt_occ = nullptr, *h_mt_occ = nullptr; /* occupancy of the local-map matcher */
    int mt_rescans = 0;
    /* monocular-initialisation matcher (M3): chooser lists: nodes [2][mt_cap] + heads [2][K], result [mt_cap] + F2's mvKeysUn */
    uint32_t *d_mt_ent = nullptr;
    int *d_mt_cnt = nullptr, *d_mt_m12 = nullptr, *h_mt_m12 = nullptr;
    float *h_mt_unxy = nullptr;
    BowState *bow = nullptr; /* bag of words of one frame + SearchByBoW staging (borb_bow.inl) */
    int mt_rounds = 0;

    std::vector<GraphEntry> graphs;
    int graph_miss_streak = 0;
    bool have_last_miss = false;
    GraphKey last_miss = {};
    int last_n = 0;          /* frames of the last call */
    int last_launches = 0;   /* kernels the last call's shape enqueues */
    bool results_on_host = false, pyramid_on_host = false;
    int nout_frame_on_host = -1; /* a matcher call has downloaded this frame's keypoint count since the last extract call */
    float last_ms = -1.f;
};

namespace {

/* ---- constructor tables: src/ORBextractor.cc:437-520, same float/double mix ---- */
void build_tables(borb_extractor *h)
{
    const int nl = h->nlevels;
    h->scale.assign(nl, 1.f);
    h->sigma2.assign(nl, 1.f);
    for (int i = 1; i < nl; ++i) {
        h->scale[i] = (float)(h->scale[i - 1] * h->scale_factor); /* member is a double */
        h->sigma2[i] = h->scale[i] * h->scale[i];
    }
    h->inv_scale.resize(nl);
    h->inv_sigma2.resize(nl);
    for (int i = 0; i < nl; ++i) {
        h->inv_scale[i] = 1.0f / h->scale[i];
        h->inv_sigma2[i] = 1.0f / h->sigma2[i];
    }
    h->quota.assign(nl, 0);
    const float factor = (float)(1.0f / h->scale_factor);
    float want = h->nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nl));
    int sum = 0;
    for (int l = 0; l < nl - 1; ++l) {
        h->quota[l] = round_half_even(want);
        sum += h->quota[l];
        want *= factor;
    }
    h->quota[nl - 1] = std::max(h->nfeatures - sum, 0);
    /* half-widths of the rows of the 31-px orientation disc */
    const int half = 15;
    const int vmax = (int)std::floor((double)(half * std::sqrt(2.f) / 2 + 1));
    const int vmin = (int)std::ceil((double)(half * std::sqrt(2.f) / 2));
    const double hp2 = half * half;
    for (int v = 0; v <= vmax; ++v) h->umax[v] = round_half_even_d(std::sqrt(hp2 - v * v));
    for (int v = half, v0 = 0; v >= vmin; --v) {
        while (h->umax[v0] == h->umax[v0 + 1]) ++v0;
        h->umax[v] = v0;
        ++v0;
    }
}

/* taps of cv::resize INTER_LINEAR for one axis.  OpenCV clamps the horizontal
 * taps in the table but clips the vertical ROW INDEX at run time; clamp_taps selects which. */
void resize_axis(int sn, int dn, bool clamp_taps, int *ofs, short *coef)
{
    const double scale = 1.0 / ((double)dn / (double)sn);
    for (int d = 0; d < dn; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = floor_f(f);
        f -= (float)s;
        if (clamp_taps) {
            if (s < 0) { f = 0.f; s = 0; }
            if (s >= sn - 1) { f = 0.f; s = sn - 1; }
        }
        ofs[d] = s;
        coef[2 * d] = sat16(round_half_even((1.f - f) * 2048.f));
        coef[2 * d + 1] = sat16(round_half_even(f * 2048.f));
    }
}

void free_geometry(borb_extractor *h)
{
    for (GraphEntry &g : h->graphs) cudaGraphExecDestroy(g.exec);
    h->graphs.clear();
    void *dev[] = {h->d_plan, h->d_cells, h->d_tiles, h->d_tab_i, h->d_tab_s, h->d_in, h->d_pyr,
                   h->d_blur, h->d_cand_xy, h->d_cand_score, h->d_cand_node, h->d_ncand, h->d_nlkp,
                   h->d_res, h->d_lkp_xy, h->d_lkp_score,
                   h->d_st_u, h->d_st_z, h->d_st_sad, h->d_st_kept, h->d_st_rows, h->d_st_ent, h->d_gray,
                   h->d_fr, h->d_cell_of,
                   h->d_chain, h->d_chain_tabs, h->d_mt_last, h->d_mt_choice, h->d_mt_taker, h->d_mt_match, h->d_mt_cache, h->d_mt_more, h->d_mt_occ, h->d_mt_ent, h->d_mt_cnt, h->d_mt_m12};
    for (void *p : dev) if (p) cudaFree(p);
    void *host[] = {h->h_in, h->h_pyr, h->h_res,
                    h->h_st_u, h->h_st_z, h->h_st_kept, h->h_fr,
                    h->h_mt_last, h->h_mt_match, h->h_mt_occ, h->h_mt_m12, h->h_mt_unxy};
    for (void *p : host) if (p) cudaFreeHost(p);
    bow_release(h->bow);
    h->d_mt_last = h->h_mt_last = nullptr;
    h->mt_cap = 0;
    h->d_mt_choice = h->d_mt_taker = h->d_mt_match = h->h_mt_match = nullptr;
    h->d_mt_cache = nullptr; h->d_mt_more = nullptr; h->d_mt_occ = h->h_mt_occ = nullptr;
    h->d_mt_ent = nullptr; h->d_mt_cnt = h->d_mt_m12 = h->h_mt_m12 = nullptr; h->h_mt_unxy = nullptr;
    h->d_gray = nullptr;
    h->d_chain = nullptr; h->d_chain_tabs = nullptr; h->chain_tiles = 0;
    h->in_cap = 0;
    h->d_un_xy = h->h_un_xy = nullptr;
    h->d_fr = h->h_fr = nullptr; h->fr_bytes = 0;
    h->d_cell_start = h->d_cell_idx = h->d_cell_of = h->h_cell_start = h->h_cell_idx = nullptr;
    h->d_bounds = h->h_bounds = nullptr;
    h->frame_on_host = false;
    h->d_st_u = h->d_st_z = h->h_st_u = h->h_st_z = nullptr;
    h->d_st_sad = h->d_st_kept = h->h_st_kept = nullptr;
    h->d_st_rows = nullptr;
    h->d_st_ent = nullptr;
    h->stereo_on_host = false;
    h->d_plan = nullptr; h->d_cells = nullptr; h->d_tiles = nullptr; h->d_tab_i = nullptr;
    h->d_tab_s = nullptr; h->d_in = h->d_pyr = h->d_blur = nullptr;
    h->d_cand_xy = h->d_cand_score = h->d_cand_node = nullptr;
    h->d_ncand = h->d_nlkp = h->d_nout = h->d_err = nullptr;
    h->d_lkp_xy = h->d_lkp_score = nullptr; h->d_out_kp = nullptr; h->d_out_desc = nullptr;
    h->h_in = h->h_pyr = h->h_out_desc = nullptr; h->h_out_kp = nullptr;
    h->h_nout = h->h_err = nullptr;
    h->d_res = h->h_res = nullptr; h->h_res_dev = nullptr; h->res_bytes = 0;
    h->have_plan = false;
    h->results_on_host = h->pyramid_on_host = false;
    h->nout_frame_on_host = -1;
    h->last_n = 0;
}

/* Rectangles of k_pyr_chain: the levels are cut by one gx x gy grid (x cuts on multiples of 4 so that a tile's
 * rows are whole words of the padded plane); from the last level back, a tile needs its own rectangle, the
 * interior pixels its part of the reflect-101 frame mirrors, and whatever the taps of its next level's rectangle
 * reach (cv::resize's clamped x taps, run-time clipped y taps: resize_axis above). */
int build_chain(borb_extractor *h, const BorbPlan &P, const std::vector<int> &tab_i, const std::vector<short> &tab_s)
{
    const int nl = P.nlevels;
    const BorbLevel &Z = P.lv[nl - 1];
    const char *etw = getenv("BORB_PCH_TW"), *eth = getenv("BORB_PCH_TH"); /* tuning knobs */
    const int tw = etw ? std::max(4, atoi(etw)) : PCH_TW, th = eth ? std::max(2, atoi(eth)) : PCH_TH;
    const int gx = std::max(1, Z.w / tw), gy = std::max(1, Z.h / th);
    std::vector<PyrChainTile> tiles((size_t)gx * gy);
    std::memset(tiles.data(), 0, tiles.size() * sizeof(PyrChainTile));
    /* The cuts are not uniform: a tile on the image edge also writes its part of the reflect-101 frame at every level
     * (19 rows / 19 + alignment columns whatever the level's size: 57 % more pixels than a middle tile at 1241 x 376), and
     * the launch takes as long as its slowest tile (tools/probe_chain.py: 13.0 us against a median tile's 9.0), so the
     * edge rows and columns of tiles are cut narrower by these weights. */
    const char *ewx = getenv("BORB_PCH_WX"), *ewy = getenv("BORB_PCH_WY"); /* tuning knobs */
    const double wx = ewx ? atof(ewx) : PCH_EDGE_WX, wy = ewy ? atof(ewy) : PCH_EDGE_WY;
    std::vector<double> fx(gx + 1, 0.0), fy(gy + 1, 0.0);
    for (int i = 0; i < gx; ++i) fx[i + 1] = fx[i] + ((i == 0 || i == gx - 1) && gx > 2 ? wx : 1.0);
    for (int i = 0; i < gy; ++i) fy[i + 1] = fy[i] + ((i == 0 || i == gy - 1) && gy > 2 ? wy : 1.0);
    auto cut = [](const std::vector<double> &f, int i, int n) { return (int)((double)n * f[i] / f.back()); };
    long long max_a = 0, max_b = 0, max_t = 0;
    for (int ty = 0; ty < gy; ++ty)
        for (int tx = 0; tx < gx; ++tx) {
            PyrChainRect *R = tiles[(size_t)ty * gx + tx].r;
            int nx0 = 0, ny0 = 0, nx1 = 0, ny1 = 0; /* need(l + 1) */
            long long t = 0;
            for (int l = nl - 1; l >= 0; --l) {
                const BorbLevel &L = P.lv[l];
                const int ox0 = tx == 0 ? 0 : cut(fx, tx, L.w) & ~3;
                const int ox1 = tx == gx - 1 ? L.w : cut(fx, tx + 1, L.w) & ~3;
                const int oy0 = ty == 0 ? 0 : cut(fy, ty, L.h), oy1 = ty == gy - 1 ? L.h : cut(fy, ty + 1, L.h);
                int x0 = 1 << 30, y0 = 1 << 30, x1 = -1, y1 = -1; /* need(l), exclusive upper bounds */
                const bool own = ox1 > ox0 && oy1 > oy0;
                if (own) {
                    x0 = ox0; x1 = ox1; y0 = oy0; y1 = oy1;
                    if (ox0 == 0) x1 = std::max(x1, std::min(BORB_EDGE, L.w - 1) + 1);
                    if (ox1 == L.w) x0 = std::min(x0, std::max(0, L.w - 1 - BORB_EDGE));
                    if (oy0 == 0) y1 = std::max(y1, std::min(BORB_EDGE, L.h - 1) + 1);
                    if (oy1 == L.h) y0 = std::min(y0, std::max(0, L.h - 1 - BORB_EDGE));
                }
                if (l < nl - 1 && nx1 > nx0 && ny1 > ny0) {
                    const BorbLevel &D = P.lv[l + 1];
                    const int *xo = &tab_i[D.xofs_off], *yo = &tab_i[D.yofs_off];
                    auto cy = [&](int v) { return std::min(std::max(v, 0), L.h - 1); };
                    x0 = std::min(x0, xo[nx0]);
                    x1 = std::max(x1, std::min(xo[nx1 - 1] + 1, L.w - 1) + 1);
                    y0 = std::min(y0, cy(yo[ny0]));
                    y1 = std::max(y1, cy(yo[ny1 - 1] + 1) + 1);
                }
                PyrChainRect &r = R[l];
                if (x1 > x0 && y1 > y0) {
                    r.nx0 = (short)x0; r.ny0 = (short)y0; r.nw = (short)(x1 - x0); r.nh = (short)(y1 - y0);
                } else {
                    r.nx0 = r.ny0 = r.nw = r.nh = 0;
                    x0 = y0 = x1 = y1 = 0;
                }
                r.ox0 = (short)ox0; r.oy0 = (short)oy0; r.ox1 = (short)(own ? ox1 : ox0); r.oy1 = (short)(own ? oy1 : oy0);
                const long long bytes = (long long)pch_pitch(r.nw) * r.nh + 16; /* + slack: the x tap pair of the last pixel */
                if (l & 1) max_b = std::max(max_b, bytes); else max_a = std::max(max_a, bytes);
                if (l > 0) t += (long long)(pch_pitch(r.nw) + r.nh) * 8;
                nx0 = x0; ny0 = y0; nx1 = x1; ny1 = y1;
            }
            max_t = std::max(max_t, t);
        }
    h->chain_tiles = 0;
    max_t = align_up(max_t, 16);
    const long long off_b = align_up(max_a, 16), off_t = align_up(off_b + max_b, 16), total = off_t + max_t;
    if (total > 160 * 1024) return BORB_OK; /* tiles of this geometry do not fit: the per-level kernels serve every shape */
    /* per tile: table offsets, loop divisors, and the tap tables themselves relative to the source rectangles, so that
     * the kernel's prologue is plain copies with no dependent loads.  x part, three entries per OUTPUT WORD (four pixels,
     * the row padded to a multiple of 4 with copies of the last pixel): the four taps of pixels 0, 1 lie within two
     * aligned words of the staged source row and so do those of pixels 2, 3 (scale factors up to 3), so an entry holds
     * {word A | word B << 16, PRMT selector A | selector B << 16} -- a selector gathers x0, x1 of one pixel into bytes
     * 0, 1 and of the next into bytes 2, 3 -- then {a0 | a1 << 16} of pixels 0, 1 and of pixels 2, 3 (dp2a operands).
     * Then the y entries {sy0 | sy1 << 16, b0 | b1 << 16}. */
    const size_t tstride = (size_t)max_t / 8;
    std::vector<uint2> tabs(tiles.size() * tstride, make_uint2(0u, 0u));
    auto magic = [](int w) { return w > 1 ? 0xffffffffu / (uint32_t)w + 1u : 0u; };
    bool chain_ok = true;
    for (size_t t = 0; t < tiles.size(); ++t) {
        PyrChainTile &T = tiles[t];
        uint2 *tab = &tabs[t * tstride];
        int off = 0;
        for (int l = 0; l < nl; ++l) {
            const BorbLevel &L = P.lv[l];
            const PyrChainRect &r = T.r[l];
            T.toff[l] = off;
            const int ex0 = r.ox0 == 0 ? -BORB_XPAD : r.ox0, ex1 = r.ox1 == L.w ? L.pitch - BORB_XPAD : r.ox1;
            T.magic[l][0] = magic(pch_pitch(r.nw) >> 2);
            T.magic[l][1] = magic((ex1 - ex0) >> 2);
            if (l == 0 || r.nw == 0) continue;
            const BorbLevel &S = P.lv[l - 1];
            const PyrChainRect &q = T.r[l - 1];
            const int nx = pch_pitch(r.nw), nxe = 3 * (nx / 4);
            for (int w = 0; w < nx / 4; ++w) {
                int x0[4], x1[4];
                uint32_t wgt[4];
                for (int k = 0; k < 4; ++k) {
                    const int x = r.nx0 + std::min(4 * w + k, r.nw - 1), sx = tab_i[L.xofs_off + x];
                    x0[k] = sx - q.nx0;
                    x1[k] = std::min(sx + 1, S.w - 1) - q.nx0;
                    wgt[k] = (uint32_t)(uint16_t)tab_s[L.xa_off + 2 * x] | ((uint32_t)(uint16_t)tab_s[L.xa_off + 2 * x + 1] << 16);
                }
                uint32_t base[2], sel[2];
                for (int pr = 0; pr < 2; ++pr) {
                    const int k = 2 * pr, lo = std::min(std::min(x0[k], x1[k]), std::min(x0[k + 1], x1[k + 1]));
                    const int hi = std::max(std::max(x0[k], x1[k]), std::max(x0[k + 1], x1[k + 1]));
                    const int b = lo >> 2;
                    if (lo < 0 || hi - 4 * b > 7) chain_ok = false; /* a scale factor above 3: the per-level kernels serve it */
                    base[pr] = (uint32_t)b;
                    sel[pr] = (uint32_t)(x0[k] - 4 * b) | ((uint32_t)(x1[k] - 4 * b) << 4) | ((uint32_t)(x0[k + 1] - 4 * b) << 8) |
                              ((uint32_t)(x1[k + 1] - 4 * b) << 12);
                }
                tab[off + 3 * w].x = base[0] | (base[1] << 16);
                tab[off + 3 * w].y = (sel[0] & 0xffffu) | ((sel[1] & 0xffffu) << 16);
                tab[off + 3 * w + 1].x = wgt[0]; tab[off + 3 * w + 1].y = wgt[1];
                tab[off + 3 * w + 2].x = wgt[2]; tab[off + 3 * w + 2].y = wgt[3];
            }
            for (int j = 0; j < r.nh; ++j) {
                const int y = r.ny0 + j, sy = tab_i[L.yofs_off + y];
                const int sy0 = std::min(std::max(sy, 0), S.h - 1), sy1 = std::min(std::max(sy + 1, 0), S.h - 1);
                tab[off + nxe + j].x = (uint32_t)(sy0 - q.ny0) | ((uint32_t)(sy1 - q.ny0) << 16);
                tab[off + nxe + j].y = (uint32_t)(uint16_t)tab_s[L.yb_off + 2 * y] | ((uint32_t)(uint16_t)tab_s[L.yb_off + 2 * y + 1] << 16);
            }
            off += nxe + r.nh;
        }
        T.toff[nl] = off;
    }
    if (!chain_ok) return BORB_OK; /* chain_tiles stays 0 */
    CU(cudaMalloc(&h->d_chain, tiles.size() * sizeof(PyrChainTile)));
    CU(cudaMemcpy(h->d_chain, tiles.data(), tiles.size() * sizeof(PyrChainTile), cudaMemcpyHostToDevice));
    CU(cudaMalloc(&h->d_chain_tabs, tabs.size() * sizeof(uint2)));
    CU(cudaMemcpy(h->d_chain_tabs, tabs.data(), tabs.size() * sizeof(uint2), cudaMemcpyHostToDevice));
    CU(cudaFuncSetAttribute(k_pyr_chain, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)total));
    h->chain_tstride = (int)tstride;
    h->chain_tiles = gx * gy;
    h->chain_off_b = (int)off_b; h->chain_off_t = (int)off_t; h->chain_smem = (int)total;
    return BORB_OK;
}

/* everything that depends on (cols, rows) */
int build_plan(borb_extractor *h, int cols, int rows)
{
    if (h->have_plan && h->plan.cols == cols && h->plan.rows == rows) return BORB_OK;
    CU(cudaStreamSynchronize(h->stream));
    free_geometry(h);

    BorbPlan P;
    std::memset(&P, 0, sizeof P);
    P.nlevels = h->nlevels;
    P.cols = cols; P.rows = rows;
    P.ini_th = h->ini_th; P.min_th = h->min_th;
    std::memcpy(P.umax, h->umax, sizeof P.umax);
    std::vector<int> tab_i;
    std::vector<short> tab_s;
    std::vector<BorbCell> cells;
    std::vector<BorbTile> tiles;
    long long pyr_off = 0, cand_off = 0;
    int kp_off = 0, M = 0;
    for (int l = 0; l < h->nlevels; ++l) {
        BorbLevel &L = P.lv[l];
        /* level size from the ORIGINAL size, src/ORBextractor.cc:1404-1406 */
        L.w = round_half_even((float)cols * h->inv_scale[l]);
        L.h = round_half_even((float)rows * h->inv_scale[l]);
        if (L.w > 32767 || L.h > 32767)
            return fail(BORB_ERR_GEOMETRY, "level %d is %dx%d: coordinates are packed in 16 bits", l, L.w, L.h);
        /* + 3: K1 writes whole words from column -20, the last one may end 3 columns past the frame */
        L.pitch = (int)align_up(BORB_XPAD + L.w + BORB_EDGE + 3, 32);
        L.prows = L.h + 2 * BORB_EDGE;
        L.off = pyr_off;
        pyr_off = align_up(pyr_off + (long long)L.pitch * L.prows, 256);
        L.scale = h->scale[l];
        L.inv_scale = h->inv_scale[l];
        L.patch_size = (float)(int)(31 * h->scale[l]);
        L.quota = h->quota[l];
        if (l > 0) {
            const BorbLevel &S = P.lv[l - 1];
            L.xofs_off = (int)tab_i.size(); tab_i.resize(tab_i.size() + L.w);
            L.yofs_off = (int)tab_i.size(); tab_i.resize(tab_i.size() + L.h);
            L.xa_off = (int)tab_s.size(); tab_s.resize(tab_s.size() + 2 * L.w);
            L.yb_off = (int)tab_s.size(); tab_s.resize(tab_s.size() + 2 * L.h);
            resize_axis(S.w, L.w, true, &tab_i[L.xofs_off], &tab_s[L.xa_off]);
            resize_axis(S.h, L.h, false, &tab_i[L.yofs_off], &tab_s[L.yb_off]);
            /* K1 stages, per 128 x 16 output tile, the source span its taps reach: bound both spans
             * over all tile positions (reflected frame pixels only re-use interior taps) */
            int span_x = 0, span_y = 0;
            for (int x0 = 0; x0 < L.w; ++x0) {
                const int x1 = std::min(x0 + PYR_TW - 1, L.w - 1);
                const int lo = tab_i[L.xofs_off + x0] & ~15, hi = std::min(tab_i[L.xofs_off + x1] + 1, S.w - 1);
                span_x = std::max(span_x, hi - lo + 1);
            }
            for (int y0 = 0; y0 < L.h; ++y0) {
                const int y1 = std::min(y0 + PYR_TH - 1, L.h - 1);
                const int lo = std::max(tab_i[L.yofs_off + y0], 0), hi = std::min(tab_i[L.yofs_off + y1] + 1, S.h - 1);
                span_y = std::max(span_y, hi - lo + 1);
            }
            h->pyr_sp[l] = (int)align_up(span_x + 16, 16);
            h->pyr_srows[l] = span_y + 1;
            {   /* k_pyr_stream (source-row driven): an output row d must blend source rows yofs[d], yofs[d] + 1 with yofs
                 * strictly increasing and inside the source (true for every level scale > 1), and the staged window must
                 * fit the kernel's fixed pitch; else the level keeps k_pyr_resize */
                bool ok = h->pyr_sp[l] <= PYR_SP_MAX && L.h >= 2 && S.h >= 2;
                /* ... and the taps x0, x0 + 1 of the two pixels of every lane's pixel pair (an aligned group of four padded
                 * columns from -PYR_LPAD on, reflected into the image) must lie within two aligned words */
                auto refl = [&](int i) { if (i < 0) i = -i; if (i >= L.w) i = 2 * (L.w - 1) - i; return i; };
                for (int c = -PYR_LPAD; c < L.w + BORB_EDGE && ok; c += 2) {
                    const int ta = tab_i[L.xofs_off + refl(std::min(c, L.w - 1 + BORB_EDGE))];
                    const int tb = tab_i[L.xofs_off + refl(std::min(c + 1, L.w - 1 + BORB_EDGE))];
                    ok = std::max(ta, tb) + 1 - (std::min(ta, tb) & ~3) <= 7;
                }
                for (int d = 0; d < L.h && ok; ++d) {
                    const int sy = tab_i[L.yofs_off + d];
                    ok = sy >= 0 && sy + 1 <= S.h - 1 && (d == 0 || sy > tab_i[L.yofs_off + d - 1]);
                }
                h->pyr_stream[l] = ok;
                if (tab_i.size() & 1) tab_i.push_back(0); /* uint2 alignment */
                h->pyr_yinv_off[l] = (int)tab_i.size();
                tab_i.resize(tab_i.size() + 2 * (size_t)S.h, -1);
                if (ok)
                    for (int d = 0; d < L.h; ++d) {
                        const int s1 = tab_i[L.yofs_off + d] + 1;
                        tab_i[h->pyr_yinv_off[l] + 2 * s1] = d;
                        tab_i[h->pyr_yinv_off[l] + 2 * s1 + 1] = (int)((uint32_t)(uint16_t)tab_s[L.yb_off + 2 * d] |
                                                                       ((uint32_t)(uint16_t)tab_s[L.yb_off + 2 * d + 1] << 16));
                    }
            }
            if ((long long)2 * h->pyr_sp[l] * h->pyr_srows[l] > 48 * 1024)
                return fail(BORB_ERR_GEOMETRY, "scale factor too large for the pyramid tile at level %d", l);
        }
        /* FAST cell grid, src/ORBextractor.cc:974-1064 (float arithmetic as written there) */
        const int min_b = BORB_FAST_MARGIN;
        L.max_bx = L.w - BORB_EDGE + 3;
        L.max_by = L.h - BORB_EDGE + 3;
        const float width = (float)(L.max_bx - min_b), height = (float)(L.max_by - min_b);
        L.n_cols = (int)(width / 30.f);
        L.n_rows = (int)(height / 30.f);
        if (L.n_cols <= 0 || L.n_rows <= 0)
            return fail(BORB_ERR_GEOMETRY, "level %d (%dx%d) is smaller than one 30-px FAST cell", l, L.w, L.h);
        L.w_cell = (int)std::ceil(width / L.n_cols);
        L.h_cell = (int)std::ceil(height / L.n_rows);
        if (L.w_cell > BORB_MAXCELL || L.h_cell > BORB_MAXCELL || L.n_cols > 4095 || L.n_rows > 4095)
            return fail(BORB_ERR_GEOMETRY, "cell grid out of range at level %d", l);
        long long cap = 0;
        for (int i = 0; i < L.n_rows; ++i) {
            const float ini_y = (float)(min_b + i * L.h_cell);
            float max_y = ini_y + L.h_cell + 6;
            if (ini_y >= L.max_by - 3) continue;
            if (max_y > L.max_by) max_y = (float)L.max_by;
            for (int j = 0; j < L.n_cols; ++j) {
                const float ini_x = (float)(min_b + j * L.w_cell);
                float max_x = ini_x + L.w_cell + 6;
                if (ini_x >= L.max_bx - 6) continue;
                if (max_x > L.max_bx) max_x = (float)L.max_bx;
                BorbCell c;
                c.level = (short)l; c.pad = 0;
                c.x0 = (short)(int)ini_x; c.y0 = (short)(int)ini_y;
                c.ww = (short)((int)max_x - (int)ini_x);
                c.wh = (short)((int)max_y - (int)ini_y);
                c.sx = (short)(j * L.w_cell); c.sy = (short)(i * L.h_cell);
                if (c.ww < 7 || c.wh < 7) continue; /* cv::FAST finds nothing in such a window */
                c.wpr = ((c.x0 & 3) + c.ww + 3) >> 2;
                {   /* stage-1 geometry of K2: Q words cover a domain row, R = rows per warp step */
                    const int b0 = FAST_DOMAIN_BYTE(c.x0), dw = c.ww - 6;
                    const int Q = ((b0 + dw - 1) >> 2) - (b0 >> 2) + 1;
                    c.dmagic = (unsigned)Q | ((unsigned)std::min(8, 32 / Q) << 8);
                }
                c.wmagic = fast_div_magic(c.wpr);
                c.pad2 = 0;
                cells.push_back(c);
                P.fast_max_wh = std::max(P.fast_max_wh, (int)c.wh);
                P.fast_max_ww = std::max(P.fast_max_ww, (int)c.ww);
                P.fast_max_pix = std::max(P.fast_max_pix, (c.ww - 6) * (c.wh - 6));
                cap += (long long)((c.ww - 6 + 1) / 2) * ((c.wh - 6 + 1) / 2);
            }
        }
        L.cand_cap = (int)std::max<long long>(cap, 1);
        L.cand_off = cand_off;
        cand_off += align_up(L.cand_cap, 32);
        /* quadtree roots, src/ORBextractor.cc:674-680 */
        L.qt_w = L.max_bx - min_b;
        L.qt_h = L.max_by - min_b;
        L.n_ini = (int)std::round((float)L.qt_w / (float)L.qt_h);
        if (L.n_ini <= 0)
            return fail(BORB_ERR_GEOMETRY, "level %d: width/height rounds to 0 root nodes "
                        "(the reference divides by zero here)", l);
        L.hx = (float)L.qt_w / (float)L.n_ini;
        L.kp_cap = std::max(L.quota, 4 * L.n_ini) + 4;
        L.kp_off = kp_off;
        kp_off += L.kp_cap;
        M = std::max(M, L.kp_cap);
        for (int y = 0; y < L.h; y += BORB_BLUR_TH)
            for (int x = 0; x < L.w; x += BORB_BLUR_TW) {
                BorbTile t;
                t.level = (short)l; t.pad = 0; t.x0 = (short)x; t.y0 = (short)y;
                tiles.push_back(t);
            }
    }
    M = (int)align_up(M, 4); /* K3 clears a node's four quarter counters with one 16-byte store */
    if (M > BORB_QT_MAXNODES)
        return fail(BORB_ERR_CAPACITY, "%d features on one level exceed the shared-memory quadtree "
                    "capacity (%d)", M, BORB_QT_MAXNODES);
    P.pyr_bytes = pyr_off;
    P.cand_total = cand_off;
    P.kp_total = kp_off;
    P.n_cells = (int)cells.size();
    P.n_tiles = (int)tiles.size();

    const size_t B = (size_t)h->max_batch;
    CU(cudaMalloc(&h->d_plan, sizeof P));
    CU(cudaMalloc(&h->d_cells, cells.size() * sizeof(BorbCell)));
    CU(cudaMalloc(&h->d_tiles, tiles.size() * sizeof(BorbTile)));
    CU(cudaMalloc(&h->d_tab_i, std::max<size_t>(tab_i.size(), 1) * sizeof(int)));
    CU(cudaMalloc(&h->d_tab_s, std::max<size_t>(tab_s.size(), 1) * sizeof(short)));
    CU(cudaMalloc(&h->d_pyr, B * (size_t)P.pyr_bytes));
    CU(cudaMalloc(&h->d_blur, B * (size_t)P.pyr_bytes));
    CU(cudaMalloc(&h->d_cand_xy, B * (size_t)P.cand_total * 4));
    CU(cudaMalloc(&h->d_cand_score, B * (size_t)P.cand_total * 4));
    CU(cudaMalloc(&h->d_cand_node, B * (size_t)P.cand_total * 4));
    CU(cudaMalloc(&h->d_ncand, B * BORB_MAXL * 4));
    CU(cudaMalloc(&h->d_nlkp, B * BORB_MAXL * 4));
    {
        const size_t o_desc = (size_t)align_up((long long)(B * (size_t)P.kp_total * sizeof(BorbKpRec)), 256);
        const size_t o_nout = o_desc + (size_t)align_up((long long)(B * (size_t)P.kp_total * 32), 256);
        const size_t o_err = o_nout + (size_t)align_up((long long)(B * 4), 16);
        h->res_bytes = o_err + 16;
        h->res_o_desc = o_desc; h->res_o_nout = o_nout; h->res_o_err = o_err;
        CU(cudaMalloc(&h->d_res, h->res_bytes));
        CU(cudaMallocHost(&h->h_res, h->res_bytes));
        CU(cudaMemset(h->d_res, 0, h->res_bytes));
        std::memset(h->h_res, 0, h->res_bytes);
        {
            void *alias = nullptr;
            if (cudaHostGetDevicePointer(&alias, h->h_res, 0) != cudaSuccess) { cudaGetLastError(); alias = nullptr; }
            h->h_res_dev = (h->res_bytes % 16 == 0) ? static_cast<uint8_t *>(alias) : nullptr;
        }
        h->d_out_kp = reinterpret_cast<BorbKpRec *>(h->d_res); h->h_out_kp = reinterpret_cast<BorbKpRec *>(h->h_res);
        h->d_out_desc = h->d_res + o_desc; h->h_out_desc = h->h_res + o_desc;
        h->d_nout = reinterpret_cast<int *>(h->d_res + o_nout); h->h_nout = reinterpret_cast<int *>(h->h_res + o_nout);
        h->d_err = reinterpret_cast<int *>(h->d_res + o_err); h->h_err = reinterpret_cast<int *>(h->h_res + o_err);
    }
    CU(cudaMalloc(&h->d_lkp_xy, B * (size_t)P.kp_total * 4));
    CU(cudaMalloc(&h->d_lkp_score, B * (size_t)P.kp_total * 4));
    CU(cudaMemcpy(h->d_plan, &P, sizeof P, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(h->d_cells, cells.data(), cells.size() * sizeof(BorbCell), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(h->d_tiles, tiles.data(), tiles.size() * sizeof(BorbTile), cudaMemcpyHostToDevice));
    if (!tab_i.empty()) CU(cudaMemcpy(h->d_tab_i, tab_i.data(), tab_i.size() * sizeof(int), cudaMemcpyHostToDevice));
    if (!tab_s.empty()) CU(cudaMemcpy(h->d_tab_s, tab_s.data(), tab_s.size() * sizeof(short), cudaMemcpyHostToDevice));
    {
        const int rc = build_chain(h, P, tab_i, tab_s);
        if (rc != BORB_OK) return rc;
    }
    CU(cudaMemset(h->d_err, 0, 4));
    /* the blurred slab's frame/padding is never written; keep it defined */
    CU(cudaMemset(h->d_blur, 0, B * (size_t)P.pyr_bytes));
    CU(cudaMemset(h->d_pyr, 0, B * (size_t)P.pyr_bytes));

    h->fast_wp = FAST_WP_FOR(P.fast_max_ww);
    h->fast_smem = fast_smem_layout(P.fast_max_wh, P.fast_max_pix, h->fast_wp, 2).bytes;
    h->fast_smem4 = fast_smem_layout(P.fast_max_wh, P.fast_max_pix, h->fast_wp, 4).bytes;
    CU(cudaFuncSetAttribute(k_fast_cells<64, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->fast_smem));
    CU(cudaFuncSetAttribute(k_fast_cells<80, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->fast_smem));
    CU(cudaFuncSetAttribute(k_fast_cells<96, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->fast_smem));
    CU(cudaFuncSetAttribute(k_fast_cells<64, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->fast_smem4));
    CU(cudaFuncSetAttribute(k_fast_cells<80, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->fast_smem4));
    CU(cudaFuncSetAttribute(k_fast_cells<96, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->fast_smem4));
    {   /* TMA descriptors of K2's window loads: one 3-D map (column bytes, rows, frames) per padded
         * level plane; box = fast_wp bytes x fast_max_wh rows x 1 frame */
        typedef CUresult (*EncodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                        const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                        CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        CU(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
        if (!fn || qres != cudaDriverEntryPointSuccess) return fail(BORB_ERR_CUDA, "cuTensorMapEncodeTiled is not available");
        CUtensorMap *maps = h->fast_maps.m;
        std::memset(&h->fast_maps, 0, sizeof h->fast_maps);
        std::memset(&h->desc_maps, 0, sizeof h->desc_maps);
        h->fast_maps.max_wh = P.fast_max_wh; h->fast_maps.max_pix = P.fast_max_pix; h->fast_maps.n_cells = P.n_cells;
        h->fast_maps.ini_th = P.ini_th; h->fast_maps.min_th = P.min_th;
        for (int l = 0; l < P.nlevels; ++l) {
            const BorbLevel &L = P.lv[l];
            const cuuint64_t dims[3] = {(cuuint64_t)L.pitch, (cuuint64_t)L.prows, (cuuint64_t)B};
            const cuuint64_t strides[2] = {(cuuint64_t)L.pitch, (cuuint64_t)P.pyr_bytes};
            const cuuint32_t box[3] = {(cuuint32_t)h->fast_wp, (cuuint32_t)P.fast_max_wh, 1u};
            const cuuint32_t estr[3] = {1u, 1u, 1u};
            const CUresult r = ((EncodeTiled)fn)(&maps[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, h->d_pyr + L.off, dims, strides, box,
                                                 estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                                 CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) return fail(BORB_ERR_CUDA, "cuTensorMapEncodeTiled(level %d) failed: %d", l, (int)r);
            const cuuint32_t bbox[3] = {(cuuint32_t)BLUR_BOX_W, (cuuint32_t)BLUR_BOX_H, 1u};
            const CUresult rb = ((EncodeTiled)fn)(&h->blur_maps.m[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, h->d_pyr + L.off, dims,
                                                  strides, bbox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                                  CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (rb != CUDA_SUCCESS) return fail(BORB_ERR_CUDA, "cuTensorMapEncodeTiled(blur, level %d) failed: %d", l, (int)rb);
            const cuuint32_t ibox[3] = {(cuuint32_t)DESC_IB, (cuuint32_t)DESC_IR, 1u}, pbox[3] = {(cuuint32_t)DESC_PB, (cuuint32_t)DESC_PR, 1u};
            const CUresult ri = ((EncodeTiled)fn)(&h->desc_maps.pyr[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, h->d_pyr + L.off, dims,
                                                  strides, ibox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                                  CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            const CUresult rp = ((EncodeTiled)fn)(&h->desc_maps.blur[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, h->d_blur + L.off, dims,
                                                  strides, pbox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                                  CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (ri != CUDA_SUCCESS || rp != CUDA_SUCCESS)
                return fail(BORB_ERR_CUDA, "cuTensorMapEncodeTiled(describe, level %d) failed: %d / %d", l, (int)ri, (int)rp);
            h->desc_maps.kp_off[l] = L.kp_off; h->desc_maps.scale[l] = L.scale; h->desc_maps.patch_size[l] = L.patch_size;
        }
        h->desc_maps.nlevels = P.nlevels; h->desc_maps.kp_total = P.kp_total;
        std::memcpy(h->desc_maps.umax, P.umax, sizeof P.umax);
        CU(cudaFuncSetAttribute(k_orient_describe<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, DESC_SMEM_BYTES));
        CU(cudaFuncSetAttribute(k_orient_describe<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, DESC_SMEM_BYTES));
        CU(cudaFuncSetAttribute(k_blur_strips, cudaFuncAttributeMaxDynamicSharedMemorySize, BLUR_SMEM_BYTES));
    }
    h->qt_M = M;
    const size_t qt_smem = qt_smem_bytes(M);
    CU(cudaFuncSetAttribute(k_quadtree<QT_NT_LATENCY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)qt_smem));
    CU(cudaFuncSetAttribute(k_quadtree<QT_NT_THROUGHPUT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)qt_smem));
    h->plan = P;
    h->cells.swap(cells);
    h->tiles.swap(tiles);
    h->have_plan = true;
    return BORB_OK;
}

} // namespace

namespace {

bool pre_active(const borb_extractor *h) { return h->in_ch > 1 || h->d_map_xy != nullptr; }

/* A kernel launch with the programmatic-serialization attribute (pdl): the kernel may be scheduled while its
 * predecessor in the stream still runs and meets it at borb_pdl_wait() (borb_cta.cuh).  Captured into a graph the edge
 * becomes a programmatic dependency.  Used for the single-frame shape only, where the ~1.5 us between two dependent
 * kernels is a visible part of the call. */
template <typename... KArgs, typename... Args>
static cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, bool pdl, Args &&...args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1u : 0u;
    return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}


/* Enqueue the whole per-call chain for n frames whose pixels are at dev_img on the device. */
int enqueue_kernels(borb_extractor *h, const uint8_t *dev_img, int n, size_t step, size_t stride,
                    int scols, int srows, bool stage_events = false)
{
    stage_events = stage_events || h->stage_timing; /* inside a capture these become event-record nodes */
    cudaStreamCaptureStatus cap_status = cudaStreamCaptureStatusNone;
    CU(cudaStreamIsCapturing(h->stream, &cap_status));
    const bool capturing = cap_status == cudaStreamCaptureStatusActive;
/* cudaEventRecordExternal: inside a stream capture this becomes an event-record node whose event
 * can be timed from outside the graph (a plainly captured event cannot) */
#define STAGE(i) do { if (stage_events) CU(cudaEventRecordWithFlags(h->evs[i], st, capturing ? cudaEventRecordExternal : cudaEventRecordDefault)); } while (0)
    const BorbPlan &P = h->plan;
    cudaStream_t st = h->stream;
    /* Frames are independent, so a large batch runs through the chain in sub-batches of `chunk`
     * frames: a sub-batch's pyramid, blurred levels and candidates (~4 MB per frame) then still sit
     * in the 126 MB L2 when the next stage reads them.  BORB_CHUNK overrides (0 = whole batch). */
    int chunk = h->chunk > 0 ? h->chunk : n;
    if (stage_events) chunk = n; /* stage timers describe whole-batch launches */
    const int n_total = n;
    const uint8_t *const img_all = dev_img;
    const size_t in_step = step, in_stride = stride;
    for (int f0 = 0; f0 < n_total; f0 += chunk) {
    n = std::min(chunk, n_total - f0);
    step = in_step;
    stride = in_stride;
    dev_img = img_all + (size_t)f0 * stride;
    uint8_t *const d_pyr = h->d_pyr + (size_t)f0 * (size_t)P.pyr_bytes;
    uint32_t *const d_cand_xy = h->d_cand_xy + (size_t)f0 * (size_t)P.cand_total;
    uint32_t *const d_cand_score = h->d_cand_score + (size_t)f0 * (size_t)P.cand_total;
    uint32_t *const d_cand_node = h->d_cand_node + (size_t)f0 * (size_t)P.cand_total;
    int *const d_ncand = h->d_ncand + (size_t)f0 * BORB_MAXL, *const d_nlkp = h->d_nlkp + (size_t)f0 * BORB_MAXL;
    int *const d_nout = h->d_nout + f0;
    uint32_t *const d_lkp_xy = h->d_lkp_xy + (size_t)f0 * P.kp_total, *const d_lkp_score = h->d_lkp_score + (size_t)f0 * P.kp_total;
    BorbKpRec *const d_out_kp = h->d_out_kp + (size_t)f0 * P.kp_total;
    uint8_t *const d_out_desc = h->d_out_desc + (size_t)f0 * P.kp_total * 32;
    CU(cudaMemsetAsync(d_ncand, 0, (size_t)n * BORB_MAXL * 4, st));
    STAGE(0);
    if (pre_active(h)) { /* K0: rectify and/or convert to gray; K1 then reads its output */
        const int C = h->in_ch, rgb = (h->in_format == BORB_RGB8 || h->in_format == BORB_RGBA8) ? 1 : 0;
        const size_t gstride = (size_t)h->gray_pitch * P.rows;
        uint8_t *gray = h->d_gray + (size_t)f0 * gstride;
        const dim3 g((P.cols + PRE_PX - 1) / PRE_PX, P.rows, n);
        if (h->d_map_xy) {
#define PRE_REMAP(CH) k_pre_remap<CH><<<g, PRE_NT, 0, st>>>(dev_img, step, stride, scols, srows, h->d_map_xy, h->d_map_fr, \
                                                        h->map_pitch, rgb, gray, h->gray_pitch, gstride, P.cols)
            if (C == 1) PRE_REMAP(1); else if (C == 3) PRE_REMAP(3); else PRE_REMAP(4);
#undef PRE_REMAP
        } else {
            if (C == 3) k_pre_gray<3><<<g, PRE_NT, 0, st>>>(dev_img, step, stride, P.cols, rgb, gray, h->gray_pitch, gstride);
            else k_pre_gray<4><<<g, PRE_NT, 0, st>>>(dev_img, step, stride, P.cols, rgb, gray, h->gray_pitch, gstride);
        }
        dev_img = gray;
        step = (size_t)h->gray_pitch;
        stride = gstride;
        STAGE(6); /* K0 = evs[0] .. evs[6]; the pyramid stage time includes it */
    }
    static const bool chain_off = getenv("BORB_PYR_LEVELS") != nullptr; /* A/B knob: per-level kernels in every shape */
    static const bool pyr_stream_off = getenv("BORB_PYR_RESIZE_V3") != nullptr; /* A/B knob: k_pyr_resize instead of k_pyr_stream */
    const bool chain = !chain_off && h->chain_tiles > 0 && (long long)n * h->chain_tiles <= 2 * h->n_sm;
    h->last_launches = (chain ? 1 : P.nlevels + 1) + 4 + (pre_active(h) ? 1 : 0) + (h->have_cam ? 1 : 0);
    static const bool pdl_off = getenv("BORB_NO_PDL") != nullptr; /* A/B knob */
    const bool pdl = chain && !stage_events && !pdl_off; /* K2, K3 and K6 launched programmatically behind their predecessors */
    const bool pdl2 = pdl, pdl3 = pdl, pdl6 = pdl;
    if (chain) {
        /* K1, latency shape: the whole chain in one launch */
        k_pyr_chain<<<dim3(h->chain_tiles, n), PCH_NT, h->chain_smem, st>>>(h->plan, h->d_chain, h->d_chain_tabs, h->chain_tstride,
                                                                         dev_img, step, stride, d_pyr, h->chain_off_b, h->chain_off_t);
    } else {   /* K1 */
        const BorbLevel &L0 = P.lv[0];
        const int units = L0.pitch / 16, n_int = L0.w / 16;
        if (n_int > 0)
            k_pyr_level0<true><<<dim3((n_int + 31) / 32, (L0.prows + 3) / 4, n), dim3(32, 4), 0, st>>>(h->plan, dev_img, step, stride, d_pyr);
        k_pyr_level0<false><<<dim3((L0.prows + 127) / 128, units - n_int, n), dim3(32, 4), 0, st>>>(h->plan, dev_img, step, stride, d_pyr);
        for (int l = 1; l < P.nlevels; ++l) {
            const BorbLevel &L = P.lv[l];
            if (h->pyr_stream[l] && !pyr_stream_off) {
                const dim3 g((4 * pyr_words_per_row(L.w) + PYR_TW - 1) / PYR_TW, (P.lv[l - 1].h - 1 + 4 * PST_NSW - 1) / (4 * PST_NSW), n);
                const uint2 *yinv = reinterpret_cast<const uint2 *>(h->d_tab_i + h->pyr_yinv_off[l]);
                if (h->pyr_sp[l] <= 192) /* + 16: the second word of the last pixel pair of the last staged row */
                    k_pyr_stream<192><<<g, PYR_NT, PST_ROWS * 192 + 16, st>>>(h->plan, l, h->d_tab_i, h->d_tab_s, yinv, d_pyr);
                else
                    k_pyr_stream<PYR_SP_MAX><<<g, PYR_NT, PST_ROWS * PYR_SP_MAX + 16, st>>>(h->plan, l, h->d_tab_i, h->d_tab_s, yinv, d_pyr);
                continue;
            }
            const dim3 g((4 * pyr_words_per_row(L.w) + PYR_TW - 1) / PYR_TW, (L.prows + PYR_TH - 1) / PYR_TH, n);
            k_pyr_resize<<<g, PYR_NT, (size_t)2 * h->pyr_sp[l] * h->pyr_srows[l], st>>>(h->plan, l, h->d_tab_i, h->d_tab_s,
                                                                                   d_pyr, h->pyr_sp[l], h->pyr_srows[l]);
        }
    }
    STAGE(1);
    /* latency shape (a frame or two: every grid is far smaller than the machine): the blur runs beside FAST and the
     * quadtree.  Throughput shapes keep the serial chain: measured no gain there (every kernel is issue-bound and the
     * quadtree's CTAs hold an SM's shared memory), and the stage timers describe a serial chain. */
    static const bool blur_serial = getenv("BORB_BLUR_SERIAL") != nullptr; /* A/B knob */
    const bool blur_beside = !blur_serial && !stage_events && n * P.nlevels <= 2 * h->n_sm;
#define K5_ON(S_) k_blur_strips<<<dim3((P.n_tiles + BLUR_NT / 32 - 1) / (BLUR_NT / 32), n), BLUR_NT, BLUR_SMEM_BYTES, S_>>>( \
        h->plan, h->blur_maps, h->d_tiles, f0, h->d_blur)
    if (blur_beside) {
        CU(cudaEventRecord(h->ev_fork, st));
        CU(cudaStreamWaitEvent(h->stream2, h->ev_fork, 0));
        K5_ON(h->stream2);
        CU(cudaEventRecord(h->ev_join, h->stream2));
    }
    {   /* K2: persistent CTAs walking (cell, frame) items, frames first (stride gridDim.y), then cells (stride
         * gridDim.x); small jobs get one CTA per (cell, frame) */
        int per_sm = 768; /* far more CTAs than fit at once: the hardware scheduler evens out the uneven cells.
                             K2k, 512 frames, CTA rows per cell column (frames walked per CTA = 512 / rows): 29 rows 2.025 ms,
                             64 1.996, 96 1.994, 128 2.005, 256 2.058, 512 (one frame per CTA) 2.190 */
        static const int env_per_sm = getenv("BORB_FAST_CTAS_PER_SM") ? atoi(getenv("BORB_FAST_CTAS_PER_SM")) : 0; /* tuning knobs, read once */
        static const int env_grid_y = getenv("BORB_FAST_GRID_Y") ? atoi(getenv("BORB_FAST_GRID_Y")) : 0;
        if (env_per_sm > 0) per_sm = env_per_sm;
        const int ctas = h->n_sm * per_sm;
        dim3 grid(P.n_cells, n);
        if ((long long)P.n_cells * n > 2LL * ctas) {
            /* one CTA column per cell: a CTA keeps its cell and walks frames blockIdx.y, + gridDim.y, ... so that the
             * per-cell lane geometry is computed once per CTA (k_fast.cuh) */
            grid.x = std::min(P.n_cells, ctas);
            grid.y = std::max(1, std::min(std::min(n, 96), ctas / (int)grid.x));
            if (env_grid_y > 0) grid.y = std::min(n, env_grid_y);
        }
        /* a frame or two: one CTA per (cell, frame), and the time is a cell's dependent chain -> four warps per cell */
        static const bool env_nw2 = getenv("BORB_FAST_NW2") != nullptr; /* A/B knob: two warps per cell in every shape */
        const bool four = !env_nw2 && (long long)P.n_cells * n <= 2LL * h->n_sm * 16;
        /* a single frame: no more CTAs than are resident at once (8 of 128 threads per SM by registers) -- the few cells
         * beyond are walked by CTAs that already hold their places, with the next window's TMA in flight, instead of
         * waiting for a place and paying a prologue (1220 cells on 1184 places: 46.8 -> 46.3 us per call; fewer CTAs
         * than that: 47.3 us at 7 per SM, 48.8 at 6) */
        if (four && n == 1) grid.x = std::min(P.n_cells, h->n_sm * 8);
#define K2(WP, NW) CU(launch_k(k_fast_cells<WP, NW>, grid, dim3(32 * NW), (size_t)(NW == 4 ? h->fast_smem4 : h->fast_smem), st, pdl2, \
        h->plan, (const BorbCell *)h->d_cells, n, h->fast_maps, f0, d_cand_xy, d_cand_score, d_ncand))
        if (four) { if (h->fast_wp == 64) K2(64, 4); else if (h->fast_wp == 80) K2(80, 4); else K2(96, 4); }
        else { if (h->fast_wp == 64) K2(64, 2); else if (h->fast_wp == 80) K2(80, 2); else K2(96, 2); }
#undef K2
    }
    STAGE(2);
    if (n * P.nlevels <= 2 * h->n_sm) /* few CTAs: give each one a whole SM's worth of threads */
        CU(launch_k(k_quadtree<QT_NT_LATENCY>, dim3(P.nlevels, n), dim3(QT_NT_LATENCY), qt_smem_bytes(h->qt_M), st, pdl3,
                    h->plan, h->qt_M, (const uint32_t *)d_cand_xy, (const uint32_t *)d_cand_score, d_cand_node, (const int *)d_ncand,
                    d_lkp_xy, d_lkp_score, d_nlkp, h->d_err));
    else
        k_quadtree<QT_NT_THROUGHPUT><<<dim3(P.nlevels, n), QT_NT_THROUGHPUT, qt_smem_bytes(h->qt_M), st>>>(
            h->plan, h->qt_M, d_cand_xy, d_cand_score, d_cand_node, d_ncand, d_lkp_xy, d_lkp_score, d_nlkp, h->d_err);
    STAGE(3);
    /* the descriptors address the whole slab: frame index = f0 + blockIdx.y */
    if (blur_beside) CU(cudaStreamWaitEvent(st, h->ev_join, 0)); /* (joining before K3 instead: 61 us against 55 -- the blur ends after K2) */
    else K5_ON(st);
#undef K5_ON
    STAGE(4);
    /* keypoint slots per warp: 16 when there are plenty of frames, 1 for a single frame */
    const int kpw = std::max(1, std::min(DESC_KPW, n / 4));
    CU(launch_k(pdl6 ? k_orient_describe<true> : k_orient_describe<false>, dim3((P.kp_total + DESC_NW * kpw - 1) / (DESC_NW * kpw), n), dim3(DESC_NT), (size_t)DESC_SMEM_BYTES, st, pdl6,
                h->desc_maps, (const signed char *)h->d_pattern, f0, (const uint32_t *)d_lkp_xy, (const uint32_t *)d_lkp_score,
                (const int *)d_nlkp, d_out_kp, d_out_desc, d_nout));
    if (h->have_cam) { /* K8: rides in the describe stage's time slot */
#define K8(NT) k_frame_undistort_grid<NT><<<n, NT, 0, st>>>(h->d_plan, h->cam, d_out_kp, d_nout, \
        h->d_un_xy + (size_t)f0 * P.kp_total, h->d_cell_start + (size_t)f0 * (FR_CELLS + 1), \
        h->d_cell_idx + (size_t)f0 * P.kp_total, h->d_cell_of + (size_t)f0 * P.kp_total, h->d_bounds + (size_t)f0 * 4)
        if (n <= 2 * h->n_sm) K8(FR_NT_LATENCY); else K8(FR_NT_THROUGHPUT);
#undef K8
    }
    STAGE(5);
    }
    n = n_total;
#undef STAGE
    CU(cudaGetLastError());
    return BORB_OK;
}

/* The result slab of a frame or two goes to the host by a KERNEL (16-byte loads from the device slab, 16-byte stores
 * through the device alias of the page-locked host slab) instead of a copy node: the copy engine's fixed cost -- about
 * 9 of the 11 us the download of 122 KB added to a single-frame call -- is most of that call's download time.  Launched
 * programmatically behind the last kernel of the chain. */
__global__ void __launch_bounds__(256)
k_results_to_host(const uint4 *src_p, uint4 *__restrict__ dst, int n16)
{
    borb_pdl_wait();
    const uint4 *src = borb_pdl_after_wait(src_p);
    for (int i = blockIdx.x * 256 + threadIdx.x; i < n16; i += gridDim.x * 256) dst[i] = src[i];
}

int enqueue_fetch(borb_extractor *h, int n, unsigned flags)
{
    const BorbPlan &P = h->plan;
    cudaStream_t st = h->stream;
    static const bool fetch_copy = getenv("BORB_FETCH_COPY") != nullptr; /* A/B knob: always the copy engine */
    if (n == h->max_batch && n <= 2 && h->h_res_dev && !fetch_copy) {
        const int n16 = (int)(h->res_bytes / 16);
        CU(launch_k(k_results_to_host, dim3((unsigned)std::min(h->n_sm, (n16 + 255) / 256)), dim3(256), (size_t)0, st, true,
                    (const uint4 *)h->d_res, (uint4 *)h->h_res_dev, n16));
    } else if (n == h->max_batch) { /* the whole result slab in one copy */
        CU(cudaMemcpyAsync(h->h_res, h->d_res, h->res_bytes, cudaMemcpyDeviceToHost, st));
    } else {
        CU(cudaMemcpyAsync(h->h_nout, h->d_nout, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(h->h_err, h->d_err, 4, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(h->h_out_kp, h->d_out_kp, (size_t)n * P.kp_total * sizeof(BorbKpRec),
                           cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(h->h_out_desc, h->d_out_desc, (size_t)n * P.kp_total * 32,
                           cudaMemcpyDeviceToHost, st));
    }
    if (flags & BORB_WANT_PYRAMID)
        CU(cudaMemcpyAsync(h->h_pyr, h->d_pyr, (size_t)n * (size_t)P.pyr_bytes,
                           cudaMemcpyDeviceToHost, st));
    if (h->have_cam) {
        if (n == h->max_batch) {
            CU(cudaMemcpyAsync(h->h_fr, h->d_fr, h->fr_bytes, cudaMemcpyDeviceToHost, st));
        } else {
            CU(cudaMemcpyAsync(h->h_un_xy, h->d_un_xy, (size_t)n * P.kp_total * 8, cudaMemcpyDeviceToHost, st));
            CU(cudaMemcpyAsync(h->h_cell_start, h->d_cell_start, (size_t)n * (FR_CELLS + 1) * 4, cudaMemcpyDeviceToHost, st));
            CU(cudaMemcpyAsync(h->h_cell_idx, h->d_cell_idx, (size_t)n * P.kp_total * 4, cudaMemcpyDeviceToHost, st));
            CU(cudaMemcpyAsync(h->h_bounds, h->d_bounds, (size_t)n * 16, cudaMemcpyDeviceToHost, st));
        }
    }
    return BORB_OK;
}

/* body(h) enqueues work on h->stream; run it through a cached CUDA graph when enabled */
template <typename F>
int run_cached(borb_extractor *h, const GraphKey &key, F body, bool always_capture = false)
{
    if (!h->use_graph) return body();
    for (GraphEntry &g : h->graphs)
        if (g.key == key) {
            h->graph_miss_streak = 0;
            CU(cudaGraphLaunch(g.exec, h->stream));
            return BORB_OK;
        }
    /* A caller that hands over a different buffer on every call would pay a capture + instantiation each time:
     * after 8 misses in a row a call shape is only captured once it has been seen twice in a row, and launched
     * directly until then */
    if (!always_capture && h->graph_miss_streak >= 8 && !(h->have_last_miss && h->last_miss == key)) {
        h->last_miss = key;
        h->have_last_miss = true;
        return body();
    }
    ++h->graph_miss_streak;
    h->have_last_miss = false;
    cudaGraph_t graph = nullptr;
    CU(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
    const int rc = body();
    cudaError_t e = cudaStreamEndCapture(h->stream, &graph);
    if (rc != BORB_OK) {
        if (graph) cudaGraphDestroy(graph);
        return rc;
    }
    if (e != cudaSuccess) return fail(BORB_ERR_CUDA, "graph capture: %s", cudaGetErrorString(e));
    GraphEntry ge;
    ge.key = key;
    e = cudaGraphInstantiate(&ge.exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) return fail(BORB_ERR_CUDA, "graph instantiate: %s", cudaGetErrorString(e));
    if (h->graphs.size() >= 48) {
        cudaGraphExecDestroy(h->graphs.front().exec);
        h->graphs.erase(h->graphs.begin());
    }
    h->graphs.push_back(ge);
    CU(cudaGraphLaunch(ge.exec, h->stream));
    return BORB_OK;
}

/* every extract call invalidates what the previous one left on the host -- and the stereo association, whose
 * mvuRight / mvDepth are parallel to the PREVIOUS call's keypoints (stereo_gen != extract_gen from here on) */
void note_new_extract(borb_extractor *h)
{
    ++h->extract_gen;
    h->results_on_host = false;
    h->nout_frame_on_host = -1;
    h->pyramid_on_host = false;
    h->stereo_on_host = false;
    h->frame_on_host = false;
}

int check_device_err(borb_extractor *h)
{
    if (*h->h_err != 0) {
        const int e = *h->h_err;
        *h->h_err = 0;
        cudaMemsetAsync(h->d_err, 0, 4, h->stream);
        return fail(BORB_ERR_CAPACITY, "device buffer overflow (code %d)", e);
    }
    return BORB_OK;
}

int ensure_pyr_host(borb_extractor *h)
{
    if (!h->h_pyr)
        CU(cudaMallocHost(&h->h_pyr, (size_t)h->max_batch * (size_t)h->plan.pyr_bytes));
    return BORB_OK;
}

} // namespace

extern "C" {

const char *borb_last_error(void) { return g_err.c_str(); }
const char *borb_version(void) { return "orb_b200 0.1 (sm_100a)"; }

int borb_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th, int device,
                int max_batch, borb_extractor **out)
{
    if (!out) return fail(BORB_ERR_ARG, "out is null");
    *out = nullptr;
    if (nfeatures < 1 || nlevels < 1 || nlevels > BORB_MAXL || !(scale_factor > 1.0f) ||
        ini_th < 1 || min_th < 1 || ini_th > 254 || min_th > 254 || max_batch < 1)
        return fail(BORB_ERR_ARG, "unsupported parameters (nlevels 1..%d, scaleFactor > 1, "
                    "FAST thresholds 1..254, max_batch >= 1)", BORB_MAXL);
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(BORB_ERR_CUDA, "no CUDA device: this library has no CPU path");
    }
    if (device < 0) CU(cudaGetDevice(&device));
    if (device >= ndev) return fail(BORB_ERR_ARG, "device %d of %d", device, ndev);
    CU(cudaSetDevice(device));
    int n_sm = 0;
    CU(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, device));
    borb_extractor *h = new borb_extractor();
    h->n_sm = n_sm > 0 ? n_sm : 148;
    h->nfeatures = nfeatures; h->nlevels = nlevels; h->ini_th = ini_th; h->min_th = min_th;
    h->scale_factor = scale_factor;
    h->device = device; h->max_batch = max_batch;
    build_tables(h);
    if (const char *e = getenv("BORB_CHUNK")) h->chunk = std::max(0, atoi(e));
    const char *ng = getenv("BORB_NO_GRAPH");
    h->use_graph = !(ng && ng[0] == '1');
    cudaError_t e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->stream2, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreate(&h->ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&h->ev1);
    if (e == cudaSuccess) e = cudaEventCreate(&h->evt0);
    if (e == cudaSuccess) e = cudaEventCreate(&h->evt1);
    for (int i = 0; i < 8 && e == cudaSuccess; ++i) e = cudaEventCreate(&h->evs[i]);
    if (e == cudaSuccess) e = cudaMalloc(&h->d_pattern, sizeof kPattern);
    if (e == cudaSuccess) e = cudaMemcpy(h->d_pattern, kPattern, sizeof kPattern, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        borb_destroy(h);
        return fail(BORB_ERR_CUDA, "create: %s", cudaGetErrorString(e));
    }
    *out = h;
    return BORB_OK;
}

void borb_destroy(borb_extractor *h)
{
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    free_geometry(h);
    if (h->d_pattern) cudaFree(h->d_pattern);
    if (h->d_map_xy) cudaFree(h->d_map_xy);
    if (h->d_map_fr) cudaFree(h->d_map_fr);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->evt0) cudaEventDestroy(h->evt0);
    if (h->evt1) cudaEventDestroy(h->evt1);
    for (int i = 0; i < 8; ++i) if (h->evs[i]) cudaEventDestroy(h->evs[i]);
    if (h->ev_peer) cudaEventDestroy(h->ev_peer);
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    if (h->ev_join) cudaEventDestroy(h->ev_join);
    if (h->stream2) cudaStreamDestroy(h->stream2);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

int borb_levels(const borb_extractor *h) { return h ? h->nlevels : 0; }
float borb_scale_factor(const borb_extractor *h) { return h ? (float)h->scale_factor : 0.f; }

int borb_table(const borb_extractor *h, int which, float *out)
{
    if (!h || !out || which < 0 || which > 3) return fail(BORB_ERR_ARG, "borb_table arguments");
    const std::vector<float> &v = which == 0 ? h->scale : which == 1 ? h->inv_scale
                                : which == 2 ? h->sigma2 : h->inv_sigma2;
    std::memcpy(out, v.data(), v.size() * sizeof(float));
    return BORB_OK;
}
int borb_features_per_level(const borb_extractor *h, int32_t *out)
{
    if (!h || !out) return fail(BORB_ERR_ARG, "null");
    std::memcpy(out, h->quota.data(), h->quota.size() * sizeof(int));
    return BORB_OK;
}
int borb_umax(const borb_extractor *h, int32_t *out)
{
    if (!h || !out) return fail(BORB_ERR_ARG, "null");
    std::memcpy(out, h->umax, sizeof h->umax);
    return BORB_OK;
}

/* device / pinned input staging for frames of `bytes` bytes (grows on demand) */
static int ensure_input(borb_extractor *h, size_t bytes)
{
    if (h->d_in && bytes <= h->in_cap) return BORB_OK;
    CU(cudaStreamSynchronize(h->stream));
    for (GraphEntry &g : h->graphs) cudaGraphExecDestroy(g.exec); /* they hold the old pointers */
    h->graphs.clear();
    if (h->d_in) cudaFree(h->d_in);
    if (h->h_in) cudaFreeHost(h->h_in);
    h->d_in = h->h_in = nullptr;
    h->in_cap = 0;
    CU(cudaMalloc(&h->d_in, (size_t)h->max_batch * bytes));
    CU(cudaMallocHost(&h->h_in, (size_t)h->max_batch * bytes)); /* (write-combined pages: no change, 99.2 us against 98.4) */
    h->in_cap = bytes;
    return BORB_OK;
}

/* cols x rows: the SOURCE image (in_ch interleaved channels); the extractor's geometry is the
 * rectification map's when one is set */
/* K8's outputs (undistorted positions, grid, bounds): one slab on either side of the link, once a camera is set */
static int ensure_frame_buffers(borb_extractor *h)
{
    if (h->have_cam && !h->d_un_xy) {
        const size_t B = (size_t)h->max_batch, K = (size_t)h->plan.kp_total;
        const size_t o_cs = (size_t)align_up((long long)(B * K * 8), 256);
        const size_t o_ci = o_cs + (size_t)align_up((long long)(B * (FR_CELLS + 1) * 4), 256);
        const size_t o_bd = o_ci + (size_t)align_up((long long)(B * K * 4), 256);
        h->fr_bytes = o_bd + B * 16;
        h->fr_o_cs = o_cs; h->fr_o_ci = o_ci; h->fr_o_bd = o_bd;
        CU(cudaMalloc(&h->d_fr, h->fr_bytes));
        CU(cudaMallocHost(&h->h_fr, h->fr_bytes));
        CU(cudaMemset(h->d_fr, 0, h->fr_bytes));
        CU(cudaMalloc(&h->d_cell_of, B * K * 4));
        h->d_un_xy = reinterpret_cast<float2 *>(h->d_fr); h->h_un_xy = reinterpret_cast<float2 *>(h->h_fr);
        h->d_cell_start = reinterpret_cast<int *>(h->d_fr + o_cs); h->h_cell_start = reinterpret_cast<int *>(h->h_fr + o_cs);
        h->d_cell_idx = reinterpret_cast<int *>(h->d_fr + o_ci); h->h_cell_idx = reinterpret_cast<int *>(h->h_fr + o_ci);
        h->d_bounds = reinterpret_cast<float *>(h->d_fr + o_bd); h->h_bounds = reinterpret_cast<float *>(h->h_fr + o_bd);
    }
    return BORB_OK;
}

static int check_call(borb_extractor *h, const void *img, int n, int cols, int rows, size_t step,
                      size_t stride)
{
    if (!h || !img) return fail(BORB_ERR_ARG, "null handle or image");
    const size_t row_bytes = (size_t)cols * h->in_ch;
    if (n < 1 || cols < 1 || rows < 1 || step < row_bytes || (n > 1 && stride < step * (size_t)(rows - 1) + row_bytes))
        return fail(BORB_ERR_ARG, "bad image geometry");
    if (n > h->max_batch) return fail(BORB_ERR_CAPACITY, "%d frames > max_batch %d", n, h->max_batch);
    CU(cudaSetDevice(h->device));
    const int ocols = h->d_map_xy ? h->map_cols : cols, orows = h->d_map_xy ? h->map_rows : rows;
    int rc = build_plan(h, ocols, orows);
    if (rc != BORB_OK) return rc;
    if ((rc = ensure_frame_buffers(h)) != BORB_OK) return rc;
    if (pre_active(h) && !h->d_gray) {
        h->gray_pitch = (ocols + 3) & ~3;
        CU(cudaMalloc(&h->d_gray, (size_t)h->max_batch * h->gray_pitch * orows));
    }
    return BORB_OK;
}

/* A frame that arrives in ordinary pageable memory (a cv::Mat) is packed into the handle's page-locked staging area by
 * the calling thread.  With ordinary stores that copy ran at 19 GB/s -- 25 us of a 105 us single-frame call -- although
 * the same memcpy reaches 51 GB/s on the box: the copy engine's read of the previous frame leaves the staging lines out
 * of the cache, so every store first fetches its line from DRAM (tools/packbench.cu: 9 us alone, 24 us between two
 * uploads).  Streaming stores write whole lines without fetching them: 98.4-102.5 -> 92.3 us per call.  x86-64 only (SSE2
 * is part of the base ISA); BORB_PACK_MEMCPY=1 is the A/B knob. */
#if defined(__x86_64__)
#include <emmintrin.h>
#endif
static void pack_copy(uint8_t *d, const uint8_t *s, size_t n)
{
#if defined(__x86_64__)
    static const bool plain = getenv("BORB_PACK_MEMCPY") != nullptr;
    if (plain || n < 512) { std::memcpy(d, s, n); return; }
    size_t head = (16 - (reinterpret_cast<uintptr_t>(d) & 15)) & 15;
    std::memcpy(d, s, head);
    size_t i = head;
    for (; i + 64 <= n; i += 64) {
        const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i *>(s + i)), b = _mm_loadu_si128(reinterpret_cast<const __m128i *>(s + i + 16));
        const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i *>(s + i + 32)), e = _mm_loadu_si128(reinterpret_cast<const __m128i *>(s + i + 48));
        _mm_stream_si128(reinterpret_cast<__m128i *>(d + i), a); _mm_stream_si128(reinterpret_cast<__m128i *>(d + i + 16), b);
        _mm_stream_si128(reinterpret_cast<__m128i *>(d + i + 32), c); _mm_stream_si128(reinterpret_cast<__m128i *>(d + i + 48), e);
    }
    _mm_sfence(); /* the stores are globally visible before the upload is enqueued */
    std::memcpy(d + i, s + i, n - i);
#else
    std::memcpy(d, s, n);
#endif
}

int borb_submit_host(borb_extractor *h, const uint8_t *img, int n, int cols, int rows, size_t step,
                     size_t stride, unsigned flags)
{
    /* before check_call: a geometry change there frees the buffers the pending call is still using */
    if (h && h->pending) return fail(BORB_ERR_STATE, "previous submit has not been waited for");
    int rc = check_call(h, img, n, cols, rows, step, stride);
    if (rc != BORB_OK) return rc;
    if ((flags & BORB_WANT_PYRAMID) && (rc = ensure_pyr_host(h)) != BORB_OK) return rc;
    const size_t rowb = (size_t)cols * h->in_ch, fsz = rowb * rows;
    if ((rc = ensure_input(h, fsz)) != BORB_OK) return rc;
    /* Caller memory that is already page-locked (borb_host_alloc / cudaHostRegister) and densely
     * packed is copied to the device directly; anything else -- the reference's callers hand over
     * ordinary pageable cv::Mat data -- is first packed into the handle's pinned staging area. */
    const uint8_t *src = h->h_in;
    bool direct = false;
    if (step == rowb && (n == 1 || stride == fsz)) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, img) == cudaSuccess && at.type == cudaMemoryTypeHost) {
            direct = true;
            src = img;
        } else {
            cudaGetLastError();
        }
    }
    if (!direct)
        for (int f = 0; f < n; ++f) {
            const uint8_t *s = img + (size_t)f * stride;
            uint8_t *d = h->h_in + (size_t)f * fsz;
            if (step == rowb) pack_copy(d, s, fsz);
            else for (int y = 0; y < rows; ++y) pack_copy(d + (size_t)y * rowb, s + (size_t)y * step, rowb);
        }
    GraphKey key = {0, direct ? (const void *)img : nullptr, n, 0, 0, flags, cols, rows};
    rc = run_cached(h, key, [&]() -> int {
        CU(cudaMemcpyAsync(h->d_in, src, (size_t)n * fsz, cudaMemcpyHostToDevice, h->stream));
        int r = enqueue_kernels(h, h->d_in, n, rowb, fsz, cols, rows);
        if (r != BORB_OK) return r;
        return enqueue_fetch(h, n, flags);
    });
    if (rc != BORB_OK) return rc;
    h->last_n = n;
    h->pending = true;
    h->pending_flags = flags;
    note_new_extract(h);
    return BORB_OK;
}

int borb_wait(borb_extractor *h)
{
    if (!h) return fail(BORB_ERR_ARG, "null");
    if (!h->pending) return fail(BORB_ERR_STATE, "nothing submitted");
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    h->pending = false;
    h->results_on_host = true;
    h->pyramid_on_host = (h->pending_flags & BORB_WANT_PYRAMID) != 0;
    return check_device_err(h);
}

int borb_extract_host(borb_extractor *h, const uint8_t *img, int n, int cols, int rows, size_t step,
                      size_t stride, unsigned flags)
{
    const int rc = borb_submit_host(h, img, n, cols, rows, step, stride, flags);
    return rc != BORB_OK ? rc : borb_wait(h);
}

void *borb_host_alloc(size_t bytes)
{
    void *p = nullptr;
    /* BORB_HOST_WC=1: write-combined pages (the CPU only writes frames into them, the GPU reads) */
    const char *wc = getenv("BORB_HOST_WC");
    if (cudaHostAlloc(&p, bytes, wc && atoi(wc) ? cudaHostAllocWriteCombined : cudaHostAllocDefault) != cudaSuccess) {
        cudaGetLastError();
        fail(BORB_ERR_CUDA, "cudaMallocHost(%zu) failed", bytes);
        return nullptr;
    }
    return p;
}
void borb_host_free(void *p) { if (p) cudaFreeHost(p); }

int borb_timer_start(borb_extractor *h)
{
    if (!h) return fail(BORB_ERR_ARG, "null");
    CU(cudaSetDevice(h->device));
    CU(cudaEventRecord(h->evt0, h->stream));
    return BORB_OK;
}
int borb_timer_stop(borb_extractor *h)
{
    if (!h) return fail(BORB_ERR_ARG, "null");
    CU(cudaSetDevice(h->device));
    CU(cudaEventRecord(h->evt1, h->stream));
    return BORB_OK;
}
/* milliseconds between the two marks on the handle's stream; synchronises on the stop mark */
float borb_timer_ms(borb_extractor *h)
{
    float ms = -1.f;
    if (!h || cudaSetDevice(h->device) != cudaSuccess || cudaEventSynchronize(h->evt1) != cudaSuccess ||
        cudaEventElapsedTime(&ms, h->evt0, h->evt1) != cudaSuccess) {
        cudaGetLastError();
        return -1.f;
    }
    return ms;
}

/* One un-graphed pass with an event between stages; ms[0..4] = pyramid (all levels), FAST,
 * quadtree, blur, orientation+descriptor+pack. */
int borb_stage_times(borb_extractor *h, const uint8_t *dev_img, int n, int cols, int rows,
                     size_t step, size_t stride, float *ms)
{
    if (h && h->pending) return fail(BORB_ERR_STATE, "a submitted call has not been waited for");
    int rc = check_call(h, dev_img, n, cols, rows, step, stride);
    if (rc != BORB_OK) return rc;
    if (!ms) return fail(BORB_ERR_ARG, "ms is null");
    rc = enqueue_kernels(h, dev_img, n, step, stride, cols, rows, true);
    if (rc != BORB_OK) return rc;
    CU(cudaStreamSynchronize(h->stream));
    for (int i = 0; i < 5; ++i) CU(cudaEventElapsedTime(&ms[i], h->evs[i], h->evs[i + 1]));
    h->last_n = n;
    note_new_extract(h);
    return BORB_OK;
}

int borb_extract_device(borb_extractor *h, const uint8_t *dev_img, int n, int cols, int rows,
                        size_t step, size_t stride)
{
    if (h && h->pending) return fail(BORB_ERR_STATE, "a submitted call has not been waited for");
    int rc = check_call(h, dev_img, n, cols, rows, step, stride);
    if (rc != BORB_OK) return rc;
    GraphKey key = {1, dev_img, n, step, stride, 0, cols, rows};
    CU(cudaEventRecord(h->ev0, h->stream));
    rc = run_cached(h, key, [&]() -> int { return enqueue_kernels(h, dev_img, n, step, stride, cols, rows); });
    if (rc != BORB_OK) return rc;
    CU(cudaEventRecord(h->ev1, h->stream));
    h->last_n = n;
    note_new_extract(h);
    return BORB_OK;
}

/* A frame that was extracted earlier (a KeyFrame: src/KeyFrame.cc copies mvKeys, mDescriptors and the grid from the Frame it
 * was made from) put back into the handle as if an extract call had just produced it: records and descriptors are uploaded
 * into slot 0 of the result arrays and, with a camera set, K8 rebuilds mvKeysUn, mGrid and the bounds from them (the same
 * arithmetic that filled the Frame's).  Every matcher then takes the handle as that frame. */
int borb_load_frame(borb_extractor *h, const borb_keypoint *kps, const uint8_t *desc, const float *u_right, int n, int cols,
                    int rows)
{
    if (!h || n < 0 || (n > 0 && (!kps || !desc)) || cols < 1 || rows < 1) return fail(BORB_ERR_ARG, "arguments");
    if (h->pending) return fail(BORB_ERR_STATE, "a submitted call has not been waited for");
    CU(cudaSetDevice(h->device));
    int rc = build_plan(h, cols, rows);
    if (rc != BORB_OK) return rc;
    if ((rc = ensure_frame_buffers(h)) != BORB_OK) return rc;
    const BorbPlan &P = h->plan;
    if (n > P.kp_total) return fail(BORB_ERR_CAPACITY, "%d keypoints; a frame of this extractor holds at most %d", n, P.kp_total);
    cudaStream_t st = h->stream;
    CU(cudaStreamSynchronize(st)); /* the page-locked twins are about to be overwritten */
    if (n) {
        std::memcpy(h->h_out_kp, kps, (size_t)n * sizeof(BorbKpRec));
        std::memcpy(h->h_out_desc, desc, (size_t)n * 32);
        CU(cudaMemcpyAsync(h->d_out_kp, h->h_out_kp, (size_t)n * sizeof(BorbKpRec), cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(h->d_out_desc, h->h_out_desc, (size_t)n * 32, cudaMemcpyHostToDevice, st));
    }
    h->h_nout[0] = n;
    h->h_err[0] = 0;
    CU(cudaMemcpyAsync(h->d_nout, h->h_nout, 4, cudaMemcpyHostToDevice, st));
    if (h->have_cam)
        k_frame_undistort_grid<FR_NT_LATENCY><<<1, FR_NT_LATENCY, 0, st>>>(h->d_plan, h->cam, h->d_out_kp, h->d_nout, h->d_un_xy,
                                                                          h->d_cell_start, h->d_cell_idx, h->d_cell_of, h->d_bounds);
    CU(cudaGetLastError());
    h->last_n = 1;
    note_new_extract(h);
    h->nout_frame_on_host = 0;
    if (u_right) { /* the keyframe's mvuRight, where borb_stereo_match would have left it (BORB_MATCH_USE_STEREO reads it) */
        const size_t B = (size_t)h->max_batch, K = (size_t)P.kp_total;
        if (!h->d_st_u) {
            CU(cudaMalloc(&h->d_st_u, B * K * 4));
            CU(cudaMalloc(&h->d_st_z, B * K * 4));
            CU(cudaMalloc(&h->d_st_kept, B * 4));
            CU(cudaMallocHost(&h->h_st_u, B * K * 4));
            CU(cudaMallocHost(&h->h_st_z, B * K * 4));
            CU(cudaMallocHost(&h->h_st_kept, B * 4));
        }
        for (size_t i = 0; i < K; ++i) h->h_st_u[i] = i < (size_t)n ? u_right[i] : -1.f;
        CU(cudaMemcpyAsync(h->d_st_u, h->h_st_u, K * 4, cudaMemcpyHostToDevice, st));
        h->stereo_gen = h->extract_gen;
    }
    return BORB_OK;
}

int borb_fetch(borb_extractor *h, unsigned flags)
{
    if (!h || !h->have_plan || h->last_n < 1) return fail(BORB_ERR_STATE, "nothing to fetch");
    CU(cudaSetDevice(h->device));
    int rc;
    if ((flags & BORB_WANT_PYRAMID) && (rc = ensure_pyr_host(h)) != BORB_OK) return rc;
    rc = enqueue_fetch(h, h->last_n, flags);
    if (rc != BORB_OK) return rc;
    CU(cudaStreamSynchronize(h->stream));
    h->results_on_host = true;
    h->pyramid_on_host = (flags & BORB_WANT_PYRAMID) != 0;
    return check_device_err(h);
}

int borb_sync(borb_extractor *h)
{
    if (!h) return fail(BORB_ERR_ARG, "null");
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    return BORB_OK;
}

float borb_last_device_ms(const borb_extractor *h)
{
    float ms = -1.f;
    if (!h || cudaEventElapsedTime(&ms, h->ev0, h->ev1) != cudaSuccess) {
        cudaGetLastError();
        return -1.f;
    }
    return ms;
}

int borb_set_stage_timing(borb_extractor *h, int on)
{
    if (!h) return fail(BORB_ERR_ARG, "null");
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    if ((on != 0) != h->stage_timing) { /* graphs captured under the other setting are stale */
        for (GraphEntry &g : h->graphs) cudaGraphExecDestroy(g.exec);
        h->graphs.clear();
    }
    h->stage_timing = on != 0;
    return BORB_OK;
}

int borb_last_stage_ms(borb_extractor *h, float *ms)
{
    if (!h || !ms) return fail(BORB_ERR_ARG, "null");
    if (!h->stage_timing) return fail(BORB_ERR_STATE, "stage timing is off");
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    for (int i = 0; i < 5; ++i) CU(cudaEventElapsedTime(&ms[i], h->evs[i], h->evs[i + 1]));
    return BORB_OK;
}

float borb_last_prestep_ms(borb_extractor *h)
{
    float ms = -1.f;
    if (!h || !h->stage_timing || !pre_active(h)) return -1.f;
    if (cudaSetDevice(h->device) != cudaSuccess || cudaStreamSynchronize(h->stream) != cudaSuccess ||
        cudaEventElapsedTime(&ms, h->evs[0], h->evs[6]) != cudaSuccess) {
        cudaGetLastError();
        return -1.f;
    }
    return ms;
}

int borb_launches_per_call(const borb_extractor *h)
{
    if (!h) return 0;
    return h->last_launches ? h->last_launches : h->nlevels + 4 + (pre_active(h) ? 1 : 0) + (h->have_cam ? 1 : 0);
}

int borb_result(const borb_extractor *h, int frame, const borb_keypoint **kps, const uint8_t **desc, int *n)
{
    if (!h || !h->results_on_host) return fail(BORB_ERR_STATE, "no results on the host");
    if (frame < 0 || frame >= h->last_n) return fail(BORB_ERR_ARG, "frame %d of %d", frame, h->last_n);
    const size_t o = (size_t)frame * h->plan.kp_total;
    if (kps) *kps = reinterpret_cast<const borb_keypoint *>(h->h_out_kp + o);
    if (desc) *desc = h->h_out_desc + o * 32;
    if (n) *n = h->h_nout[frame];
    return BORB_OK;
}

int borb_counts(const borb_extractor *h, int32_t *out)
{
    if (!h || !out || !h->results_on_host) return fail(BORB_ERR_STATE, "no results on the host");
    std::memcpy(out, h->h_nout, (size_t)h->last_n * 4);
    return BORB_OK;
}

/* -- the per-keypoint step after the extractor (K8) -- */
int borb_set_camera(borb_extractor *h, const float *K4, const float *dist, int n_dist)
{
    if (!h) return fail(BORB_ERR_ARG, "null");
    if (h->pending) return fail(BORB_ERR_STATE, "a submitted call has not been waited for");
    if (K4 && (n_dist < 0 || n_dist > 5 || (n_dist > 0 && !dist) || !(K4[0] != 0.f) || !(K4[1] != 0.f)))
        return fail(BORB_ERR_ARG, "bad camera");
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    for (GraphEntry &g : h->graphs) cudaGraphExecDestroy(g.exec); /* the camera is a kernel argument */
    h->graphs.clear();
    h->frame_on_host = false;
    h->have_cam = K4 != nullptr;
    if (!K4) return BORB_OK;
    h->cam.fx = K4[0]; h->cam.fy = K4[1]; h->cam.cx = K4[2]; h->cam.cy = K4[3];
    for (int i = 0; i < 5; ++i) h->cam.k[i] = i < n_dist ? dist[i] : 0.f;
    return BORB_OK;
}

int borb_frame_grid(const borb_extractor *h, int frame, const float **xy_un, const int32_t **cell_start,
                    const int32_t **cell_idx, float *bounds)
{
    if (!h || !h->have_cam || !h->results_on_host || !h->h_un_xy)
        return fail(BORB_ERR_STATE, "no camera set, or no results on the host");
    if (frame < 0 || frame >= h->last_n) return fail(BORB_ERR_ARG, "frame %d of %d", frame, h->last_n);
    const size_t o = (size_t)frame * h->plan.kp_total;
    if (xy_un) *xy_un = reinterpret_cast<const float *>(h->h_un_xy + o);
    if (cell_start) *cell_start = h->h_cell_start + (size_t)frame * (FR_CELLS + 1);
    if (cell_idx) *cell_idx = h->h_cell_idx + o;
    if (bounds) std::memcpy(bounds, h->h_bounds + (size_t)frame * 4, 16);
    return BORB_OK;
}

/* -- the step before the extractor (K0) -- */
int borb_set_input_format(borb_extractor *h, int format)
{
    if (!h) return fail(BORB_ERR_ARG, "null");
    if (format < BORB_GRAY8 || format > BORB_RGBA8) return fail(BORB_ERR_ARG, "unknown pixel format %d", format);
    if (h->pending) return fail(BORB_ERR_STATE, "a submitted call has not been waited for");
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    for (GraphEntry &g : h->graphs) cudaGraphExecDestroy(g.exec);
    h->graphs.clear();
    h->in_format = format;
    h->in_ch = format == BORB_GRAY8 ? 1 : (format == BORB_BGR8 || format == BORB_RGB8) ? 3 : 4;
    return BORB_OK;
}

int borb_set_rectify_map(borb_extractor *h, const float *map_x, const float *map_y, int cols, int rows,
                         size_t map_step)
{
    if (!h) return fail(BORB_ERR_ARG, "null");
    if (h->pending) return fail(BORB_ERR_STATE, "a submitted call has not been waited for");
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    for (GraphEntry &g : h->graphs) cudaGraphExecDestroy(g.exec);
    h->graphs.clear();
    if (h->d_map_xy) cudaFree(h->d_map_xy);
    if (h->d_map_fr) cudaFree(h->d_map_fr);
    h->d_map_xy = nullptr;
    h->d_map_fr = nullptr;
    h->map_cols = h->map_rows = h->map_pitch = 0;
    if (!map_x && !map_y) return BORB_OK; /* map cleared */
    if (!map_x || !map_y || cols < 1 || rows < 1 || map_step < (size_t)cols)
        return fail(BORB_ERR_ARG, "bad rectification map");
    /* cv::remap's own first step (convertMaps to 1/32-pixel fixed point), done once here:
     * s = cvRound(map * 32) with the product in float; values outside int range (and NaN) go far
     * outside every image, where all four taps read the constant border */
    const int pitch = (cols + 3) & ~3;
    std::vector<uint32_t> xy((size_t)pitch * rows, 0x80008000u);
    std::vector<uint16_t> fr((size_t)pitch * rows, 0);
    auto fix = [](float m) -> int {
        const float v = m * 32.f;
        if (!(v > -1073741824.f)) return -1073741824; /* also NaN */
        if (v > 1073741824.f) return 1073741824;
        return (int)lrintf(v);
    };
    auto sat = [](int v) -> int { return v < -32768 ? -32768 : v > 32767 ? 32767 : v; };
    for (int y = 0; y < rows; ++y)
        for (int x = 0; x < cols; ++x) {
            const int sx = fix(map_x[(size_t)y * map_step + x]), sy = fix(map_y[(size_t)y * map_step + x]);
            xy[(size_t)y * pitch + x] = (uint32_t)(sat(sx >> 5) & 0xffff) | ((uint32_t)(sat(sy >> 5) & 0xffff) << 16);
            fr[(size_t)y * pitch + x] = (uint16_t)(((sy & 31) << 5) | (sx & 31));
        }
    CU(cudaMalloc(&h->d_map_xy, xy.size() * 4));
    CU(cudaMalloc(&h->d_map_fr, fr.size() * 2));
    CU(cudaMemcpy(h->d_map_xy, xy.data(), xy.size() * 4, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(h->d_map_fr, fr.data(), fr.size() * 2, cudaMemcpyHostToDevice));
    h->map_cols = cols;
    h->map_rows = rows;
    h->map_pitch = pitch;
    return BORB_OK;
}

namespace {

/* scratch of the stereo association on the left handle: SADs, per-row candidate tables (allocated on first use) */
int ensure_stereo_scratch(borb_extractor *l)
{
    if (l->d_st_sad) return BORB_OK;
    const BorbPlan &P = l->plan;
    const size_t B = (size_t)l->max_batch, K = (size_t)P.kp_total;
    CU(cudaMalloc(&l->d_st_sad, B * K * 4));
    l->st_span = (int)(4.f * l->scale[l->nlevels - 1]) + 3;
    CU(cudaMalloc(&l->d_st_rows, B * (size_t)(P.rows + 1) * 4));
    CU(cudaMalloc(&l->d_st_ent, B * K * (size_t)l->st_span * 2));
    CU(cudaFuncSetAttribute(k_stereo_rows, cudaFuncAttributeMaxDynamicSharedMemorySize,
                            (int)((2 * (size_t)P.rows + ST_ROWS_NT) * 4)));
    return BORB_OK;
}

int check_stereo_pair(const borb_extractor *l, const borb_extractor *r)
{
    if (l->device != r->device) return fail(BORB_ERR_ARG, "handles live on devices %d and %d", l->device, r->device);
    if (l->plan.cols != r->plan.cols || l->plan.rows != r->plan.rows || l->nlevels != r->nlevels ||
        l->scale_factor != r->scale_factor || l->plan.kp_total != r->plan.kp_total)
        return fail(BORB_ERR_ARG, "left and right extractor differ in geometry or pyramid parameters");
    /* S0 stores right-keypoint indices as 16-bit entries and S1 packs the index into the low 16 bits of its key */
    if (l->plan.kp_total > 65535)
        return fail(BORB_ERR_ARG, "the stereo association packs keypoint indices in 16 bits: %d keypoint slots per frame",
                    l->plan.kp_total);
    return BORB_OK;
}

/* S0 + S1 + S2 over the n frames whose results the two handles' (currently bound) result slabs hold; outputs to
 * u / z / kept.  Runs on l's stream, ordered after r's; r's next call is ordered after it. */
int enqueue_stereo(borb_extractor *l, borb_extractor *r, int n, float mbf, float mb, float *u, float *z, int *kept)
{
    const BorbPlan &P = l->plan;
    const size_t K = (size_t)P.kp_total;
    if (!r->ev_peer) CU(cudaEventCreateWithFlags(&r->ev_peer, cudaEventDisableTiming));
    if (!l->ev_peer) CU(cudaEventCreateWithFlags(&l->ev_peer, cudaEventDisableTiming));
    /* the right extractor's kernels run on its own stream: order the match after them */
    CU(cudaEventRecord(r->ev_peer, r->stream));
    CU(cudaStreamWaitEvent(l->stream, r->ev_peer, 0));
    const dim3 grid((unsigned)((K + ST_NT / 32 - 1) / (ST_NT / 32)), (unsigned)n);
    k_stereo_rows<<<n, ST_ROWS_NT, (2 * (size_t)P.rows + ST_ROWS_NT) * 4, l->stream>>>(
        l->d_plan, r->d_out_kp, r->d_nout, l->st_span, l->d_st_rows, l->d_st_ent);
    k_stereo_match<<<grid, ST_NT, 0, l->stream>>>(l->d_plan, l->d_out_kp, l->d_out_desc, l->d_nout, r->d_out_kp,
                                                  r->d_out_desc, r->d_nout, l->st_span, l->d_st_rows, l->d_st_ent,
                                                  l->d_pyr, r->d_pyr, mbf, mb, u, z, l->d_st_sad);
    k_stereo_cut<<<n, ST_CUT_NT, 0, l->stream>>>(l->d_plan, l->d_nout, u, z, l->d_st_sad, kept);
    CU(cudaGetLastError());
    /* the right handle must not start its next call before the match has read its buffers */
    CU(cudaEventRecord(l->ev_peer, l->stream));
    CU(cudaStreamWaitEvent(r->stream, l->ev_peer, 0));
    return BORB_OK;
}

} // namespace

/* Frame::ComputeStereoMatches (src/Frame.cc:619-834) on the two handles' device-resident results */
int borb_stereo_match(borb_extractor *l, borb_extractor *r, float mbf, float mb, unsigned flags)
{
    if (!l || !r || l == r) return fail(BORB_ERR_ARG, "two distinct handles are needed");
    if (!l->have_plan || !r->have_plan || l->last_n < 1 || l->last_n != r->last_n)
        return fail(BORB_ERR_STATE, "both handles need a finished extract call over the same number of frames");
    int rc = check_stereo_pair(l, r);
    if (rc != BORB_OK) return rc;
    if (!(mbf > 0.f) || !(mb > 0.f)) return fail(BORB_ERR_ARG, "mbf %g, mb %g", (double)mbf, (double)mb);
    CU(cudaSetDevice(l->device));
    const BorbPlan &P = l->plan;
    const size_t B = (size_t)l->max_batch, K = (size_t)P.kp_total;
    if ((rc = ensure_stereo_scratch(l)) != BORB_OK) return rc;
    if (!l->d_st_u) {
        CU(cudaMalloc(&l->d_st_u, B * K * 4));
        CU(cudaMalloc(&l->d_st_z, B * K * 4));
        CU(cudaMalloc(&l->d_st_kept, B * 4));
        CU(cudaMallocHost(&l->h_st_u, B * K * 4));
        CU(cudaMallocHost(&l->h_st_z, B * K * 4));
        CU(cudaMallocHost(&l->h_st_kept, B * 4));
    }
    const int n = l->last_n;
    if ((rc = enqueue_stereo(l, r, n, mbf, mb, l->d_st_u, l->d_st_z, l->d_st_kept)) != BORB_OK) return rc;
    l->stereo_on_host = false;
    l->stereo_gen = l->extract_gen; /* mvuRight / mvDepth are parallel to THIS call's keypoints */
    if (flags & BORB_STEREO_KEEP_ON_DEVICE) return BORB_OK;
    CU(cudaMemcpyAsync(l->h_st_u, l->d_st_u, (size_t)n * K * 4, cudaMemcpyDeviceToHost, l->stream));
    CU(cudaMemcpyAsync(l->h_st_z, l->d_st_z, (size_t)n * K * 4, cudaMemcpyDeviceToHost, l->stream));
    CU(cudaMemcpyAsync(l->h_st_kept, l->d_st_kept, (size_t)n * 4, cudaMemcpyDeviceToHost, l->stream));
    CU(cudaMemcpyAsync(l->h_nout, l->d_nout, (size_t)n * 4, cudaMemcpyDeviceToHost, l->stream));
    CU(cudaStreamSynchronize(l->stream));
    l->stereo_on_host = true;
    return BORB_OK;
}

int borb_stereo_result(const borb_extractor *l, int frame, const float **u_right, const float **depth,
                       int *n, int *n_matched)
{
    if (!l || !l->stereo_on_host) return fail(BORB_ERR_STATE, "no stereo result on the host");
    if (frame < 0 || frame >= l->last_n) return fail(BORB_ERR_ARG, "frame %d of %d", frame, l->last_n);
    const size_t o = (size_t)frame * l->plan.kp_total;
    if (u_right) *u_right = l->h_st_u + o;
    if (depth) *depth = l->h_st_z + o;
    if (n) *n = l->h_nout[frame];
    if (n_matched) *n_matched = l->h_st_kept[frame];
    return BORB_OK;
}

static int g_match_max_rounds = 0; /* 0 = MT_MAX_ROUNDS; borb_debug_set_match_rounds */

namespace {

/* Ow = -Rcw.t() * tcw as cv::MatExpr evaluates it (src/ORBmatcher.cc:1516, :1676; src/Frame.cc:334): cv::gemm(Rcw, tcw,
 * -1, noArray(), 0, GEMM_1_T) -- OpenCV's generic path, products and sum in double, left to right, times alpha, one
 * rounding to float (pinned against cv2 4.13.0 in tests/test_oracle_match.py). */
void camera_centre(const float *Tcw /* 3 x 4 row-major */, float *Ow)
{
    for (int y = 0; y < 3; ++y) {
        double s = 0.0;
        for (int k = 0; k < 3; ++k) s += (double)Tcw[4 * k + y] * (double)Tcw[4 * k + 3];
        Ow[y] = (float)(s * -1.0);
    }
}

/* Both matchers (k_match.cuh): upload the query side's packed arrays [desc 32c][geo 16c][angle 4c][octave 4c][flags c]
 * (+ the occupancy bytes of the local-map matcher), M1a over the whole GPU, M1b in one CTA, download the matches. */
int run_matcher(borb_extractor *cur, int frame, int mode, int n, const float *geo, int geo_stride, const uint8_t *desc,
                const uint8_t *qflags, const int32_t *octave, const float *angle, const uint8_t *occupied, MatchParams &mp,
                bool use_stereo, int32_t *match, int *n_cur, int *nmatches, const float *extra = nullptr /* 4 floats per query */)
{
    if (!cur->have_plan || cur->last_n < 1) return fail(BORB_ERR_STATE, "no finished extract call");
    if (!cur->have_cam || !cur->d_un_xy) return fail(BORB_ERR_STATE, "borb_set_camera must be on: the matcher reads mvKeysUn, mGrid and K");
    if (cur->pending) return fail(BORB_ERR_STATE, "a submitted call has not been waited for");
    if (frame < 0 || frame >= cur->last_n) return fail(BORB_ERR_ARG, "frame %d of %d", frame, cur->last_n);
    if (n < 0 || (n > 0 && (!geo || !desc || !qflags || !octave))) return fail(BORB_ERR_ARG, "query arrays");
    if (use_stereo && (!cur->d_st_u || cur->stereo_gen != cur->extract_gen))
        return fail(BORB_ERR_STATE, "no stereo result for this handle's last extract call (run borb_stereo_match after it)");
    for (int i = 0; i < n && mode < 4; ++i) /* modes >= 4 predict the level; their octave column carries mfMinDistance */
        if ((qflags[i] & 1) && (octave[i] < 0 || octave[i] >= cur->nlevels))
            return fail(BORB_ERR_ARG, "octave %d of query %d", octave[i], i);
    CU(cudaSetDevice(cur->device));
    const BorbPlan &P = cur->plan;
    const size_t K = (size_t)P.kp_total;
    if (K > 65535) return fail(BORB_ERR_ARG, "the matcher packs list positions in 16 bits: %zu keypoint slots per frame", K);
    if (!cur->d_mt_match) {
        CU(cudaMalloc(&cur->d_mt_taker, 2 * K * 4));
        CU(cudaMalloc(&cur->d_mt_match, (K + 4) * 4));
        CU(cudaMallocHost(&cur->h_mt_match, (K + 4) * 4));
        CU(cudaMalloc(&cur->d_mt_occ, K));
        CU(cudaMallocHost(&cur->h_mt_occ, K));
    }
    if ((size_t)n > cur->mt_cap || !cur->d_mt_last) {
        const size_t cap = (((size_t)std::max(n, 1) + 1023) / 1024) * 1024;
        CU(cudaStreamSynchronize(cur->stream));
        if (cur->d_mt_last) cudaFree(cur->d_mt_last);
        if (cur->h_mt_last) cudaFreeHost(cur->h_mt_last);
        if (cur->d_mt_choice) cudaFree(cur->d_mt_choice);
        if (cur->d_mt_cache) cudaFree(cur->d_mt_cache);
        if (cur->d_mt_more) cudaFree(cur->d_mt_more);
        if (cur->d_mt_m12) cudaFree(cur->d_mt_m12);
        if (cur->d_mt_ent) cudaFree(cur->d_mt_ent);
        cur->d_mt_ent = nullptr;
        if (cur->h_mt_m12) cudaFreeHost(cur->h_mt_m12);
        cur->d_mt_m12 = cur->h_mt_m12 = nullptr;
        cur->d_mt_last = cur->h_mt_last = nullptr; cur->d_mt_choice = nullptr; cur->mt_cap = 0;
        cur->d_mt_cache = nullptr; cur->d_mt_more = nullptr;
        CU(cudaMalloc(&cur->d_mt_last, cap * 73));
        CU(cudaMallocHost(&cur->h_mt_last, cap * 73));
        CU(cudaMalloc(&cur->d_mt_choice, cap * 8));
        CU(cudaMalloc(&cur->d_mt_m12, cap * 4));
        CU(cudaMalloc(&cur->d_mt_ent, 2 * cap * 4));
        CU(cudaMallocHost(&cur->h_mt_m12, cap * 4));
        CU(cudaMalloc(&cur->d_mt_cache, cap * MT_K * 4));
        CU(cudaMalloc(&cur->d_mt_more, cap));
        cur->mt_cap = cap;
    }
    const size_t c = cur->mt_cap, un = (size_t)n;
    unsigned char *hp = cur->h_mt_last, *dp = cur->d_mt_last;
    if (n) {
        std::memcpy(hp, desc, un * 32);
        float *g = reinterpret_cast<float *>(hp + 32 * c);
        if (geo_stride == 4) std::memcpy(g, geo, un * 16);
        else if (geo_stride == 2)
            for (size_t i = 0; i < un; ++i) { g[4 * i] = geo[2 * i]; g[4 * i + 1] = geo[2 * i + 1]; g[4 * i + 2] = 0.f; g[4 * i + 3] = 0.f; }
        else
            for (size_t i = 0; i < un; ++i) { g[4 * i] = geo[3 * i]; g[4 * i + 1] = geo[3 * i + 1]; g[4 * i + 2] = geo[3 * i + 2]; g[4 * i + 3] = 0.f; }
        if (angle) std::memcpy(hp + 48 * c, angle, un * 4);
        std::memcpy(hp + 52 * c, octave, un * 4);
        std::memcpy(hp + 56 * c, qflags, un);
        if (extra) std::memcpy(hp + 57 * c, extra, un * 16);
    }
    cudaStream_t st = cur->stream;
    /* one contiguous upload when the arrays are nearly full, else one per array */
    if (un * 2 >= c) CU(cudaMemcpyAsync(dp, hp, (extra ? 73 : 57) * c, cudaMemcpyHostToDevice, st));
    else if (n) {
        if (extra) CU(cudaMemcpyAsync(dp + 57 * c, hp + 57 * c, un * 16, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(dp, hp, un * 32, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(dp + 32 * c, hp + 32 * c, un * 16, cudaMemcpyHostToDevice, st));
        if (angle) CU(cudaMemcpyAsync(dp + 48 * c, hp + 48 * c, un * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(dp + 52 * c, hp + 52 * c, un * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(dp + 56 * c, hp + 56 * c, un, cudaMemcpyHostToDevice, st));
    }
    const uint8_t *d_occ = nullptr;
    if (occupied) { /* the caller's view of F.mvpMapPoints on entry; sized by the frame's keypoint count, which is on the host */
        if (!cur->results_on_host && cur->nout_frame_on_host != frame) { /* else the count is on the host already */
            CU(cudaMemcpyAsync(cur->h_nout + frame, cur->d_nout + frame, 4, cudaMemcpyDeviceToHost, st));
            CU(cudaStreamSynchronize(st));
            cur->nout_frame_on_host = frame;
        }
        const size_t nc0 = (size_t)std::min(cur->h_nout[frame], (int)K);
        std::memset(cur->h_mt_occ, 0, K);
        std::memcpy(cur->h_mt_occ, occupied, nc0);
        CU(cudaMemcpyAsync(cur->d_mt_occ, cur->h_mt_occ, K, cudaMemcpyHostToDevice, st));
        d_occ = cur->d_mt_occ;
    }
    mp.fx = cur->cam.fx; mp.fy = cur->cam.fy; mp.cx = cur->cam.cx; mp.cy = cur->cam.cy;
    mp.n_last = n;
    mp.max_rounds = g_match_max_rounds;
    for (int l = 0; l < BORB_MAXL; ++l) mp.scale[l] = l < cur->nlevels ? cur->scale[l] : 0.f;
    const size_t fo = (size_t)frame * K;
    const float *d_ur = use_stereo ? cur->d_st_u + fo : nullptr;
    const float *l_geo = reinterpret_cast<const float *>(dp + 32 * c), *l_angle = reinterpret_cast<const float *>(dp + 48 * c);
    const int *l_oct = reinterpret_cast<const int *>(dp + 52 * c);
    const uint4 *l_desc = reinterpret_cast<const uint4 *>(dp);
    const uint8_t *l_flags = dp + 56 * c;
    const float4 *l_extra = reinterpret_cast<const float4 *>(dp + 57 * c);
    const int *cs = cur->d_cell_start + (size_t)frame * (FR_CELLS + 1), *ci = cur->d_cell_idx + fo;
    const int cand_grid = (n + MT_CAND_NT / 32 - 1) / (MT_CAND_NT / 32);
    const size_t tk_smem = 2 * K * 4 <= 96 * 1024 ? 2 * K * 4 : 0; /* takers in shared memory when they fit */
    if (tk_smem > 48 * 1024) {
        CU(cudaFuncSetAttribute(k_match_resolve<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tk_smem));
        CU(cudaFuncSetAttribute(k_match_resolve<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tk_smem));
        CU(cudaFuncSetAttribute(k_match_resolve<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tk_smem));
        CU(cudaFuncSetAttribute(k_match_resolve<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tk_smem));
    }
#define MATCH_LAUNCH(MODE)                                                                                                   \
    do {                                                                                                                     \
        if (n > 0)                                                                                                           \
            k_match_candidates<MODE><<<cand_grid, MT_CAND_NT, 0, st>>>(                                                       \
                mp, cur->d_out_kp + fo, cur->d_out_desc + fo * 32, cur->d_un_xy + fo, cs, ci, cur->d_bounds + (size_t)frame * 4, \
                d_ur, d_occ, l_geo, l_oct, l_desc, l_flags, cur->d_mt_cache, cur->d_mt_more, l_extra);                        \
        k_match_resolve<MODE><<<1, MT_NT, tk_smem, st>>>(mp, cur->d_out_kp + fo, cur->d_out_desc + fo * 32, cur->d_un_xy + fo, cs, ci, \
                                                   cur->d_bounds + (size_t)frame * 4, cur->d_nout + frame, d_ur, d_occ, l_geo,  \
                                                   l_angle, l_oct, l_desc, l_flags, cur->d_mt_cache, cur->d_mt_more,           \
                                                   cur->d_mt_choice, cur->d_mt_taker, (int)K, cur->d_mt_match,                 \
                                                   cur->d_mt_match + K, tk_smem ? 1 : 0, l_extra);                           \
    } while (0)
    if (mode == 5 || mode == 6) {
        /* the matchers whose queries are independent (Fuse: the choice per map point; a direction of SearchBySim3): M1a
         * alone; the best cached entry of a query IS its choice -- first smallest distance in GetFeaturesInArea order */
        if (n > 0 && mode == 5)
            k_match_candidates<5><<<cand_grid, MT_CAND_NT, 0, st>>>(
                mp, cur->d_out_kp + fo, cur->d_out_desc + fo * 32, cur->d_un_xy + fo, cs, ci, cur->d_bounds + (size_t)frame * 4,
                d_ur, nullptr, l_geo, l_oct, l_desc, l_flags, cur->d_mt_cache, cur->d_mt_more, l_extra);
        else if (n > 0)
            k_match_candidates<6><<<cand_grid, MT_CAND_NT, 0, st>>>(
                mp, cur->d_out_kp + fo, cur->d_out_desc + fo * 32, cur->d_un_xy + fo, cs, ci, cur->d_bounds + (size_t)frame * 4,
                nullptr, nullptr, l_geo, l_oct, l_desc, l_flags, cur->d_mt_cache, cur->d_mt_more, l_extra);
        CU(cudaGetLastError());
        CU(cudaMemcpy2DAsync(cur->h_mt_m12, 4, cur->d_mt_cache, MT_K * 4, 4, std::max<size_t>(un, 1), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        int nm = 0;
        for (size_t i = 0; i < un; ++i) {
            const uint32_t key = (uint32_t)cur->h_mt_m12[i];
            const bool ok = key != MT_NOKEY && (int)(key >> 16) <= mp.orb_dist;
            match[i] = ok ? (int32_t)(key & 0xffffu) : -1;
            nm += ok;
        }
        if (nmatches) *nmatches = nm;
        if (n_cur) *n_cur = n;
        cur->mt_rounds = 1;
        cur->mt_rescans = 0;
        return BORB_OK;
    }
    if (mode == 1) MATCH_LAUNCH(1);
    else if (mode == 2) MATCH_LAUNCH(2);
    else if (mode == 4) MATCH_LAUNCH(4);
    else if (mode == 7) MATCH_LAUNCH(6); /* MODE 6 queries with the sequential rule: SearchByProjection(pKF, Scw, ...) */
    else {
        if (!cur->d_mt_cnt) {
            CU(cudaMalloc(&cur->d_mt_cnt, 2 * K * 4));
            CU(cudaMallocHost(&cur->h_mt_unxy, K * 8));
        }
        if (n > 0)
            k_match_candidates<3><<<cand_grid, MT_CAND_NT, 0, st>>>(
                mp, cur->d_out_kp + fo, cur->d_out_desc + fo * 32, cur->d_un_xy + fo, cs, ci, cur->d_bounds + (size_t)frame * 4,
                nullptr, nullptr, l_geo, l_oct, l_desc, l_flags, cur->d_mt_cache, cur->d_mt_more);
        k_match_resolve_init<<<1, MT_NT, 0, st>>>(mp, cur->d_out_kp + fo, cur->d_out_desc + fo * 32, cur->d_un_xy + fo, cs, ci,
                                                  cur->d_bounds + (size_t)frame * 4, l_geo, l_angle, l_oct, l_desc, l_flags,
                                                  cur->d_mt_cache, cur->d_mt_more, cur->d_mt_choice, cur->d_mt_ent, (int)c, cur->d_mt_cnt,
                                                  (int)K, cur->d_mt_match, cur->d_mt_m12, cur->d_mt_match + K);
        CU(cudaGetLastError());
        CU(cudaMemcpyAsync(cur->h_mt_m12, cur->d_mt_m12, std::max<size_t>(un, 1) * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(cur->h_mt_match + K, cur->d_mt_match + K, 16, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(cur->h_mt_unxy, cur->d_un_xy + fo, K * 8, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        cur->mt_rounds = cur->h_mt_match[K + 1];
        cur->mt_rescans = cur->h_mt_match[K + 2];
        if (n) std::memcpy(match, cur->h_mt_m12, un * 4);
        if (nmatches) *nmatches = cur->h_mt_match[K];
        if (n_cur) *n_cur = n;
        return BORB_OK;
    }
#undef MATCH_LAUNCH
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(cur->h_mt_match, cur->d_mt_match, (K + 4) * 4, cudaMemcpyDeviceToHost, st));
    if (!cur->results_on_host && cur->nout_frame_on_host != frame) /* else the frame's keypoint count is on the host already */
        CU(cudaMemcpyAsync(cur->h_nout + frame, cur->d_nout + frame, 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    cur->nout_frame_on_host = frame;
    const int nc = std::min(cur->h_nout[frame], (int)K);
    std::memcpy(match, cur->h_mt_match, (size_t)nc * 4);
    if (n_cur) *n_cur = nc;
    if (nmatches) *nmatches = cur->h_mt_match[K];
    cur->mt_rounds = cur->h_mt_match[K + 1];
    cur->mt_rescans = cur->h_mt_match[K + 2];
    return BORB_OK;
}

} // namespace

/* ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (src/ORBmatcher.cc:1503-1668) against the
 * device-resident results of frame `frame` of cur's last call */
int borb_search_by_projection(borb_extractor *cur, int frame, const borb_last_frame *last, const float *Tcw_cur,
                              float mbf, float mb, float th, unsigned flags, int32_t *match, int *n_cur,
                              int *nmatches)
{
    if (!cur || !last || !Tcw_cur || !match) return fail(BORB_ERR_ARG, "null argument");
    if (!last->Tcw || (last->n > 0 && !last->angle)) return fail(BORB_ERR_ARG, "last frame arrays");
    MatchParams mp;
    std::memset(&mp, 0, sizeof mp);
    std::memcpy(mp.Tc, Tcw_cur, 48);
    std::memcpy(mp.Tl, last->Tcw, 48);
    mp.mbf = mbf; mp.mb = mb; mp.th = th;
    mp.mono = (flags & BORB_MATCH_MONO) ? 1 : 0;
    mp.check_ori = (flags & BORB_MATCH_NO_ORIENTATION) ? 0 : 1;
    {   /* search direction (:1515-1526): twc = -Rcw.t() * tcw is cv::gemm with GEMM_1_T and alpha = -1 (the sum over k
         * in double, times alpha, one rounding: camera_centre), tlc = Rlw twc + tlw OpenCV's small float product
         * (products summed left to right); this file is compiled without contraction */
        const float *Tc = mp.Tc, *Tl = mp.Tl;
        float twc[3];
        camera_centre(Tc, twc);
        float sacc = 0.f;
        for (int k = 0; k < 3; ++k) sacc += Tl[8 + k] * twc[k];
        const float tlc_z = sacc + Tl[11];
        const bool fwd = tlc_z > mb && !mp.mono, bwd = -tlc_z > mb && !mp.mono;
        mp.dir = fwd ? 1 : bwd ? 2 : 0;
    }
    return run_matcher(cur, frame, 1, last->n, last->xyz, 3, last->desc, last->flags, last->octave, last->angle, nullptr, mp,
                       (flags & BORB_MATCH_USE_STEREO) != 0, match, n_cur, nmatches);
}

/* ORBmatcher::SearchByProjection(F, vpMapPoints, th) (src/ORBmatcher.cc:59-154), the matcher of Tracking::SearchLocalPoints */
int borb_search_local_points(borb_extractor *cur, int frame, const borb_map_points *mps, const uint8_t *occupied, float th,
                             float nnratio, unsigned flags, int32_t *match, int *n_cur, int *nmatches)
{
    if (!cur || !mps || !match) return fail(BORB_ERR_ARG, "null argument");
    MatchParams mp;
    std::memset(&mp, 0, sizeof mp);
    mp.th = th;
    mp.nnratio = nnratio;
    return run_matcher(cur, frame, 2, mps->n, mps->proj, 4, mps->desc, mps->flags, mps->level, nullptr, occupied, mp,
                       (flags & BORB_MATCH_USE_STEREO) != 0, match, n_cur, nmatches);
}

/* ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:1670-1798), the
 * relocalisation matcher, against the device-resident results of frame `frame` of cur's last call.  The keyframe's map
 * points become the queries of M1a / M1b in MODE 4 (k_match.cuh): every set slot is closed to every query, so all queries
 * carry the "holds its slot" bit and the caller's occupancy bytes close the slots set on entry. */
int borb_search_by_projection_keyframe(borb_extractor *cur, int frame, const borb_keyframe_points *kf, const float *Tcw_cur,
                                       const uint8_t *occupied, float th, int orb_dist, unsigned flags, int32_t *match,
                                       int *n_cur, int *nmatches)
{
    if (!cur || !kf || !Tcw_cur || !match) return fail(BORB_ERR_ARG, "null argument");
    if (orb_dist < 0 || orb_dist > 255) return fail(BORB_ERR_ARG, "ORBdist %d: a 256-bit Hamming distance is 0..256 and the reference's best starts at 256", orb_dist);
    if (kf->n > 0 && (!kf->xyz || !kf->desc || !kf->flags || !kf->dist || !kf->angle)) return fail(BORB_ERR_ARG, "keyframe arrays");
    MatchParams mp;
    std::memset(&mp, 0, sizeof mp);
    std::memcpy(mp.Tc, Tcw_cur, 48);
    mp.th = th;
    mp.mono = 1;
    mp.check_ori = (flags & BORB_MATCH_NO_ORIENTATION) ? 0 : 1;
    mp.orb_dist = orb_dist;
    mp.nlevels = cur->nlevels;
    mp.log_scale = orb_logf_exact((float)cur->scale_factor); /* Frame: mfLogScaleFactor = log(mfScaleFactor), the float overload */
    camera_centre(mp.Tc, mp.Ow);
    const size_t n = (size_t)std::max(kf->n, 0);
    std::vector<float> geo(4 * n);
    std::vector<int32_t> mind(n);
    std::vector<uint8_t> qf(n), occ;
    for (size_t i = 0; i < n; ++i) {
        geo[4 * i] = kf->xyz[3 * i]; geo[4 * i + 1] = kf->xyz[3 * i + 1]; geo[4 * i + 2] = kf->xyz[3 * i + 2];
        geo[4 * i + 3] = kf->dist[2 * i + 1];
        std::memcpy(&mind[i], &kf->dist[2 * i], 4);
        qf[i] = ((kf->flags[i] & 1) && !(kf->flags[i] & 6)) ? 3 : 0; /* a query, and it closes the slot it takes */
    }
    if (occupied) { /* any set slot is skipped (:1735-1736): the kernels close the slots marked 2 */
        if (!cur->have_plan || cur->last_n < 1 || frame < 0 || frame >= cur->last_n)
            return fail(BORB_ERR_STATE, "no finished extract call holds frame %d", frame);
        if (!cur->results_on_host && cur->nout_frame_on_host != frame) { /* the frame's keypoint count sizes the array */
            CU(cudaSetDevice(cur->device));
            CU(cudaMemcpyAsync(cur->h_nout + frame, cur->d_nout + frame, 4, cudaMemcpyDeviceToHost, cur->stream));
            CU(cudaStreamSynchronize(cur->stream));
            cur->nout_frame_on_host = frame;
        }
        const int nc0 = std::min(cur->h_nout[frame], cur->plan.kp_total);
        occ.assign((size_t)cur->plan.kp_total, 0);
        for (int i = 0; i < nc0; ++i) occ[i] = occupied[i] ? 2 : 0;
    }
    return run_matcher(cur, frame, 4, kf->n, geo.data(), 4, kf->desc, qf.data(), mind.data(), kf->angle,
                       occupied ? occ.data() : nullptr, mp, false, match, n_cur, nmatches);
}

/* ORBmatcher::Fuse(KeyFrame *pKF, const vector<MapPoint*> &vpMapPoints, const float th) (src/ORBmatcher.cc:968-1131): the
 * keypoint of pKF each map point is to be fused into.  pKF = frame `frame` of kf's last call -- a keyframe put there by
 * borb_load_frame.  One kernel (M1a in MODE 5, k_match.cuh); the queries are independent. */
int borb_fuse_search(borb_extractor *kf, int frame, const borb_fuse_points *mps, const float *Tcw, const float *Ow, float bf,
                     float th, unsigned flags, int32_t *best_idx, int *nfused)
{
    if (!kf || !mps || !Tcw || !Ow || !best_idx) return fail(BORB_ERR_ARG, "null argument");
    if (mps->n > 0 && (!mps->xyz || !mps->desc || !mps->flags || !mps->dist || !mps->normal)) return fail(BORB_ERR_ARG, "map point arrays");
    MatchParams mp;
    std::memset(&mp, 0, sizeof mp);
    std::memcpy(mp.Tc, Tcw, 48);
    std::memcpy(mp.Ow, Ow, 12);
    mp.th = th;
    mp.mbf = bf;
    mp.orb_dist = MT_TH_LOW;
    mp.nlevels = kf->nlevels;
    mp.log_scale = orb_logf_exact((float)kf->scale_factor);
    for (int l = 0; l < BORB_MAXL; ++l) mp.inv_sigma2[l] = l < kf->nlevels ? kf->inv_sigma2[l] : 0.f;
    const size_t n = (size_t)std::max(mps->n, 0);
    std::vector<float> geo(4 * n), extra(4 * n);
    std::vector<int32_t> mind(n);
    std::vector<uint8_t> qf(n);
    for (size_t i = 0; i < n; ++i) {
        geo[4 * i] = mps->xyz[3 * i]; geo[4 * i + 1] = mps->xyz[3 * i + 1]; geo[4 * i + 2] = mps->xyz[3 * i + 2];
        geo[4 * i + 3] = mps->dist[2 * i + 1];
        std::memcpy(&mind[i], &mps->dist[2 * i], 4);
        extra[4 * i] = mps->normal[3 * i]; extra[4 * i + 1] = mps->normal[3 * i + 1]; extra[4 * i + 2] = mps->normal[3 * i + 2];
        extra[4 * i + 3] = 0.f;
        qf[i] = ((mps->flags[i] & 1) && !(mps->flags[i] & 6)) ? 1 : 0;
    }
    return run_matcher(kf, frame, 5, mps->n, geo.data(), 4, mps->desc, qf.data(), mind.data(), nullptr, nullptr, mp,
                       (flags & BORB_MATCH_USE_STEREO) != 0, best_idx, nullptr, nfused, extra.data());
}

namespace {
/* the point arrays of the loop-closing matchers as run_matcher's query columns */
struct ProjQueries {
    std::vector<float> geo, extra;
    std::vector<int32_t> mind;
    std::vector<uint8_t> qf;
    ProjQueries(const borb_fuse_points *p, uint8_t usable)
    {
        const size_t n = (size_t)std::max(p->n, 0);
        geo.resize(4 * n); mind.resize(n); qf.resize(n);
        if (p->normal) extra.resize(4 * n);
        for (size_t i = 0; i < n; ++i) {
            geo[4 * i] = p->xyz[3 * i]; geo[4 * i + 1] = p->xyz[3 * i + 1]; geo[4 * i + 2] = p->xyz[3 * i + 2];
            geo[4 * i + 3] = p->dist[2 * i + 1];
            std::memcpy(&mind[i], &p->dist[2 * i], 4);
            if (p->normal) { extra[4 * i] = p->normal[3 * i]; extra[4 * i + 1] = p->normal[3 * i + 1]; extra[4 * i + 2] = p->normal[3 * i + 2]; extra[4 * i + 3] = 0.f; }
            qf[i] = ((p->flags[i] & 1) && !(p->flags[i] & 6)) ? usable : 0;
        }
    }
};

int proj_params(borb_extractor *kf, const borb_fuse_points *pts, const borb_projection *pr, float th, int max_dist, MatchParams &mp)
{
    if (!kf || !pts || !pr || !pr->T1) return fail(BORB_ERR_ARG, "null argument");
    if (pts->n > 0 && (!pts->xyz || !pts->desc || !pts->flags || !pts->dist)) return fail(BORB_ERR_ARG, "point arrays");
    if (max_dist < 0 || max_dist > 255) return fail(BORB_ERR_ARG, "distance threshold %d", max_dist);
    std::memset(&mp, 0, sizeof mp);
    std::memcpy(mp.Tc, pr->T1, 48);
    if (pr->T2) { std::memcpy(mp.Tl, pr->T2, 48); mp.two_step = 1; }
    if (pr->Ow) std::memcpy(mp.Ow, pr->Ow, 12); else mp.cam_norm = 1;
    mp.use_normal = pts->normal ? 1 : 0;
    if (mp.use_normal && mp.cam_norm) return fail(BORB_ERR_ARG, "the viewing-angle test needs the camera centre");
    mp.th = th;
    mp.orb_dist = max_dist;
    mp.nlevels = kf->nlevels;
    mp.log_scale = orb_logf_exact((float)kf->scale_factor);
    return BORB_OK;
}
} // namespace

/* The search inside ORBmatcher::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (src/ORBmatcher.cc:1133-1260) and one direction
 * of ORBmatcher::SearchBySim3 (:1305-1390 / :1393-1481): per point the first smallest Hamming distance among pKF's keypoints
 * of the predicted level and the one below inside the window, when <= max_dist.  M1a in MODE 6; independent queries. */
int borb_project_search(borb_extractor *kf, int frame, const borb_fuse_points *pts, const borb_projection *pr, float th,
                        int max_dist, int32_t *best_idx, int *nfound)
{
    MatchParams mp;
    int rc = proj_params(kf, pts, pr, th, max_dist, mp);
    if (rc != BORB_OK) return rc;
    if (!best_idx) return fail(BORB_ERR_ARG, "null output");
    ProjQueries q(pts, 1);
    return run_matcher(kf, frame, 6, pts->n, q.geo.data(), 4, pts->desc, q.qf.data(), q.mind.data(), nullptr, nullptr, mp, false,
                       best_idx, nullptr, nfound, pts->normal ? q.extra.data() : nullptr);
}

/* ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (src/ORBmatcher.cc:353-485): sequential in the points (a
 * keypoint whose vpMatched entry is set is closed, :437-438) -- M1a in MODE 6, M1b with the rule of the relocalisation
 * matcher and TH_LOW. */
int borb_search_by_projection_sim3(borb_extractor *kf, int frame, const borb_fuse_points *pts, const borb_projection *pr,
                                   const uint8_t *occupied, int th, int32_t *match, int *n_kf, int *nmatches)
{
    MatchParams mp;
    int rc = proj_params(kf, pts, pr, (float)th, MT_TH_LOW, mp);
    if (rc != BORB_OK) return rc;
    if (!match) return fail(BORB_ERR_ARG, "null output");
    if (!pr->Ow || !pts->normal) return fail(BORB_ERR_ARG, "this matcher tests the viewing angle: camera centre and normals");
    ProjQueries q(pts, 3); /* a query, and it closes the slot it takes */
    std::vector<uint8_t> occ;
    if (occupied) {
        if (!kf->have_plan || kf->last_n < 1 || frame < 0 || frame >= kf->last_n)
            return fail(BORB_ERR_STATE, "no finished extract call holds frame %d", frame);
        if (!kf->results_on_host && kf->nout_frame_on_host != frame) {
            CU(cudaSetDevice(kf->device));
            CU(cudaMemcpyAsync(kf->h_nout + frame, kf->d_nout + frame, 4, cudaMemcpyDeviceToHost, kf->stream));
            CU(cudaStreamSynchronize(kf->stream));
            kf->nout_frame_on_host = frame;
        }
        const int nc0 = std::min(kf->h_nout[frame], kf->plan.kp_total);
        occ.assign((size_t)kf->plan.kp_total, 0);
        for (int i = 0; i < nc0; ++i) occ[i] = occupied[i] ? 2 : 0;
    }
    return run_matcher(kf, frame, 7, pts->n, q.geo.data(), 4, pts->desc, q.qf.data(), q.mind.data(), nullptr,
                       occupied ? occ.data() : nullptr, mp, false, match, n_kf, nmatches, q.extra.data());
}

/* ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th) (src/ORBmatcher.cc:1262-1501): pKF1's points searched
 * in pKF2 and pKF2's in pKF1 (two borb_project_search calls, TH_HIGH), then the agreement check (:1484-1498). */
int borb_search_by_sim3(borb_extractor *kf1, borb_extractor *kf2, const borb_fuse_points *pts1, const borb_fuse_points *pts2,
                        const float *T1w, const float *T2w, const float *T12, const float *T21, float th, const int32_t *pre12,
                        int32_t *match12, int *nfound)
{
    if (!kf1 || !kf2 || !pts1 || !pts2 || !T1w || !T2w || !T12 || !T21 || !match12) return fail(BORB_ERR_ARG, "null argument");
    const int n1 = pts1->n, n2 = pts2->n;
    if (n1 < 0 || n2 < 0) return fail(BORB_ERR_ARG, "point counts");
    std::vector<uint8_t> f1(pts1->flags, pts1->flags + n1), f2(pts2->flags, pts2->flags + n2);
    for (int i = 0; i < n1 && pre12; ++i)
        if (pre12[i] >= 0) { /* vbAlreadyMatched1 / 2 (:1291-1302) */
            f1[i] |= 4;
            if (pre12[i] < n2) f2[pre12[i]] |= 4;
        }
    borb_fuse_points a = *pts1, b = *pts2;
    a.flags = f1.data(); a.normal = nullptr;
    b.flags = f2.data(); b.normal = nullptr;
    std::vector<int32_t> m1((size_t)std::max(n1, 1)), m2((size_t)std::max(n2, 1));
    borb_projection p12 = {T1w, T21, nullptr}, p21 = {T2w, T12, nullptr};
    int rc = borb_project_search(kf2, 0, &a, &p12, th, MT_TH_HIGH, m1.data(), nullptr);
    if (rc != BORB_OK) return rc;
    rc = borb_project_search(kf1, 0, &b, &p21, th, MT_TH_HIGH, m2.data(), nullptr);
    if (rc != BORB_OK) return rc;
    int nf = 0;
    for (int i = 0; i < n1; ++i) {
        match12[i] = pre12 ? pre12[i] : -1;
        const int idx2 = m1[i];
        if (idx2 >= 0 && idx2 < n2 && m2[idx2] == i) { match12[i] = idx2; ++nf; }
    }
    if (nfound) *nfound = nf;
    return BORB_OK;
}

/* ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) (src/ORBmatcher.cc:487-632) */
int borb_search_for_initialization(borb_extractor *cur, int frame, const borb_init_frame *f1, float *prev_matched,
                                   int window_size, float nnratio, unsigned flags, int32_t *matches12, int *nmatches)
{
    if (!cur || !f1 || !matches12 || (f1->n > 0 && (!prev_matched || !f1->octave || !f1->angle || !f1->desc)))
        return fail(BORB_ERR_ARG, "null argument");
    if (f1->n < 0 || window_size < 0) return fail(BORB_ERR_ARG, "n %d, window %d", f1->n, window_size);
    MatchParams mp;
    std::memset(&mp, 0, sizeof mp);
    mp.th = (float)window_size;
    mp.nnratio = nnratio;
    mp.check_ori = (flags & BORB_MATCH_NO_ORIENTATION) ? 0 : 1;
    std::vector<uint8_t> qf((size_t)std::max(f1->n, 1));
    std::vector<int32_t> lvl((size_t)std::max(f1->n, 1), 0);
    for (int i = 0; i < f1->n; ++i) qf[i] = f1->octave[i] > 0 ? 0 : 1; /* only level 0 is searched (:507-508) */
    int nm = 0;
    const int rc = run_matcher(cur, frame, 3, f1->n, prev_matched, 2, f1->desc, qf.data(), lvl.data(), f1->angle, nullptr, mp,
                               false, matches12, nullptr, &nm);
    if (rc != BORB_OK) return rc;
    for (int i = 0; i < f1->n; ++i) /* :621-623 */
        if (matches12[i] >= 0) { prev_matched[2 * i] = cur->h_mt_unxy[2 * matches12[i]]; prev_matched[2 * i + 1] = cur->h_mt_unxy[2 * matches12[i] + 1]; }
    if (nmatches) *nmatches = nm;
    return BORB_OK;
}

int borb_match_rounds(const borb_extractor *cur) { return cur ? cur->mt_rounds | (cur->mt_rescans << 8) : 0; }
int borb_max_keypoints(const borb_extractor *h) { return h && h->have_plan ? h->plan.kp_total : 0; }

int borb_level_size(const borb_extractor *h, int level, int *cols, int *rows)
{
    if (!h || !h->have_plan) return fail(BORB_ERR_STATE, "no geometry yet");
    if (level < 0 || level >= h->nlevels) return fail(BORB_ERR_ARG, "level");
    if (cols) *cols = h->plan.lv[level].w;
    if (rows) *rows = h->plan.lv[level].h;
    return BORB_OK;
}

int borb_pyramid(const borb_extractor *h, int frame, int level, const uint8_t **data, int *cols,
                 int *rows, size_t *step)
{
    if (!h || !h->pyramid_on_host) return fail(BORB_ERR_STATE, "pyramid was not requested");
    if (frame < 0 || frame >= h->last_n || level < 0 || level >= h->nlevels) return fail(BORB_ERR_ARG, "frame/level");
    const BorbLevel &L = h->plan.lv[level];
    if (data) *data = h->h_pyr + (size_t)frame * (size_t)h->plan.pyr_bytes + L.off + (size_t)BORB_EDGE * L.pitch + BORB_XPAD;
    if (cols) *cols = L.w;
    if (rows) *rows = L.h;
    if (step) *step = (size_t)L.pitch;
    return BORB_OK;
}

int borb_debug_candidates(borb_extractor *h, int frame, int level, int32_t *xs, int32_t *ys,
                          int32_t *scores, int cap, int *n)
{
    if (!h || !h->have_plan || frame < 0 || frame >= h->last_n || level < 0 || level >= h->nlevels)
        return fail(BORB_ERR_ARG, "debug_candidates arguments");
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    int cnt = 0;
    CU(cudaMemcpy(&cnt, h->d_ncand + frame * BORB_MAXL + level, 4, cudaMemcpyDeviceToHost));
    const BorbLevel &L = h->plan.lv[level];
    const int m = std::min(std::min(cnt, cap), L.cand_cap);
    std::vector<uint32_t> xy((size_t)std::max(m, 1)), sc((size_t)std::max(m, 1));
    const size_t o = (size_t)frame * (size_t)h->plan.cand_total + L.cand_off;
    if (m > 0) {
        CU(cudaMemcpy(xy.data(), h->d_cand_xy + o, (size_t)m * 4, cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(sc.data(), h->d_cand_score + o, (size_t)m * 4, cudaMemcpyDeviceToHost));
    }
    for (int i = 0; i < m; ++i) {
        xs[i] = (int)(xy[i] & 0xffff);
        ys[i] = (int)(xy[i] >> 16);
        scores[i] = (int)sc[i];
    }
    if (n) *n = cnt;
    return BORB_OK;
}

int borb_debug_blurred(borb_extractor *h, int frame, int level, uint8_t *out)
{
    if (!h || !h->have_plan || frame < 0 || frame >= h->last_n || level < 0 || level >= h->nlevels)
        return fail(BORB_ERR_ARG, "debug_blurred arguments");
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    const BorbLevel &L = h->plan.lv[level];
    const uint8_t *src = h->d_blur + (size_t)frame * (size_t)h->plan.pyr_bytes + L.off + (size_t)BORB_EDGE * L.pitch + BORB_XPAD;
    CU(cudaMemcpy2D(out, (size_t)L.w, src, (size_t)L.pitch, (size_t)L.w, (size_t)L.h, cudaMemcpyDeviceToHost));
    return BORB_OK;
}

/* K6 (computeOrbDescriptor, src/ORBextractor.cc:117-157) on a caller-supplied image and angle: the device function the
 * extract kernels use, on one warp.  (x, y) must keep the rotated pattern (|offset| <= 18) inside the image. */
int borb_debug_describe(borb_extractor *h, const uint8_t *img, int cols, int rows, size_t step, int x, int y, float angle_deg,
                        uint8_t *desc32)
{
    if (!h || !img || !desc32 || cols < 37 || rows < 37 || step < (size_t)cols || x < 18 || y < 18 || x + 18 >= cols || y + 18 >= rows)
        return fail(BORB_ERR_ARG, "debug_describe arguments");
    CU(cudaSetDevice(h->device));
    uint8_t *d = nullptr;
    CU(cudaMalloc(&d, step * (size_t)rows + 32));
    cudaError_t e = cudaMemcpy(d, img, step * (size_t)rows, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        k_debug_describe<<<1, 32, 0, h->stream>>>(d, (int)step, x, y, angle_deg, h->d_pattern, d + step * (size_t)rows);
        e = cudaStreamSynchronize(h->stream);
    }
    if (e == cudaSuccess) e = cudaMemcpy(desc32, d + step * (size_t)rows, 32, cudaMemcpyDeviceToHost);
    cudaFree(d);
    if (e != cudaSuccess) return fail(BORB_ERR_CUDA, "debug_describe: %s", cudaGetErrorString(e));
    return BORB_OK;
}

void borb_debug_sincos(float radians, float *s, float *c) { orb_sincosf_exact(radians, s, c); }
void borb_debug_logf(const float *x, float *y, int n)
{
    for (int i = 0; i < n; ++i) y[i] = orb_logf_exact(x[i]);
}
int borb_debug_logf_device(const float *x, float *y, int n)
{
    if (n < 1 || !x || !y) return fail(BORB_ERR_ARG, "arguments");
    float *d = nullptr;
    CU(cudaMalloc(&d, (size_t)n * 4));
    cudaError_t e = cudaMemcpy(d, x, (size_t)n * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        k_debug_logf<<<(n + 255) / 256, 256>>>(d, n);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(y, d, (size_t)n * 4, cudaMemcpyDeviceToHost);
    cudaFree(d);
    if (e != cudaSuccess) return fail(BORB_ERR_CUDA, "%s", cudaGetErrorString(e));
    return BORB_OK;
}
void borb_debug_set_match_rounds(int max_rounds) { g_match_max_rounds = max_rounds > 0 ? max_rounds : 0; }
int borb_debug_fast_score(const uint8_t *p, int pitch, int t_lo) { return fast_score_px(p, pitch, t_lo); }

} // extern "C"

#include "borb_stream.inl"
#include "borb_noalloc.inl"
#include "borb_bow.inl"

#ifdef BORB_FAST_PROBE
/* development only (tools/probe_fast.py): the probe words of the last k_fast_cells launch */
extern "C" int borb_debug_fast_probe(unsigned long long *out, int n_words)
{
    if (cudaDeviceSynchronize() != cudaSuccess) return BORB_ERR_CUDA;
    return cudaMemcpyFromSymbol(out, g_fast_probe, (size_t)n_words * 8) == cudaSuccess ? BORB_OK : BORB_ERR_CUDA;
}
#endif

#ifdef BORB_QT_PROBE
/* development only (tools/probe_qt.py): the probe log of the last k_quadtree launch */
extern "C" int borb_debug_qt_probe(unsigned long long *out, int *n_out)
{
    if (cudaDeviceSynchronize() != cudaSuccess) return BORB_ERR_CUDA;
    if (cudaMemcpyFromSymbol(out, g_qt_probe, sizeof(unsigned long long) * 8 * 128 * 2) != cudaSuccess) return BORB_ERR_CUDA;
    return cudaMemcpyFromSymbol(n_out, g_qt_probe_n, sizeof(int) * 8) == cudaSuccess ? BORB_OK : BORB_ERR_CUDA;
}
#endif

#ifdef BORB_PCH_PROBE
/* development only (tools/probe_chain.py): the probe words of the last k_pyr_chain launch */
extern "C" int borb_debug_chain_probe(unsigned long long *out, int n_words)
{
    if (cudaDeviceSynchronize() != cudaSuccess) return BORB_ERR_CUDA;
    return cudaMemcpyFromSymbol(out, g_pch_probe, (size_t)n_words * 8) == cudaSuccess ? BORB_OK : BORB_ERR_CUDA;
}
#endif
