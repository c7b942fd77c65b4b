/*
 * borb_internal.h -- device-side plan shared by the host API (borb_api.cu) and the kernels.
 *
 * HBM layout for one extractor handle (B = max_batch frames, L = nlevels):
 *   pyr    [B][pyr_bytes]   all L padded pyramid levels of a frame, level l at lv[l].off;
 *                           row pitch lv[l].pitch (multiple of 32), pixel (0,0) at column
 *                           BORB_XPAD of row BORB_EDGE -> interior rows are 32-byte aligned
 *   blur   [B][pyr_bytes]   Gaussian-blurred levels, same geometry (frame never read)
 *   cand   [B][cand_total]  FAST candidates per level at lv[l].cand_off: packed u32 x|y<<16,
 *                           u32 score, u32 current quadtree node
 *   ncand  [B][L]           candidate counters (atomics), zeroed each call
 *   lkp    [B][kp_total]    per-level survivors (x, y, score) at lv[l].kp_off, nlkp [B][L]
 *   out    [B][kp_total]    final cv::KeyPoint records + [B][kp_total][32] descriptors, nout [B]
 *   stereo [B][kp_total]    uRight, depth (float) and SAD (int) per left keypoint, kept [B]
 *                           (allocated by the first borb_stereo_match on the left handle)
 */
#ifndef BORB_INTERNAL_H
#define BORB_INTERNAL_H

#include <stdint.h>

#define BORB_MAXL 16     /* max pyramid levels                                         */
#define BORB_EDGE 19     /* EDGE_THRESHOLD, reference src/ORBextractor.cc:74           */
#define BORB_XPAD 32     /* left pad columns in our pitch (>= BORB_EDGE, 32-B aligned) */
#define BORB_FAST_MARGIN 16 /* EDGE_THRESHOLD-3: first pixel of the FAST region, :974  */
#define BORB_MAXCELL 60  /* a cell is < 60 px in either axis (width/floor(width/30))   */
#define BORB_QT_MAXNODES 2048 /* quadtree list capacity per level held in shared memory */

struct BorbLevel {
    int w, h;          /* level image size                                    */
    int pitch;         /* bytes per padded row                                */
    int prows;         /* h + 2*BORB_EDGE                                     */
    long long off;     /* byte offset of the padded level inside a frame slab */
    /* resize tables (level l from level l-1), offsets into the int/short table blobs */
    int xofs_off, yofs_off; /* into tab_i  (int)   */
    int xa_off, yb_off;     /* into tab_s  (short pairs) */
    /* FAST cell grid, reference :986-996 */
    int n_cols, n_rows, w_cell, h_cell;
    int max_bx, max_by;     /* maxBorderX/Y = w-16, h-16 */
    /* quadtree, reference :668-739 */
    int qt_w, qt_h;         /* maxBorder - minBorder */
    int n_ini;
    float hx;
    int quota;              /* mnFeaturesPerLevel[l] */
    int kp_cap;             /* survivors capacity of this level */
    int kp_off;             /* offset into per-frame keypoint arrays */
    int cand_cap;           /* candidate capacity */
    long long cand_off;     /* offset into per-frame candidate arrays */
    float scale;            /* mvScaleFactor[l] */
    float inv_scale;        /* mvInvScaleFactor[l] */
    float patch_size;       /* (float)(int)(31*scale) */
};

struct BorbCell { /* one FAST window, reference :1000-1064 */
    short level;
    short pad;
    short x0, y0;     /* window origin in level pixels */
    short ww, wh;     /* window size */
    short sx, sy;     /* j*wCell, i*hCell: shift added to window-relative coords */
    /* derived on the host: words per staged row, its fast_div() magic number, and the stage-1
     * lane geometry dmagic = Q | R << 8 (Q words cover a detection-domain row, R rows per warp step) */
    int wpr;
    unsigned dmagic, wmagic;
    int pad2;
};

struct BorbTile { /* one blur strip: BORB_BLUR_TW columns x BORB_BLUR_TH rows of a level */
    short level;
    short pad;
    short x0, y0;
};

struct BorbPlan {
    int nlevels;
    int cols, rows;
    int ini_th, min_th;
    long long pyr_bytes;    /* per frame */
    long long cand_total;   /* per frame */
    int kp_total;           /* per frame */
    int n_cells, n_tiles;
    int fast_max_wh, fast_max_pix; /* tallest FAST window, largest detection domain (pixels) */
    int fast_max_ww;               /* widest FAST window */
    int umax[16];
    BorbLevel lv[BORB_MAXL];
};

#define BORB_BLUR_TW 128 /* one warp x 4 columns */
#ifndef BORB_BLUR_TH
#define BORB_BLUR_TH 36 /* a multiple of 6: K5 unrolls its row loop by 6 */
#endif

#endif
