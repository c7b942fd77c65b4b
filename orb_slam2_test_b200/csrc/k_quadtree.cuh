/*
 * k_quadtree.cuh -- K3: DistributeOctTree on the GPU (src/ORBextractor.cc:668-951 and
 * ExtractorNode::DivideNode :537-614), one CTA per (level, frame), node list in shared memory.
 *
 * The reference mutates a std::list: multi-point nodes are replaced by their non-empty quarters
 * (pushed to the FRONT in the order UL, UR, BL, BR), first breadth-wise over the whole list,
 * then -- once one more full pass could overshoot the per-level quota N -- largest node first,
 * stopping the moment the list holds N nodes; finally each node keeps its best-response point.
 * Restated level-synchronously, with `c` the creation index of a
 * non-empty child among the C children born in a pass:
 *     new list = [child with c = C-1, ..., child with c = 0] ++ [unsplit nodes, old order]
 * Breadth pass: all nodes with > 1 points split, parents ranked in list order.
 * Ordered pass: expandable nodes ranked by (point count desc, creation order desc) -- the
 * canonical tie-break (Appendix B.4) -- and only the ranks up to the first prefix whose size
 * reaches N are split.
 * Points never move: each candidate carries the list position of its node; children are counted
 * with shared-memory atomics and positions come from block-wide exclusive scans.
 * The winner of a node is max(response), ties to the earliest candidate in the reference's input
 * order (cell row, cell col, y, x), which is recomputed from the coordinates.
 */
#ifndef K_QUADTREE_CUH
#define K_QUADTREE_CUH

#include "borb_internal.h"

/* threads per CTA is a template parameter: 256 for throughput (many (level, frame) CTAs in flight;
 * measured on B200 per 512 frames: 1024 -> 0.80 ms, 512 -> 0.51 ms, 256 -> 0.48 ms), 1024 for the
 * latency of a single frame (8 CTAs in total: 45 us vs 78 us) */
#define QT_NT_THROUGHPUT 256
#define QT_NT_LATENCY 1024

/* in-place exclusive scan of a[0..len), returns the total; all threads must call */
template <int QT_NT>
__device__ int qt_scan(int *a, int len, int *wtmp /* 33 ints */)
{
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int per = (len + QT_NT - 1) / QT_NT;
    const int b = min(tid * per, len), e = min(b + per, len);
    int s = 0;
    for (int i = b; i < e; ++i) s += a[i];
    int v = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    if (lane == 31) wtmp[w] = v;
    __syncthreads();
    if (w == 0) {
        const int t = lane < QT_NT / 32 ? wtmp[lane] : 0;
        int u = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int z = __shfl_up_sync(0xffffffffu, u, o);
            if (lane >= o) u += z;
        }
        wtmp[lane] = u - t;
        if (lane == 31) wtmp[32] = u;
    }
    __syncthreads();
    int run = wtmp[w] + v - s;
    for (int i = b; i < e; ++i) {
        const int t = a[i];
        a[i] = run;
        run += t;
    }
    const int total = wtmp[32];
    __syncthreads();
    return total;
}

struct QtSmem {
    int4 *box[2];
    int *cnt[2];
    int *rank, *par, *uf, *ccnt, *cf, *ef;
    int *exp_pos[2], *exp_size[2];
    unsigned long long *best;
    int *wtmp;
};

__device__ __forceinline__ int qt_quadrant(int x, int y, const int4 b)
{
    const int mx = b.x + ((b.z - b.x + 1) >> 1); /* x0 + ceil((x1-x0)/2) */
    const int my = b.y + ((b.w - b.y + 1) >> 1);
    return (y >= my) * 2 + (x >= mx);
}

__device__ __forceinline__ int4 qt_child_box(const int4 b, int q)
{
    const int mx = b.x + ((b.z - b.x + 1) >> 1);
    const int my = b.y + ((b.w - b.y + 1) >> 1);
    int4 c;
    c.x = (q & 1) ? mx : b.x;
    c.z = (q & 1) ? b.z : mx;
    c.y = (q & 2) ? my : b.y;
    c.w = (q & 2) ? b.w : my;
    return c;
}

/* bytes of dynamic shared memory for a list capacity of M nodes */
static inline size_t qt_smem_bytes(int M)
{
    return (size_t)M * (2 * 16 + 2 * 4 + 3 * 4 + 3 * 16 + 4 * 4 + 8) + 64 * 4;
}

/* One split pass over the current list (cur) into (1-cur).
 * On entry: rank[j] = processing rank of list node j (or -1), par[k] = node of rank k,
 * ccnt[4k+q] = points of rank k's q-th quarter, for all n_rank ranked nodes.
 * Only ranks < n_proc are actually split.  Returns the new list size; *n_exp = number of
 * children with more than one point (written to the other expandable buffer, creation order). */
template <int QT_NT>
__device__ int qt_split(const QtSmem &s, int cur, int S, int n_proc, int n_pts,
                        const uint32_t *__restrict__ xy, uint32_t *__restrict__ node, int *n_exp)
{
    const int tid = threadIdx.x;
    const int nc = 4 * n_proc;
    for (int c = tid; c < nc; c += QT_NT) {
        s.cf[c] = s.ccnt[c] > 0;
        s.ef[c] = s.ccnt[c] > 1;
    }
    for (int j = tid; j < S; j += QT_NT) s.uf[j] = !(s.rank[j] >= 0 && s.rank[j] < n_proc);
    __syncthreads();
    const int C = qt_scan<QT_NT>(s.cf, nc, s.wtmp);
    const int E = qt_scan<QT_NT>(s.ef, nc, s.wtmp);
    const int U = qt_scan<QT_NT>(s.uf, S, s.wtmp);
    const int nxt = 1 - cur;
    for (int c = tid; c < nc; c += QT_NT) {
        const int n = s.ccnt[c];
        if (n > 0) {
            const int p = C - 1 - s.cf[c];
            s.box[nxt][p] = qt_child_box(s.box[cur][s.par[c >> 2]], c & 3);
            s.cnt[nxt][p] = n;
            s.cf[c] = p;
            if (n > 1) {
                s.exp_pos[nxt][s.ef[c]] = p;
                s.exp_size[nxt][s.ef[c]] = n;
            }
        }
    }
    for (int j = tid; j < S; j += QT_NT) {
        if (!(s.rank[j] >= 0 && s.rank[j] < n_proc)) {
            const int p = C + s.uf[j];
            s.box[nxt][p] = s.box[cur][j];
            s.cnt[nxt][p] = s.cnt[cur][j];
            s.uf[j] = p;
        }
    }
    __syncthreads();
    for (int i = tid; i < n_pts; i += QT_NT) {
        const int j = (int)node[i];
        const int k = s.rank[j];
        if (k >= 0 && k < n_proc) {
            const uint32_t v = xy[i];
            node[i] = (uint32_t)s.cf[4 * k + qt_quadrant((int)(v & 0xffff), (int)(v >> 16), s.box[cur][j])];
        } else {
            node[i] = (uint32_t)s.uf[j];
        }
    }
    __syncthreads();
    *n_exp = E;
    return C + U;
}

/* count the quarters of every ranked node */
template <int QT_NT>
__device__ void qt_count(const QtSmem &s, int cur, int n_rank, int n_pts,
                         const uint32_t *__restrict__ xy, const uint32_t *__restrict__ node)
{
    const int tid = threadIdx.x;
    for (int c = tid; c < 4 * n_rank; c += QT_NT) s.ccnt[c] = 0;
    __syncthreads();
    for (int i = tid; i < n_pts; i += QT_NT) {
        const int j = (int)node[i];
        const int k = s.rank[j];
        if (k >= 0) {
            const uint32_t v = xy[i];
            atomicAdd(&s.ccnt[4 * k + qt_quadrant((int)(v & 0xffff), (int)(v >> 16), s.box[cur][j])], 1);
        }
    }
    __syncthreads();
}

template <int QT_NT>
__global__ void __launch_bounds__(QT_NT, 1024 / QT_NT)
k_quadtree(const __grid_constant__ BorbPlan plan_c, int M, const uint32_t *__restrict__ cand_xy,
           const uint32_t *__restrict__ cand_score, uint32_t *__restrict__ cand_node,
           const int *__restrict__ ncand, uint32_t *__restrict__ lkp_xy,
           uint32_t *__restrict__ lkp_score, int *__restrict__ nlkp, int *__restrict__ err)
{
    const BorbPlan *plan = &plan_c; /* the plan rides in the parameter block: constant-bank reads, no global loads */
    extern __shared__ __align__(16) unsigned char qt_raw[];
    __shared__ int sh_proc;
    QtSmem s;
    {
        unsigned char *p = qt_raw;
        s.box[0] = (int4 *)p; p += (size_t)M * 16;
        s.box[1] = (int4 *)p; p += (size_t)M * 16;
        s.best = (unsigned long long *)p; p += (size_t)M * 8;
        s.cnt[0] = (int *)p; p += (size_t)M * 4;
        s.cnt[1] = (int *)p; p += (size_t)M * 4;
        s.rank = (int *)p; p += (size_t)M * 4;
        s.par = (int *)p; p += (size_t)M * 4;
        s.uf = (int *)p; p += (size_t)M * 4;
        s.ccnt = (int *)p; p += (size_t)M * 16;
        s.cf = (int *)p; p += (size_t)M * 16;
        s.ef = (int *)p; p += (size_t)M * 16;
        s.exp_pos[0] = (int *)p; p += (size_t)M * 4;
        s.exp_pos[1] = (int *)p; p += (size_t)M * 4;
        s.exp_size[0] = (int *)p; p += (size_t)M * 4;
        s.exp_size[1] = (int *)p; p += (size_t)M * 4;
        s.wtmp = (int *)p;
    }
    const int level = blockIdx.x, frame = blockIdx.y, tid = threadIdx.x;
    const BorbLevel &L = plan->lv[level];
    const size_t cbase = (size_t)frame * (size_t)plan->cand_total + L.cand_off;
    const uint32_t *xy = cand_xy + cbase;
    const uint32_t *score = cand_score + cbase;
    uint32_t *node = cand_node + cbase;
    int n = ncand[frame * BORB_MAXL + level];
    if (n > L.cand_cap) { /* cannot happen (capacity is the NMS bound); never truncate silently */
        if (tid == 0) atomicExch(err, 1);
        n = L.cand_cap;
    }
    const int N = L.quota;

    /* ---- roots (:674-739): point -> root by FLOAT division, empty roots dropped ---- */
    const int n_ini = L.n_ini;
    const float hx = L.hx;
    for (int j = tid; j < n_ini; j += QT_NT) s.ccnt[j] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += QT_NT) {
        int r = (int)__fdiv_rn((float)(xy[i] & 0xffff), hx);
        r = min(max(r, 0), n_ini - 1);
        node[i] = (uint32_t)r;
        atomicAdd(&s.ccnt[r], 1);
    }
    __syncthreads();
    for (int j = tid; j < n_ini; j += QT_NT) s.cf[j] = s.ccnt[j] > 0;
    __syncthreads();
    int S = qt_scan<QT_NT>(s.cf, n_ini, s.wtmp);
    int cur = 0;
    for (int j = tid; j < n_ini; j += QT_NT) {
        if (s.ccnt[j] > 0) {
            const int p = s.cf[j];
            s.box[cur][p] = make_int4((int)__fmul_rn(hx, (float)j), 0,
                                      (int)__fmul_rn(hx, (float)(j + 1)), L.qt_h);
            s.cnt[cur][p] = s.ccnt[j];
        }
    }
    __syncthreads();
    for (int i = tid; i < n; i += QT_NT) node[i] = (uint32_t)s.cf[node[i]];
    __syncthreads();

    /* ---- split passes (:741-927) ---- */
    bool done = (S == 0);
    while (!done) {
        /* breadth pass: rank the multi-point nodes in list order */
        const int prev = S;
        for (int j = tid; j < S; j += QT_NT) s.ef[j] = s.cnt[cur][j] > 1;
        __syncthreads();
        const int n_split = qt_scan<QT_NT>(s.ef, S, s.wtmp);
        if (n_split == 0) break; /* list cannot grow: lNodes.size()==prevSize */
        for (int j = tid; j < S; j += QT_NT) {
            const bool sp = s.cnt[cur][j] > 1;
            s.rank[j] = sp ? s.ef[j] : -1;
            if (sp) s.par[s.ef[j]] = j;
        }
        __syncthreads();
        qt_count<QT_NT>(s, cur, n_split, n, xy, node);
        int n_exp;
        S = qt_split<QT_NT>(s, cur, S, n_split, n, xy, node, &n_exp);
        cur = 1 - cur;
        if (S >= N || S == prev) break;
        if (S + 3 * n_exp <= N) continue;

        /* ordered passes (:853-924) */
        while (true) {
            const int prev2 = S;
            const int E = n_exp;
            if (E == 0) { done = true; break; }
            for (int j = tid; j < S; j += QT_NT) s.rank[j] = -1;
            __syncthreads();
            for (int e = tid; e < E; e += QT_NT) {
                const int sz = s.exp_size[cur][e];
                int r = 0;
                for (int f = 0; f < E; ++f) {
                    const int sf = s.exp_size[cur][f];
                    r += (sf > sz) || (sf == sz && f > e);
                }
                s.rank[s.exp_pos[cur][e]] = r;
                s.par[r] = s.exp_pos[cur][e];
            }
            __syncthreads();
            qt_count<QT_NT>(s, cur, E, n, xy, node);
            /* size after processing ranks 0..k; stop at the first k reaching N */
            for (int k = tid; k < E; k += QT_NT) {
                const int g = (s.ccnt[4 * k] > 0) + (s.ccnt[4 * k + 1] > 0) +
                              (s.ccnt[4 * k + 2] > 0) + (s.ccnt[4 * k + 3] > 0) - 1;
                s.cf[k] = g;
                s.ef[k] = g;
            }
            if (tid == 0) sh_proc = E;
            __syncthreads();
            qt_scan<QT_NT>(s.ef, E, s.wtmp); /* exclusive prefix of the gains */
            for (int k = tid; k < E; k += QT_NT)
                if (S + s.ef[k] + s.cf[k] >= N) atomicMin(&sh_proc, k + 1);
            __syncthreads();
            const int n_proc = sh_proc;
            __syncthreads();
            S = qt_split<QT_NT>(s, cur, S, n_proc, n, xy, node, &n_exp);
            cur = 1 - cur;
            if (S >= N || S == prev2) { done = true; break; }
        }
    }

    /* ---- one survivor per node (:930-948) ---- */
    if (S > M) { if (tid == 0) atomicExch(err, 2); S = M; }
    for (int j = tid; j < S; j += QT_NT) s.best[j] = 0ull;
    __syncthreads();
    for (int i = tid; i < n; i += QT_NT) {
        const uint32_t v = xy[i];
        const unsigned x = v & 0xffff, y = v >> 16;
        const unsigned cr = (y - 3) / (unsigned)L.h_cell, cc = (x - 3) / (unsigned)L.w_cell;
        const unsigned long long key = ((unsigned long long)score[i] << 56) |
                                       ((unsigned long long)(4095u - cr) << 44) |
                                       ((unsigned long long)(4095u - cc) << 32) |
                                       ((unsigned long long)(65535u - y) << 16) |
                                       (unsigned long long)(65535u - x);
        atomicMax(&s.best[node[i]], key);
    }
    __syncthreads();
    uint32_t *oxy = lkp_xy + (size_t)frame * plan->kp_total + L.kp_off;
    uint32_t *osc = lkp_score + (size_t)frame * plan->kp_total + L.kp_off;
    for (int j = tid; j < S; j += QT_NT) {
        const unsigned long long b = s.best[j];
        const unsigned x = 65535u - (unsigned)(b & 0xffff), y = 65535u - (unsigned)((b >> 16) & 0xffff);
        /* back to level pixel coordinates (:1116-1117) */
        oxy[j] = (x + BORB_FAST_MARGIN) | ((y + BORB_FAST_MARGIN) << 16);
        osc[j] = (uint32_t)(b >> 56);
    }
    if (tid == 0) nlkp[frame * BORB_MAXL + level] = S;
}

#endif
