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
#include "borb_cta.cuh"

/* -DBORB_QT_PROBE (tools/probe_qt.py, never in the shipped library): (tag, %globaltimer) log of thread 0 per level */
#ifdef BORB_QT_PROBE
__device__ unsigned long long g_qt_probe[8 * 128 * 2];
__device__ int g_qt_probe_n[8];
#define QT_PROBE(tag)                                                                          \
    do {                                                                                       \
        if (threadIdx.x == 0 && blockIdx.y == 0 && blockIdx.x < 8) {                           \
            unsigned long long t_;                                                             \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                             \
            const int k_ = g_qt_probe_n[blockIdx.x];                                           \
            if (k_ < 128) {                                                                     \
                g_qt_probe[(blockIdx.x * 128 + k_) * 2] = (unsigned long long)(tag);            \
                g_qt_probe[(blockIdx.x * 128 + k_) * 2 + 1] = t_;                               \
                g_qt_probe_n[blockIdx.x] = k_ + 1;                                             \
            }                                                                                  \
        }                                                                                      \
    } while (0)
#define QT_PROBE_RESET() do { if (threadIdx.x == 0 && blockIdx.y == 0 && blockIdx.x < 8) g_qt_probe_n[blockIdx.x] = 0; } while (0)
#else
#define QT_PROBE(tag) do { } while (0)
#define QT_PROBE_RESET() do { } while (0)
#endif

/* threads per CTA is a template parameter: 256 for throughput (many (level, frame) CTAs in flight;
 * measured on B200 per 512 frames: 1024 -> 0.80 ms, 512 -> 0.51 ms, 256 -> 0.48 ms), 1024 for the
 * latency of a single frame (8 CTAs in total: 45 us vs 78 us); re-measured with the round-2 kernel: 128 -> 0.32 ms,
 * 256 -> 0.216, 512 -> 0.246 */
#ifndef QT_NT_THROUGHPUT
#define QT_NT_THROUGHPUT 256
#endif
#define QT_NT_LATENCY 1024

/* Block-wide exclusive scan of up to three 21-bit fields packed in a 64-bit word.  gen(i) is the
 * packed value of element i in [0, len); thread t owns a contiguous run of elements, so no array of
 * flags has to be written (and fenced) first.  After the two internal barriers every thread calls
 * emit(i, g, before, total) for its own elements: g = gen(i), before = packed sums of the elements in
 * front of i, total = packed sums of all.  emit must not write what gen reads.  One trailing barrier
 * publishes what emit wrote; three barriers in all.  All threads must call. */
#define QT_EBITS 11 /* creation index of an expandable node: < BORB_QT_MAXNODES = 2048 */
#define QT_F 21
#define QT_FMASK ((1u << QT_F) - 1u)
typedef unsigned long long qt_u64;

template <int QT_NT, typename Gen, typename Emit>
__device__ __forceinline__ qt_u64 qt_scan_emit(int len, qt_u64 *wtmp /* 33 words */, Gen gen, Emit emit)
{
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int per = (len + QT_NT - 1) / QT_NT;
    const int b = min(tid * per, len), e = min(b + per, len);
    qt_u64 s = 0;
#pragma unroll 1
    for (int i = b; i < e; ++i) s += gen(i);
    qt_u64 v = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const qt_u64 t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    if (lane == 31) wtmp[w] = v;
    QT_PROBE(10);
    __syncthreads();
    QT_PROBE(11);
    if (w == 0) {
        const qt_u64 t = lane < QT_NT / 32 ? wtmp[lane] : 0ull;
        qt_u64 u = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const qt_u64 z = __shfl_up_sync(0xffffffffu, u, o);
            if (lane >= o) u += z;
        }
        wtmp[lane] = u - t;
        if (lane == 31) wtmp[32] = u;
    }
    QT_PROBE(12);
    __syncthreads();
    QT_PROBE(13);
    qt_u64 run = wtmp[w] + v - s;
    const qt_u64 total = wtmp[32];
#pragma unroll 1
    for (int i = b; i < e; ++i) {
        const qt_u64 g = gen(i);
        emit(i, g, run, total);
        run += g;
    }
    QT_PROBE(14);
    __syncthreads();
    QT_PROBE(15);
    return total;
}

struct QtSmem { /* the double-buffered arrays are base + buffer * M: no pointer table indexed at run time (local memory) */
    ushort4 *box_; /* x0, y0, x1, y1 (candidate coordinates are 16-bit) */
    uint32_t *mid_; /* the node's split point mx | my << 16: all the relabelling needs of a box, in one word */
    int *cnt_, *exp_size_, *ccnt_, *cf;
    uint16_t *par_, *exp_pos_; /* list positions and ranks are < BORB_QT_MAXNODES = 2048 */
    uint32_t *ru; /* per node: rank (low half, QT_NONE = not ranked) | new position if it stays (high half) */
    unsigned long long *best;
    qt_u64 *wtmp;
    int M;
    __device__ __forceinline__ ushort4 *box(int b) const { return box_ + b * M; }
    __device__ __forceinline__ uint32_t *mid(int b) const { return mid_ + b * M; }
    __device__ __forceinline__ uint16_t &rank(int j) const { return reinterpret_cast<uint16_t *>(ru)[2 * j]; }
    __device__ __forceinline__ uint16_t &uf(int j) const { return reinterpret_cast<uint16_t *>(ru)[2 * j + 1]; }
    __device__ __forceinline__ int *cnt(int b) const { return cnt_ + b * M; }
    __device__ __forceinline__ int *ccnt(int b) const { return ccnt_ + b * 4 * M; }
    __device__ __forceinline__ uint16_t *par(int b) const { return par_ + b * M; }
    __device__ __forceinline__ uint16_t *exp_pos(int b) const { return exp_pos_ + b * M; }
    __device__ __forceinline__ int *exp_size(int b) const { return exp_size_ + b * M; }
};
#define QT_NONE 0xffffu /* rank[] of a node that is not ranked */

__device__ __forceinline__ uint32_t qt_mid(const ushort4 b)
{
    const uint32_t mx = b.x + ((b.z - b.x + 1u) >> 1); /* x0 + ceil((x1-x0)/2) */
    const uint32_t my = b.y + ((b.w - b.y + 1u) >> 1);
    return mx | (my << 16);
}

/* quarter of candidate v = x | y << 16 in a node with split point m = mx | my << 16 */
__device__ __forceinline__ int qt_quadrant(uint32_t v, uint32_t m)
{
    return ((v >> 16) >= (m >> 16)) * 2 + ((v & 0xffffu) >= (m & 0xffffu));
}

__device__ __forceinline__ ushort4 qt_child_box(const ushort4 b, int q)
{
    const uint32_t m = qt_mid(b);
    const unsigned short mx = (unsigned short)(m & 0xffffu), my = (unsigned short)(m >> 16);
    ushort4 c;
    c.x = (q & 1) ? mx : b.x;
    c.z = (q & 1) ? b.z : mx;
    c.y = (q & 2) ? my : b.y;
    c.w = (q & 2) ? b.w : my;
    return c;
}

/* bytes of dynamic shared memory for a list capacity of M nodes */
static inline size_t qt_smem_bytes(int M)
{
    /* quarter counters 2x16, child positions 16, box 2x8, winner keys 8, split point 2x4, counts 2x4, rank keys 2x4, rank | stay
     * position 4, then the 16-bit arrays: par 2x2, exp_pos 2x2 */
    return (size_t)M * (32 + 16 + 16 + 8 + 8 + 8 + 8 + 4 + 4 + 4) + 40 * 8 + 16; /* + 16: zero padding of the rank keys */
}

/* The candidates of a (level, frame): in the latency shape (1024 threads, a CTA alone on its SM) a thread keeps
 * its candidates and their node numbers in registers as long as there are at most QT_RPT per thread; otherwise
 * (and always in the throughput shape) they are read from and written back to global memory every pass. */
#define QT_RPT 8
template <int QT_NT, bool REG>
struct QtPts {
    const uint32_t *__restrict__ xy;
    uint32_t *__restrict__ node;
    int n;
    uint32_t rxy[REG ? QT_RPT : 1], rnode[REG ? QT_RPT : 1];
    __device__ __forceinline__ void load()
    {
        if (REG) {
#pragma unroll
            for (int u = 0; u < QT_RPT; ++u) {
                const int i = threadIdx.x + u * QT_NT;
                rxy[u] = i < n ? xy[i] : 0u;
            }
        }
    }
    /* first assignment of the node numbers: body(xy) -> node */
    template <typename F>
    __device__ __forceinline__ void init(F body)
    {
        if (REG) {
#pragma unroll
            for (int u = 0; u < QT_RPT; ++u) {
                const int i = threadIdx.x + u * QT_NT;
                rnode[u] = i < n ? body(rxy[u]) : 0u; /* unused slots: node word 0, a valid index for staged() */
            }
        } else {
            for (int i = threadIdx.x; i < n; i += QT_NT) node[i] = body(xy[i]);
        }
    }
    /* The per-candidate steps of a pass are chains of dependent shared-memory loads (node box -> child slot -> child
     * box -> counter).  Run candidate by candidate, with a branch around each, the register-resident shape paid
     * the whole chain once per candidate (1.2 us of a 2.1 us breadth pass, tools/probe_qt.py).  staged() runs the chain
     * stage by stage over four candidates at a time: the stages take a Slot by reference and must be safe to run on an
     * unused slot (ok == false, v == 0, nd == 0: every index they form has to stay inside its array; they must not
     * store or count), so their loads carry no branch and four of them are in flight together. */
    struct Slot {
        int i;       /* candidate index */
        uint32_t v;  /* x | y << 16 */
        uint32_t nd; /* node word (written back after the last stage) */
        bool ok;
        uint32_t m, w, t; /* scratch of the stages */
    };
    template <typename A, typename B, typename C, typename D>
    __device__ __forceinline__ void staged(A a, B b, C c, D d)
    {
        if (REG) {
#pragma unroll
            for (int h = 0; h < QT_RPT; h += 4) {
                if (h * QT_NT >= n) break; /* the same for the whole CTA */
                Slot sl[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    sl[u].i = threadIdx.x + (h + u) * QT_NT;
                    sl[u].ok = sl[u].i < n;
                    sl[u].v = rxy[h + u];
                    sl[u].nd = rnode[h + u];
                    sl[u].w = 0; sl[u].t = 0;
                    a(sl[u]);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) b(sl[u]);
#pragma unroll
                for (int u = 0; u < 4; ++u) c(sl[u]);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    d(sl[u]);
                    rnode[h + u] = sl[u].nd;
                }
            }
        } else {
            for (int i = threadIdx.x; i < n; i += QT_NT) {
                Slot sl;
                sl.i = i; sl.ok = true; sl.v = xy[i]; sl.nd = node[i]; sl.w = 0; sl.t = 0;
                const uint32_t before = sl.nd;
                a(sl); b(sl); c(sl); d(sl);
                if (sl.nd != before) node[i] = sl.nd;
            }
        }
    }
};

/* One split pass over the current list (cur) into (1-cur).
 * On entry: par(cur)[k] = list node of processing rank k, ccnt(cur)[4k+q] = points of rank k's q-th quarter, for all
 * ranked nodes.  Only ranks < n_proc are actually split.  Returns the new list size; *n_exp = number of children
 * with more than one point (written to the other expandable buffer, creation order).  One packed scan gives the
 * positions of the non-empty children (field 0), of the expandable ones (field 1) and of the nodes that stay (field 2).
 *
 * BREADTH (every multi-point node splits): a candidate's node word is  position | (rank + 1) << 16  (0 in the upper
 * half: a single-point node, which stays), so no rank table is read; and the pass prepares the NEXT breadth pass as it
 * goes: after it every multi-point node is a child born in it, the children stand at the head of the new list in
 * reverse creation order and the scan has numbered the multi-point ones e = 0 .. E-1 in creation order, so a child's
 * next rank is E-1-e -- written into its node word, par(1-cur) and the zeroed counters ccnt(1-cur) by the scan's emit
 * step, and the relabelling step counts each candidate into its new node's quarter right away.  A breadth pass is the
 * scan (three barriers) and the relabel-and-count step (one).
 * Ordered pass (!BREADTH): ranks come from rank[] (QT_NONE = not ranked), node words are plain positions. */
template <int QT_NT, bool REG, bool BREADTH>
__device__ __forceinline__ int qt_split(const QtSmem &s, int cur, int S, int n_proc, QtPts<QT_NT, REG> &pts, int *n_exp, bool packed)
{
    const int nc = 4 * n_proc;
    const int nxt = 1 - cur;
    const int *const ccnt = s.ccnt(cur);
    const qt_u64 tot = qt_scan_emit<QT_NT>(
        max(nc, S), s.wtmp,
        [&](int i) -> qt_u64 {
            qt_u64 g = 0;
            if (i < nc) {
                const int n = ccnt[i];
                g = (qt_u64)(n > 0) | ((qt_u64)(n > 1) << QT_F);
            }
            if (i < S) {
                bool stays;
                if (BREADTH) stays = s.cnt(cur)[i] <= 1;
                else stays = (int)s.rank(i) >= n_proc; /* QT_NONE is larger than any rank */
                g |= (qt_u64)stays << (2 * QT_F);
            }
            return g;
        },
        [&](int i, qt_u64 g, qt_u64 before, qt_u64 total) {
            const int C = (int)(total & QT_FMASK);
            if (g & 1) {
                const int n = ccnt[i];
                const int p = C - 1 - (int)(before & QT_FMASK);
                const ushort4 cb = qt_child_box(s.box(cur)[s.par(cur)[i >> 2]], i & 3);
                s.box(nxt)[p] = cb;
                s.mid(nxt)[p] = qt_mid(cb);
                s.cnt(nxt)[p] = n;
                int word = p;
                if (n > 1) {
                    const int e = (int)((before >> QT_F) & QT_FMASK);
                    s.exp_pos(nxt)[e] = (uint16_t)p;
                    s.exp_size(nxt)[e] = packed ? (n << QT_EBITS) | e : n; /* the ordered phase ranks by (size, creation order) */
                    if (BREADTH) {
                        const int kk = (int)((total >> QT_F) & QT_FMASK) - 1 - e;
                        s.par(nxt)[kk] = (uint16_t)p;
                        *reinterpret_cast<int4 *>(&s.ccnt(nxt)[4 * kk]) = make_int4(0, 0, 0, 0);
                        word |= (kk + 1) << 16;
                    }
                }
                s.cf[i] = word;
            }
            if ((g >> (2 * QT_F)) & 1) {
                const int p = C + (int)(before >> (2 * QT_F));
                s.box(nxt)[p] = s.box(cur)[i];
                s.mid(nxt)[p] = s.mid(cur)[i];
                s.cnt(nxt)[p] = s.cnt(cur)[i];
                s.uf(i) = (uint16_t)p;
            }
        });
    typedef typename QtPts<QT_NT, REG>::Slot Slot;
    if (BREADTH) {
        pts.staged(
            [&](Slot &p) { p.m = s.mid(cur)[p.nd & 0xffffu]; },
            [&](Slot &p) { /* ONE load: the child slot of a node that splits, or the word whose high half is the new
                              position of a node that stays */
                const int k1 = (int)(p.nd >> 16);
                const uint32_t *src = k1 ? reinterpret_cast<const uint32_t *>(s.cf) + 4 * (k1 - 1) + qt_quadrant(p.v, p.m)
                                         : s.ru + (p.nd & 0xffffu);
                const uint32_t w = *src;
                p.w = p.ok ? (k1 ? w : w >> 16) : 0u;
            },
            [&](Slot &p) { p.m = s.mid(nxt)[p.w & 0xffffu]; },
            [&](Slot &p) {
                if (p.ok) p.nd = p.w;
                if (p.w >> 16) atomicAdd(&s.ccnt(nxt)[4 * ((int)(p.w >> 16) - 1) + qt_quadrant(p.v, p.m)], 1);
            });
    } else {
        pts.staged(
            [&](Slot &p) {
                p.m = s.mid(cur)[p.nd & 0xffffu];
                p.t = s.ru[p.nd & 0xffffu]; /* rank | stay position << 16 */
            },
            [&](Slot &p) {
                const int k = (int)(p.t & 0xffffu);
                const bool split = k < n_proc;
                const uint32_t ws = (uint32_t)s.cf[split ? 4 * k + qt_quadrant(p.v, p.m) : 0];
                p.w = split ? ws : p.t >> 16;
            },
            [&](Slot &) {},
            [&](Slot &p) { if (p.ok) p.nd = p.w; });
    }
    QT_PROBE(16);
    __syncthreads();
    *n_exp = (int)((tot >> QT_F) & QT_FMASK);
    return (int)(tot & QT_FMASK) + (int)(tot >> (2 * QT_F));
}

/* ordered pass: count the quarters of every ranked node; ccnt(cur)[0 .. 4 * ranked) must be zero on entry */
template <int QT_NT, bool REG>
__device__ __forceinline__ void qt_count(const QtSmem &s, int cur, QtPts<QT_NT, REG> &pts)
{
    typedef typename QtPts<QT_NT, REG>::Slot Slot;
    pts.staged(
        [&](Slot &p) {
            p.m = s.mid(cur)[p.nd & 0xffffu];
            p.t = s.rank(p.nd & 0xffffu);
        },
        [&](Slot &) {}, [&](Slot &) {},
        [&](Slot &p) {
            if (p.ok && p.t != QT_NONE) atomicAdd(&s.ccnt(cur)[4 * p.t + qt_quadrant(p.v, p.m)], 1);
        });
    __syncthreads();
}

template <int QT_NT, bool REG>
__device__ __forceinline__ void qt_run(const BorbPlan *plan, int M, const uint32_t *__restrict__ cand_xy,
                                       const uint32_t *__restrict__ cand_score, uint32_t *__restrict__ cand_node,
                                       int n, uint32_t *__restrict__ lkp_xy, uint32_t *__restrict__ lkp_score,
                                       int *__restrict__ nlkp, int *__restrict__ err, unsigned char *qt_raw, int &sh_proc)
{
    QtSmem s;
    {
        unsigned char *p = qt_raw;
        s.M = M;
        s.ccnt_ = (int *)p; p += (size_t)M * 32;
        s.cf = (int *)p; p += (size_t)M * 16;
        s.box_ = (ushort4 *)p; p += (size_t)M * 16;
        s.best = (unsigned long long *)p; p += (size_t)M * 8;
        s.wtmp = (qt_u64 *)p; p += 40 * 8;
        s.mid_ = (uint32_t *)p; p += (size_t)M * 8;
        s.cnt_ = (int *)p; p += (size_t)M * 8;
        s.exp_size_ = (int *)p; p += (size_t)M * 8 + 16;
        s.ru = (uint32_t *)p; p += (size_t)M * 4;
        s.par_ = (uint16_t *)p; p += (size_t)M * 4;
        s.exp_pos_ = (uint16_t *)p;
    }
    const int level = blockIdx.x, frame = blockIdx.y, tid = threadIdx.x;
    const BorbLevel &L = plan->lv[level];
    const size_t cbase = (size_t)frame * (size_t)plan->cand_total + L.cand_off;
    const uint32_t *score = cand_score + cbase;
    QtPts<QT_NT, REG> pts;
    pts.xy = cand_xy + cbase;
    pts.node = cand_node + cbase;
    pts.n = n;
    QT_PROBE(1 | (n << 8));
    pts.load();
    for (int i = tid * 8; i < n; i += QT_NT * 8) /* the scores are read once, at the very end: have them in L1 by then */
        asm volatile("prefetch.global.L1 [%0];" ::"l"(score + i));
    const int N = L.quota;

    /* ---- roots (:674-739): point -> root by FLOAT division, empty roots dropped ---- */
    const int n_ini = L.n_ini;
    const float hx = L.hx;
    int *const rootc = s.ccnt(1); /* the first breadth pass counts into ccnt(0) */
#pragma unroll 1
    for (int j = tid; j < n_ini; j += QT_NT) rootc[j] = 0;
#pragma unroll 1
    for (int j = tid; j < M; j += QT_NT) s.best[j] = 0ull; /* the winner keys of the last step: no barrier of their own */
    __syncthreads();
    pts.init([&](uint32_t v) -> uint32_t {
        int r = (int)__fdiv_rn((float)(v & 0xffff), hx);
        r = min(max(r, 0), n_ini - 1);
        atomicAdd(&rootc[r], 1);
        return (uint32_t)r;
    });
    __syncthreads();
    QT_PROBE(2);
    int cur = 0;
    /* the scan that drops the empty roots also ranks the multi-point ones (list order) for the first breadth pass, and
     * the relabelling step counts their quarters: from here on every breadth pass finds its ranks and counts ready */
    const qt_u64 tot0 = qt_scan_emit<QT_NT>(
        n_ini, s.wtmp, [&](int j) -> qt_u64 { const int c = rootc[j]; return (qt_u64)(c > 0) | ((qt_u64)(c > 1) << QT_F); },
        [&](int j, qt_u64 g, qt_u64 before, qt_u64) {
            if (g & 1) {
                const int p = (int)(before & QT_FMASK);
                const ushort4 rb = make_ushort4((unsigned short)(int)__fmul_rn(hx, (float)j), 0,
                                                (unsigned short)(int)__fmul_rn(hx, (float)(j + 1)), (unsigned short)L.qt_h);
                s.box(0)[p] = rb;
                s.mid(0)[p] = qt_mid(rb);
                s.cnt(0)[p] = rootc[j];
                int word = p;
                if (g >> QT_F) {
                    const int k = (int)(before >> QT_F);
                    s.par(0)[k] = (uint16_t)p;
                    *reinterpret_cast<int4 *>(&s.ccnt(0)[4 * k]) = make_int4(0, 0, 0, 0);
                    word |= (k + 1) << 16;
                }
                s.cf[j] = word;
            }
        });
    int S = (int)(tot0 & QT_FMASK);
    int n_exp = (int)(tot0 >> QT_F); /* multi-point nodes of the current list */
    typedef typename QtPts<QT_NT, REG>::Slot Slot;
    pts.staged(
        [&](Slot &p) { p.w = p.ok ? (uint32_t)s.cf[p.nd] : 0u; },
        [&](Slot &p) { p.m = s.mid(0)[p.w & 0xffffu]; },
        [&](Slot &) {},
        [&](Slot &p) {
            if (p.ok) p.nd = p.w;
            if (p.w >> 16) atomicAdd(&s.ccnt(0)[4 * ((int)(p.w >> 16) - 1) + qt_quadrant(p.v, p.m)], 1);
        });
    __syncthreads();

    QT_PROBE(3 | (S << 8));
    /* ---- split passes (:741-927) ---- */
    static_assert(BORB_QT_MAXNODES <= (1 << QT_EBITS), "rank key");
    static_assert(BORB_QT_MAXNODES < 0xffff, "16-bit positions and ranks");
    const bool packed = n < (1 << (31 - QT_EBITS)); /* (size << 11 | creation index) fits a positive int */
    bool done = (S == 0);
    while (!done) {
        /* breadth pass: every multi-point node splits */
        const int prev = S;
        const int n_split = n_exp;
        if (n_split == 0) break; /* list cannot grow: lNodes.size()==prevSize */
        S = qt_split<QT_NT, REG, true>(s, cur, S, n_split, pts, &n_exp, packed);
        cur = 1 - cur;
        QT_PROBE(4 | (S << 8));
        if (S >= N || S == prev) break;
        if (S + 3 * n_exp <= N) continue;

        /* ordered passes (:853-924) */
        while (true) {
            const int prev2 = S;
            const int E = n_exp;
            if (E == 0) { done = true; break; }
            int *const ccnt = s.ccnt(cur);
#pragma unroll 1
            for (int j = tid; j < S; j += QT_NT) s.rank(j) = (uint16_t)QT_NONE;
#pragma unroll 1
            for (int c = tid; c < 4 * E; c += QT_NT) ccnt[c] = 0;
            if (tid == 0) sh_proc = E;
            if (tid < 4) s.exp_size(cur)[E + tid] = 0; /* padding of the rank keys (the array has 4 spare entries) */
            QT_PROBE(17);
            __syncthreads();
            QT_PROBE(18);
            if (packed) { /* rank = number of larger keys; four keys per shared-memory load (zero padding never counts).
                             (Four or eight lanes per key, so that all 1024 threads of the single-frame shape compare:
                             no faster, with eight 2.9 us against 1.5 for this step, tools/probe_qt.py.  The step is its instruction
                             count: E * E / 32 warp-level compares at three instructions each (compare, add, select) on the
                             E / 32 warps that have keys.) */
                const int4 *keys = reinterpret_cast<const int4 *>(s.exp_size(cur));
#pragma unroll 1
                for (int e = tid; e < E; e += QT_NT) {
                    const int key = s.exp_size(cur)[e];
                    int r0 = 0, r1 = 0, r2 = 0, r3 = 0; /* four chains, not one of 4 * E / 4 dependent adds */
#pragma unroll 4
                    for (int f = 0; f < (E + 3) >> 2; ++f) {
                        const int4 k = keys[f];
                        r0 += k.x > key; r1 += k.y > key; r2 += k.z > key; r3 += k.w > key;
                    }
                    const int r = (r0 + r1) + (r2 + r3);
                    s.rank(s.exp_pos(cur)[e]) = (uint16_t)r;
                    s.par(cur)[r] = s.exp_pos(cur)[e];
                }
            } else {
#pragma unroll 1
                for (int e = tid; e < E; e += QT_NT) {
                    const int sz = s.exp_size(cur)[e];
                    int r = 0;
#pragma unroll 1
                    for (int f = 0; f < E; ++f) {
                        const int sf = s.exp_size(cur)[f];
                        r += (sf > sz) || (sf == sz && f > e);
                    }
                    s.rank(s.exp_pos(cur)[e]) = (uint16_t)r;
                    s.par(cur)[r] = s.exp_pos(cur)[e];
                }
            }
            QT_PROBE(19);
            __syncthreads();
            QT_PROBE(5 | (E << 8));
            qt_count<QT_NT, REG>(s, cur, pts);
            QT_PROBE(6);
            /* size after processing ranks 0..k; stop at the first k reaching N */
            qt_scan_emit<QT_NT>(
                E, s.wtmp,
                [&](int k) -> qt_u64 {
                    const int4 c = *reinterpret_cast<const int4 *>(&ccnt[4 * k]);
                    return (qt_u64)((c.x > 0) + (c.y > 0) + (c.z > 0) + (c.w > 0) - 1);
                },
                [&](int k, qt_u64 g, qt_u64 before, qt_u64) {
                    if (S + (int)before + (int)g >= N) atomicMin(&sh_proc, k + 1);
                });
            const int n_proc = sh_proc;
            QT_PROBE(7 | (n_proc << 8));
            S = qt_split<QT_NT, REG, false>(s, cur, S, n_proc, pts, &n_exp, packed);
            cur = 1 - cur;
            QT_PROBE(8 | (S << 8));
            if (S >= N || S == prev2) { done = true; break; }
        }
    }

    /* ---- one survivor per node (:930-948) ---- */
    const uint32_t mh = 0xffffffffu / (uint32_t)L.h_cell + 1u, mw = 0xffffffffu / (uint32_t)L.w_cell + 1u; /* exact for 16-bit dividends */
    if (S > M) { if (tid == 0) atomicExch(err, 2); S = M; }
    uint32_t *oxy = lkp_xy + (size_t)frame * plan->kp_total + L.kp_off;
    uint32_t *osc = lkp_score + (size_t)frame * plan->kp_total + L.kp_off;
    /* The winner of a node is its maximum under (score, then the reference's input order: cell row, cell column, y, x).
     * When a level has fewer than 64 x 64 cells of at most 64 x 64 pixels -- every level of an image up to ~1900 pixels
     * a side -- that order fits 32 bits (score 8 | cell row 6 | cell column 6 | y in cell 6 | x in cell 6, the last four
     * complemented) and shared memory has a native 32-bit atomic max; the 64-bit key needs a compare-and-swap loop per
     * candidate (ATOMS.CAST.SPIN). */
    const bool k32 = L.h_cell <= 64 && L.w_cell <= 64 && L.n_rows <= 64 && L.n_cols <= 64;
    if (k32) {
        uint32_t *const b32 = reinterpret_cast<uint32_t *>(s.best); /* zeroed with the 64-bit entries */
        pts.staged(
            [&](Slot &p) { p.t = p.ok ? score[p.i] : 0u; },
            [&](Slot &) {}, [&](Slot &) {},
            [&](Slot &p) {
                if (!p.ok) return;
                const unsigned x3 = (p.v & 0xffff) - 3u, y3 = (p.v >> 16) - 3u;
                const unsigned cr = L.h_cell > 1 ? __umulhi(y3, mh) : y3, cc = L.w_cell > 1 ? __umulhi(x3, mw) : x3;
                const unsigned yin = y3 - cr * (unsigned)L.h_cell, xin = x3 - cc * (unsigned)L.w_cell;
                atomicMax(&b32[p.nd & 0xffffu], (p.t << 24) | ((63u - cr) << 18) | ((63u - cc) << 12) | ((63u - yin) << 6) | (63u - xin));
            });
        __syncthreads();
#pragma unroll 1
        for (int j = tid; j < S; j += QT_NT) {
            const uint32_t b = b32[j];
            const unsigned cr = 63u - ((b >> 18) & 63u), cc = 63u - ((b >> 12) & 63u), yin = 63u - ((b >> 6) & 63u), xin = 63u - (b & 63u);
            const unsigned x = 3u + cc * (unsigned)L.w_cell + xin, y = 3u + cr * (unsigned)L.h_cell + yin;
            oxy[j] = (x + BORB_FAST_MARGIN) | ((y + BORB_FAST_MARGIN) << 16); /* back to level pixel coordinates (:1116-1117) */
            osc[j] = b >> 24;
        }
    } else {
        pts.staged(
            [&](Slot &p) { p.t = p.ok ? score[p.i] : 0u; },
            [&](Slot &) {}, [&](Slot &) {},
            [&](Slot &p) {
                if (!p.ok) return;
                const unsigned x = p.v & 0xffff, y = p.v >> 16;
                const unsigned cr = L.h_cell > 1 ? __umulhi(y - 3, mh) : y - 3, cc = L.w_cell > 1 ? __umulhi(x - 3, mw) : x - 3;
                const unsigned long long key = ((unsigned long long)p.t << 56) |
                                               ((unsigned long long)(4095u - cr) << 44) |
                                               ((unsigned long long)(4095u - cc) << 32) |
                                               ((unsigned long long)(65535u - y) << 16) |
                                               (unsigned long long)(65535u - x);
                atomicMax(&s.best[p.nd & 0xffffu], key);
            });
        __syncthreads();
#pragma unroll 1
        for (int j = tid; j < S; j += QT_NT) {
            const unsigned long long b = s.best[j];
            const unsigned x = 65535u - (unsigned)(b & 0xffff), y = 65535u - (unsigned)((b >> 16) & 0xffff);
            /* back to level pixel coordinates (:1116-1117) */
            oxy[j] = (x + BORB_FAST_MARGIN) | ((y + BORB_FAST_MARGIN) << 16);
            osc[j] = (uint32_t)(b >> 56);
        }
    }
    if (tid == 0) nlkp[frame * BORB_MAXL + level] = S;
    QT_PROBE(9);
}


template <int QT_NT>
__global__ void __launch_bounds__(QT_NT, 1024 / QT_NT)
k_quadtree(const __grid_constant__ BorbPlan plan_c, int M, const uint32_t *__restrict__ cand_xy_p,
           const uint32_t *__restrict__ cand_score_p, uint32_t *__restrict__ cand_node,
           const int *__restrict__ ncand_p, uint32_t *__restrict__ lkp_xy,
           uint32_t *__restrict__ lkp_score, int *__restrict__ nlkp, int *__restrict__ err)
{
    const BorbPlan *plan = &plan_c; /* the plan rides in the parameter block: constant-bank reads, no global loads */
    extern __shared__ __align__(16) unsigned char qt_raw[];
    __shared__ int sh_proc;
    const BorbLevel &L = plan->lv[blockIdx.x];
    QT_PROBE_RESET();
    QT_PROBE(0);
    /* The 1024-thread shape is the one launched programmatically behind K2 (a frame or two per call): what K2 wrote is
     * read with ordinary loads through pointers that exist only after the wait (borb_cta.cuh); the throughput shape
     * keeps the read-only loads the compiler derives from `const __restrict__`. */
    const uint32_t *cand_xy = cand_xy_p, *cand_score = cand_score_p;
    const int *ncand = ncand_p;
    if (QT_NT == QT_NT_LATENCY) {
        borb_pdl_wait(); /* K2 is complete: the candidates and their counts */
        borb_pdl_launch_dependents();
        cand_xy = borb_pdl_after_wait(cand_xy_p);
        cand_score = borb_pdl_after_wait(cand_score_p);
        ncand = borb_pdl_after_wait(ncand_p);
    }
    int n = ncand[blockIdx.y * BORB_MAXL + blockIdx.x];
    if (n > L.cand_cap) { /* cannot happen (capacity is the NMS bound); never truncate silently */
        if (threadIdx.x == 0) atomicExch(err, 1);
        n = L.cand_cap;
    }
    if (QT_NT == QT_NT_LATENCY && n <= QT_RPT * QT_NT)
        qt_run<QT_NT, QT_NT == QT_NT_LATENCY>(plan, M, cand_xy, cand_score, cand_node, n, lkp_xy, lkp_score, nlkp, err, qt_raw, sh_proc);
    else
        qt_run<QT_NT, false>(plan, M, cand_xy, cand_score, cand_node, n, lkp_xy, lkp_score, nlkp, err, qt_raw, sh_proc);
}

#endif
