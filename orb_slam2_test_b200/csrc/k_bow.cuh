/*
 * k_bow.cuh -- the bag-of-words step of a frame and the matcher that consumes it, on the device:
 *   Frame::ComputeBoW                                   /root/reference/src/Frame.cc:532-539
 *     -> TemplatedVocabulary::transform(features, BowVector, FeatureVector, levelsup = 4)
 *                                                       Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1206
 *     -> transform(feature, word_id, weight, nid, levelsup)   :1218-1270  (descent of one descriptor)
 *     -> FORB::distance                                 Thirdparty/DBoW2/DBoW2/FORB.cpp:81-101
 *     -> BowVector::addWeight / addIfNotExist / normalize     Thirdparty/DBoW2/DBoW2/BowVector.cpp:33-94
 *   ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&)   src/ORBmatcher.cc:195-351
 *
 * The vocabulary tree lives in device memory in CHILD-SLOT order: the children of a node occupy consecutive
 * slots (in the reference's children order = file order), and everything the descent needs of a child --
 * descriptor, where ITS children start and how many there are, node id -- is indexed by slot, so one level of
 * the descent is one round of independent loads (a lane per child) and one packed (distance << 16 | child)
 * minimum over the group: the reference's strict `<` keeps the first of equal distances, which is the smaller
 * child index.  k = 10 (ORBvoc.txt) puts two descriptors in a warp (groups of 16 lanes).
 *
 * B1 k_bow_descend   a lane group per descriptor: word id, node id at level L - levelsup, weight
 * B2 k_bow_assemble  one CTA per frame: the two std::map structures as sorted arrays.  Shared-memory bitonic
 *                    sort of (node << 32 | feature) -> FeatureVector in CSR form (node ids ascending, feature
 *                    indices ascending inside a node = push_back order); sort of (word << 32 | feature) ->
 *                    unique words, each value = its weight added once per feature, left to right, as addWeight
 *                    does (TF-IDF / TF) or once (IDF / BINARY); then BowVector::normalize: the norm is summed
 *                    SEQUENTIALLY in word order by one thread, because that order is what the reference's
 *                    double rounding follows (L2: fma(v, v, norm), as the reference object contracts it).
 * B3 k_bow_search    a warp per keyframe node: the reference's loop over a node's keyframe features is
 *                    sequential (a frame feature matched by an earlier one is skipped, :243-244) but nodes are
 *                    independent (a feature belongs to one node); inside, lanes scan the frame features of the
 *                    node 32 at a time for the best and second-best distance.
 * B4 k_bow_finish    orientation histogram -> three maxima (:1800-1843) -> removal, nmatches.
 */
#ifndef K_BOW_CUH
#define K_BOW_CUH

#include "borb_internal.h"
#include "k_describe.cuh" /* BorbKpRec */
#include "k_match.cuh"    /* mt_hamming, mt_bin, MT_HISTO */

#define BOW_NT 256          /* B1 */
#define BOW_ASM_NT 1024     /* B2 */
#define BOW_MAX_FEATURES 16384 /* per frame: B2 sorts 64-bit keys in shared memory (128 KB) */
#define BOW_TH_LOW 50

struct BowVocDev {
    const uint4 *cdesc;    /* [slots][2] */
    const int2 *cinfo;     /* [slots] first child slot and child count of the node in this slot */
    const int *cnode;      /* [slots] node id */
    const int *cword;      /* [slots] word id (leaves) */
    const double *cweight; /* [slots] */
    int root_start, root_count, L, pad;
};

/* n = min(*n_p, cap) when n_p is given (a frame's keypoint count on the device), else n_val */
template <int G>
__global__ void __launch_bounds__(BOW_NT)
k_bow_descend(const BowVocDev V, const uint4 *__restrict__ desc, const int *__restrict__ n_p, int n_val, int cap,
              int levelsup, int *__restrict__ word, int *__restrict__ node, double *__restrict__ weight)
{
    const int n = n_p ? min(*n_p, cap) : n_val;
    const int lane = threadIdx.x & 31, lig = lane % G;
    const int i = blockIdx.x * (BOW_NT / G) + threadIdx.x / G;
    if (i >= n) return; /* whole group */
    const unsigned gmask = G == 32 ? 0xffffffffu : (((1u << (G & 31)) - 1u) << (lane - lig));
    const uint4 f0 = desc[2 * i], f1 = desc[2 * i + 1];
    const int nid_level = V.L - levelsup;
    int c0 = V.root_start, nc = V.root_count, level = 0, nid = 0, slot = 0;
    while (nc > 0) {
        ++level;
        uint32_t best = 0xffffffffu;
        int bx = 0, by = 0, bnode = 0;
        for (int base = 0; base < nc; base += G) {
            const int j = base + lig;
            uint32_t key = 0xffffffffu;
            int2 info = make_int2(0, 0);
            int nd = 0;
            if (j < nc) {
                const int s = c0 + j;
                key = ((uint32_t)mt_hamming(f0, f1, V.cdesc + 2 * (size_t)s) << 16) | (uint32_t)j;
                info = V.cinfo[s];
                nd = V.cnode[s];
            }
            const uint32_t m = __reduce_min_sync(gmask, key);
            if (m < best) { /* uniform over the group */
                best = m;
                const int src = (lane - lig) + (int)(m & 0xffffu) - base;
                bx = __shfl_sync(gmask, info.x, src);
                by = __shfl_sync(gmask, info.y, src);
                bnode = __shfl_sync(gmask, nd, src);
            }
        }
        slot = c0 + (int)(best & 0xffffu);
        if (level == nid_level) nid = bnode;
        c0 = bx;
        nc = by;
    }
    if (lig == 0) {
        word[i] = V.cword[slot];
        node[i] = nid;
        weight[i] = V.cweight[slot];
    }
}

/* ascending bitonic sort of np2 (a power of two) 64-bit keys in shared memory, whole CTA */
__device__ __forceinline__ void bow_bitonic(unsigned long long *k, int np2)
{
    for (int size = 2; size <= np2; size <<= 1)
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            __syncthreads();
            for (int t = threadIdx.x; t < (np2 >> 1); t += BOW_ASM_NT) {
                const int lo = 2 * t - (t & (stride - 1)), hi = lo + stride;
                const unsigned long long a = k[lo], b = k[hi];
                if ((a > b) == ((lo & size) == 0)) { k[lo] = b; k[hi] = a; }
            }
        }
    __syncthreads();
}

/* exclusive scan of one int per thread over the CTA (BOW_ASM_NT threads); *total = sum.  sw: 33 ints of shared memory */
__device__ __forceinline__ int bow_block_exscan(int v, int *sw, int *total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += up;
    }
    __syncthreads(); /* sw may still be read from a previous call */
    if (lane == 31) sw[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int w = sw[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int up = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += up;
        }
        sw[lane] = w; /* inclusive warp totals */
        if (lane == 31) sw[32] = w;
    }
    __syncthreads();
    *total = sw[32];
    return inc - v + (warp ? sw[warp - 1] : 0);
}

/* the sorted keys' segments: ids [u] = high word of the first key of segment u, start [u] = its position,
 * start [n_seg] = number of valid keys; low [t] = low word of key t (the feature index) when given.  Returns n_seg. */
__device__ __forceinline__ int bow_segments(const unsigned long long *k, int np2, int *sw, uint32_t *__restrict__ ids,
                                            int *__restrict__ start, uint32_t *__restrict__ low)
{
    const int chunk = np2 / BOW_ASM_NT, t0 = threadIdx.x * chunk; /* np2 >= BOW_ASM_NT */
    int heads = 0, valid = 0;
    for (int t = t0; t < t0 + chunk; ++t) {
        const unsigned long long key = k[t];
        if (key == ~0ull) break; /* padding sorts last */
        ++valid;
        heads += (t == 0 || (uint32_t)(k[t - 1] >> 32) != (uint32_t)(key >> 32));
    }
    int n_seg, n_valid;
    int u = bow_block_exscan(heads, sw, &n_seg);
    (void)bow_block_exscan(valid, sw, &n_valid);
    for (int t = t0; t < t0 + valid; ++t) {
        const unsigned long long key = k[t];
        if (low) low[t] = (uint32_t)key;
        if (t == 0 || (uint32_t)(k[t - 1] >> 32) != (uint32_t)(key >> 32)) {
            ids[u] = (uint32_t)(key >> 32);
            start[u] = t;
            ++u;
        }
    }
    if (threadIdx.x == 0) start[n_seg] = n_valid;
    return n_seg;
}

/* flags: bit 0 accumulate (TF-IDF, TF), bits 1-2 norm (0 none, 1 L1, 2 L2), bit 3 divide by the number of words
 * when accumulating without a norm.  counts[0] = words, counts[1] = nodes.  Dynamic shared memory: np2_cap keys. */
__global__ void __launch_bounds__(BOW_ASM_NT, 1)
k_bow_assemble(const int *__restrict__ word, const int *__restrict__ node, const double *__restrict__ weight,
               const int *__restrict__ n_p, int n_val, int cap, unsigned flags, uint32_t *__restrict__ bow_word,
               double *__restrict__ bow_val, int *__restrict__ bow_start /* scratch [cap + 1] */,
               uint32_t *__restrict__ fv_node, int *__restrict__ fv_start, uint32_t *__restrict__ fv_feat,
               int *__restrict__ counts)
{
    extern __shared__ __align__(16) unsigned long long bow_keys[];
    __shared__ int sw[33];
    __shared__ double s_norm;
    const int tid = threadIdx.x;
    const int n = n_p ? min(*n_p, cap) : n_val;
    int np2 = BOW_ASM_NT;
    while (np2 < n) np2 <<= 1;

    /* ---- FeatureVector: features grouped by node, node ids ascending (std::map order), features ascending ---- */
    for (int t = tid; t < np2; t += BOW_ASM_NT)
        bow_keys[t] = (t < n && weight[t] > 0.0) ? (((unsigned long long)(uint32_t)node[t] << 32) | (uint32_t)t) : ~0ull;
    bow_bitonic(bow_keys, np2);
    const int n_nodes = bow_segments(bow_keys, np2, sw, fv_node, fv_start, fv_feat);
    __syncthreads();

    /* ---- BowVector: unique words ascending; value = the word's weight once per feature (or once) ---- */
    for (int t = tid; t < np2; t += BOW_ASM_NT)
        bow_keys[t] = (t < n && weight[t] > 0.0) ? (((unsigned long long)(uint32_t)word[t] << 32) | (uint32_t)t) : ~0ull;
    bow_bitonic(bow_keys, np2);
    const int n_words = bow_segments(bow_keys, np2, sw, bow_word, bow_start, nullptr);
    __syncthreads(); /* bow_start is complete (global memory, this CTA only) */
    const bool accumulate = flags & 1u;
    const unsigned norm_mode = (flags >> 1) & 3u;
    for (int u = tid; u < n_words; u += BOW_ASM_NT) {
        const int s = bow_start[u], c = bow_start[u + 1] - s;
        const double w = weight[(uint32_t)bow_keys[s]];
        double v = w;
        if (accumulate)
            for (int j = 1; j < c; ++j) v = __dadd_rn(v, w); /* addWeight: += in feature order */
        if ((flags & 8u) && accumulate) v = __ddiv_rn(v, (double)n_words);
        bow_val[u] = v;
    }
    __syncthreads();
    if (norm_mode) { /* BowVector::normalize: the sum runs in word order */
        double *vals = reinterpret_cast<double *>(bow_keys); /* the keys are dead */
        for (int u = tid; u < n_words; u += BOW_ASM_NT) vals[u] = bow_val[u];
        __syncthreads();
        if (tid == 0) {
            double norm = 0.0;
            if (norm_mode == 1) for (int u = 0; u < n_words; ++u) norm = __dadd_rn(norm, fabs(vals[u]));
            else {
                for (int u = 0; u < n_words; ++u) norm = __fma_rn(vals[u], vals[u], norm);
                norm = __dsqrt_rn(norm);
            }
            s_norm = norm;
        }
        __syncthreads();
        const double norm = s_norm;
        if (norm > 0.0)
            for (int u = tid; u < n_words; u += BOW_ASM_NT) bow_val[u] = __ddiv_rn(vals[u], norm);
    }
    if (tid == 0) { counts[0] = n_words; counts[1] = n_nodes; }
}

/* ---- ORBmatcher::SearchByBoW(pKF, F, vpMapPointMatches), src/ORBmatcher.cc:195-351 ---- */
#define BOW_SEARCH_NT 128

/* match [kp_total] preset to -1, hist [MT_HISTO] preset to 0.  kf_mp: bit 0 = map point, bit 1 = isBad(). */
__global__ void __launch_bounds__(BOW_SEARCH_NT)
k_bow_search(const uint4 *__restrict__ kf_desc, const float *__restrict__ kf_angle, const uint8_t *__restrict__ kf_mp,
             const uint32_t *__restrict__ kf_node, const int *__restrict__ kf_start, const uint32_t *__restrict__ kf_feat,
             int kf_nodes, const uint4 *__restrict__ f_desc, const BorbKpRec *__restrict__ f_kp,
             const uint32_t *__restrict__ f_node, const int *__restrict__ f_start, const uint32_t *__restrict__ f_feat,
             const int *__restrict__ f_counts, float nnratio, int check_ori, int *match, uint8_t *__restrict__ bin_of,
             int *__restrict__ hist)
{
    const int lane = threadIdx.x & 31;
    const int a = blockIdx.x * (BOW_SEARCH_NT / 32) + (threadIdx.x >> 5);
    if (a >= kf_nodes) return;
    int b;
    {   /* the frame's list of the same node: both lists are sorted, the reference walks them in step (:219-338) */
        const uint32_t want = kf_node[a];
        int lo = 0, hi = f_counts[1];
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (f_node[mid] < want) lo = mid + 1;
            else hi = mid;
        }
        if (lo >= f_counts[1] || f_node[lo] != want) return;
        b = lo;
    }
    const int fs = f_start[b], fe = f_start[b + 1];
    for (int ik = kf_start[a]; ik < kf_start[a + 1]; ++ik) {
        const uint32_t rk = kf_feat[ik];
        const uint32_t fl = kf_mp[rk];
        if (!(fl & 1u) || (fl & 2u)) continue; /* no map point, or a bad one (:230-234) */
        const uint4 d0 = kf_desc[2 * rk], d1 = kf_desc[2 * rk + 1];
        uint32_t B1 = (256u << 16) | 0xffffu; /* best (distance, list position); 256 = none yet */
        int B2 = 256;
        for (int base = fs; base < fe; base += 32) {
            const int jf = base + lane;
            uint32_t key = 0xffffffffu;
            if (jf < fe) {
                const uint32_t rf = f_feat[jf];
                if (*(volatile int *)(match + rf) < 0) { /* not matched by an earlier keyframe feature (:243-244) */
                    const int d = mt_hamming(d0, d1, f_desc + 2 * (size_t)rf);
                    if (d < 256) key = ((uint32_t)d << 16) | (uint32_t)(jf - fs);
                }
            }
            const uint32_t m1 = __reduce_min_sync(0xffffffffu, key);
            if (m1 == 0xffffffffu) continue;
            const uint32_t m2 = __reduce_min_sync(0xffffffffu, key == m1 ? 0xffffffffu : key);
            const int m2d = m2 == 0xffffffffu ? 256 : (int)(m2 >> 16);
            if (m1 < B1) { B2 = min((int)(B1 >> 16), m2d); B1 = m1; }
            else B2 = min(B2, (int)(m1 >> 16));
        }
        const int best1 = (int)(B1 >> 16);
        if (best1 <= BOW_TH_LOW && (float)best1 < nnratio * (float)B2) {
            if (lane == 0) {
                const uint32_t bi = f_feat[fs + (int)(B1 & 0xffffu)];
                match[bi] = (int)rk;
                if (check_ori) {
                    const int bin = mt_bin(kf_angle[rk], f_kp[bi].angle);
                    bin_of[bi] = (uint8_t)bin;
                    atomicAdd(&hist[bin], 1);
                }
            }
        }
        __syncwarp(); /* lane 0's match[] write is visible to the warp's next scan */
    }
}

/* nm[0] = nmatches */
__global__ void __launch_bounds__(1024, 1)
k_bow_finish(const int *__restrict__ n_p, int cap, int check_ori, const int *__restrict__ hist,
             const uint8_t *__restrict__ bin_of, int *__restrict__ match, int *__restrict__ nm)
{
    __shared__ int s_keep[MT_HISTO], s_nm;
    const int tid = threadIdx.x, n = min(*n_p, cap);
    if (tid == 0) {
        s_nm = 0;
        int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1; /* ComputeThreeMaxima :1800-1843 */
        for (int b = 0; b < MT_HISTO; ++b) {
            const int s = hist[b];
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = b; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = b; }
            else if (s > max3) { max3 = s; ind3 = b; }
        }
        if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
        else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
        for (int b = 0; b < MT_HISTO; ++b) s_keep[b] = !check_ori || b == ind1 || b == ind2 || b == ind3;
    }
    __syncthreads();
    int mine = 0;
    for (int i = tid; i < cap; i += 1024) {
        if (i >= n) { match[i] = -1; continue; }
        if (match[i] < 0) continue;
        if (check_ori && !s_keep[bin_of[i]]) match[i] = -1;
        else ++mine;
    }
    mine = __reduce_add_sync(0xffffffffu, mine);
    if ((tid & 31) == 0 && mine) atomicAdd(&s_nm, mine);
    __syncthreads();
    if (tid == 0) nm[0] = s_nm;
}

#endif
