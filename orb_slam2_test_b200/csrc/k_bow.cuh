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
 * B3 k_bow_candidates / k_bow_resolve   SearchByBoW: the Hamming scans of all keyframe features in parallel, then a
 *                    warp per node walks its features in the reference's order on the cached candidates, see there.
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

/* the same for two arrays at once: one barrier per stage serves both (the sorts are barrier-bound at these sizes) */
__device__ __forceinline__ void bow_bitonic2(unsigned long long *k, unsigned long long *k2, int np2)
{
    for (int size = 2; size <= np2; size <<= 1)
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            __syncthreads();
            for (int t = threadIdx.x; t < (np2 >> 1); t += BOW_ASM_NT) {
                const int lo = 2 * t - (t & (stride - 1)), hi = lo + stride;
                const bool up = (lo & size) == 0;
                const unsigned long long a = k[lo], b = k[hi], c = k2[lo], d = k2[hi];
                if ((a > b) == up) { k[lo] = b; k[hi] = a; }
                if ((c > d) == up) { k2[lo] = d; k2[hi] = c; }
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
               int *__restrict__ counts, int dual /* shared memory holds both key arrays: one sorting pass */)
{
    extern __shared__ __align__(16) unsigned long long bow_keys[];
    __shared__ int sw[33];
    __shared__ double s_norm;
    const int tid = threadIdx.x;
    const int n = n_p ? min(*n_p, cap) : n_val;
    int np2 = BOW_ASM_NT;
    while (np2 < n) np2 <<= 1;

    /* ---- FeatureVector: features grouped by node, node ids ascending (std::map order), features ascending;
     *      BowVector: unique words ascending; value = the word's weight once per feature (or once) ---- */
    unsigned long long *const wkeys = dual ? bow_keys + np2 : bow_keys;
    for (int t = tid; t < np2; t += BOW_ASM_NT) {
        const bool on = t < n && weight[t] > 0.0;
        bow_keys[t] = on ? (((unsigned long long)(uint32_t)node[t] << 32) | (uint32_t)t) : ~0ull;
        if (dual) wkeys[t] = on ? (((unsigned long long)(uint32_t)word[t] << 32) | (uint32_t)t) : ~0ull;
    }
    if (dual) bow_bitonic2(bow_keys, wkeys, np2);
    else bow_bitonic(bow_keys, np2);
    const int n_nodes = bow_segments(bow_keys, np2, sw, fv_node, fv_start, fv_feat);
    __syncthreads();
    if (!dual) {
        for (int t = tid; t < np2; t += BOW_ASM_NT)
            bow_keys[t] = (t < n && weight[t] > 0.0) ? (((unsigned long long)(uint32_t)word[t] << 32) | (uint32_t)t) : ~0ull;
        bow_bitonic(bow_keys, np2);
    }
    const int n_words = bow_segments(wkeys, np2, sw, bow_word, bow_start, nullptr);
    __syncthreads(); /* bow_start is complete (global memory, this CTA only) */
    const bool accumulate = flags & 1u;
    const unsigned norm_mode = (flags >> 1) & 3u;
    for (int u = tid; u < n_words; u += BOW_ASM_NT) {
        const int s = bow_start[u], c = bow_start[u + 1] - s;
        const double w = weight[(uint32_t)wkeys[s]];
        double v = w;
        if (accumulate)
            for (int j = 1; j < c; ++j) v = __dadd_rn(v, w); /* addWeight: += in feature order */
        if ((flags & 8u) && accumulate) v = __ddiv_rn(v, (double)n_words);
        bow_val[u] = v;
    }
    __syncthreads();
    if (norm_mode) { /* BowVector::normalize: the sum runs in word order */
        double *vals = reinterpret_cast<double *>(bow_keys); /* the node keys are dead; the word keys have been read */
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

/* ---- ORBmatcher::SearchByBoW(pKF, F, vpMapPointMatches), src/ORBmatcher.cc:195-351 ----
 * Inside a node the reference loop is sequential in the keyframe feature: a frame feature matched by an earlier
 * keyframe feature is skipped by every later one (:243-244); nodes are independent (a feature belongs to one node).
 * Two kernels, as for the projection matchers (k_match.cuh):
 *   B3a k_bow_candidates  a warp per keyframe feature, over the whole GPU: Hamming distance to every frame feature of
 *       the same node, the BOW_K best (distance << 16 | position in the node's list) kept in ascending order -- one key
 *       per lane, an insertion is one shuffle, one max and one min -- plus "there were more".  Nothing here depends on
 *       other queries.
 *   B3b k_bow_resolve     a warp per node walks its keyframe features IN ORDER: the first two cached candidates that no
 *       earlier feature has taken are the reference's bestDist1 / bestDist2 (its `<` / `else if <` chain keeps the
 *       first of equal distances: the smaller list position); a taken-bit per frame feature of the node lives in
 *       shared memory; only a feature whose cached candidates are (all but at most one) taken AND that had more
 *       scans the node again.  A step of the walk is a shared-memory lookup and a ballot, not a Hamming scan. */
#define BOW_K 16
#define BOW_CAND_NT 256
#define BOW_RES_NT 128
#define BOW_NOKEY 0xffffffffu

/* per keyframe list position p: pos_node [p] = its node's index in the keyframe's feature vector; per node:
 * node_f [a] = (first, end) of the same node's list in the frame's feature vector, (0, 0) if the frame has none */
__global__ void __launch_bounds__(BOW_CAND_NT)
k_bow_candidates(const uint4 *__restrict__ kf_desc, const uint8_t *__restrict__ kf_mp, const uint32_t *__restrict__ kf_feat,
                 const int *__restrict__ pos_node, const int2 *__restrict__ node_f, int n_pos,
                 const uint4 *__restrict__ f_desc, const uint32_t *__restrict__ f_feat, uint32_t *__restrict__ cache,
                 uint8_t *__restrict__ more)
{
    const int lane = threadIdx.x & 31;
    const int p = blockIdx.x * (BOW_CAND_NT / 32) + (threadIdx.x >> 5);
    if (p >= n_pos) return; /* whole warp */
    uint32_t T = BOW_NOKEY; /* lane l < BOW_K: the l-th best key so far */
    int cnt = 0;
    const uint32_t rk = kf_feat[p];
    const uint32_t fl = kf_mp[rk];
    if ((fl & 1u) && !(fl & 2u)) { /* a map point, and not a bad one (:230-234) */
        const int2 nf = node_f[pos_node[p]];
        const uint4 d0 = kf_desc[2 * rk], d1 = kf_desc[2 * rk + 1];
        for (int base = nf.x; base < nf.y; base += 32) {
            const int jf = base + lane;
            uint32_t key = BOW_NOKEY;
            if (jf < nf.y) {
                const int d = mt_hamming(d0, d1, f_desc + 2 * (size_t)f_feat[jf]);
                if (d < 256) key = ((uint32_t)d << 16) | (uint32_t)(jf - nf.x); /* 256 never beats bestDist = 256 */
            }
            cnt += __popc(__ballot_sync(0xffffffffu, key != BOW_NOKEY));
            while (true) { /* move this chunk's keys that beat the cached worst into the sorted cache */
                const uint32_t m = __reduce_min_sync(0xffffffffu, key);
                if (m >= __shfl_sync(0xffffffffu, T, BOW_K - 1)) break;
                if (key == m) key = BOW_NOKEY;
                const uint32_t up = __shfl_up_sync(0xffffffffu, T, 1);
                T = min(T, max(m, lane ? up : 0u));
            }
        }
    }
    if (lane < BOW_K) cache[(size_t)p * BOW_K + lane] = T;
    if (lane == 0) more[p] = cnt > BOW_K;
}

/* match [kp_total] preset to -1, hist [MT_HISTO] preset to 0.  Dynamic shared memory: taken bits (tw words per warp),
 * then a 32 x (BOW_K + 1) tile of cached keys per warp. */
__global__ void __launch_bounds__(BOW_RES_NT)
k_bow_resolve(const uint4 *__restrict__ kf_desc, const float *__restrict__ kf_angle, const uint32_t *__restrict__ kf_feat,
              const int *__restrict__ kf_start, const int2 *__restrict__ node_f, int kf_nodes,
              const uint4 *__restrict__ f_desc, const BorbKpRec *__restrict__ f_kp, const uint32_t *__restrict__ f_feat,
              const uint32_t *__restrict__ cache, const uint8_t *__restrict__ more, float nnratio, int check_ori, int tw,
              int *choice /* scratch, one per keyframe list position */, int *__restrict__ match, uint8_t *__restrict__ bin_of, int *__restrict__ hist, int *__restrict__ rescans,
              int th_low /* accept best <= th_low: TH_LOW for (KeyFrame, Frame) :273, TH_LOW - 1 for (KeyFrame, KeyFrame) whose test is strict :712 */,
              const float *__restrict__ f_angle /* the second side's angles when it is a keyframe (no cv::KeyPoint records on the device), else NULL */)
{
    extern __shared__ uint32_t bow_taken[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int a = blockIdx.x * (BOW_RES_NT / 32) + warp;
    if (a >= kf_nodes) return; /* whole warp */
    uint32_t *taken = bow_taken + warp * tw;
    const int2 nf = node_f[a];
    const int fs = nf.x, fe = nf.y;
    if (fe <= fs) return;
    for (int i = lane; i < ((fe - fs + 31) >> 5); i += 32) taken[i] = 0u;
    __syncwarp();
    const int ks = kf_start[a], ke = kf_start[a + 1];
    /* The walk, 32 features at a time: lane j holds feature p0 + j's cached keys in registers and a mask of the ones
     * still open, and the decision those give (first two open keys); step j broadcasts lane j's decision, an accepted
     * position is struck from every lane's mask and only a lane that lost an open key decides again -- the serial
     * chain of a step is one shuffle.  Taken bits also go to shared memory, for the next block's masks and for the rare full scan. */
    uint32_t *tile = bow_taken + (BOW_RES_NT / 32) * tw + warp * (32 * (BOW_K + 1));
    for (int p0 = ks; p0 < ke; p0 += 32) {
        const int nsteps = min(32, ke - p0);
        __syncwarp();
#pragma unroll
        for (int i = 0; i < BOW_K; ++i) { /* all loads of the block in flight at once; rows padded to BOW_K + 1 words */
            const int row = 2 * i + (lane >> 4);
            tile[row * (BOW_K + 1) + (lane & 15)] = row < nsteps ? cache[(size_t)(p0 + row) * BOW_K + (lane & 15)] : BOW_NOKEY;
        }
        const int mo = lane < nsteps ? more[p0 + lane] : 0;
        __syncwarp();
        uint32_t K[BOW_K], mask = 0u;
#pragma unroll
        for (int k = 0; k < BOW_K; ++k) {
            K[k] = tile[lane * (BOW_K + 1) + k];
            if (K[k] != BOW_NOKEY && !((taken[(K[k] & 0xffffu) >> 5] >> (K[k] & 31u)) & 1u)) mask |= 1u << k;
        }
        /* every lane keeps the decision its feature would take with its current mask (position, -1 none, -2 "scan the
         * node again"): a step is then one shuffle, and only a lane that loses an open key decides again */
#define BOW_DECIDE()                                                                                          \
    do {                                                                                                      \
        dec = -1;                                                                                             \
        if (K[0] != BOW_NOKEY) { /* else: no map point, or no candidate */                                    \
            if (__popc(mask) >= 2 || !mo) { /* the cached order is (distance, list position): first two open keys */ \
                if (mask) {                                                                                   \
                    const uint32_t k1 = tile[lane * (BOW_K + 1) + __ffs(mask) - 1];                            \
                    const uint32_t m2 = mask & (mask - 1u);                                                   \
                    const int best2 = m2 ? (int)(tile[lane * (BOW_K + 1) + __ffs(m2) - 1] >> 16) : 256;        \
                    const int best1 = (int)(k1 >> 16);                                                        \
                    if (best1 <= th_low && (float)best1 < nnratio * (float)best2) dec = (int)(k1 & 0xffffu);     \
                }                                                                                             \
            } else dec = -2; /* fewer than two open among the cached ones and there were more */              \
        }                                                                                                     \
    } while (0)
        int mypos = -1, dec;
        BOW_DECIDE();
        for (int j = 0; j < nsteps; ++j) {
            int pos = __shfl_sync(0xffffffffu, dec, j);
            if (pos == -2) { /* scan the node again, all lanes */
                __syncwarp(); /* this block's taken bits */
                if (lane == 0) atomicAdd(rescans, 1);
                const uint32_t rk = kf_feat[p0 + j];
                const uint4 d0 = kf_desc[2 * rk], d1 = kf_desc[2 * rk + 1];
                uint32_t B1 = (256u << 16) | 0xffffu;
                int best2 = 256;
                for (int base = fs; base < fe; base += 32) {
                    const int jf = base + lane;
                    uint32_t k = BOW_NOKEY;
                    if (jf < fe && !((taken[(jf - fs) >> 5] >> ((jf - fs) & 31)) & 1u)) {
                        const int d = mt_hamming(d0, d1, f_desc + 2 * (size_t)f_feat[jf]);
                        if (d < 256) k = ((uint32_t)d << 16) | (uint32_t)(jf - fs);
                    }
                    const uint32_t m1 = __reduce_min_sync(0xffffffffu, k);
                    if (m1 == BOW_NOKEY) continue;
                    const uint32_t m2 = __reduce_min_sync(0xffffffffu, k == m1 ? BOW_NOKEY : k);
                    const int m2d = m2 == BOW_NOKEY ? 256 : (int)(m2 >> 16);
                    if (m1 < B1) { best2 = min((int)(B1 >> 16), m2d); B1 = m1; }
                    else best2 = min(best2, (int)(m1 >> 16));
                }
                const int best1 = (int)(B1 >> 16);
                pos = (best1 <= th_low && (float)best1 < nnratio * (float)best2) ? (int)(B1 & 0xffffu) : -1;
            }
            if (pos >= 0) {
                if (lane == j) {
                    mypos = pos;
                    taken[pos >> 5] |= 1u << (pos & 31);
                }
                const uint32_t before = mask;
#pragma unroll
                for (int k = 0; k < BOW_K; ++k)
                    if ((K[k] & 0xffffu) == (uint32_t)pos) mask &= ~(1u << k);
                if (mask != before && lane > j) BOW_DECIDE();
            }
        }
#undef BOW_DECIDE
        if (lane < nsteps) choice[p0 + lane] = mypos;
    }
    __syncwarp();
    for (int p = ks + lane; p < ke; p += 32) {
        const int pos = choice[p];
        if (pos < 0) continue;
        const uint32_t rk = kf_feat[p], bi = f_feat[fs + pos];
        match[bi] = (int)rk;
        if (check_ori) {
            const int bin = mt_bin(kf_angle[rk], f_angle ? f_angle[bi] : f_kp[bi].angle);
            bin_of[bi] = (uint8_t)bin;
            atomicAdd(&hist[bin], 1);
        }
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

/* ---- ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo), src/ORBmatcher.cc:779-966, with
 * CheckDistEpipolarLine :165-182 (LocalMapping::CreateNewMapPoints) ----
 * The reference never sets vbMatched2, so the keypoints of pKF1 are independent: a warp per pKF1 list position, the node's
 * pKF2 list 32 at a time.  The running best of the reference starts at TH_LOW and a candidate with dist <= best that
 * passes the geometry takes over (:867-888), i.e. the LAST of the smallest distances among the candidates that pass wins:
 * the minimum of (distance << 16 | 0xffff - list position).  The host has dropped from pKF2's lists what can never be a
 * candidate (a keypoint with a map point, :850-851; a monocular one under bOnlyStereo, :855-857).
 * Float forms as the reference object has them (oracle/bow_oracle.cpp); the library is built with --fmad=false, so only
 * the fmaf() written here fuses. */
struct TriParams {
    float F[9];   /* F12, row-major */
    float ex, ey; /* the epipole of pKF1's centre in pKF2 (:785-793), evaluated on the host */
    int only_stereo;
    float scale2[BORB_MAXL], sigma2[BORB_MAXL]; /* pKF2->mvScaleFactors, mvLevelSigma2 */
};
#define BOW_TH_LOW_TRI 50

__global__ void __launch_bounds__(BOW_CAND_NT)
k_bow_triangulate(const __grid_constant__ TriParams tp, const uint4 *__restrict__ desc1, const uint8_t *__restrict__ mp1,
                  const float2 *__restrict__ xy1, const float *__restrict__ ang1, const float *__restrict__ ur1,
                  const uint32_t *__restrict__ feat1, const int *__restrict__ pos_node, const int2 *__restrict__ node_f, int n_pos,
                  const uint4 *__restrict__ desc2, const float2 *__restrict__ xy2, const int *__restrict__ oct2,
                  const float *__restrict__ ang2, const float *__restrict__ ur2, const uint32_t *__restrict__ feat2,
                  int check_ori, int *__restrict__ match12, uint8_t *__restrict__ bin_of, int *__restrict__ hist)
{
    const int lane = threadIdx.x & 31;
    const int p = blockIdx.x * (BOW_CAND_NT / 32) + (threadIdx.x >> 5);
    if (p >= n_pos) return; /* whole warp */
    const uint32_t i1 = feat1[p];
    if (mp1[i1] & 1u) return; /* has a map point already (:823-824) */
    const bool stereo1 = ur1 && ur1[i1] >= 0.f;
    if (tp.only_stereo && !stereo1) return;
    const float2 P1 = xy1[i1];
    /* the epipolar line of keypoint 1 in image 2 (:168-170) */
    const float la = __fadd_rn(fmaf(P1.x, tp.F[0], __fmul_rn(P1.y, tp.F[3])), tp.F[6]);
    const float lb = __fadd_rn(fmaf(P1.x, tp.F[1], __fmul_rn(P1.y, tp.F[4])), tp.F[7]);
    const float lc = __fadd_rn(fmaf(P1.y, tp.F[5], __fmul_rn(P1.x, tp.F[2])), tp.F[8]);
    const float den = fmaf(la, la, __fmul_rn(lb, lb));
    const int2 nf = node_f[pos_node[p]];
    const uint4 d0 = desc1[2 * i1], d1 = desc1[2 * i1 + 1];
    uint32_t best = BOW_NOKEY;
    for (int base = nf.x; base < nf.y; base += 32) {
        const int jf = base + lane;
        if (jf >= nf.y) continue;
        const uint32_t i2 = feat2[jf];
        const int d = mt_hamming(d0, d1, desc2 + 2 * (size_t)i2);
        if (d > BOW_TH_LOW_TRI) continue;
        const float2 P2 = xy2[i2];
        const int o2 = oct2[i2];
        const bool stereo2 = ur2 && ur2[i2] >= 0.f;
        if (!stereo1 && !stereo2) { /* too close to the epipole (:872-879) */
            const float dx = __fsub_rn(tp.ex, P2.x), dy = __fsub_rn(tp.ey, P2.y);
            if (fmaf(dx, dx, __fmul_rn(dy, dy)) < __fmul_rn(100.f, tp.scale2[o2])) continue;
        }
        if (den == 0.f) continue;
        const float num = __fadd_rn(lc, fmaf(lb, P2.y, __fmul_rn(la, P2.x)));
        const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
        if (!((double)dsqr < __dmul_rn(3.84, (double)tp.sigma2[o2]))) continue;
        best = min(best, ((uint32_t)d << 16) | (0xffffu - (uint32_t)(jf - nf.x)));
    }
    best = __reduce_min_sync(0xffffffffu, best);
    if (lane == 0 && best != BOW_NOKEY) {
        const uint32_t i2 = feat2[nf.x + (int)(0xffffu - (best & 0xffffu))];
        match12[i1] = (int)i2;
        if (check_ori) {
            const int bin = mt_bin(ang1[i1], ang2[i2]);
            bin_of[i1] = (uint8_t)bin;
            atomicAdd(&hist[bin], 1);
        }
    }
}

#endif
