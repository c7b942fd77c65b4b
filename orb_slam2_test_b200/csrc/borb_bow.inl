/*
 * borb_bow.inl -- host side of the bag-of-words entry points (include/borb.h), part of borb_api.cu:
 *   borb_voc_load_text / borb_voc_create     TemplatedVocabulary::loadFromTextFile
 *                                            /root/reference/Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1338-1421
 *   borb_compute_bow / borb_voc_transform    Frame::ComputeBoW src/Frame.cc:532-539 -> transform :1127-1206
 *   borb_search_by_bow                       ORBmatcher::SearchByBoW(KeyFrame*, Frame&, ...) src/ORBmatcher.cc:195-351
 * Kernels: k_bow.cuh.  The vocabulary is uploaded once, in child-slot order (see k_bow.cuh); a frame's bow vector
 * and feature vector are built on the device from the descriptors the extract call left there and stay there for
 * the matcher; one slab comes back to page-locked host memory.
 */
#include "k_bow.cuh"

struct borb_vocabulary {
    int device = 0, k = 0, L = 0, scoring = 0, weighting = 0, n_nodes = 0, n_words = 0, max_children = 0;
    int n_slots = 0;
    BowVocDev dev = {};
    void *d_blob = nullptr; /* all five slot arrays */
    /* scratch of the stand-alone transform (borb_voc_transform), grown on demand */
    cudaStream_t stream = nullptr;
    int cap = 0;
    uint8_t *d_desc = nullptr, *h_desc = nullptr;
    int *d_word = nullptr, *d_node = nullptr;
    double *d_weight = nullptr;
    uint8_t *d_out = nullptr, *h_out = nullptr; /* result slab, layout bow_slab_layout */
    int *d_bstart = nullptr;
    /* scratch of the keyframe-to-keyframe matcher (borb_voc_search_by_bow), grown on demand: staging blob of both
     * sides (layout in the function), candidate cache, choices, results */
    size_t kk_c1 = 0, kk_c2 = 0, kk_blob = 0;
    uint8_t *d_kk = nullptr, *h_kk = nullptr, *d_kk_more = nullptr, *d_kk_bin = nullptr;
    uint32_t *d_kk_cache = nullptr;
    int *d_kk_choice = nullptr, *d_kk_match = nullptr, *h_kk_match = nullptr, *d_kk_hist = nullptr;
};

struct BowSlab { /* byte offsets of one frame's results inside a slab, for a capacity of K features */
    size_t word, val, node, start, feat, counts, bytes;
};
static BowSlab bow_slab_layout(size_t K)
{
    BowSlab s;
    size_t b = 0;
    s.val = b; b += K * 8;
    s.word = b; b += K * 4;
    s.node = b; b += K * 4;
    s.start = b; b += (K + 1) * 4;
    s.feat = b; b += K * 4;
    b = (b + 15) & ~(size_t)15;
    s.counts = b; b += 16;
    s.bytes = b;
    return s;
}

struct BowState { /* per extractor handle, allocated by the first borb_compute_bow */
    size_t K = 0;
    int *d_word = nullptr, *d_node = nullptr, *d_bstart = nullptr;
    double *d_weight = nullptr;
    uint8_t *d_out = nullptr, *h_out = nullptr;
    int frame = -1;
    unsigned long long gen = 0; /* extract_gen the results belong to */
    bool on_host = false;
    /* SearchByBoW: keyframe side staging [desc 32c][angle 4c][node 4c][start 4(c+1)][feat 4c][mp c], results */
    size_t kf_cap = 0;
    uint8_t *d_kf = nullptr, *h_kf = nullptr;
    int *d_match = nullptr, *h_match = nullptr; /* [K + 4]: matches, then nmatches */
    uint32_t *d_cache = nullptr; /* [kf_cap][BOW_K] candidate keys */
    int *d_choice = nullptr;     /* [kf_cap] accepted list position per keyframe feature */
    uint8_t *d_more = nullptr;
    uint8_t *d_bin = nullptr;
    int *d_hist = nullptr;
};

static void bow_release(BowState *&b)
{
    if (!b) return;
    void *dev[] = {b->d_word, b->d_node, b->d_bstart, b->d_weight, b->d_out, b->d_kf, b->d_match, b->d_bin, b->d_hist, b->d_cache, b->d_more, b->d_choice};
    for (void *p : dev) if (p) cudaFree(p);
    void *host[] = {b->h_out, b->h_kf, b->h_match};
    for (void *p : host) if (p) cudaFreeHost(p);
    delete b;
    b = nullptr;
}

namespace {

unsigned bow_flags(int scoring, int weighting)
{
    /* TF_IDF (0), TF (1): BowVector::addWeight; IDF (2), BINARY (3): addIfNotExist (TemplatedVocabulary.h:1144-1197).
     * L1 norm for L1 / chi-square / KL / Bhattacharyya scoring, L2 for L2, none for the dot product
     * (ScoringObject.h:74-89); without a norm the accumulated values are divided by the number of words (:1160-1166). */
    const bool accumulate = weighting == 0 || weighting == 1;
    const unsigned norm = scoring == 5 ? 0u : scoring == 1 ? 2u : 1u;
    return (accumulate ? 1u : 0u) | (norm << 1) | ((accumulate && norm == 0) ? 8u : 0u);
}

int bow_np2(int n)
{
    int p = BOW_ASM_NT;
    while (p < n) p <<= 1;
    return p;
}

/* B1 + B2 on `st`: descriptors desc (device, 32 B each), count from n_p (device) or n_val */
int bow_enqueue(const borb_vocabulary *voc, const uint8_t *d_desc, const int *n_p, int n_val, int cap, int levelsup,
                int *d_word, int *d_node, double *d_weight, int *d_bstart, uint8_t *d_out, cudaStream_t st)
{
    const BowSlab S = bow_slab_layout((size_t)cap);
    const int G = voc->max_children <= 8 ? 8 : voc->max_children <= 16 ? 16 : 32;
    const int per_cta = BOW_NT / G, grid = std::max(1, (cap + per_cta - 1) / per_cta);
#define B1(G_) k_bow_descend<G_><<<grid, BOW_NT, 0, st>>>(voc->dev, reinterpret_cast<const uint4 *>(d_desc), n_p, n_val, cap, \
                                                      levelsup, d_word, d_node, d_weight)
    if (G == 8) B1(8); else if (G == 16) B1(16); else B1(32);
#undef B1
    const int dual = bow_np2(cap) <= 8192 ? 1 : 0; /* both key arrays in shared memory: one sorting pass */
    const size_t smem = (size_t)bow_np2(cap) * 8 * (size_t)(1 + dual);
    if (smem > 48 * 1024) CU(cudaFuncSetAttribute(k_bow_assemble, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_bow_assemble<<<1, BOW_ASM_NT, smem, st>>>(d_word, d_node, d_weight, n_p, n_val, cap, bow_flags(voc->scoring, voc->weighting),
                                               reinterpret_cast<uint32_t *>(d_out + S.word), reinterpret_cast<double *>(d_out + S.val),
                                               d_bstart, reinterpret_cast<uint32_t *>(d_out + S.node),
                                               reinterpret_cast<int *>(d_out + S.start), reinterpret_cast<uint32_t *>(d_out + S.feat),
                                               reinterpret_cast<int *>(d_out + S.counts), dual);
    CU(cudaGetLastError());
    return BORB_OK;
}

/* scratch of the keyframe-to-keyframe matchers on the vocabulary handle: a staging blob of `blob` bytes on both sides of
 * the link, per-pKF1 arrays for c1 keypoints, per-keypoint result arrays for max(c1, c2) */
int kk_reserve(borb_vocabulary *v, size_t c1, size_t c2, size_t blob)
{
    if (c1 <= v->kk_c1 && c2 <= v->kk_c2 && blob <= v->kk_blob && v->d_kk) return BORB_OK;
    c1 = std::max(c1, v->kk_c1); c2 = std::max(c2, v->kk_c2); blob = std::max(blob, v->kk_blob);
    CU(cudaStreamSynchronize(v->stream));
    void *dev[] = {v->d_kk, v->d_kk_more, v->d_kk_bin, v->d_kk_cache, v->d_kk_choice, v->d_kk_match, v->d_kk_hist};
    for (void *p : dev) if (p) cudaFree(p);
    if (v->h_kk) cudaFreeHost(v->h_kk);
    if (v->h_kk_match) cudaFreeHost(v->h_kk_match);
    v->d_kk = v->h_kk = v->d_kk_more = v->d_kk_bin = nullptr; v->d_kk_cache = nullptr;
    v->d_kk_choice = v->d_kk_match = v->h_kk_match = v->d_kk_hist = nullptr; v->kk_c1 = v->kk_c2 = v->kk_blob = 0;
    const size_t cm = std::max(c1, c2);
    CU(cudaMalloc(&v->d_kk, blob));
    CU(cudaMallocHost(&v->h_kk, blob));
    CU(cudaMalloc(&v->d_kk_more, c1));
    CU(cudaMalloc(&v->d_kk_bin, cm));
    CU(cudaMalloc(&v->d_kk_cache, c1 * BOW_K * 4));
    CU(cudaMalloc(&v->d_kk_choice, c1 * 4));
    CU(cudaMalloc(&v->d_kk_match, (cm + 4) * 4));
    CU(cudaMallocHost(&v->h_kk_match, (cm + 4) * 4));
    CU(cudaMalloc(&v->d_kk_hist, MT_HISTO * 4));
    v->kk_c1 = c1; v->kk_c2 = c2; v->kk_blob = blob;
    return BORB_OK;
}

bool bow_next_token(char *&p, char *&tok)
{
    while (*p == ' ' || *p == '\t' || *p == '\r' || *p == '\n') ++p;
    if (!*p) return false;
    tok = p;
    while (*p && *p != ' ' && *p != '\t' && *p != '\r' && *p != '\n') ++p;
    if (*p) *p++ = 0;
    return true;
}

} // namespace

extern "C" {

/* nodes 1 .. n_nodes (node 0 is the root): parent id (< the node's own id), leaf flag (> 0: the node is a word;
 * word ids count the flagged nodes in order), 32-byte descriptor, weight */
int borb_voc_create(int k, int L, int scoring, int weighting, int n_nodes, const int32_t *parent, const uint8_t *is_leaf,
                    const uint8_t *desc, const double *weight, int device, borb_vocabulary **out)
{
    if (!out) return fail(BORB_ERR_ARG, "out is null");
    *out = nullptr;
    /* the reference's own check (TemplatedVocabulary.h:1359) */
    if (k < 0 || k > 20 || L < 1 || L > 10 || scoring < 0 || scoring > 5 || weighting < 0 || weighting > 3)
        return fail(BORB_ERR_ARG, "not a vocabulary header: k %d, L %d, scoring %d, weighting %d", k, L, scoring, weighting);
    if (n_nodes < 0 || (n_nodes > 0 && (!parent || !is_leaf || !desc || !weight))) return fail(BORB_ERR_ARG, "node arrays");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(BORB_ERR_CUDA, "no CUDA device: this library has no CPU path");
    if (device < 0) CU(cudaGetDevice(&device));
    if (device >= ndev) return fail(BORB_ERR_ARG, "device %d of %d", device, ndev);
    const int N = n_nodes + 1;
    for (int i = 1; i < N; ++i)
        if (parent[i - 1] < 0 || parent[i - 1] >= i)
            return fail(BORB_ERR_ARG, "node %d names parent %d (the reference indexes m_nodes[parent] while it reads the file)", i, parent[i - 1]);
    std::vector<int> cstart((size_t)N + 1, 0), child((size_t)std::max(n_nodes, 1)), word_id((size_t)N, 0);
    for (int i = 1; i < N; ++i) ++cstart[(size_t)parent[i - 1] + 1];
    for (int i = 0; i < N; ++i) cstart[(size_t)i + 1] += cstart[i];
    {
        std::vector<int> fill(cstart.begin(), cstart.end() - 1);
        for (int i = 1; i < N; ++i) child[(size_t)fill[parent[i - 1]]++] = i; /* children in file order */
    }
    int n_words = 0, max_children = 0;
    for (int i = 1; i < N; ++i) if (is_leaf[i - 1] > 0) word_id[i] = n_words++;
    for (int i = 0; i < N; ++i) max_children = std::max(max_children, cstart[(size_t)i + 1] - cstart[i]);
    if (max_children > 65535) return fail(BORB_ERR_ARG, "a node has %d children", max_children);

    borb_vocabulary *v = new borb_vocabulary();
    v->device = device; v->k = k; v->L = L; v->scoring = scoring; v->weighting = weighting;
    v->n_nodes = N; v->n_words = n_words; v->max_children = max_children; v->n_slots = n_nodes;
    const size_t S = (size_t)std::max(n_nodes, 1);
    /* blob: [cweight 8S][cdesc 32S][cinfo 8S][cnode 4S][cword 4S], every section on a 16-byte boundary */
    const size_t o_desc = (S * 8 + 15) & ~(size_t)15, o_info = o_desc + S * 32, o_node = (o_info + S * 8 + 15) & ~(size_t)15,
                 o_word = (o_node + S * 4 + 15) & ~(size_t)15, bytes = o_word + S * 4;
    std::vector<uint8_t> blob(bytes, 0);
    double *cw = reinterpret_cast<double *>(blob.data());
    uint8_t *cd = blob.data() + o_desc;
    int *ci = reinterpret_cast<int *>(blob.data() + o_info), *cn = reinterpret_cast<int *>(blob.data() + o_node),
        *cwd = reinterpret_cast<int *>(blob.data() + o_word);
    for (int s = 0; s < n_nodes; ++s) {
        const int id = child[s];
        cw[s] = weight[id - 1];
        std::memcpy(cd + (size_t)s * 32, desc + (size_t)(id - 1) * 32, 32);
        ci[2 * s] = cstart[id];
        ci[2 * s + 1] = cstart[(size_t)id + 1] - cstart[id];
        cn[s] = id;
        cwd[s] = word_id[id];
    }
    int prev = 0;
    cudaGetDevice(&prev);
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaMalloc(&v->d_blob, bytes);
    if (e == cudaSuccess) e = cudaMemcpy(v->d_blob, blob.data(), bytes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&v->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        if (v->d_blob) cudaFree(v->d_blob);
        delete v;
        cudaSetDevice(prev);
        return fail(BORB_ERR_CUDA, "vocabulary upload: %s", cudaGetErrorString(e));
    }
    uint8_t *b = static_cast<uint8_t *>(v->d_blob);
    v->dev.cweight = reinterpret_cast<const double *>(b);
    v->dev.cdesc = reinterpret_cast<const uint4 *>(b + o_desc);
    v->dev.cinfo = reinterpret_cast<const int2 *>(b + o_info);
    v->dev.cnode = reinterpret_cast<const int *>(b + o_node);
    v->dev.cword = reinterpret_cast<const int *>(b + o_word);
    v->dev.root_start = cstart[0];
    v->dev.root_count = cstart[1] - cstart[0];
    v->dev.L = L;
    *out = v;
    return BORB_OK;
}

/* TemplatedVocabulary::loadFromTextFile (:1338-1421): "k L scoring weighting", then one node per line:
 * parent isLeaf d0 .. d31 weight.  A blank line (the reference would parse it into a node with an indeterminate
 * descriptor) is skipped; a descriptor token that is not a number leaves a 0 byte. */
int borb_voc_load_text(const char *path, int device, borb_vocabulary **out)
{
    if (!path || !out) return fail(BORB_ERR_ARG, "null argument");
    *out = nullptr;
    FILE *f = std::fopen(path, "r");
    if (!f) return fail(BORB_ERR_ARG, "cannot open %s", path);
    char *line = nullptr;
    size_t cap = 0;
    int h[4] = {-1, -1, -1, -1};
    if (getline(&line, &cap, f) >= 0) {
        char *p = line, *tok;
        for (int i = 0; i < 4 && bow_next_token(p, tok); ++i) h[i] = (int)std::strtol(tok, nullptr, 10);
    }
    std::vector<int32_t> parent;
    std::vector<uint8_t> leaf, desc;
    std::vector<double> weight;
    while (getline(&line, &cap, f) >= 0) {
        char *p = line, *tok;
        if (!bow_next_token(p, tok)) continue;
        parent.push_back((int32_t)std::strtol(tok, nullptr, 10));
        leaf.push_back(bow_next_token(p, tok) && std::strtol(tok, nullptr, 10) > 0 ? 1 : 0);
        uint8_t d[32] = {0};
        for (int i = 0; i < 32; ++i) {
            if (!bow_next_token(p, tok)) break;
            char *end = tok;
            const long n = std::strtol(tok, &end, 10);
            if (end != tok) d[i] = (uint8_t)n;
        }
        desc.insert(desc.end(), d, d + 32);
        weight.push_back(bow_next_token(p, tok) ? std::strtod(tok, nullptr) : 0.0);
    }
    std::free(line);
    std::fclose(f);
    return borb_voc_create(h[0], h[1], h[2], h[3], (int)parent.size(), parent.data(), leaf.data(), desc.data(), weight.data(),
                           device, out);
}

void borb_voc_destroy(borb_vocabulary *v)
{
    if (!v) return;
    cudaSetDevice(v->device);
    if (v->stream) cudaStreamSynchronize(v->stream);
    void *dev[] = {v->d_blob, v->d_desc, v->d_word, v->d_node, v->d_weight, v->d_out, v->d_bstart, v->d_kk, v->d_kk_more,
                   v->d_kk_bin, v->d_kk_cache, v->d_kk_choice, v->d_kk_match, v->d_kk_hist};
    for (void *p : dev) if (p) cudaFree(p);
    if (v->h_desc) cudaFreeHost(v->h_desc);
    if (v->h_out) cudaFreeHost(v->h_out);
    if (v->h_kk) cudaFreeHost(v->h_kk);
    if (v->h_kk_match) cudaFreeHost(v->h_kk_match);
    if (v->stream) cudaStreamDestroy(v->stream);
    delete v;
}

int borb_voc_info(const borb_vocabulary *v, int *k, int *L, int *n_nodes, int *n_words, int *scoring, int *weighting)
{
    if (!v) return fail(BORB_ERR_ARG, "null vocabulary");
    if (k) *k = v->k;
    if (L) *L = v->L;
    if (n_nodes) *n_nodes = v->n_nodes;
    if (n_words) *n_words = v->n_words;
    if (scoring) *scoring = v->scoring;
    if (weighting) *weighting = v->weighting;
    return BORB_OK;
}

/* transform(features, v, fv, levelsup) for descriptors in HOST memory (KeyFrame::ComputeBoW on a keyframe whose
 * descriptors no handle holds).  Capacities: word_id / word_val / node_id / feat_idx [n], node_start [n + 1]. */
int borb_voc_transform(borb_vocabulary *v, const uint8_t *desc, int n, int levelsup, uint32_t *word_id, double *word_val,
                       int *n_words, uint32_t *node_id, int32_t *node_start, uint32_t *feat_idx, int *n_nodes)
{
    if (!v || n < 0 || (n > 0 && !desc) || !n_words || !n_nodes || !node_start) return fail(BORB_ERR_ARG, "null argument");
    if (n > 0 && (!word_id || !word_val || !node_id || !feat_idx)) return fail(BORB_ERR_ARG, "null output array");
    if (n > BOW_MAX_FEATURES) return fail(BORB_ERR_CAPACITY, "%d features; the bag-of-words step sorts up to %d per frame", n, BOW_MAX_FEATURES);
    *n_words = 0; *n_nodes = 0; node_start[0] = 0;
    if (n == 0 || v->n_words == 0) return BORB_OK; /* empty(): the reference returns empty vectors (:1136) */
    CU(cudaSetDevice(v->device));
    if (n > v->cap) {
        const int cap = bow_np2(n);
        CU(cudaStreamSynchronize(v->stream));
        void *dev[] = {v->d_desc, v->d_word, v->d_node, v->d_weight, v->d_out, v->d_bstart};
        for (void *p : dev) if (p) cudaFree(p);
        if (v->h_desc) cudaFreeHost(v->h_desc);
        if (v->h_out) cudaFreeHost(v->h_out);
        v->d_desc = v->h_desc = v->d_out = v->h_out = nullptr; v->d_word = v->d_node = v->d_bstart = nullptr; v->d_weight = nullptr;
        v->cap = 0;
        const BowSlab S = bow_slab_layout((size_t)cap);
        CU(cudaMalloc(&v->d_desc, (size_t)cap * 32));
        CU(cudaMallocHost(&v->h_desc, (size_t)cap * 32));
        CU(cudaMalloc(&v->d_word, (size_t)cap * 4));
        CU(cudaMalloc(&v->d_node, (size_t)cap * 4));
        CU(cudaMalloc(&v->d_weight, (size_t)cap * 8));
        CU(cudaMalloc(&v->d_bstart, ((size_t)cap + 1) * 4));
        CU(cudaMalloc(&v->d_out, S.bytes));
        CU(cudaMallocHost(&v->h_out, S.bytes));
        v->cap = cap;
    }
    const BowSlab S = bow_slab_layout((size_t)v->cap);
    std::memcpy(v->h_desc, desc, (size_t)n * 32);
    CU(cudaMemcpyAsync(v->d_desc, v->h_desc, (size_t)n * 32, cudaMemcpyHostToDevice, v->stream));
    int rc = bow_enqueue(v, v->d_desc, nullptr, n, v->cap, levelsup, v->d_word, v->d_node, v->d_weight, v->d_bstart, v->d_out, v->stream);
    if (rc != BORB_OK) return rc;
    CU(cudaMemcpyAsync(v->h_out, v->d_out, S.bytes, cudaMemcpyDeviceToHost, v->stream));
    CU(cudaStreamSynchronize(v->stream));
    const int *cnt = reinterpret_cast<const int *>(v->h_out + S.counts);
    const int nw = cnt[0], nn = cnt[1];
    std::memcpy(word_id, v->h_out + S.word, (size_t)nw * 4);
    std::memcpy(word_val, v->h_out + S.val, (size_t)nw * 8);
    std::memcpy(node_id, v->h_out + S.node, (size_t)nn * 4);
    std::memcpy(node_start, v->h_out + S.start, ((size_t)nn + 1) * 4);
    std::memcpy(feat_idx, v->h_out + S.feat, (size_t)node_start[nn] * 4);
    *n_words = nw; *n_nodes = nn;
    return BORB_OK;
}

/* Frame::ComputeBoW (src/Frame.cc:532-539; the reference passes levelsup = 4) on frame `frame` of h's last call */
int borb_compute_bow(borb_extractor *h, int frame, const borb_vocabulary *voc, int levelsup)
{
    if (!h || !voc) return fail(BORB_ERR_ARG, "null argument");
    if (!h->have_plan || h->last_n < 1) return fail(BORB_ERR_STATE, "no extraction yet");
    if (h->pending) return fail(BORB_ERR_STATE, "a submitted call has not been waited for");
    if (frame < 0 || frame >= h->last_n) return fail(BORB_ERR_ARG, "frame %d of %d", frame, h->last_n);
    if (voc->device != h->device) return fail(BORB_ERR_ARG, "vocabulary on device %d, extractor on %d", voc->device, h->device);
    const size_t K = (size_t)h->plan.kp_total;
    if (K > BOW_MAX_FEATURES) return fail(BORB_ERR_CAPACITY, "%zu keypoint slots per frame; the bag-of-words step sorts up to %d", K, BOW_MAX_FEATURES);
    CU(cudaSetDevice(h->device));
    if (h->bow && h->bow->K != K) { CU(cudaStreamSynchronize(h->stream)); bow_release(h->bow); }
    if (!h->bow) {
        BowState *b = new BowState();
        h->bow = b;
        b->K = K;
        const BowSlab S = bow_slab_layout(K);
        CU(cudaMalloc(&b->d_word, K * 4));
        CU(cudaMalloc(&b->d_node, K * 4));
        CU(cudaMalloc(&b->d_weight, K * 8));
        CU(cudaMalloc(&b->d_bstart, (K + 1) * 4));
        CU(cudaMalloc(&b->d_out, S.bytes));
        CU(cudaMallocHost(&b->h_out, S.bytes));
        CU(cudaMemset(b->d_out, 0, S.bytes));
        std::memset(b->h_out, 0, S.bytes);
    }
    BowState *b = h->bow;
    const BowSlab S = bow_slab_layout(K);
    b->on_host = false;
    b->frame = -1;
    if (voc->n_words == 0) { /* empty(): empty vectors */
        CU(cudaMemsetAsync(b->d_out + S.counts, 0, 16, h->stream));
        CU(cudaMemsetAsync(b->d_out + S.start, 0, 4, h->stream));
    } else {
        int rc = bow_enqueue(voc, h->d_out_desc + (size_t)frame * K * 32, h->d_nout + frame, 0, (int)K, levelsup, b->d_word,
                             b->d_node, b->d_weight, b->d_bstart, b->d_out, h->stream);
        if (rc != BORB_OK) return rc;
    }
    CU(cudaMemcpyAsync(b->h_out, b->d_out, S.bytes, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    b->frame = frame;
    b->gen = h->extract_gen;
    b->on_host = true;
    return BORB_OK;
}

/* mBowVec as (word id ascending, value) and mFeatVec in CSR form (node id ascending; the features of node u are
 * feat_idx[node_start[u] .. node_start[u+1]), ascending).  Pointers into page-locked memory owned by the handle,
 * valid until its next borb_compute_bow. */
int borb_bow_result(borb_extractor *h, int frame, const uint32_t **word_id, const double **word_val, int *n_words,
                    const uint32_t **node_id, const int32_t **node_start, const uint32_t **feat_idx, int *n_nodes)
{
    if (!h) return fail(BORB_ERR_ARG, "null handle");
    BowState *b = h->bow;
    if (!b || !b->on_host || b->frame != frame || b->gen != h->extract_gen)
        return fail(BORB_ERR_STATE, "borb_compute_bow has not run for frame %d of the current extraction", frame);
    const BowSlab S = bow_slab_layout(b->K);
    const int *cnt = reinterpret_cast<const int *>(b->h_out + S.counts);
    if (word_id) *word_id = reinterpret_cast<const uint32_t *>(b->h_out + S.word);
    if (word_val) *word_val = reinterpret_cast<const double *>(b->h_out + S.val);
    if (n_words) *n_words = cnt[0];
    if (node_id) *node_id = reinterpret_cast<const uint32_t *>(b->h_out + S.node);
    if (node_start) *node_start = reinterpret_cast<const int32_t *>(b->h_out + S.start);
    if (feat_idx) *feat_idx = reinterpret_cast<const uint32_t *>(b->h_out + S.feat);
    if (n_nodes) *n_nodes = cnt[1];
    return BORB_OK;
}

/* ORBmatcher::SearchByBoW(pKF, F, vpMapPointMatches) (src/ORBmatcher.cc:195-351): F = frame `frame` of cur's last
 * call, whose bow vectors borb_compute_bow left on the device.  match [borb_max_keypoints]: per current keypoint the
 * index of the keyframe keypoint whose map point it takes, or -1. */
int borb_search_by_bow(borb_extractor *cur, int frame, const borb_keyframe *kf, float nnratio, unsigned flags,
                       int32_t *match, int *n_cur, int *nmatches)
{
    if (!cur || !kf || !match) return fail(BORB_ERR_ARG, "null argument");
    BowState *b = cur->bow;
    if (!b || b->frame != frame || b->gen != cur->extract_gen)
        return fail(BORB_ERR_STATE, "borb_compute_bow has not run for frame %d of the current extraction", frame);
    if (cur->pending) return fail(BORB_ERR_STATE, "a submitted call has not been waited for");
    const int n = kf->n, nn = kf->n_nodes;
    if (n < 0 || nn < 0 || (n > 0 && (!kf->desc || !kf->angle || !kf->mp_flags)) || (nn > 0 && (!kf->node_id || !kf->node_start || !kf->feat_idx)))
        return fail(BORB_ERR_ARG, "keyframe arrays");
    const int nfeat = nn > 0 ? kf->node_start[nn] : 0;
    if (nfeat < 0 || nfeat > n) return fail(BORB_ERR_ARG, "keyframe feature vector lists %d features of %d", nfeat, n);
    for (int i = 0; i < nfeat; ++i)
        if (kf->feat_idx[i] >= (uint32_t)n) return fail(BORB_ERR_ARG, "keyframe feature index %u of %d", kf->feat_idx[i], n);
    CU(cudaSetDevice(cur->device));
    const size_t K = b->K;
    cudaStream_t st = cur->stream;
    if (!b->on_host) return fail(BORB_ERR_STATE, "the frame's bow vectors are not on the host");
    if ((size_t)n > b->kf_cap || !b->d_kf) {
        const size_t c = (((size_t)std::max(n, 1) + 1023) / 1024) * 1024;
        CU(cudaStreamSynchronize(st));
        if (b->d_kf) cudaFree(b->d_kf);
        if (b->h_kf) cudaFreeHost(b->h_kf);
        if (b->d_cache) cudaFree(b->d_cache);
        if (b->d_more) cudaFree(b->d_more);
        if (b->d_choice) cudaFree(b->d_choice);
        b->d_kf = b->h_kf = nullptr; b->d_cache = nullptr; b->d_more = nullptr; b->d_choice = nullptr; b->kf_cap = 0;
        CU(cudaMalloc(&b->d_choice, c * 4));
        CU(cudaMalloc(&b->d_cache, c * BOW_K * 4));
        CU(cudaMalloc(&b->d_more, c));
        CU(cudaMalloc(&b->d_kf, c * 57 + 32));
        CU(cudaMallocHost(&b->h_kf, c * 57 + 32));
        b->kf_cap = c;
    }
    if (!b->d_match) {
        CU(cudaMalloc(&b->d_match, (K + 4) * 4));
        CU(cudaMallocHost(&b->h_match, (K + 4) * 4));
        CU(cudaMalloc(&b->d_bin, K));
        CU(cudaMalloc(&b->d_hist, MT_HISTO * 4));
    }
    const size_t c = b->kf_cap, un = (size_t)n;
    /* [desc 32c][angle 4c][feat 4c][pos_node 4c][start 4(c+1) (+4)][node_f 8c][mp c] */
    const size_t o_ang = 32 * c, o_feat = 36 * c, o_pn = 40 * c, o_start = 44 * c, o_nf = 48 * c + 8, o_mp = 56 * c + 8;
    uint8_t *hp = b->h_kf, *dp = b->d_kf;
    if (n) {
        std::memcpy(hp, kf->desc, un * 32);
        std::memcpy(hp + o_ang, kf->angle, un * 4);
        std::memcpy(hp + o_mp, kf->mp_flags, un);
    }
    int max_nf = 0;
    {   /* both feature vectors are sorted by node id: walk them in step, as the reference does (:219-338) */
        const BowSlab S = bow_slab_layout(K);
        const int *cnt = reinterpret_cast<const int *>(b->h_out + S.counts);
        const uint32_t *fnode = reinterpret_cast<const uint32_t *>(b->h_out + S.node);
        const int *fstart = reinterpret_cast<const int *>(b->h_out + S.start);
        int *pn = reinterpret_cast<int *>(hp + o_pn), *nf = reinterpret_cast<int *>(hp + o_nf);
        int fb = 0;
        for (int a = 0; a < nn; ++a) {
            if (kf->node_start[a] < 0 || kf->node_start[a + 1] < kf->node_start[a] || kf->node_start[a + 1] > nfeat)
                return fail(BORB_ERR_ARG, "keyframe feature vector: node_start is not ascending");
            while (fb < cnt[1] && fnode[fb] < kf->node_id[a]) ++fb;
            const bool both = fb < cnt[1] && fnode[fb] == kf->node_id[a];
            nf[2 * a] = both ? fstart[fb] : 0;
            nf[2 * a + 1] = both ? fstart[fb + 1] : 0;
            max_nf = std::max(max_nf, nf[2 * a + 1] - nf[2 * a]);
            for (int p = kf->node_start[a]; p < kf->node_start[a + 1]; ++p) pn[p] = a;
        }
    }
    if (nn) {
        std::memcpy(hp + o_start, kf->node_start, ((size_t)nn + 1) * 4);
        std::memcpy(hp + o_feat, kf->feat_idx, (size_t)nfeat * 4);
    }
    CU(cudaMemcpyAsync(dp, hp, c * 57 + 32, cudaMemcpyHostToDevice, st));
    CU(cudaMemsetAsync(b->d_match, 0xff, K * 4, st));
    CU(cudaMemsetAsync(b->d_match + K, 0, 16, st));
    CU(cudaMemsetAsync(b->d_hist, 0, MT_HISTO * 4, st));
    const BowSlab S = bow_slab_layout(K);
    const size_t fo = (size_t)frame * K;
    const int check_ori = (flags & BORB_MATCH_NO_ORIENTATION) ? 0 : 1;
    if (nn > 0 && nfeat > 0) {
        const uint4 *f_desc = reinterpret_cast<const uint4 *>(cur->d_out_desc + fo * 32);
        const uint32_t *f_feat = reinterpret_cast<const uint32_t *>(b->d_out + S.feat);
        const uint32_t *k_feat = reinterpret_cast<const uint32_t *>(dp + o_feat);
        const int2 *node_f = reinterpret_cast<const int2 *>(dp + o_nf);
        k_bow_candidates<<<(nfeat + BOW_CAND_NT / 32 - 1) / (BOW_CAND_NT / 32), BOW_CAND_NT, 0, st>>>(
            reinterpret_cast<const uint4 *>(dp), dp + o_mp, k_feat, reinterpret_cast<const int *>(dp + o_pn), node_f, nfeat, f_desc,
            f_feat, b->d_cache, b->d_more);
        const int tw = (max_nf + 31) / 32 + 1;
        const size_t smem = (size_t)(BOW_RES_NT / 32) * ((size_t)tw + 32 * (BOW_K + 1)) * 4;
        k_bow_resolve<<<(nn + BOW_RES_NT / 32 - 1) / (BOW_RES_NT / 32), BOW_RES_NT, smem, st>>>(
            reinterpret_cast<const uint4 *>(dp), reinterpret_cast<const float *>(dp + o_ang), k_feat,
            reinterpret_cast<const int *>(dp + o_start), node_f, nn, f_desc, cur->d_out_kp + fo, f_feat, b->d_cache, b->d_more,
            nnratio, check_ori, tw, b->d_choice, b->d_match, b->d_bin, b->d_hist, b->d_match + K + 1, BOW_TH_LOW, nullptr);
    }
    k_bow_finish<<<1, 1024, 0, st>>>(cur->d_nout + frame, (int)K, check_ori, b->d_hist, b->d_bin, b->d_match, b->d_match + K);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(b->h_match, b->d_match, (K + 4) * 4, cudaMemcpyDeviceToHost, st));
    if (!cur->results_on_host && cur->nout_frame_on_host != frame)
        CU(cudaMemcpyAsync(cur->h_nout + frame, cur->d_nout + frame, 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    cur->nout_frame_on_host = frame;
    const int nc = std::min(cur->h_nout[frame], (int)K);
    std::memcpy(match, b->h_match, (size_t)nc * 4);
    if (n_cur) *n_cur = nc;
    if (nmatches) *nmatches = b->h_match[K];
    cur->mt_rounds = 1;
    cur->mt_rescans = b->h_match[K + 1]; /* borb_match_rounds: features that scanned their node again */
    return BORB_OK;
}

/* ORBmatcher::SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, vector<MapPoint*> &vpMatches12) (src/ORBmatcher.cc:634-777), the
 * matcher of LoopClosing::ComputeSim3: both sides are keyframes no extractor handle holds, so both arrive as flat arrays
 * and the vocabulary handle lends its device, stream and scratch.  The same kernels as borb_search_by_bow (B3a / B3b / B4,
 * k_bow.cuh) with pKF1 as the query side: the pKF2 features that can never be chosen -- no map point, or a bad one
 * (:686-690) -- are dropped from pKF2's node lists on the host (their order is kept, so the first-of-equals rule is
 * untouched); the distance test is strict here (bestDist1 < TH_LOW, :712), i.e. <= TH_LOW - 1; the kernels' per-pKF2
 * result (each pKF2 keypoint is matched at most once, vbMatched2) is turned into vpMatches12 on the host. */
int borb_voc_search_by_bow(borb_vocabulary *v, const borb_keyframe *kf1, const borb_keyframe *kf2, float nnratio,
                           unsigned flags, int32_t *match12, int *nmatches)
{
    if (!v || !kf1 || !kf2 || !match12) return fail(BORB_ERR_ARG, "null argument");
    const borb_keyframe *kf[2] = {kf1, kf2};
    int nfeat[2] = {0, 0};
    for (int s = 0; s < 2; ++s) {
        const borb_keyframe *k = kf[s];
        if (k->n < 0 || k->n_nodes < 0 || (k->n > 0 && (!k->desc || !k->angle || !k->mp_flags)) ||
            (k->n_nodes > 0 && (!k->node_id || !k->node_start || !k->feat_idx)))
            return fail(BORB_ERR_ARG, "keyframe %d arrays", s + 1);
        if (k->n > 65535) return fail(BORB_ERR_CAPACITY, "keyframe %d has %d keypoints; list positions are packed in 16 bits", s + 1, k->n);
        nfeat[s] = k->n_nodes > 0 ? k->node_start[k->n_nodes] : 0;
        if (nfeat[s] < 0 || nfeat[s] > k->n) return fail(BORB_ERR_ARG, "keyframe %d feature vector lists %d features of %d", s + 1, nfeat[s], k->n);
        for (int a = 0; a < k->n_nodes; ++a)
            if (k->node_start[a] < 0 || k->node_start[a + 1] < k->node_start[a] || k->node_start[a + 1] > nfeat[s])
                return fail(BORB_ERR_ARG, "keyframe %d feature vector: node_start is not ascending", s + 1);
        for (int i = 0; i < nfeat[s]; ++i)
            if (k->feat_idx[i] >= (uint32_t)k->n) return fail(BORB_ERR_ARG, "keyframe %d feature index %u of %d", s + 1, k->feat_idx[i], k->n);
    }
    const int n1 = kf1->n, n2 = kf2->n, nn1 = kf1->n_nodes, nn2 = kf2->n_nodes;
    for (int i = 0; i < n1; ++i) match12[i] = -1;
    if (nmatches) *nmatches = 0;
    if (n1 == 0 || n2 == 0 || nfeat[0] == 0 || nfeat[1] == 0) return BORB_OK;
    CU(cudaSetDevice(v->device));
    cudaStream_t st = v->stream;
    const size_t c1 = (((size_t)n1 + 1023) / 1024) * 1024, c2 = (((size_t)n2 + 1023) / 1024) * 1024;
    {
        const size_t a1 = std::max(c1, v->kk_c1), a2 = std::max(c2, v->kk_c2);
        const int rc = kk_reserve(v, a1, a2, a1 * 57 + 32 + a2 * 40 + 16);
        if (rc != BORB_OK) return rc;
    }
    const size_t A1 = v->kk_c1, A2 = v->kk_c2, K2 = A2;
    /* side 1: [desc 32][angle 4][feat 4][pos_node 4][start 4 (+8)][node_f 8 (+8)][mp 1] per A1; side 2 after it, 16-byte
     * aligned: [desc 32][angle 4][feat 4] per A2, then the keypoint count */
    const size_t o_ang = 32 * A1, o_feat = 36 * A1, o_pn = 40 * A1, o_start = 44 * A1, o_nf = 48 * A1 + 8, o_mp = 56 * A1 + 8;
    const size_t o2 = (57 * A1 + 8 + 15) & ~(size_t)15, o2_ang = o2 + 32 * A2, o2_feat = o2 + 36 * A2, o2_cnt = o2 + 40 * A2;
    uint8_t *hp = v->h_kk, *dp = v->d_kk;
    std::memcpy(hp, kf1->desc, (size_t)n1 * 32);
    std::memcpy(hp + o_ang, kf1->angle, (size_t)n1 * 4);
    std::memcpy(hp + o_mp, kf1->mp_flags, (size_t)n1);
    std::memcpy(hp + o_start, kf1->node_start, ((size_t)nn1 + 1) * 4);
    std::memcpy(hp + o_feat, kf1->feat_idx, (size_t)nfeat[0] * 4);
    std::memcpy(hp + o2, kf2->desc, (size_t)n2 * 32);
    std::memcpy(hp + o2_ang, kf2->angle, (size_t)n2 * 4);
    *reinterpret_cast<int *>(hp + o2_cnt) = n2;
    int max_nf = 0;
    {   /* pKF2's lists without the features that are never candidates; both vectors are sorted by node id: walked in
         * step, as the reference does (:658-745) */
        uint32_t *f2 = reinterpret_cast<uint32_t *>(hp + o2_feat);
        int *pn = reinterpret_cast<int *>(hp + o_pn), *nf = reinterpret_cast<int *>(hp + o_nf);
        int b = 0, w = 0;
        for (int a = 0; a < nn1; ++a) {
            while (b < nn2 && kf2->node_id[b] < kf1->node_id[a]) ++b;
            const bool both = b < nn2 && kf2->node_id[b] == kf1->node_id[a];
            nf[2 * a] = w;
            if (both)
                for (int p = kf2->node_start[b]; p < kf2->node_start[b + 1]; ++p) {
                    const uint32_t i2 = kf2->feat_idx[p];
                    if ((kf2->mp_flags[i2] & 1) && !(kf2->mp_flags[i2] & 2)) f2[w++] = i2;
                }
            nf[2 * a + 1] = w;
            max_nf = std::max(max_nf, nf[2 * a + 1] - nf[2 * a]);
            for (int p = kf1->node_start[a]; p < kf1->node_start[a + 1]; ++p) pn[p] = a;
        }
    }
    CU(cudaMemcpyAsync(dp, hp, o2_cnt + 16, cudaMemcpyHostToDevice, st));
    CU(cudaMemsetAsync(v->d_kk_match, 0xff, K2 * 4, st));
    CU(cudaMemsetAsync(v->d_kk_match + K2, 0, 16, st));
    CU(cudaMemsetAsync(v->d_kk_hist, 0, MT_HISTO * 4, st));
    const int check_ori = (flags & BORB_MATCH_NO_ORIENTATION) ? 0 : 1;
    {
        const uint4 *f_desc = reinterpret_cast<const uint4 *>(dp + o2);
        const uint32_t *f_feat = reinterpret_cast<const uint32_t *>(dp + o2_feat);
        const uint32_t *k_feat = reinterpret_cast<const uint32_t *>(dp + o_feat);
        const int2 *node_f = reinterpret_cast<const int2 *>(dp + o_nf);
        k_bow_candidates<<<(nfeat[0] + BOW_CAND_NT / 32 - 1) / (BOW_CAND_NT / 32), BOW_CAND_NT, 0, st>>>(
            reinterpret_cast<const uint4 *>(dp), dp + o_mp, k_feat, reinterpret_cast<const int *>(dp + o_pn), node_f, nfeat[0], f_desc,
            f_feat, v->d_kk_cache, v->d_kk_more);
        const int tw = (max_nf + 31) / 32 + 1;
        const size_t smem = (size_t)(BOW_RES_NT / 32) * ((size_t)tw + 32 * (BOW_K + 1)) * 4;
        if (smem > 48 * 1024) CU(cudaFuncSetAttribute(k_bow_resolve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_bow_resolve<<<(nn1 + BOW_RES_NT / 32 - 1) / (BOW_RES_NT / 32), BOW_RES_NT, smem, st>>>(
            reinterpret_cast<const uint4 *>(dp), reinterpret_cast<const float *>(dp + o_ang), k_feat,
            reinterpret_cast<const int *>(dp + o_start), node_f, nn1, f_desc, nullptr, f_feat, v->d_kk_cache, v->d_kk_more,
            nnratio, check_ori, tw, v->d_kk_choice, v->d_kk_match, v->d_kk_bin, v->d_kk_hist, v->d_kk_match + K2 + 1,
            BOW_TH_LOW - 1, reinterpret_cast<const float *>(dp + o2_ang));
    }
    k_bow_finish<<<1, 1024, 0, st>>>(reinterpret_cast<const int *>(dp + o2_cnt), (int)K2, check_ori, v->d_kk_hist, v->d_kk_bin,
                                     v->d_kk_match, v->d_kk_match + K2);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(v->h_kk_match, v->d_kk_match, (K2 + 4) * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    for (int i2 = 0; i2 < n2; ++i2) {
        const int i1 = v->h_kk_match[i2];
        if (i1 >= 0 && i1 < n1) match12[i1] = i2;
    }
    if (nmatches) *nmatches = v->h_kk_match[K2];
    return BORB_OK;
}

/* ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) (src/ORBmatcher.cc:779-966), the matcher
 * of LocalMapping::CreateNewMapPoints: one kernel (k_bow_triangulate, a warp per pKF1 keypoint) + B4 for the orientation
 * histogram, on the vocabulary's device, stream and scratch. */
int borb_voc_search_for_triangulation(borb_vocabulary *v, const borb_keyframe *kf1, const borb_keyframe *kf2,
                                      const borb_triangulation *g, unsigned flags, int32_t *match12, int *nmatches)
{
    if (!v || !kf1 || !kf2 || !g || !match12) return fail(BORB_ERR_ARG, "null argument");
    if (!g->F12 || !g->Ow1 || !g->Tcw2 || !g->scale_factors2 || !g->level_sigma2_2) return fail(BORB_ERR_ARG, "geometry arrays");
    if (g->nlevels < 1 || g->nlevels > BORB_MAXL) return fail(BORB_ERR_ARG, "nlevels %d", g->nlevels);
    const borb_keyframe *kf[2] = {kf1, kf2};
    int nfeat[2] = {0, 0};
    for (int s = 0; s < 2; ++s) {
        const borb_keyframe *k = kf[s];
        if (k->n < 0 || k->n_nodes < 0 || (k->n > 0 && (!k->desc || !k->angle || !k->mp_flags)) ||
            (k->n_nodes > 0 && (!k->node_id || !k->node_start || !k->feat_idx)))
            return fail(BORB_ERR_ARG, "keyframe %d arrays", s + 1);
        if (k->n > 65535) return fail(BORB_ERR_CAPACITY, "keyframe %d has %d keypoints; list positions are packed in 16 bits", s + 1, k->n);
        nfeat[s] = k->n_nodes > 0 ? k->node_start[k->n_nodes] : 0;
        if (nfeat[s] < 0 || nfeat[s] > k->n) return fail(BORB_ERR_ARG, "keyframe %d feature vector lists %d features of %d", s + 1, nfeat[s], k->n);
        for (int a = 0; a < k->n_nodes; ++a)
            if (k->node_start[a] < 0 || k->node_start[a + 1] < k->node_start[a] || k->node_start[a + 1] > nfeat[s])
                return fail(BORB_ERR_ARG, "keyframe %d feature vector: node_start is not ascending", s + 1);
        for (int i = 0; i < nfeat[s]; ++i)
            if (k->feat_idx[i] >= (uint32_t)k->n) return fail(BORB_ERR_ARG, "keyframe %d feature index %u of %d", s + 1, k->feat_idx[i], k->n);
    }
    const int n1 = kf1->n, n2 = kf2->n, nn1 = kf1->n_nodes, nn2 = kf2->n_nodes;
    if ((n1 > 0 && !g->xy1) || (n2 > 0 && (!g->xy2 || !g->octave2))) return fail(BORB_ERR_ARG, "keypoint positions");
    for (int i = 0; i < n2; ++i)
        if (g->octave2[i] < 0 || g->octave2[i] >= g->nlevels) return fail(BORB_ERR_ARG, "octave %d of pKF2 keypoint %d", g->octave2[i], i);
    const bool only_stereo = (flags & BORB_MATCH_ONLY_STEREO) != 0;
    for (int i = 0; i < n1; ++i) match12[i] = -1;
    if (nmatches) *nmatches = 0;
    if (n1 == 0 || n2 == 0 || nfeat[0] == 0 || nfeat[1] == 0) return BORB_OK;
    CU(cudaSetDevice(v->device));
    cudaStream_t st = v->stream;
    const size_t c1 = (((size_t)n1 + 1023) / 1024) * 1024, c2 = (((size_t)n2 + 1023) / 1024) * 1024;
    {
        const size_t a1 = std::max(c1, v->kk_c1), a2 = std::max(c2, v->kk_c2);
        const int rc = kk_reserve(v, a1, a2, a1 * 69 + 32 + a2 * 56 + 32);
        if (rc != BORB_OK) return rc;
    }
    const size_t A1 = v->kk_c1, A2 = v->kk_c2;
    /* side 1, per A1: [desc 32][angle 4][feat 4][pos_node 4][node_f 8][xy 8][u_right 4][mp 1], then the keypoint count;
     * side 2, per A2, 16-byte aligned: [desc 32][angle 4][feat 4][xy 8][octave 4][u_right 4] */
    const size_t o_ang = 32 * A1, o_feat = 36 * A1, o_pn = 40 * A1, o_nf = 44 * A1, o_xy = 52 * A1, o_ur = 60 * A1, o_mp = 64 * A1,
                 o_cnt = 65 * A1;
    const size_t o2 = (65 * A1 + 16 + 15) & ~(size_t)15, o2_ang = o2 + 32 * A2, o2_feat = o2 + 36 * A2, o2_xy = o2 + 40 * A2,
                 o2_oct = o2 + 48 * A2, o2_ur = o2 + 52 * A2, o_end = o2 + 56 * A2;
    uint8_t *hp = v->h_kk, *dp = v->d_kk;
    std::memcpy(hp, kf1->desc, (size_t)n1 * 32);
    std::memcpy(hp + o_ang, kf1->angle, (size_t)n1 * 4);
    std::memcpy(hp + o_feat, kf1->feat_idx, (size_t)nfeat[0] * 4);
    std::memcpy(hp + o_xy, g->xy1, (size_t)n1 * 8);
    if (g->u_right1) std::memcpy(hp + o_ur, g->u_right1, (size_t)n1 * 4);
    std::memcpy(hp + o_mp, kf1->mp_flags, (size_t)n1);
    *reinterpret_cast<int *>(hp + o_cnt) = n1;
    std::memcpy(hp + o2, kf2->desc, (size_t)n2 * 32);
    std::memcpy(hp + o2_ang, kf2->angle, (size_t)n2 * 4);
    std::memcpy(hp + o2_xy, g->xy2, (size_t)n2 * 8);
    std::memcpy(hp + o2_oct, g->octave2, (size_t)n2 * 4);
    if (g->u_right2) std::memcpy(hp + o2_ur, g->u_right2, (size_t)n2 * 4);
    {   /* pKF2's lists without what can never be a candidate; the two vectors walked in step (:809-919) */
        uint32_t *f2 = reinterpret_cast<uint32_t *>(hp + o2_feat);
        int *pn = reinterpret_cast<int *>(hp + o_pn), *nf = reinterpret_cast<int *>(hp + o_nf);
        int b = 0, w = 0;
        for (int a = 0; a < nn1; ++a) {
            while (b < nn2 && kf2->node_id[b] < kf1->node_id[a]) ++b;
            const bool both = b < nn2 && kf2->node_id[b] == kf1->node_id[a];
            nf[2 * a] = w;
            if (both)
                for (int p = kf2->node_start[b]; p < kf2->node_start[b + 1]; ++p) {
                    const uint32_t i2 = kf2->feat_idx[p];
                    if (kf2->mp_flags[i2] & 1) continue;
                    if (only_stereo && !(g->u_right2 && g->u_right2[i2] >= 0.f)) continue;
                    f2[w++] = i2;
                }
            nf[2 * a + 1] = w;
            for (int p = kf1->node_start[a]; p < kf1->node_start[a + 1]; ++p) pn[p] = a;
        }
    }
    TriParams tp;
    std::memset(&tp, 0, sizeof tp);
    std::memcpy(tp.F, g->F12, 36);
    {   /* C2 = R2w * Cw + t2w (the small float product), the epipole (:785-793) */
        float C2[3];
        for (int r = 0; r < 3; ++r) {
            float acc = 0.f;
            for (int k = 0; k < 3; ++k) acc += g->Tcw2[4 * r + k] * g->Ow1[k];
            C2[r] = acc + g->Tcw2[4 * r + 3];
        }
        const float invz = 1.0f / C2[2];
        tp.ex = fmaf(g->fx2 * C2[0], invz, g->cx2);
        tp.ey = fmaf(g->fy2 * C2[1], invz, g->cy2);
    }
    tp.only_stereo = only_stereo ? 1 : 0;
    for (int l = 0; l < g->nlevels; ++l) { tp.scale2[l] = g->scale_factors2[l]; tp.sigma2[l] = g->level_sigma2_2[l]; }
    const size_t cm = std::max(A1, A2);
    CU(cudaMemcpyAsync(dp, hp, o_end, cudaMemcpyHostToDevice, st));
    CU(cudaMemsetAsync(v->d_kk_match, 0xff, cm * 4, st));
    CU(cudaMemsetAsync(v->d_kk_match + cm, 0, 16, st));
    CU(cudaMemsetAsync(v->d_kk_hist, 0, MT_HISTO * 4, st));
    const int check_ori = (flags & BORB_MATCH_NO_ORIENTATION) ? 0 : 1;
    k_bow_triangulate<<<(nfeat[0] + BOW_CAND_NT / 32 - 1) / (BOW_CAND_NT / 32), BOW_CAND_NT, 0, st>>>(
        tp, reinterpret_cast<const uint4 *>(dp), dp + o_mp, reinterpret_cast<const float2 *>(dp + o_xy),
        reinterpret_cast<const float *>(dp + o_ang), g->u_right1 ? reinterpret_cast<const float *>(dp + o_ur) : nullptr,
        reinterpret_cast<const uint32_t *>(dp + o_feat), reinterpret_cast<const int *>(dp + o_pn),
        reinterpret_cast<const int2 *>(dp + o_nf), nfeat[0], reinterpret_cast<const uint4 *>(dp + o2),
        reinterpret_cast<const float2 *>(dp + o2_xy), reinterpret_cast<const int *>(dp + o2_oct),
        reinterpret_cast<const float *>(dp + o2_ang), g->u_right2 ? reinterpret_cast<const float *>(dp + o2_ur) : nullptr,
        reinterpret_cast<const uint32_t *>(dp + o2_feat), check_ori, v->d_kk_match, v->d_kk_bin, v->d_kk_hist);
    k_bow_finish<<<1, 1024, 0, st>>>(reinterpret_cast<const int *>(dp + o_cnt), (int)cm, check_ori, v->d_kk_hist, v->d_kk_bin,
                                     v->d_kk_match, v->d_kk_match + cm);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(v->h_kk_match, v->d_kk_match, (cm + 4) * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    std::memcpy(match12, v->h_kk_match, (size_t)n1 * 4);
    if (nmatches) *nmatches = v->h_kk_match[cm];
    return BORB_OK;
}

} // extern "C"


