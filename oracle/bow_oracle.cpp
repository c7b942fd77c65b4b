/*
 * oracle/bow_oracle.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement, on flat arrays, of the reference's bag-of-words step and of the matcher that consumes it:
 *   TemplatedVocabulary::loadFromTextFile   /root/reference/Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1338-1421
 *   TemplatedVocabulary::transform          :1127-1206 (bow vector + feature vector), :1218-1270 (one descriptor)
 *   FORB::distance                          Thirdparty/DBoW2/DBoW2/FORB.cpp:81-101 (256-bit Hamming)
 *   BowVector::addWeight / addIfNotExist / normalize   Thirdparty/DBoW2/DBoW2/BowVector.cpp:33-94
 *   FeatureVector::addFeature               Thirdparty/DBoW2/DBoW2/FeatureVector.cpp
 *   Frame::ComputeBoW                       src/Frame.cc:532-539 (levelsup = 4)
 *   ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&)   src/ORBmatcher.cc:195-351
 *   L1Scoring::score                        Thirdparty/DBoW2/DBoW2/ScoringObject.cpp:21-71
 * Pinned to the reference's own object code (oracle/_ref/librefbow.so, oracle/bow_driver.cpp) by
 * tests/test_oracle_bow_vs_reference.py.  The vocabulary blob ORBvoc.txt is not in the reference tree; vocabularies
 * are synthetic files in its text format (tests/golden/make_vocabulary.py).
 *
 * Where the reference's behaviour is undefined this restatement picks a definite one (the product does the same):
 *   - a blank line at the end of the file is skipped (the reference parses it into a node with an indeterminate
 *     descriptor under the root, :1377-1417);
 *   - a descriptor token that does not parse leaves a 0 byte (FORB::fromString leaves the byte uninitialised);
 *   - when a descent ends above the feature-vector level the node id is 0 (the reference's `nid` is never written).
 */
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <cstdio>
#include <map>
#include <string>
#include <vector>

namespace {

struct Voc {
    int k = 0, L = 0, scoring = 0, weighting = 0;
    std::vector<int> parent, word_id;          /* per node, node 0 = root */
    std::vector<uint8_t> leaf_flag, desc;      /* desc: 32 B per node */
    std::vector<double> weight;
    std::vector<int> child_start, child;       /* children of node i: child[child_start[i] .. child_start[i+1]) in file order */
    int n_words = 0;
};

inline int hamming256(const uint8_t *a, const uint8_t *b)
{
    int d = 0;
    for (int i = 0; i < 32; i += 8) {
        uint64_t x, y;
        std::memcpy(&x, a + i, 8);
        std::memcpy(&y, b + i, 8);
        d += __builtin_popcountll(x ^ y);
    }
    return d;
}

/* :1218-1270 */
void descend(const Voc &v, const uint8_t *f, int levelsup, int *word, double *w, int *nid)
{
    const int nid_level = v.L - levelsup;
    *nid = 0;
    int final_id = 0, level = 0;
    do {
        ++level;
        const int c0 = v.child_start[final_id], c1 = v.child_start[final_id + 1];
        final_id = v.child[c0];
        int best = hamming256(f, &v.desc[(size_t)final_id * 32]);
        for (int c = c0 + 1; c < c1; ++c) {
            const int id = v.child[c];
            const int d = hamming256(f, &v.desc[(size_t)id * 32]);
            if (d < best) { best = d; final_id = id; }
        }
        if (level == nid_level) *nid = final_id;
    } while (v.child_start[final_id + 1] > v.child_start[final_id]);
    *word = v.word_id[final_id];
    *w = v.weight[final_id];
}

} // namespace

extern "C" {

/* Plain C stdio on purpose: this library carries a static libstdc++, and iostreams from a second copy of it crash in
 * a process that also holds the shared one (numpy, torch).  Tokens are parsed as `stream >> int/double` would. */
static bool next_token(char *&p, char *&tok)
{
    while (*p == ' ' || *p == '\t' || *p == '\r' || *p == '\n') ++p;
    if (!*p) return false;
    tok = p;
    while (*p && *p != ' ' && *p != '\t' && *p != '\r' && *p != '\n') ++p;
    if (*p) *p++ = 0;
    return true;
}

void *bo_load_text(const char *path)
{
    FILE *f = std::fopen(path, "r");
    if (!f) return nullptr;
    Voc *v = new Voc();
    char *line = nullptr;
    size_t cap = 0;
    bool ok = getline(&line, &cap, f) >= 0;
    if (ok) {
        char *p = line, *tok;
        int h[4] = {-1, -1, -1, -1};
        for (int i = 0; i < 4 && next_token(p, tok); ++i) h[i] = (int)std::strtol(tok, nullptr, 10);
        v->k = h[0]; v->L = h[1]; v->scoring = h[2]; v->weighting = h[3];
        ok = !(v->k < 0 || v->k > 20 || v->L < 1 || v->L > 10 || h[2] < 0 || h[2] > 5 || h[3] < 0 || h[3] > 3);
    }
    v->parent.push_back(0); v->word_id.push_back(0); v->leaf_flag.push_back(0); v->weight.push_back(0.0);
    v->desc.resize(32, 0);
    while (ok && getline(&line, &cap, f) >= 0) {
        char *p = line, *tok;
        if (!next_token(p, tok)) continue; /* blank line */
        const int pid = (int)std::strtol(tok, nullptr, 10);
        const int is_leaf = next_token(p, tok) ? (int)std::strtol(tok, nullptr, 10) : 0;
        if (pid < 0 || pid >= (int)v->parent.size()) { ok = false; break; } /* the reference indexes m_nodes[pid] unchecked */
        uint8_t d[32] = {0};
        for (int i = 0; i < 32; ++i) {
            if (!next_token(p, tok)) break;
            char *end = tok;
            const long n = std::strtol(tok, &end, 10);
            if (end != tok) d[i] = (uint8_t)n;
        }
        const double w = next_token(p, tok) ? std::strtod(tok, nullptr) : 0.0;
        v->parent.push_back(pid);
        v->leaf_flag.push_back(is_leaf > 0);
        v->desc.insert(v->desc.end(), d, d + 32);
        v->weight.push_back(w);
        v->word_id.push_back(is_leaf > 0 ? v->n_words++ : 0);
    }
    std::free(line);
    std::fclose(f);
    if (!ok) { delete v; return nullptr; }
    const int N = (int)v->parent.size();
    v->child_start.assign(N + 1, 0);
    for (int i = 1; i < N; ++i) ++v->child_start[v->parent[i] + 1];
    for (int i = 0; i < N; ++i) v->child_start[i + 1] += v->child_start[i];
    v->child.resize(N > 1 ? N - 1 : 0);
    std::vector<int> fill(v->child_start.begin(), v->child_start.end() - 1);
    for (int i = 1; i < N; ++i) v->child[fill[v->parent[i]]++] = i; /* ascending node id = file order */
    return v;
}

void bo_free(void *p) { delete static_cast<Voc *>(p); }

void bo_info(void *p, int *k, int *L, int *n_nodes, int *n_words, int *scoring, int *weighting)
{
    const Voc *v = static_cast<Voc *>(p);
    *k = v->k; *L = v->L; *n_nodes = (int)v->parent.size(); *n_words = v->n_words; *scoring = v->scoring; *weighting = v->weighting;
}

/* the parsed tree as flat arrays (what borb_voc_create takes): parent, leaf flag, descriptor, weight of nodes 1 .. n_nodes - 1 */
void bo_arrays(void *p, int32_t *parent, uint8_t *leaf, uint8_t *desc, double *weight)
{
    const Voc *v = static_cast<Voc *>(p);
    const int N = (int)v->parent.size();
    for (int i = 1; i < N; ++i) {
        parent[i - 1] = v->parent[i];
        leaf[i - 1] = v->leaf_flag[i];
        weight[i - 1] = v->weight[i];
    }
    std::memcpy(desc, v->desc.data() + 32, (size_t)(N - 1) * 32);
}

/* per-descriptor result of the descent: word id, node id at L - levelsup, weight */
void bo_descend(void *p, const uint8_t *desc, int n, int levelsup, int32_t *word, int32_t *node, double *weight)
{
    const Voc *v = static_cast<Voc *>(p);
    for (int i = 0; i < n; ++i) {
        if (v->n_words == 0) { word[i] = 0; node[i] = 0; weight[i] = 0; continue; }
        descend(*v, desc + (size_t)i * 32, levelsup, &word[i], &weight[i], &node[i]);
    }
}

/* transform(features, BowVector, FeatureVector, levelsup) :1127-1206.  Capacities: word_* [n], node_id [n],
 * node_start [n + 1], feat_idx [n]. */
void bo_transform(void *p, const uint8_t *desc, int n, int levelsup, uint32_t *word_id, double *word_val, int *n_words,
                  uint32_t *node_id, int32_t *node_start, uint32_t *feat_idx, int *n_nodes)
{
    const Voc *v = static_cast<Voc *>(p);
    std::map<uint32_t, double> bow;
    std::map<uint32_t, std::vector<uint32_t>> fv;
    *n_words = 0; *n_nodes = 0; node_start[0] = 0;
    if (v->n_words == 0) return; /* empty(): :1136 */
    const bool accumulate = v->weighting == 0 || v->weighting == 1; /* TF_IDF, TF: addWeight; IDF, BINARY: addIfNotExist */
    for (int i = 0; i < n; ++i) {
        int word, nid;
        double w;
        descend(*v, desc + (size_t)i * 32, levelsup, &word, &w, &nid);
        if (!(w > 0)) continue; /* stopped word */
        std::map<uint32_t, double>::iterator it = bow.find((uint32_t)word);
        if (it == bow.end()) bow[(uint32_t)word] = w;
        else if (accumulate) it->second += w;
        fv[(uint32_t)nid].push_back((uint32_t)i);
    }
    const bool must = v->scoring != 5; /* DotProductScoring does not normalise */
    const bool l2 = v->scoring == 1;
    if (accumulate && !bow.empty() && !must) {
        const double nd = (double)bow.size();
        for (auto &e : bow) e.second /= nd;
    }
    if (must) { /* BowVector::normalize */
        double norm = 0.0;
        if (!l2) for (auto &e : bow) norm += std::fabs(e.second);
        else {
            /* the reference object (-O3 -march with FMA, contraction on) fuses this into fma(v, v, norm) */
            for (auto &e : bow) norm = std::fma(e.second, e.second, norm);
            norm = std::sqrt(norm);
        }
        if (norm > 0.0) for (auto &e : bow) e.second /= norm;
    }
    int w = 0;
    for (auto &e : bow) { word_id[w] = e.first; word_val[w] = e.second; ++w; }
    *n_words = w;
    int nn = 0, pos = 0;
    for (auto &e : fv) {
        node_id[nn] = e.first;
        node_start[nn] = pos;
        for (uint32_t i : e.second) feat_idx[pos++] = i;
        ++nn;
    }
    node_start[nn] = pos;
    *n_nodes = nn;
}

/* L1Scoring::score (ScoringObject.cpp:21-71): both vectors sorted by word id */
double bo_score_l1(const uint32_t *id1, const double *val1, int n1, const uint32_t *id2, const double *val2, int n2)
{
    double score = 0;
    int a = 0, b = 0;
    while (a < n1 && b < n2) {
        if (id1[a] == id2[b]) {
            score += std::fabs(val1[a] - val2[b]) - std::fabs(val1[a]) - std::fabs(val2[b]);
            ++a; ++b;
        } else if (id1[a] < id2[b]) ++a; /* lower_bound lands on the first id >= the other's; stepping reaches the same place */
        else ++b;
    }
    return -score / 2.0;
}

/* ORBmatcher::SearchByBoW(pKF, F, vpMapPointMatches) :195-351 on the two feature vectors (CSR: node ids ascending).
 * kf_mp[i]: bit 0 = the keyframe keypoint has a map point, bit 1 = it isBad().  out_match[iF] = keyframe keypoint
 * index or -1.  Returns nmatches. */
int bo_search_by_bow(int n_kf, const float *kf_angle, const uint8_t *kf_desc, const uint8_t *kf_mp,
                     const uint32_t *kf_node, const int32_t *kf_start, const uint32_t *kf_feat, int kf_nodes,
                     int n_f, const float *f_angle, const uint8_t *f_desc,
                     const uint32_t *f_node, const int32_t *f_start, const uint32_t *f_feat, int f_nodes,
                     float nnratio, int check_ori, int32_t *out_match)
{
    (void)n_kf;
    const int TH_LOW = 50, HISTO = 30;
    for (int i = 0; i < n_f; ++i) out_match[i] = -1;
    std::vector<int> hist[30];
    const float factor = 1.0f / HISTO;
    int nmatches = 0, a = 0, b = 0;
    while (a < kf_nodes && b < f_nodes) {
        if (kf_node[a] == f_node[b]) {
            for (int ik = kf_start[a]; ik < kf_start[a + 1]; ++ik) {
                const uint32_t rk = kf_feat[ik];
                if (!(kf_mp[rk] & 1) || (kf_mp[rk] & 2)) continue;
                int best1 = 256, best2 = 256, best_idx = -1;
                for (int jf = f_start[b]; jf < f_start[b + 1]; ++jf) {
                    const uint32_t rf = f_feat[jf];
                    if (out_match[rf] >= 0) continue;
                    const int d = hamming256(kf_desc + (size_t)rk * 32, f_desc + (size_t)rf * 32);
                    if (d < best1) { best2 = best1; best1 = d; best_idx = (int)rf; }
                    else if (d < best2) best2 = d;
                }
                if (best1 <= TH_LOW && (float)best1 < nnratio * (float)best2) {
                    out_match[best_idx] = (int32_t)rk;
                    if (check_ori) {
                        float rot = kf_angle[rk] - f_angle[best_idx];
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)std::round(rot * factor);
                        if (bin == HISTO) bin = 0;
                        hist[bin].push_back(best_idx);
                    }
                    ++nmatches;
                }
            }
            ++a; ++b;
        } else if (kf_node[a] < f_node[b]) ++a;
        else ++b;
    }
    if (check_ori) { /* ComputeThreeMaxima :1800-1843 */
        int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < HISTO; ++i) {
            const int s = (int)hist[i].size();
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
            else if (s > max3) { max3 = s; ind3 = i; }
        }
        if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
        else if (max3 < 0.1f * (float)max1) ind3 = -1;
        for (int i = 0; i < HISTO; ++i) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx : hist[i]) { out_match[idx] = -1; --nmatches; }
        }
    }
    return nmatches;
}

/* ORBmatcher::SearchByBoW(pKF1, pKF2, vpMatches12) :634-777 on the two feature vectors.  mp1 / mp2: bit 0 = the keypoint
 * has a map point, bit 1 = it isBad().  A pKF2 keypoint without a usable map point or matched already is skipped
 * (:686-690); acceptance is bestDist1 < TH_LOW -- strict here, unlike the keyframe-to-frame overload -- and the ratio
 * test (:712-714); the orientation histogram holds pKF1 indices (:726) and a discarded bin clears vpMatches12 there.
 * out_match12[i1] = pKF2 keypoint index or -1.  Returns nmatches. */
int bo_search_by_bow_kf(int n1, const float *angle1, const uint8_t *desc1, const uint8_t *mp1,
                        const uint32_t *node1, const int32_t *start1, const uint32_t *feat1, int nodes1,
                        int n2, const float *angle2, const uint8_t *desc2, const uint8_t *mp2,
                        const uint32_t *node2, const int32_t *start2, const uint32_t *feat2, int nodes2,
                        float nnratio, int check_ori, int32_t *out_match12)
{
    const int TH_LOW = 50, HISTO = 30;
    for (int i = 0; i < n1; ++i) out_match12[i] = -1;
    std::vector<char> matched2((size_t)std::max(n2, 1), 0);
    std::vector<int> hist[30];
    const float factor = 1.0f / HISTO;
    int nmatches = 0, a = 0, b = 0;
    while (a < nodes1 && b < nodes2) {
        if (node1[a] == node2[b]) {
            for (int i1 = start1[a]; i1 < start1[a + 1]; ++i1) {
                const uint32_t idx1 = feat1[i1];
                if (!(mp1[idx1] & 1) || (mp1[idx1] & 2)) continue;
                int best1 = 256, best2 = 256, best_idx = -1;
                for (int i2 = start2[b]; i2 < start2[b + 1]; ++i2) {
                    const uint32_t idx2 = feat2[i2];
                    if (matched2[idx2] || !(mp2[idx2] & 1)) continue;
                    if (mp2[idx2] & 2) continue;
                    const int d = hamming256(desc1 + (size_t)idx1 * 32, desc2 + (size_t)idx2 * 32);
                    if (d < best1) { best2 = best1; best1 = d; best_idx = (int)idx2; }
                    else if (d < best2) best2 = d;
                }
                if (best1 < TH_LOW && (float)best1 < nnratio * (float)best2) {
                    out_match12[idx1] = best_idx;
                    matched2[best_idx] = 1;
                    if (check_ori) {
                        float rot = angle1[idx1] - angle2[best_idx];
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)std::round(rot * factor);
                        if (bin == HISTO) bin = 0;
                        hist[bin].push_back((int)idx1);
                    }
                    ++nmatches;
                }
            }
            ++a; ++b;
        } else if (node1[a] < node2[b]) ++a;
        else ++b;
    }
    if (check_ori) { /* ComputeThreeMaxima :1800-1843 */
        int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < HISTO; ++i) {
            const int s = (int)hist[i].size();
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
            else if (s > max3) { max3 = s; ind3 = i; }
        }
        if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
        else if (max3 < 0.1f * (float)max1) ind3 = -1;
        for (int i = 0; i < HISTO; ++i) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx : hist[i]) { out_match12[idx] = -1; --nmatches; }
        }
    }
    return nmatches;
}

/* ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) :779-966 on the two feature vectors.
 * Only bit 0 of mp1 / mp2 is read (a keypoint that has ANY map point is left out on either side, :823-824, :850-851);
 * vbMatched2 is never set by the reference, so a pKF2 keypoint may be paired with several pKF1 keypoints and the
 * keypoints of pKF1 are independent of each other.  Inside a node the running best starts at TH_LOW and a candidate with
 * dist <= best that passes the geometry takes over (:867-888): the LAST of the smallest distances wins.
 * Float forms as the reference object has them (objdump of _ref/ORBmatcher.o):
 *   C2 = R2w * Cw + t2w       the small float product; invz = 1.0f / C2z
 *   ex = fmaf(fx * C2x, invz, cx), ey likewise                                            (:791-793)
 *   fmaf(distex, distex, distey * distey) < 100 * mvScaleFactors[octave2]  -> skipped      (:874-879)
 *   CheckDistEpipolarLine (:165-182): a = fmaf(x1, F00, y1 * F10) + F20, b = fmaf(x1, F01, y1 * F11) + F21,
 *   c = fmaf(y1, F12, x1 * F02) + F22, den = fmaf(a, a, b * b), num = c + fmaf(b, y2, a * x2),
 *   (double)(num * num / den) < 3.84 * (double)mvLevelSigma2[octave2]
 * xy / octave / angle come from the 28-B cv::KeyPoint records (mvKeysUn).  out_match12[i1] = pKF2 index or -1. */
struct bo_kp { float x, y, size, angle, response; int octave, class_id; };
int bo_search_for_triangulation(int n1, const uint8_t *kps1, const uint8_t *desc1, const uint8_t *mp1, const float *ur1,
                                const float *Ow1, const uint32_t *node1, const int32_t *start1, const uint32_t *feat1, int nodes1,
                                int n2, const uint8_t *kps2, const uint8_t *desc2, const uint8_t *mp2, const float *ur2,
                                const float *Tcw2, const uint32_t *node2, const int32_t *start2, const uint32_t *feat2, int nodes2,
                                const float *cam2, const float *scale2, const float *sigma2_2, const float *F, int only_stereo,
                                int check_ori, int32_t *out_match12)
{
    (void)n2;
    const int TH_LOW = 50, HISTO = 30;
    const bo_kp *k1 = reinterpret_cast<const bo_kp *>(kps1), *k2 = reinterpret_cast<const bo_kp *>(kps2);
    float C2[3];
    for (int r = 0; r < 3; ++r) {
        float s = 0.f;
        for (int k = 0; k < 3; ++k) s += Tcw2[4 * r + k] * Ow1[k];
        C2[r] = s + Tcw2[4 * r + 3];
    }
    const float invz = 1.0f / C2[2];
    const float ex = std::fmaf(cam2[0] * C2[0], invz, cam2[2]), ey = std::fmaf(cam2[1] * C2[1], invz, cam2[3]);
    for (int i = 0; i < n1; ++i) out_match12[i] = -1;
    std::vector<int> hist[30];
    const float factor = 1.0f / HISTO;
    int nmatches = 0, a = 0, b = 0;
    while (a < nodes1 && b < nodes2) {
        if (node1[a] == node2[b]) {
            for (int i1 = start1[a]; i1 < start1[a + 1]; ++i1) {
                const uint32_t idx1 = feat1[i1];
                if (mp1[idx1] & 1) continue;
                const bool stereo1 = ur1 && ur1[idx1] >= 0;
                if (only_stereo && !stereo1) continue;
                const float x1 = k1[idx1].x, y1 = k1[idx1].y;
                int best = TH_LOW, best_idx = -1;
                for (int i2 = start2[b]; i2 < start2[b + 1]; ++i2) {
                    const uint32_t idx2 = feat2[i2];
                    if (mp2[idx2] & 1) continue;
                    const bool stereo2 = ur2 && ur2[idx2] >= 0;
                    if (only_stereo && !stereo2) continue;
                    const int d = hamming256(desc1 + (size_t)idx1 * 32, desc2 + (size_t)idx2 * 32);
                    if (d > TH_LOW || d > best) continue;
                    const float x2 = k2[idx2].x, y2 = k2[idx2].y;
                    const int o2 = k2[idx2].octave;
                    if (!stereo1 && !stereo2) {
                        const float dx = ex - x2, dy = ey - y2;
                        if (std::fmaf(dx, dx, dy * dy) < 100 * scale2[o2]) continue;
                    }
                    const float la = std::fmaf(x1, F[0], y1 * F[3]) + F[6];
                    const float lb = std::fmaf(x1, F[1], y1 * F[4]) + F[7];
                    const float lc = std::fmaf(y1, F[5], x1 * F[2]) + F[8];
                    const float den = std::fmaf(la, la, lb * lb);
                    if (den == 0) continue;
                    const float num = lc + std::fmaf(lb, y2, la * x2);
                    const float dsqr = num * num / den;
                    if ((double)dsqr < 3.84 * (double)sigma2_2[o2]) { best_idx = (int)idx2; best = d; }
                }
                if (best_idx >= 0) {
                    out_match12[idx1] = best_idx;
                    ++nmatches;
                    if (check_ori) {
                        float rot = k1[idx1].angle - k2[best_idx].angle;
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)std::round(rot * factor);
                        if (bin == HISTO) bin = 0;
                        hist[bin].push_back((int)idx1);
                    }
                }
            }
            ++a; ++b;
        } else if (node1[a] < node2[b]) ++a;
        else ++b;
    }
    if (check_ori) { /* ComputeThreeMaxima :1800-1843 */
        int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < HISTO; ++i) {
            const int s = (int)hist[i].size();
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
            else if (s > max3) { max3 = s; ind3 = i; }
        }
        if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
        else if (max3 < 0.1f * (float)max1) ind3 = -1;
        for (int i = 0; i < HISTO; ++i) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx : hist[i]) { out_match12[idx] = -1; --nmatches; }
        }
    }
    return nmatches;
}

} // extern "C"


