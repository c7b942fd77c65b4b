/*
 * oracle/ref_driver.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * C-ABI wrapper around the reference's own ORB_SLAM2::ORBextractor
 * (/root/reference/include/ORBextractor.h:46-144, src/ORBextractor.cc), which oracle/Makefile
 * compiles unmodified from the read-only reference tree against the cvcompat header shim.
 * Loaded with ctypes by tests (to validate our restatement orb_oracle.cpp) and by bench.py's
 * `--impl reference` / cpu_baseline legs (to time the reference's CPU path).
 *
 * With -DREF_MONOTONE_ALLOC the library also replaces operator new for the one allocation size
 * that matters -- std::list<ExtractorNode> nodes -- by a per-thread bump arena.  The reference
 * sorts pair<int, ExtractorNode*> (src/ORBextractor.cc:864-869), i.e. breaks equal-size ties by
 * heap address; under monotone allocation address order == creation order, which is the
 * canonical tie-break the project grades against.
 */
#include <atomic>
#include <sys/mman.h>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "ORBextractor.h"

namespace {

struct Exposed : public ORB_SLAM2::ORBextractor {
    using ORB_SLAM2::ORBextractor::ORBextractor;
    const std::vector<int> &quotas() const { return mnFeaturesPerLevel; }
    const std::vector<int> &umaxv() const { return umax; }
    /* the reference's quadtree on its own (protected member, include/ORBextractor.h:116-117) */
    std::vector<cv::KeyPoint> distribute(const std::vector<cv::KeyPoint> &k, int min_x, int max_x, int min_y,
                                         int max_y, int n, int level)
    {
        return DistributeOctTree(k, min_x, max_x, min_y, max_y, n, level);
    }
};

#ifdef REF_MONOTONE_ALLOC
const size_t kNodeBytes = sizeof(std::_List_node<ORB_SLAM2::ExtractorNode>);
/* Every thread that allocates list nodes gets its own bump arena, carved from ONE address-space
 * reservation (untouched pages cost nothing), so that operator delete -- possibly running on
 * another thread, possibly after the allocating thread has exited -- recognises arena memory by a
 * range check and leaves it alone.  A thread's arena goes back to a free list when the thread exits:
 * the reference's Frame constructor starts two fresh threads per frame (src/Frame.cc:110-113), and
 * without recycling a process would run out of arenas after 128 frames and silently fall back to
 * malloc, i.e. to the raw heap-address tie-break.  Every list node an extraction allocates dies
 * inside DistributeOctTree (lNodes is a local, src/ORBextractor.cc:682), so an exited thread's arena
 * holds no live node. */
const size_t kArenaBytes = size_t(64) << 20;
const size_t kReserveBytes = size_t(16) << 30; /* 256 thread arenas per process, then plain malloc */
thread_local char *t_arena = nullptr;
thread_local size_t t_used = 0;
thread_local bool t_active = false;
/* process-wide switch for callers that run the extractor on threads of their own (the reference's
 * Frame constructor spawns two, src/Frame.cc:110-113) */
std::atomic<int> g_force_monotone{0};
std::atomic<size_t> g_next_arena{0};
std::atomic<long> g_malloc_fallbacks{0}; /* node allocations that fell through to malloc while monotone mode was on */
/* free list of recycled arenas: fixed storage, spin lock (no allocation inside operator new) */
char *g_free_arena[256];
int g_n_free = 0;
std::atomic_flag g_free_lock = ATOMIC_FLAG_INIT;
struct ArenaReturn {
    ~ArenaReturn()
    {
        if (!t_arena) return;
        while (g_free_lock.test_and_set(std::memory_order_acquire)) {}
        if (g_n_free < 256) g_free_arena[g_n_free++] = t_arena;
        g_free_lock.clear(std::memory_order_release);
        t_arena = nullptr;
    }
};
thread_local ArenaReturn t_return;

inline char *reservation()
{
    static char *base = []() -> char * {
        void *p = mmap(nullptr, kReserveBytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        return p == MAP_FAILED ? nullptr : (char *)p;
    }();
    return base;
}

inline bool in_arena(void *p)
{
    char *base = reservation();
    return base && (char *)p >= base && (char *)p < base + kReserveBytes;
}
#endif

} // namespace

#ifdef REF_MONOTONE_ALLOC
void *operator new(size_t n)
{
    if ((t_active || g_force_monotone.load(std::memory_order_relaxed)) && n == kNodeBytes) {
        if (!t_arena && reservation()) {
            (void)&t_return; /* odr-use: constructed here, destroyed (arena returned) at thread exit */
            while (g_free_lock.test_and_set(std::memory_order_acquire)) {}
            if (g_n_free > 0) t_arena = g_free_arena[--g_n_free];
            g_free_lock.clear(std::memory_order_release);
            if (!t_arena) {
                const size_t off = g_next_arena.fetch_add(kArenaBytes);
                if (off + kArenaBytes <= kReserveBytes) t_arena = reservation() + off;
            }
            t_used = 0;
        }
        const size_t need = (n + 15) & ~size_t(15);
        if (t_arena && t_used + need <= kArenaBytes) {
            void *p = t_arena + t_used;
            t_used += need;
            return p;
        }
        g_malloc_fallbacks.fetch_add(1, std::memory_order_relaxed);
    }
    void *p = std::malloc(n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
}
void operator delete(void *p) noexcept
{
    if (p && !in_arena(p)) std::free(p);
}
void operator delete(void *p, size_t) noexcept
{
    if (p && !in_arena(p)) std::free(p);
}
#endif

extern "C" {

struct reforb_kp { /* == cv::KeyPoint, 28 bytes */
    float x, y, size, angle, response;
    int octave, class_id;
};

/* node allocations that missed the arena while monotone mode was on: must stay 0 for a canonical run */
long reforb_malloc_fallbacks(void)
{
#ifdef REF_MONOTONE_ALLOC
    return g_malloc_fallbacks.load();
#else
    return -1;
#endif
}

void reforb_force_monotone(int on)
{
#ifdef REF_MONOTONE_ALLOC
    g_force_monotone.store(on);
#else
    (void)on;
#endif
}

int reforb_monotone(void)
{
#ifdef REF_MONOTONE_ALLOC
    return 1;
#else
    return 0;
#endif
}

void *reforb_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th)
{
    return new Exposed(nfeatures, scale_factor, nlevels, ini_th, min_th);
}

void reforb_destroy(void *h) { delete (Exposed *)h; }

/* Runs operator(); copies up to cap keypoints / descriptors out; returns the keypoint count, or
 * -1 when the reference took its empty-image early return. */
int reforb_extract(void *h, const uint8_t *img, int cols, int rows, size_t step,
                   reforb_kp *kps, uint8_t *desc, int cap)
{
    Exposed *e = (Exposed *)h;
    if (!img || cols <= 0 || rows <= 0) {
        std::vector<cv::KeyPoint> k;
        cv::Mat d;
        (*e)(cv::Mat(), cv::Mat(), k, d);
        return -1;
    }
#ifdef REF_MONOTONE_ALLOC
    t_used = 0;
    t_active = true;
#endif
    cv::Mat image(rows, cols, CV_8UC1, (void *)img, step);
    std::vector<cv::KeyPoint> keys;
    cv::Mat d;
    (*e)(image, cv::Mat(), keys, d);
#ifdef REF_MONOTONE_ALLOC
    t_active = false;
#endif
    const int n = (int)keys.size();
    static_assert(sizeof(cv::KeyPoint) == sizeof(reforb_kp), "KeyPoint layout");
    for (int i = 0; i < n && i < cap; ++i) {
        std::memcpy(&kps[i], &keys[i], sizeof(reforb_kp));
        std::memcpy(desc + (size_t)i * 32, d.ptr(i), 32);
    }
    return n;
}

int reforb_levels(void *h) { return ((Exposed *)h)->GetLevels(); }
float reforb_scale_factor(void *h) { return ((Exposed *)h)->GetScaleFactor(); }

/* which: 0 scale factors, 1 inverse, 2 sigma2, 3 inverse sigma2 */
void reforb_table(void *h, int which, float *out)
{
    Exposed *e = (Exposed *)h;
    std::vector<float> v = which == 0 ? e->GetScaleFactors()
                         : which == 1 ? e->GetInverseScaleFactors()
                         : which == 2 ? e->GetScaleSigmaSquares()
                                      : e->GetInverseScaleSigmaSquares();
    std::memcpy(out, v.data(), v.size() * sizeof(float));
}
void reforb_quotas(void *h, int *out)
{
    const std::vector<int> &q = ((Exposed *)h)->quotas();
    std::memcpy(out, q.data(), q.size() * sizeof(int));
}
void reforb_umax(void *h, int *out)
{
    const std::vector<int> &u = ((Exposed *)h)->umaxv();
    std::memcpy(out, u.data(), u.size() * sizeof(int));
}

/* The reference's DistributeOctTree (src/ORBextractor.cc:668-951) on caller-supplied candidates, in
 * input order; class_id carries the candidate index through, so the survivors come back as
 * indices in the reference's output order.  Returns their count. */
int reforb_distribute(void *h, const int *xs, const int *ys, const int *resp, int n, int min_x, int max_x,
                      int min_y, int max_y, int n_target, int *out_idx, int cap)
{
    Exposed *e = (Exposed *)h;
    std::vector<cv::KeyPoint> in((size_t)n);
    for (int i = 0; i < n; ++i) {
        in[i] = cv::KeyPoint((float)xs[i], (float)ys[i], 7.f, -1.f, (float)resp[i], 0, i);
    }
#ifdef REF_MONOTONE_ALLOC
    t_used = 0;
    t_active = true;
#endif
    std::vector<cv::KeyPoint> out = e->distribute(in, min_x, max_x, min_y, max_y, n_target, 0);
#ifdef REF_MONOTONE_ALLOC
    t_active = false;
#endif
    for (size_t i = 0; i < out.size() && (int)i < cap; ++i) out_idx[i] = out[i].class_id;
    return (int)out.size();
}

/* geometry of mvImagePyramid[level] after the last call */
int reforb_pyramid_info(void *h, int level, int *cols, int *rows)
{
    Exposed *e = (Exposed *)h;
    if (level < 0 || level >= (int)e->mvImagePyramid.size() || e->mvImagePyramid[level].empty()) return -1;
    *cols = e->mvImagePyramid[level].cols;
    *rows = e->mvImagePyramid[level].rows;
    return 0;
}
/* copies the (cols+2*border) x (rows+2*border) padded parent of mvImagePyramid[level]; the
 * reference allocates it with a 19-pixel frame (src/ORBextractor.cc:1407-1417) */
int reforb_pyramid_copy(void *h, int level, int border, uint8_t *out)
{
    Exposed *e = (Exposed *)h;
    if (level < 0 || level >= (int)e->mvImagePyramid.size() || e->mvImagePyramid[level].empty()) return -1;
    const cv::Mat &m = e->mvImagePyramid[level];
    const int W = m.cols + 2 * border, H = m.rows + 2 * border;
    const uint8_t *base = m.data - (size_t)border * m.step - border;
    for (int y = 0; y < H; ++y) std::memcpy(out + (size_t)y * W, base + (size_t)y * m.step, (size_t)W);
    return 0;
}

} // extern "C"
