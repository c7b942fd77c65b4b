/*
 * borb_noalloc.inl -- borb_undistort_grid_noalloc (include/borb.h), part of borb_api.cu.
 *
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  It is K8
 * (k_frame.cuh: Frame::UndistortKeyPoints + ComputeImageBounds + AssignFeaturesToGrid, src/Frame.cc:542-609,
 * :292-307) on caller-owned device buffers, with K8's one piece of scratch memory -- the cell number of every
 * keypoint between the histogram pass and the scatter pass -- in a file-scope __device__ array instead of a
 * cudaMalloc'ed slab.  The existing path (borb_set_camera + extract + borb_frame_grid) is untouched.
 *
 * The array below lives in the module's image: every GPU that loads the library holds BORB_NOALLOC_MAX_SLOTS * 4
 * bytes (1 MiB) until the process exits.  In HOST code `g_na_cell_of` is only a stand-in for the symbol, not a
 * device address: the address a kernel can use comes from cudaGetSymbolAddress, once per device, with that device
 * current (na_scratch).  Every call shares the array and finds the previous call's contents in it: nothing in it has
 * to start at zero (each entry a call reads, that call has written first; the counters that do start at zero are
 * shared memory, cleared by every CTA), and two calls that overlap on the GPU -- two streams, two host threads --
 * race on it.
 */
#include <mutex>

#define BORB_NA_MAXDEV 64

static __device__ __align__(256) int g_na_cell_of[BORB_NOALLOC_MAX_SLOTS];

namespace {

std::mutex g_na_mu;
int *g_na_ptr[BORB_NA_MAXDEV] = {};

/* device address of the scratch array on the CURRENT device (looked up once per device) */
int na_scratch(int dev, int **out)
{
    std::lock_guard<std::mutex> lock(g_na_mu);
    if (!g_na_ptr[dev]) {
        void *p = nullptr;
        CU(cudaGetSymbolAddress(&p, g_na_cell_of));
        if (!p) return fail(BORB_ERR_CUDA, "cudaGetSymbolAddress gave a null scratch address on device %d", dev);
        g_na_ptr[dev] = static_cast<int *>(p);
    }
    *out = g_na_ptr[dev];
    return BORB_OK;
}

/* K8 with the geometry by value (no device-resident plan: there is no handle) and the scratch passed in */
template <int NT>
__global__ void __launch_bounds__(NT)
k_na_undistort_grid(int cols, int rows, int kp_stride, BorbCamera cam, const BorbKpRec *__restrict__ kp,
                    const int *__restrict__ n_kp, float2 *__restrict__ un_xy, int *__restrict__ cell_start,
                    int *__restrict__ cell_idx, int *__restrict__ cell_of, float *__restrict__ bounds_out)
{
    __shared__ int cnt[FR_CELLS];
    __shared__ int cur[FR_CELLS];
    __shared__ int part[32];
    __shared__ float corner[4][2];
    __shared__ float bnd[6]; /* minX maxX minY maxY invW invH */
    const int frame = blockIdx.x, tid = threadIdx.x;
    const size_t fo = (size_t)frame * kp_stride;
    const int n = min(max(n_kp[frame], 0), kp_stride);
    const bool distorted = cam.k[0] != 0.0f;

    for (int c = tid; c < FR_CELLS; c += NT) cnt[c] = 0;
    if (tid < 4) {
        const float u = (tid & 1) ? (float)cols : 0.f, v = (tid & 2) ? (float)rows : 0.f;
        if (distorted) fr_undistort(cam, u, v, &corner[tid][0], &corner[tid][1]);
        else { corner[tid][0] = u; corner[tid][1] = v; }
    }
    __syncthreads();
    if (tid == 0) {
        bnd[0] = fminf(corner[0][0], corner[2][0]);
        bnd[1] = fmaxf(corner[1][0], corner[3][0]);
        bnd[2] = fminf(corner[0][1], corner[1][1]);
        bnd[3] = fmaxf(corner[2][1], corner[3][1]);
        bnd[4] = __fdiv_rn((float)FR_GRID_COLS, __fsub_rn(bnd[1], bnd[0]));
        bnd[5] = __fdiv_rn((float)FR_GRID_ROWS, __fsub_rn(bnd[3], bnd[2]));
        for (int i = 0; i < 4; ++i) bounds_out[frame * 4 + i] = bnd[i];
    }
    __syncthreads();
    /* pass 1: undistort, cell of every keypoint (-> scratch), histogram */
    for (int i = tid; i < n; i += NT) {
        const BorbKpRec k = kp[fo + i];
        float x = k.x, y = k.y;
        if (distorted) fr_undistort(cam, k.x, k.y, &x, &y);
        un_xy[fo + i] = make_float2(x, y);
        const int px = (int)roundf(__fmul_rn(__fsub_rn(x, bnd[0]), bnd[4]));
        const int py = (int)roundf(__fmul_rn(__fsub_rn(y, bnd[2]), bnd[5]));
        int cell = -1;
        if (px >= 0 && px < FR_GRID_COLS && py >= 0 && py < FR_GRID_ROWS) {
            cell = px * FR_GRID_ROWS + py;
            atomicAdd(&cnt[cell], 1);
        }
        cell_of[fo + i] = cell;
    }
    __syncthreads();
    /* pass 2: exclusive scan of the 3072 counts */
    const int per = FR_CELLS / NT;
    int sum = 0;
    for (int c = 0; c < per; ++c) sum += cnt[tid * per + c];
    int total;
    int base = cta_exscan<NT>(sum, part, &total);
    int *start = cell_start + (size_t)frame * (FR_CELLS + 1);
    for (int c = 0; c < per; ++c) {
        start[tid * per + c] = base;
        cur[tid * per + c] = base;
        base += cnt[tid * per + c];
    }
    if (tid == 0) start[FR_CELLS] = total;
    __syncthreads();
    /* pass 3: scatter (<- scratch), then ascending keypoint index inside each cell */
    for (int i = tid; i < n; i += NT) {
        const int c = cell_of[fo + i];
        if (c >= 0) cell_idx[fo + atomicAdd(&cur[c], 1)] = i;
    }
    __syncthreads();
    for (int c = tid * per; c < (tid + 1) * per; ++c) {
        const int s = start[c], e = s + cnt[c];
        for (int a = s + 1; a < e; ++a) {
            const int v = cell_idx[fo + a];
            int b = a - 1;
            while (b >= s && cell_idx[fo + b] > v) { cell_idx[fo + b + 1] = cell_idx[fo + b]; --b; }
            cell_idx[fo + b + 1] = v;
        }
    }
}

} // namespace

extern "C" int borb_undistort_grid_noalloc(const borb_keypoint *dev_kps, const int32_t *dev_counts, int n_frames,
                                           int kp_stride, int cols, int rows, const float *K4, const float *dist,
                                           int n_dist, float *dev_xy_un, int32_t *dev_cell_start, int32_t *dev_cell_idx,
                                           float *dev_bounds, void *stream)
{
    if (!dev_kps || !dev_counts || !dev_xy_un || !dev_cell_start || !dev_cell_idx || !dev_bounds || !K4)
        return fail(BORB_ERR_ARG, "null argument");
    if (n_frames < 1 || kp_stride < 1 || cols < 1 || rows < 1) return fail(BORB_ERR_ARG, "n_frames, kp_stride, cols, rows must be >= 1");
    if ((long long)n_frames * kp_stride > (long long)BORB_NOALLOC_MAX_SLOTS)
        return fail(BORB_ERR_CAPACITY, "%d frames x %d keypoint slots = %lld exceed the static scratch array of %d slots "
                    "(BORB_NOALLOC_MAX_SLOTS); split the call or use borb_set_camera + borb_frame_grid",
                    n_frames, kp_stride, (long long)n_frames * kp_stride, BORB_NOALLOC_MAX_SLOTS);
    if (n_dist < 0 || n_dist > 5 || (n_dist > 0 && !dist) || !(K4[0] != 0.f) || !(K4[1] != 0.f)) return fail(BORB_ERR_ARG, "bad camera");
    int dev = -1;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= BORB_NA_MAXDEV) return fail(BORB_ERR_ARG, "device %d", dev);
    int *scratch = nullptr;
    const int rc = na_scratch(dev, &scratch);
    if (rc != BORB_OK) return rc;
    BorbCamera cam;
    cam.fx = K4[0]; cam.fy = K4[1]; cam.cx = K4[2]; cam.cy = K4[3];
    for (int i = 0; i < 5; ++i) cam.k[i] = i < n_dist ? dist[i] : 0.f;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const BorbKpRec *kp = reinterpret_cast<const BorbKpRec *>(dev_kps);
    float2 *un = reinterpret_cast<float2 *>(dev_xy_un);
    if (n_frames <= 296)
        k_na_undistort_grid<FR_NT_LATENCY><<<n_frames, FR_NT_LATENCY, 0, st>>>(cols, rows, kp_stride, cam, kp, dev_counts, un,
                                                                           dev_cell_start, dev_cell_idx, scratch, dev_bounds);
    else
        k_na_undistort_grid<FR_NT_THROUGHPUT><<<n_frames, FR_NT_THROUGHPUT, 0, st>>>(cols, rows, kp_stride, cam, kp, dev_counts, un,
                                                                                 dev_cell_start, dev_cell_idx, scratch, dev_bounds);
    CU(cudaGetLastError());
    return BORB_OK;
}
