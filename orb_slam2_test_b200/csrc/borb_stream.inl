/*
 * borb_stream.inl -- the pipelined sequence feeder of include/borb.h (borb_stream_*), part of borb_api.cu.
 *
 * The reference's drivers push one frame at a time through System::TrackStereo (Examples/Stereo/stereo_kitti.cc:66-108);
 * inside, Frame::Frame runs the two extractors on two threads and then ComputeStereoMatches (src/Frame.cc:110-124).
 * A stream is that loop for a caller with many frames: a ring of batches, three CUDA streams.
 *
 *   upload stream    H2D of batch k+1            (page-locked slot buffer or the caller's own page-locked memory)
 *   compute streams  kernels of batch k          (the extractor handles' own streams; one cached graph per slot and eye;
 *                                                 for pairs the stereo association follows on the left handle's stream)
 *   download stream  D2H of batch k-1            (result slab of each eye, stereo slab, grid slab, pyramids on request)
 *
 * What is per SLOT: input buffers and result slabs, host and device -- a batch's input must survive until its pyramid
 * has been built, its results until they have been downloaded.  What is per HANDLE: pyramid, blurred levels, candidates,
 * survivors -- scratch of the kernel chain, which runs one batch at a time anyway.  The kernels find "their" result slab
 * through the handle's pointers, which are re-pointed at the slot's slabs around the enqueue (bind / unbind).
 */

struct StreamSlot {
    uint8_t *h_in[2] = {}, *d_in[2] = {};
    uint8_t *d_res[2] = {}, *h_res[2] = {};
    uint8_t *d_fr[2] = {}, *h_fr = nullptr;    /* K8 outputs; only the left eye's are downloaded */
    uint8_t *d_st = nullptr, *h_st = nullptr;  /* [u_right B*K f32][depth B*K f32][kept B i32] */
    uint8_t *h_pyr[2] = {};
    cudaEvent_t ev_up = nullptr, ev_done = nullptr, ev_down = nullptr;
    uint64_t tag = 0;
    int n = 0;
    int state = 0; /* 0 free, 1 in flight, 2 popped and held */
};

struct borb_stream {
    borb_extractor *ex[2] = {nullptr, nullptr};
    int eyes = 1, device = 0, cols = 0, rows = 0, batch = 0, slots = 0;
    unsigned flags = 0;
    float mbf = 0.f, mb = 0.f;
    bool have_cam = false;
    cudaStream_t up = nullptr, down = nullptr;
    std::vector<StreamSlot> ring;
    int head = 0, tail = 0, in_flight = 0, held = -1;
    cudaEvent_t last_down = nullptr; /* download of the most recently pushed batch (pyramid slab re-use) */
    size_t st_o_z = 0, st_o_kept = 0, st_bytes = 0;
    int launches = 0;
    /* feeder thread */
    std::thread worker;
    bool running = false;
    int worker_rc = BORB_OK;
    std::string worker_err;
};

namespace {

struct SlabSave {
    uint8_t *d_res, *h_res, *d_fr, *h_fr;
};

/* point the handle's result pointers at a result slab (same layout as its own: build_plan) */
void bind_res(borb_extractor *h, uint8_t *d_res, uint8_t *h_res)
{
    h->d_res = d_res; h->h_res = h_res;
    h->d_out_kp = reinterpret_cast<BorbKpRec *>(d_res); h->h_out_kp = reinterpret_cast<BorbKpRec *>(h_res);
    h->d_out_desc = d_res + h->res_o_desc; h->h_out_desc = h_res ? h_res + h->res_o_desc : nullptr;
    h->d_nout = reinterpret_cast<int *>(d_res + h->res_o_nout); h->h_nout = h_res ? reinterpret_cast<int *>(h_res + h->res_o_nout) : nullptr;
    h->d_err = reinterpret_cast<int *>(d_res + h->res_o_err); h->h_err = h_res ? reinterpret_cast<int *>(h_res + h->res_o_err) : nullptr;
}
void bind_fr(borb_extractor *h, uint8_t *d_fr, uint8_t *h_fr)
{
    h->d_fr = d_fr; h->h_fr = h_fr;
    h->d_un_xy = reinterpret_cast<float2 *>(d_fr); h->h_un_xy = reinterpret_cast<float2 *>(h_fr);
    h->d_cell_start = reinterpret_cast<int *>(d_fr + h->fr_o_cs); h->h_cell_start = h_fr ? reinterpret_cast<int *>(h_fr + h->fr_o_cs) : nullptr;
    h->d_cell_idx = reinterpret_cast<int *>(d_fr + h->fr_o_ci); h->h_cell_idx = h_fr ? reinterpret_cast<int *>(h_fr + h->fr_o_ci) : nullptr;
    h->d_bounds = reinterpret_cast<float *>(d_fr + h->fr_o_bd); h->h_bounds = h_fr ? reinterpret_cast<float *>(h_fr + h->fr_o_bd) : nullptr;
}
SlabSave bind_slot(borb_stream *s, const StreamSlot &sl, int e)
{
    borb_extractor *h = s->ex[e];
    SlabSave sv = {h->d_res, h->h_res, h->d_fr, h->h_fr};
    bind_res(h, sl.d_res[e], sl.h_res[e]);
    if (s->have_cam) bind_fr(h, sl.d_fr[e], e == 0 ? sl.h_fr : nullptr);
    return sv;
}
void unbind_slot(borb_stream *s, int e, const SlabSave &sv)
{
    borb_extractor *h = s->ex[e];
    bind_res(h, sv.d_res, sv.h_res);
    if (s->have_cam) bind_fr(h, sv.d_fr, sv.h_fr);
}

void stream_free(borb_stream *s)
{
    if (!s) return;
    cudaSetDevice(s->device);
    for (int e = 0; e < 2; ++e)
        if (s->ex[e] && s->ex[e]->stream) cudaStreamSynchronize(s->ex[e]->stream);
    if (s->up) cudaStreamSynchronize(s->up);
    if (s->down) cudaStreamSynchronize(s->down);
    for (StreamSlot &sl : s->ring) {
        for (int e = 0; e < 2; ++e) {
            if (sl.h_in[e]) cudaFreeHost(sl.h_in[e]);
            if (sl.d_in[e]) cudaFree(sl.d_in[e]);
            if (sl.d_res[e]) cudaFree(sl.d_res[e]);
            if (sl.h_res[e]) cudaFreeHost(sl.h_res[e]);
            if (sl.d_fr[e]) cudaFree(sl.d_fr[e]);
            if (sl.h_pyr[e]) cudaFreeHost(sl.h_pyr[e]);
        }
        if (sl.h_fr) cudaFreeHost(sl.h_fr);
        if (sl.d_st) cudaFree(sl.d_st);
        if (sl.h_st) cudaFreeHost(sl.h_st);
        if (sl.ev_up) cudaEventDestroy(sl.ev_up);
        if (sl.ev_done) cudaEventDestroy(sl.ev_done);
        if (sl.ev_down) cudaEventDestroy(sl.ev_down);
    }
    if (s->up) cudaStreamDestroy(s->up);
    if (s->down) cudaStreamDestroy(s->down);
    for (int e = 0; e < 2; ++e)
        if (s->ex[e]) borb_destroy(s->ex[e]);
    delete s;
}

/* K8's slabs per slot (allocated when a camera is set; the handles' own allocation happens in check_call) */
int stream_alloc_fr(borb_stream *s)
{
    for (int e = 0; e < s->eyes; ++e) {
        borb_extractor *h = s->ex[e];
        if (!h->d_un_xy) { /* the handle's own slab + cell scratch, as check_call does on the first call with a camera */
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
            bind_fr(h, h->d_fr, h->h_fr);
        }
        for (StreamSlot &sl : s->ring)
            if (!sl.d_fr[e]) {
                CU(cudaMalloc(&sl.d_fr[e], h->fr_bytes));
                CU(cudaMemset(sl.d_fr[e], 0, h->fr_bytes));
                if (e == 0) CU(cudaMallocHost(&sl.h_fr, h->fr_bytes));
            }
    }
    return BORB_OK;
}

bool stream_idle(const borb_stream *s) { return s->in_flight == 0 && !s->running; }

} // namespace

extern "C" {

int borb_stream_open(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th, int device, int cols,
                     int rows, int batch, int slots, unsigned flags, borb_stream **out)
{
    if (!out) return fail(BORB_ERR_ARG, "out is null");
    *out = nullptr;
    if (batch < 1 || slots < 2 || slots > 16 || cols < 1 || rows < 1)
        return fail(BORB_ERR_ARG, "stream: batch >= 1, 2 <= slots <= 16");
    if (flags & ~(BORB_STREAM_STEREO | BORB_STREAM_PYRAMID)) return fail(BORB_ERR_ARG, "unknown stream flags %u", flags);
    borb_stream *s = new borb_stream();
    s->eyes = (flags & BORB_STREAM_STEREO) ? 2 : 1;
    s->flags = flags; s->cols = cols; s->rows = rows; s->batch = batch; s->slots = slots;
    int rc = BORB_OK;
    for (int e = 0; e < s->eyes && rc == BORB_OK; ++e) {
        rc = borb_create(nfeatures, scale_factor, nlevels, ini_th, min_th, device, batch, &s->ex[e]);
        if (rc == BORB_OK) rc = build_plan(s->ex[e], cols, rows);
    }
    if (rc != BORB_OK) { stream_free(s); return rc; }
    s->device = s->ex[0]->device;
#define SCU(call)                                                                                              \
    do {                                                                                                       \
        cudaError_t e_ = (call);                                                                               \
        if (e_ != cudaSuccess) {                                                                               \
            stream_free(s);                                                                                    \
            return fail(BORB_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__);   \
        }                                                                                                      \
    } while (0)
    if (s->eyes == 2) {
        rc = check_stereo_pair(s->ex[0], s->ex[1]);
        if (rc == BORB_OK) rc = ensure_stereo_scratch(s->ex[0]);
        if (rc != BORB_OK) { stream_free(s); return rc; }
    }
    SCU(cudaStreamCreateWithFlags(&s->up, cudaStreamNonBlocking));
    SCU(cudaStreamCreateWithFlags(&s->down, cudaStreamNonBlocking));
    const borb_extractor *h0 = s->ex[0];
    const size_t B = (size_t)batch, K = (size_t)h0->plan.kp_total, fsz = (size_t)cols * rows;
    s->st_o_z = (size_t)align_up((long long)(B * K * 4), 256);
    s->st_o_kept = 2 * s->st_o_z;
    s->st_bytes = s->st_o_kept + (size_t)align_up((long long)(B * 4), 256);
    s->ring.resize((size_t)slots);
    for (StreamSlot &sl : s->ring) {
        for (int e = 0; e < s->eyes; ++e) {
            SCU(cudaMallocHost(&sl.h_in[e], B * fsz));
            SCU(cudaMalloc(&sl.d_in[e], B * fsz));
            SCU(cudaMalloc(&sl.d_res[e], h0->res_bytes));
            SCU(cudaMallocHost(&sl.h_res[e], h0->res_bytes));
            SCU(cudaMemset(sl.d_res[e], 0, h0->res_bytes));
            std::memset(sl.h_res[e], 0, h0->res_bytes);
            if (flags & BORB_STREAM_PYRAMID) SCU(cudaMallocHost(&sl.h_pyr[e], B * (size_t)h0->plan.pyr_bytes));
        }
        if (s->eyes == 2) {
            SCU(cudaMalloc(&sl.d_st, s->st_bytes));
            SCU(cudaMallocHost(&sl.h_st, s->st_bytes));
            SCU(cudaMemset(sl.d_st, 0, s->st_bytes));
        }
        SCU(cudaEventCreateWithFlags(&sl.ev_up, cudaEventDisableTiming));
        SCU(cudaEventCreateWithFlags(&sl.ev_done, cudaEventDisableTiming));
        SCU(cudaEventCreateWithFlags(&sl.ev_down, cudaEventDisableTiming));
    }
#undef SCU
    *out = s;
    return BORB_OK;
}

void borb_stream_close(borb_stream *s)
{
    if (!s) return;
    if (s->worker.joinable()) s->worker.join();
    stream_free(s);
}

int borb_stream_set_stereo(borb_stream *s, float mbf, float mb)
{
    if (!s || s->eyes != 2) return fail(BORB_ERR_ARG, "not a stereo stream");
    if (!(mbf > 0.f) || !(mb > 0.f)) return fail(BORB_ERR_ARG, "mbf %g, mb %g", (double)mbf, (double)mb);
    s->mbf = mbf; s->mb = mb;
    return BORB_OK;
}

int borb_stream_set_camera(borb_stream *s, const float *K4, const float *dist, int n_dist)
{
    if (!s) return fail(BORB_ERR_ARG, "null");
    if (!stream_idle(s)) return fail(BORB_ERR_STATE, "batches are in flight");
    for (int e = 0; e < s->eyes; ++e) {
        const int rc = borb_set_camera(s->ex[e], K4, dist, n_dist);
        if (rc != BORB_OK) return rc;
    }
    s->have_cam = K4 != nullptr;
    return s->have_cam ? stream_alloc_fr(s) : BORB_OK;
}

int borb_stream_in_flight(const borb_stream *s) { return s ? s->in_flight : 0; }
int borb_stream_slots_free(const borb_stream *s) { return s ? s->slots - s->in_flight - (s->held >= 0 ? 1 : 0) : 0; }

int borb_stream_input(borb_stream *s, int eye, uint8_t **buf, size_t *capacity)
{
    if (!s || eye < 0 || eye >= s->eyes || !buf) return fail(BORB_ERR_ARG, "stream / eye");
    if (borb_stream_slots_free(s) < 1) return fail(BORB_ERR_STATE, "ring full: %d in flight, %d held", s->in_flight, s->held >= 0);
    *buf = s->ring[(size_t)s->head].h_in[eye];
    if (capacity) *capacity = (size_t)s->batch * s->cols * s->rows;
    return BORB_OK;
}

int borb_stream_push(borb_stream *s, const uint8_t *left, const uint8_t *right, int n, size_t step, size_t stride,
                     uint64_t tag)
{
    if (!s) return fail(BORB_ERR_ARG, "null");
    if (n < 1 || n > s->batch) return fail(BORB_ERR_CAPACITY, "%d items, batch %d", n, s->batch);
    if (left && s->eyes == 2 && !right) return fail(BORB_ERR_ARG, "a stereo stream needs both eyes");
    if (s->eyes == 2 && !(s->mbf > 0.f)) return fail(BORB_ERR_STATE, "borb_stream_set_stereo has not been called");
    if (borb_stream_slots_free(s) < 1) return fail(BORB_ERR_STATE, "ring full: %d in flight, %d held", s->in_flight, s->held >= 0);
    const size_t rowb = (size_t)s->cols, fsz = rowb * s->rows;
    if (left && (step < rowb || (n > 1 && stride < step * (size_t)(s->rows - 1) + rowb))) return fail(BORB_ERR_ARG, "bad image geometry");
    CU(cudaSetDevice(s->device));
    StreamSlot &sl = s->ring[(size_t)s->head];
    const uint8_t *img[2] = {left, right};
    /* upload */
    for (int e = 0; e < s->eyes; ++e) {
        const uint8_t *src = sl.h_in[e];
        if (img[e]) {
            bool direct = false;
            if (step == rowb && (n == 1 || stride == fsz)) {
                cudaPointerAttributes at;
                if (cudaPointerGetAttributes(&at, img[e]) == cudaSuccess && at.type == cudaMemoryTypeHost) direct = true;
                else cudaGetLastError();
            }
            if (direct) src = img[e];
            else
                for (int f = 0; f < n; ++f) {
                    const uint8_t *p = img[e] + (size_t)f * stride;
                    uint8_t *d = sl.h_in[e] + (size_t)f * fsz;
                    if (step == rowb) pack_copy(d, p, fsz); /* streaming stores: borb_api.cu */
                    else for (int y = 0; y < s->rows; ++y) pack_copy(d + (size_t)y * rowb, p + (size_t)y * step, rowb);
                }
        }
        CU(cudaMemcpyAsync(sl.d_in[e], src, (size_t)n * fsz, cudaMemcpyHostToDevice, s->up));
    }
    CU(cudaEventRecord(sl.ev_up, s->up));
    /* kernels */
    SlabSave sv[2];
    int rc = BORB_OK, launches = 0;
    for (int e = 0; e < s->eyes && rc == BORB_OK; ++e) {
        borb_extractor *h = s->ex[e];
        CU(cudaStreamWaitEvent(h->stream, sl.ev_up, 0));
        /* the pyramid is downloaded from the handle's own slab: the next batch's K1 must not overwrite it earlier */
        if ((s->flags & BORB_STREAM_PYRAMID) && s->last_down) CU(cudaStreamWaitEvent(h->stream, s->last_down, 0));
        sv[e] = bind_slot(s, sl, e);
        GraphKey key = {2, sl.d_in[e], n, rowb, fsz, 0, s->cols, s->rows};
        rc = run_cached(h, key, [&]() -> int { return enqueue_kernels(h, sl.d_in[e], n, rowb, fsz, s->cols, s->rows); },
                        true /* a ring of slots is a fixed set of call shapes: capture each on first use */);
        h->last_n = n;
        launches += h->last_launches;
    }
    if (rc == BORB_OK && s->eyes == 2) {
        rc = enqueue_stereo(s->ex[0], s->ex[1], n, s->mbf, s->mb, reinterpret_cast<float *>(sl.d_st),
                            reinterpret_cast<float *>(sl.d_st + s->st_o_z), reinterpret_cast<int *>(sl.d_st + s->st_o_kept));
        launches += 3;
    }
    for (int e = 0; e < s->eyes; ++e) unbind_slot(s, e, sv[e]);
    if (rc != BORB_OK) return rc;
    s->launches = launches;
    CU(cudaEventRecord(sl.ev_done, s->ex[0]->stream));
    /* download */
    CU(cudaStreamWaitEvent(s->down, sl.ev_done, 0));
    const borb_extractor *h0 = s->ex[0];
    const size_t K = (size_t)h0->plan.kp_total, un = (size_t)n;
    for (int e = 0; e < s->eyes; ++e) {
        if (n == s->batch) {
            CU(cudaMemcpyAsync(sl.h_res[e], sl.d_res[e], h0->res_bytes, cudaMemcpyDeviceToHost, s->down));
        } else {
            CU(cudaMemcpyAsync(sl.h_res[e], sl.d_res[e], un * K * sizeof(BorbKpRec), cudaMemcpyDeviceToHost, s->down));
            CU(cudaMemcpyAsync(sl.h_res[e] + h0->res_o_desc, sl.d_res[e] + h0->res_o_desc, un * K * 32, cudaMemcpyDeviceToHost, s->down));
            CU(cudaMemcpyAsync(sl.h_res[e] + h0->res_o_nout, sl.d_res[e] + h0->res_o_nout, un * 4, cudaMemcpyDeviceToHost, s->down));
            CU(cudaMemcpyAsync(sl.h_res[e] + h0->res_o_err, sl.d_res[e] + h0->res_o_err, 4, cudaMemcpyDeviceToHost, s->down));
        }
        if (s->flags & BORB_STREAM_PYRAMID)
            CU(cudaMemcpyAsync(sl.h_pyr[e], s->ex[e]->d_pyr, un * (size_t)h0->plan.pyr_bytes, cudaMemcpyDeviceToHost, s->down));
    }
    if (s->eyes == 2) {
        if (n == s->batch) {
            CU(cudaMemcpyAsync(sl.h_st, sl.d_st, s->st_bytes, cudaMemcpyDeviceToHost, s->down));
        } else {
            CU(cudaMemcpyAsync(sl.h_st, sl.d_st, un * K * 4, cudaMemcpyDeviceToHost, s->down));
            CU(cudaMemcpyAsync(sl.h_st + s->st_o_z, sl.d_st + s->st_o_z, un * K * 4, cudaMemcpyDeviceToHost, s->down));
            CU(cudaMemcpyAsync(sl.h_st + s->st_o_kept, sl.d_st + s->st_o_kept, un * 4, cudaMemcpyDeviceToHost, s->down));
        }
    }
    if (s->have_cam) {
        if (n == s->batch) {
            CU(cudaMemcpyAsync(sl.h_fr, sl.d_fr[0], h0->fr_bytes, cudaMemcpyDeviceToHost, s->down));
        } else {
            CU(cudaMemcpyAsync(sl.h_fr, sl.d_fr[0], un * K * 8, cudaMemcpyDeviceToHost, s->down));
            CU(cudaMemcpyAsync(sl.h_fr + h0->fr_o_cs, sl.d_fr[0] + h0->fr_o_cs, un * (FR_CELLS + 1) * 4, cudaMemcpyDeviceToHost, s->down));
            CU(cudaMemcpyAsync(sl.h_fr + h0->fr_o_ci, sl.d_fr[0] + h0->fr_o_ci, un * K * 4, cudaMemcpyDeviceToHost, s->down));
            CU(cudaMemcpyAsync(sl.h_fr + h0->fr_o_bd, sl.d_fr[0] + h0->fr_o_bd, un * 16, cudaMemcpyDeviceToHost, s->down));
        }
    }
    CU(cudaEventRecord(sl.ev_down, s->down));
    s->last_down = sl.ev_down;
    sl.tag = tag; sl.n = n; sl.state = 1;
    s->head = (s->head + 1) % s->slots;
    ++s->in_flight;
    return BORB_OK;
}

int borb_stream_release(borb_stream *s)
{
    if (!s) return fail(BORB_ERR_ARG, "null");
    if (s->held >= 0) {
        s->ring[(size_t)s->held].state = 0;
        s->held = -1;
    }
    return BORB_OK;
}

int borb_stream_pop(borb_stream *s, uint64_t *tag, int *n)
{
    if (!s) return fail(BORB_ERR_ARG, "null");
    if (s->in_flight < 1) return fail(BORB_ERR_STATE, "nothing in flight");
    borb_stream_release(s);
    CU(cudaSetDevice(s->device));
    StreamSlot &sl = s->ring[(size_t)s->tail];
    CU(cudaEventSynchronize(sl.ev_down));
    sl.state = 2;
    s->held = s->tail;
    s->tail = (s->tail + 1) % s->slots;
    --s->in_flight;
    if (tag) *tag = sl.tag;
    if (n) *n = sl.n;
    for (int e = 0; e < s->eyes; ++e) {
        int *err = reinterpret_cast<int *>(sl.h_res[e] + s->ex[0]->res_o_err);
        if (*err != 0) {
            const int code = *err;
            *err = 0;
            cudaMemsetAsync(sl.d_res[e] + s->ex[0]->res_o_err, 0, 4, s->ex[e]->stream);
            return fail(BORB_ERR_CAPACITY, "device buffer overflow (code %d, eye %d)", code, e);
        }
    }
    return BORB_OK;
}

static const StreamSlot *held_slot(const borb_stream *s, int item)
{
    if (!s || s->held < 0) { fail(BORB_ERR_STATE, "no popped batch"); return nullptr; }
    const StreamSlot &sl = s->ring[(size_t)s->held];
    if (item < 0 || item >= sl.n) { fail(BORB_ERR_ARG, "item %d of %d", item, sl.n); return nullptr; }
    return &sl;
}

int borb_stream_result(const borb_stream *s, int eye, int item, const borb_keypoint **kps, const uint8_t **desc, int *n)
{
    const StreamSlot *sl = held_slot(s, item);
    if (!sl) return s && s->held >= 0 ? BORB_ERR_ARG : BORB_ERR_STATE;
    if (eye < 0 || eye >= s->eyes) return fail(BORB_ERR_ARG, "eye %d", eye);
    const borb_extractor *h = s->ex[0];
    const size_t o = (size_t)item * h->plan.kp_total;
    const uint8_t *base = sl->h_res[eye];
    if (kps) *kps = reinterpret_cast<const borb_keypoint *>(base) + o;
    if (desc) *desc = base + h->res_o_desc + o * 32;
    if (n) *n = reinterpret_cast<const int *>(base + h->res_o_nout)[item];
    return BORB_OK;
}

int borb_stream_stereo_result(const borb_stream *s, int item, const float **u_right, const float **depth, int *n, int *n_matched)
{
    const StreamSlot *sl = held_slot(s, item);
    if (!sl) return s && s->held >= 0 ? BORB_ERR_ARG : BORB_ERR_STATE;
    if (s->eyes != 2) return fail(BORB_ERR_STATE, "not a stereo stream");
    const borb_extractor *h = s->ex[0];
    const size_t o = (size_t)item * h->plan.kp_total;
    if (u_right) *u_right = reinterpret_cast<const float *>(sl->h_st) + o;
    if (depth) *depth = reinterpret_cast<const float *>(sl->h_st + s->st_o_z) + o;
    if (n) *n = reinterpret_cast<const int *>(sl->h_res[0] + h->res_o_nout)[item];
    if (n_matched) *n_matched = reinterpret_cast<const int *>(sl->h_st + s->st_o_kept)[item];
    return BORB_OK;
}

int borb_stream_frame_grid(const borb_stream *s, int item, const float **xy_un, const int32_t **cell_start,
                           const int32_t **cell_idx, float *bounds)
{
    const StreamSlot *sl = held_slot(s, item);
    if (!sl) return s && s->held >= 0 ? BORB_ERR_ARG : BORB_ERR_STATE;
    if (!s->have_cam || !sl->h_fr) return fail(BORB_ERR_STATE, "no camera set");
    const borb_extractor *h = s->ex[0];
    const size_t o = (size_t)item * h->plan.kp_total;
    if (xy_un) *xy_un = reinterpret_cast<const float *>(sl->h_fr) + 2 * o;
    if (cell_start) *cell_start = reinterpret_cast<const int32_t *>(sl->h_fr + h->fr_o_cs) + (size_t)item * (FR_CELLS + 1);
    if (cell_idx) *cell_idx = reinterpret_cast<const int32_t *>(sl->h_fr + h->fr_o_ci) + o;
    if (bounds) std::memcpy(bounds, reinterpret_cast<const float *>(sl->h_fr + h->fr_o_bd) + (size_t)item * 4, 16);
    return BORB_OK;
}

int borb_stream_pyramid(const borb_stream *s, int eye, int item, int level, const uint8_t **data, int *cols, int *rows, size_t *step)
{
    const StreamSlot *sl = held_slot(s, item);
    if (!sl) return s && s->held >= 0 ? BORB_ERR_ARG : BORB_ERR_STATE;
    if (!(s->flags & BORB_STREAM_PYRAMID)) return fail(BORB_ERR_STATE, "the stream was opened without BORB_STREAM_PYRAMID");
    const borb_extractor *h = s->ex[0];
    if (eye < 0 || eye >= s->eyes || level < 0 || level >= h->nlevels) return fail(BORB_ERR_ARG, "eye/level");
    const BorbLevel &L = h->plan.lv[level];
    if (data) *data = sl->h_pyr[eye] + (size_t)item * (size_t)h->plan.pyr_bytes + L.off + (size_t)BORB_EDGE * L.pitch + BORB_XPAD;
    if (cols) *cols = L.w;
    if (rows) *rows = L.h;
    if (step) *step = (size_t)L.pitch;
    return BORB_OK;
}

int borb_stream_link_bytes(const borb_stream *s, uint64_t *bytes)
{
    if (!s || !bytes) return fail(BORB_ERR_ARG, "null");
    const borb_extractor *h = s->ex[0];
    bytes[0] = (uint64_t)s->eyes * s->batch * s->cols * s->rows;
    bytes[1] = (uint64_t)s->eyes * h->res_bytes + (s->eyes == 2 ? s->st_bytes : 0) + (s->have_cam ? h->fr_bytes : 0) +
               ((s->flags & BORB_STREAM_PYRAMID) ? (uint64_t)s->eyes * s->batch * (uint64_t)h->plan.pyr_bytes : 0);
    return BORB_OK;
}

int borb_stream_launches_per_push(const borb_stream *s)
{
    if (!s) return 0;
    if (s->launches) return s->launches;
    return s->eyes * borb_launches_per_call(s->ex[0]) + (s->eyes == 2 ? 3 : 0);
}

int borb_stream_start(borb_stream *s, borb_produce_fn produce, borb_consume_fn consume, void *user)
{
    if (!s || !produce) return fail(BORB_ERR_ARG, "null");
    if (s->running || s->worker.joinable()) return fail(BORB_ERR_STATE, "the feeder thread is already running");
    if (s->in_flight) return fail(BORB_ERR_STATE, "batches are in flight");
    s->running = true;
    s->worker_rc = BORB_OK;
    s->worker_err.clear();
    s->worker = std::thread([s, produce, consume, user]() {
        uint64_t index = 0;
        bool end = false;
        int rc = BORB_OK;
        auto drain_one = [&]() -> int {
            uint64_t tag = 0;
            int n = 0;
            const int r = borb_stream_pop(s, &tag, &n);
            if (r == BORB_OK && consume) consume(user, s, tag, n);
            return r;
        };
        while (rc == BORB_OK && !(end && s->in_flight == 0)) {
            if (!end && borb_stream_slots_free(s) > 0) {
                uint8_t *l = nullptr, *r = nullptr;
                rc = borb_stream_input(s, 0, &l, nullptr);
                if (rc == BORB_OK && s->eyes == 2) rc = borb_stream_input(s, 1, &r, nullptr);
                if (rc != BORB_OK) break;
                const int n = produce(user, index, l, r, s->batch);
                if (n < 0) { rc = fail(BORB_ERR_ARG, "the producer aborted at batch %llu (%d)", (unsigned long long)index, n); break; }
                if (n == 0) { end = true; continue; }
                rc = borb_stream_push(s, nullptr, nullptr, n, 0, 0, index);
                ++index;
                /* keep the ring full before blocking on the oldest batch */
                if (rc == BORB_OK && borb_stream_slots_free(s) > 0) continue;
            }
            if (rc == BORB_OK && s->in_flight > 0) rc = drain_one();
        }
        if (rc != BORB_OK) {
            s->worker_err = g_err; /* thread-local message -> the joining thread */
            /* drain what is in flight so that the stream is reusable */
            while (s->in_flight > 0) {
                uint64_t tag; int n;
                if (borb_stream_pop(s, &tag, &n) != BORB_OK) break;
            }
        }
        borb_stream_release(s);
        s->worker_rc = rc;
    });
    return BORB_OK;
}

int borb_stream_join(borb_stream *s)
{
    if (!s) return fail(BORB_ERR_ARG, "null");
    if (!s->worker.joinable()) return fail(BORB_ERR_STATE, "no feeder thread");
    s->worker.join();
    s->running = false;
    if (s->worker_rc != BORB_OK) {
        g_err = s->worker_err;
        return s->worker_rc;
    }
    return BORB_OK;
}

} // extern "C"
