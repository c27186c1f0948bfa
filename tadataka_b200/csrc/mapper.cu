// Native keyframe loop: the device-resident state of examples/semi_dense_vo.py:152-207
// (increment_age -> propagate -> update_depth [-> fusion] [-> regularize] -> refframes.append) behind
// ONE host call per frame.  All device memory is one caller-owned arena (tdk_mapper_arena_bytes);
// the mapper keeps the reference images in a contiguous ring inside it, so update_depth runs on
// plain [R,H,W] blocks without pointer tables, stages the incoming frame through pinned memory
// (an 8-bit gray / RGB / RGBA frame is converted to the f64 gray image on the device), and never
// synchronises the host with the GPU inside track(): the status words of a frame's update and the
// age census that drives the pruning of unused reference frames are read back asynchronously into
// two alternating host buffers and looked at two calls later, so the host can enqueue frame k + 1
// while the GPU still works on frame k.
#include <stdlib.h>
#include <string.h>

#include <new>

#include "tdk_common.cuh"

namespace {

size_t up256(size_t v) { return (v + 255) / 256 * 256; }

// 8-bit frame -> f64 gray image.  gray = v / 255 for one channel; for RGB(A) the luma of
// skimage.color.rgb2gray, which the reference's dataset loaders apply to every frame
// (tadataka/dataset/new_tsukuba.py:85-86): ((R/255) 0.2125 + (G/255) 0.7154) + (B/255) 0.0721,
// each operation rounded on its own (no FMA), left to right; alpha is dropped.
__global__ void __launch_bounds__(256) ingest_u8_kernel(const unsigned char *__restrict__ src, int channels, size_t n,
                                                        double *__restrict__ dst)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        if (channels == 1) {
            dst[i] = (double)src[i] / 255.;
        } else {
            const unsigned char *q = src + i * channels;
            const double r = (double)q[0] / 255., g = (double)q[1] / 255., b = (double)q[2] / 255.;
            dst[i] = (r * 0.2125 + g * 0.7154) + b * 0.0721;
        }
    }
}

struct Layout {
    size_t ring, cams, poses, T10, depth[2], var[2], age[2], flag, tmp_d, tmp_v, raw, counts, ws_update, ws_prop, total;
    size_t ws_update_bytes, ws_prop_bytes;
};

Layout make_layout(int64_t H, int64_t W, int64_t cap)
{
    Layout l;
    const size_t hw = (size_t)H * W;
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t at = o; o += up256(bytes); return at; };
    l.ring = take((size_t)cap * hw * 8);
    l.cams = take((size_t)cap * 4 * 8);
    l.poses = take((size_t)cap * 16 * 8);
    l.T10 = take(16 * 8);
    for (int i = 0; i < 2; i++) { l.depth[i] = take(hw * 8); l.var[i] = take(hw * 8); l.age[i] = take(hw * 8); }
    l.flag = take(hw * 8);
    l.tmp_d = take(hw * 8);
    l.tmp_v = take(hw * 8);
    l.raw = take(hw * 4);                       // an RGBA frame before conversion
    l.counts = take((size_t)(cap + 2) * 8);
    l.ws_update_bytes = tdk_update_depth_workspace_bytes(1, cap, H, W);
    l.ws_update = take(l.ws_update_bytes);
    l.ws_prop_bytes = tdk_propagate_workspace_bytes(1, H, W);
    l.ws_prop = take(l.ws_prop_bytes);
    l.total = o;
    return l;
}

}  // namespace

struct tdk_mapper {
    int dev;
    int64_t H, W, cap;
    size_t hw;
    double params[6], default_depth, default_variance, bias;
    int fuse, regularize, prune;
    char *arena;
    Layout l;
    int64_t start, end;     // live window [start, end) of the ring: reference frames + the newest frame
    int cur;                // which of the two map sets holds the current maps
    bool has_frame;
    cudaStream_t stream;
    // pinned staging, double buffered; an event says when the copies out of a buffer are done
    unsigned char *stage[2];
    cudaEvent_t stage_done[2];
    int parity;
    // results that travel back asynchronously: frame k uses set k % 2, and track() looks at a set just
    // before it reuses it -- two frames later, so the host never waits for the frame the GPU is working on
    struct Back {
        uint64_t *counts_host;
        int *status_host;
        cudaEvent_t done;
        int64_t census_n;         // ring length the pending census belongs to (0 = none pending)
        int64_t census_first_id;  // absolute id of the oldest frame the census saw
        bool status_pending;
    } back[2];
    int64_t first_id;       // absolute id (frames since reset) of the frame in ring slot `start`
    int64_t tracked;
};

namespace {

template <typename T>
T *at(tdk_mapper *m, size_t off) { return reinterpret_cast<T *>(m->arena + off); }

size_t frame_bytes(const tdk_mapper *m, int format)
{
    switch (format) {
        case TDK_IMAGE_F64: return m->hw * 8;
        case TDK_IMAGE_U8: return m->hw;
        case TDK_IMAGE_RGB8: return m->hw * 3;
        case TDK_IMAGE_RGBA8: return m->hw * 4;
        default: return 0;
    }
}

// Put a frame into ring slot `slot` and its camera / pose beside it.
int upload_frame(tdk_mapper *m, int64_t slot, const void *image, int format, const double *cam, const double *T_wf,
                 const double *T10)
{
    const size_t nb = frame_bytes(m, format);
    if (nb == 0) { tdk_set_error("tdk_mapper: unknown image format %d", format); return TDK_ERR_INVALID; }
    const int k = m->parity;
    m->parity ^= 1;
    TDK_CUDA(cudaEventSynchronize(m->stage_done[k]));   // the copies of two frames ago; long finished
    unsigned char *st = m->stage[k];
    double *small = reinterpret_cast<double *>(st);      // [cam 4 | pose 16 | T10 16], then the frame
    memcpy(small, cam, 32);
    memcpy(small + 4, T_wf, 128);
    if (T10) memcpy(small + 20, T10, 128);
    memcpy(st + 512, image, nb);
    double *dst = at<double>(m, m->l.ring) + (size_t)slot * m->hw;
    TDK_CUDA(cudaMemcpyAsync(at<double>(m, m->l.cams) + 4 * slot, small, 32, cudaMemcpyHostToDevice, m->stream));
    TDK_CUDA(cudaMemcpyAsync(at<double>(m, m->l.poses) + 16 * slot, small + 4, 128, cudaMemcpyHostToDevice, m->stream));
    if (T10) TDK_CUDA(cudaMemcpyAsync(at<double>(m, m->l.T10), small + 20, 128, cudaMemcpyHostToDevice, m->stream));
    if (format == TDK_IMAGE_F64) {
        TDK_CUDA(cudaMemcpyAsync(dst, st + 512, nb, cudaMemcpyHostToDevice, m->stream));
    } else {
        unsigned char *raw = at<unsigned char>(m, m->l.raw);
        TDK_CUDA(cudaMemcpyAsync(raw, st + 512, nb, cudaMemcpyHostToDevice, m->stream));
        const int channels = format == TDK_IMAGE_U8 ? 1 : format == TDK_IMAGE_RGB8 ? 3 : 4;
        unsigned blocks = (unsigned)((m->hw + 1023) / 1024);
        if (blocks > 2048u) blocks = 2048u;
        ingest_u8_kernel<<<blocks, 256, 0, m->stream>>>(raw, channels, m->hw, dst);
        TDK_CUDA(cudaGetLastError());
    }
    TDK_CUDA(cudaEventRecord(m->stage_done[k], m->stream));
    return TDK_OK;
}

// Look at what an earlier frame sent back: its update's status words and, when pruning is on, the
// age census that says which reference frames are still in use.
int collect(tdk_mapper *m, tdk_mapper::Back &b, bool apply_census)
{
    if (!b.status_pending && (b.census_n == 0 || !apply_census)) return TDK_OK;
    TDK_CUDA(cudaEventSynchronize(b.done));
    int rc = TDK_OK;
    if (b.status_pending) {
        b.status_pending = false;
        const int *s = b.status_host;
        if (s[TDK_ST_SINGULAR]) { tdk_set_error("a reference pose T_wr is singular"); rc = TDK_ERR_SINGULAR; }
        else if (s[TDK_ST_AGE]) { tdk_set_error("Age exceeds the refframe size"); rc = TDK_ERR_AGE; }
        else if (s[TDK_ST_SEARCH_LEN]) {
            tdk_set_error("an epipolar search line has more than %d samples; increase ref_step_size", TDK_MAX_SEARCH_SAMPLES);
            rc = TDK_ERR_SEARCH_LEN;
        }
    }
    if (apply_census && b.census_n > 0) {
        // update_depth reads refframes[len - age] (semi_dense.rs:207): frames older than the largest
        // age on the map are dead, and stay dead -- ages and the ring grow together.  At the update
        // after the census the ages have grown by one and the ring by one, so of the n frames the
        // census saw, the one at position n - 1 - oldest is the oldest still read; the newest frame
        // always stays (the next propagation needs its pose and camera).  Frames are named by their
        // absolute id, so a census may be applied any number of frames late.
        const int64_t n = b.census_n;
        b.census_n = 0;
        int64_t oldest = 0;
        for (int64_t a = 0; a < n + 2; a++)
            if (b.counts_host[a]) oldest = a;
        int64_t keep_from = n - 1 - oldest;
        if (keep_from > n - 1) keep_from = n - 1;
        const int64_t first_live = b.census_first_id + (keep_from > 0 ? keep_from : 0);
        if (first_live > m->first_id) {
            int64_t drop = first_live - m->first_id;
            const int64_t have = m->end - m->start;
            if (drop > have - 1) drop = have - 1;
            if (drop > 0) { m->start += drop; m->first_id += drop; }
        }
    }
    return rc;
}

// The ring's window has reached the end of the arena: move it to the front.
int compact(tdk_mapper *m)
{
    const int64_t n = m->end - m->start;
    if (m->start == 0) {
        tdk_set_error("tdk_mapper: %lld frames are in use, the ring holds %lld; create the mapper with a larger "
                      "capacity or enable pruning", (long long)n, (long long)m->cap);
        return TDK_ERR_WORKSPACE;
    }
    double *ring = at<double>(m, m->l.ring), *cams = at<double>(m, m->l.cams), *poses = at<double>(m, m->l.poses);
    for (int64_t i = 0; i < n; i++) {   // frame by frame, ascending: a frame never overlaps its destination
        const int64_t s = m->start + i;
        TDK_CUDA(cudaMemcpyAsync(ring + (size_t)i * m->hw, ring + (size_t)s * m->hw, m->hw * 8, cudaMemcpyDeviceToDevice, m->stream));
        TDK_CUDA(cudaMemcpyAsync(cams + 4 * i, cams + 4 * s, 32, cudaMemcpyDeviceToDevice, m->stream));
        TDK_CUDA(cudaMemcpyAsync(poses + 16 * i, poses + 16 * s, 128, cudaMemcpyDeviceToDevice, m->stream));
    }
    m->start = 0;
    m->end = n;
    return TDK_OK;
}

}  // namespace

extern "C" {

size_t tdk_mapper_arena_bytes(int64_t H, int64_t W, int64_t capacity)
{
    if (H <= 0 || W <= 0 || capacity < 2) return 0;
    return make_layout(H, W, capacity).total;
}

int tdk_mapper_create(void *arena, size_t arena_bytes, int64_t H, int64_t W, int64_t capacity, const double *params,
                      double default_depth, double default_variance, double uncertaintity_bias, int fuse,
                      int regularize, int prune, void *stream, tdk_mapper **out)
{
    if (!arena || !params || !out || H <= 0 || W <= 0 || capacity < 2 || capacity > 65535) {
        tdk_set_error("tdk_mapper_create: bad argument (capacity must be in [2, 65535])");
        return TDK_ERR_INVALID;
    }
    const Layout l = make_layout(H, W, capacity);
    if (arena_bytes < l.total) {
        tdk_set_error("tdk_mapper_create: arena %zu < required %zu", arena_bytes, l.total);
        return TDK_ERR_WORKSPACE;
    }
    tdk_mapper *m = new (std::nothrow) tdk_mapper();
    if (!m) { tdk_set_error("tdk_mapper_create: out of host memory"); return TDK_ERR_INVALID; }
    memset(m, 0, sizeof(*m));
    TDK_CUDA(cudaGetDevice(&m->dev));
    m->H = H; m->W = W; m->cap = capacity; m->hw = (size_t)H * W;
    memcpy(m->params, params, sizeof(m->params));
    m->default_depth = default_depth; m->default_variance = default_variance; m->bias = uncertaintity_bias;
    m->fuse = fuse ? 1 : 0; m->regularize = regularize ? 1 : 0; m->prune = prune ? 1 : 0;
    m->arena = (char *)arena;
    m->l = l;
    m->stream = (cudaStream_t)stream;
    int rc = TDK_OK;
    const size_t stage_bytes = 512 + m->hw * 8;
    for (int i = 0; i < 2 && rc == TDK_OK; i++) {
        rc = tdk_check_cuda(cudaMallocHost((void **)&m->stage[i], stage_bytes), "cudaMallocHost");
        if (rc == TDK_OK) rc = tdk_check_cuda(cudaEventCreateWithFlags(&m->stage_done[i], cudaEventDisableTiming), "cudaEventCreate");
    }
    for (int i = 0; i < 2 && rc == TDK_OK; i++) {
        rc = tdk_check_cuda(cudaMallocHost((void **)&m->back[i].counts_host, (size_t)(capacity + 2) * 8 + 64), "cudaMallocHost");
        if (rc == TDK_OK) rc = tdk_check_cuda(cudaEventCreateWithFlags(&m->back[i].done, cudaEventDisableTiming), "cudaEventCreate");
        if (rc == TDK_OK) m->back[i].status_host = reinterpret_cast<int *>(m->back[i].counts_host + capacity + 2);
    }
    if (rc != TDK_OK) { tdk_mapper_destroy(m); return rc; }
    *out = m;
    return TDK_OK;
}

void tdk_mapper_destroy(tdk_mapper *m)
{
    if (!m) return;
    cudaStreamSynchronize(m->stream);
    for (int i = 0; i < 2; i++) {
        if (m->stage[i]) cudaFreeHost(m->stage[i]);
        if (m->stage_done[i]) cudaEventDestroy(m->stage_done[i]);
    }
    for (int i = 0; i < 2; i++) {
        if (m->back[i].counts_host) cudaFreeHost(m->back[i].counts_host);
        if (m->back[i].done) cudaEventDestroy(m->back[i].done);
    }
    delete m;
}

int tdk_mapper_reset(tdk_mapper *m, const void *image, int image_format, const double *cam, const double *T_wf,
                     const double *depth_map, const double *variance_map)
{
    if (!m || !image || !cam || !T_wf) { tdk_set_error("tdk_mapper_reset: null pointer argument"); return TDK_ERR_INVALID; }
    TDK_CUDA(cudaStreamSynchronize(m->stream));
    m->start = 0; m->end = 0; m->cur = 0; m->tracked = 0; m->first_id = 0;
    for (int i = 0; i < 2; i++) { m->back[i].census_n = 0; m->back[i].status_pending = false; }
    int rc = upload_frame(m, 0, image, image_format, cam, T_wf, nullptr);
    if (rc != TDK_OK) return rc;
    m->end = 1;
    m->has_frame = true;
    // first frame: age 0 everywhere, flag NotProcessed; depth / variance from the caller or the defaults
    // (examples/semi_dense_vo.py:210-219)
    double *host = (double *)malloc(m->hw * 8);
    if (!host) { tdk_set_error("tdk_mapper_reset: out of host memory"); return TDK_ERR_INVALID; }
    for (int which = 0; which < 2 && rc == TDK_OK; which++) {
        const double *src = which == 0 ? depth_map : variance_map;
        const double fill = which == 0 ? m->default_depth : m->default_variance;
        if (!src) { for (size_t i = 0; i < m->hw; i++) host[i] = fill; src = host; }
        double *dst = at<double>(m, which == 0 ? m->l.depth[0] : m->l.var[0]);
        rc = tdk_check_cuda(cudaMemcpyAsync(dst, src, m->hw * 8, cudaMemcpyHostToDevice, m->stream), "H2D map");
        if (rc == TDK_OK) rc = tdk_check_cuda(cudaStreamSynchronize(m->stream), "cudaStreamSynchronize");
    }
    if (rc == TDK_OK) {
        int64_t *fl = (int64_t *)host;
        for (size_t i = 0; i < m->hw; i++) fl[i] = TDK_FLAG_NOT_PROCESSED;
        rc = tdk_check_cuda(cudaMemcpyAsync(at<int64_t>(m, m->l.flag), fl, m->hw * 8, cudaMemcpyHostToDevice, m->stream), "H2D flag");
        if (rc == TDK_OK) rc = tdk_check_cuda(cudaMemsetAsync(at<uint64_t>(m, m->l.age[0]), 0, m->hw * 8, m->stream), "memset age");
        if (rc == TDK_OK) rc = tdk_check_cuda(cudaStreamSynchronize(m->stream), "cudaStreamSynchronize");
    }
    free(host);
    return rc;
}

int tdk_mapper_track(tdk_mapper *m, const void *image, int image_format, const double *cam, const double *T_wf,
                     const double *T10)
{
    TdkRange range("tdk_mapper_track");
    if (!m || !image || !cam || !T_wf || !T10) { tdk_set_error("tdk_mapper_track: null pointer argument"); return TDK_ERR_INVALID; }
    if (!m->has_frame) { tdk_set_error("tdk_mapper_track: call tdk_mapper_reset with the first frame"); return TDK_ERR_INVALID; }
    tdk_mapper::Back &b = m->back[m->tracked & 1];
    int rc = collect(m, b, true);  // status and census of the frame before the previous one
    if (rc != TDK_OK) return rc;
    if (m->end == m->cap && (rc = compact(m)) != TDK_OK) return rc;
    const int64_t slot = m->end, R = m->end - m->start;
    if ((rc = upload_frame(m, slot, image, image_format, cam, T_wf, T10)) != TDK_OK) return rc;
    double *ring = at<double>(m, m->l.ring), *cams = at<double>(m, m->l.cams), *poses = at<double>(m, m->l.poses);
    const int c = m->cur, n = c ^ 1;
    double *depth0 = at<double>(m, m->l.depth[c]), *var0 = at<double>(m, m->l.var[c]);
    uint64_t *age0 = at<uint64_t>(m, m->l.age[c]), *age1 = at<uint64_t>(m, m->l.age[n]);
    double *pd = at<double>(m, m->l.tmp_d), *pv = at<double>(m, m->l.tmp_v);
    // increment_age + propagate through one warp of the map (examples/semi_dense_vo.py:185-194)
    rc = tdk_propagate_age_dev(at<double>(m, m->l.T10), cams + 4 * (slot - 1), cams + 4 * slot, depth0, var0, age0, 1,
                               m->H, m->W, m->default_depth, m->default_variance, m->bias, pd, pv, age1,
                               m->arena + m->l.ws_prop, m->l.ws_prop_bytes, m->stream);
    if (rc != TDK_OK) return rc;
    double *depth1 = at<double>(m, m->l.depth[n]), *var1 = at<double>(m, m->l.var[n]);
    int64_t *flag = at<int64_t>(m, m->l.flag);
    rc = tdk_update_depth_dev(ring + (size_t)slot * m->hw, cams + 4 * slot, poses + 16 * slot, ring + (size_t)m->start * m->hw,
                              cams + 4 * m->start, poses + 16 * m->start, age1, pd, pv, m->params, 1, R, m->H, m->W,
                              depth1, var1, flag, nullptr, nullptr, m->fuse, m->arena + m->l.ws_update,
                              m->l.ws_update_bytes, m->stream);
    if (rc != TDK_OK) return rc;
    if (m->regularize) {
        rc = tdk_regularize_dev(depth1, var1, flag, 1, m->H, m->W, pd, m->stream);
        if (rc != TDK_OK) return rc;
        const size_t t = m->l.depth[n];   // the smoothed map becomes the depth map; its old place is scratch
        m->l.depth[n] = m->l.tmp_d;
        m->l.tmp_d = t;
    }
    m->cur = n;
    m->end = slot + 1;
    m->tracked++;
    // what the next call wants to know: the status words and, when pruning, the age census
    TDK_CUDA(cudaMemcpyAsync(b.status_host, m->arena + m->l.ws_update, sizeof(int) * TDK_ST_WORDS, cudaMemcpyDeviceToHost, m->stream));
    b.status_pending = true;
    if (m->prune) {
        const int64_t len = m->end - m->start;
        uint64_t *counts = at<uint64_t>(m, m->l.counts);
        rc = tdk_age_usage_dev(at<uint64_t>(m, m->l.age[m->cur]), 1, m->H, m->W, len, counts, m->stream);
        if (rc != TDK_OK) return rc;
        TDK_CUDA(cudaMemcpyAsync(b.counts_host, counts, (size_t)(len + 2) * 8, cudaMemcpyDeviceToHost, m->stream));
        b.census_n = len;
        b.census_first_id = m->first_id;
    }
    TDK_CUDA(cudaEventRecord(b.done, m->stream));
    return TDK_OK;
}

int tdk_mapper_sync(tdk_mapper *m)
{
    if (!m) { tdk_set_error("tdk_mapper_sync: null mapper"); return TDK_ERR_INVALID; }
    TDK_CUDA(cudaStreamSynchronize(m->stream));
    // status words only: a pending census stays pending, pruning is applied by track()
    int rc = collect(m, m->back[m->tracked & 1], false);
    const int rc2 = collect(m, m->back[(m->tracked & 1) ^ 1], false);
    return rc != TDK_OK ? rc : rc2;
}

int tdk_mapper_prune(tdk_mapper *m)
{
    if (!m) { tdk_set_error("tdk_mapper_prune: null mapper"); return TDK_ERR_INVALID; }
    // whatever is still on its way back is older than the census taken now: apply it first
    int rc = collect(m, m->back[m->tracked & 1], true);
    const int rc2 = collect(m, m->back[(m->tracked & 1) ^ 1], true);
    if (rc == TDK_OK) rc = rc2;
    if (rc != TDK_OK) return rc;
    tdk_mapper::Back &b = m->back[0];
    const int64_t len = m->end - m->start;
    uint64_t *counts = at<uint64_t>(m, m->l.counts);
    rc = tdk_age_usage_dev(at<uint64_t>(m, m->l.age[m->cur]), 1, m->H, m->W, len, counts, m->stream);
    if (rc != TDK_OK) return rc;
    TDK_CUDA(cudaMemcpyAsync(b.counts_host, counts, (size_t)(len + 2) * 8, cudaMemcpyDeviceToHost, m->stream));
    b.census_n = len;
    b.census_first_id = m->first_id;
    TDK_CUDA(cudaEventRecord(b.done, m->stream));
    return collect(m, b, true);
}

int tdk_mapper_state(tdk_mapper *m, int64_t *n_frames, size_t *depth_off, size_t *variance_off, size_t *age_off,
                     size_t *flag_off, size_t *ring_off, int64_t *ring_start)
{
    if (!m) { tdk_set_error("tdk_mapper_state: null mapper"); return TDK_ERR_INVALID; }
    if (n_frames) *n_frames = m->end - m->start;
    if (depth_off) *depth_off = m->l.depth[m->cur];
    if (variance_off) *variance_off = m->l.var[m->cur];
    if (age_off) *age_off = m->l.age[m->cur];
    if (flag_off) *flag_off = m->l.flag;
    if (ring_off) *ring_off = m->l.ring;
    if (ring_start) *ring_start = m->start;
    return TDK_OK;
}

}  // extern "C"
