// Host-side plumbing of the C-ABI: error text, version, and the host-pointer entry points
// (copy in -> launch -> copy out), which stage through a grow-only per-device arena.  The batched
// update_depth forms pipeline chunks of keyframes over several streams so that H2D, kernels and D2H
// overlap; a small pool of host threads narrows / widens the byte-wide maps and moves pageable
// caller memory through pinned rings.
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <functional>
#include <mutex>
#include <thread>
#include <unordered_map>
#include <vector>

#include "tdk_common.cuh"

static thread_local char g_err[512] = "";

void tdk_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int tdk_check_cuda(cudaError_t e, const char *what)
{
    if (e == cudaSuccess) return TDK_OK;
    tdk_set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
    return TDK_ERR_CUDA;
}

namespace {

// Chunking of the host-pointer entry points (measured on the C2 step, PCIe 5 x16):
//  tdk_update_depth_host         -- H2D dominated (every keyframe ships 8 references): 96 MB chunks in 3 slots
//                                   (36.0 ms per step; 38.2 with 24-48 MB chunks in 6-8 slots)
//  tdk_update_depth_frames_host  -- both directions busy: 32 MB chunks (two VGA keyframes, 4.9 MB per copy) in 6
//                                   slots (13.3 ms; 13.9 with one keyframe or with four per chunk, 13.9 in 4 slots;
//                                   tools/pcie_copy_size.py: 81 GB/s both ways at 2.5 MB per copy, 90 from 4.9 MB on)
#ifndef TDK_FRAMES_SLOTS
#define TDK_FRAMES_SLOTS 6
#endif
#ifndef TDK_FRAMES_CHUNK_MB
#define TDK_FRAMES_CHUNK_MB 32
#endif
constexpr int NSTREAM = TDK_FRAMES_SLOTS > 6 ? TDK_FRAMES_SLOTS : 6;
constexpr int HOST_SLOTS = 3;
constexpr size_t HOST_CHUNK_BYTES = (size_t)96 << 20;
constexpr int FRAMES_SLOTS = TDK_FRAMES_SLOTS;
constexpr size_t FRAMES_CHUNK_BYTES = (size_t)TDK_FRAMES_CHUNK_MB << 20;
constexpr int MAX_DEV = 64;

struct Arena {
    void *base = nullptr;
    size_t bytes = 0;
    cudaStream_t streams[NSTREAM] = {};
    cudaStream_t upload = nullptr;   // images of tdk_update_depth_frames_host: never queued behind a chunk's read-back
    bool streams_ready = false;
    int *status_host = nullptr;      // pinned, grow-only: the status words of every chunk of a call
    size_t status_words = 0;
    std::vector<cudaEvent_t> events; // grow-only pool (tdk_update_depth_frames_host: three per chunk)
    unsigned char *wire_host = nullptr;   // pinned, grow-only: 8-bit age maps going up, 8-bit flag maps coming back
    size_t wire_bytes = 0;
    unsigned char *stage_host = nullptr;  // pinned ring of the pageable-memory path (struct Stage): 2 x 32 slots of 512 KB
    cudaEvent_t stage_ev[64] = {};
};

// ---- host workers of the narrow-wire path (tdk_update_depth_frames_host)
// The reference's layout spends 8 bytes per pixel on an age (<= R) and 8 on a flag (-9 .. 0).  Over
// PCIe they travel as one byte each: worker threads narrow the caller's age maps into pinned memory
// ahead of the uploads and widen the flag maps into the caller's array behind the read-backs; two
// small kernels do the opposite on the device.  The pool is created on first use and never
// destroyed (its threads sleep on a condition variable; a heap object, so no static destructor
// runs under them).  TDK_HOST_THREADS sets its size (default: min(6, hardware threads): 3-6 serve one
// numpy call best, 8 contend on the page faults of fresh output arrays; the batched call is indifferent from 4 up).
class HostPool {
public:
    static HostPool &get()
    {
        // one pool per process: a child of fork() has none of the parent's threads, so it makes its own
        static std::mutex guard;
        static HostPool *pool = nullptr;
        static pid_t owner = 0;
        std::lock_guard<std::mutex> lk(guard);
        if (!pool || owner != getpid()) {
            pool = new HostPool;
            owner = getpid();
        }
        return *pool;
    }
    void submit(std::function<void()> fn)
    {
        {
            std::lock_guard<std::mutex> lk(m_);
            q_.push_back(std::move(fn));
            queued_.fetch_add(1, std::memory_order_release);
        }
        cv_.notify_one();
    }

private:
    HostPool()
    {
        int n = 0;
        if (const char *e = getenv("TDK_HOST_THREADS")) n = atoi(e);
        if (n <= 0) {
            n = (int)std::thread::hardware_concurrency();
            if (n > 6) n = 6;
        }
        if (n < 1) n = 1;
        for (int i = 0; i < n; i++) std::thread([this] { loop(); }).detach();
    }
    void loop()
    {
        for (;;) {
            std::function<void()> fn;
            // stay hot for a moment after the last task: pieces of one call arrive microseconds apart
            for (int spins = 0; queued_.load(std::memory_order_acquire) == 0 && spins < 5000; spins++) {
#if defined(__x86_64__) || defined(__i386__)
                __builtin_ia32_pause();
#endif
            }
            {
                std::unique_lock<std::mutex> lk(m_);
                cv_.wait(lk, [this] { return !q_.empty(); });
                fn = std::move(q_.front());
                q_.pop_front();
                queued_.fetch_sub(1, std::memory_order_relaxed);
            }
            fn();
        }
    }
    std::atomic<int> queued_{0};
    std::mutex m_;
    std::condition_variable cv_;
    std::deque<std::function<void()>> q_;
};

// what one call's chunks wait on: "chunk c's ages are narrowed", "chunk c's flags are in the caller's array"
struct ChunkSync {
    std::mutex m;
    std::condition_variable cv;
    std::vector<int> narrow_left, widen_left;   // pieces of chunk c still with the workers; -1 = not handed out yet
    int64_t pending = 0;                        // pieces handed to the pool and not finished
    explicit ChunkSync(int64_t n) : narrow_left((size_t)n, -1), widen_left((size_t)n, -1) {}
    void begin(std::vector<int> &left, int64_t c, int pieces)
    {
        std::lock_guard<std::mutex> lk(m);
        left[(size_t)c] = pieces;
        pending += pieces;
    }
    void end(std::vector<int> &left, int64_t c)
    {
        {
            std::lock_guard<std::mutex> lk(m);
            left[(size_t)c]--;
            pending--;
        }
        cv.notify_all();
    }
    void wait(const std::vector<int> &left, int64_t c)
    {
        std::unique_lock<std::mutex> lk(m);
        cv.wait(lk, [&] { return left[(size_t)c] == 0; });
    }
    void drain()
    {
        std::unique_lock<std::mutex> lk(m);
        cv.wait(lk, [&] { return pending == 0; });
    }
};

constexpr size_t HOST_PIECE = 76800;   // pixels per worker task: a chunk's map is converted by several threads

#ifdef TDK_HOST_TIMING
double host_now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
#endif

void narrow_age_host(const uint64_t *src, unsigned char *dst, size_t n)
{
    for (size_t i = 0; i < n; i++) dst[i] = src[i] > 255u ? (unsigned char)255 : (unsigned char)src[i];
}

void widen_flag_host(const signed char *src, int64_t *dst, size_t n)
{
    for (size_t i = 0; i < n; i++) dst[i] = src[i];
}

// device side of the narrow wire; eight values per thread, the tail one by one
__global__ void __launch_bounds__(256) widen_age_kernel(const unsigned char *__restrict__ in, uint64_t *__restrict__ out,
                                                         size_t n)
{
    const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
    if (i + 8 <= n) {
        const uint2 v = *reinterpret_cast<const uint2 *>(in + i);
        ulonglong2 *o = reinterpret_cast<ulonglong2 *>(out + i);
        o[0] = make_ulonglong2(v.x & 255u, (v.x >> 8) & 255u);
        o[1] = make_ulonglong2((v.x >> 16) & 255u, v.x >> 24);
        o[2] = make_ulonglong2(v.y & 255u, (v.y >> 8) & 255u);
        o[3] = make_ulonglong2((v.y >> 16) & 255u, v.y >> 24);
    } else {
        for (size_t j = i; j < n; j++) out[j] = in[j];
    }
}

__global__ void __launch_bounds__(256) narrow_flag_kernel(const int64_t *__restrict__ in, signed char *__restrict__ out,
                                                           size_t n, const int *__restrict__ status_src,
                                                           int *__restrict__ status_dst)
{
    // the chunk's status words move to the call's table on the way (one read-back per call, not per chunk)
    if (blockIdx.x == 0 && threadIdx.x < TDK_ST_WORDS) status_dst[threadIdx.x] = status_src[threadIdx.x];
    const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
    if (i + 8 <= n) {
        const longlong2 *q = reinterpret_cast<const longlong2 *>(in + i);
        const longlong2 a = q[0], b = q[1], c = q[2], d = q[3];
        uint2 v;
        v.x = ((unsigned)a.x & 255u) | (((unsigned)a.y & 255u) << 8) | (((unsigned)b.x & 255u) << 16) | (((unsigned)b.y & 255u) << 24);
        v.y = ((unsigned)c.x & 255u) | (((unsigned)c.y & 255u) << 8) | (((unsigned)d.x & 255u) << 16) | (((unsigned)d.y & 255u) << 24);
        *reinterpret_cast<uint2 *>(out + i) = v;
    } else {
        for (size_t j = i; j < n; j++) out[j] = (signed char)in[j];
    }
}

int wire_host_get(Arena *ar, size_t bytes)
{
    if (ar->wire_bytes < bytes) {
        if (ar->wire_host) { TDK_CUDA(cudaFreeHost(ar->wire_host)); ar->wire_host = nullptr; ar->wire_bytes = 0; }
        TDK_CUDA(cudaMallocHost((void **)&ar->wire_host, bytes));
        ar->wire_bytes = bytes;
    }
    return TDK_OK;
}

int status_host_get(Arena *ar, int64_t nchunks)
{
    if (ar->status_words < (size_t)(TDK_ST_WORDS * nchunks)) {
        if (ar->status_host) { TDK_CUDA(cudaFreeHost(ar->status_host)); ar->status_host = nullptr; ar->status_words = 0; }
        const size_t want = (size_t)TDK_ST_WORDS * (nchunks < 64 ? 64 : nchunks);
        TDK_CUDA(cudaMallocHost((void **)&ar->status_host, sizeof(int) * want));
        ar->status_words = want;
    }
    return TDK_OK;
}

int status_report(const int (*st)[TDK_ST_WORDS], int64_t nchunks)
{
    for (int64_t c = 0; c < nchunks; c++) {
        if (st[c][TDK_ST_SINGULAR]) { tdk_set_error("a reference pose T_wr is singular"); return TDK_ERR_SINGULAR; }
        if (st[c][TDK_ST_AGE]) { tdk_set_error("Age exceeds the refframe size"); return TDK_ERR_AGE; }
        if (st[c][TDK_ST_SEARCH_LEN]) {
            tdk_set_error("an epipolar search line has more than %d samples; increase ref_step_size", TDK_MAX_SEARCH_SAMPLES);
            return TDK_ERR_SEARCH_LEN;
        }
    }
    return TDK_OK;
}
Arena g_arena[MAX_DEV];
std::mutex g_mutex[MAX_DEV];   // one per device: host-pointer calls on different GPUs do not serialise

int current_device(int *dev)
{
    TDK_CUDA(cudaGetDevice(dev));
    if (*dev < 0 || *dev >= MAX_DEV) { tdk_set_error("device index %d out of range", *dev); return TDK_ERR_INVALID; }
    return TDK_OK;
}

size_t up256(size_t v) { return (v + 255) / 256 * 256; }

int arena_get(size_t bytes, Arena **out)
{
    int dev = 0;
    TDK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEV) { tdk_set_error("device index %d out of range", dev); return TDK_ERR_INVALID; }
    Arena &ar = g_arena[dev];
    if (!ar.streams_ready) {
        for (int i = 0; i < NSTREAM; i++) TDK_CUDA(cudaStreamCreateWithFlags(&ar.streams[i], cudaStreamNonBlocking));
        TDK_CUDA(cudaStreamCreateWithFlags(&ar.upload, cudaStreamNonBlocking));
        ar.streams_ready = true;
    }
    if (ar.bytes < bytes) {
        if (ar.base) { TDK_CUDA(cudaDeviceSynchronize()); TDK_CUDA(cudaFree(ar.base)); ar.base = nullptr; ar.bytes = 0; }
        TDK_CUDA(cudaMalloc(&ar.base, bytes));
        ar.bytes = bytes;
    }
    *out = &ar;
    return TDK_OK;
}

// per-chunk device buffers for the host path
struct Slot {
    double *key_imgs, *key_cams, *T_wk, *ref_imgs, *ref_cams, *T_wr, *pd, *pv, *od, *ov;
    uint64_t *age;
    int64_t *of;
    int32_t *dbg;
    uint64_t *hist;
    void *ws;
    size_t ws_bytes;
};

size_t carve_slot(char *base, size_t off, int64_t C, int64_t R, int64_t H, int64_t W, bool dbg, bool hist, Slot *s)
{
    const size_t hw = (size_t)H * W;
    auto take = [&](size_t bytes) { char *p = base ? base + off : nullptr; off += up256(bytes); return p; };
    Slot t;
    t.key_imgs = (double *)take(C * hw * 8);
    t.key_cams = (double *)take(C * 4 * 8);
    t.T_wk = (double *)take(C * 16 * 8);
    t.ref_imgs = (double *)take(C * R * hw * 8);
    t.ref_cams = (double *)take(C * R * 4 * 8);
    t.T_wr = (double *)take(C * R * 16 * 8);
    t.age = (uint64_t *)take(C * hw * 8);
    t.pd = (double *)take(C * hw * 8);
    t.pv = (double *)take(C * hw * 8);
    t.od = (double *)take(C * hw * 8);
    t.ov = (double *)take(C * hw * 8);
    t.of = (int64_t *)take(C * hw * 8);
    t.dbg = dbg ? (int32_t *)take(C * hw * 4) : nullptr;
    t.hist = hist ? (uint64_t *)take(C * 10 * 8) : nullptr;
    t.ws_bytes = tdk_update_depth_workspace_bytes(C, R, H, W);
    t.ws = take(t.ws_bytes);
    if (s) *s = t;
    return off;
}

}  // namespace

extern "C" {

const char *tdk_last_error(void) { return g_err; }

const char *tdk_version(void) { return "tadataka-b200 0.1.0 sm_100a"; }

int tdk_update_depth_host(const double *key_imgs, const double *key_cams, const double *T_wk,
                          const double *ref_imgs, const double *ref_cams, const double *T_wr,
                          const uint64_t *age, const double *prior_depth, const double *prior_variance,
                          const double *params, int64_t P, int64_t R, int64_t H, int64_t W,
                          double *out_depth, double *out_variance, int64_t *out_flag,
                          int32_t *dbg_argmin, uint64_t *out_hist, int fuse_on_success)
{
    TdkRange range("tdk_update_depth_host");
    if (!key_imgs || !ref_imgs || !key_cams || !T_wk || !ref_cams || !T_wr || !age || !prior_depth ||
        !prior_variance || !params || !out_depth || !out_variance || !out_flag) {
        tdk_set_error("tdk_update_depth_host: null pointer argument");
        return TDK_ERR_INVALID;
    }
    if (P <= 0 || R <= 0 || H <= 0 || W <= 0) {
        tdk_set_error("tdk_update_depth_host: bad sizes");
        return TDK_ERR_INVALID;
    }
    int dev = 0;
    int rc = current_device(&dev);
    if (rc != TDK_OK) return rc;
    std::lock_guard<std::mutex> lock(g_mutex[dev]);
    const size_t hw = (size_t)H * W;
    // chunk = as many keyframes as fit ~96 MB of staged input, so copies and kernels overlap
    const size_t in_per_key = (size_t)(1 + R + 3) * hw * 8;
    int64_t C = (int64_t)(HOST_CHUNK_BYTES / in_per_key);
    if (C < 1) C = 1;
    if (C > P) C = P;
    const int nslots = (int)((P + C - 1) / C < HOST_SLOTS ? (P + C - 1) / C : HOST_SLOTS);
    const size_t slot_bytes = carve_slot(nullptr, 0, C, R, H, W, dbg_argmin != nullptr, out_hist != nullptr, nullptr);
    Arena *ar = nullptr;
    rc = arena_get(slot_bytes * nslots, &ar);
    if (rc != TDK_OK) return rc;
    Slot slots[NSTREAM];
    for (int i = 0; i < nslots; i++)
        carve_slot((char *)ar->base, slot_bytes * i, C, R, H, W, dbg_argmin != nullptr, out_hist != nullptr, &slots[i]);

    const int64_t nchunks = (P + C - 1) / C;
    // pinned landing place for the chunks' status words, kept between calls (a cudaMallocHost per
    // call costs a device-wide synchronisation)
    if (ar->status_words < (size_t)(TDK_ST_WORDS * nchunks)) {
        if (ar->status_host) { TDK_CUDA(cudaFreeHost(ar->status_host)); ar->status_host = nullptr; ar->status_words = 0; }
        const size_t want = (size_t)TDK_ST_WORDS * (nchunks < 64 ? 64 : nchunks);
        TDK_CUDA(cudaMallocHost((void **)&ar->status_host, sizeof(int) * want));
        ar->status_words = want;
    }
    int(*st_host)[TDK_ST_WORDS] = reinterpret_cast<int(*)[TDK_ST_WORDS]>(ar->status_host);
    rc = TDK_OK;
    for (int64_t c = 0; c < nchunks && rc == TDK_OK; c++) {
        const int si = (int)(c % nslots);
        Slot &s = slots[si];
        cudaStream_t st = ar->streams[si];
        const int64_t p0 = c * C, n = (p0 + C <= P) ? C : P - p0;
#define H2D(dst, src, bytes) \
    if (rc == TDK_OK) rc = tdk_check_cuda(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st), "H2D " #src)
#define D2H(dst, src, bytes) \
    if (rc == TDK_OK) rc = tdk_check_cuda(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, st), "D2H " #src)
        H2D(s.key_imgs, key_imgs + p0 * hw, n * hw * 8);
        H2D(s.key_cams, key_cams + p0 * 4, n * 4 * 8);
        H2D(s.T_wk, T_wk + p0 * 16, n * 16 * 8);
        H2D(s.ref_imgs, ref_imgs + p0 * R * hw, n * R * hw * 8);
        H2D(s.ref_cams, ref_cams + p0 * R * 4, n * R * 4 * 8);
        H2D(s.T_wr, T_wr + p0 * R * 16, n * R * 16 * 8);
        H2D(s.age, age + p0 * hw, n * hw * 8);
        H2D(s.pd, prior_depth + p0 * hw, n * hw * 8);
        H2D(s.pv, prior_variance + p0 * hw, n * hw * 8);
        if (s.hist && rc == TDK_OK) rc = tdk_check_cuda(cudaMemsetAsync(s.hist, 0, n * 10 * 8, st), "memset hist");
        if (rc == TDK_OK)
            rc = tdk_update_depth_dev(s.key_imgs, s.key_cams, s.T_wk, s.ref_imgs, s.ref_cams, s.T_wr, s.age,
                                      s.pd, s.pv, params, n, R, H, W, s.od, s.ov, s.of, s.dbg, s.hist,
                                      fuse_on_success, s.ws, s.ws_bytes, st);
        D2H(out_depth + p0 * hw, s.od, n * hw * 8);
        D2H(out_variance + p0 * hw, s.ov, n * hw * 8);
        D2H(out_flag + p0 * hw, s.of, n * hw * 8);
        if (dbg_argmin) D2H(dbg_argmin + p0 * hw, s.dbg, n * hw * 4);
        if (out_hist) D2H(out_hist + p0 * 10, s.hist, n * 10 * 8);
        D2H(st_host[c], s.ws, sizeof(int) * TDK_ST_WORDS);
#undef H2D
#undef D2H
    }
    for (int i = 0; i < nslots; i++) {
        const int rc2 = tdk_check_cuda(cudaStreamSynchronize(ar->streams[i]), "cudaStreamSynchronize");
        if (rc == TDK_OK) rc = rc2;
    }
    if (rc == TDK_OK) {
        for (int64_t c = 0; c < nchunks; c++) {
            if (st_host[c][TDK_ST_SINGULAR]) { tdk_set_error("a reference pose T_wr is singular"); rc = TDK_ERR_SINGULAR; break; }
            if (st_host[c][TDK_ST_AGE]) { tdk_set_error("Age exceeds the refframe size"); rc = TDK_ERR_AGE; break; }
            if (st_host[c][TDK_ST_SEARCH_LEN]) {
                tdk_set_error("an epipolar search line has more than %d samples; increase ref_step_size",
                              TDK_MAX_SEARCH_SAMPLES);
                rc = TDK_ERR_SEARCH_LEN;
                break;
            }
        }
    }
    return rc;
}

// update_depth from host memory with frames passed by POINTER: an image that several keyframes
// name -- the reference frames of the reference's own loop, examples/semi_dense_vo.py:196-203, where
// every keyframe is matched against the frames before it -- crosses PCIe once.  Chunks of keyframes
// rotate over FRAMES_SLOTS streams; the images a chunk needs that are not on the device yet go up on
// a stream of their own just before it, so image upload, kernels and the copy back of earlier chunks
// overlap.  Cameras, poses and pointer tables go up once; ages and flags travel as bytes (R < 255).
int tdk_update_depth_frames_host(const double *const *key_img_ptrs, const double *key_cams, const double *T_wk,
                                 const double *const *ref_img_ptrs, const double *ref_cams, const double *T_wr,
                                 const uint64_t *age, const double *prior_depth, const double *prior_variance,
                                 const double *params, int64_t P, int64_t R, int64_t H, int64_t W, double *out_depth,
                                 double *out_variance, int64_t *out_flag, int32_t *dbg_argmin, uint64_t *out_hist,
                                 int fuse_on_success)
{
    TdkRange range("tdk_update_depth_frames_host");
    if (!key_img_ptrs || !ref_img_ptrs || !key_cams || !T_wk || !ref_cams || !T_wr || !age || !prior_depth ||
        !prior_variance || !params || !out_depth || !out_variance || !out_flag) {
        tdk_set_error("tdk_update_depth_frames_host: null pointer argument");
        return TDK_ERR_INVALID;
    }
    if (P <= 0 || R <= 0 || H <= 0 || W <= 0) {
        tdk_set_error("tdk_update_depth_frames_host: bad sizes");
        return TDK_ERR_INVALID;
    }
    int dev = 0;
    int rc = current_device(&dev);
    if (rc != TDK_OK) return rc;
    std::lock_guard<std::mutex> lock(g_mutex[dev]);
    const size_t hw = (size_t)H * W;
    // distinct images, in the order the keyframes need them
    std::unordered_map<const double *, int64_t> index;
    std::vector<const double *> unique;
    auto slot_of = [&](const double *img) {
        auto it = index.find(img);
        if (it != index.end()) return it->second;
        const int64_t k = (int64_t)unique.size();
        index.emplace(img, k);
        unique.push_back(img);
        return k;
    };
    std::vector<int64_t> key_slot((size_t)P), ref_slot((size_t)(P * R));
    for (int64_t p = 0; p < P; p++) {
        if (!key_img_ptrs[p]) { tdk_set_error("tdk_update_depth_frames_host: null image pointer"); return TDK_ERR_INVALID; }
        key_slot[p] = slot_of(key_img_ptrs[p]);
        for (int64_t r = 0; r < R; r++) {
            if (!ref_img_ptrs[p * R + r]) { tdk_set_error("tdk_update_depth_frames_host: null image pointer"); return TDK_ERR_INVALID; }
            ref_slot[p * R + r] = slot_of(ref_img_ptrs[p * R + r]);
        }
    }
    const int64_t U = (int64_t)unique.size();
    // chunk = keyframes whose maps (3 in, 3 out) make up ~32 MB
    int64_t C = (int64_t)(FRAMES_CHUNK_BYTES / (6 * hw * 8));
    if (C < 1) C = 1;
    if (C > P) C = P;
    const int64_t nchunks = (P + C - 1) / C;
    const int nslots = (int)(nchunks < FRAMES_SLOTS ? nchunks : FRAMES_SLOTS);
    // ages (<= R, anything larger is an error whatever its value) and flags (-9 .. 0) cross the link as bytes
    const bool narrow = R < 255;
    struct FSlot { double *pd, *pv, *od, *ov; uint64_t *age; int64_t *of; unsigned char *age8; signed char *of8; int32_t *dbg; uint64_t *hist; void *ws; size_t ws_bytes; };
    auto carve = [&](char *base, size_t off, FSlot *f) {
        auto take = [&](size_t bytes) { char *q = base ? base + off : nullptr; off += up256(bytes); return q; };
        FSlot t;
        t.age = (uint64_t *)take(C * hw * 8); t.pd = (double *)take(C * hw * 8); t.pv = (double *)take(C * hw * 8);
        t.od = (double *)take(C * hw * 8); t.ov = (double *)take(C * hw * 8); t.of = (int64_t *)take(C * hw * 8);
        t.age8 = narrow ? (unsigned char *)take(C * hw) : nullptr;
        t.of8 = narrow ? (signed char *)take(C * hw) : nullptr;
        t.dbg = dbg_argmin ? (int32_t *)take(C * hw * 4) : nullptr;
        t.hist = out_hist ? (uint64_t *)take(C * 10 * 8) : nullptr;
        t.ws_bytes = tdk_update_depth_workspace_bytes(C, R, H, W);
        t.ws = take(t.ws_bytes);
        if (f) *f = t;
        return off;
    };
    const size_t img_bytes = up256((size_t)U * hw * 8);
    const size_t slot_bytes = carve(nullptr, 0, nullptr);
    // what the whole call needs once: cameras, poses, the two tables of device image pointers, a status row per chunk
    struct Tables { double *cams, *T_wk, *ref_cams, *T_wr; const double **kptr, **rptr; int *status; } tab;
    auto carve_tables = [&](char *base, size_t off) {
        auto take = [&](size_t bytes) { char *q = base ? base + off : nullptr; off += up256(bytes); return q; };
        tab.cams = (double *)take(P * 4 * 8); tab.T_wk = (double *)take(P * 16 * 8);
        tab.ref_cams = (double *)take(P * R * 4 * 8); tab.T_wr = (double *)take(P * R * 16 * 8);
        tab.kptr = (const double **)take(P * sizeof(void *)); tab.rptr = (const double **)take(P * R * sizeof(void *));
        tab.status = (int *)take(nchunks * TDK_ST_WORDS * sizeof(int));
        return off;
    };
    const size_t tab_bytes = carve_tables(nullptr, 0);
    Arena *ar = nullptr;
    rc = arena_get(img_bytes + tab_bytes + slot_bytes * nslots, &ar);
    if (rc != TDK_OK) return rc;
    if ((rc = status_host_get(ar, nchunks)) != TDK_OK) return rc;
    // events per chunk: [c] its images are up, [nchunks + c] its ages are up, [2 nchunks + c] its results are down
    while ((int64_t)ar->events.size() < 3 * nchunks) {
        cudaEvent_t e;
        TDK_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        ar->events.push_back(e);
    }
    cudaEvent_t *ev_img = ar->events.data(), *ev_in = ev_img + nchunks, *ev_out = ev_in + nchunks;
    // pinned byte maps: a ring of RING chunks each way; a slot is taken again only when its last user is through
    const int64_t RING = nchunks < 64 ? nchunks : 64;
    const int64_t AHEAD = RING < 16 ? RING : 16;           // chunks the narrowing runs ahead of the uploads
    const size_t wire_chunk = up256((size_t)C * hw);
    unsigned char *age_wire = nullptr;
    signed char *flag_wire = nullptr;
    if (narrow) {
        if ((rc = wire_host_get(ar, 2 * wire_chunk * (size_t)RING)) != TDK_OK) return rc;
        age_wire = ar->wire_host;
        flag_wire = (signed char *)(ar->wire_host + wire_chunk * (size_t)RING);
    }
    double *images = (double *)ar->base;
    FSlot slots[NSTREAM];
    carve_tables((char *)ar->base, img_bytes);
    for (int i = 0; i < nslots; i++) carve((char *)ar->base, img_bytes + tab_bytes + slot_bytes * i, &slots[i]);
    int(*st_host)[TDK_ST_WORDS] = reinterpret_cast<int(*)[TDK_ST_WORDS]>(ar->status_host);
    std::vector<char> uploaded((size_t)U, 0);
    std::vector<const double *> kp((size_t)P), rp((size_t)(P * R));   // device addresses of every keyframe's images
    for (int64_t i = 0; i < P; i++) kp[i] = images + (size_t)key_slot[i] * hw;
    for (int64_t i = 0; i < P * R; i++) rp[i] = images + (size_t)ref_slot[i] * hw;
    cudaStream_t up_stream = ar->upload;
#ifdef TDK_HOST_TIMING
    const double t_begin = host_now();
#endif
    ChunkSync sync(nchunks);
    auto chunk_len = [&](int64_t c) { return (c * C + C <= P) ? C : P - c * C; };
    rc = TDK_OK;
#define GO(call, what) if (rc == TDK_OK) rc = tdk_check_cuda((call), what)
    // the upload stream carries the tables first, then the images in the order the chunks need them
    GO(cudaMemcpyAsync(tab.cams, key_cams, P * 4 * 8, cudaMemcpyHostToDevice, up_stream), "H2D key_cams");
    GO(cudaMemcpyAsync(tab.T_wk, T_wk, P * 16 * 8, cudaMemcpyHostToDevice, up_stream), "H2D T_wk");
    GO(cudaMemcpyAsync(tab.ref_cams, ref_cams, P * R * 4 * 8, cudaMemcpyHostToDevice, up_stream), "H2D ref_cams");
    GO(cudaMemcpyAsync(tab.T_wr, T_wr, P * R * 16 * 8, cudaMemcpyHostToDevice, up_stream), "H2D T_wr");
    GO(cudaMemcpyAsync(tab.kptr, kp.data(), P * sizeof(void *), cudaMemcpyHostToDevice, up_stream), "H2D key pointers");
    GO(cudaMemcpyAsync(tab.rptr, rp.data(), P * R * sizeof(void *), cudaMemcpyHostToDevice, up_stream), "H2D ref pointers");
    int64_t next_narrow = 0, next_widen = 0;
    // hand chunk k's age maps to the workers (its ring slot is free once chunk k - RING has been uploaded)
    auto narrow_upto = [&](int64_t upto) {
        for (; narrow && next_narrow < nchunks && next_narrow < upto && rc == TDK_OK; next_narrow++) {
            const int64_t k = next_narrow;
            if (k >= RING) GO(cudaEventSynchronize(ev_in[k - RING]), "cudaEventSynchronize");
            if (rc != TDK_OK) break;
            const uint64_t *src = age + (size_t)k * C * hw;
            unsigned char *dst = age_wire + wire_chunk * (size_t)(k % RING);
            const size_t cnt = (size_t)chunk_len(k) * hw;
            const int pieces = (int)((cnt + HOST_PIECE - 1) / HOST_PIECE);
            sync.begin(sync.narrow_left, k, pieces);
            for (size_t o = 0; o < cnt; o += HOST_PIECE) {
                const size_t len = cnt - o < HOST_PIECE ? cnt - o : HOST_PIECE;
                HostPool::get().submit([src, dst, o, len, k, &sync] { narrow_age_host(src + o, dst + o, len); sync.end(sync.narrow_left, k); });
            }
        }
    };
    // wait for chunk k's read-back and hand its flag map to the workers
    auto widen_upto = [&](int64_t upto) {
        for (; narrow && next_widen < upto && rc == TDK_OK; next_widen++) {
            const int64_t k = next_widen;
            GO(cudaEventSynchronize(ev_out[k]), "cudaEventSynchronize");
            if (rc != TDK_OK) break;
            const signed char *src = flag_wire + wire_chunk * (size_t)(k % RING);
            int64_t *dst = out_flag + (size_t)k * C * hw;
            const size_t cnt = (size_t)chunk_len(k) * hw;
            const int pieces = (int)((cnt + HOST_PIECE - 1) / HOST_PIECE);
            sync.begin(sync.widen_left, k, pieces);
            for (size_t o = 0; o < cnt; o += HOST_PIECE) {
                const size_t len = cnt - o < HOST_PIECE ? cnt - o : HOST_PIECE;
                HostPool::get().submit([src, dst, o, len, k, &sync] { widen_flag_host(src + o, dst + o, len); sync.end(sync.widen_left, k); });
            }
        }
    };
    for (int64_t c = 0; c < nchunks && rc == TDK_OK; c++) {
        const int64_t p0 = c * C, n = chunk_len(c);
        FSlot &f = slots[c % nslots];
        cudaStream_t st = ar->streams[c % nslots];
        if (narrow && c >= RING) {   // the flag ring comes round: chunk c - RING must have left its slot
            widen_upto(c - RING + 1);
            if (rc == TDK_OK) sync.wait(sync.widen_left, c - RING);
        }
        narrow_upto(c + AHEAD + 1);
        auto need = [&](int64_t u) {
            if (!uploaded[u]) {
                GO(cudaMemcpyAsync(images + (size_t)u * hw, unique[u], hw * 8, cudaMemcpyHostToDevice, up_stream), "H2D image");
                uploaded[u] = 1;
            }
        };
        for (int64_t i = 0; i < n; i++) {
            need(key_slot[p0 + i]);
            for (int64_t r = 0; r < R; r++) need(ref_slot[(p0 + i) * R + r]);
        }
        GO(cudaEventRecord(ev_img[c], up_stream), "cudaEventRecord");
        GO(cudaStreamWaitEvent(st, ev_img[c], 0), "cudaStreamWaitEvent");
        const size_t cnt = (size_t)n * hw;
        const unsigned blocks = (unsigned)((cnt + 2047) / 2048);
        // the two f64 maps first: the widening kernel then runs while they are still on the way in other streams
        GO(cudaMemcpyAsync(f.pd, prior_depth + p0 * hw, cnt * 8, cudaMemcpyHostToDevice, st), "H2D prior_depth");
        GO(cudaMemcpyAsync(f.pv, prior_variance + p0 * hw, cnt * 8, cudaMemcpyHostToDevice, st), "H2D prior_variance");
        if (narrow) {
            if (rc == TDK_OK) sync.wait(sync.narrow_left, c);
            GO(cudaMemcpyAsync(f.age8, age_wire + wire_chunk * (size_t)(c % RING), cnt, cudaMemcpyHostToDevice, st), "H2D age");
            GO(cudaEventRecord(ev_in[c], st), "cudaEventRecord");
            if (rc == TDK_OK) {
                widen_age_kernel<<<blocks, 256, 0, st>>>(f.age8, f.age, cnt);
                rc = tdk_check_cuda(cudaGetLastError(), "widen_age_kernel");
            }
        } else {
            GO(cudaMemcpyAsync(f.age, age + p0 * hw, cnt * 8, cudaMemcpyHostToDevice, st), "H2D age");
        }
        if (f.hist) GO(cudaMemsetAsync(f.hist, 0, n * 10 * 8, st), "memset hist");
        if (rc == TDK_OK)
            rc = tdk_update_depth_tables_dev(tab.kptr + p0, tab.cams + p0 * 4, tab.T_wk + p0 * 16, tab.rptr + p0 * R,
                                             tab.ref_cams + p0 * R * 4, tab.T_wr + p0 * R * 16, f.age, f.pd, f.pv,
                                             params, n, R, H, W, f.od, f.ov, f.of, f.dbg, f.hist, fuse_on_success, f.ws,
                                             f.ws_bytes, st);
        if (narrow && rc == TDK_OK) {
            narrow_flag_kernel<<<blocks, 256, 0, st>>>(f.of, f.of8, cnt, (const int *)f.ws, tab.status + c * TDK_ST_WORDS);
            rc = tdk_check_cuda(cudaGetLastError(), "narrow_flag_kernel");
        }
        GO(cudaMemcpyAsync(out_depth + p0 * hw, f.od, cnt * 8, cudaMemcpyDeviceToHost, st), "D2H depth");
        GO(cudaMemcpyAsync(out_variance + p0 * hw, f.ov, cnt * 8, cudaMemcpyDeviceToHost, st), "D2H variance");
        if (narrow)
            GO(cudaMemcpyAsync(flag_wire + wire_chunk * (size_t)(c % RING), f.of8, cnt, cudaMemcpyDeviceToHost, st), "D2H flag");
        else
            GO(cudaMemcpyAsync(out_flag + p0 * hw, f.of, cnt * 8, cudaMemcpyDeviceToHost, st), "D2H flag");
        if (dbg_argmin) GO(cudaMemcpyAsync(dbg_argmin + p0 * hw, f.dbg, cnt * 4, cudaMemcpyDeviceToHost, st), "D2H argmin");
        if (out_hist) GO(cudaMemcpyAsync(out_hist + p0 * 10, f.hist, n * 10 * 8, cudaMemcpyDeviceToHost, st), "D2H hist");
        if (!narrow) GO(cudaMemcpyAsync(st_host[c], f.ws, sizeof(int) * TDK_ST_WORDS, cudaMemcpyDeviceToHost, st), "D2H status");
        GO(cudaEventRecord(ev_out[c], st), "cudaEventRecord");
    }
#ifdef TDK_HOST_TIMING
    const double t_enq = host_now();
#endif
    widen_upto(nchunks);
#undef GO
#ifdef TDK_HOST_TIMING
    const double t_wid = host_now();
#endif
    sync.drain();   // no worker may still hold a pointer into this frame or the caller's arrays
    for (int i = 0; i <= NSTREAM; i++) {
        const int rc2 = tdk_check_cuda(cudaStreamSynchronize(i < NSTREAM ? ar->streams[i] : up_stream), "cudaStreamSynchronize");
        if (rc == TDK_OK) rc = rc2;
    }
    if (narrow && rc == TDK_OK)
        rc = tdk_check_cuda(cudaMemcpy(st_host, tab.status, sizeof(int) * TDK_ST_WORDS * nchunks, cudaMemcpyDeviceToHost), "D2H status");
#ifdef TDK_HOST_TIMING
    fprintf(stderr, "frames_host: enqueue %.2f ms, read-backs done +%.2f ms, end +%.2f ms (%lld chunks)\n",
            1e3 * (t_enq - t_begin), 1e3 * (t_wid - t_enq), 1e3 * (host_now() - t_wid), (long long)nchunks);
#endif
    if (rc == TDK_OK) rc = status_report(st_host, nchunks);
    return rc;
}

// Host half of the byte wire on its own (no device needed): ages -> bytes -> back, flags -> bytes ->
// back through the worker pool, in quarter-frame tasks as tdk_update_depth_frames_host issues them.
int tdk_selftest_host_wire(const uint64_t *age, const int64_t *flag, int64_t n, uint64_t *age_back, int64_t *flag_back)
{
    if (!age || !flag || !age_back || !flag_back || n < 0) {
        tdk_set_error("tdk_selftest_host_wire: bad argument");
        return TDK_ERR_INVALID;
    }
    std::vector<unsigned char> a8((size_t)n);
    std::vector<signed char> f8((size_t)n);
    for (int64_t i = 0; i < n; i++) f8[i] = (signed char)flag[i];   // what narrow_flag_kernel stores
    ChunkSync sync(2);
    const int pieces = (int)(((size_t)n + HOST_PIECE - 1) / HOST_PIECE);
    if (pieces == 0) return TDK_OK;
    sync.begin(sync.narrow_left, 0, pieces);
    sync.begin(sync.widen_left, 1, pieces);
    for (size_t o = 0; o < (size_t)n; o += HOST_PIECE) {
        const size_t len = (size_t)n - o < HOST_PIECE ? (size_t)n - o : HOST_PIECE;
        unsigned char *a = a8.data();
        const signed char *f = f8.data();
        HostPool::get().submit([=, &sync] { narrow_age_host(age + o, a + o, len); sync.end(sync.narrow_left, 0); });
        HostPool::get().submit([=, &sync] { widen_flag_host(f + o, flag_back + o, len); sync.end(sync.widen_left, 1); });
    }
    sync.wait(sync.narrow_left, 0);
    sync.drain();
    for (int64_t i = 0; i < n; i++) age_back[i] = a8[i];   // what widen_age_kernel stores
    return TDK_OK;
}

int tdk_estimate_debug_host(const int64_t *u_key, double prior_depth, double prior_variance,
                            const double *key_img, const double *key_cam, const double *T_wk,
                            const double *ref_img, const double *ref_cam, const double *T_wr,
                            const double *params, int64_t H, int64_t W, double *out, int64_t *out_flag,
                            int64_t *dbg)
{
    if (!u_key || !key_img || !key_cam || !T_wk || !ref_img || !ref_cam || !T_wr || !params || !out || !out_flag) {
        tdk_set_error("tdk_estimate_debug_host: null pointer argument");
        return TDK_ERR_INVALID;
    }
    if (H <= 0 || W <= 0 || !(params[4] > 0.)) {
        tdk_set_error("tdk_estimate_debug_host: bad image size or ref_step_size");
        return TDK_ERR_INVALID;
    }
    if (u_key[0] < 0 || u_key[0] >= W || u_key[1] < 0 || u_key[1] >= H) {
        // The reference evaluates estimate() for any u_key (src/py/semi_dense.rs:126-155): outside the
        // image it ends at KEY_OUT_OF_RANGE unless an earlier test fails; the prior is echoed.
        int dev = 0, rc = current_device(&dev);
        if (rc != TDK_OK) return rc;
        std::lock_guard<std::mutex> lock(g_mutex[dev]);
        Arena *ar = nullptr;
        if ((rc = arena_get(4096, &ar)) != TDK_OK) return rc;
        char *base = (char *)ar->base;
        double *d_in = (double *)base;              // key_cam 4, T_wk 16, ref_cam 4, T_wr 16
        int *d_flag = (int *)(base + 512);
        cudaStream_t st = ar->streams[0];
        double h_in[40];
        memcpy(h_in, key_cam, 32); memcpy(h_in + 4, T_wk, 128); memcpy(h_in + 20, ref_cam, 32); memcpy(h_in + 24, T_wr, 128);
        TDK_CUDA(cudaMemcpyAsync(d_in, h_in, sizeof(h_in), cudaMemcpyHostToDevice, st));
        rc = tdk_estimate_outside_dev(d_in, d_in + 4, d_in + 20, d_in + 24, params, H, W, u_key[0], u_key[1],
                                      prior_depth, prior_variance, d_flag, base + 1024, st);
        if (rc != TDK_OK) return rc;
        int h[1 + TDK_ST_WORDS];
        TDK_CUDA(cudaMemcpyAsync(h, d_flag, sizeof(int), cudaMemcpyDeviceToHost, st));
        TDK_CUDA(cudaMemcpyAsync(h + 1, base + 1024, sizeof(int) * TDK_ST_WORDS, cudaMemcpyDeviceToHost, st));
        TDK_CUDA(cudaStreamSynchronize(st));
        if (h[1 + TDK_ST_SINGULAR]) { tdk_set_error("a reference pose T_wr is singular"); return TDK_ERR_SINGULAR; }
        *out_flag = h[0];
        out[0] = prior_depth;
        out[1] = prior_variance;
        if (dbg) dbg[0] = -1;
        return TDK_OK;
    }
    // One-pixel update: age 1 at u_key, 0 elsewhere, through the same kernel as the full map.
    const size_t hw = (size_t)H * W, at = (size_t)u_key[1] * W + u_key[0];
    uint64_t *age = (uint64_t *)calloc(hw, 8);
    double *pd = (double *)malloc(hw * 8), *pv = (double *)malloc(hw * 8);
    double *od = (double *)malloc(hw * 8), *ov = (double *)malloc(hw * 8);
    int64_t *of = (int64_t *)malloc(hw * 8);
    int32_t *arg = (int32_t *)malloc(hw * 4);
    int rc = TDK_OK;
    if (!age || !pd || !pv || !od || !ov || !of || !arg) {
        tdk_set_error("tdk_estimate_debug_host: out of host memory");
        rc = TDK_ERR_INVALID;
    } else {
        for (size_t i = 0; i < hw; i++) { pd[i] = prior_depth; pv[i] = prior_variance; }
        age[at] = 1;
        rc = tdk_update_depth_host(key_img, key_cam, T_wk, ref_img, ref_cam, T_wr, age, pd, pv, params, 1, 1,
                                   H, W, od, ov, of, arg, nullptr, 0);
        if (rc == TDK_OK) {
            *out_flag = of[at];
            // py/semi_dense.rs:138-154: the prior is echoed untouched on any failure
            out[0] = of[at] == TDK_FLAG_SUCCESS ? od[at] : prior_depth;
            out[1] = of[at] == TDK_FLAG_SUCCESS ? ov[at] : prior_variance;
            if (dbg) dbg[0] = arg[at];
        }
    }
    free(age); free(pd); free(pv); free(od); free(ov); free(of); free(arg);
    return rc;
}

// ---- host-pointer forms of the companions: copy in, run the *_dev entry point, copy out, one
// stream of the per-device arena, returns when the results are in host memory.
namespace {
// Copies between the caller's host arrays and the device.  Pinned memory goes straight to
// cudaMemcpyAsync.  Pageable memory (a numpy array) would be staged by the driver on the calling
// thread at 10-15 GB/s; instead the worker pool copies it through a ring of pinned 512 KB slots,
// piece by piece, and the DMA of a piece starts as soon as it is in its slot (uploads) or the
// workers pick a piece up as soon as its DMA is through (read-backs).
constexpr size_t STAGE_PIECE = (size_t)512 << 10;
constexpr int STAGE_SLOTS = 32;                  // per direction: 16 MB in flight each way
constexpr size_t STAGE_MIN = (size_t)128 << 10;  // smaller copies are not worth a hand-over

// A short wait is cheaper spun than slept: a futex wake-up costs tens of microseconds, a 512 KB memcpy less.
inline void spin_pause(int &spins)
{
    if (++spins < 4096) {
#if defined(__x86_64__) || defined(__i386__)
        __builtin_ia32_pause();
#endif
    } else {
        std::this_thread::yield();
    }
}

struct Tickets {   // completion of tasks handed to the pool
    std::deque<std::atomic<int>> done;   // a deque never moves its elements: workers hold pointers into it
    std::atomic<int64_t> pending{0};
    std::atomic<int> *issue()
    {
        done.emplace_back(0);
        pending.fetch_add(1, std::memory_order_relaxed);
        return &done.back();
    }
    void finish(std::atomic<int> *t)
    {
        t->store(1, std::memory_order_release);
        pending.fetch_sub(1, std::memory_order_release);
    }
    void wait(std::atomic<int> *t)
    {
        int spins = 0;
        while (!t->load(std::memory_order_acquire)) spin_pause(spins);
    }
    void drain()
    {
        int spins = 0;
        while (pending.load(std::memory_order_acquire) != 0) spin_pause(spins);
    }
};

bool is_pageable(const void *p)
{
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
        cudaGetLastError();
        return true;
    }
    return at.type == cudaMemoryTypeUnregistered;
}

// fills `len` bytes of a pinned slot with what belongs at byte offset `off` of the device buffer
typedef std::function<void(unsigned char *slot, size_t off, size_t len)> FillFn;
// takes `len` bytes of a pinned slot that came from byte offset `off` of the device buffer
typedef std::function<void(const unsigned char *slot, size_t off, size_t len)> DrainFn;

struct Stage {
    Arena *ar = nullptr;
    cudaStream_t st = nullptr;
    size_t off = 0;
    Tickets tickets;
    int up_next = 0, down_next = 0;
    bool up_used[STAGE_SLOTS] = {}, down_used[STAGE_SLOTS] = {};
    struct Pending { int slot; size_t off, len; DrainFn fn; };
    std::deque<Pending> pend;   // read-backs whose DMA is queued, oldest first
    void *take(size_t bytes) { void *p = (char *)ar->base + off; off += up256(bytes); return p; }
    ~Stage() { tickets.drain(); }   // no worker may outlive the frame its task points into
    unsigned char *up_slot(int i) { return ar->stage_host + (size_t)i * STAGE_PIECE; }
    unsigned char *down_slot(int i) { return ar->stage_host + (size_t)(STAGE_SLOTS + i) * STAGE_PIECE; }
    cudaEvent_t up_ev(int i) { return ar->stage_ev[i]; }
    cudaEvent_t down_ev(int i) { return ar->stage_ev[STAGE_SLOTS + i]; }
};

int stage_ring_get(Arena *ar)
{
    if (!ar->stage_host) {
        TDK_CUDA(cudaMallocHost((void **)&ar->stage_host, 2 * STAGE_SLOTS * STAGE_PIECE));
        for (int i = 0; i < 2 * STAGE_SLOTS; i++) TDK_CUDA(cudaEventCreateWithFlags(&ar->stage_ev[i], cudaEventDisableTiming));
    }
    return TDK_OK;
}

int stage_open(size_t bytes, Stage *s)
{
    int rc = arena_get(bytes, &s->ar);
    if (rc != TDK_OK) return rc;
    s->st = s->ar->streams[0];
    s->off = 0;
    return TDK_OK;
}

// upload of `bytes` device bytes produced piecewise by `fill` on the worker pool
int up_pieces(Stage &s, void *dst, size_t bytes, const FillFn &fill, size_t piece = STAGE_PIECE)
{
    int rc = stage_ring_get(s.ar);
    if (rc != TDK_OK) return rc;
    for (size_t o = 0; o < bytes;) {
        struct Piece { int slot; size_t off, len; std::atomic<int> *ticket; } win[STAGE_SLOTS];
        int w = 0;
        for (; w < STAGE_SLOTS && o < bytes; w++) {
            Piece &q = win[w];
            q.slot = s.up_next++ % STAGE_SLOTS;
            q.off = o;
            q.len = bytes - o < piece ? bytes - o : piece;
            o += q.len;
            if (s.up_used[q.slot]) TDK_CUDA(cudaEventSynchronize(s.up_ev(q.slot)));   // its last DMA has read it
            q.ticket = s.tickets.issue();
            unsigned char *slot = s.up_slot(q.slot);
            const size_t off = q.off, len = q.len;
            std::atomic<int> *t = q.ticket;
            Tickets *tk = &s.tickets;
            HostPool::get().submit([=] { fill(slot, off, len); tk->finish(t); });
        }
        for (int i = 0; i < w; i++) {
            s.tickets.wait(win[i].ticket);
            TDK_CUDA(cudaMemcpyAsync((char *)dst + win[i].off, s.up_slot(win[i].slot), win[i].len, cudaMemcpyHostToDevice, s.st));
            TDK_CUDA(cudaEventRecord(s.up_ev(win[i].slot), s.st));
            s.up_used[win[i].slot] = true;
        }
    }
    return TDK_OK;
}

// hands the oldest queued read-back to the workers; wait = also until they are through with it
int retire_one(Stage &s, bool wait)
{
    Stage::Pending q = s.pend.front();
    s.pend.pop_front();
    TDK_CUDA(cudaEventSynchronize(s.down_ev(q.slot)));
    std::atomic<int> *t = s.tickets.issue();
    const unsigned char *slot = s.down_slot(q.slot);
    Tickets *tk = &s.tickets;
    HostPool::get().submit([=] { q.fn(slot, q.off, q.len); tk->finish(t); });
    if (wait) s.tickets.wait(t);
    return TDK_OK;
}

// read-back of `bytes` device bytes consumed piecewise by `drain` on the worker pool (after stage_finish)
int down_pieces(Stage &s, const void *src, size_t bytes, const DrainFn &drain, size_t piece = STAGE_PIECE)
{
    int rc = stage_ring_get(s.ar);
    if (rc != TDK_OK) return rc;
    for (size_t o = 0; o < bytes;) {
        const int slot = s.down_next++ % STAGE_SLOTS;
        const size_t len = bytes - o < piece ? bytes - o : piece;
        if (s.down_used[slot]) {   // the ring came round: everything up to its last user has to leave first
            while (!s.pend.empty()) {
                const bool last = s.pend.front().slot == slot;
                if ((rc = retire_one(s, true)) != TDK_OK) return rc;
                if (last) break;
            }
        }
        TDK_CUDA(cudaMemcpyAsync(s.down_slot(slot), (const char *)src + o, len, cudaMemcpyDeviceToHost, s.st));
        TDK_CUDA(cudaEventRecord(s.down_ev(slot), s.st));
        s.down_used[slot] = true;
        s.pend.push_back({slot, o, len, drain});
        o += len;
    }
    return TDK_OK;
}

int up(Stage &s, void *dst, const void *src, size_t bytes)
{
    if (bytes >= STAGE_MIN && is_pageable(src)) {
        const unsigned char *from = (const unsigned char *)src;
        return up_pieces(s, dst, bytes, [from](unsigned char *slot, size_t off, size_t len) { memcpy(slot, from + off, len); });
    }
    return tdk_check_cuda(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, s.st), "H2D");
}

int down(Stage &s, void *dst, const void *src, size_t bytes)
{
    if (bytes >= STAGE_MIN && is_pageable(dst)) {
        unsigned char *to = (unsigned char *)dst;
        return down_pieces(s, src, bytes, [to](const unsigned char *slot, size_t off, size_t len) { memcpy(to + off, slot, len); });
    }
    return tdk_check_cuda(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, s.st), "D2H");
}

// everything queued on the stage is done and every read-back is in the caller's memory
int stage_finish(Stage &s)
{
    int rc = TDK_OK;
    while (!s.pend.empty() && rc == TDK_OK) rc = retire_one(s, false);
    s.tickets.drain();
    const int rc2 = tdk_check_cuda(cudaStreamSynchronize(s.st), "cudaStreamSynchronize");
    return rc != TDK_OK ? rc : rc2;
}
#define TDK_RC(call) do { rc = (call); if (rc != TDK_OK) return rc; } while (0)
}  // namespace

// One call of the numpy drop-in's update_depth (rust_bindings.semi_dense.update_depth,
// src/py/semi_dense.rs:157-187) for a caller that keeps its frames on the device: the images are
// DEVICE pointers (host tables, as tdk_update_depth_frames_dev), cameras, poses and the maps are HOST
// arrays, pageable or pinned.  Ages and flags cross the link as bytes when R < 255.
int tdk_update_depth_maps_host(const double *const *key_img_ptrs, const double *key_cams, const double *T_wk,
                               const double *const *ref_img_ptrs, const double *ref_cams, const double *T_wr,
                               const uint64_t *age, const double *prior_depth, const double *prior_variance,
                               const double *params, int64_t P, int64_t R, int64_t H, int64_t W, double *out_depth,
                               double *out_variance, int64_t *out_flag, int32_t *dbg_argmin, uint64_t *out_hist,
                               int fuse_on_success)
{
    TdkRange range("tdk_update_depth_maps_host");
    if (!key_img_ptrs || !ref_img_ptrs || !key_cams || !T_wk || !ref_cams || !T_wr || !age || !prior_depth ||
        !prior_variance || !params || !out_depth || !out_variance || !out_flag) {
        tdk_set_error("tdk_update_depth_maps_host: null pointer argument");
        return TDK_ERR_INVALID;
    }
    if (P <= 0 || R <= 0 || H <= 0 || W <= 0) {
        tdk_set_error("tdk_update_depth_maps_host: bad sizes");
        return TDK_ERR_INVALID;
    }
    int dev = 0, rc = current_device(&dev);
    if (rc != TDK_OK) return rc;
    std::lock_guard<std::mutex> lock(g_mutex[dev]);
    const size_t n = (size_t)P * H * W, wsb = tdk_update_depth_workspace_bytes(P, R, H, W);
    const bool narrow = R < 255;
    Stage s;
    TDK_RC(stage_open(up256(n * 8) * 6 + up256(n) * 2 + up256(n * 4) + up256(P * 80) + up256(P * 32) + up256(P * 128) +
                          up256(P * R * 32) + up256(P * R * 128) + 512 + up256(wsb), &s));
    if ((rc = status_host_get(s.ar, 1)) != TDK_OK) return rc;
    uint64_t *d_age = (uint64_t *)s.take(n * 8);
    double *d_pd = (double *)s.take(n * 8), *d_pv = (double *)s.take(n * 8);
    double *d_od = (double *)s.take(n * 8), *d_ov = (double *)s.take(n * 8);
    int64_t *d_of = (int64_t *)s.take(n * 8);
    unsigned char *d_age8 = (unsigned char *)s.take(n);
    signed char *d_of8 = (signed char *)s.take(n);
    int32_t *d_dbg = dbg_argmin ? (int32_t *)s.take(n * 4) : nullptr;
    uint64_t *d_hist = out_hist ? (uint64_t *)s.take(P * 80) : nullptr;
    double *d_kc = (double *)s.take(P * 32), *d_Twk = (double *)s.take(P * 128);
    double *d_rc = (double *)s.take(P * R * 32), *d_Twr = (double *)s.take(P * R * 128);
    int *d_status = (int *)s.take(256);
    void *ws = s.take(wsb);
    const unsigned blocks = (unsigned)((n + 2047) / 2048);
#ifdef TDK_HOST_TIMING
    const double t0 = host_now();
#endif
    TDK_RC(up(s, d_kc, key_cams, P * 32));
    TDK_RC(up(s, d_Twk, T_wk, P * 128));
    TDK_RC(up(s, d_rc, ref_cams, P * R * 32));
    TDK_RC(up(s, d_Twr, T_wr, P * R * 128));
    if (narrow) {
        TDK_RC(up_pieces(s, d_age8, n, [age](unsigned char *slot, size_t off, size_t len) { narrow_age_host(age + off, slot, len); }, STAGE_PIECE / 8));   // a piece = 512 KB of the CALLER's array
        widen_age_kernel<<<blocks, 256, 0, s.st>>>(d_age8, d_age, n);
        TDK_RC(tdk_check_cuda(cudaGetLastError(), "widen_age_kernel"));
    } else {
        TDK_RC(up(s, d_age, age, n * 8));
    }
    TDK_RC(up(s, d_pd, prior_depth, n * 8));
    TDK_RC(up(s, d_pv, prior_variance, n * 8));
    if (d_hist) TDK_RC(tdk_check_cuda(cudaMemsetAsync(d_hist, 0, P * 80, s.st), "memset hist"));
#ifdef TDK_HOST_TIMING
    const double t1 = host_now();
#endif
    TDK_RC(tdk_update_depth_frames_dev(key_img_ptrs, d_kc, d_Twk, ref_img_ptrs, d_rc, d_Twr, d_age, d_pd, d_pv, params, P, R,
                                       H, W, d_od, d_ov, d_of, d_dbg, d_hist, fuse_on_success, ws, wsb, s.st));
    if (narrow) {
        narrow_flag_kernel<<<blocks, 256, 0, s.st>>>(d_of, d_of8, n, (const int *)ws, d_status);
        TDK_RC(tdk_check_cuda(cudaGetLastError(), "narrow_flag_kernel"));
    }
#ifdef TDK_HOST_TIMING
    const double t2 = host_now();
#endif
    TDK_RC(down(s, out_depth, d_od, n * 8));
    TDK_RC(down(s, out_variance, d_ov, n * 8));
    if (narrow) {
        TDK_RC(down_pieces(s, d_of8, n, [out_flag](const unsigned char *slot, size_t off, size_t len) {
            widen_flag_host((const signed char *)slot, out_flag + off, len);
        }, STAGE_PIECE / 8));
    } else {
        TDK_RC(down(s, out_flag, d_of, n * 8));
    }
    if (dbg_argmin) TDK_RC(down(s, dbg_argmin, d_dbg, n * 4));
    if (out_hist) TDK_RC(down(s, out_hist, d_hist, P * 80));
    TDK_RC(tdk_check_cuda(cudaMemcpyAsync(s.ar->status_host, ws, sizeof(int) * TDK_ST_WORDS, cudaMemcpyDeviceToHost, s.st), "D2H status"));
#ifdef TDK_HOST_TIMING
    const double t3 = host_now();
#endif
    TDK_RC(stage_finish(s));
#ifdef TDK_HOST_TIMING
    fprintf(stderr, "maps_host: uploads %.0f us, launches %.0f us, read-backs queued %.0f us, finish %.0f us\n",
            1e6 * (t1 - t0), 1e6 * (t2 - t1), 1e6 * (t3 - t2), 1e6 * (host_now() - t3));
#endif
    return status_report(reinterpret_cast<const int(*)[TDK_ST_WORDS]>(s.ar->status_host), 1);
}

int tdk_increment_age_host(const uint64_t *age0, const double *cam0, const double *cam1, const double *T10,
                           const double *depth0, int64_t P, int64_t H, int64_t W, uint64_t *age1)
{
    TdkRange range("tdk_increment_age_host");
    if (!age0 || !cam0 || !cam1 || !T10 || !depth0 || !age1 || P <= 0 || H <= 0 || W <= 0) {
        tdk_set_error("tdk_increment_age_host: bad argument");
        return TDK_ERR_INVALID;
    }
    int dev = 0, rc = current_device(&dev);
    if (rc != TDK_OK) return rc;
    std::lock_guard<std::mutex> lock(g_mutex[dev]);
    const size_t n = (size_t)P * H * W, wsb = tdk_increment_age_workspace_bytes(P, H, W);
    Stage s;
    TDK_RC(stage_open(up256(n * 8) * 3 + up256(P * 32) * 2 + up256(P * 128) + up256(wsb), &s));
    uint64_t *d_age0 = (uint64_t *)s.take(n * 8), *d_age1 = (uint64_t *)s.take(n * 8);
    double *d_depth = (double *)s.take(n * 8), *d_c0 = (double *)s.take(P * 32), *d_c1 = (double *)s.take(P * 32);
    double *d_T = (double *)s.take(P * 128);
    void *ws = s.take(wsb);
    TDK_RC(up(s, d_age0, age0, n * 8));
    TDK_RC(up(s, d_depth, depth0, n * 8));
    TDK_RC(up(s, d_c0, cam0, P * 32));
    TDK_RC(up(s, d_c1, cam1, P * 32));
    TDK_RC(up(s, d_T, T10, P * 128));
    TDK_RC(tdk_increment_age_dev(d_age0, d_c0, d_c1, d_T, d_depth, P, H, W, d_age1, ws, wsb, s.st));
    TDK_RC(down(s, age1, d_age1, n * 8));
    return stage_finish(s);
}

int tdk_propagate_host(const double *T10, const double *cam0, const double *cam1, const double *depth0,
                       const double *variance0, int64_t P, int64_t H, int64_t W, double default_depth,
                       double default_variance, double uncertaintity_bias, double *depth1, double *variance1)
{
    TdkRange range("tdk_propagate_host");
    if (!T10 || !cam0 || !cam1 || !depth0 || !variance0 || !depth1 || !variance1 || P <= 0 || H <= 0 || W <= 0) {
        tdk_set_error("tdk_propagate_host: bad argument");
        return TDK_ERR_INVALID;
    }
    int dev = 0, rc = current_device(&dev);
    if (rc != TDK_OK) return rc;
    std::lock_guard<std::mutex> lock(g_mutex[dev]);
    const size_t n = (size_t)P * H * W, wsb = tdk_propagate_workspace_bytes(P, H, W);
    Stage s;
    TDK_RC(stage_open(up256(n * 8) * 4 + up256(P * 32) * 2 + up256(P * 128) + up256(wsb), &s));
    double *d_d0 = (double *)s.take(n * 8), *d_v0 = (double *)s.take(n * 8);
    double *d_d1 = (double *)s.take(n * 8), *d_v1 = (double *)s.take(n * 8);
    double *d_c0 = (double *)s.take(P * 32), *d_c1 = (double *)s.take(P * 32), *d_T = (double *)s.take(P * 128);
    void *ws = s.take(wsb);
    TDK_RC(up(s, d_d0, depth0, n * 8));
    TDK_RC(up(s, d_v0, variance0, n * 8));
    TDK_RC(up(s, d_c0, cam0, P * 32));
    TDK_RC(up(s, d_c1, cam1, P * 32));
    TDK_RC(up(s, d_T, T10, P * 128));
    TDK_RC(tdk_propagate_dev(d_T, d_c0, d_c1, d_d0, d_v0, P, H, W, default_depth, default_variance,
                             uncertaintity_bias, d_d1, d_v1, ws, wsb, s.st));
    TDK_RC(down(s, depth1, d_d1, n * 8));
    TDK_RC(down(s, variance1, d_v1, n * 8));
    return stage_finish(s);
}

int tdk_fusion_host(const double *mu1, const double *mu2, const double *var1, const double *var2,
                    const int64_t *flag, int64_t n, int depth_domain, double *mu, double *var)
{
    TdkRange range("tdk_fusion_host");
    if (!mu1 || !mu2 || !var1 || !var2 || !mu || !var || n <= 0) {
        tdk_set_error("tdk_fusion_host: bad argument");
        return TDK_ERR_INVALID;
    }
    int dev = 0, rc = current_device(&dev);
    if (rc != TDK_OK) return rc;
    std::lock_guard<std::mutex> lock(g_mutex[dev]);
    const size_t b = (size_t)n * 8;
    Stage s;
    TDK_RC(stage_open(up256(b) * 7, &s));
    double *d_m1 = (double *)s.take(b), *d_m2 = (double *)s.take(b), *d_v1 = (double *)s.take(b), *d_v2 = (double *)s.take(b);
    double *d_mu = (double *)s.take(b), *d_var = (double *)s.take(b);
    int64_t *d_f = flag ? (int64_t *)s.take(b) : nullptr;
    TDK_RC(up(s, d_m1, mu1, b));
    TDK_RC(up(s, d_m2, mu2, b));
    TDK_RC(up(s, d_v1, var1, b));
    TDK_RC(up(s, d_v2, var2, b));
    if (flag) TDK_RC(up(s, d_f, flag, b));
    TDK_RC(tdk_fusion_dev(d_m1, d_m2, d_v1, d_v2, d_f, n, depth_domain, d_mu, d_var, s.st));
    TDK_RC(down(s, mu, d_mu, b));
    TDK_RC(down(s, var, d_var, b));
    return stage_finish(s);
}

int tdk_regularize_host(const double *depth, const double *variance, const int64_t *flag, int64_t P, int64_t H,
                        int64_t W, double *out_depth)
{
    TdkRange range("tdk_regularize_host");
    if (!depth || !variance || !flag || !out_depth || P <= 0 || H <= 0 || W <= 0) {
        tdk_set_error("tdk_regularize_host: bad argument");
        return TDK_ERR_INVALID;
    }
    int dev = 0, rc = current_device(&dev);
    if (rc != TDK_OK) return rc;
    std::lock_guard<std::mutex> lock(g_mutex[dev]);
    const size_t b = (size_t)P * H * W * 8;
    Stage s;
    TDK_RC(stage_open(up256(b) * 4, &s));
    double *d_d = (double *)s.take(b), *d_v = (double *)s.take(b), *d_o = (double *)s.take(b);
    int64_t *d_f = (int64_t *)s.take(b);
    TDK_RC(up(s, d_d, depth, b));
    TDK_RC(up(s, d_v, variance, b));
    TDK_RC(up(s, d_f, flag, b));
    TDK_RC(tdk_regularize_dev(d_d, d_v, d_f, P, H, W, d_o, s.st));
    TDK_RC(down(s, out_depth, d_o, b));
    return stage_finish(s);
}

}  // extern "C"

// ---- host-pointer forms of the helpers on the drop-in surface (src/py/{warp,interpolation,projection,
// transform,triangulation}.rs: numpy in, numpy out)
namespace {
struct In { const void *host; size_t bytes; };
struct Out { void *host; size_t bytes; };
// copy the inputs up, run `body(device inputs, device outputs, stream)`, copy the outputs back
template <int NI, int NO, typename Body>
int run_staged(const char *who, const In (&in)[NI], const Out (&out)[NO], Body body)
{
    TdkRange range(who);
    int dev = 0, rc = current_device(&dev);
    if (rc != TDK_OK) return rc;
    std::lock_guard<std::mutex> lock(g_mutex[dev]);
    size_t total = 0;
    for (int i = 0; i < NI; i++) total += up256(in[i].bytes);
    for (int i = 0; i < NO; i++) total += up256(out[i].bytes);
    Stage s;
    TDK_RC(stage_open(total + 256, &s));
    void *din[NI], *dout[NO];
    for (int i = 0; i < NI; i++) {
        din[i] = s.take(in[i].bytes);
        if (in[i].bytes) TDK_RC(up(s, din[i], in[i].host, in[i].bytes));
    }
    for (int i = 0; i < NO; i++) dout[i] = s.take(out[i].bytes);
    TDK_RC(body(din, dout, s.st));
    for (int i = 0; i < NO; i++)
        if (out[i].bytes) TDK_RC(down(s, out[i].host, dout[i], out[i].bytes));
    return stage_finish(s);
}
}  // namespace

extern "C" {

int tdk_warp_vecs_host(const double *T10, const double *xs0, const double *depths0, int64_t n, double *xs1, double *depths1)
{
    if (!T10 || !xs0 || !depths0 || !xs1 || !depths1 || n < 0) { tdk_set_error("tdk_warp_vecs_host: bad argument"); return TDK_ERR_INVALID; }
    if (n == 0) return TDK_OK;
    const In in[3] = {{T10, 128}, {xs0, (size_t)n * 16}, {depths0, (size_t)n * 8}};
    const Out out[2] = {{xs1, (size_t)n * 16}, {depths1, (size_t)n * 8}};
    return run_staged("tdk_warp_vecs_host", in, out, [&](void **i, void **o, cudaStream_t st) {
        return tdk_warp_vecs_dev((const double *)i[0], (const double *)i[1], (const double *)i[2], n, (double *)o[0], (double *)o[1], st);
    });
}

int tdk_interpolate_host(const double *image, int64_t H, int64_t W, const double *coords, int64_t n, double *out_values)
{
    if (!image || !coords || !out_values || H <= 0 || W <= 0 || n < 0) { tdk_set_error("tdk_interpolate_host: bad argument"); return TDK_ERR_INVALID; }
    if (n == 0) return TDK_OK;
    int32_t status = 0;
    const In in[2] = {{image, (size_t)H * W * 8}, {coords, (size_t)n * 16}};
    const Out out[2] = {{out_values, (size_t)n * 8}, {&status, 4}};
    int rc = run_staged("tdk_interpolate_host", in, out, [&](void **i, void **o, cudaStream_t st) {
        int r = tdk_check_cuda(cudaMemsetAsync(o[1], 0, 4, st), "cudaMemsetAsync");
        if (r != TDK_OK) return r;
        return tdk_interpolate_dev((const double *)i[0], H, W, (const double *)i[1], n, (double *)o[0], (int32_t *)o[1], st);
    });
    if (rc == TDK_OK && status != 0) {
        // tadataka/interpolation/__init__.py:23-29 raises ValueError for such coordinates
        tdk_set_error("Coordinates out of image range");
        return TDK_ERR_RANGE;
    }
    return rc;
}

int tdk_project_vecs_host(const double *ps, int64_t n, double *xs)
{
    if (!ps || !xs || n < 0) { tdk_set_error("tdk_project_vecs_host: bad argument"); return TDK_ERR_INVALID; }
    if (n == 0) return TDK_OK;
    const In in[1] = {{ps, (size_t)n * 24}};
    const Out out[1] = {{xs, (size_t)n * 16}};
    return run_staged("tdk_project_vecs_host", in, out, [&](void **i, void **o, cudaStream_t st) {
        return tdk_project_vecs_dev((const double *)i[0], n, (double *)o[0], st);
    });
}

int tdk_inv_project_vecs_host(const double *xs, const double *depths, int64_t n, double *ps)
{
    if (!xs || !depths || !ps || n < 0) { tdk_set_error("tdk_inv_project_vecs_host: bad argument"); return TDK_ERR_INVALID; }
    if (n == 0) return TDK_OK;
    const In in[2] = {{xs, (size_t)n * 16}, {depths, (size_t)n * 8}};
    const Out out[1] = {{ps, (size_t)n * 24}};
    return run_staged("tdk_inv_project_vecs_host", in, out, [&](void **i, void **o, cudaStream_t st) {
        return tdk_inv_project_vecs_dev((const double *)i[0], (const double *)i[1], n, (double *)o[0], st);
    });
}

int tdk_transform_host(const double *T10, const double *ps0, int64_t n, double *ps1)
{
    if (!T10 || !ps0 || !ps1 || n < 0) { tdk_set_error("tdk_transform_host: bad argument"); return TDK_ERR_INVALID; }
    if (n == 0) return TDK_OK;
    const In in[2] = {{T10, 128}, {ps0, (size_t)n * 24}};
    const Out out[1] = {{ps1, (size_t)n * 24}};
    return run_staged("tdk_transform_host", in, out, [&](void **i, void **o, cudaStream_t st) {
        return tdk_transform_dev((const double *)i[0], (const double *)i[1], n, (double *)o[0], st);
    });
}

int tdk_calc_depth0_host(const double *T10, const double *xs0, const double *xs1, int64_t n, double *depths0)
{
    if (!T10 || !xs0 || !xs1 || !depths0 || n < 0) { tdk_set_error("tdk_calc_depth0_host: bad argument"); return TDK_ERR_INVALID; }
    if (n == 0) return TDK_OK;
    const In in[3] = {{T10, 128}, {xs0, (size_t)n * 16}, {xs1, (size_t)n * 16}};
    const Out out[1] = {{depths0, (size_t)n * 8}};
    return run_staged("tdk_calc_depth0_host", in, out, [&](void **i, void **o, cudaStream_t st) {
        return tdk_calc_depth0_dev((const double *)i[0], (const double *)i[1], (const double *)i[2], n, (double *)o[0], st);
    });
}

}  // extern "C"
