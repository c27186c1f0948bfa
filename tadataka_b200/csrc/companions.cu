// Companion kernels of the semi-dense path: Gaussian fusion, 3x3 validity-weighted
// regularisation, age-map increment and depth/variance propagation.
// All of them are streaming or scatter kernels bound by HBM; see DESIGN.md for bytes/pixel.
#include <math.h>

#include "tdk_common.cuh"

namespace {

constexpr int TPB = 256;

// ------------------------------------------------------------------------------------ fusion
// fusion.rs:3-11 : mu = (mu1*v2 + mu2*v1)/(v1+v2), var = v1*v2/(v1+v2)
__global__ void __launch_bounds__(TPB) fusion_kernel(const double *__restrict__ mu1,
                                                     const double *__restrict__ mu2,
                                                     const double *__restrict__ var1,
                                                     const double *__restrict__ var2,
                                                     const long long *__restrict__ flag, size_t n,
                                                     int depth_domain, double *__restrict__ mu,
                                                     double *__restrict__ var)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const double m1 = mu1[i], v1 = var1[i];
        if (flag && flag[i] != TDK_FLAG_SUCCESS) {
            mu[i] = m1;
            var[i] = v1;
            continue;
        }
        const double m2 = mu2[i], v2 = var2[i];
        const double v = v1 + v2;
        if (depth_domain) {   // propagation.rs:27-34: fuse inverse depths, return a depth
            const double ia = tdk::inv(m1), ib = tdk::inv(m2);
            mu[i] = tdk::inv((ia * v2 + ib * v1) / v);
        } else {
            mu[i] = (m1 * v2 + m2 * v1) / v;
        }
        var[i] = (v1 * v2) / v;
    }
}

// -------------------------------------------------------------------------------- regularize
// regularization.rs:29-64 : sum over the zero-padded 3x3 (row-major) of id*iv / sum of iv where
// flag == Success; no Success cell -> the input depth is kept.
// A CTA stages id*iv and iv of its 64x16 tile plus a one-pixel halo in shared memory (two
// inversions per pixel instead of two per neighbour visit; 18 rows read for 16 written); cells the
// reference skips hold exact zeros, which leave its left-to-right sums unchanged.
#ifndef TDK_SCATTER_PPT
#define TDK_SCATTER_PPT 4   // increment_age 4.37 / 3.70 / 3.45 us per VGA frame at 1 / 2 / 4 pixels per thread, propagate 8.18 / 7.88 / 7.79
#endif
constexpr int RT_W = 64, RT_ROWS = 4, SC_PPT = TDK_SCATTER_PPT;   // source tiles of the forward-warp kernels below:
constexpr int RT_H = RT_ROWS * SC_PPT;                            // 64 x (4 PPT) pixels, PPT pixels per thread (rows ly, ly + 4, ...)
#ifndef TDK_SCATTER_MIN_CTAS
#define TDK_SCATTER_MIN_CTAS 8   // these kernels wait on loads and atomics: 64 resident warps beat 32 (9.5 vs 11.3 us per frame)
#endif
constexpr int SCATTER_MIN_CTAS = TDK_SCATTER_MIN_CTAS;
#ifndef TDK_REG_H
#define TDK_REG_H 8    // 2.47 us per VGA frame against 2.57 at 16 rows and 2.80 at 32: more, smaller CTAs hide the staging latency better
#endif
constexpr int REG_W = 64, REG_H = TDK_REG_H;   // regularize tile
__global__ void __launch_bounds__(TPB) regularize_kernel(const double *__restrict__ depth,
                                                         const double *__restrict__ variance,
                                                         const long long *__restrict__ flag, int P, int H,
                                                         int W, double *__restrict__ out)
{
    __shared__ double s_prod[REG_H + 2][REG_W + 2], s_iv[REG_H + 2][REG_W + 2];
    const int x0 = blockIdx.x * REG_W, y0 = blockIdx.y * REG_H;
    const size_t base = (size_t)blockIdx.z * H * W;
    // Staging in two batched rounds per thread: first the flags of all its cells, then depth and
    // variance of the Success cells only -- every load of a round is in flight at once, and a sparse
    // flag map (3-15 % Success in this pipeline) costs 8 bytes per pixel instead of 24.
    constexpr int CELLS = (REG_H + 2) * (REG_W + 2), PER = (CELLS + TPB - 1) / TPB;
    long long f[PER];
    int idx[PER];
#pragma unroll
    for (int k = 0; k < PER; k++) {
        const int e = threadIdx.x + k * TPB;
        const int hy = e / (REG_W + 2), hx = e - hy * (REG_W + 2);
        const int yy = y0 - 1 + hy, xx = x0 - 1 + hx;
        idx[k] = (e < CELLS && yy >= 0 && yy < H && xx >= 0 && xx < W) ? yy * W + xx : -1;
        f[k] = idx[k] >= 0 ? __ldg(flag + base + (unsigned)idx[k]) : (long long)TDK_FLAG_NOT_PROCESSED;
    }
    double d[PER], v[PER];
#pragma unroll
    for (int k = 0; k < PER; k++) {
        const bool ok = f[k] == TDK_FLAG_SUCCESS;
        d[k] = ok ? __ldg(depth + base + (unsigned)idx[k]) : 0.;
        v[k] = ok ? __ldg(variance + base + (unsigned)idx[k]) : 0.;
    }
#pragma unroll
    for (int k = 0; k < PER; k++) {
        const int e = threadIdx.x + k * TPB;
        if (e >= CELLS) break;
        const int hy = e / (REG_W + 2), hx = e - hy * (REG_W + 2);
        double prod = 0., iv = 0.;
        if (f[k] == TDK_FLAG_SUCCESS) {
            const double id = tdk::inv(d[k]);
            iv = tdk::inv(v[k]);
            prod = id * iv;
        }
        s_prod[hy][hx] = prod;
        s_iv[hy][hx] = iv;
    }
    __syncthreads();
    const int lx = threadIdx.x & (REG_W - 1);
    const int x = x0 + lx;
    if (x >= W) return;
    for (int ly = threadIdx.x / REG_W; ly < REG_H; ly += TPB / REG_W) {
        const int y = y0 + ly;
        if (y >= H) break;
        double num = 0., den = 0.;
#pragma unroll
        for (int dy = 0; dy < 3; dy++) {
#pragma unroll
            for (int dx = 0; dx < 3; dx++) {
                num = num + s_prod[ly + dy][lx + dx];
                den = den + s_iv[ly + dy][lx + dx];
            }
        }
        const size_t c = base + (size_t)y * W + x;
        out[c] = den == 0. ? depth[c] : tdk::inv(num / den);
    }
}

// ------------------------------------------------------------------- forward warp of a pixel
// PerspectiveWarp (warp.rs:72-85): K1 * pi(T10 * pi^-1(K0^-1 u0, depth0)); returns the target's
// linear index inside the frame or -1 when the float coordinates fall outside the image
// (image_range.rs:10-17; the truncation is propagation.rs:73 / age.rs:27).
struct FrameConst {
    double T[12];
    double f0x, f0y, o0x, o0y, f1x, f1y, o1x, o1y;
};

__device__ __forceinline__ FrameConst load_frame_const(const double *__restrict__ T10,
                                                       const double *__restrict__ cam0,
                                                       const double *__restrict__ cam1, int p)
{
    FrameConst c;
#pragma unroll
    for (int i = 0; i < 12; i++) c.T[i] = __ldg(T10 + 16 * (size_t)p + i);
    c.f0x = __ldg(cam0 + 4 * p); c.f0y = __ldg(cam0 + 4 * p + 1);
    c.o0x = __ldg(cam0 + 4 * p + 2); c.o0y = __ldg(cam0 + 4 * p + 3);
    c.f1x = __ldg(cam1 + 4 * p); c.f1y = __ldg(cam1 + 4 * p + 1);
    c.o1x = __ldg(cam1 + 4 * p + 2); c.o1y = __ldg(cam1 + 4 * p + 3);
    return c;
}

__device__ __forceinline__ int warp_target_n(const FrameConst &c, double qx, double qy, double depth0, int H,
                                             int W, double &depth1)
{
    double x1, y1;
    tdk::warp_point(c.T, qx, qy, depth0, x1, y1, depth1);
    const double u = x1 * c.f1x + c.o1x;
    const double v = y1 * c.f1y + c.o1y;
    if (!tdk::in_range(u, v, (double)W - 1., (double)H - 1.)) return -1;
    return (int)v * W + (int)u;
}


// ----------------------------------------------------------------------------- increment_age
// age.rs:6-32.  The raster-order "last writer wins" becomes: the winner of a target is the source
// with the largest linear index (atomicMax), then one gather pass writes age0[winner] + 1.
// Source pixels are visited in 64x4 tiles; K0^-1 u0 depends on the column (row) only, so a tile
// computes 64 + 4 divisions instead of two per pixel (camera.rs:55-57).
__device__ __forceinline__ void tile_tables(const FrameConst &c, double *s_qx, double *s_qy)
{
    const int tid = threadIdx.x;
    if (tid < RT_W) s_qx[tid] = ((double)(blockIdx.x * RT_W + tid) - c.o0x) / c.f0x;
    else if (tid < RT_W + RT_H) s_qy[tid - RT_W] = ((double)(blockIdx.y * RT_H + tid - RT_W) - c.o0y) / c.f0y;
    __syncthreads();
}

__global__ void __launch_bounds__(TPB, SCATTER_MIN_CTAS) age_scatter_kernel(const double *__restrict__ T10,
                                                          const double *__restrict__ cam0,
                                                          const double *__restrict__ cam1,
                                                          const double *__restrict__ depth0, int H, int W,
                                                          unsigned int *__restrict__ winner)
{
    __shared__ double s_qx[RT_W], s_qy[RT_H];
    const int p = blockIdx.z;
    const int hw = H * W;
    const FrameConst c = load_frame_const(T10, cam0, cam1, p);
    const int lx = threadIdx.x & (RT_W - 1), ly = threadIdx.x / RT_W;
    const int x = blockIdx.x * RT_W + lx;
    double d0[SC_PPT];
#pragma unroll
    for (int k = 0; k < SC_PPT; k++) {   // every depth load of the thread is in flight before the tables are built
        const int y = blockIdx.y * RT_H + ly + k * RT_ROWS;
        d0[k] = (x < W && y < H) ? depth0[(size_t)p * hw + (unsigned)(y * W + x)] : 0.;
    }
    tile_tables(c, s_qx, s_qy);
    if (x >= W) return;
    const double qx = s_qx[lx];
#pragma unroll
    for (int k = 0; k < SC_PPT; k++) {
        const int y = blockIdx.y * RT_H + ly + k * RT_ROWS;
        if (y >= H) break;
        double d1;
        const int t = warp_target_n(c, qx, s_qy[ly + k * RT_ROWS], d0[k], H, W, d1);
        if (t >= 0) atomicMax(winner + (size_t)p * hw + t, (unsigned int)(y * W + x) + 1u);
    }
}

__global__ void __launch_bounds__(TPB) age_gather_kernel(const uint64_t *__restrict__ age0,
                                                         const unsigned int *__restrict__ winner, int hw,
                                                         uint64_t *__restrict__ age1)
{
    const int p = blockIdx.y;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= hw) return;
    const unsigned int w = winner[(size_t)p * hw + t];
    age1[(size_t)p * hw + t] = w ? age0[(size_t)p * hw + (w - 1u)] + 1ull : 0ull;
}

// ---- propagate (propagation.rs:9-92, stat.rs:5-26) ---------------------------------------------
// The reference folds the sources of every target pixel in raster order of the SOURCE, and the fold
// is order dependent (gated fusion, nearer-depth-wins).  Here every source pushes itself onto a
// linked list at its target (one atomicExch on the target's head word); prop_fold_kernel walks the
// list of its target, puts the sources in raster order -- up to four in registers through a sorting
// network, which covers every target of an ordinary camera motion; longer lists in a scratch
// segment, insertion or heap sort -- and replays the reference's fold.  Two kernels and one memset
// per call, linear in the number of sources whatever the warp does.
__global__ void __launch_bounds__(TPB, SCATTER_MIN_CTAS) prop_scatter_kernel(const double *__restrict__ T10,
                                                           const double *__restrict__ cam0,
                                                           const double *__restrict__ cam1,
                                                           const double *__restrict__ depth0,
                                                           const double *__restrict__ variance0, double bias,
                                                           int H, int W, unsigned int *__restrict__ head,
                                                           unsigned int *__restrict__ next,
                                                           double *__restrict__ src_d1, double *__restrict__ src_v1)
{
    __shared__ double s_qx[RT_W], s_qy[RT_H];
    const int p = blockIdx.z;
    const int hw = H * W;
    const FrameConst c = load_frame_const(T10, cam0, cam1, p);
    const int lx = threadIdx.x & (RT_W - 1), ly = threadIdx.x / RT_W;
    const int x = blockIdx.x * RT_W + lx;
    double d0[SC_PPT], v0[SC_PPT];
#pragma unroll
    for (int k = 0; k < SC_PPT; k++) {   // every map load of the thread is in flight before the tables are built
        const int y = blockIdx.y * RT_H + ly + k * RT_ROWS;
        const bool in = x < W && y < H;
        d0[k] = in ? depth0[(size_t)p * hw + (unsigned)(y * W + x)] : 0.;
        v0[k] = in ? variance0[(size_t)p * hw + (unsigned)(y * W + x)] : 0.;
    }
    tile_tables(c, s_qx, s_qy);
    if (x >= W) return;
    const double qx = s_qx[lx];
#pragma unroll
    for (int k = 0; k < SC_PPT; k++) {
        const int y = blockIdx.y * RT_H + ly + k * RT_ROWS;
        if (y >= H) break;
        const unsigned int src = (unsigned int)(y * W + x);
        const size_t s = (size_t)p * hw + src;
        double d1;
        const int t = warp_target_n(c, qx, s_qy[ly + k * RT_ROWS], d0[k], H, W, d1);
        if (t < 0) continue;
        // what this source carries to its target: propagate_variance, propagation.rs:9-19 ; powi(4) = (r*r)*(r*r)
        const double ratio = tdk::inv(d1) / tdk::inv(d0[k]);
        const double r2 = ratio * ratio;
        src_d1[s] = d1;
        src_v1[s] = (r2 * r2) * v0[k] + bias;
        next[s] = atomicExch(head + (size_t)p * hw + t, src + 1u);   // 0 terminates a list
    }
}

// handle_collision(a = new, b = stored), propagation.rs:21-46 ; stat.rs:5-26 ; fusion.rs:3-11
__device__ __forceinline__ void fold_step(double d1a, double v1a, double &acc_d, double &acc_v)
{
    const double ia = tdk::inv(d1a), ib = tdk::inv(acc_d);
    const double ds = (ia - ib) * (ia - ib);
    const double fs = TDK_STAT_FACTOR * TDK_STAT_FACTOR;
    if (ds <= fs * v1a && ds <= fs * acc_v) {
        const double v = v1a + acc_v;
        const double mu = (ia * acc_v + ib * v1a) / v;
        acc_d = tdk::inv(mu);
        acc_v = (v1a * acc_v) / v;
    } else if (d1a < acc_d) {
        acc_d = d1a;
        acc_v = v1a;
    }
}

__device__ __forceinline__ void sift_down(unsigned int *a, unsigned int start, unsigned int end)
{
    unsigned int root = start;
    while (2 * root + 1 <= end) {
        unsigned int child = 2 * root + 1, sw = root;
        if (a[sw] < a[child]) sw = child;
        if (child + 1 <= end && a[sw] < a[child + 1]) sw = child + 1;
        if (sw == root) return;
        const unsigned int tmp = a[root];
        a[root] = a[sw];
        a[sw] = tmp;
        root = sw;
    }
}

#ifndef TDK_FOLD_MIN_CTAS
#define TDK_FOLD_MIN_CTAS 8
#endif
__global__ void __launch_bounds__(TPB, TDK_FOLD_MIN_CTAS) prop_fold_kernel(const double *__restrict__ src_d1,
                                                        const double *__restrict__ src_v1, int hw,
                                                        const unsigned int *__restrict__ head,
                                                        const unsigned int *__restrict__ next,
                                                        unsigned int *__restrict__ list,
                                                        unsigned int *__restrict__ list_cursor, double def_d,
                                                        double def_v, double *__restrict__ depth1,
                                                        double *__restrict__ variance1,
                                                        const uint64_t *__restrict__ age0,
                                                        uint64_t *__restrict__ age1)
{
    const int p = blockIdx.y;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= hw) return;
    const size_t g = (size_t)p * hw + t;
    const unsigned int *__restrict__ nx = next + (size_t)p * hw;
    const double *__restrict__ sd = src_d1 + (size_t)p * hw, *__restrict__ sv = src_v1 + (size_t)p * hw;
    const unsigned int h = head[g];
    if (h == 0) {
        depth1[g] = def_d;
        variance1[g] = def_v;
        if (age1) age1[g] = 0ull;
        return;
    }
    // the first four sources of the list, 0xffffffff where it is shorter
    unsigned int a = h - 1u, b = 0xffffffffu, c = 0xffffffffu, d = 0xffffffffu, more = 0u;
    unsigned int link = nx[a];
    if (link) { b = link - 1u; link = nx[b]; }
    if (link) { c = link - 1u; link = nx[c]; }
    if (link) { d = link - 1u; more = nx[d]; }
    double acc_d, acc_v;
    unsigned int last;
    if (more == 0u) {
        unsigned int lo, hi;
        lo = min(a, b); hi = max(a, b); a = lo; b = hi;
        lo = min(c, d); hi = max(c, d); c = lo; d = hi;
        lo = min(a, c); hi = max(a, c); a = lo; c = hi;
        lo = min(b, d); hi = max(b, d); b = lo; d = hi;
        lo = min(b, c); hi = max(b, c); b = lo; c = hi;
        acc_d = sd[a]; acc_v = sv[a];
        if (b != 0xffffffffu) fold_step(sd[b], sv[b], acc_d, acc_v);
        if (c != 0xffffffffu) fold_step(sd[c], sv[c], acc_d, acc_v);
        if (d != 0xffffffffu) fold_step(sd[d], sv[d], acc_d, acc_v);
        last = d != 0xffffffffu ? d : c != 0xffffffffu ? c : b != 0xffffffffu ? b : a;
    } else {
        // longer list: copy it into a segment of the scratch list (the lists of different targets are
        // disjoint, so the segments handed out per frame fit hw), sort, fold
        unsigned int k = 4;
        for (unsigned int l2 = more; l2; l2 = nx[l2 - 1u]) k++;
        unsigned int *seg = list + (size_t)p * hw + atomicAdd(list_cursor + p, k);
        seg[0] = a; seg[1] = b; seg[2] = c; seg[3] = d;
        unsigned int j = 4;
        for (unsigned int l2 = more; l2; l2 = nx[l2 - 1u]) seg[j++] = l2 - 1u;
        if (k <= 16) {   // insertion sort
            for (unsigned int i = 1; i < k; i++) {
                const unsigned int key = seg[i];
                unsigned int jj = i;
                while (jj > 0 && seg[jj - 1] > key) { seg[jj] = seg[jj - 1]; jj--; }
                seg[jj] = key;
            }
        } else {         // heap sort keeps degenerate many-to-one warps O(k log k)
            for (int start = (int)(k - 2) / 2; start >= 0; start--) sift_down(seg, (unsigned int)start, k - 1);
            for (unsigned int end = k - 1; end > 0; end--) {
                const unsigned int tmp = seg[end];
                seg[end] = seg[0];
                seg[0] = tmp;
                sift_down(seg, 0, end - 1);
            }
        }
        acc_d = sd[seg[0]]; acc_v = sv[seg[0]];
        for (unsigned int i = 1; i < k; i++) fold_step(sd[seg[i]], sv[seg[i]], acc_d, acc_v);
        last = seg[k - 1];
    }
    depth1[g] = acc_d;
    variance1[g] = acc_v;
    // age.rs:6-32 through the same warp: the last raster writer of a target is its largest source
    if (age1) age1[g] = age0[(size_t)p * hw + last] + 1ull;
}

size_t up256(size_t v) { return (v + 255) / 256 * 256; }

struct PropLayout {
    size_t head, list_cursor, clear_bytes, next, list, src_d1, src_v1, total;
};
PropLayout prop_layout(int64_t P, int64_t H, int64_t W)
{
    const size_t n = (size_t)P * H * W;
    PropLayout l;
    size_t o = 0;
    // cleared by one memset per call: the list heads and the per-frame cursors of the scratch list
    l.head = o; o = o + n * 4;
    l.list_cursor = o; o = o + (size_t)P * 4;
    l.clear_bytes = o;
    o = up256(o);
    l.next = o; o = up256(o + n * 4);
    l.list = o; o = up256(o + n * 4);
    l.src_d1 = o; o = up256(o + n * 8);
    l.src_v1 = o; o = up256(o + n * 8);
    l.total = o;
    return l;
}

int check_frames(const char *who, int64_t P, int64_t H, int64_t W)
{
    if (P <= 0 || H <= 0 || W <= 0 || P > 65535 || H * W > 0x7fffffffLL / 2 ||
        (double)P * (double)H * (double)W >= 4294967295.0) {
        tdk_set_error("%s: bad sizes P=%lld H=%lld W=%lld", who, (long long)P, (long long)H, (long long)W);
        return TDK_ERR_INVALID;
    }
    return TDK_OK;
}


// How many pixels of each keyframe carry each age, ages above max_age pooled in the last bin:
// the information the caller's "remove unused reference frames" step needs
// (examples/semi_dense_vo.py:202, a TODO there).  Warp-aggregated shared-memory counts for small
// bin counts, global atomics beyond.
constexpr int AGE_BINS_SMEM = 1024;
__global__ void __launch_bounds__(256) age_usage_kernel(const uint64_t *__restrict__ age, int hw, int bins,
                                                       unsigned long long *__restrict__ counts)
{
    __shared__ unsigned int s_cnt[AGE_BINS_SMEM];
    const int p = blockIdx.y;
    const bool in_smem = bins <= AGE_BINS_SMEM;
    if (in_smem) {
        for (int i = threadIdx.x; i < bins; i += blockDim.x) s_cnt[i] = 0;
        __syncthreads();
    }
    unsigned long long *out = counts + (size_t)p * bins;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < hw; i += gridDim.x * blockDim.x) {
        const uint64_t a = age[(size_t)p * hw + i];
        const int b = a < (uint64_t)(bins - 1) ? (int)a : bins - 1;
        if (in_smem) atomicAdd(&s_cnt[b], 1u);
        else atomicAdd(&out[b], 1ULL);
    }
    if (in_smem) {
        __syncthreads();
        for (int i = threadIdx.x; i < bins; i += blockDim.x)
            if (s_cnt[i]) atomicAdd(&out[i], (unsigned long long)s_cnt[i]);
    }
}

}  // namespace

extern "C" {

int tdk_fusion_dev(const double *mu1, const double *mu2, const double *var1, const double *var2,
                   const int64_t *flag, int64_t n, int depth_domain, double *mu, double *var, void *stream)
{
    if (!mu1 || !mu2 || !var1 || !var2 || !mu || !var || n < 0) {
        tdk_set_error("tdk_fusion_dev: bad argument");
        return TDK_ERR_INVALID;
    }
    if (n == 0) return TDK_OK;
    long long blocks = (n + TPB - 1) / TPB;
    if (blocks > 148LL * 32) blocks = 148LL * 32;
    fusion_kernel<<<(unsigned)blocks, TPB, 0, (cudaStream_t)stream>>>(mu1, mu2, var1, var2, (const long long *)flag,
                                                                     (size_t)n, depth_domain, mu, var);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

int tdk_regularize_dev(const double *depth, const double *variance, const int64_t *flag, int64_t P, int64_t H,
                       int64_t W, double *out_depth, void *stream)
{
    if (!depth || !variance || !flag || !out_depth) {
        tdk_set_error("tdk_regularize_dev: null pointer argument");
        return TDK_ERR_INVALID;
    }
    int rc = check_frames("tdk_regularize_dev", P, H, W);
    if (rc != TDK_OK) return rc;
    dim3 grid((unsigned)((W + REG_W - 1) / REG_W), (unsigned)((H + REG_H - 1) / REG_H), (unsigned)P);
    regularize_kernel<<<grid, TPB, 0, (cudaStream_t)stream>>>(depth, variance, (const long long *)flag, (int)P,
                                                              (int)H, (int)W, out_depth);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

size_t tdk_increment_age_workspace_bytes(int64_t P, int64_t H, int64_t W)
{
    if (P <= 0 || H <= 0 || W <= 0) return 0;
    return up256((size_t)P * H * W * 4);
}

int tdk_increment_age_dev(const uint64_t *age0, const double *cam0, const double *cam1, const double *T10,
                          const double *depth0, int64_t P, int64_t H, int64_t W, uint64_t *age1,
                          void *workspace, size_t workspace_bytes, void *stream)
{
    if (!age0 || !cam0 || !cam1 || !T10 || !depth0 || !age1 || !workspace) {
        tdk_set_error("tdk_increment_age_dev: null pointer argument");
        return TDK_ERR_INVALID;
    }
    int rc = check_frames("tdk_increment_age_dev", P, H, W);
    if (rc != TDK_OK) return rc;
    if (workspace_bytes < tdk_increment_age_workspace_bytes(P, H, W)) {
        tdk_set_error("tdk_increment_age_dev: workspace too small");
        return TDK_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int hw = (int)(H * W);
    unsigned int *winner = (unsigned int *)workspace;
    TDK_CUDA(cudaMemsetAsync(winner, 0, (size_t)P * hw * 4, st));
    dim3 grid((unsigned)((hw + TPB - 1) / TPB), (unsigned)P);
    dim3 tgrid((unsigned)((W + RT_W - 1) / RT_W), (unsigned)((H + RT_H - 1) / RT_H), (unsigned)P);
    age_scatter_kernel<<<tgrid, TPB, 0, st>>>(T10, cam0, cam1, depth0, (int)H, (int)W, winner);
    TDK_CUDA(cudaGetLastError());
    age_gather_kernel<<<grid, TPB, 0, st>>>(age0, winner, hw, age1);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

int tdk_age_usage_dev(const uint64_t *age, int64_t P, int64_t H, int64_t W, int64_t max_age, uint64_t *counts,
                      void *stream)
{
    if (!age || !counts) { tdk_set_error("tdk_age_usage_dev: null pointer argument"); return TDK_ERR_INVALID; }
    int rc = check_frames("tdk_age_usage_dev", P, H, W);
    if (rc != TDK_OK) return rc;
    if (max_age < 0 || max_age > (1 << 24)) {
        tdk_set_error("tdk_age_usage_dev: max_age %lld not in [0, 2^24]", (long long)max_age);
        return TDK_ERR_INVALID;
    }
    const int bins = (int)max_age + 2;   // ages 0..max_age, then "older"
    cudaStream_t st = (cudaStream_t)stream;
    TDK_CUDA(cudaMemsetAsync(counts, 0, (size_t)P * bins * sizeof(uint64_t), st));
    const int hw = (int)(H * W);
    int blocks = (hw + 256 * 8 - 1) / (256 * 8);
    if (blocks > 296) blocks = 296;
    age_usage_kernel<<<dim3((unsigned)blocks, (unsigned)P), 256, 0, st>>>(age, hw, bins, (unsigned long long *)counts);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

size_t tdk_propagate_workspace_bytes(int64_t P, int64_t H, int64_t W)
{
    if (P <= 0 || H <= 0 || W <= 0) return 0;
    return prop_layout(P, H, W).total;
}

static int propagate_launch(const char *who, const double *T10, const double *cam0, const double *cam1,
                            const double *depth0, const double *variance0, const uint64_t *age0, int64_t P,
                            int64_t H, int64_t W, double default_depth, double default_variance,
                            double uncertaintity_bias, double *depth1, double *variance1, uint64_t *age1,
                            void *workspace, size_t workspace_bytes, void *stream)
{
    TdkRange range("tdk_propagate");
    if (!T10 || !cam0 || !cam1 || !depth0 || !variance0 || !depth1 || !variance1 || !workspace ||
        ((age0 == nullptr) != (age1 == nullptr))) {
        tdk_set_error("%s: null pointer argument", who);
        return TDK_ERR_INVALID;
    }
    int rc = check_frames(who, P, H, W);
    if (rc != TDK_OK) return rc;
    const PropLayout l = prop_layout(P, H, W);
    if (workspace_bytes < l.total) {
        tdk_set_error("%s: workspace %zu < required %zu", who, workspace_bytes, l.total);
        return TDK_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    char *ws = (char *)workspace;
    unsigned int *head = (unsigned int *)(ws + l.head), *next = (unsigned int *)(ws + l.next);
    unsigned int *list_cursor = (unsigned int *)(ws + l.list_cursor), *list = (unsigned int *)(ws + l.list);
    double *src_d1 = (double *)(ws + l.src_d1), *src_v1 = (double *)(ws + l.src_v1);
    const int hw = (int)(H * W);
    TDK_CUDA(cudaMemsetAsync(ws, 0, l.clear_bytes, st));
    dim3 grid((unsigned)((hw + TPB - 1) / TPB), (unsigned)P);
    dim3 tgrid((unsigned)((W + RT_W - 1) / RT_W), (unsigned)((H + RT_H - 1) / RT_H), (unsigned)P);
    prop_scatter_kernel<<<tgrid, TPB, 0, st>>>(T10, cam0, cam1, depth0, variance0, uncertaintity_bias, (int)H,
                                               (int)W, head, next, src_d1, src_v1);
    TDK_CUDA(cudaGetLastError());
    prop_fold_kernel<<<grid, TPB, 0, st>>>(src_d1, src_v1, hw, head, next, list, list_cursor, default_depth,
                                           default_variance, depth1, variance1, age0, age1);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

int tdk_propagate_dev(const double *T10, const double *cam0, const double *cam1, const double *depth0,
                      const double *variance0, int64_t P, int64_t H, int64_t W, double default_depth,
                      double default_variance, double uncertaintity_bias, double *depth1, double *variance1,
                      void *workspace, size_t workspace_bytes, void *stream)
{
    return propagate_launch("tdk_propagate_dev", T10, cam0, cam1, depth0, variance0, nullptr, P, H, W, default_depth,
                            default_variance, uncertaintity_bias, depth1, variance1, nullptr, workspace,
                            workspace_bytes, stream);
}

int tdk_propagate_age_dev(const double *T10, const double *cam0, const double *cam1, const double *depth0,
                          const double *variance0, const uint64_t *age0, int64_t P, int64_t H, int64_t W,
                          double default_depth, double default_variance, double uncertaintity_bias,
                          double *depth1, double *variance1, uint64_t *age1, void *workspace,
                          size_t workspace_bytes, void *stream)
{
    if (!age0 || !age1) { tdk_set_error("tdk_propagate_age_dev: null pointer argument"); return TDK_ERR_INVALID; }
    return propagate_launch("tdk_propagate_age_dev", T10, cam0, cam1, depth0, variance0, age0, P, H, W,
                            default_depth, default_variance, uncertaintity_bias, depth1, variance1, age1, workspace,
                            workspace_bytes, stream);
}

}  // extern "C"
