// Direct (photometric) pose-change estimation: the per-pixel halves of the reference's
// tadataka/vo/dvo/__init__.py:46-110 and tadataka/metric.py:13-39.
//
// One Gauss-Newton iteration of the reference builds an [N,6] Jacobian with numpy, solves a
// least-squares problem and re-evaluates the photometric error with a second warp.  Here the two
// per-pixel passes are kernels that reduce straight to the 6x6 normal equations / the error sum;
// the 6x6 solve, the SE(3) exponential and the loop control stay on the host (tadataka_b200/dvo.py).
// Sums are floating-point reductions in a different order than numpy's, so parity with the oracle
// is to tolerance (tests state it), not bit for bit.
#include <math.h>

#include "tdk_common.cuh"

namespace {

constexpr int TPB = 256;
constexpr int NACC = 30;   // 21 (upper J^T W J) + 6 (J^T W r) + count + sum w r^2 + sum r^2

struct DvoConst {
    double T[12];
    double f0x, f0y, o0x, o0y, f1x, f1y, o1x, o1y;
};

// np.gradient(image) (tadataka/vo/dvo/jacobian.py:27-29): central differences inside, one-sided
// differences on the border; returns DX (axis 1) and DY (axis 0).
__global__ void __launch_bounds__(TPB) image_gradient_kernel(const double *__restrict__ img, int H, int W,
                                                             double *__restrict__ gx, double *__restrict__ gy)
{
    const int x = blockIdx.x * 64 + (threadIdx.x & 63);
    const int y = blockIdx.y * 4 + (threadIdx.x >> 6);
    if (x >= W || y >= H) return;
    const size_t i = (size_t)y * W + x;
    double dx = 0., dy = 0.;
    if (W > 1) {
        if (x == 0) dx = (img[i + 1] - img[i]) / 1.;
        else if (x == W - 1) dx = (img[i] - img[i - 1]) / 1.;
        else dx = (img[i + 1] - img[i - 1]) / 2.;
    }
    if (H > 1) {
        if (y == 0) dy = (img[i + W] - img[i]) / 1.;
        else if (y == H - 1) dy = (img[i] - img[i - W]) / 1.;
        else dy = (img[i + W] - img[i - W]) / 2.;
    }
    gx[i] = dx;
    gy[i] = dy;
}

// warp of pixel (x, y): K1 * pi(T10 * pi^-1(K0^-1 u0, d))  (warp.py:78-88 / dvo __init__.py:93-96)
__device__ __forceinline__ void warp_pixel(const DvoConst &c, int x, int y, double d, double &u, double &v,
                                           double &px, double &py, double &pz)
{
    const double x0 = ((double)x - c.o0x) / c.f0x, y0 = ((double)y - c.o0y) / c.f0y;
    const double p0 = x0 * d, p1 = y0 * d, p2 = 1. * d;
    px = c.T[0] * p0 + c.T[1] * p1 + c.T[2] * p2 + c.T[3];
    py = c.T[4] * p0 + c.T[5] * p1 + c.T[6] * p2 + c.T[7];
    pz = c.T[8] * p0 + c.T[9] * p1 + c.T[10] * p2 + c.T[11];
    const double z = pz + TDK_EPS;
    u = (px / z) * c.f1x + c.o1x;
    v = (py / z) * c.f1y + c.o1y;
}

template <int N>
__device__ __forceinline__ void block_accumulate(double (&acc)[N], double *s_acc, double *__restrict__ out)
{
#pragma unroll
    for (int k = 0; k < N; k++) {
        double v = acc[k];
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
        acc[k] = v;
    }
    if (threadIdx.x < N) s_acc[threadIdx.x] = 0.;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < N; k++) atomicAdd(&s_acc[k], acc[k]);
    }
    __syncthreads();
    if (threadIdx.x < N) atomicAdd(&out[threadIdx.x], s_acc[threadIdx.x]);
}

enum { W_NONE = 0, W_ARRAY = 1, W_HUBER = 2, W_STUDENT_T = 3, W_TUKEY = 4 };
constexpr double TUKEY_BETA = 4.6851;   // weights.py:34

// effective weight of the normal equations.  The reference scales rows by sqrt(weights)
// (math.py:45-49), i.e. the normal equations carry `weights` itself:
//   huber      weights.py:38-43  w = 1 if |r| <= k else k / |r|
//   student-t  weights.py:4-20   the function returns sqrt((nu+1) / (nu + r^2 / variance))
//   tukey      weights.py:23-35  (1 - (u / beta)^2)^2 for |u| <= beta else 0, u = r / (c * MAD)
__device__ __forceinline__ double row_weight(int mode, double param, double nu, double r, const double *wmap, size_t i)
{
    if (mode == W_ARRAY) return wmap[i];
    if (mode == W_HUBER) { const double a = fabs(r); return a > param ? param / a : 1.; }
    if (mode == W_STUDENT_T) return sqrt((nu + 1.) / (nu + (r * r) / param));
    if (mode == W_TUKEY) {   // weights.py:23-35: param = c * MAD (tdk_dvo_tukey_scale_dev)
        const double u = r / param;
        if (!(fabs(u) <= TUKEY_BETA)) return 0.;
        const double q = u / TUKEY_BETA, t = 1. - q * q;
        return t * t;
    }
    return 1.;
}

// calc_pose_update (dvo __init__.py:46-73): accumulates J^T W J, J^T W r over the pixels whose warped
// position is inside image 1 and in front of camera 1.
__global__ void __launch_bounds__(TPB) dvo_normal_equations_kernel(
    const double *__restrict__ I0, const double *__restrict__ D0, const double *__restrict__ I1,
    const double *__restrict__ GX1, const double *__restrict__ GY1, const double *__restrict__ wmap, int mode,
    double wparam, double nu, const DvoConst c, int H, int W, double *__restrict__ out)
{
    __shared__ double s_acc[NACC];
    double acc[NACC];
#pragma unroll
    for (int k = 0; k < NACC; k++) acc[k] = 0.;
    const int hw = H * W;
    for (int i = blockIdx.x * TPB + threadIdx.x; i < hw; i += gridDim.x * TPB) {
        const int y = i / W, x = i - y * W;
        double u, v, px, py, pz;
        warp_pixel(c, x, y, D0[i], u, v, px, py, pz);
        if (!(tdk::in_range(u, v, (double)W - 1., (double)H - 1.) && pz > 0.)) continue;
        const double gx = tdk::bilinear(GX1, W, H, u, v), gy = tdk::bilinear(GY1, W, H, u, v);
        const double r = I0[i] - I1[i];   // dvo __init__.py:90: residuals = (I0 - I1).flatten()
        // calc_jacobian, jacobian.py:8-24
        const double fgx = c.f1x * gx, fgy = c.f1y * gy;
        const double z2 = pz * pz, xy = px * py;
        double J[6];
        J[0] = fgx / pz;
        J[1] = fgy / pz;
        J[2] = -(fgx * px + fgy * py) / (pz * pz);
        J[3] = -(fgx * xy + fgy * (z2 + py * py)) / z2;
        J[4] = (fgx * (z2 + px * px) + fgy * xy) / z2;
        J[5] = (-fgx * py + fgy * px) / pz;
        const double w = row_weight(mode, wparam, nu, r, wmap, (size_t)i);
        int k = 0;
#pragma unroll
        for (int a = 0; a < 6; a++) {
#pragma unroll
            for (int b = a; b < 6; b++) acc[k++] += w * J[a] * J[b];
        }
#pragma unroll
        for (int a = 0; a < 6; a++) acc[21 + a] += w * J[a] * r;
        acc[27] += 1.;
        acc[28] += w * r * r;
        acc[29] += r * r;
    }
    block_accumulate<NACC>(acc, s_acc, out);
}

// one fixed-point step of the student-t scale (weights.py:13-18): sum over the masked residuals of
// s * (nu+1) / (nu + s/variance), and their count
__global__ void __launch_bounds__(TPB) dvo_student_t_kernel(const double *__restrict__ I0,
                                                            const double *__restrict__ D0,
                                                            const double *__restrict__ I1, double variance,
                                                            double nu, const DvoConst c, int H, int W,
                                                            double *__restrict__ out)
{
    __shared__ double s_acc[2];
    double acc[2] = {0., 0.};
    const int hw = H * W;
    for (int i = blockIdx.x * TPB + threadIdx.x; i < hw; i += gridDim.x * TPB) {
        const int y = i / W, x = i - y * W;
        double u, v, px, py, pz;
        warp_pixel(c, x, y, D0[i], u, v, px, py, pz);
        if (!(tdk::in_range(u, v, (double)W - 1., (double)H - 1.) && pz > 0.)) continue;
        const double r = I0[i] - I1[i];
        const double s = r * r;
        acc[0] += s * ((nu + 1.) / (nu + s / variance));
        acc[1] += 1.;
    }
    block_accumulate<2>(acc, s_acc, out);
}

// photometric_error (metric.py:13-28): sum of (I0[u0] - I1(warp(u0)))^2 and the number of pixels whose
// warped position is inside image 1 (no test on the sign of the depth)
__global__ void __launch_bounds__(TPB) dvo_photometric_error_kernel(const double *__restrict__ I0,
                                                                    const double *__restrict__ D0,
                                                                    const double *__restrict__ I1,
                                                                    const DvoConst c, int H, int W,
                                                                    double *__restrict__ out)
{
    __shared__ double s_acc[2];
    double acc[2] = {0., 0.};
    const int hw = H * W;
    for (int i = blockIdx.x * TPB + threadIdx.x; i < hw; i += gridDim.x * TPB) {
        const int y = i / W, x = i - y * W;
        double u, v, px, py, pz;
        warp_pixel(c, x, y, D0[i], u, v, px, py, pz);
        if (!tdk::in_range(u, v, (double)W - 1., (double)H - 1.)) continue;
        const double d = I0[i] - tdk::bilinear(I1, W, H, u, v);
        acc[0] += d * d;
        acc[1] += 1.;
    }
    block_accumulate<2>(acc, s_acc, out);
}

// ---- Tukey scale: c * median(|r - median(r)|) over the residuals of the usable pixels
// (weights.py:29-35).  residual_compact_kernel gathers them (order is irrelevant to a median);
// median_select_kernel, one CTA, finds the middle element(s) by an 8-pass radix select over
// order-preserving 64-bit keys and averages the two middle ones for an even count, as np.median does.
__global__ void __launch_bounds__(TPB) residual_compact_kernel(const double *__restrict__ I0, const double *__restrict__ D0,
                                                               const double *__restrict__ I1, const DvoConst c, int H,
                                                               int W, double *__restrict__ out, unsigned int *count)
{
    const int hw = H * W;
    for (int i0 = blockIdx.x * TPB; i0 < hw; i0 += gridDim.x * TPB) {
        const int i = i0 + threadIdx.x;
        bool ok = false;
        double r = 0.;
        if (i < hw) {
            const int y = i / W, x = i - y * W;
            double u, v, px, py, pz;
            warp_pixel(c, x, y, D0[i], u, v, px, py, pz);
            ok = tdk::in_range(u, v, (double)W - 1., (double)H - 1.) && pz > 0.;
            r = I0[i] - I1[i];
        }
        const unsigned m = __ballot_sync(0xffffffffu, ok);
        if (m) {
            const int lane = threadIdx.x & 31;
            unsigned base = 0;
            if (lane == 0) base = atomicAdd(count, (unsigned)__popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (ok) out[base + __popc(m & ((1u << lane) - 1u))] = r;
        }
    }
}

__device__ __forceinline__ unsigned long long order_key(double v)
{
    const unsigned long long b = (unsigned long long)__double_as_longlong(v);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double key_value(unsigned long long k)
{
    const unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
    return __longlong_as_double((long long)b);
}

// values[0..n): *result = median (np.median semantics).  If abs_dev != 0 the median is taken of
// |values[i] - *centre| instead and *result = scale * that.  One CTA of 1024 threads.
__global__ void __launch_bounds__(1024) median_select_kernel(const double *__restrict__ values, const unsigned int *count,
                                                             int abs_dev, const double *centre, double scale,
                                                             double *result)
{
    __shared__ unsigned int s_hist[256];
    __shared__ unsigned long long s_prefix;
    __shared__ unsigned int s_rank;
    __shared__ double s_found[2];
    const unsigned int n = *count;
    if (n == 0) { if (threadIdx.x == 0) *result = __longlong_as_double(0x7ff8000000000000ll); return; }
    const double ctr = abs_dev ? *centre : 0.;
    const unsigned int ranks[2] = {(n - 1) / 2, n / 2};
    for (int which = 0; which < 2; which++) {
        if (which == 1 && ranks[1] == ranks[0]) { if (threadIdx.x == 0) s_found[1] = s_found[0]; break; }
        if (threadIdx.x == 0) { s_prefix = 0ull; s_rank = ranks[which]; }
        for (int pass = 0; pass < 8; pass++) {
            const int shift = 56 - 8 * pass;
            if (threadIdx.x < 256) s_hist[threadIdx.x] = 0;
            __syncthreads();
            const unsigned long long prefix = s_prefix;
            const unsigned long long mask = pass == 0 ? 0ull : ~0ull << (shift + 8);
            for (unsigned int i = threadIdx.x; i < n; i += blockDim.x) {
                const double v = abs_dev ? fabs(values[i] - ctr) : values[i];
                const unsigned long long k = order_key(v);
                if ((k & mask) == prefix) atomicAdd(&s_hist[(unsigned)(k >> shift) & 255u], 1u);
            }
            __syncthreads();
            if (threadIdx.x == 0) {
                unsigned int r = s_rank, b = 0;
                while (b < 255 && r >= s_hist[b]) { r -= s_hist[b]; b++; }
                s_rank = r;
                s_prefix = prefix | ((unsigned long long)b << shift);
            }
            __syncthreads();
        }
        if (threadIdx.x == 0) s_found[which] = key_value(s_prefix);
        __syncthreads();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const double med = ranks[0] == ranks[1] ? s_found[0] : (s_found[0] + s_found[1]) / 2.;
        *result = abs_dev ? scale * med : med;
    }
}

DvoConst make_const(const double *cam0, const double *cam1, const double *T10)
{
    DvoConst c;
    for (int i = 0; i < 12; i++) c.T[i] = T10[i];
    c.f0x = cam0[0]; c.f0y = cam0[1]; c.o0x = cam0[2]; c.o0y = cam0[3];
    c.f1x = cam1[0]; c.f1y = cam1[1]; c.o1x = cam1[2]; c.o1y = cam1[3];
    return c;
}

int check_image(const char *who, int64_t H, int64_t W, bool ptrs_ok)
{
    if (!ptrs_ok || H <= 0 || W <= 0 || H > 32767 || W > 32767) {
        tdk_set_error("%s: bad argument (H=%lld, W=%lld)", who, (long long)H, (long long)W);
        return TDK_ERR_INVALID;
    }
    return TDK_OK;
}

unsigned reduce_grid(int64_t hw)
{
    const int64_t blocks = (hw + TPB - 1) / TPB;
    return (unsigned)(blocks < 148 * 4 ? blocks : 148 * 4);
}

}  // namespace

extern "C" {

int tdk_image_gradient_dev(const double *image, int64_t H, int64_t W, double *gx, double *gy, void *stream)
{
    int rc = check_image("tdk_image_gradient_dev", H, W, image && gx && gy);
    if (rc != TDK_OK) return rc;
    dim3 grid((unsigned)((W + 63) / 64), (unsigned)((H + 3) / 4));
    image_gradient_kernel<<<grid, TPB, 0, (cudaStream_t)stream>>>(image, (int)H, (int)W, gx, gy);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

int tdk_dvo_normal_equations_dev(const double *I0, const double *D0, const double *I1, const double *GX1,
                                 const double *GY1, const double *weights, int weight_mode, double weight_param,
                                 double nu, const double *cam0, const double *cam1, const double *T10, int64_t H,
                                 int64_t W, double *out, void *stream)
{
    int rc = check_image("tdk_dvo_normal_equations_dev", H, W, I0 && D0 && I1 && GX1 && GY1 && cam0 && cam1 && T10 && out);
    if (rc != TDK_OK) return rc;
    if (weight_mode < W_NONE || weight_mode > W_TUKEY || (weight_mode == W_ARRAY && !weights)) {
        tdk_set_error("tdk_dvo_normal_equations_dev: bad weight mode %d", weight_mode);
        return TDK_ERR_INVALID;
    }
    cudaStream_t st = (cudaStream_t)stream;
    TDK_CUDA(cudaMemsetAsync(out, 0, NACC * sizeof(double), st));
    dvo_normal_equations_kernel<<<reduce_grid(H * W), TPB, 0, st>>>(I0, D0, I1, GX1, GY1, weights, weight_mode,
                                                                   weight_param, nu, make_const(cam0, cam1, T10),
                                                                   (int)H, (int)W, out);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

int tdk_dvo_student_t_dev(const double *I0, const double *D0, const double *I1, double variance, double nu,
                          const double *cam0, const double *cam1, const double *T10, int64_t H, int64_t W,
                          double *out, void *stream)
{
    int rc = check_image("tdk_dvo_student_t_dev", H, W, I0 && D0 && I1 && cam0 && cam1 && T10 && out);
    if (rc != TDK_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    TDK_CUDA(cudaMemsetAsync(out, 0, 2 * sizeof(double), st));
    dvo_student_t_kernel<<<reduce_grid(H * W), TPB, 0, st>>>(I0, D0, I1, variance, nu, make_const(cam0, cam1, T10),
                                                            (int)H, (int)W, out);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

size_t tdk_dvo_tukey_workspace_bytes(int64_t H, int64_t W)
{
    if (H <= 0 || W <= 0) return 0;
    return 256 + (size_t)H * W * sizeof(double);
}

int tdk_dvo_tukey_scale_dev(const double *I0, const double *D0, const double *I1, double c_mad, const double *cam0,
                            const double *cam1, const double *T10, int64_t H, int64_t W, double *out,
                            void *workspace, size_t workspace_bytes, void *stream)
{
    int rc = check_image("tdk_dvo_tukey_scale_dev", H, W, I0 && D0 && I1 && cam0 && cam1 && T10 && out && workspace);
    if (rc != TDK_OK) return rc;
    if (workspace_bytes < tdk_dvo_tukey_workspace_bytes(H, W)) {
        tdk_set_error("tdk_dvo_tukey_scale_dev: workspace too small");
        return TDK_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    unsigned int *count = (unsigned int *)workspace;
    double *median = (double *)((char *)workspace + 64);
    double *values = (double *)((char *)workspace + 256);
    TDK_CUDA(cudaMemsetAsync(workspace, 0, 256, st));
    residual_compact_kernel<<<reduce_grid(H * W), TPB, 0, st>>>(I0, D0, I1, make_const(cam0, cam1, T10), (int)H, (int)W,
                                                               values, count);
    median_select_kernel<<<1, 1024, 0, st>>>(values, count, 0, nullptr, 1., median);
    median_select_kernel<<<1, 1024, 0, st>>>(values, count, 1, median, c_mad, out);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

int tdk_dvo_photometric_error_dev(const double *I0, const double *D0, const double *I1, const double *cam0,
                                  const double *cam1, const double *T10, int64_t H, int64_t W, double *out,
                                  void *stream)
{
    int rc = check_image("tdk_dvo_photometric_error_dev", H, W, I0 && D0 && I1 && cam0 && cam1 && T10 && out);
    if (rc != TDK_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    TDK_CUDA(cudaMemsetAsync(out, 0, 2 * sizeof(double), st));
    dvo_photometric_error_kernel<<<reduce_grid(H * W), TPB, 0, st>>>(I0, D0, I1, make_const(cam0, cam1, T10), (int)H,
                                                                    (int)W, out);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

}  // extern "C"
