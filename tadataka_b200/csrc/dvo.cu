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
__device__ __forceinline__ void warp_pixel(const DvoConst &c, const double (&T)[12], int x, int y, double d,
                                           double &u, double &v, double &px, double &py, double &pz)
{
    const double x0 = ((double)x - c.o0x) / c.f0x, y0 = ((double)y - c.o0y) / c.f0y;
    const double p0 = x0 * d, p1 = y0 * d, p2 = 1. * d;
    px = T[0] * p0 + T[1] * p1 + T[2] * p2 + T[3];
    py = T[4] * p0 + T[5] * p1 + T[6] * p2 + T[7];
    pz = T[8] * p0 + T[9] * p1 + T[10] * p2 + T[11];
    const double z = pz + TDK_EPS;
    u = (px / z) * c.f1x + c.o1x;
    v = (py / z) * c.f1y + c.o1y;
}

// the pose a kernel works with: the launch's own (by value) or, in the device-resident loop, the one
// in device memory.  Read into an array of its own: overwriting part of a local copy of the by-value
// struct lost element 0 (nvcc 12.9 kept reading T[0] from the parameter).
__device__ __forceinline__ void load_pose(const DvoConst &c, const double *__restrict__ T_dev, double (&T)[12])
{
#pragma unroll
    for (int k = 0; k < 12; k++) T[k] = T_dev ? T_dev[k] : c.T[k];
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
    double wparam, double nu, const DvoConst c0, int H, int W, double *__restrict__ out,
    const double *__restrict__ T_dev = nullptr, const double *__restrict__ done = nullptr)
{
    if (done && *done != 0.) return;   // device-resident loop (tdk_dvo_iterate_dev): the iteration has stopped
    const DvoConst &c = c0;
    double T[12];
    load_pose(c0, T_dev, T);
    __shared__ double s_acc[NACC];
    double acc[NACC];
#pragma unroll
    for (int k = 0; k < NACC; k++) acc[k] = 0.;
    const int hw = H * W;
    for (int i = blockIdx.x * TPB + threadIdx.x; i < hw; i += gridDim.x * TPB) {
        const int y = i / W, x = i - y * W;
        double u, v, px, py, pz;
        warp_pixel(c, T, x, y, D0[i], u, v, px, py, pz);
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
        warp_pixel(c, c.T, x, y, D0[i], u, v, px, py, pz);
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
                                                                    const DvoConst c0, int H, int W,
                                                                    double *__restrict__ out,
                                                                    const double *__restrict__ T_dev = nullptr,
                                                                    const double *__restrict__ done = nullptr)
{
    if (done && *done != 0.) return;
    const DvoConst &c = c0;
    double T[12];
    load_pose(c0, T_dev, T);
    __shared__ double s_acc[2];
    double acc[2] = {0., 0.};
    const int hw = H * W;
    for (int i = blockIdx.x * TPB + threadIdx.x; i < hw; i += gridDim.x * TPB) {
        const int y = i / W, x = i - y * W;
        double u, v, px, py, pz;
        warp_pixel(c, T, x, y, D0[i], u, v, px, py, pz);
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
            warp_pixel(c, c.T, x, y, D0[i], u, v, px, py, pz);
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


// ---- the Gauss-Newton loop of _PoseChangeEstimator.__call__ (dvo __init__.py:76-110) without a host
// round trip per iteration.  State, in device doubles:
enum { ST_T = 0, ST_CAND = 16, ST_NE = 32, ST_ERR = 64, ST_STATUS = 66, ST_DONE = 67, ST_ITERS = 68, ST_PREV = 69, ST_LAST_S = 70, ST_LAST_N = 71,
       ST_WORDS = 128 };

struct Pose16 { double m[16]; };

__global__ void dvo_loop_init_kernel(double *__restrict__ st, const Pose16 T)
{
    const int i = threadIdx.x;
    if (i < ST_WORDS) st[i] = i < 16 ? T.m[i] : 0.;
}

// after the first error evaluation: prev_error = mean over the usable pixels (nan when there is none, as np.mean)
__global__ void dvo_loop_first_error_kernel(double *__restrict__ st)
{
    st[ST_PREV] = st[ST_ERR + 1] > 0. ? st[ST_ERR] / st[ST_ERR + 1] : nan("");
    st[ST_ERR] = 0.;
    st[ST_ERR + 1] = 0.;
}

// xi = A^-1 b by Cholesky of the diagonally scaled A (A = J^T W J is symmetric, positive definite unless the scene is degenerate:
// then status 2 and the host takes over with its minimum-norm solve), candidate = exp(xi) * T10
// (pose.py:44-53, se3.py:16-31; the closed form of tadataka_b200/dvo.py exp_se3)
__global__ void dvo_loop_solve_kernel(double *__restrict__ st)
{
    if (st[ST_DONE] != 0.) return;
    const double *ne = st + ST_NE;
    if (ne[27] == 0.) {   // no usable pixel: "Camera pose change is too large."
        st[ST_STATUS] = 1.;
        st[ST_DONE] = 1.;
        return;
    }
    // Jacobi scaling first (translation and rotation columns differ by orders of magnitude): solve
    // (S A S) z = S b with S = diag(A)^-1/2, xi = S z
    double L[6][6], xi[6], y[6], sc[6];
    int k = 0;
    for (int a = 0; a < 6; a++)
        for (int b = a; b < 6; b++) L[b][a] = ne[k++];   // lower triangle
    bool ok = true;
    for (int a = 0; a < 6; a++) {
        if (!(L[a][a] > 0.) || !(L[a][a] < 1e300)) ok = false;
        sc[a] = ok ? 1. / sqrt(L[a][a]) : 0.;
    }
    for (int a = 0; a < 6 && ok; a++)
        for (int b = 0; b <= a; b++) L[a][b] = a == b ? 1. : L[a][b] * sc[a] * sc[b];
    for (int j = 0; j < 6 && ok; j++) {
        double d = L[j][j];
        for (int m = 0; m < j; m++) d -= L[j][m] * L[j][m];
        if (!(d > 1e-14)) { ok = false; break; }   // the scaled matrix has a unit diagonal
        const double r = sqrt(d);
        L[j][j] = r;
        for (int i = j + 1; i < 6; i++) {
            double v = L[i][j];
            for (int m = 0; m < j; m++) v -= L[i][m] * L[j][m];
            L[i][j] = v / r;
        }
    }
    if (!ok) {
        st[ST_STATUS] = 2.;
        st[ST_DONE] = 1.;
        return;
    }
    for (int i = 0; i < 6; i++) {
        double v = ne[21 + i] * sc[i];
        for (int m = 0; m < i; m++) v -= L[i][m] * y[m];
        y[i] = v / L[i][i];
    }
    for (int i = 5; i >= 0; i--) {
        double v = y[i];
        for (int m = i + 1; m < 6; m++) v -= L[m][i] * xi[m];
        xi[i] = v / L[i][i];
    }
    for (int i = 0; i < 6; i++) xi[i] *= sc[i];
    // exp_se3: R = I + a W + b W^2, V = I + b W + c W^2 on the un-normalised skew matrix W = [w]x
    const double wx = xi[3], wy = xi[4], wz = xi[5];
    const double t2 = wx * wx + wy * wy + wz * wz;
    double a, b, c;
    if (t2 < 1e-6) {
        a = 1. - t2 / 6. * (1. - t2 / 20.);
        b = 0.5 - t2 / 24. * (1. - t2 / 30.);
        c = 1. / 6. - t2 / 120. * (1. - t2 / 42.);
    } else {
        const double t = sqrt(t2), sn = sin(t), cs = cos(t);
        a = sn / t;
        b = (1. - cs) / t2;
        c = (t - sn) / (t2 * t);
    }
    const double Wm[3][3] = {{0., -wz, wy}, {wz, 0., -wx}, {-wy, wx, 0.}};
    double W2[3][3], E[4][4];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) W2[i][j] = Wm[i][0] * Wm[0][j] + Wm[i][1] * Wm[1][j] + Wm[i][2] * Wm[2][j];
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) E[i][j] = (i == j ? 1. : 0.) + a * Wm[i][j] + b * W2[i][j];
        const double Wv = Wm[i][0] * xi[0] + Wm[i][1] * xi[1] + Wm[i][2] * xi[2];
        const double W2v = W2[i][0] * xi[0] + W2[i][1] * xi[1] + W2[i][2] * xi[2];
        E[i][3] = xi[i] + b * Wv + c * W2v;
    }
    E[3][0] = E[3][1] = E[3][2] = 0.;
    E[3][3] = 1.;
    const double *T = st + ST_T;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++)
            st[ST_CAND + i * 4 + j] = E[i][0] * T[j] + E[i][1] * T[4 + j] + E[i][2] * T[8 + j] + E[i][3] * T[12 + j];
}

// the accept / stop rule (dvo __init__.py:100-108): stop when the error rises, else take the candidate
__global__ void dvo_loop_accept_kernel(double *__restrict__ st)
{
    const int i = threadIdx.x;
    const bool done = st[ST_DONE] != 0.;
    const double curr = st[ST_ERR + 1] > 0. ? st[ST_ERR] / st[ST_ERR + 1] : nan("");
    const bool stop = curr > st[ST_PREV];   // false for nan, as in the reference
    __syncthreads();
    if (!done && !stop && i < 16) st[ST_T + i] = st[ST_CAND + i];
    if (i >= 32 && i < 32 + NACC) st[ST_NE + (i - 32)] = 0.;   // accumulators of the next iteration
    if (i == 0) {
        if (!done) {
            st[ST_LAST_S] = st[ST_ERR];
            st[ST_LAST_N] = st[ST_ERR + 1];
            if (stop) st[ST_DONE] = 1.;
            else { st[ST_PREV] = curr; st[ST_ITERS] += 1.; }
        }
        st[ST_ERR] = 0.;
        st[ST_ERR + 1] = 0.;
    }
}

// ---- image pyramid of PoseChangeEstimator.__call__ (dvo __init__.py:139-148) on the device: the
// Gaussian pre-filter and the bilinear resampling of tadataka_b200/dvo.py `rescale` (scipy's
// gaussian_filter(mode="reflect", truncate=4) and map_coordinates(order=1, mode="reflect")).
constexpr int GAUSS_MAX_RADIUS = 32;
struct GaussKernel { int radius; double w[2 * GAUSS_MAX_RADIUS + 1]; };

__device__ __forceinline__ int reflect_index(int j, int n)   // d c b a | a b c d | d c b a
{
    while (j < 0 || j >= n) j = j < 0 ? -j - 1 : 2 * n - 1 - j;
    return j;
}

// one separable pass along axis 0 (rows: stride W) or axis 1
__global__ void __launch_bounds__(TPB) gauss_blur_axis_kernel(const double *__restrict__ in, double *__restrict__ out,
                                                              int H, int W, int axis, const GaussKernel g)
{
    const int x = blockIdx.x * 64 + (threadIdx.x & 63);
    const int y = blockIdx.y * 4 + (threadIdx.x >> 6);
    if (x >= W || y >= H) return;
    double acc = 0.;
    for (int k = -g.radius; k <= g.radius; k++) {
        const double v = axis == 0 ? in[(size_t)reflect_index(y + k, H) * W + x] : in[(size_t)y * W + reflect_index(x + k, W)];
        acc += g.w[k + g.radius] * v;
    }
    out[(size_t)y * W + x] = acc;
}

// out[i, j] = image sampled bilinearly at ((i + 0.5) / scale - 0.5, (j + 0.5) / scale - 0.5)
__global__ void __launch_bounds__(TPB) resample_bilinear_kernel(const double *__restrict__ in, int H, int W,
                                                                double *__restrict__ out, int h, int w, double scale)
{
    const int j = blockIdx.x * 64 + (threadIdx.x & 63);
    const int i = blockIdx.y * 4 + (threadIdx.x >> 6);
    if (j >= w || i >= h) return;
    const double cy = ((double)i + 0.5) / scale - 0.5, cx = ((double)j + 0.5) / scale - 0.5;
    const double fy = floor(cy), fx = floor(cx), ty = cy - fy, tx = cx - fx;
    const int y0 = reflect_index((int)fy, H), y1 = reflect_index((int)fy + 1, H);
    const int x0 = reflect_index((int)fx, W), x1 = reflect_index((int)fx + 1, W);
    const double a = in[(size_t)y0 * W + x0], b = in[(size_t)y0 * W + x1];
    const double c = in[(size_t)y1 * W + x0], d = in[(size_t)y1 * W + x1];
    out[(size_t)i * w + j] = (1. - ty) * ((1. - tx) * a + tx * b) + ty * ((1. - tx) * c + tx * d);
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

int tdk_dvo_iterate_dev(const double *I0, const double *D0, const double *I1, const double *GX1, const double *GY1,
                        const double *weights, int weight_mode, double weight_param, const double *cam0,
                        const double *cam1, const double *T10, int64_t H, int64_t W, int max_iter, int resume,
                        double *state, void *stream)
{
    int rc = check_image("tdk_dvo_iterate_dev", H, W, I0 && D0 && I1 && GX1 && GY1 && cam0 && cam1 && T10 && state);
    if (rc != TDK_OK) return rc;
    if (!(weight_mode == W_NONE || weight_mode == W_HUBER || (weight_mode == W_ARRAY && weights)) || max_iter < 0) {
        tdk_set_error("tdk_dvo_iterate_dev: weight mode %d needs a scale per iteration (student-t, tukey): use the "
                      "per-iteration entry points", weight_mode);
        return TDK_ERR_INVALID;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned grid = reduce_grid(H * W);
    Pose16 T;
    for (int i = 0; i < 16; i++) T.m[i] = T10[i];
    const DvoConst c = make_const(cam0, cam1, T10);
    if (!resume) {   // resume: `state` holds an unfinished loop of the same inputs; max_iter more iterations
        dvo_loop_init_kernel<<<1, ST_WORDS, 0, st>>>(state, T);
        dvo_photometric_error_kernel<<<grid, TPB, 0, st>>>(I0, D0, I1, c, (int)H, (int)W, state + ST_ERR, state + ST_T, state + ST_DONE);
        dvo_loop_first_error_kernel<<<1, 1, 0, st>>>(state);
    }
    for (int it = 0; it < max_iter; it++) {
        dvo_normal_equations_kernel<<<grid, TPB, 0, st>>>(I0, D0, I1, GX1, GY1, weights, weight_mode, weight_param, 0., c,
                                                          (int)H, (int)W, state + ST_NE, state + ST_T, state + ST_DONE);
        dvo_loop_solve_kernel<<<1, 1, 0, st>>>(state);
        dvo_photometric_error_kernel<<<grid, TPB, 0, st>>>(I0, D0, I1, c, (int)H, (int)W, state + ST_ERR, state + ST_CAND,
                                                           state + ST_DONE);
        dvo_loop_accept_kernel<<<1, 64, 0, st>>>(state);
    }
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

int tdk_gaussian_blur_dev(const double *image, int64_t H, int64_t W, const double *kernel, int radius, double *tmp,
                          double *out, void *stream)
{
    int rc = check_image("tdk_gaussian_blur_dev", H, W, image && kernel && tmp && out);
    if (rc != TDK_OK) return rc;
    if (radius < 0 || radius > GAUSS_MAX_RADIUS) {
        tdk_set_error("tdk_gaussian_blur_dev: radius %d outside 0..%d", radius, GAUSS_MAX_RADIUS);
        return TDK_ERR_INVALID;
    }
    GaussKernel g;
    g.radius = radius;
    for (int i = 0; i < 2 * radius + 1; i++) g.w[i] = kernel[i];
    cudaStream_t st = (cudaStream_t)stream;
    dim3 grid((unsigned)((W + 63) / 64), (unsigned)((H + 3) / 4));
    gauss_blur_axis_kernel<<<grid, TPB, 0, st>>>(image, tmp, (int)H, (int)W, 0, g);
    gauss_blur_axis_kernel<<<grid, TPB, 0, st>>>(tmp, out, (int)H, (int)W, 1, g);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

int tdk_resample_bilinear_dev(const double *image, int64_t H, int64_t W, double scale, double *out, int64_t h, int64_t w,
                              void *stream)
{
    int rc = check_image("tdk_resample_bilinear_dev", H, W, image && out);
    if (rc != TDK_OK) return rc;
    if (!(scale > 0.) || h <= 0 || w <= 0 || h > 32767 || w > 32767) {
        tdk_set_error("tdk_resample_bilinear_dev: bad scale or output size");
        return TDK_ERR_INVALID;
    }
    dim3 grid((unsigned)((w + 63) / 64), (unsigned)((h + 3) / 4));
    resample_bilinear_kernel<<<grid, TPB, 0, (cudaStream_t)stream>>>(image, (int)H, (int)W, out, (int)h, (int)w, scale);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

}  // extern "C"
