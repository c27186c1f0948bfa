// update_depth: per-pixel epipolar search + inverse-depth update, batched over keyframes.
// Replaces src/semi_dense/semi_dense.rs:91-234 of the reference (see include/tdk.h).
//
// One launch = pair_setup_kernel (P*R threads: the per-pair constants the reference recomputes
// for every pixel) + update_depth_kernel.  A CTA owns a 64x4 pixel tile of one keyframe and runs
// three phases over it:
//   A  one thread per pixel : prior checks, epipolar geometry, 5 key samples, gradient test,
//                             reference-line end tests.  ~85-98 % of pixels finish here.
//   B  G lanes per survivor : bilinear samples along the reference epipolar line, the
//                             normalised 5-sample SSD for every window, shuffle argmin
//                             (first minimum wins, as intensities.rs:24 `e < min_error`).
//   C  one thread per survivor: triangulation, alpha / geometric / photometric variance,
//                             final hypothesis check.
// Survivors are compacted into a shared-memory queue between A and B, so no lane idles on a
// finished pixel during the search, and no global scratch is needed for the queue.
// Results are staged in shared memory and written once, coalesced.
#include <math.h>

#include <mutex>

#include "tdk_common.cuh"

namespace {

constexpr int TILE_W = 64;
constexpr int TILE_H = 4;
constexpr int BLOCK = TILE_W * TILE_H;

struct UpdArgs {
    const TdkKeyConst *keys;
    const TdkPairConst *pairs;
    const uint64_t *age;
    const double *prior_d, *prior_v;
    double *out_d, *out_v;
    long long *out_f;
    int *dbg;
    unsigned long long *hist;
    int *status;
    double vmin, vmax, geo, photo, step, min_grad;
    int P, R, H, W, tiles_x, tiles_y, fuse;
};

// ---------------------------------------------------------------------------------------------
// 4x4 inverse: LU with partial pivoting, inverse of U, then inv(A)*L = inv(U) and the column
// interchanges -- the netlib getrf/getri sequence ndarray-linalg's inv() runs for the reference
// (semi_dense.rs:87).  Returns 0 or (j+1) for an exactly singular U[j][j].
__device__ int mat4_inverse(double *A)
{
    int ipiv[4];
    int info = 0;
    for (int j = 0; j < 4; j++) {
        int p = j;
        double best = fabs(A[4 * j + j]);
        for (int i = j + 1; i < 4; i++) {
            const double c = fabs(A[4 * i + j]);
            if (c > best) { best = c; p = i; }
        }
        ipiv[j] = p;
        if (A[4 * p + j] != 0.) {
            if (p != j)
                for (int k = 0; k < 4; k++) {
                    const double t = A[4 * j + k];
                    A[4 * j + k] = A[4 * p + k];
                    A[4 * p + k] = t;
                }
            if (fabs(A[4 * j + j]) >= 2.2250738585072014e-308) {
                const double r = 1. / A[4 * j + j];
                for (int i = j + 1; i < 4; i++) A[4 * i + j] = A[4 * i + j] * r;
            } else {
                for (int i = j + 1; i < 4; i++) A[4 * i + j] = A[4 * i + j] / A[4 * j + j];
            }
        } else if (info == 0) {
            info = j + 1;
        }
        for (int k = j + 1; k < 4; k++)
            for (int i = j + 1; i < 4; i++) A[4 * i + k] = A[4 * i + k] - A[4 * i + j] * A[4 * j + k];
    }
    if (info) return info;
    for (int j = 0; j < 4; j++) {  // inverse of the upper factor, column by column
        A[4 * j + j] = 1. / A[4 * j + j];
        const double ajj = -A[4 * j + j];
        for (int jj = 0; jj < j; jj++) {
            if (A[4 * jj + j] != 0.) {
                const double temp = A[4 * jj + j];
                for (int i = 0; i < jj; i++) A[4 * i + j] = A[4 * i + j] + temp * A[4 * i + jj];
                A[4 * jj + j] = A[4 * jj + j] * A[4 * jj + jj];
            }
        }
        for (int i = 0; i < j; i++) A[4 * i + j] = ajj * A[4 * i + j];
    }
    for (int j = 3; j >= 0; j--) {
        double work[4] = {0., 0., 0., 0.};
        for (int i = j + 1; i < 4; i++) { work[i] = A[4 * i + j]; A[4 * i + j] = 0.; }
        for (int jj = j + 1; jj < 4; jj++) {
            const double temp = -1. * work[jj];
            for (int i = 0; i < 4; i++) A[4 * i + j] = A[4 * i + j] + temp * A[4 * i + jj];
        }
    }
    for (int j = 2; j >= 0; j--) {
        const int jp = ipiv[j];
        if (jp != j)
            for (int i = 0; i < 4; i++) {
                const double t = A[4 * i + j];
                A[4 * i + j] = A[4 * i + jp];
                A[4 * i + jp] = t;
            }
    }
    return 0;
}

// One thread per (keyframe, reference): T_rk = inv(T_wr) * T_wk, the key epipole, pi(t_rk).
// Image pointers come either from contiguous [P,H,W] / [P,R,H,W] blocks or from pointer tables.
__global__ void pair_setup_kernel(const double *key_imgs, const double *const *key_ptrs,
                                  const double *__restrict__ key_cams, const double *__restrict__ T_wk,
                                  const double *ref_imgs, const double *const *ref_ptrs,
                                  const double *__restrict__ ref_cams, const double *__restrict__ T_wr,
                                  int P, int R, size_t hw, TdkKeyConst *keys, TdkPairConst *pairs,
                                  int *status)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P * R) return;
    const int p = idx / R, r = idx % R;
    if (r == 0) {
        TdkKeyConst k;
        k.fx = key_cams[4 * p + 0]; k.fy = key_cams[4 * p + 1];
        k.ox = key_cams[4 * p + 2]; k.oy = key_cams[4 * p + 3];
        k.img = key_ptrs ? key_ptrs[p] : key_imgs + (size_t)p * hw;
        k.pad = 0;
        keys[p] = k;
    }
    double Twk[16], Trw[16];
    for (int i = 0; i < 16; i++) { Twk[i] = T_wk[16 * (size_t)p + i]; Trw[i] = T_wr[16 * (size_t)idx + i]; }
    const double twr[3] = {Trw[3], Trw[7], Trw[11]};
    if (mat4_inverse(Trw) != 0) atomicExch(&status[TDK_ST_SINGULAR], 1);
    TdkPairConst c;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 4; j++)
            c.T[4 * i + j] = Trw[4 * i + 0] * Twk[0 + j] + Trw[4 * i + 1] * Twk[4 + j] +
                             Trw[4 * i + 2] * Twk[8 + j] + Trw[4 * i + 3] * Twk[12 + j];
    // e_key = pi(R_wk^T (t_wr - t_wk))   (epipolar.rs:9-20)
    const double dx = twr[0] - Twk[3], dy = twr[1] - Twk[7], dz = twr[2] - Twk[11];
    const double px = Twk[0] * dx + Twk[4] * dy + Twk[8] * dz;
    const double py = Twk[1] * dx + Twk[5] * dy + Twk[9] * dz;
    const double pz = Twk[2] * dx + Twk[6] * dy + Twk[10] * dz;
    const double ez = pz + TDK_EPS;
    c.ex = px / ez;
    c.ey = py / ez;
    const double tz = c.T[11] + TDK_EPS;   // pi(t_rk)  (variance.rs:51)
    c.ptx = c.T[3] / tz;
    c.pty = c.T[7] / tz;
    c.rfx = ref_cams[4 * (size_t)idx + 0]; c.rfy = ref_cams[4 * (size_t)idx + 1];
    c.rox = ref_cams[4 * (size_t)idx + 2]; c.roy = ref_cams[4 * (size_t)idx + 3];
    c.ref = ref_ptrs ? ref_ptrs[idx] : ref_imgs + (size_t)idx * hw;
    c.pad = 0;
    pairs[idx] = c;
}

// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int check_args(double id, double var, double vmin, double vmax)
{
    // hypothesis.rs:21-37
    if (id <= 0.) return TDK_FLAG_NEGATIVE_PRIOR_DEPTH;
    const double lo = id - TDK_VARIANCE_FACTOR * var;
    const double hi = id + TDK_VARIANCE_FACTOR * var;
    if (hi <= vmin || vmax <= lo) return TDK_FLAG_HYPOTHESIS_OUT_OF_SEARCH_RANGE;
    return TDK_FLAG_SUCCESS;
}

// 3x3 correlation at an interior pixel, summed the way ndarray's sum() folds a contiguous 3x3
// (gradient.rs:4-26, convolution.rs:27-51): (e0+e4)+(e1+e5)+(e2+e6)+(e3+e7)+e8.
__device__ __forceinline__ double fold9(const double *e)
{
    double acc = 0. + (e[0] + e[4]);
    acc = acc + (e[1] + e[5]);
    acc = acc + (e[2] + e[6]);
    acc = acc + (e[3] + e[7]);
    acc = acc + e[8];
    return acc;
}

template <int G>
__global__ void __launch_bounds__(BLOCK, 2) update_depth_kernel(const UpdArgs a)
{
    __shared__ double s_depth[BLOCK], s_var[BLOCK];
    __shared__ double q_xminx[BLOCK], q_xminy[BLOCK], q_ux[BLOCK], q_uy[BLOCK];
    __shared__ double q_rdx[BLOCK], q_rdy[BLOCK], q_kstep[BLOCK], q_g[BLOCK];
    __shared__ double q_k[5][BLOCK];
    __shared__ int q_n[BLOCK], q_arg[BLOCK], s_arg[BLOCK];
    __shared__ unsigned short q_pix[BLOCK], q_ref[BLOCK];
    __shared__ signed char s_flag[BLOCK];
    __shared__ int s_nq;
    __shared__ unsigned int s_hist[10];

    const int tid = threadIdx.x;
    const int tiles = a.tiles_x * a.tiles_y;
    const int p = blockIdx.x / tiles;
    const int t = blockIdx.x - p * tiles;
    const int ty = t / a.tiles_x, tx = t - ty * a.tiles_x;
    const int W = a.W, H = a.H;
    const size_t map_base = (size_t)p * H * W;
    const TdkKeyConst key = a.keys[p];
    const double wm1 = (double)W - 1., hm1 = (double)H - 1.;

    if (tid == 0) s_nq = 0;
    if (tid < 10) s_hist[tid] = 0;
    __syncthreads();

    // ------------------------------------------------------------------ phase A
    const int x = tx * TILE_W + (tid & (TILE_W - 1));
    const int y = ty * TILE_H + (tid / TILE_W);
    const bool valid = x < W && y < H;
    const size_t pix = map_base + (size_t)y * W + x;
    int flag = TDK_FLAG_NOT_PROCESSED;
    s_arg[tid] = -1;
    if (valid) {
        const uint64_t age = a.age[pix];
        const double d = a.prior_d[pix], v = a.prior_v[pix];
        double out_d = d, out_v = v;
        do {
            if (age == 0) break;   // semi_dense.rs:194-200
            if (age > (uint64_t)a.R) { atomicExch(&a.status[TDK_ST_AGE], 1); break; }
            const double idp = tdk::inv(d);
            flag = check_args(idp, v, a.vmin, a.vmax);   // semi_dense.rs:209-214
            if (flag != TDK_FLAG_SUCCESS) break;
            // from here on a failure stores inv(inv(d))  (semi_dense.rs:220-225)
            out_d = tdk::inv(idp);
            const int r = a.R - (int)age;
            const TdkPairConst *__restrict__ pc = a.pairs + (size_t)p * a.R + r;
            const double *__restrict__ T = pc->T;

            // hypothesis.rs:54-61, depth.rs:25-30
            const double lo = tdk::clampd(idp - TDK_VARIANCE_FACTOR * v, a.vmin, a.vmax);
            const double hi = tdk::clampd(idp + TDK_VARIANCE_FACTOR * v, a.vmin, a.vmax);
            const double dmin = tdk::inv(hi), dmax = tdk::inv(lo);
            const double xk = ((double)x - key.ox) / key.fx;   // camera.rs:55-57
            const double yk = ((double)y - key.oy) / key.fy;

            // step_ratio, semi_dense.rs:27-40 + depth.rs:6-15
            const double kd = tdk::inv(idp);
            const double rz = T[8] * (xk * kd) + T[9] * (yk * kd) + T[10] * (1. * kd) + T[11];
            if (rz <= 0.) { flag = TDK_FLAG_NEGATIVE_REF_DEPTH; break; }
            const double kstep = (idp / tdk::inv(rz)) * a.step;

            // calc_ref_ends, semi_dense.rs:51-60
            double xminx, xminy, xmaxx, xmaxy, zz;
            tdk::warp_point(T, xk, yk, dmin, xminx, xminy, zz);
            tdk::warp_point(T, xk, yk, dmax, xmaxx, xmaxy, zz);
            const double rdx = xmaxx - xminx, rdy = xmaxy - xminy;

            // calc_key_direction, semi_dense.rs:42-49 ; key_coordinates, epipolar.rs:22-36
            double kdx = xk - pc->ex, kdy = yk - pc->ey;
            if (!(rdx * kdx + rdy * kdy > 0.)) { kdx = -kdx; kdy = -kdy; }
            const double knorm = sqrt(kdx * kdx + kdy * kdy);
            if (knorm != 0.) { kdx = kdx / knorm; kdy = kdy / knorm; }
            double ki[5], kcx[5], kcy[5];
            bool key_ok = true;
#pragma unroll
            for (int k = 0; k < 5; k++) {
                const double s = kstep * (double)(k - 2);
                kcx[k] = (xk + s * kdx) * key.fx + key.ox;
                kcy[k] = (yk + s * kdy) * key.fy + key.oy;
                key_ok = key_ok && tdk::in_range(kcx[k], kcy[k], wm1, hm1);
            }
            if (!key_ok) { flag = TDK_FLAG_KEY_OUT_OF_RANGE; break; }
#pragma unroll
            for (int k = 0; k < 5; k++) ki[k] = tdk::bilinear(key.img, W, H, kcx[k], kcy[k]);
            // gradient1d + norm, gradient.rs:28-35, semi_dense.rs:130
            const double g0 = ki[1] - ki[0], g1 = ki[2] - ki[1], g2 = ki[3] - ki[2], g3 = ki[4] - ki[3];
            const double g = sqrt(g0 * g0 + g1 * g1 + g2 * g2 + g3 * g3);
            if (g < a.min_grad) { flag = TDK_FLAG_INSUFFICIENT_GRADIENT; break; }

            // ref_coordinates + check_us_ref, epipolar.rs:38-54, semi_dense.rs:62-81
            const double L = sqrt(rdx * rdx + rdy * rdy);
            const double ux = rdx / (L + TDK_EPS), uy = rdy / (L + TDK_EPS);
            const double nf = L / a.step;
            int n = 0;
            if (nf > 0.) n = nf >= 2147483647. ? 2147483647 : (int)nf;   // Rust `as usize`
            if (n < 5) { flag = TDK_FLAG_REF_EPIPOLAR_TOO_SHORT; break; }
            {
                const double fx0 = (xminx + (0. * a.step) * ux) * pc->rfx + pc->rox;
                const double fy0 = (xminy + (0. * a.step) * uy) * pc->rfy + pc->roy;
                if (!tdk::in_range(fx0, fy0, wm1, hm1)) { flag = TDK_FLAG_REF_CLOSE_OUT_OF_RANGE; break; }
                const double sl = (double)(n - 1) * a.step;
                const double fx1 = (xminx + sl * ux) * pc->rfx + pc->rox;
                const double fy1 = (xminy + sl * uy) * pc->rfy + pc->roy;
                if (!tdk::in_range(fx1, fy1, wm1, hm1)) { flag = TDK_FLAG_REF_FAR_OUT_OF_RANGE; break; }
            }
            if (n > TDK_MAX_SEARCH_SAMPLES) {
                atomicExch(&a.status[TDK_ST_SEARCH_LEN], 1);
                n = TDK_MAX_SEARCH_SAMPLES;
            }
            // survivor: normalise the key window once (intensities.rs:20 does it per window)
            const double kn = sqrt(ki[0] * ki[0] + ki[1] * ki[1] + ki[2] * ki[2] + ki[3] * ki[3] + ki[4] * ki[4]);
            const int slot = atomicAdd(&s_nq, 1);
#pragma unroll
            for (int k = 0; k < 5; k++) q_k[k][slot] = kn == 0. ? ki[k] : ki[k] / kn;
            q_xminx[slot] = xminx; q_xminy[slot] = xminy;
            q_ux[slot] = ux; q_uy[slot] = uy;
            q_rdx[slot] = rdx; q_rdy[slot] = rdy;
            q_kstep[slot] = kstep; q_g[slot] = g;
            q_n[slot] = n;
            q_pix[slot] = (unsigned short)tid;
            q_ref[slot] = (unsigned short)r;
            out_v = v;
            flag = 1;   // pending
        } while (false);
        s_depth[tid] = out_d;
        s_var[tid] = out_v;
    }
    s_flag[tid] = (signed char)flag;
    __syncthreads();

    // ------------------------------------------------------------------ phase B
    const int nq = s_nq;
    {
        constexpr int NGRP = BLOCK / G;
        const int lane = tid & (G - 1);
        const int grp = tid / G;
        const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << ((tid & 31) & ~(G - 1)));
        for (int q = grp; q < nq; q += NGRP) {
            const int n = q_n[q];
            const TdkPairConst *__restrict__ pc = a.pairs + (size_t)p * a.R + q_ref[q];
            const double *__restrict__ ref = pc->ref;
            const double rfx = pc->rfx, rfy = pc->rfy, rox = pc->rox, roy = pc->roy;
            const double xminx = q_xminx[q], xminy = q_xminy[q], ux = q_ux[q], uy = q_uy[q];
            const double k0 = q_k[0][q], k1 = q_k[1][q], k2 = q_k[2][q], k3 = q_k[3][q], k4 = q_k[4][q];
            const int nwin = n - 4;
            double best_e = INFINITY;
            int best_i = 0x7fffffff;
            for (int base = 0; base < nwin; base += G - 4) {
                const int i = base + lane;
                double s0 = 0.;
                if (i < n) {
                    const double s = (double)i * a.step;
                    const double cx = (xminx + s * ux) * rfx + rox;
                    const double cy = (xminy + s * uy) * rfy + roy;
                    s0 = tdk::bilinear(ref, W, H, cx, cy);
                }
                const double s1 = __shfl_down_sync(gmask, s0, 1, G);
                const double s2 = __shfl_down_sync(gmask, s0, 2, G);
                const double s3 = __shfl_down_sync(gmask, s0, 3, G);
                const double s4 = __shfl_down_sync(gmask, s0, 4, G);
                if (lane < G - 4 && i < nwin) {
                    // intensities.rs:6-29: || normalize(window) - normalize(kernel) ||^2
                    const double wn = sqrt(s0 * s0 + s1 * s1 + s2 * s2 + s3 * s3 + s4 * s4);
                    double w0 = s0, w1 = s1, w2 = s2, w3 = s3, w4 = s4;
                    if (wn != 0.) { w0 = s0 / wn; w1 = s1 / wn; w2 = s2 / wn; w3 = s3 / wn; w4 = s4 / wn; }
                    const double d0 = w0 - k0, d1 = w1 - k1, d2 = w2 - k2, d3 = w3 - k3, d4 = w4 - k4;
                    const double e = d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3 + d4 * d4;
                    if (e < best_e) { best_e = e; best_i = i; }
                }
            }
#pragma unroll
            for (int off = G / 2; off > 0; off >>= 1) {
                const double oe = __shfl_xor_sync(gmask, best_e, off, G);
                const int oi = __shfl_xor_sync(gmask, best_i, off, G);
                if (oe < best_e || (oe == best_e && oi < best_i)) { best_e = oe; best_i = oi; }
            }
            if (lane == 0) q_arg[q] = (best_i == 0x7fffffff ? 0 : best_i) + 2;   // intensities.rs:32-36
        }
    }
    __syncthreads();

    // ------------------------------------------------------------------ phase C
    if (tid < nq) {
        const int q = tid;
        const int owner = q_pix[q];
        const int arg = q_arg[q];
        const int sx = tx * TILE_W + (owner & (TILE_W - 1));
        const int sy = ty * TILE_H + (owner / TILE_W);
        const TdkPairConst *__restrict__ pc = a.pairs + (size_t)p * a.R + q_ref[q];
        const double *__restrict__ T = pc->T;
        const double xk = ((double)sx - key.ox) / key.fx;
        const double yk = ((double)sy - key.oy) / key.fy;
        const double s = (double)arg * a.step;
        const double xr = q_xminx[q] + s * q_ux[q];
        const double yr = q_xminy[q] + s * q_uy[q];

        // calc_depth0, triangulation.rs:8-36
        const double rzy = T[8] * xk + T[9] * yk + T[10] * 1.;
        const double r0y = T[0] * xk + T[1] * yk + T[2] * 1.;
        const double r1y = T[4] * xk + T[5] * yk + T[6] * 1.;
        const double tzz = T[11];
        const bool ax0 = fabs(T[3]) > fabs(T[7]);
        const double x1i = ax0 ? xr : yr, ti = ax0 ? T[3] : T[7], riy = ax0 ? r0y : r1y;
        const double depth = (ti - tzz * x1i) / ((rzy * x1i - riy) + TDK_EPS);

        // calc_alpha, variance.rs:55-106
        const double rdx = q_rdx[q], rdy = q_rdy[q];
        const double rn = sqrt(rdx * rdx + rdy * rdy);
        const double dhx = rn == 0. ? rdx : rdx / rn, dhy = rn == 0. ? rdy : rdy / rn;
        double xrefx, xrefy, zz;
        tdk::warp_point(T, xk, yk, depth, xrefx, xrefy, zz);
        const bool bx0 = fabs(dhx) > fabs(dhy);
        const double xref_i = bx0 ? xrefx : xrefy, dir_i = bx0 ? dhx : dhy;
        const double tj = bx0 ? T[3] : T[7], rjy = bx0 ? r0y : r1y;
        const double ad = rzy * tj - rjy * tzz;
        const double an = xref_i * tzz - tj;
        const double alpha = dir_i * ad / (an * an);

        // geo_var, variance.rs:30-53 ; Sobel at the integer pixel, border = 0
        double gx = 0., gy = 0.;
        if (sx > 0 && sy > 0 && sx < W - 1 && sy < H - 1) {
            const double *c = key.img + (size_t)sy * W + sx;
            const double m00 = __ldg(c - W - 1), m01 = __ldg(c - W), m02 = __ldg(c - W + 1);
            const double m10 = __ldg(c - 1), m11 = __ldg(c), m12 = __ldg(c + 1);
            const double m20 = __ldg(c + W - 1), m21 = __ldg(c + W), m22 = __ldg(c + W + 1);
            const double ex[9] = {1. * m00, 0. * m01, -1. * m02, 2. * m10, 0. * m11, -2. * m12,
                                  1. * m20, 0. * m21, -1. * m22};
            const double ey[9] = {1. * m00, 2. * m01, 1. * m02, 0. * m10, 0. * m11, 0. * m12,
                                  -1. * m20, -2. * m21, -1. * m22};
            gx = fold9(ex);
            gy = fold9(ey);
        }
        double lx = xk - pc->ptx, ly = yk - pc->pty;
        const double ln = sqrt(lx * lx + ly * ly);
        if (ln != 0.) { lx = lx / ln; ly = ly / ln; }
        const double gn = sqrt(gx * gx + gy * gy);
        if (gn != 0.) { gx = gx / gn; gy = gy / gn; }
        const double pp = lx * gx + ly * gy;
        const double geo_var = pp == 0. ? 1. / TDK_EPS : 1. / (pp * pp);
        const double photo_var = 2. / (q_g[q] / q_kstep[q]);
        const double variance = (alpha * alpha) * ((a.geo * a.geo) * geo_var + (a.photo * a.photo) * photo_var);

        const double idn = tdk::inv(depth);
        const int f = check_args(idn, variance, a.vmin, a.vmax);
        if (f == TDK_FLAG_SUCCESS) {
            double nd = tdk::inv(idn), nv = variance;
            if (a.fuse) {   // fusion.rs:3-11 in inverse depth, as propagation.rs:27-34
                const size_t pg = map_base + (size_t)sy * W + sx;
                const double pd = a.prior_d[pg], pv = a.prior_v[pg];
                const double ia = tdk::inv(pd), ib = tdk::inv(nd);
                const double vs = pv + nv;
                nd = tdk::inv((ia * nv + ib * pv) / vs);
                nv = (pv * nv) / vs;
            }
            s_depth[owner] = nd;
            s_var[owner] = nv;
        }
        s_flag[owner] = (signed char)f;
        s_arg[owner] = arg;
    }
    __syncthreads();

    // ------------------------------------------------------------------ store
    if (valid) {
        const int f = s_flag[tid];
        a.out_d[pix] = s_depth[tid];
        a.out_v[pix] = s_var[tid];
        a.out_f[pix] = (long long)f;
        if (a.dbg) a.dbg[pix] = s_arg[tid];
        if (a.hist) atomicAdd(&s_hist[-f], 1u);
    }
    if (a.hist) {
        __syncthreads();
        if (tid < 10 && s_hist[tid]) atomicAdd(&a.hist[(size_t)p * 10 + tid], (unsigned long long)s_hist[tid]);
    }
}

// ---------------------------------------------------------------------------------------------
size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct WsLayout {
    size_t status, keys, pairs, key_ptrs, ref_ptrs, total;
};

WsLayout ws_layout(int64_t P, int64_t R)
{
    WsLayout l;
    size_t off = 0;
    l.status = off; off = align_up(off + TDK_ST_WORDS * sizeof(int), 256);
    l.keys = off; off = align_up(off + (size_t)P * sizeof(TdkKeyConst), 256);
    l.pairs = off; off = align_up(off + (size_t)P * R * sizeof(TdkPairConst), 256);
    l.key_ptrs = off; off = align_up(off + (size_t)P * sizeof(void *), 256);
    l.ref_ptrs = off; off = align_up(off + (size_t)P * R * sizeof(void *), 256);
    l.total = off;
    return l;
}

int g_search_lanes = 16;

int launch(const double *key_imgs, const double *const *key_ptrs_host, const double *key_cams,
           const double *T_wk, const double *ref_imgs, const double *const *ref_ptrs_host,
           const double *ref_cams, const double *T_wr, const uint64_t *age, const double *prior_depth,
           const double *prior_variance, const double *params, int64_t P, int64_t R, int64_t H,
           int64_t W, double *out_depth, double *out_variance, int64_t *out_flag, int32_t *dbg_argmin,
           uint64_t *out_hist, int fuse, void *workspace, size_t workspace_bytes, cudaStream_t stream)
{
    if (!key_cams || !T_wk || !ref_cams || !T_wr || !age || !prior_depth || !prior_variance ||
        !params || !out_depth || !out_variance || !out_flag || !workspace ||
        (!key_imgs && !key_ptrs_host) || (!ref_imgs && !ref_ptrs_host)) {
        tdk_set_error("tdk_update_depth: null pointer argument");
        return TDK_ERR_INVALID;
    }
    if (P <= 0 || R <= 0 || H <= 0 || W <= 0 || R > 65535 || H > 32767 || W > 32767) {
        tdk_set_error("tdk_update_depth: bad sizes P=%lld R=%lld H=%lld W=%lld (need 1<=R<=65535, H,W<=32767)",
                      (long long)P, (long long)R, (long long)H, (long long)W);
        return TDK_ERR_INVALID;
    }
    if (!(params[4] > 0.)) {
        tdk_set_error("tdk_update_depth: ref_step_size must be > 0 (got %g)", params[4]);
        return TDK_ERR_INVALID;
    }
    const WsLayout l = ws_layout(P, R);
    if (workspace_bytes < l.total) {
        tdk_set_error("tdk_update_depth: workspace %zu < required %zu", workspace_bytes, l.total);
        return TDK_ERR_WORKSPACE;
    }
    const int tiles_x = (int)((W + TILE_W - 1) / TILE_W), tiles_y = (int)((H + TILE_H - 1) / TILE_H);
    const long long nblocks = (long long)P * tiles_x * tiles_y;
    if (nblocks > 2147483647LL || P * R > 2147483647LL) {
        tdk_set_error("tdk_update_depth: batch too large for one launch (%lld tiles)", nblocks);
        return TDK_ERR_INVALID;
    }
    char *ws = (char *)workspace;
    int *status = (int *)(ws + l.status);
    TdkKeyConst *keys = (TdkKeyConst *)(ws + l.keys);
    TdkPairConst *pairs = (TdkPairConst *)(ws + l.pairs);
    const double *const *key_ptrs = nullptr, *const *ref_ptrs = nullptr;
    TDK_CUDA(cudaMemsetAsync(status, 0, TDK_ST_WORDS * sizeof(int), stream));
    if (key_ptrs_host) {
        TDK_CUDA(cudaMemcpyAsync(ws + l.key_ptrs, key_ptrs_host, (size_t)P * sizeof(void *),
                                 cudaMemcpyHostToDevice, stream));
        key_ptrs = (const double *const *)(ws + l.key_ptrs);
    }
    if (ref_ptrs_host) {
        TDK_CUDA(cudaMemcpyAsync(ws + l.ref_ptrs, ref_ptrs_host, (size_t)P * R * sizeof(void *),
                                 cudaMemcpyHostToDevice, stream));
        ref_ptrs = (const double *const *)(ws + l.ref_ptrs);
    }
    const int npairs = (int)(P * R);
    pair_setup_kernel<<<(npairs + 127) / 128, 128, 0, stream>>>(
        key_imgs, key_ptrs, key_cams, T_wk, ref_imgs, ref_ptrs, ref_cams, T_wr, (int)P, (int)R,
        (size_t)H * W, keys, pairs, status);
    TDK_CUDA(cudaGetLastError());

    UpdArgs a;
    a.keys = keys; a.pairs = pairs; a.age = age; a.prior_d = prior_depth; a.prior_v = prior_variance;
    a.out_d = out_depth; a.out_v = out_variance; a.out_f = (long long *)out_flag; a.dbg = dbg_argmin;
    a.hist = (unsigned long long *)out_hist; a.status = status;
    // Params::new stores (inv(max_depth), inv(min_depth))  (py/semi_dense.rs:102-107)
    a.vmin = 1. / (params[1] + TDK_EPS_INV);
    a.vmax = 1. / (params[0] + TDK_EPS_INV);
    a.geo = params[2]; a.photo = params[3]; a.step = params[4]; a.min_grad = params[5];
    a.P = (int)P; a.R = (int)R; a.H = (int)H; a.W = (int)W; a.tiles_x = tiles_x; a.tiles_y = tiles_y; a.fuse = fuse ? 1 : 0;
    switch (g_search_lanes) {
        case 8: update_depth_kernel<8><<<(unsigned)nblocks, BLOCK, 0, stream>>>(a); break;
        case 32: update_depth_kernel<32><<<(unsigned)nblocks, BLOCK, 0, stream>>>(a); break;
        default: update_depth_kernel<16><<<(unsigned)nblocks, BLOCK, 0, stream>>>(a); break;
    }
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// Built-in scratch for tdk_update_depth_static_dev: the same bytes tdk_update_depth_dev expects in
// its caller-provided workspace (status words, per-keyframe and per-pair constants), for batches
// of at most TDK_STATIC_MAX_PAIRS keyframe/reference pairs.  Every device that loads this library
// carries one copy of the array until the process exits.
static __device__ __align__(256) unsigned char g_static_scratch[TDK_STATIC_SCRATCH_BYTES];

namespace {

std::mutex g_static_mutex;
void *g_static_scratch_addr[64] = {};   // per device ordinal; filled on the first call there

// Host code cannot hand `g_static_scratch` to a kernel: on the host that name only identifies the
// symbol to the runtime.  The address the device sees is asked for once per device, with that
// device current, and cached.
int static_scratch_address(void **out)
{
    int dev = -1;
    TDK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) {
        tdk_set_error("tdk_update_depth_static_dev: device ordinal %d not supported", dev);
        return TDK_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_scratch_addr[dev]) {
        void *addr = nullptr;
        TDK_CUDA(cudaGetSymbolAddress(&addr, g_static_scratch));
        if (!addr) {
            tdk_set_error("tdk_update_depth_static_dev: cudaGetSymbolAddress returned a null address");
            return TDK_ERR_CUDA;
        }
        g_static_scratch_addr[dev] = addr;
    }
    *out = g_static_scratch_addr[dev];
    return TDK_OK;
}

}  // namespace

extern "C" {

size_t tdk_update_depth_workspace_bytes(int64_t P, int64_t R)
{
    if (P <= 0 || R <= 0) return 0;
    return ws_layout(P, R).total;
}

// tuning knob (not part of the reference surface): lanes that cooperate on one epipolar line
int tdk_set_search_lanes(int lanes)
{
    if (lanes != 8 && lanes != 16 && lanes != 32) {
        tdk_set_error("tdk_set_search_lanes: %d not in {8,16,32}", lanes);
        return TDK_ERR_INVALID;
    }
    g_search_lanes = lanes;
    return TDK_OK;
}

int tdk_update_depth_dev(const double *key_imgs, const double *key_cams, const double *T_wk,
                         const double *ref_imgs, const double *ref_cams, const double *T_wr,
                         const uint64_t *age, const double *prior_depth, const double *prior_variance,
                         const double *params, int64_t P, int64_t R, int64_t H, int64_t W,
                         double *out_depth, double *out_variance, int64_t *out_flag,
                         int32_t *dbg_argmin, uint64_t *out_hist, int fuse_on_success, void *workspace,
                         size_t workspace_bytes, void *stream)
{
    return launch(key_imgs, nullptr, key_cams, T_wk, ref_imgs, nullptr, ref_cams, T_wr, age,
                  prior_depth, prior_variance, params, P, R, H, W, out_depth, out_variance, out_flag,
                  dbg_argmin, out_hist, fuse_on_success, workspace, workspace_bytes, (cudaStream_t)stream);
}

int tdk_update_depth_frames_dev(const double *const *key_img_ptrs, const double *key_cams,
                                const double *T_wk, const double *const *ref_img_ptrs,
                                const double *ref_cams, const double *T_wr, const uint64_t *age,
                                const double *prior_depth, const double *prior_variance,
                                const double *params, int64_t P, int64_t R, int64_t H, int64_t W,
                                double *out_depth, double *out_variance, int64_t *out_flag,
                                int32_t *dbg_argmin, uint64_t *out_hist, int fuse_on_success,
                                void *workspace, size_t workspace_bytes, void *stream)
{
    return launch(nullptr, key_img_ptrs, key_cams, T_wk, nullptr, ref_img_ptrs, ref_cams, T_wr, age,
                  prior_depth, prior_variance, params, P, R, H, W, out_depth, out_variance, out_flag,
                  dbg_argmin, out_hist, fuse_on_success, workspace, workspace_bytes, (cudaStream_t)stream);
}

int tdk_update_depth_status(const void *workspace, void *stream)
{
    if (!workspace) { tdk_set_error("tdk_update_depth_status: null workspace"); return TDK_ERR_INVALID; }
    int st[TDK_ST_WORDS];
    TDK_CUDA(cudaMemcpyAsync(st, workspace, sizeof(st), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    TDK_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    if (st[TDK_ST_SINGULAR]) { tdk_set_error("a reference pose T_wr is singular"); return TDK_ERR_SINGULAR; }
    if (st[TDK_ST_AGE]) { tdk_set_error("Age exceeds the refframe size"); return TDK_ERR_AGE; }
    if (st[TDK_ST_SEARCH_LEN]) {
        tdk_set_error("an epipolar search line has more than %d samples; increase ref_step_size",
                      TDK_MAX_SEARCH_SAMPLES);
        return TDK_ERR_SEARCH_LEN;
    }
    return TDK_OK;
}

int tdk_update_depth_static_dev(const double *key_imgs, const double *key_cams, const double *T_wk,
                                const double *ref_imgs, const double *ref_cams, const double *T_wr,
                                const uint64_t *age, const double *prior_depth,
                                const double *prior_variance, const double *params, int64_t P, int64_t R,
                                int64_t H, int64_t W, double *out_depth, double *out_variance,
                                int64_t *out_flag, int32_t *dbg_argmin, uint64_t *out_hist,
                                int fuse_on_success, int wait_and_check, void *stream)
{
    if (P <= 0 || R <= 0 || P * R > TDK_STATIC_MAX_PAIRS ||
        ws_layout(P, R).total > (size_t)TDK_STATIC_SCRATCH_BYTES) {
        tdk_set_error("tdk_update_depth_static_dev: P*R = %lld x %lld exceeds the built-in scratch; this entry "
                      "point takes at most %d keyframe/reference pairs per call (use tdk_update_depth_dev "
                      "with a caller-owned workspace for larger batches)",
                      (long long)P, (long long)R, TDK_STATIC_MAX_PAIRS);
        return TDK_ERR_INVALID;
    }
    void *scratch = nullptr;
    int rc = static_scratch_address(&scratch);
    if (rc != TDK_OK) return rc;
    // launch() zeroes the status words at the head of the scratch on `stream` before the kernels run,
    // so nothing a previous call left behind is read; everything else in it is rewritten by
    // pair_setup_kernel.
    rc = launch(key_imgs, nullptr, key_cams, T_wk, ref_imgs, nullptr, ref_cams, T_wr, age, prior_depth,
                prior_variance, params, P, R, H, W, out_depth, out_variance, out_flag, dbg_argmin, out_hist,
                fuse_on_success, scratch, (size_t)TDK_STATIC_SCRATCH_BYTES, (cudaStream_t)stream);
    if (rc != TDK_OK || !wait_and_check) return rc;
    return tdk_update_depth_status(scratch, stream);
}

}  // extern "C"
