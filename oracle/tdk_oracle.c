/*
 * tdk_oracle.c -- CPU restatement of Tadataka's semi-dense inverse-depth path.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the parity oracle: a plain-C,
 * function-by-function restatement of the reference's Rust sources
 * (src/semi_dense/ *.rs and the src/ *.rs helpers they call).  Only tests/,
 * __graft_entry__.smoke() and the cpu_baseline / --impl reference legs of
 * bench.py may load it.  The product (tadataka_b200/) never does.
 *
 * Parity status: PINNED against the reference's own known-answer tests
 * (every #[test] vector on the path, replayed in tests/test_oracle_golden.py)
 * and against the five replayable flag cases of
 * tests/vo/semi_dense/test_semi_dense.py on the NewTsukuba fixture
 * (tests/golden/).  The reference itself (Rust nightly + pyo3 + OpenBLAS) cannot
 * be built in this image, so there is no oracle/_ref for the hot path.
 *
 * Arithmetic: IEEE f64, no FMA contraction (-ffp-contract=off), sums evaluated
 * in the order the reference's libraries evaluate them:
 *   - ndarray 0.13 `dot` on short vectors and `Norm::norm_l2` : left-to-right;
 *   - ndarray 0.13 `sum()` on a contiguous 3x3 (Sobel)         : the 8-lane
 *     unrolled fold (p0+p4)+(p1+p5)+(p2+p6)+(p3+p7)+e8;
 *   - the 4x4 `inv()` (LAPACK getrf/getri via ndarray-linalg 0.12, OpenBLAS
 *     0.9 -- a dependency that is not under /root/reference): restated from
 *     the netlib reference algorithm (dgetf2 + dtrti2 + dgetri unblocked);
 *   - 4x4 * 4 products that the reference hands to BLAS dgemv: left-to-right.
 * Where OpenBLAS' internal order is unknowable the result can differ from the
 * Rust build in the last ulp; the north star's 1e-4 tolerance covers that.
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORA_EPS_INV 2.220446049250313e-16 /* src/numeric.rs:1 (f64::EPSILON) */
#define ORA_EPS 1e-16 /* projection.rs:4, epipolar.rs:7, triangulation.rs:5, variance.rs:8 */
#define ORA_VARIANCE_FACTOR 2.0 /* src/semi_dense/hypothesis.rs:5 */
#define ORA_STAT_FACTOR 2.0 /* src/semi_dense/stat.rs:3 */

/* src/semi_dense/flag.rs:3-14 */
enum {
    ORA_SUCCESS = 0,
    ORA_HYPOTHESIS_OUT_OF_SEARCH_RANGE = -1,
    ORA_KEY_OUT_OF_RANGE = -2,
    ORA_REF_CLOSE_OUT_OF_RANGE = -3,
    ORA_REF_FAR_OUT_OF_RANGE = -4,
    ORA_REF_EPIPOLAR_TOO_SHORT = -5,
    ORA_INSUFFICIENT_GRADIENT = -6,
    ORA_NEGATIVE_PRIOR_DEPTH = -7,
    ORA_NEGATIVE_REF_DEPTH = -8,
    ORA_NOT_PROCESSED = -9
};

/* camera = {fx, fy, ox, oy}  (src/camera.rs:4-15)
 * params = {vmin, vmax, geo, photo, ref_step, min_gradient}
 *          (src/semi_dense/params.rs:5-11; vmin/vmax already inverted as
 *           src/py/semi_dense.rs:102-107 does) */

/* ---------------------------------------------------------------- numeric */

/* src/numeric.rs:3-5 */
double ora_safe_invert(double v) { return 1. / (v + ORA_EPS_INV); }

/* src/cmp.rs:3-12 */
double ora_clamp(double v, double lo, double hi)
{
    if (v < lo) return lo;
    if (v > hi) return hi;
    return v;
}

/* ndarray `dot` on a short 1-D pair (len < 8): sum = 0; sum += a[i]*b[i] */
static double dot_n(int n, const double *a, int sa, const double *b, int sb)
{
    double s = 0.;
    for (int i = 0; i < n; i++) s = s + a[i * sa] * b[i * sb];
    return s;
}

/* ndarray_linalg Norm::norm (L2): sqrt(sum of squares, left to right) */
double ora_norm(int n, const double *v)
{
    double s = 0.;
    for (int i = 0; i < n; i++) s = s + v[i] * v[i];
    return sqrt(s);
}

/* src/vector.rs:4-11 */
void ora_normalize(int n, const double *v, double *out)
{
    double norm = ora_norm(n, v);
    if (norm == 0.) {
        for (int i = 0; i < n; i++) out[i] = v[i];
        return;
    }
    for (int i = 0; i < n; i++) out[i] = v[i] / norm;
}

/* ------------------------------------------------- projection / transform */

/* src/projection.rs:11-14 */
void ora_project(const double *p, double *x)
{
    double z = p[2] + ORA_EPS;
    x[0] = p[0] / z;
    x[1] = p[1] / z;
}

/* src/projection.rs:16-18, src/homogeneous.rs:14-16 */
void ora_inv_project(const double *x, double depth, double *p)
{
    p[0] = x[0] * depth;
    p[1] = x[1] * depth;
    p[2] = 1. * depth;
}

/* src/transform.rs:17-23 : T * (p, 1), last row dropped */
void ora_transform(const double *T, const double *p, double *q)
{
    double h[4] = {p[0], p[1], p[2], 1.};
    for (int i = 0; i < 3; i++) q[i] = dot_n(4, T + 4 * i, 1, h, 1);
}

/* src/warp.rs:18-28 */
void ora_warp(const double *T, const double *x0, double depth0, double *x1, double *depth1)
{
    double p0[3], p1[3];
    ora_inv_project(x0, depth0, p0);
    ora_transform(T, p0, p1);
    ora_project(p1, x1);
    *depth1 = p1[2];
}

/* src/camera.rs:55-61 */
void ora_cam_normalize(const double *cam, const double *u, double *x)
{
    x[0] = (u[0] - cam[2]) / cam[0];
    x[1] = (u[1] - cam[3]) / cam[1];
}
void ora_cam_unnormalize(const double *cam, const double *x, double *u)
{
    u[0] = x[0] * cam[0] + cam[2];
    u[1] = x[1] * cam[1] + cam[3];
}

/* src/warp.rs:72-85 */
void ora_perspective_warp(const double *T10, const double *cam0, const double *cam1,
                          const double *u0, double depth0, double *u1, double *depth1)
{
    double x0[2], x1[2];
    ora_cam_normalize(cam0, u0, x0);
    ora_warp(T10, x0, depth0, x1, depth1);
    ora_cam_unnormalize(cam1, x1, u1);
}

/* General 4x4 inverse, restating what ndarray_linalg's `inv()` asks LAPACK for
 * (call site src/semi_dense/semi_dense.rs:87): LU with partial pivoting
 * (netlib dgetf2), inverse of U (dtrti2), then inv(A)*L = inv(U) (dgetri,
 * unblocked) and the column interchanges.  Returns 0, or j+1 if U[j][j] == 0. */
int ora_mat4_inverse(const double *A_in, double *A)
{
    int ipiv[4], info = 0;
    memcpy(A, A_in, 16 * sizeof(double));
    for (int j = 0; j < 4; j++) {
        int p = j;
        double best = fabs(A[4 * j + j]);
        for (int i = j + 1; i < 4; i++)
            if (fabs(A[4 * i + j]) > best) { best = fabs(A[4 * i + j]); p = i; }
        ipiv[j] = p;
        if (A[4 * p + j] != 0.) {
            if (p != j)
                for (int k = 0; k < 4; k++) {
                    double t = A[4 * j + k]; A[4 * j + k] = A[4 * p + k]; A[4 * p + k] = t;
                }
            if (fabs(A[4 * j + j]) >= 2.2250738585072014e-308) {
                double r = 1. / A[4 * j + j];
                for (int i = j + 1; i < 4; i++) A[4 * i + j] = A[4 * i + j] * r;
            } else {
                for (int i = j + 1; i < 4; i++) A[4 * i + j] = A[4 * i + j] / A[4 * j + j];
            }
        } else if (info == 0) {
            info = j + 1;
        }
        for (int k = j + 1; k < 4; k++)          /* dger, column by column */
            for (int i = j + 1; i < 4; i++)
                A[4 * i + k] = A[4 * i + k] - A[4 * i + j] * A[4 * j + k];
    }
    if (info) return info;
    /* dtrti2, upper, non-unit */
    for (int j = 0; j < 4; j++) {
        A[4 * j + j] = 1. / A[4 * j + j];
        double ajj = -A[4 * j + j];
        /* dtrmv('U','N','N') on the leading j x j block applied to column j */
        for (int jj = 0; jj < j; jj++) {
            if (A[4 * jj + j] != 0.) {
                double temp = A[4 * jj + j];
                for (int i = 0; i < jj; i++) A[4 * i + j] = A[4 * i + j] + temp * A[4 * i + jj];
                A[4 * jj + j] = A[4 * jj + j] * A[4 * jj + jj];
            }
        }
        for (int i = 0; i < j; i++) A[4 * i + j] = ajj * A[4 * i + j];
    }
    /* dgetri unblocked: solve inv(A)*L = inv(U) */
    for (int j = 3; j >= 0; j--) {
        double work[4] = {0., 0., 0., 0.};
        for (int i = j + 1; i < 4; i++) { work[i] = A[4 * i + j]; A[4 * i + j] = 0.; }
        for (int jj = j + 1; jj < 4; jj++) {    /* dgemv('N', alpha=-1, beta=1) */
            double temp = -1. * work[jj];
            for (int i = 0; i < 4; i++) A[4 * i + j] = A[4 * i + j] + temp * A[4 * i + jj];
        }
    }
    for (int j = 2; j >= 0; j--) {
        int jp = ipiv[j];
        if (jp != j)
            for (int i = 0; i < 4; i++) {
                double t = A[4 * i + j]; A[4 * i + j] = A[4 * i + jp]; A[4 * i + jp] = t;
            }
    }
    return 0;
}

/* ndarray general 4x4 * 4x4 `dot` */
void ora_mat4_mul(const double *A, const double *B, double *C)
{
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) C[4 * i + j] = dot_n(4, A + 4 * i, 1, B + j, 4);
}

/* src/semi_dense/semi_dense.rs:83-89 */
int ora_transform_rk(const double *T_wk, const double *T_wr, double *T_rk)
{
    double T_rw[16];
    int info = ora_mat4_inverse(T_wr, T_rw);
    ora_mat4_mul(T_rw, T_wk, T_rk);
    return info;
}

/* src/transform.rs:71-79 : [R^T | -(R^T t)] */
void ora_inv_transform(const double *T, double *out)
{
    double rt[3];
    for (int i = 0; i < 3; i++) {
        double col[3] = {T[0 * 4 + i], T[1 * 4 + i], T[2 * 4 + i]};
        double t[3] = {T[3], T[7], T[11]};
        rt[i] = dot_n(3, col, 1, t, 1);
    }
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) out[4 * i + j] = T[4 * j + i];
        out[4 * i + 3] = -rt[i];
    }
    out[12] = 0.; out[13] = 0.; out[14] = 0.; out[15] = 1.;
}

/* ---------------------------------------------------------------- epipolar */

/* src/semi_dense/epipolar.rs:9-20 */
void ora_calc_key_epipole(const double *T_wk, const double *T_wr, double *e_key)
{
    double T_kw[16], diff[3], p[3];
    ora_inv_transform(T_wk, T_kw);
    for (int i = 0; i < 3; i++) diff[i] = T_wr[4 * i + 3] - T_wk[4 * i + 3];
    for (int i = 0; i < 3; i++) p[i] = dot_n(3, T_kw + 4 * i, 1, diff, 1);
    ora_project(p, e_key);
}

/* src/semi_dense/epipolar.rs:22-36 ; out is [5][2] */
void ora_key_coordinates(const double *direction, const double *x_key, double step_size,
                         double *out)
{
    static const double steps[5] = {-2., -1., 0., 1., 2.};
    double d[2];
    ora_normalize(2, direction, d);
    for (int i = 0; i < 5; i++) {
        double s = step_size * steps[i];
        out[2 * i + 0] = x_key[0] + s * d[0];
        out[2 * i + 1] = x_key[1] + s * d[1];
    }
}

/* src/semi_dense/epipolar.rs:38-54, split so that no [n,2] array is needed:
 * ora_ref_line gives n and the unit direction, ora_ref_coordinate one row. */
static int64_t trunc_to_usize(double v)
{
    /* Rust `as usize`: NaN -> 0, negative -> 0, saturating */
    if (!(v > 0.)) return 0;
    if (v >= 9.2e18) return INT64_MAX;
    return (int64_t)v;
}
int64_t ora_ref_line(const double *direction, double step_size, double *unit)
{
    double norm = ora_norm(2, direction);
    unit[0] = direction[0] / (norm + ORA_EPS);
    unit[1] = direction[1] / (norm + ORA_EPS);
    return trunc_to_usize(norm / step_size);
}
void ora_ref_coordinate(const double *x_min, const double *unit, double step_size, int64_t i,
                        double *x)
{
    double s = (double)i * step_size;
    x[0] = x_min[0] + s * unit[0];
    x[1] = x_min[1] + s * unit[1];
}
/* array form used by the golden test (epipolar.rs:138-158); returns n */
int64_t ora_ref_coordinates(const double *x_min, const double *direction, double step_size,
                            double *out, int64_t cap)
{
    double unit[2];
    int64_t n = ora_ref_line(direction, step_size, unit);
    for (int64_t i = 0; i < n && i < cap; i++)
        ora_ref_coordinate(x_min, unit, step_size, i, out + 2 * i);
    return n;
}

/* ------------------------------------------------------- range / sampling */

/* src/image_range.rs:10-17 */
int ora_is_in_range(double x, double y, int64_t H, int64_t W)
{
    double h = (double)H, w = (double)W;
    return 0. <= x && x <= w - 1. && 0. <= y && y <= h - 1.;
}

/* src/interpolation.rs:9-43 ; caller guarantees the point is in range */
double ora_interpolate(const double *img, int64_t W, double cx, double cy)
{
    double lx = floor(cx), ly = floor(cy);
    int64_t lxi = (int64_t)lx, lyi = (int64_t)ly;
    if (lx == cx && ly == cy) return img[lyi * W + lxi];
    double ux = lx + 1., uy = ly + 1.;
    int64_t uxi = (int64_t)ux, uyi = (int64_t)uy;
    if (lx == cx)
        return img[lyi * W + lxi] * (ux - cx) * (uy - cy)
             + img[uyi * W + lxi] * (ux - cx) * (cy - ly);
    if (ly == cy)
        return img[lyi * W + lxi] * (ux - cx) * (uy - cy)
             + img[lyi * W + uxi] * (cx - lx) * (uy - cy);
    return img[lyi * W + lxi] * (ux - cx) * (uy - cy)
         + img[lyi * W + uxi] * (cx - lx) * (uy - cy)
         + img[uyi * W + lxi] * (ux - cx) * (cy - ly)
         + img[uyi * W + uxi] * (cx - lx) * (cy - ly);
}
void ora_interpolate_n(const double *img, int64_t W, const double *coords, int64_t n, double *out)
{
    for (int64_t i = 0; i < n; i++) out[i] = ora_interpolate(img, W, coords[2 * i], coords[2 * i + 1]);
}

/* src/gradient.rs:4-26 + src/convolution.rs:7-51 : 3x3 correlation, border 0.
 * The per-pixel sum is ndarray's `sum()` over the contiguous 3x3 product. */
static double fold9(const double *e)
{
    double p[8], acc = 0.;
    for (int i = 0; i < 8; i++) p[i] = 0. + e[i];
    acc = acc + (p[0] + p[4]);
    acc = acc + (p[1] + p[5]);
    acc = acc + (p[2] + p[6]);
    acc = acc + (p[3] + p[7]);
    acc = acc + e[8];
    return acc;
}
static const double SOBEL_X[9] = {1., 0., -1., 2., 0., -2., 1., 0., -1.};
static const double SOBEL_Y[9] = {1., 2., 1., 0., 0., 0., -1., -2., -1.};
static double correlate3(const double *img, int64_t W, int64_t y, int64_t x, const double *k)
{
    double e[9];
    for (int dy = 0; dy < 3; dy++)
        for (int dx = 0; dx < 3; dx++)
            e[3 * dy + dx] = k[3 * dy + dx] * img[(y - 1 + dy) * W + (x - 1 + dx)];
    return fold9(e);
}
void ora_sobel(const double *img, int64_t H, int64_t W, double *gx, double *gy)
{
    memset(gx, 0, sizeof(double) * H * W);
    memset(gy, 0, sizeof(double) * H * W);
    for (int64_t y = 1; y + 1 < H; y++)
        for (int64_t x = 1; x + 1 < W; x++) {
            gx[y * W + x] = correlate3(img, W, y, x, SOBEL_X);
            gy[y * W + x] = correlate3(img, W, y, x, SOBEL_Y);
        }
}

/* src/gradient.rs:28-35 */
void ora_gradient1d(const double *x, int n, double *g)
{
    for (int i = 0; i + 1 < n; i++) g[i] = x[i + 1] - x[i];
}

/* ------------------------------------------------------------- hypothesis */

/* src/semi_dense/hypothesis.rs:21-37 */
int ora_check_args(double inv_depth, double variance, double vmin, double vmax)
{
    if (inv_depth <= 0.) return ORA_NEGATIVE_PRIOR_DEPTH;
    double lo = inv_depth - ORA_VARIANCE_FACTOR * variance;
    double hi = inv_depth + ORA_VARIANCE_FACTOR * variance;
    if (hi <= vmin || vmax <= lo) return ORA_HYPOTHESIS_OUT_OF_SEARCH_RANGE;
    return ORA_SUCCESS;
}

/* src/semi_dense/hypothesis.rs:54-61 */
void ora_hypothesis_range(double inv_depth, double variance, double vmin, double vmax,
                          double *lo, double *hi)
{
    *lo = ora_clamp(inv_depth - ORA_VARIANCE_FACTOR * variance, vmin, vmax);
    *hi = ora_clamp(inv_depth + ORA_VARIANCE_FACTOR * variance, vmin, vmax);
}

/* src/semi_dense/depth.rs:25-30 */
void ora_depth_search_range(double min_inv, double max_inv, double *dmin, double *dmax)
{
    *dmin = ora_safe_invert(max_inv);
    *dmax = ora_safe_invert(min_inv);
}

/* src/semi_dense/depth.rs:6-15 */
double ora_calc_ref_depth(const double *T_rk, const double *x_key, double depth_key)
{
    double p[3];
    ora_inv_project(x_key, depth_key, p);
    return dot_n(3, T_rk + 8, 1, p, 1) + T_rk[11];
}

/* src/semi_dense/semi_dense.rs:27-40 ; returns flag, ratio through *out */
int ora_step_ratio(const double *T_rk, const double *x_key, double key_inv_depth, double *out)
{
    double key_depth = ora_safe_invert(key_inv_depth);
    double ref_depth = ora_calc_ref_depth(T_rk, x_key, key_depth);
    if (ref_depth <= 0.) return ORA_NEGATIVE_REF_DEPTH;
    double ref_inv_depth = ora_safe_invert(ref_depth);
    *out = key_inv_depth / ref_inv_depth;
    return ORA_SUCCESS;
}

/* src/semi_dense/semi_dense.rs:42-49 */
void ora_calc_key_direction(const double *x_key, const double *e_key, const double *ref_dir,
                            double *out)
{
    double d[2] = {x_key[0] - e_key[0], x_key[1] - e_key[1]};
    if (dot_n(2, ref_dir, 1, d, 1) > 0.) { out[0] = d[0]; out[1] = d[1]; }
    else { out[0] = -d[0]; out[1] = -d[1]; }
}

/* src/semi_dense/semi_dense.rs:51-60 */
void ora_calc_ref_ends(const double *T_rk, const double *x_key, double dmin, double dmax,
                       double *x_min_ref, double *x_max_ref)
{
    double z;
    ora_warp(T_rk, x_key, dmin, x_min_ref, &z);
    ora_warp(T_rk, x_key, dmax, x_max_ref, &z);
}

/* src/semi_dense/semi_dense.rs:62-81 ; us_ref given as its first and last row */
int ora_check_us_ref(int64_t n, const double *first, const double *last, int64_t us_key_size,
                     int64_t H, int64_t W)
{
    if (n < us_key_size) return ORA_REF_EPIPOLAR_TOO_SHORT;
    if (!ora_is_in_range(first[0], first[1], H, W)) return ORA_REF_CLOSE_OUT_OF_RANGE;
    if (!ora_is_in_range(last[0], last[1], H, W)) return ORA_REF_FAR_OUT_OF_RANGE;
    return ORA_SUCCESS;
}

/* ------------------------------------------------------- intensity search */

/* src/semi_dense/intensities.rs:6-36 */
int64_t ora_search(const double *sequence, int64_t n, const double *kernel, int k)
{
    double min_error = INFINITY;
    int64_t argmin = 0;
    double kn[8], wn[8];
    for (int64_t i = 0; i < n - k + 1; i++) {
        ora_normalize(k, sequence + i, wn);
        ora_normalize(k, kernel, kn);
        double e = 0.;
        for (int j = 0; j < k; j++) { double d = wn[j] - kn[j]; e = e + d * d; }
        if (e < min_error) { min_error = e; argmin = i; }
    }
    return argmin + (k / 2);
}

/* ------------------------------------------------------------ triangulation */

/* src/triangulation.rs:8-36 */
double ora_calc_depth0(const double *T10, const double *x0, const double *x1)
{
    int i = fabs(T10[3]) > fabs(T10[7]) ? 0 : 1;
    double y0[3] = {x0[0], x0[1], 1.};
    double t_i = T10[4 * i + 3], t_z = T10[11];
    double n = t_i - t_z * x1[i];
    double d = dot_n(3, T10 + 8, 1, y0, 1) * x1[i] - dot_n(3, T10 + 4 * i, 1, y0, 1);
    return n / (d + ORA_EPS);
}

/* ---------------------------------------------------------------- variance */

/* src/semi_dense/variance.rs:15-24 */
double ora_calc_variance(double alpha, double geo_var, double photo_var, double geo, double photo)
{
    double a2 = alpha * alpha, g2 = geo * geo, p2 = photo * photo;
    return a2 * (g2 * geo_var + p2 * photo_var);
}
/* src/semi_dense/variance.rs:26-28 */
double ora_photo_var(double gradient) { return 2. / gradient; }

/* src/semi_dense/variance.rs:30-44 */
double ora_geo_var_(const double *direction, const double *image_grad)
{
    double d[2], g[2];
    ora_normalize(2, direction, d);
    ora_normalize(2, image_grad, g);
    double p = dot_n(2, d, 1, g, 1);
    if (p == 0.) return 1. / ORA_EPS;
    return 1. / (p * p);
}
/* src/semi_dense/variance.rs:46-53 */
double ora_geo_var(const double *x_key, const double *t_rk, const double *image_grad)
{
    double pt[2];
    ora_project(t_rk, pt);
    double dir[2] = {x_key[0] - pt[0], x_key[1] - pt[1]};
    return ora_geo_var_(dir, image_grad);
}
/* src/semi_dense/variance.rs:55-70 */
double ora_alpha_(const double *x_key, double x_ref_i, double direction_i, const double *ri,
                  const double *rz, double ti, double tz)
{
    double y[3] = {x_key[0], x_key[1], 1.};
    double d = dot_n(3, rz, 1, y, 1) * ti - dot_n(3, ri, 1, y, 1) * tz;
    double n = x_ref_i * tz - ti;
    return direction_i * d / (n * n);
}
/* src/semi_dense/variance.rs:72-81 */
void ora_ref_epipolar_direction(const double *T_rk, const double *x_key, double dmin, double dmax,
                                double *out)
{
    double a[2], b[2], d[2];
    ora_calc_ref_ends(T_rk, x_key, dmin, dmax, a, b);
    d[0] = b[0] - a[0];
    d[1] = b[1] - a[1];
    ora_normalize(2, d, out);
}
/* src/semi_dense/variance.rs:83-96 */
double ora_calc_alpha_(const double *T_rk, const double *x_key, const double *direction,
                       double prior_depth)
{
    double x_ref[2], z;
    ora_warp(T_rk, x_key, prior_depth, x_ref, &z);
    int i = fabs(direction[0]) > fabs(direction[1]) ? 0 : 1;
    return ora_alpha_(x_key, x_ref[i], direction[i], T_rk + 4 * i, T_rk + 8, T_rk[4 * i + 3],
                      T_rk[11]);
}
/* src/semi_dense/variance.rs:98-106 */
double ora_calc_alpha(const double *T_rk, const double *x_key, double dmin, double dmax,
                      double prior_depth)
{
    double d[2];
    ora_ref_epipolar_direction(T_rk, x_key, dmin, dmax, d);
    return ora_calc_alpha_(T_rk, x_key, d, prior_depth);
}

/* ---------------------------------------------------------------- estimate */

/* src/semi_dense/semi_dense.rs:91-158.  gx/gy are the full Sobel maps of the
 * key image (src/semi_dense/gradient.rs:11-25).  dbg (optional, 2 values)
 * receives the search result {argmin+2, n} for the parity tests. */
int ora_estimate(const double *u_key, double prior_inv_depth, double prior_variance,
                 const double *key_img, const double *key_cam, const double *T_wk,
                 const double *ref_img, const double *ref_cam, const double *T_wr,
                 const double *gx, const double *gy, int64_t H, int64_t W, const double *params,
                 double *out_inv_depth, double *out_variance, int64_t *dbg)
{
    const double vmin = params[0], vmax = params[1], geo = params[2], photo = params[3];
    const double ref_step = params[4], min_gradient = params[5];
    double T_rk[16];
    ora_transform_rk(T_wk, T_wr, T_rk);

    double lo, hi, dmin, dmax;
    ora_hypothesis_range(prior_inv_depth, prior_variance, vmin, vmax, &lo, &hi);
    ora_depth_search_range(lo, hi, &dmin, &dmax);
    double x_key[2];
    ora_cam_normalize(key_cam, u_key, x_key);

    double ratio;
    int flag = ora_step_ratio(T_rk, x_key, prior_inv_depth, &ratio);
    if (flag != ORA_SUCCESS) return flag;
    double key_step_size = ratio * ref_step;

    double x_min_ref[2], x_max_ref[2], ref_direction[2];
    ora_calc_ref_ends(T_rk, x_key, dmin, dmax, x_min_ref, x_max_ref);
    ref_direction[0] = x_max_ref[0] - x_min_ref[0];
    ref_direction[1] = x_max_ref[1] - x_min_ref[1];

    double e_key[2], key_direction[2];
    ora_calc_key_epipole(T_wk, T_wr, e_key);
    ora_calc_key_direction(x_key, e_key, ref_direction, key_direction);

    double xs_key[10], us_key[10];
    ora_key_coordinates(key_direction, x_key, key_step_size, xs_key);
    for (int i = 0; i < 5; i++) ora_cam_unnormalize(key_cam, xs_key + 2 * i, us_key + 2 * i);
    for (int i = 0; i < 5; i++)
        if (!ora_is_in_range(us_key[2 * i], us_key[2 * i + 1], H, W)) return ORA_KEY_OUT_OF_RANGE;

    double key_intensities[5], g1[4];
    ora_interpolate_n(key_img, W, us_key, 5, key_intensities);
    ora_gradient1d(key_intensities, 5, g1);
    double key_gradient = ora_norm(4, g1);
    if (key_gradient < min_gradient) return ORA_INSUFFICIENT_GRADIENT;

    double unit[2], first[2], last[2], u_first[2], u_last[2];
    int64_t n = ora_ref_line(ref_direction, ref_step, unit);
    if (dbg) { dbg[0] = -1; dbg[1] = n; }
    if (n > 0) {
        ora_ref_coordinate(x_min_ref, unit, ref_step, 0, first);
        ora_ref_coordinate(x_min_ref, unit, ref_step, n - 1, last);
        ora_cam_unnormalize(ref_cam, first, u_first);
        ora_cam_unnormalize(ref_cam, last, u_last);
    } else {
        u_first[0] = u_first[1] = u_last[0] = u_last[1] = 0.;
    }
    flag = ora_check_us_ref(n, u_first, u_last, 5, H, W);
    if (flag != ORA_SUCCESS) return flag;

    double *ref_intensities = (double *)malloc(sizeof(double) * (size_t)n);
    if (!ref_intensities) return -100;
    for (int64_t i = 0; i < n; i++) {
        double x[2], u[2];
        ora_ref_coordinate(x_min_ref, unit, ref_step, i, x);
        ora_cam_unnormalize(ref_cam, x, u);
        ref_intensities[i] = ora_interpolate(ref_img, W, u[0], u[1]);
    }
    int64_t argmin = ora_search(ref_intensities, n, key_intensities, 5);
    free(ref_intensities);
    if (dbg) dbg[0] = argmin;

    double x_ref[2];
    ora_ref_coordinate(x_min_ref, unit, ref_step, argmin, x_ref);
    double key_depth = ora_calc_depth0(T_rk, x_key, x_ref);

    double alpha = ora_calc_alpha(T_rk, x_key, dmin, dmax, key_depth);
    double t_rk[3] = {T_rk[3], T_rk[7], T_rk[11]};
    int64_t px = (int64_t)u_key[0], py = (int64_t)u_key[1];
    double grad[2] = {gx[py * W + px], gy[py * W + px]};
    double geo_var = ora_geo_var(x_key, t_rk, grad);
    double photo_var = ora_photo_var(key_gradient / key_step_size);
    double variance = ora_calc_variance(alpha, geo_var, photo_var, geo, photo);

    double new_inv = ora_safe_invert(key_depth);
    flag = ora_check_args(new_inv, variance, vmin, vmax);
    if (flag != ORA_SUCCESS) return flag;
    *out_inv_depth = new_inv;
    *out_variance = variance;
    return ORA_SUCCESS;
}

/* src/py/semi_dense.rs:126-155 */
int ora_estimate_debug(const int64_t *u_key_i, double prior_depth, double prior_variance,
                       const double *key_img, const double *key_cam, const double *T_wk,
                       const double *ref_img, const double *ref_cam, const double *T_wr,
                       int64_t H, int64_t W, const double *params, double *out_depth,
                       double *out_variance, int64_t *dbg)
{
    *out_depth = prior_depth;
    *out_variance = prior_variance;
    if (dbg) { dbg[0] = -1; dbg[1] = -1; }
    int flag = ora_check_args(ora_safe_invert(prior_depth), prior_variance, params[0], params[1]);
    if (flag != ORA_SUCCESS) return flag;
    double *gx = (double *)malloc(sizeof(double) * H * W);
    double *gy = (double *)malloc(sizeof(double) * H * W);
    ora_sobel(key_img, H, W, gx, gy);
    double u_key[2] = {(double)u_key_i[0], (double)u_key_i[1]};
    double id, var;
    flag = ora_estimate(u_key, ora_safe_invert(prior_depth), prior_variance, key_img, key_cam,
                        T_wk, ref_img, ref_cam, T_wr, gx, gy, H, W, params, &id, &var, dbg);
    free(gx);
    free(gy);
    if (flag == ORA_SUCCESS) { *out_depth = ora_safe_invert(id); *out_variance = var; }
    return flag;
}

/* src/semi_dense/semi_dense.rs:160-234 for ONE keyframe with R reference
 * frames (ref_imgs [R,H,W], ref_cams [R,4], T_wrs [R,16]).  Returns 0, or -1
 * if some age exceeds R (the reference prints and exits, :202-205).
 * dbg_argmin / dbg_n are optional [H,W] int64 maps (-1 where no search ran). */
int ora_update_depth_one(const double *key_img, const double *key_cam, const double *T_wk,
                         const double *ref_imgs, const double *ref_cams, const double *T_wrs,
                         int64_t R, const uint64_t *age_map, const double *prior_depth,
                         const double *prior_variance, const double *params, int64_t H, int64_t W,
                         double *out_depth, double *out_variance, int64_t *out_flag,
                         int64_t *dbg_argmin, int64_t *dbg_n)
{
    double *gx = (double *)malloc(sizeof(double) * H * W);
    double *gy = (double *)malloc(sizeof(double) * H * W);
    ora_sobel(key_img, H, W, gx, gy);
    int bad_age = 0;
    for (int64_t y = 0; y < H; y++) {
        for (int64_t x = 0; x < W; x++) {
            int64_t i = y * W + x;
            uint64_t age = age_map[i];
            double d = prior_depth[i], v = prior_variance[i];
            if (dbg_argmin) dbg_argmin[i] = -1;
            if (dbg_n) dbg_n[i] = -1;
            if (age == 0) {
                out_depth[i] = d; out_variance[i] = v; out_flag[i] = ORA_NOT_PROCESSED;
                continue;
            }
            if ((uint64_t)R < age) { bad_age = 1; out_depth[i] = d; out_variance[i] = v;
                                     out_flag[i] = ORA_NOT_PROCESSED; continue; }
            int64_t r = R - (int64_t)age;
            int f = ora_check_args(ora_safe_invert(d), v, params[0], params[1]);
            if (f != ORA_SUCCESS) {
                out_depth[i] = d; out_variance[i] = v; out_flag[i] = f;
                continue;
            }
            double u_key[2] = {(double)x, (double)y};
            double prior_id = ora_safe_invert(d), id = 0., var = 0.;
            int64_t dbg[2] = {-1, -1};
            f = ora_estimate(u_key, prior_id, v, key_img, key_cam, T_wk, ref_imgs + r * H * W,
                             ref_cams + 4 * r, T_wrs + 16 * r, gx, gy, H, W, params, &id, &var,
                             dbg);
            if (f != ORA_SUCCESS) { id = prior_id; var = v; }
            out_depth[i] = ora_safe_invert(id);     /* :225 -- inv(inv(d)) on failure */
            out_variance[i] = var;
            out_flag[i] = f;
            if (dbg_argmin) dbg_argmin[i] = dbg[0];
            if (dbg_n) dbg_n[i] = dbg[1];
        }
    }
    free(gx);
    free(gy);
    return bad_age ? -1 : 0;
}

/* P independent keyframes (each with its own R refs), OpenMP over keyframes:
 * the pair-level parallelism the CPU baseline is allowed. */
int ora_update_depth_batch(const double *key_imgs, const double *key_cams, const double *T_wks,
                           const double *ref_imgs, const double *ref_cams, const double *T_wrs,
                           int64_t P, int64_t R, const uint64_t *age_maps,
                           const double *prior_depth, const double *prior_variance,
                           const double *params, int64_t H, int64_t W, double *out_depth,
                           double *out_variance, int64_t *out_flag, int64_t *dbg_argmin,
                           int64_t *dbg_n, int threads)
{
    int rc = 0;
    int64_t hw = H * W;
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads) reduction(| : rc)
    for (int64_t p = 0; p < P; p++) {
        int r = ora_update_depth_one(
            key_imgs + p * hw, key_cams + 4 * p, T_wks + 16 * p, ref_imgs + p * R * hw,
            ref_cams + 4 * p * R, T_wrs + 16 * p * R, R, age_maps + p * hw, prior_depth + p * hw,
            prior_variance + p * hw, params, H, W, out_depth + p * hw, out_variance + p * hw,
            out_flag + p * hw, dbg_argmin ? dbg_argmin + p * hw : 0, dbg_n ? dbg_n + p * hw : 0);
        rc |= (r != 0);
    }
    return rc ? -1 : 0;
}

/* ------------------------------------------------------------- companions */

/* src/semi_dense/fusion.rs:3-11 */
void ora_fusion(double mu1, double mu2, double var1, double var2, double *mu, double *var)
{
    double v = var1 + var2;
    *mu = (mu1 * var2 + mu2 * var1) / v;
    *var = (var1 * var2) / v;
}
/* src/semi_dense/fusion.rs:13-42 */
void ora_fusion_arrays(const double *mu1, const double *mu2, const double *var1,
                       const double *var2, int64_t n, double *mu, double *var)
{
    for (int64_t i = 0; i < n; i++) ora_fusion(mu1[i], mu2[i], var1[i], var2[i], mu + i, var + i);
}

/* src/semi_dense/stat.rs:5-26 */
int ora_is_statically_same(double id1, double id2, double variance)
{
    double ds = (id1 - id2) * (id1 - id2);
    double fs = ORA_STAT_FACTOR * ORA_STAT_FACTOR;
    return ds <= fs * variance;
}
int ora_are_statically_same(double id1, double id2, double v1, double v2)
{
    return ora_is_statically_same(id1, id2, v1) && ora_is_statically_same(id1, id2, v2);
}

/* src/semi_dense/propagation.rs:9-19 ; powi(4) == (r*r)*(r*r) */
double ora_propagate_variance(double depth0, double depth1, double variance0, double bias)
{
    double ratio = ora_safe_invert(depth1) / ora_safe_invert(depth0);
    double r2 = ratio * ratio;
    return (r2 * r2) * variance0 + bias;
}

/* src/semi_dense/propagation.rs:21-46 */
void ora_handle_collision(double depth_a, double depth_b, double var_a, double var_b,
                          double *depth, double *var)
{
    double ia = ora_safe_invert(depth_a), ib = ora_safe_invert(depth_b);
    if (ora_are_statically_same(ia, ib, var_a, var_b)) {
        double id, v;
        ora_fusion(ia, ib, var_a, var_b, &id, &v);
        *depth = ora_safe_invert(id);
        *var = v;
        return;
    }
    if (depth_a < depth_b) { *depth = depth_a; *var = var_a; }
    else { *depth = depth_b; *var = var_b; }
}

/* src/semi_dense/propagation.rs:48-92 ; the HashMap is a dense occupancy map */
void ora_propagate(const double *T10, const double *cam0, const double *cam1,
                   const double *depth_map0, const double *variance_map0, int64_t H, int64_t W,
                   double default_depth, double default_variance, double bias,
                   double *depth_map1, double *variance_map1)
{
    unsigned char *occupied = (unsigned char *)calloc((size_t)(H * W), 1);
    for (int64_t i = 0; i < H * W; i++) { depth_map1[i] = default_depth; variance_map1[i] = default_variance; }
    for (int64_t y0 = 0; y0 < H; y0++) {
        for (int64_t x0 = 0; x0 < W; x0++) {
            double u0[2] = {(double)x0, (double)y0}, u1[2], depth1a;
            double depth0 = depth_map0[y0 * W + x0];
            ora_perspective_warp(T10, cam0, cam1, u0, depth0, u1, &depth1a);
            if (!ora_is_in_range(u1[0], u1[1], H, W)) continue;
            double variance0 = variance_map0[y0 * W + x0];
            double variance1a = ora_propagate_variance(depth0, depth1a, variance0, bias);
            int64_t t = (int64_t)u1[1] * W + (int64_t)u1[0];
            if (occupied[t]) {
                double d, v;
                ora_handle_collision(depth1a, depth_map1[t], variance1a, variance_map1[t], &d, &v);
                depth_map1[t] = d;
                variance_map1[t] = v;
            } else {
                occupied[t] = 1;
                depth_map1[t] = depth1a;
                variance_map1[t] = variance1a;
            }
        }
    }
    free(occupied);
}

/* src/semi_dense/age.rs:6-32 */
void ora_increment_age(const uint64_t *age_map0, const double *cam0, const double *cam1,
                       const double *T10, const double *depth_map0, int64_t H, int64_t W,
                       uint64_t *age_map1)
{
    memset(age_map1, 0, sizeof(uint64_t) * H * W);
    for (int64_t y0 = 0; y0 < H; y0++) {
        for (int64_t x0 = 0; x0 < W; x0++) {
            double p0[2] = {(double)x0, (double)y0}, q0[2], q1[2], p1[2], z;
            ora_cam_normalize(cam0, p0, q0);
            ora_warp(T10, q0, depth_map0[y0 * W + x0], q1, &z);
            ora_cam_unnormalize(cam1, q1, p1);
            if (!ora_is_in_range(p1[0], p1[1], H, W)) continue;
            age_map1[(int64_t)p1[1] * W + (int64_t)p1[0]] = age_map0[y0 * W + x0] + 1;
        }
    }
}

/* src/semi_dense/regularization.rs:5-64 (disabled in the reference build, F7;
 * the code as written is the only spec). */
void ora_regularize(const double *depth_map, const double *variance_map, const int64_t *flag_map,
                    int64_t H, int64_t W, double *out)
{
    for (int64_t y = 0; y < H; y++) {
        for (int64_t x = 0; x < W; x++) {
            double numerator = 0., denominator = 0.;
            for (int dy = -1; dy <= 1; dy++) {
                for (int dx = -1; dx <= 1; dx++) {
                    int64_t yy = y + dy, xx = x + dx;
                    if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue; /* pad: NotProcessed */
                    if (flag_map[yy * W + xx] != ORA_SUCCESS) continue;
                    double id = ora_safe_invert(depth_map[yy * W + xx]);
                    double iv = ora_safe_invert(variance_map[yy * W + xx]);
                    numerator = numerator + id * iv;
                    denominator = denominator + iv;
                }
            }
            if (denominator == 0.) out[y * W + x] = depth_map[y * W + x];
            else out[y * W + x] = ora_safe_invert(numerator / denominator);
        }
    }
}

/* ------------------------------------------ batch helpers on the drop-in surface */

/* src/warp.rs:31-50 */
void ora_warp_vecs(const double *T10, const double *xs0, const double *depths0, int64_t n,
                   double *xs1, double *depths1)
{
    for (int64_t i = 0; i < n; i++) ora_warp(T10, xs0 + 2 * i, depths0[i], xs1 + 2 * i, depths1 + i);
}
/* src/transform.rs:9-28 ([N,3] form) */
void ora_transform_vecs(const double *T10, const double *ps0, int64_t n, double *ps1)
{
    for (int64_t i = 0; i < n; i++) ora_transform(T10, ps0 + 3 * i, ps1 + 3 * i);
}
/* src/projection.rs:34-60 */
void ora_project_vecs(const double *ps, int64_t n, double *xs)
{
    for (int64_t i = 0; i < n; i++) ora_project(ps + 3 * i, xs + 2 * i);
}
void ora_inv_project_vecs(const double *xs, const double *depths, int64_t n, double *ps)
{
    for (int64_t i = 0; i < n; i++) ora_inv_project(xs + 2 * i, depths[i], ps + 3 * i);
}
