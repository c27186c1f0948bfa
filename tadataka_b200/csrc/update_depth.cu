// update_depth: per-pixel epipolar search + inverse-depth update, batched over keyframes.
// Replaces src/semi_dense/semi_dense.rs:91-234 of the reference (see include/tdk.h).
//
// pair_setup_kernel computes, once per (keyframe, reference), what the reference recomputes for
// every pixel.  The per-pixel pipeline is cut into phases (A0 certified sieve, A1 exact prefix up to
// the gradient test, A2 line set-up, B search, C triangulation + variance; see the block comments
// further down) and arranged in three plans:
//   sorted (default for batches): key_prep_kernel (f32 copy of the key images), sieve_kernel (A0 on
//       every pixel: 80-96 % of the pixels get their final values there, the undecided ones go to a
//       per-keyframe queue), lines_kernel (A1 + A2 per queued pixel, a 128-byte line record per
//       survivor), order_kernel (counting sort by line length), match_kernel (B + C);
//   two kernels (single small keyframes): key_prep, sieve, then search_kernel takes 256 queued pixels
//       per CTA and runs A1 / A2 / B / C;
//   one kernel: a CTA owns a 64x4 tile, runs the exact A1 on every pixel, compacts its own candidates
//       in shared memory and runs all phases; needs no queue (what the workspace-free entry point uses).
// All three produce bit-identical results.
#include <math.h>

#include <atomic>
#include <mutex>

#include "tdk_common.cuh"

namespace {

constexpr int TILE_W = 64;
constexpr int TILE_H = 4;
constexpr int BLOCK = TILE_W * TILE_H;
#ifndef TDK_MIN_CTAS
#define TDK_MIN_CTAS 4
#endif
#ifndef TDK_SIEVE_MIN_CTAS
#define TDK_SIEVE_MIN_CTAS 8
#endif
constexpr int MIN_CTAS = TDK_MIN_CTAS;       // resident CTAs per SM the register allocation is tuned for
constexpr int SIEVE_MIN_CTAS = TDK_SIEVE_MIN_CTAS;
constexpr int SORT_BUCKETS = 64;          // phase B sorts a CTA's lines by min(length, 63)
#ifndef TDK_LONG_LINE
#define TDK_LONG_LINE 32
#endif
constexpr int LONG_LINE = TDK_LONG_LINE;  // lines with more samples than this are searched by a whole warp

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
    // queue plans, all indexed by the keyframe's position in the current sub-batch (blockIdx.z):
    unsigned int *cand;         // [S, H*W] queue of the pixels the sieve could not decide
    unsigned int *cand_count;   // [S] queue lengths
    const float *img32;         // [S, H*W] f32 copy of the key images (key_prep_kernel)
    const unsigned int *amax;   // [S] bits of a float >= max |key image| (+inf if a value is not finite)
    // sorted plan:
    double2 *rec;               // [S, H*W] one 128-byte line record per queue slot (lines_kernel)
    unsigned char *len_class;   // [S, H*W] min(line length, 63) per queue slot, 0 = dropped in A2
    unsigned int *order;        // [S, H*W] queue slots of the live lines, longest class first
    unsigned int *len_hist;     // [S, 64] lines per length class
    unsigned int *len_cursor;   // [S, 64] slots handed out per class while `order` is filled
    size_t rec_stride;          // slots per keyframe in rec: H*W rounded up to whole chunks of 256
    int warp_budget;            // match_kernel: warps per keyframe that can be resident at once
    int p0;                     // first keyframe of the sub-batch (blockIdx.z + p0 indexes maps, keys, pairs)
    double vmin, vmax, geo, photo, step, min_grad;
    double min_grad_sq;         // smallest s with sqrt(s) >= min_grad (NaN if min_grad is NaN)
    // phase A0 (certified sieve), see phase_a0()
    double cert_m;              // min_grad * (1 - 1e-5), or -1 if min_grad is not a positive number
    double cert_e;              // bound on |f32 difference - reference difference| per unit of max |image|
    float cert_lo, cert_xhi, cert_yhi;      // a coordinate in [lo, hi] is inside the image for certain
    float cert_out, cert_xout, cert_yout;   // a coordinate < out or > [xy]out is outside for certain
    int P, R, H, W, tiles_x, tiles_y, fuse;
    int subpix;                 // opt-in sub-sample refinement of the search result (not in the reference)
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
                                  int *status, unsigned int *zero_words, int n_zero_words)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    // the queue plans' counters (queue lengths, line-length histograms) start at zero; clearing them
    // here saves two memset nodes per launch
    for (int i = idx; i < n_zero_words; i += gridDim.x * blockDim.x) zero_words[i] = 0u;
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

// =============================================================================================
// The per-pixel pipeline, cut into phases that the kernels below arrange in two ways.
//   A1  thread per pixel      prior checks, step ratio, 5 key samples, gradient test
//   A2  thread per candidate  reference-line ends, direction sign, line-length / end-point tests
//   B   whole CTA             the search: samples and windows of all the CTA's lines, laid end to end
//   C   thread per survivor   triangulation, variances, final check, optional fusion with the prior
// Candidate state lives in a shared-memory queue (structure of arrays, BLOCK slots).
// =============================================================================================
struct CtaQueue {
    double k[5][BLOCK];                       // key intensities (A1), then the normalised key window (A2)
    double kstep[BLOCK], g[BLOCK], grev[BLOCK], idp[BLOCK], v[BLOCK], xk[BLOCK], yk[BLOCK];
    double xminx[BLOCK], xminy[BLOCK], ux[BLOCK], uy[BLOCK], rdx[BLOCK], rdy[BLOCK];
    int n[BLOCK], arg[BLOCK], bucket[SORT_BUCKETS], nlong;
    unsigned short perm[BLOCK];               // queue slots ordered by line length (phase B)
    int px[BLOCK];                            // x | y << 16 of the pixel
    unsigned short ref[BLOCK], owner[BLOCK];  // reference index; slot of the owning thread (one-kernel plan)
};

constexpr size_t CTA_SMEM_BYTES = sizeof(CtaQueue);   // dynamic shared memory of the tile and search kernels

struct A1Result {
    double ki[5];
    double kstep, gf, gr, idp;
    int r;
};

// Phase A1 for pixel (x, y) with prior (d, v) and age > 0 already known to be <= R.
// Returns a final flag (<= 0) or 1 = candidate.  out_d is what a failed pixel stores:
// d before check_args passes, inv(inv(d)) after (semi_dense.rs:209-225).
__device__ __forceinline__ int phase_a1(const UpdArgs &a, const TdkKeyConst &key, int p, double xk, double yk,
                                        uint64_t age, double d, double v, double wm1, double hm1,
                                        double &out_d, A1Result &o)
{
    const int W = a.W, H = a.H;
    out_d = d;
    const double idp = tdk::inv(d);
    int flag = check_args(idp, v, a.vmin, a.vmax);   // semi_dense.rs:209-214
    if (flag != TDK_FLAG_SUCCESS) return flag;
    const double kd = tdk::inv(idp);
    out_d = kd;
    const int r = a.R - (int)age;
    const TdkPairConst *__restrict__ pc = a.pairs + (size_t)p * a.R + r;

    // step_ratio, semi_dense.rs:27-40 + depth.rs:6-15
    const double rz = pc->T[8] * (xk * kd) + pc->T[9] * (yk * kd) + pc->T[10] * (1. * kd) + pc->T[11];
    if (rz <= 0.) return TDK_FLAG_NEGATIVE_REF_DEPTH;
    const double kstep = (idp / tdk::inv(rz)) * a.step;

    // key_coordinates, epipolar.rs:22-36.  The sign of the key direction (semi_dense.rs:42-49) needs
    // the reference-line ends, i.e. six more divisions; it is decided in A2, for candidates only.
    // Flipping the sign maps sample k onto sample 4-k exactly (IEEE negation is exact), so the range
    // test is sign independent and the gradient norm only changes its summation order: both orders
    // are tested here.
    double kdx = xk - pc->ex, kdy = yk - pc->ey;
    const double knorm = sqrt(kdx * kdx + kdy * kdy);
    if (knorm != 0.) tdk::div2(kdx, kdy, knorm, kdx, kdy);
    double kcx[5], kcy[5];
#pragma unroll
    for (int k = 0; k < 5; k++) {
        const double s = kstep * (double)(k - 2);
        kcx[k] = (xk + s * kdx) * key.fx + key.ox;
        kcy[k] = (yk + s * kdy) * key.fy + key.oy;
    }
    // all_in_range, image_range.rs:45-57.  Every operation above is monotone in k (rounding is
    // monotone), so the five points are ordered along the line and lie between the two outer ones:
    // testing those two decides the test for all five.  A NaN anywhere shows up in both.
    if (!(tdk::in_range(kcx[0], kcy[0], wm1, hm1) && tdk::in_range(kcx[4], kcy[4], wm1, hm1)))
        return TDK_FLAG_KEY_OUT_OF_RANGE;
#pragma unroll
    for (int k = 0; k < 5; k++) o.ki[k] = tdk::bilinear(key.img, W, H, kcx[k], kcy[k]);
    // gradient1d + norm, gradient.rs:28-35, semi_dense.rs:130
    const double g0 = o.ki[1] - o.ki[0], g1 = o.ki[2] - o.ki[1], g2 = o.ki[3] - o.ki[2], g3 = o.ki[4] - o.ki[3];
    const double q0 = g0 * g0, q1 = g1 * g1, q2 = g2 * g2, q3 = g3 * g3;
    const double sf = q0 + q1 + q2 + q3, sr = q3 + q2 + q1 + q0;
    // sqrt(s) < min_gradient <=> s < min_grad_sq (launch() finds that threshold exactly): the two
    // square roots are only taken for the pixels that pass
    if (sf < a.min_grad_sq && sr < a.min_grad_sq) return TDK_FLAG_INSUFFICIENT_GRADIENT;
    const double gf = sqrt(sf);
    const double gr = sr == sf ? gf : sqrt(sr);
    o.kstep = kstep; o.gf = gf; o.gr = gr; o.idp = idp; o.r = r;
    return 1;
}

// =============================================================================================
// Phase A0: the certified sieve.  Most pixels (80-96 %) end at the gradient test of phase A1 after
// paying for five exact f64 bilinear samples, two square roots and eight IEEE divisions.  A0 decides
// them from a cheap approximation whose distance to the reference's f64 values is bounded, and
// finishes a pixel only when the bound leaves no doubt; everything else goes to the exact path, so
// the outputs stay bit-identical.
//
// Exact, as in A1: inv(d), check_args, inv(inv(d)) (what a failed pixel stores), ref_z and its sign.
// Approximate from there:
//   key_step  idp * (rz + eps) * step instead of (idp / inv(rz)) * step: relative distance <= 2^-50;
//   direction (x_key - e_key) * rsqrt(|.|^2), rsqrt = MUFU.RSQ64H + two Newton steps: <= 2^-30;
//   a = (float)(key_step * dir * f): the step between two samples in pixels, per axis;
//   sample coordinates fmaf(k - 2, a, (float)pixel): for half segments below CERT_HMAX pixels each is
//       within delta = (2 CERT_HMAX + max(W, H)) * 2^-22 pixels of the reference's f64 coordinate
//       (2^-24 relative for the conversion of a, half an ulp of a float below W + CERT_HMAX for the
//       sum, 2^-40 (W + CERT_HMAX) for everything rounded in f64; the bound used is 4x that sum);
//   samples   f32 bilinear on the f32 copy of the key image.  With A = max |image|: moving a
//       coordinate by delta moves the (continuous, piecewise bilinear) interpolant by at most 2 A delta
//       per axis; the f32 copy is within 2^-24 A; the f32 lerps round 12 times at most 2^-24 A; the
//       reference's own f64 rounding is below 2^-49 A.  So |s32 - s_ref| <= (4 delta + 16 * 2^-24) A,
//       and a difference of two neighbouring samples is within e = 2 (4 delta + 16 * 2^-24) A + 2^-22 A
//       of the reference's (cert_e = e / A is computed by launch()).
// Decisions:
//   range test (image_range.rs:45-57; flag -2 precedes -6): the outer points 0 and 4 bound the other
//       three.  Both inside [max(delta, 1e-3), size - 1 - max(delta, 1e-3)]: in range for certain.
//       One of them beyond the image by more than that margin: KEY_OUT_OF_RANGE for certain.
//   gradient test (semi_dense.rs:129-136): the reference's sum of squared differences, in either
//       summation order, is at most (sqrt(S32) (1 + 2^-21) + 2 e)^2 (1 + 2^-49) (triangle inequality
//       in R^4), so S32 < ((min_grad (1 - 1e-5) - 2 e)^2 (1 - 1e-5), rounded down to f32) certifies
//       INSUFFICIENT_GRADIENT.  NaN or Inf anywhere fails every comparison and falls through.
// tools/sieve_model.py restates this in numpy and checks it against the oracle on the CPU.
constexpr float CERT_HMAX = 128.f;   // longest half segment (2 key steps, in pixels) the cheap path takes

__device__ __forceinline__ double rsqrt_2newton(double x)
{
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));   // MUFU.RSQ64H
    const double h = 0.5 * x;
    y = y * __fma_rn(-h, y * y, 1.5);
    y = y * __fma_rn(-h, y * y, 1.5);
    return y;
}

// f32 bilinear sample; (cx, cy) inside [0, W-1) x [0, H-1) so that all four taps exist
__device__ __forceinline__ float sample32(const float *__restrict__ img, int W, float cx, float cy)
{
    const int ix = __float2int_rd(cx), iy = __float2int_rd(cy);
    const float wx = cx - (float)ix, wy = cy - (float)iy;
    // unsigned 32-bit offsets: one IMAD.WIDE.U32 per address instead of a sign-extended 64-bit chain
    const unsigned o0 = (unsigned)(iy * W + ix), o1 = o0 + (unsigned)W;
    const float a = __ldg(img + o0), b = __ldg(img + o0 + 1u), c = __ldg(img + o1), d = __ldg(img + o1 + 1u);
    const float top = __fmaf_rn(wx, b - a, a), bot = __fmaf_rn(wx, d - c, c);
    return __fmaf_rn(wy, bot - top, top);
}

// Key tile staged in shared memory by the TMA engine (sieve variant, see update_depth_sieve_kernel):
// the f32 key pixels of the CTA's 64 x (4 PPT) tile plus ST_HALO pixels on every side.
constexpr int ST_HALO = 16;
constexpr int ST_W = TILE_W + 2 * ST_HALO;
struct KeyTile {
    const float *s;    // shared memory, row pitch ST_W; nullptr = no tile, sample global memory
    int x0, y0, rows;  // image coordinates of the tile's first element, number of rows
};

// f32 bilinear sample out of the staged tile; (ix, iy) = floor of the coordinates, already known to lie
// inside the tile with their +1 neighbours
__device__ __forceinline__ float sample32_tile(const KeyTile &t, float cx, float cy)
{
    const int ix = __float2int_rd(cx), iy = __float2int_rd(cy);
    const float wx = cx - (float)ix, wy = cy - (float)iy;
    const float *p = t.s + (iy - t.y0) * ST_W + (ix - t.x0);
    const float a = p[0], b = p[1], c = p[ST_W], d = p[ST_W + 1];
    const float top = __fmaf_rn(wx, b - a, a), bot = __fmaf_rn(wx, d - c, c);
    return __fmaf_rn(wy, bot - top, top);
}

__device__ __forceinline__ bool tile_holds(const KeyTile &t, float cx, float cy)
{
    const int ix = __float2int_rd(cx) - t.x0, iy = __float2int_rd(cy) - t.y0;
    return ix >= 0 && ix < ST_W - 1 && iy >= 0 && iy < t.rows - 1;
}

// Phase A0 for pixel (x, y) with prior (d, v) and 0 < age <= R.  Returns a final flag (<= 0) or
// 1 = undecided (the exact path has to look at it).  out_d as in phase_a1.  thr2: the keyframe's
// threshold on the f32 sum of squared differences (negative: certify nothing).
__device__ __forceinline__ int phase_a0(const UpdArgs &a, const TdkKeyConst &key, const float *__restrict__ img32,
                                        int p, int x, int y, double xk, double yk, uint64_t age, double d,
                                        double v, float thr2, double &out_d, const KeyTile tile = KeyTile{nullptr, 0, 0, 0})
{
    out_d = d;
    const double idp = tdk::inv(d);
    const int flag = check_args(idp, v, a.vmin, a.vmax);   // semi_dense.rs:209-214
    if (flag != TDK_FLAG_SUCCESS) return flag;
    const double kd = tdk::inv(idp);
    out_d = kd;
    const TdkPairConst *__restrict__ pc = a.pairs + (size_t)p * a.R + (a.R - (int)age);
    const double rz = pc->T[8] * (xk * kd) + pc->T[9] * (yk * kd) + pc->T[10] * (1. * kd) + pc->T[11];
    if (rz <= 0.) return TDK_FLAG_NEGATIVE_REF_DEPTH;
    // ---- approximate from here on
    const double kstep = (idp * (rz + TDK_EPS_INV)) * a.step;
    const double kdx = xk - pc->ex, kdy = yk - pc->ey;
    const double rn = rsqrt_2newton(kdx * kdx + kdy * kdy);
    const float ax = (float)((kstep * (kdx * rn)) * key.fx);
    const float ay = (float)((kstep * (kdy * rn)) * key.fy);
    if (!(fabsf(ax) < 0.5f * CERT_HMAX && fabsf(ay) < 0.5f * CERT_HMAX)) return 1;   // too long, or not a number
    const float xf = (float)x, yf = (float)y;
    const float x0 = __fmaf_rn(-2.f, ax, xf), y0 = __fmaf_rn(-2.f, ay, yf);
    const float x4 = __fmaf_rn(2.f, ax, xf), y4 = __fmaf_rn(2.f, ay, yf);
    const float xlo = fminf(x0, x4), xhi = fmaxf(x0, x4), ylo = fminf(y0, y4), yhi = fmaxf(y0, y4);
    if (!(xlo >= a.cert_lo && xhi <= a.cert_xhi && ylo >= a.cert_lo && yhi <= a.cert_yhi)) {
        const bool out = xlo < a.cert_out || xhi > a.cert_xout || ylo < a.cert_out || yhi > a.cert_yout;
        return out ? TDK_FLAG_KEY_OUT_OF_RANGE : 1;
    }
    const int W = a.W;
    float s0, s1, s2, s3, s4;
    if (tile.s && tile_holds(tile, x0, y0) && tile_holds(tile, x4, y4)) {
        // the segment is straight: with both ends inside the tile every sample is
        s0 = sample32_tile(tile, x0, y0);
        s1 = sample32_tile(tile, xf - ax, yf - ay);
        s2 = tile.s[(y - tile.y0) * ST_W + (x - tile.x0)];
        s3 = sample32_tile(tile, xf + ax, yf + ay);
        s4 = sample32_tile(tile, x4, y4);
    } else {
        s0 = sample32(img32, W, x0, y0);
        s1 = sample32(img32, W, xf - ax, yf - ay);
        s2 = __ldg(img32 + (unsigned)(y * W + x));
        s3 = sample32(img32, W, xf + ax, yf + ay);
        s4 = sample32(img32, W, x4, y4);
    }
    const float d0 = s1 - s0, d1 = s2 - s1, d2 = s3 - s2, d3 = s4 - s3;
    const float S = __fmaf_rn(d3, d3, __fmaf_rn(d2, d2, __fmaf_rn(d1, d1, d0 * d0)));
    return S < thr2 ? TDK_FLAG_INSUFFICIENT_GRADIENT : 1;
}

// The keyframe's threshold for phase_a0 from A = max |key image| (see the block comment above).
__device__ __forceinline__ float cert_threshold(const UpdArgs &a, unsigned int amax_bits)
{
    const double A = (double)__uint_as_float(amax_bits);
    const double thr = a.cert_m - 2. * (a.cert_e * A);
    return thr > 0. ? __double2float_rd((thr * thr) * (1. - 1e-5)) : -1.f;   // NaN -> -1
}

__device__ __forceinline__ void queue_push(CtaQueue &Q, int slot, const A1Result &c, double v, double xk, double yk,
                                           int x, int y, int owner)
{
#pragma unroll
    for (int k = 0; k < 5; k++) Q.k[k][slot] = c.ki[k];
    Q.kstep[slot] = c.kstep; Q.g[slot] = c.gf; Q.grev[slot] = c.gr;
    Q.idp[slot] = c.idp; Q.v[slot] = v; Q.xk[slot] = xk; Q.yk[slot] = yk;
    Q.px[slot] = x | (y << 16);
    Q.ref[slot] = (unsigned short)c.r;
    Q.owner[slot] = (unsigned short)owner;
}

// What phases B and C need of a candidate that passed A2.  Lives in registers, in the shared-memory
// queue (one-kernel plan) or in a 128-byte record of the workspace (sorted plan).
struct LineRec {
    double k0, k1, k2, k3, k4;      // normalised key window, in the direction the reference sampled it
    double xminx, xminy, ux, uy;    // near end and unit direction of the reference line
    double rdx, rdy;                // x_max_ref - x_min_ref
    double g, kstep;                // key gradient norm, key step
    int n;                          // samples on the reference line
};

// Phase A2 in registers.  Returns 1 (survivor, o filled) or a final flag (o.n = 0).
__device__ __forceinline__ int phase_a2_core(const UpdArgs &a, const TdkPairConst *__restrict__ pc, double xk,
                                             double yk, double v, const A1Result &c, double wm1, double hm1,
                                             LineRec &o)
{
    const double *__restrict__ T = pc->T;
    const double idp = c.idp;
    o.n = 0;
    // hypothesis.rs:54-61, depth.rs:25-30
    const double lo = tdk::clampd(idp - TDK_VARIANCE_FACTOR * v, a.vmin, a.vmax);
    const double hi = tdk::clampd(idp + TDK_VARIANCE_FACTOR * v, a.vmin, a.vmax);
    const double dmin = tdk::inv(hi), dmax = tdk::inv(lo);
    // calc_ref_ends, semi_dense.rs:51-60
    double xminx, xminy, xmaxx, xmaxy, zz;
    tdk::warp_point(T, xk, yk, dmin, xminx, xminy, zz);
    tdk::warp_point(T, xk, yk, dmax, xmaxx, xmaxy, zz);
    const double rdx = xmaxx - xminx, rdy = xmaxy - xminy;
    // calc_key_direction, semi_dense.rs:42-49: d if ref_dir . d > 0 else -d
    const double kdx = xk - pc->ex, kdy = yk - pc->ey;
    const bool keep = rdx * kdx + rdy * kdy > 0.;
    double k0 = c.ki[0], k1 = c.ki[1], k2 = c.ki[2], k3 = c.ki[3], k4 = c.ki[4];
    double g = c.gf;
    if (!keep) {   // the reference sampled along -d: same five values in reverse order
        double tmp = k0; k0 = k4; k4 = tmp;
        tmp = k1; k1 = k3; k3 = tmp;
        g = c.gr;
    }
    if (g < a.min_grad) return TDK_FLAG_INSUFFICIENT_GRADIENT;
    // ref_coordinates + check_us_ref, epipolar.rs:38-54, semi_dense.rs:62-81
    const double L = sqrt(rdx * rdx + rdy * rdy);
    double ux, uy;
    tdk::div2(rdx, rdy, L + TDK_EPS, ux, uy);
    const double nf = L / a.step;
    int n = 0;
    if (nf > 0.) n = nf >= 2147483647. ? 2147483647 : (int)nf;   // Rust `as usize`
    if (n < 5) return TDK_FLAG_REF_EPIPOLAR_TOO_SHORT;
    const double fx0 = (xminx + (0. * a.step) * ux) * pc->rfx + pc->rox;
    const double fy0 = (xminy + (0. * a.step) * uy) * pc->rfy + pc->roy;
    if (!tdk::in_range(fx0, fy0, wm1, hm1)) return TDK_FLAG_REF_CLOSE_OUT_OF_RANGE;
    const double sl = (double)(n - 1) * a.step;
    const double fx1 = (xminx + sl * ux) * pc->rfx + pc->rox;
    const double fy1 = (xminy + sl * uy) * pc->rfy + pc->roy;
    if (!tdk::in_range(fx1, fy1, wm1, hm1)) return TDK_FLAG_REF_FAR_OUT_OF_RANGE;
    if (n > TDK_MAX_SEARCH_SAMPLES) {
        atomicExch(&a.status[TDK_ST_SEARCH_LEN], 1);
        n = TDK_MAX_SEARCH_SAMPLES;
    }
    // normalise the key window once (intensities.rs:20 does it for every window)
    const double kn = sqrt(k0 * k0 + k1 * k1 + k2 * k2 + k3 * k3 + k4 * k4);
    if (kn != 0.) tdk::div5(k0, k1, k2, k3, k4, kn);
    o.k0 = k0; o.k1 = k1; o.k2 = k2; o.k3 = k3; o.k4 = k4;
    o.xminx = xminx; o.xminy = xminy; o.ux = ux; o.uy = uy; o.rdx = rdx; o.rdy = rdy;
    o.g = g; o.kstep = c.kstep; o.n = n;
    return 1;
}

// Phase A2 for queue slot q.  Returns 1 (survivor; Q.n[q] = line length) or a final flag (Q.n[q] = 0).
__device__ __forceinline__ int phase_a2(const UpdArgs &a, CtaQueue &Q, int p, int q, double wm1, double hm1)
{
    const TdkPairConst *__restrict__ pc = a.pairs + (size_t)p * a.R + Q.ref[q];
    A1Result c;
#pragma unroll
    for (int k = 0; k < 5; k++) c.ki[k] = Q.k[k][q];
    c.kstep = Q.kstep[q]; c.gf = Q.g[q]; c.gr = Q.grev[q]; c.idp = Q.idp[q]; c.r = Q.ref[q];
    LineRec o;
    const int f = phase_a2_core(a, pc, Q.xk[q], Q.yk[q], Q.v[q], c, wm1, hm1, o);
    if (f == 1) {
        Q.k[0][q] = o.k0; Q.k[1][q] = o.k1; Q.k[2][q] = o.k2; Q.k[3][q] = o.k3; Q.k[4][q] = o.k4;
        Q.g[q] = o.g;
        Q.xminx[q] = o.xminx; Q.xminy[q] = o.xminy;
        Q.ux[q] = o.ux; Q.uy[q] = o.uy;
        Q.rdx[q] = o.rdx; Q.rdy[q] = o.rdy;
    }
    Q.n[q] = o.n;
    return f;
}

// One reference-line sample: bilinear(ref, K_r * (x_min + (i * step) * dir))
// (epipolar.rs:38-54, camera.rs:59-61, interpolation.rs:9-43).
struct LineGeom {
    const double *ref;
    double rfx, rfy, rox, roy, xminx, xminy, ux, uy;
};
__device__ __forceinline__ LineGeom line_geom(const UpdArgs &a, const CtaQueue &Q, int p, int q)
{
    const TdkPairConst *__restrict__ pc = a.pairs + (size_t)p * a.R + Q.ref[q];
    return LineGeom{pc->ref, pc->rfx, pc->rfy, pc->rox, pc->roy, Q.xminx[q], Q.xminy[q], Q.ux[q], Q.uy[q]};
}
__device__ __forceinline__ double line_sample(const LineGeom &g, int i, double step, int W, int H)
{
    const double s = (double)i * step;
    const double cx = (g.xminx + s * g.ux) * g.rfx + g.rox;
    const double cy = (g.xminy + s * g.uy) * g.rfy + g.roy;
    return tdk::bilinear(g.ref, W, H, cx, cy);
}

// || normalize(window) - normalize(kernel) ||^2   (intensities.rs:6-29, vector.rs:4-11)
__device__ __forceinline__ double window_error(double s0, double s1, double s2, double s3, double s4,
                                               double k0, double k1, double k2, double k3, double k4)
{
#ifdef TDK_GATE_SEARCH
    // Experiment (not shipped; DESIGN.md "gate-tolerance build"): with both windows normalised the
    // error is 2 - 2 cos, so the argmin is the argmax of <s, k^> / |s|; an approximate reciprocal
    // square root (one Newton step, ~2^-40) instead of sqrt + five IEEE divisions.
    const double n2 = s0 * s0 + s1 * s1 + s2 * s2 + s3 * s3 + s4 * s4;
    const double dot = s0 * k0 + s1 * k1 + s2 * k2 + s3 * k3 + s4 * k4;
    if (!(n2 > 0.)) return n2 == 0. ? 0.5 : n2;   // empty window; NaN stays NaN
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(n2));
    y = y * __fma_rn(-0.5 * n2, y * y, 1.5);
    return 1. - dot * y;
#else
    const double wn = sqrt(s0 * s0 + s1 * s1 + s2 * s2 + s3 * s3 + s4 * s4);
    double w0 = s0, w1 = s1, w2 = s2, w3 = s3, w4 = s4;
    if (wn != 0.) tdk::div5(w0, w1, w2, w3, w4, wn);
    const double d0 = w0 - k0, d1 = w1 - k1, d2 = w2 - k2, d3 = w3 - k3, d4 = w4 - k4;
    return d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3 + d4 * d4;
#endif
}

// Opt-in sub-sample refinement (north star step 3; the reference has none: intensities.rs:11-36
// returns an integer index, so this is OFF unless tdk_set_subpixel_refinement(1) was called and
// results then no longer match the reference).  A parabola through the errors of the best window
// and its two neighbours; returns the offset of its minimum in samples, within [-0.5, 0.5], or 0
// when the best window sits at an end of the line or the three errors do not bracket a minimum.
__device__ __forceinline__ double subpixel_offset(const LineGeom &g, int n, int best, double step, int W, int H,
                                                  double k0, double k1, double k2, double k3, double k4)
{
    if (best < 1 || best + 1 > n - 5) return 0.;
    double s[7];
#pragma unroll
    for (int j = 0; j < 7; j++) s[j] = line_sample(g, best - 1 + j, step, W, H);
    const double em = window_error(s[0], s[1], s[2], s[3], s[4], k0, k1, k2, k3, k4);
    const double e0 = window_error(s[1], s[2], s[3], s[4], s[5], k0, k1, k2, k3, k4);
    const double ep = window_error(s[2], s[3], s[4], s[5], s[6], k0, k1, k2, k3, k4);
    const double den = (em - e0) + (ep - e0);
    if (!(den > 0.) || !(em >= e0) || !(ep >= e0)) return 0.;
    const double off = 0.5 * (em - ep) / den;
    return off < -0.5 ? -0.5 : off > 0.5 ? 0.5 : off;
}

// Phase C in registers for the pixel (sx, sy) whose search returned `arg` (= argmin + 2).  Returns the
// final flag; on SUCCESS nd / nv hold the new (or fused) depth and variance.  pv = prior variance.
__device__ __forceinline__ int phase_c_core(const UpdArgs &a, const TdkKeyConst &key,
                                            const TdkPairConst *__restrict__ pc, int arg, int sx, int sy, double xk,
                                            double yk, double xminx, double xminy, double ux, double uy,
                                            double rdx, double rdy, double g, double kstep, double pv,
                                            size_t map_base, double &nd, double &nv, double sub = 0.)
{
    const int W = a.W, H = a.H;
    const double *__restrict__ T = pc->T;
    const double s = ((double)arg + sub) * a.step;   // sub = 0 unless the opt-in refinement ran: x + 0. == x
    const double xr = xminx + s * ux;
    const double yr = xminy + s * uy;

    // calc_depth0, triangulation.rs:8-36
    const double rzy = T[8] * xk + T[9] * yk + T[10] * 1.;
    const double r0y = T[0] * xk + T[1] * yk + T[2] * 1.;
    const double r1y = T[4] * xk + T[5] * yk + T[6] * 1.;
    const double tzz = T[11];
    const bool ax0 = fabs(T[3]) > fabs(T[7]);
    const double x1i = ax0 ? xr : yr, ti = ax0 ? T[3] : T[7], riy = ax0 ? r0y : r1y;
    const double depth = (ti - tzz * x1i) / ((rzy * x1i - riy) + TDK_EPS);

    // calc_alpha, variance.rs:55-106
    const double rn = sqrt(rdx * rdx + rdy * rdy);
    double dhx = rdx, dhy = rdy;
    if (rn != 0.) tdk::div2(rdx, rdy, rn, dhx, dhy);
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
    double lx2 = xk - pc->ptx, ly2 = yk - pc->pty;
    const double ln = sqrt(lx2 * lx2 + ly2 * ly2);
    if (ln != 0.) tdk::div2(lx2, ly2, ln, lx2, ly2);
    const double gn = sqrt(gx * gx + gy * gy);
    if (gn != 0.) tdk::div2(gx, gy, gn, gx, gy);
    const double pp = lx2 * gx + ly2 * gy;
    const double geo_var = pp == 0. ? 1. / TDK_EPS : 1. / (pp * pp);
    const double photo_var = 2. / (g / kstep);
    const double variance = (alpha * alpha) * ((a.geo * a.geo) * geo_var + (a.photo * a.photo) * photo_var);

    const double idn = tdk::inv(depth);
    const int f = check_args(idn, variance, a.vmin, a.vmax);
    if (f == TDK_FLAG_SUCCESS) {
        nd = tdk::inv(idn);
        nv = variance;
        if (a.fuse) {   // fusion.rs:3-11 in inverse depth, as propagation.rs:27-34
            const double pd = a.prior_d[map_base + (size_t)sy * W + sx];
            const double ia = tdk::inv(pd), ib = tdk::inv(nd);
            const double vs = pv + nv;
            nd = tdk::inv((ia * nv + ib * pv) / vs);
            nv = (pv * nv) / vs;
        }
    }
    return f;
}

// Phase C for a live queue slot q.
template <bool SUBPIX>
__device__ __forceinline__ int phase_c(const UpdArgs &a, const CtaQueue &Q, const TdkKeyConst &key, int p, int q,
                                       size_t map_base, double &nd, double &nv)
{
    const TdkPairConst *__restrict__ pc = a.pairs + (size_t)p * a.R + Q.ref[q];
    double sub = 0.;
    if (SUBPIX)
        sub = subpixel_offset(line_geom(a, Q, p, q), Q.n[q], Q.arg[q] - 2, a.step, a.W, a.H, Q.k[0][q], Q.k[1][q],
                              Q.k[2][q], Q.k[3][q], Q.k[4][q]);
    return phase_c_core(a, key, pc, Q.arg[q], Q.px[q] & 0xffff, Q.px[q] >> 16, Q.xk[q], Q.yk[q], Q.xminx[q],
                        Q.xminy[q], Q.ux[q], Q.uy[q], Q.rdx[q], Q.rdy[q], Q.g[q], Q.kstep[q], Q.v[q], map_base,
                        nd, nv, sub);
}

// Search of ONE line by a full warp: lane i takes sample base+i, windows come from shfl_down, the
// argmin is an (error, index) shuffle reduction with the lowest index winning ties.  Returns argmin
// (0 if no window beat +inf, as intensities.rs:16-29 does) on every lane.
__device__ __forceinline__ int search_line_warp(const LineGeom &g, int n, double step, int W, int H,
                                                double k0, double k1, double k2, double k3, double k4, int lane)
{
    const int nwin = n - 4;
    double best_e = INFINITY;
    int best_i = 0x7fffffff;
    for (int base = 0; base < nwin; base += 28) {
        const int i = base + lane;
        const double s0 = i < n ? line_sample(g, i, step, W, H) : 0.;
        const double s1 = __shfl_down_sync(0xffffffffu, s0, 1);
        const double s2 = __shfl_down_sync(0xffffffffu, s0, 2);
        const double s3 = __shfl_down_sync(0xffffffffu, s0, 3);
        const double s4 = __shfl_down_sync(0xffffffffu, s0, 4);
        if (lane < 28 && i < nwin) {
            const double e = window_error(s0, s1, s2, s3, s4, k0, k1, k2, k3, k4);
            if (e < best_e) { best_e = e; best_i = i; }
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        const double oe = __shfl_xor_sync(0xffffffffu, best_e, off);
        const int oi = __shfl_xor_sync(0xffffffffu, best_i, off);
        if (oe < best_e || (oe == best_e && oi < best_i)) { best_e = oe; best_i = oi; }
    }
    return best_i == 0x7fffffff ? 0 : best_i;
}

// Phase B for the nq slots of the queue (whole CTA; contains __syncthreads, so every thread must
// call it with the same nq).  Writes Q.arg[q] = argmin + 2 (intensities.rs:32-36) for each live line
// and then calls finish(q) -- phase C and whatever the kernel does with its result -- from the
// thread that searched the line, so no CTA barrier separates the search from phase C and a warp
// that is done with its lines never waits for the warp holding the longest ones.
//
// The CTA counting-sorts its lines by length (longest first).  Lines of more than LONG_LINE samples
// are searched by a whole warp each (search_line_warp).  Every other line is walked by one thread
// from its first sample to its last, exactly as intensities.rs:16-29 does: a sliding window of five
// samples in registers, the reference's error for every window, the first minimum kept by
// `e < best`; thanks to the sort the 32 lines of a warp have (nearly) the same length.
template <typename Finish>
__device__ __forceinline__ void phase_b(const UpdArgs &a, CtaQueue &Q, int p, int nq, Finish finish)
{
    const int tid = threadIdx.x, W = a.W, H = a.H;
    if (tid < SORT_BUCKETS) Q.bucket[tid] = 0;
    __syncthreads();
    const int n_mine = tid < nq ? Q.n[tid] : 0;
    const int b_mine = SORT_BUCKETS - 1 - min(n_mine, SORT_BUCKETS - 1);   // long lines first
    if (tid < nq) atomicAdd(&Q.bucket[b_mine], 1);
    __syncthreads();
    if (tid < 32) {   // exclusive prefix sum over the 64 buckets by one warp
        const int c0 = Q.bucket[2 * tid], c1 = Q.bucket[2 * tid + 1];
        int incl = c0 + c1;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int o = __shfl_up_sync(0xffffffffu, incl, off);
            if (tid >= off) incl += o;
        }
        Q.bucket[2 * tid] = incl - c0 - c1;
        Q.bucket[2 * tid + 1] = incl - c1;
        // lines with more than LONG_LINE samples occupy buckets [0, SORT_BUCKETS - 1 - LONG_LINE)
        constexpr int B0 = SORT_BUCKETS - 1 - LONG_LINE;
        if (2 * tid == B0) Q.nlong = incl - c0 - c1;
        if (2 * tid + 1 == B0) Q.nlong = incl - c1;
    }
    __syncthreads();
    if (tid < nq) Q.perm[atomicAdd(&Q.bucket[b_mine], 1)] = (unsigned short)tid;
    __syncthreads();
    const int nlong = Q.nlong;
    for (int r = tid >> 5; r < nlong; r += BLOCK / 32) {   // long lines: one warp each
        const int q = Q.perm[r];
        const LineGeom g = line_geom(a, Q, p, q);
        const int arg = search_line_warp(g, Q.n[q], a.step, W, H, Q.k[0][q], Q.k[1][q], Q.k[2][q], Q.k[3][q],
                                         Q.k[4][q], tid & 31);
        if ((tid & 31) == 0) Q.arg[q] = arg + 2;
    }
    if (nlong > 0) {   // phase C of the long lines this warp searched, one lane per line
        __syncwarp();
        for (int r = (tid >> 5) + (BLOCK / 32) * (tid & 31); r < nlong; r += BLOCK) finish((int)Q.perm[r]);
    }
    const int r = nlong + tid;   // the other lines: one thread each
    if (r >= nq) return;
    const int q = Q.perm[r];
    const int n = Q.n[q];
    if (n == 0) return;   // dropped in A2
    const LineGeom g = line_geom(a, Q, p, q);
    const double k0 = Q.k[0][q], k1 = Q.k[1][q], k2 = Q.k[2][q], k3 = Q.k[3][q], k4 = Q.k[4][q];
    double s0 = line_sample(g, 0, a.step, W, H), s1 = line_sample(g, 1, a.step, W, H);
    double s2 = line_sample(g, 2, a.step, W, H), s3 = line_sample(g, 3, a.step, W, H);
    double best = INFINITY;
    int arg = 0;
    for (int i = 0; i + 4 < n; i++) {
        const double s4 = line_sample(g, i + 4, a.step, W, H);
        const double e = window_error(s0, s1, s2, s3, s4, k0, k1, k2, k3, k4);
        if (e < best) { best = e; arg = i; }
        s0 = s1; s1 = s2; s2 = s3; s3 = s4;
    }
    Q.arg[q] = arg + 2;
    finish(q);
}

// ---------------------------------------------------------------------------------------------
// One-kernel plan: a CTA owns a 64x4 tile, compacts its own candidates and runs every phase.
// Needs no global queue (this is what tdk_update_depth_static_dev runs), but only the tile's own
// ~5-15 % of candidates populate phases A2/B/C.
template <bool SUBPIX>
__global__ void __launch_bounds__(BLOCK, MIN_CTAS) update_depth_tile_kernel(const UpdArgs a)
{
    extern __shared__ __align__(16) unsigned char s_dyn[];
    CtaQueue &Q = *reinterpret_cast<CtaQueue *>(s_dyn);
    __shared__ double s_depth[BLOCK], s_var[BLOCK];
    __shared__ int s_arg[BLOCK];
    __shared__ signed char s_flag[BLOCK];
    __shared__ double s_xk[TILE_W], s_yk[TILE_H];    // normalised key coordinates of the tile
    __shared__ int s_nq;
    __shared__ unsigned int s_hist[10];

    const int tid = threadIdx.x;
    const int p = (int)blockIdx.z + a.p0;
    const int ty = blockIdx.y, tx = blockIdx.x;
    const int W = a.W, H = a.H;
    const size_t map_base = (size_t)p * H * W;
    const TdkKeyConst key = a.keys[p];
    const double wm1 = (double)W - 1., hm1 = (double)H - 1.;

    if (tid == 0) s_nq = 0;
    if (tid < 10) s_hist[tid] = 0;
    // camera.rs:55-57: x_key depends on the column (row) only -- one division per column (row) of
    // the tile instead of two per pixel
    if (tid < TILE_W) s_xk[tid] = ((double)(tx * TILE_W + tid) - key.ox) / key.fx;
    else if (tid < TILE_W + TILE_H) s_yk[tid - TILE_W] = ((double)(ty * TILE_H + tid - TILE_W) - key.oy) / key.fy;
    __syncthreads();

    // ---- phase A1
    const int lx = tid & (TILE_W - 1), ly = tid / TILE_W;
    const int x = tx * TILE_W + lx, y = ty * TILE_H + ly;
    const bool valid = x < W && y < H;
    const size_t pix = map_base + (size_t)y * W + x;
    int flag = TDK_FLAG_NOT_PROCESSED;
    s_arg[tid] = -1;
    if (valid) {
        const uint64_t age = a.age[pix];
        const double d = a.prior_d[pix], v = a.prior_v[pix];
        double out_d = d;
        if (age != 0) {   // semi_dense.rs:194-200
            if (age > (uint64_t)a.R) {
                atomicExch(&a.status[TDK_ST_AGE], 1);
            } else {
                A1Result c;
                flag = phase_a1(a, key, p, s_xk[lx], s_yk[ly], age, d, v, wm1, hm1, out_d, c);
                if (flag == 1) queue_push(Q, atomicAdd(&s_nq, 1), c, v, s_xk[lx], s_yk[ly], x, y, tid);
            }
        }
        s_depth[tid] = out_d;
        s_var[tid] = v;
    }
    s_flag[tid] = (signed char)flag;
    __syncthreads();

    // ---- phase A2
    const int nq = s_nq;
    if (tid < nq) {
        const int f = phase_a2(a, Q, p, tid, wm1, hm1);
        if (f != 1) s_flag[Q.owner[tid]] = (signed char)f;
    }
    __syncthreads();

    // ---- phases B and C
    phase_b(a, Q, p, nq, [&](int q) {
        double nd, nv;
        const int f = phase_c<SUBPIX>(a, Q, key, p, q, map_base, nd, nv);
        const int owner = Q.owner[q];
        if (f == TDK_FLAG_SUCCESS) { s_depth[owner] = nd; s_var[owner] = nv; }
        s_flag[owner] = (signed char)f;
        s_arg[owner] = Q.arg[q];
    });
    __syncthreads();

    // ---- store
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
// Queue plans.  key_prep_kernel leaves an f32 copy of every key image of the sub-batch and a bound on
// its largest magnitude in the workspace; sieve_kernel runs phase A0 on every pixel, stores the
// final values of every pixel it can decide (and the provisional "failed estimate" values of the
// others) and appends the undecided pixels of its tile to the keyframe's queue in global memory.
__device__ __forceinline__ float prep_max(float m, double v)
{
    const float av = __double2float_ru(fabs(v));
    return av == av ? fmaxf(m, av) : INFINITY;   // a NaN in the image switches the cheap path off
}

__global__ void __launch_bounds__(256) key_prep_kernel(const TdkKeyConst *__restrict__ keys, int p0, size_t hw,
                                                       float *__restrict__ img32, unsigned int *amax)
{
    const int pl = blockIdx.y;
    const double *__restrict__ src = keys[pl + p0].img;
    float *__restrict__ dst = img32 + (size_t)pl * hw;
    float m = 0.f;
    constexpr int PER = 4;   // independent 16-byte loads in flight per thread
    const size_t stride = (size_t)gridDim.x * blockDim.x, first = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    // pairs of pixels with 16-byte loads and 8-byte stores where the image allows it (the f32 copy
    // of keyframe pl starts at an even element iff pl * hw is even)
    const bool wide = ((reinterpret_cast<uintptr_t>(src) & 15) == 0) && ((reinterpret_cast<uintptr_t>(dst) & 7) == 0);
    const size_t npair = wide ? hw / 2 : 0;
    const double2 *__restrict__ src2 = reinterpret_cast<const double2 *>(src);
    float2 *__restrict__ dst2 = reinterpret_cast<float2 *>(dst);
    for (size_t i0 = first; i0 < npair; i0 += stride * PER) {
        double2 v[PER];
#pragma unroll
        for (int k = 0; k < PER; k++) {
            const size_t i = i0 + k * stride;
            v[k] = i < npair ? __ldg(src2 + i) : make_double2(0., 0.);
        }
#pragma unroll
        for (int k = 0; k < PER; k++) {
            const size_t i = i0 + k * stride;
            if (i < npair) dst2[i] = make_float2((float)v[k].x, (float)v[k].y);
            m = prep_max(prep_max(m, v[k].x), v[k].y);
        }
    }
    for (size_t i = 2 * npair + first; i < hw; i += stride) {   // the odd last pixel, or the whole image if unaligned
        const double v = __ldg(src + i);
        dst[i] = (float)v;
        m = prep_max(m, v);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
    // non-negative floats order like their bit patterns
    if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(amax + pl, __float_as_uint(m));
}

#ifndef TDK_SIEVE_PPT
#define TDK_SIEVE_PPT 2
#endif
constexpr int SIEVE_PPT = TDK_SIEVE_PPT;              // pixels per thread: rows ly, ly + 4, ... of a 64 x (4 PPT) tile
constexpr int SIEVE_TILE_H = TILE_H * SIEVE_PPT;
static_assert(TILE_W + SIEVE_TILE_H + 1 <= BLOCK, "the table set-up needs one thread per column, row and threshold");

__global__ void __launch_bounds__(BLOCK, SIEVE_MIN_CTAS) update_depth_sieve_kernel(const UpdArgs a)
{
    __shared__ double s_xk[TILE_W], s_yk[SIEVE_TILE_H];
    __shared__ unsigned int s_hist[10];
    __shared__ float s_thr2;
#ifdef TDK_SIEVE_TMA
    // the CTA's key tile + halo, fetched row by row by the TMA engine (cp.async.bulk, one elected
    // thread, completion counted in bytes on an mbarrier) while the threads load their maps and run
    // the exact part of phase A0
    constexpr int ST_H = SIEVE_TILE_H + 2 * ST_HALO;
    __shared__ __align__(128) float s_tile[ST_H * ST_W];
    __shared__ __align__(8) unsigned long long s_bar;
#endif

    const int tid = threadIdx.x;
    const int pl = blockIdx.z, p = pl + a.p0;
    const int ty = blockIdx.y, tx = blockIdx.x;
    const int W = a.W, H = a.H;
    const size_t hw = (size_t)H * W, map_base = (size_t)p * hw;
#ifdef TDK_SIEVE_TMA
    const int rx0 = tx * TILE_W - ST_HALO, ry0 = ty * SIEVE_TILE_H - ST_HALO;
    const bool use_tile = (W & 3) == 0;   // 16-byte granularity of the bulk copies
    if (use_tile && tid == 0) {
        const unsigned bar = (unsigned)__cvta_generic_to_shared(&s_bar);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        const int cx0 = rx0 < 0 ? 0 : rx0, cx1 = min(W, rx0 + ST_W);
        const int gy0 = ry0 < 0 ? 0 : ry0, gy1 = min(H, ry0 + ST_H);
        const unsigned row_bytes = (unsigned)(cx1 - cx0) * 4u;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(row_bytes * (unsigned)(gy1 - gy0)) : "memory");
        const float *src = a.img32 + (size_t)pl * hw + (size_t)gy0 * W + cx0;
        unsigned dst = (unsigned)__cvta_generic_to_shared(s_tile + (gy0 - ry0) * ST_W + (cx0 - rx0));
        for (int gy = gy0; gy < gy1; gy++, src += W, dst += ST_W * 4)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                         "r"(row_bytes), "r"(bar) : "memory");
    }
#endif
    const TdkKeyConst key = a.keys[p];

    const int lx = tid & (TILE_W - 1), ly = tid / TILE_W;
    const int x = tx * TILE_W + lx, y0 = ty * SIEVE_TILE_H + ly;
    // per-keyframe base pointers are uniform over the CTA; a pixel is addressed by its unsigned 32-bit
    // offset inside the keyframe (H * W < 2^31), one IMAD.WIDE.U32 per access
    const unsigned long long *__restrict__ g_age = reinterpret_cast<const unsigned long long *>(a.age) + map_base;
    const double *__restrict__ g_pd = a.prior_d + map_base, *__restrict__ g_pv = a.prior_v + map_base;
    // every map load of the thread's pixels goes out first: their DRAM latency overlaps the table
    // set-up below and each other
    uint64_t age[SIEVE_PPT];
    double d[SIEVE_PPT], v[SIEVE_PPT];
#pragma unroll
    for (int k = 0; k < SIEVE_PPT; k++) {
        const int y = y0 + k * TILE_H;
        age[k] = 0; d[k] = 0.; v[k] = 0.;
        if (x < W && y < H) {
            const unsigned o = (unsigned)(y * W + x);
            age[k] = g_age[o]; d[k] = g_pd[o]; v[k] = g_pv[o];
        }
    }

    if (tid < 10) s_hist[tid] = 0;
    // camera.rs:55-57: x_key depends on the column (row) only -- one division per column (row) of
    // the tile instead of two per pixel
    if (tid < TILE_W) s_xk[tid] = ((double)(tx * TILE_W + tid) - key.ox) / key.fx;
    else if (tid < TILE_W + SIEVE_TILE_H) s_yk[tid - TILE_W] = ((double)(ty * SIEVE_TILE_H + tid - TILE_W) - key.oy) / key.fy;
    else if (tid == TILE_W + SIEVE_TILE_H) s_thr2 = cert_threshold(a, a.amax[pl]);
    __syncthreads();

    const float *__restrict__ img32 = a.img32 + (size_t)pl * hw;
    const float thr2 = s_thr2;
    const double xk = s_xk[lx];
    const int lane = tid & 31;
#ifdef TDK_SIEVE_TMA
    KeyTile tile{nullptr, rx0, ry0, ST_H};
    if (use_tile) {   // the barrier was initialised before the __syncthreads above; phase 0 completes when every row has landed
        const unsigned bar = (unsigned)__cvta_generic_to_shared(&s_bar);
        asm volatile("{\n .reg .pred p;\n TDK_WAIT:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n @p bra TDK_DONE;\n bra TDK_WAIT;\n TDK_DONE:\n}" ::"r"(bar) : "memory");
        tile.s = s_tile;
    }
#else
    const KeyTile tile{nullptr, 0, 0, 0};
#endif
#pragma unroll
    for (int k = 0; k < SIEVE_PPT; k++) {
        const int y = y0 + k * TILE_H;
        const bool valid = x < W && y < H;
        int flag = TDK_FLAG_NOT_PROCESSED;
        if (valid) {
            const unsigned o = (unsigned)(y * W + x);
            double out_d = d[k];
            if (age[k] != 0) {
                if (age[k] > (uint64_t)a.R) {
                    atomicExch(&a.status[TDK_ST_AGE], 1);
                } else {
                    flag = phase_a0(a, key, img32, p, x, y, xk, s_yk[ly + k * TILE_H], age[k], d[k], v[k], thr2, out_d, tile);
                }
            }
            // undecided pixels get the values a failed estimate keeps; the later kernels overwrite the
            // flag if the exact path fails them and all three values on success
            (a.out_d + map_base)[o] = out_d;
            (a.out_v + map_base)[o] = v[k];
            (a.out_f + map_base)[o] = (long long)(flag == 1 ? TDK_FLAG_SUCCESS : flag);
            if (a.dbg) (a.dbg + map_base)[o] = -1;
            if (a.hist && flag != 1) atomicAdd(&s_hist[-flag], 1u);
        }
        // append this warp's undecided pixels to the keyframe's queue: one global atomic per warp, no CTA barrier
        const unsigned cmask = __ballot_sync(0xffffffffu, flag == 1);
        if (cmask) {
            unsigned int base = 0;
            if (lane == 0) base = atomicAdd(a.cand_count + pl, (unsigned int)__popc(cmask));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (flag == 1) (a.cand + (size_t)pl * hw)[base + __popc(cmask & ((1u << lane) - 1u))] = (unsigned int)(y * W + x);
        }
    }
    if (a.hist) {
        __syncthreads();
        if (tid < 10 && s_hist[tid]) atomicAdd(&a.hist[(size_t)p * 10 + tid], (unsigned long long)s_hist[tid]);
    }
}

template <bool SUBPIX>
__global__ void __launch_bounds__(BLOCK, MIN_CTAS) update_depth_search_kernel(const UpdArgs a)
{
    extern __shared__ __align__(16) unsigned char s_dyn[];
    CtaQueue &Q = *reinterpret_cast<CtaQueue *>(s_dyn);
    __shared__ unsigned int s_hist[10];

    const int tid = threadIdx.x;
    const int pl = blockIdx.z, p = pl + a.p0;
    const int total = (int)a.cand_count[pl];
    const int W = a.W, H = a.H;
    const size_t hw = (size_t)H * W, map_base = (size_t)p * hw;
    const TdkKeyConst key = a.keys[p];
    const double wm1 = (double)W - 1., hm1 = (double)H - 1.;
    if (tid < 10) s_hist[tid] = 0;
    __syncthreads();

    // the grid holds a quarter of the keyframe's tiles per keyframe (usually 5-20 % of the pixels
    // are queued); CTAs stride over the queue in chunks of BLOCK pixels
    for (int first = (int)blockIdx.x * BLOCK; first < total; first += (int)gridDim.x * BLOCK) {
        const int nq = min(BLOCK, total - first);
        // ---- the exact phase A1 for this chunk's pixels (the sieve could not decide them), then A2
        if (tid < nq) {
            const unsigned int li = a.cand[(size_t)pl * hw + first + tid];
            const int y = (int)(li / (unsigned int)W), x = (int)(li - (unsigned int)y * W);
            const size_t pix = map_base + li;
            const uint64_t age = a.age[pix];
            const double d = a.prior_d[pix], v = a.prior_v[pix];
            const double xk = ((double)x - key.ox) / key.fx, yk = ((double)y - key.oy) / key.fy;
            A1Result c;
            double fail_d;
            int flag = phase_a1(a, key, p, xk, yk, age, d, v, wm1, hm1, fail_d, c);
            Q.n[tid] = 0;
            if (flag == 1) {
                queue_push(Q, tid, c, v, xk, yk, x, y, tid);
                flag = phase_a2(a, Q, p, tid, wm1, hm1);
            }
            if (flag != 1) {   // failed in A1 / A2: the sieve's provisional depth / variance stand, only the flag changes
                a.out_f[pix] = (long long)flag;
                if (a.hist) atomicAdd(&s_hist[-flag], 1u);
            }
        }
        __syncthreads();

        // ---- phases B and C, then the scattered store, all by the thread that searched the line
        phase_b(a, Q, p, nq, [&](int q) {
            double nd, nv;
            const int f = phase_c<SUBPIX>(a, Q, key, p, q, map_base, nd, nv);
            const size_t at = map_base + (size_t)(Q.px[q] >> 16) * W + (Q.px[q] & 0xffff);
            if (f == TDK_FLAG_SUCCESS) { a.out_d[at] = nd; a.out_v[at] = nv; }
            else a.out_f[at] = (long long)f;
            if (a.dbg) a.dbg[at] = Q.arg[q];
            if (a.hist) atomicAdd(&s_hist[-f], 1u);
        });
        __syncthreads();   // the queue is reused by the next chunk
    }
    if (a.hist) {
        __syncthreads();
        if (tid < 10 && s_hist[tid]) atomicAdd(&a.hist[(size_t)p * 10 + tid], (unsigned long long)s_hist[tid]);
    }
}

// ---------------------------------------------------------------------------------------------
// Sorted plan (default).  After the sieve, three kernels per sub-batch of keyframes:
//   lines_kernel  thread per queued pixel: the exact phases A1 and A2, a 128-byte line record into the
//                 workspace, the line's length class and a per-keyframe histogram;
//   order_kernel  counting sort by length class: fills `order` with the live queue slots, longest first;
//   match_kernel  a warp takes 32 consecutive entries of `order` -- lines of (nearly) equal length --
//                 and every lane walks its own line exactly as intensities.rs:16-29 does, then runs
//                 phase C and stores.  No shared-memory queue and no CTA barrier: a warp never waits
//                 for another warp's longer lines.  Lines of LEN_CLASSES-1 samples or more take a whole
//                 warp each (search_line_warp).
constexpr int LEN_CLASSES = 64;
#ifndef TDK_LINES_MIN_CTAS
#define TDK_LINES_MIN_CTAS 4
#endif
#ifndef TDK_MATCH_MIN_CTAS
#define TDK_MATCH_MIN_CTAS 4
#endif
constexpr int LINES_MIN_CTAS = TDK_LINES_MIN_CTAS, MATCH_MIN_CTAS = TDK_MATCH_MIN_CTAS;
constexpr int REC_DOUBLE2 = 8;   // 128 bytes per queue slot

struct RecTail { int n, ref, px, pad; };

__device__ __forceinline__ void record_store(double2 *__restrict__ r, const LineRec &l, int ref, int px, double v)
{
    r[0] = make_double2(l.k0, l.k1);
    r[1] = make_double2(l.k2, l.k3);
    r[2] = make_double2(l.k4, l.xminx);
    r[3] = make_double2(l.xminy, l.ux);
    r[4] = make_double2(l.uy, l.rdx);
    r[5] = make_double2(l.rdy, l.g);
    r[6] = make_double2(l.kstep, v);
    RecTail t{l.n, ref, px, 0};
    r[7] = *reinterpret_cast<const double2 *>(&t);
}

__global__ void __launch_bounds__(BLOCK, LINES_MIN_CTAS) update_depth_lines_kernel(const UpdArgs a)
{
    __shared__ unsigned int s_hist[10], s_len[LEN_CLASSES];
    constexpr int REC_PAD = REC_DOUBLE2 + 1;   // 144-byte pitch: 16-byte accesses of a quarter warp hit 32 distinct banks
    __shared__ double2 s_rec[BLOCK * REC_PAD];
    const int tid = threadIdx.x;
    const int pl = blockIdx.z, p = pl + a.p0;
    const int total = (int)a.cand_count[pl];
    if ((int)blockIdx.x * BLOCK >= total) return;
    const int W = a.W, H = a.H;
    const size_t hw = (size_t)H * W, map_base = (size_t)p * hw, ws_base = (size_t)pl * hw;
    const TdkKeyConst key = a.keys[p];
    const double wm1 = (double)W - 1., hm1 = (double)H - 1.;
    const size_t rec_base = (size_t)pl * a.rec_stride;
    if (tid < 10) s_hist[tid] = 0;
    if (tid < LEN_CLASSES) s_len[tid] = 0;
    __syncthreads();

    for (int first = (int)blockIdx.x * BLOCK; first < total; first += (int)gridDim.x * BLOCK) {
        const int i = first + tid;
        if (i < total) {
            const unsigned int li = a.cand[ws_base + i];
            const int y = (int)(li / (unsigned int)W), x = (int)(li - (unsigned int)y * W);
            const size_t pix = map_base + li;
            const uint64_t age = a.age[pix];
            const double d = a.prior_d[pix], v = a.prior_v[pix];
            const double xk = ((double)x - key.ox) / key.fx, yk = ((double)y - key.oy) / key.fy;
            A1Result c;
            double fail_d;
            LineRec l;
            int flag = phase_a1(a, key, p, xk, yk, age, d, v, wm1, hm1, fail_d, c);
            if (flag == 1) flag = phase_a2_core(a, a.pairs + (size_t)p * a.R + c.r, xk, yk, v, c, wm1, hm1, l);
            int cls = 0;
            if (flag == 1) {
                cls = min(l.n, LEN_CLASSES - 1);
                record_store(s_rec + tid * REC_PAD, l, c.r, x | (y << 16), v);
                atomicAdd(&s_len[cls], 1u);
            } else {   // failed in A1 / A2: the sieve's provisional depth / variance stand, only the flag changes
                a.out_f[pix] = (long long)flag;
                if (a.hist) atomicAdd(&s_hist[-flag], 1u);
            }
            a.len_class[ws_base + i] = (unsigned char)cls;
        }
        __syncthreads();
        // the chunk's 256 records are 32 KB of consecutive workspace: copy them out with consecutive
        // threads on consecutive 16-byte pieces (records of dropped lines are never read)
        double2 *__restrict__ out = a.rec + (rec_base + first) * REC_DOUBLE2;
#pragma unroll
        for (int it = 0; it < REC_DOUBLE2; it++) {
            const int idx = it * BLOCK + tid;
            out[idx] = s_rec[(idx >> 3) * REC_PAD + (idx & 7)];
        }
        __syncthreads();   // s_rec is reused by the next chunk
    }
    if (tid < LEN_CLASSES && s_len[tid]) atomicAdd(&a.len_hist[pl * LEN_CLASSES + tid], s_len[tid]);
    if (a.hist && tid < 10 && s_hist[tid]) atomicAdd(&a.hist[(size_t)p * 10 + tid], (unsigned long long)s_hist[tid]);
}

// first position in `order` of each length class, longest class first: start[c] = sum of hist[c' > c]
__device__ __forceinline__ void class_starts(const unsigned int *__restrict__ hist, unsigned int *s_start, int tid)
{
    if (tid < 32) {
        const unsigned int c_hi = hist[LEN_CLASSES - 1 - 2 * tid], c_lo = hist[LEN_CLASSES - 2 - 2 * tid];
        unsigned int incl = c_hi + c_lo;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const unsigned int o = __shfl_up_sync(0xffffffffu, incl, off);
            if (tid >= off) incl += o;
        }
        s_start[LEN_CLASSES - 1 - 2 * tid] = incl - c_hi - c_lo;
        s_start[LEN_CLASSES - 2 - 2 * tid] = incl - c_lo;
    }
}

constexpr int ORDER_ITEMS = 4;   // queue slots per thread and round of order_kernel

__global__ void __launch_bounds__(BLOCK) update_depth_order_kernel(const UpdArgs a)
{
    __shared__ unsigned int s_start[LEN_CLASSES], s_cnt[LEN_CLASSES], s_off[LEN_CLASSES];
    const int tid = threadIdx.x;
    const int pl = blockIdx.z;
    const int total = (int)a.cand_count[pl];
    constexpr int ROUND = BLOCK * ORDER_ITEMS;
    if ((int)blockIdx.x * ROUND >= total) return;
    const size_t ws_base = (size_t)pl * a.H * a.W;   // a multiple of 4 is not guaranteed: byte loads below
    class_starts(a.len_hist + pl * LEN_CLASSES, s_start, tid);
    for (int first = (int)blockIdx.x * ROUND; first < total; first += (int)gridDim.x * ROUND) {
        if (tid < LEN_CLASSES) s_cnt[tid] = 0;
        __syncthreads();
        int cls[ORDER_ITEMS];
        unsigned int rank[ORDER_ITEMS];
#pragma unroll
        for (int k = 0; k < ORDER_ITEMS; k++) {
            const int i = first + k * BLOCK + tid;
            cls[k] = i < total ? (int)a.len_class[ws_base + i] : 0;
            rank[k] = cls[k] ? atomicAdd(&s_cnt[cls[k]], 1u) : 0u;
        }
        __syncthreads();
        if (tid < LEN_CLASSES && tid > 0 && s_cnt[tid])
            s_off[tid] = atomicAdd(&a.len_cursor[pl * LEN_CLASSES + tid], s_cnt[tid]);
        __syncthreads();
#pragma unroll
        for (int k = 0; k < ORDER_ITEMS; k++)
            if (cls[k]) a.order[ws_base + s_start[cls[k]] + s_off[cls[k]] + rank[k]] = (unsigned int)(first + k * BLOCK + tid);
    }
}

#ifndef TDK_MATCH_BLOCK
#define TDK_MATCH_BLOCK 128   // four warps per CTA: finer-grained scheduling of the longest-first item order (18.1 vs 18.3 us per keyframe)
#endif
constexpr int MATCH_BLOCK = TDK_MATCH_BLOCK;
template <bool SUBPIX>
__global__ void __launch_bounds__(MATCH_BLOCK, MATCH_MIN_CTAS * BLOCK / MATCH_BLOCK) update_depth_match_kernel(const UpdArgs a)
{
    __shared__ unsigned int s_hist[10], s_start[LEN_CLASSES];
    const int tid = threadIdx.x, lane = tid & 31;
    const int pl = blockIdx.z, p = pl + a.p0;
    const int W = a.W, H = a.H;
    const size_t hw = (size_t)H * W, map_base = (size_t)p * hw, ws_base = (size_t)pl * hw;
    if (tid < 10) s_hist[tid] = 0;
    class_starts(a.len_hist + pl * LEN_CLASSES, s_start, tid);
    __syncthreads();
    const int nlive = (int)s_start[0];                 // class 0 (dropped lines) would start here
    // Lines of class >= whole come first in `order` and take a whole warp each.  Class 63 always
    // does; when the launch is small enough to leave warps idle (a.warp_budget warps per keyframe
    // can be resident at once), shorter lines do too, longest first, as far as the budget goes:
    // the search of a single keyframe is a latency chain, and a 40-sample line walked by one lane
    // is 20 times longer than the same line walked by a warp.
    int whole = LEN_CLASSES - 1;
    while (whole > 8) {
        const int nl = (int)s_start[whole - 2];   // lines of class >= whole - 1
        if (nl + (nlive - nl + 31) / 32 > a.warp_budget) break;
        whole--;
    }
    const int nlong = (int)s_start[whole - 1];
    const int nitems = nlong + (nlive - nlong + 31) / 32;
    const TdkKeyConst key = a.keys[p];
    const int warps = MATCH_BLOCK / 32;

    for (int item = (int)blockIdx.x * warps + (tid >> 5); item < nitems; item += (int)gridDim.x * warps) {
        __syncwarp();   // `item` is warp-uniform: lanes that left the previous item early wait here
        const bool whole_warp = item < nlong;
        const int pos = whole_warp ? item : nlong + (item - nlong) * 32 + lane;
        const bool active = pos < nlive;
        const unsigned int slot = active ? a.order[ws_base + pos] : 0u;
        if (!active) continue;
        const double2 *__restrict__ r = a.rec + ((size_t)pl * a.rec_stride + slot) * REC_DOUBLE2;
        const double2 r0 = r[0], r1 = r[1], r2 = r[2], r3 = r[3], r4 = r[4];
        const double2 r7 = r[7];
        const RecTail t = *reinterpret_cast<const RecTail *>(&r7);
        const TdkPairConst *__restrict__ pc = a.pairs + (size_t)p * a.R + t.ref;
        const LineGeom g{pc->ref, pc->rfx, pc->rfy, pc->rox, pc->roy, r2.y, r3.x, r3.y, r4.x};
        const double k0 = r0.x, k1 = r0.y, k2 = r1.x, k3 = r1.y, k4 = r2.x;
        int arg = 0;
        if (whole_warp) {
            arg = search_line_warp(g, t.n, a.step, W, H, k0, k1, k2, k3, k4, lane);
            if (lane != 0) continue;
        } else {
            double s0 = line_sample(g, 0, a.step, W, H), s1 = line_sample(g, 1, a.step, W, H);
            double s2 = line_sample(g, 2, a.step, W, H), s3 = line_sample(g, 3, a.step, W, H);
            double best = INFINITY;
            for (int i = 0; i + 4 < t.n; i++) {
                const double s4 = line_sample(g, i + 4, a.step, W, H);
                const double e = window_error(s0, s1, s2, s3, s4, k0, k1, k2, k3, k4);
                if (e < best) { best = e; arg = i; }
                s0 = s1; s1 = s2; s2 = s3; s3 = s4;
            }
        }
        arg += 2;   // intensities.rs:32-36
        // ---- phase C and the scattered store
        const int sx = t.px & 0xffff, sy = t.px >> 16;
        const double xk = ((double)sx - key.ox) / key.fx, yk = ((double)sy - key.oy) / key.fy;
        const double2 r5 = r[5], r6 = r[6];
        double nd, nv;
        const double sub = SUBPIX ? subpixel_offset(g, t.n, arg - 2, a.step, W, H, k0, k1, k2, k3, k4) : 0.;
        const int f = phase_c_core(a, key, pc, arg, sx, sy, xk, yk, r2.y, r3.x, r3.y, r4.x, r4.y, r5.x, r5.y,
                                   r6.x, r6.y, map_base, nd, nv, sub);
        const size_t at = map_base + (size_t)sy * W + sx;
        if (f == TDK_FLAG_SUCCESS) { a.out_d[at] = nd; a.out_v[at] = nv; }
        else a.out_f[at] = (long long)f;
        if (a.dbg) a.dbg[at] = arg;
        if (a.hist) atomicAdd(&s_hist[-f], 1u);
    }
    if (a.hist) {
        __syncthreads();
        if (tid < 10 && s_hist[tid]) atomicAdd(&a.hist[(size_t)p * 10 + tid], (unsigned long long)s_hist[tid]);
    }
}

// ---------------------------------------------------------------------------------------------
// estimate_debug_ for a pixel OUTSIDE the image (src/py/semi_dense.rs:126-155 accepts any u_key): the
// exact phase A1 on that one pixel.  Its centre sample is the pixel itself, so the range test
// (flag -2) fails before any image access unless an earlier test (-7, -1, -8) already did.
__global__ void estimate_outside_kernel(const UpdArgs a, long long x, long long y, double d, double v, int *out_flag)
{
    const TdkKeyConst key = a.keys[0];
    const double xk = ((double)x - key.ox) / key.fx, yk = ((double)y - key.oy) / key.fy;
    A1Result c;
    double fail_d;
    const int f = phase_a1(a, key, 0, xk, yk, 1, d, v, (double)a.W - 1., (double)a.H - 1., fail_d, c);
    *out_flag = f == 1 ? TDK_FLAG_KEY_OUT_OF_RANGE : f;
}

// ---------------------------------------------------------------------------------------------
size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct WsLayout {
    size_t status, keys, pairs, key_ptrs, ref_ptrs, total_tile;
    size_t counters, counters_end, cand, img32, total_queue, len_class, order, rec, total;
    int64_t sub;   // keyframes per sub-batch of the queue plans
};

// The queue plans' per-keyframe arrays (8 bytes per pixel for the queue and the f32 key image, plus
// 133 bytes per pixel for the sorted plan's line records: worst case, every pixel queued) are sized
// for a sub-batch of keyframes and reused by the sub-batches of a larger launch; the kernels of
// successive sub-batches run in stream order.  Every sub-batch pays its kernels' tails, so bigger is
// faster.  tdk_update_depth_workspace_bytes() asks for min(P, SUB_BATCH) keyframes' worth; launch()
// uses as many as the workspace it is given holds, down to one.
#ifndef TDK_SUB_BATCH
#define TDK_SUB_BATCH 64
#endif
constexpr int64_t SUB_BATCH = TDK_SUB_BATCH;
constexpr size_t COUNTER_WORDS = 2 + 2 * LEN_CLASSES;   // per sub-batch keyframe: queue length, max |image|, length histogram, cursor

// total_tile: what the one-kernel plan needs; total_queue: plus the queues and f32 key images of the
// two-kernel plan; total: plus the line records of the sorted plan
WsLayout ws_layout(int64_t P, int64_t R, int64_t H, int64_t W, int64_t sub = SUB_BATCH)
{
    WsLayout l;
    size_t off = 0;
    l.status = off; off = align_up(off + TDK_ST_WORDS * sizeof(int), 256);
    l.keys = off; off = align_up(off + (size_t)P * sizeof(TdkKeyConst), 256);
    l.pairs = off; off = align_up(off + (size_t)P * R * sizeof(TdkPairConst), 256);
    l.key_ptrs = off; off = align_up(off + (size_t)P * sizeof(void *), 256);
    l.ref_ptrs = off; off = align_up(off + (size_t)P * R * sizeof(void *), 256);
    l.total_tile = off;
    l.sub = P < sub ? P : sub;
    const size_t S = (size_t)l.sub, hw = (size_t)H * W;
    // counters, contiguous so that one sweep clears them: [S] queue lengths, [S] max |image| bits,
    // [S,64] length histogram, [S,64] cursor
    l.counters = off; off = align_up(off + S * COUNTER_WORDS * sizeof(unsigned int), 256);
    l.counters_end = off;
    l.cand = off; off = align_up(off + S * hw * sizeof(unsigned int), 256);
    l.img32 = off; off = align_up(off + S * hw * sizeof(float), 256);
    l.total_queue = off;
    l.len_class = off; off = align_up(off + S * hw, 256);
    l.order = off; off = align_up(off + S * hw * sizeof(unsigned int), 256);
    // records: whole chunks of BLOCK slots (lines_kernel writes them chunk by chunk)
    l.rec = off; off = align_up(off + S * align_up(hw, BLOCK) * REC_DOUBLE2 * sizeof(double2), 256);
    l.total = off;
    return l;
}

// The smallest double s >= 0 with sqrt(s) >= m, so that `sqrt(s) < m` and `s < threshold` are the same
// test for every s (IEEE sqrt is correctly rounded, hence monotone; the host's sqrt and the
// device's agree).  m <= 0: sqrt(s) < m never holds -> 0; m = NaN -> NaN (both tests always false).
double sqrt_threshold(double m)
{
    if (m != m) return m;
    if (m <= 0.) return 0.;
    if (m > 1.7976931348623157e308) return m;   // +inf: sqrt(s) < inf <=> s < inf
    double s = m * m;
    if (s > 1.7976931348623157e308) s = 1.7976931348623157e308;
    while (s > 0. && sqrt(s) >= m) s = nextafter(s, 0.);
    while (sqrt(s) < m) s = nextafter(s, INFINITY);
    return s;
}

enum { PLAN_AUTO = -1, PLAN_TWO_KERNELS = 0, PLAN_ONE_KERNEL = 1, PLAN_SORTED = 2 };
std::atomic<int> g_plan{PLAN_AUTO};
std::atomic<int> g_subpixel{0};

// per-device one-time set-up (kernel attributes, SM count), safe against concurrent first calls
struct DeviceInfo { std::once_flag once; int sm_count = 148; cudaError_t err = cudaSuccess; };
DeviceInfo g_device_info[64];

float float_down(double v) { float f = (float)v; return (double)f > v ? nextafterf(f, -INFINITY) : f; }
float float_up(double v) { float f = (float)v; return (double)f < v ? nextafterf(f, INFINITY) : f; }

// The sorted plan pays two more launches and a trip of the line records through the workspace; a
// launch of a single VGA-sized keyframe is too small to earn that back (50 against 57 us).
int resolve_plan(int plan, int64_t P, int64_t H, int64_t W)
{
    if (plan != PLAN_AUTO) return plan;
    return P * H * W <= 640 * 480 ? PLAN_TWO_KERNELS : PLAN_SORTED;
}


int launch(const double *key_imgs, const double *const *key_ptrs_host, const double *key_cams,
           const double *T_wk, const double *ref_imgs, const double *const *ref_ptrs_host,
           const double *ref_cams, const double *T_wr, const uint64_t *age, const double *prior_depth,
           const double *prior_variance, const double *params, int64_t P, int64_t R, int64_t H,
           int64_t W, double *out_depth, double *out_variance, int64_t *out_flag, int32_t *dbg_argmin,
           uint64_t *out_hist, int fuse, int plan, void *workspace, size_t workspace_bytes,
           cudaStream_t stream, bool tables_on_device = false)
{
    TdkRange range("tdk_update_depth");
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
    WsLayout l = ws_layout(P, R, H, W);
    auto need_of = [&](const WsLayout &w) {
        return plan == PLAN_ONE_KERNEL ? w.total_tile : plan == PLAN_TWO_KERNELS ? w.total_queue : w.total;
    };
    if (plan != PLAN_ONE_KERNEL)   // as many keyframes per sub-batch as the workspace holds
        while (l.sub > 1 && workspace_bytes < need_of(l)) l = ws_layout(P, R, H, W, l.sub - 1);
    const size_t need = need_of(l);
    if (workspace_bytes < need) {
        tdk_set_error("tdk_update_depth: workspace %zu < required %zu", workspace_bytes, need);
        return TDK_ERR_WORKSPACE;
    }
    const int tiles_x = (int)((W + TILE_W - 1) / TILE_W), tiles_y = (int)((H + TILE_H - 1) / TILE_H);
    const long long nblocks = (long long)P * tiles_x * tiles_y;
    if (nblocks > 2147483647LL || P * R > 2147483647LL || P > 65535 || tiles_y > 65535) {
        tdk_set_error("tdk_update_depth: batch too large for one launch (P=%lld, %lld tiles; at most 65535 "
                      "keyframes per call)", (long long)P, nblocks);
        return TDK_ERR_INVALID;
    }
    int dev = 0;
    TDK_CUDA(cudaGetDevice(&dev));
    int sm_count = 148;
    if (dev >= 0 && dev < 64) {
        DeviceInfo &di = g_device_info[dev];
        std::call_once(di.once, [&di, dev]() {
            di.err = cudaDeviceGetAttribute(&di.sm_count, cudaDevAttrMultiProcessorCount, dev);
            if (di.err == cudaSuccess)
                di.err = cudaFuncSetAttribute(update_depth_tile_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)CTA_SMEM_BYTES);
            if (di.err == cudaSuccess)
                di.err = cudaFuncSetAttribute(update_depth_tile_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)CTA_SMEM_BYTES);
            if (di.err == cudaSuccess)
                di.err = cudaFuncSetAttribute(update_depth_search_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)CTA_SMEM_BYTES);
            if (di.err == cudaSuccess)
                di.err = cudaFuncSetAttribute(update_depth_search_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)CTA_SMEM_BYTES);
        });
        TDK_CUDA(di.err);
        sm_count = di.sm_count;
    } else {
        tdk_set_error("tdk_update_depth: device ordinal %d not supported", dev);
        return TDK_ERR_INVALID;
    }
    char *ws = (char *)workspace;
    int *status = (int *)(ws + l.status);
    TdkKeyConst *keys = (TdkKeyConst *)(ws + l.keys);
    TdkPairConst *pairs = (TdkPairConst *)(ws + l.pairs);
    const double *const *key_ptrs = nullptr, *const *ref_ptrs = nullptr;
    TDK_CUDA(cudaMemsetAsync(status, 0, TDK_ST_WORDS * sizeof(int), stream));
    if (tables_on_device) {   // the caller's tables are device memory already
        key_ptrs = key_ptrs_host;
        ref_ptrs = ref_ptrs_host;
    } else {
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
    }
    const int npairs = (int)(P * R);
    const size_t counter_words = (size_t)l.sub * COUNTER_WORDS;
    pair_setup_kernel<<<(npairs + 127) / 128, 128, 0, stream>>>(
        key_imgs, key_ptrs, key_cams, T_wk, ref_imgs, ref_ptrs, ref_cams, T_wr, (int)P, (int)R,
        (size_t)H * W, keys, pairs, status,
        plan == PLAN_ONE_KERNEL ? nullptr : (unsigned int *)(ws + l.counters),
        plan == PLAN_ONE_KERNEL ? 0 : (int)counter_words);
    TDK_CUDA(cudaGetLastError());

    UpdArgs a;
    a.keys = keys; a.pairs = pairs; a.age = age; a.prior_d = prior_depth; a.prior_v = prior_variance;
    a.out_d = out_depth; a.out_v = out_variance; a.out_f = (long long *)out_flag; a.dbg = dbg_argmin;
    a.hist = (unsigned long long *)out_hist; a.status = status;
    // Params::new stores (inv(max_depth), inv(min_depth))  (py/semi_dense.rs:102-107)
    a.vmin = 1. / (params[1] + TDK_EPS_INV);
    a.vmax = 1. / (params[0] + TDK_EPS_INV);
    a.geo = params[2]; a.photo = params[3]; a.step = params[4]; a.min_grad = params[5];
    a.min_grad_sq = sqrt_threshold(a.min_grad);
    a.p0 = 0; a.warp_budget = 0; a.rec_stride = 0; a.rec = nullptr; a.len_class = nullptr; a.order = nullptr;
    a.len_hist = nullptr; a.len_cursor = nullptr; a.cand = nullptr; a.cand_count = nullptr; a.img32 = nullptr; a.amax = nullptr;
    a.P = (int)P; a.R = (int)R; a.H = (int)H; a.W = (int)W; a.tiles_x = tiles_x; a.tiles_y = tiles_y; a.fuse = fuse ? 1 : 0;
    a.subpix = g_subpixel.load();
    {   // constants of the certified sieve (phase_a0): coordinate error bound delta, margins, e / max |image|
        const double two22 = 1. / 4194304., two24 = 1. / 16777216.;
        const double delta = (2. * (double)CERT_HMAX + (double)(W > H ? W : H)) * two22;
        const double dm = delta > 1e-3 ? delta : 1e-3;
        a.cert_e = 2. * (4. * delta + 16. * two24) + two22;
        a.cert_m = a.min_grad > 0. ? a.min_grad * (1. - 1e-5) : -1.;   // NaN -> -1
        a.cert_lo = float_up(dm);
        a.cert_xhi = float_down((double)W - 1. - dm);
        a.cert_yhi = float_down((double)H - 1. - dm);
        a.cert_out = float_down(-dm);
        a.cert_xout = float_up((double)W - 1. + dm);
        a.cert_yout = float_up((double)H - 1. + dm);
    }
    if (plan == PLAN_ONE_KERNEL) {
        const dim3 tgrid((unsigned)tiles_x, (unsigned)tiles_y, (unsigned)P);
        if (a.subpix) update_depth_tile_kernel<true><<<tgrid, BLOCK, CTA_SMEM_BYTES, stream>>>(a);
        else update_depth_tile_kernel<false><<<tgrid, BLOCK, CTA_SMEM_BYTES, stream>>>(a);
    } else {
        const size_t hw = (size_t)H * W;
        unsigned int *counters = (unsigned int *)(ws + l.counters);
        a.cand_count = counters;
        a.amax = counters + l.sub;
        a.len_hist = counters + 2 * l.sub;
        a.len_cursor = a.len_hist + l.sub * LEN_CLASSES;
        a.cand = (unsigned int *)(ws + l.cand);
        a.img32 = (const float *)(ws + l.img32);
        if (plan == PLAN_SORTED) {
            a.len_class = (unsigned char *)(ws + l.len_class);
            a.order = (unsigned int *)(ws + l.order);
            a.rec = (double2 *)(ws + l.rec);
            a.rec_stride = align_up(hw, BLOCK);
        }
        const unsigned qx = (unsigned)((tiles_x * tiles_y + 3) / 4);
#ifndef TDK_PREP_PX_PER_BLOCK
#define TDK_PREP_PX_PER_BLOCK 8192   // four sweeps of eight pixels per thread: a CTA's first load (the image pointer out of
                                     // keys[]) is amortised (68 -> 42 us per 64 VGA keyframes, 5.7 TB/s)
#endif
        unsigned prep_x = (unsigned)((hw + TDK_PREP_PX_PER_BLOCK - 1) / TDK_PREP_PX_PER_BLOCK);
        if (prep_x > 4096u) prep_x = 4096u;
        for (int64_t p0 = 0; p0 < P; p0 += l.sub) {
            const unsigned ns = (unsigned)(P - p0 < l.sub ? P - p0 : l.sub);
            a.p0 = (int)p0;
            if (p0 > 0)   // pair_setup_kernel cleared them for the first sub-batch
                TDK_CUDA(cudaMemsetAsync(counters, 0, counter_words * sizeof(unsigned int), stream));
            key_prep_kernel<<<dim3(prep_x, ns), 256, 0, stream>>>(keys, (int)p0, hw, (float *)(ws + l.img32),
                                                                  counters + l.sub);
            update_depth_sieve_kernel<<<dim3((unsigned)tiles_x, (unsigned)((H + SIEVE_TILE_H - 1) / SIEVE_TILE_H), ns), BLOCK, 0, stream>>>(a);
            if (plan == PLAN_TWO_KERNELS) {
                if (a.subpix) update_depth_search_kernel<true><<<dim3(qx, 1u, ns), BLOCK, CTA_SMEM_BYTES, stream>>>(a);
                else update_depth_search_kernel<false><<<dim3(qx, 1u, ns), BLOCK, CTA_SMEM_BYTES, stream>>>(a);
            } else {
                // 32 resident warps per SM at match_kernel's 64 registers.  Only launches of one or two
                // keyframes leave warps idle; in a batch a sparse keyframe's spare warps belong to its
                // neighbours, and whole-warp searches of short lines would cost throughput there.
                a.warp_budget = ns <= 2 ? (int)((int64_t)sm_count * 32 / ns) : 0;
                const unsigned mx = (unsigned)((a.warp_budget + MATCH_BLOCK / 32 - 1) / (MATCH_BLOCK / 32));
                const unsigned mq = qx * (BLOCK / MATCH_BLOCK);
                update_depth_lines_kernel<<<dim3(qx, 1u, ns), BLOCK, 0, stream>>>(a);
                update_depth_order_kernel<<<dim3((qx + ORDER_ITEMS - 1) / ORDER_ITEMS, 1u, ns), BLOCK, 0, stream>>>(a);
                const dim3 mgrid(mx > mq ? mx : mq, 1u, ns);
                if (a.subpix) update_depth_match_kernel<true><<<mgrid, MATCH_BLOCK, 0, stream>>>(a);
                else update_depth_match_kernel<false><<<mgrid, MATCH_BLOCK, 0, stream>>>(a);
            }
            TDK_CUDA(cudaGetLastError());
        }
    }
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

}  // namespace

// internal (capi.cu): the flag estimate_debug_ returns for a pixel outside the image; all pointers
// but params are device pointers, scratch holds at least 1024 bytes
int tdk_estimate_outside_dev(const double *key_cam, const double *T_wk, const double *ref_cam, const double *T_wr,
                             const double *params, int64_t H, int64_t W, int64_t x, int64_t y, double prior_depth,
                             double prior_variance, int *out_flag, void *scratch, cudaStream_t stream)
{
    char *ws = (char *)scratch;
    int *status = (int *)ws;
    TdkKeyConst *keys = (TdkKeyConst *)(ws + 256);
    TdkPairConst *pairs = (TdkPairConst *)(ws + 512);
    TDK_CUDA(cudaMemsetAsync(status, 0, TDK_ST_WORDS * sizeof(int), stream));
    // the image pointers are stored, never followed: the range test fails first
    pair_setup_kernel<<<1, 32, 0, stream>>>(key_cam, nullptr, key_cam, T_wk, key_cam, nullptr, ref_cam, T_wr, 1, 1,
                                            (size_t)H * W, keys, pairs, status, nullptr, 0);
    UpdArgs a = {};
    a.keys = keys; a.pairs = pairs; a.status = status;
    a.vmin = 1. / (params[1] + TDK_EPS_INV);
    a.vmax = 1. / (params[0] + TDK_EPS_INV);
    a.geo = params[2]; a.photo = params[3]; a.step = params[4]; a.min_grad = params[5];
    a.min_grad_sq = sqrt_threshold(a.min_grad);
    a.P = 1; a.R = 1; a.H = (int)H; a.W = (int)W;
    estimate_outside_kernel<<<1, 1, 0, stream>>>(a, (long long)x, (long long)y, prior_depth, prior_variance, out_flag);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

namespace {
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

size_t tdk_update_depth_workspace_bytes(int64_t P, int64_t R, int64_t H, int64_t W)
{
    if (P <= 0 || R <= 0 || H <= 0 || W <= 0) return 0;
    return ws_layout(P, R, H, W).total;
}

size_t tdk_update_depth_workspace_bytes_min(int64_t P, int64_t R, int64_t H, int64_t W)
{
    if (P <= 0 || R <= 0 || H <= 0 || W <= 0) return 0;
    return ws_layout(P, R, H, W, 1).total;
}

// test hook: the threshold the sieve compares the squared gradient norm with (host arithmetic only)
double tdk_sqrt_threshold(double min_gradient) { return sqrt_threshold(min_gradient); }

// tuning knob (not part of the reference surface): how update_depth is mapped onto kernels
int tdk_set_kernel_plan(int plan)
{
    if (plan != PLAN_AUTO && plan != PLAN_TWO_KERNELS && plan != PLAN_ONE_KERNEL && plan != PLAN_SORTED) {
        tdk_set_error("tdk_set_kernel_plan: %d not in {-1,0,1,2}", plan);
        return TDK_ERR_INVALID;
    }
    g_plan.store(plan);
    return TDK_OK;
}

int tdk_set_subpixel_refinement(int on)
{
    g_subpixel.store(on ? 1 : 0);
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
                  dbg_argmin, out_hist, fuse_on_success, resolve_plan(g_plan.load(), P, H, W), workspace, workspace_bytes,
                  (cudaStream_t)stream);
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
                  dbg_argmin, out_hist, fuse_on_success, resolve_plan(g_plan.load(), P, H, W), workspace, workspace_bytes,
                  (cudaStream_t)stream);
}

int tdk_update_depth_tables_dev(const double *const *key_img_ptrs_dev, const double *key_cams,
                                const double *T_wk, const double *const *ref_img_ptrs_dev,
                                const double *ref_cams, const double *T_wr, const uint64_t *age,
                                const double *prior_depth, const double *prior_variance,
                                const double *params, int64_t P, int64_t R, int64_t H, int64_t W,
                                double *out_depth, double *out_variance, int64_t *out_flag,
                                int32_t *dbg_argmin, uint64_t *out_hist, int fuse_on_success,
                                void *workspace, size_t workspace_bytes, void *stream)
{
    if (!key_img_ptrs_dev || !ref_img_ptrs_dev) {
        tdk_set_error("tdk_update_depth_tables_dev: null pointer table");
        return TDK_ERR_INVALID;
    }
    return launch(nullptr, key_img_ptrs_dev, key_cams, T_wk, nullptr, ref_img_ptrs_dev, ref_cams, T_wr, age,
                  prior_depth, prior_variance, params, P, R, H, W, out_depth, out_variance, out_flag,
                  dbg_argmin, out_hist, fuse_on_success, resolve_plan(g_plan.load(), P, H, W), workspace, workspace_bytes,
                  (cudaStream_t)stream, true);
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
    if (P <= 0 || R <= 0 || P > TDK_STATIC_MAX_PAIRS || R > TDK_STATIC_MAX_PAIRS || P * R > TDK_STATIC_MAX_PAIRS ||
        ws_layout(P, R, 1, 1).total_tile > (size_t)TDK_STATIC_SCRATCH_BYTES) {
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
                fuse_on_success, PLAN_ONE_KERNEL, scratch, (size_t)TDK_STATIC_SCRATCH_BYTES, (cudaStream_t)stream);
    if (rc != TDK_OK || !wait_and_check) return rc;
    return tdk_update_depth_status(scratch, stream);
}

}  // extern "C"
