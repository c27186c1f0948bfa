// Vectorised geometry/image helpers on the drop-in surface (the reference's
// rust_bindings.{warp,interpolation,projection,transform,triangulation}); one thread per element.
#include <math.h>

#include "tdk_common.cuh"

namespace {

constexpr int TPB = 256;

// warp.rs:31-50
__global__ void warp_vecs_kernel(const double *__restrict__ T10, const double *__restrict__ xs0,
                                 const double *__restrict__ depths0, size_t n, double *__restrict__ xs1,
                                 double *__restrict__ depths1)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double T[12];
#pragma unroll
    for (int k = 0; k < 12; k++) T[k] = __ldg(T10 + k);
    double x1, y1, z1;
    tdk::warp_point(T, xs0[2 * i], xs0[2 * i + 1], depths0[i], x1, y1, z1);
    xs1[2 * i] = x1;
    xs1[2 * i + 1] = y1;
    depths1[i] = z1;
}

// interpolation.rs:57-73 ; the range test of tadataka/interpolation/__init__.py:23-29 is done
// here per coordinate (status = 1 and NaN output instead of a ValueError, raised by the caller)
__global__ void interpolate_kernel(const double *__restrict__ img, int H, int W,
                                   const double *__restrict__ coords, size_t n, double *__restrict__ out,
                                   int *__restrict__ status)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double cx = coords[2 * i], cy = coords[2 * i + 1];
    if (!tdk::in_range(cx, cy, (double)W - 1., (double)H - 1.)) {
        if (status) atomicExch(status, 1);
        out[i] = nan("");
        return;
    }
    out[i] = tdk::bilinear(img, W, H, cx, cy);
}

// projection.rs:11-14
__global__ void project_kernel(const double *__restrict__ ps, size_t n, double *__restrict__ xs)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double z = ps[3 * i + 2] + TDK_EPS;
    xs[2 * i] = ps[3 * i] / z;
    xs[2 * i + 1] = ps[3 * i + 1] / z;
}

// projection.rs:16-18
__global__ void inv_project_kernel(const double *__restrict__ xs, const double *__restrict__ depths,
                                   size_t n, double *__restrict__ ps)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double d = depths[i];
    ps[3 * i] = xs[2 * i] * d;
    ps[3 * i + 1] = xs[2 * i + 1] * d;
    ps[3 * i + 2] = 1. * d;
}

// transform.rs:9-28
__global__ void transform_kernel(const double *__restrict__ T10, const double *__restrict__ ps0, size_t n,
                                 double *__restrict__ ps1)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double p0 = ps0[3 * i], p1 = ps0[3 * i + 1], p2 = ps0[3 * i + 2];
#pragma unroll
    for (int r = 0; r < 3; r++)
        ps1[3 * i + r] = __ldg(T10 + 4 * r) * p0 + __ldg(T10 + 4 * r + 1) * p1 + __ldg(T10 + 4 * r + 2) * p2 +
                         __ldg(T10 + 4 * r + 3) * 1.;
}

// triangulation.rs:8-36
__global__ void calc_depth0_kernel(const double *__restrict__ T10, const double *__restrict__ xs0,
                                   const double *__restrict__ xs1, size_t n, double *__restrict__ depths0)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int ax = fabs(__ldg(T10 + 3)) > fabs(__ldg(T10 + 7)) ? 0 : 1;
    const double x = xs0[2 * i], y = xs0[2 * i + 1];
    const double x1i = xs1[2 * i + ax];
    const double ti = __ldg(T10 + 4 * ax + 3), tz = __ldg(T10 + 11);
    const double rzy = __ldg(T10 + 8) * x + __ldg(T10 + 9) * y + __ldg(T10 + 10) * 1.;
    const double riy = __ldg(T10 + 4 * ax) * x + __ldg(T10 + 4 * ax + 1) * y + __ldg(T10 + 4 * ax + 2) * 1.;
    depths0[i] = (ti - tz * x1i) / ((rzy * x1i - riy) + TDK_EPS);
}

unsigned blocks_for(int64_t n) { return (unsigned)((n + TPB - 1) / TPB); }

int check_n(const char *who, int64_t n, bool ptrs_ok)
{
    if (!ptrs_ok || n < 0 || n > 2147483647LL * TPB) {
        tdk_set_error("%s: bad argument", who);
        return TDK_ERR_INVALID;
    }
    return TDK_OK;
}


// ---- self-test of tdk::div2 / tdk::div5 against the plain `/` (tdk_common.cuh) -----------------
__device__ __forceinline__ unsigned long long mix64(unsigned long long x)
{
    x += 0x9e3779b97f4a7c15ULL;
    x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ULL;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebULL;
    return x ^ (x >> 31);
}

// operand classes: 0 = image-like magnitudes, 1 = any bit pattern (NaN, Inf, subnormals included),
// 2 = exponents next to the limits of the fast path
__device__ __forceinline__ double test_operand(unsigned long long h, int cls)
{
    const unsigned long long mant = h & 0x000fffffffffffffULL, sign = h & 0x8000000000000000ULL;
    unsigned long long ex;
    if (cls == 0) ex = 1023 - 40 + (h >> 52) % 80;
    else if (cls == 1) return __longlong_as_double((long long)h);
    else {
        const unsigned long long edges[8] = {0, 1, 2, 54, 0x360, 0x7f7, 0x7fd, 0x7fe};
        ex = edges[(h >> 52) & 7] + ((h >> 56) & 3);
        if (ex > 0x7fe) ex = 0x7fe;
    }
    return __longlong_as_double((long long)(sign | (ex << 52) | mant));
}

__global__ void shared_division_selftest_kernel(unsigned long long n, unsigned long long seed,
                                                unsigned long long *mismatches)
{
    unsigned long long bad = 0;
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned long long h = mix64(seed + i * 7);
        const int cls = (int)(h % 3), cls_b = (int)((h >> 8) % 3);
        double a[5];
        for (int j = 0; j < 5; j++) a[j] = test_operand(mix64(h + j + 1), cls);
        const double b = test_operand(mix64(h + 6), cls_b);
        double q0, q1;
        tdk::div2(a[0], a[1], b, q0, q1);
        double r0 = a[0], r1 = a[1], r2 = a[2], r3 = a[3], r4 = a[4];
        tdk::div5(r0, r1, r2, r3, r4, b);
        const double got[7] = {q0, q1, r0, r1, r2, r3, r4};
        const double want[7] = {a[0] / b, a[1] / b, a[0] / b, a[1] / b, a[2] / b, a[3] / b, a[4] / b};
        for (int j = 0; j < 7; j++) {
            const bool nan_g = got[j] != got[j], nan_w = want[j] != want[j];
            if (nan_g || nan_w) bad += nan_g != nan_w;   // NaN payloads are not compared
            else bad += __double_as_longlong(got[j]) != __double_as_longlong(want[j]);
        }
    }
    if (bad) atomicAdd(mismatches, bad);
}

}  // namespace

extern "C" {

int tdk_warp_vecs_dev(const double *T10, const double *xs0, const double *depths0, int64_t n, double *xs1,
                      double *depths1, void *stream)
{
    int rc = check_n("tdk_warp_vecs_dev", n, T10 && xs0 && depths0 && xs1 && depths1);
    if (rc != TDK_OK || n == 0) return rc;
    warp_vecs_kernel<<<blocks_for(n), TPB, 0, (cudaStream_t)stream>>>(T10, xs0, depths0, (size_t)n, xs1, depths1);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

int tdk_interpolate_dev(const double *image, int64_t H, int64_t W, const double *coords, int64_t n,
                        double *out, int32_t *status, void *stream)
{
    int rc = check_n("tdk_interpolate_dev", n, image && coords && out && H > 0 && W > 0 && H <= 2147483647LL && W <= 2147483647LL);
    if (rc != TDK_OK || n == 0) return rc;
    interpolate_kernel<<<blocks_for(n), TPB, 0, (cudaStream_t)stream>>>(image, (int)H, (int)W, coords, (size_t)n,
                                                                        out, status);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

int tdk_project_vecs_dev(const double *ps, int64_t n, double *xs, void *stream)
{
    int rc = check_n("tdk_project_vecs_dev", n, ps && xs);
    if (rc != TDK_OK || n == 0) return rc;
    project_kernel<<<blocks_for(n), TPB, 0, (cudaStream_t)stream>>>(ps, (size_t)n, xs);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

int tdk_inv_project_vecs_dev(const double *xs, const double *depths, int64_t n, double *ps, void *stream)
{
    int rc = check_n("tdk_inv_project_vecs_dev", n, xs && depths && ps);
    if (rc != TDK_OK || n == 0) return rc;
    inv_project_kernel<<<blocks_for(n), TPB, 0, (cudaStream_t)stream>>>(xs, depths, (size_t)n, ps);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

int tdk_transform_dev(const double *T10, const double *ps0, int64_t n, double *ps1, void *stream)
{
    int rc = check_n("tdk_transform_dev", n, T10 && ps0 && ps1);
    if (rc != TDK_OK || n == 0) return rc;
    transform_kernel<<<blocks_for(n), TPB, 0, (cudaStream_t)stream>>>(T10, ps0, (size_t)n, ps1);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

int tdk_calc_depth0_dev(const double *T10, const double *xs0, const double *xs1, int64_t n, double *depths0,
                        void *stream)
{
    int rc = check_n("tdk_calc_depth0_dev", n, T10 && xs0 && xs1 && depths0);
    if (rc != TDK_OK || n == 0) return rc;
    calc_depth0_kernel<<<blocks_for(n), TPB, 0, (cudaStream_t)stream>>>(T10, xs0, xs1, (size_t)n, depths0);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

int tdk_selftest_shared_division_dev(int64_t n, uint64_t seed, uint64_t *mismatches, void *stream)
{
    if (n < 0 || !mismatches) { tdk_set_error("tdk_selftest_shared_division_dev: bad arguments"); return TDK_ERR_INVALID; }
    TDK_CUDA(cudaMemsetAsync(mismatches, 0, sizeof(uint64_t), (cudaStream_t)stream));
    if (n == 0) return TDK_OK;
    shared_division_selftest_kernel<<<148 * 8, 256, 0, (cudaStream_t)stream>>>(
        (unsigned long long)n, (unsigned long long)seed, (unsigned long long *)mismatches);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}

}  // extern "C"
