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

}  // extern "C"
