// Synthetic piecewise-planar scenes rendered on the device (SURVEY 8d workload, 8f rank 3).
// The CUDA counterpart of tadataka_b200/scene.py::World.render: an analytic ray-cast, one thread
// per pixel, no graphics API.  It exists so that the inputs of large configurations (C5: thousands
// of image pairs) are produced where they are consumed instead of being rendered by numpy and
// pushed through PCIe.  No counterpart in the reference, which reads its frames from datasets.
#include <math.h>

#include "tdk_common.cuh"

namespace {

constexpr int PLANE_WORDS = 14;   // p0[3], normal[3], a[3], b[3], half_a, half_b
constexpr int TEX_WORDS = 5;      // fa, fb, p1, p2, fn
constexpr int LATTICE = 64;

struct SceneArgs {
    const double *planes, *tex, *lattice;   // [K,14], [K,5], [K,64,64]
    const double *T_wc, *cams;              // [N,16], [N,4]
    double *image, *depth;                  // [N,H,W]
    double z0;
    int K, N, H, W;
};

// value noise on a 64x64 periodic lattice with smoothstep weights (scene.py::World._noise)
__device__ __forceinline__ double lattice_noise(const double *__restrict__ lat, double s, double t)
{
    const double s0 = floor(s), t0 = floor(t);
    double fs = s - s0, ft = t - t0;
    fs = fs * fs * (3. - 2. * fs);
    ft = ft * ft * (3. - 2. * ft);
    // numpy's % on int64 is a floor modulo
    long long i0 = (long long)s0 % LATTICE, j0 = (long long)t0 % LATTICE;
    if (i0 < 0) i0 += LATTICE;
    if (j0 < 0) j0 += LATTICE;
    const long long i1 = (i0 + 1) % LATTICE, j1 = (j0 + 1) % LATTICE;
    return lat[j0 * LATTICE + i0] * (1. - fs) * (1. - ft) + lat[j0 * LATTICE + i1] * fs * (1. - ft) +
           lat[j1 * LATTICE + i0] * (1. - fs) * ft + lat[j1 * LATTICE + i1] * fs * ft;
}

__device__ __forceinline__ double clamp01(double v) { return v < 0. ? 0. : (v > 1. ? 1. : v); }

__global__ void __launch_bounds__(256) render_scene_kernel(const SceneArgs a)
{
    const int x = blockIdx.x * 64 + (threadIdx.x & 63), y = blockIdx.y * 4 + (threadIdx.x >> 6);
    const int n = blockIdx.z;
    if (x >= a.W || y >= a.H) return;
    const double *__restrict__ T = a.T_wc + 16 * (size_t)n;
    const double *__restrict__ cam = a.cams + 4 * (size_t)n;
    const double X = ((double)x - cam[2]) / cam[0], Y = ((double)y - cam[3]) / cam[1];
    // ray direction in the world: R * (X, Y, 1); its parameter is the camera-frame depth
    const double dx = X * T[0] + Y * T[1] + T[2], dy = X * T[4] + Y * T[5] + T[6], dz = X * T[8] + Y * T[9] + T[10];
    const double ox = T[3], oy = T[7], oz = T[11];
    double best_t = INFINITY, value = 0.;
    for (int k = 0; k < a.K; k++) {
        const double *__restrict__ p = a.planes + PLANE_WORDS * k;
        const double denom = dx * p[3] + dy * p[4] + dz * p[5];
        const double t = ((p[0] - ox) * p[3] + (p[1] - oy) * p[4] + (p[2] - oz) * p[5]) / denom;
        const double rx = ox + dx * t - p[0], ry = oy + dy * t - p[1], rz = oz + dz * t - p[2];
        const double s = rx * p[6] + ry * p[7] + rz * p[8], tt = rx * p[9] + ry * p[10] + rz * p[11];
        const bool ok = t > 1e-9 && isfinite(t) && fabs(s) <= p[12] && fabs(tt) <= p[13] && t < best_t;
        if (!ok) continue;
        const double *__restrict__ tx = a.tex + TEX_WORDS * k;
        const double *__restrict__ lat = a.lattice + (size_t)k * LATTICE * LATTICE;
        const double fn = tx[4];
        const double mask = clamp01((lattice_noise(lat, s * fn / 12. + 7.3, tt * fn / 12. + 1.9) - 0.52) * 8.);
        const double fine = (lattice_noise(lat, s * fn, tt * fn) - 0.5) * 2.0;
        const double coarse = (lattice_noise(lat, s * fn / 4. + 3.1, tt * fn / 4. + 5.7) - 0.5) * 2.0;
        const double two_pi = 2. * 3.141592653589793;
        const double wave = sin(two_pi * tx[0] * s + tx[2]) * sin(two_pi * tx[1] * tt + tx[3]);
        value = clamp01(0.5 + mask * (0.35 * fine + 0.1 * wave) + 0.08 * coarse);
        best_t = t;
    }
    const size_t at = ((size_t)n * a.H + y) * a.W + x;
    a.image[at] = value;
    a.depth[at] = isfinite(best_t) ? best_t : a.z0;
}

}  // namespace

extern "C" int tdk_render_scene_dev(const double *planes, const double *tex, const double *lattice, int64_t K,
                                    double z0, const double *T_wc, const double *cams, int64_t N, int64_t H,
                                    int64_t W, double *image, double *depth, void *stream)
{
    if (!planes || !tex || !lattice || !T_wc || !cams || !image || !depth) {
        tdk_set_error("tdk_render_scene_dev: null pointer argument");
        return TDK_ERR_INVALID;
    }
    if (K <= 0 || K > 1024 || N < 0 || N > 65535 || H <= 0 || W <= 0 || H > 32767 || W > 32767) {
        tdk_set_error("tdk_render_scene_dev: bad sizes K=%lld N=%lld H=%lld W=%lld", (long long)K, (long long)N,
                      (long long)H, (long long)W);
        return TDK_ERR_INVALID;
    }
    if (N == 0) return TDK_OK;
    SceneArgs a{planes, tex, lattice, T_wc, cams, image, depth, z0, (int)K, (int)N, (int)H, (int)W};
    const dim3 grid((unsigned)((W + 63) / 64), (unsigned)((H + 3) / 4), (unsigned)N);
    render_scene_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(a);
    TDK_CUDA(cudaGetLastError());
    return TDK_OK;
}
