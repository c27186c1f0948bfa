// Micro-benchmark: scattered 2x2 footprints of f64 pixels through the LSU (__ldg) and through the
// texture unit (tex2D<int2> point fetches on a pitch-linear texture).  Development tool.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_bench gather_bench.cu && ./gather_bench
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

constexpr int W = 640, H = 480, NIMG = 64, SAMPLES = 12;

__device__ __forceinline__ unsigned hash(unsigned x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }

template <int MODE>
__global__ void __launch_bounds__(256, 4) gather(const double *imgs, const cudaTextureObject_t *tex, double *out, int coherent)
{
    const unsigned gid = blockIdx.x * blockDim.x + threadIdx.x;
    const int img = (gid >> 12) % NIMG;
    // a line start per thread: scattered over the image (coherent = 0) or next to its neighbour lane (1)
    unsigned h = hash(coherent ? (gid >> 5) : gid);
    int x0 = h % (W - 80), y0 = (h >> 12) % (H - 20);
    if (coherent) x0 += (gid & 31) % 8, y0 += (gid & 31) / 8;
    double acc = 0.;
    const double *base = imgs + (size_t)img * W * H;
    const cudaTextureObject_t t = MODE == 1 ? tex[img] : 0;
    for (int s = 0; s < SAMPLES; s++) {
        const int x = x0 + 5 * s, y = y0 + s;
        double a, b, c, d;
        if (MODE == 0) {
            a = __ldg(base + y * W + x); b = __ldg(base + y * W + x + 1);
            c = __ldg(base + (y + 1) * W + x); d = __ldg(base + (y + 1) * W + x + 1);
        } else if (MODE == 2) {   // 256-bit aligned loads: elements x&~3 .. +3 of each row, a second one if x%4 == 3
            const int xa = x & ~3, k = x & 3;
            double r[2][2];
#pragma unroll
            for (int row = 0; row < 2; row++) {
                const double *q = base + (y + row) * W + xa;
                double e0, e1, e2, e3;
                asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(e0), "=d"(e1), "=d"(e2), "=d"(e3) : "l"(q));
                double lo = k == 0 ? e0 : k == 1 ? e1 : k == 2 ? e2 : e3;
                double hi = k == 0 ? e1 : k == 1 ? e2 : e3;
                if (k == 3) hi = __ldg(q + 4);
                r[row][0] = lo; r[row][1] = hi;
            }
            a = r[0][0]; b = r[0][1]; c = r[1][0]; d = r[1][1];
        } else if (MODE == 3) {   // 128-bit aligned loads
            const int xa = x & ~1, k = x & 1;
            double r[2][2];
#pragma unroll
            for (int row = 0; row < 2; row++) {
                const double2 *q = reinterpret_cast<const double2 *>(base + (y + row) * W + xa);
                const double2 e = __ldg(q);
                double lo = k ? e.y : e.x, hi = e.y;
                if (k) hi = __ldg(reinterpret_cast<const double *>(q) + 2);
                r[row][0] = lo; r[row][1] = hi;
            }
            a = r[0][0]; b = r[0][1]; c = r[1][0]; d = r[1][1];
        } else {
            const int2 ta = tex2D<int2>(t, x + 0.5f, y + 0.5f), tb = tex2D<int2>(t, x + 1.5f, y + 0.5f);
            const int2 tc = tex2D<int2>(t, x + 0.5f, y + 1.5f), td = tex2D<int2>(t, x + 1.5f, y + 1.5f);
            a = __hiloint2double(ta.y, ta.x); b = __hiloint2double(tb.y, tb.x);
            c = __hiloint2double(tc.y, tc.x); d = __hiloint2double(td.y, td.x);
        }
        acc += a * 0.25 + b * 0.25 + c * 0.25 + d * 0.25;
    }
    out[gid] = acc;
}

int main()
{
    double *imgs, *out;
    const size_t n = (size_t)NIMG * W * H;
    CK(cudaMalloc(&imgs, n * 8));
    double *h = (double *)malloc(n * 8);
    for (size_t i = 0; i < n; i++) h[i] = (double)(i % 1000) * 0.001 + 1e-9 * (double)(i % 7);
    CK(cudaMemcpy(imgs, h, n * 8, cudaMemcpyHostToDevice));
    const int threads = 1 << 20;
    CK(cudaMalloc(&out, threads * 8));
    cudaTextureObject_t ht[NIMG], *dt;
    for (int i = 0; i < NIMG; i++) {
        cudaResourceDesc rd = {};
        rd.resType = cudaResourceTypePitch2D;
        rd.res.pitch2D.devPtr = imgs + (size_t)i * W * H;
        rd.res.pitch2D.desc = cudaCreateChannelDesc<int2>();
        rd.res.pitch2D.width = W; rd.res.pitch2D.height = H; rd.res.pitch2D.pitchInBytes = W * 8;
        cudaTextureDesc td = {};
        td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
        td.filterMode = cudaFilterModePoint; td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
        CK(cudaCreateTextureObject(&ht[i], &rd, &td, nullptr));
    }
    CK(cudaMalloc(&dt, sizeof(ht)));
    CK(cudaMemcpy(dt, ht, sizeof(ht), cudaMemcpyHostToDevice));
    double *r0 = (double *)malloc(threads * 8), *r1 = (double *)malloc(threads * 8);
    for (int coherent = 0; coherent < 2; coherent++) {
        for (int mode = 0; mode < 4; mode++) {
            cudaEvent_t e0, e1;
            cudaEventCreate(&e0); cudaEventCreate(&e1);
            for (int it = 0; it < 3; it++) {
                if (mode == 0) gather<0><<<threads / 256, 256>>>(imgs, dt, out, coherent);
                else if (mode == 1) gather<1><<<threads / 256, 256>>>(imgs, dt, out, coherent);
                else if (mode == 2) gather<2><<<threads / 256, 256>>>(imgs, dt, out, coherent);
                else gather<3><<<threads / 256, 256>>>(imgs, dt, out, coherent);
            }
            cudaEventRecord(e0);
            for (int it = 0; it < 10; it++) {
                if (mode == 0) gather<0><<<threads / 256, 256>>>(imgs, dt, out, coherent);
                else if (mode == 1) gather<1><<<threads / 256, 256>>>(imgs, dt, out, coherent);
                else if (mode == 2) gather<2><<<threads / 256, 256>>>(imgs, dt, out, coherent);
                else gather<3><<<threads / 256, 256>>>(imgs, dt, out, coherent);
            }
            cudaEventRecord(e1);
            CK(cudaDeviceSynchronize());
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            CK(cudaMemcpy(mode ? r1 : r0, out, threads * 8, cudaMemcpyDeviceToHost));
            const char *names[4] = {"__ldg x4     ", "tex2D<int2>  ", "ld.v4.f64 x2+", "ld.v2.f64 x2+"};
            size_t bad = 0;
            if (mode) for (int i = 0; i < threads; i++) bad += r0[i] != r1[i];
            printf("coherent=%d %s: %.1f us per launch, %.2f ns per 1000 samples, %zu mismatches\n", coherent, names[mode],
                   ms * 100., ms * 1e6 / 10. / ((double)threads * SAMPLES) * 1000., bad);
        }
    }
    return 0;
}
