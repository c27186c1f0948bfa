// Measured L2 copy peak: a buffer that fits the L2 (default 32 MB read + 32 MB written per pass)
// copied over and over with 16-byte accesses; GB/s = (bytes read + bytes written) / time.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/micro/l2_peak tools/micro/l2_peak.cu
//   tools/micro/l2_peak [MB per buffer = 32] [passes = 200]
// For comparison the same kernel on buffers far larger than the L2 (HBM copy rate).
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

// `reps` sweeps over the buffer inside ONE launch, so that a 32 MB sweep (a few microseconds) is not
// timed through launch overhead
__global__ void copy16(const uint4 *__restrict__ src, uint4 *__restrict__ dst, size_t n, int reps)
{
    for (int r = 0; r < reps; r++)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
            dst[i] = __ldcg(src + i);   // ld.global.cg: L2 only, so that repeated sweeps do not hit in L1
}

__global__ void read16(const uint4 *__restrict__ src, size_t n, unsigned int *sink, int reps)
{
    unsigned int acc = 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int r = 0; r < reps; r++) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 7 * stride < n; i += 8 * stride) {   // eight independent 16-byte loads in flight per thread
        uint4 v[8];
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = __ldcg(src + i + k * stride);
#pragma unroll
        for (int k = 0; k < 8; k++) acc ^= v[k].x ^ v[k].y ^ v[k].z ^ v[k].w;
    }
    for (; i < n; i += stride) {
        const uint4 v = __ldcg(src + i);
        acc ^= v.x ^ v.y ^ v.z ^ v.w;
    }
    asm volatile("" ::: "memory");
    }
    if (acc == 0x12345678u) *sink = acc;
}

static double run(size_t bytes, int passes, bool read_only, int reps = 1)
{
    uint4 *a, *b;
    unsigned int *sink;
    cudaMalloc(&a, bytes); cudaMalloc(&b, bytes); cudaMalloc(&sink, 4);
    cudaMemset(a, 1, bytes); cudaMemset(b, 2, bytes);
    const size_t n = bytes / 16;
    const int blocks = 148 * 4, threads = 512;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < 5; i++) {
        if (read_only) read16<<<blocks, threads>>>(a, n, sink, reps); else copy16<<<blocks, threads>>>(a, b, n, reps);
    }
    cudaEventRecord(e0);
    for (int i = 0; i < passes; i++) {
        if (read_only) read16<<<blocks, threads>>>(a, n, sink, reps); else copy16<<<blocks, threads>>>(a, b, n, reps);
    }
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaFree(a); cudaFree(b); cudaFree(sink);
    return (read_only ? 1.0 : 2.0) * bytes * passes * reps / (ms * 1e-3) / 1e9;
}

int main(int argc, char **argv)
{
    const size_t mb = argc > 1 ? atoi(argv[1]) : 32;
    const int passes = argc > 2 ? atoi(argv[2]) : 200;
    printf("{\"l2_copy_gbs\": %.1f, \"l2_read_gbs\": %.1f, \"buffer_mb\": %zu, ", run(mb << 20, passes / 10, false, 20),
           run(mb << 20, passes / 10, true, 20), mb);
    printf("\"hbm_copy_gbs\": %.1f, \"hbm_read_gbs\": %.1f, \"hbm_buffer_mb\": 4096}\n", run((size_t)4096 << 20, 10, false),
           run((size_t)4096 << 20, 10, true));
    return 0;
}
