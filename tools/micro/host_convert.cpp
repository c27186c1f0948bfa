// Micro-benchmark for the host side of the narrow-wire host path: u64 -> u8 (age), i8 -> i64 (flags)
// and a plain memcpy, on T threads over 64 VGA maps.
//   g++ -O3 -march=native -pthread -o host_convert host_convert.cpp && ./host_convert
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main(int argc, char **argv)
{
    const size_t hw = 640 * 480, P = 64, n = hw * P;
    uint64_t *age = (uint64_t *)aligned_alloc(4096, n * 8);
    int64_t *flag = (int64_t *)aligned_alloc(4096, n * 8);
    uint8_t *a8 = (uint8_t *)aligned_alloc(4096, n);
    int8_t *f8 = (int8_t *)aligned_alloc(4096, n);
    double *src = (double *)aligned_alloc(4096, n * 8), *dst = (double *)aligned_alloc(4096, n * 8);
    for (size_t i = 0; i < n; i++) { age[i] = i % 9; f8[i] = -(int8_t)(i % 10); src[i] = (double)i; a8[i] = 0; flag[i] = 0; dst[i] = 0; }
    for (int T : {1, 2, 4, 8, 16}) {
        auto run = [&](const char *name, auto fn, double bytes) {
            double best = 1e9;
            for (int rep = 0; rep < 5; rep++) {
                const double t0 = now();
                std::vector<std::thread> th;
                for (int t = 0; t < T; t++) th.emplace_back([&, t] {
                    for (size_t p = t; p < P; p += T) fn(p);
                });
                for (auto &x : th) x.join();
                best = std::min(best, now() - t0);
            }
            printf("T=%2d %-12s %.2f ms  %.1f GB/s\n", T, name, best * 1e3, bytes / best / 1e9);
        };
        run("narrow age", [&](size_t p) { const uint64_t *s = age + p * hw; uint8_t *d = a8 + p * hw; for (size_t i = 0; i < hw; i++) d[i] = s[i] > 255 ? 255 : (uint8_t)s[i]; }, n * 9.);
        run("expand flag", [&](size_t p) { const int8_t *s = f8 + p * hw; int64_t *d = flag + p * hw; for (size_t i = 0; i < hw; i++) d[i] = s[i]; }, n * 9.);
        run("memcpy f64", [&](size_t p) { memcpy(dst + p * hw, src + p * hw, hw * 8); }, n * 16.);
    }
    return (int)(a8[5] + flag[7] + (int)dst[3]);
}
