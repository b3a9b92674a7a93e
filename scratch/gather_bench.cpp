// host-side row gather rate: T threads copy 4096 random 8 KB rows (of a 512 MB array) into a staging buffer
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
int main() {
    const size_t rows = 65536, rb = 8192, B = 4096;
    uint8_t *src = (uint8_t *)aligned_alloc(4096, rows * rb), *dst = (uint8_t *)aligned_alloc(4096, 4 * B * rb);
    memset(src, 1, rows * rb); memset(dst, 0, 4 * B * rb);
    std::vector<uint32_t> idx(B * 64);
    uint64_t s = 88172645463325252ull;
    for (auto &v : idx) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; v = (uint32_t)(s % rows); }
    printf("hardware_concurrency %u\n", std::thread::hardware_concurrency());
    for (int T : {1, 2, 4, 6, 8, 12, 16}) {
        const int steps = 40;
        auto t0 = std::chrono::steady_clock::now();
        std::vector<std::thread> th;
        for (int t = 0; t < T; t++)
            th.emplace_back([&, t] {
                for (int st = 0; st < steps; st++) {
                    const uint32_t *ix = idx.data() + (size_t)(st % 64) * B;
                    uint8_t *d = dst + (size_t)(st % 4) * B * rb;
                    for (size_t r = t; r < B; r += T) memcpy(d + r * rb, src + (size_t)ix[r] * rb, rb);
                }
            });
        for (auto &x : th) x.join();
        double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        printf("T=%2d  %.3f ms/step  %.1f GB/s\n", T, dt / steps * 1e3, steps * B * rb / dt / 1e9);
    }
    return 0;
}
