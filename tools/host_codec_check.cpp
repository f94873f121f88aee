#include "../microrts_py_b200/csrc/mrts_host.h"
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>
int main() {
    using namespace mrts_host;
    for (int planes : {29, 31}) for (int hw : {64, 256, 100}) {
    size_t N = 2048;
    std::vector<uint32_t> w(N * hw), pm(N * hw * 3, 0);
    for (auto& v : w) v = (uint32_t)rand() * 2654435761u;
    for (size_t i = 0; i < pm.size(); i += 37) pm[i] = rand();
    Pool* p = pool_create(0);
    int32_t* out = (int32_t*)aligned_alloc(4096, N * hw * planes * 4 + 4096);
    int32_t* mk = (int32_t*)aligned_alloc(4096, N * hw * 78 * 4);
    for (int off : {0, 1}) {
        int32_t* o = out + off;  // off=1: unaligned path
        auto t0 = std::chrono::steady_clock::now();
        for (int r = 0; r < 5; r++) expand_obs(p, w.data(), N, hw, planes, 4, o);
        auto t1 = std::chrono::steady_clock::now();
        size_t bad = 0;
        for (size_t c = 0; c < N * hw; c++) for (int k = 0; k < planes; k++) bad += o[c * planes + k] != (int)((w[c] >> k) & 1);
        printf("planes %d hw %d off %d: obs bad %zu  %.1f GB/s (%d threads)\n", planes, hw, off, bad, 5.0 * N * hw * planes * 4 / std::chrono::duration<double>(t1 - t0).count() / 1e9, pool_threads(p));
    }
    auto t0 = std::chrono::steady_clock::now();
    for (int r = 0; r < 5; r++) expand_mask(p, pm.data(), N, hw, 4, mk);
    auto t1 = std::chrono::steady_clock::now();
    size_t bad = 0;
    for (size_t c = 0; c < N * hw; c++) for (int k = 0; k < 78; k++) bad += mk[c * 78 + k] != (int)((pm[3 * c + (k >> 5)] >> (k & 31)) & 1);
    printf("  mask bad %zu  %.1f GB/s\n", bad, 5.0 * N * hw * 78 * 4 / std::chrono::duration<double>(t1 - t0).count() / 1e9);
    pool_destroy(p); free(out); free(mk);
    }
}
