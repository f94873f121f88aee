// Internal interface between the translation units of libmrts_b200.so.  The rules are compiled in
// three units (light games, heavy games, the full-capacity / debug kernels) and the expansion
// kernel in a fourth, so that `make -j` builds them side by side; mrts_api.cu holds the C ABI and
// the launch logic.
#pragma once
#include <cuda_runtime.h>

#include "mrts_step.h"

namespace mrts {

constexpr int GPB_MAX = 4;  // games (warps) per block of the warp kernels, chosen per environment

// ---- rules, one warp per game (mrts_rules_light.cu / mrts_rules_heavy.cu).  regs = 64 / 96 / 128: the
// register budget variant (__launch_bounds__), chosen per environment from the map size.
#define MRTS_RULES_API(K)                                                                                          \
    cudaError_t K##_launch_warp(int regs, const EnvParams& P, int gpb, int smem_per_game, cudaStream_t st);          \
    cudaError_t K##_launch_block(const EnvParams& P, int threads, int smem, cudaStream_t st);                         \
    cudaError_t K##_set_smem(int warp_bytes, int block_bytes);                                                        \
    int K##_warp_regs(int regs); /* registers per thread of the variant (cudaFuncGetAttributes), for introspection */
MRTS_RULES_API(light)
MRTS_RULES_API(heavy)
#undef MRTS_RULES_API
inline cudaError_t launch_rules_warp(int heavy, int regs, const EnvParams& P, int gpb, int smem_per_game, cudaStream_t st) {
    return heavy ? heavy_launch_warp(regs, P, gpb, smem_per_game, st) : light_launch_warp(regs, P, gpb, smem_per_game, st);
}
// ---- rules + direct outputs, one 256-thread block per game (large maps, direct mode)
inline cudaError_t launch_rules_block(int heavy, const EnvParams& P, int threads, int smem, cudaStream_t st) {
    return heavy ? heavy_launch_block(P, threads, smem, st) : light_launch_block(P, threads, smem, st);
}

// ---- full-capacity pass over a list of games (deferred games; map rotation resets), any game class
cudaError_t launch_overflow(const EnvParams& P, const int32_t* count, const int32_t* list, int grid, int smem, cudaStream_t st);
// ---- raw engine access (trace replay)
struct DebugArgs {
    int game;
    int op;  // 0 = issue_safe(rows), 1 = cycle
    const int32_t* rows7;
    int n;
    int* gameover;
};
cudaError_t launch_debug(const EnvParams& P, const DebugArgs& A, int smem, cudaStream_t st);
cudaError_t any_set_smem(int bytes);

// ---- expansion of the packed outputs into the caller's tensors (mrts_expand.cu)
struct ExpandParams {
    const uint32_t* pobs;  // (N, HW) observation words
    const uint32_t* ment;  // (N, 1 + cap, 4) mask entries
    void* obs;             // (N, H, W, planes) or null
    void* mask;            // (N, HW, 78) or null
    int32_t env_lo, env_hi;
    int32_t HW, planes, elem, cap;
    uint32_t empty_word;   // observation word of an empty cell (mrts_encode.h: obs_empty_word)
    int32_t* counter;      // work counter of this launch (zeroed by the caller): views are handed out one at a time
};
// true when the shape can take the bulk-store path (every view's obs and mask blocks are multiples of 16 bytes)
bool expand_supported(int HW, int planes, int elem, const void* obs, const void* mask);
cudaError_t launch_expand(const ExpandParams& X, int blocks, cudaStream_t st);
cudaError_t expand_init();
int expand_smem_bytes();
int expand_threads();

// ---- delta compaction of the packed outputs against a result-buffer slot's shadow (mrts_delta.cu)
constexpr int DELTA_ENTRY_WORDS = 6;  // {flat cell, obs word, mask words 0..2, span}: 24 bytes
struct DeltaParams {
    const uint32_t* pobs;   // (cells) observation words of this step
    const uint32_t* pmask;  // (cells, 3) mask words of this step
    uint32_t* shadow;       // (cells, 4) {obs word, mask words} last expanded into the slot's arrays; brought up to date
    uint32_t* count;        // number of cells that differ (may exceed cap: then the entries are incomplete)
    uint32_t* bitmap;       // ceil(cells / 32) words: bit = mask bit 0 of the cell
    uint32_t* entries;      // (cap, DELTA_ENTRY_WORDS)
    long long cells;
    uint32_t cap;
    int init;               // 1 = only fill the shadow and the bitmap (the host expands the whole arrays this time)
};
cudaError_t launch_delta(const DeltaParams& D, int n_sm, cudaStream_t st);
// (n, 2) {flat cell, packed row} -> pact[cell] (clear != 0: zeros instead)
cudaError_t launch_scatter_actions(const uint32_t* pairs, int n, uint16_t* pact, int clear, cudaStream_t st);

}  // namespace mrts
