// Rules kernels (sm_100a), included by mrts_rules_light.cu (HEAVY = 0) and mrts_rules_heavy.cu (HEAVY = 1).
//
// Execution model: one WARP per game.  The warp stages the game's state blob (header + unit table
// [+ scripted-bot table]) and the map's wall grid in shared memory, runs the lock-step rules
// (mrts_core.h: every lane executes the same code, per-unit scans are spread over the lanes), then
// writes the outputs (mrts_step.h) and puts the blob back.  There is no tensor-core work on this
// path (no dense contraction).  Light games (self-play, PassiveAI) and heavy games (scripted bots
// built on AbstractionLayerAI: table + path searches) are separate kernels with their own register
// budgets and instruction footprints; a slab holding both kinds launches both.
#pragma once
#include "mrts_launch.h"

namespace mrts {

struct BlockSync {
    __device__ __forceinline__ void operator()() const { __syncthreads(); }
};
struct WarpSync {
    __device__ __forceinline__ void operator()() const { __syncwarp(); }
};

// launch slot -> game (EnvParams::order)
__device__ __forceinline__ int slot_game(const EnvParams& P, int slot) {
    if (!P.order) return P.game0 + slot;
    int k = slot, c = 0;
    if (P.order_cnt)
        while (c < 3 && k >= P.order_cnt[c]) k -= P.order_cnt[c++];
    return P.order[(size_t)c * P.order_stride + k];
}

// MINB = resident 128-thread blocks per SM the register budget must allow: 8 -> 64 registers, 6 -> 80, 5 -> 96, 4 -> 128
template <int HEAVY, int MINB>
__global__ void __launch_bounds__(32 * GPB_MAX, MINB) mrts_rules_warp_kernel(const __grid_constant__ EnvParams P, const int smem_per_game) {
    extern __shared__ uint4 smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int slot = blockIdx.x * (blockDim.x >> 5) + warp;
    // direct mode: the block's tile of zeros of the bulk mask writer sits in front of the games' regions
    const int zb = P.bulk_out ? BULK_TILES_BYTES : 0;
    if (zb) bulk::zero_tile_init((uint8_t*)smem_raw, BULK_ZERO_BYTES, (int)threadIdx.x, (int)blockDim.x);
    if (slot >= P.nslots) return;
    game_pass<HEAVY>(P, slot_game(P, slot), (uint8_t*)smem_raw + zb + (size_t)warp * smem_per_game, lane, 32, WarpSync(), zb ? (uint8_t*)smem_raw : nullptr);
}

// one block per game: every thread takes part in the staging and the output phases (direct mode on large maps)
template <int HEAVY>
__global__ void __launch_bounds__(256) mrts_rules_block_kernel(const __grid_constant__ EnvParams P) {
    extern __shared__ uint4 smem_raw[];
    const int slot = blockIdx.x;
    if (slot >= P.nslots) return;
    game_pass<HEAVY>(P, slot_game(P, slot), (uint8_t*)smem_raw, (int)threadIdx.x, (int)blockDim.x, BlockSync());
}

template <int HEAVY>
static cudaError_t launch_warp_t(int regs, const EnvParams& P, int gpb, int smem_per_game, cudaStream_t st) {
    const dim3 grid((P.nslots + gpb - 1) / gpb);
    const int smem = gpb * smem_per_game + (P.bulk_out ? BULK_TILES_BYTES : 0);
    if (regs <= 64) mrts_rules_warp_kernel<HEAVY, 8><<<grid, 32 * gpb, smem, st>>>(P, smem_per_game);
    else if (regs <= 80) mrts_rules_warp_kernel<HEAVY, 6><<<grid, 32 * gpb, smem, st>>>(P, smem_per_game);
    else if (regs <= 96) mrts_rules_warp_kernel<HEAVY, 5><<<grid, 32 * gpb, smem, st>>>(P, smem_per_game);
    else mrts_rules_warp_kernel<HEAVY, 4><<<grid, 32 * gpb, smem, st>>>(P, smem_per_game);
    return cudaGetLastError();
}
template <int HEAVY>
static cudaError_t set_smem_t(int warp_bytes, int block_bytes) {
    cudaError_t e = cudaSuccess;
    if (warp_bytes > 48 * 1024) {
        if ((e = cudaFuncSetAttribute(mrts_rules_warp_kernel<HEAVY, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, warp_bytes)) != cudaSuccess) return e;
        if ((e = cudaFuncSetAttribute(mrts_rules_warp_kernel<HEAVY, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, warp_bytes)) != cudaSuccess) return e;
        if ((e = cudaFuncSetAttribute(mrts_rules_warp_kernel<HEAVY, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, warp_bytes)) != cudaSuccess) return e;
        if ((e = cudaFuncSetAttribute(mrts_rules_warp_kernel<HEAVY, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, warp_bytes)) != cudaSuccess) return e;
    }
    if (block_bytes > 48 * 1024) e = cudaFuncSetAttribute(mrts_rules_block_kernel<HEAVY>, cudaFuncAttributeMaxDynamicSharedMemorySize, block_bytes);
    return e;
}
template <int HEAVY>
static int warp_regs_t(int regs) {
    cudaFuncAttributes a;
    cudaError_t e;
    if (regs <= 64) e = cudaFuncGetAttributes(&a, mrts_rules_warp_kernel<HEAVY, 8>);
    else if (regs <= 80) e = cudaFuncGetAttributes(&a, mrts_rules_warp_kernel<HEAVY, 6>);
    else if (regs <= 96) e = cudaFuncGetAttributes(&a, mrts_rules_warp_kernel<HEAVY, 5>);
    else e = cudaFuncGetAttributes(&a, mrts_rules_warp_kernel<HEAVY, 4>);
    return e == cudaSuccess ? a.numRegs : -1;
}

}  // namespace mrts
