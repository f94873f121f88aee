// Expansion of the packed step outputs into the caller's tensors: the HBM-bound half of a step.
//
//   observation words (N, HW) uint32, bit k = plane k      ->  obs  (N, H, W, P)  one element per plane
//   mask entries (N, 1 + cap) x {cell, 78 bits}            ->  mask (N, HW, 78)   > 99 % zeros
//
// in the reference's layout (GameState.getVectorObservation + MicroRTSGridModeVecEnv._encode_obs,
// J/rts/GameState.java:922-969, gym_microrts/envs/vec_env.py:311-321; JNIGridnetClient.getMasks,
// J/tests/JNIGridnetClient.java:203-216).  The rules kernels (mrts_rules.cuh) produce ~1.4 KB per
// view; this kernel turns it into the 27 KB - 1.75 MB per view the reference's tensors take.
//
// Blackwell store path: nothing is written with per-thread global stores.  A warp expands a tile of
// the observation into shared memory and hands it to the bulk-copy engine
// (cp.async.bulk.global.shared::cta, UBLKCP in SASS); the mask is written from ONE static tile of zeros
// per block -- no expansion work at all for the all-zero runs -- and the few rows that hold a unit are
// staged in 624-byte groups (the smallest 16-byte-aligned run of whole 78-element rows) and bulk-stored
// in between, so every output byte is written exactly once.  The copies are asynchronous: while the
// engine drains a tile the warp expands the next one, and the load/store pipe stays free for the
// rules warps of the next slab chunk that share the SM.  Views are handed out one at a time by an
// atomic counter (persistent blocks, one per SM).
#include "mrts_bulk.cuh"
#include "mrts_launch.h"

namespace mrts {
namespace {

constexpr int EXP_WARPS = 8, EXP_THREADS = 32 * EXP_WARPS;
using bulk::GROUP_BYTES;
constexpr int ZERO_BYTES = 16 * GROUP_BYTES;  // 9,984 bytes of zeros per block
constexpr int SEG_CELLS = 1024;    // words of a view staged at a time
constexpr int TILES_BYTES = ZERO_BYTES + bulk::EMPTY_BYTES_MAX;
constexpr int WARP_BYTES = SEG_CELLS * 4 + bulk::STAGE_BYTES;
constexpr int SMEM_BYTES = TILES_BYTES + EXP_WARPS * WARP_BYTES;
static_assert(WARP_BYTES % 16 == 0 && TILES_BYTES % 16 == 0, "128-bit alignment of the staging buffers");
constexpr unsigned FULL = 0xFFFFFFFFu;

template <int ELEM, int PLANES>
__global__ void __launch_bounds__(EXP_THREADS) mrts_expand_kernel(const __grid_constant__ ExpandParams X) {
    extern __shared__ uint4 smem_raw[];
    uint8_t* const sm = (uint8_t*)smem_raw;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    bulk::tiles_init(sm, ZERO_BYTES, ELEM, PLANES, X.empty_word, threadIdx.x, EXP_THREADS);
    uint8_t* const wm = sm + TILES_BYTES + warp * WARP_BYTES;
    uint32_t* const wseg = (uint32_t*)wm;
    const int HW = X.HW;
    const size_t obs_env = (size_t)HW * PLANES * ELEM, mask_env = (size_t)HW * 78 * ELEM;
    const size_t ment_stride = ment_stride_words(X.cap);
    bulk::Stage stage;
    stage.init(wm + SEG_CELLS * 4, lane);
    bulk::ObsWriter ow;
    ow.init(&stage, sm + ZERO_BYTES, X.empty_word, ELEM, PLANES);
    bulk::MaskWriter mw;
    mw.init(&stage, sm, ZERO_BYTES, ELEM);

    for (;;) {
        int v = 0;
        if (lane == 0) v = X.env_lo + atomicAdd(X.counter, 1);
        v = __shfl_sync(FULL, v, 0);
        if (v >= X.env_hi) break;

        // ---------------- observation: runs of empty groups from the block's tile, the rest staged (lane 0 owns the bulk groups)
        if (X.obs) {
            const uint32_t* words = X.pobs + (size_t)v * HW;
            ow.begin((uint8_t*)X.obs + (size_t)v * obs_env);
#pragma unroll 1
            for (int c0 = 0; c0 < HW; c0 += SEG_CELLS) {
                const int cells = HW - c0 < SEG_CELLS ? HW - c0 : SEG_CELLS;
                __syncwarp();  // the previous segment's words are no longer read
                for (int k = lane * 4; k < cells; k += 128) *(uint4*)(wseg + k) = __ldg((const uint4*)(words + c0 + k));
                __syncwarp();
                ow.segment(wseg, c0 / (16 / ELEM), cells / (16 / ELEM));
            }
            ow.finish(HW / (16 / ELEM));
        }

        // ---------------- mask: zero runs from the block's zero tile, rows of acting units staged per group (lane 1 owns the bulk groups)
        if (X.mask) {
            const uint32_t* ent = X.ment + (size_t)v * ment_stride;
            const int count = (int)__ldg(ent);
            mw.begin((uint8_t*)X.mask + (size_t)v * mask_env);
#pragma unroll 1
            for (int b0 = 0; b0 < count; b0 += 32) {
                uint4 e = make_uint4(0, 0, 0, 0);
                if (b0 + lane < count) e = __ldg((const uint4*)(ent + 4 * (size_t)(1 + b0 + lane)));
                const int nb = count - b0 < 32 ? count - b0 : 32;
#pragma unroll 1
                for (int i = 0; i < nb; i++)
                    mw.add(__shfl_sync(FULL, e.x, i), __shfl_sync(FULL, e.y, i), __shfl_sync(FULL, e.z, i), __shfl_sync(FULL, e.w, i));
            }
            mw.finish(HW / (8 / ELEM));
        }
    }
    // shared memory must outlive the copies that read it
    stage.drain();
}

}  // namespace

bool expand_supported(int HW, int planes, int elem, const void* obs, const void* mask) {
    if (planes != 29 && planes != 31) return false;
    if (elem != 4 && elem != 1) return false;
    if ((HW * elem) % 16 != 0) return false;  // every view's obs (HW x P x elem) and mask (HW x 78 x elem) block starts on a 16-byte boundary
    if ((((uintptr_t)obs) | ((uintptr_t)mask)) & 15) return false;
    return true;
}
int expand_smem_bytes() { return SMEM_BYTES; }
int expand_threads() { return EXP_THREADS; }

cudaError_t expand_init() {
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(mrts_expand_kernel<4, 29>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(mrts_expand_kernel<4, 31>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(mrts_expand_kernel<1, 29>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES)) != cudaSuccess) return e;
    return cudaFuncSetAttribute(mrts_expand_kernel<1, 31>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
}

cudaError_t launch_expand(const ExpandParams& X, int blocks, cudaStream_t st) {
    if (X.elem == 4 && X.planes == 29) mrts_expand_kernel<4, 29><<<blocks, EXP_THREADS, SMEM_BYTES, st>>>(X);
    else if (X.elem == 4) mrts_expand_kernel<4, 31><<<blocks, EXP_THREADS, SMEM_BYTES, st>>>(X);
    else if (X.planes == 29) mrts_expand_kernel<1, 29><<<blocks, EXP_THREADS, SMEM_BYTES, st>>>(X);
    else mrts_expand_kernel<1, 31><<<blocks, EXP_THREADS, SMEM_BYTES, st>>>(X);
    return cudaGetLastError();
}

}  // namespace mrts
