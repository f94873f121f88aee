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
#include "mrts_launch.h"

namespace mrts {
namespace {

constexpr int EXP_WARPS = 8, EXP_THREADS = 32 * EXP_WARPS;
constexpr int GROUP_BYTES = 624;  // 78 elements x (2 cells x 4 B, or 8 cells x 1 B)
constexpr int SLOT_BYTES = 640;   // one staged group, padded to a multiple of 128 bits x 8
constexpr int NSLOTS = 4;         // staged groups in flight per warp
constexpr int ZERO_GROUPS = 16, ZERO_BYTES = ZERO_GROUPS * GROUP_BYTES;  // 9,984 bytes of zeros per block
constexpr int OBS_TILE_BYTES = 4096;  // >= 128 x 31
constexpr int WBUF_WORDS = 136;       // the words of one tile (<= 128) + the one that follows
constexpr int WARP_BYTES = 2 * OBS_TILE_BYTES + NSLOTS * SLOT_BYTES + WBUF_WORDS * 4;
constexpr int SMEM_BYTES = ZERO_BYTES + EXP_WARPS * WARP_BYTES;
static_assert(WARP_BYTES % 16 == 0 && ZERO_BYTES % 16 == 0, "128-bit alignment of the staging buffers");
constexpr unsigned FULL = 0xFFFFFFFFu;

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
// shared -> global bulk copy of `bytes` (multiple of 16, both sides 16-byte aligned), tracked by the issuing thread's bulk group
__device__ __forceinline__ void bulk_store(void* gdst, const void* ssrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_addr(ssrc)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {  // all but the N most recent groups of this thread have read their source
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// generic-proxy writes to shared memory -> visible to the async proxy (the copy engine)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

template <int ELEM, int PLANES>
__global__ void __launch_bounds__(EXP_THREADS) mrts_expand_kernel(const __grid_constant__ ExpandParams X) {
    extern __shared__ uint4 smem_raw[];
    uint8_t* const sm = (uint8_t*)smem_raw;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < ZERO_BYTES / 16; i += EXP_THREADS) ((uint4*)sm)[i] = make_uint4(0, 0, 0, 0);
    fence_async_smem();
    __syncthreads();
    uint8_t* const wm = sm + ZERO_BYTES + warp * WARP_BYTES;
    uint8_t* const obuf = wm;
    uint8_t* const slots = wm + 2 * OBS_TILE_BYTES;
    uint32_t* const wbuf = (uint32_t*)(slots + NSLOTS * SLOT_BYTES);

    constexpr int EPC = 16 / ELEM;   // elements per 16-byte chunk
    constexpr int TC = 128 / ELEM;   // cells per observation tile: 128 x PLANES bytes
    constexpr int WPL = (TC + 1 + 31) / 32;  // tile words (+ the one that follows) per lane
    constexpr int GC = 8 / ELEM;     // cells per mask group
    const int HW = X.HW;
    const size_t obs_env = (size_t)HW * PLANES * ELEM, mask_env = (size_t)HW * 78 * ELEM;
    const size_t ment_stride = ment_stride_words(X.cap);
    unsigned tile_no = 0, slot_no = 0;  // running counters of this warp's staging buffers

    for (;;) {
        int v = 0;
        if (lane == 0) v = X.env_lo + atomicAdd(X.counter, 1);
        v = __shfl_sync(FULL, v, 0);
        if (v >= X.env_hi) break;

        // ---------------- observation: tiles of TC cells, double buffered (lane 0 owns the bulk groups)
        if (X.obs) {
            const uint32_t* words = X.pobs + (size_t)v * HW;
            uint8_t* const out = (uint8_t*)X.obs + (size_t)v * obs_env;
            uint32_t pre[WPL];
#pragma unroll
            for (int k = 0; k < WPL; k++) {
                const int idx = k * 32 + lane;
                pre[k] = (idx <= TC && idx < HW) ? __ldg(words + idx) : 0u;
            }
#pragma unroll 1
            for (int c0 = 0; c0 < HW; c0 += TC) {
                uint8_t* const buf = obuf + (tile_no & 1u) * OBS_TILE_BYTES;
                if (lane == 0) bulk_wait_read<1>();  // the copy that last read this buffer (two tiles ago) is done with it
                __syncwarp();
#pragma unroll
                for (int k = 0; k < WPL; k++) {
                    const int idx = k * 32 + lane;
                    if (idx <= TC) wbuf[idx] = pre[k];
                }
#pragma unroll
                for (int k = 0; k < WPL; k++) {  // the next tile's words travel while this one is expanded
                    const int idx = k * 32 + lane, c = c0 + TC + idx;
                    pre[k] = (idx <= TC && c < HW) ? __ldg(words + c) : 0u;
                }
                __syncwarp();
                const int cells = HW - c0 < TC ? HW - c0 : TC;
                const int nchunks = cells * PLANES * ELEM / 16;
#pragma unroll 2
                for (int j = lane; j < nchunks; j += 32) {
                    const int e0 = j * EPC;
                    const int q = e0 / PLANES;
                    const int o = e0 - q * PLANES;
                    const uint64_t lo = wbuf[q], hi = wbuf[q + 1];
                    const uint32_t bits = (uint32_t)(((hi << PLANES) | lo) >> o);
                    uint4 val;
                    if (ELEM == 4) {
                        val.x = bits & 1u;
                        val.y = (bits >> 1) & 1u;
                        val.z = (bits >> 2) & 1u;
                        val.w = (bits >> 3) & 1u;
                    } else {
                        val.x = spread4(bits & 0xF);
                        val.y = spread4((bits >> 4) & 0xF);
                        val.z = spread4((bits >> 8) & 0xF);
                        val.w = spread4((bits >> 12) & 0xF);
                    }
                    *(uint4*)(buf + (size_t)j * 16) = val;
                }
                fence_async_smem();
                __syncwarp();
                if (lane == 0) {
                    bulk_store(out + (size_t)c0 * PLANES * ELEM, buf, (uint32_t)nchunks * 16u);
                    bulk_commit();
                }
                tile_no++;
            }
        }

        // ---------------- mask: zero runs from the block's zero tile, rows of acting units staged per group (lane 1 owns the bulk groups)
        if (X.mask) {
            uint8_t* const out = (uint8_t*)X.mask + (size_t)v * mask_env;
            const uint32_t* ent = X.ment + (size_t)v * ment_stride;
            const int count = (int)__ldg(ent);
            const int ngroups = HW / GC;
            int cur_g = -1, run = 0;  // the open group; the first group not written yet
            uint8_t* slot = slots;
            auto zero_run = [&](int a, int b) {  // lane 1: groups [a, b) are all zero
                size_t off = (size_t)a * GROUP_BYTES;
                long rem = (long)(b - a) * GROUP_BYTES;
                while (rem > 0) {
                    const uint32_t n = rem < ZERO_BYTES ? (uint32_t)rem : (uint32_t)ZERO_BYTES;
                    bulk_store(out + off, sm, n);
                    off += n;
                    rem -= n;
                }
            };
            auto close_group = [&]() {
                fence_async_smem();
                __syncwarp();
                if (lane == 1) {
                    bulk_store(out + (size_t)cur_g * GROUP_BYTES, slot, GROUP_BYTES);
                    bulk_commit();
                }
            };
#pragma unroll 1
            for (int b0 = 0; b0 < count; b0 += 32) {
                uint4 e = make_uint4(0, 0, 0, 0);
                if (b0 + lane < count) e = __ldg((const uint4*)(ent + 4 * (size_t)(1 + b0 + lane)));
                const int nb = count - b0 < 32 ? count - b0 : 32;
#pragma unroll 1
                for (int i = 0; i < nb; i++) {
                    const uint32_t cell = __shfl_sync(FULL, e.x, i), w0 = __shfl_sync(FULL, e.y, i), w1 = __shfl_sync(FULL, e.z, i),
                                   w2 = __shfl_sync(FULL, e.w, i);
                    const int g = (int)(cell / GC);
                    if (g != cur_g) {
                        if (cur_g >= 0) close_group();
                        slot = slots + (slot_no % NSLOTS) * SLOT_BYTES;
                        slot_no++;
                        if (lane == 1) {
                            zero_run(run, g);
                            bulk_wait_read<NSLOTS - 1>();  // the copy that last read this slot (NSLOTS groups ago) is done with it
                        }
                        __syncwarp();
                        for (int j = lane; j < GROUP_BYTES / 16; j += 32) ((uint4*)slot)[j] = make_uint4(0, 0, 0, 0);
                        __syncwarp();
                        cur_g = g;
                        run = g + 1;
                    }
                    const int r0 = ((int)cell - g * GC) * 78;
                    for (int b = lane; b < 78; b += 32) {
                        const uint32_t w = b < 32 ? w0 : (b < 64 ? w1 : w2);
                        const uint32_t bit = (w >> (b & 31)) & 1u;
                        if (ELEM == 4) ((uint32_t*)slot)[r0 + b] = bit;
                        else slot[r0 + b] = (uint8_t)bit;
                    }
                }
            }
            if (cur_g >= 0) close_group();
            if (lane == 1) {
                zero_run(run, ngroups);
                bulk_commit();
            }
        }
    }
    // shared memory must outlive the copies that read it, and the writes must have landed before the grid ends
    if (lane < 2) bulk_wait_all();
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
