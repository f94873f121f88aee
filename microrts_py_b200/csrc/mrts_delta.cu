// Delta compaction of the packed step outputs for the host-buffer entry points (mrts_step_host_slot).
//
// The reference returns NumPy arrays (gym_microrts/envs/vec_env.py:1001-1061, 1091-1101): 110 KB per 16x16 env-step
// of 0/1 values of which a handful of cells differ from what the same result array held two steps earlier.  The
// rules pass leaves one observation word and three mask words per cell in HBM; this kernel compares them with the
// words last expanded into the caller's array pair (`shadow`, one per result-buffer slot, also in HBM) and emits
//   entries  {flat cell index, obs word, mask words 0..2, span} of the cells that differ, in any order (each cell once);
//            span = first | last << 5 changed observation plane, first << 10 | last << 17 changed mask entry
//            (first > last: that tensor's row is unchanged)
//   bitmap   one bit per cell: mask bit 0 ("a unit of the player may act here"), what the host needs to pick the
//            action rows it packs for the next step (vec_env.py:972-974)
// so that tens of KB cross PCIe instead of 16 bytes per cell, and the host rewrites only those cells.
// HBM-bound: 32 B read + (changed cells) x 40 B written per cell.
#include "mrts_launch.h"

namespace mrts {
namespace {

constexpr unsigned FULL = 0xFFFFFFFFu;
constexpr int DELTA_THREADS = 256;

__global__ void __launch_bounds__(DELTA_THREADS) mrts_delta_kernel(const DeltaParams D) {
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * (DELTA_THREADS / 32);
    const long long warp0 = (long long)blockIdx.x * (DELTA_THREADS / 32) + (threadIdx.x >> 5);
#pragma unroll 1
    for (long long base = warp0 * 32; base < D.cells; base += warps * 32) {
        const long long g = base + lane;
        const bool ok = g < D.cells;
        uint32_t w = 0, m0 = 0, m1 = 0, m2 = 0, span = 0;
        bool changed = false;
        if (ok) {
            w = __ldg(D.pobs + g);
            const uint32_t* pm = D.pmask + 3 * g;
            m0 = __ldg(pm), m1 = __ldg(pm + 1), m2 = __ldg(pm + 2);
            uint4* sh = (uint4*)D.shadow + g;
            if (!D.init) {
                const uint4 s = *sh;
                const uint32_t xo = s.x ^ w, x0 = s.y ^ m0, x1 = s.z ^ m1, x2 = s.w ^ m2;
                changed = (xo | x0 | x1 | x2) != 0;
                if (changed) {
                    // first / last element that differs, per tensor: the host rewrites only that range of the cell's row
                    // (a unit that starts or ends an action flips a few planes: one cache line instead of the row's two to six)
                    const uint32_t ko0 = xo ? (uint32_t)__ffs((int)xo) - 1u : 31u, ko1 = xo ? 31u - (uint32_t)__clz((int)xo) : 0u;
                    uint32_t kb0 = 127u, kb1 = 0u;
                    if (x0 | x1 | x2) {
                        kb0 = x0 ? (uint32_t)__ffs((int)x0) - 1u : (x1 ? 31u + (uint32_t)__ffs((int)x1) : 63u + (uint32_t)__ffs((int)x2));
                        kb1 = x2 ? 95u - (uint32_t)__clz((int)x2) : (x1 ? 63u - (uint32_t)__clz((int)x1) : 31u - (uint32_t)__clz((int)x0));
                    }
                    span = ko0 | (ko1 << 5) | (kb0 << 10) | (kb1 << 17);
                }
            }
            if (changed || D.init) *sh = make_uint4(w, m0, m1, m2);
        }
        const unsigned src = __ballot_sync(FULL, ok && (m0 & 1u));
        if (lane == 0) D.bitmap[base >> 5] = src;
        const unsigned chm = __ballot_sync(FULL, changed);
        if (chm) {
            unsigned b = 0;
            if (lane == 0) b = atomicAdd(D.count, (unsigned)__popc(chm));
            b = __shfl_sync(FULL, b, 0);
            if (changed) {
                const unsigned idx = b + __popc(chm & ((1u << lane) - 1u));
                if (idx < D.cap) {
                    uint32_t* e = D.entries + DELTA_ENTRY_WORDS * (size_t)idx;
                    e[0] = (uint32_t)g, e[1] = w, e[2] = m0, e[3] = m1, e[4] = m2, e[5] = span;
                }
            }
        }
    }
}

// packed action rows of the acting cells, as a list -> the dense (N, HW) array the rules kernels read; `value` < 0: write zeros
// (the cells of the previous step's list: a cell that no longer acts must read as NOOP, like a row the reference never forwards)
__global__ void __launch_bounds__(256) mrts_scatter_actions_kernel(const uint2* __restrict__ pairs, int n, uint16_t* __restrict__ pact, int clear) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const uint2 pr = __ldg(pairs + i);
    pact[pr.x] = clear ? (uint16_t)0 : (uint16_t)pr.y;
}

}  // namespace

cudaError_t launch_scatter_actions(const uint32_t* pairs, int n, uint16_t* pact, int clear, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    mrts_scatter_actions_kernel<<<(n + 255) / 256, 256, 0, st>>>((const uint2*)pairs, n, pact, clear);
    return cudaGetLastError();
}

cudaError_t launch_delta(const DeltaParams& D, int n_sm, cudaStream_t st) {
    long long blocks = (D.cells + DELTA_THREADS - 1) / DELTA_THREADS;
    if (blocks > (long long)n_sm * 8) blocks = (long long)n_sm * 8;
    if (blocks < 1) blocks = 1;
    mrts_delta_kernel<<<(int)blocks, DELTA_THREADS, 0, st>>>(D);
    return cudaGetLastError();
}

}  // namespace mrts
