// Rules kernels that take any game class: the full-capacity pass over a list of games and the raw
// engine access used to replay recorded games.  Both are off the throughput path.
#include "mrts_rules.cuh"

namespace mrts {

// Games the warp kernels deferred because their unit table may outgrow the staged capacity
// (EnvParams::scap), and the games map rotation restarts: same pass, one block per game, shared
// memory sized for the full capacity.  P.ovf is null here (nothing can be deferred again);
// `list` = *count game indices.
__global__ void __launch_bounds__(128) mrts_overflow_kernel(const __grid_constant__ EnvParams P, const int32_t* count_p, const int32_t* list) {
    extern __shared__ uint4 smem_raw[];
    const int count = *count_p;
    for (int k = blockIdx.x; k < count; k += gridDim.x) {
        game_pass_any(P, list[k], (uint8_t*)smem_raw, (int)threadIdx.x, (int)blockDim.x, BlockSync());
        __syncthreads();
    }
}

// ---- raw engine access (trace replay): one block, one game.  GameState.issueSafe of a list of unit actions,
// or GameState.cycle, on the stored state (J/test/microrts/TestTracesIntegrity.java:72-126).
__global__ void mrts_debug_kernel(const __grid_constant__ EnvParams P, const DebugArgs A) {
    extern __shared__ uint4 smem_raw[];
    uint8_t* mem = (uint8_t*)smem_raw;
    const int tid = threadIdx.x, nth = blockDim.x;
    const Layout L = layout_of(P);
    Sim s;
    bind_sim(s, P, L, mem, A.game);
    s.partial = 0;
    s.nl = 32;
    s.lane = tid & 31;
    const MapTpl tpl = P.maps[P.game_map[A.game]];
    uint8_t* gblob = P.games + (size_t)A.game * P.game_stride;
    copy16<Q16>(mem + L.blob, gblob, 32 + P.cap * 32, tid, nth);
    copy16<uint32_t>(s.grid, P.grid_pool + tpl.grid_off, align16(P.HW * 2), tid, nth);
    __syncthreads();
    for (int i = tid; i < s.hdr->nunits; i += nth) s.grid[s.u[i].x + s.u[i].y * P.W] = (uint16_t)i;
    __syncthreads();
    if (tid == 0) {
        St z;
        z.npa = z.npa2 = z.astar_gen = z.gameover = z.reset = z.nprod = z.timing_on = z.rows_ready = z.nidle = z.explore = z.unobs_ready = 0;
        z.winner = -1;
        for (int k = 0; k < 12; k++) z.tp[k] = 0;
        for (int p = 0; p < 2; p++)
            for (int k = 0; k < 6; k++) z.rew[p][k] = 0;
        *s.st = z;
    }
    __syncthreads();
    if (tid < 32) {  // the rules run on one warp, every lane executing the same code (see Sim::lane)
        if (A.op == 0) {
            int cnt = 0;
            for (int k = 0; k < A.n && cnt < P.cap; k++) {
                const int32_t* r = A.rows7 + 7 * k;
                int i = s.unit_at(r[0], r[1]);
                if (i < 0) continue;
                PAct e;
                e.unit = (uint16_t)i;
                e.type = (int8_t)r[2];
                e.param = (int16_t)r[3];
                e.x = (int16_t)r[4];
                e.y = (int16_t)r[5];
                e.utype = (int8_t)r[6];
                e.pad = 0;
                s.pa[cnt++] = e;
            }
            s.issue_safe(s.pa, cnt);
        } else {
            s.cycle();
            s.eval_gameover();
            s.compact();
            if (A.gameover) *A.gameover = s.st->gameover;
        }
    }
    __syncthreads();
    copy16<Q16>(gblob, mem + L.blob, 32 + P.cap * 32, tid, nth);
}

cudaError_t launch_overflow(const EnvParams& P, const int32_t* count, const int32_t* list, int grid, int smem, cudaStream_t st) {
    mrts_overflow_kernel<<<grid, 128, smem, st>>>(P, count, list);
    return cudaGetLastError();
}
cudaError_t launch_debug(const EnvParams& P, const DebugArgs& A, int smem, cudaStream_t st) {
    mrts_debug_kernel<<<1, 64, smem, st>>>(P, A);
    return cudaGetLastError();
}
cudaError_t any_set_smem(int bytes) {
    if (bytes <= 48 * 1024) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(mrts_overflow_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(mrts_debug_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

}  // namespace mrts
