// CUDA kernels (sm_100a) and the C ABI of include/mrts_b200.h.
//
// Execution model: one thread block per game.  The block stages the game's state blob
// (header + unit table [+ scripted-bot table]) and the map's wall grid in shared memory, thread 0
// runs the lock-step rules (mrts_core.h), then all threads encode and stream out the
// observation and mask tensors with 128-bit stores (mrts_encode.h) and write the blob back.
// There is no tensor-core work on this path (no dense contraction); the bound is HBM write
// bandwidth of the obs + mask tensors.  No CPU fallback exists: every entry point launches a kernel.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/mrts_b200.h"
#include "mrts_host.h"
#include "mrts_setup.h"

using namespace mrts;

namespace {

thread_local std::string g_err;
int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}
#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return fail(MRTS_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

struct BlockSync {
    __device__ __forceinline__ void operator()() const { __syncthreads(); }
};

template <int ELEM, int PLANES>
__global__ void mrts_pass_kernel(const EnvParams P) {
    extern __shared__ uint4 smem_raw[];
    const int g = blockIdx.x;
    if (g >= P.n_games) return;
    game_pass<ELEM, PLANES>(P, g, (uint8_t*)smem_raw, (int)threadIdx.x, (int)blockDim.x, BlockSync());
}

// One warp per game, GPB games per block: the group barrier is __syncwarp(), and a block's
// shared memory is GPB independent per-game regions.  Small maps use this variant: it doubles
// (vs 64-thread blocks) the number of games whose serial rules phase is in flight per SM.
struct WarpSync {
    __device__ __forceinline__ void operator()() const { __syncwarp(); }
};
constexpr int GPB_MAX = 4;  // games per block is chosen per environment (mrts_env::gpb), see choose_gpb
#ifndef MRTS_WARP_MIN_BLOCKS
#define MRTS_WARP_MIN_BLOCKS 5
#endif
// MINB = resident 128-thread blocks per SM the register budget must allow: 5 (96 registers) is
// the best trade for 16x16 and larger maps; 8x8 maps (small state, many games per SM) run faster
// with 8 (64 registers, 32+ games per SM).
template <int ELEM, int PLANES, int MINB>
__global__ void __launch_bounds__(32 * GPB_MAX, MINB) mrts_pass_warp_kernel(const EnvParams P, const int smem_per_game) {
    extern __shared__ uint4 smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int slot = blockIdx.x * (blockDim.x >> 5) + warp;
    if (slot >= P.n_games) return;
    int g = slot;
    if (P.order) {  // longest games first: blocks are dispatched in index order
        int k = slot, c = 0;
        while (c < 3 && k >= P.order_n[c]) k -= P.order_n[c++];
        g = P.order[c * P.n_games + k];
    }
    game_pass<ELEM, PLANES>(P, g, (uint8_t*)smem_raw + (size_t)warp * smem_per_game, lane, 32, WarpSync());
}

// Launch order of a slab with scripted-bot games.  A bot game's step time has a heavy tail: one path
// search per idle bot unit, several in one frame every ~10 frames.  The number of idle bot units is
// known from the previous pass, so games are binned by it and launched slowest class first;
// otherwise the launch ends with a few long games running alone.
constexpr uint32_t ORDER_T0 = 4, ORDER_T1 = 2, ORDER_T2 = 0;  // idle bot units: >= 5, 3-4, 1-2, none
__global__ void mrts_order_kernel(const uint32_t* cost, int n, int32_t* order, int32_t* order_n) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    const uint32_t c = cost[g];
    const int cls = c > ORDER_T0 ? 0 : (c > ORDER_T1 ? 1 : (c > ORDER_T2 ? 2 : 3));
    order[cls * n + atomicAdd(&order_n[cls], 1)] = g;
}

// Games the warp kernel deferred because their unit table may outgrow the staged capacity
// (EnvParams::scap): same pass, one block per game, shared memory sized for the full capacity.
// P.ovf is null here (nothing can be deferred again); `list` = [count, game indices...].
template <int ELEM, int PLANES>
__global__ void __launch_bounds__(128) mrts_overflow_kernel(const EnvParams P, const int32_t* list) {
    extern __shared__ uint4 smem_raw[];
    const int count = list[0];
    for (int k = blockIdx.x; k < count; k += gridDim.x) {
        game_pass<ELEM, PLANES>(P, list[1 + k], (uint8_t*)smem_raw, (int)threadIdx.x, (int)blockDim.x, BlockSync());
        __syncthreads();
    }
}

// cycle_maps (gym_microrts/envs/vec_env.py:132-133, 1038-1056): after a step, every client whose episode
// ended gets `mapPath = next(cycle)` and is reset again, in ascending env order.  One block: each thread
// owns a contiguous run of games (order kept), counts its ended games, a block scan gives each its rank,
// the k-th ended game takes rotation entry (pos + k) mod n_cyc.  list = [count, games...] feeds the reset launch.
constexpr int CYCLE_THREADS = 1024;
__global__ void __launch_bounds__(CYCLE_THREADS) mrts_cycle_kernel(const uint8_t* ended, int n_games, const int32_t* cyc, int n_cyc, int32_t* pos,
                                                                   int32_t* game_map, int32_t* list) {
    __shared__ int part[CYCLE_THREADS / 32];
    const int t = threadIdx.x, per = (n_games + CYCLE_THREADS - 1) / CYCLE_THREADS;
    const int g0 = t * per, g1 = min(n_games, g0 + per);
    int mine = 0;
    for (int g = g0; g < g1; g++) mine += ended[g] != 0;
    int incl = mine;  // inclusive scan: warp shuffle, then the warp totals
    for (int d = 1; d < 32; d <<= 1) {
        const int v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if ((t & 31) >= d) incl += v;
    }
    if ((t & 31) == 31) part[t >> 5] = incl;
    __syncthreads();
    int before = incl - mine, total = 0;
    for (int w = 0; w < CYCLE_THREADS / 32; w++) {
        if (w < (t >> 5)) before += part[w];
        total += part[w];
    }
    const int p0 = *pos;
    for (int g = g0; g < g1; g++)
        if (ended[g]) {
            game_map[g] = cyc[(p0 + before) % n_cyc];
            list[1 + before] = g;
            before++;
        }
    __syncthreads();
    if (t == 0) {
        list[0] = total;
        *pos = (p0 + total) % n_cyc;
    }
}

// ---- raw engine access (trace replay): one block, one game
struct DebugArgs {
    int game;
    int op;  // 0 = issue_safe(rows), 1 = cycle
    const int32_t* rows7;
    int n;
    int* gameover;
};
__global__ void mrts_debug_kernel(const EnvParams P, const DebugArgs A) {
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
        *(Sim*)(mem + L.sim) = s;
        St z;
        z.npa = z.npa2 = z.astar_gen = z.gameover = z.reset = z.nprod = z.timing_on = z.rows_ready = z.nidle = z.explore = z.unobs_ready = 0;
        z.winner = -1;
        for (int k = 0; k < 8; k++) z.tp[k] = 0;
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
                // a re-issued unit replaces its assignment (LinkedHashMap.put keeps the slot): drop the old one first
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

}  // namespace

struct mrts_env {
    EnvParams P;
    int device = 0;
    int threads = 64;
    int gpb = 1;  // warp kernel: games (warps) per block
    int minb = MRTS_WARP_MIN_BLOCKS;  // warp kernel: register budget variant
    int smem = 0;       // bytes per game at the staged capacity (P.scap)
    int smem_full = 0;  // bytes per game at the full capacity (overflow / debug launches)
    int32_t* d_ovf = nullptr;
    long long* d_timing = nullptr;  // profiling aid (mrts_debug_timing)
    // ---- host-buffer entry points (mrts_reset_host / mrts_step_host): packed transport, see mrts_host.h
    struct HostIO {
        cudaStream_t stream = nullptr;
        cudaEvent_t ev_obs = nullptr;
        uint16_t *h_pact = nullptr, *d_pact = nullptr;
        uint32_t *h_pobs = nullptr, *d_pobs = nullptr, *h_pmask = nullptr, *d_pmask = nullptr;
        double *h_rew = nullptr, *d_rew = nullptr;
        uint8_t *h_done = nullptr, *d_done = nullptr;
        mrts_host::Pool* pool = nullptr;
        bool ready = false, have_mask = false;
        // result-array pairs the caller hands back unmodified (mrts_*_host_slot): the words last expanded into them
        struct Slot {
            void *obs = nullptr, *mask = nullptr;
            std::vector<uint32_t> pobs, pmask;
        } slots[MRTS_HOST_SLOTS];
    } hio;
    const uint16_t* cur_pact = nullptr;  // per-launch packed pointers (set by the host entry points)
    uint32_t *cur_pobs = nullptr, *cur_pmask = nullptr;
    uint32_t* d_cost = nullptr;     // bot slabs: last step cycles per game
    int32_t* d_order = nullptr;     // 4 * n_games launch order + 4 class counts
    // cycle_maps (mrts_set_cycle_maps): rotation list, its position, per-game ended flags, list of games to restart
    int n_cycle = 0;
    int32_t *d_cycle = nullptr, *d_cycle_pos = nullptr, *d_cycle_list = nullptr;
    uint8_t* d_ended = nullptr;
    int N = 0;
    int n_maps = 0;
    bool all_scripted = false;  // every game is bot vs bot (mrts_config::bot_ai1): mrts_step takes no actions
    std::vector<void*> allocs;
    int* d_flag = nullptr;
    int32_t* d_rows = nullptr;
    int rows_cap = 0;
    uint32_t sample_counter = 0;
};

namespace {

int choose_threads(int HW) {  // 32 = one warp per game (mrts_pass_warp_kernel), otherwise one block per game
    if (HW <= 1024) return 32;
    return 256;
}

template <class T>
int upload(mrts_env* e, const std::vector<T>& h, const T** out) {
    void* d = nullptr;
    size_t bytes = (h.size() ? h.size() : 1) * sizeof(T);
    CK(cudaMalloc(&d, bytes));
    e->allocs.push_back(d);
    if (!h.empty()) CK(cudaMemcpy(d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
    *out = (const T*)d;
    return MRTS_OK;
}

int launch(mrts_env* e, int mode, const int32_t* actions, void* obs, void* mask, double* reward, uint8_t* done, void* stream,
           int32_t* next_actions = nullptr, uint64_t seed = 0) {
    if (!e) return fail(MRTS_EINVAL, "null env");
    CK(cudaSetDevice(e->device));
    EnvParams P = e->P;
    P.mode = mode;
    P.actions = actions;
    P.obs = obs;
    P.mask = mask;
    P.reward = reward;
    P.done = done;
    P.next_actions = next_actions;
    P.seed = seed;
    P.counter = next_actions ? e->sample_counter++ : 0;
    P.pact = e->cur_pact;
    P.pobs = e->cur_pobs;
    P.pmask = e->cur_pmask;
    e->cur_pact = nullptr;
    e->cur_pobs = e->cur_pmask = nullptr;
    P.timing = e->d_timing;
    const bool rotate = e->n_cycle > 0 && mode == MODE_STEP;
    P.ended = rotate ? e->d_ended : nullptr;
    const int GPB = e->gpb;
    const dim3 wgrid((P.n_games + GPB - 1) / GPB), bgrid(P.n_games);
    const cudaStream_t st = (cudaStream_t)stream;
    const bool defer = e->threads == 32 && P.scap < P.cap;
    EnvParams F = P;  // full-capacity parameters of the overflow launch
    F.scap = P.cap;
    F.ovf = nullptr;
    P.cost = e->d_cost;
    P.order = nullptr;
    P.order_n = nullptr;
    if (e->d_order && e->threads == 32 && mode == MODE_STEP) {
        int32_t* cnt = e->d_order + 4 * (size_t)P.n_games;
        CK(cudaMemsetAsync(cnt, 0, 4 * sizeof(int32_t), st));
        mrts_order_kernel<<<(P.n_games + 255) / 256, 256, 0, st>>>(e->d_cost, P.n_games, e->d_order, cnt);
        P.order = e->d_order;
        P.order_n = cnt;
    }
    F.cost = P.cost;
    if (defer) {
        P.ovf = e->d_ovf;
        CK(cudaMemsetAsync(e->d_ovf, 0, sizeof(int32_t), st));
    } else {
        P.ovf = nullptr;
    }
    const int ogrid = P.n_games < 296 ? P.n_games : 296;
    // map rotation: the games that ended restart once more from their new template; only their obs / mask /
    // sampled-action rows are rewritten (reward and done keep the step's values, vec_env.py:1057)
    EnvParams R = F;
    R.mode = MODE_RESET;
    R.actions = nullptr;
    R.pact = nullptr;
    R.reward = nullptr;
    R.done = nullptr;
    R.ended = nullptr;
    R.cost = nullptr;
#define MRTS_LAUNCH(E, PL)                                                                       \
    do {                                                                                         \
        if (e->threads == 32 && e->minb == 8) mrts_pass_warp_kernel<E, PL, 8><<<wgrid, 32 * GPB, GPB * e->smem, st>>>(P, e->smem); \
        else if (e->threads == 32) mrts_pass_warp_kernel<E, PL, MRTS_WARP_MIN_BLOCKS><<<wgrid, 32 * GPB, GPB * e->smem, st>>>(P, e->smem); \
        else mrts_pass_kernel<E, PL><<<bgrid, e->threads, e->smem, st>>>(P);                     \
        if (defer) mrts_overflow_kernel<E, PL><<<ogrid, 128, e->smem_full, st>>>(F, e->d_ovf);   \
        if (rotate) {                                                                            \
            mrts_cycle_kernel<<<1, CYCLE_THREADS, 0, st>>>(e->d_ended, P.n_games, e->d_cycle, e->n_cycle, e->d_cycle_pos, \
                                                           const_cast<int32_t*>(P.game_map), e->d_cycle_list); \
            mrts_overflow_kernel<E, PL><<<ogrid, 128, e->smem_full, st>>>(R, e->d_cycle_list);   \
        }                                                                                        \
    } while (0)
    if (P.elem == 4 && P.planes == 29) MRTS_LAUNCH(4, 29);
    else if (P.elem == 4) MRTS_LAUNCH(4, 31);
    else if (P.planes == 29) MRTS_LAUNCH(1, 29);
    else MRTS_LAUNCH(1, 31);
#undef MRTS_LAUNCH
    CK(cudaGetLastError());
    return MRTS_OK;
}

}  // namespace

extern "C" {

const char* mrts_last_error(void) { return g_err.c_str(); }
const char* mrts_version(void) { return "mrts-b200 0.1 (sm_100a)"; }

int mrts_create(const mrts_config* cfg, mrts_env** out) {
    if (!cfg || !out) return fail(MRTS_EINVAL, "null argument");
    *out = nullptr;
    HostSetup S;
    std::string err;
    int rc = build_setup(cfg, S, err);
    if (rc) return fail(rc, err);
    mrts_env* e = new mrts_env();
    e->device = cfg->device;
    e->N = S.N;
    e->n_maps = cfg->n_maps;
    e->all_scripted = S.N > 0;
    for (int8_t a1 : S.game_ai1) e->all_scripted = e->all_scripted && a1 != AI_NONE;
    e->P = S.P;
    EnvParams& P = e->P;
    const int W = P.W, HW = P.HW, cap = P.cap;
    cudaError_t ce = cudaSetDevice(cfg->device);
    if (ce != cudaSuccess) {
        delete e;
        return fail(MRTS_ECUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(ce));
    }
    if ((rc = upload(e, S.tpls, &P.maps)) || (rc = upload(e, S.grids, &P.grid_pool)) || (rc = upload(e, S.units, &P.unit_pool)) ||
        (rc = upload(e, S.game_map, &P.game_map)) || (rc = upload(e, S.game_ai, &P.game_ai)) ||
        (rc = upload(e, S.game_ai1, &P.game_ai1))) {
        mrts_destroy(e);
        return rc;
    }
    void* d = nullptr;
    ce = cudaMalloc(&d, (size_t)P.n_games * P.game_stride);
    if (ce != cudaSuccess) {
        mrts_destroy(e);
        return fail(MRTS_ENOMEM, std::string("cudaMalloc(games): ") + cudaGetErrorString(ce));
    }
    e->allocs.push_back(d);
    P.games = (uint8_t*)d;
    cudaMemset(d, 0, (size_t)P.n_games * P.game_stride);
    ce = cudaMalloc(&d, sizeof(int));
    if (ce == cudaSuccess) {
        e->allocs.push_back(d);
        e->d_flag = (int*)d;
    }

    // threads_per_game: 0 = choose; n > 0 = n threads per game; n < 0 = warp kernel with -n games per block (tuning)
    e->threads = cfg->threads_per_game > 0 ? cfg->threads_per_game : (cfg->threads_per_game < 0 ? 32 : choose_threads(HW));
    if (e->threads > 1024) e->threads = 1024;
    if (e->threads < 32 || (e->threads & 31)) {
        mrts_destroy(e);
        return fail(MRTS_EINVAL, "threads_per_game must be a multiple of 32");
    }
    {
        EnvParams F = P;
        F.scap = P.cap;
        e->smem_full = layout_of(F).total;
    }
    // Staged unit capacity of the warp kernel: the blob in HBM holds `cap` units (the proven bound:
    // initial units + every resource on the map), shared memory only `scap`; a game that could
    // outgrow it during a step is deferred to mrts_overflow_kernel.  cfg->staged_units overrides.
    P.scap = P.cap;
    if (e->threads == 32 && (P.cap > 64 || cfg->staged_units > 0)) {
        int most = 0;
        for (const MapTpl& t : S.tpls) most = t.nunits > most ? t.nunits : most;
        int want = cfg->staged_units > 0 ? cfg->staged_units : (HW <= 256 ? 48 : 64);
        if (cfg->staged_units <= 0 && want < most + 16) want = most + 16;
        if (want < most) want = most;  // a reset must always fit
        if (want < P.cap) P.scap = want;
    }
    e->smem = layout_of(P).total;
    if (P.has_bot && e->threads == 32 && P.n_games >= 2048) {
        void *c = nullptr, *o = nullptr;
        if (cudaMalloc(&c, sizeof(uint32_t) * (size_t)P.n_games) == cudaSuccess && cudaMalloc(&o, sizeof(int32_t) * (4 * (size_t)P.n_games + 4)) == cudaSuccess) {
            e->allocs.push_back(c);
            e->allocs.push_back(o);
            cudaMemset(c, 0, sizeof(uint32_t) * (size_t)P.n_games);
            e->d_cost = (uint32_t*)c;
            e->d_order = (int32_t*)o;
        } else if (c) {
            cudaFree(c);
        }
    }
    if (P.scap < P.cap) {
        void* o = nullptr;
        ce = cudaMalloc(&o, sizeof(int32_t) * (size_t)(P.n_games + 1));
        if (ce != cudaSuccess) {
            mrts_destroy(e);
            return fail(MRTS_ENOMEM, std::string("cudaMalloc(overflow list): ") + cudaGetErrorString(ce));
        }
        e->allocs.push_back(o);
        e->d_ovf = (int32_t*)o;
    }
    int maxsm = 0, smem_sm = 0, max_blocks = 32;
    cudaDeviceGetAttribute(&maxsm, cudaDevAttrMaxSharedMemoryPerBlockOptin, cfg->device);
    cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, cfg->device);
    cudaDeviceGetAttribute(&max_blocks, cudaDevAttrMaxBlocksPerMultiprocessor, cfg->device);
    if (e->smem > maxsm || e->smem_full > maxsm) {
        mrts_destroy(e);
        return fail(MRTS_EINVAL, "map too large for shared-memory staging");
    }
    if (e->threads == 32) {
        // Games per block: a block's shared memory and warp slots are only released when its slowest
        // game is done and game times vary widely (scripted-bot path finding), so take the smallest
        // count that still reaches the most resident games per SM.
        e->minb = HW <= 64 ? 8 : MRTS_WARP_MIN_BLOCKS;
        const int by_regs = 4 * e->minb;  // warps per SM the register budget of the variant allows (__launch_bounds__(128, minb))
        int best = 1, best_w = 0;
        for (int g = 1; g <= GPB_MAX; g *= 2) {
            if (g * e->smem > maxsm) break;
            int blocks = smem_sm / (g * e->smem + 1024);  // 1 KB per-block reservation
            if (blocks > max_blocks) blocks = max_blocks;
            int w = blocks * g;
            if (w > by_regs) w = by_regs;
            if (w > best_w) best_w = w, best = g;
        }
        e->gpb = cfg->threads_per_game < 0 ? -cfg->threads_per_game : best;
        if (e->gpb > GPB_MAX) e->gpb = GPB_MAX;
        const int wbytes = e->gpb * e->smem;
        if (wbytes > 48 * 1024) {
            cudaFuncSetAttribute(mrts_pass_warp_kernel<4, 29, MRTS_WARP_MIN_BLOCKS>, cudaFuncAttributeMaxDynamicSharedMemorySize, wbytes);
            cudaFuncSetAttribute(mrts_pass_warp_kernel<4, 31, MRTS_WARP_MIN_BLOCKS>, cudaFuncAttributeMaxDynamicSharedMemorySize, wbytes);
            cudaFuncSetAttribute(mrts_pass_warp_kernel<1, 29, MRTS_WARP_MIN_BLOCKS>, cudaFuncAttributeMaxDynamicSharedMemorySize, wbytes);
            cudaFuncSetAttribute(mrts_pass_warp_kernel<1, 31, MRTS_WARP_MIN_BLOCKS>, cudaFuncAttributeMaxDynamicSharedMemorySize, wbytes);
            cudaFuncSetAttribute(mrts_pass_warp_kernel<4, 29, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, wbytes);
            cudaFuncSetAttribute(mrts_pass_warp_kernel<4, 31, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, wbytes);
            cudaFuncSetAttribute(mrts_pass_warp_kernel<1, 29, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, wbytes);
            cudaFuncSetAttribute(mrts_pass_warp_kernel<1, 31, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, wbytes);
        }
    }
    if (e->smem_full > 48 * 1024) {
        cudaFuncSetAttribute(mrts_overflow_kernel<4, 29>, cudaFuncAttributeMaxDynamicSharedMemorySize, e->smem_full);
        cudaFuncSetAttribute(mrts_overflow_kernel<4, 31>, cudaFuncAttributeMaxDynamicSharedMemorySize, e->smem_full);
        cudaFuncSetAttribute(mrts_overflow_kernel<1, 29>, cudaFuncAttributeMaxDynamicSharedMemorySize, e->smem_full);
        cudaFuncSetAttribute(mrts_overflow_kernel<1, 31>, cudaFuncAttributeMaxDynamicSharedMemorySize, e->smem_full);
        cudaFuncSetAttribute(mrts_debug_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, e->smem_full);
    }
    if (e->smem > 48 * 1024) {
        cudaFuncSetAttribute(mrts_pass_kernel<4, 29>, cudaFuncAttributeMaxDynamicSharedMemorySize, e->smem);
        cudaFuncSetAttribute(mrts_pass_kernel<4, 31>, cudaFuncAttributeMaxDynamicSharedMemorySize, e->smem);
        cudaFuncSetAttribute(mrts_pass_kernel<1, 29>, cudaFuncAttributeMaxDynamicSharedMemorySize, e->smem);
        cudaFuncSetAttribute(mrts_pass_kernel<1, 31>, cudaFuncAttributeMaxDynamicSharedMemorySize, e->smem);
    }
    // start every game from its template so that the state is defined before the first reset()
    rc = launch(e, MODE_RESET, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr);
    if (rc) {
        mrts_destroy(e);
        return rc;
    }
    ce = cudaDeviceSynchronize();
    if (ce != cudaSuccess) {
        mrts_destroy(e);
        return fail(MRTS_ECUDA, std::string("initial reset: ") + cudaGetErrorString(ce));
    }
    *out = e;
    return MRTS_OK;
}

void mrts_destroy(mrts_env* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    {
        auto& h = e->hio;
        if (h.stream) cudaStreamSynchronize(h.stream);
        for (void* p : {(void*)h.h_pact, (void*)h.h_pobs, (void*)h.h_pmask, (void*)h.h_rew, (void*)h.h_done})
            if (p) cudaFreeHost(p);
        for (void* p : {(void*)h.d_pact, (void*)h.d_pobs, (void*)h.d_pmask, (void*)h.d_rew, (void*)h.d_done})
            if (p) cudaFree(p);
        if (h.ev_obs) cudaEventDestroy(h.ev_obs);
        if (h.stream) cudaStreamDestroy(h.stream);
        mrts_host::pool_destroy(h.pool);
    }
    for (void* p : e->allocs) cudaFree(p);
    if (e->d_rows) cudaFree(e->d_rows);
    delete e;
}

int mrts_reset(mrts_env* e, void* obs, void* mask, void* stream) { return launch(e, MODE_RESET, nullptr, obs, mask, nullptr, nullptr, stream); }

int mrts_step(mrts_env* e, const int32_t* actions, void* obs, void* mask, double* reward, uint8_t* done, void* stream) {
    if (!actions && !(e && e->all_scripted)) return fail(MRTS_EINVAL, "actions is null");
    return launch(e, MODE_STEP, actions, obs, mask, reward, done, stream);
}

int mrts_get_masks(mrts_env* e, void* mask, void* stream) {
    if (!mask) return fail(MRTS_EINVAL, "mask is null");
    return launch(e, MODE_OBSERVE, nullptr, nullptr, mask, nullptr, nullptr, stream);
}

int mrts_step_ex(mrts_env* e, const mrts_step_io* io, void* stream) {
    if (!io || !io->actions) return fail(MRTS_EINVAL, "actions is null");
    return launch(e, MODE_STEP, io->actions, io->obs, io->mask, io->reward, io->done, stream, io->next_actions, io->seed);
}

int mrts_sample_actions(mrts_env* e, int32_t* next_actions, uint64_t seed, void* stream) {
    if (!next_actions) return fail(MRTS_EINVAL, "next_actions is null");
    return launch(e, MODE_OBSERVE, nullptr, nullptr, nullptr, nullptr, nullptr, stream, next_actions, seed);
}

static int hostio_init(mrts_env* e) {
    auto& h = e->hio;
    if (h.ready) return MRTS_OK;
    CK(cudaSetDevice(e->device));
    const size_t N = (size_t)e->N, HW = (size_t)e->P.HW;
    CK(cudaStreamCreateWithFlags(&h.stream, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&h.ev_obs, cudaEventDisableTiming));
    CK(cudaMallocHost((void**)&h.h_pact, N * HW * sizeof(uint16_t)));
    CK(cudaMallocHost((void**)&h.h_pobs, N * HW * sizeof(uint32_t)));
    CK(cudaMallocHost((void**)&h.h_pmask, N * HW * 3 * sizeof(uint32_t)));
    CK(cudaMallocHost((void**)&h.h_rew, N * 6 * sizeof(double)));
    CK(cudaMallocHost((void**)&h.h_done, N));
    CK(cudaMalloc((void**)&h.d_pact, N * HW * sizeof(uint16_t)));
    CK(cudaMalloc((void**)&h.d_pobs, N * HW * sizeof(uint32_t)));
    CK(cudaMalloc((void**)&h.d_pmask, N * HW * 3 * sizeof(uint32_t)));
    CK(cudaMalloc((void**)&h.d_rew, N * 6 * sizeof(double)));
    CK(cudaMalloc((void**)&h.d_done, N));
    h.pool = mrts_host::pool_create(0);
    h.ready = true;
    return MRTS_OK;
}

// device -> pinned staging -> the caller's arrays; the observation is widened while the mask words are still in flight
static int hostio_finish(mrts_env* e, void* h_obs, void* h_mask, double* h_reward, uint8_t* h_done, int slot = -1) {
    auto& h = e->hio;
    const size_t N = (size_t)e->N, HW = (size_t)e->P.HW;
    CK(cudaMemcpyAsync(h.h_pobs, h.d_pobs, N * HW * 4, cudaMemcpyDeviceToHost, h.stream));
    CK(cudaEventRecord(h.ev_obs, h.stream));
    CK(cudaMemcpyAsync(h.h_pmask, h.d_pmask, N * HW * 12, cudaMemcpyDeviceToHost, h.stream));
    if (h_reward) CK(cudaMemcpyAsync(h.h_rew, h.d_rew, N * 48, cudaMemcpyDeviceToHost, h.stream));
    if (h_done) CK(cudaMemcpyAsync(h.h_done, h.d_done, N, cudaMemcpyDeviceToHost, h.stream));
    CK(cudaEventSynchronize(h.ev_obs));
    mrts_env::HostIO::Slot* S = slot >= 0 ? &h.slots[slot] : nullptr;
    if (h_obs) {
        if (S && S->obs == h_obs && S->pobs.size() == N * HW) {
            mrts_host::expand_obs_delta(h.pool, h.h_pobs, S->pobs.data(), N, (int)HW, e->P.planes, e->P.elem, h_obs);
        } else {
            mrts_host::expand_obs(h.pool, h.h_pobs, N, (int)HW, e->P.planes, e->P.elem, h_obs);
            if (S) {
                S->pobs.assign(h.h_pobs, h.h_pobs + N * HW);
                S->obs = h_obs;
            }
        }
    }
    CK(cudaStreamSynchronize(h.stream));
    h.have_mask = true;
    if (h_mask) {
        if (S && S->mask == h_mask && S->pmask.size() == N * HW * 3) {
            mrts_host::expand_mask_delta(h.pool, h.h_pmask, S->pmask.data(), N, (int)HW, e->P.elem, h_mask);
        } else {
            mrts_host::expand_mask(h.pool, h.h_pmask, N, (int)HW, e->P.elem, h_mask);
            if (S) {
                S->pmask.assign(h.h_pmask, h.h_pmask + N * HW * 3);
                S->mask = h_mask;
            }
        }
    }
    if (h_reward) memcpy(h_reward, h.h_rew, N * 48);
    if (h_done) memcpy(h_done, h.h_done, N);
    return MRTS_OK;
}

int mrts_reset_host(mrts_env* e, void* h_obs, void* h_mask) { return mrts_reset_host_slot(e, h_obs, h_mask, -1); }
int mrts_step_host(mrts_env* e, const void* h_actions, int action_elem_bytes, void* h_obs, void* h_mask, double* h_reward, uint8_t* h_done) {
    return mrts_step_host_slot(e, h_actions, action_elem_bytes, h_obs, h_mask, h_reward, h_done, -1);
}
int mrts_host_forget_slot(mrts_env* e, int slot) {
    if (!e || slot < 0 || slot >= MRTS_HOST_SLOTS) return fail(MRTS_EINVAL, "bad slot");
    e->hio.slots[slot] = mrts_env::HostIO::Slot();
    return MRTS_OK;
}

int mrts_reset_host_slot(mrts_env* e, void* h_obs, void* h_mask, int slot) {
    if (!e) return fail(MRTS_EINVAL, "null env");
    if (slot >= MRTS_HOST_SLOTS) return fail(MRTS_EINVAL, "bad slot");
    int rc = hostio_init(e);
    if (rc) return rc;
    e->cur_pobs = e->hio.d_pobs;
    e->cur_pmask = e->hio.d_pmask;
    rc = launch(e, MODE_RESET, nullptr, nullptr, nullptr, nullptr, nullptr, e->hio.stream);
    if (rc) return rc;
    return hostio_finish(e, h_obs, h_mask, nullptr, nullptr, slot);
}

int mrts_step_host_slot(mrts_env* e, const void* h_actions, int action_elem_bytes, void* h_obs, void* h_mask, double* h_reward, uint8_t* h_done, int slot) {
    if (!e || !h_actions) return fail(MRTS_EINVAL, "null argument");
    if (slot >= MRTS_HOST_SLOTS) return fail(MRTS_EINVAL, "bad slot");
    if (action_elem_bytes != 4 && action_elem_bytes != 8) return fail(MRTS_EINVAL, "actions must be int32 or int64");
    int rc = hostio_init(e);
    if (rc) return rc;
    auto& h = e->hio;
    const size_t N = (size_t)e->N, HW = (size_t)e->P.HW;
    if (!h.have_mask) {  // first call without a preceding mrts_reset_host: fetch the mask words of the current state
        e->cur_pobs = h.d_pobs;
        e->cur_pmask = h.d_pmask;
        rc = launch(e, MODE_OBSERVE, nullptr, nullptr, nullptr, nullptr, nullptr, h.stream);
        if (rc) return rc;
        rc = hostio_finish(e, nullptr, nullptr, nullptr, nullptr);
        if (rc) return rc;
    }
    mrts_host::pack_actions(h.pool, h_actions, action_elem_bytes, h.h_pmask, N, (int)HW, h.h_pact);
    CK(cudaMemcpyAsync(h.d_pact, h.h_pact, N * HW * 2, cudaMemcpyHostToDevice, h.stream));
    e->cur_pact = h.d_pact;
    e->cur_pobs = h.d_pobs;
    e->cur_pmask = h.d_pmask;
    // `actions` is unused when the packed rows are given; pass a valid non-null pointer for the argument check
    rc = launch(e, MODE_STEP, (const int32_t*)h.d_pact, nullptr, nullptr, h.d_rew, h.d_done, h.stream);
    if (rc) return rc;
    return hostio_finish(e, h_obs, h_mask, h_reward, h_done, slot);
}

int mrts_launches_per_step(const mrts_env* e) {
    if (!e) return MRTS_EINVAL;
    return 1 + (e->d_order && e->threads == 32 ? 1 : 0) + (e->threads == 32 && e->P.scap < e->P.cap ? 1 : 0) + (e->n_cycle > 0 ? 2 : 0);
}
int mrts_staged_units(const mrts_env* e) { return e ? e->P.scap : MRTS_EINVAL; }
// the transport codec on its own (host arrays in, host arrays out): lets the CPU test-suite check it without a GPU
int mrts_host_expand_obs(const uint32_t* words, int64_t n_env, int hw, int planes, int elem_bytes, void* out) {
    if (!words || !out || (elem_bytes != 4 && elem_bytes != 1) || planes < 1 || planes > 31) return fail(MRTS_EINVAL, "bad argument");
    mrts_host::Pool* p = mrts_host::pool_create(0);
    mrts_host::expand_obs(p, words, (size_t)n_env, hw, planes, elem_bytes, out);
    mrts_host::pool_destroy(p);
    return MRTS_OK;
}
int mrts_host_expand_mask(const uint32_t* mask_words, int64_t n_env, int hw, int elem_bytes, void* out) {
    if (!mask_words || !out || (elem_bytes != 4 && elem_bytes != 1)) return fail(MRTS_EINVAL, "bad argument");
    mrts_host::Pool* p = mrts_host::pool_create(0);
    mrts_host::expand_mask(p, mask_words, (size_t)n_env, hw, elem_bytes, out);
    mrts_host::pool_destroy(p);
    return MRTS_OK;
}
int mrts_host_expand_obs_delta(const uint32_t* words, uint32_t* shadow, int64_t n_env, int hw, int planes, int elem_bytes, void* out) {
    if (!words || !shadow || !out || (elem_bytes != 4 && elem_bytes != 1) || planes < 1 || planes > 31) return fail(MRTS_EINVAL, "bad argument");
    mrts_host::Pool* p = mrts_host::pool_create(0);
    mrts_host::expand_obs_delta(p, words, shadow, (size_t)n_env, hw, planes, elem_bytes, out);
    mrts_host::pool_destroy(p);
    return MRTS_OK;
}
int mrts_host_expand_mask_delta(const uint32_t* mask_words, uint32_t* shadow, int64_t n_env, int hw, int elem_bytes, void* out) {
    if (!mask_words || !shadow || !out || (elem_bytes != 4 && elem_bytes != 1)) return fail(MRTS_EINVAL, "bad argument");
    mrts_host::Pool* p = mrts_host::pool_create(0);
    mrts_host::expand_mask_delta(p, mask_words, shadow, (size_t)n_env, hw, elem_bytes, out);
    mrts_host::pool_destroy(p);
    return MRTS_OK;
}
int mrts_host_pack_actions(const void* actions, int action_elem_bytes, const uint32_t* mask_words, int64_t n_env, int hw, uint16_t* out) {
    if (!actions || !mask_words || !out || (action_elem_bytes != 4 && action_elem_bytes != 8)) return fail(MRTS_EINVAL, "bad argument");
    mrts_host::Pool* p = mrts_host::pool_create(0);
    mrts_host::pack_actions(p, actions, action_elem_bytes, mask_words, (size_t)n_env, hw, out);
    mrts_host::pool_destroy(p);
    return MRTS_OK;
}
int mrts_host_threads(const mrts_env* e) { return e && e->hio.pool ? mrts_host::pool_threads(e->hio.pool) : 0; }

int mrts_num_envs(const mrts_env* e) { return e ? e->N : MRTS_EINVAL; }
int mrts_num_games(const mrts_env* e) { return e ? e->P.n_games : MRTS_EINVAL; }
int mrts_planes(const mrts_env* e) { return e ? e->P.planes : MRTS_EINVAL; }
int mrts_unit_capacity(const mrts_env* e) { return e ? e->P.cap : MRTS_EINVAL; }
int mrts_threads_per_game(const mrts_env* e) { return e ? e->threads : MRTS_EINVAL; }
int mrts_smem_bytes(const mrts_env* e) { return e ? e->smem : MRTS_EINVAL; }

int mrts_export_state(mrts_env* e, int game, int32_t* hdr5, int32_t* rows12, int cap_rows, void* stream) {
    if (!e || game < 0 || game >= e->P.n_games) return fail(MRTS_EINVAL, "bad game index");
    CK(cudaSetDevice(e->device));
    CK(cudaStreamSynchronize((cudaStream_t)stream));
    std::vector<uint8_t> blob(e->P.game_stride);
    CK(cudaMemcpy(blob.data(), e->P.games + (size_t)game * e->P.game_stride, blob.size(), cudaMemcpyDeviceToHost));
    export_blob(blob.data(), hdr5, rows12, cap_rows);
    return ((const GameHdr*)blob.data())->nunits;
}

static int debug_launch(mrts_env* e, int game, int op, const int32_t* rows7, int n, int* gameover, void* stream) {
    if (!e || game < 0 || game >= e->P.n_games) return fail(MRTS_EINVAL, "bad game index");
    CK(cudaSetDevice(e->device));
    DebugArgs A;
    A.game = game;
    A.op = op;
    A.rows7 = nullptr;
    A.n = n;
    A.gameover = e->d_flag;
    if (op == 0 && n > 0) {
        if (n > e->rows_cap) {
            if (e->d_rows) cudaFree(e->d_rows);
            e->d_rows = nullptr;
            CK(cudaMalloc((void**)&e->d_rows, (size_t)n * 7 * sizeof(int32_t)));
            e->rows_cap = n;
        }
        CK(cudaMemcpyAsync(e->d_rows, rows7, (size_t)n * 7 * sizeof(int32_t), cudaMemcpyHostToDevice, (cudaStream_t)stream));
        A.rows7 = e->d_rows;
    }
    EnvParams F = e->P;
    F.scap = F.cap;
    F.ovf = nullptr;
    mrts_debug_kernel<<<1, 64, e->smem_full, (cudaStream_t)stream>>>(F, A);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize((cudaStream_t)stream));
    if (op == 1 && gameover) CK(cudaMemcpy(gameover, e->d_flag, sizeof(int), cudaMemcpyDeviceToHost));
    return MRTS_OK;
}
int mrts_set_cycle_maps(mrts_env* e, const int32_t* template_idx, int n) {
    if (!e || n < 0 || (n > 0 && !template_idx)) return fail(MRTS_EINVAL, "bad argument");
    if (n > 0 && e->P.num_selfplay > 0 && e->P.num_bot > 0)
        return fail(MRTS_EINVAL, "cycle maps need all-bot or all-self-play envs: with both kinds the reference's index test (vec_env.py:1041) addresses the wrong clients");
    for (int k = 0; k < n; k++)
        if (template_idx[k] < 0 || template_idx[k] >= e->n_maps) return fail(MRTS_EINVAL, "cycle map index outside the templates given to mrts_create");
    CK(cudaSetDevice(e->device));
    CK(cudaDeviceSynchronize());
    e->n_cycle = 0;
    if (n == 0) return MRTS_OK;
    if (!e->d_ended) {
        void *a = nullptr, *b = nullptr, *c = nullptr;
        if (cudaMalloc(&a, (size_t)e->P.n_games) != cudaSuccess || cudaMalloc(&b, sizeof(int32_t)) != cudaSuccess ||
            cudaMalloc(&c, sizeof(int32_t) * ((size_t)e->P.n_games + 1)) != cudaSuccess) {
            cudaFree(a);
            cudaFree(b);
            cudaFree(c);
            return fail(MRTS_ENOMEM, "cudaMalloc(cycle maps)");
        }
        e->allocs.push_back(a);
        e->allocs.push_back(b);
        e->allocs.push_back(c);
        e->d_ended = (uint8_t*)a;
        e->d_cycle_pos = (int32_t*)b;
        e->d_cycle_list = (int32_t*)c;
    }
    void* d = nullptr;
    CK(cudaMalloc(&d, sizeof(int32_t) * (size_t)n));
    e->allocs.push_back(d);
    e->d_cycle = (int32_t*)d;
    CK(cudaMemcpy(d, template_idx, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice));
    CK(cudaMemset(e->d_cycle_pos, 0, sizeof(int32_t)));  // itertools.cycle starts at the first entry (vec_env.py:133)
    CK(cudaMemset(e->d_ended, 0, (size_t)e->P.n_games));
    e->n_cycle = n;
    return MRTS_OK;
}
int mrts_debug_timing(mrts_env* e, long long* device_buf) {
    if (!e) return fail(MRTS_EINVAL, "null env");
    e->d_timing = device_buf;
    return MRTS_OK;
}
int mrts_debug_issue(mrts_env* e, int game, const int32_t* rows7, int n, void* stream) { return debug_launch(e, game, 0, rows7, n, nullptr, stream); }
int mrts_debug_cycle(mrts_env* e, int game, int* gameover, void* stream) { return debug_launch(e, game, 1, nullptr, 0, gameover, stream); }

}  // extern "C"
