// C ABI of include/mrts_b200.h and the launch logic behind it.
//
// A step is two kinds of work with opposite bounds:
//   rules   (mrts_rules.cuh)  latency-bound: one warp per game, ~1.4 KB of packed output per view
//   expand  (mrts_expand.cu)  HBM-bound: 27 KB - 1.75 MB per view written with bulk async stores
// The SPLIT pipeline cuts the slab into chunks and runs them through the two kernels on separate
// streams, so that the store stream of chunk k overlaps the rules of chunk k + 1 on the same SMs.
// The DIRECT pipeline (small slabs, shapes the bulk path cannot take) writes the tensors from the
// rules pass itself, one launch.  No CPU fallback exists: every entry point launches a kernel.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <string>
#include <vector>

#include "../../include/mrts_b200.h"
#include "mrts_host.h"
#include "mrts_launch.h"
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

int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return e && *e ? atoi(e) : dflt;
}

}  // namespace

// error reporting for the other translation units that hold entry points (mrts_expand_static.cu)
namespace mrts {
int api_fail(int code, const std::string& msg) { return fail(code, msg); }
}  // namespace mrts

constexpr int MAX_CHUNKS = 16;

struct mrts_env {
    EnvParams P;
    int device = 0;
    int N = 0, n_maps = 0, n_sm = 148;
    bool all_scripted = false;  // every game is bot vs bot (mrts_config::bot_ai1): mrts_step takes no actions
    // ---- kernels
    int threads = 32;           // direct mode: 32 = warp kernels, otherwise threads of the block kernel
    int regs[2] = {64, 96};     // register budget variant of the warp kernels (light, heavy)
    int gpb[2] = {1, 1};        // games (warps) per block (light, heavy)
    int smem[2] = {0, 0};       // bytes per game at the staged capacity (light, heavy)
    int smem_block = 0;         // block kernel (staged = full capacity, heavy layout)
    int smem_full = 0;          // bytes per game at the full capacity (overflow / debug launches)
    // ---- pipeline
    int pipeline = MRTS_PIPELINE_DIRECT;  // what mrts_step uses when the shape allows
    int K = 1;                  // slab chunks of the split pipeline
    int exp_blocks = 148;       // persistent blocks of the expansion kernel
    int chunk_g0[MAX_CHUNKS + 1] = {0};       // game ranges
    int chunk_nlight[MAX_CHUNKS] = {0}, chunk_nheavy[MAX_CHUNKS] = {0}, chunk_light_off[MAX_CHUNKS] = {0};
    int32_t* d_light = nullptr;  // light games, chunk after chunk
    int32_t* d_order[2] = {nullptr, nullptr};  // heavy games: 4 class lists of n_games entries, this step's / next step's
    int order_cur = 0;
    int32_t* d_ctr = nullptr;    // [order counts 0 (4K) | deferred counts (K) | expansion work counters (K) | order counts 1 (4K)]
    int32_t* d_ovf_list = nullptr;
    uint32_t *d_pobs = nullptr, *d_ment = nullptr;  // packed outputs of the split pipeline
    cudaStream_t s_rules[2] = {nullptr, nullptr}, s_exp = nullptr;
    // direct pipeline with both game classes: the light (bandwidth-bound) launch runs on its own low-priority stream
    // beside the heavy (latency-bound) one instead of behind its tail
    cudaStream_t s_light = nullptr;
    cudaEvent_t ev_lfork = nullptr, ev_ljoin = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_rules[MAX_CHUNKS] = {nullptr};
    long long* d_timing = nullptr;  // profiling aid (mrts_debug_timing)
    // ---- host-buffer entry points (mrts_reset_host / mrts_step_host): packed transport, see mrts_host.h
    struct HostIO {
        cudaStream_t stream = nullptr;
        cudaEvent_t ev_obs = nullptr, ev_user = nullptr;
        uint16_t *h_pact = nullptr, *d_pact = nullptr;
        uint32_t *h_pobs = nullptr, *d_pobs = nullptr, *h_pmask = nullptr, *d_pmask = nullptr;
        double *h_rew = nullptr, *d_rew = nullptr;
        uint8_t *h_done = nullptr, *d_done = nullptr;
        mrts_host::Pool* pool = nullptr;
        bool ready = false, have_mask = false;
        // result-array pairs the caller hands back unmodified (mrts_*_host_slot): the words last expanded into them
        struct Slot {
            void *obs = nullptr, *mask = nullptr;
            std::vector<uint32_t> pobs, pmask;  // host-side shadows (calls that pass only one of the two arrays)
            bool on_device = false;             // d_shadow[slot] holds the words expanded into (obs, mask)
        } slots[MRTS_HOST_SLOTS];
        // delta transport of the slot calls (mrts_delta.cu): ONE device buffer and its pinned mirror
        //   [header 64 words: word 0 = changed cells | reward N x 6 doubles | done N bytes | bitmap | entries cap x 6 words]
        // so that a step's results leave with a single copy
        uint32_t *d_delta = nullptr, *h_delta = nullptr;
        size_t off_rew = 0, off_done = 0, off_bitmap = 0, off_entries = 0, delta_cap = 0;  // byte offsets; capacity in entries
        uint32_t* d_shadow[MRTS_HOST_SLOTS] = {nullptr};
        // action rows go up as a list of the acting cells {flat cell, packed row} and are scattered on the device
        uint32_t *h_pairs = nullptr, *d_pairs[2] = {nullptr, nullptr};
        size_t pairs_cap = 0, pairs_prev_n = 0;
        int pairs_cur = 0;
        bool pact_clean = false;  // d_pact is zero except at the cells of d_pairs[pairs_cur]
        size_t last_count = 0;
        bool bitmap_valid = false;  // h_delta's bitmap describes the mask of the current state (else h_pmask does)
        double t_phase[8] = {0};    // accumulated wall seconds: pack, upload + launch, wait, patch / expand + copy-out; [4] = bytes host->device, [5] = changed cells, [6] = bytes device->host, [7] = calls
    } hio;
    const uint16_t* cur_pact = nullptr;  // per-launch packed pointers (set by the host / packed entry points)
    uint32_t *cur_pobs = nullptr, *cur_pmask = nullptr, *cur_ment = nullptr;
    cudaStream_t last_stream = nullptr;  // stream of the last launch and an event recorded behind it: a launch on
    cudaEvent_t ev_last = nullptr;       // another stream waits for it (the game blobs and scratch lists are shared)
    bool have_last = false;
    // cycle_maps (mrts_set_cycle_maps): rotation list, its position, per-game ended flags, list of games to restart
    int n_cycle = 0;
    int32_t *d_cycle = nullptr, *d_cycle_pos = nullptr, *d_cycle_list = nullptr;
    uint8_t* d_ended = nullptr;
    std::vector<void*> allocs;
    int* d_flag = nullptr;
    int32_t* d_rows = nullptr;
    int rows_cap = 0;
    uint32_t sample_counter = 0;
    int last_launches = 0;  // kernels enqueued by the last launch()
    int diag_skip = 0;      // MRTS_DIAG_SKIP bits: 1 / 2 = heavy games write no mask / obs, 4 / 8 = light games (timing experiments only)
    int bulk_mask = 1;      // direct mode: the mask takes the bulk writer (MRTS_BULK_MASK overrides, tuning)
};

namespace {

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
template <class T>
int dalloc(mrts_env* e, T** out, size_t count, bool zero = true) {
    void* d = nullptr;
    const size_t bytes = (count ? count : 1) * sizeof(T);
    cudaError_t ce = cudaMalloc(&d, bytes);
    if (ce != cudaSuccess) return fail(MRTS_ENOMEM, std::string("cudaMalloc: ") + cudaGetErrorString(ce));
    e->allocs.push_back(d);
    if (zero) CK(cudaMemset(d, 0, bytes));
    *out = (T*)d;
    return MRTS_OK;
}

int env0_of(const EnvParams& P, int g) {  // first env view of game g (self-play games come first, two views each)
    const int nsp = P.num_selfplay / 2;
    return g < nsp ? 2 * g : P.num_selfplay + (g - nsp);
}

// Launches on different streams share the game blobs and the per-env scratch: order them with an event
int order_after_last(mrts_env* e, cudaStream_t st) {
    if (e->have_last && e->last_stream != st) CK(cudaStreamWaitEvent(st, e->ev_last, 0));
    return MRTS_OK;
}
int mark_last(mrts_env* e, cudaStream_t st) {
    CK(cudaEventRecord(e->ev_last, st));
    e->last_stream = st;
    e->have_last = true;
    return MRTS_OK;
}

int launch(mrts_env* e, int mode, const int32_t* actions, void* obs, void* mask, double* reward, uint8_t* done, void* stream,
           int32_t* next_actions = nullptr, uint64_t seed = 0) {
    if (!e) return fail(MRTS_EINVAL, "null env");
    CK(cudaSetDevice(e->device));
    const cudaStream_t st = (cudaStream_t)stream;
    int rc = order_after_last(e, st);
    if (rc) return rc;
    EnvParams P = e->P;
    P.mode = mode;
    P.actions = actions;
    P.reward = reward;
    P.done = done;
    P.next_actions = next_actions;
    P.seed = seed;
    P.counter = next_actions ? e->sample_counter++ : 0;
    P.pact = e->cur_pact;
    P.pobs = e->cur_pobs;
    P.pmask = e->cur_pmask;
    P.ment = e->cur_ment;
    e->cur_pact = nullptr;
    e->cur_pobs = e->cur_pmask = e->cur_ment = nullptr;
    P.timing = e->d_timing;
    const bool rotate = e->n_cycle > 0 && mode == MODE_STEP;
    P.ended = rotate ? e->d_ended : nullptr;

    // ---- which pipeline
    const bool want = obs != nullptr || mask != nullptr;
    const bool split = e->pipeline == MRTS_PIPELINE_SPLIT && want && e->d_pobs && expand_supported(P.HW, P.planes, P.elem, obs, mask);
    const bool warp = split || e->threads == 32;
    const int K = split && !rotate ? e->K : 1;
    if (split) {
        if (obs) P.pobs = e->d_pobs;
        if (mask) P.ment = e->d_ment;
        P.obs = nullptr;
        P.mask = nullptr;
    } else {
        P.obs = obs;
        P.mask = mask;
    }
    if (!warp) P.scap = P.cap;  // the block kernel stages the full capacity
    // direct mode on the warp kernels: the tensors go through the copy engine when their rows line up with 16-byte groups
    P.bulk_out = !split && warp && mask && e->bulk_mask && (P.HW * P.elem) % 8 == 0 && (((uintptr_t)mask) & 15) == 0 ? 1 : 0;
    const bool defer = warp && P.scap < P.cap;

    // ---- per-step counters: next step's class counts, deferred-game counts, expansion work counters
    const int Kc = e->K;
    int32_t* const cnt_cur = e->d_ctr + (e->order_cur == 0 ? 0 : 6 * Kc);
    int32_t* const cnt_next = e->d_ctr + (e->order_cur == 0 ? 6 * Kc : 0);
    int32_t* const ovf_cnt = e->d_ctr + 4 * Kc;
    int32_t* const exp_ctr = e->d_ctr + 5 * Kc;
    CK(cudaMemsetAsync(e->order_cur == 0 ? ovf_cnt : e->d_ctr, 0, sizeof(int32_t) * 6 * Kc, st));
    const int32_t* const ord_cur = e->d_order[e->order_cur];
    int32_t* const ord_next = e->d_order[e->order_cur ^ 1];

    const bool fork = K > 1;
    if (fork) {
        CK(cudaEventRecord(e->ev_fork, st));
        CK(cudaStreamWaitEvent(e->s_rules[0], e->ev_fork, 0));
        CK(cudaStreamWaitEvent(e->s_rules[1], e->ev_fork, 0));
        CK(cudaStreamWaitEvent(e->s_exp, e->ev_fork, 0));
    }
    int launches = 0;
    const int ogrid = 2 * e->n_sm;
    for (int c = 0; c < K; c++) {
        // chunk c of the pipeline = chunks [c, c + 1) of the static partition, or all of them when K == 1
        const int c0 = K == 1 ? 0 : c, c1 = K == 1 ? Kc : c + 1;
        const cudaStream_t sr = fork ? e->s_rules[c & 1] : st;
        for (int q = c0; q < c1; q++) {
            EnvParams Q = P;
            Q.game0 = e->chunk_g0[q];
            Q.game1 = e->chunk_g0[q + 1];
            Q.order_stride = P.n_games;
            if (defer) {
                Q.ovf_cnt = ovf_cnt + q;
                Q.ovf_list = e->d_ovf_list + Q.game0;
            }
            if (mode == MODE_STEP) {
                Q.order_next = ord_next + Q.game0;
                Q.order_cnt_next = cnt_next + 4 * q;
            }
            const bool beside = e->s_light && !fork && e->chunk_nheavy[q] > 0 && e->chunk_nlight[q] > 0;
            const cudaStream_t sl = beside ? e->s_light : sr;
            if (beside) {
                CK(cudaEventRecord(e->ev_lfork, sr));  // behind the counter memset and everything earlier on the caller's stream
                CK(cudaStreamWaitEvent(sl, e->ev_lfork, 0));
            }
            if (e->chunk_nheavy[q] > 0) {
                EnvParams H = Q;
                H.order = ord_cur + Q.game0;
                H.order_cnt = cnt_cur + 4 * q;
                H.nslots = e->chunk_nheavy[q];
                if (e->diag_skip & 1) H.mask = nullptr, H.bulk_out = 0;  // diagnostics only (MRTS_DIAG_SKIP): outputs are then incomplete
                if (e->diag_skip & 2) H.obs = nullptr;
                if (warp) CK(launch_rules_warp(1, e->regs[1], H, e->gpb[1], e->smem[1], sr));
                else CK(launch_rules_block(1, H, e->threads, e->smem_block, sr));
                launches++;
            }
            if (e->chunk_nlight[q] > 0) {
                EnvParams L = Q;
                if (warp) L.has_bot = 0;  // light games stage no scripted-bot table / search scratch (smaller shared-memory plan)
                L.order = e->d_light + e->chunk_light_off[q];
                L.order_cnt = nullptr;
                L.nslots = e->chunk_nlight[q];
                if (e->diag_skip & 4) L.mask = nullptr, L.bulk_out = 0;
                if (e->diag_skip & 8) L.obs = nullptr;
                if (warp) CK(launch_rules_warp(0, e->regs[0], L, e->gpb[0], e->smem[0], sl));
                else CK(launch_rules_block(0, L, e->threads, e->smem_block, sl));
                launches++;
            }
            if (beside) {
                CK(cudaEventRecord(e->ev_ljoin, sl));
                CK(cudaStreamWaitEvent(sr, e->ev_ljoin, 0));
            }
            if (defer) {
                EnvParams F = Q;  // full-capacity parameters of the overflow launch
                F.scap = P.cap;
                F.ovf_cnt = nullptr;
                F.ovf_list = nullptr;
                CK(launch_overflow(F, ovf_cnt + q, e->d_ovf_list + Q.game0, ogrid, e->smem_full, sr));
                launches++;
            }
        }
        if (rotate) {
            // map rotation: the games that ended restart once more from their new template; only their packed / direct
            // outputs and sampled-action rows are rewritten (reward and done keep the step's values, vec_env.py:1057)
            EnvParams R = P;
            R.scap = P.cap;
            R.ovf_cnt = nullptr;
            R.ovf_list = nullptr;
            R.mode = MODE_RESET;
            R.actions = nullptr;
            R.pact = nullptr;
            R.reward = nullptr;
            R.done = nullptr;
            R.ended = nullptr;
            R.order_next = nullptr;
            mrts_cycle_kernel<<<1, CYCLE_THREADS, 0, sr>>>(e->d_ended, P.n_games, e->d_cycle, e->n_cycle, e->d_cycle_pos, const_cast<int32_t*>(P.game_map),
                                                           e->d_cycle_list);
            CK(cudaGetLastError());
            CK(launch_overflow(R, e->d_cycle_list, e->d_cycle_list + 1, ogrid, e->smem_full, sr));
            launches += 2;
        }
        if (split) {
            ExpandParams X;
            X.pobs = e->d_pobs;
            X.ment = e->d_ment;
            X.obs = obs;
            X.mask = mask;
            X.env_lo = env0_of(P, e->chunk_g0[c0]);
            X.env_hi = e->chunk_g0[c1] >= P.n_games ? e->N : env0_of(P, e->chunk_g0[c1]);
            X.HW = P.HW;
            X.planes = P.planes;
            X.elem = P.elem;
            X.cap = P.cap;
            X.empty_word = obs_empty_word(P.partial);
            X.counter = exp_ctr + c0;
            cudaStream_t se = st;
            if (fork) {
                CK(cudaEventRecord(e->ev_rules[c], sr));
                CK(cudaStreamWaitEvent(e->s_exp, e->ev_rules[c], 0));
                se = e->s_exp;
            }
            const int views = X.env_hi - X.env_lo;
            int blocks = (views + 7) / 8;
            if (blocks > e->exp_blocks) blocks = e->exp_blocks;
            if (blocks < 1) blocks = 1;
            CK(launch_expand(X, blocks, se));
            launches++;
        }
    }
    if (fork) {
        CK(cudaEventRecord(e->ev_join, e->s_exp));
        CK(cudaStreamWaitEvent(st, e->ev_join, 0));
    }
    if (mode == MODE_STEP) e->order_cur ^= 1;
    e->last_launches = launches;
    return mark_last(e, st);
}

// direct mode: 32 = one warp per game up to 16x16; above it a block per game so that every thread takes part in the
// store loops (446 KB per 32x32 partial view, 1.87 MB per 64x64 view) and in the per-cell visibility tests.  Measured at
// 32x32 partial (cfg4, M env-steps/s): warp kernel 8.8, blocks of 64 / 128 / 256 threads 8.8 / 10.6 / 8.4; at 64x64 256
// threads.  cfg->threads_per_game overrides.
int choose_threads(int HW) { return HW <= 256 ? 32 : (HW <= 1024 ? 128 : 256); }

}  // namespace

extern "C" {

const char* mrts_last_error(void) { return g_err.c_str(); }
const char* mrts_version(void) { return "mrts-b200 0.2 (sm_100a)"; }

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
    const int HW = P.HW;
#define CREATE_CK(call)                                                                 \
    do {                                                                                \
        cudaError_t ce_ = (call);                                                       \
        if (ce_ != cudaSuccess) {                                                       \
            mrts_destroy(e);                                                            \
            return fail(MRTS_ECUDA, std::string(#call) + ": " + cudaGetErrorString(ce_)); \
        }                                                                               \
    } while (0)
#define CREATE_RC(call)       \
    do {                      \
        if ((rc = (call))) {  \
            mrts_destroy(e);  \
            return rc;        \
        }                     \
    } while (0)
    CREATE_CK(cudaSetDevice(cfg->device));
    CREATE_CK(cudaDeviceGetAttribute(&e->n_sm, cudaDevAttrMultiProcessorCount, cfg->device));
    CREATE_RC(upload(e, S.tpls, &P.maps));
    CREATE_RC(upload(e, S.grids, &P.grid_pool));
    CREATE_RC(upload(e, S.units, &P.unit_pool));
    CREATE_RC(upload(e, S.game_map, &P.game_map));
    CREATE_RC(upload(e, S.game_ai, &P.game_ai));
    CREATE_RC(upload(e, S.game_ai1, &P.game_ai1));
    CREATE_RC(dalloc(e, &P.games, (size_t)P.n_games * P.game_stride));
    CREATE_RC(dalloc(e, &e->d_flag, 1));
    CREATE_CK(cudaEventCreateWithFlags(&e->ev_last, cudaEventDisableTiming));

    // ---- kernels.  threads_per_game: 0 = choose; n > 0 = n threads per game (block kernel, direct mode);
    // n < 0 = warp kernels with -n games per block (tuning)
    e->threads = cfg->threads_per_game > 0 ? cfg->threads_per_game : (cfg->threads_per_game < 0 ? 32 : choose_threads(HW));
    if (cfg->threads_per_game == 0 && env_int("MRTS_THREADS_PER_GAME", 0) > 0) e->threads = env_int("MRTS_THREADS_PER_GAME", 0);  // tuning experiments
    if (e->threads > 256) e->threads = 256;
    if (e->threads < 32 || (e->threads & 31)) {
        mrts_destroy(e);
        return fail(MRTS_EINVAL, "threads_per_game must be a multiple of 32");
    }
    {
        EnvParams F = P;
        F.scap = P.cap;
        e->smem_full = layout_of(F).total;
        e->smem_block = e->smem_full;
    }
    // Staged unit capacity of the warp kernels: the blob in HBM holds `cap` units (the proven bound:
    // initial units + every resource on the map), shared memory only `scap`; a game that could
    // outgrow it during a step is deferred to mrts_overflow_kernel.  cfg->staged_units overrides.
    P.scap = P.cap;
    if (P.cap > 64 || cfg->staged_units > 0) {
        int most = 0;
        for (const MapTpl& t : S.tpls) most = t.nunits > most ? t.nunits : most;
        int want = cfg->staged_units > 0 ? cfg->staged_units : (HW <= 256 ? 48 : 64);
        if (cfg->staged_units <= 0 && want < most + 16) want = most + 16;
        if (want < most) want = most;  // a reset must always fit
        if (want < P.cap) P.scap = want;
    }
    {
        EnvParams L = P;
        L.has_bot = 0;
        e->smem[0] = layout_of(L).total;
        e->smem[1] = layout_of(P).total;
    }
    int maxsm = 0, smem_sm = 0, max_blocks = 32;
    CREATE_CK(cudaDeviceGetAttribute(&maxsm, cudaDevAttrMaxSharedMemoryPerBlockOptin, cfg->device));
    CREATE_CK(cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, cfg->device));
    CREATE_CK(cudaDeviceGetAttribute(&max_blocks, cudaDevAttrMaxBlocksPerMultiprocessor, cfg->device));
    if (e->smem[1] > maxsm || e->smem_full > maxsm) {
        mrts_destroy(e);
        return fail(MRTS_EINVAL, "map too large for shared-memory staging");
    }
    // ---- pipeline
    e->pipeline = cfg->pipeline == MRTS_PIPELINE_DIRECT || cfg->pipeline == MRTS_PIPELINE_SPLIT ? cfg->pipeline : 0;
    {
        const int forced = env_int("MRTS_PIPELINE", 0);
        if (forced == MRTS_PIPELINE_DIRECT || forced == MRTS_PIPELINE_SPLIT) e->pipeline = forced;
    }
    if (e->pipeline == 0) e->pipeline = MRTS_PIPELINE_DIRECT;  // measured faster than SPLIT on every BASELINE config
    e->bulk_mask = env_int("MRTS_BULK_MASK", -1);  // 1 / 0 = force the bulk mask writer on / off; -1 = choose below
    // Register budget: light games (self-play, PassiveAI) fit 64 registers without spills -> 32 warps per SM;
    // the scripted-bot kernels take 80 when every game is one (44 bytes of spills; 25 instead of 21 resident warps: +7 % on WorkerRush).
    e->regs[0] = env_int("MRTS_REGS_LIGHT", 64);
    {
        // With both game classes the light launch runs beside the heavy one (s_light): 96 registers leave the light warps
        // room on every SM.  Measured on cfg2 (M env-steps/s, int32 / uint8): 96 -> 40.7 / 48.5, 80 -> 39.3 / 40.9.
        int n_heavy = 0;
        for (int g = 0; g < P.n_games; g++) n_heavy += ai_plans(S.game_ai[g]) || ai_plans(S.game_ai1[g]);
        const bool mixed = n_heavy > 0 && n_heavy < P.n_games && env_int("MRTS_OVERLAP", 1);
        e->regs[1] = env_int("MRTS_REGS_HEAVY", mixed ? 96 : 80);
    }
    // Games per block: a block's shared memory and warp slots are only released when its slowest
    // game is done and game times vary widely (scripted-bot path finding), so take the smallest
    // count that still reaches the most resident games per SM.  The bulk mask writer adds one tile of zeros per
    // block (see the choice below).
    auto choose_gpb = [&](int k, int tile_bytes, int* warps) {
        const int by_regs = 65536 / (32 * (e->regs[k] <= 64 ? 64 : (e->regs[k] <= 80 ? 80 : (e->regs[k] <= 96 ? 96 : 128))));
        int best = 1, best_w = 0;
        for (int g = 1; g <= GPB_MAX; g *= 2) {
            if (g * e->smem[k] + tile_bytes > maxsm) break;
            int blocks = smem_sm / (g * e->smem[k] + 1024 + tile_bytes);  // 1 KB per-block reservation
            if (blocks > max_blocks) blocks = max_blocks;
            int w = blocks * g;
            if (w > by_regs) w = by_regs;
            if (w > best_w) best_w = w, best = g;
        }
        *warps = best_w;
        return best;
    };
    {
        int w_plain[2], w_bulk[2], g_plain[2], g_bulk[2];
        for (int k = 0; k < 2; k++) {
            g_plain[k] = choose_gpb(k, 0, &w_plain[k]);
            g_bulk[k] = choose_gpb(k, 4992, &w_bulk[k]);
        }
        // Measured (profiles/ab_r02_bulk_mask.md): up to 16x16 the bulk writer wins although its tile costs an eighth
        // of the resident light games (28 of 32 warps); on 32x32 (16 of 18 warps) it loses.  Larger maps keep the
        // occupancy test.
        if (e->bulk_mask < 0) {
            const bool keeps_occupancy = 20 * w_bulk[0] >= 19 * w_plain[0] && 20 * w_bulk[1] >= 19 * w_plain[1];
            e->bulk_mask = P.HW <= 256 || keeps_occupancy ? 1 : 0;
        }
        if (e->pipeline != MRTS_PIPELINE_DIRECT) e->bulk_mask = e->bulk_mask > 0 ? 1 : 0;
        for (int k = 0; k < 2; k++) {
            e->gpb[k] = cfg->threads_per_game < 0 ? -cfg->threads_per_game : (e->bulk_mask ? g_bulk[k] : g_plain[k]);
            const int forced = env_int(k ? "MRTS_GPB_HEAVY" : "MRTS_GPB_LIGHT", 0);
            if (forced > 0) e->gpb[k] = forced;
            if (e->gpb[k] > GPB_MAX) e->gpb[k] = GPB_MAX;
            while (e->gpb[k] > 1 && e->gpb[k] * e->smem[k] + 4992 > maxsm) e->gpb[k] /= 2;
        }
    }
    CREATE_CK(light_set_smem(e->gpb[0] * e->smem[0] + 4992, e->smem_block));
    CREATE_CK(heavy_set_smem(e->gpb[1] * e->smem[1] + 4992, e->smem_block));
    CREATE_CK(any_set_smem(e->smem_full));
    CREATE_CK(expand_init());

    // ---- chunks of the slab, light / heavy game lists per chunk
    int K = cfg->chunks > 0 ? cfg->chunks : env_int("MRTS_CHUNKS", 0);
    if (K <= 0) {
        // a chunk should fill the GPU a few times over: ~20 games per SM are resident at once
        K = P.n_games / (e->n_sm * 40);
        if (K > 4) K = 4;
    }
    if (e->pipeline != MRTS_PIPELINE_SPLIT || K < 1) K = 1;
    if (K > MAX_CHUNKS) K = MAX_CHUNKS;
    if (K > P.n_games) K = P.n_games;
    e->K = K;
    e->exp_blocks = e->n_sm * (env_int("MRTS_EXPAND_BLOCKS_PER_SM", 1) > 0 ? env_int("MRTS_EXPAND_BLOCKS_PER_SM", 1) : 1);
    std::vector<int32_t> light, order0((size_t)4 * P.n_games, 0), ctr((size_t)10 * K, 0);
    for (int q = 0; q <= K; q++) e->chunk_g0[q] = (int)((long long)P.n_games * q / K);
    for (int q = 0; q < K; q++) {
        e->chunk_light_off[q] = (int)light.size();
        int nh = 0;
        for (int g = e->chunk_g0[q]; g < e->chunk_g0[q + 1]; g++) {
            const bool heavy = ai_plans(S.game_ai[g]) || ai_plans(S.game_ai1[g]);
            if (heavy) order0[(size_t)3 * P.n_games + e->chunk_g0[q] + nh++] = g;  // class 3 (no idle bot unit known yet) of buffer 0
            else light.push_back(g);
        }
        e->chunk_nheavy[q] = nh;
        e->chunk_nlight[q] = (int)light.size() - e->chunk_light_off[q];
        ctr[4 * q + 3] = nh;
    }
    CREATE_RC(dalloc(e, &e->d_light, light.size()));
    if (!light.empty()) CREATE_CK(cudaMemcpy(e->d_light, light.data(), light.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    CREATE_RC(dalloc(e, &e->d_order[0], order0.size()));
    CREATE_RC(dalloc(e, &e->d_order[1], order0.size()));
    CREATE_CK(cudaMemcpy(e->d_order[0], order0.data(), order0.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    CREATE_RC(dalloc(e, &e->d_ctr, ctr.size()));
    CREATE_CK(cudaMemcpy(e->d_ctr, ctr.data(), ctr.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    CREATE_RC(dalloc(e, &e->d_ovf_list, (size_t)P.n_games));
    if (e->pipeline == MRTS_PIPELINE_SPLIT) {
        CREATE_RC(dalloc(e, &e->d_pobs, (size_t)e->N * HW, false));
        CREATE_RC(dalloc(e, &e->d_ment, (size_t)e->N * ment_stride_words(P.cap), false));
        if (K > 1) {
            int lo = 0, hi = 0;
            CREATE_CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));  // lo = least priority (numerically greatest)
            // the store stream goes first whenever an SM frees a slot: it needs few warps and the rules fill the rest
            CREATE_CK(cudaStreamCreateWithPriority(&e->s_exp, cudaStreamNonBlocking, hi));
            CREATE_CK(cudaStreamCreateWithPriority(&e->s_rules[0], cudaStreamNonBlocking, lo));
            CREATE_CK(cudaStreamCreateWithPriority(&e->s_rules[1], cudaStreamNonBlocking, lo));
            CREATE_CK(cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming));
            CREATE_CK(cudaEventCreateWithFlags(&e->ev_join, cudaEventDisableTiming));
            for (int q = 0; q < K; q++) CREATE_CK(cudaEventCreateWithFlags(&e->ev_rules[q], cudaEventDisableTiming));
        }
    }
    e->diag_skip = env_int("MRTS_DIAG_SKIP", 0);
    if (env_int("MRTS_OVERLAP", 1)) {
        bool both = false;
        for (int q = 0; q < K; q++) both = both || (e->chunk_nheavy[q] > 0 && e->chunk_nlight[q] > 0);
        if (both) {
            int lo = 0, hi = 0;
            CREATE_CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
            CREATE_CK(cudaStreamCreateWithPriority(&e->s_light, cudaStreamNonBlocking, lo));
            CREATE_CK(cudaEventCreateWithFlags(&e->ev_lfork, cudaEventDisableTiming));
            CREATE_CK(cudaEventCreateWithFlags(&e->ev_ljoin, cudaEventDisableTiming));
        }
    }
    // start every game from its template so that the state is defined before the first reset()
    CREATE_RC(launch(e, MODE_RESET, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr));
    CREATE_CK(cudaDeviceSynchronize());
#undef CREATE_CK
#undef CREATE_RC
    *out = e;
    return MRTS_OK;
}

void mrts_destroy(mrts_env* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    cudaDeviceSynchronize();
    {
        auto& h = e->hio;
        for (void* p : {(void*)h.h_pact, (void*)h.h_pobs, (void*)h.h_pmask, (void*)h.h_rew, (void*)h.h_done})
            if (p) cudaFreeHost(p);
        if (h.h_pairs) cudaFreeHost(h.h_pairs);
        for (uint32_t* p : h.d_pairs)
            if (p) cudaFree(p);
        if (h.h_delta) cudaFreeHost(h.h_delta);
        if (h.d_delta) cudaFree(h.d_delta);
        for (uint32_t* p : h.d_shadow)
            if (p) cudaFree(p);
        for (void* p : {(void*)h.d_pact, (void*)h.d_pobs, (void*)h.d_pmask, (void*)h.d_rew, (void*)h.d_done})
            if (p) cudaFree(p);
        if (h.ev_obs) cudaEventDestroy(h.ev_obs);
        if (h.ev_user) cudaEventDestroy(h.ev_user);
        if (h.stream) cudaStreamDestroy(h.stream);
        mrts_host::pool_destroy(h.pool);
    }
    for (cudaEvent_t ev : {e->ev_fork, e->ev_join, e->ev_last})
        if (ev) cudaEventDestroy(ev);
    for (cudaEvent_t ev : e->ev_rules)
        if (ev) cudaEventDestroy(ev);
    for (cudaEvent_t ev : {e->ev_lfork, e->ev_ljoin})
        if (ev) cudaEventDestroy(ev);
    for (cudaStream_t s : {e->s_rules[0], e->s_rules[1], e->s_exp, e->s_light})
        if (s) cudaStreamDestroy(s);
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

// ---- packed form (the rules half of a step on its own) and its expansion
int mrts_packed_sizes(const mrts_env* e, int64_t* obs_words, int64_t* mask_entry_words) {
    if (!e) return fail(MRTS_EINVAL, "null env");
    if (obs_words) *obs_words = (int64_t)e->N * e->P.HW;
    if (mask_entry_words) *mask_entry_words = (int64_t)e->N * (int64_t)ment_stride_words(e->P.cap);
    return MRTS_OK;
}
int mrts_reset_packed(mrts_env* e, uint32_t* obs_words, uint32_t* mask_entries, void* stream) {
    if (!e) return fail(MRTS_EINVAL, "null env");
    e->cur_pobs = obs_words;
    e->cur_ment = mask_entries;
    return launch(e, MODE_RESET, nullptr, nullptr, nullptr, nullptr, nullptr, stream);
}
int mrts_step_packed(mrts_env* e, const mrts_packed_io* io, void* stream) {
    if (!e || !io) return fail(MRTS_EINVAL, "null argument");
    if (!io->actions && !e->all_scripted) return fail(MRTS_EINVAL, "actions is null");
    e->cur_pobs = io->obs_words;
    e->cur_ment = io->mask_entries;
    return launch(e, MODE_STEP, io->actions, nullptr, nullptr, io->reward, io->done, stream, io->next_actions, io->seed);
}
int mrts_expand(mrts_env* e, const uint32_t* obs_words, const uint32_t* mask_entries, void* obs, void* mask, void* stream) {
    if (!e) return fail(MRTS_EINVAL, "null env");
    if ((obs && !obs_words) || (mask && !mask_entries)) return fail(MRTS_EINVAL, "packed input missing");
    if (!obs && !mask) return MRTS_OK;
    if (!expand_supported(e->P.HW, e->P.planes, e->P.elem, obs, mask) || ((((uintptr_t)obs_words) | ((uintptr_t)mask_entries)) & 15))
        return fail(MRTS_EINVAL, "mrts_expand needs H*W*element_bytes to be a multiple of 16 and 16-byte aligned tensors");
    CK(cudaSetDevice(e->device));
    const cudaStream_t st = (cudaStream_t)stream;
    int rc = order_after_last(e, st);
    if (rc) return rc;
    int32_t* ctr = e->d_ctr + 5 * e->K;
    CK(cudaMemsetAsync(ctr, 0, sizeof(int32_t), st));
    ExpandParams X;
    X.pobs = obs_words;
    X.ment = mask_entries;
    X.obs = obs;
    X.mask = mask;
    X.env_lo = 0;
    X.env_hi = e->N;
    X.HW = e->P.HW;
    X.planes = e->P.planes;
    X.elem = e->P.elem;
    X.cap = e->P.cap;
    X.empty_word = obs_empty_word(e->P.partial);
    X.counter = ctr;
    int blocks = (e->N + 7) / 8;
    if (blocks > e->exp_blocks) blocks = e->exp_blocks;
    CK(launch_expand(X, blocks, st));
    return mark_last(e, st);
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
    // one pool per process would oversubscribe the cores when every rank of a node builds its own:
    // the threads this rank may use = visible cores / ranks on the node (LOCAL_WORLD_SIZE, torchrun)
    int threads = env_int("MRTS_HOST_THREADS", 0);
    if (threads <= 0) {
        const int ranks = env_int("LOCAL_WORLD_SIZE", 1);
        threads = mrts_host::visible_cores() / (ranks > 0 ? ranks : 1);
        if (threads < 1) threads = 1;
    }
    h.pool = mrts_host::pool_create(threads);
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
    if (S) S->on_device = false;  // this path keeps its shadows on the host: the device shadow no longer matches the arrays
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

constexpr size_t DELTA_ENTRY_BYTES = DELTA_ENTRY_WORDS * sizeof(uint32_t);
static double now_s() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

// buffers of the delta transport: the shared result buffer on first use, one shadow per slot
static int hostio_delta_init(mrts_env* e, int slot) {
    auto& h = e->hio;
    CK(cudaSetDevice(e->device));
    const size_t N = (size_t)e->N, cells = N * (size_t)e->P.HW;
    if (!h.d_delta) {
        auto up = [](size_t v) { return (v + 255) / 256 * 256; };
        h.off_rew = 256;
        h.off_done = h.off_rew + up(N * 48);
        h.off_bitmap = h.off_done + up(N);
        h.off_entries = h.off_bitmap + up((cells + 31) / 32 * 4);
        h.delta_cap = cells;  // every cell may change (a reset of all games): the list can never overflow
        const size_t bytes = h.off_entries + h.delta_cap * DELTA_ENTRY_BYTES;
        CK(cudaMalloc((void**)&h.d_delta, bytes));
        CK(cudaMallocHost((void**)&h.h_delta, bytes));
        CK(cudaMemset(h.d_delta, 0, h.off_entries));
        memset(h.h_delta, 0, h.off_entries);
    }
    if (!h.h_pairs) {
        h.pairs_cap = cells / 8 > 65536 ? cells / 8 : 65536;
        CK(cudaMallocHost((void**)&h.h_pairs, h.pairs_cap * 8));
        CK(cudaMalloc((void**)&h.d_pairs[0], h.pairs_cap * 8));
        CK(cudaMalloc((void**)&h.d_pairs[1], h.pairs_cap * 8));
    }
    if (!h.d_shadow[slot]) CK(cudaMalloc((void**)&h.d_shadow[slot], cells * 16));
    return MRTS_OK;
}

// Slot form with both arrays given: the device compares this step's words with the slot's shadow and only the cells that
// differ cross PCIe (one copy: count, rewards, dones, source bitmap, entries); the host pool rewrites those cells.
// The first call on a slot (or after its arrays changed) expands everything and seeds the shadow.
static int hostio_finish_slot(mrts_env* e, void* h_obs, void* h_mask, double* h_reward, uint8_t* h_done, int slot) {
    auto& h = e->hio;
    auto& S = h.slots[slot];
    const size_t N = (size_t)e->N, HW = (size_t)e->P.HW, cells = N * HW;
    const bool delta = S.on_device && S.obs == h_obs && S.mask == h_mask;
    uint8_t* const db = (uint8_t*)h.d_delta;
    uint8_t* const hb = (uint8_t*)h.h_delta;
    CK(cudaMemsetAsync(h.d_delta, 0, 256, h.stream));
    DeltaParams D;
    D.pobs = h.d_pobs;
    D.pmask = h.d_pmask;
    D.shadow = h.d_shadow[slot];
    D.count = h.d_delta;
    D.bitmap = (uint32_t*)(db + h.off_bitmap);
    D.entries = (uint32_t*)(db + h.off_entries);
    D.cells = (long long)cells;
    D.cap = (uint32_t)h.delta_cap;
    D.init = delta ? 0 : 1;
    CK(launch_delta(D, e->n_sm, h.stream));
    e->last_launches++;
    const double t0 = now_s();
    double t1, t2;
    if (!delta) {
        CK(cudaMemcpyAsync(h.h_pobs, h.d_pobs, cells * 4, cudaMemcpyDeviceToHost, h.stream));
        CK(cudaEventRecord(h.ev_obs, h.stream));
        CK(cudaMemcpyAsync(h.h_pmask, h.d_pmask, cells * 12, cudaMemcpyDeviceToHost, h.stream));
        CK(cudaMemcpyAsync(hb, db, h.off_entries, cudaMemcpyDeviceToHost, h.stream));
        CK(cudaEventSynchronize(h.ev_obs));
        mrts_host::expand_obs(h.pool, h.h_pobs, N, (int)HW, e->P.planes, e->P.elem, h_obs);
        CK(cudaStreamSynchronize(h.stream));
        t1 = now_s();
        mrts_host::expand_mask(h.pool, h.h_pmask, N, (int)HW, e->P.elem, h_mask);
        S.obs = h_obs, S.mask = h_mask, S.on_device = true;
        S.pobs.clear(), S.pmask.clear();
        t2 = now_s();
    } else {
        size_t guess = h.last_count + h.last_count / 4 + 1024;  // the count moves slowly from step to step; a short second copy covers a miss
        if (guess > h.delta_cap) guess = h.delta_cap;
        CK(cudaMemcpyAsync(hb, db, h.off_entries + guess * DELTA_ENTRY_BYTES, cudaMemcpyDeviceToHost, h.stream));
        CK(cudaStreamSynchronize(h.stream));
        const size_t count = h.h_delta[0];
        if (count > h.delta_cap) return fail(MRTS_ECUDA, "delta list overflow");
        if (count > guess) {
            CK(cudaMemcpyAsync(hb + h.off_entries + guess * DELTA_ENTRY_BYTES, db + h.off_entries + guess * DELTA_ENTRY_BYTES, (count - guess) * DELTA_ENTRY_BYTES, cudaMemcpyDeviceToHost, h.stream));
            CK(cudaStreamSynchronize(h.stream));
        }
        t1 = now_s();
        mrts_host::patch_cells(h.pool, (const uint32_t*)(hb + h.off_entries), count, e->P.planes, e->P.elem, h_obs, h_mask);
        h.last_count = count;
        h.t_phase[5] += (double)count, h.t_phase[6] += (double)(h.off_entries + (count > guess ? count : guess) * DELTA_ENTRY_BYTES);
        t2 = now_s();
    }
    h.have_mask = true;
    h.bitmap_valid = true;
    if (h_reward) memcpy(h_reward, hb + h.off_rew, N * 48);
    if (h_done) memcpy(h_done, hb + h.off_done, N);
    h.t_phase[2] += t1 - t0, h.t_phase[3] += now_s() - t1;
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
    const bool slot_path = slot >= 0 && h_obs && h_mask;
    if (slot_path && (rc = hostio_delta_init(e, slot))) return rc;
    rc = launch(e, MODE_RESET, nullptr, nullptr, nullptr, nullptr, nullptr, e->hio.stream);
    if (rc) return rc;
    if (slot_path) return hostio_finish_slot(e, h_obs, h_mask, nullptr, nullptr, slot);
    e->hio.bitmap_valid = false;
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
    const bool slot_path = slot >= 0 && h_obs && h_mask;
    if (slot_path && (rc = hostio_delta_init(e, slot))) return rc;
    const double t0 = now_s();
    const uint32_t* const bitmap = (const uint32_t*)((const uint8_t*)h.h_delta + h.off_bitmap);
    bool sparse = false;
    size_t n_pairs = 0;
    if (h.bitmap_valid && h.h_pairs && env_int("MRTS_HOST_SPARSE_ACTIONS", 1)) {
        n_pairs = mrts_host::pack_actions_sparse(h.pool, h_actions, action_elem_bytes, bitmap, N * HW, h.h_pairs, h.pairs_cap);
        sparse = n_pairs <= h.pairs_cap;
    }
    if (!sparse) {
        if (h.bitmap_valid) mrts_host::pack_actions_bitmap(h.pool, h_actions, action_elem_bytes, bitmap, N * HW, h.h_pact);
        else mrts_host::pack_actions(h.pool, h_actions, action_elem_bytes, h.h_pmask, N, (int)HW, h.h_pact);
    }
    const double t1 = now_s();
    if (sparse) {
        // the cells of the previous list read as NOOP again (a row the reference never forwards), then this step's rows
        if (!h.pact_clean) {
            CK(cudaMemsetAsync(h.d_pact, 0, N * HW * 2, h.stream));
            h.pact_clean = true;
        } else {
            CK(launch_scatter_actions(h.d_pairs[h.pairs_cur], (int)h.pairs_prev_n, h.d_pact, 1, h.stream));
        }
        h.pairs_cur ^= 1;
        CK(cudaMemcpyAsync(h.d_pairs[h.pairs_cur], h.h_pairs, n_pairs * 8, cudaMemcpyHostToDevice, h.stream));
        CK(launch_scatter_actions(h.d_pairs[h.pairs_cur], (int)n_pairs, h.d_pact, 0, h.stream));
        h.pairs_prev_n = n_pairs;
    } else {
        CK(cudaMemcpyAsync(h.d_pact, h.h_pact, N * HW * 2, cudaMemcpyHostToDevice, h.stream));
        h.pact_clean = false;
        h.pairs_prev_n = 0;
    }
    e->cur_pact = h.d_pact;
    e->cur_pobs = h.d_pobs;
    e->cur_pmask = h.d_pmask;
    double* const d_rew = slot_path ? (double*)((uint8_t*)h.d_delta + h.off_rew) : h.d_rew;
    uint8_t* const d_done = slot_path ? (uint8_t*)h.d_delta + h.off_done : h.d_done;
    // `actions` is unused when the packed rows are given; pass a valid non-null pointer for the argument check
    rc = launch(e, MODE_STEP, (const int32_t*)h.d_pact, nullptr, nullptr, d_rew, d_done, h.stream);
    if (rc) return rc;
    h.t_phase[0] += t1 - t0, h.t_phase[1] += now_s() - t1, h.t_phase[7] += 1;
    h.t_phase[4] += sparse ? (double)(n_pairs * 8) : (double)(N * HW * 2);
    if (slot_path) return hostio_finish_slot(e, h_obs, h_mask, h_reward, h_done, slot);
    h.bitmap_valid = false;
    return hostio_finish(e, h_obs, h_mask, h_reward, h_done, slot);
}

int mrts_launches_per_step(const mrts_env* e) {
    if (!e) return MRTS_EINVAL;
    return e->last_launches;
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
int mrts_host_pack_actions_bitmap(const void* actions, int action_elem_bytes, const uint32_t* bitmap, int64_t cells, uint16_t* out) {
    if (!actions || !bitmap || !out || cells < 0 || (action_elem_bytes != 4 && action_elem_bytes != 8)) return fail(MRTS_EINVAL, "bad argument");
    mrts_host::Pool* p = mrts_host::pool_create(0);
    mrts_host::pack_actions_bitmap(p, actions, action_elem_bytes, bitmap, (size_t)cells, out);
    mrts_host::pool_destroy(p);
    return MRTS_OK;
}
int mrts_host_patch_cells(const uint32_t* entries, int64_t count, int planes, int elem_bytes, void* obs, void* mask) {
    if (!entries || count < 0 || (elem_bytes != 4 && elem_bytes != 1) || planes < 1 || planes > 31) return fail(MRTS_EINVAL, "bad argument");
    mrts_host::Pool* p = mrts_host::pool_create(0);
    mrts_host::patch_cells(p, entries, (size_t)count, planes, elem_bytes, obs, mask);
    mrts_host::pool_destroy(p);
    return MRTS_OK;
}
int mrts_host_timing(mrts_env* e, double* out8, int reset) {
    if (!e || !out8) return fail(MRTS_EINVAL, "null argument");
    for (int k = 0; k < 8; k++) out8[k] = e->hio.t_phase[k];
    if (reset)
        for (double& t : e->hio.t_phase) t = 0;
    return MRTS_OK;
}
int mrts_host_threads(const mrts_env* e) { return e && e->hio.pool ? mrts_host::pool_threads(e->hio.pool) : 0; }

int mrts_num_envs(const mrts_env* e) { return e ? e->N : MRTS_EINVAL; }
int mrts_num_games(const mrts_env* e) { return e ? e->P.n_games : MRTS_EINVAL; }
int mrts_planes(const mrts_env* e) { return e ? e->P.planes : MRTS_EINVAL; }
int mrts_unit_capacity(const mrts_env* e) { return e ? e->P.cap : MRTS_EINVAL; }
int mrts_threads_per_game(const mrts_env* e) { return e ? (e->pipeline == MRTS_PIPELINE_SPLIT ? 32 : e->threads) : MRTS_EINVAL; }
int mrts_smem_bytes(const mrts_env* e) { return e ? e->smem[1] : MRTS_EINVAL; }
int mrts_pipeline(const mrts_env* e) { return e ? e->pipeline : MRTS_EINVAL; }
int mrts_chunks(const mrts_env* e) { return e ? e->K : MRTS_EINVAL; }
int mrts_kernel_info(const mrts_env* e, int heavy, int* regs, int* games_per_block, int* smem_per_game) {
    if (!e || heavy < 0 || heavy > 1) return fail(MRTS_EINVAL, "bad argument");
    if (regs) *regs = heavy ? heavy_warp_regs(e->regs[1]) : light_warp_regs(e->regs[0]);
    if (games_per_block) *games_per_block = e->gpb[heavy];
    if (smem_per_game) *smem_per_game = e->smem[heavy];
    return MRTS_OK;
}

int mrts_export_state(mrts_env* e, int game, int32_t* hdr5, int32_t* rows12, int cap_rows, void* stream) {
    if (!e || game < 0 || game >= e->P.n_games) return fail(MRTS_EINVAL, "bad game index");
    CK(cudaSetDevice(e->device));
    CK(cudaDeviceSynchronize());
    (void)stream;
    std::vector<uint8_t> blob(e->P.game_stride);
    CK(cudaMemcpy(blob.data(), e->P.games + (size_t)game * e->P.game_stride, blob.size(), cudaMemcpyDeviceToHost));
    export_blob(blob.data(), hdr5, rows12, cap_rows);
    return ((const GameHdr*)blob.data())->nunits;
}

int mrts_import_state(mrts_env* e, int game, const int32_t* hdr5, const int32_t* rows12, int n, const int32_t* issue_order, void* stream) {
    if (!e || game < 0 || game >= e->P.n_games) return fail(MRTS_EINVAL, "bad game index");
    if (!hdr5 || (n > 0 && !rows12) || n < 0) return fail(MRTS_EINVAL, "null argument");
    if (n > e->P.cap) return fail(MRTS_EINVAL, "more units than the capacity the env was created with (mrts_unit_capacity)");
    CK(cudaSetDevice(e->device));
    CK(cudaDeviceSynchronize());
    (void)stream;
    std::vector<uint8_t> blob(e->P.game_stride, 0);
    std::string err;
    if (!import_blob(blob.data(), hdr5, rows12, n, issue_order, e->P.W, e->P.H, err)) return fail(MRTS_EINVAL, err);
    CK(cudaMemcpy(e->P.games + (size_t)game * e->P.game_stride, blob.data(), blob.size(), cudaMemcpyHostToDevice));
    return MRTS_OK;
}

static int debug_launch(mrts_env* e, int game, int op, const int32_t* rows7, int n, int* gameover, void* stream) {
    if (!e || game < 0 || game >= e->P.n_games) return fail(MRTS_EINVAL, "bad game index");
    CK(cudaSetDevice(e->device));
    const cudaStream_t st = (cudaStream_t)stream;
    int rc = order_after_last(e, st);
    if (rc) return rc;
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
        CK(cudaMemcpyAsync(e->d_rows, rows7, (size_t)n * 7 * sizeof(int32_t), cudaMemcpyHostToDevice, st));
        A.rows7 = e->d_rows;
    }
    EnvParams F = e->P;
    F.scap = F.cap;
    CK(launch_debug(F, A, e->smem_full, st));
    CK(cudaStreamSynchronize(st));
    if (op == 1 && gameover) CK(cudaMemcpy(gameover, e->d_flag, sizeof(int), cudaMemcpyDeviceToHost));
    return mark_last(e, st);
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
    int rc;
    if (!e->d_ended) {
        if ((rc = dalloc(e, &e->d_ended, (size_t)e->P.n_games)) || (rc = dalloc(e, &e->d_cycle_pos, 1)) ||
            (rc = dalloc(e, &e->d_cycle_list, (size_t)e->P.n_games + 1)))
            return rc;
    }
    if ((rc = dalloc(e, &e->d_cycle, (size_t)n, false))) return rc;
    CK(cudaMemcpy(e->d_cycle, template_idx, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice));
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
