// One game's pass through a launch: stage the state blob in (shared) memory, run the rules on
// the group's first warp, then have the whole group write the outputs and put the blob back.
// Shared by the CUDA kernels (mrts_rules_*.cu) and the host debug build (tests/hostsim).
//
// Replaces JNIGridnetVecClient.reset / gameStep / getMasks (J/tests/JNIGridnetVecClient.java:
// 179-211, 213-297, 307-316) for one client; env index layout as in :117-124 (self-play game g
// = envs 2g and 2g+1, then the bot envs).
//
// Outputs come in two forms.  PACKED (the split pipeline and the host transport): one 32-bit word
// per cell for the observation (bit k = plane k) and one 16-byte entry per unit that may act for the
// mask (cell, 78 bits), which mrts_expand.cuh widens into the caller's tensors with bulk async
// stores while the next slab chunk's rules run.  DIRECT: the same values streamed straight into the
// caller's tensors by this pass (one launch; used for small slabs).
#pragma once
#include "mrts_encode.h"

namespace mrts {

enum { MODE_STEP = 0, MODE_RESET = 1, MODE_OBSERVE = 2 };

struct EnvParams {
    int32_t num_selfplay, num_bot;  // player views, as in the reference constructor
    int32_t n_games;                // num_selfplay / 2 + num_bot
    int32_t W, H, HW, cap, planes;  // cap = unit capacity of a game blob in HBM
    int32_t scap;                   // unit capacity staged in shared memory by this launch (<= cap); games that may exceed it go to `ovf`
    // Launch order.  A launch covers `nslots` games: slot k -> game0 + k, or through `order` -- one list, or (order_cnt)
    // four class lists of order_stride entries each, walked class after class.  A bot game's time is dominated by
    // one path search per idle bot unit (a heavy tail: several searches in one frame every ~10 frames) and that
    // count is known when its pass ends, so every STEP pass of a heavy game appends the game to the class list
    // of the NEXT step (order_next / order_cnt_next): most idle bot units first, otherwise a launch ends with a
    // few long games running alone.
    const int32_t* order;
    const int32_t* order_cnt;
    int32_t* order_next;
    int32_t* order_cnt_next;
    int32_t order_stride, nslots;
    int32_t* ovf_cnt;               // games deferred to the full-capacity launch: their count and indices; null = no deferral
    int32_t* ovf_list;
    int32_t partial, none_mode, max_steps;
    int32_t has_bot;                // any game with an AbstractionLayerAI bot: blobs carry the abstract-action table
    int32_t any_bot;                // any bot env at all: the bot's PlayerAction (pa2) is part of the shared-memory plan
    int32_t game_stride;            // bytes per game blob
    uint8_t* games;                 // n_games blobs
    const MapTpl* maps;
    const uint16_t* grid_pool;      // wall grids of all templates
    const Unit* unit_pool;          // initial unit lists of all templates
    const int32_t* game_map;        // template id per game
    const int8_t* game_ai;          // AI_* per game (AI_NONE = self-play)
    const int8_t* game_ai1;         // AI_* playing player 0 (bot-vs-bot games), AI_NONE = an agent; may be null
    // per-launch
    int32_t mode;
    int32_t game0, game1;           // this launch covers the slots / games [game0, game1)
    int32_t elem;                   // bytes per obs/mask element: 4 (int32, reference dtype) or 1 (uint8)
    int32_t bulk_out;               // direct mode, warp kernels: the mask goes through the copy engine (zero tile + staged rows, mrts_bulk.cuh)
    const int32_t* actions;         // (N, HW, 7) int32
    void* obs;                      // (N, H, W, planes) or null
    void* mask;                     // (N, HW, 78) or null
    double* reward;                 // (N, 6) or null
    uint8_t* done;                  // (N,) or null
    // packed outputs
    const uint16_t* pact;           // (N, HW) action rows packed by Sim::pack_action, replaces `actions` when non-null
    uint32_t* pobs;                 // (N, HW) observation words (bit k = plane k) or null
    uint32_t* pmask;                // (N, HW, 3) mask words (bit k of the 78), dense, or null (host transport)
    uint32_t* ment;                 // (N, 1 + cap, 4) mask entries: [count,0,0,0] then (cell, w0, w1, w2) in ascending cell order, or null
    int32_t* next_actions;          // (N, HW, 7) or null: masked-uniform sample for the resulting state
    uint64_t seed;                  // sampler key
    uint32_t counter;               // launch counter (sampler stream position)
    uint8_t* ended;                 // cycle_maps: per game, 1 when the step ended its episode (read by the map-rotation pass)
    long long* timing;              // profiling aid, normally null: per game 16 values {rules cycles, total cycles, St::tp[0..7], start ns, SM id, St::tp[8..11]}
};

MRTS_HD int align16(int v) { return (v + 15) & ~15; }
MRTS_HD int blob_bytes(int cap, int has_bot) { return 32 + cap * 32 + (has_bot ? align16(cap * 12) : 0); }
MRTS_HD size_t ment_stride_words(int cap) { return (size_t)(1 + cap) * 4; }

struct Layout {
    int blob, grid, words, umask, slots, pa, pa2, bits, astar, misc, total, nbits_words;
};
// Shared-memory plan of one game.  The rules scratch (PlayerActions, reservation bitmaps, A*
// arrays) is dead once the rules are done, the output scratch (observation words, per-unit
// legal-action sets) is only written after that: the two share one region.
MRTS_HD Layout make_layout(int W, int HW, int cap, int has_bot, int any_bot) {
    Layout L;
    int o = 0;
    L.misc = o;  // St
    o += align16((int)sizeof(St));
    L.blob = o;
    o += blob_bytes(cap, has_bot);
    L.grid = o;
    o += align16(HW * 2);
    const int scratch = o;
    // ---- rules scratch
    L.pa = o;
    o += align16(cap * (int)sizeof(PAct));
    L.pa2 = o;
    o += any_bot ? align16(cap * (int)sizeof(PAct)) : 0;
    L.nbits_words = (HW + 2 * W + 31) / 32;
    L.bits = o;
    o += align16(3 * L.nbits_words * 4);
    L.astar = o;
    o += has_bot ? astar_bytes(W, HW) : 0;
    const int end_rules = o;
    // ---- output scratch (aliases the rules scratch)
    o = scratch;
    // The observation words (one view at a time) are dead when the last view's observation is stored; the staging
    // buffers of the bulk mask writer (bulk::STAGE_BYTES, mrts_bulk.cuh) are first written after that: same bytes.
    // At 16x16 this takes a light game from 6.3 to 5.3 KB: 8 instead of 7 resident blocks of four games per SM.
    L.words = o;
    L.slots = o;
    o += align16(HW * 4) > 2 * 1248 ? align16(HW * 4) : 2 * 1248;
    L.umask = o;
    o += align16(cap * 12);
    L.total = o > end_rules ? o : end_rules;
    return L;
}

MRTS_HD Layout layout_of(const EnvParams& P) { return make_layout(P.W, P.HW, P.scap, P.has_bot, P.any_bot); }

// point a Sim at one game's shared-memory regions
MRTS_HD void bind_sim(Sim& s, const EnvParams& P, const Layout& L, uint8_t* mem, int g) {
    s.W = P.W;
    s.H = P.H;
    s.HW = P.HW;
    s.cap = P.scap;
    s.none_mode = P.none_mode;
    s.partial = P.partial;
    s.max_steps = P.max_steps;
    s.ai = P.game_ai[g];
    s.ai1 = P.game_ai1 ? (int)P.game_ai1[g] : (int)AI_NONE;
    s.hdr = (GameHdr*)(mem + L.blob);
    s.u = (Unit*)(mem + L.blob + 32);
    s.aa = (AbstractAct*)(mem + L.blob + 32 + P.scap * 32);
    s.grid = (uint16_t*)(mem + L.grid);
    s.pa = (PAct*)(mem + L.pa);
    s.pa2 = (PAct*)(mem + L.pa2);
    s.nbits_words = L.nbits_words;
    s.bitsA = (uint32_t*)(mem + L.bits);
    s.bitsB = s.bitsA + L.nbits_words;
    s.bitsC = s.bitsB + L.nbits_words;
    s.astar = (int16_t*)(mem + L.astar);
    s.lane = 0;
    s.nl = 1;
    s.st = (St*)(mem + L.misc);
}

template <class T>
MRTS_HD void copy16(void* dst, const void* src, int bytes, int tid, int nth) {  // bytes multiple of 16, both 16-aligned
    const T* s = (const T*)src;
    T* d = (T*)dst;
    const int n = bytes / (int)sizeof(T);
#pragma unroll 1
    for (int i = tid; i < n; i += nth) d[i] = s[i];
}
struct alignas(16) Q16 {
    uint32_t a, b, c, d;
};
// the game blob <-> shared memory: the blob is re-read every step, its lines are kept in the L2 (evict-last) while the
// write-once output stream passes through (evict-first, mrts_bulk.cuh)
MRTS_HD void blob_load(void* smem_dst, const void* gsrc, int bytes, int tid, int nth) {
#if defined(__CUDA_ARCH__)
    const uint64_t pol = bulk::policy_evict_last();
    const int n = bytes / 16;
#pragma unroll 1
    for (int i = tid; i < n; i += nth) ((uint4*)smem_dst)[i] = bulk::ldg_hint((const uint4*)gsrc + i, pol);
#else
    copy16<Q16>(smem_dst, gsrc, bytes, tid, nth);
#endif
}
MRTS_HD void blob_store(void* gdst, const void* smem_src, int bytes, int tid, int nth) {
#if defined(__CUDA_ARCH__)
    const uint64_t pol = bulk::policy_evict_last();
    const int n = bytes / 16;
#pragma unroll 1
    for (int i = tid; i < n; i += nth) bulk::stg_hint((uint4*)gdst + i, ((const uint4*)smem_src)[i], pol);
#else
    copy16<Q16>(gdst, smem_src, bytes, tid, nth);
#endif
}

MRTS_HD long long cycles_now() {
#if defined(__CUDA_ARCH__)
    return clock64();
#else
    return 0;
#endif
}
MRTS_HD void group_add(int* p, int v) {
#if defined(__CUDA_ARCH__)
    atomicAdd(p, v);
#else
    *p += v;
#endif
}
// queue game g for the full-capacity launch (mrts_overflow_kernel)
MRTS_HD void defer_game(const EnvParams& P, int g) {
#if defined(__CUDA_ARCH__)
    const int k = atomicAdd(P.ovf_cnt, 1);
    P.ovf_list[k] = g;
#else
    (void)P;
    (void)g;
#endif
}

// idle mobile bot units after a pass -> launch class of the game's next step: >= 5, 3-4, 1-2, none
MRTS_HD int order_class(int nidle) { return nidle > 4 ? 0 : (nidle > 2 ? 1 : (nidle > 0 ? 2 : 3)); }
MRTS_HD int order_append(int32_t* cnt) {
#if defined(__CUDA_ARCH__)
    return atomicAdd(cnt, 1);
#else
    return (*cnt)++;
#endif
}

// is game g one whose rules need the HEAVY instantiation (a scripted player built on AbstractionLayerAI)?
MRTS_HD bool game_is_heavy(const EnvParams& P, int g) { return ai_plans(P.game_ai[g]) || (P.game_ai1 && ai_plans(P.game_ai1[g])); }

// SYNC: functor executing a barrier over the game's thread group.  HEAVY: see Sim::step_game.
// `blk`: shared memory of the whole block (the bulk writer's tile of zeros lives at its start), null when unused.
template <int HEAVY, class SYNC>
MRTS_HD void game_pass(const EnvParams& P, int g, uint8_t* mem, int tid, int nth, SYNC sync, uint8_t* blk = nullptr) {
    const long long t_start = P.timing ? cycles_now() : 0;
#if defined(__CUDA_ARCH__)
    if (P.timing && tid == 0) {
        unsigned long long ns;
        unsigned sm;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(ns));
        asm volatile("mov.u32 %0, %smid;" : "=r"(sm));
        P.timing[16 * g + 10] = (long long)ns;
        P.timing[16 * g + 11] = (long long)sm;
    }
#endif
    const Layout L = layout_of(P);
    Sim s;
    bind_sim(s, P, L, mem, g);
    s.nl = nth < 32 ? nth : 32;  // lanes of the rules phase: the group's first warp
    s.lane = tid & 31;
    const bool rules_lane = tid < s.nl;
    const bool plans = HEAVY && s.plans();
    uint32_t* words = (uint32_t*)(mem + L.words);
    uint32_t* umask = (uint32_t*)(mem + L.umask);

    const MapTpl tpl = P.maps[P.game_map[g]];
    const uint16_t* grid0 = P.grid_pool + tpl.grid_off;
    uint8_t* gblob = P.games + (size_t)g * P.game_stride;
    const bool selfplay = s.ai == AI_NONE;
    const int nsp_games = P.num_selfplay / 2;
    const int env0 = selfplay ? 2 * g : P.num_selfplay + (g - nsp_games);
    const int nviews = selfplay ? 2 : 1;

    // ---------------- load: template (reset) or the stored blob
    // keep_steps: an explicit reset() leaves envSteps alone (JNIGridnetVecClient.reset, :179-211, never touches
    // it); only the terminal / max_steps path of gameStep zeroes it (:245-266, :270-286)
    auto load_template = [&](int steps) {
        if (tid == 0) {
            GameHdr h;
            h.nunits = tpl.nunits;
            h.time = 0;
            h.res[0] = tpl.res[0];
            h.res[1] = tpl.res[1];
            h.steps = steps;
            h.next_seq = 0;
            h.next_uid = (uint16_t)tpl.nunits;
            h.n_aa = 0;
            h.pad = 0;
            *s.hdr = h;
        }
        copy16<Q16>(s.u, P.unit_pool + tpl.unit_off, tpl.nunits * 32, tid, nth);
        copy16<uint32_t>(s.grid, grid0, align16(P.HW * 2), tid, nth);
    };
    auto scatter_grid = [&]() {
        const int n = s.hdr->nunits;
        for (int i = tid; i < n; i += nth) s.grid[s.u[i].x + s.u[i].y * P.W] = (uint16_t)i;
    };
    if (P.mode == MODE_RESET) {
        load_template(((const GameHdr*)gblob)->steps);
    } else {
        const GameHdr* gh = (const GameHdr*)gblob;
        const int n = gh->nunits, naa = gh->n_aa;
        if (n > P.scap) {  // does not fit the staged capacity: leave the game untouched for the full-capacity launch
            if (tid == 0) defer_game(P, g);
            return;
        }
        blob_load(mem + L.blob, gblob, 32 + n * 32, tid, nth);
        if (HEAVY && P.has_bot && naa > 0) blob_load(s.aa, gblob + 32 + P.cap * 32, align16(naa * 12), tid, nth);
        copy16<uint32_t>(s.grid, grid0, align16(P.HW * 2), tid, nth);
    }
    if (P.mode == MODE_STEP && plans) s.astar_init(tid, nth);
    if (tid == 0) {
        St z;
        z.npa = z.npa2 = z.astar_gen = z.gameover = z.reset = z.nprod = z.rows_ready = z.nidle = z.explore = z.unobs_ready = 0;
        z.winner = -1;
        z.timing_on = P.timing != nullptr;
        for (int k = 0; k < 12; k++) z.tp[k] = 0;
        for (int p = 0; p < 2; p++)
            for (int k = 0; k < 6; k++) z.rew[p][k] = 0;
        *s.st = z;
    }
    sync();
    if (P.mode == MODE_STEP && P.scap < P.cap) {
        // units can only be born from PRODUCE assignments pending at load (one issued now cannot complete
        // in this cycle: produce times are >= 50): if they might not fit, defer before anything is written
        const int n = s.hdr->nunits;
        int mine = 0;
        for (int i = tid; i < n; i += nth) mine += s.u[i].a_type == T_PRODUCE;
        if (mine) group_add(&s.st->nprod, mine);
        sync();
        if (n + ((volatile St*)s.st)->nprod > P.scap) {
            if (tid == 0) defer_game(P, g);
            return;
        }
    }
    scatter_grid();
    sync();
    if (P.mode == MODE_STEP && plans && !P.partial) s.flood_base_rows(-1, tid, nth);
    // the agents' action rows of the units that may act, fetched by the whole group (the rules never wait on HBM)
    if (P.mode == MODE_STEP && s.ai1 == AI_NONE)  // a scripted player 0 reads no action rows
        s.prefetch_actions(P.pact ? nullptr : P.actions + (size_t)env0 * P.HW * 7, P.pact ? P.pact + (size_t)env0 * P.HW : nullptr, tid, nth);
    sync();

    // ---------------- rules (first warp, uniform)
    if (rules_lane && P.mode == MODE_STEP) {
        const long long t_rules = P.timing ? cycles_now() : 0;
        const int r = s.template step_game<HEAVY>() ? 1 : 0;
        Sim::wsync();
        s.st->reset = r;
        if (P.timing && tid == 0) {
            P.timing[16 * g] = cycles_now() - t_rules;
            s.st->tp[11] = (int)(t_rules - t_start);
            for (int k = 0; k < 8; k++) P.timing[16 * g + 2 + k] = s.st->tp[k];
            for (int k = 8; k < 12; k++) P.timing[16 * g + 4 + k] = s.st->tp[k];
        }
    }
    sync();
    const int reset = ((volatile St*)s.st)->reset;
    if (P.mode == MODE_STEP) {
        // six reward terms per view (exact small integers) and done[0] of WinLoss, forced true by the
        // vec client on terminal / max_steps: 12 + 2 scalars, one per lane
        for (int k = tid; k < nviews * 6; k += nth)
            if (P.reward) P.reward[(size_t)env0 * 6 + k] = (double)s.st->rew[k / 6][k % 6];
        for (int v = tid; v < nviews; v += nth)
            if (P.done) P.done[env0 + v] = (uint8_t)reset;
        if (P.ended && tid == 0) P.ended[g] = (uint8_t)reset;
    }
    if (reset) {  // JNIGridnetClient.reset: fresh game from the map, scripted opponent re-created
        load_template(0);
        sync();
        scatter_grid();
        sync();
    }
    // a step that did not end the episode took its views inside step_game (before the issue); every other
    // observation (reset, getMasks) is of a state nobody has acted on yet: the views are taken now
    if (P.partial && (P.mode != MODE_STEP || reset)) {
        if (rules_lane)
            for (int v = 0; v < nviews; v++) s.make_view(v);
        sync();
    }

    // ---------------- observations, one view at a time (before compaction: stale views keep the dead)
#if defined(__CUDA_ARCH__)
    // direct mode on the warp kernels: the mask goes through the copy engine (mrts_bulk.cuh) -- 73 % of a step's bytes,
    // > 99 % of them zeros that leave from the block's zero tile without touching the load/store pipe; the staged rows
    // drain while the warp goes on (packed outputs, blob write-back).  The observation keeps its 128-bit store loop:
    // measured faster than staging it (every cell has to be expanded anyway).
    const bool bulk_out = P.bulk_out && blk != nullptr && nth == 32;
    static_assert(bulk::STAGE_BYTES == 2 * 1248, "Layout::slots");
    bulk::Stage stage;
    bulk::MaskWriter mw;
    if (bulk_out) {
        stage.init(mem + L.slots, tid);
        mw.init(&stage, blk, BULK_ZERO_BYTES, P.elem);
    }
#else
    const bool bulk_out = false;
    (void)blk;
#endif
    const size_t obs_env = (size_t)P.HW * P.planes * P.elem;
    const size_t mask_env = (size_t)P.HW * 78 * P.elem;
    if (P.obs || P.pobs) {
        for (int v = 0; v < nviews; v++) {
            obs_words_cells(s, v, P.partial ? v : -1, grid0, words, tid, nth);
            if (P.partial) {
                sync();
                obs_words_dead(s, v, v, words, tid, nth);
            }
            sync();
            if (P.obs) store_obs_rt(P.elem, P.planes, words, P.HW, (uint8_t*)P.obs + (size_t)(env0 + v) * obs_env, tid, nth);
            if (P.pobs) copy16<uint32_t>(P.pobs + (size_t)(env0 + v) * P.HW, words, P.HW * 4, tid, nth);
            sync();
        }
    }
    if (P.mode == MODE_STEP && !reset) {
        if (rules_lane) s.compact();
        sync();
    }
    if (P.mask || P.pmask || P.ment || P.next_actions) unit_masks(s, umask, tid, nth);

    // ---------------- mask: zero fill, then the 78 entries of every source unit
    const bool bulk_mask = P.mask && bulk_out;
    if (P.mask && !bulk_mask)
        for (int v = 0; v < nviews; v++) zero_mask_rt(P.elem, P.HW, (uint8_t*)P.mask + (size_t)(env0 + v) * mask_env, tid, nth);
    if (P.pmask) {
        uint32_t* pm = P.pmask + (size_t)env0 * P.HW * 3;
        for (int k = tid; k < nviews * P.HW * 3; k += nth) pm[k] = 0;
    }
    if (HEAVY && P.order_next && P.mode == MODE_STEP) {
        const int n = s.hdr->nunits;
        int mine = 0;
        for (int i = tid; i < n; i += nth) mine += s.u[i].player >= 0 && ai_plans(s.u[i].player == 0 ? s.ai1 : s.ai) && s.u[i].a_type < 0 && ut_can_move(s.u[i].type);
        if (mine) group_add(&s.st->nidle, mine);
    }
    sync();  // zero-fill before patches (and umask complete)
    if (HEAVY && P.order_next && P.mode == MODE_STEP && tid == 0) {
        const int c = order_class(((volatile St*)s.st)->nidle);
        P.order_next[(size_t)c * P.order_stride + order_append(&P.order_cnt_next[c])] = g;
    }
    if (P.next_actions)
        for (int v = 0; v < nviews; v++)
            sample_actions(s, v, umask, P.next_actions + (size_t)(env0 + v) * P.HW * 7, P.seed, P.counter, (uint32_t)(env0 + v), tid, nth);
    if (P.mask && !bulk_mask) {
        for (int v = 0; v < nviews; v++) {
            void* m = (uint8_t*)P.mask + (size_t)(env0 + v) * mask_env;
            patch_mask_rt(P.elem, s, v, umask, m, tid, nth);
        }
    }
#if defined(__CUDA_ARCH__)
    if (bulk_mask)
        for (int v = 0; v < nviews; v++) bulk_mask_view(s, v, umask, (uint8_t*)P.mask + (size_t)(env0 + v) * mask_env, mw, tid);
#endif
    if (P.ment)
        for (int v = 0; v < nviews; v++) emit_mask_entries(s, v, umask, P.ment + (size_t)(env0 + v) * ment_stride_words(P.cap), tid, nth);
    if (P.pmask) {
        const int n = s.hdr->nunits;
        for (int i = tid; i < n; i += nth) {
            const Unit& un = s.u[i];
            if (un.player < 0 || un.player >= nviews || un.a_type >= 0 || (un.flags & UF_DEAD)) continue;
            uint32_t* pm = P.pmask + ((size_t)(env0 + un.player) * P.HW + (un.x + un.y * s.W)) * 3;
            pm[0] = umask[3 * i];
            pm[1] = umask[3 * i + 1];
            pm[2] = umask[3 * i + 2];
        }
    }

    // ---------------- put the blob back
    if (P.mode != MODE_OBSERVE || P.partial) {
        const int n = s.hdr->nunits, naa = s.hdr->n_aa;
        blob_store(gblob, mem + L.blob, 32 + n * 32, tid, nth);
        if (HEAVY && P.has_bot && naa > 0) blob_store(gblob + 32 + P.cap * 32, s.aa, align16(naa * 12), tid, nth);
    }
#if defined(__CUDA_ARCH__)
    if (bulk_out) stage.drain();  // the staging buffers and the block's static tiles must outlive the copies that read them
#endif
    if (P.timing && tid == 0) P.timing[16 * g + 1] = cycles_now() - t_start;
}

// runtime choice of the instantiation (host build, overflow / debug launches)
template <class SYNC>
MRTS_HD void game_pass_any(const EnvParams& P, int g, uint8_t* mem, int tid, int nth, SYNC sync) {
    if (game_is_heavy(P, g)) game_pass<1>(P, g, mem, tid, nth, sync);
    else game_pass<0>(P, g, mem, tid, nth, sync);
}

}  // namespace mrts
