// Game rules of the B200-native MicroRTS simulator: the serial ("leader thread") part of one
// environment step.  One thread group owns one game whose state is staged in shared memory;
// the group's leader runs the functions below, the whole group runs the output phases in
// mrts_encode.h.  Compiles for device and host (see mrts_types.h).
//
// What each function replaces in the reference (relative to /root/reference/,
// J/ = gym_microrts/microrts/src/):
//   unit_mask            Unit.getUnitActions                J/rts/units/Unit.java:368-522
//                        + UnitAction.getValidActionArray   J/rts/UnitAction.java:717-757
//   decode_actions       PlayerAction.fromVectorAction      J/rts/PlayerAction.java:384-417
//                        + UnitAction.fromVectorAction      J/rts/UnitAction.java:681-715
//                        + fillWithNones                    J/rts/PlayerAction.java:217-235
//   issue_safe           GameState.issueSafe / issue        J/rts/GameState.java:338-408, 249-328
//   cycle                GameState.cycle / UnitAction.execute  J/rts/GameState.java:553-571, J/rts/UnitAction.java:344-471
//   make_views           PartiallyObservableGameState       J/rts/PartiallyObservableGameState.java:23-58
//   step_game            JNIGridnetClient.gameStep / JNIGridnetClientSelfPlay.gameStep /
//                        JNIGridnetVecClient.gameStep       J/tests/JNIGridnet*.java
//   rewards              J/ai/reward/{WinLoss,ResourceGather,ProduceWorker,ProduceBuilding,Attack,ProduceCombatUnit}RewardFunction.java
#pragma once
#include <stddef.h>

#include "mrts_types.h"

namespace mrts {

struct PAct {  // one entry of a PlayerAction (unit, UnitAction)
    uint16_t unit;  // unit index
    int8_t type;
    int8_t utype;
    int16_t param;
    int16_t x, y;   // attack target (may be off-map before the legality rewrite)
    int16_t pad;
};

// 78-bit legal-action set of one unit, in the mask layout of getValidActionArray:
// [0..5] action types, [6..9] move dir, [10..13] harvest dir, [14..17] return dir,
// [18..21] produce dir, [22..28] produce type, [29..77] attack offset (3+dy)*7+(3+dx).
struct Mask78 {
    uint32_t w[3];
};
#if defined(MRTS_STATS) && !defined(__CUDA_ARCH__)
#include <cstdio>
#include <cstdlib>
#include <vector>
struct MrtsStats { long searches = 0, pre_fail = 0, guard = 0, guard_fail = 0, expanded = 0; ~MrtsStats() { fprintf(stderr, "[mrts stats] searches %ld precheck-fail %ld guard %ld guard-fail %ld expansions %ld\n", searches, pre_fail, guard, guard_fail, expanded); } };
inline MrtsStats& mrts_stats() { static MrtsStats s; return s; }
#define MRTS_STAT(x) (mrts_stats().x)
#else
#define MRTS_STAT(x) ((void)0)
#endif
enum { ASTAR_GUARD_AFTER = 24 };  // expansions before a search is checked for reachability
enum { MB_TYPE = 0, MB_MOVE = 6, MB_HARVEST = 10, MB_RETURN = 14, MB_PDIR = 18, MB_PTYPE = 22, MB_ATTACK = 29 };
// no dynamic indexing of w[] (that would put the set in local memory); constant b folds away
MRTS_HD void mset(Mask78& m, int b) {
    const uint32_t bit = 1u << (b & 31);
    const int k = b >> 5;
    m.w[0] |= k == 0 ? bit : 0u;
    m.w[1] |= k == 1 ? bit : 0u;
    m.w[2] |= k == 2 ? bit : 0u;
}
MRTS_HD bool mget(const Mask78& m, int b) {
    const int k = b >> 5;
    const uint32_t w = k == 0 ? m.w[0] : (k == 1 ? m.w[1] : m.w[2]);
    return (w >> (b & 31)) & 1u;
}

// cost/parent/next per cell + one stack head per f value + stamps, then 3 row bitmaps (uint64 per map row) of the reachability guard
MRTS_HD int ceil_log2(int v) {
    int s = 0;
    while ((1 << s) < v) s++;
    return s;
}
// Maps up to 64 wide: cells are addressed as (y << ceil_log2(W)) | x (no division in the search) and
// the enterable / visited / flood sets are one uint64 per map row.  Wider maps keep linear
// positions and a per-cell stamp array.
MRTS_HD int astar_cells(int W, int HW) { return W <= 64 ? ((HW / W) << ceil_log2(W)) : HW; }
MRTS_HD int astar_core_bytes(int W, int HW) {
    return ((3 * astar_cells(W, HW) + (HW + W + (HW / W) + 2)) * 2 + (W <= 64 ? 0 : HW * 2) + 15) & ~15;
}
MRTS_HD int astar_bytes(int W, int HW) { return astar_core_bytes(W, HW) + (W <= 64 ? 4 * (HW / W) * 8 : 0); }

// Per-step scalars the rules mutate; they live in the game's shared-memory region so that the
// immutable view below (Sim) can be copied freely into the registers of out-of-line functions.
struct St {
    int npa, npa2;   // lengths of pa / pa2
    int astar_gen;   // search counter (stamp generation)
    int gameover, winner;
    int reset;
    int rows_ready;  // the enterable-cell rows of this bot call are built (see astar_rows)
    int nidle;       // idle mobile units of the scripted bot after this step (launch-order hint)
    int nprod;       // pending PRODUCE assignments at load (capacity check)
    int explore;     // the bot acting now is a PO* variant (bot_action)
    int unobs_ready; // bitsC holds this bot call's not-observable cell bitmap (bot_melee)
    int timing_on;   // profiling aid (mrts_debug_timing): fill tp[]
    int tp[12];      // cycles: decode, bot census/puts, bot translate, issue agent, issue bot, cycle; then #searches, #expansions;
                     // cycles inside path searches, cycles inside their reachability guard, #guards run, cycles of the staging phase
    int rew[2][6];   // integer reward terms per player view (exact small integers)
};

struct Sim {
    // ---- configuration
    int W, H, HW, cap;
    int none_mode, partial, max_steps;
    int ai;         // scripted opponent of a bot game (AI_*), AI_NONE for a self-play game
    int ai1;        // bot-vs-bot game (JNIBotClient): the scripted player 0, AI_NONE when an agent plays
    MRTS_HD bool plans() const { return ai_plans(ai) || ai_plans(ai1); }  // abstract-action table and path searches in use
    // ---- state staged for this step
    GameHdr* hdr;
    Unit* u;
    AbstractAct* aa;
    uint16_t* grid;  // HW cells: unit index, G_NONE or G_WALL
    // ---- leader scratch
    PAct* pa;        // cap entries: PlayerAction being issued
    PAct* pa2;       // cap entries: the bot's PlayerAction
    uint32_t* bitsA; // reservation bitmaps over linear positions -W .. HW+W-1 (bit index pos+W)
    uint32_t* bitsB;
    uint32_t* bitsC;
    int nbits_words;
    int16_t* astar;  // bot games: A* scratch (astar_bytes)
    St* st;          // per-step mutable scalars
    int lane, nl;    // the rules run on the game's first warp, every lane executing the same (uniform) code on the
                     // shared state; lane / lane count (32 on the device, 1 in the host build) drive the per-unit scans

    // ------------------------------------------------------------------ small helpers
    // ---- warp primitives of the rules phase (no-ops / identities with one lane)
    static MRTS_HD unsigned wballot(bool p) {
#if defined(__CUDA_ARCH__)
        return __ballot_sync(0xFFFFFFFFu, p);
#else
        return p ? 1u : 0u;
#endif
    }
    static MRTS_HD int wmin(int v) {
#if defined(__CUDA_ARCH__)
        return __reduce_min_sync(0xFFFFFFFFu, v);
#else
        return v;
#endif
    }
    static MRTS_HD int wadd(int v) {
#if defined(__CUDA_ARCH__)
        return __reduce_add_sync(0xFFFFFFFFu, v);
#else
        return v;
#endif
    }
    static MRTS_HD void wsync() {
#if defined(__CUDA_ARCH__)
        __syncwarp();
#endif
    }
    static MRTS_HD int ffs32(unsigned v) {  // 1-based index of the lowest set bit, 0 if none
#if defined(__CUDA_ARCH__)
        return __ffs((int)v);
#else
        return __builtin_ffs((int)v);
#endif
    }
    static MRTS_HD int popc(unsigned v) {
#if defined(__CUDA_ARCH__)
        return __popc(v);
#else
        return __builtin_popcount(v);
#endif
    }
    MRTS_HD unsigned lanes_below() const { return (1u << lane) - 1u; }
    static MRTS_HD int lane_now() {
#if defined(__CUDA_ARCH__)
        return (int)(threadIdx.x & 31);
#else
        return 0;
#endif
    }
    MRTS_HD int nunits() const { return hdr->nunits; }
    MRTS_HD int tick() const {
#if defined(__CUDA_ARCH__)
        return st->timing_on ? (int)clock64() : 0;
#else
        return 0;
#endif
    }
    MRTS_HD bool alive(int i) const { return !(u[i].flags & UF_DEAD); }
    // membership of unit i in the state the caller looks at: vp < 0 = the true state,
    // vp = p = player p's partially observable view taken this step
    MRTS_HD bool inview(int i, int vp) const { return vp < 0 || ((u[i].vis >> vp) & 1); }
    MRTS_HD int unit_at(int x, int y) const {  // PhysicalGameState.getUnitAt; -1 = none
        if (x < 0 || y < 0 || x >= W || y >= H) return -1;
        uint16_t g = grid[x + y * W];
        return g < G_WALL ? (int)g : -1;
    }
    MRTS_HD int unit_at_v(int x, int y, int vp) const {
        int i = unit_at(x, y);
        return (i >= 0 && inview(i, vp)) ? i : -1;
    }
    MRTS_HD bool wall_at(int x, int y) const { return grid[x + y * W] == G_WALL; }
    MRTS_HD int eta(int ut, int type, int param, int prod) const {  // UnitAction.ETA, UnitAction.java:313-335
        switch (type) {
            case T_NONE: return param;
            case T_MOVE: return ut_move_time(ut);
            case T_ATTACK: return ut_attack_time(ut);
            case T_HARVEST: return ut_harvest_time(ut);
            case T_RETURN: return ut_move_time(ut);  // sic, UnitAction.java:327-328
            case T_PRODUCE: return ut_produce_time(prod);
        }
        return 0;
    }
    MRTS_HD int none_param(int duration) const { return none_mode == NONE_FORK ? -1 : duration; }
    // UnitAction.resourceUsage (UnitAction.java:252-302): linear target index for MOVE/PRODUCE, else -0x7fff
    MRTS_HD int usage_pos(const Unit& un, int type, int param) const {
        if (type != T_MOVE && type != T_PRODUCE) return -0x7fff;
        int pos = un.x + un.y * W;
        switch (param) {
            case D_UP: pos -= W; break;
            case D_RIGHT: pos++; break;
            case D_DOWN: pos += W; break;
            case D_LEFT: pos--; break;
        }
        return pos;
    }
    MRTS_HD void bits_clear(uint32_t* b) const {  // all lanes together
        wsync();
#pragma unroll 1
        for (int i = lane; i < nbits_words; i += nl) b[i] = 0;
        wsync();
    }
    // set from a lane-parallel scan (lanes hold different positions)
    MRTS_HD void bit_set_par(uint32_t* b, int pos) const {
        int k = pos + W;
        if (k < 0 || k >= nbits_words * 32) return;
#if defined(__CUDA_ARCH__)
        atomicOr(&b[k >> 5], 1u << (k & 31));
#else
        b[k >> 5] |= 1u << (k & 31);
#endif
    }
    MRTS_HD bool bit_get(const uint32_t* b, int pos) const {
        int k = pos + W;
        if (k < 0 || k >= nbits_words * 32) return false;
        return (b[k >> 5] >> (k & 31)) & 1u;
    }
    MRTS_HD void bit_set(uint32_t* b, int pos) const {
        int k = pos + W;
        if (k < 0 || k >= nbits_words * 32) return;
        b[k >> 5] |= 1u << (k & 31);
    }

    // ------------------------------------------------------------------ legal actions
    // Unit.getUnitActions on the TRUE state, folded into the mask layout.  Legality of a
    // concrete action (Unit.canExecuteAction, Unit.java:531-534) is a bit test on this set:
    // getUnitActions emits the full product {affordable types} x {free directions} for
    // PRODUCE, so (dir bit AND type bit) is exact.
    MRTS_HD Mask78 unit_mask(int i) const {
        Mask78 m;
        m.w[0] = m.w[1] = m.w[2] = 0;
        const Unit& un = u[i];
        const int t = un.type, x = un.x, y = un.y, pl = un.player;
        int nb[4];      // neighbour unit index or -1
        bool inb[4];    // neighbour cell on the map
        bool freec[4];  // on the map, no wall
#pragma unroll
        for (int d = 0; d < 4; d++) {
            int nx = x + dir_dx(d), ny = y + dir_dy(d);
            inb[d] = nx >= 0 && ny >= 0 && nx < W && ny < H;
            uint16_t g = inb[d] ? grid[nx + ny * W] : (uint16_t)G_WALL;
            nb[d] = g < G_WALL ? (int)g : -1;
            freec[d] = inb[d] && g != G_WALL;
        }
        if (ut_can_attack(t)) {
            if (ut_range(t) == 1) {
#pragma unroll
                for (int d = 0; d < 4; d++)
                    if (nb[d] >= 0 && u[nb[d]].player != pl && u[nb[d]].player >= 0) {
                        mset(m, MB_TYPE + T_ATTACK);
                        mset(m, MB_ATTACK + (3 + dir_dy(d)) * 7 + 3 + dir_dx(d));
                    }
            } else {
                const int r = ut_range(t), sq = r * r;
#pragma unroll 1
                for (int dy = -r; dy <= r; dy++)
#pragma unroll 1
                    for (int dx = -r; dx <= r; dx++) {
                        if (dx * dx + dy * dy > sq) continue;
                        int o = unit_at(x + dx, y + dy);
                        if (o >= 0 && u[o].player >= 0 && u[o].player != pl) {
                            mset(m, MB_TYPE + T_ATTACK);
                            mset(m, MB_ATTACK + (3 + dy) * 7 + 3 + dx);
                        }
                    }
            }
        }
        if (ut_can_harvest(t)) {
#pragma unroll
            for (int d = 0; d < 4; d++) {
                if (nb[d] < 0) continue;
                if (un.res == 0 && ut_is_resource(u[nb[d]].type)) {
                    mset(m, MB_TYPE + T_HARVEST);
                    mset(m, MB_HARVEST + d);
                }
                if (un.res > 0 && ut_is_stockpile(u[nb[d]].type) && u[nb[d]].player == pl) {
                    mset(m, MB_TYPE + T_RETURN);
                    mset(m, MB_RETURN + d);
                }
            }
        }
        const int prod = ut_produces(t);
        if (prod && pl >= 0) {
            int afford = 0;
#pragma unroll 1
            for (int k = 1; k < N_UTYPES; k++)
                if (((prod >> k) & 1) && hdr->res[pl] >= ut_cost(k)) afford |= 1 << k;
            int dirs = 0;
#pragma unroll
            for (int d = 0; d < 4; d++)
                if (freec[d] && nb[d] < 0) dirs |= 1 << d;
            if (afford && dirs) {
                mset(m, MB_TYPE + T_PRODUCE);
                m.w[0] |= (uint32_t)dirs << MB_PDIR;
                m.w[0] |= (uint32_t)afford << MB_PTYPE;
            }
        }
        if (ut_can_move(t)) {
#pragma unroll
            for (int d = 0; d < 4; d++)
                if (freec[d] && nb[d] < 0) {
                    mset(m, MB_TYPE + T_MOVE);
                    mset(m, MB_MOVE + d);
                }
        }
        mset(m, MB_TYPE + T_NONE);
        return m;
    }

    // Unit.canExecuteAction (Unit.java:531-534): is `a` in getUnitActions?  Evaluated directly for the
    // one action instead of enumerating the whole set (unit_mask); tests/test_engine_units.py checks
    // (hs_can_execute_disagreements, tests/hostsim/hostsim.cpp) that the two agree on every (unit, action) of sampled states.
    MRTS_HD bool can_execute(int i, const PAct& a) const {
        if (a.type == T_NONE) return true;  // NONE(noneDuration = ETA(a)) is always in the list
        const Unit& un = u[i];
        const int t = un.type, pl = un.player, p = a.param;
        if (a.type == T_ATTACK) {
            if (!ut_can_attack(t)) return false;
            const int dx = a.x - un.x, dy = a.y - un.y, r = ut_range(t);
            if (dx * dx + dy * dy > r * r) return false;
            const int o = unit_at(a.x, a.y);
            return o >= 0 && u[o].player >= 0 && u[o].player != pl;
        }
        if (p < 0 || p > 3) return false;
        const int nx = un.x + dir_dx(p), ny = un.y + dir_dy(p);
        if (nx < 0 || ny < 0 || nx >= W || ny >= H) return false;
        const uint16_t g = grid[nx + ny * W];
        switch (a.type) {
            case T_MOVE: return ut_can_move(t) && g == G_NONE;
            case T_HARVEST: return ut_can_harvest(t) && un.res == 0 && g < G_WALL && ut_is_resource(u[g].type);
            case T_RETURN: return ut_can_harvest(t) && un.res > 0 && g < G_WALL && ut_is_stockpile(u[g].type) && u[g].player == pl;
            case T_PRODUCE:
                return pl >= 0 && a.utype >= 0 && a.utype < N_UTYPES && ((ut_produces(t) >> a.utype) & 1) && hdr->res[pl] >= ut_cost(a.utype) &&
                       g == G_NONE;
        }
        return false;
    }

    // ------------------------------------------------------------------ resource usage of pending assignments
    // GameState.getResourceUsage / the base_ru loop of fromVectorAction (PlayerAction.java:387-394):
    // positions into `bits`, costs into cost[2]; vp selects the state the caller sees.
    MRTS_HD void pending_usage(uint32_t* bits, int* cost, int vp) const {
        bits_clear(bits);
        int c0 = 0, c1 = 0;  // scalars: a dynamically indexed cost[] would live in local memory
        const int n = nunits();
#pragma unroll 1
        for (int i = lane; i < n; i += nl) {  // one unit per lane
            const Unit& un = u[i];
            if (un.a_type != T_MOVE && un.a_type != T_PRODUCE) continue;
            if (!inview(i, vp)) continue;
            bit_set_par(bits, usage_pos(un, un.a_type, un.a_param));
            if (un.a_type == T_PRODUCE) {
                const int c = ut_cost(un.a_utype);
                if (un.player == 0) c0 += c;
                else c1 += c;
            }
        }
        cost[0] = wadd(c0);
        cost[1] = wadd(c1);
        wsync();
    }
    // ResourceUsage.consistentWith(other) for a single-action `other` against an accumulated base
    // (ResourceUsage.java:31-50)
    MRTS_HD bool consistent_single(const uint32_t* bits, int used, int pos, int pl, int c) const {
        if (pos != -0x7fff && bit_get(bits, pos)) return false;
        if (c != 0) {
            int s = used + c;
            if (s > 0 && s > (pl == 0 ? hdr->res[0] : hdr->res[1])) return false;
        }
        return true;
    }

    // ------------------------------------------------------------------ fromVectorAction + fillWithNones
    // UnitAction.fromVectorAction (UnitAction.java:681-715) of one (7,) action row, packed into 10
    // bits: [0..2] type, [3..8] payload, [9] parameter out of range.  Payload: direction for
    // MOVE/HARVEST/RETURN, direction | produced type << 2 for PRODUCE, the 7x7 offset index for
    // ATTACK (63 = out of range -> off the window -> illegal).  Out-of-range values only matter
    // through "the action is illegal" and "which cell / cost it reserves", both kept.
    static MRTS_HD uint16_t pack_action(const int32_t* a) {
        const int t = a[0];
        if (t < T_MOVE || t > T_ATTACK) return T_NONE;  // NONE, or an out-of-range type: a no-op
        int pay = 0, bad = 0;
        if (t == T_PRODUCE) {
            if (a[5] < 0 || a[5] >= N_UTYPES) return T_NONE;  // utt.getUnitType would throw: treated as a no-op
            bad = a[4] < 0 || a[4] > 3;
            pay = (bad ? 0 : a[4]) | (a[5] << 2);
        } else if (t == T_ATTACK) {
            pay = (a[6] < 0 || a[6] > 48) ? 63 : a[6];
        } else {
            const int d = a[t];  // a[1] move, a[2] harvest, a[3] return direction
            bad = d < 0 || d > 3;
            pay = bad ? 0 : d;
        }
        return (uint16_t)(t | (pay << 3) | (bad << 9));
    }
    // whole thread group, before the rules: every idle owned unit fetches its cell's row of its
    // player view's (HW,7) block (self-play: view = player; bot game: only player 0 is driven)
    // a0 = the game's first view's (HW,7) int32 block, or (host transport) p0 = its (HW,) block of rows
    // already packed by pack_action on the host
    MRTS_HD void prefetch_actions(const int32_t* a0, const uint16_t* p0, int tid, int nth) {
        const int n = nunits();
        for (int i = tid; i < n; i += nth) {
            Unit& un = u[i];
            if (un.player < 0 || un.a_type >= 0) continue;
            if (ai != AI_NONE && un.player != 0) continue;
            const size_t cell = (size_t)un.player * HW + (un.x + un.y * W);
            un.pre = p0 ? p0[cell] : pack_action(a0 + cell * 7);
        }
    }

    // The reference visits the cells flagged by the previous get_action_mask in ascending index
    // (vec_env.py:974) and Java keeps a cell iff it holds an idle unit of the player; both reduce to
    // "own idle units in ascending cell order".
    MRTS_HD void decode_actions(int player, int vp) {
        int cost0, cost1;
        {
            int cost[2];
            pending_usage(bitsA, cost, vp);
            cost0 = cost[0];
            cost1 = cost[1];
        }
        int cnt = 0;
        const int n = nunits();
        // own idle units in ascending cell order: one unit per lane marks its cell in a bitmap (bitsC is
        // free here), then the set bits are walked in order
        bits_clear(bitsC);
#pragma unroll 1
        for (int i = lane; i < n; i += nl) {
            const Unit& un = u[i];
            if (un.player != player || un.a_type >= 0 || !inview(i, vp)) continue;
            const int pos = un.x + un.y * W;
#if defined(__CUDA_ARCH__)
            atomicOr(&bitsC[pos >> 5], 1u << (pos & 31));
#else
            bitsC[pos >> 5] |= 1u << (pos & 31);
#endif
        }
        wsync();
        const int nw = (HW + 31) >> 5;
#pragma unroll 1
        for (int base = 0; base < nw; base += nl) {
            const int wi = base + lane;
            unsigned nz = wballot(wi < nw && bitsC[wi] != 0);
#pragma unroll 1
            while (nz) {
                const int k = ffs32(nz) - 1;
                nz &= nz - 1;
                uint32_t m = bitsC[base + k];
#pragma unroll 1
                while (m) {
                    const int b = ffs32(m) - 1;
                    m &= m - 1;
                    pa[cnt++].unit = grid[32 * (base + k) + b];
                }
            }
        }
        wsync();
        int kept = 0;
        const int have = player == 0 ? hdr->res[0] : hdr->res[1];
        int used = player == 0 ? cost0 : cost1;
#pragma unroll 1
        for (int k = 0; k < cnt; k++) {
            const int i = pa[k].unit;
            const Unit& un = u[i];
            const int w = un.pre, pay = (w >> 3) & 63, bad = (w >> 9) & 1;
            PAct e;
            e.unit = (uint16_t)i;
            e.type = (int8_t)(w & 7);
            e.param = -1;  // DIRECTION_NONE, UnitAction.java:114
            e.utype = -1;
            e.x = e.y = 0;
            e.pad = 0;
            if (e.type == T_PRODUCE) {
                e.param = (int16_t)(bad ? -1 : (pay & 3));
                e.utype = (int8_t)(pay >> 2);
            } else if (e.type == T_ATTACK) {
                e.x = (int16_t)(un.x + (pay % 7 - 3));
                e.y = (int16_t)(un.y + (pay / 7 - 3));
            } else if (e.type != T_NONE) {
                e.param = (int16_t)(bad ? -1 : pay);
            }
            const int pos = usage_pos(un, e.type, e.param);
            const int c = e.type == T_PRODUCE ? ut_cost(e.utype) : 0;
            // ResourceUsage.consistentWith (ResourceUsage.java:31-50) against the accumulated base
            bool ok = !(pos != -0x7fff && bit_get(bitsA, pos));
            if (ok && c != 0 && used + c > 0 && used + c > have) ok = false;
            if (ok) {
                if (pos != -0x7fff) bit_set(bitsA, pos);
                used += c;
                pa[kept++] = e;  // kept <= k: never overwrites an unread candidate
            }
        }
        fill_with_nones(pa, kept, player, 1, vp);
        st->npa = kept;
    }

    // PlayerAction.fillWithNones: idle own units not yet in the action, in unit order
    MRTS_HD void fill_with_nones(PAct* list, int& cnt, int player, int duration, int vp) {
        const int n = nunits();
        const int have = cnt;
        wsync();
#pragma unroll 1
        for (int k = lane; k < have; k += nl) u[list[k].unit].flags |= UF_LISTED;  // distinct units: no two lanes share a record
        wsync();
        int total = have;
#pragma unroll 1
        for (int base = 0; base < n; base += nl) {  // one unit per lane; the ballot keeps unit order
            const int i = base + lane;
            bool take = false;
            if (i < n) {
                Unit& un = u[i];
                if (un.flags & UF_LISTED) un.flags &= (uint8_t)~UF_LISTED;
                else take = un.player == player && un.a_type < 0 && inview(i, vp);
            }
            const unsigned m = wballot(take);
            if (take) {
                PAct e;
                e.unit = (uint16_t)i;
                e.type = T_NONE;
                e.param = (int16_t)none_param(duration);
                e.utype = -1;
                e.x = e.y = 0;
                e.pad = 0;
                list[total + popc(m & lanes_below())] = e;
            }
            total += popc(m);
        }
        wsync();
        cnt = total;
    }

    // ------------------------------------------------------------------ issueSafe + issue
    MRTS_HD void reward_for_issued(const PAct& a) {
        const Unit& un = u[a.unit];
        const int p = un.player;
        if (p < 0) return;
        if (a.type == T_HARVEST || a.type == T_RETURN) st->rew[p][1] += 1;  // ResourceGatherRewardFunction.java:22-31
        if (a.type == T_PRODUCE && a.utype >= 0) {
            if (a.utype == UT_WORKER) st->rew[p][2] += 1;                                      // ProduceWorker :20-30
            if (a.utype == UT_BASE || a.utype == UT_BARRACKS) st->rew[p][3] += 1;              // ProduceBuilding :20-30
            if (a.utype == UT_LIGHT || a.utype == UT_HEAVY || a.utype == UT_RANGED) st->rew[p][5] += 1;  // ProduceCombatUnit
        }
        if (a.type == T_ATTACK) {  // AttackRewardFunction.java:20-36: occupant of the target cell in the pre-cycle snapshot
            int o = unit_at(a.x, a.y);
            if (o >= 0) {
                if (u[o].player == 1 - p) st->rew[p][4] += 1;
                else if (u[o].player == p) st->rew[p][4] -= 1;
            }
        }
    }

    MRTS_HD void issue_safe(PAct* list, int cnt) {
        wsync();
        const int time = hdr->time;
        // target cells of every pending MOVE / PRODUCE (a superset is fine: it only gates the exact scan)
        bits_clear(bitsB);
        {
            const int n = nunits();
#pragma unroll 1
            for (int j = lane; j < n; j += nl) {  // one unit per lane
                const Unit& ou = u[j];
                if (ou.a_type == T_MOVE || ou.a_type == T_PRODUCE) bit_set_par(bitsB, usage_pos(ou, ou.a_type, ou.a_param));
            }
            wsync();
        }
#pragma unroll 1
        for (int k = 0; k < cnt; k++) {
            PAct a = list[k];
            Unit& un = u[a.unit];
            // ---- legality rewrite (GameState.java:347-354); mutates the PlayerAction
            if (!can_execute(a.unit, a)) {
                int l = eta(un.type, a.type, a.param, a.utype >= 0 ? a.utype : 0);
                a.type = T_NONE;
                a.param = (int16_t)none_param(l);
                a.utype = -1;
            }
            // (second check of GameState.java:384-400 -- target wall/occupied -- is implied by legality)
            // ---- issue (GameState.java:249-328)
            int8_t ctype = a.type;        // the local `p` that CANCEL_BOTH rebinds
            int16_t cparam = a.param;
            int8_t cutype = a.utype;
            bool mutated = false;         // GameState.java:316 mutates the Pair itself
            const int pos = usage_pos(un, a.type, a.param);
            if (pos != -0x7fff) {
                const int c = a.type == T_PRODUCE ? ut_cost(a.utype) : 0;
                const int pl = un.player;
                const int n = nunits();
                // quick test: any conflicting assignment at all?  A MOVE conflicts through its target cell
                // only; a PRODUCE also through the players' resources, which needs the scan.
                bool any = bit_get(bitsB, pos);
                if (c != 0)
#pragma unroll 1
                    for (int j = 0; j < n && !any; j++) any = conflicts(j, pos, pl, c);
                if (any) {
                    // slow path, in issue order (LinkedHashMap iteration)
                    uint32_t last = 0;
                    bool first = true;
#pragma unroll 1
                    for (;;) {
                        int best = -1;
                        uint32_t bs = 0xFFFFFFFFu;
#pragma unroll 1
                        for (int j = 0; j < n; j++) {
                            if (u[j].a_type < 0) continue;
                            if (!first && u[j].a_seq <= last) continue;
                            if (u[j].a_seq < bs) {
                                bs = u[j].a_seq;
                                best = j;
                            }
                        }
                        if (best < 0) break;
                        first = false;
                        last = bs;
                        if (!conflicts(best, pos, pl, c)) continue;
                        Unit& ou = u[best];
                        if (ou.a_time == time) {
                            int d1 = eta(ou.type, ou.a_type, ou.a_param, ou.a_utype >= 0 ? ou.a_utype : 0);
                            int d2 = eta(un.type, ctype, cparam, cutype >= 0 ? cutype : 0);
                            int mn = d1 < d2 ? d1 : d2;
                            ou.a_type = T_NONE;  // slot (a_seq) and a_time kept, GameState.java:292
                            ou.a_param = (int16_t)none_param(mn);
                            ou.a_utype = -1;
                            for (int v = 0; v < 2; v++)
                                if (ou.vtype[v] >= 0) ou.vtype[v] = T_NONE;  // views share the assignment object
                            ctype = T_NONE;
                            cparam = (int16_t)none_param(mn);
                            cutype = -1;
                        } else {
                            ctype = T_NONE;  // new UnitAction(TYPE_NONE): parameter -1
                            cparam = -1;
                            cutype = -1;
                            mutated = true;
                        }
                    }
                }
            }
            if (mutated) {
                a.type = T_NONE;
                a.param = -1;
                a.utype = -1;
            }
            reward_for_issued(a);  // rewards see the PlayerAction as left by issueSafe
            un.a_type = ctype;
            un.a_param = cparam;
            un.a_utype = cutype;
            un.a_x = (uint8_t)a.x;
            un.a_y = (uint8_t)a.y;
            un.a_time = time;
            un.a_seq = hdr->next_seq++;
            if (pos != -0x7fff) bit_set(bitsB, pos);
        }
    }
    // !consistent(usage(assignment of unit j), ru) with ru = {pos, cost c of player pl}
    MRTS_HD bool conflicts(int j, int pos, int pl, int c) const {
        const Unit& ou = u[j];
        if (ou.a_type != T_MOVE && ou.a_type != T_PRODUCE) return false;
        if (usage_pos(ou, ou.a_type, ou.a_param) == pos) return true;
        if (c != 0) {
            int oc = (ou.a_type == T_PRODUCE && ou.player == pl) ? ut_cost(ou.a_utype) : 0;
            int s = oc + c;
            if (s > 0 && s > hdr->res[pl]) return true;
        }
        return false;
    }

    // ------------------------------------------------------------------ cycle
    MRTS_HD void kill_unit(int o) {  // GameState.removeUnit: unit and its assignment leave the state
        Unit& ou = u[o];
        ou.flags |= UF_DEAD;  // the record (incl. a collected action) stays until compact()
        grid[ou.x + ou.y * W] = G_NONE;
    }
    MRTS_HD void execute(int i, int type, int param, int ax, int ay, int prod) {
        Unit& un = u[i];
        const bool dead = un.flags & UF_DEAD;  // removed earlier this cycle: still executes (ready list is a snapshot)
        switch (type) {
            case T_MOVE: {
                if (!dead) grid[un.x + un.y * W] = G_NONE;
                un.x = (uint8_t)(un.x + dir_dx(param));
                un.y = (uint8_t)(un.y + dir_dy(param));
                if (!dead) grid[un.x + un.y * W] = (uint16_t)i;
                break;
            }
            case T_ATTACK: {
                int o = unit_at(ax, ay);
                if (o >= 0) {
                    u[o].hp = (int16_t)(u[o].hp - ut_damage(un.type));
                    if (u[o].hp <= 0) kill_unit(o);
                }
                break;
            }
            case T_HARVEST: {
                int o = (param >= 0 && param < 4) ? unit_at(un.x + dir_dx(param), un.y + dir_dy(param)) : -1;
                if (o >= 0 && ut_is_resource(u[o].type) && ut_can_harvest(un.type) && un.res == 0) {
                    u[o].res -= 1;  // harvestAmount
                    if (u[o].res <= 0) kill_unit(o);
                    un.res = 1;
                }
                break;
            }
            case T_RETURN: {
                int o = (param >= 0 && param < 4) ? unit_at(un.x + dir_dx(param), un.y + dir_dy(param)) : -1;
                if (o >= 0 && ut_is_stockpile(u[o].type) && un.res > 0) {  // owner NOT checked, UnitAction.java:430
                    hdr->res[un.player] += un.res;
                    un.res = 0;
                }
                break;
            }
            case T_PRODUCE: {
                int tx = un.x + dir_dx(param), ty = un.y + dir_dy(param);
                if (hdr->res[un.player] - ut_cost(prod) >= 0 && hdr->nunits < cap) {
                    int k = hdr->nunits++;
                    Unit& nu = u[k];
                    nu.x = (uint8_t)tx;
                    nu.y = (uint8_t)ty;
                    nu.type = (uint8_t)prod;
                    nu.player = un.player;
                    nu.hp = (int16_t)ut_hp(prod);
                    nu.res = 0;
                    nu.a_type = -1;
                    nu.a_param = -1;
                    nu.a_utype = -1;
                    nu.a_x = nu.a_y = 0;
                    nu.a_time = 0;
                    nu.a_seq = 0;
                    nu.uid = hdr->next_uid++;
                    nu.flags = 0;
                    nu.vis = 0;  // born after the views were taken
                    nu.vtype[0] = nu.vtype[1] = -1;
                    nu.pre = 0;
                    grid[tx + ty * W] = (uint16_t)k;
                    hdr->res[un.player] -= ut_cost(prod);
                }
                break;
            }
        }
    }

    // GameState.cycle.  Dead units stay in the array (flagged) until compact().
    MRTS_HD void cycle() {
        wsync();
        hdr->time++;
        const int time = hdr->time;
        const int n = nunits();  // units born this cycle have no assignment
        // ready non-NONE assignments, sorted by issue order; NONEs just vanish (no effect).
        // One unit per lane finds the ready ones (ballot keeps unit order), then the few are sorted.
        int nr = 0;
        uint16_t* ready = (uint16_t*)pa;  // pa is free again
        wsync();
#pragma unroll 1
        for (int base = 0; base < n; base += nl) {
            const int i = base + lane;
            bool rdy = false;
            if (i < n) {
                Unit& un = u[i];
                if (un.a_type >= 0) {
                    const int e = eta(un.type, un.a_type, un.a_param, un.a_utype >= 0 ? un.a_utype : 0);
                    if (e + un.a_time <= time) {
                        if (un.a_type == T_NONE) un.a_type = -1;
                        else rdy = true;
                    }
                }
            }
            const unsigned m = wballot(rdy);
            if (rdy) ready[nr + popc(m & lanes_below())] = (uint16_t)i;
            nr += popc(m);
        }
        wsync();
#pragma unroll 1
        for (int a = 1; a < nr; a++) {  // insertion sort by issue order (uniform)
            const uint16_t v = ready[a];
            const uint32_t sv = u[v].a_seq;
            int k = a;
#pragma unroll 1
            while (k > 0 && u[ready[k - 1]].a_seq > sv) {
                ready[k] = ready[k - 1];
                k--;
            }
            ready[k] = v;
        }
#pragma unroll 1
        for (int k = 0; k < nr; k++) {
            const int i = ready[k];
            Unit& un = u[i];
            // a unit killed earlier this cycle lost its map entry but still executes the collected action
            const int type = un.a_type, param = un.a_param, ax = un.a_x, ay = un.a_y, prod = un.a_utype;
            un.a_type = -1;
            execute(i, type, param, ax, ay, prod);
        }
    }

    // stable compaction after cycle (PhysicalGameState.removeUnit keeps relative order) + grid fix-up
    MRTS_HD void compact() {
        const int n = nunits();
        wsync();
        int k = 0;
#pragma unroll 1
        for (int base = 0; base < n; base += nl) {  // one unit per lane; a chunk's records are read before any is moved
            const int i = base + lane;
            const bool live = i < n && !(u[i].flags & UF_DEAD);
            const unsigned m = wballot(live);
            const int dst = k + popc(m & lanes_below());
            Unit rec;
            if (live) rec = u[i];
            wsync();
            if (live && dst != i) {
                u[dst] = rec;
                grid[rec.x + rec.y * W] = (uint16_t)dst;
            }
            k += popc(m);
            wsync();
        }
        hdr->nunits = k;
        wsync();
    }

    // PhysicalGameState.gameover / st->winner (PhysicalGameState.java:319-372) over live units
    MRTS_HD void eval_gameover() {
        int c0 = 0, c1 = 0;
        const int n = nunits();
        wsync();
#pragma unroll 1
        for (int i = lane; i < n; i += nl)  // one unit per lane
            if (alive(i)) {
                c0 += u[i].player == 0;
                c1 += u[i].player == 1;
            }
        c0 = wadd(c0);
        c1 = wadd(c1);
        if (c0 + c1 == 0) {
            st->gameover = 1;
            st->winner = -1;
        } else if (c0 > 0 && c1 > 0) {
            st->gameover = 0;
            st->winner = -1;
        } else {
            st->gameover = 1;
            st->winner = c0 > 0 ? 0 : 1;
        }
    }

    // PartiallyObservableGameState(gs, p): membership + assignment copy, recorded per unit
    MRTS_HD void make_view(int p) {
        const int n = nunits();
#pragma unroll 1
        for (int i = 0; i < n; i++) {
            Unit& un = u[i];
            bool in = un.player == p;
            if (!in) {
#pragma unroll 1
                for (int j = 0; j < n && !in; j++) {
                    const Unit& w = u[j];
                    if (w.player != p) continue;
                    int dx = w.x - un.x, dy = w.y - un.y, r = ut_sight(w.type);
                    in = dx * dx + dy * dy <= r * r;  // sqrt(d2) <= r, exact on integers
                }
            }
            if (in) {
                un.vis |= (uint8_t)(1 << p);
                un.vtype[p] = un.a_type;
            } else {
                un.vis &= (uint8_t)~(1 << p);
                un.vtype[p] = -1;
            }
        }
    }

    // ================================================================== scripted opponents
    // AbstractionLayerAI table helpers (J/ai/abstraction/AbstractionLayerAI.java:36,115-141)
    MRTS_HD int aa_find(uint16_t uid) const {  // one table entry per lane
        const int na = hdr->n_aa;
        wsync();
#pragma unroll 1
        for (int base = 0; base < na; base += nl) {
            const int k = base + lane;
            const unsigned m = wballot(k < na && aa[k].unit == uid);
            if (m) return base + ffs32(m) - 1;
        }
        return -1;
    }
    // arg-min over the units of a warp-parallel scan: key = (value << 16) | unit index, so that among
    // equal values the lowest index wins -- what a strict `<` scan in unit order keeps.  -1 if no candidate.
    static MRTS_HD int argmin_key(int value, int index) { return (value << 16) | index; }
    enum { NO_KEY = 0x7FFFFFFF };
    MRTS_HD void aa_put(const AbstractAct& r, int i, int k) {  // LinkedHashMap.put: re-put keeps the slot; i = the unit's index, k = aa_find(r.unit)
        if (k < 0) {
            if (hdr->n_aa >= cap) return;
            k = hdr->n_aa++;
        }
        wsync();
        aa[k] = r;
        // bot-vs-bot games keep both bots' tables in this one table: entries of player 0's bot carry AA_OWNER0
        // (each bot only ever sees its own entries, in their insertion order -- two LinkedHashMaps)
        if (u[i].player == 0) aa[k].kind = (uint8_t)(r.kind | AA_OWNER0);
        // UF_AA_ATTACK mirrors "this unit's table entry is an Attack" (see worker_rush_jobs)
        if (r.kind == AA_ATTACK) u[i].flags |= UF_AA_ATTACK;
        else u[i].flags &= (uint8_t)~UF_AA_ATTACK;
        wsync();
    }
    // pgs.getUnits().contains(u) on the view: one unit per lane compares its uid (uids are unique inside an episode)
    MRTS_HD int find_uid(uint16_t uid, int vp) const {
        if (uid == NO_UID) return -1;
        const int n = nunits();
#pragma unroll 1
        for (int base = 0; base < n; base += nl) {
            const int i = base + lane;
            const unsigned m = wballot(i < n && u[i].uid == uid);
            if (m) {
                const int k = base + ffs32(m) - 1;
                return inview(k, vp) ? k : -1;
            }
        }
        return -1;
    }
    MRTS_HD static int iabs(int v) { return v < 0 ? -v : v; }

    // gs.free(x,y) on the bot's view, J/rts/GameState.java:191-207; bitsB = pending MOVE/PRODUCE targets
    MRTS_HD bool free_cell(int x, int y, int vp) const {
        const int pos = x + y * W;
        uint16_t g = grid[pos];
        if (g == G_WALL) return false;
        if (g != G_NONE && inview(g, vp)) return false;
        return !bit_get(bitsB, pos);
    }
    // GameState.isUnitActionAllowed, J/rts/GameState.java:434-457 (argument order of consistentWith kept)
    MRTS_HD bool action_allowed(int i, int type, int param, int prod, int vp, const int* costB) const {
        const Unit& un = u[i];
        if (type == T_MOVE) {
            int x2 = un.x + dir_dx(param), y2 = un.y + dir_dy(param);
            if (x2 < 0 || y2 < 0 || x2 >= W || y2 >= H) return false;
            if (wall_at(x2, y2) || unit_at_v(x2, y2, vp) >= 0) return false;
        }
        const int pos = usage_pos(un, type, param);
        if (pos != -0x7fff && bit_get(bitsB, pos)) return false;
#pragma unroll
        for (int p = 0; p < 2; p++) {
            if (costB[p] == 0) continue;
            int s = ((type == T_PRODUCE && un.player == p) ? ut_cost(prod) : 0) + costB[p];
            if (s > 0 && s > hdr->res[p]) return false;
        }
        return true;
    }

    // ---- A*, J/ai/abstraction/pathfinding/AStarPathFinding.java:52-79,104-138,175-295
    // The reference keeps the open list as an array sorted by f (largest first), pops from the
    // tail and inserts a new node right after the last entry (scanning from the tail) whose f is
    // >= the new one (addToOpen, :104-138).  That order is exactly "smallest f first, newest first
    // among equal f", i.e. one LIFO stack per f value, and a position is pushed at most once
    // (inOpenOrClosed, :60).  The stacks are intrusive lists threaded through `next`:
    // push and pop are O(1) where the reference shifts the arrays.
    struct AStarBuf {
        int16_t *cost, *parent, *next, *head;  // cost/parent/next per cell; head per f value (-1 = empty)
        uint16_t* stamp;                       // == generation: cell was pushed during this search
    };
    MRTS_HD int astar_nf() const { return HW + W + H + 2; }
    MRTS_HD AStarBuf astar_buf() const {
        AStarBuf b;
        const int nc = astar_cells(W, HW);
        b.cost = astar;
        b.parent = b.cost + nc;
        b.next = b.parent + nc;
        b.head = b.next + nc;
        b.stamp = (uint16_t*)(b.head + astar_nf());
        return b;
    }
    // ---- row bitmaps of the search on maps up to 64 wide (bit x of row y): F0 = cells free of walls and
    // of the units the bot sees (constant during a step), F = cells a search may enter during this
    // bot call (F0 minus pending MOVE/PRODUCE targets, bitsB, minus the desires accepted so far,
    // bitsC), V = cells pushed by the current search, R = flood-fill scratch
    MRTS_HD uint64_t* rows_F0() const { return (uint64_t*)((uint8_t*)astar + astar_core_bytes(W, HW)); }
    MRTS_HD uint64_t* rows_F() const { return rows_F0() + H; }
    MRTS_HD uint64_t* rows_V() const { return rows_F0() + 2 * H; }
    MRTS_HD uint64_t* rows_R() const { return rows_F0() + 3 * H; }
    // built once per pass by the whole group; under partial observability the bot's view hides
    // units, so the leader rebuilds the rows for its view when a search needs them
    MRTS_HD void flood_base_rows(int vp, int tid, int nth) const {
        if (W > 64) return;
        uint64_t* F0 = rows_F0();
#pragma unroll 1
        for (int y = tid; y < H; y += nth) {
            uint64_t row = 0;
#pragma unroll 1
            for (int x = 0; x < W; x++) {
                const uint16_t g = grid[x + y * W];
                if (g == G_NONE || (g < G_WALL && !inview(g, vp))) row |= 1ull << x;
            }
            F0[y] = row;
        }
    }
    // W bits of a reservation bitmap (bit index = linear position + W) starting at the row's first cell
    MRTS_HD uint64_t bits_row(const uint32_t* b, int y) const {
        const int k0 = (y + 1) * W;
        uint64_t row = 0;
        int got = 0;
#pragma unroll 1
        while (got < W) {
            const int k = k0 + got, w = k >> 5, o = k & 31;
            const int take = (32 - o) < (W - got) ? (32 - o) : (W - got);
            const uint32_t v = w < nbits_words ? b[w] : 0u;
            row |= (uint64_t)((v >> o) & (take == 32 ? 0xFFFFFFFFu : ((1u << take) - 1u))) << got;
            got += take;
        }
        return row;
    }
    // F for this bot call, on the first search (translate_actions keeps it current afterwards)
    MRTS_HD void rows_prepare(int vp) const {
        if (st->rows_ready) return;
        wsync();
        if (partial) flood_base_rows(vp, lane, nl);
        wsync();
        const uint64_t* F0 = rows_F0();
        uint64_t* F = rows_F();
#pragma unroll 1
        for (int y = lane; y < H; y += nl) F[y] = F0[y] & ~bits_row(bitsB, y) & ~bits_row(bitsC, y);  // one row per lane
        wsync();
        st->rows_ready = 1;
    }
    // a desire was accepted (its target is now in bitsC): later searches of this call may not enter it
    MRTS_HD void rows_reserve(int pos) const {
        if (W > 64 || !st->rows_ready || pos < 0 || pos >= HW) return;
        rows_F()[pos / W] &= ~(1ull << (pos % W));
    }
    static MRTS_HD uint64_t brev64(uint64_t v) {
#if defined(__CUDA_ARCH__)
        return __brevll(v);
#else
        uint64_t r = 0;
        for (int k = 0; k < 64; k++) r |= ((v >> k) & 1ull) << (63 - k);
        return r;
#endif
    }
    // seeds S (subset of F) spread along the runs of ones of F, both ways (carry trick)
    static MRTS_HD uint64_t hfill(uint64_t S, uint64_t F) {
        const uint64_t up = (((F + S) ^ F) & F) | S;
        const uint64_t Fr = brev64(F), Sr = brev64(S);
        const uint64_t dn = (((Fr + Sr) ^ Fr) & Fr) | Sr;
        return up | brev64(dn);
    }
    // Can a search from (sx,sy) ever pop a cell within `range` of (tx,ty)?  Flood fill of F from the
    // start.  Only used to cut a search short that is bound to fail: the search result is unchanged.
    // Device, maps up to 32 rows: lane y keeps row y in a register and every round each row takes in its
    // two neighbours (shuffles) and spreads along its runs of ones -- a few dozen instructions per round, one
    // round per row the fill has to travel, instead of serial sweeps over shared memory.
    MRTS_HD bool flood_reaches(int sx, int sy, int tx, int ty, int range) const {
        const uint64_t* F = rows_F();
        uint64_t* R = rows_R();
        const uint64_t sbit = 1ull << sx;
#if defined(__CUDA_ARCH__)
        if (H <= 32) {
            const uint64_t Fy = lane < H ? F[lane] : 0ull;
            // the start cell itself is occupied: seed its enterable neighbours
            uint64_t Ry = 0;
            if (lane == sy - 1 || lane == sy + 1) Ry = sbit & Fy;
            if (lane == sy) Ry = hfill(((sbit << 1) | (sbit >> 1)) & Fy, Fy);
#pragma unroll 1
            for (;;) {
                uint64_t up = __shfl_up_sync(0xFFFFFFFFu, Ry, 1), dn = __shfl_down_sync(0xFFFFFFFFu, Ry, 1);
                if (lane == 0) up = 0;
                if (lane == 31) dn = 0;
                const uint64_t filled = hfill((Ry | up | dn) & Fy, Fy);
                const unsigned changed = wballot(filled != Ry);
                Ry = filled;
                if (!changed) break;
            }
            // any filled cell within range of the target?
            const int dy = lane - ty, sq = range * range;
            bool hit = false;
            if (lane < H && dy >= -range && dy <= range) {
#pragma unroll 1
                for (int dx = -range; dx <= range; dx++) {
                    const int x = tx + dx;
                    if (x < 0 || x >= W || dx * dx + dy * dy > sq) continue;
                    hit = hit || ((Ry >> x) & 1ull);
                }
            }
            return wballot(hit) != 0;
        }
#endif
#pragma unroll 1
        for (int y = 0; y < H; y++) R[y] = 0;
        // the start cell itself is occupied: seed its enterable neighbours
        if (sy > 0) R[sy - 1] = sbit & F[sy - 1];
        if (sy + 1 < H) R[sy + 1] = sbit & F[sy + 1];
        R[sy] = hfill(((sbit << 1) | (sbit >> 1)) & F[sy], F[sy]);
        bool changed = true;
#pragma unroll 1
        while (changed) {
            changed = false;
#pragma unroll 1
            for (int pass = 0; pass < 2; pass++) {
#pragma unroll 1
                for (int k = 0; k < H; k++) {
                    const int y = pass == 0 ? k : H - 1 - k;
                    uint64_t seeds = R[y];
                    if (y > 0) seeds |= R[y - 1];
                    if (y + 1 < H) seeds |= R[y + 1];
                    seeds &= F[y];
                    const uint64_t filled = hfill(seeds, F[y]);
                    if (filled != R[y]) {
                        R[y] = filled;
                        changed = true;
                    }
                }
            }
        }
        return rows_any_in_range(R, tx, ty, range);
    }
    MRTS_HD bool rows_any_in_range(const uint64_t* rows, int tx, int ty, int range) const {
        const int sq = range * range;
#pragma unroll 1
        for (int dy = -range; dy <= range; dy++) {
            const int y = ty + dy;
            if (y < 0 || y >= H) continue;
#pragma unroll 1
            for (int dx = -range; dx <= range; dx++) {
                const int x = tx + dx;
                if (x < 0 || x >= W || dx * dx + dy * dy > sq) continue;
                if ((rows[y] >> x) & 1ull) return true;
            }
        }
        return false;
    }

    // findPathToPositionInRange on a map up to 64 wide.  Same open-list order as astar_first_move
    // (the generic form below, which documents it), with cells addressed as (y << ws) | x and the
    // neighbour tests "enterable and not yet pushed" done on the row bitmaps.
    //
    // One expansion = one pop + up to four pushes.  The four neighbours (up, right, down, left) are handled by
    // four lanes at once on the device (the host build walks d = 0..3): with a Manhattan heuristic a neighbour's
    // f is either the popped node's f (it is closer to the target) or f + 2, so the pushes fall into two stacks and
    // a ballot tells every lane who else pushes onto its stack -- the lane links its node to the next lower
    // pusher (or to the old head), the highest pusher becomes the new head: exactly the order d = 0..3 of
    // sequential pushes.  Each node also records the first move of its path (`parent` array: the root
    // direction), so the answer is read off the node that ends the search instead of walking back.
    // pusher d's link: the node below it on its stack after the pushes of this expansion (reads only)
    MRTS_HD int astar_link(const AStarBuf& b, int d, unsigned okm, unsigned cm, int x, int y, int ws, int fcur) const {
        if (!((okm >> d) & 1u)) return -1;
        const bool closer = (cm >> d) & 1u;
        const unsigned grp = closer ? cm : (okm & ~cm);
        const unsigned lower = grp & ((1u << d) - 1u);
        if (lower) {
            const int dp = 31 - clz32(lower);  // the pusher right below this one on the same stack
            return ((y + dir_dy(dp)) << ws) | (x + dir_dx(dp));
        }
        return b.head[closer ? fcur : fcur + 2];
    }
    // pusher d's stores (every lane / every d of an expansion reads its link before any of these run)
    MRTS_HD void astar_store(const AStarBuf& b, uint64_t* V, int d, int link, unsigned okm, unsigned cm, int x, int y, int ws, int g1, int fcur, int root) const {
        if (!((okm >> d) & 1u)) {
            if (d == D_RIGHT && ((okm >> D_LEFT) & 1u)) V[y] |= (1ull << x) >> 1;  // the lane of `right` marks both cells of row y
            return;
        }
        const int nx = x + dir_dx(d), ny = y + dir_dy(d);
        const int np = (ny << ws) | nx;
        const bool closer = (cm >> d) & 1u;
        const unsigned grp = closer ? cm : (okm & ~cm);
        b.cost[np] = (int16_t)g1;
        b.parent[np] = (int16_t)(root >= 0 ? root : d);  // first move of the path to np
        b.next[np] = (int16_t)link;
        if ((grp >> (d + 1)) == 0) b.head[closer ? fcur : fcur + 2] = (int16_t)np;  // the last pusher of this stack
        if (d == D_UP || d == D_DOWN) V[ny] |= 1ull << x;
        else if (d == D_RIGHT) V[y] |= (2ull << x) | (((okm >> D_LEFT) & 1u) ? ((1ull << x) >> 1) : 0ull);
        // (left: marked by the lane of `right`)
    }
    static MRTS_HD int clz32(unsigned v) {
#if defined(__CUDA_ARCH__)
        return __clz((int)v);
#else
        return v ? __builtin_clz(v) : 32;
#endif
    }
    MRTS_HD int astar_rows(int i, int targetpos, int range, int vp) {
        const AStarBuf b = astar_buf();
        const Unit& su = u[i];
        const int ws = ceil_log2(W), WP = 1 << ws;
        const int tx = targetpos % W, ty = targetpos / W;
        const int sq = range * range;
        const int sx = su.x, sy = su.y;
        rows_prepare(vp);
        const uint64_t* F = rows_F();
        uint64_t* V = rows_V();
        // A search only ever pops its start or cells it accepted as enterable neighbours.  If the
        // start is out of range and no enterable cell is in range the reference's search exhausts
        // the reachable map and returns null -- the usual case of a worker queueing behind others
        // around an enclosed target, retried on every frame: same answer here without the search.
        if ((sx - tx) * (sx - tx) + (sy - ty) * (sy - ty) > sq && !rows_any_in_range(F, tx, ty, range)) {
            MRTS_STAT(pre_fail++);
            return -1;
        }
        MRTS_STAT(searches++);
        wsync();
#pragma unroll 1
        for (int y = lane; y < H; y += nl) V[y] = 0;
        wsync();
        const int startPk = (sy << ws) | sx;
        const int h0 = iabs(sx - tx) + iabs(sy - ty);
        if (lane == 0) {
            V[sy] = 1ull << sx;
            b.parent[startPk] = -1;  // no move yet
            b.cost[startPk] = 0;
            b.next[startPk] = -1;
            b.head[h0] = (int16_t)startPk;
        }
        wsync();
        int fmin = h0, fmax = h0, nopen = 1, expanded = 0;
        int result = -1;
#pragma unroll 1
        while (nopen > 0) {
            int pk = b.head[fmin];
#pragma unroll 1
            while (pk < 0 && fmin < fmax) pk = b.head[++fmin];
            if (pk < 0) break;  // (cannot happen while nopen > 0)
            const int nxt = b.next[pk], gcur = b.cost[pk], root = b.parent[pk];
            const int x = pk & (WP - 1), y = pk >> ws;
            wsync();  // every lane has read the stack before its head moves
            if (lane == 0) b.head[fmin] = (int16_t)nxt;
            nopen--;
            if ((x - tx) * (x - tx) + (y - ty) * (y - ty) <= sq) {
                result = root;  // -1 when the start itself is in range (null)
                break;
            }
            MRTS_STAT(expanded++);
            // a search that keeps growing is checked once against a flood fill: unreachable -> null right away
            if (++expanded == ASTAR_GUARD_AFTER) {
                MRTS_STAT(guard++);
                const int t_gd = tick();
                const bool reach = flood_reaches(sx, sy, tx, ty, range);
                st->tp[9] += tick() - t_gd;
                st->tp[10]++;
                if (!reach) {
                    MRTS_STAT(guard_fail++);
                    break;
                }
            }
            const int g1 = gcur + 1, hcur = iabs(x - tx) + iabs(y - ty), fcur = gcur + hcur;
            // enterable and not yet pushed, in the reference's neighbour order: up, right, down, left (:236-262)
            unsigned okm = 0, cm = 0;
#if defined(__CUDA_ARCH__)
            {
                const int d = lane & 3, nx = x + dir_dx(d), ny = y + dir_dy(d);
                bool ok = lane < 4 && nx >= 0 && ny >= 0 && nx < W && ny < H;
                if (ok) ok = ((F[ny] & ~V[ny]) >> nx) & 1ull;
                okm = wballot(ok) & 0xFu;
                cm = wballot(ok && iabs(nx - tx) + iabs(ny - ty) < hcur) & 0xFu;
                const int link = lane < 4 ? astar_link(b, d, okm, cm, x, y, ws, fcur) : -1;
                wsync();  // all links are read before the stacks change
                if (lane < 4) astar_store(b, V, d, link, okm, cm, x, y, ws, g1, fcur, root);
            }
#else
            for (int d = 0; d < 4; d++) {
                const int nx = x + dir_dx(d), ny = y + dir_dy(d);
                if (nx < 0 || ny < 0 || nx >= W || ny >= H) continue;
                if (!(((F[ny] & ~V[ny]) >> nx) & 1ull)) continue;
                okm |= 1u << d;
                if (iabs(nx - tx) + iabs(ny - ty) < hcur) cm |= 1u << d;
            }
            {
                int link[4];
                for (int d = 0; d < 4; d++) link[d] = astar_link(b, d, okm, cm, x, y, ws, fcur);
                for (int d = 0; d < 4; d++) astar_store(b, V, d, link[d], okm, cm, x, y, ws, g1, fcur, root);
            }
#endif
            if (okm & ~cm) fmax = fmax > fcur + 2 ? fmax : fcur + 2;
            nopen += popc(okm);
            wsync();  // the pushes are visible to every lane before the next pop
        }
        st->tp[6]++;
        st->tp[7] += expanded;
        wsync();
#pragma unroll 1
        for (int f = h0 + lane; f <= fmax; f += nl) b.head[f] = -1;  // leave the stacks empty for the next search
        wsync();
        return result;
    }

    // once per game pass, by the whole thread group, before the leader runs (the region is scratch)
    MRTS_HD void astar_init(int tid, int nth) {
        AStarBuf b = astar_buf();
#pragma unroll 1
        for (int k = tid; k < astar_nf(); k += nth) b.head[k] = -1;
        if (W > 64) {
#pragma unroll 1
            for (int k = tid; k < HW; k += nth) b.stamp[k] = 0;
        }
        st->astar_gen = 0;
    }
    // findPathToPositionInRange: direction of the first move, or -1 (null)
#if defined(MRTS_STATS) && !defined(__CUDA_ARCH__)
    int astar_check(int i, int targetpos, int range, int vp) {  // host-only cross-check: the generic search on private arrays
        std::vector<int> cost(HW), parent(HW), next(HW), head(HW + W + H + 2, -1), seen(HW, 0);
        const Unit& su = u[i];
        const int tx = targetpos % W, ty = targetpos / W, sq = range * range;
        const int startPos = su.y * W + su.x, h0 = iabs(su.x - tx) + iabs(su.y - ty);
        parent[startPos] = startPos; cost[startPos] = 0; seen[startPos] = 1; next[startPos] = -1; head[h0] = startPos;
        int fmin = h0, nopen = 1, pos = startPos;
        bool found = false;
        while (nopen > 0) {
            while (head[fmin] < 0) fmin++;
            pos = head[fmin]; head[fmin] = next[pos]; nopen--;
            const int x = pos % W, y = pos / W;
            if ((x - tx) * (x - tx) + (y - ty) * (y - ty) <= sq) { found = true; break; }
            for (int d = 0; d < 4; d++) {
                const int nx = x + dir_dx(d), ny = y + dir_dy(d);
                if (nx < 0 || ny < 0 || nx >= W || ny >= H) continue;
                const int np = nx + ny * W;
                if (seen[np] || !free_cell(nx, ny, vp) || bit_get(bitsC, np)) continue;
                const int f = iabs(nx - tx) + iabs(ny - ty) + cost[pos] + 1;
                cost[np] = cost[pos] + 1; parent[np] = pos; seen[np] = 1; next[np] = head[f]; head[f] = np;
                if (f < fmin) fmin = f;
                nopen++;
            }
        }
        if (!found) return -1;
        int last = pos, par = parent[pos];
        while (par != pos) { last = pos; pos = par; par = parent[pos]; }
        if (last == pos + W) return D_DOWN;
        if (last == pos - 1) return D_LEFT;
        if (last == pos - W) return D_UP;
        if (last == pos + 1) return D_RIGHT;
        return -1;
    }
#endif
    MRTS_HD int astar_first_move(int i, int targetpos, int range, int vp) {
#if defined(MRTS_STATS) && !defined(__CUDA_ARCH__)
        if (W <= 64) {
            const int want = astar_check(i, targetpos, range, vp);
            const int got = astar_rows(i, targetpos, range, vp);
            if (want != got) fprintf(stderr, "[astar mismatch] t=%d unit (%d,%d) target (%d,%d) range %d: rows %d generic %d\n", hdr->time, u[i].x, u[i].y, targetpos % W, targetpos / W, range, got, want);
            return got;
        }
#endif
        if (W <= 64) return astar_rows(i, targetpos, range, vp);
        const AStarBuf b = astar_buf();
        const Unit& su = u[i];
        const uint16_t gen = (uint16_t)++st->astar_gen;
        const int tx = targetpos % W, ty = targetpos / W;
        const int sq = range * range;
        const int startPos = su.y * W + su.x;
        const int h0 = iabs(su.x - tx) + iabs(su.y - ty);
        // A search only ever pops its start or cells it accepted as enterable neighbours.  If the
        // start is out of range and no enterable cell is in range the reference's search exhausts
        // the reachable map and returns null -- the usual case of a worker queueing behind others
        // around an enclosed target, retried on every frame: same answer here without the search.
        if ((su.x - tx) * (su.x - tx) + (su.y - ty) * (su.y - ty) > sq) {
            bool any = false;
#pragma unroll 1
            for (int dy = -range; dy <= range && !any; dy++) {
#pragma unroll 1
                for (int dx = -range; dx <= range && !any; dx++) {
                    const int x = tx + dx, y = ty + dy;
                    if (dx * dx + dy * dy > sq || x < 0 || y < 0 || x >= W || y >= H) continue;
                    any = free_cell(x, y, vp) && !bit_get(bitsC, x + y * W);
                }
            }
            if (!any) {
                MRTS_STAT(pre_fail++);
                return -1;
            }
        }
        MRTS_STAT(searches++);
        st->tp[6]++;
        b.parent[startPos] = (int16_t)startPos;
        b.cost[startPos] = 0;
        b.stamp[startPos] = gen;
        b.next[startPos] = -1;
        b.head[h0] = (int16_t)startPos;
        int fmin = h0, fmax = h0, flo = h0, nopen = 1;
        bool found = false;
        int pos = startPos;
#pragma unroll 1
        while (nopen > 0) {
            while (b.head[fmin] < 0) fmin++;
            pos = b.head[fmin];
            b.head[fmin] = b.next[pos];
            nopen--;
            const int x = pos % W, y = pos / W;
            if ((x - tx) * (x - tx) + (y - ty) * (y - ty) <= sq) {
                found = true;
                break;
            }
            MRTS_STAT(expanded++);
            st->tp[7]++;
            const int g1 = b.cost[pos] + 1;
            // neighbours in the reference's order: up, right, down, left (:236-262)
#pragma unroll
            for (int d = 0; d < 4; d++) {
                const int nx = x + dir_dx(d), ny = y + dir_dy(d);
                if (nx < 0 || ny < 0 || nx >= W || ny >= H) continue;
                const int np = nx + ny * W;
                if (b.stamp[np] == gen) continue;
                // free(nbr) = gs.free() unless pre-seeded false by `ru` (bitsC); both are constant during a search
                if (!free_cell(nx, ny, vp) || bit_get(bitsC, np)) continue;
                const int f = iabs(nx - tx) + iabs(ny - ty) + g1;
                b.cost[np] = (int16_t)g1;
                b.parent[np] = (int16_t)pos;
                b.stamp[np] = gen;
                b.next[np] = b.head[f];
                b.head[f] = (int16_t)np;
                if (f < fmin) fmin = f;
                if (f < flo) flo = f;
                if (f > fmax) fmax = f;
                nopen++;
            }
        }
#pragma unroll 1
        for (int f = flo; f <= fmax; f++) b.head[f] = -1;  // leave the stacks empty for the next search
        if (!found) return -1;
        int last = pos, parent = b.parent[pos];
#pragma unroll 1
        while (parent != pos) {
            last = pos;
            pos = parent;
            parent = b.parent[pos];
        }
        if (last == pos + W) return D_DOWN;
        if (last == pos - 1) return D_LEFT;
        if (last == pos - W) return D_UP;
        if (last == pos + 1) return D_RIGHT;
        return -1;
    }

    // Train.score, J/ai/abstraction/Train.java:98-127
    MRTS_HD int train_score(int x, int y, int ut, int player, int vp) const {
        const int n = nunits();
        const bool harv = ut_can_harvest(ut);
        int best = NO_KEY;
        wsync();
#pragma unroll 1
        for (int i = lane; i < n; i += nl) {  // one unit per lane
            if (!inview(i, vp)) continue;
            const Unit& o = u[i];
            const bool cand = harv ? ut_is_resource(o.type) : (o.player >= 0 && o.player != player);
            if (!cand) continue;
            const int d = iabs(o.x - x) + iabs(o.y - y);
            if (d < best) best = d;
        }
        best = wmin(best);
        return best == NO_KEY ? 0 : -best;
    }

    // AbstractAction.execute: true + primitive action in `out`, or false (null).  ti = index of the entry's
    // target unit in the bot's view (Harvest: the resource, or the base when carrying; Attack: the target), -1 = gone.
    // Every kind that walks first decides (target cell, range); the ONE path search below serves them all
    // (the rules are instruction-fetch bound: each big function is inlined exactly once).
    MRTS_HD bool aa_execute(AbstractAct& r, int i, int ti, int vp, const int* costB, PAct& out) {
        const Unit& un = u[i];
        out.unit = (uint16_t)i;
        out.utype = -1;
        out.param = -1;
        out.x = out.y = 0;
        out.pad = 0;
        const int kind = r.kind;
        int tpos = 0, range = 0, tx = 0, ty = 0;
        bool search = false;
        switch (kind) {
            case AA_TRAIN: {  // Train.java:48-96
                const int x = un.x, y = un.y;
                int best_dir = -1, best_score = -1;
#pragma unroll 1
                for (int d = 0; d < 4; d++) {
                    int nx = x + dir_dx(d), ny = y + dir_dy(d);
                    if (nx < 0 || ny < 0 || nx >= W || ny >= H) continue;
                    if (!free_cell(nx, ny, vp)) continue;
                    int score = train_score(nx, ny, r.utype, un.player, vp);
                    if (score > best_score || best_dir == -1) {
                        best_score = score;
                        best_dir = d;
                    }
                }
                r.target = 1;  // completed = true
                if (best_dir == -1) return false;
                out.type = T_PRODUCE;
                out.param = (int16_t)best_dir;
                out.utype = r.utype;
                break;
            }
            case AA_BUILD:  // Build.java:56-78
                tx = r.x, ty = r.y, range = 1, search = true;
                break;
            case AA_MOVE:  // Move.java:53-59: findPath = findPathToPositionInRange(.., 0, ..)
                tx = r.x, ty = r.y, range = 0, search = true;
                break;
            case AA_HARVEST:  // Harvest.java:71-114
                if (ti < 0) return false;
                tx = u[ti].x, ty = u[ti].y, range = 1, search = true;
                break;
            case AA_ATTACK: {  // Attack.java:50-65
                if (ti < 0) return false;
                tx = u[ti].x, ty = u[ti].y;
                const int dx = tx - un.x, dy = ty - un.y;
                range = ut_range(un.type);
                if (dx * dx + dy * dy <= range * range) {
                    out.type = T_ATTACK;
                    out.x = (int16_t)tx;
                    out.y = (int16_t)ty;
                    return true;
                }
                search = true;
                break;
            }
            default: return false;
        }
        if (search) {
            tpos = tx + ty * W;
            const int t_as = tick();
            const int mv = astar_first_move(i, tpos, range, vp);
            st->tp[8] += tick() - t_as;
            if (mv >= 0) {
                out.type = T_MOVE;
                out.param = (int16_t)mv;
            } else {
                if (kind == AA_MOVE || kind == AA_ATTACK) return false;
                // Build / Harvest: act on the adjacent target
                int d = -1;
                if (tx == un.x && ty == un.y - 1) d = D_UP;
                else if (tx == un.x + 1 && ty == un.y) d = D_RIGHT;
                else if (tx == un.x && ty == un.y + 1) d = D_DOWN;
                else if (tx == un.x - 1 && ty == un.y) d = D_LEFT;
                if (d < 0) return false;
                out.param = (int16_t)d;
                if (kind == AA_HARVEST) {  // Harvest.java:97-113: no isUnitActionAllowed test
                    out.type = un.res != 0 ? T_RETURN : T_HARVEST;
                    return true;
                }
                out.type = T_PRODUCE;
                out.utype = r.utype;
            }
        }
        return action_allowed(i, out.type, out.param, out.utype, vp, costB);
    }

    // AbstractionLayerAI.translateActions, AbstractionLayerAI.java:58-113 -> pa2
    MRTS_HD void translate_actions(int player, int vp, bool enemy_seen) {
        wsync();
        int costB[2];
        pending_usage(bitsB, costB, vp);
        bits_clear(bitsC);
        st->rows_ready = 0;
        st->npa2 = 0;
        int na = hdr->n_aa, keep = 0;
#pragma unroll 1
        for (int k = 0; k < na; k++) {
            AbstractAct r = aa[k];
            const int owner0 = r.kind & AA_OWNER0;
            if ((owner0 != 0) != (player == 0)) {  // the other bot's entry (bot-vs-bot games)
                aa[keep++] = r;
                continue;
            }
            r.kind &= AA_KIND_MASK;
            const int i = find_uid(r.unit, vp);
            bool del = false;
            if (i < 0) del = true;
            else if (enemy_seen && u[i].a_type >= 0 && (u[i].flags & UF_AA_ATTACK) && ut_can_harvest(u[i].type)) {
                // busy worker with an Attack entry: the reference's per-frame re-put keeps it alive (see worker_rush)
            } else {
                // the unit the entry refers to (Harvest.java:42-48, Attack.java:30-33): gone from the view = completed
                int ti = -1;
                if (r.kind == AA_HARVEST) ti = find_uid(u[i].res > 0 ? r.base : r.target, vp);
                else if (r.kind == AA_ATTACK) ti = find_uid(r.target, vp);
                bool completed;
                switch (r.kind) {
                    case AA_TRAIN: completed = r.target != 0; break;                 // Train.java:30-32
                    case AA_BUILD: completed = unit_at_v(r.x, r.y, vp) >= 0; break;  // Build.java:33-37
                    case AA_HARVEST:
                    case AA_ATTACK: completed = ti < 0; break;
                    case AA_MOVE: completed = u[i].x == r.x && u[i].y == r.y; break;  // Move.java:32-34
                    default: completed = true;
                }
                if (completed) del = true;
                else if (u[i].a_type < 0) {
                    PAct out;
                    if (aa_execute(r, i, ti, vp, costB, out)) {
                        pa2[st->npa2++] = out;
                        int pos = usage_pos(u[i], out.type, out.param);
                        if (pos != -0x7fff) {
                            bit_set(bitsC, pos);
                            rows_reserve(pos);
                        }
                    }
                }
            }
            if (!del) {
                r.kind |= (uint8_t)owner0;
                aa[keep++] = r;  // r may carry Train's completed flag
            }
            else if (i >= 0) u[i].flags &= (uint8_t)~UF_AA_ATTACK;
        }
        hdr->n_aa = (uint16_t)keep;
        // accept desires first-come-first-served against the pending usage
#pragma unroll 1
        for (int w = 0; w < nbits_words; w++) bitsA[w] = bitsB[w];
        int used0 = costB[0], used1 = costB[1];
        int kept = 0;
#pragma unroll 1
        for (int k = 0; k < st->npa2; k++) {
            PAct d = pa2[k];
            const Unit& un = u[d.unit];
            int pos = usage_pos(un, d.type, d.param);
            int c = d.type == T_PRODUCE ? ut_cost(d.utype) : 0;
            if (consistent_single(bitsA, un.player == 0 ? used0 : used1, pos, un.player, c)) {
                if (pos != -0x7fff) bit_set(bitsA, pos);
                if (un.player == 0) used0 += c;
                else used1 += c;
                pa2[kept++] = d;
            }
        }
        fill_with_nones(pa2, kept, player, 10, vp);
        st->npa2 = kept;
    }

    // AbstractionLayerAI.findBuildingPosition, AbstractionLayerAI.java:143-228.  `reserved` = the one position an
    // earlier build of this frame took (LightRush: base, then barracks), NO_POS when the list is empty.
    enum { NO_POS = -2 };
    MRTS_HD int find_building_position(int dX, int dY, int vp, int reserved = NO_POS) const {
        const int mx = H > W ? H : W;
#pragma unroll 1
        for (int l = 1; l < mx; l++) {
            for (int side = 0; side < 4; side++) {
                const bool horiz = (side == 0 || side == 2);
                const int fixed = side == 0 ? dY - l : side == 1 ? dX + l : side == 2 ? dY + l : dX - l;
                if (fixed < 0 || fixed >= (horiz ? H : W)) continue;
#pragma unroll 1
                for (int dd = -l; dd <= l; dd++) {
                    int x = horiz ? dX + dd : fixed;
                    int y = horiz ? fixed : dY + dd;
                    if (x < 0 || y < 0 || x >= W || y >= H) continue;
                    if (x + y * W != reserved && !wall_at(x, y) && unit_at_v(x, y, vp) < 0) return x + y * W;
                }
            }
        }
        return -1;
    }

    // ---- the puts of one getAction call.  A bot's getAction makes its puts in a fixed order (the table's insertion
    // order decides who wins a resource conflict in translateActions); here the call first QUEUES them as jobs in
    // that order (worker_rush_jobs / light_rush_jobs) and one loop (bot_jobs) executes the queue, so that each
    // decision routine below is inlined exactly once.  The queue lives in pa2 (free until translate_actions):
    // at most two jobs per unit (build or harvest, then attack when it stays free), 4 bytes each.
    enum { JOB_TRAIN = 0, JOB_MELEE = 1, JOB_BUILD = 2, JOB_HARVEST = 3 };
    struct Job {
        uint16_t unit;  // unit index
        uint8_t kind;
        int8_t utype;   // Train / Build
    };
    static MRTS_HD Job make_job(int unit, int kind, int utype) {
        Job j;
        j.unit = (uint16_t)unit;
        j.kind = (uint8_t)kind;
        j.utype = (int8_t)utype;
        return j;
    }
    // one unit per lane: the picked ones join the queue in unit order
    MRTS_HD void jobs_append(Job* q, int& nq, bool pick, int i, int kind, int utype) const {
        const unsigned m = wballot(pick);
        if (pick) q[nq + popc(m & lanes_below())] = make_job(i, kind, utype);
        nq += popc(m);
    }

    // WorkerRush.meleeUnitBehavior, WorkerRush.java:106-122: Attack(closest enemy); the PO* variants explore when
    // none is in view.  true = `r` holds the put.
    MRTS_HD bool melee_target(int i, int player, int vp, AbstractAct& r) {
        const int n = nunits();
        const int ux = u[i].x, uy = u[i].y;
        int best = NO_KEY;
        wsync();
#pragma unroll 1
        for (int j = lane; j < n; j += nl) {  // one unit per lane
            if (!inview(j, vp)) continue;
            const Unit& o = u[j];
            if (o.player < 0 || o.player == player) continue;
            const int key = argmin_key(iabs(o.x - ux) + iabs(o.y - uy), j);
            if (key < best) best = key;
        }
        best = wmin(best);
        const int closest = best == NO_KEY ? -1 : (best & 0xFFFF);
        if (closest >= 0) {
            r.kind = AA_ATTACK;
            r.target = u[closest].uid;
            return true;
        }
        if (!(st->explore && vp >= 0)) return false;
        // PO*.meleeUnitBehavior (partialobservability/POWorkerRush.java:56-76): the state is a
        // PartiallyObservableGameState (vp >= 0) and shows no enemy -> Move to the nearest cell that is not
        // observable (row-major scan with strict <: lowest cell index among equal distances)
        if (!st->unobs_ready) {  // bitsC (free until translate_actions): bit c = cell c is NOT within sight of an own unit
#pragma unroll 1
            for (int base = 0; base < HW; base += nl) {
                const int c = base + lane;
                bool seen = true;
                if (c < HW) {
                    const int cx = c % W, cy = c / W;
                    seen = false;
#pragma unroll 1
                    for (int j = 0; j < n && !seen; j++) {
                        const Unit& o = u[j];
                        if (o.player != player) continue;
                        const int dx = o.x - cx, dy = o.y - cy, rr = ut_sight(o.type);
                        seen = dx * dx + dy * dy <= rr * rr;
                    }
                }
#if defined(__CUDA_ARCH__)
                const unsigned word = wballot(!seen);
                if (lane == 0) bitsC[base >> 5] = word;
#else
                if ((c & 31) == 0) bitsC[c >> 5] = 0;
                if (!seen) bitsC[c >> 5] |= 1u << (c & 31);
#endif
            }
            st->unobs_ready = 1;
            wsync();
        }
        int bestc = NO_KEY;
#pragma unroll 1
        for (int c = lane; c < HW; c += nl) {
            if (!((bitsC[c >> 5] >> (c & 31)) & 1u)) continue;
            const int dx = ux - c % W, dy = uy - c / W;
            const int key = ((dx * dx + dy * dy) << 14) | c;
            if (key < bestc) bestc = key;
        }
        bestc = wmin(bestc);
        if (bestc == NO_KEY) return false;
        const int c = bestc & 0x3FFF;
        r.kind = AA_MOVE;
        r.x = (int16_t)(c % W);
        r.y = (int16_t)(c / W);
        return true;
    }
    // WorkerRush.workersBehavior tail, WorkerRush.java:148-197 (LightRush.java:213-244): nearest resource / own
    // stockpile; `put` = (re)assign Harvest in `r`; returns true when the worker stays free.  ke = its table entry.
    MRTS_HD bool harvest_target(int w, int player, int vp, int ke, AbstractAct& r, bool& put) {
        const int n = nunits();
        const int wx = u[w].x, wy = u[w].y;
        int kres = NO_KEY, kbase = NO_KEY;
        wsync();
#pragma unroll 1
        for (int j = lane; j < n; j += nl) {  // one unit per lane: nearest resource and nearest own stockpile
            if (!inview(j, vp)) continue;
            const int d = iabs(u[j].x - wx) + iabs(u[j].y - wy);
            if (ut_is_resource(u[j].type)) {
                const int key = argmin_key(d, j);
                if (key < kres) kres = key;
            }
            if (ut_is_stockpile(u[j].type) && u[j].player == player) {
                const int key = argmin_key(d, j);
                if (key < kbase) kbase = key;
            }
        }
        kres = wmin(kres);
        kbase = wmin(kbase);
        const int cres = kres == NO_KEY ? -1 : (kres & 0xFFFF), cbase = kbase == NO_KEY ? -1 : (kbase & 0xFFFF);
        const bool isharv = ke >= 0 && (aa[ke].kind & AA_KIND_MASK) == AA_HARVEST;
        r.kind = AA_HARVEST;
        if (u[w].res > 0) {
            if (cbase < 0) return true;
            if (!isharv || aa[ke].base != u[cbase].uid) {
                r.base = u[cbase].uid;  // Harvest(null, base)
                put = true;
            }
            return false;
        }
        if (cres < 0 || cbase < 0) return true;
        if (!isharv || aa[ke].target != u[cres].uid || aa[ke].base != u[cbase].uid) {
            r.target = u[cres].uid;
            r.base = u[cbase].uid;
            put = true;
        }
        return false;
    }

    // WorkerRush.getAction, J/ai/abstraction/WorkerRush.java:63-204 -> job queue.
    //
    // The reference re-puts Attack(closest enemy) for EVERY free worker on every frame, busy or not
    // (workersBehavior, :198-203).  For a worker that is busy and whose table entry already is an
    // Attack the re-put is unobservable: the slot is kept (LinkedHashMap.put), the target is only read
    // when the worker is idle (that frame's put is executed here as well), and because the fresh
    // target is a unit of this frame's view the entry is never "completed" in translateActions.  So
    // such workers are skipped and translate_actions keeps their entry -- provided an enemy is in
    // view at all (otherwise the reference puts nothing and the stale entry is checked as usual).
    MRTS_HD int worker_rush_jobs(int player, int vp, Job* q, bool& enemy_seen_out) {
        wsync();
        const int n = nunits();
        const int pres = hdr->res[player];
        // census (:124-137) + is any enemy in view: one unit per lane
        int nbases = 0, nworkers = 0, kw0 = NO_KEY, kw1 = NO_KEY;
        bool enemy_seen = false;
#pragma unroll 1
        for (int base = 0; base < n; base += nl) {
            const int i = base + lane;
            bool is_base = false, is_worker = false, is_enemy = false;
            if (i < n) {
                const Unit& un = u[i];
                if (un.player == player) {
                    is_base = un.type == UT_BASE;
                    is_worker = ut_can_harvest(un.type);
                } else {
                    is_enemy = un.player >= 0 && inview(i, vp);
                }
            }
            const unsigned mw = wballot(is_worker);
            nbases += popc(wballot(is_base));
            enemy_seen = enemy_seen || wballot(is_enemy) != 0;
            if (mw) {  // first and second worker in unit order
                const int f = base + ffs32(mw) - 1;
                if (kw0 == NO_KEY) {
                    kw0 = f;
                    const unsigned rest = mw & (mw - 1);
                    if (rest) kw1 = base + ffs32(rest) - 1;
                } else if (kw1 == NO_KEY) {
                    kw1 = f;
                }
            }
            nworkers += popc(mw);
        }
        enemy_seen_out = enemy_seen;
        const int w0 = kw0 == NO_KEY ? -1 : kw0, w1 = kw1 == NO_KEY ? -1 : kw1;
        int nq = 0;
#pragma unroll 1
        for (int base = 0; base < n; base += nl) {  // idle bases train (:102-104)
            const int i = base + lane;
            jobs_append(q, nq, i < n && u[i].type == UT_BASE && u[i].player == player && u[i].a_type < 0 && pres >= ut_cost(UT_WORKER), i, JOB_TRAIN, UT_WORKER);
        }
#pragma unroll 1
        for (int base = 0; base < n; base += nl) {  // idle melee units attack (:106-122)
            const int i = base + lane;
            jobs_append(q, nq, i < n && ut_can_attack(u[i].type) && !ut_can_harvest(u[i].type) && u[i].player == player && u[i].a_type < 0, i, JOB_MELEE, -1);
        }
        // workersBehavior, WorkerRush.java:124-204
        if (nworkers > 0) {
            const bool want_base = nbases == 0 && pres >= ut_cost(UT_BASE);
            const int builder = want_base ? w0 : -1;
            const int harvester = want_base ? w1 : w0;
            wsync();
            if (builder >= 0) q[nq++] = make_job(builder, JOB_BUILD, UT_BASE);      // buildIfNotAlreadyBuilding, AbstractionLayerAI.java:230-245
            if (harvester >= 0) q[nq++] = make_job(harvester, JOB_HARVEST, -1);     // a harvester that stays free attacks last (bot_jobs)
            wsync();
#pragma unroll 1
            for (int base = 0; base < n; base += nl) {  // every other worker attacks, in unit order
                const int i = base + lane;
                bool need = false;
                if (i < n) {
                    const Unit& un = u[i];
                    need = un.player == player && ut_can_harvest(un.type) && i != builder && i != harvester &&
                           !(enemy_seen && un.a_type >= 0 && (un.flags & UF_AA_ATTACK));  // unobservable re-put, see above
                }
                jobs_append(q, nq, need, i, JOB_MELEE, -1);
            }
        }
        wsync();
        return nq;
    }

    // LightRush.getAction, J/ai/abstraction/LightRush.java:77-254 -> job queue.  Every put of the reference is made,
    // in its order: bases, barracks, melee units, then the workers (builders, harvesters, the rest attack).
    MRTS_HD int light_rush_jobs(int player, int vp, int rush_ut, Job* q) {  // rush_ut: what the barracks train (UT_LIGHT / UT_HEAVY / UT_RANGED)
        wsync();
        const int n = nunits();
        const int pres = hdr->res[player];
        int nbases = 0, nbarracks = 0, nworkers = 0;
#pragma unroll 1
        for (int base = 0; base < n; base += nl) {  // census, one unit per lane (a bot sees only its view)
            const int i = base + lane;
            const bool mine = i < n && u[i].player == player && inview(i, vp);
            nbases += popc(wballot(mine && u[i].type == UT_BASE));
            nbarracks += popc(wballot(mine && u[i].type == UT_BARRACKS));
            nworkers += popc(wballot(mine && u[i].type == UT_WORKER));
        }
        int nq = 0;
        // three passes in the reference's order (:93-119): idle bases, idle barracks, idle melee units
#pragma unroll 1
        for (int pass = 0; pass < 3; pass++) {
            if (pass == 0 && !(nworkers < 1 && pres >= ut_cost(UT_WORKER))) continue;  // baseBehavior, :131-142
            if (pass == 1 && pres < ut_cost(rush_ut)) continue;                       // barracksBehavior, :144-148
#pragma unroll 1
            for (int base = 0; base < n; base += nl) {
                const int i = base + lane;
                bool pick = false;
                if (i < n && u[i].player == player && u[i].a_type < 0 && inview(i, vp)) {
                    const int t = u[i].type;
                    pick = pass == 0 ? t == UT_BASE : pass == 1 ? t == UT_BARRACKS : (ut_can_attack(t) && !ut_can_harvest(t));
                }
                jobs_append(q, nq, pick, i, pass == 2 ? JOB_MELEE : JOB_TRAIN, pass == 0 ? UT_WORKER : rush_ut);
            }
        }
        // workersBehavior, :160-254: the workers in unit order; the first may build a base, the next a barracks,
        // every other one harvests (and attacks when it stays free: queued by bot_jobs behind all of these)
        int used = 0, taken = 0;
        const bool want_base = nbases == 0 && pres >= ut_cost(UT_BASE);
        wsync();
#pragma unroll 1
        for (int base = 0; base < n; base += nl) {
            const int i = base + lane;
            unsigned m = wballot(i < n && u[i].player == player && ut_can_harvest(u[i].type) && inview(i, vp));
#pragma unroll 1
            while (m) {
                const int j = base + ffs32(m) - 1;
                m &= m - 1;
                if (taken == 0 && want_base) {
                    q[nq++] = make_job(j, JOB_BUILD, UT_BASE);
                    used += ut_cost(UT_BASE);
                    taken = 1;
                    continue;
                }
                if (taken <= 1 && nbarracks == 0 && pres >= ut_cost(UT_BARRACKS) + used) {
                    if (taken == 0 || want_base) {  // at most one barracks builder: the first worker left after the base builder
                        q[nq++] = make_job(j, JOB_BUILD, UT_BARRACKS);
                        used += ut_cost(UT_BARRACKS);
                        taken = 2;
                        continue;
                    }
                }
                if (taken < 2) taken = 2;  // the builder slots are decided by the first workers only
                q[nq++] = make_job(j, JOB_HARVEST, -1);
            }
        }
        wsync();
        return nq;
    }

    // executes the queue; a harvester that stays free joins its tail as an attacker (WorkerRush.java:198-203,
    // LightRush.java:246-253).  `reserved` = the position an earlier build of this frame took (LightRush: base, then barracks).
    MRTS_HD void bot_jobs(Job* q, int nq, int player, int vp) {
        int reserved = NO_POS;
#pragma unroll 1
        for (int k = 0; k < nq; k++) {
            wsync();
            const Job jb = q[k];
            const int j = jb.unit;
            const int ke = aa_find(u[j].uid);  // the unit's entry in the table, -1 = none
            AbstractAct r;
            r.unit = u[j].uid;
            r.kind = AA_TRAIN;
            r.utype = -1;
            r.x = r.y = 0;
            r.target = NO_UID;
            r.base = NO_UID;
            bool put = false;
            switch (jb.kind) {
                case JOB_TRAIN:
                    r.utype = jb.utype;
                    r.target = 0;  // Train's `completed` flag
                    put = true;
                    break;
                case JOB_MELEE: put = melee_target(j, player, vp, r); break;
                case JOB_BUILD: {  // buildIfNotAlreadyBuilding, AbstractionLayerAI.java:230-245
                    if (ke >= 0 && (aa[ke].kind & AA_KIND_MASK) == AA_BUILD && aa[ke].utype == jb.utype) break;
                    const int pos = find_building_position(u[j].x, u[j].y, vp, reserved);
                    r.kind = AA_BUILD;
                    r.utype = jb.utype;
                    r.x = (int16_t)(pos % W);  // pos == -1 -> (-1, 0): Java % and / truncate like C++
                    r.y = (int16_t)(pos / W);
                    reserved = pos;  // reservedPositions.add(pos), also when it is -1
                    put = true;
                    break;
                }
                case JOB_HARVEST:
                    if (harvest_target(j, player, vp, ke, r, put)) {
                        wsync();
                        q[nq++] = make_job(j, JOB_MELEE, -1);
                        wsync();
                    }
                    break;
            }
            if (put) aa_put(r, j, ke);
        }
        wsync();
    }

    // the scripted opponent's PlayerAction for `player` -> pa2 (heavy kernels only; PassiveAI.getAction,
    // J/ai/PassiveAI.java:41-45, has no table entries: translate_actions then only fills NONE(10))
    MRTS_HD void bot_action(int which, int player, int vp) {
        st->npa2 = 0;
        st->explore = ai_explores(which);
        st->unobs_ready = 0;
        wsync();
        const int t_puts = tick();
        Job* const q = (Job*)pa2;
        int nq = 0;
        bool enemy_seen = false;
        if (which == AI_WORKER_RUSH || which == AI_PO_WORKER_RUSH) nq = worker_rush_jobs(player, vp, q, enemy_seen);
        else if (ai_barracks_rush(which)) nq = light_rush_jobs(player, vp, ai_rush_unit(which), q);
        bot_jobs(q, nq, player, vp);
        const int t_tr = tick();
        st->tp[1] = t_tr - t_puts;
        translate_actions(player, vp, enemy_seen);
        st->tp[2] = tick() - t_tr;
    }

    // ================================================================== one environment step
    // The agents' actions were staged per unit by prefetch_actions.
    // On return: st->rew[][] holds the six reward terms per view, st->gameover/st->winner are set, the unit
    // array still contains units killed this cycle (flagged) so that stale partially observable
    // views can be encoded; the caller then runs compact() or reloads the map template.
    //
    // HEAVY = the game has a scripted player built on AbstractionLayerAI (table + path searches); light kernels
    // (self-play, PassiveAI) are compiled without that code.  The three clients of the reference differ only in
    // the order of four operations -- {player 0, player 1} x {make the PlayerAction, issueSafe it} -- so one loop
    // runs them and every rule function below has a single call site:
    //   JNIGridnetClientSelfPlay.gameStep (:159-189)   act 0, issue 0, act 1, issue 1  (player 1 decodes after player 0's issue)
    //   JNIGridnetClient.gameStep (:163-196)           act 0, act 1, issue 0, issue 1  (views of both taken before any issue)
    //   JNIBotClient.gameStep (JNIBotClient.java:108-113)   same order, both players scripted, full state
    template <int HEAVY>
    MRTS_HD bool step_game() {
        wsync();
        const bool selfplay = ai == AI_NONE;
        const bool views = partial && ai1 == AI_NONE;  // bot-vs-bot clients ignore partialObs
#pragma unroll 1
        for (int k = 0; k < 4; k++) {
            const int p = selfplay ? (k >> 1) : (k & 1);
            const bool act = selfplay ? (k & 1) == 0 : k < 2;
            const bool first_list = selfplay || p == 0;  // self-play games reuse pa for player 1 (pa2 belongs to bot games)
            PAct* const list = first_list ? pa : pa2;
            const int t0 = tick();
            if (act) {
                if (views && (selfplay || k == 0)) {
                    const int vlast = selfplay ? p : 1;
#pragma unroll 1
                    for (int v = selfplay ? p : 0; v <= vlast; v++) make_view(v);
                }
                const int vp = views ? p : -1;
                const int which = p == 0 ? ai1 : ai;
                if (which == AI_NONE) {
                    decode_actions(p, vp);  // -> pa
                    st->tp[0] = tick() - t0;
                } else if (HEAVY) {
                    bot_action(which, p, vp);  // -> pa2
                    if (p == 0) {              // a scripted player 0: its action is issued first, from pa
                        wsync();
                        const int n0 = st->npa2;
#pragma unroll 1
                        for (int e = lane; e < n0; e += nl) pa[e] = pa2[e];
                        st->npa = n0;
                        wsync();
                    }
                } else {  // PassiveAI.getAction, J/ai/PassiveAI.java:41-45
                    int cnt = 0;
                    fill_with_nones(list, cnt, p, 10, vp);
                    if (first_list) st->npa = cnt;
                    else st->npa2 = cnt;
                }
            } else {
                issue_safe(list, first_list ? st->npa : st->npa2);
                st->tp[3 + p] = tick() - t0;
            }
        }
        const int tc = tick();
        cycle();
        eval_gameover();
        st->tp[5] = tick() - tc;
        if (st->gameover) {  // WinLossRewardFunction.java:16-24
            st->rew[0][0] = st->winner == 0 ? 1 : -1;
            st->rew[1][0] = st->winner == 1 ? 1 : -1;
        }
        hdr->steps++;
        return st->gameover || hdr->steps >= max_steps;  // JNIGridnetVecClient.java:245-266 / :270-286
    }
};

}  // namespace mrts
