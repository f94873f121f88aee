// TEST INFRASTRUCTURE -- NOT PRODUCT CODE, NOT A FALLBACK.
//
// Host build of the very same rule / encode headers the CUDA kernels are compiled from
// (microrts_py_b200/csrc/mrts_{core,encode,step}.h), driven with a thread group of size 1 and a
// no-op barrier.  It exists so that the game logic can be debugged against the oracle in the
// build container, which has no GPU.  Nothing in microrts_py_b200/ loads it; the product path
// fails loudly when libmrts_b200.so or a CUDA device is missing.
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../microrts_py_b200/csrc/mrts_setup.h"

using namespace mrts;

struct hs_env {
    HostSetup S;
    std::vector<uint8_t> games;
    std::vector<uint8_t> mem;
    std::vector<int32_t> cycle;  // map rotation (the host loop standing in for mrts_cycle_kernel)
    std::vector<uint8_t> ended;
    int cycle_pos = 0, n_maps = 0;
    // MRTS_HOSTSIM_SPLIT=1: outputs go through the packed form (observation words + mask entries) and are
    // expanded afterwards, as the split pipeline of the CUDA library does (mrts_expand.cuh)
    bool split = false;
    std::vector<uint32_t> pobs, ment;
};
struct NoSync {
    void operator()() const {}
};
static std::string g_err;

static void pass(hs_env* e, int mode, const int32_t* actions, void* obs, void* mask, double* reward, uint8_t* done) {
    EnvParams P = e->S.P;
    P.games = e->games.data();
    P.maps = e->S.tpls.data();
    P.grid_pool = e->S.grids.data();
    P.unit_pool = e->S.units.data();
    P.game_map = e->S.game_map.data();
    P.game_ai = e->S.game_ai.data();
    P.game_ai1 = e->S.game_ai1.data();
    P.mode = mode;
    P.actions = actions;
    P.obs = obs;
    P.mask = mask;
    P.reward = reward;
    P.done = done;
    const bool rotate = mode == MODE_STEP && !e->cycle.empty();
    P.ended = rotate ? e->ended.data() : nullptr;
    auto one = [&](const EnvParams& Q0, int g) {
        EnvParams Q = Q0;
        if (!e->split || (!Q.obs && !Q.mask)) {
            game_pass_any(Q, g, e->mem.data(), 0, 1, NoSync());
            return;
        }
        const size_t ms = ment_stride_words(Q.cap);
        Q.obs = nullptr, Q.mask = nullptr;
        Q.pobs = Q0.obs ? e->pobs.data() : nullptr;
        Q.ment = Q0.mask ? e->ment.data() : nullptr;
        game_pass_any(Q, g, e->mem.data(), 0, 1, NoSync());
        const bool sp = Q.game_ai[g] == AI_NONE;
        const int env0 = sp ? 2 * g : Q.num_selfplay + (g - Q.num_selfplay / 2);
        for (int v = 0; v < (sp ? 2 : 1); v++) {
            const size_t env = (size_t)(env0 + v);
            if (Q0.obs) store_obs_rt(Q.elem, Q.planes, e->pobs.data() + env * Q.HW, Q.HW, (uint8_t*)Q0.obs + env * Q.HW * Q.planes * Q.elem, 0, 1);
            if (Q0.mask) {
                void* m = (uint8_t*)Q0.mask + env * Q.HW * 78 * Q.elem;
                if (Q.elem == 4) expand_mask_entries<4>(e->ment.data() + env * ms, Q.HW, m);
                else expand_mask_entries<1>(e->ment.data() + env * ms, Q.HW, m);
            }
        }
    };
    for (int g = 0; g < P.n_games; g++) one(P, g);
    if (rotate) {  // same rule as mrts_cycle_kernel + the reset launch (mrts_api.cu)
        EnvParams R = P;
        R.mode = MODE_RESET;
        R.actions = nullptr;
        R.reward = nullptr;
        R.done = nullptr;
        R.ended = nullptr;
        for (int g = 0; g < P.n_games; g++)
            if (e->ended[g]) {
                e->S.game_map[g] = e->cycle[e->cycle_pos];
                e->cycle_pos = (e->cycle_pos + 1) % (int)e->cycle.size();
                one(R, g);
            }
    }
}

extern "C" {
const char* hs_last_error() { return g_err.c_str(); }
int hs_create(const mrts_config* cfg, hs_env** out) {
    hs_env* e = new hs_env();
    int rc = build_setup(cfg, e->S, g_err);
    if (rc) {
        delete e;
        return rc;
    }
    const EnvParams& P = e->S.P;
    e->n_maps = cfg->n_maps;
    e->ended.assign(P.n_games, 0);
    e->games.assign((size_t)P.n_games * P.game_stride + 64, 0);
    e->mem.assign(layout_of(P).total + 64, 0);
    const char* sp = getenv("MRTS_HOSTSIM_SPLIT");
    e->split = sp && sp[0] == '1';
    e->pobs.assign((size_t)e->S.N * P.HW, 0);
    e->ment.assign((size_t)e->S.N * ment_stride_words(P.cap), 0);
    pass(e, MODE_RESET, nullptr, nullptr, nullptr, nullptr, nullptr);
    *out = e;
    return 0;
}
void hs_destroy(hs_env* e) { delete e; }
int hs_reset(hs_env* e, void* obs, void* mask) {
    pass(e, MODE_RESET, nullptr, obs, mask, nullptr, nullptr);
    return 0;
}
int hs_step(hs_env* e, const int32_t* actions, void* obs, void* mask, double* reward, uint8_t* done) {
    pass(e, MODE_STEP, actions, obs, mask, reward, done);
    return 0;
}
int hs_get_masks(hs_env* e, void* mask) {
    pass(e, MODE_OBSERVE, nullptr, nullptr, mask, nullptr, nullptr);
    return 0;
}
int hs_set_cycle_maps(hs_env* e, const int32_t* idx, int n) {
    if (n > 0 && e->S.P.num_selfplay > 0 && e->S.P.num_bot > 0) return -1;
    for (int k = 0; k < n; k++)
        if (idx[k] < 0 || idx[k] >= e->n_maps) return -1;
    e->cycle.assign(idx, idx + n);
    e->cycle_pos = 0;
    return 0;
}
int hs_planes(hs_env* e) { return e->S.P.planes; }
int hs_unit_capacity(hs_env* e) { return e->S.P.cap; }
int hs_import_state(hs_env* e, int game, const int32_t* hdr5, const int32_t* rows12, int n, const int32_t* issue_order) {
    if (game < 0 || game >= e->S.P.n_games || n > e->S.P.cap) return -1;
    uint8_t* blob = e->games.data() + (size_t)game * e->S.P.game_stride;
    memset(blob, 0, e->S.P.game_stride);
    return import_blob(blob, hdr5, rows12, n, issue_order, e->S.P.W, e->S.P.H, g_err) ? 0 : -1;
}
// Sim::can_execute against Sim::unit_mask on the stored state of `game`: every unit x every action the action space can
// name (NONE; MOVE / HARVEST / RETURN x 4 directions; PRODUCE x 4 directions x 7 types; ATTACK x 49 relative cells).
// Returns the number of (unit, action) pairs on which the two disagree; *checked = pairs compared.
int hs_can_execute_disagreements(hs_env* e, int game, int64_t* checked) {
    EnvParams P = e->S.P;
    P.game_ai = e->S.game_ai.data();
    P.game_ai1 = e->S.game_ai1.data();
    P.scap = P.cap;
    const Layout L = layout_of(P);
    std::vector<uint8_t> mem((size_t)L.total + 64, 0);
    Sim s;
    bind_sim(s, P, L, mem.data(), game);
    const uint8_t* blob = e->games.data() + (size_t)game * P.game_stride;
    memcpy(mem.data() + L.blob, blob, 32 + (size_t)P.cap * 32);
    const MapTpl tpl = e->S.tpls[e->S.game_map[game]];
    memcpy(s.grid, e->S.grids.data() + tpl.grid_off, (size_t)P.HW * 2);
    for (int i = 0; i < s.hdr->nunits; i++) s.grid[s.u[i].x + s.u[i].y * P.W] = (uint16_t)i;
    int bad = 0;
    int64_t n = 0;
    auto cmp = [&](int i, int type, int param, int x, int y, int utype, bool want) {
        PAct a;
        a.unit = (uint16_t)i, a.type = (int8_t)type, a.param = (int16_t)param, a.x = (int16_t)x, a.y = (int16_t)y, a.utype = (int8_t)utype, a.pad = 0;
        n++;
        if (s.can_execute(i, a) != want) bad++;
    };
    for (int i = 0; i < s.hdr->nunits; i++) {
        const Mask78 m = s.unit_mask(i);
        const Unit& un = s.u[i];
        cmp(i, T_NONE, 0, 0, 0, 0, mget(m, MB_TYPE + T_NONE));
        for (int d = 0; d < 4; d++) {
            cmp(i, T_MOVE, d, 0, 0, 0, mget(m, MB_MOVE + d));
            cmp(i, T_HARVEST, d, 0, 0, 0, mget(m, MB_HARVEST + d));
            cmp(i, T_RETURN, d, 0, 0, 0, mget(m, MB_RETURN + d));
            for (int k = 0; k < N_UTYPES; k++) cmp(i, T_PRODUCE, d, 0, 0, k, mget(m, MB_TYPE + T_PRODUCE) && mget(m, MB_PDIR + d) && mget(m, MB_PTYPE + k));
        }
        for (int dy = -3; dy <= 3; dy++)
            for (int dx = -3; dx <= 3; dx++) cmp(i, T_ATTACK, 0, un.x + dx, un.y + dy, 0, mget(m, MB_ATTACK + (3 + dy) * 7 + 3 + dx));
    }
    if (checked) *checked = n;
    return bad;
}
int hs_export_state(hs_env* e, int game, int32_t* hdr5, int32_t* rows12, int cap_rows) {
    const uint8_t* blob = e->games.data() + (size_t)game * e->S.P.game_stride;
    export_blob(blob, hdr5, rows12, cap_rows);
    return ((const GameHdr*)blob)->nunits;
}
}
