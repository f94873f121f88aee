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
    P.mode = mode;
    P.actions = actions;
    P.obs = obs;
    P.mask = mask;
    P.reward = reward;
    P.done = done;
    for (int g = 0; g < P.n_games; g++) {
        if (P.elem == 4 && P.planes == 29) game_pass<4, 29>(P, g, e->mem.data(), 0, 1, NoSync());
        else if (P.elem == 4) game_pass<4, 31>(P, g, e->mem.data(), 0, 1, NoSync());
        else if (P.planes == 29) game_pass<1, 29>(P, g, e->mem.data(), 0, 1, NoSync());
        else game_pass<1, 31>(P, g, e->mem.data(), 0, 1, NoSync());
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
    e->games.assign((size_t)P.n_games * P.game_stride + 64, 0);
    e->mem.assign(layout_of(P).total + 64, 0);
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
int hs_planes(hs_env* e) { return e->S.P.planes; }
int hs_unit_capacity(hs_env* e) { return e->S.P.cap; }
int hs_export_state(hs_env* e, int game, int32_t* hdr5, int32_t* rows12, int cap_rows) {
    const uint8_t* blob = e->games.data() + (size_t)game * e->S.P.game_stride;
    export_blob(blob, hdr5, rows12, cap_rows);
    return ((const GameHdr*)blob)->nunits;
}
}
