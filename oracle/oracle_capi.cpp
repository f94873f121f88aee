// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.  See mrts_oracle.hpp for the header note.
// Plain C ABI over the oracle so tests / bench.py can drive it through ctypes.
#include <thread>
#include <cstring>

#include "mrts_oracle.hpp"

using namespace orc;

namespace {
MapTemplate make_template(int W, int H, const uint8_t* terrain, int res0, int res1, int nunits, const int32_t* units7) {
    MapTemplate m;
    m.W = W;
    m.H = H;
    m.terrain.assign(terrain, terrain + (size_t)W * H);
    m.res[0] = res0;
    m.res[1] = res1;
    for (int i = 0; i < nunits; i++) {
        const int32_t* r = units7 + 7 * i;  // id, type, player, x, y, resources, hitpoints
        Unit u;
        u.id = r[0];
        u.type = r[1];
        u.player = r[2];
        u.x = r[3];
        u.y = r[4];
        u.res = r[5];
        u.hp = r[6];
        m.units.push_back(u);
    }
    return m;
}

struct GameHandle {
    GameState gs;
    long long next_id = 0;
};
struct BotHandle {
    std::unique_ptr<BotAI> bot;
};

// rows of 7: ux, uy, type, param, ax, ay, utype
PlayerAction rows_to_pa(const GameState& gs, int n, const int32_t* rows) {
    PlayerAction pa;
    for (int i = 0; i < n; i++) {
        const int32_t* r = rows + 7 * i;
        UnitP u;
        for (auto& s : gs.units)
            if (s->x == r[0] && s->y == r[1]) {
                u = s;
                break;
            }
        if (!u) continue;
        Action a;
        a.type = r[2];
        a.param = r[3];
        a.x = r[4];
        a.y = r[5];
        a.utype = r[6];
        pa.push_back({u, a});
    }
    return pa;
}
int pa_to_rows(const PlayerAction& pa, int32_t* rows, int cap) {
    int n = 0;
    for (auto& p : pa) {
        if (n >= cap) break;
        int32_t* r = rows + 7 * n++;
        r[0] = p.first->x;
        r[1] = p.first->y;
        r[2] = p.second.type;
        r[3] = p.second.param;
        r[4] = p.second.x;
        r[5] = p.second.y;
        r[6] = p.second.utype;
    }
    return n;
}
}  // namespace

extern "C" {

// ---- vectorised env
// host threads the independent envs are spread over (the timed CPU baseline runs at 1 and at all cores)
void orc_vec_set_threads(void* h, int threads) { ((VecEnv*)h)->threads = threads < 1 ? 1 : threads; }
void* orc_vec_create(int num_selfplay, int num_bot, int max_steps, int partial_obs, int none_mode, int threads) {
    VecEnv* v = new VecEnv();
    v->num_selfplay = num_selfplay;
    v->num_bot = num_bot;
    v->max_steps = max_steps;
    v->partial_obs = partial_obs != 0;
    v->none_mode = none_mode;
    v->threads = threads < 1 ? 1 : threads;
    return v;
}
// one call per env, in env order; ai_id < 0 for self-play views
void orc_vec_add_env(void* h, int W, int H, const uint8_t* terrain, int res0, int res1, int nunits, const int32_t* units7, int ai_id) {
    VecEnv* v = (VecEnv*)h;
    v->maps.push_back(make_template(W, H, terrain, res0, res1, nunits, units7));
    if (ai_id >= 0) v->ai_ids.push_back(ai_id);
}
// JNIBotClient games: scripted player 0 of bot env b (call before orc_vec_init)
void orc_vec_set_ai1(void* h, int b, int ai1_id) {
    VecEnv* v = (VecEnv*)h;
    if ((int)v->ai1_ids.size() <= b) v->ai1_ids.resize(b + 1, -1);
    v->ai1_ids[b] = ai1_id;
}
void orc_vec_init(void* h) { ((VecEnv*)h)->init(); }
void orc_vec_reset(void* h, int32_t* obs) { ((VecEnv*)h)->reset(obs); }
void orc_vec_masks(void* h, int32_t* mask78, int32_t* source) { ((VecEnv*)h)->get_masks(mask78, source); }
// Policy stand-in of the timed CPU baseline (bench.py): one masked-uniform action per cell that may act, drawn from the
// mask the caller just got from orc_vec_masks -- the recipe of the reference's hello_world.py:27-67, spread over the same
// host threads as the env.  Not part of the restated path; it only feeds it valid actions quickly (NumPy takes ~10x the
// env's own time at 8,192 envs).  actions (N, HW, 7) int32; cells that may not act get zeros.
void orc_vec_sample_actions(void* h, const int32_t* mask78, uint64_t seed, int32_t* actions) {
    VecEnv* v = (VecEnv*)h;
    const int HW = v->H * v->W, N = v->N;
    static const int lo[8] = {0, 6, 10, 14, 18, 22, 29, 78};
    auto mix = [](uint64_t z) {
        z += 0x9E3779B97F4A7C15ULL;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
        return z ^ (z >> 31);
    };
    auto work = [&](int e0, int e1) {
        for (int e = e0; e < e1; e++)
            for (int c = 0; c < HW; c++) {
                const int32_t* m = mask78 + ((size_t)e * HW + c) * 78;
                int32_t* a = actions + ((size_t)e * HW + c) * 7;
                if (!m[0]) {
                    for (int k = 0; k < 7; k++) a[k] = 0;
                    continue;
                }
                uint64_t r = mix(seed ^ ((uint64_t)e << 20) ^ (uint64_t)c);
                for (int k = 0; k < 7; k++) {
                    int n = 0;
                    for (int b = lo[k]; b < lo[k + 1]; b++) n += m[b] != 0;
                    int pick = 0;
                    if (n > 0) {
                        r = mix(r);
                        int want = (int)(r % (uint64_t)n);
                        for (int b = lo[k]; b < lo[k + 1]; b++)
                            if (m[b] && want-- == 0) {
                                pick = b - lo[k];
                                break;
                            }
                    }
                    a[k] = pick;
                }
            }
    };
    const int T = v->threads;
    if (T <= 1 || N < 2 * T) {
        work(0, N);
        return;
    }
    std::vector<std::thread> th;
    for (int t = 0; t < T; t++) th.emplace_back(work, (int)((long long)N * t / T), (int)((long long)N * (t + 1) / T));
    for (auto& x : th) x.join();
}
void orc_vec_step(void* h, const int32_t* actions, int32_t* obs, double* rewards, uint8_t* done) {
    ((VecEnv*)h)->step(actions, obs, rewards, done);
}
int orc_vec_planes(void* h) { return ((VecEnv*)h)->planes; }
void orc_vec_swap_map(void* h, int env, int W, int H, const uint8_t* terrain, int res0, int res1, int nunits, const int32_t* units7, int32_t* obs) {
    ((VecEnv*)h)->swap_map_and_reset(env, make_template(W, H, terrain, res0, res1, nunits, units7), obs);
}
// state of one env's game: returns number of units; rows of 12:
// type, player, x, y, res, hp, a_type, a_param, a_x, a_y, a_utype, a_time (a_type = -1 when idle)
int orc_vec_game_state(void* h, int env, int32_t* hdr3 /*time,res0,res1*/, int32_t* rows, int cap);
void orc_vec_destroy(void* h) { delete (VecEnv*)h; }

// ---- single game (trace replay / bot regeneration)
void* orc_game_create(int W, int H, const uint8_t* terrain, int res0, int res1, int nunits, const int32_t* units7, int none_mode) {
    GameHandle* g = new GameHandle();
    MapTemplate m = make_template(W, H, terrain, res0, res1, nunits, units7);
    g->gs = make_game(m, none_mode, &g->next_id);
    return g;
}
void orc_game_destroy(void* h) { delete (GameHandle*)h; }
int orc_game_issue_safe(void* h, int n, const int32_t* rows7) {
    GameHandle* g = (GameHandle*)h;
    PlayerAction pa = rows_to_pa(g->gs, n, rows7);
    return g->gs.issue_safe(pa) ? 1 : 0;
}
int orc_game_cycle(void* h) { return ((GameHandle*)h)->gs.cycle() ? 1 : 0; }
int orc_game_time(void* h) { return ((GameHandle*)h)->gs.time; }
int orc_game_winner(void* h) { return ((GameHandle*)h)->gs.winner(); }

static int dump_state(const GameState& gs, int32_t* hdr3, int32_t* rows, int cap) {
    hdr3[0] = gs.time;
    hdr3[1] = gs.players->res[0];
    hdr3[2] = gs.players->res[1];
    int n = 0;
    for (auto& u : gs.units) {
        if (n >= cap) break;
        int32_t* r = rows + 12 * n++;
        r[0] = u->type;
        r[1] = u->player;
        r[2] = u->x;
        r[3] = u->y;
        r[4] = u->res;
        r[5] = u->hp;
        AssignP a = gs.assignment(u.get());
        if (a) {
            r[6] = a->action.type;
            r[7] = a->action.param;
            r[8] = a->action.x;
            r[9] = a->action.y;
            r[10] = a->action.utype;
            r[11] = a->time;
        } else {
            r[6] = -1;
            r[7] = r[8] = r[9] = r[10] = r[11] = 0;
        }
    }
    return (int)gs.units.size();
}
// position of every unit's assignment in the unitActions map (issue order), -1 = idle; one entry per unit in pgs.units order
int orc_game_issue_order(void* h, int32_t* out, int cap) {
    const GameState& gs = ((GameHandle*)h)->gs;
    int n = 0;
    for (auto& u : gs.units) {
        if (n >= cap) break;
        int pos = -1;
        for (size_t k = 0; k < gs.acts.size(); k++)
            if (gs.acts[k]->unit.get() == u.get()) pos = (int)k;
        out[n++] = pos;
    }
    return n;
}
int orc_game_state(void* h, int32_t* hdr3, int32_t* rows, int cap) { return dump_state(((GameHandle*)h)->gs, hdr3, rows, cap); }
int orc_vec_game_state(void* h, int env, int32_t* hdr3, int32_t* rows, int cap) {
    VecEnv* v = (VecEnv*)h;
    const GameState& gs = env < v->num_selfplay ? v->selfplay[env / 2].gs : v->botgames[env - v->num_selfplay].gs;
    return dump_state(gs, hdr3, rows, cap);
}
// masks / raw obs of a single game (known-answer tests)
void orc_game_masks(void* h, int player, int32_t* mask79) { build_masks(((GameHandle*)h)->gs, player, mask79); }

void* orc_bot_create(int ai_id) {
    BotHandle* b = new BotHandle();
    b->bot = make_bot(ai_id);
    if (!b->bot) {
        delete b;
        return nullptr;
    }
    b->bot->reset();
    return b;
}
void orc_bot_destroy(void* b) { delete (BotHandle*)b; }
// the bot's PlayerAction for `player` as rows of 7 (not issued)
int orc_bot_action(void* gh, void* bh, int player, int32_t* rows7, int cap) {
    GameHandle* g = (GameHandle*)gh;
    BotHandle* b = (BotHandle*)bh;
    PlayerAction pa = b->bot->get_action(player, g->gs);
    return pa_to_rows(pa, rows7, cap);
}

const char* orc_version() { return "mrts-oracle-1"; }
}
