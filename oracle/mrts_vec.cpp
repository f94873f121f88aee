// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.  See mrts_oracle.hpp for the header note.
// JNIGridnetVecClient / JNIGridnetClient / JNIGridnetClientSelfPlay + the NumPy glue of
// MicroRTSGridModeVecEnv (gym_microrts/envs/vec_env.py:278-321,968-1057,1091-1101).
#include <algorithm>
#include <atomic>
#include <cstring>
#include <thread>

#include "mrts_oracle.hpp"

namespace orc {

void VecEnv::init() {
    N = num_selfplay + num_bot;
    H = maps[0].H;
    W = maps[0].W;
    planes = partial_obs ? 31 : 29;
    selfplay.resize(num_selfplay / 2);
    botgames.resize(num_bot);
    env_steps.assign(N, 0);
    source_mask.assign((size_t)N * H * W, 0);
    source_mask_valid = false;
    // JNIGridnetClient / SelfPlay constructors only load the map; gs exists after reset().
}

void VecEnv::reset_game(Game& g, const MapTemplate& m, int ai_id, int ai1_id) {
    // PhysicalGameState.load + new GameState (JNIGridnetClient.java:232-233); Unit.next_ID is a
    // JVM-wide static in the reference but never observable -- kept per game here.
    g.gs = make_game(m, none_mode, &g.next_id);
    if (ai_id >= 0) {
        g.bot = make_bot(ai_id);  // ai2 = ai2.clone(); ai2.reset()  (JNIGridnetClient.java:230-231)
        g.bot->reset();
    }
    if (ai1_id >= 0) {  // JNIBotClient.reset, JNIBotClient.java:146-150
        g.bot1 = make_bot(ai1_id);
        g.bot1->reset();
    }
}

// obs for one player view: 29 one-hot planes from `view` (gs or the stale POGS) and, under
// partial_obs, the 2 visibility planes -- the reference's Python declares them but its Java
// never produces them; defined README-literally: [29]=1 iff the cell is NOT
// within sight of any unit of the observer's opponent in the true post-step state, [30]=1 iff it is.
void VecEnv::observe(const GameState& view, const GameState& truth, int player, int32_t* out) const {
    std::vector<int32_t> raw;
    view.vector_observation(player, raw);
    encode_obs(raw, H, W, out, planes);
    if (partial_obs) {
        for (int y = 0; y < H; y++)
            for (int x = 0; x < W; x++) {
                bool vis = false;
                for (auto& u : truth.units)
                    if (u->player == 1 - player) {
                        int dx = u->x - x, dy = u->y - y, r = utype(u->type).sightRadius;
                        if (dx * dx + dy * dy <= r * r) {
                            vis = true;
                            break;
                        }
                    }
                int32_t* o = out + (size_t)(y * W + x) * planes;
                o[29] = vis ? 0 : 1;
                o[30] = vis ? 1 : 0;
            }
    }
}

void VecEnv::reset(int32_t* obs) {
    const size_t stride = (size_t)H * W * planes;
    for (size_t g = 0; g < selfplay.size(); g++) {
        reset_game(selfplay[g], maps[2 * g], -1);
        for (int i = 0; i < 2; i++) {
            if (partial_obs) observe(selfplay[g].gs.pogs(i), selfplay[g].gs, i, obs + (2 * g + i) * stride);
            else observe(selfplay[g].gs, selfplay[g].gs, i, obs + (2 * g + i) * stride);
        }
    }
    for (int b = 0; b < num_bot; b++) {
        reset_game(botgames[b], maps[num_selfplay + b], ai_ids[b], ai1_of(b));
        const GameState& gs = botgames[b].gs;
        if (partial_obs) observe(gs.pogs(0), gs, 0, obs + (num_selfplay + b) * stride);
        else observe(gs, gs, 0, obs + (num_selfplay + b) * stride);
    }
}

void VecEnv::swap_map_and_reset(int env, const MapTemplate& m, int32_t* obs) {
    const size_t stride = (size_t)H * W * planes;
    if (env < num_selfplay) {  // selfPlayClients[g].mapPath = ...; reset(); getResponse(0), getResponse(1)
        const int g = env / 2;
        maps[2 * g] = m;
        reset_game(selfplay[g], maps[2 * g], -1);
        for (int i = 0; i < 2; i++) {
            const GameState& gs = selfplay[g].gs;
            if (partial_obs) observe(gs.pogs(i), gs, i, obs + (2 * g + i) * stride);
            else observe(gs, gs, i, obs + (2 * g + i) * stride);
        }
    } else {  // clients[b].mapPath = ...; reset(0)  (JNIGridnetClient.java:228-247)
        const int b = env - num_selfplay;
        maps[env] = m;
        reset_game(botgames[b], maps[env], ai_ids[b], ai1_of(b));
        const GameState& gs = botgames[b].gs;
        if (partial_obs) observe(gs.pogs(0), gs, 0, obs + env * stride);
        else observe(gs, gs, 0, obs + env * stride);
    }
}

void VecEnv::get_masks(int32_t* mask78, int32_t* source_unit_mask) {
    const int HW = H * W;
    auto run = [&](int e0, int e1) {
        std::vector<int32_t> m79((size_t)HW * 79);
        for (int e = e0; e < e1; e++) {
            const GameState& gs = e < num_selfplay ? selfplay[e / 2].gs : botgames[e - num_selfplay].gs;
            int player = e < num_selfplay ? (e % 2) : 0;
            build_masks(gs, player, m79.data());  // always the TRUE state (JNIGridnetClient.java:203-216)
            for (int c = 0; c < HW; c++) {
                source_mask[(size_t)e * HW + c] = m79[(size_t)c * 79];
                if (mask78) std::memcpy(mask78 + ((size_t)e * HW + c) * 78, &m79[(size_t)c * 79 + 1], 78 * sizeof(int32_t));
            }
        }
    };
    if (threads <= 1 || N < 2 * threads) {
        run(0, N);
    } else {  // envs are independent: the CPU baseline spreads them over host threads
        std::atomic<int> next(0);
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; t++)
            pool.emplace_back([&]() {
                for (;;) {
                    int e0 = next.fetch_add(16);
                    if (e0 >= N) break;
                    run(e0, std::min(N, e0 + 16));
                }
            });
        for (auto& th : pool) th.join();
    }
    if (source_unit_mask) std::memcpy(source_unit_mask, source_mask.data(), source_mask.size() * sizeof(int32_t));
    source_mask_valid = true;
}

// vec_env.py:969-984: rows [cell, a0..a6] of the cells whose source_unit_mask (from the LAST
// get_action_mask call) is 1, ascending cell index.  Without a previous get_action_mask the
// reference raises AttributeError; here every cell is passed (Java filters by unit anyway).
std::vector<int32_t> VecEnv::gather_actions(int env, const int32_t* actions, const GameState&, int) const {
    const int HW = H * W;
    std::vector<int32_t> rows;
    for (int c = 0; c < HW; c++) {
        if (source_mask_valid && source_mask[(size_t)env * HW + c] != 1) continue;
        rows.push_back(c);
        const int32_t* a = actions + ((size_t)env * HW + c) * 7;
        for (int k = 0; k < 7; k++) rows.push_back(a[k]);
    }
    return rows;
}

static void snapshot_units(const GameState& gs, TraceSnapshot& te) {
    te.units.clear();
    for (auto& u : gs.units) te.units.push_back(*u);
}
static void add_player_action(TraceSnapshot& te, const PlayerAction& pa) {
    for (auto& p : pa) te.actions.push_back({*p.first, p.second});
}

void VecEnv::step_selfplay(int g, const int32_t* actions, int32_t* obs, double* rewards, uint8_t* done) {
    Game& game = selfplay[g];
    GameState& gs = game.gs;
    const size_t stride = (size_t)H * W * planes;
    // JNIGridnetClientSelfPlay.gameStep, :159-189
    TraceSnapshot te;
    snapshot_units(gs, te);  // BEFORE issue (:160)
    GameState views[2];
    for (int i = 0; i < 2; i++) {
        if (partial_obs) views[i] = gs.pogs(i);  // built after player 0's issue for i == 1
        const GameState& view = partial_obs ? views[i] : gs;
        std::vector<int32_t> rows = gather_actions(2 * g + i, actions, view, i);
        PlayerAction pa = from_vector_action(view, rows.data(), (int)rows.size() / 8, i);
        view.fill_with_nones(pa, i, 1);  // JNIAI.java:53
        gs.issue_safe(pa);
        add_player_action(te, pa);
    }
    gs.cycle();
    double rw[2][6];
    bool dn[2][6];
    for (int i = 0; i < 2; i++) {
        compute_rewards(i, te, gs, rw[i], dn[i]);
        observe(partial_obs ? views[i] : gs, gs, i, obs + (2 * g + i) * stride);
    }
    // JNIGridnetVecClient.gameStep, :241-265
    env_steps[2 * g] += 1;
    env_steps[2 * g + 1] += 1;
    if (dn[0][0] || env_steps[2 * g] >= max_steps) {
        reset_game(game, maps[2 * g], -1);
        for (int i = 0; i < 2; i++) {
            if (partial_obs) observe(gs.pogs(i), gs, i, obs + (2 * g + i) * stride);
            else observe(gs, gs, i, obs + (2 * g + i) * stride);
        }
        dn[0][0] = dn[1][0] = true;
        env_steps[2 * g] = env_steps[2 * g + 1] = 0;
    }
    for (int i = 0; i < 2; i++) {
        for (int j = 0; j < 6; j++) rewards[(2 * g + i) * 6 + j] = rw[i][j];
        done[2 * g + i] = dn[i][0] ? 1 : 0;
    }
}

void VecEnv::step_bot(int b, const int32_t* actions, int32_t* obs, double* rewards, uint8_t* done) {
    Game& game = botgames[b];
    GameState& gs = game.gs;
    const int env = num_selfplay + b;
    const size_t stride = (size_t)H * W * planes;
    env_steps[env] += 1;
    // JNIGridnetClient.gameStep, :163-196
    GameState v1, v2;
    if (partial_obs) {
        v1 = gs.pogs(0);
        v2 = gs.pogs(1);
    }
    const GameState& p1gs = partial_obs ? v1 : gs;
    const GameState& p2gs = partial_obs ? v2 : gs;
    PlayerAction pa1, pa2;
    if (game.bot1) {  // JNIBotClient.gameStep, JNIBotClient.java:108-111: both bots act on the full state
        pa1 = game.bot1->get_action(0, gs);
        pa2 = game.bot->get_action(1, gs);
    } else {
        std::vector<int32_t> rows = gather_actions(env, actions, p1gs, 0);
        pa1 = from_vector_action(p1gs, rows.data(), (int)rows.size() / 8, 0);
        p1gs.fill_with_nones(pa1, 0, 1);
        pa2 = game.bot->get_action(1, p2gs);
    }
    gs.issue_safe(pa1);
    gs.issue_safe(pa2);
    TraceSnapshot te;
    snapshot_units(gs, te);  // AFTER issue, before cycle (:175)
    add_player_action(te, pa1);
    add_player_action(te, pa2);
    gs.cycle();
    double rw[6];
    bool dn[6];
    compute_rewards(0, te, gs, rw, dn);
    observe(p1gs, gs, 0, obs + env * stride);
    // JNIGridnetVecClient.gameStep, :268-286
    if (dn[0] || env_steps[env] >= max_steps) {
        reset_game(game, maps[env], ai_ids[b], ai1_of(b));
        if (partial_obs) observe(gs.pogs(0), gs, 0, obs + env * stride);
        else observe(gs, gs, 0, obs + env * stride);
        dn[0] = true;
        env_steps[env] = 0;
    }
    for (int j = 0; j < 6; j++) rewards[env * 6 + j] = rw[j];
    done[env] = dn[0] ? 1 : 0;
}

void VecEnv::step(const int32_t* actions, int32_t* obs, double* rewards, uint8_t* done) {
    const int G = (int)selfplay.size();
    const int total = G + num_bot;
    auto run = [&](int k) {
        if (k < G) step_selfplay(k, actions, obs, rewards, done);
        else step_bot(k - G, actions, obs, rewards, done);
    };
    if (threads <= 1 || total < 2 * threads) {
        for (int k = 0; k < total; k++) run(k);
        return;
    }
    // games are independent (JNIGridnetVecClient.java:241-287 is a plain sequential loop);
    // the CPU baseline spreads them over host threads.
    std::atomic<int> next(0);
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; t++)
        pool.emplace_back([&]() {
            for (;;) {
                int k0 = next.fetch_add(16);
                if (k0 >= total) break;
                for (int k = k0; k < std::min(total, k0 + 16); k++) run(k);
            }
        });
    for (auto& th : pool) th.join();
}

}  // namespace orc
