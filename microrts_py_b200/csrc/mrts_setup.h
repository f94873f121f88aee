// Host-side preparation of a vec env: validates the constructor arguments
// (JNIGridnetVecClient, J/tests/JNIGridnetVecClient.java:106-142) and flattens the map templates
// (PhysicalGameState.load, J/rts/PhysicalGameState.java:685-711) into the pools the kernels read.
#pragma once
#include <algorithm>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/mrts_b200.h"
#include "mrts_step.h"

namespace mrts {

struct HostSetup {
    EnvParams P;  // scalar fields filled; pointer fields left null
    int N = 0;
    std::vector<MapTpl> tpls;
    std::vector<uint16_t> grids;
    std::vector<Unit> units;
    std::vector<int32_t> game_map;
    std::vector<int8_t> game_ai;
    std::vector<int8_t> game_ai1;
};

inline int build_setup(const mrts_config* cfg, HostSetup& S, std::string& err) {
    if (!cfg) {
        err = "null config";
        return MRTS_EINVAL;
    }
    if (cfg->num_selfplay_envs < 0 || cfg->num_bot_envs < 0 || (cfg->num_selfplay_envs & 1)) {
        err = "num_selfplay_envs must be even and counts non-negative";
        return MRTS_EINVAL;
    }
    const int N = cfg->num_selfplay_envs + cfg->num_bot_envs;
    if (N <= 0) {
        err = "no environments";
        return MRTS_EINVAL;
    }
    if (cfg->n_maps <= 0 || !cfg->maps || !cfg->env_map) {
        err = "no maps";
        return MRTS_EINVAL;
    }
    if (cfg->num_bot_envs > 0 && !cfg->bot_ai) {
        err = "bot_ai missing";
        return MRTS_EINVAL;
    }
    if (cfg->elem_bytes != 4 && cfg->elem_bytes != 1) {
        err = "elem_bytes must be 4 or 1";
        return MRTS_EINVAL;
    }
    if (cfg->max_steps < 1) {
        err = "max_steps must be positive";
        return MRTS_EINVAL;
    }
    const int W = cfg->maps[0].width, H = cfg->maps[0].height;
    if (W < 1 || H < 1 || W > 128 || H > 128) {
        err = "map size must be within 1..128";
        return MRTS_EINVAL;
    }
    const int HW = W * H;
    int cap = 0;
    for (int m = 0; m < cfg->n_maps; m++) {
        const mrts_map_template& t = cfg->maps[m];
        if (t.width != W || t.height != H) {
            err = "all maps of one vec env must have the same size";
            return MRTS_EINVAL;
        }
        if (!t.terrain || (t.n_units > 0 && !t.units7)) {
            err = "map template has null arrays";
            return MRTS_EINVAL;
        }
        // units ever alive <= initial units + every resource on the map (each unit costs >= 1), and <= free cells
        long total = (long)t.n_units + t.resources[0] + t.resources[1];
        int walls = 0;
        for (int c = 0; c < HW; c++) walls += t.terrain[c] ? 1 : 0;
        for (int k = 0; k < t.n_units; k++) {
            const int32_t* r = t.units7 + 7 * k;
            if (r[1] < 0 || r[1] >= N_UTYPES) {
                err = "unknown unit type in map";
                return MRTS_EINVAL;
            }
            if (r[3] < 0 || r[3] >= W || r[4] < 0 || r[4] >= H) {
                err = "unit off the map";
                return MRTS_EINVAL;
            }
            if (r[2] < -1 || r[2] > 1) {
                err = "only players -1, 0, 1 are supported";
                return MRTS_EINVAL;
            }
            total += r[5];
        }
        long lim = HW - walls;
        if (lim < t.n_units) lim = t.n_units;
        if (total > lim) total = lim;
        if (total > cap) cap = (int)total;
    }
    if (cap < 1) cap = 1;
    if (cap >= G_WALL) {
        err = "unit capacity too large";
        return MRTS_EINVAL;
    }
    S.N = N;
    EnvParams& P = S.P;
    memset(&P, 0, sizeof(P));
    P.num_selfplay = cfg->num_selfplay_envs;
    P.num_bot = cfg->num_bot_envs;
    P.n_games = cfg->num_selfplay_envs / 2 + cfg->num_bot_envs;
    P.W = W;
    P.H = H;
    P.HW = HW;
    P.cap = cap;
    P.scap = cap;  // the CUDA library may stage fewer (mrts_create)
    P.ovf_cnt = nullptr;
    P.ovf_list = nullptr;
    P.partial = cfg->partial_obs ? 1 : 0;
    P.planes = P.partial ? 31 : 29;
    P.none_mode = cfg->none_mode ? NONE_FORK : NONE_UPSTREAM;
    P.max_steps = cfg->max_steps;
    P.elem = cfg->elem_bytes;
    P.has_bot = 0;
    P.any_bot = cfg->num_bot_envs > 0 ? 1 : 0;
    for (int b = 0; b < cfg->num_bot_envs; b++) {
        if (cfg->bot_ai[b] < 0 || cfg->bot_ai[b] >= AI_COUNT) {
            err = "unsupported scripted opponent id";
            return MRTS_EINVAL;
        }
        if (ai_plans(cfg->bot_ai[b])) P.has_bot = 1;
        if (cfg->bot_ai1 && ai_plans(cfg->bot_ai1[b])) P.has_bot = 1;
    }
    P.game_stride = blob_bytes(cap, P.has_bot);

    const int grid_pad = align16(HW * 2) / 2;
    S.tpls.assign(cfg->n_maps, MapTpl());
    S.grids.assign((size_t)cfg->n_maps * grid_pad, (uint16_t)G_NONE);
    S.units.clear();
    for (int m = 0; m < cfg->n_maps; m++) {
        const mrts_map_template& t = cfg->maps[m];
        MapTpl& d = S.tpls[m];
        d.W = W;
        d.H = H;
        d.res[0] = t.resources[0];
        d.res[1] = t.resources[1];
        d.nunits = t.n_units;
        d.grid_off = m * grid_pad;
        d.unit_off = (int)S.units.size();
        d.pad = 0;
        for (int c = 0; c < HW; c++) S.grids[(size_t)m * grid_pad + c] = t.terrain[c] ? (uint16_t)G_WALL : (uint16_t)G_NONE;
        for (int k = 0; k < t.n_units; k++) {
            const int32_t* r = t.units7 + 7 * k;
            Unit u;
            memset(&u, 0, sizeof(u));
            u.type = (uint8_t)r[1];
            u.player = (int8_t)r[2];
            u.x = (uint8_t)r[3];
            u.y = (uint8_t)r[4];
            u.res = r[5];
            u.hp = (int16_t)r[6];
            u.a_type = -1;
            u.a_param = -1;
            u.a_utype = -1;
            u.uid = (uint16_t)k;
            u.vtype[0] = u.vtype[1] = -1;
            S.units.push_back(u);
        }
    }
    S.game_map.assign(P.n_games, 0);
    S.game_ai.assign(P.n_games, (int8_t)AI_NONE);
    S.game_ai1.assign(P.n_games, (int8_t)AI_NONE);
    const int nsp = P.num_selfplay / 2;
    for (int g = 0; g < P.n_games; g++) {
        const int env = g < nsp ? 2 * g : P.num_selfplay + (g - nsp);
        const int m = cfg->env_map[env];
        if (m < 0 || m >= cfg->n_maps) {
            err = "env_map entry out of range";
            return MRTS_EINVAL;
        }
        S.game_map[g] = m;
        S.game_ai[g] = g < nsp ? (int8_t)AI_NONE : (int8_t)cfg->bot_ai[g - nsp];
        if (g >= nsp && cfg->bot_ai1 && cfg->bot_ai1[g - nsp] >= 0) {
            const int a1 = cfg->bot_ai1[g - nsp];
            if (a1 >= AI_COUNT) {
                err = "unsupported scripted player-0 id";
                return MRTS_EINVAL;
            }
            if (cfg->partial_obs) {
                err = "bot-vs-bot envs ignore partial_obs in the reference (JNIBotClient acts on the full state): create them with partial_obs = 0";
                return MRTS_EINVAL;
            }
            S.game_ai1[g] = (int8_t)a1;
        }
    }
    return MRTS_OK;
}

inline void export_blob(const uint8_t* blob, int32_t* hdr5, int32_t* rows12, int cap_rows) {
    const GameHdr* h = (const GameHdr*)blob;
    const Unit* u = (const Unit*)(blob + 32);
    if (hdr5) {
        hdr5[0] = h->time;
        hdr5[1] = h->res[0];
        hdr5[2] = h->res[1];
        hdr5[3] = h->steps;
        hdr5[4] = h->nunits;
    }
    for (int i = 0; i < h->nunits && i < cap_rows && rows12; i++) {
        int32_t* r = rows12 + 12 * i;
        r[0] = u[i].type;
        r[1] = u[i].player;
        r[2] = u[i].x;
        r[3] = u[i].y;
        r[4] = u[i].res;
        r[5] = u[i].hp;
        if (u[i].a_type >= 0) {
            r[6] = u[i].a_type;
            r[7] = u[i].a_param;
            r[8] = u[i].a_type == T_ATTACK ? u[i].a_x : 0;
            r[9] = u[i].a_type == T_ATTACK ? u[i].a_y : 0;
            r[10] = u[i].a_utype;
            r[11] = u[i].a_time;
        } else {
            r[6] = -1;
            r[7] = r[8] = r[9] = r[10] = r[11] = 0;
        }
    }
}

// Inverse of export_blob (GameState.fromXML, J/rts/GameState.java:763-915): header + unit rows -> blob.  Derived
// fields are rebuilt: uids follow the row order (= pgs.units order), the issue order of the pending assignments
// (the unitActions LinkedHashMap) is `issue_order[i]` ascending when given, else (time of issue, row).  A scripted
// bot's abstract-action table is not part of a GameState: it starts empty, as after a reset.
inline bool import_blob(uint8_t* blob, const int32_t* hdr5, const int32_t* rows12, int n, const int32_t* issue_order, int W, int H, std::string& err) {
    GameHdr* h = (GameHdr*)blob;
    Unit* u = (Unit*)(blob + 32);
    memset(h, 0, sizeof(*h));
    h->time = hdr5[0];
    h->res[0] = hdr5[1];
    h->res[1] = hdr5[2];
    h->steps = hdr5[3];
    h->nunits = n;
    h->next_uid = (uint16_t)n;
    std::vector<int> issued;
    for (int i = 0; i < n; i++) {
        const int32_t* r = rows12 + 12 * i;
        if (r[0] < 0 || r[0] >= N_UTYPES || r[1] < -1 || r[1] > 1 || r[2] < 0 || r[2] >= W || r[3] < 0 || r[3] >= H) {
            err = "unit row out of range";
            return false;
        }
        Unit& d = u[i];
        memset(&d, 0, sizeof(d));
        d.type = (uint8_t)r[0];
        d.player = (int8_t)r[1];
        d.x = (uint8_t)r[2];
        d.y = (uint8_t)r[3];
        d.res = r[4];
        d.hp = (int16_t)r[5];
        d.uid = (uint16_t)i;
        d.vtype[0] = d.vtype[1] = -1;
        d.a_type = -1;
        d.a_param = -1;
        d.a_utype = -1;
        if (r[6] >= 0) {
            if (r[6] > T_ATTACK) {
                err = "unknown action type";
                return false;
            }
            d.a_type = (int8_t)r[6];
            d.a_param = (int16_t)r[7];
            d.a_x = (uint8_t)r[8];
            d.a_y = (uint8_t)r[9];
            d.a_utype = (int8_t)r[10];
            d.a_time = r[11];
            issued.push_back(i);
        }
    }
    std::stable_sort(issued.begin(), issued.end(), [&](int a, int b) {
        if (issue_order) return issue_order[a] < issue_order[b];
        return u[a].a_time < u[b].a_time;
    });
    for (size_t k = 0; k < issued.size(); k++) u[issued[k]].a_seq = (uint32_t)k;
    h->next_seq = (uint32_t)issued.size();
    return true;
}

}  // namespace mrts
