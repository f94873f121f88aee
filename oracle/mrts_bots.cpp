// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.  See mrts_oracle.hpp for the header note.
// Scripted opponents: PassiveAI, AbstractionLayerAI (+Train/Build/Harvest/Attack), A*,
// WorkerRush and LightRush (the latter only as the vehicle that pins the shared machinery
// against the 140 LightRush golden traces).
#include <algorithm>
#include <cstdlib>

#include "mrts_oracle.hpp"

namespace orc {

// ---------------------------------------------------------------- A*
// J/ai/abstraction/pathfinding/AStarPathFinding.java:52-79,104-138,175-295
struct AStar {
    std::vector<int8_t> freec;  // -1 unknown (Boolean null), 0 false, 1 true; indexed pos
    std::vector<int> closed, open, heur, parents, cost, inoc;
    int openinsert = 0;

    void add_to_open(int newPos, int oldPos, int h) {
        cost[newPos] = cost[oldPos] + 1;
        for (int i = openinsert - 1; i >= 0; i--) {
            if (heur[i] + cost[open[i]] >= h + cost[newPos]) {
                for (int j = openinsert; j > i + 1; j--) {
                    open[j] = open[j - 1];
                    heur[j] = heur[j - 1];
                    parents[j] = parents[j - 1];
                }
                open[i + 1] = newPos;
                heur[i + 1] = h;
                parents[i + 1] = oldPos;
                openinsert++;
                inoc[newPos] = 1;
                return;
            }
        }
        for (int j = openinsert; j > 0; j--) {
            open[j] = open[j - 1];
            heur[j] = heur[j - 1];
            parents[j] = parents[j - 1];
        }
        open[0] = newPos;
        heur[0] = h;
        parents[0] = oldPos;
        openinsert++;
        inoc[newPos] = 1;
    }

    bool run(const Unit& start, int targetpos, int range, const GameState& gs, const Usage* ru) {
        const int w = gs.W, h = gs.H, n = w * h;
        freec.assign(n, -1);
        closed.assign(n, -1);
        inoc.assign(n, 0);
        open.resize(n + 1);
        heur.resize(n + 1);
        parents.resize(n + 1);
        cost.resize(n);
        if (ru)
            for (int pos : ru->pos) freec[pos] = 0;
        const int tx = targetpos % w, ty = targetpos / w;
        const int sq = range * range;
        const int startPos = start.y * w + start.x;
        openinsert = 0;
        open[0] = startPos;
        heur[0] = std::abs(start.x - tx) + std::abs(start.y - ty);
        parents[0] = startPos;
        inoc[startPos] = 1;
        cost[startPos] = 0;
        openinsert++;
        while (openinsert > 0) {
            openinsert--;
            int pos = open[openinsert];
            int parent = parents[openinsert];
            if (closed[pos] != -1) continue;
            closed[pos] = parent;
            int x = pos % w, y = pos / w;
            if ((x - tx) * (x - tx) + (y - ty) * (y - ty) <= sq) return true;
            auto probe = [&](int nx, int ny, int npos) {
                if (freec[npos] == -1) freec[npos] = gs.free_cell(nx, ny) ? 1 : 0;
                if (freec[npos] == 1) add_to_open(npos, pos, std::abs(nx - tx) + std::abs(ny - ty));
            };
            if (y > 0 && inoc[pos - w] == 0) probe(x, y - 1, pos - w);
            if (x < w - 1 && inoc[pos + 1] == 0) probe(x + 1, y, pos + 1);
            if (y < h - 1 && inoc[pos + w] == 0) probe(x, y + 1, pos + w);
            if (x > 0 && inoc[pos - 1] == 0) probe(x - 1, y, pos - 1);
        }
        return false;
    }

    // findPathToPositionInRange: returns true and the first move, or false (null)
    bool find_path(const Unit& start, int targetpos, int range, const GameState& gs, const Usage* ru, Action& out) {
        if (!run(start, targetpos, range, gs, ru)) return false;
        const int w = gs.W;
        int pos = open[openinsert];
        int parent = parents[openinsert];
        int last = pos;
        while (parent != pos) {
            last = pos;
            pos = parent;
            parent = closed[pos];
        }
        if (last == pos + w) { out = act_dir(T_MOVE, D_DOWN); return true; }
        if (last == pos - 1) { out = act_dir(T_MOVE, D_LEFT); return true; }
        if (last == pos - w) { out = act_dir(T_MOVE, D_UP); return true; }
        if (last == pos + 1) { out = act_dir(T_MOVE, D_RIGHT); return true; }
        return false;
    }
};

// ---------------------------------------------------------------- abstract actions
enum { AA_TRAIN, AA_BUILD, AA_HARVEST, AA_ATTACK, AA_MOVE };
struct AbstractAction {
    int kind;
    UnitP unit;
    int utype = -1;          // Train / Build
    bool completed_flag = false;  // Train
    int x = 0, y = 0;        // Build / Move
    UnitP target, base;      // Harvest (target may be null) / Attack (target)
};
using AAP = std::shared_ptr<AbstractAction>;

struct AbstractionLayerAI : BotAI {
    std::vector<AAP> actions;  // LinkedHashMap<Unit, AbstractAction>: insertion order, re-put keeps slot
    bool explore = false;      // the PO* subclasses (ai/abstraction/partialobservability/): their meleeUnitBehavior override
    AStar pf;

    void reset() override { actions.clear(); }  // AbstractionLayerAI.java:53-55

    AAP get_abstract(const Unit* u) const {
        for (auto& a : actions)
            if (a->unit.get() == u) return a;
        return nullptr;
    }
    void put(const AAP& a) {
        for (auto& e : actions)
            if (e->unit.get() == a->unit.get()) {
                e = a;
                return;
            }
        actions.push_back(a);
    }
    void train(const UnitP& u, int ut) {
        auto a = std::make_shared<AbstractAction>();
        a->kind = AA_TRAIN;
        a->unit = u;
        a->utype = ut;
        put(a);
    }
    void build(const UnitP& u, int ut, int x, int y) {
        auto a = std::make_shared<AbstractAction>();
        a->kind = AA_BUILD;
        a->unit = u;
        a->utype = ut;
        a->x = x;
        a->y = y;
        put(a);
    }
    void harvest(const UnitP& u, const UnitP& target, const UnitP& base) {
        auto a = std::make_shared<AbstractAction>();
        a->kind = AA_HARVEST;
        a->unit = u;
        a->target = target;
        a->base = base;
        put(a);
    }
    void move(const UnitP& u, int x, int y) {  // AbstractionLayerAI.java:119-121
        auto a = std::make_shared<AbstractAction>();
        a->kind = AA_MOVE;
        a->unit = u;
        a->x = x;
        a->y = y;
        put(a);
    }
    void attack(const UnitP& u, const UnitP& target) {
        auto a = std::make_shared<AbstractAction>();
        a->kind = AA_ATTACK;
        a->unit = u;
        a->target = target;
        put(a);
    }

    static bool completed(const AbstractAction& aa, const GameState& gs) {
        switch (aa.kind) {
            case AA_TRAIN: return aa.completed_flag;                      // Train.java:30-32
            case AA_BUILD: return gs.unit_at(aa.x, aa.y) != nullptr;       // Build.java:33-37
            case AA_HARVEST:                                               // Harvest.java:42-48
                if (aa.unit->res > 0) return !gs.contains(aa.base.get());
                return !gs.contains(aa.target.get());  // contains(null) is false
            case AA_ATTACK: return !gs.contains(aa.target.get());         // Attack.java:30-33
            case AA_MOVE: return aa.unit->x == aa.x && aa.unit->y == aa.y; // Move.java:32-34
        }
        return true;
    }

    // Train.score, Train.java:98-127
    static int train_score(int x, int y, int ut, int player, const GameState& gs) {
        int distance = 0;
        bool first = true;
        if (utype(ut).canHarvest) {
            for (auto& u : gs.units)
                if (utype(u->type).isResource) {
                    int d = std::abs(u->x - x) + std::abs(u->y - y);
                    if (first || d < distance) {
                        distance = d;
                        first = false;
                    }
                }
        } else {
            for (auto& u : gs.units)
                if (u->player >= 0 && u->player != player) {
                    int d = std::abs(u->x - x) + std::abs(u->y - y);
                    if (first || d < distance) {
                        distance = d;
                        first = false;
                    }
                }
        }
        return -distance;
    }

    // returns true and sets `out` when the abstract action yields a primitive action (non-null)
    bool execute(AbstractAction& aa, const GameState& gs, const Usage& ru, Action& out) {
        const Unit& unit = *aa.unit;
        switch (aa.kind) {
            case AA_TRAIN: {  // Train.java:48-96
                int x = unit.x, y = unit.y;
                int best_dir = -1, best_score = -1;
                auto consider = [&](int nx, int ny, int dir) {
                    int score = train_score(nx, ny, aa.utype, unit.player, gs);
                    if (score > best_score || best_dir == -1) {
                        best_score = score;
                        best_dir = dir;
                    }
                };
                if (y > 0 && gs.free_cell(x, y - 1)) consider(x, y - 1, D_UP);
                if (x < gs.W - 1 && gs.free_cell(x + 1, y)) consider(x + 1, y, D_RIGHT);
                if (y < gs.H - 1 && gs.free_cell(x, y + 1)) consider(x, y + 1, D_DOWN);
                if (x > 0 && gs.free_cell(x - 1, y)) consider(x - 1, y, D_LEFT);
                aa.completed_flag = true;
                if (best_dir != -1) {
                    Action ua = act_produce(best_dir, aa.utype);
                    if (gs.action_allowed(unit, ua)) {
                        out = ua;
                        return true;
                    }
                }
                return false;
            }
            case AA_BUILD: {  // Build.java:56-78
                Action mv;
                if (pf.find_path(unit, aa.x + aa.y * gs.W, 1, gs, &ru, mv)) {
                    if (gs.action_allowed(unit, mv)) {
                        out = mv;
                        return true;
                    }
                    return false;
                }
                bool have = false;
                Action ua;
                if (aa.x == unit.x && aa.y == unit.y - 1) { ua = act_produce(D_UP, aa.utype); have = true; }
                if (aa.x == unit.x + 1 && aa.y == unit.y) { ua = act_produce(D_RIGHT, aa.utype); have = true; }
                if (aa.x == unit.x && aa.y == unit.y + 1) { ua = act_produce(D_DOWN, aa.utype); have = true; }
                if (aa.x == unit.x - 1 && aa.y == unit.y) { ua = act_produce(D_LEFT, aa.utype); have = true; }
                if (have && gs.action_allowed(unit, ua)) {
                    out = ua;
                    return true;
                }
                return false;
            }
            case AA_HARVEST: {  // Harvest.java:71-114
                const bool carrying = unit.res != 0;
                const Unit* tgt = carrying ? aa.base.get() : aa.target.get();
                if (!tgt) return false;
                Action mv;
                if (pf.find_path(unit, tgt->x + tgt->y * gs.W, 1, gs, &ru, mv)) {
                    if (gs.action_allowed(unit, mv)) {
                        out = mv;
                        return true;
                    }
                    return false;
                }
                int ty = carrying ? T_RETURN : T_HARVEST;
                if (tgt->x == unit.x && tgt->y == unit.y - 1) { out = act_dir(ty, D_UP); return true; }
                if (tgt->x == unit.x + 1 && tgt->y == unit.y) { out = act_dir(ty, D_RIGHT); return true; }
                if (tgt->x == unit.x && tgt->y == unit.y + 1) { out = act_dir(ty, D_DOWN); return true; }
                if (tgt->x == unit.x - 1 && tgt->y == unit.y) { out = act_dir(ty, D_LEFT); return true; }
                return false;
            }
            case AA_MOVE: {  // Move.java:53-59: findPath = findPathToPositionInRange(..., 0, ...)
                Action mv;
                if (pf.find_path(unit, aa.x + aa.y * gs.W, 0, gs, &ru, mv) && gs.action_allowed(unit, mv)) {
                    out = mv;
                    return true;
                }
                return false;
            }
            case AA_ATTACK: {  // Attack.java:50-65
                const Unit& t = *aa.target;
                int dx = t.x - unit.x, dy = t.y - unit.y;
                int r = utype(unit.type).attackRange;
                if (dx * dx + dy * dy <= r * r) {
                    out = act_attack(t.x, t.y);
                    return true;
                }
                Action mv;
                if (pf.find_path(unit, t.x + t.y * gs.W, r, gs, &ru, mv) && gs.action_allowed(unit, mv)) {
                    out = mv;
                    return true;
                }
                return false;
            }
        }
        return false;
    }

    // AbstractionLayerAI.java:58-113
    PlayerAction translate_actions(int player, const GameState& gs) {
        PlayerAction pa;
        std::vector<std::pair<UnitP, Action>> desires;
        std::vector<const Unit*> to_delete;
        Usage ru;
        for (auto& aa : actions) {
            if (!gs.contains(aa->unit.get())) {
                to_delete.push_back(aa->unit.get());
            } else if (completed(*aa, gs)) {
                to_delete.push_back(aa->unit.get());
            } else if (!gs.assignment(aa->unit.get())) {
                Action ua;
                if (execute(*aa, gs, ru, ua)) {
                    desires.push_back({aa->unit, ua});
                    ru.merge(gs.usage(*aa->unit, ua));
                }
            }
        }
        for (auto* u : to_delete)
            for (size_t i = 0; i < actions.size(); i++)
                if (actions[i]->unit.get() == u) {
                    actions.erase(actions.begin() + i);
                    break;
                }
        Usage r = gs.resource_usage();
        for (auto& d : desires) {
            Usage r2 = gs.usage(*d.first, d.second);
            if (gs.consistent(r, r2)) {
                pa.push_back(d);
                r.merge(r2);
            }
        }
        gs.fill_with_nones(pa, player, 10);
        return pa;
    }

    // AbstractionLayerAI.java:143-228 (pgs.getAllFree(): no wall and no unit)
    int find_building_position(const std::vector<int>& reserved, int dX, int dY, const GameState& gs) {
        auto ok = [&](int x, int y) {
            int pos = x + y * gs.W;
            if (std::find(reserved.begin(), reserved.end(), pos) != reserved.end()) return false;
            return gs.terrain_at(x, y) == 0 && gs.unit_at(x, y) == nullptr;
        };
        int mx = std::max(gs.H, gs.W);
        for (int l = 1; l < mx; l++) {
            for (int side = 0; side < 4; side++) {
                if (side == 0) {
                    int y = dY - l;
                    if (y < 0) continue;
                    for (int dx = -l; dx <= l; dx++) {
                        int x = dX + dx;
                        if (x < 0 || x >= gs.W) continue;
                        if (ok(x, y)) return x + y * gs.W;
                    }
                } else if (side == 1) {
                    int x = dX + l;
                    if (x >= gs.W) continue;
                    for (int dy = -l; dy <= l; dy++) {
                        int y = dY + dy;
                        if (y < 0 || y >= gs.H) continue;
                        if (ok(x, y)) return x + y * gs.W;
                    }
                } else if (side == 2) {
                    int y = dY + l;
                    if (y >= gs.H) continue;
                    for (int dx = -l; dx <= l; dx++) {
                        int x = dX + dx;
                        if (x < 0 || x >= gs.W) continue;
                        if (ok(x, y)) return x + y * gs.W;
                    }
                } else {
                    int x = dX - l;
                    if (x < 0) continue;
                    for (int dy = -l; dy <= l; dy++) {
                        int y = dY + dy;
                        if (y < 0 || y >= gs.H) continue;
                        if (ok(x, y)) return x + y * gs.W;
                    }
                }
            }
        }
        return -1;
    }

    // AbstractionLayerAI.java:230-245
    void build_if_not_already_building(const UnitP& u, int ut, int dX, int dY, std::vector<int>& reserved, const GameState& gs) {
        AAP a = get_abstract(u.get());
        if (!a || a->kind != AA_BUILD || a->utype != ut) {
            int pos = find_building_position(reserved, dX, dY, gs);
            build(u, ut, pos % gs.W, pos / gs.W);  // pos == -1 -> (-1, 0): Java % and / truncate like C++
            reserved.push_back(pos);
        }
    }

    // WorkerRush.meleeUnitBehavior / LightRush.meleeUnitBehavior
    void melee_behavior(const UnitP& u, int player, const GameState& gs) {
        UnitP closest;
        int cd = 0;
        for (auto& u2 : gs.units)
            if (u2->player >= 0 && u2->player != player) {
                int d = std::abs(u2->x - u->x) + std::abs(u2->y - u->y);
                if (!closest || d < cd) {
                    closest = u2;
                    cd = d;
                }
            }
        if (closest) {
            attack(u, closest);
        } else if (explore && gs.pogs_player >= 0) {
            // POWorkerRush / POLightRush.meleeUnitBehavior (partialobservability/POWorkerRush.java:56-76): no enemy
            // in view -> walk to the nearest cell that is not observable (row-major scan, strict <)
            int cx = 0, cy = 0, best = -1;
            for (int i = 0; i < gs.H; i++)
                for (int j = 0; j < gs.W; j++)
                    if (!gs.observable(j, i)) {
                        int d = (u->x - j) * (u->x - j) + (u->y - i) * (u->y - i);
                        if (best == -1 || d < best) {
                            cx = j;
                            cy = i;
                            best = d;
                        }
                    }
            if (best != -1) move(u, cx, cy);
        }
    }

    // shared tail of WorkerRush.workersBehavior:148-197 / LightRush.workersBehavior:201-244.
    // returns true when the worker stays free
    bool harvest_behavior(const UnitP& w, int player, const GameState& gs) {
        UnitP cbase, cres;
        int cd = 0;
        for (auto& u2 : gs.units)
            if (utype(u2->type).isResource) {
                int d = std::abs(u2->x - w->x) + std::abs(u2->y - w->y);
                if (!cres || d < cd) {
                    cres = u2;
                    cd = d;
                }
            }
        cd = 0;
        for (auto& u2 : gs.units)
            if (utype(u2->type).isStockpile && u2->player == player) {
                int d = std::abs(u2->x - w->x) + std::abs(u2->y - w->y);
                if (!cbase || d < cd) {
                    cbase = u2;
                    cd = d;
                }
            }
        bool still_free = true;
        if (w->res > 0) {
            if (cbase) {
                AAP aa = get_abstract(w.get());
                if (aa && aa->kind == AA_HARVEST) {
                    if (aa->base.get() != cbase.get()) harvest(w, nullptr, cbase);
                } else {
                    harvest(w, nullptr, cbase);
                }
                still_free = false;
            }
        } else {
            if (cres && cbase) {
                AAP aa = get_abstract(w.get());
                if (aa && aa->kind == AA_HARVEST) {
                    if (aa->target.get() != cres.get() || aa->base.get() != cbase.get()) harvest(w, cres, cbase);
                } else {
                    harvest(w, cres, cbase);
                }
                still_free = false;
            }
        }
        return still_free;
    }
};

// ---------------------------------------------------------------- PassiveAI (J/ai/PassiveAI.java:41-45)
struct PassiveAI : BotAI {
    void reset() override {}
    PlayerAction get_action(int player, const GameState& gs) override {
        PlayerAction pa;
        gs.fill_with_nones(pa, player, 10);
        return pa;
    }
};

// ---------------------------------------------------------------- WorkerRush (J/ai/abstraction/WorkerRush.java:63-204)
struct WorkerRush : AbstractionLayerAI {
    PlayerAction get_action(int player, const GameState& gs) override {
        const int pres = gs.players->res[player];
        for (auto& u : gs.units)
            if (u->type == UT_BASE && u->player == player && !gs.assignment(u.get()))
                if (pres >= utype(UT_WORKER).cost) train(u, UT_WORKER);
        for (auto& u : gs.units)
            if (utype(u->type).canAttack && !utype(u->type).canHarvest && u->player == player && !gs.assignment(u.get()))
                melee_behavior(u, player, gs);
        std::vector<UnitP> workers;
        for (auto& u : gs.units)
            if (utype(u->type).canHarvest && u->player == player) workers.push_back(u);
        workers_behavior(workers, player, gs);
        return translate_actions(player, gs);
    }
    void workers_behavior(const std::vector<UnitP>& workers, int player, const GameState& gs) {
        int nbases = 0, resourcesUsed = 0;
        std::vector<UnitP> free_w(workers);
        if (workers.empty()) return;
        for (auto& u2 : gs.units)
            if (u2->type == UT_BASE && u2->player == player) nbases++;
        std::vector<int> reserved;
        if (nbases == 0 && !free_w.empty()) {
            if (gs.players->res[player] >= utype(UT_BASE).cost + resourcesUsed) {
                UnitP u = free_w.front();
                free_w.erase(free_w.begin());
                build_if_not_already_building(u, UT_BASE, u->x, u->y, reserved, gs);
                resourcesUsed += utype(UT_BASE).cost;
            }
        }
        UnitP hw;
        if (!free_w.empty()) {
            hw = free_w.front();
            free_w.erase(free_w.begin());
        }
        if (hw) {
            if (harvest_behavior(hw, player, gs)) free_w.push_back(hw);
        }
        for (auto& u : free_w) melee_behavior(u, player, gs);
    }
};

// ---------------------------------------------------------------- LightRush (J/ai/abstraction/LightRush.java:77-254)
// HeavyRush.java / RangedRush.java are the same class with heavyType / rangedType in barracksBehavior
// (HeavyRush.java:136-137, RangedRush.java:124-125; RangedRush.java:181-183 only reorders the && of the
// barracks-building test): rush_type names the unit the barracks train.
struct LightRush : AbstractionLayerAI {
    int rush_type = UT_LIGHT;
    PlayerAction get_action(int player, const GameState& gs) override {
        const int pres = gs.players->res[player];
        for (auto& u : gs.units)
            if (u->type == UT_BASE && u->player == player && !gs.assignment(u.get())) {
                int nworkers = 0;
                for (auto& u2 : gs.units)
                    if (u2->type == UT_WORKER && u2->player == player) nworkers++;
                if (nworkers < 1 && pres >= utype(UT_WORKER).cost) train(u, UT_WORKER);
            }
        for (auto& u : gs.units)
            if (u->type == UT_BARRACKS && u->player == player && !gs.assignment(u.get()))
                if (pres >= utype(rush_type).cost) train(u, rush_type);
        for (auto& u : gs.units)
            if (utype(u->type).canAttack && !utype(u->type).canHarvest && u->player == player && !gs.assignment(u.get()))
                melee_behavior(u, player, gs);
        std::vector<UnitP> workers;
        for (auto& u : gs.units)
            if (utype(u->type).canHarvest && u->player == player) workers.push_back(u);
        workers_behavior(workers, player, gs);
        return translate_actions(player, gs);
    }
    void workers_behavior(const std::vector<UnitP>& workers, int player, const GameState& gs) {
        int nbases = 0, nbarracks = 0, resourcesUsed = 0;
        std::vector<UnitP> free_w(workers);
        if (workers.empty()) return;
        for (auto& u2 : gs.units) {
            if (u2->type == UT_BASE && u2->player == player) nbases++;
            if (u2->type == UT_BARRACKS && u2->player == player) nbarracks++;
        }
        std::vector<int> reserved;
        const int pres = gs.players->res[player];
        if (nbases == 0 && !free_w.empty()) {
            if (pres >= utype(UT_BASE).cost + resourcesUsed) {
                UnitP u = free_w.front();
                free_w.erase(free_w.begin());
                build_if_not_already_building(u, UT_BASE, u->x, u->y, reserved, gs);
                resourcesUsed += utype(UT_BASE).cost;
            }
        }
        if (nbarracks == 0) {
            if (pres >= utype(UT_BARRACKS).cost + resourcesUsed && !free_w.empty()) {
                UnitP u = free_w.front();
                free_w.erase(free_w.begin());
                build_if_not_already_building(u, UT_BARRACKS, u->x, u->y, reserved, gs);
                resourcesUsed += utype(UT_BARRACKS).cost;
            }
        }
        std::vector<UnitP> still_free;
        for (auto& u : free_w)
            if (harvest_behavior(u, player, gs)) still_free.push_back(u);
        for (auto& u : still_free) melee_behavior(u, player, gs);
    }
};

std::unique_ptr<BotAI> make_bot(int ai_id) {
    switch (ai_id) {
        case AI_PASSIVE: return std::unique_ptr<BotAI>(new PassiveAI());
        case AI_WORKER_RUSH: return std::unique_ptr<BotAI>(new WorkerRush());
        case AI_LIGHT_RUSH: return std::unique_ptr<BotAI>(new LightRush());
        case AI_PO_WORKER_RUSH: {
            auto* b = new WorkerRush();
            b->explore = true;
            return std::unique_ptr<BotAI>(b);
        }
        case AI_PO_LIGHT_RUSH:
        case AI_PO_HEAVY_RUSH:
        case AI_PO_RANGED_RUSH: {  // partialobservability/PO{Light,Heavy,Ranged}Rush.java:23-78
            auto* b = new LightRush();
            b->explore = true;
            b->rush_type = ai_id == AI_PO_HEAVY_RUSH ? UT_HEAVY : (ai_id == AI_PO_RANGED_RUSH ? UT_RANGED : UT_LIGHT);
            return std::unique_ptr<BotAI>(b);
        }
    }
    return nullptr;
}

}  // namespace orc
