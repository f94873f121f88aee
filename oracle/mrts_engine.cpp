// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.  See mrts_oracle.hpp for the header note.
// Engine restatement: UnitTypeTable, UnitAction, ResourceUsage, GameState, PhysicalGameState,
// PartiallyObservableGameState, PlayerAction.fromVectorAction, masks, rewards, obs encode.
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "mrts_oracle.hpp"

namespace orc {

const int DIR_DX[4] = {0, 1, 0, -1};   // J/rts/UnitAction.java:96
const int DIR_DY[4] = {-1, 0, 1, 0};   // J/rts/UnitAction.java:101

// J/rts/units/UnitTypeTable.java:105-290 with VERSION_ORIGINAL; unset fields keep the
// defaults of J/rts/units/UnitType.java:41-108.
static const UnitType kTypes[N_UTYPES] = {
    //id name       cost hp mnD mxD rng prodT movT atkT hrvT retT hAmt sight isRes  isStock canHarv canMove canAtk nprod produces
    {0, "Resource", 1, 1, 1, 1, 1, 10, 10, 10, 10, 10, 1, 0, true, false, false, false, false, 0, {0, 0, 0}},
    {1, "Base", 10, 10, 1, 1, 1, 250, 10, 10, 10, 10, 1, 5, false, true, false, false, false, 1, {UT_WORKER, 0, 0}},
    {2, "Barracks", 5, 4, 1, 1, 1, 200, 10, 10, 10, 10, 1, 3, false, false, false, false, false, 3, {UT_LIGHT, UT_HEAVY, UT_RANGED}},
    {3, "Worker", 1, 1, 1, 1, 1, 50, 10, 5, 20, 10, 1, 3, false, false, true, true, true, 2, {UT_BASE, UT_BARRACKS, 0}},
    {4, "Light", 2, 4, 2, 2, 1, 80, 8, 5, 10, 10, 1, 2, false, false, false, true, true, 0, {0, 0, 0}},
    {5, "Heavy", 2, 4, 4, 4, 1, 120, 12, 5, 10, 10, 1, 2, false, false, false, true, true, 0, {0, 0, 0}},
    {6, "Ranged", 2, 1, 1, 1, 3, 100, 10, 5, 10, 10, 1, 3, false, false, false, true, true, 0, {0, 0, 0}},
};

const UnitType& utype(int id) { return kTypes[id]; }

int utype_by_name(const std::string& name) {
    for (int i = 0; i < N_UTYPES; i++)
        if (name == kTypes[i].name) return i;
    return -1;
}

// ---------------------------------------------------------------- UnitAction
Action act_none_default() { return Action(); }

Action act_none(int duration, int none_mode) {
    Action a;
    a.type = T_NONE;
    a.param = (none_mode == NONE_FORK) ? -1 : duration;  // UnitAction.java:150-151 (fork quirk)
    return a;
}

Action act_dir(int type, int dir) {
    Action a;
    a.type = type;
    a.param = dir;
    return a;
}

Action act_produce(int dir, int ut) {
    Action a;
    a.type = T_PRODUCE;
    a.param = dir;
    a.utype = ut;
    return a;
}

Action act_attack(int x, int y) {
    Action a;
    a.type = T_ATTACK;
    a.x = x;
    a.y = y;
    return a;
}

bool act_equals(const Action& a, const Action& b) {
    if (a.type != b.type) return false;
    if (a.type == T_NONE || a.type == T_MOVE || a.type == T_HARVEST || a.type == T_RETURN) return a.param == b.param;
    if (a.type == T_ATTACK) return a.x == b.x && a.y == b.y;
    return a.param == b.param && a.utype == b.utype;
}

int act_eta(const Unit& u, const Action& a) {
    const UnitType& t = utype(u.type);
    switch (a.type) {
        case T_NONE: return a.param;
        case T_MOVE: return t.moveTime;
        case T_ATTACK: return t.attackTime;
        case T_HARVEST: return t.harvestTime;
        case T_RETURN: return t.moveTime;  // sic, UnitAction.java:327-328
        case T_PRODUCE: return utype(a.utype).produceTime;
    }
    return 0;
}

void Usage::merge(const Usage& o) {
    pos.insert(pos.end(), o.pos.begin(), o.pos.end());
    res[0] += o.res[0];
    res[1] += o.res[1];
}

// ---------------------------------------------------------------- GameState basics
Unit* GameState::unit_at(int x, int y) const {
    for (auto& u : units)
        if (u->x == x && u->y == y) return u.get();
    return nullptr;
}

bool GameState::contains(const Unit* u) const {
    for (auto& v : units)
        if (v.get() == u) return true;
    return false;
}

AssignP GameState::assignment(const Unit* u) const {
    for (auto& a : acts)
        if (a->unit.get() == u) return a;
    return nullptr;
}

void GameState::put_assignment(const AssignP& a) {
    for (auto& e : acts)
        if (e->unit.get() == a->unit.get()) {
            e = a;
            return;
        }
    acts.push_back(a);
}

void GameState::remove_assignment(const Unit* u) {
    for (size_t i = 0; i < acts.size(); i++)
        if (acts[i]->unit.get() == u) {
            acts.erase(acts.begin() + i);
            return;
        }
}

void GameState::remove_unit(const Unit* u) {
    for (size_t i = 0; i < units.size(); i++)
        if (units[i].get() == u) {
            units.erase(units.begin() + i);
            break;
        }
    remove_assignment(u);
}

Usage GameState::usage(const Unit& u, const Action& a) const {
    Usage r;
    if (a.type == T_MOVE || a.type == T_PRODUCE) {
        if (a.type == T_PRODUCE) r.res[u.player] += utype(a.utype).cost;
        int pos = u.x + u.y * W;
        switch (a.param) {
            case D_UP: pos -= W; break;
            case D_RIGHT: pos++; break;
            case D_DOWN: pos += W; break;
            case D_LEFT: pos--; break;
        }
        r.pos.push_back(pos);
    }
    return r;
}

bool GameState::consistent(const Usage& base, const Usage& other) const {
    for (int p : other.pos)
        if (std::find(base.pos.begin(), base.pos.end(), p) != base.pos.end()) return false;
    for (int i = 0; i < 2; i++) {
        if (other.res[i] == 0) continue;
        if (base.res[i] + other.res[i] > 0 && base.res[i] + other.res[i] > players->res[i]) return false;
    }
    return true;
}

std::vector<Action> GameState::unit_actions(const Unit& u, int noneDuration) const {
    std::vector<Action> l;
    const UnitType& t = utype(u.type);
    const int x = u.x, y = u.y;
    // neighbours: the LAST unit in list order wins (Unit.java:389-406)
    const Unit *uup = nullptr, *uright = nullptr, *udown = nullptr, *uleft = nullptr;
    for (auto& o : units) {
        if (o->x == x) {
            if (o->y == y - 1) uup = o.get();
            else if (o->y == y + 1) udown = o.get();
        } else if (o->y == y) {
            if (o->x == x - 1) uleft = o.get();
            else if (o->x == x + 1) uright = o.get();
        }
    }
    if (t.canAttack) {
        if (t.attackRange == 1) {
            if (y > 0 && uup && uup->player != u.player && uup->player >= 0) l.push_back(act_attack(uup->x, uup->y));
            if (x < W - 1 && uright && uright->player != u.player && uright->player >= 0) l.push_back(act_attack(uright->x, uright->y));
            if (y < H - 1 && udown && udown->player != u.player && udown->player >= 0) l.push_back(act_attack(udown->x, udown->y));
            if (x > 0 && uleft && uleft->player != u.player && uleft->player >= 0) l.push_back(act_attack(uleft->x, uleft->y));
        } else {
            int sq = t.attackRange * t.attackRange;
            for (auto& o : units) {
                if (o->player < 0 || o->player == u.player) continue;
                int dx = o->x - x, dy = o->y - y;
                if (dx * dx + dy * dy <= sq) l.push_back(act_attack(o->x, o->y));
            }
        }
    }
    if (t.canHarvest) {
        if (u.res == 0) {
            if (y > 0 && uup && utype(uup->type).isResource) l.push_back(act_dir(T_HARVEST, D_UP));
            if (x < W - 1 && uright && utype(uright->type).isResource) l.push_back(act_dir(T_HARVEST, D_RIGHT));
            if (y < H - 1 && udown && utype(udown->type).isResource) l.push_back(act_dir(T_HARVEST, D_DOWN));
            if (x > 0 && uleft && utype(uleft->type).isResource) l.push_back(act_dir(T_HARVEST, D_LEFT));
        }
        if (u.res > 0) {
            if (y > 0 && uup && utype(uup->type).isStockpile && uup->player == u.player) l.push_back(act_dir(T_RETURN, D_UP));
            if (x < W - 1 && uright && utype(uright->type).isStockpile && uright->player == u.player) l.push_back(act_dir(T_RETURN, D_RIGHT));
            if (y < H - 1 && udown && utype(udown->type).isStockpile && udown->player == u.player) l.push_back(act_dir(T_RETURN, D_DOWN));
            if (x > 0 && uleft && utype(uleft->type).isStockpile && uleft->player == u.player) l.push_back(act_dir(T_RETURN, D_LEFT));
        }
    }
    const int tup = (y > 0 ? terrain_at(x, y - 1) : 1);
    const int tright = (x < W - 1 ? terrain_at(x + 1, y) : 1);
    const int tdown = (y < H - 1 ? terrain_at(x, y + 1) : 1);
    const int tleft = (x > 0 ? terrain_at(x - 1, y) : 1);
    for (int k = 0; k < t.nproduces; k++) {
        int ut = t.produces[k];
        if (players->res[u.player] >= utype(ut).cost) {
            if (tup == 0 && unit_at(x, y - 1) == nullptr) l.push_back(act_produce(D_UP, ut));
            if (tright == 0 && unit_at(x + 1, y) == nullptr) l.push_back(act_produce(D_RIGHT, ut));
            if (tdown == 0 && unit_at(x, y + 1) == nullptr) l.push_back(act_produce(D_DOWN, ut));
            if (tleft == 0 && unit_at(x - 1, y) == nullptr) l.push_back(act_produce(D_LEFT, ut));
        }
    }
    if (t.canMove) {
        if (tup == 0 && !uup) l.push_back(act_dir(T_MOVE, D_UP));
        if (tright == 0 && !uright) l.push_back(act_dir(T_MOVE, D_RIGHT));
        if (tdown == 0 && !udown) l.push_back(act_dir(T_MOVE, D_DOWN));
        if (tleft == 0 && !uleft) l.push_back(act_dir(T_MOVE, D_LEFT));
    }
    l.push_back(act_none(noneDuration, none_mode));
    return l;
}

bool GameState::can_execute(const Unit& u, const Action& a) const {
    std::vector<Action> l = unit_actions(u, act_eta(u, a));
    for (auto& b : l)
        if (act_equals(b, a)) return true;  // List.contains -> UnitAction.equals
    return false;
}

bool GameState::integrity_check() const {
    for (size_t i = 0; i < acts.size(); i++) {
        if (!contains(acts[i]->unit.get())) return false;
        for (size_t j = 0; j < i; j++)
            if (acts[j]->unit.get() == acts[i]->unit.get()) return false;
    }
    return true;
}

bool GameState::issue(PlayerAction& pa) {
    bool ret = false;
    for (size_t k = 0; k < pa.size(); k++) {
        UnitP unit = pa[k].first;
        // `p` is a local that CANCEL_BOTH rebinds (GameState.java:296): the PlayerAction keeps the
        // original action in that case, but GameState.java:316 mutates the Pair itself.
        Action cur = pa[k].second;
        const Usage ru = usage(*unit, cur);  // computed once, before the scan (GameState.java:264)
        for (auto& uaa : acts) {
            if (!consistent(usage(*uaa->unit, uaa->action), ru)) {
                if (uaa->time == time) {
                    int d1 = act_eta(*uaa->unit, uaa->action);
                    int d2 = act_eta(*unit, cur);
                    int m = std::min(d1, d2);
                    uaa->action = act_none(m, none_mode);  // cancel old (slot kept)
                    cur = act_none(m, none_mode);          // cancel new (local only)
                } else {
                    cur = act_none_default();  // p.m_b = new UnitAction(TYPE_NONE)
                    pa[k].second = cur;        // mutates the Pair
                }
            }
        }
        auto na = std::make_shared<Assign>();
        na->unit = unit;
        na->action = cur;
        na->time = time;
        put_assignment(na);
        if (cur.type != T_NONE) ret = true;
    }
    return ret;
}

bool GameState::issue_safe(PlayerAction& pa) {
    // PlayerAction.integrityCheck (PlayerAction.java:244-259): single owner
    for (size_t k = 1; k < pa.size(); k++)
        if (pa[k].first->player != pa[0].first->player) abort();
    if (!integrity_check()) abort();
    for (auto& p : pa) {
        if (!can_execute(*p.first, p.second)) {
            int l = act_eta(*p.first, p.second);
            p.second = act_none(l, none_mode);  // GameState.java:352-353
        }
        // (the substitute-unit search of GameState.java:357-382 is a no-op: units are never cloned here)
        Usage r = usage(*p.first, p.second);
        for (int position : r.pos) {
            // Java / and % truncate toward zero like C++
            int y = position / W;
            int x = position % W;
            if (terrain_at(x, y) != 0 || unit_at(x, y) != nullptr) {
                p.second = act_none(act_eta(*p.first, p.second), none_mode);  // GameState.java:392-397
            }
        }
    }
    bool ret = issue(pa);
    if (!integrity_check()) abort();
    return ret;
}

void GameState::execute(const UnitP& up, const Action& a) {
    Unit& u = *up;
    switch (a.type) {
        case T_NONE: break;
        case T_MOVE:
            switch (a.param) {
                case D_UP: u.y--; break;
                case D_RIGHT: u.x++; break;
                case D_DOWN: u.y++; break;
                case D_LEFT: u.x--; break;
            }
            break;
        case T_ATTACK: {
            Unit* other = unit_at(a.x, a.y);
            if (other) {
                int damage = utype(u.type).minDamage;  // min == max in VERSION_ORIGINAL
                other->hp -= damage;
                if (other->hp <= 0) remove_unit(other);
            }
            break;
        }
        case T_HARVEST: {
            Unit* r = nullptr;
            if (a.param >= 0 && a.param < 4) r = unit_at(u.x + DIR_DX[a.param], u.y + DIR_DY[a.param]);
            if (r && utype(r->type).isResource && utype(u.type).canHarvest && u.res == 0) {
                r->res -= utype(u.type).harvestAmount;
                if (r->res <= 0) remove_unit(r);
                u.res = utype(u.type).harvestAmount;
            }
            break;
        }
        case T_RETURN: {
            Unit* b = nullptr;
            if (a.param >= 0 && a.param < 4) b = unit_at(u.x + DIR_DX[a.param], u.y + DIR_DY[a.param]);
            if (b && utype(b->type).isStockpile && u.res > 0) {  // owner is NOT checked (UnitAction.java:430)
                players->res[u.player] += u.res;
                u.res = 0;
            }
            break;
        }
        case T_PRODUCE: {
            int tx = u.x, ty = u.y;
            if (a.param >= 0 && a.param < 4) {
                tx += DIR_DX[a.param];
                ty += DIR_DY[a.param];
            }
            const UnitType& nt = utype(a.utype);
            if (players->res[u.player] - nt.cost >= 0) {
                // pgs.addUnit throws if the cell is occupied (PhysicalGameState.java:174-182);
                // reservations make that unreachable -- assert it.
                if (unit_at(tx, ty) != nullptr) abort();
                auto nu = std::make_shared<Unit>();
                nu->id = next_id ? (*next_id)++ : 0;
                nu->type = a.utype;
                nu->player = u.player;
                nu->x = tx;
                nu->y = ty;
                nu->res = 0;
                nu->hp = nt.hp;
                units.push_back(nu);
                players->res[u.player] -= nt.cost;
            }
            break;
        }
    }
}

bool GameState::cycle() {
    time++;
    std::vector<AssignP> ready;
    for (auto& uaa : acts)
        if (act_eta(*uaa->unit, uaa->action) + uaa->time <= time) ready.push_back(uaa);
    for (auto& uaa : ready) {
        remove_assignment(uaa->unit.get());
        execute(uaa->unit, uaa->action);  // a unit removed earlier this cycle still executes
    }
    return gameover();
}

bool GameState::gameover() const {
    int counts[2] = {0, 0}, total = 0;
    for (auto& u : units)
        if (u->player >= 0) {
            counts[u->player]++;
            total++;
        }
    if (total == 0) return true;
    int w = -1;
    for (int i = 0; i < 2; i++)
        if (counts[i] > 0) {
            if (w == -1) w = i;
            else return false;
        }
    return w != -1;
}

int GameState::winner() const {
    int counts[2] = {0, 0};
    for (auto& u : units)
        if (u->player >= 0) counts[u->player]++;
    int w = -1;
    for (int i = 0; i < 2; i++)
        if (counts[i] > 0) {
            if (w == -1) w = i;
            else return -1;
        }
    return w;
}

bool GameState::free_cell(int x, int y) const {
    if (terrain_at(x, y) != 0) return false;
    for (auto& u : units)
        if (u->x == x && u->y == y) return false;
    for (auto& ua : acts) {
        if (ua->action.type == T_MOVE || ua->action.type == T_PRODUCE) {
            const Unit& u = *ua->unit;
            int d = ua->action.param;
            if (d == D_UP && u.x == x && u.y == y + 1) return false;
            if (d == D_RIGHT && u.x == x - 1 && u.y == y) return false;
            if (d == D_DOWN && u.x == x && u.y == y - 1) return false;
            if (d == D_LEFT && u.x == x + 1 && u.y == y) return false;
        }
    }
    return true;
}

Usage GameState::resource_usage() const {
    Usage base;
    for (auto& u : units) {
        AssignP uaa = assignment(u.get());
        if (uaa) base.merge(usage(*u, uaa->action));
    }
    return base;
}

bool GameState::action_allowed(const Unit& u, const Action& a) const {
    if (a.type == T_MOVE) {
        int x2 = u.x + DIR_DX[a.param], y2 = u.y + DIR_DY[a.param];
        if (x2 < 0 || y2 < 0 || x2 >= W || y2 >= H || terrain_at(x2, y2) == 1 || unit_at(x2, y2) != nullptr) return false;
    }
    Usage empty = resource_usage();
    return consistent(usage(u, a), empty);  // ua.resourceUsage(u).consistentWith(empty.r)
}

void GameState::fill_with_nones(PlayerAction& pa, int player, int duration) const {
    for (auto& u : units) {
        if (u->player != player) continue;
        if (assignment(u.get())) continue;
        bool found = false;
        for (auto& p : pa)
            if (p.first.get() == u.get()) {
                found = true;
                break;
            }
        if (!found) pa.push_back({u, act_none(duration, none_mode)});
    }
}

void GameState::vector_observation(int player, std::vector<int32_t>& out) const {
    const int HW = H * W;
    out.assign(6 * HW, 0);
    for (auto& u : units) {
        int c = u->y * W + u->x;
        AssignP uaa = assignment(u.get());
        out[0 * HW + c] = u->hp;
        out[1 * HW + c] = u->res;
        if (u->player >= 0) out[2 * HW + c] = ((u->player + player) % 2) + 1;
        out[3 * HW + c] = u->type + 1;
        if (uaa) out[4 * HW + c] = uaa->action.type;
    }
    for (int c = 0; c < HW; c++) out[5 * HW + c] = (*terrain)[c];
}

GameState GameState::pogs(int player) const {
    GameState v = *this;  // shares Unit / Assign / Players objects; own lists (cloneKeepingUnits + putAll)
    std::vector<const Unit*> del;
    for (auto& u : v.units) {
        if (u->player != player) {
            bool obs = false;
            for (auto& w : v.units) {
                if (w->player == player) {
                    int dx = w->x - u->x, dy = w->y - u->y;
                    int r = utype(w->type).sightRadius;
                    if (dx * dx + dy * dy <= r * r) {  // sqrt(d2) <= r, exact on integers
                        obs = true;
                        break;
                    }
                }
            }
            if (!obs) del.push_back(u.get());
        }
    }
    for (auto* u : del) v.remove_unit(u);
    v.pogs_player = player;
    return v;
}

bool GameState::observable(int x, int y) const {
    for (auto& u : units)
        if (u->player == pogs_player) {
            int dx = u->x - x, dy = u->y - y, r = utype(u->type).sightRadius;
            if (dx * dx + dy * dy <= r * r) return true;  // sqrt(d2) <= r, exact on integers
        }
    return false;
}

GameState make_game(const MapTemplate& m, int none_mode, long long* next_id) {
    GameState gs;
    gs.W = m.W;
    gs.H = m.H;
    gs.terrain = std::make_shared<const std::vector<uint8_t>>(m.terrain);
    gs.players = std::make_shared<Players>();
    gs.players->res[0] = m.res[0];
    gs.players->res[1] = m.res[1];
    gs.none_mode = none_mode;
    gs.next_id = next_id;
    for (auto& u : m.units) {
        gs.units.push_back(std::make_shared<Unit>(u));
        if (next_id && u.id >= *next_id) *next_id = u.id + 1;  // Unit.java:52-54
    }
    return gs;
}

// ---------------------------------------------------------------- fromVectorAction
PlayerAction from_vector_action(const GameState& view, const int32_t* actions, int n, int player) {
    PlayerAction pa;
    Usage base = view.resource_usage();  // PlayerAction.java:387-394
    const int W = view.W;
    for (int i = 0; i < n; i++) {
        const int32_t* a = actions + 8 * i;
        Unit* up = view.unit_at(a[0] % W, a[0] / W);
        if (!up || up->player != player || view.assignment(up)) continue;
        UnitP u;
        for (auto& s : view.units)
            if (s.get() == up) u = s;
        Action ua;  // UnitAction.fromVectorAction, UnitAction.java:681-715
        ua.type = a[1];
        switch (ua.type) {
            case T_NONE: break;
            case T_MOVE: ua.param = a[2]; break;
            case T_HARVEST: ua.param = a[3]; break;
            case T_RETURN: ua.param = a[4]; break;
            case T_PRODUCE:
                ua.param = a[5];
                ua.utype = a[6];
                break;
            case T_ATTACK:
                ua.x = u->x + (a[7] % 7 - 3);
                ua.y = u->y + (a[7] / 7 - 3);
                break;
        }
        Usage ru = view.usage(*u, ua);
        if (view.consistent(base, ru)) {
            base.merge(ru);
            pa.push_back({u, ua});
        }
    }
    return pa;
}

// ---------------------------------------------------------------- masks
void build_masks(const GameState& gs, int player, int32_t* mask79) {
    const int HW = gs.H * gs.W;
    std::memset(mask79, 0, sizeof(int32_t) * HW * 79);
    for (auto& up : gs.units) {
        const Unit& u = *up;
        if (u.player != player || gs.assignment(&u)) continue;
        int32_t* m = mask79 + (u.y * gs.W + u.x) * 79;
        m[0] = 1;
        // UnitAction.getValidActionArray(u, gs, utt, mask, 7, 1), UnitAction.java:717-757
        for (auto& ua : gs.unit_actions(u, 10)) {
            m[1 + ua.type] = 1;
            switch (ua.type) {
                case T_MOVE: m[1 + 6 + ua.param] = 1; break;
                case T_HARVEST: m[1 + 6 + 4 + ua.param] = 1; break;
                case T_RETURN: m[1 + 6 + 8 + ua.param] = 1; break;
                case T_PRODUCE:
                    m[1 + 6 + 12 + ua.param] = 1;
                    m[1 + 6 + 16 + ua.utype] = 1;
                    break;
                case T_ATTACK: {
                    int rx = ua.x - u.x, ry = ua.y - u.y;
                    m[1 + 6 + 16 + 7 + (3 + ry) * 7 + (3 + rx)] = 1;
                    break;
                }
            }
        }
    }
}

// ---------------------------------------------------------------- rewards
Unit* TraceSnapshot::unit_at(int x, int y) {
    for (auto& u : units)
        if (u.x == x && u.y == y) return &u;
    return nullptr;
}

void compute_rewards(int maxplayer, const TraceSnapshot& te_c, const GameState& after, double* reward, bool* done) {
    TraceSnapshot& te = const_cast<TraceSnapshot&>(te_c);
    const int minplayer = 1 - maxplayer;
    for (int i = 0; i < 6; i++) {
        reward[i] = 0.0;
        done[i] = false;
    }
    // [0] WinLossRewardFunction.java:16-24
    if (after.gameover()) {
        done[0] = true;
        reward[0] = after.winner() == maxplayer ? 1.0 : -1.0;
    }
    for (auto& p : te.actions) {
        const Unit& u = p.first;
        const Action& a = p.second;
        if (u.player != maxplayer) continue;
        // [1] ResourceGatherRewardFunction.java:22-31
        if (a.type == T_HARVEST) reward[1] += 1.0;
        else if (a.type == T_RETURN) reward[1] += 1.0;
        if (a.type == T_PRODUCE && a.utype >= 0) {
            if (a.utype == UT_WORKER) reward[2] += 1.0;                              // ProduceWorker :20-30
            if (a.utype == UT_BARRACKS || a.utype == UT_BASE) reward[3] += 1.0;      // ProduceBuilding :20-30
            if (a.utype == UT_LIGHT || a.utype == UT_HEAVY || a.utype == UT_RANGED) reward[5] += 1.0;  // ProduceCombatUnit :20-30
        }
        // [4] AttackRewardFunction.java:20-36
        if (a.type == T_ATTACK) {
            Unit* other = te.unit_at(a.x, a.y);
            if (other) {
                if (other->player == minplayer) reward[4] += 1.0;
                else if (other->player == maxplayer) reward[4] -= 1.0;
            }
        }
    }
    // ResourceGather done flag (:32-41): true unless some Resource unit still holds resources
    done[1] = true;
    for (auto& u : after.units)
        if (u->type == UT_RESOURCE && u->res > 0) {
            done[1] = false;
            break;
        }
}

// ---------------------------------------------------------------- obs encode (vec_env.py:311-321)
void encode_obs(const std::vector<int32_t>& raw, int H, int W, int32_t* out, int planes) {
    static const int np_[6] = {5, 5, 3, 8, 6, 2};
    static const int off[6] = {0, 5, 10, 13, 21, 27};
    const int HW = H * W;
    for (int c = 0; c < HW; c++) {
        int32_t* o = out + (size_t)c * planes;
        for (int k = 0; k < 29; k++) o[k] = 0;  // visibility planes (if any) are written by the caller
        for (int f = 0; f < 6; f++) {
            int v = raw[f * HW + c];
            if (v < 0) v = 0;
            if (v > np_[f] - 1) v = np_[f] - 1;
            o[off[f] + v] = 1;
        }
    }
}

}  // namespace orc
