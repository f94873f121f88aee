// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.
//
// CPU restatement ("oracle") of the hot path of adFrej/MicroRTS-Py: the Java game engine
// (submodule MicroRTS-KG @ 0c694d4d) plus the Python glue of MicroRTSGridModeVecEnv.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may build, load or call anything in this directory.  The product (the CUDA
// library under microrts_py_b200/csrc) never links or calls it.
//
// Parity pinning: the dynamics (issueSafe / issue / cycle / execute / getUnitActions) and
// the scripted-bot machinery (AbstractionLayerAI, Train/Build/Harvest/Attack, A*) are
// pinned by the reference's 280 golden traces (tests/golden/traces.bin.gz, generated from
// gym_microrts/microrts/data/traces by tests/golden/make_golden.py); observations and masks
// by the known-answer vectors of tests/test_observation.py and tests/test_mask.py.  Parts
// with no golden in the reference (fromVectorAction ordering, reward magnitudes, auto-reset,
// WorkerRush's own policy, partial observability) follow the Java source only: "parity
// unpinned" for those, see DESIGN.md.
//
// All file:line citations are relative to /root/reference/; J/ = gym_microrts/microrts/src/.
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

namespace orc {

// J/rts/UnitAction.java:31-102
enum { T_NONE = 0, T_MOVE = 1, T_HARVEST = 2, T_RETURN = 3, T_PRODUCE = 4, T_ATTACK = 5 };
enum { D_UP = 0, D_RIGHT = 1, D_DOWN = 2, D_LEFT = 3 };
extern const int DIR_DX[4];
extern const int DIR_DY[4];

// How `new UnitAction(TYPE_NONE, d)` behaves.
//  NONE_FORK     : the reference fork forces parameter = -1 for every NONE built through the
//                  (type, direction) constructors (J/rts/UnitAction.java:149-154,163-169), so
//                  every NONE has ETA -1 and completes on the next cycle.  This is what
//                  MicroRTSGridModeVecEnv runs on.
//  NONE_UPSTREAM : parameter = d (upstream microrts).  The golden traces under
//                  data/traces were generated with this behaviour; used only to regenerate
//                  the LightRush traces action for action.
enum { NONE_UPSTREAM = 0, NONE_FORK = 1 };

// J/rts/units/UnitType.java:30-52, values J/rts/units/UnitTypeTable.java:105-290 (VERSION_ORIGINAL)
struct UnitType {
    int id;
    const char* name;
    int cost, hp, minDamage, maxDamage, attackRange;
    int produceTime, moveTime, attackTime, harvestTime, returnTime;
    int harvestAmount, sightRadius;
    bool isResource, isStockpile, canHarvest, canMove, canAttack;
    int nproduces;
    int produces[3];
};
enum { UT_RESOURCE = 0, UT_BASE = 1, UT_BARRACKS = 2, UT_WORKER = 3, UT_LIGHT = 4, UT_HEAVY = 5, UT_RANGED = 6, N_UTYPES = 7 };
const UnitType& utype(int id);
int utype_by_name(const std::string& name);

struct Unit {  // J/rts/units/Unit.java:23-40
    long long id;
    int type;
    int player;
    int x, y;
    int res;
    int hp;
};
using UnitP = std::shared_ptr<Unit>;

struct Action {  // J/rts/UnitAction.java:112-118
    int type = T_NONE;
    int param = -1;  // DIRECTION_NONE
    int x = 0, y = 0;
    int utype = -1;  // null
};
Action act_none_default();                   // new UnitAction(TYPE_NONE)          UnitAction.java:139-141
Action act_none(int duration, int none_mode);  // new UnitAction(TYPE_NONE, d)       UnitAction.java:149-154
Action act_dir(int type, int dir);           // new UnitAction(type, dir)
Action act_produce(int dir, int ut);         // new UnitAction(TYPE_PRODUCE, dir, ut)
Action act_attack(int x, int y);             // new UnitAction(TYPE_ATTACK_LOCATION, x, y)
bool act_equals(const Action& a, const Action& b);  // UnitAction.java:197-214
int act_eta(const Unit& u, const Action& a);        // UnitAction.java:313-335

struct Usage {  // J/rts/ResourceUsage.java
    std::vector<int> pos;
    int res[2] = {0, 0};
    void merge(const Usage& o);
};

struct Assign {  // J/rts/UnitActionAssignment.java:10-26
    UnitP unit;
    Action action;
    int time;
};
using AssignP = std::shared_ptr<Assign>;

struct Players {
    int res[2];
};

struct MapTemplate {  // parsed map XML, J/rts/PhysicalGameState.java:685-711
    int W = 0, H = 0;
    std::vector<uint8_t> terrain;  // row-major, 0 free / 1 wall
    int res[2] = {0, 0};
    std::vector<Unit> units;  // XML order
};

using PlayerAction = std::vector<std::pair<UnitP, Action>>;  // J/rts/PlayerAction.java:29

// GameState + PhysicalGameState folded together (J/rts/GameState.java, PhysicalGameState.java).
// Also represents a PartiallyObservableGameState (shallow copy sharing Unit, Assign and
// Players objects, J/rts/PartiallyObservableGameState.java:23-44).
struct GameState {
    int W = 0, H = 0;
    std::shared_ptr<const std::vector<uint8_t>> terrain;
    std::shared_ptr<Players> players;
    std::vector<UnitP> units;   // pgs.units, a LinkedList in creation order
    std::vector<AssignP> acts;  // unitActions, a LinkedHashMap in issue order
    int time = 0;
    int none_mode = NONE_FORK;
    long long* next_id = nullptr;  // Unit.next_ID stand-in (never observable)
    int pogs_player = -1;          // >= 0: this state is a PartiallyObservableGameState of that player (`instanceof` test of the PO bots)
    bool observable(int x, int y) const;  // PartiallyObservableGameState.observable, PartiallyObservableGameState.java:48-58

    int terrain_at(int x, int y) const { return (*terrain)[x + y * W]; }
    Unit* unit_at(int x, int y) const;            // PhysicalGameState.java:248-255 (first match)
    bool contains(const Unit* u) const;           // pgs.getUnits().contains(u)
    AssignP assignment(const Unit* u) const;      // unitActions.get(u)
    void put_assignment(const AssignP& a);        // LinkedHashMap.put (re-put keeps the slot)
    void remove_assignment(const Unit* u);
    void remove_unit(const Unit* u);              // GameState.java:79-82
    Usage usage(const Unit& u, const Action& a) const;               // UnitAction.java:252-302
    bool consistent(const Usage& base, const Usage& other) const;    // ResourceUsage.java:31-50
    std::vector<Action> unit_actions(const Unit& u, int noneDuration) const;  // Unit.java:382-522
    bool can_execute(const Unit& u, const Action& a) const;          // Unit.java:531-534
    bool issue(PlayerAction& pa);                 // GameState.java:249-328
    bool issue_safe(PlayerAction& pa);            // GameState.java:338-408
    bool cycle();                                 // GameState.java:553-571
    void execute(const UnitP& u, const Action& a);  // UnitAction.java:344-471
    bool gameover() const;                        // PhysicalGameState.java:346-372
    int winner() const;                           // PhysicalGameState.java:319-338
    bool free_cell(int x, int y) const;           // GameState.java:191-207
    bool action_allowed(const Unit& u, const Action& a) const;  // GameState.java:434-457
    Usage resource_usage() const;                 // GameState.java:652-664
    bool integrity_check() const;                 // GameState.java:703-719
    void fill_with_nones(PlayerAction& pa, int player, int duration) const;  // PlayerAction.java:217-235
    // raw 6 feature maps [6][H][W], GameState.java:922-969
    void vector_observation(int player, std::vector<int32_t>& out) const;
    GameState pogs(int player) const;             // PartiallyObservableGameState.java:23-58
};

GameState make_game(const MapTemplate& m, int none_mode, long long* next_id);

// PlayerAction.fromVectorAction, J/rts/PlayerAction.java:384-417 (+ UnitAction.java:681-715).
// `actions` rows are [cell, type, p_move, p_harvest, p_return, p_prod_dir, p_prod_type, p_attack].
PlayerAction from_vector_action(const GameState& view, const int32_t* actions, int n, int player);

// JNIGridnetClient.getMasks / UnitAction.getValidActionArray: mask79 is [H][W][79].
void build_masks(const GameState& gs, int player, int32_t* mask79);

// six reward terms in vec_env.py:185-195 order; te = actions issued this step (post issueSafe
// rewrite) and the pre-cycle unit snapshot (only positions/owners matter).
struct TraceSnapshot {
    std::vector<Unit> units;                       // te.getPhysicalGameState() clone
    std::vector<std::pair<Unit, Action>> actions;  // te.getActions() (unit copied by value: owner only)
    Unit* unit_at(int x, int y);
};
void compute_rewards(int maxplayer, const TraceSnapshot& te, const GameState& after, double* reward6, bool* done6);

// vec_env.py:311-321 one-hot encode of the 6 raw maps (+2 visibility planes when partial_obs)
void encode_obs(const std::vector<int32_t>& raw, int H, int W, int32_t* out, int planes);

// ---------------------------------------------------------------- scripted bots
struct AbstractAction;  // J/ai/abstraction/AbstractAction.java
struct BotAI {
    virtual ~BotAI() {}
    virtual PlayerAction get_action(int player, const GameState& gs) = 0;
    virtual void reset() = 0;
};
enum { AI_PASSIVE = 0, AI_WORKER_RUSH = 1, AI_LIGHT_RUSH = 2, AI_PO_WORKER_RUSH = 3, AI_PO_LIGHT_RUSH = 4, AI_PO_HEAVY_RUSH = 5, AI_PO_RANGED_RUSH = 6 };
std::unique_ptr<BotAI> make_bot(int ai_id);

// ---------------------------------------------------------------- vectorised client
// JNIGridnetVecClient + JNIGridnetClient + JNIGridnetClientSelfPlay + the Python class glue.
struct VecEnv {
    int num_selfplay = 0, num_bot = 0, N = 0, H = 0, W = 0, max_steps = 0;
    bool partial_obs = false;
    int none_mode = NONE_FORK;
    int planes = 29;
    std::vector<MapTemplate> maps;  // one per env (self-play uses maps[2g])
    std::vector<int> ai_ids;
    std::vector<int> ai1_ids;  // JNIBotClient games (bot vs bot): the scripted player 0 per bot env, -1 = the agent plays
    struct Game {
        GameState gs;
        std::unique_ptr<BotAI> bot;  // bot envs only
        std::unique_ptr<BotAI> bot1; // JNIBotClient games only: ai1
        long long next_id = 0;
    };
    std::vector<Game> selfplay;  // num_selfplay/2
    std::vector<Game> botgames;  // num_bot
    std::vector<int> env_steps;
    std::vector<uint8_t> have_source_mask;
    std::vector<int32_t> source_mask;  // (N, HW) of the last get_masks call (vec_env.py:1099)
    bool source_mask_valid = false;
    int threads = 1;

    void init();
    int ai1_of(int b) const { return b < (int)ai1_ids.size() ? ai1_ids[b] : -1; }
    void reset(int32_t* obs);
    void get_masks(int32_t* mask78, int32_t* source_unit_mask);
    void step(const int32_t* actions, int32_t* obs, double* rewards, uint8_t* done);
    // cycle_maps (vec_env.py:1038-1056): `client.mapPath = next(self.next_map); client.reset(...)` for the
    // client of env `env` (a bot env, or view 0 of a self-play game); the rows of its view(s) in `obs`
    // (the full (N,H,W,P) array) are replaced as the Python loop does.
    void swap_map_and_reset(int env, const MapTemplate& m, int32_t* obs);

   private:
    void observe(const GameState& view, const GameState& truth, int player, int32_t* obs_out) const;
    void reset_game(Game& g, const MapTemplate& m, int ai_id, int ai1_id = -1);
    void step_selfplay(int g, const int32_t* actions, int32_t* obs, double* rewards, uint8_t* done);
    void step_bot(int b, const int32_t* actions, int32_t* obs, double* rewards, uint8_t* done);
    std::vector<int32_t> gather_actions(int env, const int32_t* actions, const GameState& view, int player) const;
};

}  // namespace orc
