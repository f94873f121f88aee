// Device-resident game state and rule tables of the B200-native MicroRTS simulator.
//
// Everything in this file compiles both as device code (nvcc, sm_100a) and as plain host
// C++ (tests/hostsim builds the very same logic for CPU-side debugging; the product never
// runs it).  Citations are relative to /root/reference/, J/ = gym_microrts/microrts/src/.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define MRTS_HD __host__ __device__ __forceinline__
#define MRTS_HDN __host__ __device__ __noinline__
#else
#define MRTS_HD inline
#define MRTS_HDN inline
#endif

namespace mrts {

// J/rts/UnitAction.java:31-102
enum { T_NONE = 0, T_MOVE = 1, T_HARVEST = 2, T_RETURN = 3, T_PRODUCE = 4, T_ATTACK = 5 };
enum { D_UP = 0, D_RIGHT = 1, D_DOWN = 2, D_LEFT = 3 };
enum { UT_RESOURCE = 0, UT_BASE = 1, UT_BARRACKS = 2, UT_WORKER = 3, UT_LIGHT = 4, UT_HEAVY = 5, UT_RANGED = 6, N_UTYPES = 7 };
// how `new UnitAction(TYPE_NONE, d)` behaves: the fork forces parameter -1 (J/rts/UnitAction.java:149-154)
enum { NONE_UPSTREAM = 0, NONE_FORK = 1 };
enum { AI_NONE = -1, AI_PASSIVE = 0, AI_WORKER_RUSH = 1, AI_LIGHT_RUSH = 2, AI_PO_WORKER_RUSH = 3, AI_PO_LIGHT_RUSH = 4, AI_PO_HEAVY_RUSH = 5, AI_PO_RANGED_RUSH = 6, AI_COUNT = 7 };
// scripted opponents built on AbstractionLayerAI: abstract-action table in the blob, path searches
MRTS_HD bool ai_plans(int ai) { return ai >= AI_WORKER_RUSH && ai < AI_COUNT; }
// the PO* variants (ai/abstraction/partialobservability/): same bot, melee units explore when no enemy is in view
MRTS_HD bool ai_explores(int ai) { return ai >= AI_PO_WORKER_RUSH && ai <= AI_PO_RANGED_RUSH; }
// LightRush / HeavyRush / RangedRush (ai/abstraction/{Light,Heavy,Ranged}Rush.java) are one bot that differs in the
// unit type its barracks train (LightRush.java:137-138, HeavyRush.java:136-137, RangedRush.java:124-125)
MRTS_HD bool ai_barracks_rush(int ai) { return ai == AI_LIGHT_RUSH || (ai >= AI_PO_LIGHT_RUSH && ai <= AI_PO_RANGED_RUSH); }
MRTS_HD int ai_rush_unit(int ai) { return ai == AI_PO_HEAVY_RUSH ? UT_HEAVY : (ai == AI_PO_RANGED_RUSH ? UT_RANGED : UT_LIGHT); }

MRTS_HD int dir_dx(int d) { return d == D_RIGHT ? 1 : (d == D_LEFT ? -1 : 0); }
MRTS_HD int dir_dy(int d) { return d == D_DOWN ? 1 : (d == D_UP ? -1 : 0); }

// ---------------------------------------------------------------------------------------
// Unit type table, VERSION_ORIGINAL (J/rts/units/UnitTypeTable.java:105-290, defaults
// J/rts/units/UnitType.java:41-108), packed into immediates so that no memory is touched:
// index = type id 0..6 = Resource, Base, Barracks, Worker, Light, Heavy, Ranged.
// ---------------------------------------------------------------------------------------
MRTS_HD int nib(uint32_t packed, int t) { return (packed >> (4 * t)) & 0xF; }
MRTS_HD int byt(uint64_t packed, int t) { return (int)((packed >> (8 * t)) & 0xFF); }
//                                         type:  6 5 4 3 2 1 0
MRTS_HD int ut_cost(int t) { return nib(0x22215A1u, t); }    // 1,10,5,1,2,2,2
MRTS_HD int ut_hp(int t) { return nib(0x14414A1u, t); }      // 1,10,4,1,4,4,1
MRTS_HD int ut_damage(int t) { return nib(0x1421111u, t); }  // 1,1,1,1,2,4,1 (min == max)
MRTS_HD int ut_range(int t) { return t == UT_RANGED ? 3 : 1; }
MRTS_HD int ut_sight(int t) { return nib(0x3223350u, t); }   // 0,5,3,3,2,2,3
MRTS_HD int ut_produce_time(int t) { return byt(0x0064785032C8FA0AULL, t); }  // 10,250,200,50,80,120,100
MRTS_HD int ut_move_time(int t) { return byt(0x000A0C080A0A0A0AULL, t); }     // 10,10,10,10,8,12,10
MRTS_HD int ut_attack_time(int t) { return byt(0x00050505050A0A0AULL, t); }   // 10,10,10,5,5,5,5
MRTS_HD int ut_harvest_time(int t) { return t == UT_WORKER ? 20 : 10; }
MRTS_HD bool ut_is_resource(int t) { return t == UT_RESOURCE; }
MRTS_HD bool ut_is_stockpile(int t) { return t == UT_BASE; }
MRTS_HD bool ut_can_harvest(int t) { return t == UT_WORKER; }
MRTS_HD bool ut_can_move(int t) { return t >= UT_WORKER; }
MRTS_HD bool ut_can_attack(int t) { return t >= UT_WORKER; }
// bit k set = this type produces type k: Base->Worker, Barracks->Light/Heavy/Ranged, Worker->Base/Barracks
MRTS_HD int ut_produces(int t) { return byt(0x0000000006700800ULL, t); }

// ---------------------------------------------------------------------------------------
// State records.  One game = header + Unit[cap] (+ AbstractAct[cap] for scripted-bot games),
// laid out contiguously ("blob"), 16-byte aligned so a thread group moves it with uint4 copies.
// Unit array order IS pgs.units order (creation order; removal compacts, J/rts/PhysicalGameState.java:185-195).
// Assignment (unitActions LinkedHashMap, J/rts/GameState.java:42) is stored inside the unit;
// a_seq is the monotone issue counter that restores the map's iteration order.
// ---------------------------------------------------------------------------------------
struct Unit {  // 32 bytes
    uint8_t x, y, type;
    int8_t player;   // -1 neutral
    int16_t hp;
    int16_t a_param;  // direction, or NONE duration
    int32_t res;
    int8_t a_type;    // -1 = no assignment
    int8_t a_utype;   // produced type or -1
    uint8_t a_x, a_y; // attack target
    int32_t a_time;   // time of issue
    uint32_t a_seq;   // issue order
    uint16_t uid;     // stable identity inside one episode (abstract actions refer to it)
    uint8_t flags;    // UF_*
    uint8_t vis;      // bit p: unit is in player p's partially observable view (taken this step)
    int8_t vtype[2];  // action type recorded in view p's assignment copy, -1 none
    uint16_t pre;     // this step's decoded agent action for the unit (pack_action), staged by the whole group before the rules run
};
static_assert(sizeof(Unit) == 32, "Unit must be 32 bytes");
enum { UF_DEAD = 1, UF_THIS_STEP = 2, UF_LISTED = 4 /* scratch mark of fill_with_nones, always cleared again */,
       UF_AA_ATTACK = 8 /* scripted bot: the unit's abstract-action entry is an Attack */ };

// abstract action of the scripted bots (J/ai/abstraction/{Train,Build,Harvest,Attack}.java)
enum { AA_TRAIN = 0, AA_BUILD = 1, AA_HARVEST = 2, AA_ATTACK = 3, AA_MOVE = 4, AA_KIND_MASK = 15, AA_OWNER0 = 16 /* stored entries only */ };
struct AbstractAct {  // 12 bytes; table order = LinkedHashMap insertion order (AbstractionLayerAI.java:36)
    uint16_t unit;    // uid
    uint8_t kind;
    int8_t utype;     // Train / Build
    int16_t x, y;     // Build target (findBuildingPosition may yield (-1,0))
    uint16_t target;  // uid or NO_UID: Harvest resource / Attack target
    uint16_t base;    // uid or NO_UID: Harvest base
    // Train's `completed` flag lives in `target` (0/1)
};
static_assert(sizeof(AbstractAct) == 12, "AbstractAct must be 12 bytes");
enum : uint16_t { NO_UID = 0xFFFF };

struct GameHdr {  // 32 bytes
    int32_t nunits;
    int32_t time;
    int32_t res[2];
    int32_t steps;      // envSteps (JNIGridnetVecClient.java:245,270)
    uint32_t next_seq;
    uint16_t next_uid;
    uint16_t n_aa;      // abstract-action table length
    int32_t pad;
};
static_assert(sizeof(GameHdr) == 32, "GameHdr must be 32 bytes");

enum : uint16_t { G_NONE = 0xFFFF, G_WALL = 0xFFFE };  // grid codes; anything below = unit index

// per-map template on the device (PhysicalGameState.load result, J/rts/PhysicalGameState.java:685-711)
struct MapTpl {
    int32_t W, H;
    int32_t res[2];
    int32_t nunits;
    int32_t grid_off;   // offset (in uint16) of the wall grid inside the template pool
    int32_t unit_off;   // offset (in Unit) of the initial unit list inside the unit pool
    int32_t pad;
};

// issued-action record of one step (what TraceEntry.addPlayerAction keeps, J/rts/TraceEntry.java:51-72)
struct PAEnt {  // 8 bytes
    uint16_t unit;  // index into the unit array
    int8_t type;
    int8_t utype;
    int16_t param;
    uint8_t x, y;
};

}  // namespace mrts
