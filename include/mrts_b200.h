/* C ABI of the B200-native vectorised MicroRTS simulator (libmrts_b200.so).
 *
 * This is the boundary the reference crosses with JPype today: gym_microrts/envs/vec_env.py
 * builds a Java `tests.JNIGridnetVecClient` (vec_env.py:256-276) and calls its reset / gameStep /
 * getMasks / close methods.  Each entry point below names the Java method it replaces
 * (J/ = gym_microrts/microrts/src/).  All tensor pointers are DEVICE pointers owned by the
 * caller; all work is enqueued on the given CUDA stream (cudaStream_t passed as void*, NULL =
 * legacy default stream) and nothing synchronises unless stated.  Every function returns 0 on
 * success or a negative MRTS_E* code; mrts_last_error() gives the message for the calling thread.
 *
 * Layouts (C-contiguous):
 *   obs      (N, H, W, P)   P = 29, or 31 with partial_obs; element = int32 (reference dtype) or uint8
 *   mask     (N, H*W, 78)   same element type as obs; 78 = [6,4,4,4,4,7,49]
 *   actions  (N, H*W, 7)    int32, MultiDiscrete [6,4,4,4,4,7,49] per cell
 *   reward   (N, 6)         float64: WinLoss, ResourceGather, ProduceWorker, ProduceBuilding, Attack, ProduceCombatUnit
 *   done     (N,)           uint8
 * N = num_selfplay_envs + num_bot_envs; env order as in JNIGridnetVecClient.java:117-124.
 */
#ifndef MRTS_B200_H
#define MRTS_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MRTS_OK 0
#define MRTS_EINVAL (-1)
#define MRTS_ECUDA (-2)
#define MRTS_ENOMEM (-3)

#define MRTS_AI_PASSIVE 0     /* ai.PassiveAI      (gym_microrts/microrts_ai.py:13-16) */
#define MRTS_AI_WORKER_RUSH 1 /* ai.abstraction.WorkerRush (microrts_ai.py:19-22)      */
#define MRTS_AI_LIGHT_RUSH 2  /* ai.abstraction.LightRush  (microrts_ai.py:25-28)      */
#define MRTS_AI_PO_WORKER_RUSH 3 /* ai.abstraction.partialobservability.POWorkerRush (microrts_ai.py:37-40) */
#define MRTS_AI_PO_LIGHT_RUSH 4  /* ai.abstraction.partialobservability.POLightRush  (microrts_ai.py:31-34) */
#define MRTS_AI_PO_HEAVY_RUSH 5  /* ai.abstraction.partialobservability.POHeavyRush  (microrts_ai.py:43-46) */
#define MRTS_AI_PO_RANGED_RUSH 6 /* ai.abstraction.partialobservability.PORangedRush (microrts_ai.py:49-52) */

/* How mrts_step produces the tensors (mrts_config::pipeline); see DESIGN.md section 4.  SPLIT: the rules kernels write a
 * packed form (one word per cell, one entry per acting unit) and an expansion kernel streams it into the tensors with bulk
 * async stores, the slab cut into `chunks` so that the two overlap.  DIRECT: the rules pass writes the tensors itself. */
#define MRTS_PIPELINE_AUTO 0
#define MRTS_PIPELINE_DIRECT 1
#define MRTS_PIPELINE_SPLIT 2

/* One parsed map: what PhysicalGameState.load returns (J/rts/PhysicalGameState.java:54-65,685-711). */
typedef struct mrts_map_template {
    int32_t width, height;
    const uint8_t* terrain;  /* width*height, row-major, 0 free / 1 wall (host pointer)            */
    int32_t resources[2];    /* players' starting resources                                        */
    int32_t n_units;
    const int32_t* units7;   /* n_units rows: id, type, player, x, y, resources, hitpoints (host)  */
} mrts_map_template;

/* Constructor arguments of JNIGridnetVecClient (J/tests/JNIGridnetVecClient.java:106-142). */
typedef struct mrts_config {
    int32_t num_selfplay_envs; /* even                                                             */
    int32_t num_bot_envs;
    int32_t max_steps;
    int32_t partial_obs;
    int32_t none_mode;         /* 1 = this fork (every NONE has parameter -1, UnitAction.java:149-154), 0 = upstream */
    int32_t device;            /* CUDA device ordinal                                              */
    int32_t n_maps;
    const mrts_map_template* maps;
    const int32_t* env_map;    /* N entries: template index per env (self-play game g uses entry 2g) */
    const int32_t* bot_ai;     /* num_bot_envs entries: MRTS_AI_*                                  */
    int32_t elem_bytes;        /* 4 = int32 obs/mask (reference dtype), 1 = uint8                  */
    int32_t threads_per_game;  /* 0 = choose from the map size; n > 0 = n threads per game; n < 0 = one warp per game, -n games per block */
    int32_t staged_units;      /* 0 = choose; units per game staged in shared memory by the warp kernel (games that may
                                  outgrow it take the full-capacity launch); tuning / tests                          */
    const int32_t* bot_ai1;    /* NULL, or num_bot_envs entries: MRTS_AI_* playing player 0 of that bot env (bot vs bot,
                                  J/tests/JNIBotClient.java:108-134, MicroRTSBotVecEnv), -1 = the agent plays it.  The
                                  action rows of such envs are ignored; both bots act on the full state             */
    int32_t pipeline;          /* MRTS_PIPELINE_*: 0 = choose from the slab size                                      */
    int32_t chunks;            /* slab chunks of the split pipeline, 0 = choose                                       */
} mrts_config;

typedef struct mrts_env mrts_env;

/* new JNIGridnetVecClient(...) */
int mrts_create(const mrts_config* cfg, mrts_env** out);
/* JNIGridnetVecClient.close (:318-329) */
void mrts_destroy(mrts_env* env);

/* JNIGridnetVecClient.reset (:179-211): every game restarts from its map; obs / mask may be NULL. */
int mrts_reset(mrts_env* env, void* obs, void* mask, void* stream);
/* JNIGridnetVecClient.gameStep (:213-297) fused with getMasks (:307-316) of the resulting state:
 * any of obs / mask / reward / done may be NULL; actions may be NULL when every env is bot vs bot (mrts_config::bot_ai1). */
int mrts_step(mrts_env* env, const int32_t* actions, void* obs, void* mask, double* reward, uint8_t* done, void* stream);
/* JNIGridnetVecClient.getMasks (:307-316) of the current state. */
int mrts_get_masks(mrts_env* env, void* mask, void* stream);

/* cycle_maps of MicroRTSGridModeVecEnv (gym_microrts/envs/vec_env.py:132-133, 1038-1056): after every
 * mrts_step, each game whose episode ended takes the next entry of a round-robin list shared by the whole
 * env (ascending env order, starting at entry 0) as its map and restarts from it; the obs / mask /
 * next_actions rows of its view(s) are those of the new map, reward and done stay the step's.
 * template_idx[k] indexes cfg->maps of mrts_create (so rotation maps are part of that table and share its
 * size).  n = 0 turns the rotation off.  Only for all-bot or all-self-play envs: with both kinds the
 * reference's `done_idx < num_bot_envs` test addresses the wrong clients.  Synchronises the device. */
int mrts_set_cycle_maps(mrts_env* env, const int32_t* template_idx, int n);

/* gameStep + getMasks + (optionally) a masked-uniform random policy for the resulting state,
 * the sampling recipe of the reference's hello_world.py:27-67 moved onto the device so that a
 * random-policy rollout never touches the host.  next_actions (N, H*W, 7) int32 receives, at the
 * cell of every unit that may act next step, one action drawn uniformly per component from the
 * unit's mask (0 where a component has no legal value); other cells are left as they are.
 * next_actions may alias actions.  Reproducible from (seed, number of sampled launches so far). */
typedef struct mrts_step_io {
    const int32_t* actions; /* (N, H*W, 7) int32, required                                      */
    void* obs;              /* nullable                                                         */
    void* mask;             /* nullable                                                         */
    double* reward;         /* (N, 6), nullable                                                 */
    uint8_t* done;          /* (N,), nullable                                                   */
    int32_t* next_actions;  /* nullable                                                         */
    uint64_t seed;
} mrts_step_io;
int mrts_step_ex(mrts_env* env, const mrts_step_io* io, void* stream);
/* The same sampler on the current state (e.g. after mrts_reset). */
int mrts_sample_actions(mrts_env* env, int32_t* next_actions, uint64_t seed, void* stream);

/* The packed form of the same outputs -- what the rules kernels produce before the expansion, ~1.4 KB per 16x16 view
 * instead of 110 KB -- for consumers that unpack on the fly (or not at all):
 *   obs_words    (N, H*W) uint32         bit k of a cell's word = plane k of the observation
 *   mask_entries (N, 1 + cap, 4) uint32  row 0 = {n, 0, 0, 0}; rows 1..n = {cell, bits 0-31, bits 32-63, bits 64-77} of the
 *                                        mask rows that are not all zero, in ascending cell order; cap = mrts_unit_capacity
 * mrts_packed_sizes gives the element counts of the two arrays.  mrts_expand turns them into the obs / mask tensors of
 * mrts_step (either may be NULL); it needs H*W*element_bytes to be a multiple of 16. */
typedef struct mrts_packed_io {
    const int32_t* actions;  /* (N, H*W, 7) int32; NULL only when every env is bot vs bot */
    uint32_t* obs_words;     /* nullable */
    uint32_t* mask_entries;  /* nullable */
    double* reward;          /* (N, 6), nullable */
    uint8_t* done;           /* (N,), nullable */
    int32_t* next_actions;   /* nullable */
    uint64_t seed;
} mrts_packed_io;
int mrts_packed_sizes(const mrts_env* env, int64_t* obs_words, int64_t* mask_entry_words);
int mrts_reset_packed(mrts_env* env, uint32_t* obs_words, uint32_t* mask_entries, void* stream);
int mrts_step_packed(mrts_env* env, const mrts_packed_io* io, void* stream);
int mrts_expand(mrts_env* env, const uint32_t* obs_words, const uint32_t* mask_entries, void* obs, void* mask, void* stream);

/* mrts_expand without an env and without allocating: for code under CUDA graph capture, serving loops and harnesses that
 * may not call cudaMalloc.  This entry point has NO counterpart in the reference.
 * The shape that mrts_expand takes from the env comes with the call: n_views views of height x width
 * cells, mask_entries with 1 + unit_capacity rows per view, 31 planes when partial_obs else 29, element_bytes 4 (int32)
 * or 1 (uint8).  The CURRENT device is used; all four arrays live on it.  Scratch memory (a work counter and the order in
 * which views are handed out, heaviest first) is a file-scope __device__ array of MRTS_EXPAND_STATIC_SCRATCH_BYTES that
 * every GPU loading the library holds until exit; the call makes no cudaMalloc / cudaFree / allocating library call.
 * Limits and rules:
 *   - n_views <= MRTS_EXPAND_STATIC_MAX_VIEWS, else MRTS_EINVAL (split the call, or use mrts_expand);
 *   - H*W*element_bytes a multiple of 16, arrays 16-byte aligned (as mrts_expand);
 *   - every call on a device shares the one scratch array: calls whose kernels overlap on the GPU (different streams
 *     with no ordering between them, or two host threads) RACE; calls on one stream, or ordered by events, are safe.
 *     Each call rewrites the scratch it uses; nothing needs clearing between calls;
 *   - the first call on a device does the one-time set-up (cudaGetSymbolAddress, kernel attributes): make it outside
 *     graph capture.
 * Output is identical to mrts_expand's (tests/test_zz_no_alloc.py). */
#define MRTS_EXPAND_STATIC_MAX_VIEWS 65536
#define MRTS_EXPAND_STATIC_SCRATCH_BYTES (256 + 4 * MRTS_EXPAND_STATIC_MAX_VIEWS)
int mrts_expand_static(const uint32_t* obs_words, const uint32_t* mask_entries, void* obs, void* mask, int n_views, int height, int width,
                       int unit_capacity, int partial_obs, int element_bytes, void* stream);

/* Host-buffer forms of reset / gameStep + getMasks: every pointer is a HOST pointer (pageable is fine),
 * the call returns when the arrays are filled -- what the reference's NumPy-returning methods do
 * (gym_microrts/envs/vec_env.py:278-282, 968-1061, 1091-1101).  obs (N,H,W,P) and mask (N,H*W,78) have
 * the element type of the env (int32 / uint8); reward (N,6) float64; done (N,) uint8; actions
 * (N,H*W,7) int32 or int64 (action_elem_bytes 4 / 8) -- only the rows of cells whose mask entry 0 was 1 in
 * the mask last returned are read, as in vec_env.py:972-974.  Any output pointer may be NULL.  The
 * values cross PCIe as one bit per element and are widened on the host by a thread pool (csrc/mrts_host.h);
 * the library owns its own stream and pinned staging buffers for these calls. */
int mrts_reset_host(mrts_env* env, void* obs, void* mask);
int mrts_step_host(mrts_env* env, const void* actions, int action_elem_bytes, void* obs, void* mask, double* reward, uint8_t* done);
/* Incremental forms for callers that alternate between a few result buffers (as vec envs usually do).
 * `slot` (0 .. MRTS_HOST_SLOTS-1) names one (obs, mask) pair of caller arrays; the caller promises that the
 * arrays it passes with a slot are the ones it passed with that slot before, NOT MODIFIED since.  The
 * library keeps the words it last widened into them and rewrites only the cells whose value changed
 * (between two steps most of a map does not change): the host then writes a few hundred bytes per
 * env-step instead of ~110 KB at 16x16.  First use of a slot, or a different pointer, = full write;
 * slot < 0 = the plain forms above.  mrts_host_forget_slot drops what is remembered about a slot.
 * With both arrays given the comparison runs on the DEVICE (csrc/mrts_delta.cu): the slot's words stay in HBM, a
 * kernel emits the list of cells that differ plus one bit per cell for the source-unit mask, and one copy of a few
 * tens of KB (count, rewards, dones, bitmap, entries) replaces the 16 bytes per cell of the plain forms. */
#define MRTS_HOST_SLOTS 4
int mrts_reset_host_slot(mrts_env* env, void* obs, void* mask, int slot);
int mrts_step_host_slot(mrts_env* env, const void* actions, int action_elem_bytes, void* obs, void* mask, double* reward, uint8_t* done, int slot);
int mrts_host_forget_slot(mrts_env* env, int slot);
/* The transport codec alone, host arrays in and out (no device needed): observation words (n_env, hw) ->
 * one-hot (n_env, hw, planes); mask words (n_env, hw, 3) -> (n_env, hw, 78); action grid (n_env, hw, 7) ->
 * 16-bit rows at the cells whose mask word 0 has bit 0 set. */
int mrts_host_expand_obs(const uint32_t* words, int64_t n_env, int hw, int planes, int elem_bytes, void* out);
int mrts_host_expand_mask(const uint32_t* mask_words, int64_t n_env, int hw, int elem_bytes, void* out);
int mrts_host_pack_actions(const void* actions, int action_elem_bytes, const uint32_t* mask_words, int64_t n_env, int hw, uint16_t* out);
/* the incremental widening alone: `shadow` = the words `out` currently holds the expansion of (updated in place) */
int mrts_host_expand_obs_delta(const uint32_t* words, uint32_t* shadow, int64_t n_env, int hw, int planes, int elem_bytes, void* out);
int mrts_host_expand_mask_delta(const uint32_t* mask_words, uint32_t* shadow, int64_t n_env, int hw, int elem_bytes, void* out);
/* the codec of the device-delta slot calls: action rows picked by a bitmap over the flat cell index (only set cells of `out`
 * are written), and entries (count, 6) {flat cell, obs word, mask words 0..2, span} patched into obs (cells, planes) / mask (cells, 78);
 * span = first | last << 5 changed observation plane, first << 10 | last << 17 changed mask entry (first > last: row unchanged) */
int mrts_host_pack_actions_bitmap(const void* actions, int action_elem_bytes, const uint32_t* bitmap, int64_t cells, uint16_t* out);
int mrts_host_patch_cells(const uint32_t* entries, int64_t count, int planes, int elem_bytes, void* obs, void* mask);
/* the action rows of the set cells as a list: pairs (n, 2) {flat cell, packed row}, any order; *n_pairs = n (pairs beyond `cap` are not written) */
int mrts_host_pack_actions_sparse(const void* actions, int action_elem_bytes, const uint32_t* bitmap, int64_t cells, uint32_t* pairs, int64_t cap, int64_t* n_pairs);
int mrts_host_threads(const mrts_env* env); /* worker threads of the widening pool (0 before the first host call) */
/* profiling aid: accumulated over the host-buffer step calls {seconds packing actions, uploading + launching, waiting for the device,
 * widening / patching + copying out; bytes host->device, changed cells, bytes device->host, calls}; reset != 0 clears them */
int mrts_host_timing(mrts_env* env, double* out8, int reset);

/* Introspection (synchronises the stream). */
int mrts_num_envs(const mrts_env* env);
int mrts_num_games(const mrts_env* env);
int mrts_planes(const mrts_env* env);
int mrts_unit_capacity(const mrts_env* env);
int mrts_threads_per_game(const mrts_env* env);
int mrts_smem_bytes(const mrts_env* env);
int mrts_staged_units(const mrts_env* env);     /* units per game the warp kernel stages in shared memory */
int mrts_launches_per_step(const mrts_env* env); /* kernels the last launch enqueued: rules per game class and slab chunk, the
                                                    full-capacity launch when staged < capacity, the expansion per chunk, map rotation */
int mrts_pipeline(const mrts_env* env);          /* MRTS_PIPELINE_DIRECT or MRTS_PIPELINE_SPLIT */
int mrts_chunks(const mrts_env* env);            /* slab chunks of the split pipeline */
/* registers per thread, games per block and shared memory per game of the light (0) / heavy (1) rules kernel in use */
int mrts_kernel_info(const mrts_env* env, int heavy, int* regs, int* games_per_block, int* smem_per_game);
/* State of one game for cross-checks against the oracle: hdr5 = time, res0, res1, steps, nunits;
 * rows12 = per unit: type, player, x, y, res, hp, a_type (-1 idle), a_param, a_x, a_y, a_utype, a_time.
 * Returns the number of units (or a negative error); host pointers. */
int mrts_export_state(mrts_env* env, int game, int32_t* hdr5, int32_t* rows12, int cap_rows, void* stream);
/* The inverse (GameState.fromXML, J/rts/GameState.java:763-915): game `game` continues from the given state; n = rows
 * (units, in pgs.units order).  The pending assignments are taken to have been issued in ascending issue_order[i]
 * (n entries, one per row; NULL = by time of issue, then row).  The scripted bot of the game starts with an empty
 * abstract-action table, as after a reset.  Synchronises the device. */
int mrts_import_state(mrts_env* env, int game, const int32_t* hdr5, const int32_t* rows12, int n, const int32_t* issue_order, void* stream);

/* Raw engine access for replaying recorded games (J/test/microrts/TestTracesIntegrity.java:72-126):
 * GameState.issueSafe of one player's actions on one game, rows7 = ux, uy, type, param, ax, ay, utype (host). */
int mrts_debug_issue(mrts_env* env, int game, const int32_t* rows7, int n, void* stream);
/* GameState.cycle on one game; *gameover (host, may be NULL) receives the result. */
int mrts_debug_cycle(mrts_env* env, int game, int* gameover, void* stream);

/* Profiling aid: when device_buf (16 * num_games int64, device) is non-null every launch records per game
 * {cycles of the rules phase, cycles of the whole pass, 6 sub-phase cycle counts, #path searches, #expansions, start time in ns (globaltimer), SM id}; NULL switches it off again. */
int mrts_debug_timing(mrts_env* env, long long* device_buf);

const char* mrts_last_error(void);
const char* mrts_version(void);

#ifdef __cplusplus
}
#endif
#endif
