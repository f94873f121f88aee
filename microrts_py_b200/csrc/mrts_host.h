// Host side of the packed transport used by the host-buffer entry points (mrts_reset_host /
// mrts_step_host, include/mrts_b200.h).
//
// The reference hands NumPy int32 arrays to its caller (gym_microrts/envs/vec_env.py:278-282,
// 1001-1061, 1091-1101): 109,568 B per env-step at 16x16, all of it 0/1 values.  Moving that
// over PCIe costs more than computing it, so the kernels emit one BIT per element (mrts_step.h:
// EnvParams::pobs / pmask), the bits cross the bus, and the functions below only widen them to
// the int32 (or uint8) layout of the caller's arrays -- a transport codec, no game logic: every
// value is computed on the device.  The other direction packs the caller's (N, HW, 7) action
// grid into 16 bits per source cell (same encoding as Sim::pack_action).
#pragma once
#include <stddef.h>
#include <stdint.h>

namespace mrts_host {

// persistent worker pool; parallel_for(n, f) runs f(begin, end) over [0, n) split into contiguous blocks
struct Pool;
Pool* pool_create(int threads /* 0 = all cores this process may use */);
void pool_destroy(Pool* p);
int pool_threads(const Pool* p);
int visible_cores();  // cores this process may run on (sched_getaffinity)

// obs words (n_env, hw) -> one-hot (n_env, hw, planes), elem = 4 (int32) or 1 (uint8)
void expand_obs(Pool* p, const uint32_t* words, size_t n_env, int hw, int planes, int elem, void* out);
// mask words (n_env, hw, 3) -> (n_env, hw, 78), elem = 4 or 1
void expand_mask(Pool* p, const uint32_t* pmask, size_t n_env, int hw, int elem, void* out);
// Incremental forms for arrays the caller hands back unmodified (alternating result buffers): `shadow`
// holds the words last expanded into `out` (same shape as the words); only the cells whose words differ
// are rewritten and `shadow` is brought up to date.  Between two steps most of a map does not change, so
// this replaces ~110 KB of stores per 16x16 env-step by a few hundred bytes plus a compare of 8 KB.
void expand_obs_delta(Pool* p, const uint32_t* words, uint32_t* shadow, size_t n_env, int hw, int planes, int elem, void* out);
void expand_mask_delta(Pool* p, const uint32_t* pmask, uint32_t* shadow, size_t n_env, int hw, int elem, void* out);
// actions (n_env, hw, 7) int32 or int64 -> packed (n_env, hw) uint16, only at cells whose mask words
// (the mask the caller was given for this state) are non-zero; other cells 0
void pack_actions(Pool* p, const void* actions, int action_elem /* 4 or 8 */, const uint32_t* pmask, size_t n_env, int hw, uint16_t* out);

// The same with the source cells given as a bitmap over the flat cell index (bit = mask bit 0, mrts_delta.cu): only the
// rows of set bits are read and only those cells of `out` are written (the device reads no other cell).
void pack_actions_bitmap(Pool* p, const void* actions, int action_elem, const uint32_t* bitmap, size_t cells, uint16_t* out);
// The same as a list: pairs (n, 2) {flat cell, packed row} of the set cells, in any order; returns n (pairs beyond `cap` are
// not written: the caller then falls back to the dense form).  ~4 acting cells per env instead of H*W 16-bit codes cross PCIe.
size_t pack_actions_sparse(Pool* p, const void* actions, int action_elem, const uint32_t* bitmap, size_t cells, uint32_t* pairs, size_t cap);
// entries (count, 6) {flat cell, obs word, mask words 0..2, span} -> those cells of obs (cells, planes) and mask (cells, 78).
// span = first | last << 5 changed plane of the cell's observation row, first << 10 | last << 17 changed entry of its mask
// row (mrts_delta.cu): only those ranges are rewritten -- usually one cache line per tensor instead of the rows' two to six,
// and it is the cache lines touched (each one a miss in arrays of hundreds of MB) that this costs.  Every cell appears at
// most once, either array may be null.
void patch_cells(Pool* p, const uint32_t* entries, size_t count, int planes, int elem, void* obs, void* mask);

}  // namespace mrts_host
