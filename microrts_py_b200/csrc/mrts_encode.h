// Output phases of one environment step, run by the whole thread group that owns a game:
// the one-hot observation (GameState.getVectorObservation, J/rts/GameState.java:922-969, +
// MicroRTSGridModeVecEnv._encode_obs, gym_microrts/envs/vec_env.py:311-321) and the per-cell
// action mask (JNIGridnetClient.getMasks, J/tests/JNIGridnetClient.java:203-216 +
// vec_env.py:1091-1101), written with 128-bit stores straight into the caller's tensors.
//
// Every function takes (tid, nth): the calling thread's rank in its group and the group size.
// The host build (tests/hostsim) calls them with (0, 1).
#pragma once
#include "mrts_core.h"
#if defined(__CUDACC__)
#include "mrts_bulk.cuh"
#endif

namespace mrts {

// observation bit layout inside one cell word (= channel index): hp 0-4, resources 5-9,
// owner 10-12, unit type 13-20, action 21-26, terrain 27-28, visibility 29-30 (partial_obs)
enum { OB_HP = 0, OB_RES = 5, OB_OWNER = 10, OB_TYPE = 13, OB_ACT = 21, OB_TERRAIN = 27, OB_VIS = 29 };
MRTS_HD int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// word of a unit as seen by `viewer` (owner planes are relative: 1 = mine, 2 = opponent's)
MRTS_HD uint32_t unit_obs_word(const Unit& un, int viewer, int act_type) {
    uint32_t w = 1u << (OB_HP + clampi(un.hp, 0, 4));
    w |= 1u << (OB_RES + clampi(un.res, 0, 4));
    w |= 1u << (OB_OWNER + (un.player >= 0 ? ((un.player + viewer) & 1) + 1 : 0));
    w |= 1u << (OB_TYPE + un.type + 1);
    w |= 1u << (OB_ACT + clampi(act_type, 0, 5));
    return w;
}
enum : uint32_t { OBS_EMPTY = (1u << OB_HP) | (1u << OB_RES) | (1u << OB_OWNER) | (1u << OB_TYPE) | (1u << OB_ACT) };
// the word of most cells of a map: nothing on it, no wall, (partial_obs) not seen by the opponent
MRTS_HD uint32_t obs_empty_word(int partial) { return OBS_EMPTY | (1u << OB_TERRAIN) | (partial ? (1u << OB_VIS) : 0u); }

// Phase A (cells): words[c] for every cell.  vp < 0: the true state; vp = p: player p's stale
// partially observable view (membership + assignment copy recorded in Unit::vis / vtype).
// grid0 = the map's wall grid (terrain plane).
MRTS_HD void obs_words_cells(const Sim& s, int viewer, int vp, const uint16_t* grid0, uint32_t* words, int tid, int nth) {
    const int n = s.hdr->nunits;
    for (int c = tid; c < s.HW; c += nth) {
        uint32_t w = OBS_EMPTY;
        const uint16_t g = s.grid[c];
        if (g < G_WALL && s.inview(g, vp)) {
            const Unit& un = s.u[g];
            w = unit_obs_word(un, viewer, vp < 0 ? un.a_type : un.vtype[vp]);
        }
        w |= 1u << (OB_TERRAIN + (grid0[c] == G_WALL ? 1 : 0));
        if (s.partial) {
            // README-literal visibility planes: is this cell within sight of any unit
            // of the viewer's opponent in the TRUE post-step state
            const int x = c % s.W, y = c / s.W;
            bool seen = false;
            for (int j = 0; j < n && !seen; j++) {
                const Unit& o = s.u[j];
                if ((o.flags & UF_DEAD) || o.player != 1 - viewer) continue;
                int dx = o.x - x, dy = o.y - y, r = ut_sight(o.type);
                seen = dx * dx + dy * dy <= r * r;
            }
            w |= 1u << (OB_VIS + (seen ? 1 : 0));
        }
        words[c] = w;
    }
}
// Phase B (units, after a barrier; partial_obs only): units killed this cycle still belong to
// the stale view (hp <= 0 clips to plane 0).
MRTS_HD void obs_words_dead(const Sim& s, int viewer, int vp, uint32_t* words, int tid, int nth) {
    if (vp < 0) return;
    const int n = s.hdr->nunits;
    for (int i = tid; i < n; i += nth) {
        const Unit& un = s.u[i];
        if (!(un.flags & UF_DEAD) || !s.inview(i, vp)) continue;
        const int c = un.x + un.y * s.W;
        const uint32_t keep = words[c] & ((3u << OB_TERRAIN) | (3u << OB_VIS));
        words[c] = keep | unit_obs_word(un, viewer, un.vtype[vp]);
    }
}

// ---- element expansion: 4 (int32) or 16 (uint8) consecutive one-hot elements per 16-byte store
struct Vec16 {
    uint32_t a, b, c, d;
};
MRTS_HD uint32_t spread4(uint32_t nibble) { return (nibble * 0x00204081u) & 0x01010101u; }  // bit k -> byte k

#if defined(__CUDA_ARCH__)
MRTS_HD void store16(void* p, Vec16 v) {
    // streaming 128-bit store: the outputs are written once and never re-read by this kernel
    asm volatile("st.global.cs.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.a), "r"(v.b), "r"(v.c), "r"(v.d) : "memory");
}
#else
MRTS_HD void store16(void* p, Vec16 v) { *(Vec16*)p = v; }
#endif

// Phase C: one env's observation.  ELEM = 4 (int32) or 1 (uint8).  P = planes per cell.
template <int ELEM, int P>
MRTS_HD void store_obs(const uint32_t* words, int HW, void* out, int tid, int nth) {
    const bool vec = (((uintptr_t)out) & 15) == 0;  // odd map sizes: env rows are not 16-byte aligned -> scalar stores
    const int total = HW * P;                 // elements
    constexpr int EPC = 16 / ELEM;            // elements per 16-byte chunk
    const int nchunks = vec ? total / EPC : 0;
    uint8_t* base = (uint8_t*)out;
    for (int j = tid; j < nchunks; j += nth) {
        const int e0 = j * EPC;
        const int q = e0 / P;
        const int o = e0 - q * P;
        const uint64_t lo = words[q];
        const uint64_t hi = (q + 1 < HW) ? words[q + 1] : 0;
        const uint32_t bits = (uint32_t)(((hi << P) | lo) >> o);
        Vec16 v;
        if (ELEM == 4) {
            v.a = bits & 1u;
            v.b = (bits >> 1) & 1u;
            v.c = (bits >> 2) & 1u;
            v.d = (bits >> 3) & 1u;
        } else {
            v.a = spread4(bits & 0xF);
            v.b = spread4((bits >> 4) & 0xF);
            v.c = spread4((bits >> 8) & 0xF);
            v.d = spread4((bits >> 12) & 0xF);
        }
        store16(base + (size_t)j * 16, v);
    }
    // tail (total not a multiple of the chunk): scalar stores
    for (int e = nchunks * EPC + tid; e < total; e += nth) {
        const int q = e / P, o = e - q * P;
        const uint32_t bit = (words[q] >> o) & 1u;
        if (ELEM == 4) ((int32_t*)out)[e] = (int32_t)bit;
        else ((uint8_t*)out)[e] = (uint8_t)bit;
    }
}

// Phase C': zero one env's mask (>99% of it stays zero)
template <int ELEM>
MRTS_HD void zero_mask(int HW, void* out, int tid, int nth) {
    const int total = HW * 78;
    constexpr int EPC = 16 / ELEM;
    const int nchunks = ((((uintptr_t)out) & 15) == 0) ? total / EPC : 0;
    uint8_t* base = (uint8_t*)out;
    Vec16 z;
    z.a = z.b = z.c = z.d = 0;
    for (int j = tid; j < nchunks; j += nth) store16(base + (size_t)j * 16, z);
    for (int e = nchunks * EPC + tid; e < total; e += nth) {
        if (ELEM == 4) ((int32_t*)out)[e] = 0;
        else ((uint8_t*)out)[e] = 0;
    }
}

// Phase M (units): legal-action sets of every idle owned unit (true state), into umask[3*i..]
MRTS_HD void unit_masks(const Sim& s, uint32_t* umask, int tid, int nth) {
    const int n = s.hdr->nunits;
    for (int i = tid; i < n; i += nth) {
        const Unit& un = s.u[i];
        if (un.player < 0 || un.a_type >= 0 || (un.flags & UF_DEAD)) continue;
        Mask78 m = s.unit_mask(i);
        umask[3 * i] = m.w[0];
        umask[3 * i + 1] = m.w[1];
        umask[3 * i + 2] = m.w[2];
    }
}

// Phase D (after a barrier): the 78 entries of every source unit of `player`
template <int ELEM>
MRTS_HD void patch_mask(const Sim& s, int player, const uint32_t* umask, void* out, int tid, int nth) {
    const int n = s.hdr->nunits;
    // item = (unit, entry) flattened over 96 slots per unit so that a warp stays on one unit
    const int items = n * 96;
    for (int it = tid; it < items; it += nth) {
        const int i = it / 96, b = it - i * 96;
        if (b >= 78) continue;
        const Unit& un = s.u[i];
        if (un.player != player || un.a_type >= 0 || (un.flags & UF_DEAD)) continue;
        const uint32_t bit = (umask[3 * i + (b >> 5)] >> (b & 31)) & 1u;
        const size_t e = (size_t)(un.x + un.y * s.W) * 78 + b;
        if (ELEM == 4) ((int32_t*)out)[e] = (int32_t)bit;
        else ((uint8_t*)out)[e] = (uint8_t)bit;
    }
}

// ---- run-time choice of the element size / plane count (the pass itself is not templated on them:
// the store loops are a small part of its code, the rules are not)
MRTS_HD void store_obs_rt(int elem, int planes, const uint32_t* words, int HW, void* out, int tid, int nth) {
    if (elem == 4) {
        if (planes == 29) store_obs<4, 29>(words, HW, out, tid, nth);
        else store_obs<4, 31>(words, HW, out, tid, nth);
    } else {
        if (planes == 29) store_obs<1, 29>(words, HW, out, tid, nth);
        else store_obs<1, 31>(words, HW, out, tid, nth);
    }
}
MRTS_HD void zero_mask_rt(int elem, int HW, void* out, int tid, int nth) {
    if (elem == 4) zero_mask<4>(HW, out, tid, nth);
    else zero_mask<1>(HW, out, tid, nth);
}
MRTS_HD void patch_mask_rt(int elem, const Sim& s, int player, const uint32_t* umask, void* out, int tid, int nth) {
    if (elem == 4) patch_mask<4>(s, player, umask, out, tid, nth);
    else patch_mask<1>(s, player, umask, out, tid, nth);
}

#if defined(__CUDACC__)
constexpr int BULK_ZERO_BYTES = 8 * bulk::GROUP_BYTES;                           // 4,992 bytes of zeros per block
constexpr int BULK_TILES_BYTES = BULK_ZERO_BYTES;                                // per block, in front of the games' regions
#endif
#if defined(__CUDA_ARCH__)
// Phase D, bulk form (warp kernels in direct mode, device only): one view's mask through the copy engine
// (mrts_bulk.cuh) -- runs of zero rows straight from the block's zero tile, only the rows of units that may act staged.
// The warp walks the grid 32 cells at a time so that the rows arrive in ascending cell order.
__device__ __forceinline__ void bulk_mask_view(const Sim& s, int player, const uint32_t* umask, void* out, bulk::MaskWriter& mw, int lane) {
    mw.begin(out);
#pragma unroll 1
    for (int base = 0; base < s.HW; base += 32) {
        const int c = base + lane;
        bool take = false;
        int i = 0;
        if (c < s.HW) {
            const uint16_t g = s.grid[c];
            if (g < G_WALL) {
                const Unit& un = s.u[g];
                i = g;
                take = un.player == player && un.a_type < 0 && !(un.flags & UF_DEAD);
            }
        }
        unsigned m = __ballot_sync(0xFFFFFFFFu, take);
#pragma unroll 1
        while (m) {
            const int src = __ffs((int)m) - 1;
            m &= m - 1;
            const int k = __shfl_sync(0xFFFFFFFFu, i, src);
            mw.add((uint32_t)(base + src), umask[3 * k], umask[3 * k + 1], umask[3 * k + 2]);
        }
    }
    mw.finish(s.HW / mw.cells_per_group());
}
#endif

// Phase E (after unit_masks + barrier): the packed form of one view's mask -- one 16-byte entry
// (cell, 78 bits) per unit of `player` that may act, in ASCENDING CELL ORDER (the expansion kernel builds
// the mask rows of neighbouring cells together), behind a header word holding their count.  The group's
// first warp walks the grid 32 cells at a time; a ballot keeps the order.
MRTS_HD void emit_mask_entries(const Sim& s, int player, const uint32_t* umask, uint32_t* out, int tid, int nth) {
    if (tid >= 32) return;
    const int lanes = nth < 32 ? nth : 32;
    int total = 0;
#pragma unroll 1
    for (int base = 0; base < s.HW; base += lanes) {
        const int c = base + tid;
        bool take = false;
        int i = 0;
        if (c < s.HW) {
            const uint16_t g = s.grid[c];
            if (g < G_WALL) {
                const Unit& un = s.u[g];
                i = g;
                take = un.player == player && un.a_type < 0 && !(un.flags & UF_DEAD);
            }
        }
        const unsigned m = Sim::wballot(take);
        if (take) {
            uint32_t* e = out + 4 * (size_t)(1 + total + Sim::popc(m & ((1u << tid) - 1u)));
            e[0] = (uint32_t)c;
            e[1] = umask[3 * i];
            e[2] = umask[3 * i + 1];
            e[3] = umask[3 * i + 2];
        }
        total += Sim::popc(m);
    }
    if (tid == 0) {
        out[0] = (uint32_t)total;
        out[1] = out[2] = out[3] = 0;
    }
}
// host-side / reference expansion of the packed mask entries into (HW, 78) elements (what mrts_expand_kernel
// does with bulk stores on the device): zero fill + the 78 entries of every listed cell
template <int ELEM>
MRTS_HD void expand_mask_entries(const uint32_t* ent, int HW, void* out) {
    zero_mask<ELEM>(HW, out, 0, 1);
    const int n = (int)ent[0];
    for (int k = 0; k < n; k++) {
        const uint32_t* e = ent + 4 * (size_t)(1 + k);
        for (int b = 0; b < 78; b++) {
            const uint32_t bit = (e[1 + (b >> 5)] >> (b & 31)) & 1u;
            const size_t idx = (size_t)e[0] * 78 + b;
            if (ELEM == 4) ((int32_t*)out)[idx] = (int32_t)bit;
            else ((uint8_t*)out)[idx] = (uint8_t)bit;
        }
    }
}

// ---- masked-uniform action sampler (the recipe of the reference's hello_world.py:27-67: an
// independent categorical per action component restricted to mask == 1; an all-zero group
// yields 0), driven by a counter-based hash so a run is reproducible from (seed, launch counter).
MRTS_HD uint64_t mix64(uint64_t z) {  // splitmix64 finaliser
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
MRTS_HD int popc64(uint64_t v) {
#if defined(__CUDA_ARCH__)
    return __popcll(v);
#else
    return __builtin_popcountll(v);
#endif
}
// index of the (r mod popcount)-th set bit of `bits`, 0 if none is set
MRTS_HD int pick_bit(uint64_t bits, uint32_t r) {
    const int n = popc64(bits);
    if (n == 0) return 0;
    int k = (int)(r % (uint32_t)n);
#if defined(__CUDA_ARCH__)
    const uint32_t lo = (uint32_t)bits, hi = (uint32_t)(bits >> 32);
    const int nlo = __popc(lo);
    return k < nlo ? (int)__fns(lo, 0, k + 1) : 32 + (int)__fns(hi, 0, k - nlo + 1);
#else
    for (int b = 0; b < 64; b++)
        if ((bits >> b) & 1) {
            if (k == 0) return b;
            k--;
        }
    return 0;
#endif
}
MRTS_HD void sample_unit_action(const uint32_t* um, uint64_t key, int32_t* out7) {
    const uint64_t lo = (uint64_t)um[0] | ((uint64_t)um[1] << 32);
    const uint64_t att = (lo >> 29) | ((uint64_t)um[2] << 35);  // bits 29..77
    const uint64_t r0 = mix64(key), r1 = mix64(key ^ 0xD1B54A32D192ED03ULL);
    out7[0] = pick_bit(lo & 0x3F, (uint32_t)r0 & 0xFFFF);
    out7[1] = pick_bit((lo >> 6) & 0xF, (uint32_t)(r0 >> 16) & 0xFFFF);
    out7[2] = pick_bit((lo >> 10) & 0xF, (uint32_t)(r0 >> 32) & 0xFFFF);
    out7[3] = pick_bit((lo >> 14) & 0xF, (uint32_t)(r0 >> 48) & 0xFFFF);
    out7[4] = pick_bit((lo >> 18) & 0xF, (uint32_t)r1 & 0xFFFF);
    out7[5] = pick_bit((lo >> 22) & 0x7F, (uint32_t)(r1 >> 16) & 0xFFFF);
    out7[6] = pick_bit(att & 0x1FFFFFFFFFFFFULL, (uint32_t)(r1 >> 32));
}
// Phase S (units, after unit_masks + barrier): write the sampled action of every source unit of
// view `player` at its cell of that view's (HW,7) block; other cells are left untouched (the
// step only ever reads the cells of source units, PlayerAction.java:407).
MRTS_HD void sample_actions(const Sim& s, int player, const uint32_t* umask, int32_t* out, uint64_t seed, uint32_t counter, uint32_t env,
                            int tid, int nth) {
    const int n = s.hdr->nunits;
    for (int i = tid; i < n; i += nth) {
        const Unit& un = s.u[i];
        if (un.player != player || un.a_type >= 0 || (un.flags & UF_DEAD)) continue;
        const int c = un.x + un.y * s.W;
        const uint64_t key = mix64(seed ^ ((uint64_t)counter << 32 | env)) ^ (uint64_t)c * 0x9E3779B97F4A7C15ULL;
        int32_t a[7];
        sample_unit_action(umask + 3 * i, key, a);
        int32_t* o = out + (size_t)c * 7;
        for (int k = 0; k < 7; k++) o[k] = a[k];
    }
}

}  // namespace mrts
