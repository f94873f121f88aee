// Bulk asynchronous stores (cp.async.bulk.global.shared::cta, UBLKCP in SASS) and the warp-level writers of
// one view's observation and action mask built on them ("run-length bulk stores").  Device code only (sm_100a).
//
// Both tensors are almost entirely made of two constant patterns -- the one-hot code of an EMPTY cell and the
// all-zero mask row -- so a warp hands RUNS of them to the copy engine straight from static tiles in shared
// memory and expands only the few cells that hold something.  Per 16x16 view that is ~40 copy commands and ~35
// staged groups instead of ~6,800 128-bit stores with the expansion arithmetic in front of each.
//
// Alignment decides the unit of work.  A bulk copy moves multiples of 16 bytes between 16-byte aligned addresses:
//   mask row = 78 elements  -> the smallest aligned run of whole rows is 624 bytes  (2 cells of int32, 8 of uint8)
//   obs cell = P elements   -> 16 x P bytes = 464 / 496                            (4 cells of int32, 16 of uint8)
// such a run is a "group".  Walking a view's groups in ascending order the writers alternate
//   constant run | staged group | constant run | ...
// so that every byte of the view is written exactly once.  The copies complete asynchronously; the issuing lane
// only waits (wait_group.read) before a staging slot is reused and before the shared memory behind it goes away.
// Element size and plane count are run-time members: kernels that know them at compile time (mrts_expand.cu)
// get them folded, the rules kernels carry one copy of this code for all four combinations.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mrts {
namespace bulk {

constexpr int GROUP_BYTES = 624;      // one mask group
constexpr int EMPTY_GROUPS = 8;       // observation groups in the empty-cell tile
constexpr int EMPTY_BYTES_MAX = EMPTY_GROUPS * 16 * 31;  // 3,968
constexpr unsigned FULL = 0xFFFFFFFFu;

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
// shared -> global bulk copy of `bytes` (multiple of 16, both sides 16-byte aligned), tracked by the issuing thread's bulk group
__device__ __forceinline__ void store(void* gdst, const void* ssrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_addr(ssrc)), "r"(bytes) : "memory");
}
// L2 residency.  A step streams gigabytes of write-once output through the L2 while the few megabytes that ARE
// re-read every step (game blobs, packed words) would be evicted in between: outputs are stored evict-first,
// state is loaded / stored evict-last.
__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t policy_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ void store_hint(void* gdst, const void* ssrc, uint32_t bytes, uint64_t pol) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(gdst), "r"(smem_addr(ssrc)), "r"(bytes), "l"(pol)
                 : "memory");
}
__device__ __forceinline__ uint4 ldg_hint(const void* p, uint64_t pol) {
    uint4 v;
    asm volatile("ld.global.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ void stg_hint(void* p, uint4 v, uint64_t pol) {
    asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(pol) : "memory");
}
__device__ __forceinline__ void commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void wait_read() {  // all but the N most recent groups of this thread have read their source
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// generic-proxy writes to shared memory -> visible to the async proxy (the copy engine)
__device__ __forceinline__ void fence_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// a run of `bytes` copies of a constant pattern: pieces of at most `tile_bytes` from the static tile
__device__ __forceinline__ void store_run(uint8_t* gdst, const uint8_t* tile, int tile_bytes, long bytes, uint64_t pol) {
#pragma unroll 1
    while (bytes > 0) {
        const uint32_t n = bytes < tile_bytes ? (uint32_t)bytes : (uint32_t)tile_bytes;
        store_hint(gdst, tile, n, pol);
        gdst += n;
        bytes -= n;
    }
}

__device__ __forceinline__ uint32_t spread_nibble(uint32_t nibble) { return (nibble * 0x00204081u) & 0x01010101u; }  // bit k -> byte k
// element index -> cell inside a group (e / planes for e < 1,024; planes = 29 or 31)
__device__ __forceinline__ int div_planes(int e, int planes) { return (int)(((uint32_t)e * ((1u << 20) / (uint32_t)planes + 1u)) >> 20); }
// 16 bytes of one-hot elements: chunk j of a group whose cells have the words w[0 .. 16 / elem) (or all the word `same`)
__device__ __forceinline__ uint4 expand_chunk(const uint32_t* w, int j, int elem, int planes, bool all_same, uint32_t same) {
    const int epc = 16 / elem, e0 = j * epc, q = div_planes(e0, planes), o = e0 - q * planes;
    const uint32_t lo = all_same ? same : w[q];
    const uint32_t hi = all_same ? same : (q + 1 < epc ? w[q + 1] : 0u);
    const uint32_t bits = (uint32_t)(((((uint64_t)hi) << planes) | lo) >> o);
    uint4 v;
    if (elem == 4) {
        v.x = bits & 1u;
        v.y = (bits >> 1) & 1u;
        v.z = (bits >> 2) & 1u;
        v.w = (bits >> 3) & 1u;
    } else {
        v.x = spread_nibble(bits & 0xF);
        v.y = spread_nibble((bits >> 4) & 0xF);
        v.z = spread_nibble((bits >> 8) & 0xF);
        v.w = spread_nibble((bits >> 12) & 0xF);
    }
    return v;
}

// a block's tile of zeros alone (every thread of the block, once; ends with a block barrier)
__device__ __forceinline__ void zero_tile_init(uint8_t* tile, int zero_bytes, int tid, int nth) {
#pragma unroll 1
    for (int i = tid; i < zero_bytes / 16; i += nth) ((uint4*)tile)[i] = make_uint4(0, 0, 0, 0);
    fence_smem();
    __syncthreads();
}
// The block's static tiles: `zero_bytes` of zeros at `tiles`, then EMPTY_GROUPS observation groups of the empty cell.
// Every thread of the block calls this once, before any writer runs (it ends with a block barrier).
__device__ __forceinline__ void tiles_init(uint8_t* tiles, int zero_bytes, int elem, int planes, uint32_t empty_word, int tid, int nth) {
#pragma unroll 1
    for (int i = tid; i < zero_bytes / 16; i += nth) ((uint4*)tiles)[i] = make_uint4(0, 0, 0, 0);
    uint4* e = (uint4*)(tiles + zero_bytes);
#pragma unroll 1
    for (int i = tid; i < EMPTY_GROUPS * planes; i += nth) e[i] = expand_chunk(nullptr, i % planes, elem, planes, true, empty_word);
    fence_smem();
    __syncthreads();
}

// ---- staging.  Groups that hold something are expanded into one of TWO staging buffers and leave with one bulk
// copy per RUN of adjacent groups (units cluster: a base, its workers and the resources next to them share a few
// groups; under partial observability whole regions differ from the empty pattern).  Both writers of a warp share
// the buffers and LANE 0 issues every copy, so one bulk-group sequence orders all reuse: before a buffer is
// written again the copy that read it two flushes ago must be done (wait_group.read 1).
constexpr int STAGE_BUF_BYTES = 1248;  // 2 mask groups (2 x 624) or 2 observation groups (2 x 464 / 496): shared memory is what bounds the resident games
constexpr int STAGE_BYTES = 2 * STAGE_BUF_BYTES;
struct Stage {
    uint8_t* buf0;
    int cur;  // buffer being filled
    int lane;
    uint64_t pol;  // L2 policy of the output stores (evict first)
    __device__ __forceinline__ void init(uint8_t* region, int lane_) {
        buf0 = region;
        cur = 0;
        lane = lane_;
        pol = policy_evict_first();
    }
    __device__ __forceinline__ uint8_t* buf() const { return buf0 + cur * STAGE_BUF_BYTES; }
    // the buffer about to be filled is free again (lane 0 waits, the warp learns of it)
    __device__ __forceinline__ void acquire() {
        if (lane == 0) wait_read<1>();
        __syncwarp();
    }
    // `bytes` staged in the current buffer -> global; the other buffer is filled next
    __device__ __forceinline__ void flush(uint8_t* gdst, int bytes) {
        fence_smem();
        __syncwarp();
        if (lane == 0) {
            store_hint(gdst, buf(), (uint32_t)bytes, pol);
            commit();
        }
        cur ^= 1;
    }
    // before the shared memory behind the buffers / the block's static tiles is reused or released
    __device__ __forceinline__ void drain() {
        if (lane == 0) {
            commit();  // constant runs issued after the last flush
            wait_read<0>();
        }
        __syncwarp();
    }
};

// One warp writes one view's observation from its words (bit k of a cell's word = plane k).  All 32 lanes call every
// method with the same arguments.
struct ObsWriter {
    Stage* st;
    const uint8_t* etile;
    uint32_t empty_word;
    int elem, planes;
    uint8_t* out;
    int run;      // the first group not written yet
    int gs, nf;   // staged run: first group, groups

    __device__ __forceinline__ int cells_per_group() const { return 16 / elem; }
    __device__ __forceinline__ int group_bytes() const { return 16 * planes; }
    __device__ __forceinline__ void init(Stage* st_, const uint8_t* etile_, uint32_t empty_word_, int elem_, int planes_) {
        st = st_;
        etile = etile_;
        empty_word = empty_word_;
        elem = elem_;
        planes = planes_;
    }
    __device__ __forceinline__ void begin(void* out_view) {
        out = (uint8_t*)out_view;
        run = 0;
        nf = 0;
        gs = 0;
    }
    __device__ __forceinline__ void flush() {
        st->flush(out + (size_t)gs * group_bytes(), nf * group_bytes());
        nf = 0;
    }
    // words of the groups [g0, g0 + n): n * cells_per_group() words the warp can read (shared memory)
    __device__ __forceinline__ void segment(const uint32_t* w, int g0, int n) {
        const int og = cells_per_group(), gb = group_bytes();
        const int cap = STAGE_BUF_BYTES / gb;
#pragma unroll 1
        for (int base = 0; base < n; base += 32) {
            const int gl = base + st->lane;
            bool dirty = false;
            if (gl < n) {
#pragma unroll 4
                for (int k = 0; k < og; k++) dirty = dirty || w[gl * og + k] != empty_word;
            }
            unsigned m = __ballot_sync(FULL, dirty);
#pragma unroll 1
            while (m) {
                const int src = __ffs((int)m) - 1;
                m &= m - 1;
                const int g = g0 + base + src;
                if (nf > 0 && (g != gs + nf || nf == cap)) flush();
                if (nf == 0) {  // a new run: the empty groups in front of it leave from the tile
                    if (st->lane == 0) store_run(out + (size_t)run * gb, etile, EMPTY_GROUPS * gb, (long)(g - run) * gb, st->pol);
                    st->acquire();
                    gs = g;
                }
                if (st->lane < planes) ((uint4*)(st->buf() + nf * gb))[st->lane] = expand_chunk(w + (base + src) * og, st->lane, elem, planes, false, 0u);  // 16 x planes bytes = planes chunks
                nf++;
                run = g + 1;
            }
        }
    }
    __device__ __forceinline__ void finish(int ngroups) {
        if (nf > 0) flush();
        const int gb = group_bytes();
        if (st->lane == 0) store_run(out + (size_t)run * gb, etile, EMPTY_GROUPS * gb, (long)(ngroups - run) * gb, st->pol);
    }
};

// One warp writes one view's mask.  All 32 lanes call every method with the same arguments.
struct MaskWriter {
    Stage* st;
    const uint8_t* ztile;
    int zbytes;  // size of the zero tile (a multiple of GROUP_BYTES)
    int elem;
    uint8_t* out;
    int run;          // the first group not written yet
    int gs, nf;       // staged run: first group, groups (the last one is the open group)

    __device__ __forceinline__ int cells_per_group() const { return 8 / elem; }
    __device__ __forceinline__ void init(Stage* st_, const uint8_t* ztile_, int zbytes_, int elem_) {
        st = st_;
        ztile = ztile_;
        zbytes = zbytes_;
        elem = elem_;
    }
    __device__ __forceinline__ void begin(void* out_view) {
        out = (uint8_t*)out_view;
        run = 0;
        nf = 0;
        gs = 0;
    }
    __device__ __forceinline__ void flush() {
        st->flush(out + (size_t)gs * GROUP_BYTES, nf * GROUP_BYTES);
        nf = 0;
    }
    // the 78 mask bits of the unit at `cell` (cells arrive in ascending order)
    __device__ __forceinline__ void add(uint32_t cell, uint32_t w0, uint32_t w1, uint32_t w2) {
        const int gc = cells_per_group();
        const int g = (int)cell / gc;
        if (nf == 0 || g != gs + nf - 1) {  // not the open group: a new one
            if (nf > 0 && (g != gs + nf || nf == STAGE_BUF_BYTES / GROUP_BYTES)) flush();
            if (nf == 0) {
                if (st->lane == 0) store_run(out + (size_t)run * GROUP_BYTES, ztile, zbytes, (long)(g - run) * GROUP_BYTES, st->pol);
                st->acquire();
                gs = g;
            }
            uint4* z = (uint4*)(st->buf() + nf * GROUP_BYTES);
            for (int j = st->lane; j < GROUP_BYTES / 16; j += 32) z[j] = make_uint4(0, 0, 0, 0);
            __syncwarp();
            nf++;
            run = g + 1;
        }
        uint8_t* const slot = st->buf() + (nf - 1) * GROUP_BYTES;
        const int r0 = ((int)cell - g * gc) * 78;
#pragma unroll 1
        for (int b = st->lane; b < 78; b += 32) {
            const uint32_t w = b < 32 ? w0 : (b < 64 ? w1 : w2);
            const uint32_t bit = (w >> (b & 31)) & 1u;
            if (elem == 4) ((uint32_t*)slot)[r0 + b] = bit;
            else slot[r0 + b] = (uint8_t)bit;
        }
    }
    __device__ __forceinline__ void finish(int ngroups) {
        if (nf > 0) flush();
        if (st->lane == 0) store_run(out + (size_t)run * GROUP_BYTES, ztile, zbytes, (long)(ngroups - run) * GROUP_BYTES, st->pol);
    }
};

}  // namespace bulk
}  // namespace mrts
