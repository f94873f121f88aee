// See mrts_host.h.  Plain C++17 + AVX2 where the CPU has it (checked at run time).
#include "mrts_host.h"

#include <sched.h>
#include <stdlib.h>
#include <string.h>

#include <time.h>

#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#if defined(__x86_64__)
#include <immintrin.h>
#endif

namespace mrts_host {
constexpr int ENTRY_WORDS = 6;  // mrts_launch.h: DELTA_ENTRY_WORDS
#if defined(__x86_64__)
static void mask_row_avx2_32(const uint32_t* pm, int32_t* o);  // defined with the incremental forms below
#endif

// ------------------------------------------------------------------------------------------------
// Workers SPIN for a short while after a job (MRTS_HOST_SPIN_US, default 300) before they block on the condition
// variable: a host-buffer step runs two short jobs ~0.2 ms apart (pack the actions, patch the results) and a
// futex wake-up of every worker cost as much as the jobs themselves.  The pool never has more threads than
// the cores this rank may use (mrts_api.cu: visible cores / LOCAL_WORLD_SIZE), so the spinning starves nobody.
struct Pool {
    std::vector<std::thread> workers;
    std::mutex mu;
    std::condition_variable cv_work;
    std::function<void(size_t, size_t)> fn;
    size_t n = 0;
    std::atomic<int> generation{0}, pending{0}, sleepers{0};
    std::atomic<bool> stop{false};
    int nthreads = 1;
    long spin_ns = 300000;
};

static inline void cpu_relax() {
#if defined(__x86_64__)
    _mm_pause();
#endif
}
static inline long mono_ns() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1000000000L + ts.tv_nsec;
}

static void worker_main(Pool* p, int rank) {
    int seen = 0;
    for (;;) {
        long t0 = mono_ns();
        int spins = 0;
        while (p->generation.load() == seen && !p->stop.load()) {
            cpu_relax();
            if ((++spins & 255) == 0 && mono_ns() - t0 > p->spin_ns) {
                std::unique_lock<std::mutex> lk(p->mu);
                p->sleepers.fetch_add(1);
                p->cv_work.wait(lk, [&] { return p->stop.load() || p->generation.load() != seen; });
                p->sleepers.fetch_sub(1);
                break;
            }
        }
        if (p->stop.load()) return;
        seen = p->generation.load();
        const size_t n = p->n;
        const size_t per = (n + p->nthreads - 1) / p->nthreads;
        const size_t b = per * rank, e = b + per < n ? b + per : n;
        if (b < e) p->fn(b, e);
        p->pending.fetch_sub(1);
    }
}

int visible_cores() {
    cpu_set_t set;
    CPU_ZERO(&set);
    int n = sched_getaffinity(0, sizeof(set), &set) == 0 ? CPU_COUNT(&set) : (int)std::thread::hardware_concurrency();
    return n > 0 ? n : 1;
}

Pool* pool_create(int threads) {
    if (threads <= 0) threads = visible_cores();
    Pool* p = new Pool();
    p->nthreads = threads;
    if (const char* v = getenv("MRTS_HOST_SPIN_US")) p->spin_ns = atol(v) * 1000L;
    for (int r = 1; r < threads; r++) p->workers.emplace_back(worker_main, p, r);  // rank 0 = the caller
    return p;
}
void pool_destroy(Pool* p) {
    if (!p) return;
    {
        std::lock_guard<std::mutex> lk(p->mu);
        p->stop.store(true);
    }
    p->cv_work.notify_all();
    for (auto& t : p->workers) t.join();
    delete p;
}
int pool_threads(const Pool* p) { return p ? p->nthreads : 1; }

static void parallel_for(Pool* p, size_t n, const std::function<void(size_t, size_t)>& fn) {
    if (!p || p->nthreads == 1 || n < 2) {
        fn(0, n);
        return;
    }
    p->fn = fn;
    p->n = n;
    p->pending.store((int)p->workers.size());
    p->generation.fetch_add(1);  // publishes fn / n (sequentially consistent, as is the sleepers test below)
    if (p->sleepers.load() > 0) {
        std::lock_guard<std::mutex> lk(p->mu);
        p->cv_work.notify_all();
    }
    const size_t per = (n + p->nthreads - 1) / p->nthreads;
    if (per > 0) fn(0, per < n ? per : n);
    int spins = 0;
    while (p->pending.load() != 0) {
        cpu_relax();
        if ((++spins & 4095) == 0) sched_yield();
    }
}

// ------------------------------------------------------------------------------------------------
static void obs_scalar32(const uint32_t* w, size_t cells, int planes, int32_t* out) {
    for (size_t c = 0; c < cells; c++) {
        const uint32_t v = w[c];
        int32_t* o = out + c * planes;
        for (int k = 0; k < planes; k++) o[k] = (int32_t)((v >> k) & 1u);
    }
}
static void obs_scalar8(const uint32_t* w, size_t cells, int planes, uint8_t* out) {
    for (size_t c = 0; c < cells; c++) {
        const uint32_t v = w[c];
        uint8_t* o = out + c * planes;
        for (int k = 0; k < planes; k++) o[k] = (uint8_t)((v >> k) & 1u);
    }
}

#if defined(__x86_64__)
// 8 planes per 256-bit store; a cell's last store spills into the next cell, which then overwrites it
// (cells are written in increasing order); the final cell of the range is finished with scalar code.
__attribute__((target("avx2"))) static void obs_avx2_32(const uint32_t* w, size_t cells, int planes, int32_t* out) {
    if (cells == 0) return;
    const __m256i one = _mm256_set1_epi32(1);
    const __m256i s0 = _mm256_setr_epi32(0, 1, 2, 3, 4, 5, 6, 7);
    const __m256i s1 = _mm256_setr_epi32(8, 9, 10, 11, 12, 13, 14, 15);
    const __m256i s2 = _mm256_setr_epi32(16, 17, 18, 19, 20, 21, 22, 23);
    const __m256i s3 = _mm256_setr_epi32(24, 25, 26, 27, 28, 29, 30, 31);
    for (size_t c = 0; c + 1 < cells; c++) {
        const __m256i v = _mm256_set1_epi32((int)w[c]);
        int32_t* o = out + c * planes;
        _mm256_storeu_si256((__m256i*)(o), _mm256_and_si256(_mm256_srlv_epi32(v, s0), one));
        _mm256_storeu_si256((__m256i*)(o + 8), _mm256_and_si256(_mm256_srlv_epi32(v, s1), one));
        _mm256_storeu_si256((__m256i*)(o + 16), _mm256_and_si256(_mm256_srlv_epi32(v, s2), one));
        _mm256_storeu_si256((__m256i*)(o + 24), _mm256_and_si256(_mm256_srlv_epi32(v, s3), one));  // planes >= 24 + 5; spill <= 3 ints
    }
    obs_scalar32(w + cells - 1, 1, planes, out + (cells - 1) * planes);
}
// Aligned output (the usual case: NumPy's large arrays are page aligned and an env's block is a
// multiple of 32 bytes): the env's elements are produced 8 at a time from the continuous bit stream
// (as the device kernels do) and written with non-temporal stores -- the arrays are far larger
// than the caches and are only read later, by the consumer.
__attribute__((target("avx2"))) static void obs_avx2_stream32(const uint32_t* w, size_t cells, int planes, int32_t* out) {
    const __m256i one = _mm256_set1_epi32(1);
    const __m256i s0 = _mm256_setr_epi32(0, 1, 2, 3, 4, 5, 6, 7);
    const size_t total = cells * (size_t)planes, n8 = total / 8;
    size_t q = 0;
    int o = 0;
    for (size_t j = 0; j < n8; j++) {
        const uint64_t keep = (1ull << planes) - 1;
        const uint64_t lo = w[q] & keep, hi = q + 1 < cells ? (w[q + 1] & keep) : 0;
        const uint32_t bits = (uint32_t)(((hi << planes) | lo) >> o);
        _mm256_stream_si256((__m256i*)(out + 8 * j), _mm256_and_si256(_mm256_srlv_epi32(_mm256_set1_epi32((int)bits), s0), one));
        o += 8;
        if (o >= planes) {
            o -= planes;
            q++;
        }
    }
    for (size_t e = n8 * 8; e < total; e++) out[e] = (int32_t)((w[e / planes] >> (e % planes)) & 1u);
    _mm_sfence();
}
__attribute__((target("avx2"))) static void zero_stream(void* p, size_t bytes) {  // p 32-byte aligned
    const __m256i z = _mm256_setzero_si256();
    uint8_t* b = (uint8_t*)p;
    size_t k = 0;
    for (; k + 128 <= bytes; k += 128) {
        _mm256_stream_si256((__m256i*)(b + k), z);
        _mm256_stream_si256((__m256i*)(b + k + 32), z);
        _mm256_stream_si256((__m256i*)(b + k + 64), z);
        _mm256_stream_si256((__m256i*)(b + k + 96), z);
    }
    for (; k + 32 <= bytes; k += 32) _mm256_stream_si256((__m256i*)(b + k), z);
    if (k < bytes) memset(b + k, 0, bytes - k);
    _mm_sfence();
}
static bool has_avx2() {
    static const bool v = __builtin_cpu_supports("avx2");
    return v;
}
// Non-temporal stores are OFF unless MRTS_HOST_NT=1: on the B200 hosts measured (16-core Xeon VM) they
// were slower than regular stores (561k vs 636k env-steps/s end to end at 16x16, 8,192 envs).
static bool use_stream() {
    static const bool v = [] {
        const char* e = getenv("MRTS_HOST_NT");
        return e && e[0] == '1';
    }();
    return v;
}
#endif

void expand_obs(Pool* p, const uint32_t* words, size_t n_env, int hw, int planes, int elem, void* out) {
    parallel_for(p, n_env, [=](size_t b, size_t e) {
        const size_t cells = (e - b) * (size_t)hw;
        const uint32_t* w = words + b * hw;
        if (elem == 4) {
            int32_t* o = (int32_t*)out + b * (size_t)hw * planes;
#if defined(__x86_64__)
            if (has_avx2() && use_stream() && (((uintptr_t)o) & 31) == 0 && planes <= 31) {
                obs_avx2_stream32(w, cells, planes, o);
                return;
            }
            if (has_avx2() && planes >= 29) {
                obs_avx2_32(w, cells, planes, o);
                return;
            }
#endif
            obs_scalar32(w, cells, planes, o);
        } else {
            obs_scalar8(w, cells, planes, (uint8_t*)out + b * (size_t)hw * planes);
        }
    });
}

void expand_mask(Pool* p, const uint32_t* pmask, size_t n_env, int hw, int elem, void* out) {
    parallel_for(p, n_env, [=](size_t b, size_t e) {
        for (size_t env = b; env < e; env++) {
            const uint32_t* pm = pmask + env * (size_t)hw * 3;
            uint8_t* o = (uint8_t*)out + env * (size_t)hw * 78 * elem;
#if defined(__x86_64__)
            if (has_avx2() && use_stream() && (((uintptr_t)o) & 31) == 0) zero_stream(o, (size_t)hw * 78 * elem);
            else
#endif
                memset(o, 0, (size_t)hw * 78 * elem);
            for (int c = 0; c < hw; c++) {
                const uint32_t m0 = pm[3 * c], m1 = pm[3 * c + 1], m2 = pm[3 * c + 2];
                if ((m0 | m1 | m2) == 0) continue;
#if defined(__x86_64__)
                if (elem == 4 && has_avx2()) {
                    mask_row_avx2_32(pm + 3 * c, (int32_t*)o + (size_t)c * 78);
                    continue;
                }
#endif
                for (int k = 0; k < 78; k++) {
                    const uint32_t word = k < 32 ? m0 : (k < 64 ? m1 : m2);
                    const uint32_t bit = (word >> (k & 31)) & 1u;
                    if (elem == 4) ((int32_t*)o)[(size_t)c * 78 + k] = (int32_t)bit;
                    else o[(size_t)c * 78 + k] = (uint8_t)bit;
                }
            }
        }
    });
}

// ------------------------------------------------------------------------------------------------ incremental forms
static inline void obs_cell(uint32_t v, int planes, int elem, void* out) {
    if (elem == 4) {
        int32_t* o = (int32_t*)out;
        for (int k = 0; k < planes; k++) o[k] = (int32_t)((v >> k) & 1u);
    } else {
        uint8_t* o = (uint8_t*)out;
        for (int k = 0; k < planes; k++) o[k] = (uint8_t)((v >> k) & 1u);
    }
}
static inline void mask_cell(const uint32_t* pm, int elem, void* out) {
    if ((pm[0] | pm[1] | pm[2]) == 0) {
        memset(out, 0, (size_t)78 * elem);
        return;
    }
    for (int k = 0; k < 78; k++) {
        const uint32_t bit = (pm[k >> 5] >> (k & 31)) & 1u;
        if (elem == 4) ((int32_t*)out)[k] = (int32_t)bit;
        else ((uint8_t*)out)[k] = (uint8_t)bit;
    }
}
#if defined(__x86_64__)
// one cell, int32, planes in 24..32: three aligned-to-cell stores and one overlapping the tail (no spill into the next cell)
__attribute__((target("avx2"))) static inline void obs_cell_avx2_32(uint32_t w, int planes, int32_t* o) {
    const __m256i one = _mm256_set1_epi32(1), v = _mm256_set1_epi32((int)w);
    const __m256i s0 = _mm256_setr_epi32(0, 1, 2, 3, 4, 5, 6, 7);
    _mm256_storeu_si256((__m256i*)(o), _mm256_and_si256(_mm256_srlv_epi32(v, s0), one));
    _mm256_storeu_si256((__m256i*)(o + 8), _mm256_and_si256(_mm256_srlv_epi32(v, _mm256_add_epi32(s0, _mm256_set1_epi32(8))), one));
    _mm256_storeu_si256((__m256i*)(o + 16), _mm256_and_si256(_mm256_srlv_epi32(v, _mm256_add_epi32(s0, _mm256_set1_epi32(16))), one));
    _mm256_storeu_si256((__m256i*)(o + planes - 8), _mm256_and_si256(_mm256_srlv_epi32(v, _mm256_add_epi32(s0, _mm256_set1_epi32(planes - 8))), one));
}
// one mask row, int32: nine aligned-to-row stores of 8 elements and one overlapping the tail (elements 70..77)
__attribute__((target("avx2"))) static inline void mask_cell_avx2_32(const uint32_t* pm, int32_t* o) {
    const __m256i one = _mm256_set1_epi32(1), s0 = _mm256_setr_epi32(0, 1, 2, 3, 4, 5, 6, 7);
    for (int k = 0; k < 9; k++) {
        const uint32_t chunk = (pm[k >> 2] >> ((k & 3) * 8)) & 0xFFu;
        _mm256_storeu_si256((__m256i*)(o + 8 * k), _mm256_and_si256(_mm256_srlv_epi32(_mm256_set1_epi32((int)chunk), s0), one));
    }
    _mm256_storeu_si256((__m256i*)(o + 70), _mm256_and_si256(_mm256_srlv_epi32(_mm256_set1_epi32((int)(pm[2] >> 6)), s0), one));
}
// elements [k0, k1] of a row of `n` int32 one-hot values whose bits are bits[0..2] (n <= 96): whole 8-element chunks, the last
// one pulled back so that it ends with the row (values outside the changed range are rewritten with what they already hold)
__attribute__((target("avx2"))) static inline void row_range_avx2_32(const uint32_t* bits, uint32_t k0, uint32_t k1, uint32_t n, int32_t* o) {
    const __m256i one = _mm256_set1_epi32(1), s0 = _mm256_setr_epi32(0, 1, 2, 3, 4, 5, 6, 7);
    for (uint32_t c = k0 & ~7u; c <= k1; c += 8) {
        const uint32_t e0 = c + 8 <= n ? c : n - 8;  // first element of this chunk
        const uint64_t two = (uint64_t)bits[e0 >> 5] | ((e0 >> 5) + 1 < 3 ? (uint64_t)bits[(e0 >> 5) + 1] << 32 : 0ull);
        const uint32_t chunk = (uint32_t)(two >> (e0 & 31)) & 0xFFu;
        _mm256_storeu_si256((__m256i*)(o + e0), _mm256_and_si256(_mm256_srlv_epi32(_mm256_set1_epi32((int)chunk), s0), one));
    }
}
__attribute__((target("avx2"))) static void obs_delta_avx2(const uint32_t* w, uint32_t* sh, size_t cells, int planes, int elem, uint8_t* out) {
    const size_t cell_bytes = (size_t)planes * elem;
    const bool wide = elem == 4 && planes >= 24 && planes <= 32;
    size_t c = 0;
    for (; c + 8 <= cells; c += 8) {
        const __m256i a = _mm256_loadu_si256((const __m256i*)(w + c)), b = _mm256_loadu_si256((const __m256i*)(sh + c));
        unsigned neq = ~(unsigned)_mm256_movemask_ps(_mm256_castsi256_ps(_mm256_cmpeq_epi32(a, b))) & 0xFFu;
        if (!neq) continue;
        _mm256_storeu_si256((__m256i*)(sh + c), a);
        while (neq) {
            const int k = __builtin_ctz(neq);
            neq &= neq - 1;
            if (wide) obs_cell_avx2_32(w[c + k], planes, (int32_t*)(out + (c + k) * cell_bytes));
            else obs_cell(w[c + k], planes, elem, out + (c + k) * cell_bytes);
        }
    }
    for (; c < cells; c++)
        if (w[c] != sh[c]) {
            sh[c] = w[c];
            obs_cell(w[c], planes, elem, out + c * cell_bytes);
        }
}
__attribute__((target("avx2"))) static void mask_delta_avx2(const uint32_t* pm, uint32_t* sh, size_t cells, int elem, uint8_t* out) {
    const size_t cell_bytes = (size_t)78 * elem;
    size_t c = 0;
    for (; c + 8 <= cells; c += 8) {  // 8 cells = 24 words = three 256-bit compares
        const __m256i* a = (const __m256i*)(pm + 3 * c);
        const __m256i* b = (const __m256i*)(sh + 3 * c);
        const __m256i d = _mm256_or_si256(_mm256_or_si256(_mm256_xor_si256(_mm256_loadu_si256(a), _mm256_loadu_si256(b)),
                                                          _mm256_xor_si256(_mm256_loadu_si256(a + 1), _mm256_loadu_si256(b + 1))),
                                          _mm256_xor_si256(_mm256_loadu_si256(a + 2), _mm256_loadu_si256(b + 2)));
        if (_mm256_testz_si256(d, d)) continue;
        for (size_t k = c; k < c + 8; k++) {
            const uint32_t *n = pm + 3 * k;
            uint32_t* o = sh + 3 * k;
            if (n[0] == o[0] && n[1] == o[1] && n[2] == o[2]) continue;
            o[0] = n[0], o[1] = n[1], o[2] = n[2];
            mask_cell(n, elem, out + k * cell_bytes);
        }
    }
    for (; c < cells; c++) {
        const uint32_t* n = pm + 3 * c;
        uint32_t* o = sh + 3 * c;
        if (n[0] == o[0] && n[1] == o[1] && n[2] == o[2]) continue;
        o[0] = n[0], o[1] = n[1], o[2] = n[2];
        mask_cell(n, elem, out + c * cell_bytes);
    }
}
#endif

#if defined(__x86_64__)
__attribute__((target("avx2"))) static void mask_row_avx2_32(const uint32_t* pm, int32_t* o) { mask_cell_avx2_32(pm, o); }
#endif

void expand_obs_delta(Pool* p, const uint32_t* words, uint32_t* shadow, size_t n_env, int hw, int planes, int elem, void* out) {
    parallel_for(p, n_env, [=](size_t b, size_t e) {
        const size_t c0 = b * (size_t)hw, cells = (e - b) * (size_t)hw;
        uint8_t* o = (uint8_t*)out + c0 * (size_t)planes * elem;
#if defined(__x86_64__)
        if (has_avx2()) {
            obs_delta_avx2(words + c0, shadow + c0, cells, planes, elem, o);
            return;
        }
#endif
        for (size_t c = 0; c < cells; c++)
            if (words[c0 + c] != shadow[c0 + c]) {
                shadow[c0 + c] = words[c0 + c];
                obs_cell(words[c0 + c], planes, elem, o + c * (size_t)planes * elem);
            }
    });
}

void expand_mask_delta(Pool* p, const uint32_t* pmask, uint32_t* shadow, size_t n_env, int hw, int elem, void* out) {
    parallel_for(p, n_env, [=](size_t b, size_t e) {
        const size_t c0 = b * (size_t)hw, cells = (e - b) * (size_t)hw;
        uint8_t* o = (uint8_t*)out + c0 * (size_t)78 * elem;
#if defined(__x86_64__)
        if (has_avx2()) {
            mask_delta_avx2(pmask + 3 * c0, shadow + 3 * c0, cells, elem, o);
            return;
        }
#endif
        for (size_t c = 0; c < cells; c++) {
            const uint32_t* n = pmask + 3 * (c0 + c);
            uint32_t* s = shadow + 3 * (c0 + c);
            if (n[0] == s[0] && n[1] == s[1] && n[2] == s[2]) continue;
            s[0] = n[0], s[1] = n[1], s[2] = n[2];
            mask_cell(n, elem, o + c * (size_t)78 * elem);
        }
    });
}

// same encoding as mrts::Sim::pack_action (mrts_core.h)
template <class T>
static inline uint16_t pack_row(const T* a) {
    const long long t = (long long)a[0];
    if (t < 1 || t > 5) return 0;
    int pay = 0, bad = 0;
    if (t == 4) {
        if (a[5] < 0 || a[5] >= 7) return 0;
        bad = a[4] < 0 || a[4] > 3;
        pay = (bad ? 0 : (int)a[4]) | ((int)a[5] << 2);
    } else if (t == 5) {
        pay = (a[6] < 0 || a[6] > 48) ? 63 : (int)a[6];
    } else {
        const long long d = (long long)a[t];
        bad = d < 0 || d > 3;
        pay = bad ? 0 : (int)d;
    }
    return (uint16_t)(t | (pay << 3) | (bad << 9));
}

void pack_actions(Pool* p, const void* actions, int action_elem, const uint32_t* pmask, size_t n_env, int hw, uint16_t* out) {
    parallel_for(p, n_env, [=](size_t b, size_t e) {
        for (size_t env = b; env < e; env++) {
            const uint32_t* pm = pmask + env * (size_t)hw * 3;
            uint16_t* o = out + env * (size_t)hw;
            for (int c = 0; c < hw; c++) {
                if ((pm[3 * c] & 1u) == 0) {  // NOOP is legal exactly at the cells of units that may act
                    o[c] = 0;
                    continue;
                }
                const size_t row = (env * (size_t)hw + c) * 7;
                o[c] = action_elem == 8 ? pack_row((const int64_t*)actions + row) : pack_row((const int32_t*)actions + row);
            }
        }
    });
}

// Both loops below touch a few scattered cache lines per item of a large array: the addresses of a small batch are
// computed and prefetched first so that the misses overlap instead of queueing behind one another.
void pack_actions_bitmap(Pool* p, const void* actions, int action_elem, const uint32_t* bitmap, size_t cells, uint16_t* out) {
    const size_t nwords = (cells + 31) / 32;
    parallel_for(p, nwords, [=](size_t b, size_t e) {
        constexpr int BATCH = 64;
        size_t idx[BATCH + 32];
        const size_t row_bytes = (size_t)7 * action_elem;
        size_t i = b;
        while (i < e) {
            int k = 0;
            while (i < e && k < BATCH) {
                uint32_t bits = bitmap[i];
                while (bits) {
                    const size_t g = i * 32 + (size_t)__builtin_ctz(bits);
                    bits &= bits - 1;
                    if (g < cells) idx[k++] = g;
                }
                i++;
            }
            for (int j = 0; j < k; j++) {
                const char* row = (const char*)actions + idx[j] * row_bytes;
                __builtin_prefetch(row, 0, 0);
                __builtin_prefetch(row + row_bytes - 1, 0, 0);
            }
            for (int j = 0; j < k; j++) {
                const size_t g = idx[j];
                out[g] = action_elem == 8 ? pack_row((const int64_t*)actions + g * 7) : pack_row((const int32_t*)actions + g * 7);
            }
        }
    });
}

size_t pack_actions_sparse(Pool* p, const void* actions, int action_elem, const uint32_t* bitmap, size_t cells, uint32_t* pairs, size_t cap) {
    const size_t nwords = (cells + 31) / 32;
    std::atomic<size_t> cursor{0};
    std::atomic<size_t>* cur = &cursor;
    parallel_for(p, nwords, [=](size_t b, size_t e) {
        constexpr int BATCH = 64;
        size_t idx[BATCH + 32];
        const size_t row_bytes = (size_t)7 * action_elem;
        size_t i = b;
        while (i < e) {
            int k = 0;
            while (i < e && k < BATCH) {
                uint32_t bits = bitmap[i];
                while (bits) {
                    const size_t g = i * 32 + (size_t)__builtin_ctz(bits);
                    bits &= bits - 1;
                    if (g < cells) idx[k++] = g;
                }
                i++;
            }
            if (k == 0) continue;
            for (int j = 0; j < k; j++) {
                const char* row = (const char*)actions + idx[j] * row_bytes;
                __builtin_prefetch(row, 0, 0);
                __builtin_prefetch(row + row_bytes - 1, 0, 0);
            }
            const size_t at = cur->fetch_add((size_t)k);  // a run of k slots of the list (the order of the list is irrelevant)
            for (int j = 0; j < k; j++) {
                const size_t g = idx[j];
                const uint32_t v = action_elem == 8 ? pack_row((const int64_t*)actions + g * 7) : pack_row((const int32_t*)actions + g * 7);
                if (at + j < cap) pairs[2 * (at + j)] = (uint32_t)g, pairs[2 * (at + j) + 1] = v;
            }
        }
    });
    return cursor.load();
}

void patch_cells(Pool* p, const uint32_t* entries, size_t count, int planes, int elem, void* obs, void* mask) {
    parallel_for(p, count, [=](size_t b, size_t e) {
        const size_t obs_cell_bytes = (size_t)planes * elem, mask_cell_bytes = (size_t)78 * elem;
        constexpr size_t AHEAD = 8;
#if defined(__x86_64__)
        const bool avx = has_avx2(), wide = avx && planes >= 8 && planes <= 32;
#endif
        // the changed range of an entry's two rows: byte offsets [lo, hi) inside the tensors, or lo >= hi
        auto obs_range = [&](const uint32_t* en, size_t& lo, size_t& hi) {
            const uint32_t k0 = en[5] & 31u, k1 = (en[5] >> 5) & 31u;
            lo = en[0] * obs_cell_bytes + (size_t)k0 * elem;
            hi = k0 <= k1 ? en[0] * obs_cell_bytes + (size_t)(k1 + 1) * elem : lo;
        };
        auto mask_range = [&](const uint32_t* en, size_t& lo, size_t& hi) {
            const uint32_t b0 = (en[5] >> 10) & 127u, b1 = (en[5] >> 17) & 127u;
            lo = en[0] * mask_cell_bytes + (size_t)b0 * elem;
            hi = b0 <= b1 ? en[0] * mask_cell_bytes + (size_t)(b1 + 1) * elem : lo;
        };
        auto prefetch = [&](size_t i) {
            const uint32_t* en = entries + ENTRY_WORDS * i;
            size_t lo, hi;
            if (obs) {
                obs_range(en, lo, hi);
                for (size_t a = lo & ~(size_t)63; a < hi; a += 64) __builtin_prefetch((const char*)obs + a, 1, 0);
            }
            if (mask) {
                mask_range(en, lo, hi);
                for (size_t a = lo & ~(size_t)63; a < hi; a += 64) __builtin_prefetch((const char*)mask + a, 1, 0);
            }
        };
        for (size_t i = b; i < e && i < b + AHEAD; i++) prefetch(i);
        for (size_t i = b; i < e; i++) {
            if (i + AHEAD < e) prefetch(i + AHEAD);
            const uint32_t* en = entries + ENTRY_WORDS * i;
            const size_t g = en[0];
            if (obs) {
                const uint32_t k0 = en[5] & 31u, k1 = (en[5] >> 5) & 31u, w = en[1];
                if (elem == 4) {
                    int32_t* o = (int32_t*)obs + g * planes;
#if defined(__x86_64__)
                    if (wide && k0 <= k1 && k1 < (uint32_t)planes) {
                        const uint32_t bits3[3] = {w, 0u, 0u};
                        row_range_avx2_32(bits3, k0, k1, (uint32_t)planes, o);
                    } else
#endif
                        for (uint32_t k = k0; k <= k1 && k < (uint32_t)planes; k++) o[k] = (int32_t)((w >> k) & 1u);
                } else {
                    uint8_t* o = (uint8_t*)obs + g * planes;
                    for (uint32_t k = k0; k <= k1 && k < (uint32_t)planes; k++) o[k] = (uint8_t)((w >> k) & 1u);
                }
            }
            if (mask) {
                const uint32_t b0 = (en[5] >> 10) & 127u, b1 = (en[5] >> 17) & 127u;
                const uint32_t* pm = en + 2;
                if (elem == 4) {
                    int32_t* m = (int32_t*)mask + g * 78;
#if defined(__x86_64__)
                    if (avx && b0 <= b1 && b1 < 78u) row_range_avx2_32(pm, b0, b1, 78u, m);
                    else
#endif
                        for (uint32_t k = b0; k <= b1 && k < 78u; k++) m[k] = (int32_t)((pm[k >> 5] >> (k & 31)) & 1u);
                } else {
                    uint8_t* m = (uint8_t*)mask + g * 78;
                    for (uint32_t k = b0; k <= b1 && k < 78u; k++) m[k] = (uint8_t)((pm[k >> 5] >> (k & 31)) & 1u);
                }
            }
        }
    });
}

}  // namespace mrts_host

// the list form of the action packer on its own, for the CPU test-suite (declared in include/mrts_b200.h with the other codec entry points)
namespace mrts {
int api_fail(int code, const std::string& msg);  // mrts_api.cu
}
extern "C" int mrts_host_pack_actions_sparse(const void* actions, int action_elem_bytes, const uint32_t* bitmap, int64_t cells, uint32_t* pairs, int64_t cap,
                                             int64_t* n_pairs) {
    if (!actions || !bitmap || !pairs || !n_pairs || cells < 0 || cap < 0 || (action_elem_bytes != 4 && action_elem_bytes != 8))
        return mrts::api_fail(-1, "bad argument");
    mrts_host::Pool* p = mrts_host::pool_create(0);
    *n_pairs = (int64_t)mrts_host::pack_actions_sparse(p, actions, action_elem_bytes, bitmap, (size_t)cells, pairs, (size_t)cap);
    mrts_host::pool_destroy(p);
    return 0;
}
