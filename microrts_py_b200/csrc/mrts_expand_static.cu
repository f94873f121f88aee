// mrts_expand_static: the expansion of mrts_expand.cu for callers that must not allocate (code under CUDA graph
// capture, serving loops, harnesses that forbid cudaMalloc).  It needs no mrts_env: the shape comes with the call
// and the scratch memory lives in ONE file-scope __device__ array that every GPU loading this library holds until
// exit (MRTS_EXPAND_STATIC_SCRATCH_BYTES, 262,400 bytes).  No counterpart in the reference.
//
// What the scratch is for.  mrts_expand hands views out in index order from a work counter the env owns.  Here
//   scratch[0]                      the work counter of the launch
//   scratch[64 .. 64 + n_views)     the order in which views are handed out: most mask entries first
// (a one-block counting sort by entry count; views with many acting units take the longest, so the persistent
// blocks end together instead of waiting for a late heavy view).  Both are rewritten by every call -- nothing has
// to be zero beforehand, nothing survives that a later call relies on -- and EVERY call on a device shares them:
// two calls whose kernels overlap on the GPU (different streams without an ordering between them) race.
//
// Host code must not pass the array's NAME to a kernel (in host code that is a host-side stand-in, not a device
// address): the device address is taken once per device with cudaGetSymbolAddress, with that device current.
#include <mutex>
#include <string>

#include "../../include/mrts_b200.h"
#include "mrts_bulk.cuh"
#include "mrts_encode.h"
#include "mrts_launch.h"

namespace mrts {
int api_fail(int code, const std::string& msg);  // mrts_api.cu: sets mrts_last_error for this thread

namespace {

constexpr int HEADER_WORDS = 64;  // 256 bytes: the counter has a line of its own
constexpr int MAX_VIEWS = MRTS_EXPAND_STATIC_MAX_VIEWS;
static_assert((HEADER_WORDS + MAX_VIEWS) * 4 == MRTS_EXPAND_STATIC_SCRATCH_BYTES, "header and scratch array disagree");
static __device__ __align__(256) int32_t g_expand_scratch[HEADER_WORDS + MAX_VIEWS];

constexpr int EXP_WARPS = 8, EXP_THREADS = 32 * EXP_WARPS;
constexpr int ZERO_BYTES = 16 * bulk::GROUP_BYTES;
constexpr int SEG_CELLS = 1024;
constexpr int TILES_BYTES = ZERO_BYTES + bulk::EMPTY_BYTES_MAX;
constexpr int WARP_BYTES = SEG_CELLS * 4 + bulk::STAGE_BYTES;
constexpr int SMEM_BYTES = TILES_BYTES + EXP_WARPS * WARP_BYTES;
constexpr int ORDER_THREADS = 1024, BUCKETS = 64;
constexpr unsigned FULL = 0xFFFFFFFFu;

// One block: counting sort of the views by mask-entry count, descending; resets the work counter.
__global__ void __launch_bounds__(ORDER_THREADS) mrts_expand_static_order_kernel(const uint32_t* ment, int n, size_t stride, int32_t* scratch) {
    __shared__ int hist[BUCKETS], base[BUCKETS];
    const int tid = threadIdx.x;
    if (tid < BUCKETS) hist[tid] = 0;
    __syncthreads();
    for (int v = tid; v < n; v += ORDER_THREADS) {
        const uint32_t c = ment ? __ldg(ment + (size_t)v * stride) : 0u;
        atomicAdd(&hist[c < BUCKETS - 1 ? c : BUCKETS - 1], 1);
    }
    __syncthreads();
    if (tid == 0) {
        int run = 0;
        for (int b = BUCKETS - 1; b >= 0; b--) base[b] = run, run += hist[b];
        scratch[0] = 0;
    }
    __syncthreads();
    int32_t* order = scratch + HEADER_WORDS;
    for (int v = tid; v < n; v += ORDER_THREADS) {
        const uint32_t c = ment ? __ldg(ment + (size_t)v * stride) : 0u;
        order[atomicAdd(&base[c < BUCKETS - 1 ? c : BUCKETS - 1], 1)] = v;
    }
}

// mrts_expand_kernel (mrts_expand.cu) with the views taken from the order list in the scratch array.
template <int ELEM, int PLANES>
__global__ void __launch_bounds__(EXP_THREADS) mrts_expand_static_kernel(const __grid_constant__ ExpandParams X, const int32_t* __restrict__ order) {
    extern __shared__ uint4 smem_raw[];
    uint8_t* const sm = (uint8_t*)smem_raw;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    bulk::tiles_init(sm, ZERO_BYTES, ELEM, PLANES, X.empty_word, threadIdx.x, EXP_THREADS);
    uint8_t* const wm = sm + TILES_BYTES + warp * WARP_BYTES;
    uint32_t* const wseg = (uint32_t*)wm;
    const int HW = X.HW;
    const size_t obs_env = (size_t)HW * PLANES * ELEM, mask_env = (size_t)HW * 78 * ELEM;
    const size_t ment_stride = ment_stride_words(X.cap);
    bulk::Stage stage;
    stage.init(wm + SEG_CELLS * 4, lane);
    bulk::ObsWriter ow;
    ow.init(&stage, sm + ZERO_BYTES, X.empty_word, ELEM, PLANES);
    bulk::MaskWriter mw;
    mw.init(&stage, sm, ZERO_BYTES, ELEM);

    for (;;) {
        int v = -1;
        if (lane == 0) {
            const int t = atomicAdd(X.counter, 1);
            if (t < X.env_hi) v = order[t];
        }
        v = __shfl_sync(FULL, v, 0);
        if (v < 0) break;
        if (X.obs) {
            const uint32_t* words = X.pobs + (size_t)v * HW;
            ow.begin((uint8_t*)X.obs + (size_t)v * obs_env);
#pragma unroll 1
            for (int c0 = 0; c0 < HW; c0 += SEG_CELLS) {
                const int cells = HW - c0 < SEG_CELLS ? HW - c0 : SEG_CELLS;
                __syncwarp();
                for (int k = lane * 4; k < cells; k += 128) *(uint4*)(wseg + k) = __ldg((const uint4*)(words + c0 + k));
                __syncwarp();
                ow.segment(wseg, c0 / (16 / ELEM), cells / (16 / ELEM));
            }
            ow.finish(HW / (16 / ELEM));
        }
        if (X.mask) {
            const uint32_t* ent = X.ment + (size_t)v * ment_stride;
            int count = (int)__ldg(ent);
            if (count > X.cap) count = X.cap;  // entries come from the caller here, not from our own kernels
            mw.begin((uint8_t*)X.mask + (size_t)v * mask_env);
#pragma unroll 1
            for (int b0 = 0; b0 < count; b0 += 32) {
                uint4 e = make_uint4(0, 0, 0, 0);
                if (b0 + lane < count) e = __ldg((const uint4*)(ent + 4 * (size_t)(1 + b0 + lane)));
                const int nb = count - b0 < 32 ? count - b0 : 32;
#pragma unroll 1
                for (int i = 0; i < nb; i++)
                    mw.add(__shfl_sync(FULL, e.x, i), __shfl_sync(FULL, e.y, i), __shfl_sync(FULL, e.z, i), __shfl_sync(FULL, e.w, i));
            }
            mw.finish(HW / (8 / ELEM));
        }
    }
    stage.drain();
}

struct DeviceSetup {
    bool ready = false;
    int32_t* scratch = nullptr;  // device address of g_expand_scratch on this device
    int n_sm = 0;
};
constexpr int MAX_DEVICES = 64;
DeviceSetup g_setup[MAX_DEVICES];
std::mutex g_setup_mu;

#define SCK(call)                                                                                         \
    do {                                                                                                  \
        cudaError_t e_ = (call);                                                                          \
        if (e_ != cudaSuccess) return api_fail(MRTS_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

// once per device, with that device current
int setup_device(int dev, DeviceSetup** out) {
    std::lock_guard<std::mutex> lock(g_setup_mu);
    DeviceSetup& d = g_setup[dev];
    if (!d.ready) {
        int32_t* p = nullptr;
        SCK(cudaGetSymbolAddress((void**)&p, g_expand_scratch));
        if (!p) return api_fail(MRTS_ECUDA, "cudaGetSymbolAddress returned a null address for the scratch array");
        SCK(cudaFuncSetAttribute(mrts_expand_static_kernel<4, 29>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        SCK(cudaFuncSetAttribute(mrts_expand_static_kernel<4, 31>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        SCK(cudaFuncSetAttribute(mrts_expand_static_kernel<1, 29>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        SCK(cudaFuncSetAttribute(mrts_expand_static_kernel<1, 31>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        SCK(cudaDeviceGetAttribute(&d.n_sm, cudaDevAttrMultiProcessorCount, dev));
        d.scratch = p;
        d.ready = true;
    }
    *out = &d;
    return MRTS_OK;
}

}  // namespace
}  // namespace mrts

using namespace mrts;

extern "C" int mrts_expand_static(const uint32_t* obs_words, const uint32_t* mask_entries, void* obs, void* mask, int n_views, int height, int width,
                                  int unit_capacity, int partial_obs, int element_bytes, void* stream) {
    if (n_views <= 0 || height <= 0 || width <= 0 || unit_capacity < 0) return api_fail(MRTS_EINVAL, "mrts_expand_static: n_views, height, width must be positive");
    if (n_views > MAX_VIEWS)
        return api_fail(MRTS_EINVAL, "mrts_expand_static: n_views = " + std::to_string(n_views) + " exceeds MRTS_EXPAND_STATIC_MAX_VIEWS = " + std::to_string(MAX_VIEWS) +
                                         " (the static scratch array holds one word per view); split the call or use mrts_expand");
    if ((long long)height * width > (1 << 20)) return api_fail(MRTS_EINVAL, "mrts_expand_static: map too large");
    if ((obs && !obs_words) || (mask && !mask_entries)) return api_fail(MRTS_EINVAL, "mrts_expand_static: packed input missing");
    if (!obs && !mask) return MRTS_OK;
    const int HW = height * width, planes = partial_obs ? 31 : 29;
    if (!expand_supported(HW, planes, element_bytes, obs, mask) || ((((uintptr_t)obs_words) | ((uintptr_t)mask_entries)) & 15))
        return api_fail(MRTS_EINVAL, "mrts_expand_static needs element_bytes 1 or 4, H*W*element_bytes a multiple of 16 and 16-byte aligned arrays");
    int dev = -1;
    SCK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEVICES) return api_fail(MRTS_EINVAL, "mrts_expand_static: device index out of range");
    {
        // the scratch address belongs to the CURRENT device: refuse tensors that live on another one
        cudaPointerAttributes at;
        SCK(cudaPointerGetAttributes(&at, obs ? obs : mask));
        if (at.type == cudaMemoryTypeDevice && at.device != dev)
            return api_fail(MRTS_EINVAL, "mrts_expand_static: the tensors are on device " + std::to_string(at.device) + " but device " + std::to_string(dev) + " is current");
    }
    DeviceSetup* d = nullptr;
    int rc = setup_device(dev, &d);
    if (rc) return rc;
    const cudaStream_t st = (cudaStream_t)stream;
    mrts_expand_static_order_kernel<<<1, ORDER_THREADS, 0, st>>>(mask ? mask_entries : nullptr, n_views, ment_stride_words(unit_capacity), d->scratch);
    SCK(cudaGetLastError());
    ExpandParams X;
    X.pobs = obs_words;
    X.ment = mask_entries;
    X.obs = obs;
    X.mask = mask;
    X.env_lo = 0;
    X.env_hi = n_views;
    X.HW = HW;
    X.planes = planes;
    X.elem = element_bytes;
    X.cap = unit_capacity;
    X.empty_word = obs_empty_word(partial_obs ? 1 : 0);
    X.counter = d->scratch;
    const int32_t* order = d->scratch + HEADER_WORDS;
    int blocks = (n_views + EXP_WARPS - 1) / EXP_WARPS;
    if (blocks > d->n_sm) blocks = d->n_sm;
    if (element_bytes == 4 && planes == 29) mrts_expand_static_kernel<4, 29><<<blocks, EXP_THREADS, SMEM_BYTES, st>>>(X, order);
    else if (element_bytes == 4) mrts_expand_static_kernel<4, 31><<<blocks, EXP_THREADS, SMEM_BYTES, st>>>(X, order);
    else if (planes == 29) mrts_expand_static_kernel<1, 29><<<blocks, EXP_THREADS, SMEM_BYTES, st>>>(X, order);
    else mrts_expand_static_kernel<1, 31><<<blocks, EXP_THREADS, SMEM_BYTES, st>>>(X, order);
    SCK(cudaGetLastError());
    return MRTS_OK;
}
