// Microbenchmark: what one warp pays to write a mostly-constant 80 KB block (one 16x16 mask view) in different ways.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bulk_store_probe bulk_store_probe.cu && ./bulk_store_probe
// mode 0: lane 0 issues cp.async.bulk pieces of S bytes from a zero tile in shared memory (what MaskWriter does)
// mode 1: every lane stores 16 bytes per instruction (st.global.v4)
// mode 2: as 0, but four lanes issue every fourth piece each
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void bulk_store(void* g, const void* s, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"((uint32_t)__cvta_generic_to_shared(s)), "r"(bytes) : "memory");
}

__global__ void probe(uint8_t* out, size_t view_bytes, int piece, int mode, int views_per_warp, long long* cycles) {
    extern __shared__ uint4 tile[];
    const int lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < piece / 16; i += blockDim.x) tile[i] = make_uint4(0, 0, 0, 0);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    const long long t0 = clock64();
    for (int v = 0; v < views_per_warp; v++) {
        uint8_t* dst = out + ((size_t)blockIdx.x * views_per_warp + v) * view_bytes;
        if (mode == 1) {
            for (size_t o = (size_t)lane * 16; o < view_bytes; o += 512) *(uint4*)(dst + o) = make_uint4(0, 0, 0, 0);
        } else {
            const int nl = mode == 2 ? 4 : 1;
            if (lane < nl) {
                int k = 0;
                for (size_t o = 0; o < view_bytes; o += piece, k++) {
                    if (k % nl != lane) continue;
                    const size_t n = view_bytes - o < (size_t)piece ? view_bytes - o : (size_t)piece;
                    bulk_store(dst + o, tile, (uint32_t)n);
                }
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        }
    }
    const long long t1 = clock64();
    if (mode != 1 && lane < 4) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    const long long t2 = clock64();
    if (lane == 0) {
        atomicAdd((unsigned long long*)&cycles[0], (unsigned long long)(t1 - t0));
        atomicAdd((unsigned long long*)&cycles[1], (unsigned long long)(t2 - t0));
    }
}

int main() {
    const size_t view = 79872;  // 256 cells x 78 x 4
    const int blocks_per_sm[] = {1, 8, 20};
    int n_sm = 148;
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, 0);
    const int vpw = 4;
    uint8_t* out;
    long long* cyc;
    const size_t max_blocks = (size_t)n_sm * 20;
    cudaMalloc(&out, max_blocks * vpw * view);
    cudaMalloc(&cyc, 16);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int pieces[] = {624, 1248, 2496, 4992, 9984, 19968, 39936};
    for (int bps : blocks_per_sm) {
        const int blocks = n_sm * bps;
        for (int mode = 0; mode < 3; mode++)
            for (int piece : pieces) {
                if (mode == 1 && piece != 4992) continue;
                if (bps * (piece + 1024) > 220 * 1024) continue;
                for (int rep = 0; rep < 2; rep++) {
                    cudaMemset(cyc, 0, 16);
                    cudaEventRecord(e0);
                    probe<<<blocks, 32, piece, 0>>>(out, view, piece, mode, vpw, cyc);
                    cudaEventRecord(e1);
                    cudaDeviceSynchronize();
                }
                long long h[2];
                cudaMemcpy(h, cyc, 16, cudaMemcpyDeviceToHost);
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                printf("warps/SM %2d mode %d piece %5d: issue %8.0f cycles/view, issue+drain %8.0f cycles/view, kernel %.3f ms = %.0f GB/s  (%s)\n", bps, mode, piece,
                       (double)h[0] / blocks / vpw, (double)h[1] / blocks / vpw, ms, blocks * vpw * view / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
            }
    }
    return 0;
}
