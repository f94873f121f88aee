// Rules kernels of the heavy game class (see mrts_rules.cuh).
#include "mrts_rules.cuh"

namespace mrts {

cudaError_t heavy_launch_warp(int regs, const EnvParams& P, int gpb, int smem_per_game, cudaStream_t st) { return launch_warp_t<1>(regs, P, gpb, smem_per_game, st); }
cudaError_t heavy_launch_block(const EnvParams& P, int threads, int smem, cudaStream_t st) {
    mrts_rules_block_kernel<1><<<P.nslots, threads, smem, st>>>(P);
    return cudaGetLastError();
}
cudaError_t heavy_set_smem(int warp_bytes, int block_bytes) { return set_smem_t<1>(warp_bytes, block_bytes); }
int heavy_warp_regs(int regs) { return warp_regs_t<1>(regs); }

}  // namespace mrts
