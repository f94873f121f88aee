// Rules kernels of the light game class (see mrts_rules.cuh).
#include "mrts_rules.cuh"

namespace mrts {

cudaError_t light_launch_warp(int regs, const EnvParams& P, int gpb, int smem_per_game, cudaStream_t st) { return launch_warp_t<0>(regs, P, gpb, smem_per_game, st); }
cudaError_t light_launch_block(const EnvParams& P, int threads, int smem, cudaStream_t st) {
    mrts_rules_block_kernel<0><<<P.nslots, threads, smem, st>>>(P);
    return cudaGetLastError();
}
cudaError_t light_set_smem(int warp_bytes, int block_bytes) { return set_smem_t<0>(warp_bytes, block_bytes); }
int light_warp_regs(int regs) { return warp_regs_t<0>(regs); }

}  // namespace mrts
