// One kernel family of the fused tcgen05 path per translation unit (parallel compilation); see fused.cu.
#include "fused_kernels.cuh"

namespace uqk {

// statistics sweeps of the exact batch-statistic mode: two-tile CTA-pair kernel, 8-slot weight ring (the statistics
// accumulators take the shared memory of one ring slot)
int fused_launch_pair_stat(const FusedParams& p, cudaStream_t s) { return launch_pair2_stat<8>(p, s); }

}  // namespace uqk
