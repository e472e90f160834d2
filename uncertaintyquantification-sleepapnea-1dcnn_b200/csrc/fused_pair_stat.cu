// One kernel family of the fused tcgen05 path per translation unit (parallel compilation); see fused.cu.
#include "fused_kernels.cuh"

namespace uqk {

// statistics sweeps of the exact batch-statistic mode: two-tile CTA-pair kernel, 4 ring slots of 16 KB per CTA (the
// statistics accumulators overlay the head partials / probability buffer the sweeps do not use)
int fused_launch_pair_stat(const FusedParams& p, cudaStream_t s) { return launch_pair2_stat<4>(p, s); }

}  // namespace uqk
