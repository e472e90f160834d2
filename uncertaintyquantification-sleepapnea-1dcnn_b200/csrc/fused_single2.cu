// One kernel family of the fused tcgen05 path per translation unit (parallel compilation); see fused.cu.
#include "fused_kernels.cuh"

namespace uqk {

int fused_launch_single2(const FusedParams& p, cudaStream_t s) { return launch_drop<2, 4>(p, s); }

}  // namespace uqk
