// One kernel family of the fused tcgen05 path per translation unit (parallel compilation); see fused.cu.
#include "fused_kernels.cuh"

namespace uqk {

#ifndef UQK_PAIR2_SLOTS
#define UQK_PAIR2_SLOTS 4
#endif
int fused_launch_pair2(const FusedParams& p, cudaStream_t s) { return launch_pair2<UQK_PAIR2_SLOTS>(p, s); }

}  // namespace uqk
