// K4: stand-alone per-window UQ reductions over an existing (T, N) float32 probability tensor.
// Replaces uq_evaluation_dist (reference uq_techniques.py:40-112) and the drivers' inline metrics
// (analyze_mcd_patient_level.py:109-117).  HBM-bound: 4*T bytes read + <=28 bytes written per window.
//
// Layout: probs[t*N + n].  A block stages a (T x W) tile in shared memory with coalesced, vectorised loads so the
// tensor is read from HBM exactly once although the variance needs two passes over it; each thread then walks
// its window's column in pass order (the evaluation order of np.mean/np.var(axis=0)).
#include "common.cuh"

namespace uqk {

constexpr int kReduceThreads = 128;

__global__ void __launch_bounds__(kReduceThreads)
reduce_kernel_smem(const float* __restrict__ probs, int n_pass, int64_t n, MetricsOut o) {
  extern __shared__ float tile[];  // [n_pass][kReduceThreads]
  const int64_t n_tiles = (n + kReduceThreads - 1) / kReduceThreads;
  for (int64_t tix = blockIdx.x; tix < n_tiles; tix += gridDim.x) {
    const int64_t n0 = tix * kReduceThreads;
    const int w = (int)min((int64_t)kReduceThreads, n - n0);
    const bool vec_ok = ((n & 3) == 0) && (w == kReduceThreads) && ((reinterpret_cast<uintptr_t>(probs) & 15) == 0);
    if (vec_ok) {
      // 32 lanes x float4 = one 512-byte row segment per warp instruction
      for (int i = threadIdx.x; i < n_pass * (kReduceThreads / 4); i += kReduceThreads) {
        const int t = i / (kReduceThreads / 4), c = i % (kReduceThreads / 4);
        const float4 v = __ldcs(reinterpret_cast<const float4*>(probs + (size_t)t * n + n0) + c);
        reinterpret_cast<float4*>(tile + t * kReduceThreads)[c] = v;
      }
    } else {
      for (int i = threadIdx.x; i < n_pass * kReduceThreads; i += kReduceThreads) {
        const int t = i / kReduceThreads, c = i % kReduceThreads;
        tile[i] = c < w ? __ldcs(probs + (size_t)t * n + n0 + c) : 0.f;
      }
    }
    __syncthreads();
    if ((int)threadIdx.x < w) {
      const float* col = tile + threadIdx.x;
      window_metrics([&](int t) { return col[t * kReduceThreads]; }, n_pass, n0 + threadIdx.x, o);
    }
    __syncthreads();
  }
}

// Fallback for very large T (tile would not fit in shared memory): re-reads the column (second read hits L2).
__global__ void __launch_bounds__(kReduceThreads)
reduce_kernel_direct(const float* __restrict__ probs, int n_pass, int64_t n, MetricsOut o) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float* col = probs + i;
    window_metrics([&](int t) { return __ldg(col + (size_t)t * n); }, n_pass, i, o);
  }
}

int launch_reduce(const float* probs, int n_pass, int64_t n, const MetricsOut& o, cudaStream_t s) {
  if (n == 0) return UQK_OK;
  int dev = 0, sms = 148;
  UQK_CUDA(cudaGetDevice(&dev));
  UQK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const size_t smem = (size_t)n_pass * kReduceThreads * sizeof(float);
  const int64_t n_tiles = (n + kReduceThreads - 1) / kReduceThreads;
  if (smem <= 200 * 1024) {
    if (smem > 48 * 1024)
      UQK_CUDA(cudaFuncSetAttribute(reduce_kernel_smem, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int per_sm = (int)max((size_t)1, min((size_t)8, (size_t)(220 * 1024) / max(smem, (size_t)1)));
    const int grid = (int)min(n_tiles, (int64_t)sms * per_sm);
    reduce_kernel_smem<<<grid, kReduceThreads, smem, s>>>(probs, n_pass, n, o);
  } else {
    const int grid = (int)min(n_tiles, (int64_t)sms * 8);
    reduce_kernel_direct<<<grid, kReduceThreads, 0, s>>>(probs, n_pass, n, o);
  }
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}

}  // namespace uqk
