// K4: stand-alone per-window UQ reductions over an existing (T, N) float32 probability tensor.
// Replaces uq_evaluation_dist (reference uq_techniques.py:40-112) and the drivers' inline metrics
// (analyze_mcd_patient_level.py:109-117).  HBM-bound: 4*T bytes read + <=28 bytes written per window.
//
// Layout: probs[t*N + n].  A block stages a (T x W) tile in shared memory with coalesced, vectorised loads so the
// tensor is read from HBM exactly once although the variance needs two passes over it; each thread then walks
// its window's column in pass order (the evaluation order of np.mean/np.var(axis=0)).
#include "common.cuh"

#include <algorithm>

namespace uqk {

constexpr int kReduceThreads = 128;

__global__ void __launch_bounds__(kReduceThreads)
reduce_kernel_smem(const float* __restrict__ probs, int n_pass, int64_t n, MetricsOut o, const int64_t* __restrict__ out_idx) {
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
      const int64_t w_in = n0 + threadIdx.x;
      window_metrics([&](int t) { return col[t * kReduceThreads]; }, n_pass, out_idx ? __ldg(out_idx + w_in) : w_in, o);
    }
    __syncthreads();
  }
}

// Fallback for very large T (tile would not fit in shared memory): re-reads the column (second read hits L2).
__global__ void __launch_bounds__(kReduceThreads)
reduce_kernel_direct(const float* __restrict__ probs, int n_pass, int64_t n, MetricsOut o, const int64_t* __restrict__ out_idx) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float* col = probs + i;
    window_metrics([&](int t) { return __ldg(col + (size_t)t * n); }, n_pass, out_idx ? __ldg(out_idx + i) : i, o);
  }
}

// out_idx (optional): window i of `probs` is written to row out_idx[i] of the outputs (guard-band re-evaluation).
int launch_reduce(const float* probs, int n_pass, int64_t n, const MetricsOut& o, cudaStream_t s, const int64_t* out_idx) {
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
    reduce_kernel_smem<<<grid, kReduceThreads, smem, s>>>(probs, n_pass, n, o, out_idx);
  } else {
    const int grid = (int)min(n_tiles, (int64_t)sms * 8);
    reduce_kernel_direct<<<grid, kReduceThreads, 0, s>>>(probs, n_pass, n, o, out_idx);
  }
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}

// ---- guard band of the labels (analyze_mcd_patient_level.py:117: label = mean_p > 0.5) --------------------------------
// Windows whose mean probability from the bf16 tensor-core path lies within `delta` of 0.5 are listed (any order) for
// re-evaluation on the fp32 path; count[0] receives their number.
__global__ void select_band_kernel(const float* __restrict__ mean_pred, int64_t n, float delta, int64_t* __restrict__ idx,
                                   unsigned long long* __restrict__ count) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    if (fabsf(mean_pred[i] - 0.5f) < delta) idx[atomicAdd(count, 1ull)] = i;
  }
}
int launch_select_band(const float* mean_pred, int64_t n, float delta, int64_t* idx, unsigned long long* count, cudaStream_t s) {
  UQK_CUDA(cudaMemsetAsync(count, 0, sizeof(unsigned long long), s));
  if (n == 0) return UQK_OK;
  select_band_kernel<<<(int)std::min<int64_t>((n + 255) / 256, 148 * 8), 256, 0, s>>>(mean_pred, n, delta, idx, count);
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}
// probs_sub (T, n_sub) -> columns idx[i] of probs (T, n)
__global__ void scatter_probs_kernel(const float* __restrict__ sub, const int64_t* __restrict__ idx, int n_pass, int64_t n_sub,
                                     int64_t n, float* __restrict__ probs) {
  const int64_t total = (int64_t)n_pass * n_sub;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t t = i / n_sub, w = i - t * n_sub;
    probs[t * n + __ldg(idx + w)] = sub[i];
  }
}
int launch_scatter_probs(const float* sub, const int64_t* idx, int n_pass, int64_t n_sub, int64_t n, float* probs, cudaStream_t s) {
  const int64_t total = (int64_t)n_pass * n_sub;
  if (total == 0) return UQK_OK;
  scatter_probs_kernel<<<(int)std::min<int64_t>((total + 255) / 256, 148 * 8), 256, 0, s>>>(sub, idx, n_pass, n_sub, n, probs);
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}

}  // namespace uqk

// ---- N3: per-window standardisation (reference data_prepocessing/prepare_numpy_datasets.py:83-95) -----------------
// out = (x - mean_t) / (std_t + 1e-8), mean / population std over the 60 time steps of each (window, channel), in
// float64 like NumPy on the float64 CSV data.  One warp per window: lane l owns rows l and l+32 (all 4 channels);
// HBM-bound: 1920 B read + 1920 B written per float64 window.
namespace uqk {
__device__ __forceinline__ void load_row4(const double* p, bool ok, double (&v)[4]) {
  if (ok) {
    const double2 a = __ldcs(reinterpret_cast<const double2*>(p)), b = __ldcs(reinterpret_cast<const double2*>(p) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
  } else { v[0] = v[1] = v[2] = v[3] = 0.0; }
}
__device__ __forceinline__ void load_row4(const float* p, bool ok, double (&v)[4]) {
  if (ok) {
    const float4 a = __ldcs(reinterpret_cast<const float4*>(p));
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
  } else { v[0] = v[1] = v[2] = v[3] = 0.0; }
}
__device__ __forceinline__ void store_row4(double* p, const double (&z)[4]) {
  __stcs(reinterpret_cast<double2*>(p), make_double2(z[0], z[1]));
  __stcs(reinterpret_cast<double2*>(p) + 1, make_double2(z[2], z[3]));
}
__device__ __forceinline__ void store_row4(float* p, const double (&z)[4]) {
  __stcs(reinterpret_cast<float4*>(p), make_float4((float)z[0], (float)z[1], (float)z[2], (float)z[3]));
}

template <typename T>
__global__ void __launch_bounds__(256)
standardize_kernel(const T* __restrict__ x, T* __restrict__ out, int64_t n, double eps) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t w = warp; w < n; w += nwarps) {
    const T* xw = x + (size_t)w * kL * kF;
    double v[2][kF];
    double s[kF] = {0, 0, 0, 0};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int r = lane + 32 * h;
      load_row4(xw + r * kF, r < kL, v[h]);     // one row = 4 channels = one 16 / 2 x 16 byte access
#pragma unroll
      for (int c = 0; c < kF; ++c) s[c] += v[h][c];
    }
#pragma unroll
    for (int c = 0; c < kF; ++c) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s[c] += __shfl_xor_sync(0xffffffffu, s[c], o);
      s[c] /= (double)kL;
    }
    double q[kF] = {0, 0, 0, 0};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int r = lane + 32 * h;
#pragma unroll
      for (int c = 0; c < kF; ++c) {
        const double d = r < kL ? v[h][c] - s[c] : 0.0;
        q[c] += d * d;
      }
    }
#pragma unroll
    for (int c = 0; c < kF; ++c) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) q[c] += __shfl_xor_sync(0xffffffffu, q[c], o);
      q[c] = sqrt(q[c] / (double)kL) + eps;
    }
    T* ow = out + (size_t)w * kL * kF;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int r = lane + 32 * h;
      if (r < kL) {
        double z[kF];
#pragma unroll
        for (int c = 0; c < kF; ++c) z[c] = (v[h][c] - s[c]) / q[c];
        store_row4(ow + r * kF, z);
      }
    }
  }
}
}  // namespace uqk

extern "C" int uqk_standardize(const void* x_dev, int x_dtype, int64_t n_windows, void* out_dev, void* stream) {
  using namespace uqk;
  if (n_windows < 0) { set_error("n_windows must be >= 0"); return UQK_E_BADARG; }
  if (n_windows == 0) return UQK_OK;
  if (!x_dev || !out_dev) { set_error("NULL argument"); return UQK_E_BADARG; }
  if (x_dtype != UQK_F32 && x_dtype != UQK_F64) { set_error("x_dtype must be UQK_F32 or UQK_F64"); return UQK_E_BADARG; }
  if ((reinterpret_cast<uintptr_t>(x_dev) | reinterpret_cast<uintptr_t>(out_dev)) & 15) {
    set_error("x_dev and out_dev must be 16-byte aligned (rows are moved with 16-byte accesses)"); return UQK_E_BADARG;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int grid = (int)std::min<int64_t>((n_windows * 32 + 255) / 256, 148 * 16);
  if (x_dtype == UQK_F64)
    standardize_kernel<double><<<grid, 256, 0, s>>>((const double*)x_dev, (double*)out_dev, n_windows, 1e-8);
  else
    standardize_kernel<float><<<grid, 256, 0, s>>>((const float*)x_dev, (float*)out_dev, n_windows, 1e-8);
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}
