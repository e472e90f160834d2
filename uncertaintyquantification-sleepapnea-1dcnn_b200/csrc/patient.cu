// K5: per-patient segmented reduction (replaces the pandas groupby of reference
//     uncertainty_quantification/aggregate_patient_uq_metrics.py:34-46), and
// K6: bootstrap aggregates as gather-means (replaces the resample + recompute loop of
//     uq_techniques.py:137-158; the per-window metrics of preds[:, idx] are metrics[idx]).
// Both are HBM/L2-bound integer+float work; no tensor cores.
#include "common.cuh"
#include "internal.h"

#include <mutex>

namespace uqk {

constexpr int kPatThreads = 256;
constexpr int kPatMaxSeg = 8192;  // windows per patient that fit the in-smem sort (a 24 h night is 1440)

__device__ __forceinline__ double block_sum(double v, double* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  double t = 0;
  for (int i = 0; i < kPatThreads / 32; ++i) t += red[i];
  return t;
}

// bitonic sort of `m` (power of two) floats in shared memory
__device__ void bitonic_sort(float* a, int m) {
  for (int k = 2; k <= m; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      __syncthreads();
      for (int i = threadIdx.x; i < m; i += kPatThreads) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const float x = a[i], y = a[ixj];
          const bool up = (i & k) == 0;
          if ((x > y) == up) { a[i] = y; a[ixj] = x; }
        }
      }
    }
  }
  __syncthreads();
}

__device__ void column_stats(const float* __restrict__ col, int64_t s, int n, float* sbuf, double* red, double* mean_out,
                             double* median_out, double* std_out) {
  double acc = 0;
  for (int i = threadIdx.x; i < n; i += kPatThreads) acc += (double)col[s + i];
  const double mean = block_sum(acc, red) / (double)n;
  acc = 0;
  for (int i = threadIdx.x; i < n; i += kPatThreads) { const double d = (double)col[s + i] - mean; acc += d * d; }
  const double ss = block_sum(acc, red);
  int m = 1;
  while (m < n) m <<= 1;
  for (int i = threadIdx.x; i < m; i += kPatThreads) sbuf[i] = i < n ? col[s + i] : __int_as_float(0x7f800000);
  bitonic_sort(sbuf, m);
  const double med = (n & 1) ? (double)sbuf[n / 2] : 0.5 * ((double)sbuf[n / 2 - 1] + (double)sbuf[n / 2]);
  if (threadIdx.x == 0) {
    *mean_out = mean;
    *median_out = med;
    *std_out = n > 1 ? sqrt(ss / (double)(n - 1)) : 0.0;   // ddof=1; forced to 0 for single-window patients (:45-46)
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kPatThreads)
patient_reduce_kernel(const float* __restrict__ variance, const float* __restrict__ entropy,
                      const int32_t* __restrict__ y_true, const int32_t* __restrict__ y_pred,
                      const int64_t* __restrict__ seg, int64_t n_patients, double* __restrict__ out, int* err) {
  __shared__ float sbuf[kPatMaxSeg];
  __shared__ double red[kPatThreads / 32];
  for (int64_t p = blockIdx.x; p < n_patients; p += gridDim.x) {
    const int64_t s = seg[p], e = seg[p + 1];
    const int64_t len = e - s;
    double* o = out + p * 8;
    if (len <= 0 || len > kPatMaxSeg) {
      if (threadIdx.x == 0) {
        if (len > kPatMaxSeg) atomicExch(err, 1);
        for (int i = 0; i < 8; ++i) o[i] = len == 0 ? 0.0 : __longlong_as_double(0x7ff8000000000000ll);
      }
      continue;
    }
    const int n = (int)len;
    column_stats(variance, s, n, sbuf, red, o + 0, o + 1, o + 2);
    column_stats(entropy, s, n, sbuf, red, o + 3, o + 4, o + 5);
    double c = 0;
    for (int i = threadIdx.x; i < n; i += kPatThreads) c += (y_true[s + i] == y_pred[s + i]) ? 1.0 : 0.0;   // :34
    c = block_sum(c, red);
    if (threadIdx.x == 0) { o[6] = c / (double)n; o[7] = (double)n; }
    __syncthreads();
  }
}

// ---- bootstrap gather-means ---------------------------------------------------------------------------------------
constexpr int kBootThreads = 256;
__global__ void __launch_bounds__(kBootThreads)
bootstrap_kernel(const float* __restrict__ var, const float* __restrict__ tot, const float* __restrict__ expc,
                 const float* __restrict__ mi, const int32_t* __restrict__ y, const int64_t* __restrict__ idx,
                 int64_t n, double* __restrict__ sums /* (n_boot, 8) */) {
  const int b = blockIdx.y;
  const int64_t* ib = idx + (size_t)b * n;
  double a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int64_t i = (int64_t)blockIdx.x * kBootThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kBootThreads) {
    const int64_t j = __ldcs(ib + i);
    const double v = (double)__ldg(var + j);
    const int yy = __ldg(y + j);
    a[0] += v;
    if (yy == 0) { a[1] += v; a[2] += 1.0; }
    if (yy == 1) { a[3] += v; a[4] += 1.0; }
    a[5] += (double)__ldg(tot + j);
    a[6] += (double)__ldg(expc + j);
    a[7] += (double)__ldg(mi + j);
  }
  __shared__ double red[8][kBootThreads / 32];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    double v = a[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (l == 0) red[k][w] = v;
  }
  __syncthreads();
  if (threadIdx.x < 8) {
    double t = 0;
    for (int i = 0; i < kBootThreads / 32; ++i) t += red[threadIdx.x][i];
    atomicAdd(&sums[(size_t)b * 8 + threadIdx.x], t);
  }
}

// Array-of-structs staging for uqk_bootstrap: the five per-window values a resample needs share one 32-byte L2 sector,
// so a random gather costs ONE sector instead of five (the gathers, not the index stream, bound this kernel: 5.7 MB
// arrays stay in L2 and every 4-byte random access moves a whole sector).
struct __align__(32) BootRow { float var, tot, expc, mi; int32_t y; int32_t pad[3]; };

__global__ void bootstrap_pack_kernel(const float* __restrict__ var, const float* __restrict__ tot,
                                      const float* __restrict__ expc, const float* __restrict__ mi,
                                      const int32_t* __restrict__ y, int64_t n, BootRow* __restrict__ rows) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    BootRow r;
    r.var = var[i]; r.tot = tot[i]; r.expc = expc[i]; r.mi = mi[i]; r.y = y[i]; r.pad[0] = r.pad[1] = r.pad[2] = 0;
    rows[i] = r;
  }
}

__global__ void __launch_bounds__(kBootThreads)
bootstrap_rows_kernel(const BootRow* __restrict__ rows, const int64_t* __restrict__ idx, int64_t n,
                      double* __restrict__ sums /* (n_boot, 8) */) {
  const int b = blockIdx.y;
  const int64_t* ib = idx + (size_t)b * n;
  double a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int64_t i = (int64_t)blockIdx.x * kBootThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kBootThreads) {
    const int64_t j = __ldcs(ib + i);
    const float4 f = __ldg(reinterpret_cast<const float4*>(rows + j));
    const int yy = __ldg(&rows[j].y);
    const double v = (double)f.x;
    a[0] += v;
    if (yy == 0) { a[1] += v; a[2] += 1.0; }
    if (yy == 1) { a[3] += v; a[4] += 1.0; }
    a[5] += (double)f.y;
    a[6] += (double)f.z;
    a[7] += (double)f.w;
  }
  __shared__ double red[8][kBootThreads / 32];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    double v = a[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (l == 0) red[k][w] = v;
  }
  __syncthreads();
  if (threadIdx.x < 8) {
    double t = 0;
    for (int i = 0; i < kBootThreads / 32; ++i) t += red[threadIdx.x][i];
    atomicAdd(&sums[(size_t)b * 8 + threadIdx.x], t);
  }
}

__global__ void bootstrap_finalize_kernel(const double* __restrict__ sums, int n_boot, int64_t n, double* __restrict__ out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n_boot) return;
  const double* s = sums + (size_t)b * 8;
  double* o = out + (size_t)b * 6;
  o[0] = s[0] / (double)n;
  o[1] = s[2] > 0 ? s[1] / s[2] : 0.0;   // class absent -> 0.0 (uq_techniques.py:100-101)
  o[2] = s[4] > 0 ? s[3] / s[4] : 0.0;
  o[3] = s[5] / (double)n;
  o[4] = s[6] / (double)n;
  o[5] = s[7] / (double)n;
}

}  // namespace uqk

using namespace uqk;

extern "C" {

int uqk_patient_reduce(const float* variance_dev, const float* entropy_dev, const int32_t* true_label_dev,
                       const int32_t* pred_label_dev, const int64_t* seg_offsets_dev, int64_t n_patients,
                       double* out_dev, void* stream) {
  if (n_patients < 0) { set_error("n_patients must be >= 0"); return UQK_E_BADARG; }
  if (n_patients == 0) return UQK_OK;
  if (!variance_dev || !entropy_dev || !true_label_dev || !pred_label_dev || !seg_offsets_dev || !out_dev) {
    set_error("NULL argument"); return UQK_E_BADARG;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int* d_err = nullptr;
  UQK_CUDA(cudaMallocAsync(&d_err, sizeof(int), s));
  int herr = 0;
  cudaError_t e = cudaMemsetAsync(d_err, 0, sizeof(int), s);
  if (e == cudaSuccess) {
    const int grid = (int)std::min<int64_t>(n_patients, 148 * 8);
    patient_reduce_kernel<<<grid, kPatThreads, 0, s>>>(variance_dev, entropy_dev, true_label_dev, pred_label_dev,
                                                       seg_offsets_dev, n_patients, out_dev, d_err);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    e = cudaGetLastError();
  }
  // SYNCHRONOUS by contract (include/uqk.h): the oversize-segment flag is read back before the call returns
  if (e == cudaSuccess) e = cudaMemcpyAsync(&herr, d_err, sizeof(int), cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  cudaFreeAsync(d_err, s);                      // on every path
  if (e != cudaSuccess) { set_error("uqk_patient_reduce: %s", cudaGetErrorString(e)); return UQK_E_CUDA; }
  if (herr) { set_error("a patient has more than %d windows; segment does not fit the in-kernel sort", kPatMaxSeg); return UQK_E_SHAPE; }
  return UQK_OK;
}

int uqk_bootstrap(const float* variance_dev, const float* total_entropy_dev, const float* expected_entropy_dev,
                  const float* mutual_info_dev, const int32_t* y_dev, const int64_t* idx_dev, int n_boot,
                  int64_t n_windows, double* out_dev, void* stream) {
  if (n_boot < 1 || n_windows < 1) { set_error("need n_boot >= 1 and n_windows >= 1"); return UQK_E_BADARG; }
  if (!variance_dev || !total_entropy_dev || !expected_entropy_dev || !mutual_info_dev || !y_dev || !idx_dev || !out_dev) {
    set_error("NULL argument"); return UQK_E_BADARG;
  }
  if (n_boot > 65535) { set_error("n_boot=%d exceeds 65535 iterations per call (grid dimension); split the call", n_boot); return UQK_E_SHAPE; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  double* sums = nullptr;
  BootRow* rows = nullptr;
  UQK_CUDA(cudaMallocAsync(&sums, sizeof(double) * 8 * n_boot, s));
  cudaError_t e = cudaMallocAsync(reinterpret_cast<void**>(&rows), sizeof(BootRow) * (size_t)n_windows, s);
  if (e == cudaSuccess) e = cudaMemsetAsync(sums, 0, sizeof(double) * 8 * n_boot, s);
  if (e == cudaSuccess) {
    bootstrap_pack_kernel<<<(int)std::min<int64_t>((n_windows + 255) / 256, 148 * 8), 256, 0, s>>>(
        variance_dev, total_entropy_dev, expected_entropy_dev, mutual_info_dev, y_dev, n_windows, rows);
    const int gx = (int)std::max<int64_t>(1, std::min<int64_t>((n_windows + kBootThreads * 8 - 1) / (kBootThreads * 8), 148));
    bootstrap_rows_kernel<<<dim3(gx, n_boot), kBootThreads, 0, s>>>(rows, idx_dev, n_windows, sums);
    bootstrap_finalize_kernel<<<(n_boot + 127) / 128, 128, 0, s>>>(sums, n_boot, n_windows, out_dev);
    g_launches.fetch_add(3, std::memory_order_relaxed);
    e = cudaGetLastError();
  }
  if (rows) cudaFreeAsync(rows, s);             // the stream-ordered scratch is released on every path
  cudaFreeAsync(sums, s);
  if (e != cudaSuccess) { set_error("uqk_bootstrap: %s", cudaGetErrorString(e)); return UQK_E_CUDA; }
  return UQK_OK;
}

}  // extern "C"

// ---- uqk_bootstrap_noalloc: the same gather-means with the per-iteration sums kept in a fixed device array ----------
// Who needs this: a caller that must not allocate device memory while it runs.  Typical cases are a serving process
// that hands the whole GPU to its own arena / a framework's caching allocator at start-up (so cudaMalloc fails or
// fragments), and a latency-critical loop in which cudaMalloc/cudaFree are banned because they may synchronise the
// device.  uqk_bootstrap() takes its (n_boot x 8 doubles) scratch from the stream-ordered allocator; this variant
// keeps it in the file-scope array below instead: 64 KiB that every GPU which loads libuqk.so holds until exit.
namespace uqk {
constexpr int kBootStaticMax = 1024;                        // largest n_boot the fixed scratch can hold
static __device__ __align__(256) double g_boot_scratch[kBootStaticMax * 8];
constexpr int kMaxDevices = 64;
static double* g_boot_scratch_addr[kMaxDevices] = {};       // device address of the array, per device
static std::mutex g_boot_mu;
}  // namespace uqk

extern "C" int uqk_bootstrap_noalloc(const float* variance_dev, const float* total_entropy_dev,
                                     const float* expected_entropy_dev, const float* mutual_info_dev,
                                     const int32_t* y_dev, const int64_t* idx_dev, int n_boot, int64_t n_windows,
                                     double* out_dev, void* stream) {
  if (n_boot < 1 || n_windows < 1) { set_error("need n_boot >= 1 and n_windows >= 1"); return UQK_E_BADARG; }
  if (n_boot > kBootStaticMax) {
    set_error("uqk_bootstrap_noalloc: n_boot=%d exceeds the fixed scratch (max %d iterations per call); "
              "split the call or use uqk_bootstrap", n_boot, kBootStaticMax);
    return UQK_E_SHAPE;
  }
  if (!variance_dev || !total_entropy_dev || !expected_entropy_dev || !mutual_info_dev || !y_dev || !idx_dev || !out_dev) {
    set_error("NULL argument"); return UQK_E_BADARG;
  }
  int dev = 0;
  UQK_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) { set_error("device index %d not supported", dev); return UQK_E_BADARG; }
  double* sums = nullptr;
  {
    // In host code `g_boot_scratch` is only a handle for the symbol, not a device address: resolve it once per
    // device, with that device current (it is: we just asked which one is), and hand the result to the kernels.
    std::lock_guard<std::mutex> lock(g_boot_mu);
    sums = g_boot_scratch_addr[dev];
    if (!sums) {
      UQK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&sums), g_boot_scratch));
      g_boot_scratch_addr[dev] = sums;
    }
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // the array keeps the previous call's sums: clear the part this call accumulates into
  UQK_CUDA(cudaMemsetAsync(sums, 0, sizeof(double) * 8 * (size_t)n_boot, s));
  const int gx = (int)std::max<int64_t>(1, std::min<int64_t>((n_windows + kBootThreads * 8 - 1) / (kBootThreads * 8), 148));
  bootstrap_kernel<<<dim3(gx, n_boot), kBootThreads, 0, s>>>(variance_dev, total_entropy_dev, expected_entropy_dev,
                                                            mutual_info_dev, y_dev, idx_dev, n_windows, sums);
  UQK_LAUNCH_CHECK();
  bootstrap_finalize_kernel<<<(n_boot + 127) / 128, 128, 0, s>>>(sums, n_boot, n_windows, out_dev);
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}
