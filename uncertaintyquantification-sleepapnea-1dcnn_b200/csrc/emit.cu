// N2: the per-window detail CSV that follows the hot path.
//
// The reference builds a DataFrame of seven columns and calls `detailed_results_df.to_csv(path, index=False)`
// (uncertainty_quantification/analyze_mcd_patient_level.py:134-152; analyze_de_patient_level.py has the same block):
//   Patient_ID (int64), Window_Index (0..N-1), True_Label (int), Predicted_Label (int),
//   Predicted_Probability, Predictive_Variance, Predictive_Entropy (float32, straight from NumPy).
// pandas formats 1.4 M rows x 7 columns in a single Python-driven loop (seconds); this writer produces the SAME BYTES
// with a pool of host threads.  Host-only code: no device work, so it is tested on the CPU against pandas itself.
//
// Float text = NumPy's repr of a float32 scalar, which is what pandas emits for float32 columns: the shortest digit
// string that round-trips (std::to_chars gives the same digits), positional notation for 1e-4 <= |x| < 1e6 (compared
// in double, so float32(1e-4) = 9.9999997e-05 prints as "1e-04") with at least one digit after the point, otherwise
// scientific with a sign and at least two exponent digits ("1e-07", "1.6777216e+07").
#include "common.cuh"
#include "internal.h"

#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

namespace uqk {

// Appends NumPy's repr of `v` to out; returns the number of characters written (<= 24).
static int format_f32_numpy(float v, char* out) {
  char* p = out;
  if (std::isnan(v)) { memcpy(p, "nan", 3); return 3; }          // pandas writes an empty field for NaN (na_rep='')
  if (std::isinf(v)) {
    if (v < 0) *p++ = '-';
    memcpy(p, "inf", 3);
    return (int)(p - out) + 3;
  }
  if (v == 0.0f) {
    if (std::signbit(v)) *p++ = '-';
    memcpy(p, "0.0", 3);
    return (int)(p - out) + 3;
  }
  // shortest round-trip digits in scientific form: d[.ddd]e[+-]XX
  char sci[32];
  auto r = std::to_chars(sci, sci + sizeof(sci), v, std::chars_format::scientific);
  *r.ptr = '\0';
  const char* s = sci;
  if (*s == '-') { *p++ = '-'; ++s; }
  const char* e = strchr(s, 'e');
  char digits[16];
  int nd = 0;
  for (const char* q = s; q < e; ++q)
    if (*q != '.') digits[nd++] = *q;
  const int exp10 = atoi(e + 1);                                  // value = d.ddd * 10^exp10
  const double av = std::fabs((double)v);
  if (av >= 1e-4 && av < 1e6) {                                   // NumPy's float32 rule, on the value itself
    if (exp10 < 0) {                                              // 0.000ddd
      *p++ = '0'; *p++ = '.';
      for (int i = 0; i < -exp10 - 1; ++i) *p++ = '0';
      memcpy(p, digits, nd); p += nd;
    } else {
      const int int_digits = exp10 + 1;
      for (int i = 0; i < int_digits; ++i) *p++ = i < nd ? digits[i] : '0';
      *p++ = '.';
      if (nd > int_digits) { memcpy(p, digits + int_digits, nd - int_digits); p += nd - int_digits; }
      else *p++ = '0';
    }
  } else {
    *p++ = digits[0];
    if (nd > 1) { *p++ = '.'; memcpy(p, digits + 1, nd - 1); p += nd - 1; }
    *p++ = 'e';
    *p++ = exp10 < 0 ? '-' : '+';
    const int a = exp10 < 0 ? -exp10 : exp10;
    if (a < 10) { *p++ = '0'; *p++ = (char)('0' + a); }
    else { auto r2 = std::to_chars(p, p + 4, a); p = r2.ptr; }
  }
  return (int)(p - out);
}

static inline char* put_i64(char* p, long long v) {
  auto r = std::to_chars(p, p + 24, v);
  return r.ptr;
}

static void format_rows(int64_t r0, int64_t r1, const int64_t* pid, int64_t index_base, const int64_t* y,
                        const int32_t* label, const float* prob, const float* var, const float* ent, std::string* out) {
  out->resize((size_t)(r1 - r0) * 128);
  char* base = &(*out)[0];
  char* p = base;
  for (int64_t i = r0; i < r1; ++i) {
    p = put_i64(p, pid[i]); *p++ = ',';
    p = put_i64(p, index_base + i); *p++ = ',';
    p = put_i64(p, y[i]); *p++ = ',';
    p = put_i64(p, label[i]); *p++ = ',';
    const float f[3] = {prob[i], var[i], ent[i]};
    for (int k = 0; k < 3; ++k) {
      if (!std::isnan(f[k])) p += format_f32_numpy(f[k], p);
      *p++ = k < 2 ? ',' : '\n';
    }
  }
  out->resize((size_t)(p - base));
}

}  // namespace uqk

using namespace uqk;

extern "C" int uqk_write_detail_csv(const char* path, int64_t n, const int64_t* patient_id, int64_t window_index_base,
                                    const int64_t* true_label, const int32_t* pred_label, const float* prob,
                                    const float* variance, const float* entropy, int write_header, int n_threads) {
  if (!path || n < 0) { set_error("path is NULL or n < 0"); return UQK_E_BADARG; }
  if (n > 0 && (!patient_id || !true_label || !pred_label || !prob || !variance || !entropy)) {
    set_error("NULL column"); return UQK_E_BADARG;
  }
  FILE* f = fopen(path, write_header ? "wb" : "ab");
  if (!f) { set_error("cannot open %s for writing", path); return UQK_E_STATE; }
  if (write_header)
    fputs("Patient_ID,Window_Index,True_Label,Predicted_Label,Predicted_Probability,Predictive_Variance,Predictive_Entropy\n", f);
  if (n_threads < 1) n_threads = (int)std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
  const int64_t chunk = 1 << 16;                          // rows per task
  bool ok = true;
  for (int64_t base = 0; base < n && ok; base += chunk * n_threads) {
    const int tasks = (int)std::min<int64_t>(n_threads, (n - base + chunk - 1) / chunk);
    std::vector<std::string> bufs(tasks);
    std::vector<std::thread> th;
    for (int t = 0; t < tasks; ++t) {
      const int64_t r0 = base + (int64_t)t * chunk, r1 = std::min(n, r0 + chunk);
      th.emplace_back(format_rows, r0, r1, patient_id, window_index_base, true_label, pred_label, prob, variance, entropy,
                      &bufs[t]);
    }
    for (auto& t : th) t.join();
    for (auto& b : bufs) ok = ok && fwrite(b.data(), 1, b.size(), f) == b.size();
  }
  if (fclose(f) != 0) ok = false;
  if (!ok) { set_error("short write to %s", path); return UQK_E_STATE; }
  return UQK_OK;
}

// =====================================================================================================================
// GPU twin of analyze_window_level_uncertainty.py (reference lines 30-64): overall accuracy, describe() of
// Predictive_Entropy / Predictive_Variance for correctly and incorrectly classified windows, and the 10-bin
// accuracy-vs-entropy table (pd.cut over np.linspace(min, max + 1e-9, bins + 1), right=False).
// All statistics are float64 over the float32 per-window arrays the hot path left on the device; quartiles use
// pandas' linear interpolation between the two neighbouring order statistics, found by an 8-bit radix select that
// serves all 24 (class, metric, rank) queries in the same four passes over the data.
// =====================================================================================================================
namespace uqk {

constexpr int kSumQueries = 24;      // 2 classes x 2 metrics x 3 quantiles x (lower, upper) order statistic
constexpr int kSumMaxBins = 64;

struct SummaryScratch {
  unsigned long long count[2];                 // per class (0 = correct, 1 = incorrect)
  double sum[2][2], ssd[2][2];                 // [class][metric]: sum, sum of squared deviations from the mean
  unsigned int kmin[2][2], kmax[2][2];         // order-preserving keys of min / max
  double mean[2][2];
  double edges[kSumMaxBins + 1];
  unsigned long long bin_count[kSumMaxBins], bin_correct[kSumMaxBins];
  // radix select state per query
  unsigned int q_prefix[kSumQueries];
  long long q_rank[kSumQueries];               // remaining rank inside the current prefix; -1 = query inactive
  unsigned int hist[kSumQueries][256];
};

__device__ __forceinline__ unsigned int f32_key(float x) {
  const unsigned int u = __float_as_uint(x);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_f32(unsigned int k) {
  return __uint_as_float((k & 0x80000000u) ? (k ^ 0x80000000u) : ~k);
}

__global__ void summary_init_kernel(SummaryScratch* s) {     // after the memset: minima start at the largest key
  if (threadIdx.x < 4) s->kmin[threadIdx.x >> 1][threadIdx.x & 1] = 0xFFFFFFFFu;
}

// pass 1: per class count, per (class, metric) sum / min / max
__global__ void summary_pass1_kernel(const float* __restrict__ ent, const float* __restrict__ var,
                                     const int32_t* __restrict__ y, const int32_t* __restrict__ yhat, int64_t n,
                                     SummaryScratch* s) {
  unsigned long long cnt[2] = {0, 0};
  double sum[2][2] = {{0, 0}, {0, 0}};
  unsigned int kmin[2][2] = {{~0u, ~0u}, {~0u, ~0u}}, kmax[2][2] = {{0, 0}, {0, 0}};
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (y[i] == yhat[i]) ? 0 : 1;
    const float v[2] = {ent[i], var[i]};
    ++cnt[c];
#pragma unroll
    for (int m = 0; m < 2; ++m) {
      sum[c][m] += (double)v[m];
      const unsigned int k = f32_key(v[m]);
      kmin[c][m] = min(kmin[c][m], k);
      kmax[c][m] = max(kmax[c][m], k);
    }
  }
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    for (int o = 16; o > 0; o >>= 1) cnt[c] += __shfl_xor_sync(0xffffffffu, cnt[c], o);
#pragma unroll
    for (int m = 0; m < 2; ++m) {
      for (int o = 16; o > 0; o >>= 1) {
        sum[c][m] += __shfl_xor_sync(0xffffffffu, sum[c][m], o);
        kmin[c][m] = min(kmin[c][m], __shfl_xor_sync(0xffffffffu, kmin[c][m], o));
        kmax[c][m] = max(kmax[c][m], __shfl_xor_sync(0xffffffffu, kmax[c][m], o));
      }
    }
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      if (cnt[c] == 0) continue;
      atomicAdd(&s->count[c], cnt[c]);
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        atomicAdd(&s->sum[c][m], sum[c][m]);
        atomicMin(&s->kmin[c][m], kmin[c][m]);
        atomicMax(&s->kmax[c][m], kmax[c][m]);
      }
    }
  }
}

// one thread: means, bin edges (np.linspace semantics: start + i*step without fused multiply-add, last edge = stop),
// and the order-statistic ranks of the quartile queries ((n-1)*q, pandas' linear interpolation)
__global__ void summary_prepare_kernel(SummaryScratch* s, int n_bins) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  for (int c = 0; c < 2; ++c)
    for (int m = 0; m < 2; ++m) s->mean[c][m] = s->count[c] ? s->sum[c][m] / (double)s->count[c] : 0.0;
  unsigned int kmin = ~0u, kmax = 0u;
  for (int c = 0; c < 2; ++c)
    if (s->count[c]) { kmin = min(kmin, s->kmin[c][0]); kmax = max(kmax, s->kmax[c][0]); }
  const double lo = (double)key_f32(kmin), hi = __dadd_rn((double)key_f32(kmax), 1e-9);
  const double step = __ddiv_rn(__dadd_rn(hi, -lo), (double)n_bins);
  for (int i = 0; i < n_bins; ++i) s->edges[i] = __dadd_rn(__dmul_rn((double)i, step), lo);
  s->edges[n_bins] = hi;
  const double qs[3] = {0.25, 0.5, 0.75};
  for (int c = 0; c < 2; ++c)
    for (int m = 0; m < 2; ++m)
      for (int q = 0; q < 3; ++q) {
        const int j = ((c * 2 + m) * 3 + q) * 2;
        if (s->count[c] == 0) { s->q_rank[j] = s->q_rank[j + 1] = -1; continue; }
        const double pos = (double)(s->count[c] - 1) * qs[q];
        const long long lo_r = (long long)floor(pos);
        s->q_rank[j] = lo_r;
        s->q_rank[j + 1] = min(lo_r + 1, (long long)s->count[c] - 1);
      }
}

// pass 2: squared deviations (two-pass variance) and the entropy bins
__global__ void summary_pass2_kernel(const float* __restrict__ ent, const float* __restrict__ var,
                                     const int32_t* __restrict__ y, const int32_t* __restrict__ yhat, int64_t n,
                                     int n_bins, SummaryScratch* s) {
  __shared__ unsigned int sb[2 * kSumMaxBins];
  __shared__ double sedges[kSumMaxBins + 1];
  for (int i = threadIdx.x; i < 2 * kSumMaxBins; i += blockDim.x) sb[i] = 0u;
  for (int i = threadIdx.x; i <= n_bins; i += blockDim.x) sedges[i] = s->edges[i];
  __syncthreads();
  double ssd[2][2] = {{0, 0}, {0, 0}};
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (y[i] == yhat[i]) ? 0 : 1;
    const double e = (double)ent[i], v = (double)var[i];
    const double de = e - s->mean[c][0], dv = v - s->mean[c][1];
    ssd[c][0] += de * de;
    ssd[c][1] += dv * dv;
    int b = -1;                                           // searchsorted(edges, e, side='right') - 1
    for (int k = 0; k <= n_bins; ++k) b += (sedges[k] <= e) ? 1 : 0;
    if (b >= 0 && b < n_bins) {
      atomicAdd(&sb[b], 1u);
      if (c == 0) atomicAdd(&sb[kSumMaxBins + b], 1u);
    }
  }
#pragma unroll
  for (int c = 0; c < 2; ++c)
#pragma unroll
    for (int m = 0; m < 2; ++m) {
      for (int o = 16; o > 0; o >>= 1) ssd[c][m] += __shfl_xor_sync(0xffffffffu, ssd[c][m], o);
      if ((threadIdx.x & 31) == 0 && ssd[c][m] != 0.0) atomicAdd(&s->ssd[c][m], ssd[c][m]);
    }
  __syncthreads();
  for (int i = threadIdx.x; i < n_bins; i += blockDim.x) {
    if (sb[i]) atomicAdd(&s->bin_count[i], (unsigned long long)sb[i]);
    if (sb[kSumMaxBins + i]) atomicAdd(&s->bin_correct[i], (unsigned long long)sb[kSumMaxBins + i]);
  }
}

// radix-select pass over bits [shift, shift+8): histogram of the next digit among keys matching each query's prefix
__global__ void summary_radix_hist_kernel(const float* __restrict__ ent, const float* __restrict__ var,
                                          const int32_t* __restrict__ y, const int32_t* __restrict__ yhat, int64_t n,
                                          int shift, SummaryScratch* s) {
  __shared__ unsigned int sh[kSumQueries][256];
  __shared__ unsigned int spre[kSumQueries];
  __shared__ int sact[kSumQueries];
  for (int i = threadIdx.x; i < kSumQueries * 256; i += blockDim.x) (&sh[0][0])[i] = 0u;
  if (threadIdx.x < kSumQueries) { spre[threadIdx.x] = s->q_prefix[threadIdx.x]; sact[threadIdx.x] = s->q_rank[threadIdx.x] >= 0; }
  __syncthreads();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (y[i] == yhat[i]) ? 0 : 1;
    const unsigned int key[2] = {f32_key(ent[i]), f32_key(var[i])};
#pragma unroll
    for (int m = 0; m < 2; ++m) {
      const unsigned int hi = shift == 24 ? 0u : (key[m] >> (shift + 8));
      const unsigned int digit = (key[m] >> shift) & 255u;
#pragma unroll
      for (int t = 0; t < 6; ++t) {
        const int j = (c * 2 + m) * 6 + t;
        if (sact[j] && (shift == 24 || hi == spre[j])) atomicAdd(&sh[j][digit], 1u);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kSumQueries * 256; i += blockDim.x) {
    const unsigned int v = (&sh[0][0])[i];
    if (v) atomicAdd(&(&s->hist[0][0])[i], v);
  }
}

// one thread per query: pick the digit whose cumulative count crosses the rank, extend the prefix, clear the histogram
__global__ void summary_radix_step_kernel(SummaryScratch* s) {
  const int j = threadIdx.x;
  if (j >= kSumQueries) return;
  long long r = s->q_rank[j];
  if (r >= 0) {
    unsigned int d = 0;
    for (; d < 255u; ++d) {
      const unsigned int h = s->hist[j][d];
      if (r < (long long)h) break;
      r -= h;
    }
    s->q_prefix[j] = (s->q_prefix[j] << 8) | d;
    s->q_rank[j] = r;
  }
  for (int d = 0; d < 256; ++d) s->hist[j][d] = 0u;
}

__global__ void summary_final_kernel(const SummaryScratch* s, int n_bins, double* __restrict__ out) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  const double total = (double)(s->count[0] + s->count[1]);
  out[0] = total;
  out[1] = total > 0 ? (double)s->count[0] / total : nan;
  const double qs[3] = {0.25, 0.5, 0.75};
  for (int c = 0; c < 2; ++c)
    for (int m = 0; m < 2; ++m) {
      double* o = out + 2 + 16 * c + 8 * m;
      const unsigned long long cnt = s->count[c];
      o[0] = (double)cnt;
      if (cnt == 0) { for (int k = 1; k < 8; ++k) o[k] = nan; continue; }
      o[1] = s->mean[c][m];
      o[2] = cnt > 1 ? sqrt(s->ssd[c][m] / (double)(cnt - 1)) : nan;
      o[3] = (double)key_f32(s->kmin[c][m]);
      o[7] = (double)key_f32(s->kmax[c][m]);
      for (int q = 0; q < 3; ++q) {
        const int j = ((c * 2 + m) * 3 + q) * 2;
        const double pos = (double)(cnt - 1) * qs[q];
        const double frac = pos - floor(pos);
        const double a = (double)key_f32(s->q_prefix[j]), b = (double)key_f32(s->q_prefix[j + 1]);
        o[4 + q] = a + (b - a) * frac;
      }
    }
  double* e = out + 34;
  for (int i = 0; i <= n_bins; ++i) e[i] = s->edges[i];
  double* bc = e + n_bins + 1;
  double* ba = bc + n_bins;
  for (int i = 0; i < n_bins; ++i) {
    bc[i] = (double)s->bin_count[i];
    ba[i] = s->bin_count[i] ? (double)s->bin_correct[i] / (double)s->bin_count[i] : nan;
  }
}

}  // namespace uqk

extern "C" int uqk_window_summary(const float* entropy_dev, const float* variance_dev, const int32_t* true_label_dev,
                                  const int32_t* pred_label_dev, int64_t n, int n_bins, double* out_dev, void* stream) {
  if (n < 1 || !entropy_dev || !variance_dev || !true_label_dev || !pred_label_dev || !out_dev) {
    set_error("uqk_window_summary needs n >= 1 and non-NULL arrays"); return UQK_E_BADARG;
  }
  if (n_bins < 1 || n_bins > kSumMaxBins) { set_error("n_bins must be in 1..%d", kSumMaxBins); return UQK_E_BADARG; }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  SummaryScratch* sc = nullptr;
  UQK_CUDA(cudaMallocAsync(reinterpret_cast<void**>(&sc), sizeof(SummaryScratch), s));
  UQK_CUDA(cudaMemsetAsync(sc, 0, sizeof(SummaryScratch), s));
  const int grid = (int)std::min<int64_t>((n + 255) / 256, 148 * 8);
  summary_init_kernel<<<1, 32, 0, s>>>(sc);
  UQK_LAUNCH_CHECK();
  summary_pass1_kernel<<<grid, 256, 0, s>>>(entropy_dev, variance_dev, true_label_dev, pred_label_dev, n, sc);
  UQK_LAUNCH_CHECK();
  summary_prepare_kernel<<<1, 32, 0, s>>>(sc, n_bins);
  UQK_LAUNCH_CHECK();
  summary_pass2_kernel<<<grid, 256, 0, s>>>(entropy_dev, variance_dev, true_label_dev, pred_label_dev, n, n_bins, sc);
  UQK_LAUNCH_CHECK();
  for (int shift = 24; shift >= 0; shift -= 8) {
    summary_radix_hist_kernel<<<grid, 256, 0, s>>>(entropy_dev, variance_dev, true_label_dev, pred_label_dev, n, shift, sc);
    UQK_LAUNCH_CHECK();
    summary_radix_step_kernel<<<1, 32, 0, s>>>(sc);
    UQK_LAUNCH_CHECK();
  }
  summary_final_kernel<<<1, 32, 0, s>>>(sc, n_bins, out_dev);
  UQK_LAUNCH_CHECK();
  UQK_CUDA(cudaFreeAsync(sc, s));
  return UQK_OK;
}
