// Shared host/device declarations for libuqk.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <atomic>

#include "../../include/uqk.h"

namespace uqk {

// ---- architecture (reference models/cnn_baseline_train.py:55-94) -------------------------------------------
constexpr int kL = 60;        // time steps per window
constexpr int kF = 4;         // input features
constexpr int kBlocks = 6;
constexpr int kMaxCh = 256;
constexpr int kDenseIn = 96;
constexpr float kBnEps = 1e-3f;

struct BlockDef { int cin, cout, k, halo; float rate; };
__host__ __device__ constexpr BlockDef block_def(int l) {
  return l == 0 ? BlockDef{4, 128, 7, 3, 0.3f}
       : l == 1 ? BlockDef{128, 192, 5, 2, 0.3f}
       : l == 2 ? BlockDef{192, 224, 3, 1, 0.4f}
       : l == 3 ? BlockDef{224, 96, 7, 3, 0.2f}
       : l == 4 ? BlockDef{96, 256, 9, 4, 0.3f}
                : BlockDef{256, 96, 9, 4, 0.5f};
}
// 16-bit dropout threshold round(rate * 65536): an element is kept iff its 16-bit uniform draw >= thr
// (TF: keep iff u >= rate).  Rates 0.3, 0.3, 0.4, 0.2, 0.3, 0.5 (cnn_baseline_train.py:61,65,70,75,80,86).
__host__ __device__ constexpr uint32_t drop_thr16(int l) {
  return l == 0 ? 19661u : l == 1 ? 19661u : l == 2 ? 26214u : l == 3 ? 13107u : l == 4 ? 19661u : 32768u;
}
// rate exactly 0.5: one random bit per element is an exact Bernoulli(1/2) draw
__host__ __device__ constexpr bool drop_is_half(int l) { return drop_thr16(l) == 32768u; }

// Offsets (in floats) into one member's Keras-order blob.
struct BlobOffsets {
  size_t conv_w[kBlocks], conv_b[kBlocks], gamma[kBlocks], beta[kBlocks], mmean[kBlocks], mvar[kBlocks];
  size_t dense_w, dense_b, total;
};
inline BlobOffsets blob_offsets() {
  BlobOffsets o{};
  size_t p = 0;
  for (int l = 0; l < kBlocks; ++l) {
    BlockDef b = block_def(l);
    o.conv_w[l] = p; p += (size_t)b.k * b.cin * b.cout;
    o.conv_b[l] = p; p += b.cout;
    o.gamma[l] = p; p += b.cout;
    o.beta[l] = p; p += b.cout;
    o.mmean[l] = p; p += b.cout;
    o.mvar[l] = p; p += b.cout;
  }
  o.dense_w = p; p += kDenseIn;
  o.dense_b = p; p += 1;
  o.total = p;
  return o;
}
constexpr size_t kBlobFloats = 853441;
constexpr size_t kBlobStride = 853444;   // device stride between members: keeps every member 16-byte aligned (float4 loads)

// Per-layer channel offsets into the per-member affine tables (sum of Cout = 992).
__host__ __device__ constexpr int ch_off(int l) {
  return l == 0 ? 0 : l == 1 ? 128 : l == 2 ? 320 : l == 3 ? 544 : l == 4 ? 640 : 896;
}
constexpr int kChTotal = 992;

// ---- error plumbing --------------------------------------------------------------------------------------
void set_error(const char* fmt, ...);
extern std::atomic<long long> g_launches;

#define UQK_CUDA(expr)                                                                            \
  do {                                                                                            \
    cudaError_t _e = (expr);                                                                      \
    if (_e != cudaSuccess) {                                                                      \
      ::uqk::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return UQK_E_CUDA;                                                                          \
    }                                                                                             \
  } while (0)

#define UQK_LAUNCH_CHECK()                                                                        \
  do {                                                                                            \
    ::uqk::g_launches.fetch_add(1, std::memory_order_relaxed);                                    \
    cudaError_t _e = cudaGetLastError();                                                          \
    if (_e != cudaSuccess) {                                                                      \
      ::uqk::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
      return UQK_E_CUDA;                                                                          \
    }                                                                                             \
  } while (0)

// ---- Philox4x32-10 (must match oracle/philox.py) ---------------------------------------------------------
struct Philox4 { uint32_t x, y, z, w; };
__device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                 uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    c1 = (uint32_t)p1;
    c3 = (uint32_t)p0;
    c0 = n0;
    c2 = n2;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return Philox4{c0, c1, c2, c3};
}
// ---- dropout keep-flags (must match oracle/philox.py) ---------------------------------------------------------------
// The flags are generated BIT-SLICED, 32 elements per 32-bit word, so that a keep-mask costs a handful of logic
// instructions per 32 elements instead of a compare per element:
//   * the flags of (window, pass, layer, row) are grouped by column half h = (c >> 4) & 1 (the 16-column chunks a
//     fused-kernel epilogue thread owns alternate between two warps) and flag word i = c >> 6; inside the word channel
//     c sits at bit 8*(c & 3) + 4*((c >> 5) & 1) + ((c >> 2) & 3)   (the byte-lane layout of the fused epilogue);
//   * rate 0.5 (layer 6): flag word i = output word i of ONE Philox4x32-10 call, counter
//     (window, pass, layer << 16 | row, h) -- one random bit per element, exact;
//   * other rates: the element's 16-bit uniform U = (hi8 << 8) | lo8 is compared with thr16 = round(rate * 65536);
//     bit k of hi8 is the element's bit of plane word P_k, the planes P_0..P_7 being the 2 x 4 output words of the calls
//     with counter (window, pass, layer << 16 | row, h | i << 1 | q << 3), q = 0, 1 (8 fresh bits per element);
//     lo8 is a tie-break byte SHARED by the elements of (row, h, layer): byte `layer` of the call with counter
//     (window, pass, 6 << 16 | row, h) (words x, y).  Every element's keep probability is exactly 1 - thr16 / 65536;
//     two elements of one thread are correlated only when both tie on the 8 fresh bits (probability 2^-16).
//   keep <=> U >= thr16 is evaluated for 32 elements at once, from the least significant bit up:
//     g = (lo8 >= thr_lo) ? ~0 : 0;   g = thr_hi bit k ? (P_k & g) : (P_k | g),  k = 0..7.
__host__ __device__ constexpr int drop_flag_words(int cout) { return (cout + 63) / 64; }
__host__ __device__ constexpr uint32_t drop_bitpos(int c) { return 8u * (c & 3) + 4u * ((c >> 5) & 1) + ((c >> 2) & 3); }

struct DropTie { uint32_t x, y; };     // tie-break bytes of one (window, pass, row, h): layers 0..3 in x, layer 4 in y
__device__ __forceinline__ DropTie drop_tie(uint32_t c0w, uint32_t pass, uint32_t row, uint32_t h, uint32_t k0, uint32_t k1) {
  const Philox4 r = philox4x32_10(c0w, pass, (6u << 16) | row, h, k0, k1);
  return DropTie{r.x, r.y};
}
__device__ __forceinline__ uint32_t drop_tie_byte(const DropTie& t, int layer) {
  return ((layer < 4 ? (t.x >> (8 * layer)) : t.y)) & 0xFFu;
}
// flag word i of (window, pass, layer, row, h); thr16 < 32768 + 32768 (any rate); for rate 0.5 use drop_half_words
__device__ __forceinline__ uint32_t drop_keep_word(uint32_t c0w, uint32_t pass, uint32_t layer, uint32_t row, uint32_t h,
                                                   uint32_t i, uint32_t thr16, uint32_t tie_byte, uint32_t k0, uint32_t k1) {
  const uint32_t c2 = (layer << 16) | row;
  const Philox4 a = philox4x32_10(c0w, pass, c2, h | (i << 1), k0, k1);
  const Philox4 b = philox4x32_10(c0w, pass, c2, h | (i << 1) | 8u, k0, k1);
  const uint32_t P[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
  const uint32_t thi = thr16 >> 8;
  uint32_t g = tie_byte >= (thr16 & 0xFFu) ? 0xFFFFFFFFu : 0u;
#pragma unroll
  for (int k = 0; k < 8; ++k) g = ((thi >> k) & 1u) ? (P[k] & g) : (P[k] | g);
  return g;
}
__device__ __forceinline__ Philox4 drop_half_words(uint32_t c0w, uint32_t pass, uint32_t layer, uint32_t row, uint32_t h,
                                                   uint32_t k0, uint32_t k1) {
  return philox4x32_10(c0w, pass, (layer << 16) | row, h, k0, k1);
}

// ---- per-window reductions, exact evaluation order of the reference's NumPy code ----------------------------
// H[(1-p, p)] as scipy.stats.entropy(clip(.,1e-10,1-1e-10)) computes it in float32 (uq_techniques.py:35-38)
__device__ __forceinline__ float safe_entropy2(float p) {
  const float eps = 1e-10f;
  float q0 = fminf(fmaxf(1.0f - p, eps), 1.0f);
  float q1 = fminf(fmaxf(p, eps), 1.0f);
  const float s = q0 + q1;
  // scipy renormalises the row; in float32 (1 - p) + p rounds to exactly 1 for almost every p, and x / 1 == x, so the
  // two IEEE divisions run only for the rare rows whose sum is not 1 (same result bit for bit, ~20 % fewer instructions)
  if (s != 1.0f) {
    q0 = q0 / s;
    q1 = q1 / s;
  }
  return (-q0 * logf(q0)) + (-q1 * logf(q1));
}
// log2 entropy of the drivers (analyze_mcd_patient_level.py:113-115), float32 with NumPy-2 weak scalars
__device__ __forceinline__ float entropy_bits(float m) {
  const float eps = 1e-9f;
  return -(m * log2f(m + eps) + (1.0f - m) * log2f((1.0f - m) + eps));
}

struct MetricsOut {
  float* mean_pred; float* pred_variance; float* total_pred_entropy; float* expected_aleatoric_entropy;
  float* mutual_info; float* entropy_bits; int32_t* label;
};
inline MetricsOut to_metrics(const uqk_metrics_out* o) {
  MetricsOut m{};
  if (o) {
    m.mean_pred = o->mean_pred; m.pred_variance = o->pred_variance; m.total_pred_entropy = o->total_pred_entropy;
    m.expected_aleatoric_entropy = o->expected_aleatoric_entropy; m.mutual_info = o->mutual_info;
    m.entropy_bits = o->entropy_bits; m.label = o->label;
  }
  return m;
}

// p points at n_pass probabilities with stride `stride`; sequential float32 accumulation along the pass axis,
// exactly like np.mean/np.var(axis=0) on a C-contiguous (T,N) float32 array.
template <typename Load>
__device__ __forceinline__ void window_metrics(Load load, int n_pass, int64_t n, const MetricsOut& o) {
  float sum = 0.f;
  for (int t = 0; t < n_pass; ++t) sum += load(t);
  const float mean = sum / (float)n_pass;
  float ss = 0.f, hs = 0.f;
  for (int t = 0; t < n_pass; ++t) {
    const float p = load(t);
    const float d = p - mean;
    ss += d * d;
    hs += safe_entropy2(p);
  }
  const float var = ss / (float)n_pass;
  const float total = safe_entropy2(mean);
  const float expected = hs / (float)n_pass;
  if (o.mean_pred) o.mean_pred[n] = mean;
  if (o.pred_variance) o.pred_variance[n] = var;
  if (o.total_pred_entropy) o.total_pred_entropy[n] = total;
  if (o.expected_aleatoric_entropy) o.expected_aleatoric_entropy[n] = expected;
  if (o.mutual_info) o.mutual_info[n] = fmaxf(total - expected, 0.f);
  if (o.entropy_bits) o.entropy_bits[n] = entropy_bits(mean);
  if (o.label) o.label[n] = mean > 0.5f ? 1 : 0;
}

// ---- launchers implemented in the other translation units ---------------------------------------------------
struct Handle;  // api.cu

int launch_reduce(const float* probs, int n_pass, int64_t n, const MetricsOut& o, cudaStream_t s, const int64_t* out_idx = nullptr);
int launch_select_band(const float* mean_pred, int64_t n, float delta, int64_t* idx, unsigned long long* count, cudaStream_t s);
int launch_scatter_probs(const float* sub, const int64_t* idx, int n_pass, int64_t n_sub, int64_t n, float* probs, cudaStream_t s);

}  // namespace uqk
