// Layer-synchronous fp32 forward path (CUDA cores): one kernel per layer with activations in HBM.
//
// This is the EXACT restatement of Keras `model(x, training=True)` (reference uq_techniques.py:22): with
// BatchNormalization in batch-statistics mode a layer cannot be normalised before the post-ReLU output of ALL
// N windows of that layer exists, so layers are separated by grid-wide dependencies.
// It is also the fp32 high-precision path for the deterministic / ensemble forward, and the calibration pass
// that produces the frozen statistics of the fused tensor-core kernel (fused.cu).
//
// Layer semantics: cnn_baseline_train.py:59-94 -- Conv1D(same)+bias+ReLU -> BN -> Dropout, x6, GAP, Dense, sigmoid.
#include "common.cuh"
#include "internal.h"

namespace uqk {

// ---- input cast: (N,60,4) float64/float32 -> float32 --------------------------------------------------------
template <typename T>
__global__ void cast_input_kernel(const T* __restrict__ x, float* __restrict__ out, int64_t n_elems) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_elems; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = (float)x[i];
}

// rows idx[i] of (N,60,4) float64/float32 -> float32 (n_idx,60,4): the windows a guard-band re-evaluation visits
template <typename T>
__global__ void gather_input_kernel(const T* __restrict__ x, const int64_t* __restrict__ idx, float* __restrict__ out, int64_t n_idx) {
  const int64_t total = n_idx * kL * kF;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t w = i / (kL * kF);
    out[i] = (float)x[(size_t)__ldg(idx + w) * (kL * kF) + (i - w * (kL * kF))];
  }
}

int ls_gather_input(const void* x, int x_dtype, const int64_t* idx, float* out, int64_t n_idx, cudaStream_t s) {
  const int64_t ne = n_idx * kL * kF;
  if (ne == 0) return UQK_OK;
  const int grid = (int)std::min<int64_t>((ne + 255) / 256, 148 * 16);
  if (x_dtype == UQK_F64) gather_input_kernel<double><<<grid, 256, 0, s>>>((const double*)x, idx, out, n_idx);
  else gather_input_kernel<float><<<grid, 256, 0, s>>>((const float*)x, idx, out, n_idx);
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}

// ---- Conv1D(same) + bias + ReLU, optional per-channel sum / sum-of-squares for batch-statistic BN -----------
// One block per window, 128 threads = 8 channel quads x 8 row octets x 2 channel tiles: a thread owns 8 rows x 4 channels
// of a 32-channel tile, two tiles (64 channels) are in flight per iteration.  The window (+ zero halo) is staged in
// shared memory TRANSPOSED ([channel][row], odd row pitch), so that for one input channel a thread loads its K + 7
// consecutive rows once and reuses them for all K taps from registers: K x 32 FMAs per (K + 7) shared loads + K
// weight loads.
constexpr int kConvThreads = 128;
constexpr double kStatScale = 16777216.0;   // 2^24 fixed point of the batch-statistic accumulators

template <int K>
__host__ __device__ constexpr int conv_row_pitch() { return ((64 + K - 1) | 1) + 2; }   // odd: conflict-free transposed staging

template <int CIN, int COUT, int K>
__global__ void __launch_bounds__(kConvThreads)
conv_relu_kernel(const float* __restrict__ in, const float* __restrict__ w, const float* __restrict__ bias,
                 float* __restrict__ out, double* __restrict__ stats, int64_t n) {
  constexpr int H = (K - 1) / 2;
  constexpr int RP = conv_row_pitch<K>();
  constexpr int NT = COUT / 32;              // 32-channel tiles
  extern __shared__ float smem[];
  float* xs = smem;                          // [CIN][RP]: xs[c][r] = x[r - H][c], zero outside the window
  // per-channel (sum, sum of squares) of the block, 2^-24 fixed point: integer accumulation is associative, so the
  // batch statistics -- and everything computed from them -- are bit-reproducible run to run and across GPUs
  unsigned long long* ssum = reinterpret_cast<unsigned long long*>(xs + ((CIN * RP + 1) & ~1));   // [64][2]
  const int tx = threadIdx.x & 7;            // channel quad within the 32-channel tile
  const int ro = (threadIdx.x >> 3) & 7;     // row octet: rows 8*ro .. 8*ro+7
  const int th = threadIdx.x >> 6;           // which of the two tiles in flight

  for (int64_t win = blockIdx.x; win < n; win += gridDim.x) {
    __syncthreads();
    for (int i = threadIdx.x; i < RP * CIN; i += kConvThreads) {
      const int r = i / CIN, c = i - r * CIN;
      const int src = r - H;
      xs[c * RP + r] = (src >= 0 && src < kL) ? __ldg(in + ((size_t)win * kL + src) * CIN + c) : 0.f;
    }
    __syncthreads();
    for (int ct0 = 0; ct0 < NT; ct0 += 2) {
      const int ct = ct0 + th;
      const bool active = ct < NT;
      const int ch0 = (active ? ct : 0) * 32 + tx * 4;
      float acc[8][4];
      const float4 b4 = __ldg(reinterpret_cast<const float4*>(bias + ch0));
#pragma unroll
      for (int i = 0; i < 8; ++i) { acc[i][0] = b4.x; acc[i][1] = b4.y; acc[i][2] = b4.z; acc[i][3] = b4.w; }
      if (active) {
#pragma unroll 2
        for (int c = 0; c < CIN; ++c) {
          float xw[K + 7];
          const float* xr = xs + c * RP + 8 * ro;
#pragma unroll
          for (int i = 0; i < K + 7; ++i) xw[i] = xr[i];
          const float* wr = w + (size_t)c * COUT + ch0;
#pragma unroll
          for (int j = 0; j < K; ++j) {
            const float4 wv = __ldg(reinterpret_cast<const float4*>(wr + (size_t)j * CIN * COUT));
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              acc[i][0] = fmaf(xw[i + j], wv.x, acc[i][0]);
              acc[i][1] = fmaf(xw[i + j], wv.y, acc[i][1]);
              acc[i][2] = fmaf(xw[i + j], wv.z, acc[i][2]);
              acc[i][3] = fmaf(xw[i + j], wv.w, acc[i][3]);
            }
          }
        }
      }
      float s[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
      if (active) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int row = 8 * ro + i;
          if (row < kL) {
            float4 v;
            v.x = fmaxf(acc[i][0], 0.f); v.y = fmaxf(acc[i][1], 0.f);
            v.z = fmaxf(acc[i][2], 0.f); v.w = fmaxf(acc[i][3], 0.f);
            *reinterpret_cast<float4*>(out + ((size_t)win * kL + row) * COUT + ch0) = v;
            s[0] += v.x; s[1] += v.y; s[2] += v.z; s[3] += v.w;
            s2[0] += v.x * v.x; s2[1] += v.y * v.y; s2[2] += v.z * v.z; s2[3] += v.w * v.w;
          }
        }
      }
      if (stats != nullptr) {
        ssum[threadIdx.x] = 0ull;
        __syncthreads();
        if (active) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            atomicAdd(&ssum[((th * 32) + tx * 4 + q) * 2 + 0], (unsigned long long)__double2ll_rn((double)s[q] * kStatScale));
            atomicAdd(&ssum[((th * 32) + tx * 4 + q) * 2 + 1], (unsigned long long)__double2ll_rn((double)s2[q] * kStatScale));
          }
        }
        __syncthreads();
        if (ct0 * 32 + (int)(threadIdx.x >> 1) < COUT)
          atomicAdd(reinterpret_cast<unsigned long long*>(stats) + (size_t)(ct0 * 32) * 2 + threadIdx.x, ssum[threadIdx.x]);
        __syncthreads();
      }
    }
  }
}

// ---- batch statistics -> affine (scale, shift); optionally accumulate mean/var for calibration ----------------
// y = gamma * (x - mu) / sqrt(var + eps) + beta, then dropout scale 1/(1-rate) folded in when inv_keep != 1.
__global__ void bn_finalize_kernel(const double* __restrict__ stats, double count, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, float inv_keep, int cout,
                                   float* __restrict__ scale, float* __restrict__ shift,
                                   double* __restrict__ accum_mean, double* __restrict__ accum_var, int* __restrict__ err_flag) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cout) return;
  const long long* q = reinterpret_cast<const long long*>(stats);
  // post-ReLU sums are non-negative: a set sign bit means the 2^-24 fixed point wrapped
  if ((q[2 * c] < 0 || q[2 * c + 1] < 0) && err_flag) *reinterpret_cast<volatile int*>(err_flag) = 50;
  const double mu = (double)q[2 * c] / kStatScale / count;
  double var = (double)q[2 * c + 1] / kStatScale / count - mu * mu;  // biased variance (Keras training mode)
  if (var < 0) var = 0;
  const float inv = (float)((double)gamma[c] / sqrt(var + (double)kBnEps));
  scale[c] = inv * inv_keep;
  shift[c] = (beta[c] - (float)mu * inv) * inv_keep;
  if (accum_mean) { accum_mean[c] += mu; accum_var[c] += var; }
}

// affine from given (mean, var) arrays: moving statistics (inference) or frozen calibrated statistics
__global__ void bn_affine_kernel(const float* __restrict__ mean, const float* __restrict__ var,
                                 const float* __restrict__ gamma, const float* __restrict__ beta, float inv_keep,
                                 int cout, float* __restrict__ scale, float* __restrict__ shift) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cout) return;
  const float inv = (float)((double)gamma[c] / sqrt((double)var[c] + (double)kBnEps));
  scale[c] = inv * inv_keep;
  shift[c] = (beta[c] - mean[c] * inv) * inv_keep;
}

// ---- BN affine + dropout, in place ----------------------------------------------------------------------------
// One thread per (window, row, 8-channel chunk); the keep-flags come from the bit-sliced generator of common.cuh
// (the thread regenerates the flag word its 8 channels live in).
__global__ void bn_dropout_kernel(float* __restrict__ act, const float* __restrict__ scale,
                                  const float* __restrict__ shift, int cout, int64_t n, int layer, int dropout,
                                  uint32_t k0, uint32_t k1, uint32_t pass, int64_t window_base,
                                  const uint32_t* __restrict__ mask_bits /* (N,6,60,8) for this pass or null */,
                                  const int64_t* __restrict__ win_ids /* position of row i in the call, or null = i */) {
  const int nq = cout / 8;
  const int64_t total = n * kL * nq;
  const uint32_t thr16 = drop_thr16(layer);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int q = (int)(i % nq);
    const int64_t wr = i / nq;
    const int row = (int)(wr % kL);
    const int64_t win = win_ids ? __ldg(win_ids + wr / kL) : wr / kL;    // the masks are keyed by the window's position in the call
    float4* p = reinterpret_cast<float4*>(act + (size_t)wr * cout + q * 8);
    float4 a = p[0], b = p[1];
    float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    uint32_t keep = 0xFFu;
    if (dropout) {
      if (mask_bits) {
        const uint32_t word = __ldg(mask_bits + (((size_t)win * kBlocks + layer) * kL + row) * 8 + (q >> 2));
        keep = (word >> ((q & 3) * 8)) & 0xFFu;
      } else {
        const uint32_t c0w = (uint32_t)(window_base + win), h = (uint32_t)(q >> 1) & 1u, wi = (uint32_t)(q >> 3);
        uint32_t word;
        if (thr16 == 32768u) {
          const Philox4 r = drop_half_words(c0w, pass, (uint32_t)layer, (uint32_t)row, h, k0, k1);
          word = wi == 0 ? r.x : wi == 1 ? r.y : wi == 2 ? r.z : r.w;
        } else {
          const DropTie tie = drop_tie(c0w, pass, (uint32_t)row, h, k0, k1);
          word = drop_keep_word(c0w, pass, (uint32_t)layer, (uint32_t)row, h, wi, thr16, drop_tie_byte(tie, layer), k0, k1);
        }
        keep = 0;
#pragma unroll
        for (int e = 0; e < 8; ++e) keep |= ((word >> drop_bitpos(q * 8 + e)) & 1u) << e;
      }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int c = q * 8 + e;
      const float y = fmaf(v[e], __ldg(scale + c), __ldg(shift + c));
      v[e] = ((keep >> e) & 1u) ? y : 0.f;
    }
    p[0] = make_float4(v[0], v[1], v[2], v[3]);
    p[1] = make_float4(v[4], v[5], v[6], v[7]);
  }
}

// ---- GlobalAveragePooling1D + Dense(96->1) + sigmoid: one warp per window ------------------------------------
__global__ void head_kernel(const float* __restrict__ act, const float* __restrict__ wd, const float* __restrict__ bd,
                            float* __restrict__ probs, int64_t n) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t win = warp; win < n; win += nwarps) {
    const float* a = act + (size_t)win * kL * kDenseIn;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f;
    for (int r = 0; r < kL; ++r) {
      s0 += a[r * kDenseIn + lane];
      s1 += a[r * kDenseIn + lane + 32];
      s2 += a[r * kDenseIn + lane + 64];
    }
    const float inv = 1.0f / (float)kL;
    float z = (s0 * inv) * __ldg(wd + lane) + (s1 * inv) * __ldg(wd + lane + 32) + (s2 * inv) * __ldg(wd + lane + 64);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) z += __shfl_xor_sync(0xffffffffu, z, o);
    if (lane == 0) probs[win] = 1.0f / (1.0f + expf(-(z + __ldg(bd))));
  }
}

// ---- host orchestration -------------------------------------------------------------------------------------
template <int CIN, int COUT, int K>
static int launch_conv(const float* in, const float* w, const float* b, float* out, double* stats, int64_t n,
                       cudaStream_t s) {
  const size_t smem = ((size_t)CIN * conv_row_pitch<K>() + 2 + 256) * sizeof(float);
  if (smem > 48 * 1024)
    UQK_CUDA(cudaFuncSetAttribute(conv_relu_kernel<CIN, COUT, K>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)smem));
  if (n == 0) return UQK_OK;   // empty shard of a multi-GPU call: nothing to convolve, statistics stay zero
  const int grid = (int)std::min<int64_t>(n, 148 * 64);
  conv_relu_kernel<CIN, COUT, K><<<grid, kConvThreads, smem, s>>>(in, w, b, out, stats, n);
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}

static int launch_conv_layer(int l, const float* in, const float* w, const float* b, float* out, double* stats,
                             int64_t n, cudaStream_t s) {
  switch (l) {
    case 0: return launch_conv<4, 128, 7>(in, w, b, out, stats, n, s);
    case 1: return launch_conv<128, 192, 5>(in, w, b, out, stats, n, s);
    case 2: return launch_conv<192, 224, 3>(in, w, b, out, stats, n, s);
    case 3: return launch_conv<224, 96, 7>(in, w, b, out, stats, n, s);
    case 4: return launch_conv<96, 256, 9>(in, w, b, out, stats, n, s);
    default: return launch_conv<256, 96, 9>(in, w, b, out, stats, n, s);
  }
}

int ls_cast_input(const void* x, int x_dtype, float* out, int64_t n_windows, cudaStream_t s) {
  const int64_t ne = n_windows * kL * kF;
  if (ne == 0) return UQK_OK;
  const int grid = (int)std::min<int64_t>((ne + 255) / 256, 148 * 16);
  if (x_dtype == UQK_F64) cast_input_kernel<double><<<grid, 256, 0, s>>>((const double*)x, out, ne);
  else cast_input_kernel<float><<<grid, 256, 0, s>>>((const float*)x, out, ne);
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}

// One forward pass of `n` windows (already float32, (n,60,4)) through the six blocks + head.
// p: see LsPassParams in internal.h.
int ls_forward_pass(const LsPassParams& p, cudaStream_t s) {
  const BlobOffsets off = blob_offsets();
  const float* blob = p.blob_dev;
  const float* in = p.x_f32;
  float* bufs[2] = {p.act0, p.act1};
  int cur = 0;
  for (int l = 0; l < kBlocks; ++l) {
    const BlockDef b = block_def(l);
    float* out = bufs[cur];
    double* stats = nullptr;
    if (p.bn_mode == UQK_BN_BATCH) {
      stats = p.stats_scratch + (size_t)l * kMaxCh * 2;
      UQK_CUDA(cudaMemsetAsync(stats, 0, sizeof(double) * 2 * b.cout, s));
    }
    int rc = launch_conv_layer(l, in, blob + off.conv_w[l], blob + off.conv_b[l], out, stats, p.n, s);
    if (rc) return rc;
    int64_t stat_windows = p.n;
    if (stats != nullptr && p.exchange != nullptr) {
      // sum the fixed-point (sum, sum of squares) over the ranks that hold the other windows of this call
      if (p.exchange(p.exchange_user, reinterpret_cast<long long*>(stats), 2 * b.cout, static_cast<void*>(s)) != 0) {
        set_error("batch-statistics exchange callback failed at layer %d, pass %d", l, p.pass);
        return UQK_E_STATE;
      }
      stat_windows = p.global_n;
    }
    const float inv_keep = p.dropout ? 1.0f / (1.0f - b.rate) : 1.0f;
    const float* scale = p.affine_scratch + (size_t)l * kMaxCh * 2;
    const float* shift = scale + kMaxCh;
    const int tb = 128, gb = (b.cout + tb - 1) / tb;
    if (p.affine_tab) {
      // the affine of this pass was fixed elsewhere (per-pass batch statistics of the swept tensor-core run)
      scale = p.affine_tab + kChTotal + ch_off(l);
      shift = p.affine_tab + 2 * kChTotal + ch_off(l);
    } else if (p.bn_mode == UQK_BN_BATCH) {
      bn_finalize_kernel<<<gb, tb, 0, s>>>(stats, (double)stat_windows * kL, blob + off.gamma[l], blob + off.beta[l], inv_keep,
                                           b.cout, const_cast<float*>(scale), const_cast<float*>(shift),
                                           p.accum_mean ? p.accum_mean + (size_t)l * kMaxCh : nullptr,
                                           p.accum_var ? p.accum_var + (size_t)l * kMaxCh : nullptr, p.err_flag);
    } else {
      const float* mean = p.bn_mode == UQK_BN_CALIBRATED ? p.frozen_mean + (size_t)l * kMaxCh : blob + off.mmean[l];
      const float* var = p.bn_mode == UQK_BN_CALIBRATED ? p.frozen_var + (size_t)l * kMaxCh : blob + off.mvar[l];
      bn_affine_kernel<<<gb, tb, 0, s>>>(mean, var, blob + off.gamma[l], blob + off.beta[l], inv_keep, b.cout,
                                         const_cast<float*>(scale), const_cast<float*>(shift));
    }
    if (!p.affine_tab) UQK_LAUNCH_CHECK();
    if (p.n > 0) {
      const int64_t total = p.n * kL * (b.cout / 8);
      const int grid = (int)std::min<int64_t>((total + 255) / 256, 148 * 32);
      bn_dropout_kernel<<<grid, 256, 0, s>>>(out, scale, shift, b.cout, p.n, l, p.dropout ? 1 : 0,
                                             (uint32_t)(p.seed & 0xFFFFFFFFull), (uint32_t)(p.seed >> 32),
                                             (uint32_t)p.pass, p.window_base, p.mask_bits, p.win_ids);
      UQK_LAUNCH_CHECK();
    }
    in = out;
    cur ^= 1;
  }
  if (p.n > 0) {
    const int64_t threads = p.n * 32;
    const int grid = (int)std::min<int64_t>((threads + 255) / 256, 148 * 32);
    head_kernel<<<grid, 256, 0, s>>>(in, blob + off.dense_w, blob + off.dense_b, p.probs_out, p.n);
    UQK_LAUNCH_CHECK();
  }
  return UQK_OK;
}

}  // namespace uqk
