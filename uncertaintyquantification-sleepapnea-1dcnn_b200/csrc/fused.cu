// K1+K2+K3: fused forward of the whole 1D-CNN on 5th-generation tensor cores (tcgen05 / TMEM), sm_100a only.
//
// What it replaces: the T x `model(x, training=True)` loop of mc_dropout_predict (reference uq_techniques.py:22)
// and the M x `model.predict` loop of deep_ensembles_predict (:29-30), the network being
// cnn_baseline_train.py:55-94, fused with the per-window reductions of uq_evaluation_dist (:69-91).
//
// Mapping (see DESIGN.md for the derivation):
//  * A work UNIT is 2*TILES consecutive windows; a GROUP g of a unit is pass g (MC-Dropout) or member g (ensemble) of
//    those windows.  One CTA owns a unit, so all T (or M) probabilities of a window meet in shared memory and the
//    (T,N) tensor never has to exist in HBM.
//  * A TILE is one UMMA M=128 accumulator = 2 windows x 64 rows (60 time steps + 4 zero rows).  The zero rows are the
//    `same` padding of the NEXT layer: activations live in shared memory as bf16 in the UMMA canonical K-major,
//    no-swizzle layout with all rows of an 8-channel chunk contiguous ("plane"), so tap j of a Conv1D is the same
//    plane addressed j*16 bytes further -- an implicit GEMM with no im2col copy:
//         Y[t, o] = sum_j sum_c X[t + j - h, c] * W[j, c, o]   ==   sum over (j, 16-channel slab) of one tcgen05.mma.
//  * Activations are updated IN PLACE: the accumulator of layer l lives in TMEM until every MMA of layer l has
//    retired (tcgen05.commit -> mbarrier), then the epilogue warps overwrite the layer's input with
//    bias+ReLU -> BN affine -> Philox dropout -> bf16.  One 71.8 KB buffer per tile instead of two.
//  * Weights (1.7 MB bf16 per member, L2 resident) stream through a shared-memory ring with cp.async.bulk (TMA
//    bulk copies, mbarrier complete_tx); they are pre-packed on the host into the exact shared-memory image.
//  * Warp roles: warp 0 = weight producer, then one warp per accumulator = TMEM allocator / MMA issuer (one elected
//    lane), 8 warps per tile = epilogue (TMEM lane quarter = warp_id % 4, two warps per quarter on alternate 16-column
//    chunks).
//  * Kernels (fused_kernels.cuh; one translation unit per family so that they compile in parallel), all with this
//    mapping: fused_forward_kernel<TILES,...> (one CTA per unit, no clusters; fused_single1.cu / fused_single2.cu),
//    fused_pair_kernel (cta_group::2 CTA pairs, one tile per CTA, two CTAs per SM; fused_pair.cu) and
//    fused_pair2_kernel (CTA pairs with two tile-pairs per pair sharing one weight stream -- the default,
//    fused_pair2.cu; its STAT instantiation runs the per-layer sweeps of the exact batch-statistic mode,
//    fused_pair_stat.cu).  This file holds the host side: weight packing, epilogue tables, the exact mode's small
//    kernels and the dispatcher.
#include "fused_kernels.cuh"

namespace uqk {

size_t fused_packed_bytes() { return kPackedBytes; }

static inline uint16_t f32_to_bf16_rne(float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  if ((u & 0x7F800000u) == 0x7F800000u) return (uint16_t)(u >> 16);  // inf / nan
  u += 0x7FFFu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
static inline float bf16_to_f32(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  std::memcpy(&f, &u, 4);
  return f;
}

// Host-side packing of one member into the shared-memory image the kernel streams:
//   layer l, step s, plane p (K elements 8p..8p+7 of the step), output channel n, element kk:
//   dst[(w_off/2) + ((s*2 + p)*cout + n)*8 + kk]
//   layers 2..6: slab i = s / k, tap j = s % k (slab-major: the steps that read the first input channels come first, so
//                a layer can start while the previous layer's epilogue is still writing the later channels),
//                channel c = 16 i + 8 p + kk                                               -> W[j][c][n]
//   layer 1    : tap j = 2 s + p, channel kk (<4)                                        -> W[j][kk][n], else 0
//   bias step (s = nconv): plane 0, kk = 0 -> bf16(bias[n]), kk = 1 -> bf16(bias[n] - hi)  (hi + lo split: the MMA
//   adds the bias with 16 mantissa bits because the A operand of this step is the constant [1, 1, 0, ...] plane)
void fused_pack_member(const float* blob, uint16_t* dst) {
  const BlobOffsets off = blob_offsets();
  for (int l = 0; l < kBlocks; ++l) {
    const FusedLayer fl = fused_layer(l);
    const BlockDef b = block_def(l);
    const float* w = blob + off.conv_w[l];  // (k, cin, cout)
    const float* bias = blob + off.conv_b[l];
    uint16_t* d = dst + fl.w_off / 2;
    for (int s = 0; s < fl.nsteps; ++s)
      for (int p = 0; p < 2; ++p)
        for (int n = 0; n < fl.cout; ++n)
          for (int kk = 0; kk < 8; ++kk) {
            uint16_t o = 0;
            if (s == fl.nconv) {
              if (p == 0 && kk == 0) o = f32_to_bf16_rne(bias[n]);
              if (p == 0 && kk == 1) o = f32_to_bf16_rne(bias[n] - bf16_to_f32(f32_to_bf16_rne(bias[n])));
            } else if (l == 0) {
              const int j = 2 * s + p;
              if (j < b.k && kk < b.cin) o = f32_to_bf16_rne(w[((size_t)j * b.cin + kk) * b.cout + n]);
            } else {
              const int j = s % b.k, i = s / b.k;
              const int c = 16 * i + 8 * p + kk;
              o = f32_to_bf16_rne(w[((size_t)j * b.cin + c) * b.cout + n]);
            }
            d[((size_t)(s * 2 + p) * fl.cout + n) * 8 + kk] = o;
          }
  }
}

// Pair layout (cta_group::2): CTA r of a pair holds output channels [r*N/2, (r+1)*N/2) of every step, i.e. each CTA
// streams HALF of the weights and the tensor cores of the two SMs exchange the halves.  Image = [half 0][half 1], each
// half laid out like the single-CTA image with N/2 rows per plane.
void fused_pack_member_pair(const float* blob, uint16_t* dst) {
  std::vector<uint16_t> full(kPackedBytes / 2);
  fused_pack_member(blob, full.data());
  for (int r = 0; r < 2; ++r)
    for (int l = 0; l < kBlocks; ++l) {
      const FusedLayer fl = fused_layer(l);
      const int nh = fl.cout / 2;
      const uint16_t* src = full.data() + fl.w_off / 2;
      uint16_t* d = dst + (size_t)r * (kPackedBytes / 4) + fl.w_off / 4;
      for (int sp = 0; sp < fl.nsteps * 2; ++sp)
        std::memcpy(d + (size_t)sp * nh * 8, src + ((size_t)sp * fl.cout + (size_t)r * nh) * 8, (size_t)nh * 16);
    }
}

// ---- tables builder ---------------------------------------------------------------------------------------------
__global__ void build_tables_kernel(const float* __restrict__ blob, BlobOffsets off, const float* __restrict__ mean,
                                    const float* __restrict__ var, int fold_dropout, float* __restrict__ table) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < kChTotal) {
    int l = 0;
    while (l < kBlocks - 1 && i >= ch_off(l + 1)) ++l;
    const int c = i - ch_off(l);
    const BlockDef b = block_def(l);
    const float mu = mean ? mean[l * kMaxCh + c] : blob[off.mmean[l] + c];
    const float vr = var ? var[l * kMaxCh + c] : blob[off.mvar[l] + c];
    const float inv_keep = fold_dropout ? 1.0f / (1.0f - b.rate) : 1.0f;
    const float inv = (float)((double)blob[off.gamma[l] + c] / sqrt((double)vr + (double)kBnEps));
    table[i] = blob[off.conv_b[l] + c];
    table[kChTotal + i] = inv * inv_keep;
    table[2 * kChTotal + i] = (blob[off.beta[l] + c] - mu * inv) * inv_keep;
  }
  if (i < kDenseIn) table[3 * kChTotal + i] = blob[off.dense_w + i];
  if (i == 0) table[3 * kChTotal + kDenseIn] = blob[off.dense_b];
}


int build_tables(const float* blob_dev, const float* mean, const float* var, int fold_dropout, float* table_dev,
                 cudaStream_t s) {
  build_tables_kernel<<<(kChTotal + 127) / 128, 128, 0, s>>>(blob_dev, blob_offsets(), mean, var, fold_dropout,
                                                             table_dev);
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}

// Exact batch-statistic mode: after the statistics sweep of `layer`, turn the fixed-point (sum, sum of squares) of every
// pass g into that pass's BatchNorm affine (dropout scale folded) inside its epilogue table.  Same arithmetic as
// bn_finalize_kernel of the layer-synchronous path.  grid = n_groups, block >= cout.
__global__ void exact_layer_tables_kernel(const float* __restrict__ blob, BlobOffsets off,
                                          const unsigned long long* __restrict__ stats, double count, int layer,
                                          float* __restrict__ tables, int* __restrict__ err_flag) {
  const int g = blockIdx.x, c = threadIdx.x;
  const BlockDef b = block_def(layer);
  if (c >= b.cout) return;
  const long long* q = reinterpret_cast<const long long*>(stats) + (size_t)g * 2 * kMaxCh;
  // sums of post-ReLU values are non-negative: a set sign bit means the 2^-24 fixed point wrapped (sum of squares > 5.5e11)
  if ((q[2 * c] < 0 || q[2 * c + 1] < 0) && err_flag) *reinterpret_cast<volatile int*>(err_flag) = 50;
  const double mu = (double)q[2 * c] / kFusedStatScale / count;
  double var = (double)q[2 * c + 1] / kFusedStatScale / count - mu * mu;   // biased variance (Keras training mode)
  if (var < 0) var = 0;
  const float inv_keep = 1.0f / (1.0f - b.rate);
  const float inv = (float)((double)blob[off.gamma[layer] + c] / sqrt(var + (double)kBnEps));
  float* t = tables + (size_t)g * kTableFloats;
  t[kChTotal + ch_off(layer) + c] = inv * inv_keep;
  t[2 * kChTotal + ch_off(layer) + c] = (blob[off.beta[layer] + c] - (float)mu * inv) * inv_keep;
}

__global__ void replicate_table_kernel(const float* __restrict__ base, float* __restrict__ tables) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < kTableFloats) tables[(size_t)blockIdx.y * kTableFloats + i] = base[i];
}

// Calibration on the tensor-core sweeps: average the per-pass batch (mean, biased variance) of `layer` over the passes
// into the frozen statistics (frozen[0..] = means 6 x 256, frozen[6*256..] = variances).  One thread per channel.
__global__ void exact_layer_average_kernel(const unsigned long long* __restrict__ stats, double count, int layer, int cout,
                                           int n_groups, float* __restrict__ frozen) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cout) return;
  double am = 0.0, av = 0.0;
  for (int g = 0; g < n_groups; ++g) {
    const long long* q = reinterpret_cast<const long long*>(stats) + (size_t)g * 2 * kMaxCh;
    const double mu = (double)q[2 * c] / kFusedStatScale / count;
    double var = (double)q[2 * c + 1] / kFusedStatScale / count - mu * mu;
    if (var < 0) var = 0;
    am += mu;
    av += var;
  }
  frozen[layer * kMaxCh + c] = (float)(am / (double)n_groups);
  frozen[(kBlocks + layer) * kMaxCh + c] = (float)(av / (double)n_groups);
}

int exact_layer_average(const unsigned long long* stats, double count, int layer, int n_groups, float* frozen, cudaStream_t s) {
  exact_layer_average_kernel<<<1, kMaxCh, 0, s>>>(stats, count, layer, block_def(layer).cout, n_groups, frozen);
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}

// Exact mode, last step: probabilities from the head partials of the sixth sweep.  With A = masked column sum and
// B = kept-row count of a window's layer-6 output, GAP + Dense give z = (1/60) * sum_c w_c (scale_c A_c + shift_c B_c) + b
// (scale / shift = this pass's BatchNorm affine with the dropout scale folded).  One warp per (pass, window).
__global__ void __launch_bounds__(256)
exact_head_kernel(const float* __restrict__ head_sum, const unsigned char* __restrict__ head_cnt,
                  const float* __restrict__ tables, int n_groups, int64_t n, float* __restrict__ probs /* (n_groups, n) */) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  constexpr int kL6 = kBlocks - 1;
  for (int64_t i = warp; i < (int64_t)n_groups * n; i += nwarps) {
    const int g = (int)(i / n);
    const float* t = tables + (size_t)g * kTableFloats;
    const float* hs = head_sum + (size_t)i * 2 * kDenseIn;
    const unsigned char* hc = head_cnt + (size_t)i * 2 * kDenseIn;
    float z = 0.f;
#pragma unroll
    for (int k = 0; k < kDenseIn / 32; ++k) {
      const int c = lane + 32 * k;
      const float A = __ldcs(hs + c) + __ldcs(hs + kDenseIn + c);
      const float B = (float)((int)hc[c] + (int)hc[kDenseIn + c]);
      const float sc = __ldg(t + kChTotal + ch_off(kL6) + c), sh = __ldg(t + 2 * kChTotal + ch_off(kL6) + c);
      z = fmaf(__ldg(t + 3 * kChTotal + c), fmaf(sc, A, sh * B), z);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) z += __shfl_xor_sync(0xffffffffu, z, o);
    if (lane == 0) probs[i] = 1.0f / (1.0f + expf(-(z * (1.0f / (float)kL) + __ldg(t + 3 * kChTotal + kDenseIn))));
  }
}

int exact_head(const float* head_sum, const unsigned char* head_cnt, const float* tables, int n_groups, int64_t n, float* probs,
               cudaStream_t s) {
  if (n == 0 || n_groups == 0) return UQK_OK;
  const int64_t warps = (int64_t)n_groups * n;
  exact_head_kernel<<<(int)std::min<int64_t>((warps * 32 + 255) / 256, 148 * 16), 256, 0, s>>>(head_sum, head_cnt, tables, n_groups,
                                                                                              n, probs);
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}

__global__ void replicate_stats_kernel(unsigned long long* __restrict__ stats) {   // row 0 -> rows 1..gridDim.y
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < 2 * kMaxCh) stats[(size_t)(blockIdx.y + 1) * 2 * kMaxCh + i] = stats[i];
}

int exact_replicate_stats(unsigned long long* stats, int n_groups, cudaStream_t s) {
  if (n_groups < 2) return UQK_OK;
  replicate_stats_kernel<<<dim3(2, n_groups - 1), 256, 0, s>>>(stats);
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}

int exact_replicate_tables(const float* base_table, int n_groups, float* tables, cudaStream_t s) {
  replicate_table_kernel<<<dim3((kTableFloats + 255) / 256, n_groups), 256, 0, s>>>(base_table, tables);
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}

int exact_layer_tables(const float* blob_dev, const unsigned long long* stats, double count, int layer, int n_groups,
                       float* tables, int* err_flag, cudaStream_t s) {
  exact_layer_tables_kernel<<<n_groups, kMaxCh, 0, s>>>(blob_dev, blob_offsets(), stats, count, layer, tables, err_flag);
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}

int fused_launch(const FusedParams& p, cudaStream_t s) {
  if (p.n == 0) return UQK_OK;
  // variant >> 4: 0 / 3 = CTA pairs with two tiles per CTA (default), 2 = CTA pairs with one tile per CTA and two CTAs
  // per SM, 4 = two tiles per CTA, 1 = one tile per CTA.  Every configuration falls through to a smaller one when T
  // (or M) probabilities per window do not fit its shared memory.
  const int v = (p.variant >> 4) & 15;
  if (p.stats) return fused_launch_pair_stat(p, s);   // statistics sweeps exist in the CTA-pair kernel only
  int rc = UQK_E_SHAPE;
  if (v == 0 || v == 3) rc = fused_launch_pair2(p, s);
  if (rc == UQK_E_SHAPE && (v == 0 || v == 2 || v == 3)) rc = fused_launch_pair(p, s);
  if (rc == UQK_E_SHAPE && v != 1) rc = fused_launch_single2(p, s);
  if (rc == UQK_E_SHAPE) rc = fused_launch_single1(p, s);
  if (rc == UQK_E_SHAPE) set_error("fused kernel: %d passes/members do not fit in shared memory", p.n_groups);
  return rc;
}

}  // namespace uqk