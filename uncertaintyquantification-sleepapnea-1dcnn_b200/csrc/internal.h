// Internal declarations shared by the translation units of libuqk.so.
#pragma once
#include "common.cuh"

namespace uqk {

// ---- layer-synchronous fp32 path (layersync.cu) -------------------------------------------------------------
struct LsPassParams {
  const float* blob_dev;    // one member's Keras-order blob on the device
  const float* x_f32;       // (n,60,4) float32
  float* act0;              // 2 activation buffers of n*60*256 floats
  float* act1;
  int64_t n;
  int bn_mode;              // UQK_BN_*
  int dropout;              // 1: apply dropout masks (training=True)
  uint64_t seed;
  int pass;
  int64_t window_base;
  const uint32_t* mask_bits;  // (n,6,60,8) keep bits for THIS pass, or null -> Philox
  const float* frozen_mean;   // 6 x 256 (UQK_BN_CALIBRATED)
  const float* frozen_var;
  double* stats_scratch;      // 6 x 256 x 2 doubles
  float* affine_scratch;      // 6 x 2 x 256 floats
  double* accum_mean;         // optional 6 x 256 doubles: += batch mean (calibration)
  double* accum_var;
  float* probs_out;           // (n,)
  // cross-rank exchange of the batch statistics (uqk_set_stats_exchange); null = this rank holds the whole batch
  int (*exchange)(void* user, long long* stats_dev, int count, void* stream);
  void* exchange_user;
  int64_t global_n;           // windows of the whole call over all ranks (divisor of the statistics) when exchanging
  const int64_t* win_ids;     // optional (n,): position of row i in the call (masks are keyed by it); null = i
  int* err_flag;              // self-check flag (mapped host memory): 50 = statistics accumulator overflow
  const float* affine_tab;    // optional epilogue table (kTableFloats) whose scale / shift rows replace the BatchNorm affine
};
int ls_cast_input(const void* x, int x_dtype, float* out, int64_t n_windows, cudaStream_t s);
int ls_gather_input(const void* x, int x_dtype, const int64_t* idx, float* out, int64_t n_idx, cudaStream_t s);
int ls_forward_pass(const LsPassParams& p, cudaStream_t s);

// ---- fused tcgen05 path (fused.cu) ----------------------------------------------------------------------------
// Per-set epilogue tables (a "set" is an ensemble member, or the single MC-Dropout network):
//   float bias[992], scale[992], shift[992], dense_w[96], dense_b, pad[3]   -> kTableFloats floats
constexpr int kTableFloats = 3 * kChTotal + kDenseIn + 4;
// Packed bf16 weights of one member for the fused kernel (UMMA K-major, no swizzle), see fused.cu.
size_t fused_packed_bytes();                      // bytes per member
void fused_pack_member(const float* blob_host, uint16_t* packed_host);   // host-side packing
void fused_pack_member_pair(const float* blob_host, uint16_t* packed_host);   // [half 0][half 1] for CTA pairs

struct FusedParams {
  const void* x;              // (n,60,4) f32 or f64, device
  int x_dtype;
  int64_t n;
  int n_groups;               // T passes (MCD) or M members (DE)
  int mode_de;                // 1: group g uses member g's weights/tables, no dropout; 0: member 0, dropout
  int dropout;
  uint64_t seed;
  int64_t window_base;
  const uint32_t* mask_bits;  // (T,N,6,60,8) or null
  const uint16_t* packed;     // n_members x packed bytes
  const uint16_t* packed_pair;  // same, CTA-pair layout
  const float* tables;        // n_sets x kTableFloats
  float* probs;               // (n_groups, n) or null
  MetricsOut out;
  float* dbg_act;             // optional debug dump (6,128,256) floats of unit 0 / group 0 / tile 0
  int* err_flag;              // device int: set !=0 by the in-kernel watchdog
  int variant;                // kernel variant selector (tests)
  int no_smem_probs;          // 1: too many passes for shared memory: probabilities go to `probs` only (K4 reduces them)
  // exact batch-statistic mode (see run_fused_exact in api.cu)
  int tab_per_group;          // 1: group g reads tables[g] (per-pass BatchNorm affine) with member 0's weights
  int n_layers;               // statistics sweep: layers [0, n_layers) are computed ...
  unsigned long long* stats;  // ... and (sum, sum of squares) of layer n_layers-1 go to stats[g][2*256]; null = normal run
  int pass_base;              // Philox pass counter of group 0
  float* head_sum;            // sixth sweep: (n_groups, n, 2, 96) masked column sums of the layer-6 output, or null
  unsigned char* head_cnt;    // ... and kept-row counts
  // exact mode, per-layer hand-off through HBM: post-ReLU output of a layer, bf16, (n_groups, n, Cout/8, 60) x 16 B
  void* ckpt_out;             // written by this sweep (output of layer n_layers-1), or null
  const void* ckpt_in;        // output of layer first_layer-1, read by a sweep that starts at first_layer >= 2, or null
  int first_layer;            // first layer the sweep computes (0 = from the windows)
  int pipelined;              // option exact_pipelined: three-stage pipelined variant of the resumed sweeps
};
// bytes of a layer's stored output per window-pass
constexpr size_t exact_layer_bytes(int l) { return (size_t)(block_def(l).cout / 8) * kL * 16; }
int fused_launch(const FusedParams& p, cudaStream_t s);
// one kernel family per translation unit (fused_pair2.cu, fused_pair.cu, fused_pair_stat.cu, fused_single2.cu, fused_single1.cu)
int fused_launch_pair2(const FusedParams& p, cudaStream_t s);
int fused_launch_pair(const FusedParams& p, cudaStream_t s);
int fused_launch_pair_stat(const FusedParams& p, cudaStream_t s);
int fused_launch_single2(const FusedParams& p, cudaStream_t s);
int fused_launch_single1(const FusedParams& p, cudaStream_t s);

// tables builder: bias/scale/shift/dense for one set.  mean/var: 6x256 floats (frozen) or null -> moving stats.
int build_tables(const float* blob_dev, const float* mean, const float* var, int fold_dropout, float* table_dev,
                 cudaStream_t s);

// exact batch-statistic mode helpers (fused.cu)
int exact_replicate_tables(const float* base_table, int n_groups, float* tables, cudaStream_t s);
int exact_layer_tables(const float* blob_dev, const unsigned long long* stats, double count, int layer, int n_groups,
                       float* tables, int* err_flag, cudaStream_t s);

int exact_head(const float* head_sum, const unsigned char* head_cnt, const float* tables, int n_groups, int64_t n, float* probs,
               cudaStream_t s);
int exact_replicate_stats(unsigned long long* stats, int n_groups, cudaStream_t s);
int exact_layer_average(const unsigned long long* stats, double count, int layer, int n_groups, float* frozen, cudaStream_t s);

}  // namespace uqk
