// C-ABI of libuqk.so (declared in include/uqk.h): handle life cycle, model upload, and the entry points that
// stand in for the reference's mc_dropout_predict / deep_ensembles_predict / uq_evaluation_dist
// (uncertainty_quantification/uq_techniques.py:12-112).
#include "common.cuh"
#include "internal.h"

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <vector>
#include <algorithm>

namespace uqk {

static thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

struct Handle {
  int device = 0;
  int n_members = 0;
  float* d_blobs = nullptr;       // n_members x kBlobFloats
  uint16_t* d_packed = nullptr;   // n_members x packed image (bf16)
  uint16_t* d_packed_pair = nullptr;  // n_members x packed image, CTA-pair layout
  float* d_tables_inf = nullptr;  // n_members x kTableFloats: inference BN, no dropout (ensemble / deterministic)
  float* d_table_mcd = nullptr;   // kTableFloats: MC-Dropout table (dropout scale folded), rebuilt per call
  float* d_frozen = nullptr;      // 2 x 6 x 256: calibrated mean, var
  bool have_frozen = false;
  double* d_stats = nullptr;      // 6 x 256 x 2
  float* d_affine = nullptr;      // 6 x 2 x 256
  double* d_accum = nullptr;      // 2 x 6 x 256
  int* d_err = nullptr;           // device alias of h_err
  int* h_err = nullptr;           // watchdog / self-check flag in MAPPED PINNED HOST memory: readable after a device trap
  // grow-on-demand workspace of the layer-synchronous path
  float* d_x32 = nullptr; size_t x32_cap = 0;
  float* d_act[2] = {nullptr, nullptr}; size_t act_cap = 0;
  float* d_probs_tmp = nullptr; size_t probs_cap = 0;
  float* d_tables_pass = nullptr; size_t tables_pass_cap = 0;   // exact batch-statistic mode: one epilogue table per pass
  float* d_stats_pass = nullptr; size_t stats_pass_cap = 0;     // ... and n_passes x 2 x 256 int64 statistics (as floats x2)
  float* d_head_sum = nullptr; size_t head_sum_cap = 0;         // exact mode: head partials of one chunk of passes
  float* d_head_cnt = nullptr; size_t head_cnt_cap = 0;         // ... kept-row counts (bytes, allocated in floats)
  float* d_ckpt = nullptr; size_t ckpt_cap = 0;                 // ... per-layer hand-off buffer A (outputs of layers 2 and 4)
  float* d_ckpt2 = nullptr; size_t ckpt2_cap = 0;               // ... and B (outputs of layers 3 and 5)
  float* d_tables_all = nullptr; size_t tables_all_cap = 0;     // exact mode: the per-pass epilogue tables of ALL passes of the last call
  int exact_passes = 0; uint64_t exact_seed = 0; int64_t exact_n = -1;   // ... which call they belong to (guard-band re-evaluation)
  unsigned long long* d_stats_l0 = nullptr;                     // exact mode: layer 1's (exchanged) statistics of the current call
  float* d_band_idx = nullptr; size_t band_idx_cap = 0;         // guard band: int64 positions of the windows to re-evaluate (as floats x2)
  unsigned long long* d_band_count = nullptr;
  float* d_dbg = nullptr;         // 6 x 128 x 256 debug dump of the fused kernel (tests)
  int variant = 0;
  int dbg_dump = 0;
  int64_t calib_fp32_max = 8192;  // calibration sets up to this size use the fp32 layer-synchronous kernels
  int exact_chunk_max = 0;        // tests: cap on the passes per chunk of the exact mode (0 = memory budget only)
  int watchdog = 1;               // 0: the kernels wait on their barriers without a time budget (debuggers, sanitizers)
  int exact_pipelined = 1;        // resumed sweeps of the exact mode as a three-stage pipeline over 4-window units (0: both tiles in lock-step, 8-window units)
  // cross-rank exchange of batch statistics (uqk_set_stats_exchange)
  uqk_stats_exchange_fn xchg = nullptr;
  void* xchg_user = nullptr;
  int64_t xchg_global_n = 0;
};

static int ensure(float** p, size_t* cap, size_t need_floats) {
  if (*cap >= need_floats) return UQK_OK;
  if (*p) cudaFree(*p);
  *p = nullptr; *cap = 0;
  cudaError_t e = cudaMalloc(p, need_floats * sizeof(float));
  if (e != cudaSuccess) {
    cudaGetLastError();
    set_error("workspace allocation of %zu bytes failed: %s", need_floats * sizeof(float), cudaGetErrorString(e));
    return UQK_E_NOMEM;
  }
  *cap = need_floats;
  return UQK_OK;
}

// Device-side self-checks report through h->h_err (mapped host memory): barrier watchdog codes 1..23 (the kernel then
// traps: the CUDA context is lost), 50 = a batch-statistics accumulator overflowed its 2^-24 fixed point.
static int poll_kernel_flag(Handle* h) {
  const int code = h->h_err ? *reinterpret_cast<volatile int*>(h->h_err) : 0;
  if (code == 0) return UQK_OK;
  if (code == 50) {
    *reinterpret_cast<volatile int*>(h->h_err) = 0;        // reported once; the handle stays usable (no trap happened)
    set_error("device self-check %d: a batch-statistics sum overflowed its fixed-point accumulator (activations too large); "
              "results of that call are invalid", code);
  } else
    set_error("device watchdog %d: a kernel waited ~2 s on barrier class %d and trapped (protocol fault); the CUDA context of this "
              "process is lost -- restart the process", code, code);
  return UQK_E_KERNEL;
}

// waits ~2 s on a barrier nobody signals: exercises the watchdog path (tests only; the context is lost afterwards)
__global__ void trip_watchdog_kernel(int* err_flag) {
  __shared__ unsigned long long bar;
  const uint32_t b = (uint32_t)__cvta_generic_to_shared(&bar);
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b));
  long long t0 = clock64();
  for (;;) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.b32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(b) : "memory");
    if (ok) return;
    if (clock64() - t0 > 4000000000ll) {
      *reinterpret_cast<volatile int*>(err_flag) = 99;
      __threadfence_system();
      __trap();
    }
  }
}

__global__ void average_stats_kernel(const double* __restrict__ accum, int n_passes, float* __restrict__ frozen) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < 2 * kBlocks * kMaxCh) frozen[i] = (float)(accum[i] / (double)n_passes);
}

static int check_x(const void* x, int dtype, int64_t n) {
  if (n < 0) { set_error("n_windows must be >= 0"); return UQK_E_BADARG; }
  if (n > 0 && x == nullptr) { set_error("x_dev is NULL"); return UQK_E_BADARG; }
  if (dtype != UQK_F32 && dtype != UQK_F64) { set_error("x_dtype must be UQK_F32 or UQK_F64"); return UQK_E_BADARG; }
  if (n > 0 && (reinterpret_cast<uintptr_t>(x) & 31)) { set_error("x_dev must be 32-byte aligned"); return UQK_E_BADARG; }
  return UQK_OK;
}

// Layer-synchronous T passes (or M members) over all windows into probs (n_groups, n).
// win_ids (optional, device): the n rows are rows win_ids[i] of a call of n_call windows (guard-band re-evaluation):
// x is gathered through them, masks are keyed / indexed by them; frozen-statistics modes only.
static int run_layersync(Handle* h, int member0, const void* x, int x_dtype, int64_t n, int n_groups, bool de, bool dropout,
                         int bn_mode, uint64_t seed, int64_t window_base, const uint32_t* mask_bits, float* probs,
                         double* accum_mean, double* accum_var, cudaStream_t s, const int64_t* win_ids = nullptr,
                         int64_t n_call = 0, const float* affine_tabs = nullptr) {
  const bool exchanging = bn_mode == UQK_BN_BATCH && h->xchg != nullptr;
  if (n == 0 && !exchanging) return UQK_OK;   // an empty shard still takes part in every exchange of the call
  const int64_t chunk = bn_mode == UQK_BN_BATCH ? std::max<int64_t>(n, 1) : std::min<int64_t>(n, 16384);
  int rc;
  if ((rc = ensure(&h->d_x32, &h->x32_cap, (size_t)n * kL * kF))) return rc;
  size_t cap0 = h->act_cap, cap1 = h->act_cap;
  if ((rc = ensure(&h->d_act[0], &cap0, (size_t)chunk * kL * kMaxCh))) return rc;
  if ((rc = ensure(&h->d_act[1], &cap1, (size_t)chunk * kL * kMaxCh))) { h->act_cap = 0; return rc; }
  h->act_cap = std::min(cap0, cap1);
  if (win_ids) rc = ls_gather_input(x, x_dtype, win_ids, h->d_x32, n, s);
  else rc = ls_cast_input(x, x_dtype, h->d_x32, n, s);
  if (rc) return rc;
  for (int g = 0; g < n_groups; ++g) {
    for (int64_t c0 = 0; c0 < std::max<int64_t>(n, 1); c0 += chunk) {
      const int64_t cn = std::min(chunk, n - c0);
      LsPassParams p{};
      p.blob_dev = h->d_blobs + (size_t)(member0 + (de ? g : 0)) * kBlobStride;
      p.x_f32 = h->d_x32 + (size_t)c0 * kL * kF;
      p.act0 = h->d_act[0]; p.act1 = h->d_act[1];
      p.n = cn; p.bn_mode = bn_mode; p.dropout = dropout ? 1 : 0; p.seed = seed; p.pass = g;
      p.window_base = window_base + c0;
      p.mask_bits = mask_bits ? mask_bits + ((size_t)g * (win_ids ? n_call : n) + (win_ids ? 0 : c0)) * kBlocks * kL * 8 : nullptr;
      p.win_ids = win_ids ? win_ids + c0 : nullptr;
      if (win_ids) p.window_base = window_base;
      p.err_flag = h->d_err;
      p.affine_tab = affine_tabs ? affine_tabs + (size_t)g * kTableFloats : nullptr;
      p.frozen_mean = h->d_frozen; p.frozen_var = h->d_frozen + kBlocks * kMaxCh;
      p.stats_scratch = h->d_stats; p.affine_scratch = h->d_affine;
      p.accum_mean = accum_mean; p.accum_var = accum_var;
      p.probs_out = probs + (size_t)g * n + c0;
      if (exchanging) { p.exchange = h->xchg; p.exchange_user = h->xchg_user; p.global_n = h->xchg_global_n; }
      if ((rc = ls_forward_pass(p, s))) return rc;
    }
  }
  return UQK_OK;
}

// Passes (or members) per window that still fit the fused kernel's shared-memory probability buffer in its default
// configuration; beyond it the probabilities go through HBM and the stand-alone reduction kernel (K4).
constexpr int kMaxFusedGroups = 256;

static int run_fused(Handle* h, int member0, const void* x, int x_dtype, int64_t n, int n_groups, bool de, bool dropout,
                     uint64_t seed, int64_t window_base, const uint32_t* mask_bits, const float* tables, float* probs,
                     const MetricsOut& out, bool want_metrics, cudaStream_t s) {
  FusedParams p{};
  const size_t pb = fused_packed_bytes();
  p.x = x; p.x_dtype = x_dtype; p.n = n; p.n_groups = n_groups; p.mode_de = de ? 1 : 0; p.dropout = dropout ? 1 : 0;
  p.seed = seed; p.window_base = window_base; p.mask_bits = mask_bits;
  p.packed = reinterpret_cast<const uint16_t*>(reinterpret_cast<const uint8_t*>(h->d_packed) + (size_t)member0 * pb);
  p.packed_pair = reinterpret_cast<const uint16_t*>(reinterpret_cast<const uint8_t*>(h->d_packed_pair) + (size_t)member0 * pb);
  p.tables = tables; p.err_flag = h->watchdog ? h->d_err : nullptr; p.variant = h->variant;
  p.dbg_act = h->dbg_dump ? h->d_dbg : nullptr;
  if (n_groups <= kMaxFusedGroups) {
    p.probs = probs; p.out = out;
    return fused_launch(p, s);
  }
  int rc;
  if (!probs) {
    if ((rc = ensure(&h->d_probs_tmp, &h->probs_cap, (size_t)n * n_groups))) return rc;
    probs = h->d_probs_tmp;
  }
  p.probs = probs; p.out = MetricsOut{}; p.no_smem_probs = 1;
  if ((rc = fused_launch(p, s))) return rc;
  return want_metrics ? launch_reduce(probs, n_groups, n, out, s) : UQK_OK;
}

// The six statistics sweeps over passes [pass0, pass0 + n_passes): leaves every pass's BatchNorm affine in
// h->d_tables_pass and, when `frozen` is given, the pass-averaged (mean, variance) of every layer there (tensor-core
// calibration).  When head buffers are given the sixth sweep also emits the head partials (see exact_head_kernel).
// Every layer is computed ONCE: sweep k (k >= 3) reads layer k-1's post-ReLU output from HBM (bf16), normalises it with
// the statistics the previous sweep produced, runs layer k and stores ITS post-ReLU output; sweep 2 recomputes the
// cheap, pass-invariant layer 1 from the windows (0.4 % of the flops) instead of storing it.  bufA / bufB: ping-pong
// buffers of exact_buf_bytes(0 / 1) bytes per window-pass.
constexpr size_t exact_buf_bytes(int which) {
  return which == 0 ? std::max(exact_layer_bytes(1), exact_layer_bytes(3)) : std::max(exact_layer_bytes(2), exact_layer_bytes(4));
}
constexpr size_t kExactBytesPerWindowPass = exact_buf_bytes(0) + exact_buf_bytes(1) + 2 * kDenseIn * 5;

static int run_exact_sweeps(Handle* h, const void* x, int x_dtype, int64_t n, int pass0, int n_passes, uint64_t seed,
                            int64_t window_base, const uint32_t* mask_bits, float* frozen, float* head_sum,
                            unsigned char* head_cnt, void* bufA, void* bufB, cudaStream_t s, bool reuse_l0 = false) {
  // reuse_l0: layer 1 sees neither dropout nor another layer's normalisation, so its statistics are the same for every
  // pass AND every chunk of passes of a call: later chunks take them from the first chunk (already exchanged over the
  // ranks) instead of sweeping the windows again
  int rc;
  if (!h->d_stats_l0) UQK_CUDA(cudaMalloc(&h->d_stats_l0, sizeof(unsigned long long) * 2 * kMaxCh));
  if ((rc = ensure(&h->d_tables_pass, &h->tables_pass_cap, (size_t)n_passes * kTableFloats))) return rc;
  if ((rc = ensure(&h->d_stats_pass, &h->stats_pass_cap, (size_t)n_passes * 2 * kMaxCh * 2))) return rc;
  unsigned long long* stats = reinterpret_cast<unsigned long long*>(h->d_stats_pass);
  // bias / dense rows of every pass's table; the scale / shift rows are filled layer by layer below
  if ((rc = build_tables(h->d_blobs, nullptr, nullptr, 1, h->d_table_mcd, s))) return rc;
  if ((rc = exact_replicate_tables(h->d_table_mcd, n_passes, h->d_tables_pass, s))) return rc;
  const bool exchanging = h->xchg != nullptr;
  const int64_t stat_windows = exchanging ? h->xchg_global_n : n;
  FusedParams p{};
  p.x = x; p.x_dtype = x_dtype; p.n = n; p.n_groups = n_passes; p.mode_de = 0; p.dropout = 1;
  p.seed = seed; p.window_base = window_base; p.pass_base = pass0;
  p.mask_bits = mask_bits ? mask_bits + (size_t)pass0 * n * kBlocks * kL * 8 : nullptr;
  p.packed = h->d_packed; p.packed_pair = h->d_packed_pair;
  p.tables = h->d_tables_pass; p.tab_per_group = 1; p.err_flag = h->watchdog ? h->d_err : nullptr; p.variant = h->variant;
  p.pipelined = h->exact_pipelined;
  if (stat_windows < 1) return UQK_OK;
  for (int k = 0; k < kBlocks; ++k) {
    UQK_CUDA(cudaMemsetAsync(stats, 0, sizeof(unsigned long long) * (size_t)n_passes * 2 * kMaxCh, s));
    p.stats = stats;
    // nothing is random before the first BatchNorm: layer 1's statistics are the same for every pass, so its sweep runs
    // one pass and the sums are copied to the other passes' rows
    p.n_groups = k == 0 ? 1 : n_passes;
    p.head_sum = k == kBlocks - 1 ? head_sum : nullptr;
    p.head_cnt = k == kBlocks - 1 ? head_cnt : nullptr;
    void* const buf_of[kBlocks] = {nullptr, bufA, bufB, bufA, bufB, nullptr};    // where layer k's output lives
    p.first_layer = k <= 1 ? 0 : k;
    p.n_layers = k + 1;
    p.ckpt_in = k >= 2 ? buf_of[k - 1] : nullptr;
    p.ckpt_out = buf_of[k];
    const bool reuse = k == 0 && reuse_l0;
    if (reuse) UQK_CUDA(cudaMemcpyAsync(stats, h->d_stats_l0, sizeof(unsigned long long) * 2 * kMaxCh, cudaMemcpyDeviceToDevice, s));
    else if ((rc = fused_launch(p, s))) return rc;      // n == 0 (empty shard): no launch, statistics stay zero
    if (k == 0 && n_passes > 1 && (rc = exact_replicate_stats(stats, n_passes, s))) return rc;
    if (exchanging && !reuse &&
        h->xchg(h->xchg_user, reinterpret_cast<long long*>(stats), n_passes * 2 * kMaxCh, static_cast<void*>(s)) != 0) {
      set_error("batch-statistics exchange callback failed at layer %d", k);
      return UQK_E_STATE;
    }
    if (k == 0 && !reuse) UQK_CUDA(cudaMemcpyAsync(h->d_stats_l0, stats, sizeof(unsigned long long) * 2 * kMaxCh, cudaMemcpyDeviceToDevice, s));
    if ((rc = exact_layer_tables(h->d_blobs, stats, (double)stat_windows * kL, k, n_passes, h->d_tables_pass,
                                 (h->variant & 14) ? nullptr : h->d_err /* perf-decomposition switches compute garbage */, s))) return rc;
    if (frozen && (rc = exact_layer_average(stats, (double)stat_windows * kL, k, n_passes, frozen, s))) return rc;
  }
  return UQK_OK;
}

// Exact Keras `training=True` semantics on the tensor-core kernel (uq_techniques.py:22: BatchNormalization uses the
// statistics of the whole call).  The statistics of layer k depend on the normalised output of layers < k over ALL
// windows, so the network is swept layer by layer (run_exact_sweeps): sweep k accumulates layer k's per-pass
// (sum, sum of squares) and hands its un-normalised output to sweep k+1 through HBM (96 KB written + 96 KB read per
// window-pass in total, bf16).  The sixth sweep leaves, per (pass, window), the dropout-masked column sums of the
// layer-6 output; BatchNorm-6, GlobalAveragePooling and Dense are linear in them, so a small kernel turns them into
// the probabilities once layer 6's statistics are known.  The passes are processed in chunks so that the two
// hand-off buffers and the head partials stay within kExactScratchBudget.
constexpr size_t kExactScratchBudget = (size_t)120 << 30;

static int exact_scratch(Handle* h, int64_t n, int chunk, bool with_head) {
  int rc = ensure(&h->d_ckpt, &h->ckpt_cap, ((size_t)chunk * n * exact_buf_bytes(0) + 3) / 4);
  if (!rc) rc = ensure(&h->d_ckpt2, &h->ckpt2_cap, ((size_t)chunk * n * exact_buf_bytes(1) + 3) / 4);
  if (!rc && with_head) rc = ensure(&h->d_head_sum, &h->head_sum_cap, (size_t)chunk * n * 2 * kDenseIn);
  if (!rc && with_head) rc = ensure(&h->d_head_cnt, &h->head_cnt_cap, ((size_t)chunk * n * 2 * kDenseIn + 3) / 4);
  return rc;
}

static int run_fused_exact(Handle* h, const void* x, int x_dtype, int64_t n, int n_passes, uint64_t seed, int64_t window_base,
                           const uint32_t* mask_bits, float* probs, const MetricsOut& out, bool want_metrics, cudaStream_t s) {
  int rc;
  const bool exchanging = h->xchg != nullptr;
  // every rank of a sharded call must cut the passes identically: size the chunks by the windows of the whole call
  const int64_t n_for_chunks = std::max<int64_t>(1, exchanging ? h->xchg_global_n : n);
  const size_t per_pass = (size_t)n_for_chunks * kExactBytesPerWindowPass;
  int chunk = (int)std::max<size_t>(1, std::min<size_t>((size_t)n_passes, std::min<size_t>(kExactScratchBudget / per_pass, kMaxFusedGroups)));
  if (h->exact_chunk_max > 0) chunk = std::min(chunk, h->exact_chunk_max);
  if (n > 0) {
    for (;;) {
      // the result does not depend on the chunk size, so a GPU short of memory gets smaller chunks instead of an error
      // (not when the passes are cut jointly with other ranks)
      rc = exact_scratch(h, n, chunk, true);
      if (rc != UQK_E_NOMEM || chunk == 1 || exchanging) break;
      chunk = (chunk + 1) / 2;
    }
    if (rc) return rc;
    if (!probs) {
      if ((rc = ensure(&h->d_probs_tmp, &h->probs_cap, (size_t)n * n_passes))) return rc;
      probs = h->d_probs_tmp;
    }
  }
  unsigned char* head_cnt = reinterpret_cast<unsigned char*>(h->d_head_cnt);
  h->exact_passes = 0;
  if ((rc = ensure(&h->d_tables_all, &h->tables_all_cap, (size_t)n_passes * kTableFloats))) return rc;
  for (int g0 = 0; g0 < n_passes; g0 += chunk) {
    const int gc = std::min(chunk, n_passes - g0);
    if ((rc = run_exact_sweeps(h, x, x_dtype, n, g0, gc, seed, window_base, mask_bits, nullptr, h->d_head_sum, head_cnt, h->d_ckpt, h->d_ckpt2, s,
                               /*reuse_l0=*/g0 > 0))) return rc;
    // keep every pass's BatchNorm affine: uqk_refine_band re-evaluates guard-band windows with exactly these statistics
    UQK_CUDA(cudaMemcpyAsync(h->d_tables_all + (size_t)g0 * kTableFloats, h->d_tables_pass, sizeof(float) * (size_t)gc * kTableFloats,
                             cudaMemcpyDeviceToDevice, s));
    if (n > 0 && (rc = exact_head(h->d_head_sum, head_cnt, h->d_tables_pass, gc, n, probs + (size_t)g0 * n, s))) return rc;
  }
  h->exact_passes = n_passes; h->exact_seed = seed; h->exact_n = n;
  if (n == 0) return UQK_OK;
  return want_metrics ? launch_reduce(probs, n_passes, n, out, s) : UQK_OK;
}

#ifdef UQK_TIMELINE
long long* g_uqk_timeline = nullptr;   // timeline build (tools/timeline_probe.py) only
#endif
}  // namespace uqk

using namespace uqk;
#ifdef UQK_TIMELINE
extern "C" void uqk_debug_set_timeline(long long* dev_buf) { uqk::g_uqk_timeline = dev_buf; }
#endif

extern "C" {

int uqk_version(void) { return UQK_VERSION; }
const char* uqk_last_error(void) { return g_err; }
int64_t uqk_launch_count(uqk_handle*) { return (int64_t)g_launches.load(); }

int uqk_create(int device, uqk_handle** out) {
  if (!out) { set_error("out is NULL"); return UQK_E_BADARG; }
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    cudaGetLastError();
    set_error("no CUDA device available (%s); libuqk has no CPU fallback", e == cudaSuccess ? "count=0" : cudaGetErrorString(e));
    return UQK_E_CUDA;
  }
  if (device < 0 || device >= count) { set_error("device %d out of range (0..%d)", device, count - 1); return UQK_E_BADARG; }
  cudaDeviceProp prop;
  UQK_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    set_error("device %d is sm_%d%d; libuqk is built for sm_100a (B200) only", device, prop.major, prop.minor);
    return UQK_E_ARCH;
  }
  UQK_CUDA(cudaSetDevice(device));
  Handle* h = new Handle();
  h->device = device;
  UQK_CUDA(cudaMalloc(&h->d_table_mcd, sizeof(float) * kTableFloats));
  UQK_CUDA(cudaMalloc(&h->d_frozen, sizeof(float) * 2 * kBlocks * kMaxCh));
  UQK_CUDA(cudaMalloc(&h->d_stats, sizeof(double) * kBlocks * kMaxCh * 2));
  UQK_CUDA(cudaMalloc(&h->d_affine, sizeof(float) * kBlocks * 2 * kMaxCh));
  UQK_CUDA(cudaMalloc(&h->d_accum, sizeof(double) * 2 * kBlocks * kMaxCh));
  UQK_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&h->h_err), sizeof(int), cudaHostAllocMapped));
  *h->h_err = 0;
  UQK_CUDA(cudaHostGetDevicePointer(reinterpret_cast<void**>(&h->d_err), h->h_err, 0));
  UQK_CUDA(cudaMalloc(&h->d_dbg, sizeof(float) * kBlocks * 128 * 256));
  UQK_CUDA(cudaMemset(h->d_dbg, 0, sizeof(float) * kBlocks * 128 * 256));
  *out = reinterpret_cast<uqk_handle*>(h);
  return UQK_OK;
}

int uqk_destroy(uqk_handle* hh) {
  if (!hh) return UQK_OK;
  Handle* h = reinterpret_cast<Handle*>(hh);
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  void* ptrs[] = {h->d_blobs, h->d_packed, h->d_packed_pair, h->d_tables_inf, h->d_table_mcd, h->d_frozen, h->d_stats, h->d_affine,
                  h->d_accum, h->d_tables_all, h->d_stats_l0, h->d_band_idx, h->d_band_count, h->d_x32, h->d_act[0], h->d_act[1], h->d_probs_tmp, h->d_dbg, h->d_tables_pass, h->d_stats_pass, h->d_head_sum, h->d_head_cnt, h->d_ckpt, h->d_ckpt2};
  for (void* p : ptrs) if (p) cudaFree(p);
  if (h->h_err) cudaFreeHost(h->h_err);
  cudaGetLastError();
  delete h;
  return UQK_OK;
}

int uqk_set_option(uqk_handle* hh, const char* key, int64_t value) {
  if (!hh || !key) { set_error("NULL argument"); return UQK_E_BADARG; }
  Handle* h = reinterpret_cast<Handle*>(hh);
  if (!strcmp(key, "variant")) { h->variant = (int)value; return UQK_OK; }
  if (!strcmp(key, "dbg_dump")) { h->dbg_dump = (int)value; return UQK_OK; }
  if (!strcmp(key, "calib_fp32_max")) { h->calib_fp32_max = value; return UQK_OK; }   // see uqk_calibrate
  if (!strcmp(key, "watchdog")) { h->watchdog = value ? 1 : 0; return UQK_OK; }
  if (!strcmp(key, "exact_pipelined")) { h->exact_pipelined = (int)value; return UQK_OK; }
  if (!strcmp(key, "exact_chunk_max")) { h->exact_chunk_max = (int)value; return UQK_OK; }   // passes per chunk of the exact mode (0 = by memory budget)
  set_error("unknown option '%s'", key);
  return UQK_E_BADARG;
}

int uqk_set_model(uqk_handle* hh, int n_members, const float* blobs_host, size_t floats_per_member) {
  if (!hh || !blobs_host) { set_error("NULL argument"); return UQK_E_BADARG; }
  if (n_members < 1 || n_members > 4096) { set_error("n_members must be in 1..4096"); return UQK_E_BADARG; }
  if (floats_per_member != kBlobFloats) {
    set_error("weight blob has %zu floats per member, the architecture needs %zu", floats_per_member, (size_t)kBlobFloats);
    return UQK_E_SHAPE;
  }
  Handle* h = reinterpret_cast<Handle*>(hh);
  UQK_CUDA(cudaSetDevice(h->device));
  UQK_CUDA(cudaDeviceSynchronize());
  for (void* p : {(void*)h->d_blobs, (void*)h->d_packed, (void*)h->d_packed_pair, (void*)h->d_tables_inf}) if (p) cudaFree(p);
  h->d_blobs = nullptr; h->d_packed = nullptr; h->d_packed_pair = nullptr; h->d_tables_inf = nullptr; h->n_members = 0;
  const size_t pb = fused_packed_bytes();
  UQK_CUDA(cudaMalloc(&h->d_blobs, sizeof(float) * kBlobStride * n_members));
  UQK_CUDA(cudaMalloc(&h->d_packed, pb * n_members));
  UQK_CUDA(cudaMalloc(&h->d_packed_pair, pb * n_members));
  UQK_CUDA(cudaMalloc(&h->d_tables_inf, sizeof(float) * kTableFloats * n_members));
  UQK_CUDA(cudaMemcpy2D(h->d_blobs, sizeof(float) * kBlobStride, blobs_host, sizeof(float) * kBlobFloats,
                        sizeof(float) * kBlobFloats, n_members, cudaMemcpyHostToDevice));
  std::vector<uint16_t> packed(pb / 2);
  for (int m = 0; m < n_members; ++m) {
    fused_pack_member(blobs_host + (size_t)m * kBlobFloats, packed.data());
    UQK_CUDA(cudaMemcpy(reinterpret_cast<uint8_t*>(h->d_packed) + (size_t)m * pb, packed.data(), pb, cudaMemcpyHostToDevice));
    fused_pack_member_pair(blobs_host + (size_t)m * kBlobFloats, packed.data());
    UQK_CUDA(cudaMemcpy(reinterpret_cast<uint8_t*>(h->d_packed_pair) + (size_t)m * pb, packed.data(), pb, cudaMemcpyHostToDevice));
    int rc = build_tables(h->d_blobs + (size_t)m * kBlobStride, nullptr, nullptr, 0, h->d_tables_inf + (size_t)m * kTableFloats, 0);
    if (rc) return rc;
  }
  UQK_CUDA(cudaDeviceSynchronize());
  h->n_members = n_members;
  h->have_frozen = false;
  return UQK_OK;
}

int uqk_forward_det(uqk_handle* hh, int member, const void* x_dev, int x_dtype, int64_t n, int precision,
                    float* probs_dev, void* stream) {
  if (!hh) { set_error("handle is NULL"); return UQK_E_BADARG; }
  Handle* h = reinterpret_cast<Handle*>(hh);
  if (int rc_flag = poll_kernel_flag(h)) return rc_flag;
  if (h->n_members == 0) { set_error("uqk_set_model has not been called"); return UQK_E_STATE; }
  if (member < 0 || member >= h->n_members) { set_error("member %d out of range", member); return UQK_E_BADARG; }
  int rc = check_x(x_dev, x_dtype, n);
  if (rc) return rc;
  if (n > 0 && !probs_dev) { set_error("probs_dev is NULL"); return UQK_E_BADARG; }
  UQK_CUDA(cudaSetDevice(h->device));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (precision == UQK_PREC_FP32)
    return run_layersync(h, member, x_dev, x_dtype, n, 1, true, false, UQK_BN_INFERENCE, 0, 0, nullptr, probs_dev, nullptr,
                         nullptr, s);
  if (precision != UQK_PREC_BF16_TC) { set_error("unknown precision %d", precision); return UQK_E_BADARG; }
  return run_fused(h, member, x_dev, x_dtype, n, 1, true, false, 0, 0, nullptr, h->d_tables_inf + (size_t)member * kTableFloats,
                   probs_dev, MetricsOut{}, false, s);
}

int uqk_calibrate(uqk_handle* hh, const void* x_dev, int x_dtype, int64_t n, int n_passes, uint64_t seed, int64_t window_base,
                  void* stream) {
  if (!hh) { set_error("handle is NULL"); return UQK_E_BADARG; }
  Handle* h = reinterpret_cast<Handle*>(hh);
  if (int rc_flag = poll_kernel_flag(h)) return rc_flag;
  if (h->n_members == 0) { set_error("uqk_set_model has not been called"); return UQK_E_STATE; }
  const bool exchanging = h->xchg != nullptr;
  const int64_t n_all = exchanging ? h->xchg_global_n : n;      // windows the statistics are taken over (all ranks)
  if (n < (exchanging ? 0 : 1) || n_all < 1 || n_passes < 1) { set_error("calibration needs n_windows >= 1 and n_passes >= 1"); return UQK_E_BADARG; }
  int rc = check_x(x_dev, x_dtype, n);
  if (rc) return rc;
  UQK_CUDA(cudaSetDevice(h->device));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (n_all > h->calib_fp32_max) {
    // larger calibration sets go through the statistics sweeps of the tensor-core kernel (same definition: per-pass
    // batch statistics of the exact training=True forward, averaged over the passes)
    if (n_passes > kMaxFusedGroups) { set_error("calibration on more than %lld windows takes at most %d passes",
                                                (long long)h->calib_fp32_max, kMaxFusedGroups); return UQK_E_BADARG; }
    if (n > 0 && (rc = exact_scratch(h, n, n_passes, false))) return rc;
    if ((rc = run_exact_sweeps(h, x_dev, x_dtype, n, 0, n_passes, seed, window_base, nullptr, h->d_frozen, nullptr, nullptr, h->d_ckpt, h->d_ckpt2, s))) return rc;
    h->have_frozen = true;
    return UQK_OK;
  }
  if ((rc = ensure(&h->d_probs_tmp, &h->probs_cap, (size_t)n * n_passes))) return rc;
  UQK_CUDA(cudaMemsetAsync(h->d_accum, 0, sizeof(double) * 2 * kBlocks * kMaxCh, s));
  rc = run_layersync(h, 0, x_dev, x_dtype, n, n_passes, false, true, UQK_BN_BATCH, seed, window_base, nullptr, h->d_probs_tmp,
                     h->d_accum, h->d_accum + kBlocks * kMaxCh, s);
  if (rc) return rc;
  average_stats_kernel<<<(2 * kBlocks * kMaxCh + 255) / 256, 256, 0, s>>>(h->d_accum, n_passes, h->d_frozen);
  UQK_LAUNCH_CHECK();
  h->have_frozen = true;
  return UQK_OK;
}

int uqk_get_bn_stats(uqk_handle* hh, float* mean_host, float* var_host) {
  if (!hh || !mean_host || !var_host) { set_error("NULL argument"); return UQK_E_BADARG; }
  Handle* h = reinterpret_cast<Handle*>(hh);
  if (!h->have_frozen) { set_error("no calibrated statistics (call uqk_calibrate or uqk_set_bn_stats)"); return UQK_E_STATE; }
  UQK_CUDA(cudaSetDevice(h->device));
  UQK_CUDA(cudaDeviceSynchronize());
  UQK_CUDA(cudaMemcpy(mean_host, h->d_frozen, sizeof(float) * kBlocks * kMaxCh, cudaMemcpyDeviceToHost));
  UQK_CUDA(cudaMemcpy(var_host, h->d_frozen + kBlocks * kMaxCh, sizeof(float) * kBlocks * kMaxCh, cudaMemcpyDeviceToHost));
  return UQK_OK;
}

int uqk_set_bn_stats(uqk_handle* hh, const float* mean_host, const float* var_host) {
  if (!hh || !mean_host || !var_host) { set_error("NULL argument"); return UQK_E_BADARG; }
  Handle* h = reinterpret_cast<Handle*>(hh);
  UQK_CUDA(cudaSetDevice(h->device));
  UQK_CUDA(cudaDeviceSynchronize());
  UQK_CUDA(cudaMemcpy(h->d_frozen, mean_host, sizeof(float) * kBlocks * kMaxCh, cudaMemcpyHostToDevice));
  UQK_CUDA(cudaMemcpy(h->d_frozen + kBlocks * kMaxCh, var_host, sizeof(float) * kBlocks * kMaxCh, cudaMemcpyHostToDevice));
  h->have_frozen = true;
  return UQK_OK;
}

int uqk_set_stats_exchange(uqk_handle* hh, uqk_stats_exchange_fn fn, void* user, int64_t global_windows) {
  if (!hh) { set_error("handle is NULL"); return UQK_E_BADARG; }
  Handle* h = reinterpret_cast<Handle*>(hh);
  if (fn != nullptr && global_windows < 1) {
    set_error("global_windows must be >= 1 (the window count of the whole call over all ranks), got %lld",
              (long long)global_windows);
    return UQK_E_BADARG;
  }
  h->xchg = fn; h->xchg_user = fn ? user : nullptr; h->xchg_global_n = fn ? global_windows : 0;
  return UQK_OK;
}

int uqk_mcd(uqk_handle* hh, const void* x_dev, int x_dtype, int64_t n, int n_passes, uint64_t seed, int bn_mode,
            int precision, int64_t window_base, const uint32_t* mask_bits_dev, float* probs_dev,
            const uqk_metrics_out* out, void* stream) {
  if (!hh) { set_error("handle is NULL"); return UQK_E_BADARG; }
  Handle* h = reinterpret_cast<Handle*>(hh);
  if (int rc_flag = poll_kernel_flag(h)) return rc_flag;
  if (h->n_members == 0) { set_error("uqk_set_model has not been called"); return UQK_E_STATE; }
  if (n_passes < 1 || n_passes > 8192) { set_error("n_passes must be in 1..8192"); return UQK_E_BADARG; }
  if (bn_mode != UQK_BN_INFERENCE && bn_mode != UQK_BN_BATCH && bn_mode != UQK_BN_CALIBRATED) {
    set_error("unknown bn_mode %d", bn_mode); return UQK_E_BADARG;
  }
  if (bn_mode == UQK_BN_CALIBRATED && !h->have_frozen) {
    set_error("UQK_BN_CALIBRATED needs uqk_calibrate() or uqk_set_bn_stats() first"); return UQK_E_STATE;
  }
  int rc = check_x(x_dev, x_dtype, n);
  if (rc) return rc;
  UQK_CUDA(cudaSetDevice(h->device));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const MetricsOut mo = to_metrics(out);
  if (precision == UQK_PREC_BF16_TC) {
    if (bn_mode == UQK_BN_BATCH)
      return run_fused_exact(h, x_dev, x_dtype, n, n_passes, seed, window_base, mask_bits_dev, probs_dev, mo,
                             out != nullptr, s);
    const float* mean = bn_mode == UQK_BN_CALIBRATED ? h->d_frozen : nullptr;
    const float* var = bn_mode == UQK_BN_CALIBRATED ? h->d_frozen + kBlocks * kMaxCh : nullptr;
    if ((rc = build_tables(h->d_blobs, mean, var, 1, h->d_table_mcd, s))) return rc;
    return run_fused(h, 0, x_dev, x_dtype, n, n_passes, false, true, seed, window_base, mask_bits_dev, h->d_table_mcd,
                     probs_dev, mo, out != nullptr, s);
  }
  if (precision != UQK_PREC_FP32) { set_error("unknown precision %d", precision); return UQK_E_BADARG; }
  float* probs = probs_dev;
  if (!probs) {
    if ((rc = ensure(&h->d_probs_tmp, &h->probs_cap, (size_t)n * n_passes))) return rc;
    probs = h->d_probs_tmp;
  }
  rc = run_layersync(h, 0, x_dev, x_dtype, n, n_passes, false, true, bn_mode, seed, window_base, mask_bits_dev, probs,
                     nullptr, nullptr, s);
  if (rc) return rc;
  if (out) return launch_reduce(probs, n_passes, n, mo, s);
  return UQK_OK;
}

int uqk_de(uqk_handle* hh, const void* x_dev, int x_dtype, int64_t n, int n_members, int precision, float* probs_dev,
           const uqk_metrics_out* out, void* stream) {
  if (!hh) { set_error("handle is NULL"); return UQK_E_BADARG; }
  Handle* h = reinterpret_cast<Handle*>(hh);
  if (int rc_flag = poll_kernel_flag(h)) return rc_flag;
  if (h->n_members == 0) { set_error("uqk_set_model has not been called"); return UQK_E_STATE; }
  if (n_members < 1 || n_members > h->n_members) {
    set_error("n_members=%d but the handle holds %d members", n_members, h->n_members); return UQK_E_BADARG;
  }
  int rc = check_x(x_dev, x_dtype, n);
  if (rc) return rc;
  UQK_CUDA(cudaSetDevice(h->device));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const MetricsOut mo = to_metrics(out);
  if (precision == UQK_PREC_BF16_TC)
    return run_fused(h, 0, x_dev, x_dtype, n, n_members, true, false, 0, 0, nullptr, h->d_tables_inf, probs_dev, mo, out != nullptr, s);
  if (precision != UQK_PREC_FP32) { set_error("unknown precision %d", precision); return UQK_E_BADARG; }
  float* probs = probs_dev;
  if (!probs) {
    if ((rc = ensure(&h->d_probs_tmp, &h->probs_cap, (size_t)n * n_members))) return rc;
    probs = h->d_probs_tmp;
  }
  rc = run_layersync(h, 0, x_dev, x_dtype, n, n_members, true, false, UQK_BN_INFERENCE, 0, 0, nullptr, probs, nullptr,
                     nullptr, s);
  if (rc) return rc;
  if (out) return launch_reduce(probs, n_members, n, mo, s);
  return UQK_OK;
}

int uqk_refine_band(uqk_handle* hh, int mode_de, const void* x_dev, int x_dtype, int64_t n, int n_groups, uint64_t seed,
                    int bn_mode, int64_t window_base, const uint32_t* mask_bits_dev, float delta, float* probs_dev,
                    const uqk_metrics_out* out, int64_t* n_refined_host, void* stream) {
  if (!hh) { set_error("handle is NULL"); return UQK_E_BADARG; }
  Handle* h = reinterpret_cast<Handle*>(hh);
  if (int rc_flag = poll_kernel_flag(h)) return rc_flag;
  if (n_refined_host) *n_refined_host = 0;
  if (h->n_members == 0) { set_error("uqk_set_model has not been called"); return UQK_E_STATE; }
  if (!out || !out->mean_pred) { set_error("uqk_refine_band needs out->mean_pred (the result of the bf16 call)"); return UQK_E_BADARG; }
  if (!(delta >= 0.f)) { set_error("delta must be >= 0"); return UQK_E_BADARG; }
  if (mode_de) {
    if (n_groups < 1 || n_groups > h->n_members) { set_error("n_groups=%d but the handle holds %d members", n_groups, h->n_members); return UQK_E_BADARG; }
    bn_mode = UQK_BN_INFERENCE;
  } else {
    if (n_groups < 1 || n_groups > 8192) { set_error("n_groups must be in 1..8192"); return UQK_E_BADARG; }
    if (bn_mode == UQK_BN_BATCH && (h->exact_passes != n_groups || h->exact_seed != seed || h->exact_n != n)) {
      set_error("uqk_refine_band(UQK_BN_BATCH) must follow the uqk_mcd(UQK_BN_BATCH, UQK_PREC_BF16_TC) call it refines "
                "(same windows, passes and seed): it reuses that call's per-pass batch statistics");
      return UQK_E_STATE;
    }
    if (bn_mode != UQK_BN_INFERENCE && bn_mode != UQK_BN_CALIBRATED && bn_mode != UQK_BN_BATCH) { set_error("unknown bn_mode %d", bn_mode); return UQK_E_BADARG; }
    if (bn_mode == UQK_BN_CALIBRATED && !h->have_frozen) { set_error("UQK_BN_CALIBRATED needs uqk_calibrate() or uqk_set_bn_stats() first"); return UQK_E_STATE; }
  }
  int rc = check_x(x_dev, x_dtype, n);
  if (rc) return rc;
  if (n == 0) return UQK_OK;
  UQK_CUDA(cudaSetDevice(h->device));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if ((rc = ensure(&h->d_band_idx, &h->band_idx_cap, (size_t)n * 2))) return rc;
  if (!h->d_band_count) UQK_CUDA(cudaMalloc(&h->d_band_count, sizeof(unsigned long long)));
  int64_t* idx = reinterpret_cast<int64_t*>(h->d_band_idx);
  if ((rc = launch_select_band(out->mean_pred, n, delta, idx, h->d_band_count, s))) return rc;
  unsigned long long cnt = 0;
  UQK_CUDA(cudaMemcpyAsync(&cnt, h->d_band_count, sizeof(cnt), cudaMemcpyDeviceToHost, s));
  UQK_CUDA(cudaStreamSynchronize(s));
  const int64_t m = (int64_t)cnt;
  if (n_refined_host) *n_refined_host = m;
  if (m == 0) return UQK_OK;
  if ((rc = ensure(&h->d_probs_tmp, &h->probs_cap, (size_t)m * n_groups))) return rc;
  // UQK_BN_BATCH: the statistics are those of the whole call, measured by the swept run (its per-pass tables)
  const float* tabs = (!mode_de && bn_mode == UQK_BN_BATCH) ? h->d_tables_all : nullptr;
  rc = run_layersync(h, 0, x_dev, x_dtype, m, n_groups, mode_de != 0, mode_de == 0, tabs ? UQK_BN_INFERENCE : bn_mode, seed, window_base,
                     mask_bits_dev, h->d_probs_tmp, nullptr, nullptr, s, idx, n, tabs);
  if (rc) return rc;
  if ((rc = launch_reduce(h->d_probs_tmp, n_groups, m, to_metrics(out), s, idx))) return rc;
  if (probs_dev && (rc = launch_scatter_probs(h->d_probs_tmp, idx, n_groups, m, n, probs_dev, s))) return rc;
  return UQK_OK;
}

int uqk_poll_error(uqk_handle* hh) {
  if (!hh) { set_error("handle is NULL"); return UQK_E_BADARG; }
  return poll_kernel_flag(reinterpret_cast<Handle*>(hh));
}

int uqk_debug_trip_watchdog(uqk_handle* hh) {
  if (!hh) { set_error("handle is NULL"); return UQK_E_BADARG; }
  Handle* h = reinterpret_cast<Handle*>(hh);
  UQK_CUDA(cudaSetDevice(h->device));
  trip_watchdog_kernel<<<1, 1>>>(h->d_err);
  cudaDeviceSynchronize();      // returns an error after the trap; the flag in host memory tells why
  cudaGetLastError();
  return poll_kernel_flag(h);
}

int uqk_reduce(const float* probs_dev, int n_passes, int64_t n, const uqk_metrics_out* out, void* stream) {
  if (n < 0 || n_passes < 1) { set_error("need n_windows >= 0 and n_passes >= 1"); return UQK_E_BADARG; }
  if (n > 0 && (!probs_dev || !out)) { set_error("NULL argument"); return UQK_E_BADARG; }
  if (n == 0) return UQK_OK;
  return launch_reduce(probs_dev, n_passes, n, to_metrics(out), static_cast<cudaStream_t>(stream));
}

// debug: copy the fused kernel's activation dump (6,128,256) to the host
int uqk_debug_read_dump(uqk_handle* hh, float* host) {
  if (!hh || !host) { set_error("NULL argument"); return UQK_E_BADARG; }
  Handle* h = reinterpret_cast<Handle*>(hh);
  UQK_CUDA(cudaSetDevice(h->device));
  UQK_CUDA(cudaDeviceSynchronize());
  UQK_CUDA(cudaMemcpy(host, h->d_dbg, sizeof(float) * kBlocks * 128 * 256, cudaMemcpyDeviceToHost));
  return UQK_OK;
}

}  // extern "C"
