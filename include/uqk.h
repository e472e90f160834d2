/*
 * uqk.h -- C-ABI of libuqk.so: the B200 (sm_100a) UQ-inference hot path of
 * TrondVQ/UncertaintyQuantification-SleepApnea-1DCNN.
 *
 * The reference has no FFI layer of its own: the path sits behind the Python module
 * uncertainty_quantification/uq_techniques.py, which delegates the arithmetic to TensorFlow/Keras, NumPy and
 * SciPy.  Each entry point below cites the reference call site(s) it replaces.  All pointers named *_dev are
 * device pointers owned by the caller; the library allocates only its private workspace inside the handle
 * (packed weights, BatchNorm statistics, activation scratch for the layer-synchronous path) and frees it in
 * uqk_destroy().  Every function returns 0 on success or a negative UQK_E_* code; uqk_last_error() gives the
 * text (thread-local).  Nothing throws across the boundary.  There is no CPU fallback: without an sm_100
 * device every compute entry point fails with UQK_E_ARCH / UQK_E_CUDA.
 *
 * A handle is bound to one device and is not thread-safe.  Compute entry points enqueue work on the given
 * stream and return without synchronising, except where noted.
 */
#ifndef UQK_H
#define UQK_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define UQK_VERSION 101

#define UQK_OK 0
#define UQK_E_BADARG (-1)
#define UQK_E_SHAPE (-2)
#define UQK_E_ARCH (-3)   /* device is not sm_100 (B200) */
#define UQK_E_CUDA (-4)
#define UQK_E_STATE (-5)  /* e.g. model not set, statistics not calibrated */
#define UQK_E_NOMEM (-6)
#define UQK_E_KERNEL (-7) /* device-side watchdog / self-check failed, see uqk_poll_error */

/* dtype of the window tensor x (N,60,4): prepare_numpy_datasets.py:93,233 saves float64 */
#define UQK_F32 0
#define UQK_F64 1

/* BatchNormalization statistics used by the stochastic (training=True) passes */
#define UQK_BN_INFERENCE 0  /* moving statistics (frozen BN; what model.predict uses, uq_techniques.py:29) */
#define UQK_BN_BATCH 1      /* exact Keras training=True semantics: batch mean / biased variance over (N,L) of the
                               whole call, per pass (uq_techniques.py:22).  UQK_PREC_FP32: layer-synchronous kernels
                               (activations of the call in HBM).  UQK_PREC_BF16_TC: one tensor-core sweep per
                               layer: sweep k reads layer k-1's post-ReLU output (bf16) from HBM, normalises it
                               with the statistics sweep k-1 produced, computes layer k once, collects its per-pass
                               statistics and stores its output; the sixth sweep leaves what the linear head needs */
#define UQK_BN_CALIBRATED 2 /* batch statistics estimated once by uqk_calibrate() and frozen for all passes */

/* arithmetic of the forward passes */
#define UQK_PREC_BF16_TC 0  /* fused tcgen05 kernel: bf16 operands, fp32 accumulate in TMEM */
#define UQK_PREC_FP32 1     /* layer-synchronous CUDA-core kernels, fp32 throughout */

typedef struct uqk_handle uqk_handle;

/* Per-window outputs, each a device array of N elements; any pointer may be NULL (not produced).
 * Replaces the arrays of uq_evaluation_dist (uq_techniques.py:69-91,103-108) and the drivers' inline
 * per-window metrics (analyze_mcd_patient_level.py:109-117; analyze_de_patient_level.py:124-132). */
typedef struct uqk_metrics_out {
  float* mean_pred;                  /* uq_techniques.py:69 */
  float* pred_variance;              /* :72, ddof=0, two-pass */
  float* total_pred_entropy;         /* :78-79  H[(1-mean, mean)] in nats, clip 1e-10, renormalised */
  float* expected_aleatoric_entropy; /* :83-87  mean_t H[(1-p_t, p_t)] */
  float* mutual_info;                /* :91     max(total - expected, 0) */
  float* entropy_bits;               /* analyze_mcd_patient_level.py:113-115  log2 form with +1e-9 */
  int32_t* label;                    /* analyze_mcd_patient_level.py:117      mean > 0.5 */
} uqk_metrics_out;

int uqk_version(void);
const char* uqk_last_error(void);

/* Handle life cycle.  Synchronous. */
int uqk_create(int device, uqk_handle** out);
int uqk_destroy(uqk_handle* h);

/* Install n_members networks.  blobs_host: n_members x 853441 float32, each in Keras get_weights() order
 * (per block: conv kernel (k,Cin,Cout), bias, gamma, beta, moving_mean, moving_variance; then dense kernel
 * (96,1), bias) -- the arrays tensorflow.keras.models.load_model() would hand to the reference
 * (analyze_mcd_patient_level.py:199; analyze_de_patient_level.py:48).  Synchronous. */
int uqk_set_model(uqk_handle* h, int n_members, const float* blobs_host, size_t floats_per_member);

/* Deterministic forward, BN in inference mode, no dropout: replaces cnn_model(X, training=False)
 * (analyze_mcd_patient_level.py:204) and model.predict (evaluation/evaluate_classification.py:30-31).
 * probs_dev: (N,) float32. */
int uqk_forward_det(uqk_handle* h, int member, const void* x_dev, int x_dtype, int64_t n_windows, int precision,
                    float* probs_dev, void* stream);

/* Estimate the per-layer batch statistics that Keras training=True would use, on n_windows windows and
 * n_passes stochastic passes, average them over the passes and freeze them for UQK_BN_CALIBRATED.  Up to 8192
 * windows (option "calib_fp32_max") the exact fp32 layer-synchronous kernels compute them; larger sets go through
 * the statistics sweeps of the tensor-core kernel (same definition, bf16 operands).  window_index_base = position of the
 * first window in the call (dropout masks are keyed by it, as in uqk_mcd).  With a statistics exchange installed
 * (uqk_set_stats_exchange) every rank passes ITS shard and the frozen statistics are those of the windows of all ranks
 * (the fp32 / tensor-core choice then follows the global window count, so all ranks take the same path).
 * Enqueued on stream. */
int uqk_calibrate(uqk_handle* h, const void* x_dev, int x_dtype, int64_t n_windows, int n_passes, uint64_t seed,
                  int64_t window_index_base, void* stream);
/* Read back / install the frozen statistics: mean and var are host arrays of 6 x 256 floats (layer-major,
 * only the first Cout entries of each layer are used).  Synchronous on the handle's device. */
int uqk_get_bn_stats(uqk_handle* h, float* mean_host, float* var_host);
int uqk_set_bn_stats(uqk_handle* h, const float* mean_host, const float* var_host);

/* Exact batch-statistic BatchNorm over windows that are sharded across GPUs ("the exception").
 * Keras normalises `model(x, training=True)` (uq_techniques.py:22) with the statistics of the WHOLE call, so when
 * the windows of one call live on several GPUs every layer of every pass needs the per-channel sums of all ranks.
 * With an exchange installed, every UQK_BN_BATCH computation of this handle (uqk_mcd, uqk_calibrate) calls, on the
 * host thread of that call,
 *       fn(user, stats_dev, count, stream)
 * after a layer's statistics have been enqueued -- once per layer and pass with count = 2*Cout on the
 * layer-synchronous path, once per layer with count = n_passes*512 on the swept tensor-core path.  stats_dev holds
 * `count` int64 values on the device ((sum, sum of squares) per channel in 2^-24 fixed point) and fn must enqueue on
 * `stream` an in-place SUM
 * all-reduce over the ranks (e.g. ncclAllReduce(ncclInt64, ncclSum) / torch.distributed.all_reduce) and return 0.
 * Integer sums are order-independent, so the sharded result is bit-identical to one GPU holding all windows.
 * global_windows = number of windows of the whole call (the divisor of the statistics); every rank must make the
 * same sequence of calls, including ranks whose shard is empty.  fn == NULL removes the exchange. */
typedef int (*uqk_stats_exchange_fn)(void* user, long long* stats_dev, int count, void* stream);
int uqk_set_stats_exchange(uqk_handle* h, uqk_stats_exchange_fn fn, void* user, int64_t global_windows);

/* MC-Dropout: n_passes stochastic passes of member 0 over all windows + the per-window reductions.
 * Replaces mc_dropout_predict (uq_techniques.py:12-24) fused with uq_evaluation_dist (:40-112).
 *   seed              Philox4x32-10 key; masks are a pure function of (seed, window_index_base + n, pass,
 *                     layer, row, channel), hence independent of sharding.
 *   mask_bits_dev     NULL, or explicit keep-masks (n_passes, N, 6, 60, 8) uint32, bit c%32 of word c/32 --
 *                     the mask-injection mode used for exact parity with the oracle.
 *   probs_dev         NULL, or (n_passes, N) float32: the tensor mc_dropout_predict returns.
 *   out               NULL, or the per-window metrics. */
int uqk_mcd(uqk_handle* h, const void* x_dev, int x_dtype, int64_t n_windows, int n_passes, uint64_t seed,
            int bn_mode, int precision, int64_t window_index_base, const uint32_t* mask_bits_dev,
            float* probs_dev, const uqk_metrics_out* out, void* stream);

/* Deep Ensemble: members [0, n_members) in inference mode + the per-window reductions.
 * Replaces deep_ensembles_predict (uq_techniques.py:26-32) fused with uq_evaluation_dist.
 * probs_dev: NULL or (n_members, N). */
int uqk_de(uqk_handle* h, const void* x_dev, int x_dtype, int64_t n_windows, int n_members, int precision,
           float* probs_dev, const uqk_metrics_out* out, void* stream);

/* Guard band of the labels.  The reference's label is `mean_p > 0.5` in fp32 (analyze_mcd_patient_level.py:117,
 * analyze_de_patient_level.py:132, evaluate_classification.py:49); with bf16 operands a window whose mean probability
 * lies within the arithmetic error of 0.5 can land on the other side.  After uqk_mcd / uqk_de on the bf16 path, this
 * call re-evaluates every window with |out->mean_pred - 0.5| < delta on the fp32 layer-synchronous path -- same
 * windows, same dropout masks (seed, window_index_base + position, or mask_bits_dev), same BatchNorm statistics
 * (UQK_BN_CALIBRATED / UQK_BN_INFERENCE: the frozen ones; UQK_BN_BATCH: the per-pass batch statistics of the whole call
 * as the preceding uqk_mcd(UQK_BN_BATCH, UQK_PREC_BF16_TC) measured them -- it must be the handle's last such call) -- and
 * overwrites ALL of that window's outputs (and its column of probs_dev, when given) with the fp32 results.
 *   mode_de = 0: the n_groups passes of uqk_mcd;  mode_de = 1: the n_groups members of uqk_de (seed, bn_mode, masks unused).
 *   n_refined_host: receives the number of re-evaluated windows.
 * SYNCHRONOUS (the band population is read back to size the launches).  out->mean_pred must be non-NULL. */
int uqk_refine_band(uqk_handle* h, int mode_de, const void* x_dev, int x_dtype, int64_t n_windows, int n_groups,
                    uint64_t seed, int bn_mode, int64_t window_index_base, const uint32_t* mask_bits_dev, float delta,
                    float* probs_dev, const uqk_metrics_out* out, int64_t* n_refined_host, void* stream);

/* Device-side self-checks.  The kernels report through a flag in mapped pinned host memory, so the reason survives even
 * a device trap.  Every compute entry point of the handle checks the flag first and returns UQK_E_KERNEL when it is
 * set; call uqk_poll_error after synchronising a stream to learn about a failure of work that was still in flight.
 *   codes 1..23: barrier watchdog -- a warp of a fused kernel waited ~2 s (4e9 clock ticks) at one mbarrier, which
 *     only a protocol fault can cause; the kernel then TRAPS, the CUDA context of the process is lost (every later
 *     CUDA call fails) and the process must be restarted.  The budget is fixed: under cuda-gdb, a sanitizer or heavy
 *     time-slicing a healthy kernel could exceed it -- uqk_set_option(h, "watchdog", 0) switches the budget off (the
 *     kernels then wait indefinitely, as ordinary CUDA kernels do).
 *   code 50: a batch-statistics accumulator (2^-24 fixed point in 64 bits) overflowed -- sum of squares of one
 *     channel above 5.5e11; no trap, the flag is cleared when reported, the results of that call are invalid.
 * uqk_debug_trip_watchdog (tests): runs a kernel that waits on a barrier nobody signals; returns UQK_E_KERNEL and
 * leaves the context unusable -- call it only in a process that exits afterwards. */
int uqk_poll_error(uqk_handle* h);
int uqk_debug_trip_watchdog(uqk_handle* h);

/* Stand-alone reductions over an existing (n_passes, N) float32 probability tensor.
 * Replaces uq_evaluation_dist (uq_techniques.py:40-112) and the inline metrics. */
int uqk_reduce(const float* probs_dev, int n_passes, int64_t n_windows, const uqk_metrics_out* out, void* stream);

/* Per-patient aggregation of per-window rows.  Replaces the pandas groupby of
 * aggregate_patient_uq_metrics.py:34-46.  Rows of patient p are seg_offsets_dev[p] .. seg_offsets_dev[p+1]-1
 * of the (already patient-sorted) columns.  out_dev: (n_patients, 8) float64 rows =
 * [mean_var, median_var, std_var(ddof=1; 0 if n<=1), mean_ent, median_ent, std_ent, accuracy, num_windows].
 * SYNCHRONOUS: a patient with more than 8192 windows does not fit the in-kernel sort; the kernel's flag is read back
 * and the call returns UQK_E_SHAPE (the rows of such patients are NaN) instead of failing later. */
int uqk_patient_reduce(const float* variance_dev, const float* entropy_dev, const int32_t* true_label_dev,
                       const int32_t* pred_label_dev, const int64_t* seg_offsets_dev, int64_t n_patients,
                       double* out_dev, void* stream);

/* Bootstrap aggregates behind bootstrap_metrics (uq_techniques.py:116-172): for each of n_boot index vectors
 * idx_dev[b, :] (drawn on the host with np.random.choice exactly as :142), the six aggregate means of
 * :150-157 computed as gather-means of the per-window metrics.  out_dev: (n_boot, 6) float64 in the order
 * overall_mean_variance, mean_variance_class_0, mean_variance_class_1, mean_total_pred_entropy,
 * mean_expected_aleatoric_entropy, mean_mutual_info. */
int uqk_bootstrap(const float* variance_dev, const float* total_entropy_dev, const float* expected_entropy_dev,
                  const float* mutual_info_dev, const int32_t* y_dev, const int64_t* idx_dev, int n_boot,
                  int64_t n_windows, double* out_dev, void* stream);

/* Per-window standardisation (x - mean_t) / (std_t + 1e-8), population std over the 60 time steps of every
 * (window, channel), computed in float64.  Replaces standardize_per_window
 * (data_prepocessing/prepare_numpy_datasets.py:83-95), the step immediately before the hot path.
 * x_dev / out_dev: (N,60,4) of x_dtype (out may alias x). */
int uqk_standardize(const void* x_dev, int x_dtype, int64_t n_windows, void* out_dev, void* stream);

/* Allocation-free variant of uqk_bootstrap (no counterpart in the reference: it exists for callers of THIS library).
 * For processes that must not allocate device memory while running -- e.g. a server that gave the whole GPU to its
 * own arena or to a framework's caching allocator at start-up, or a latency-critical loop where cudaMalloc/cudaFree
 * are banned because they can synchronise the device.  Same arguments and results as uqk_bootstrap, but the
 * (n_boot x 8 doubles) scratch lives in a fixed 64 KiB device array inside libuqk.so (held by every GPU that loads
 * the library until the process exits); the call makes no cudaMalloc/cudaFree/cudaMallocAsync.
 *   - n_boot <= 1024 per call; larger values are rejected with UQK_E_SHAPE (split the iterations over several calls).
 *   - The first call on each device resolves the array's device address (cudaGetSymbolAddress); make that first
 *     call before starting any stream capture.
 *   - All calls on a device share the one array: it is cleared at the start of every call, and two calls that
 *     overlap on the GPU (different streams / threads) race on it -- serialise them. */
int uqk_bootstrap_noalloc(const float* variance_dev, const float* total_entropy_dev, const float* expected_entropy_dev,
                          const float* mutual_info_dev, const int32_t* y_dev, const int64_t* idx_dev, int n_boot,
                          int64_t n_windows, double* out_dev, void* stream);

/* Diagnostics: number of kernel launches the library has issued in this PROCESS since it was loaded (one counter for all
 * handles; the handle argument is accepted for symmetry and may be NULL). */
int64_t uqk_launch_count(uqk_handle* h);
/* N2 -- the per-window detail CSV (analyze_mcd_patient_level.py:134-152 builds the DataFrame and calls
 * to_csv(path, index=False); analyze_de_patient_level.py does the same).  HOST function: writes, byte for byte, what
 * pandas writes for int64 Patient_ID / Window_Index / True_Label, int Predicted_Label and float32 Predicted_Probability,
 * Predictive_Variance, Predictive_Entropy (NumPy float32 repr: shortest round-trip digits, scientific below 1e-4 and
 * from 1e6, NaN -> empty field), formatting row blocks on n_threads host threads (0 = all cores, at most 32).
 * Window_Index = window_index_base + row.  write_header = 0 appends rows to an existing file (shards written in turn).
 * All pointers are HOST arrays of n elements. */
int uqk_write_detail_csv(const char* path, int64_t n, const int64_t* patient_id, int64_t window_index_base,
                         const int64_t* true_label, const int32_t* pred_label, const float* prob, const float* variance,
                         const float* entropy, int write_header, int n_threads);

/* N2 -- GPU twin of analyze_window_level_uncertainty.py:30-64 on the device arrays the hot path produced: overall
 * accuracy, pandas describe() (count, mean, std ddof=1, min, 25%, 50%, 75%, max) of Predictive_Entropy and
 * Predictive_Variance for correctly / incorrectly classified windows, and the accuracy-vs-entropy table over
 * n_bins (<= 64) equal-width bins [np.linspace(min, max + 1e-9, n_bins + 1), right-open].  Float64 arithmetic.
 * out_dev: 35 + 3*n_bins doubles on the device:
 *   [0] windows  [1] accuracy
 *   [2 + 16*c + 8*m + k]   c: 0 correct, 1 incorrect;  m: 0 entropy, 1 variance;  k: the eight describe() rows
 *   [34 .. 34+n_bins]      bin edges;  then n_bins window counts;  then n_bins accuracies (NaN for an empty bin)
 * Stream-ordered scratch (cudaMallocAsync); enqueued on stream. */
int uqk_window_summary(const float* entropy_dev, const float* variance_dev, const int32_t* true_label_dev,
                       const int32_t* pred_label_dev, int64_t n_windows, int n_bins, double* out_dev, void* stream);

/* Handle options (all integers):
 *   "variant"          fused-kernel selector for tests and probes: value >> 4 = 0 two-tile CTA pairs (default), 1 one tile per
 *                      CTA (the only kernel with the activation dump), 2 one-tile CTA pairs, 4 two tiles per CTA;
 *                      bit 0 = release/acquire cluster signalling; bits 1..3 = perf-decomposition switches (results WRONG)
 *   "watchdog"         1 (default) / 0: time budget of the kernels' barrier waits, see uqk_poll_error
 *   "calib_fp32_max"   calibration sets up to this many windows use the fp32 kernels (default 8192)
 *   "exact_chunk_max"  cap on the passes per chunk of the exact mode (0 = by memory budget)
 *   "exact_pipelined"  1 (default): resumed sweeps of the exact mode as a three-stage pipeline over 4-window units;
 *                      0: both tiles of a CTA in lock-step over 8-window units (bit-identical, 5 % slower)
 *   "dbg_dump"         1: the one-tile kernel dumps the activations of its first unit (uqk_debug_read_dump) */
int uqk_set_option(uqk_handle* h, const char* key, int64_t value);

#ifdef __cplusplus
}
#endif
#endif /* UQK_H */
