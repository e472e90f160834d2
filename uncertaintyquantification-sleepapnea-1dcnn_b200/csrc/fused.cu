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
//    bias+ReLU -> BN affine -> Philox dropout -> bf16.  One 68 KB buffer per tile instead of two.
//  * Weights (1.7 MB bf16 per member, L2 resident) stream through a shared-memory ring with cp.async.bulk (TMA
//    bulk copies, mbarrier complete_tx); they are pre-packed on the host into the exact shared-memory image.
//  * Warp roles: warp 0 = weight producer, warp 1 = TMEM allocator + MMA issuer (one thread), 4 warps per tile =
//    epilogue (TMEM lane quarter = warp_id % 4).  Two CTAs are resident per SM (TILES=1), so one CTA's epilogue
//    overlaps the other's MMAs.
#include "common.cuh"
#include "internal.h"

#include <vector>
#include <cstring>

namespace uqk {

// ---- geometry ---------------------------------------------------------------------------------------------------
constexpr int kRowsPad = 136;                      // rows -4 .. 131 of a tile (4 zero rows above and below)
constexpr int kPlaneBytes = kRowsPad * 16;         // one 8-channel chunk, all rows: 2176 B
constexpr int kActBytes = (kMaxCh / 8) * kPlaneBytes;  // 69632 B per tile
constexpr int kSlotBytes = 16384;                  // weight ring slot

struct FusedLayer {
  int cinp;        // input channels as seen by the MMA (layer 1: 4 real + 4 zero, two taps per K=16 step)
  int cout;
  int taps;
  int halo;
  int nsteps;      // K=16 MMA steps per tile
  int step_bytes;  // bytes of weights per step = cout * 32
  int sps;         // steps per ring stage
  int nstages;
  uint32_t w_off;  // byte offset of the layer inside one member's packed image
};

__host__ __device__ constexpr FusedLayer fused_layer(int l) {
  // nsteps: layer 1 = 4 tap pairs; others taps * cin / 16
  return l == 0 ? FusedLayer{8, 128, 7, 3, 4, 4096, 4, 1, 0u}
       : l == 1 ? FusedLayer{128, 192, 5, 2, 40, 6144, 2, 20, 16384u}
       : l == 2 ? FusedLayer{192, 224, 3, 1, 36, 7168, 2, 18, 16384u + 245760u}
       : l == 3 ? FusedLayer{224, 96, 7, 3, 98, 3072, 5, 20, 16384u + 245760u + 258048u}
       : l == 4 ? FusedLayer{96, 256, 9, 4, 54, 8192, 2, 27, 16384u + 245760u + 258048u + 301056u}
                : FusedLayer{256, 96, 9, 4, 144, 3072, 5, 29, 16384u + 245760u + 258048u + 301056u + 442368u};
}
constexpr uint32_t kPackedBytes = 16384u + 245760u + 258048u + 301056u + 442368u + 442368u;  // 1,705,984

size_t fused_packed_bytes() { return kPackedBytes; }

static inline uint16_t f32_to_bf16_rne(float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  if ((u & 0x7F800000u) == 0x7F800000u) return (uint16_t)(u >> 16);  // inf / nan
  u += 0x7FFFu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

// Host-side packing of one member into the shared-memory image the kernel streams:
//   layer l, step s, plane p (K elements 8p..8p+7 of the step), output channel n, element kk:
//   dst[(w_off/2) + ((s*2 + p)*cout + n)*8 + kk]
//   layers 2..6: tap j = s / (cin/16), slab i = s % (cin/16), channel c = 16 i + 8 p + kk  -> W[j][c][n]
//   layer 1    : tap j = 2 s + p, channel kk (<4)                                        -> W[j][kk][n], else 0
void fused_pack_member(const float* blob, uint16_t* dst) {
  const BlobOffsets off = blob_offsets();
  for (int l = 0; l < kBlocks; ++l) {
    const FusedLayer fl = fused_layer(l);
    const BlockDef b = block_def(l);
    const float* w = blob + off.conv_w[l];  // (k, cin, cout)
    uint16_t* d = dst + fl.w_off / 2;
    for (int s = 0; s < fl.nsteps; ++s)
      for (int p = 0; p < 2; ++p)
        for (int n = 0; n < fl.cout; ++n)
          for (int kk = 0; kk < 8; ++kk) {
            float v = 0.f;
            if (l == 0) {
              const int j = 2 * s + p;
              if (j < b.k && kk < b.cin) v = w[((size_t)j * b.cin + kk) * b.cout + n];
            } else {
              const int slabs = b.cin / 16;
              const int j = s / slabs, i = s % slabs;
              const int c = 16 * i + 8 * p + kk;
              v = w[((size_t)j * b.cin + c) * b.cout + n];
            }
            d[((size_t)(s * 2 + p) * fl.cout + n) * 8 + kk] = f32_to_bf16_rne(v);
          }
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

// ---- PTX wrappers -----------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug must end the kernel with an error, never hang the GPU.
__device__ __noinline__ void mbar_timeout(int* err_flag, int code) {
  if (err_flag) atomicExch(err_flag, code);
  __threadfence_system();
  __trap();
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, int* err_flag, int code) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 4000000000ll) mbar_timeout(err_flag, code);
  }
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// UMMA shared-memory descriptor, K-major, SWIZZLE_NONE (cute::UMMA::SmemDescriptor bit layout):
//   [0,14) start>>4   [16,30) LBO>>4 (K-direction core-matrix stride)   [32,46) SBO>>4 (8-row group stride)
//   [46,48) version = 1 (Blackwell)   [61,64) layout type = 0
__device__ __forceinline__ uint64_t umma_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
// Instruction descriptor, kind::f16: D=f32 (bit 4), A=B=bf16 (bits 7,10), K-major A and B, N>>3 at 17, M>>4 at 24
__device__ __forceinline__ uint32_t umma_idesc(int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---- the kernel -------------------------------------------------------------------------------------------------
struct FusedArgs {
  const void* x;
  int x_f64;
  long long n;
  int n_groups;
  int mode_de;
  int dropout;
  uint32_t k0, k1;
  long long window_base;
  const uint32_t* mask_bits;
  const uint8_t* packed;
  const float* tables;
  float* probs;
  MetricsOut out;
  float* dbg_act;
  int* err_flag;
  int variant;
  long long n_units;
};

template <int TILES, int NSLOTS>
struct FusedCfg {
  static constexpr int kWarps = 2 + 4 * TILES;
  static constexpr int kThreads = 32 * kWarps;
  static constexpr int kWpu = 2 * TILES;  // windows per unit
  static constexpr int kTmemCols = 256 * TILES;
  // shared-memory map (bytes)
  static constexpr int oAct = 0;
  static constexpr int oRing = oAct + TILES * kActBytes;
  static constexpr int oTab = oRing + NSLOTS * kSlotBytes;           // per tile: bias/scale/shift[256] + dense[100]
  static constexpr int kTabBytes = (3 * 256 + 128) * 4;
  static constexpr int oBar = oTab + TILES * kTabBytes;              // mbarriers
  static constexpr int kNumBars = 2 * NSLOTS + 2 * TILES;
  static constexpr int oMisc = oBar + 8 * kNumBars;                  // tmem ptr, zsum[TILES][2]
  static constexpr int oProbs = oMisc + 16 + 8 * TILES;              // probs[kWpu][n_groups]
};

template <int TILES, int NSLOTS>
__global__ void __launch_bounds__(FusedCfg<TILES, NSLOTS>::kThreads, TILES == 1 ? 2 : 1)
fused_forward_kernel(const FusedArgs a) {
  using Cfg = FusedCfg<TILES, NSLOTS>;
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t s_base = smem_u32(smem);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const uint32_t bar_full = s_base + Cfg::oBar;                 // [NSLOTS]
  const uint32_t bar_empty = bar_full + 8 * NSLOTS;             // [NSLOTS]
  const uint32_t bar_acc = bar_empty + 8 * NSLOTS;              // [TILES]  accumulator complete (MMA -> epilogue)
  const uint32_t bar_act = bar_acc + 8 * TILES;                 // [TILES]  activations ready (epilogue -> MMA)
  volatile uint32_t* tmem_ptr = reinterpret_cast<volatile uint32_t*>(smem + Cfg::oMisc);
  float* zsum = reinterpret_cast<float*>(smem + Cfg::oMisc + 16);  // [TILES][2]
  float* probs_s = reinterpret_cast<float*>(smem + Cfg::oProbs);   // [kWpu][n_groups]

  // ---- one-time set-up
  for (int i = threadIdx.x; i < (TILES * kActBytes) / 16; i += Cfg::kThreads)
    reinterpret_cast<uint4*>(smem + Cfg::oAct)[i] = make_uint4(0, 0, 0, 0);   // halo rows stay zero for ever
  if (threadIdx.x < 2 * TILES) zsum[threadIdx.x] = 0.f;
  if (threadIdx.x == 0) {
    for (int i = 0; i < NSLOTS; ++i) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, 1); }
    for (int i = 0; i < TILES; ++i) { mbar_init(bar_acc + 8 * i, 1); mbar_init(bar_act + 8 * i, 128); }
    fence_barrier_init();
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(const_cast<uint32_t*>(tmem_ptr))),
                 "r"((uint32_t)Cfg::kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  fence_proxy_async();   // zero-filled activations must be visible to the tensor-core (async) proxy
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  // perf-decomposition switches (tests/tools only): results are WRONG when any is set
  const bool dbg_no_copy = (a.variant & 2) != 0;   // producer does not fetch weights
  const bool dbg_raw_epi = (a.variant & 4) != 0;   // epilogue skips bias/ReLU/BN/dropout math
  const bool dbg_no_mma = (a.variant & 8) != 0;    // issuer skips the MMAs

  if (warp == 0) {
    // =========================== weight producer (one thread) ===========================
    if (lane == 0) {
      uint32_t it = 0;
      for (long long unit = blockIdx.x; unit < a.n_units; unit += gridDim.x) {
        for (int g = 0; g < a.n_groups; ++g) {
          const uint8_t* wmem = a.packed + (size_t)(a.mode_de ? g : 0) * kPackedBytes;
          for (int l = 0; l < kBlocks; ++l) {
            const FusedLayer fl = fused_layer(l);
            int step = 0;
            for (int st = 0; st < fl.nstages; ++st, ++it) {
              const int slot = it % NSLOTS;
              const uint32_t ph = (it / NSLOTS) & 1;
              mbar_wait(bar_empty + 8 * slot, ph ^ 1, a.err_flag, 1);
              const int nst = min(fl.sps, fl.nsteps - step);
              const uint32_t bytes = (uint32_t)nst * fl.step_bytes;
              if (dbg_no_copy) { mbar_arrive(bar_full + 8 * slot); }
              else {
                mbar_arrive_expect_tx(bar_full + 8 * slot, bytes);
                bulk_g2s(s_base + Cfg::oRing + slot * kSlotBytes, wmem + fl.w_off + (size_t)step * fl.step_bytes, bytes,
                         bar_full + 8 * slot);
              }
              step += nst;
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer (one thread) ===========================
    if (lane == 0) {
      uint32_t it = 0, n_act = 0;
      for (long long unit = blockIdx.x; unit < a.n_units; unit += gridDim.x) {
        for (int g = 0; g < a.n_groups; ++g) {
          for (int l = 0; l < kBlocks; ++l) {
            const FusedLayer fl = fused_layer(l);
            const uint32_t idesc = umma_idesc(fl.cout);
            const int slabs = l == 0 ? 1 : fl.cinp / 16;
            for (int tl = 0; tl < TILES; ++tl) mbar_wait(bar_act + 8 * tl, n_act & 1, a.err_flag, 2);
            ++n_act;
            tc_fence_after();
            int step = 0, tap = 0, slab = 0;
            for (int st = 0; st < fl.nstages; ++st, ++it) {
              const int slot = it % NSLOTS;
              const uint32_t ph = (it / NSLOTS) & 1;
              mbar_wait(bar_full + 8 * slot, ph, a.err_flag, 3);
              tc_fence_after();
              const int nst = min(fl.sps, fl.nsteps - step);
              const uint32_t bslot = s_base + Cfg::oRing + slot * kSlotBytes;
              for (int sl = 0; sl < nst; ++sl, ++step) {
                // A: rows shifted by the tap; B: this step's two K planes
                uint32_t a_off, a_lbo;
                if (l == 0) { a_off = (uint32_t)(2 * step + 1) * 16u; a_lbo = 16u; }
                else { a_off = (uint32_t)(2 * slab) * kPlaneBytes + (uint32_t)(tap - fl.halo + 4) * 16u; a_lbo = kPlaneBytes; }
                const uint32_t b_addr = bslot + (uint32_t)sl * fl.step_bytes;
                const uint32_t b_lbo = (uint32_t)fl.cout * 16u;
                const uint64_t bdesc = umma_desc(b_addr, b_lbo, 128u);
#pragma unroll
                for (int tl = 0; tl < TILES; ++tl) {
                  const uint32_t a_addr = s_base + Cfg::oAct + tl * kActBytes + a_off;
                  const uint64_t adesc = umma_desc(a_addr, a_lbo, 128u);
                  if (!dbg_no_mma) umma_bf16(tmem_base + tl * 256, adesc, bdesc, idesc, step > 0 ? 1u : 0u);
                }
                if (++slab == slabs) { slab = 0; ++tap; }
              }
              umma_commit(bar_empty + 8 * slot);   // frees the ring slot when these MMAs have read it
            }
            for (int tl = 0; tl < TILES; ++tl) umma_commit(bar_acc + 8 * tl);   // accumulator(s) complete
          }
        }
      }
    }
  } else {
    // =========================== epilogue warps: 4 per tile ===========================
    const int tl = (warp - 2) >> 2;
    const int quarter = warp & 3;                 // TMEM lanes [32*quarter, +32) are accessible to this warp
    const int row = quarter * 32 + lane;          // accumulator row = TMEM lane
    const int slot_w = row >> 6;                  // which of the tile's two windows
    const int t = row & 63;                       // time step (60..63 = zero rows)
    const int tid_t = (warp - 2 - 4 * tl) * 32 + lane;  // 0..127 within the tile's epilogue group
    float* tab = reinterpret_cast<float*>(smem + Cfg::oTab + tl * Cfg::kTabBytes);
    const uint32_t act_row = s_base + Cfg::oAct + tl * kActBytes + (uint32_t)(row + 4) * 16u;
    const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(tl * 256);
    uint32_t n_acc = 0;

    for (long long unit = blockIdx.x; unit < a.n_units; unit += gridDim.x) {
      const long long win = unit * Cfg::kWpu + 2 * tl + slot_w;
      const bool row_valid = (t < kL) && (win < a.n);
      // this row's 4 input features as bf16 (kept in registers for all groups of the unit)
      uint32_t xin0 = 0, xin1 = 0;
      if (row_valid) {
        float f0, f1, f2, f3;
        if (a.x_f64) {
          const double2* p = reinterpret_cast<const double2*>(static_cast<const double*>(a.x) + ((size_t)win * kL + t) * kF);
          const double2 u = __ldg(p), v = __ldg(p + 1);
          f0 = (float)u.x; f1 = (float)u.y; f2 = (float)v.x; f3 = (float)v.y;
        } else {
          const float4 u = __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(a.x) + ((size_t)win * kL + t) * kF));
          f0 = u.x; f1 = u.y; f2 = u.z; f3 = u.w;
        }
        xin0 = pack_bf16x2(f0, f1);
        xin1 = pack_bf16x2(f2, f3);
      }
      const uint32_t c0w = (uint32_t)(a.window_base + win);

      for (int g = 0; g < a.n_groups; ++g) {
        // ---- layer-1 input: channels 0..3 real, 4..7 zero, in plane 0
        st_shared_v4(act_row, xin0, xin1, 0u, 0u);
        fence_proxy_async();
        mbar_arrive(bar_act + 8 * tl);

        const float* tab_g = a.tables + (size_t)(a.mode_de ? g : 0) * kTableFloats;
        for (int l = 0; l < kBlocks; ++l) {
          const FusedLayer fl = fused_layer(l);
          const int cout = fl.cout;
          // stage this layer's bias / scale / shift (and the dense layer for l == 5) while the MMAs run
          named_bar_sync(1 + tl, 128);   // previous layer's readers of `tab` are done
          for (int i = tid_t; i < cout; i += 128) {
            tab[i] = __ldg(tab_g + ch_off(l) + i);
            tab[256 + i] = __ldg(tab_g + kChTotal + ch_off(l) + i);
            tab[512 + i] = __ldg(tab_g + 2 * kChTotal + ch_off(l) + i);
          }
          if (l == kBlocks - 1 && tid_t < kDenseIn + 1) tab[768 + tid_t] = __ldg(tab_g + 3 * kChTotal + tid_t);
          named_bar_sync(1 + tl, 128);

          mbar_wait(bar_acc + 8 * tl, n_acc & 1, a.err_flag, 4);
          ++n_acc;
          tc_fence_after();

          const uint32_t thr = drop_thr(l);
          const uint32_t* mrow = nullptr;
          if (a.dropout && a.mask_bits && row_valid)
            mrow = a.mask_bits + ((((size_t)g * a.n + win) * kBlocks + l) * kL + t) * 8;
          float* dbg = nullptr;
          if (a.dbg_act && blockIdx.x == 0 && unit == blockIdx.x && g == 0 && tl == 0)
            dbg = a.dbg_act + ((size_t)l * 128 + row) * 256;
          float zacc = 0.f;

          for (int c0 = 0; c0 < cout; c0 += 32) {
            uint32_t v[32];
            tmem_ld32(taddr + (uint32_t)c0, v);
            tmem_ld_wait();
            uint32_t mword = 0xFFFFFFFFu;
            if (a.dropout && a.mask_bits) mword = mrow ? __ldg(mrow + (c0 >> 5)) : 0u;
            if (dbg_raw_epi) {
#pragma unroll
              for (int q = 0; q < 4; ++q)
                st_shared_v4(act_row + (uint32_t)((c0 >> 3) + q) * kPlaneBytes, v[8 * q] & 0x3F803F80u, v[8 * q + 2] & 0x3F803F80u,
                             v[8 * q + 4] & 0x3F803F80u, v[8 * q + 6] & 0x3F803F80u);
              continue;
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const int cb = c0 + 8 * q;
              uint32_t keep = 0xFFu;
              if (a.dropout) {
                if (a.mask_bits) {
                  keep = (mword >> (8 * q)) & 0xFFu;
                } else {
                  const Philox4 r = philox4x32_10(c0w, (uint32_t)g, ((uint32_t)l << 16) | (uint32_t)t,
                                                  (uint32_t)(cb >> 3), a.k0, a.k1);
                  keep = (uint32_t)((r.x & 0xFFFFu) >= thr) | ((uint32_t)((r.x >> 16) >= thr) << 1) |
                         ((uint32_t)((r.y & 0xFFFFu) >= thr) << 2) | ((uint32_t)((r.y >> 16) >= thr) << 3) |
                         ((uint32_t)((r.z & 0xFFFFu) >= thr) << 4) | ((uint32_t)((r.z >> 16) >= thr) << 5) |
                         ((uint32_t)((r.w & 0xFFFFu) >= thr) << 6) | ((uint32_t)((r.w >> 16) >= thr) << 7);
                }
              }
              if (!row_valid) keep = 0u;   // zero rows (the next layer's `same` padding) and out-of-range windows
              const float4 b0 = *reinterpret_cast<const float4*>(tab + cb), b1 = *reinterpret_cast<const float4*>(tab + cb + 4);
              const float4 s0 = *reinterpret_cast<const float4*>(tab + 256 + cb), s1 = *reinterpret_cast<const float4*>(tab + 256 + cb + 4);
              const float4 h0 = *reinterpret_cast<const float4*>(tab + 512 + cb), h1 = *reinterpret_cast<const float4*>(tab + 512 + cb + 4);
              const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
              const float ss[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
              const float hh[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
              float y[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) {
                const float r0 = fmaxf(__uint_as_float(v[8 * q + e]) + bb[e], 0.f);   // bias + ReLU
                const float y0 = fmaf(r0, ss[e], hh[e]);                                // BatchNorm affine (* 1/keep)
                y[e] = ((keep >> e) & 1u) ? y0 : 0.f;                                   // dropout
              }
              if (dbg) {
#pragma unroll
                for (int e = 0; e < 8; ++e) dbg[cb + e] = y[e];
              }
              if (l < kBlocks - 1) {
                st_shared_v4(act_row + (uint32_t)(cb >> 3) * kPlaneBytes, pack_bf16x2(y[0], y[1]), pack_bf16x2(y[2], y[3]),
                             pack_bf16x2(y[4], y[5]), pack_bf16x2(y[6], y[7]));
              } else {
                const float4 w0 = *reinterpret_cast<const float4*>(tab + 768 + cb), w1 = *reinterpret_cast<const float4*>(tab + 768 + cb + 4);
                zacc = fmaf(y[0], w0.x, zacc); zacc = fmaf(y[1], w0.y, zacc); zacc = fmaf(y[2], w0.z, zacc); zacc = fmaf(y[3], w0.w, zacc);
                zacc = fmaf(y[4], w1.x, zacc); zacc = fmaf(y[5], w1.y, zacc); zacc = fmaf(y[6], w1.z, zacc); zacc = fmaf(y[7], w1.w, zacc);
              }
            }
          }

          if (l < kBlocks - 1) {
            // activations of the next layer are in place: publish to the tensor-core proxy, hand TMEM back
            fence_proxy_async();
            tc_fence_before();
            mbar_arrive(bar_act + 8 * tl);
          } else {
            // ---- head: GlobalAveragePooling1D + Dense(96->1) + sigmoid; rows of one window = 2 warps
            tc_fence_before();
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) zacc += __shfl_xor_sync(0xffffffffu, zacc, o);
            if (lane == 0) atomicAdd(&zsum[tl * 2 + slot_w], zacc);
            named_bar_sync(1 + tl, 128);
            if ((row & 63) == 0) {
              const float z = zsum[tl * 2 + slot_w] * (1.0f / (float)kL) + tab[768 + kDenseIn];
              const float p = 1.0f / (1.0f + expf(-z));
              probs_s[(2 * tl + slot_w) * a.n_groups + g] = p;
              zsum[tl * 2 + slot_w] = 0.f;
              if (a.probs && win < a.n) a.probs[(size_t)g * a.n + win] = p;
            }
          }
        }
      }
      // ---- K3: per-window reductions over the T (or M) probabilities, in the reference's evaluation order
      if ((row & 63) == 0 && win < a.n) {
        const float* pw = probs_s + (2 * tl + slot_w) * a.n_groups;
        window_metrics([&](int tt) { return pw[tt]; }, a.n_groups, win, a.out);
      }
    }
  }

  // ---- teardown
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)Cfg::kTmemCols)
                 : "memory");
  }
}

// ---- host launcher ----------------------------------------------------------------------------------------------
template <int TILES, int NSLOTS>
static int launch_variant(const FusedParams& p, cudaStream_t s) {
  using Cfg = FusedCfg<TILES, NSLOTS>;
  FusedArgs a{};
  a.x = p.x; a.x_f64 = p.x_dtype == UQK_F64; a.n = p.n; a.n_groups = p.n_groups; a.mode_de = p.mode_de;
  a.dropout = p.dropout; a.k0 = (uint32_t)(p.seed & 0xFFFFFFFFull); a.k1 = (uint32_t)(p.seed >> 32);
  a.window_base = p.window_base; a.mask_bits = p.mask_bits; a.packed = reinterpret_cast<const uint8_t*>(p.packed);
  a.tables = p.tables; a.probs = p.probs; a.out = p.out; a.dbg_act = p.dbg_act; a.err_flag = p.err_flag;
  a.variant = p.variant;
  a.n_units = (p.n + Cfg::kWpu - 1) / Cfg::kWpu;
  const size_t smem = (size_t)Cfg::oProbs + (size_t)Cfg::kWpu * p.n_groups * sizeof(float);
  if (smem > 227 * 1024) { set_error("fused kernel: %d passes/members do not fit in shared memory", p.n_groups); return UQK_E_SHAPE; }
  int dev = 0, sms = 148;
  UQK_CUDA(cudaGetDevice(&dev));
  UQK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  UQK_CUDA(cudaFuncSetAttribute(fused_forward_kernel<TILES, NSLOTS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 1;
  UQK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fused_forward_kernel<TILES, NSLOTS>, Cfg::kThreads, smem));
  if (per_sm < 1) { set_error("fused kernel does not fit on an SM (smem %zu)", smem); return UQK_E_CUDA; }
  if (per_sm * Cfg::kTmemCols > 512) per_sm = 512 / Cfg::kTmemCols;   // TMEM: 512 columns per SM
  const long long grid = std::min<long long>(a.n_units, (long long)sms * per_sm);
  fused_forward_kernel<TILES, NSLOTS><<<(unsigned)grid, Cfg::kThreads, smem, s>>>(a);
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}

int fused_launch(const FusedParams& p, cudaStream_t s) {
  if (p.n == 0) return UQK_OK;
  switch (p.variant >> 4) {
    case 1: return launch_variant<1, 2>(p, s);    // one tile per CTA, two CTAs per SM
    default: return launch_variant<2, 4>(p, s);   // two tiles per CTA share one weight stream
  }
}

}  // namespace uqk
