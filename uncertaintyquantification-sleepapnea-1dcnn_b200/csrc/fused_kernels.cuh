// Device code of the fused tcgen05 kernels (see fused.cu for the overview).  Included by the translation units
// fused.cu (host side, tables), fused_pair2.cu, fused_pair.cu, fused_pair_stat.cu, fused_single2.cu, fused_single1.cu:
// every kernel family is instantiated in its own file so that the families compile in parallel.
#pragma once
#include "common.cuh"
#include "internal.h"

#include <algorithm>
#include <vector>
#include <cstring>
#include <cstdio>
#include <cstdlib>

namespace uqk {

// ---- geometry ---------------------------------------------------------------------------------------------------
constexpr int kRowsPad = 136;                      // rows -4 .. 131 of a tile (4 zero rows above and below)
constexpr int kPlaneBytes = kRowsPad * 16;         // one 8-channel chunk, all rows: 2176 B
constexpr int kOnesPlane = kMaxCh / 8;             // plane 32: constant [1, 1, 0, ...] rows -> bias added by the MMA
constexpr int kActBytes = (kMaxCh / 8 + 1) * kPlaneBytes;  // 71808 B per tile
constexpr int kSlotBytes = 16384;                  // weight ring slot

struct FusedLayer {
  int cinp;        // input channels as seen by the MMA (layer 1: 4 real + 4 zero, two taps per K=16 step)
  int cout;
  int taps;
  int halo;
  int nconv;       // K=16 MMA steps of the convolution per tile
  int nsteps;      // nconv + 1: the last step multiplies the ones-plane with the bias row
  int step_bytes;  // bytes of weights per step = cout * 32
  int sps;         // steps per ring stage
  int nstages;
  uint32_t w_off;  // byte offset of the layer inside one member's packed image
};

__host__ __device__ constexpr int fl_cout(int l) { return l == 0 ? 128 : l == 1 ? 192 : l == 2 ? 224 : l == 3 ? 96 : l == 4 ? 256 : 96; }
__host__ __device__ constexpr int fl_nconv(int l) { return l == 0 ? 4 : l == 1 ? 40 : l == 2 ? 36 : l == 3 ? 98 : l == 4 ? 54 : 144; }
__host__ __device__ constexpr uint32_t fl_bytes(int l) { return (uint32_t)(fl_nconv(l) + 1) * (uint32_t)fl_cout(l) * 32u; }
__host__ __device__ constexpr uint32_t fl_off(int l) { return l == 0 ? 0u : fl_off(l - 1) + fl_bytes(l - 1); }
__host__ __device__ constexpr FusedLayer fused_layer(int l, int slot_bytes = kSlotBytes) {
  const int cout = fl_cout(l), nsteps = fl_nconv(l) + 1, sb = cout * 32, sps = slot_bytes / sb;
  return FusedLayer{l == 0 ? 8 : l == 1 ? 128 : l == 2 ? 192 : l == 3 ? 224 : l == 4 ? 96 : 256,
                    cout,
                    l == 0 ? 7 : l == 1 ? 5 : l == 2 ? 3 : l == 3 ? 7 : 9,
                    l == 0 ? 3 : l == 1 ? 2 : l == 2 ? 1 : l == 3 ? 3 : 4,
                    fl_nconv(l), nsteps, sb, sps, (nsteps + sps - 1) / sps, fl_off(l)};
}
constexpr uint32_t kPackedBytes = fl_off(5) + fl_bytes(5);
// Run-time indexable copies (the constexpr helpers above recurse, which is only free at compile time).
static __constant__ FusedLayer c_layers[kBlocks] = {fused_layer(0), fused_layer(1), fused_layer(2),
                                             fused_layer(3), fused_layer(4), fused_layer(5)};
// The two-tile pair kernel streams stages of twice the size (32 KB per CTA pair, 16 KB per CTA): its MMA issuers pay
// ~300 clk of waits / commit / loop latency per stage, which 2-3 MMAs per tile-pair (192-256 clk) do not cover.
constexpr int kSlotBytesBig = 2 * kSlotBytes;
static __constant__ FusedLayer c_layers_big[kBlocks] = {fused_layer(0, kSlotBytesBig), fused_layer(1, kSlotBytesBig), fused_layer(2, kSlotBytesBig),
                                                 fused_layer(3, kSlotBytesBig), fused_layer(4, kSlotBytesBig), fused_layer(5, kSlotBytesBig)};

constexpr double kFusedStatScale = 16777216.0;   // 2^24 fixed point of batch statistics, as in layersync.cu

#ifndef UQK_WAIT_HINT_NS
#define UQK_WAIT_HINT_NS 20000u
#endif
constexpr uint32_t kWaitHintNs = UQK_WAIT_HINT_NS;   // suspend-time hint of mbarrier.try_wait

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
// err_flag points into MAPPED PINNED HOST memory (api.cu): the code survives the trap, which poisons the CUDA context.
static __device__ __noinline__ void mbar_timeout(int* err_flag, int code) {
  if (err_flag) *reinterpret_cast<volatile int*>(err_flag) = code;
  __threadfence_system();
  __trap();
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, int* err_flag, int code) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (err_flag && clock64() - t0 > 4000000000ll) mbar_timeout(err_flag, code);
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
// bf16 pair of (max(lo, 0), max(hi, 0)): the ReLU rides on the conversion
__device__ __forceinline__ uint32_t pack_bf16x2_relu(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---- cluster / CTA-pair helpers (cta_group::2 variant) -------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_shared(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
// Peer -> leader signal of a CTA pair.  strong = false (default): relaxed at cluster scope -- the data this signal
// publishes lives in the peer's shared memory, was fenced to the async proxy (fence.proxy.async) or written by TMA
// before the arrive, and is read by the PEER SM's own tensor core; a release at cluster scope costs a MEMBAR.ALL.GPU
// per arrive.  strong = true (variant bit 0): release here / acquire in the leader's wait, the pairing the PTX memory
// model spells out; tests/test_gpu_cluster_sync.py compares the two bit for bit and DESIGN.md records the cost.
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cluster_addr, bool strong = false) {
  if (strong) asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
  else asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void umma_bf16_2cta(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                               uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// commit of the pair's MMAs: arrives on the barrier at the same CTA-relative offset in BOTH CTAs (mask 0b11)
__device__ __forceinline__ void umma_commit_2cta(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
               "h"((uint16_t)3)
               : "memory");
}

// ---- the kernel -------------------------------------------------------------------------------------------------
struct FusedArgs {
  const void* x;
  int x_f64;
  long long n;
  int n_groups;
  int mode_de;
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
  int smem_probs;   // 1: per-window probabilities are kept in shared memory for the fused K3 reductions
  long long n_units;
  // statistics sweep of the exact batch-statistic mode (two-tile pair kernel, STAT instantiation): run layers [first_layer, n_layers) and
  // accumulate the per-channel (sum, sum of squares) of layer n_layers-1's post-ReLU output into stats[g][2*256]
  int n_layers;
  int tab_per_group;              // 1: epilogue tables of group g are tables[g] although the weights are member 0's
  unsigned long long* stats;
  int pass_base;                  // Philox pass counter of group 0 (exact mode processes the passes in chunks)
  // last statistics sweep (n_layers == 6): also emit, per (pass, window, row half, channel), the dropout-masked column
  // sum of the post-ReLU layer-6 output and the number of kept rows -- BatchNorm + GlobalAveragePooling + Dense are
  // linear in them, so the probabilities follow without another pass over the network (exact_head_kernel)
  float* head_sum;                // (n_groups, n, 2, 96) float32, or null
  unsigned char* head_cnt;        // (n_groups, n, 2, 96) uint8
  // per-layer activation hand-off of the exact mode: a sweep computes ONE layer (two for the sweep that starts at the
  // windows) and leaves its post-ReLU output, rounded to bf16 and NOT yet normalised, in HBM as planes
  // (n_groups, n, Cout/8 planes, 60 rows) x 16 B; the next sweep applies this layer's BatchNorm affine (now known for
  // all windows) + dropout while it stages the planes into shared memory
  uint4* ckpt_out;                // post-ReLU output of layer n_layers-1, or null
  const uint4* ckpt_in;           // post-ReLU output of layer first_layer-1 (first_layer >= 2), or null
  int first_layer;                // first layer this sweep computes; 0 = start from the windows
  int pipelined;                  // 1: resumed sweeps run the three-stage pipeline over 4-window units (option exact_pipelined)
#ifdef UQK_TIMELINE
  long long* tl_buf;              // tools/timeline_probe.py: clock64 stamps [tile-pair][16 passes][6 layers][8 events] of one unit
#endif
};
#ifdef UQK_TIMELINE
extern long long* g_uqk_timeline;   // api.cu, set through uqk_debug_set_timeline (timeline build only, not part of the C-ABI)
#define UQK_TL(buf, tl, g, L, ev) do { if ((buf) && (g) < 16) (buf)[((((tl) * 16 + (g)) * 6 + (L)) << 3) + (ev)] = clock64(); } while (0)
#else
#define UQK_TL(buf, tl, g, L, ev) do { } while (0)
#endif
#ifdef UQK_TIMELINE
#define UQK_TLE(c, g, L, ev) UQK_TL((c).tlb, (c).tl, g, L, ev)
#else
#define UQK_TLE(c, g, L, ev) do { } while (0)
#endif

// dropout source of the epilogue
constexpr int kDropNone = 0, kDropPhilox = 1, kDropBits = 2;

template <int TILES, int NSLOTS>
struct FusedCfg {
  static constexpr int kEpiWarps = 8;                       // per tile: 2 warps per TMEM lane quarter
  static constexpr int kEpiThreads = 32 * kEpiWarps;
  static constexpr int kWarps = 2 + kEpiWarps * TILES;
  static constexpr int kThreads = 32 * kWarps;
  static constexpr int kWpu = 2 * TILES;  // windows per unit
  static constexpr int kTmemCols = 256 * TILES;
  // shared-memory map (bytes)
  static constexpr int oAct = 0;
  static constexpr int oRing = oAct + TILES * kActBytes;
  static constexpr int kTabFloats = 2 * 512 + 128;                   // 2 x (scale[256], shift[256]) + dense w[96] + b
  static constexpr int oTab = oRing + NSLOTS * kSlotBytes;
  static constexpr int oBar = oTab + TILES * kTabFloats * 4;         // mbarriers
  static constexpr int kNumBars = 2 * NSLOTS + 2 * TILES;
  static constexpr int oMisc = oBar + 8 * kNumBars;                  // tmem ptr (16 B), zpart[TILES][2][4]
  static constexpr int oAlo = oMisc + 16 + 32 * TILES;               // A-descriptor low words, one per (layer, step)
  static constexpr int oProbs = oAlo + 4 * 392;                      // probs[kWpu][n_groups]
};

// Bounded mbarrier wait.  `hint_ns` lets the hardware suspend the polling thread inside try_wait instead of
// spinning; the watchdog turns a protocol bug into a trap (never a hung GPU).
__device__ __forceinline__ void mbar_wait2(uint32_t bar, uint32_t parity, int* err_flag, int code) {
  // The polling loop lives inside one asm block (try_wait, branch, counter: ~6 SASS instructions per wake-up instead
  // of the ~15 the compiler emits around a C++ loop -- pollers are woken by every arrival / transaction update, and
  // their instructions cost issue slots and power under the cap).  Every 2048 unsuccessful tries control returns to
  // C++ for the watchdog.
  long long t0 = 0;
  for (;;) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b32 n;\n\t"
        "mov.u32 n, 2048;\n"
        "UQK_TRY:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "@p bra UQK_DONE;\n\t"
        "sub.u32 n, n, 1;\n\t"
        "setp.ne.u32 p, n, 0;\n\t"
        "@p bra UQK_TRY;\n\t"
        "mov.u32 %0, 0;\n\t"
        "bra UQK_END;\n"
        "UQK_DONE:\n\t"
        "mov.u32 %0, 1;\n"
        "UQK_END:\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity), "r"(kWaitHintNs)
        : "memory");
    if (ok) return;
    const long long now = clock64();                  // watchdog: ~2 s of waiting is a protocol bug
    if (t0 == 0) t0 = now;
    else if (err_flag && now - t0 > 4000000000ll) mbar_timeout(err_flag, code);   // err_flag == null: watchdog off
  }
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
// tcgen05.wait::ld with the loaded registers as in/out operands: every use of v[] depends on the wait.
__device__ __forceinline__ void tmem_wait16(uint32_t (&v)[16]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]),
                 "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15])
               :
               : "memory");
}

// bits 15 / 31 of d -> 0xFFFF / 0 per 16-bit half (prmt selector bit 3 = replicate the selected byte's sign bit;
// the __byte_perm intrinsic ignores that bit, hence PTX)
__device__ __forceinline__ uint32_t prmt_sign_halves(uint32_t d) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %1, 0xBB99;" : "=r"(r) : "r"(d));
  return r;
}

template <uint32_t SEL>
__device__ __forceinline__ uint32_t prmt_sign_bytes(uint32_t z) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %1, %2;" : "=r"(r) : "r"(z), "n"(SEL));
  return r;
}

// Keep-flag store of one thread and layer: the flags of 4 consecutive channels ("quad" t = channel_in_thread / 4) sit
// in bit (t & 7) of the four bytes of word t >> 3 -- 32 flags per register, Cout/2 flags = at most 4 registers.
// Insert = shift the byte-sign form into place (SHF + LOP3); extract = shift back to the byte signs (SHF), which is
// what PRMT's sign replication reads when it widens a flag into the 0xFFFF / 0 mask of a bf16 half.
__device__ __forceinline__ void keep_insert(uint32_t (&kb)[4], int t, uint32_t g) {   // t is a constant after unrolling
  const int s = t & 7;
  kb[t >> 3] |= (s == 7 ? g : (g >> (7 - s))) & (0x01010101u << s);
}
__device__ __forceinline__ uint32_t keep_extract(const uint32_t (&kb)[4], int t) {
  const int s = t & 7;
  return s == 7 ? kb[t >> 3] : (kb[t >> 3] << (7 - s));
}

// One 16-column chunk of the epilogue: ReLU -> BatchNorm affine (dropout scale folded) -> dropout -> bf16 store
// (layers 1-5) or dot with the dense kernel (layer 6).  The conv bias is already in the accumulator (bias MMA step).
//
// Dropout: the thread's keep-flags of the layer arrive as flag words (make_keep_flags: bit-sliced Philox words or
// injected bits, byte-lane layout); zq[] holds the flags of this chunk's four channel quads in the byte sign bits.
// PRMT with sign replication turns two flags into the 0xFFFF / 0 masks of one bf16 pair and one LOP3 applies the mask
// (and the zero-row mask) to the packed pair: no per-element compare, no predicates.
template <int DROP, bool LAST, bool DBG, bool RELU = true>
__device__ __forceinline__ void epi_chunk16(const uint32_t (&v)[16], int c0, const float* __restrict__ tab_s,
                                            const float* __restrict__ tab_h, const float* __restrict__ tab_w,
                                            uint32_t act_row, bool row_valid, uint32_t row_mask, uint32_t thr2,
                                            uint32_t c0w, uint32_t pass, uint32_t c2, uint32_t k0, uint32_t k1,
                                            const uint32_t (&zq)[4], float& zacc, float* dbg) {
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const int cb = c0 + 8 * q;
    const float4 s0 = *reinterpret_cast<const float4*>(tab_s + cb), s1 = *reinterpret_cast<const float4*>(tab_s + cb + 4);
    const float4 h0 = *reinterpret_cast<const float4*>(tab_h + cb), h1 = *reinterpret_cast<const float4*>(tab_h + cb + 4);
    const float ss[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
    const float hh[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
    float y[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) y[e] = fmaf(RELU ? fmaxf(__uint_as_float(v[8 * q + e]), 0.f) : __uint_as_float(v[8 * q + e]), ss[e], hh[e]);
    uint32_t m[4] = {row_mask, row_mask, row_mask, row_mask};     // per bf16 pair: 0xFFFF per kept half
    if (DROP != kDropNone) {
      // zq[2q], zq[2q+1]: keep-flags of channels 0..3 / 4..7 of this half in the byte sign bits (zero for padding
      // rows); PRMT with sign replication widens two byte signs into the 0xFFFF / 0 masks of one bf16 pair
      m[0] = prmt_sign_bytes<0x9988>(zq[2 * q]); m[1] = prmt_sign_bytes<0xBBAA>(zq[2 * q]);
      m[2] = prmt_sign_bytes<0x9988>(zq[2 * q + 1]); m[3] = prmt_sign_bytes<0xBBAA>(zq[2 * q + 1]);
    }
    if (DBG || LAST) {
      // float path (debug dump and the head): apply the same masks to the fp32 values
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        y[2 * i] = (m[i] & 0x0000FFFFu) ? y[2 * i] : 0.f;
        y[2 * i + 1] = (m[i] & 0xFFFF0000u) ? y[2 * i + 1] : 0.f;
      }
    }
    if (DBG) {
      if (dbg) {
#pragma unroll
        for (int e = 0; e < 8; ++e) dbg[cb + e] = y[e];
      }
    }
    if (!LAST) {
      const uint32_t p0 = pack_bf16x2(y[0], y[1]) & m[0], p1 = pack_bf16x2(y[2], y[3]) & m[1];
      const uint32_t p2 = pack_bf16x2(y[4], y[5]) & m[2], p3 = pack_bf16x2(y[6], y[7]) & m[3];
      st_shared_v4(act_row + (uint32_t)(cb >> 3) * kPlaneBytes, p0, p1, p2, p3);
    } else {
      const float4 w0 = *reinterpret_cast<const float4*>(tab_w + cb), w1 = *reinterpret_cast<const float4*>(tab_w + cb + 4);
      zacc = fmaf(y[0], w0.x, zacc); zacc = fmaf(y[1], w0.y, zacc); zacc = fmaf(y[2], w0.z, zacc); zacc = fmaf(y[3], w0.w, zacc);
      zacc = fmaf(y[4], w1.x, zacc); zacc = fmaf(y[5], w1.y, zacc); zacc = fmaf(y[6], w1.z, zacc); zacc = fmaf(y[7], w1.w, zacc);
    }
  }
}

__host__ __device__ constexpr int fl_step_base(int l) { return l == 0 ? 0 : fl_step_base(l - 1) + fl_nconv(l - 1) + 1; }
constexpr int kTotalSteps = fl_step_base(kBlocks);   // 382 convolution steps + 6 bias steps
constexpr int kStagesPerGroup = fused_layer(0).nstages + fused_layer(1).nstages + fused_layer(2).nstages +
                                fused_layer(3).nstages + fused_layer(4).nstages + fused_layer(5).nstages;
static __constant__ int c_ch_off[kBlocks] = {ch_off(0), ch_off(1), ch_off(2), ch_off(3), ch_off(4), ch_off(5)};
static __constant__ int c_step_base[kBlocks + 1] = {fl_step_base(0), fl_step_base(1), fl_step_base(2), fl_step_base(3),
                                             fl_step_base(4), fl_step_base(5), fl_step_base(6)};

__device__ __forceinline__ uint64_t mk64(uint32_t lo, uint32_t hi) { return ((uint64_t)hi << 32) | (uint64_t)lo; }

// ---- MMA issue of one layer (warp 1; the whole warp walks the schedule, one elected lane issues) -----------------
// A-descriptor low words are precomputed per (layer, step) in shared memory (s_alo); the high word of every
// descriptor is the constant SBO=128 | version=1.
template <int L, int TILES, int NSLOTS, typename Cfg>
__device__ __forceinline__ void issue_layer(uint32_t s_base, uint32_t tmem_base, const uint32_t* s_alo, uint32_t bar_full,
                                            uint32_t bar_empty, uint32_t bar_acc, uint32_t bar_act, uint32_t& it,
                                            uint32_t& n_act, int* err_flag, bool no_mma) {
  constexpr FusedLayer fl = fused_layer(L);
  constexpr uint32_t kHi = (128u >> 4) | (1u << 14);
  constexpr uint32_t kBLbo = ((uint32_t)(fl.cout * 16) >> 4) << 16;
  constexpr uint32_t kIdesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(fl.cout >> 3) << 17) | ((128u >> 4) << 24);
  constexpr int kLast = fl.nsteps - (fl.nstages - 1) * fl.sps;
  for (int tl = 0; tl < TILES; ++tl) mbar_wait2(bar_act + 8 * tl, n_act & 1, err_flag, 2);
  ++n_act;
  tc_fence_after();
#pragma unroll 1
  for (int st = 0; st < fl.nstages; ++st, ++it) {
    const int slot = it % NSLOTS;
    const uint32_t ph = (it / NSLOTS) & 1;
    mbar_wait2(bar_full + 8 * slot, ph, err_flag, 3);
    tc_fence_after();
    const int nst = (st == fl.nstages - 1) ? kLast : fl.sps;
    const uint32_t b_lo0 = (((s_base + Cfg::oRing + slot * kSlotBytes) >> 4) & 0x3FFFu) | kBLbo;
    if (elect_one()) {
      const uint32_t* alo = s_alo + fl_step_base(L) + st * fl.sps;
      if (!no_mma) {
#pragma unroll
        for (int sl = 0; sl < fl.sps; ++sl) {
          if (sl < nst) {
            const uint32_t a0 = alo[sl];
            const uint64_t bdesc = mk64(b_lo0 + (uint32_t)sl * (uint32_t)(fl.step_bytes >> 4), kHi);
            const uint32_t acc = (st > 0 || sl > 0) ? 1u : 0u;
            umma_bf16(tmem_base, mk64(a0, kHi), bdesc, kIdesc, acc);
            if (TILES == 2) umma_bf16(tmem_base + 256, mk64(a0 + (uint32_t)(kActBytes >> 4), kHi), bdesc, kIdesc, acc);
          }
        }
      }
      umma_commit(bar_empty + 8 * slot);   // frees the ring slot when these MMAs have read it
    }
    __syncwarp();
  }
  if (elect_one()) {
    for (int tl = 0; tl < TILES; ++tl) umma_commit(bar_acc + 8 * tl);   // accumulator(s) complete
  }
  __syncwarp();
}

// ---- epilogue of one layer (8 warps per tile) -----------------------------------------------------------------------
struct EpiCtx {
  uint32_t act_row, taddr, c0w, k0, k1, bar_acc, bar_act;
  DropTie tie = {0u, 0u};    // dropout tie-break bytes of this thread's (window, pass, row, column half), see common.cuh
  bool strong_sync = false;  // pair kernels: release / acquire cluster signalling (variant bit 0)
  uint32_t remote_act = 0;   // pair kernels, peer CTA: cluster address of the leader's "peer activations ready" barrier
  uint32_t bar_half = 0, remote_half = 0;   // two-tile pair kernel: "first half of the output channels is in place" (early_half)
  uint32_t xin0 = 0, xin1 = 0;              // this thread's row of the window as bf16 (column half 0), re-staged for every pass
  // EARLY: layer 1's keep-flags and the tie-break bytes of the NEXT pass, generated while the last layer's MMAs run
  mutable uint32_t kb1[4] = {0u, 0u, 0u, 0u};
  mutable DropTie tie_next = {0u, 0u};
  int t, colhalf, tid_t, we, lane, quarter, slot_w, tl;
  bool row_valid;
  float* tab;
  float* tab_w;
  float* zp;
  int* err_flag;
#ifdef UQK_TIMELINE
  long long* tlb = nullptr;  // non-null in the one unit whose schedule is recorded (lane 0 of the tile's first and last warp)
#endif
};

// "this tile's activations are in place": local arrive (the issuer lives in this CTA), or -- peer CTA of a pair -- one
// remote arrive on the leader's barrier after all 256 threads have fenced their stores.
template <int EPI_THREADS>
__device__ __forceinline__ void epi_signal_act(const EpiCtx& c) {
  if (c.remote_act == 0) {
    // one arrival per warp (the barrier counts EPI_THREADS / 32): every lane has fenced its own stores, __syncwarp
    // orders them before lane 0's release-arrive -- 8 mbarrier operations per hand-off instead of 256
    __syncwarp();
    if (c.lane == 0) mbar_arrive(c.bar_act);
  } else {
    named_bar_sync(1 + c.tl, EPI_THREADS);
    if (c.tid_t == 0) mbar_arrive_remote(c.remote_act, c.strong_sync);
  }
}

// Early start of the next layer (two-tile pair kernel).  The K steps of a layer are slab-major (all taps of input
// channels 0..15, then 16..31, ...), and every epilogue warp walks its 16-column chunks in ascending order, so after
// half_chunks(L) chunks per warp the first 2 * half_chunks(L) slabs of the next layer's input are complete: the epilogue
// signals bar_half there, and the next layer's issuer starts on those slabs while the remaining chunks are still being
// written (different planes).  The accumulator columns the early MMAs overwrite must already have been read:
//   layer 1 -> 2: layer 1 accumulates at column 128 ([128,256); read first: [128,192)), layer 2 writes [0,192)
//   layer 3 -> 4: layer 3 at [32,256), read first [32,96); layer 4 writes [0,96)
//   layer 5 -> 6: layer 5 at [0,256), read first [0,128); layer 6 writes [0,96)
// Layers 2 -> 3 and 4 -> 5 do not fit (the unread half plus the next accumulator exceed the 256 columns of a tile-pair).
__host__ __device__ constexpr bool early_producer(int l) { return l == 0 || l == 2 || l == 4; }
__host__ __device__ constexpr int half_chunks(int l) { return l == 2 ? 2 : (((fl_cout(l) / 32 + 1) / 2) + 1) & ~1; }   // per warp: 2, 2, 4
__host__ __device__ constexpr int acc_col(int l, bool early) { return !early ? 0 : l == 0 ? 128 : l == 2 ? 32 : 0; }
template <int EPI_THREADS>
__device__ __forceinline__ void epi_signal_half(const EpiCtx& c) {
  if (c.remote_half == 0) {
    __syncwarp();
    if (c.lane == 0) mbar_arrive(c.bar_half);
  } else {
    named_bar_sync(1 + c.tl, EPI_THREADS);
    if (c.tid_t == 0) mbar_arrive_remote(c.remote_half, c.strong_sync);
  }
}

// Start of pass g for this thread's (window, row, column half): the dropout tie-break bytes of the pass (one Philox call).
template <int DROP>
__device__ __forceinline__ void epi_pass_setup(EpiCtx& c, const FusedArgs& a, int g) {
  if (DROP == kDropPhilox) c.tie = drop_tie(c.c0w, (uint32_t)(a.pass_base + g), (uint32_t)c.t, (uint32_t)c.colhalf, c.k0, c.k1);
}

// Keep-flags of this thread's row for layer L (byte-lane layout of keep_insert), all zero for padding rows.
//  - kDropPhilox: the dropout masks do not depend on the data, so they are generated while the layer's MMAs run and
//    the epilogue warps would otherwise sleep on the accumulator barrier (10 Philox rounds per 8 channels leave the
//    critical path of the epilogue);
//  - kDropBits (mask injection, parity tests): explicit keep-bits from global memory, bit c%32 of word c/32.
template <int L, int DROP>
__device__ __forceinline__ void make_keep_flags(const EpiCtx& c, const FusedArgs& a, int g, long long win, uint32_t (&kb)[4]) {
  constexpr int my = fl_cout(L) / 32;
  kb[0] = kb[1] = kb[2] = kb[3] = 0u;
  if (DROP == kDropBits) {
    const uint32_t* mrow = a.mask_bits + ((((size_t)g * a.n + (c.row_valid ? win : 0)) * kBlocks + L) * kL + (c.row_valid ? c.t : 0)) * 8;
#pragma unroll
    for (int j = 0; j < my; ++j) {
      const int k = c.colhalf + 2 * j;
      const uint32_t w16 = (__ldg(mrow + (k >> 1)) >> (16 * (k & 1))) & 0xFFFFu;
#pragma unroll
      for (int u = 0; u < 4; ++u)      // 4 bits -> bit 7 of 4 bytes (bit k lands on 8k + 7; the partial products never collide)
        keep_insert(kb, 4 * j + u, (((w16 >> (4 * u)) & 0xFu) * 0x10204080u) & 0x80808080u);
    }
  }
  if (DROP == kDropPhilox) {
    // bit-sliced flags (common.cuh): the byte-lane layout of keep_insert IS the layout the generator produces, so a
    // flag word is used as generated -- two Philox calls and four LOP3 per 32 elements, one call per 128 at rate 0.5
    constexpr int nw = drop_flag_words(fl_cout(L));
    const uint32_t pass = (uint32_t)(a.pass_base + g);
    if constexpr (drop_is_half(L)) {
      const Philox4 r = drop_half_words(c.c0w, pass, (uint32_t)L, (uint32_t)c.t, (uint32_t)c.colhalf, c.k0, c.k1);
      kb[0] = r.x; kb[1] = r.y;
      if (nw > 2) { kb[2] = r.z; kb[3] = r.w; }
    } else {
      const uint32_t tb = drop_tie_byte(c.tie, L);
#pragma unroll
      for (int i = 0; i < nw; ++i)
        kb[i] = drop_keep_word(c.c0w, pass, (uint32_t)L, (uint32_t)c.t, (uint32_t)c.colhalf, (uint32_t)i, drop_thr16(L), tb, c.k0, c.k1);
    }
  }
  if (DROP != kDropNone) {
#pragma unroll
    for (int i = 0; i < 4; ++i) kb[i] = c.row_valid ? kb[i] : 0u;   // padding rows keep nothing
  }
}

template <int L, int DROP, bool DBG, int EPI_THREADS, bool EARLY = false>
__device__ __forceinline__ void epi_layer(const EpiCtx& c, const FusedArgs& a, const float* __restrict__ tab_g, int g,
                                          long long win, uint32_t& n_acc, float* probs_s, float* dbg_base) {
  constexpr int cout = fl_cout(L);
  constexpr int my = cout / 32;            // 16-column chunks owned by this warp (chunks colhalf, colhalf+2, ...)
  constexpr int kHalfJ = (EARLY && early_producer(L)) ? half_chunks(L) : 0;   // signal bar_half after this many chunks
  const uint32_t taddr = c.taddr + (uint32_t)acc_col(L, EARLY);
  constexpr bool kLastLayer = (L == kBlocks - 1);
  if (c.we == 0) UQK_TLE(c, g, L, 6);
  // stage this layer's scale / shift (double buffered) while the MMAs run
  float* tab_s = c.tab + (L & 1) * 512;
  float* tab_h = tab_s + 256;
  for (int i = c.tid_t; i < cout; i += EPI_THREADS) {
    tab_s[i] = __ldg(tab_g + kChTotal + ch_off(L) + i);
    tab_h[i] = __ldg(tab_g + 2 * kChTotal + ch_off(L) + i);
  }
  if (kLastLayer && c.tid_t < kDenseIn + 1) c.tab_w[c.tid_t] = __ldg(tab_g + 3 * kChTotal + c.tid_t);

  // keep-flags of this thread's row for this layer (see keep_insert): produced while the layer's MMAs run
  uint32_t kb[4];
  if (EARLY && L == 0 && DROP != kDropNone && g > 0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) kb[i] = c.kb1[i];            // made during the previous pass's last layer
  } else {
    make_keep_flags<L, DROP>(c, a, g, win, kb);
  }
  if (EARLY && kLastLayer && DROP != kDropNone && g + 1 < a.n_groups) {
    // next pass, layer 1: its MMAs take no time, so its flags would be exposed -- make them here, in this layer's shadow
    EpiCtx cn = c;
    if (DROP == kDropPhilox) cn.tie = drop_tie(c.c0w, (uint32_t)(a.pass_base + g + 1), (uint32_t)c.t, (uint32_t)c.colhalf, c.k0, c.k1);
    make_keep_flags<0, DROP>(cn, a, g + 1, win, c.kb1);
    c.tie_next = cn.tie;
  }
  constexpr uint32_t thr2 = 0u;
  const uint32_t c2 = ((uint32_t)L << 16) | (uint32_t)c.t;
  const uint32_t row_mask = c.row_valid ? 0xFFFFFFFFu : 0u;   // zero rows = the next layer's `same` padding

  // one warp polls the accumulator barrier; the others sleep in the named barrier (no issue slots burnt)
  if (c.we == 0) {
    UQK_TLE(c, g, L, 2);
    mbar_wait2(c.bar_acc, n_acc & 1, c.err_flag, 4);
    tc_fence_after();
    tc_fence_before();
  }
  ++n_acc;
  named_bar_sync(1 + c.tl, EPI_THREADS);
  tc_fence_after();
  if (c.we == 0) UQK_TLE(c, g, L, 3);
  if (EARLY && kLastLayer && g + 1 < a.n_groups) {
    // the last layer's MMAs have retired: the planes are free, so the next pass's input goes in now and its layer 1
    // (accumulator at column 128, this layer's at [0,96)) runs while the head below is evaluated
    if (c.colhalf == 0) st_shared_v4(c.act_row, c.xin0, c.xin1, 0u, 0u);
    fence_proxy_async();
    epi_signal_act<EPI_THREADS>(c);
  }

  float* dbg = nullptr;
  if (DBG) {
    if (dbg_base) dbg = dbg_base + (size_t)L * 128 * 256;
  }
  float zacc = 0.f;
  // software pipeline: the TMEM load of the next chunk is in flight while the current one is processed
  uint32_t va[16], vb[16];
  tmem_ld16(taddr + (uint32_t)(16 * c.colhalf), va);
#pragma unroll
  for (int j = 0; j < my; j += 2) {
    tmem_wait16(va);
    if (j + 1 < my) tmem_ld16(taddr + (uint32_t)(16 * (c.colhalf + 2 * (j + 1))), vb);
    {
      const uint32_t zq[4] = {keep_extract(kb, 4 * j), keep_extract(kb, 4 * j + 1), keep_extract(kb, 4 * j + 2), keep_extract(kb, 4 * j + 3)};
      epi_chunk16<DROP, kLastLayer, DBG>(va, 16 * (c.colhalf + 2 * j), tab_s, tab_h, c.tab_w, c.act_row, c.row_valid, row_mask,
                                               thr2, c.c0w, (uint32_t)g, c2, c.k0, c.k1, zq, zacc, dbg);
    }
    if (j + 1 < my) {
      tmem_wait16(vb);
      if (j + 2 < my) tmem_ld16(taddr + (uint32_t)(16 * (c.colhalf + 2 * (j + 2))), va);
      const uint32_t zq[4] = {keep_extract(kb, 4 * j + 4), keep_extract(kb, 4 * j + 5), keep_extract(kb, 4 * j + 6), keep_extract(kb, 4 * j + 7)};
      epi_chunk16<DROP, kLastLayer, DBG>(vb, 16 * (c.colhalf + 2 * (j + 1)), tab_s, tab_h, c.tab_w, c.act_row, c.row_valid,
                                               row_mask, thr2, c.c0w, (uint32_t)g, c2, c.k0, c.k1, zq, zacc, dbg);
    }
    if (kHalfJ != 0 && j + 2 == kHalfJ) {
      // the first 2 * kHalfJ slabs of the next layer's input are in place and their accumulator columns have been read
      fence_proxy_async();
      tc_fence_before();
      epi_signal_half<EPI_THREADS>(c);
    }
  }

  if (!kLastLayer) {
    // activations of the next layer are in place: publish to the tensor-core proxy, hand TMEM back
    fence_proxy_async();
    tc_fence_before();
    epi_signal_act<EPI_THREADS>(c);
  } else {
    // ---- head: GlobalAveragePooling1D + Dense(96->1) + sigmoid; one window = 2 lane quarters x 2 column halves
    tc_fence_before();
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) zacc += __shfl_xor_sync(0xffffffffu, zacc, o);
    if (c.lane == 0) c.zp[c.slot_w * 4 + (c.quarter & 1) * 2 + c.colhalf] = zacc;
    named_bar_sync(1 + c.tl, EPI_THREADS);
    if (c.colhalf == 0 && c.t == 0 && (c.quarter & 1) == 0) {
      const float* z4 = c.zp + c.slot_w * 4;
      const float z = (((z4[0] + z4[1]) + z4[2]) + z4[3]) * (1.0f / (float)kL) + c.tab_w[kDenseIn];   // fixed order
      const float p = 1.0f / (1.0f + expf(-z));
      if (a.smem_probs) probs_s[(2 * c.tl + c.slot_w) * a.n_groups + g] = p;
      if (a.probs && win < a.n) a.probs[(size_t)g * a.n + win] = p;
    }
  }
  UQK_TLE(c, g, L, c.we == 0 ? 4 : 5);
}

// ---- statistics epilogue (exact batch-statistic mode) -------------------------------------------------------------
// The accumulator is read with the 16x256b TMEM load shape (the MMA C-fragment layout): lane (r = lane >> 2, q = lane & 3)
// receives rows {r, r+8} of a 16-row block and columns {2q, 2q+1, 8+2q, 9+2q} of a 16-column chunk; two loads (rows
// 0..15 and 16..31 of the warp's lane quarter) give every thread 4 rows x 4 columns.  A column sum over the warp's 32
// rows is then 3 in-thread additions and a reduction over the 8 lanes that share q -- 7 shuffles per chunk for the sums
// AND the sums of squares together (recursive halving), against 32 with one row per lane.
__device__ __forceinline__ void tmem_ld_16x256b_x2(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr)
               : "memory");
}
__device__ __forceinline__ void tmem_wait8x2(uint32_t (&a)[8], uint32_t (&b)[8]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(a[4]), "+r"(a[5]), "+r"(a[6]), "+r"(a[7]),
                 "+r"(b[0]), "+r"(b[1]), "+r"(b[2]), "+r"(b[3]), "+r"(b[4]), "+r"(b[5]), "+r"(b[6]), "+r"(b[7])
               :
               : "memory");
}
// Recursive halving over the lane bits 4, 3, 2 (the 8 lanes that share lane & 3): NV values per lane in, NV / 8 out.
// Lane bit 4 selects the upper / lower half of the values, bit 3 the quarter, bit 2 the eighth; fixed summation tree.
template <int NV>
__device__ __forceinline__ void quad_halve(float (&v)[NV], int lane) {
  static_assert(NV % 8 == 0, "NV must be a multiple of 8");
  const bool b4 = (lane & 16) != 0, b3 = (lane & 8) != 0, b2 = (lane & 4) != 0;
#pragma unroll
  for (int i = 0; i < NV / 2; ++i) {
    const float keep = b4 ? v[NV / 2 + i] : v[i], send = b4 ? v[i] : v[NV / 2 + i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
#pragma unroll
  for (int i = 0; i < NV / 4; ++i) {
    const float keep = b3 ? v[NV / 4 + i] : v[i], send = b3 ? v[i] : v[NV / 4 + i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
  }
#pragma unroll
  for (int i = 0; i < NV / 8; ++i) {
    const float keep = b2 ? v[NV / 8 + i] : v[i], send = b2 ? v[i] : v[NV / 8 + i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  }
}

// Adds a non-negative partial sum to a 2^-24 fixed-point 64-bit accumulator in shared memory with two native 32-bit
// atomics (the 64-bit shared-memory atomic is a compare-and-swap loop, slow under the 8-way contention of the epilogue
// warps): the low word's add reports the wrap-around that carries into the high word.  v * 2^24 is exact in float, so
// the integer is the same one the double-precision conversion of the fp32 kernels produces.
__device__ __forceinline__ void stat_add(unsigned long long* s_stats, int idx, float v) {
  const unsigned long long f = __float2ull_rn(v * 16777216.0f);
  uint32_t* w = reinterpret_cast<uint32_t*>(s_stats + idx);
  const uint32_t lo = (uint32_t)f, hi = (uint32_t)(f >> 32);
  const uint32_t old = atomicAdd(w, lo);
  const uint32_t up = hi + ((uint32_t)(old + lo) < old ? 1u : 0u);
  if (up) atomicAdd(w + 1, up);
}

// Layer L is the (last) layer of a statistics sweep: ReLU of the accumulator (bias already added by the MMA), then the
// per-channel sum / sum of squares over the valid rows.  One warp's partial covers 32 rows of one window; it is
// converted to fixed point BEFORE it meets any other partial, so the totals are sums of integers that depend only on
// (window, row half, channel) -- independent of the grid, of scheduling and of how windows are sharded over GPUs.
// Layers 2..5 (L = 1..4) also leave their post-ReLU output in HBM, rounded to bf16 (a.ckpt_out, see FusedArgs); their
// statistics are those of the ROUNDED values, i.e. of exactly the numbers the next sweep normalises.
// `after_acc()` runs once every MMA of the layer has retired: the activation planes are free again, the next unit's
// input can be fetched into them while this epilogue works out of TMEM.
// `in_shadow()` runs before the wait for the accumulator, i.e. while the layer's MMAs are in flight.
template <int L, int DROP, int EPI_THREADS, typename InShadow, typename AfterAcc>
__device__ __forceinline__ void epi_stats(const EpiCtx& c, const FusedArgs& a, int g, long long win, uint32_t& n_acc,
                                          unsigned long long* s_stats, InShadow&& in_shadow, AfterAcc&& after_acc) {
  constexpr int cout = fl_cout(L);
  constexpr int my = cout / 32;
  constexpr bool kHead = (L == kBlocks - 1);
  constexpr bool kStore = (L >= 1 && L <= kBlocks - 2);
  constexpr int NV = kHead ? 16 : 8;        // values reduced per chunk and thread: 4 columns x {sum, sum of squares [, masked sum, kept count]}
  const bool head = kHead && a.head_sum != nullptr;
  uint32_t kb[4] = {0u, 0u, 0u, 0u};
  if (kHead) {
    if (head) make_keep_flags<L, DROP>(c, a, g, win, kb);      // layer-6 dropout flags of row c.t (one row per lane), in the MMA shadow
  }
  in_shadow();
  if (c.we == 0) {
    mbar_wait2(c.bar_acc, n_acc & 1, c.err_flag, 4);
    tc_fence_after();
    tc_fence_before();
  }
  ++n_acc;
  named_bar_sync(1 + c.tl, EPI_THREADS);
  tc_fence_after();
  after_acc();
  const int q = c.lane & 3, r = c.lane >> 2;
  const int t0 = (c.quarter & 1) * 32 + r;                       // time steps of this thread's four rows: t0 + 8k
  const bool win_ok = win < a.n;
  bool ok[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) ok[k] = win_ok && (t0 + 8 * k < kL);
  // layer-6 flags of this thread's four rows: row t0 + 8k is lane r + 8k of this warp in the one-row-per-lane mapping
  uint32_t fw[4][2];
  if (kHead) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      fw[k][0] = __shfl_sync(0xffffffffu, kb[0], r + 8 * k);
      fw[k][1] = __shfl_sync(0xffffffffu, kb[1], r + 8 * k);
    }
  }
  // (pass, window, row half) block of the head partials; this warp covers rows [32*(quarter&1), +32) of its window
  const size_t head_row = (((size_t)g * a.n + (win_ok ? win : 0)) * 2 + (size_t)(c.quarter & 1)) * kDenseIn;
  uint32_t* out32 = nullptr;
  if (kStore) {
    if (a.ckpt_out && win_ok)
      out32 = reinterpret_cast<uint32_t*>(a.ckpt_out + (((size_t)g * a.n + win) * (cout / 8) + 2 * c.colhalf) * kL + t0) + q;
  }
  // software pipeline: the TMEM loads of chunk j + 1 are in flight while chunk j is processed
  uint32_t lo[8], hi[8], lo2[8], hi2[8];
  tmem_ld_16x256b_x2(c.taddr + (uint32_t)(16 * c.colhalf), lo);
  tmem_ld_16x256b_x2(c.taddr + (16u << 16) + (uint32_t)(16 * c.colhalf), hi);
#pragma unroll
  for (int j = 0; j < my; ++j) {
    const int col0 = 16 * (c.colhalf + 2 * j);
    tmem_wait8x2(lo, hi);
    if (j + 1 < my) {
      tmem_ld_16x256b_x2(c.taddr + (uint32_t)(col0 + 32), lo2);
      tmem_ld_16x256b_x2(c.taddr + (16u << 16) + (uint32_t)(col0 + 32), hi2);
    }
    // x[k][cc]: row t0 + 8k, column col0 + {2q, 2q+1, 8+2q, 9+2q}[cc]
    float x[4][4];
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) {
      const int o = (cc & 1) + 4 * (cc >> 1);
      x[0][cc] = __uint_as_float(lo[o]); x[1][cc] = __uint_as_float(lo[o + 2]);
      x[2][cc] = __uint_as_float(hi[o]); x[3][cc] = __uint_as_float(hi[o + 2]);
    }
    if (kStore) {
      // ReLU + rounding to bf16 in one conversion; padding rows are masked to zero; the statistics below are those of
      // the rounded values.  Plane base out32 already points at this thread's first chunk (2 * colhalf planes in).
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t okm = ok[k] ? 0xFFFFFFFFu : 0u;
        const uint32_t p0 = pack_bf16x2_relu(x[k][0], x[k][1]) & okm, p1 = pack_bf16x2_relu(x[k][2], x[k][3]) & okm;
        x[k][0] = __uint_as_float(p0 << 16); x[k][1] = __uint_as_float(p0 & 0xFFFF0000u);
        x[k][2] = __uint_as_float(p1 << 16); x[k][3] = __uint_as_float(p1 & 0xFFFF0000u);
        if (out32 && ok[k]) {
          // a quad of lanes writes the 16 bytes (8 channels) of one row; 8 rows per store instruction = 128 contiguous bytes
          __stcs(out32 + ((4 * j) * kL + 8 * k) * 4, p0);
          __stcs(out32 + ((4 * j + 1) * kL + 8 * k) * 4, p1);
        }
      }
    } else {
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) x[k][cc] = ok[k] ? fmaxf(x[k][cc], 0.f) : 0.f;
    }
    float v[NV];
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) {
      v[cc] = (x[0][cc] + x[1][cc]) + (x[2][cc] + x[3][cc]);
      v[4 + cc] = fmaf(x[3][cc], x[3][cc], fmaf(x[2][cc], x[2][cc], fmaf(x[1][cc], x[1][cc], x[0][cc] * x[0][cc])));
    }
    if (kHead) {
      // dropout-masked column sums and kept-row counts (flags are zero on padding rows)
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) {
        const int cl = (cc & 1) + 8 * (cc >> 1);                   // column offset inside the chunk, before + 2q
        float ms = 0.f, cnt = 0.f;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int ch = col0 + cl + 2 * q;
          const uint32_t w = ((col0 + cl) >> 6) ? fw[k][1] : fw[k][0];
          const bool keep = ((w >> drop_bitpos(ch)) & 1u) != 0u && ok[k];
          ms += keep ? x[k][cc] : 0.f;
          cnt += keep ? 1.f : 0.f;
        }
        v[8 + cc] = ms;
        v[12 + cc] = cnt;
      }
    }
    quad_halve<NV>(v, c.lane);
    const int b4 = (c.lane >> 4) & 1, b3 = (c.lane >> 3) & 1, b2 = (c.lane >> 2) & 1;
    if (!kHead) {
      // NV = 8: lane bit 4 chose sums / sums of squares, bits 3 and 2 the column cc = 2 * b3 + b2
      const int ch = col0 + b2 + 8 * b3 + 2 * q;
      stat_add(s_stats, 2 * ch + b4, v[0]);
    } else {
      // NV = 16: bit 4 chose plain / dropout-masked, bit 3 chose (sum | sum of squares) resp. (masked sum | kept count),
      // bit 2 the column pair; v[0], v[1] are the two columns of the pair
      const int ch = col0 + 8 * b2 + 2 * q;
      if (!b4) {
        stat_add(s_stats, 2 * ch + b3, v[0]);
        stat_add(s_stats, 2 * (ch + 1) + b3, v[1]);
      } else if (head && win_ok) {
        if (!b3) {
          a.head_sum[head_row + ch] = v[0];
          a.head_sum[head_row + ch + 1] = v[1];
        } else {
          a.head_cnt[head_row + ch] = (unsigned char)__float2int_rn(v[0]);
          a.head_cnt[head_row + ch + 1] = (unsigned char)__float2int_rn(v[1]);
        }
      }
    }
    if (j + 1 < my) {
#pragma unroll
      for (int i = 0; i < 8; ++i) { lo[i] = lo2[i]; hi[i] = hi2[i]; }
    }
  }
  tc_fence_before();   // TMEM reads are complete before this thread lets the next unit's MMAs start
}

// Load stage of a resumed statistics sweep: the raw (post-ReLU, bf16) output of layer L0 has landed in this tile's
// activation planes (bulk copies); every thread turns ITS row of ITS 16-column chunks into layer L0+1's input in place:
// BatchNorm affine of pass g (dropout scale folded) -> dropout -> bf16, exactly the arithmetic of the fused epilogue.
// kb: the row's keep-flags of layer L0 (make_keep_flags), produced earlier in the shadow of the previous unit's MMAs.
__device__ __forceinline__ void ld_shared_v4_nc(uint32_t addr, uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d) {
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr));
}
template <int L0, int DROP>
__device__ __forceinline__ void load_stage(const EpiCtx& c, const uint32_t (&kb)[4], const float* tab_s, const float* tab_h) {
  constexpr int my = fl_cout(L0) / 32;
  const uint32_t row_mask = c.row_valid ? 0xFFFFFFFFu : 0u;
  float zdummy = 0.f;
  // a thread reads and rewrites only its own 16 bytes of every plane, so the next chunk's loads may pass this chunk's
  // stores: they are issued one chunk ahead
  uint32_t u[8], un[8];
  ld_shared_v4_nc(c.act_row + (uint32_t)(2 * c.colhalf) * kPlaneBytes, u[0], u[1], u[2], u[3]);
  ld_shared_v4_nc(c.act_row + (uint32_t)(2 * c.colhalf + 1) * kPlaneBytes, u[4], u[5], u[6], u[7]);
#pragma unroll
  for (int j = 0; j < my; ++j) {
    const int c0 = 16 * (c.colhalf + 2 * j);
    if (j + 1 < my) {
      ld_shared_v4_nc(c.act_row + (uint32_t)((c0 >> 3) + 4) * kPlaneBytes, un[0], un[1], un[2], un[3]);
      ld_shared_v4_nc(c.act_row + (uint32_t)((c0 >> 3) + 5) * kPlaneBytes, un[4], un[5], un[6], un[7]);
    }
    uint32_t v[16];
#pragma unroll
    for (int i = 0; i < 8; ++i) { v[2 * i] = u[i] << 16; v[2 * i + 1] = u[i] & 0xFFFF0000u; }
    const uint32_t zq[4] = {keep_extract(kb, 4 * j), keep_extract(kb, 4 * j + 1), keep_extract(kb, 4 * j + 2), keep_extract(kb, 4 * j + 3)};
    epi_chunk16<DROP, false, false, false>(v, c0, tab_s, tab_h, nullptr, c.act_row, c.row_valid, row_mask, 0u, c.c0w, 0u, 0u,
                                           c.k0, c.k1, zq, zdummy, nullptr);
    if (j + 1 < my) {
#pragma unroll
      for (int i = 0; i < 8; ++i) u[i] = un[i];
    }
  }
}

// ---- epilogue loop of a statistics sweep (exact batch-statistic mode), one tile of the two-tile CTA-pair kernel ------
// The epilogue warps walk the passes in the OUTER loop so that one pass's statistics are accumulated in shared memory
// and flushed once.  Resumed sweep (first_layer >= 2): the input of a unit is the raw post-ReLU output of layer
// L0 = first_layer - 1, fetched by bulk copies straight into the tile's activation planes ([plane][60 rows] x 16 B per
// window in HBM, the same order as a plane in shared memory) and normalised in place by load_stage.  The copies of
// iteration q + 1 are issued as soon as the MMAs of iteration q have retired, so they overlap the tile's epilogue.
struct StatGeo {
  uint32_t act_base;     // shared-memory address of this tile's activation planes
  uint32_t bar_in;       // mbarrier: this tile's input planes have landed
  uint32_t rank;         // CTA rank in the pair
  long long pair_id, n_pairs;
  int all_threads;       // epilogue threads of BOTH tiles (named barrier 3)
};

template <int DROP, int EPI_THREADS, int WPU>
__device__ __forceinline__ void stat_epilogue_loop(EpiCtx& c, const FusedArgs& a, const StatGeo& geo,
                                                   unsigned long long* s_stats, uint32_t& n_acc) {
  for (int i = c.tl * EPI_THREADS + c.tid_t; i < 2 * kMaxCh; i += geo.all_threads) s_stats[i] = 0ull;
  named_bar_sync(3, geo.all_threads);
  const int L0 = a.first_layer - 1;
  const int planes_in = L0 >= 1 ? c_layers[L0].cout / 8 : 0;
  const long long my_units = geo.pair_id < a.n_units ? (a.n_units - geo.pair_id + geo.n_pairs - 1) / geo.n_pairs : 0;
  const long long total_it = my_units * a.n_groups;
  const long long tile_win0 = 4 * c.tl + 2 * (long long)geo.rank;     // first window of this tile inside a unit
  auto issue_in = [&](long long q) {            // whole warp we == 0 of the tile
    if (planes_in == 0 || q >= total_it) return;
    const long long gq = q / my_units, uq = geo.pair_id + (q % my_units) * geo.n_pairs;
    const long long w0 = uq * WPU + tile_win0;
    const bool dbg_no_in = (a.variant & 4) != 0;      // perf-decomposition switch (tools only): no input copies, results WRONG
    const int valid = dbg_no_in ? 0 : (w0 < a.n ? 1 : 0) + (w0 + 1 < a.n ? 1 : 0);
    if (c.lane == 0) mbar_arrive_expect_tx(geo.bar_in, (uint32_t)(valid * planes_in * kL * 16));
    __syncwarp();
    for (int idx = c.lane; idx < 2 * planes_in; idx += 32) {
      const int slot = idx >= planes_in ? 1 : 0, pl = idx - slot * planes_in;
      if (w0 + slot < a.n && !dbg_no_in)
        bulk_g2s(geo.act_base + (uint32_t)pl * kPlaneBytes + (uint32_t)(4 + 64 * slot) * 16u,
                 a.ckpt_in + (((size_t)gq * a.n + (size_t)(w0 + slot)) * planes_in + pl) * kL, (uint32_t)(kL * 16), geo.bar_in);
    }
  };
  if (c.we == 0) issue_in(0);
  // keep-flags of layer L0 for the load stage of (pass g2, unit u2): data-independent, so they are produced one
  // iteration ahead, while the MMAs of the current unit run
  auto flags_for = [&](int g2, long long u2, uint32_t (&kb)[4]) {
    EpiCtx cn = c;
    const long long win2 = u2 * WPU + tile_win0 + c.slot_w;
    cn.row_valid = (c.t < kL) && (win2 < a.n);
    cn.c0w = (uint32_t)(a.window_base + win2);
    epi_pass_setup<DROP>(cn, a, g2);
    switch (L0) {
      case 1: make_keep_flags<1, DROP>(cn, a, g2, win2, kb); break;
      case 2: make_keep_flags<2, DROP>(cn, a, g2, win2, kb); break;
      case 3: make_keep_flags<3, DROP>(cn, a, g2, win2, kb); break;
      default: make_keep_flags<4, DROP>(cn, a, g2, win2, kb); break;
    }
  };
  uint32_t kb_cur[4] = {0u, 0u, 0u, 0u}, kb_next[4] = {0u, 0u, 0u, 0u};
  if (planes_in > 0 && total_it > 0) flags_for(0, geo.pair_id, kb_cur);
  uint32_t n_in = 0;
  long long q = 0;
  float* tab_s = c.tab + (L0 & 1) * 512;
  float* tab_h = tab_s + 256;
  for (int g = 0; g < a.n_groups; ++g) {
    const float* tab_g = a.tables + (size_t)g * kTableFloats;
    if (planes_in > 0) {
      // BatchNorm affine of layer L0 for pass g (dropout scale folded), used by the load stage of every unit
      for (int i = c.tid_t; i < 8 * planes_in; i += EPI_THREADS) {
        tab_s[i] = __ldg(tab_g + kChTotal + c_ch_off[L0] + i);
        tab_h[i] = __ldg(tab_g + 2 * kChTotal + c_ch_off[L0] + i);
      }
      named_bar_sync(1 + c.tl, EPI_THREADS);
    }
    for (long long unit = geo.pair_id; unit < a.n_units; unit += geo.n_pairs, ++q) {
      const long long win = unit * WPU + tile_win0 + c.slot_w;
      c.row_valid = (c.t < kL) && (win < a.n);
      c.c0w = (uint32_t)(a.window_base + win);
      if (c.colhalf == 1) st_shared_v4(c.act_row + (uint32_t)kOnesPlane * kPlaneBytes, c.row_valid ? 0x3F803F80u : 0u, 0u, 0u, 0u);
      if (a.first_layer == 0) {
        epi_pass_setup<DROP>(c, a, g);
        if (c.colhalf == 0) {
          // this row's 4 input features as bf16: channels 0..3 real, 4..7 zero, in plane 0
          uint32_t xin0 = 0, xin1 = 0;
          if (c.row_valid) {
            float f0, f1, f2, f3;
            if (a.x_f64) {
              const double2* px = reinterpret_cast<const double2*>(static_cast<const double*>(a.x) + ((size_t)win * kL + c.t) * kF);
              const double2 u = __ldg(px), v = __ldg(px + 1);
              f0 = (float)u.x; f1 = (float)u.y; f2 = (float)v.x; f3 = (float)v.y;
            } else {
              const float4 u = __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(a.x) + ((size_t)win * kL + c.t) * kF));
              f0 = u.x; f1 = u.y; f2 = u.z; f3 = u.w;
            }
            xin0 = pack_bf16x2(f0, f1);
            xin1 = pack_bf16x2(f2, f3);
          }
          st_shared_v4(c.act_row, xin0, xin1, 0u, 0u);
        }
      } else {
        mbar_wait2(geo.bar_in, n_in & 1, a.err_flag, 5);
        ++n_in;
        switch (L0) {
          case 1: load_stage<1, DROP>(c, kb_cur, tab_s, tab_h); break;
          case 2: load_stage<2, DROP>(c, kb_cur, tab_s, tab_h); break;
          case 3: load_stage<3, DROP>(c, kb_cur, tab_s, tab_h); break;
          default: load_stage<4, DROP>(c, kb_cur, tab_s, tab_h); break;
        }
      }
      fence_proxy_async();
      epi_signal_act<EPI_THREADS>(c);
      auto after_acc = [&]() { if (c.we == 0) issue_in(q + 1); };
      auto in_shadow = [&]() {
        if (planes_in > 0 && q + 1 < total_it) {
          const bool same_pass = unit + geo.n_pairs < a.n_units;
          flags_for(same_pass ? g : g + 1, same_pass ? unit + geo.n_pairs : geo.pair_id, kb_next);
        }
      };
      switch (a.first_layer) {
        case 0:
          if (a.n_layers == 1) {
            epi_stats<0, DROP, EPI_THREADS>(c, a, g, win, n_acc, s_stats, in_shadow, after_acc);
          } else {
            epi_layer<0, DROP, false, EPI_THREADS>(c, a, tab_g, g, win, n_acc, nullptr, nullptr);
            epi_stats<1, DROP, EPI_THREADS>(c, a, g, win, n_acc, s_stats, in_shadow, after_acc);
          }
          break;
        case 2: epi_stats<2, DROP, EPI_THREADS>(c, a, g, win, n_acc, s_stats, in_shadow, after_acc); break;
        case 3: epi_stats<3, DROP, EPI_THREADS>(c, a, g, win, n_acc, s_stats, in_shadow, after_acc); break;
        case 4: epi_stats<4, DROP, EPI_THREADS>(c, a, g, win, n_acc, s_stats, in_shadow, after_acc); break;
        default: epi_stats<5, DROP, EPI_THREADS>(c, a, g, win, n_acc, s_stats, in_shadow, after_acc); break;
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) kb_cur[i] = kb_next[i];
    }
    // this pass is complete in both tiles of this CTA: add its fixed-point sums to the global accumulators of pass g
    named_bar_sync(3, geo.all_threads);
    for (int i = c.tl * EPI_THREADS + c.tid_t; i < 2 * kMaxCh; i += geo.all_threads) {
      const unsigned long long v = s_stats[i];
      if (v != 0ull) { atomicAdd(a.stats + (size_t)g * 2 * kMaxCh + i, v); s_stats[i] = 0ull; }
    }
    named_bar_sync(3, geo.all_threads);
  }
}

// ---- pipelined statistics sweep (resumed sweeps, first_layer >= 2; option "exact_pipelined", OFF by default) ----------
// Measured (262 k windows x 4 passes): 133 ms against 129 ms for the default two-tile sweep -- the load stage and the
// statistics epilogue are latency-bound instruction streams (IPC 0.12 per scheduler at 4 resident warps), so overlapping
// them with the MMAs through two buffers does not pay for the halved weight reuse (R = 4).  Kept as a tested variant
// (bit-identical results, tests/test_gpu_exact.py).
// One stream of 4-window units per CTA pair, software-pipelined over TWO activation buffers and TWO accumulators:
// while the tensor cores run the MMAs of unit q (buffer q & 1), the 8 LOADER warps normalise the input of unit q + 1 in
// the other buffer and the 8 EPILOGUE warps reduce / store the accumulator of unit q - 1.  The three stages only meet
// at mbarriers:
//   in[b]    bulk copies of a unit's raw input have landed in buffer b        (TMA -> loaders)
//   act[b]   buffer b holds normalised activations                           (loaders, both CTAs -> issuer)
//   acc[b]   the MMAs reading buffer b / writing accumulator b have retired   (issuer -> epilogue, both CTAs)
//   tfree[b] accumulator b has been read out                                  (epilogue, both CTAs -> issuer)
// The epilogue warp that observes acc[b] of unit q also issues the input copies of unit q + 2 into buffer b.
struct PipeGeo {
  uint32_t act_base;     // shared-memory address of activation buffer 0 (buffer 1 = + kActBytes)
  uint32_t bar_in, bar_acc, bar_act, bar_tfree;   // barrier pairs [2]
  uint32_t tmem_base;
  uint32_t rank;
  long long pair_id, n_pairs;
};

template <int DROP, int EPI_THREADS>
__device__ __forceinline__ void stat_pipelined_loop(EpiCtx& c, const FusedArgs& a, const PipeGeo& geo, unsigned long long* s_stats) {
  constexpr int WPU = 4;
  const int L0 = a.first_layer - 1;
  const int planes_in = c_layers[L0].cout / 8;
  const long long my_units = geo.pair_id < a.n_units ? (a.n_units - geo.pair_id + geo.n_pairs - 1) / geo.n_pairs : 0;
  const long long total_it = my_units * a.n_groups;
  const int row = c.quarter * 32 + c.lane;
  const bool leader = geo.rank == 0;
  const bool strong = c.strong_sync;
  auto unit_of = [&](long long q, int& g, long long& unit) { g = (int)(q / my_units); unit = geo.pair_id + (q % my_units) * geo.n_pairs; };
  if (c.tl == 0) {
    // =========================== loaders ===========================
    uint32_t n_in0 = 0u, n_in1 = 0u;       // phase counters of in[0], in[1] (scalars: a runtime-indexed array would live in local memory)
    float* tab_s = c.tab + (L0 & 1) * 512;
    float* tab_h = tab_s + 256;
    const uint32_t r_act = leader ? 0u : mapa_shared(geo.bar_act, 0);
    int g_tab = -1;
    for (long long q = 0; q < total_it; ++q) {
      const int b = (int)(q & 1);
      int g; long long unit;
      unit_of(q, g, unit);
      const long long win = unit * WPU + 2 * (long long)geo.rank + c.slot_w;
      c.row_valid = (c.t < kL) && (win < a.n);
      c.c0w = (uint32_t)(a.window_base + win);
      c.act_row = geo.act_base + (uint32_t)b * kActBytes + (uint32_t)(row + 4) * 16u;
      if (g != g_tab) {
        // BatchNorm affine of layer L0 for pass g (dropout scale folded); the previous pass's last load stage is done
        named_bar_sync(1, EPI_THREADS);
        const float* tab_g = a.tables + (size_t)g * kTableFloats;
        for (int i = c.tid_t; i < 8 * planes_in; i += EPI_THREADS) {
          tab_s[i] = __ldg(tab_g + kChTotal + c_ch_off[L0] + i);
          tab_h[i] = __ldg(tab_g + 2 * kChTotal + c_ch_off[L0] + i);
        }
        named_bar_sync(1, EPI_THREADS);
        g_tab = g;
      }
      // the row's keep-flags of layer L0 (data independent: produced while the copies are in flight)
      uint32_t kb[4];
      epi_pass_setup<DROP>(c, a, g);
      switch (L0) {
        case 1: make_keep_flags<1, DROP>(c, a, g, win, kb); break;
        case 2: make_keep_flags<2, DROP>(c, a, g, win, kb); break;
        case 3: make_keep_flags<3, DROP>(c, a, g, win, kb); break;
        default: make_keep_flags<4, DROP>(c, a, g, win, kb); break;
      }
      mbar_wait2(geo.bar_in + 8 * b, (b ? n_in1 : n_in0) & 1, a.err_flag, 5);      // implies: the MMAs of unit q - 2 have left this buffer
      if (b) ++n_in1; else ++n_in0;
      if (c.colhalf == 1) st_shared_v4(c.act_row + (uint32_t)kOnesPlane * kPlaneBytes, c.row_valid ? 0x3F803F80u : 0u, 0u, 0u, 0u);
      switch (L0) {
        case 1: load_stage<1, DROP>(c, kb, tab_s, tab_h); break;
        case 2: load_stage<2, DROP>(c, kb, tab_s, tab_h); break;
        case 3: load_stage<3, DROP>(c, kb, tab_s, tab_h); break;
        default: load_stage<4, DROP>(c, kb, tab_s, tab_h); break;
      }
      fence_proxy_async();
      if (leader) {
        __syncwarp();
        if (c.lane == 0) mbar_arrive(geo.bar_act + 8 * b);
      } else {
        named_bar_sync(1, EPI_THREADS);
        if (c.tid_t == 0) mbar_arrive_remote(r_act + 8 * b, strong);
      }
    }
  } else {
    // =========================== epilogue ===========================
    for (int i = c.tid_t; i < 2 * kMaxCh; i += EPI_THREADS) s_stats[i] = 0ull;
    named_bar_sync(2, EPI_THREADS);
    auto issue_in = [&](long long q) {            // whole warp we == 0: input copies of unit q into buffer q & 1
      if (q >= total_it) return;
      const int b = (int)(q & 1);
      int gq; long long uq;
      unit_of(q, gq, uq);
      const long long w0 = uq * WPU + 2 * (long long)geo.rank;
      const bool dbg_no_in = (a.variant & 4) != 0;
      const int valid = dbg_no_in ? 0 : (w0 < a.n ? 1 : 0) + (w0 + 1 < a.n ? 1 : 0);
      if (c.lane == 0) mbar_arrive_expect_tx(geo.bar_in + 8 * b, (uint32_t)(valid * planes_in * kL * 16));
      __syncwarp();
      for (int idx = c.lane; idx < 2 * planes_in; idx += 32) {
        const int slot = idx >= planes_in ? 1 : 0, pl = idx - slot * planes_in;
        if (w0 + slot < a.n && !dbg_no_in)
          bulk_g2s(geo.act_base + (uint32_t)b * kActBytes + (uint32_t)pl * kPlaneBytes + (uint32_t)(4 + 64 * slot) * 16u,
                   a.ckpt_in + (((size_t)gq * a.n + (size_t)(w0 + slot)) * planes_in + pl) * kL, (uint32_t)(kL * 16), geo.bar_in + 8 * b);
      }
    };
    if (c.we == 0) { issue_in(0); issue_in(1); }
    uint32_t n_acc0 = 0u, n_acc1 = 0u;
    const uint32_t r_tfree = leader ? 0u : mapa_shared(geo.bar_tfree, 0);
    const uint32_t taddr0 = geo.tmem_base + ((uint32_t)(c.quarter * 32) << 16);
    auto nothing = []() {};
    for (long long q = 0; q < total_it; ++q) {
      const int b = (int)(q & 1);
      int g; long long unit;
      unit_of(q, g, unit);
      const long long win = unit * WPU + 2 * (long long)geo.rank + c.slot_w;
      c.row_valid = (c.t < kL) && (win < a.n);
      c.c0w = (uint32_t)(a.window_base + win);
      c.taddr = taddr0 + (uint32_t)(b * 256);
      c.bar_acc = geo.bar_acc + 8 * b;
      auto after_acc = [&]() { if (c.we == 0) issue_in(q + 2); };
      uint32_t na = b ? n_acc1 : n_acc0;
      switch (a.first_layer) {
        case 2: epi_stats<2, DROP, EPI_THREADS>(c, a, g, win, na, s_stats, nothing, after_acc); break;
        case 3: epi_stats<3, DROP, EPI_THREADS>(c, a, g, win, na, s_stats, nothing, after_acc); break;
        case 4: epi_stats<4, DROP, EPI_THREADS>(c, a, g, win, na, s_stats, nothing, after_acc); break;
        default: epi_stats<5, DROP, EPI_THREADS>(c, a, g, win, na, s_stats, nothing, after_acc); break;
      }
      if (b) n_acc1 = na; else n_acc0 = na;
      // accumulator b has been read out (tcgen05.wait::ld + fence in epi_stats): the MMAs of unit q + 2 may overwrite it
      if (leader) {
        __syncwarp();
        if (c.lane == 0) mbar_arrive(geo.bar_tfree + 8 * b);
      } else {
        named_bar_sync(2, EPI_THREADS);
        if (c.tid_t == 0) mbar_arrive_remote(r_tfree + 8 * b, strong);
      }
      if (q + 1 == total_it || (q + 1) % my_units == 0) {
        // pass g is complete in this CTA: add its fixed-point sums to the global accumulators of pass g
        named_bar_sync(2, EPI_THREADS);
        for (int i = c.tid_t; i < 2 * kMaxCh; i += EPI_THREADS) {
          const unsigned long long v = s_stats[i];
          if (v != 0ull) { atomicAdd(a.stats + (size_t)g * 2 * kMaxCh + i, v); s_stats[i] = 0ull; }
        }
        named_bar_sync(2, EPI_THREADS);
      }
    }
  }
}

template <int TILES, int NSLOTS, int DROP, bool DBG>
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
  float* zpart = reinterpret_cast<float*>(smem + Cfg::oMisc + 16);  // [TILES][2 windows][4 partial sums]
  uint32_t* s_alo = reinterpret_cast<uint32_t*>(smem + Cfg::oAlo);  // [kTotalSteps] A-descriptor low words (tile 0)
  float* probs_s = reinterpret_cast<float*>(smem + Cfg::oProbs);    // [kWpu][n_groups]

  // ---- one-time set-up
  for (int i = threadIdx.x; i < (TILES * kActBytes) / 16; i += Cfg::kThreads)
    reinterpret_cast<uint4*>(smem + Cfg::oAct)[i] = make_uint4(0, 0, 0, 0);   // halo rows stay zero for ever
  for (int i = threadIdx.x; i < kTotalSteps; i += Cfg::kThreads) {
    int l = 0, sidx = i;
    while (sidx >= c_layers[l].nsteps) { sidx -= c_layers[l].nsteps; ++l; }
    const FusedLayer fl = c_layers[l];
    uint32_t a_off, a_lbo;
    if (sidx == fl.nconv) { a_off = (uint32_t)kOnesPlane * kPlaneBytes + 4u * 16u; a_lbo = 16u; }          // bias step
    else if (l == 0) { a_off = (uint32_t)(2 * sidx + 1) * 16u; a_lbo = 16u; }                               // tap pair
    else {
      const int tap = sidx % fl.taps, slab = sidx / fl.taps;   // slab-major K order (see fused_pack_member)
      a_off = (uint32_t)(2 * slab) * kPlaneBytes + (uint32_t)(tap - fl.halo + 4) * 16u;                     // shifted rows
      a_lbo = kPlaneBytes;
    }
    s_alo[i] = (((s_base + Cfg::oAct + a_off) >> 4) & 0x3FFFu) | ((a_lbo >> 4) << 16);
  }
  if (threadIdx.x == 0) {
    for (int i = 0; i < NSLOTS; ++i) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, 1); }
    for (int i = 0; i < TILES; ++i) { mbar_init(bar_acc + 8 * i, 1); mbar_init(bar_act + 8 * i, Cfg::kEpiThreads / 32); }
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

  // perf-decomposition switches (tools only): results are WRONG when any is set
  const bool dbg_no_copy = (a.variant & 2) != 0;   // producer does not fetch weights
  const bool dbg_no_mma = (a.variant & 8) != 0;    // issuer skips the MMAs

  if (warp == 0) {
    // =========================== weight producer (one thread) ===========================
    if (lane == 0) {
      uint32_t it = 0;
      for (long long unit = blockIdx.x; unit < a.n_units; unit += gridDim.x) {
        for (int g = 0; g < a.n_groups; ++g) {
          const uint8_t* wmem = a.packed + (size_t)(a.mode_de ? g : 0) * kPackedBytes;
#pragma unroll 1
          for (int l = 0; l < kBlocks; ++l) {
            const FusedLayer fl = c_layers[l];
            int step = 0;
            for (int st = 0; st < fl.nstages; ++st, ++it) {
              const int slot = it % NSLOTS;
              const uint32_t ph = (it / NSLOTS) & 1;
              mbar_wait2(bar_empty + 8 * slot, ph ^ 1, a.err_flag, 1);
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
    // =========================== MMA issuer ===========================
    uint32_t it = 0, n_act = 0;
    for (long long unit = blockIdx.x; unit < a.n_units; unit += gridDim.x) {
      for (int g = 0; g < a.n_groups; ++g) {
        issue_layer<0, TILES, NSLOTS, Cfg>(s_base, tmem_base, s_alo, bar_full, bar_empty, bar_acc, bar_act, it, n_act, a.err_flag, dbg_no_mma);
        issue_layer<1, TILES, NSLOTS, Cfg>(s_base, tmem_base, s_alo, bar_full, bar_empty, bar_acc, bar_act, it, n_act, a.err_flag, dbg_no_mma);
        issue_layer<2, TILES, NSLOTS, Cfg>(s_base, tmem_base, s_alo, bar_full, bar_empty, bar_acc, bar_act, it, n_act, a.err_flag, dbg_no_mma);
        issue_layer<3, TILES, NSLOTS, Cfg>(s_base, tmem_base, s_alo, bar_full, bar_empty, bar_acc, bar_act, it, n_act, a.err_flag, dbg_no_mma);
        issue_layer<4, TILES, NSLOTS, Cfg>(s_base, tmem_base, s_alo, bar_full, bar_empty, bar_acc, bar_act, it, n_act, a.err_flag, dbg_no_mma);
        issue_layer<5, TILES, NSLOTS, Cfg>(s_base, tmem_base, s_alo, bar_full, bar_empty, bar_acc, bar_act, it, n_act, a.err_flag, dbg_no_mma);
      }
    }
  } else {
    // =========================== epilogue warps: 8 per tile ===========================
    const int ew = warp - 2;
    EpiCtx c;
    c.tl = ew >> 3;
    c.we = ew & 7;
    c.lane = lane;
    c.quarter = warp & 3;                         // TMEM lanes [32*quarter, +32) are accessible to this warp
    c.colhalf = c.we >> 2;                        // which half of the 16-column chunks this warp owns
    const int row = c.quarter * 32 + lane;        // accumulator row = TMEM lane
    c.slot_w = row >> 6;                          // which of the tile's two windows
    c.t = row & 63;                               // time step (60..63 = zero rows)
    c.tid_t = c.we * 32 + lane;                   // 0..255 within the tile's epilogue group
    c.tab = reinterpret_cast<float*>(smem + Cfg::oTab) + c.tl * Cfg::kTabFloats;
    c.tab_w = c.tab + 1024;
    c.act_row = s_base + Cfg::oAct + c.tl * kActBytes + (uint32_t)(row + 4) * 16u;
    c.taddr = tmem_base + ((uint32_t)(c.quarter * 32) << 16) + (uint32_t)(c.tl * 256);
    c.zp = zpart + c.tl * 8;
    c.bar_acc = bar_acc + 8 * c.tl;
    c.bar_act = bar_act + 8 * c.tl;
    c.k0 = a.k0; c.k1 = a.k1;
    c.err_flag = a.err_flag;
    c.strong_sync = (a.variant & 1) != 0;
    uint32_t n_acc = 0;

    for (long long unit = blockIdx.x; unit < a.n_units; unit += gridDim.x) {
      const long long win = unit * Cfg::kWpu + 2 * c.tl + c.slot_w;
      c.row_valid = (c.t < kL) && (win < a.n);
      // this row's 4 input features as bf16 (kept in registers for all groups of the unit)
      uint32_t xin0 = 0, xin1 = 0;
      if (c.colhalf == 0) {
        if (c.row_valid) {
          float f0, f1, f2, f3;
          if (a.x_f64) {
            const double2* p = reinterpret_cast<const double2*>(static_cast<const double*>(a.x) + ((size_t)win * kL + c.t) * kF);
            const double2 u = __ldg(p), v = __ldg(p + 1);
            f0 = (float)u.x; f1 = (float)u.y; f2 = (float)v.x; f3 = (float)v.y;
          } else {
            const float4 u = __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(a.x) + ((size_t)win * kL + c.t) * kF));
            f0 = u.x; f1 = u.y; f2 = u.z; f3 = u.w;
          }
          xin0 = pack_bf16x2(f0, f1);
          xin1 = pack_bf16x2(f2, f3);
        }
      } else {
        // constant plane [1, 1, 0, ...] on valid rows: the bias step of every layer reads it as its A operand
        st_shared_v4(c.act_row + (uint32_t)kOnesPlane * kPlaneBytes, c.row_valid ? 0x3F803F80u : 0u, 0u, 0u, 0u);
      }
      c.c0w = (uint32_t)(a.window_base + win);
      float* dbg_base = nullptr;
      if (DBG) {
        if (a.dbg_act && blockIdx.x == 0 && unit == blockIdx.x && c.tl == 0) dbg_base = a.dbg_act + (size_t)row * 256;
      }

      for (int g = 0; g < a.n_groups; ++g) {
        epi_pass_setup<DROP>(c, a, g);
        // ---- layer-1 input: channels 0..3 real, 4..7 zero, in plane 0
        if (c.colhalf == 0) st_shared_v4(c.act_row, xin0, xin1, 0u, 0u);
        fence_proxy_async();
        epi_signal_act<Cfg::kEpiThreads>(c);
        const float* tab_g = a.tables + (size_t)((a.mode_de || a.tab_per_group) ? g : 0) * kTableFloats;
        float* dbg_g = (DBG && g == 0) ? dbg_base : nullptr;
        epi_layer<0, DROP, DBG, Cfg::kEpiThreads>(c, a, tab_g, g, win, n_acc, probs_s, dbg_g);
        epi_layer<1, DROP, DBG, Cfg::kEpiThreads>(c, a, tab_g, g, win, n_acc, probs_s, dbg_g);
        epi_layer<2, DROP, DBG, Cfg::kEpiThreads>(c, a, tab_g, g, win, n_acc, probs_s, dbg_g);
        epi_layer<3, DROP, DBG, Cfg::kEpiThreads>(c, a, tab_g, g, win, n_acc, probs_s, dbg_g);
        epi_layer<4, DROP, DBG, Cfg::kEpiThreads>(c, a, tab_g, g, win, n_acc, probs_s, dbg_g);
        epi_layer<5, DROP, DBG, Cfg::kEpiThreads>(c, a, tab_g, g, win, n_acc, probs_s, dbg_g);
      }
      // ---- K3: per-window reductions over the T (or M) probabilities, in the reference's evaluation order
      if (a.smem_probs && c.colhalf == 0 && c.t == 0 && (c.quarter & 1) == 0 && win < a.n) {
        const float* pw = probs_s + (2 * c.tl + c.slot_w) * a.n_groups;
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

// =====================================================================================================================
// CTA-pair variant (cta_group::2): a cluster of two CTAs on the two SMs of a TPC owns a unit of 4 windows; each CTA
// holds ONE tile (its 2 windows) and HALF of every weight stage; the leader CTA issues M=256 MMAs that span both SMs.
// Two pairs are co-resident per SM pair (2 CTAs / SM), so one pair's epilogue overlaps the other pair's MMAs, while
// the weight traffic per SM stays that of the two-tile kernel (R = 4 window-passes per stream).
// Peer -> leader signalling ("my half of stage s has landed", "my activations are ready") is relayed by the peer's
// otherwise idle warp 1 with remote mbarrier arrives; leader -> both signalling uses tcgen05.commit multicast.
// =====================================================================================================================
template <int NSLOTS>
struct PairCfg {
  static constexpr int kEpiThreads = 256;
  static constexpr int kWarps = 10;
  static constexpr int kThreads = 32 * kWarps;
  static constexpr int kWpu = 4;                   // windows per pair unit
  static constexpr int kTmemCols = 256;
  static constexpr int kSlotPair = kSlotBytes;
  static constexpr int kSlot = kSlotPair / 2;      // each CTA keeps half of a stage
  static constexpr int oAct = 0;
  static constexpr int oRing = oAct + kActBytes;
  static constexpr int kTabFloats = 2 * 512 + 128;
  static constexpr int oTab = oRing + NSLOTS * kSlot;
  static constexpr int oBar = oTab + kTabFloats * 4;
  static constexpr int kNumBars = 3 * NSLOTS + 3;  // full, empty, peer_full [NSLOTS]; acc, act, peer_act
  static constexpr int oMisc = oBar + 8 * kNumBars;
  static constexpr int oAlo = oMisc + 16 + 32;
  static constexpr int oProbs = oAlo + 4 * 392;    // probs[2][n_groups]
};

static __device__ __noinline__ void mbar_wait_cluster_slow(uint32_t bar, uint32_t parity, int* err_flag, int code, bool strong) {
  long long t0 = 0;
  for (;;) {
    uint32_t ok;
    if (strong) {
      asm volatile(
          "{\n\t.reg .pred p;\n\t.reg .b32 n;\n\t"
          "mov.u32 n, 2048;\n"
          "UQK_TRYA:\n\t"
          "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2, %3;\n\t"
          "@p bra UQK_DONEA;\n\t"
          "sub.u32 n, n, 1;\n\t"
          "setp.ne.u32 p, n, 0;\n\t"
          "@p bra UQK_TRYA;\n\t"
          "mov.u32 %0, 0;\n\t"
          "bra UQK_ENDA;\n"
          "UQK_DONEA:\n\t"
          "mov.u32 %0, 1;\n"
          "UQK_ENDA:\n\t}"
          : "=r"(ok)
          : "r"(bar), "r"(parity), "r"(kWaitHintNs)
          : "memory");
    } else {
      asm volatile(
          "{\n\t.reg .pred p;\n\t.reg .b32 n;\n\t"
          "mov.u32 n, 2048;\n"
          "UQK_TRY:\n\t"
          "mbarrier.try_wait.parity.relaxed.cluster.shared::cta.b64 p, [%1], %2, %3;\n\t"
          "@p bra UQK_DONE;\n\t"
          "sub.u32 n, n, 1;\n\t"
          "setp.ne.u32 p, n, 0;\n\t"
          "@p bra UQK_TRY;\n\t"
          "mov.u32 %0, 0;\n\t"
          "bra UQK_END;\n"
          "UQK_DONE:\n\t"
          "mov.u32 %0, 1;\n"
          "UQK_END:\n\t}"
          : "=r"(ok)
          : "r"(bar), "r"(parity), "r"(kWaitHintNs)
          : "memory");
    }
    if (ok) return;
    const long long now = clock64();
    if (t0 == 0) t0 = now;
    else if (err_flag && now - t0 > 4000000000ll) mbar_timeout(err_flag, code);   // err_flag == null: watchdog off
  }
}

// Cluster-scope wait of the MMA issuers.  Fast path: ONE try_wait (it suspends inside the instruction for up to
// kWaitHintNs, so in practice it returns exactly when the phase completes); the polling loop with the watchdog is a
// separate function, which keeps the issuer's per-stage instruction stream short (it is the latency-critical warp:
// every stage of 2-3 MMAs passes through it).
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity, int* err_flag, int code, bool strong = false) {
  uint32_t ok;
  if (strong)
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2, %3;\n\t"
                 "selp.b32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity), "r"(kWaitHintNs) : "memory");
  else
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "mbarrier.try_wait.parity.relaxed.cluster.shared::cta.b64 p, [%1], %2, %3;\n\t"
                 "selp.b32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity), "r"(kWaitHintNs) : "memory");
  if (!ok) mbar_wait_cluster_slow(bar, parity, err_flag, code, strong);
}

// ONE = true: called by a single thread (the two-tile pair kernel's leader: lane 0 of the issuer warp walks the schedule
// alone -- no election, no warp re-convergence per stage); ONE = false: called by a converged warp, one elected lane issues.
// EARLY = true (two-tile pair kernel, all six layers in one launch): layers 2, 4 and 6 start on the first half of their
// input slabs as soon as the previous layer's epilogue has signalled bar_half (see early_producer) and wait for the full
// hand-off (bar_act) only before the first step that reads a later slab; layer 1 accumulates at column 128.
template <int L, int NSLOTS, typename Cfg, bool ONE = false, bool EARLY = false>
__device__ __forceinline__ void issue_layer_pair(uint32_t s_base, uint32_t tmem_base, const uint32_t* s_alo,
                                                 uint32_t bar_full, uint32_t bar_empty, uint32_t bar_peer_full,
                                                 uint32_t bar_acc, uint32_t bar_act, uint32_t bar_peer_act, uint32_t& it,
                                                 uint32_t& n_act, int* err_flag, int flags, uint32_t a_tile = 0u,
                                                 long long* tlp = nullptr, uint32_t bar_half = 0u, uint32_t* n_half = nullptr) {
  const bool no_mma = (flags & 8) != 0, strong = (flags & 1) != 0;   // debug switch / release-acquire cluster signalling
  constexpr FusedLayer fl = fused_layer(L, Cfg::kSlotPair);
  constexpr uint32_t kHi = (128u >> 4) | (1u << 14);
  constexpr uint32_t kBLbo = ((uint32_t)((fl.cout / 2) * 16) >> 4) << 16;
  constexpr uint32_t kIdesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(fl.cout >> 3) << 17) | ((256u >> 4) << 24);
  constexpr int kLast = fl.nsteps - (fl.nstages - 1) * fl.sps;
  // ring position of the layer's first stage; advanced incrementally (no division per stage)
  uint32_t slot = it % NSLOTS, ph = (it / NSLOTS) & 1;
  const uint32_t ring_lo = ((s_base + Cfg::oRing) >> 4) | kBLbo;    // shared-memory addresses are < 2^18: no carry into the LBO field
  const uint32_t* alo = s_alo + fl_step_base(L);
  // early start: this layer's first kHalfSt stages read only the slabs the previous epilogue signals with bar_half
  constexpr bool kEarly = EARLY && L >= 1 && early_producer(L - 1);
  constexpr int kHalfSt = kEarly ? (2 * half_chunks(L >= 1 ? L - 1 : 0) * fl.taps) / fl.sps : 0;
  static_assert(!kEarly || (kHalfSt > 0 && kHalfSt < fl.nstages), "early start: the split must fall inside the layer");
  tmem_base += (uint32_t)acc_col(L, EARLY);
  if (kEarly) {
    mbar_wait_cluster(bar_half, *n_half & 1, err_flag, 7, strong);
    ++*n_half;
  } else {
    mbar_wait_cluster(bar_act, n_act & 1, err_flag, 2, strong);      // 8 local warp arrivals + 1 relayed from the peer
    ++n_act;
  }
  tc_fence_after();
#ifdef UQK_TIMELINE
  if (tlp) tlp[0] = clock64();
#endif
#ifdef UQK_TIMELINE
  long long tseg0 = 0, tseg1 = 0, tseg2 = 0, tq = clock64();
#define UQK_SEG(acc) do { const long long tn_ = clock64(); acc += tn_ - tq; tq = tn_; } while (0)
#else
#define UQK_SEG(acc) do { } while (0)
#endif
#pragma unroll 1
  for (int st = 0; st < fl.nstages; ++st) {
    // A-descriptor low words of the stage: read before the wait, so that the load latency hides behind it
    uint32_t av[fl.sps];
#pragma unroll
    for (int sl = 0; sl < fl.sps; ++sl) av[sl] = alo[st * fl.sps + sl] + a_tile;   // s_alo has 392 entries: a partial last stage reads past its layer, never past the array
    if (kEarly && st == kHalfSt) {
      mbar_wait_cluster(bar_act, n_act & 1, err_flag, 2, strong);    // the rest of the input is in place
      ++n_act;
      tc_fence_after();
    }
    UQK_SEG(tseg2);
    mbar_wait_cluster(bar_full + 8 * slot, ph, err_flag, 3, strong);   // own half landed (tx) + peer's half (relayed arrive)
    tc_fence_after();
    UQK_SEG(tseg0);
    const int nst = (st == fl.nstages - 1) ? kLast : fl.sps;
    const uint32_t b_lo0 = ring_lo + slot * (uint32_t)(Cfg::kSlot >> 4);
    if (ONE || elect_one()) {
      if (!no_mma) {
#pragma unroll
        for (int sl = 0; sl < fl.sps; ++sl) {
          if (sl < nst) {
            const uint64_t bdesc = mk64(b_lo0 + (uint32_t)sl * (uint32_t)(fl.step_bytes >> 5), kHi);
            umma_bf16_2cta(tmem_base, mk64(av[sl], kHi), bdesc, kIdesc, (st > 0 || sl > 0) ? 1u : 0u);
          }
        }
      }
      umma_commit_2cta(bar_empty + 8 * slot);   // both CTAs' ring slots are free once these MMAs have read them
    }
    if (!ONE) __syncwarp();
    UQK_SEG(tseg1);
    if (++slot == NSLOTS) { slot = 0; ph ^= 1u; }
  }
  it += fl.nstages;
  if (ONE || elect_one()) umma_commit_2cta(bar_acc);   // both CTAs' accumulators complete
  if (!ONE) __syncwarp();
#ifdef UQK_TIMELINE
  if (tlp) { tlp[1] = clock64(); tlp[1536] = tseg0; tlp[1537] = tseg1; tlp[1538] = tseg2; }   // second half of the buffer: time per segment
#endif
}

template <int L, int NSLOTS>
__device__ __forceinline__ void relay_layer(uint32_t bar_full, uint32_t bar_act, uint32_t r_peer_full, uint32_t r_peer_act,
                                            uint32_t& it, uint32_t& n_act, int* err_flag, bool strong) {
  constexpr FusedLayer fl = fused_layer(L);
  mbar_wait2(bar_act, n_act & 1, err_flag, 22);
  ++n_act;
  mbar_arrive_remote(r_peer_act, strong);
#pragma unroll 1
  for (int st = 0; st < fl.nstages; ++st, ++it) {
    const int slot = it % NSLOTS;
    mbar_wait2(bar_full + 8 * slot, (it / NSLOTS) & 1, err_flag, 23);
    mbar_arrive_remote(r_peer_full + 8 * slot, strong);
  }
}

template <int NSLOTS, int DROP, bool DBG>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(PairCfg<NSLOTS>::kThreads, 2)
fused_pair_kernel(const FusedArgs a) {
  using Cfg = PairCfg<NSLOTS>;
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t s_base = smem_u32(smem);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const long long pair_id = blockIdx.x >> 1;
  const long long n_pairs = gridDim.x >> 1;

  const uint32_t bar_full = s_base + Cfg::oBar;
  const uint32_t bar_empty = bar_full + 8 * NSLOTS;
  const uint32_t bar_peer_full = bar_empty + 8 * NSLOTS;        // used in the leader only
  const uint32_t bar_acc = bar_peer_full + 8 * NSLOTS;
  const uint32_t bar_act = bar_acc + 8;
  const uint32_t bar_peer_act = bar_act + 8;                    // used in the leader only
  volatile uint32_t* tmem_ptr = reinterpret_cast<volatile uint32_t*>(smem + Cfg::oMisc);
  float* zpart = reinterpret_cast<float*>(smem + Cfg::oMisc + 16);
  uint32_t* s_alo = reinterpret_cast<uint32_t*>(smem + Cfg::oAlo);
  float* probs_s = reinterpret_cast<float*>(smem + Cfg::oProbs);

  for (int i = threadIdx.x; i < kActBytes / 16; i += Cfg::kThreads)
    reinterpret_cast<uint4*>(smem + Cfg::oAct)[i] = make_uint4(0, 0, 0, 0);
  for (int i = threadIdx.x; i < kTotalSteps; i += Cfg::kThreads) {
    int l = 0, sidx = i;
    while (sidx >= c_layers[l].nsteps) { sidx -= c_layers[l].nsteps; ++l; }
    const FusedLayer fl = c_layers[l];
    uint32_t a_off, a_lbo;
    if (sidx == fl.nconv) { a_off = (uint32_t)kOnesPlane * kPlaneBytes + 4u * 16u; a_lbo = 16u; }
    else if (l == 0) { a_off = (uint32_t)(2 * sidx + 1) * 16u; a_lbo = 16u; }
    else {
      const int tap = sidx % fl.taps, slab = sidx / fl.taps;   // slab-major K order (see fused_pack_member)
      a_off = (uint32_t)(2 * slab) * kPlaneBytes + (uint32_t)(tap - fl.halo + 4) * 16u;
      a_lbo = kPlaneBytes;
    }
    s_alo[i] = (((s_base + Cfg::oAct + a_off) >> 4) & 0x3FFFu) | ((a_lbo >> 4) << 16);
  }
  if (threadIdx.x == 0) {
    // leader: every "full" / "act" barrier also counts one arrival relayed from the peer CTA
    const uint32_t extra = rank == 0 ? 1u : 0u;
    for (int i = 0; i < NSLOTS; ++i) {
      mbar_init(bar_full + 8 * i, 1 + extra); mbar_init(bar_empty + 8 * i, 1); mbar_init(bar_peer_full + 8 * i, 1);
    }
    mbar_init(bar_acc, 1); mbar_init(bar_act, Cfg::kEpiThreads / 32 + extra); mbar_init(bar_peer_act, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(const_cast<uint32_t*>(tmem_ptr))),
                 "r"((uint32_t)Cfg::kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  fence_proxy_async();
  tc_fence_before();
  cluster_sync_all();          // barriers of BOTH CTAs are initialised before anyone signals across the pair
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const bool dbg_no_copy = (a.variant & 2) != 0;
  const bool dbg_no_mma = (a.variant & 8) != 0;

  if (warp == 0) {
    // ---- weight producer: this CTA's half of every stage
    if (lane == 0) {
      uint32_t it = 0;
      for (long long unit = pair_id; unit < a.n_units; unit += n_pairs) {
        for (int g = 0; g < a.n_groups; ++g) {
          const uint8_t* wmem = a.packed + (size_t)(a.mode_de ? g : 0) * kPackedBytes + (size_t)rank * (kPackedBytes / 2);
#pragma unroll 1
          for (int l = 0; l < kBlocks; ++l) {
            const FusedLayer fl = c_layers[l];
            const uint32_t sb = (uint32_t)fl.step_bytes / 2;
            int step = 0;
            for (int st = 0; st < fl.nstages; ++st, ++it) {
              const int slot = it % NSLOTS;
              const uint32_t ph = (it / NSLOTS) & 1;
              mbar_wait2(bar_empty + 8 * slot, ph ^ 1, a.err_flag, 1);
              const int nst = min(fl.sps, fl.nsteps - step);
              const uint32_t bytes = (uint32_t)nst * sb;
              if (dbg_no_copy) { mbar_arrive(bar_full + 8 * slot); }
              else {
                mbar_arrive_expect_tx(bar_full + 8 * slot, bytes);
                bulk_g2s(s_base + Cfg::oRing + slot * Cfg::kSlot, wmem + fl.w_off / 2 + (size_t)step * sb, bytes,
                         bar_full + 8 * slot);
              }
              step += nst;
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    uint32_t it = 0, n_act = 0;
    if (rank == 0) {
      // ---- leader: MMA issuer for the pair
      for (long long unit = pair_id; unit < a.n_units; unit += n_pairs) {
        for (int g = 0; g < a.n_groups; ++g) {
          issue_layer_pair<0, NSLOTS, Cfg>(s_base, tmem_base, s_alo, bar_full, bar_empty, bar_peer_full, bar_acc, bar_act, bar_peer_act, it, n_act, a.err_flag, a.variant);
          issue_layer_pair<1, NSLOTS, Cfg>(s_base, tmem_base, s_alo, bar_full, bar_empty, bar_peer_full, bar_acc, bar_act, bar_peer_act, it, n_act, a.err_flag, a.variant);
          issue_layer_pair<2, NSLOTS, Cfg>(s_base, tmem_base, s_alo, bar_full, bar_empty, bar_peer_full, bar_acc, bar_act, bar_peer_act, it, n_act, a.err_flag, a.variant);
          issue_layer_pair<3, NSLOTS, Cfg>(s_base, tmem_base, s_alo, bar_full, bar_empty, bar_peer_full, bar_acc, bar_act, bar_peer_act, it, n_act, a.err_flag, a.variant);
          issue_layer_pair<4, NSLOTS, Cfg>(s_base, tmem_base, s_alo, bar_full, bar_empty, bar_peer_full, bar_acc, bar_act, bar_peer_act, it, n_act, a.err_flag, a.variant);
          issue_layer_pair<5, NSLOTS, Cfg>(s_base, tmem_base, s_alo, bar_full, bar_empty, bar_peer_full, bar_acc, bar_act, bar_peer_act, it, n_act, a.err_flag, a.variant);
        }
      }
    } else if (lane == 0) {
      // ---- peer: relay "stage landed" / "activations ready" to the leader's barriers
      const uint32_t r_peer_full = mapa_shared(bar_full, 0);
      const uint32_t r_peer_act = mapa_shared(bar_act, 0);
      for (long long unit = pair_id; unit < a.n_units; unit += n_pairs) {
        for (int g = 0; g < a.n_groups; ++g) {
          relay_layer<0, NSLOTS>(bar_full, bar_act, r_peer_full, r_peer_act, it, n_act, a.err_flag, (a.variant & 1) != 0);
          relay_layer<1, NSLOTS>(bar_full, bar_act, r_peer_full, r_peer_act, it, n_act, a.err_flag, (a.variant & 1) != 0);
          relay_layer<2, NSLOTS>(bar_full, bar_act, r_peer_full, r_peer_act, it, n_act, a.err_flag, (a.variant & 1) != 0);
          relay_layer<3, NSLOTS>(bar_full, bar_act, r_peer_full, r_peer_act, it, n_act, a.err_flag, (a.variant & 1) != 0);
          relay_layer<4, NSLOTS>(bar_full, bar_act, r_peer_full, r_peer_act, it, n_act, a.err_flag, (a.variant & 1) != 0);
          relay_layer<5, NSLOTS>(bar_full, bar_act, r_peer_full, r_peer_act, it, n_act, a.err_flag, (a.variant & 1) != 0);
        }
      }
    }
  } else {
    // ---- epilogue: 8 warps, this CTA's tile
    EpiCtx c;
    c.tl = 0;
    c.we = warp - 2;
    c.lane = lane;
    c.quarter = warp & 3;
    c.colhalf = c.we >> 2;
    const int row = c.quarter * 32 + lane;
    c.slot_w = row >> 6;
    c.t = row & 63;
    c.tid_t = c.we * 32 + lane;
    c.tab = reinterpret_cast<float*>(smem + Cfg::oTab);
    c.tab_w = c.tab + 1024;
    c.act_row = s_base + Cfg::oAct + (uint32_t)(row + 4) * 16u;
    c.taddr = tmem_base + ((uint32_t)(c.quarter * 32) << 16);
    c.zp = zpart;
    c.bar_acc = bar_acc;
    c.bar_act = bar_act;
    c.k0 = a.k0; c.k1 = a.k1;
    c.err_flag = a.err_flag;
    c.strong_sync = (a.variant & 1) != 0;
    uint32_t n_acc = 0;
    // per-unit prologue: which window this thread's row belongs to, its input row (bf16), the ones-plane row
    auto unit_prologue = [&](long long unit, long long& win, uint32_t& xin0, uint32_t& xin1) {
      win = unit * Cfg::kWpu + 2 * (long long)rank + c.slot_w;
      c.row_valid = (c.t < kL) && (win < a.n);
      xin0 = 0; xin1 = 0;
      if (c.colhalf == 0) {
        if (c.row_valid) {
          float f0, f1, f2, f3;
          if (a.x_f64) {
            const double2* p = reinterpret_cast<const double2*>(static_cast<const double*>(a.x) + ((size_t)win * kL + c.t) * kF);
            const double2 u = __ldg(p), v = __ldg(p + 1);
            f0 = (float)u.x; f1 = (float)u.y; f2 = (float)v.x; f3 = (float)v.y;
          } else {
            const float4 u = __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(a.x) + ((size_t)win * kL + c.t) * kF));
            f0 = u.x; f1 = u.y; f2 = u.z; f3 = u.w;
          }
          xin0 = pack_bf16x2(f0, f1);
          xin1 = pack_bf16x2(f2, f3);
        }
      } else {
        st_shared_v4(c.act_row + (uint32_t)kOnesPlane * kPlaneBytes, c.row_valid ? 0x3F803F80u : 0u, 0u, 0u, 0u);
      }
      c.c0w = (uint32_t)(a.window_base + win);
    };
    for (long long unit = pair_id; unit < a.n_units; unit += n_pairs) {
      long long win; uint32_t xin0, xin1;
      unit_prologue(unit, win, xin0, xin1);
      float* dbg_base = nullptr;
      if (DBG) {
        if (a.dbg_act && blockIdx.x == 0 && unit == pair_id) dbg_base = a.dbg_act + (size_t)row * 256;
      }
      for (int g = 0; g < a.n_groups; ++g) {
        epi_pass_setup<DROP>(c, a, g);
        if (c.colhalf == 0) st_shared_v4(c.act_row, xin0, xin1, 0u, 0u);
        fence_proxy_async();
        epi_signal_act<Cfg::kEpiThreads>(c);
        const float* tab_g = a.tables + (size_t)((a.mode_de || a.tab_per_group) ? g : 0) * kTableFloats;
        float* dbg_g = (DBG && g == 0) ? dbg_base : nullptr;
        // probs_s rows are indexed by (2*tl + slot_w) = slot_w here
        epi_layer<0, DROP, DBG, Cfg::kEpiThreads>(c, a, tab_g, g, win, n_acc, probs_s, dbg_g);
        epi_layer<1, DROP, DBG, Cfg::kEpiThreads>(c, a, tab_g, g, win, n_acc, probs_s, dbg_g);
        epi_layer<2, DROP, DBG, Cfg::kEpiThreads>(c, a, tab_g, g, win, n_acc, probs_s, dbg_g);
        epi_layer<3, DROP, DBG, Cfg::kEpiThreads>(c, a, tab_g, g, win, n_acc, probs_s, dbg_g);
        epi_layer<4, DROP, DBG, Cfg::kEpiThreads>(c, a, tab_g, g, win, n_acc, probs_s, dbg_g);
        epi_layer<5, DROP, DBG, Cfg::kEpiThreads>(c, a, tab_g, g, win, n_acc, probs_s, dbg_g);
      }
      if (a.smem_probs && c.colhalf == 0 && c.t == 0 && (c.quarter & 1) == 0 && win < a.n) {
        const float* pw = probs_s + c.slot_w * a.n_groups;
        window_metrics([&](int tt) { return pw[tt]; }, a.n_groups, win, a.out);
      }
    }
  }

  // ---- teardown: nobody may leave while the pair can still touch its shared memory / TMEM
  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)Cfg::kTmemCols)
                 : "memory");
  }
}

template <int NSLOTS, int DROP, bool DBG>
static int launch_pair_variant(const FusedParams& p, cudaStream_t s) {
  using Cfg = PairCfg<NSLOTS>;
  FusedArgs a{};
  a.x = p.x; a.x_f64 = p.x_dtype == UQK_F64; a.n = p.n; a.n_groups = p.n_groups; a.mode_de = p.mode_de; a.tab_per_group = p.tab_per_group; a.pass_base = p.pass_base;
  a.k0 = (uint32_t)(p.seed & 0xFFFFFFFFull); a.k1 = (uint32_t)(p.seed >> 32);
  a.window_base = p.window_base; a.mask_bits = p.mask_bits; a.packed = reinterpret_cast<const uint8_t*>(p.packed_pair);
  a.tables = p.tables; a.probs = p.probs; a.out = p.out; a.dbg_act = p.dbg_act; a.err_flag = p.err_flag;
  a.variant = p.variant;
  a.n_units = (p.n + Cfg::kWpu - 1) / Cfg::kWpu;
  a.smem_probs = p.no_smem_probs ? 0 : 1;
  a.n_layers = kBlocks;
  const size_t smem = (size_t)Cfg::oProbs + (p.no_smem_probs ? 0 : (size_t)2 * p.n_groups * sizeof(float));
  if (smem > 113 * 1024) return UQK_E_SHAPE;   // must leave room for two CTAs per SM
  int dev = 0, sms = 148;
  UQK_CUDA(cudaGetDevice(&dev));
  UQK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  auto kern = fused_pair_kernel<NSLOTS, DROP, DBG>;
  UQK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  UQK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  int per_sm = 1;
  UQK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, Cfg::kThreads, smem));
  if (per_sm < 1) { set_error("pair kernel does not fit on an SM (smem %zu)", smem); return UQK_E_CUDA; }
  // The occupancy API reports 1 block/SM for tcgen05 kernels on this driver although the hardware limits
  // (registers, shared memory, 2 x 256 TMEM columns) admit two -- ncu's launch__occupancy_limit_* agree with two.
  per_sm = (2 * (smem + 1024) <= 233472) ? 2 : 1;
  const long long pairs = std::min<long long>(a.n_units, (long long)(sms / 2) * per_sm);
  kern<<<(unsigned)(2 * pairs), Cfg::kThreads, smem, s>>>(a);
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}

template <int NSLOTS>
static int launch_pair(const FusedParams& p, cudaStream_t s) {
  const int drop = !p.dropout ? kDropNone : (p.mask_bits ? kDropBits : kDropPhilox);
  if (p.dbg_act) { set_error("the activation dump exists in the one-tile kernel only (variant 16)"); return UQK_E_BADARG; }
  switch (drop) {
    case kDropNone: return launch_pair_variant<NSLOTS, kDropNone, false>(p, s);
    case kDropPhilox: return launch_pair_variant<NSLOTS, kDropPhilox, false>(p, s);
    default: return launch_pair_variant<NSLOTS, kDropBits, false>(p, s);
  }
}

// =====================================================================================================================
// CTA-pair variant with TWO tiles per CTA ("P2"): 1 CTA per SM, the pair owns a unit of 8 windows = tile-pairs A and B
// (each an M=256 accumulator spanning both SMs).  Both tile-pairs consume the SAME weight stream (R = 8 window-passes
// per stream, half the L2->SM traffic of the one-tile pair kernel) through a ring of 4 slots of 16 KB per CTA (32 KB
// stages: the issuer's ~300 clk of wait / commit / loop latency per stage then covers 4-10 MMAs per tile-pair).  The
// leader CTA runs ONE MMA-ISSUER WARP PER TILE-PAIR.  A ring slot is free when BOTH tile-pairs have consumed it
// (mbarrier count 2, tcgen05.commit multicast to both CTAs), which keeps the two tile-pairs in lock-step (measured,
// tools/timeline_probe.py); the epilogues are taken off the critical path by the early hand-off instead (early_producer).
// =====================================================================================================================
template <int NSLOTS>
struct Pair2Cfg {
  static constexpr int kEpiThreads = 256;
  static constexpr int kCtrlWarps = 3;             // producer, issuer of tile-pair 0 (peer: relay), issuer of tile-pair 1
  static constexpr int kWarps = kCtrlWarps + 16;
  static constexpr int kThreads = 32 * kWarps;
  static constexpr int kWpu = 8;                   // windows per pair unit
  static constexpr int kTmemCols = 512;
  static constexpr int kSlotPair = kSlotBytesBig;  // 32 KB stages: 16 KB per CTA
  static constexpr int kSlot = kSlotPair / 2;
  static constexpr int oAct = 0;
  static constexpr int oRing = oAct + 2 * kActBytes;
  static constexpr int kTabFloats = 2 * 512 + 128;
  static constexpr int oTab = oRing + NSLOTS * kSlot;
  static constexpr int oBar = oTab + 2 * kTabFloats * 4;
  static constexpr int kNumBars = 2 * NSLOTS + 8;  // full, empty [NSLOTS]; acc[2], act[2], in[2], tfree[2] (the last two: sweeps)
  static constexpr int oMisc = oBar + 8 * kNumBars;            // tmem ptr (16 B)
  static constexpr int oAlo = oMisc + 16;                      // A-descriptor low words, one per (layer, step)
  static constexpr int oZ = oAlo + 4 * 400;                    // zpart[2][2][4] head partials (64 B); s_alo: 388 used, stages of up to 10 steps read a few past
  static constexpr int oProbs = (oZ + 64 + 15) & ~15;          // probs[4][n_groups]
  // statistics sweeps use neither zpart nor probs: their [2 x 256] u64 accumulators start at oZ
  static constexpr int oStats = oZ;
  static_assert(oZ % 8 == 0, "statistics accumulators are 64-bit");
};

// STAT = true: statistics sweep of the exact batch-statistic mode (see FusedArgs::n_layers, stat_epilogue_loop): the
// kernel computes layers [first_layer, n_layers) only; both tiles of a CTA still share every weight stage.
template <int NSLOTS, int DROP, bool DBG, bool STAT = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(Pair2Cfg<NSLOTS>::kThreads, 1)
fused_pair2_kernel(const FusedArgs a) {
  using Cfg = Pair2Cfg<NSLOTS>;
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t s_base = smem_u32(smem);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const long long pair_id = blockIdx.x >> 1;
  const long long n_pairs = gridDim.x >> 1;

  const uint32_t bar_full = s_base + Cfg::oBar;
  const uint32_t bar_empty = bar_full + 8 * NSLOTS;
  const uint32_t bar_acc = bar_empty + 8 * NSLOTS;              // [2]
  const uint32_t bar_act = bar_acc + 16;                        // [2] leader only: its own epilogue threads
  const uint32_t bar_in = bar_act + 16;                         // [2] statistics sweeps: the tile's input planes have landed
  const uint32_t bar_tfree = bar_in + 16;                       // [2] pipelined sweeps: accumulator read out (leader: + peer relay)
  const bool pipe = STAT && a.first_layer >= 2 && a.pipelined;  // pipelined single-stream sweep (stat_pipelined_loop)
  volatile uint32_t* tmem_ptr = reinterpret_cast<volatile uint32_t*>(smem + Cfg::oMisc);
  float* zpart = reinterpret_cast<float*>(smem + Cfg::oZ);
  uint32_t* s_alo = reinterpret_cast<uint32_t*>(smem + Cfg::oAlo);
  float* probs_s = reinterpret_cast<float*>(smem + Cfg::oProbs);

  for (int i = threadIdx.x; i < (2 * kActBytes) / 16; i += Cfg::kThreads)
    reinterpret_cast<uint4*>(smem + Cfg::oAct)[i] = make_uint4(0, 0, 0, 0);
  for (int i = threadIdx.x; i < kTotalSteps; i += Cfg::kThreads) {
    int l = 0, sidx = i;
    while (sidx >= c_layers[l].nsteps) { sidx -= c_layers[l].nsteps; ++l; }
    const FusedLayer fl = c_layers[l];
    uint32_t a_off, a_lbo;
    if (sidx == fl.nconv) { a_off = (uint32_t)kOnesPlane * kPlaneBytes + 4u * 16u; a_lbo = 16u; }
    else if (l == 0) { a_off = (uint32_t)(2 * sidx + 1) * 16u; a_lbo = 16u; }
    else {
      const int tap = sidx % fl.taps, slab = sidx / fl.taps;   // slab-major K order (see fused_pack_member)
      a_off = (uint32_t)(2 * slab) * kPlaneBytes + (uint32_t)(tap - fl.halo + 4) * 16u;
      a_lbo = kPlaneBytes;
    }
    s_alo[i] = (((s_base + Cfg::oAct + a_off) >> 4) & 0x3FFFu) | ((a_lbo >> 4) << 16);
  }
  if (threadIdx.x == 0) {
    // leader: every "full" / "act" barrier also counts one arrival relayed from the peer CTA; a ring slot is free when
    // BOTH tile-pairs have consumed it (two commits)
    const uint32_t extra = rank == 0 ? 1u : 0u;
    for (int i = 0; i < NSLOTS; ++i) {
      mbar_init(bar_full + 8 * i, 1 + extra); mbar_init(bar_empty + 8 * i, pipe ? 1 : 2);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_acc + 8 * i, 1); mbar_init(bar_act + 8 * i, Cfg::kEpiThreads / 32 + extra);
      mbar_init(bar_in + 8 * i, STAT ? 1 : Cfg::kEpiThreads / 32 + extra);   // sweeps: input planes landed; else: bar_half (early start)
      mbar_init(bar_tfree + 8 * i, Cfg::kEpiThreads / 32 + extra);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(const_cast<uint32_t*>(tmem_ptr))),
                 "r"((uint32_t)Cfg::kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  fence_proxy_async();
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const bool dbg_no_copy = (a.variant & 2) != 0;
  const bool dbg_no_mma = (a.variant & 8) != 0;

  if (warp == 0) {
    // ---- weight producer: this CTA's half of every stage
    if (lane == 0) {
      uint32_t it = 0;
      for (long long unit = pair_id; unit < a.n_units; unit += n_pairs) {
        for (int g = 0; g < a.n_groups; ++g) {
          const uint8_t* wmem = a.packed + (size_t)(a.mode_de ? g : 0) * kPackedBytes + (size_t)rank * (kPackedBytes / 2);
          const int l_end = STAT ? a.n_layers : kBlocks;
#pragma unroll 1
          for (int l = STAT ? a.first_layer : 0; l < l_end; ++l) {
            const FusedLayer fl = c_layers_big[l];
            const uint32_t sb = (uint32_t)fl.step_bytes / 2;
            int step = 0;
            for (int st = 0; st < fl.nstages; ++st, ++it) {
              const int slot = it % NSLOTS;
              const uint32_t ph = (it / NSLOTS) & 1;
              mbar_wait2(bar_empty + 8 * slot, ph ^ 1, a.err_flag, 1);
              const int nst = min(fl.sps, fl.nsteps - step);
              const uint32_t bytes = (uint32_t)nst * sb;
              if (dbg_no_copy) { mbar_arrive(bar_full + 8 * slot); }
              else {
                mbar_arrive_expect_tx(bar_full + 8 * slot, bytes);
                bulk_g2s(s_base + Cfg::oRing + slot * Cfg::kSlot, wmem + fl.w_off / 2 + (size_t)step * sb, bytes,
                         bar_full + 8 * slot);
              }
              step += nst;
            }
          }
        }
      }
    }
  } else if (warp == 1 || warp == 2) {
    const int tl = warp - 1;
    if (pipe) {
      if (rank == 0 && warp == 1) {
        // ---- leader, pipelined sweep: ONE issuer walks the units, alternating between the two buffers / accumulators
        const long long my_units = pair_id < a.n_units ? (a.n_units - pair_id + n_pairs - 1) / n_pairs : 0;
        const long long total_it = my_units * a.n_groups;
        uint32_t it = 0, n_act0 = 0u, n_act1 = 0u, n_tf0 = 0u, n_tf1 = 0u;
        for (long long q = 0; q < total_it; ++q) {
          const int b = (int)(q & 1);
          if (q >= 2) {
            mbar_wait_cluster(bar_tfree + 8 * b, (b ? n_tf1 : n_tf0) & 1, a.err_flag, 6, (a.variant & 1) != 0);
            if (b) ++n_tf1; else ++n_tf0;
          }
          const uint32_t tm = tmem_base + (uint32_t)b * 256u, a_tile = (uint32_t)b * (uint32_t)(kActBytes >> 4);
          uint32_t na = b ? n_act1 : n_act0;
          switch (a.first_layer) {
            case 2: issue_layer_pair<2, NSLOTS, Cfg>(s_base, tm, s_alo, bar_full, bar_empty, 0u, bar_acc + 8 * b, bar_act + 8 * b, 0u, it, na, a.err_flag, a.variant, a_tile); break;
            case 3: issue_layer_pair<3, NSLOTS, Cfg>(s_base, tm, s_alo, bar_full, bar_empty, 0u, bar_acc + 8 * b, bar_act + 8 * b, 0u, it, na, a.err_flag, a.variant, a_tile); break;
            case 4: issue_layer_pair<4, NSLOTS, Cfg>(s_base, tm, s_alo, bar_full, bar_empty, 0u, bar_acc + 8 * b, bar_act + 8 * b, 0u, it, na, a.err_flag, a.variant, a_tile); break;
            default: issue_layer_pair<5, NSLOTS, Cfg>(s_base, tm, s_alo, bar_full, bar_empty, 0u, bar_acc + 8 * b, bar_act + 8 * b, 0u, it, na, a.err_flag, a.variant, a_tile); break;
          }
          if (b) n_act1 = na; else n_act0 = na;
        }
      } else if (rank != 0 && warp == 1 && lane == 0) {
        const uint32_t r_full = mapa_shared(bar_full, 0);
        uint32_t it = 0;
        const long long my_units = pair_id < a.n_units ? (a.n_units - pair_id + n_pairs - 1) / n_pairs : 0;
        const int ns = c_layers_big[a.first_layer].nstages;
        for (long long q = 0; q < my_units * a.n_groups; ++q)
          for (int st = 0; st < ns; ++st, ++it) {
            const int slot = it % NSLOTS;
            mbar_wait2(bar_full + 8 * slot, (it / NSLOTS) & 1, a.err_flag, 23);
            mbar_arrive_remote(r_full + 8 * slot, (a.variant & 1) != 0);
          }
      }
    } else if (rank == 0) {
      // ---- leader: one MMA-issuer warp per tile-pair.  Both walk the same weight stream; they drift apart at layer
      //      boundaries (one waits for its epilogue while the other keeps the tensor pipe busy), bounded by the ring.
      uint32_t it = 0, n_act = 0, n_half = 0;
      const uint32_t b_half = bar_in + 8 * tl;
      const uint32_t tm = tmem_base + (uint32_t)tl * 256u;
      const uint32_t a_tile = (uint32_t)tl * (uint32_t)(kActBytes >> 4);
      const uint32_t b_acc = bar_acc + 8 * tl, b_act = bar_act + 8 * tl;
      for (long long unit = pair_id; unit < a.n_units; unit += n_pairs) {
        for (int g = 0; g < a.n_groups; ++g) {
#ifdef UQK_TIMELINE
          if (!STAT) {
            long long* tb = (a.tl_buf && blockIdx.x == 0 && unit == pair_id + (long long)((a.variant >> 12) & 15) * n_pairs && g < 16 && lane == 0) ? a.tl_buf + ((tl * 16 + g) * 6 << 3) : nullptr;
            issue_layer_pair<0, NSLOTS, Cfg, false, true>(s_base, tm, s_alo, bar_full, bar_empty, 0u, b_acc, b_act, 0u, it, n_act, a.err_flag, a.variant, a_tile, tb, b_half, &n_half);
            issue_layer_pair<1, NSLOTS, Cfg, false, true>(s_base, tm, s_alo, bar_full, bar_empty, 0u, b_acc, b_act, 0u, it, n_act, a.err_flag, a.variant, a_tile, tb ? tb + 8 : tb, b_half, &n_half);
            issue_layer_pair<2, NSLOTS, Cfg, false, true>(s_base, tm, s_alo, bar_full, bar_empty, 0u, b_acc, b_act, 0u, it, n_act, a.err_flag, a.variant, a_tile, tb ? tb + 16 : tb, b_half, &n_half);
            issue_layer_pair<3, NSLOTS, Cfg, false, true>(s_base, tm, s_alo, bar_full, bar_empty, 0u, b_acc, b_act, 0u, it, n_act, a.err_flag, a.variant, a_tile, tb ? tb + 24 : tb, b_half, &n_half);
            issue_layer_pair<4, NSLOTS, Cfg, false, true>(s_base, tm, s_alo, bar_full, bar_empty, 0u, b_acc, b_act, 0u, it, n_act, a.err_flag, a.variant, a_tile, tb ? tb + 32 : tb, b_half, &n_half);
            issue_layer_pair<5, NSLOTS, Cfg, false, true>(s_base, tm, s_alo, bar_full, bar_empty, 0u, b_acc, b_act, 0u, it, n_act, a.err_flag, a.variant, a_tile, tb ? tb + 40 : tb, b_half, &n_half);
            continue;
          }
#endif
          if (!STAT || a.first_layer <= 0) issue_layer_pair<0, NSLOTS, Cfg, false, !STAT>(s_base, tm, s_alo, bar_full, bar_empty, 0u, b_acc, b_act, 0u, it, n_act, a.err_flag, a.variant, a_tile, nullptr, b_half, &n_half);
          if (!STAT || (a.n_layers > 1 && a.first_layer <= 1)) issue_layer_pair<1, NSLOTS, Cfg, false, !STAT>(s_base, tm, s_alo, bar_full, bar_empty, 0u, b_acc, b_act, 0u, it, n_act, a.err_flag, a.variant, a_tile, nullptr, b_half, &n_half);
          if (!STAT || (a.n_layers > 2 && a.first_layer <= 2)) issue_layer_pair<2, NSLOTS, Cfg, false, !STAT>(s_base, tm, s_alo, bar_full, bar_empty, 0u, b_acc, b_act, 0u, it, n_act, a.err_flag, a.variant, a_tile, nullptr, b_half, &n_half);
          if (!STAT || (a.n_layers > 3 && a.first_layer <= 3)) issue_layer_pair<3, NSLOTS, Cfg, false, !STAT>(s_base, tm, s_alo, bar_full, bar_empty, 0u, b_acc, b_act, 0u, it, n_act, a.err_flag, a.variant, a_tile, nullptr, b_half, &n_half);
          if (!STAT || (a.n_layers > 4 && a.first_layer <= 4)) issue_layer_pair<4, NSLOTS, Cfg, false, !STAT>(s_base, tm, s_alo, bar_full, bar_empty, 0u, b_acc, b_act, 0u, it, n_act, a.err_flag, a.variant, a_tile, nullptr, b_half, &n_half);
          if (!STAT || (a.n_layers > 5 && a.first_layer <= 5)) issue_layer_pair<5, NSLOTS, Cfg, false, !STAT>(s_base, tm, s_alo, bar_full, bar_empty, 0u, b_acc, b_act, 0u, it, n_act, a.err_flag, a.variant, a_tile, nullptr, b_half, &n_half);
        }
      }
    } else if (warp == 1 && lane == 0) {
      // ---- peer: relay "my half of stage q has landed" to the leader's full barriers, in stream order
      const uint32_t r_full = mapa_shared(bar_full, 0);
      uint32_t it = 0;
      for (long long unit = pair_id; unit < a.n_units; unit += n_pairs)
        for (int g = 0; g < a.n_groups; ++g)
#pragma unroll 1
          for (int l = STAT ? a.first_layer : 0; l < (STAT ? a.n_layers : kBlocks); ++l) {
            const int ns = c_layers_big[l].nstages;
            for (int st = 0; st < ns; ++st, ++it) {
              const int slot = it % NSLOTS;
              mbar_wait2(bar_full + 8 * slot, (it / NSLOTS) & 1, a.err_flag, 23);
              mbar_arrive_remote(r_full + 8 * slot, (a.variant & 1) != 0);
            }
          }
    }
  } else {
    // ---- epilogue: 8 warps per tile
    const int ew = warp - Cfg::kCtrlWarps;
    EpiCtx c;
    c.tl = ew >> 3;
    c.we = ew & 7;
    c.lane = lane;
    c.quarter = warp & 3;
    c.colhalf = c.we >> 2;
    const int row = c.quarter * 32 + lane;
    c.slot_w = row >> 6;
    c.t = row & 63;
    c.tid_t = c.we * 32 + lane;
    c.tab = reinterpret_cast<float*>(smem + Cfg::oTab) + c.tl * Cfg::kTabFloats;
    c.tab_w = c.tab + 1024;
    c.act_row = s_base + Cfg::oAct + c.tl * kActBytes + (uint32_t)(row + 4) * 16u;
    c.taddr = tmem_base + ((uint32_t)(c.quarter * 32) << 16) + (uint32_t)(c.tl * 256);
    c.zp = zpart + c.tl * 8;
    c.bar_acc = bar_acc + 8 * c.tl;
    c.bar_act = bar_act + 8 * c.tl;
    c.remote_act = rank == 0 ? 0u : mapa_shared(bar_act + 8 * c.tl, 0);
    c.bar_half = bar_in + 8 * c.tl;
    c.remote_half = rank == 0 ? 0u : mapa_shared(bar_in + 8 * c.tl, 0);
    c.k0 = a.k0; c.k1 = a.k1;
    c.err_flag = a.err_flag;
    c.strong_sync = (a.variant & 1) != 0;
    uint32_t n_acc = 0;
    if (STAT && pipe) {
      PipeGeo pg;
      pg.act_base = s_base + Cfg::oAct; pg.bar_in = bar_in; pg.bar_acc = bar_acc; pg.bar_act = bar_act; pg.bar_tfree = bar_tfree;
      pg.tmem_base = tmem_base; pg.rank = rank; pg.pair_id = pair_id; pg.n_pairs = n_pairs;
      stat_pipelined_loop<DROP, Cfg::kEpiThreads>(c, a, pg, reinterpret_cast<unsigned long long*>(smem + Cfg::oStats));
    } else if constexpr (STAT) {
      StatGeo geo;
      geo.act_base = s_base + Cfg::oAct + c.tl * kActBytes;
      geo.bar_in = bar_in + 8 * c.tl;
      geo.rank = rank; geo.pair_id = pair_id; geo.n_pairs = n_pairs; geo.all_threads = 2 * Cfg::kEpiThreads;
      stat_epilogue_loop<DROP, Cfg::kEpiThreads, Cfg::kWpu>(c, a, geo, reinterpret_cast<unsigned long long*>(smem + Cfg::oStats), n_acc);
    } else
    for (long long unit = pair_id; unit < a.n_units; unit += n_pairs) {
      // unit = 8 windows: tile-pair tl covers windows 4*tl .. 4*tl+3; this CTA (rank) takes 2 of them
      const long long win = unit * Cfg::kWpu + 4 * c.tl + 2 * (long long)rank + c.slot_w;
      c.row_valid = (c.t < kL) && (win < a.n);
      uint32_t xin0 = 0, xin1 = 0;
      if (c.colhalf == 0) {
        if (c.row_valid) {
          float f0, f1, f2, f3;
          if (a.x_f64) {
            const double2* p = reinterpret_cast<const double2*>(static_cast<const double*>(a.x) + ((size_t)win * kL + c.t) * kF);
            const double2 u = __ldg(p), v = __ldg(p + 1);
            f0 = (float)u.x; f1 = (float)u.y; f2 = (float)v.x; f3 = (float)v.y;
          } else {
            const float4 u = __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(a.x) + ((size_t)win * kL + c.t) * kF));
            f0 = u.x; f1 = u.y; f2 = u.z; f3 = u.w;
          }
          xin0 = pack_bf16x2(f0, f1);
          xin1 = pack_bf16x2(f2, f3);
        }
      } else {
        st_shared_v4(c.act_row + (uint32_t)kOnesPlane * kPlaneBytes, c.row_valid ? 0x3F803F80u : 0u, 0u, 0u, 0u);
      }
      c.c0w = (uint32_t)(a.window_base + win);
      c.xin0 = xin0; c.xin1 = xin1;
#ifdef UQK_TIMELINE
      c.tlb = (a.tl_buf && blockIdx.x == 0 && unit == pair_id + (long long)((a.variant >> 12) & 15) * n_pairs && lane == 0 && (c.we == 0 || c.we == 7)) ? a.tl_buf : nullptr;
#endif
      float* dbg_base = nullptr;
      if (DBG) {
        if (a.dbg_act && blockIdx.x == 0 && unit == pair_id && c.tl == 0) dbg_base = a.dbg_act + (size_t)row * 256;
      }
      for (int g = 0; g < a.n_groups; ++g) {
        if (g == 0) epi_pass_setup<DROP>(c, a, g); else c.tie = c.tie_next;
        if (g == 0) {   // later passes: staged by the previous pass's last epilogue (epi_layer, EARLY)
          if (c.colhalf == 0) st_shared_v4(c.act_row, xin0, xin1, 0u, 0u);
          fence_proxy_async();
          epi_signal_act<Cfg::kEpiThreads>(c);
        }
        if (c.we == 0) UQK_TLE(c, g, 0, 7);
        const float* tab_g = a.tables + (size_t)((a.mode_de || a.tab_per_group) ? g : 0) * kTableFloats;
        float* dbg_g = (DBG && g == 0) ? dbg_base : nullptr;
        epi_layer<0, DROP, DBG, Cfg::kEpiThreads, true>(c, a, tab_g, g, win, n_acc, probs_s, dbg_g);
        epi_layer<1, DROP, DBG, Cfg::kEpiThreads, true>(c, a, tab_g, g, win, n_acc, probs_s, dbg_g);
        epi_layer<2, DROP, DBG, Cfg::kEpiThreads, true>(c, a, tab_g, g, win, n_acc, probs_s, dbg_g);
        epi_layer<3, DROP, DBG, Cfg::kEpiThreads, true>(c, a, tab_g, g, win, n_acc, probs_s, dbg_g);
        epi_layer<4, DROP, DBG, Cfg::kEpiThreads, true>(c, a, tab_g, g, win, n_acc, probs_s, dbg_g);
        epi_layer<5, DROP, DBG, Cfg::kEpiThreads, true>(c, a, tab_g, g, win, n_acc, probs_s, dbg_g);
      }
      if (a.smem_probs && c.colhalf == 0 && c.t == 0 && (c.quarter & 1) == 0 && win < a.n) {
        const float* pw = probs_s + (2 * c.tl + c.slot_w) * a.n_groups;
        window_metrics([&](int tt) { return pw[tt]; }, a.n_groups, win, a.out);
      }
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)Cfg::kTmemCols)
                 : "memory");
  }
}

template <int NSLOTS, int DROP, bool DBG, bool STAT = false>
static int launch_pair2_variant(const FusedParams& p, cudaStream_t s) {
  using Cfg = Pair2Cfg<NSLOTS>;
  FusedArgs a{};
  a.x = p.x; a.x_f64 = p.x_dtype == UQK_F64; a.n = p.n; a.n_groups = p.n_groups; a.mode_de = p.mode_de; a.tab_per_group = p.tab_per_group; a.pass_base = p.pass_base;
  a.k0 = (uint32_t)(p.seed & 0xFFFFFFFFull); a.k1 = (uint32_t)(p.seed >> 32);
  a.window_base = p.window_base; a.mask_bits = p.mask_bits; a.packed = reinterpret_cast<const uint8_t*>(p.packed_pair);
  a.tables = p.tables; a.probs = p.probs; a.out = p.out; a.dbg_act = p.dbg_act; a.err_flag = p.err_flag;
  a.variant = p.variant;
#ifdef UQK_TIMELINE
  a.tl_buf = STAT ? nullptr : g_uqk_timeline;
#endif
  a.n_layers = STAT ? p.n_layers : kBlocks; a.stats = STAT ? p.stats : nullptr; a.first_layer = STAT ? p.first_layer : 0;
  // option exact_pipelined: resumed sweeps run pipelined over 4-window units (stat_pipelined_loop); default: 8-window units
  a.pipelined = (STAT && p.pipelined && a.first_layer >= 2) ? 1 : 0;
  const int wpu = a.pipelined ? 4 : Cfg::kWpu;
  a.n_units = (p.n + wpu - 1) / wpu;
  a.smem_probs = (p.no_smem_probs || STAT) ? 0 : 1;
  a.head_sum = STAT ? p.head_sum : nullptr; a.head_cnt = STAT ? p.head_cnt : nullptr;
  a.ckpt_out = STAT ? reinterpret_cast<uint4*>(p.ckpt_out) : nullptr;
  a.ckpt_in = STAT ? reinterpret_cast<const uint4*>(p.ckpt_in) : nullptr;
  if (STAT) {
    // a sweep computes one layer: [0,1) or [0,2) from the windows, [k,k+1) for k >= 2 from layer k-1's stored output
    const bool from_x = p.first_layer == 0 && (p.n_layers == 1 || p.n_layers == 2) && p.ckpt_in == nullptr;
    const bool resumed = p.first_layer >= 2 && p.first_layer < kBlocks && p.n_layers == p.first_layer + 1 && p.ckpt_in != nullptr;
    if (!(from_x || resumed) || p.stats == nullptr) {
      set_error("statistics sweep: unsupported layer range [%d, %d)", p.first_layer, p.n_layers); return UQK_E_BADARG;
    }
  }
  // the probability buffer of the fused reductions doubles as the statistics accumulator [2 x 256] u64 of a sweep
  const size_t smem = STAT ? (size_t)Cfg::oStats + (size_t)2 * kMaxCh * sizeof(unsigned long long)
                           : (size_t)Cfg::oProbs + (p.no_smem_probs ? 0 : (size_t)4 * p.n_groups * sizeof(float));
  if (smem > 227 * 1024) return UQK_E_SHAPE;
  int dev = 0, sms = 148;
  UQK_CUDA(cudaGetDevice(&dev));
  UQK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  auto kern = fused_pair2_kernel<NSLOTS, DROP, DBG, STAT>;
  UQK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  UQK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  const long long pairs = std::min<long long>(a.n_units, (long long)(sms / 2));
  kern<<<(unsigned)(2 * pairs), Cfg::kThreads, smem, s>>>(a);
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}

// statistics sweep of the exact batch-statistic mode (training=True: dropout is always on)
template <int NSLOTS>
static int launch_pair2_stat(const FusedParams& p, cudaStream_t s) {
  const int drop = !p.dropout ? kDropNone : (p.mask_bits ? kDropBits : kDropPhilox);
  if (drop == kDropNone) { set_error("statistics sweeps exist for the dropout (training=True) passes only"); return UQK_E_BADARG; }
  return drop == kDropPhilox ? launch_pair2_variant<NSLOTS, kDropPhilox, false, true>(p, s)
                             : launch_pair2_variant<NSLOTS, kDropBits, false, true>(p, s);
}

template <int NSLOTS>
static int launch_pair2(const FusedParams& p, cudaStream_t s) {
  const int drop = !p.dropout ? kDropNone : (p.mask_bits ? kDropBits : kDropPhilox);
  if (p.dbg_act) { set_error("the activation dump exists in the one-tile kernel only (variant 16)"); return UQK_E_BADARG; }
  switch (drop) {
    case kDropNone: return launch_pair2_variant<NSLOTS, kDropNone, false>(p, s);
    case kDropPhilox: return launch_pair2_variant<NSLOTS, kDropPhilox, false>(p, s);
    default: return launch_pair2_variant<NSLOTS, kDropBits, false>(p, s);
  }
}

// ---- host launcher ----------------------------------------------------------------------------------------------
template <int TILES, int NSLOTS, int DROP, bool DBG>
static int launch_variant(const FusedParams& p, cudaStream_t s) {
  using Cfg = FusedCfg<TILES, NSLOTS>;
  FusedArgs a{};
  a.x = p.x; a.x_f64 = p.x_dtype == UQK_F64; a.n = p.n; a.n_groups = p.n_groups; a.mode_de = p.mode_de; a.tab_per_group = p.tab_per_group; a.pass_base = p.pass_base;
  a.k0 = (uint32_t)(p.seed & 0xFFFFFFFFull); a.k1 = (uint32_t)(p.seed >> 32);
  a.window_base = p.window_base; a.mask_bits = p.mask_bits; a.packed = reinterpret_cast<const uint8_t*>(p.packed);
  a.tables = p.tables; a.probs = p.probs; a.out = p.out; a.dbg_act = p.dbg_act; a.err_flag = p.err_flag;
  a.variant = p.variant;
  a.n_units = (p.n + Cfg::kWpu - 1) / Cfg::kWpu;
  a.smem_probs = p.no_smem_probs ? 0 : 1;
  const size_t smem = (size_t)Cfg::oProbs + (p.no_smem_probs ? 0 : (size_t)Cfg::kWpu * p.n_groups * sizeof(float));
  if (smem > 227 * 1024) return UQK_E_SHAPE;   // caller falls back to the smaller configuration
  int dev = 0, sms = 148;
  UQK_CUDA(cudaGetDevice(&dev));
  UQK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  auto kern = fused_forward_kernel<TILES, NSLOTS, DROP, DBG>;
  UQK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  UQK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  int per_sm = 1;
  UQK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, Cfg::kThreads, smem));
  if (per_sm < 1) { set_error("fused kernel does not fit on an SM (smem %zu)", smem); return UQK_E_CUDA; }
  if (per_sm * Cfg::kTmemCols > 512) per_sm = 512 / Cfg::kTmemCols;   // TMEM: 512 columns per SM
  const long long grid = std::min<long long>(a.n_units, (long long)sms * per_sm);
  kern<<<(unsigned)grid, Cfg::kThreads, smem, s>>>(a);
  UQK_LAUNCH_CHECK();
  return UQK_OK;
}

template <int TILES, int NSLOTS>
static int launch_drop(const FusedParams& p, cudaStream_t s) {
  const int drop = !p.dropout ? kDropNone : (p.mask_bits ? kDropBits : kDropPhilox);
  if (p.dbg_act) {
    if constexpr (TILES == 1) {
      switch (drop) {
        case kDropNone: return launch_variant<TILES, NSLOTS, kDropNone, true>(p, s);
        case kDropPhilox: return launch_variant<TILES, NSLOTS, kDropPhilox, true>(p, s);
        default: return launch_variant<TILES, NSLOTS, kDropBits, true>(p, s);
      }
    } else {
      set_error("the activation dump exists in the one-tile kernel only (variant 16)");
      return UQK_E_BADARG;
    }
  }
  switch (drop) {
    case kDropNone: return launch_variant<TILES, NSLOTS, kDropNone, false>(p, s);
    case kDropPhilox: return launch_variant<TILES, NSLOTS, kDropPhilox, false>(p, s);
    default: return launch_variant<TILES, NSLOTS, kDropBits, false>(p, s);
  }
}

}  // namespace uqk
