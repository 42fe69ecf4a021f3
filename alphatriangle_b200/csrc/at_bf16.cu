// tcgen05 / TMEM bf16 path (sm_100a): convolution tower + trunk FC as implicit GEMMs.
//
// tower_kernel  persistent, one CTA per SM, two boards in flight per CTA. A board's
//               activations never leave shared memory between layers: they live as a
//               zero-haloed 16-wide board image, [channel/8][17 zero + 128 rows][8 ch] bf16 (+17 zero), so the
//               nine taps of a 3x3 convolution are nine ROW-SHIFTED views of one image and
//               each tap/channel-block is one tcgen05.mma (M=128 padded positions, N=Cout,
//               K=16) whose A descriptor merely starts 16*(dy*16+dx) bytes further on.
//               Weights (BatchNorm folded, bf16, pre-tiled as core matrices) stream from L2
//               through a bulk-copy (TMA engine) ring and are shared by both boards;
//               accumulators sit in TMEM (2 x 128 columns); eight epilogue warps apply
//               bias / residual / activation and write the next layer's image.
//               Layer 0 (Cin = 1, K = 9) is CUDA-core work in the prologue, straight from
//               the PackedState bitmasks -- the occupancy grid is never materialised.
// fc_kernel     trunk FC [B, H*W*C] x [H*W*C, N] as a split-K tcgen05 GEMM over the tower's
//               output (written by the tower directly as A-operand tiles), fp32 partials.
//
// Reference semantics: alphatriangle/nn/model.py:15-60 (conv_block, ResidualBlock),
// :232-240 (shared_fc), :268-297 (forward). Parity contract for this path: policy
// KL <= 1e-3 against the fp32 reference (north_star).

#include <cuda_bf16.h>

#include <cstdint>
#include <cstdlib>
#include <mutex>
#include <type_traits>
#include <vector>

#include "at_bf16.cuh"
#include "at_fp32.cuh"
#include "at_prepjobs.cuh"
#include "at_umma.cuh"

namespace at {

namespace {

using namespace umma;

// ---- geometry of the on-chip board image --------------------------------------------
constexpr int kPadW = 16;                    // board row stride: cols + at least one shared zero column
constexpr int kM = 128;                      // MMA rows = padded positions y*16 + x
constexpr int kHalo = 17;                    // zero rows before / after: max |dy*16 + dx|
constexpr int kPlaneRows = kM + kHalo;       // 145: the trailing halo of one 8-channel plane is the leading halo of the next
constexpr int kPlaneBytes = kPlaneRows * 16; // 2,320
constexpr int kMaxCh = 128;
constexpr int kBufBytes = (kMaxCh / 8) * kPlaneBytes + kHalo * 16;  // 37,392: 16 planes + the last plane's trailing halo
constexpr int kStageBytes = 16384;
// AT_TOWER_OVERLAP = 1: the next pair's first image is produced before the current pair's last accumulator is drained (image
// buffers and accumulators swap roles from pair to pair), so the tensor pipe restarts ~10k cycles earlier per pair.
#ifndef AT_TOWER_OVERLAP
#define AT_TOWER_OVERLAP 1
#endif
#ifndef AT_TOWER_STAGES
#define AT_TOWER_STAGES 4
#endif
constexpr int kStages = AT_TOWER_STAGES;  // weight ring depth of tower_kernel
constexpr int kMaxLayers = 12;
constexpr int kMaxSlabs = 36;                // weight slabs per layer (9 taps x 128 channels / 64 = 18 at the defaults)
constexpr int kBoards = 2;                   // boards in flight per CTA
constexpr int kEpiSplit = 2;                 // epilogue warps per (board, TMEM lane quarter): each takes 1/kEpiSplit of the channels
constexpr int kEpiWarps = kBoards * 4 * kEpiSplit;
constexpr int kMmaWarpB = 2 + kEpiWarps;      // MMA-issuing warp of board 1; warp 1 issues for board 0
constexpr int kTowerThreads = 64 + kEpiWarps * 32 + 32;
constexpr int kTowerTmemCols = AT_TOWER_OVERLAP ? 512 : kBoards * 128;  // per board one accumulator, or two used by alternate pairs
#ifndef AT_TOWER_PAIR_DRAIN
#define AT_TOWER_PAIR_DRAIN 1
#endif
constexpr bool kPairDrain = AT_TOWER_PAIR_DRAIN != 0 && AT_TOWER_OVERLAP != 0;  // last layer: 32 channels of both boards per thread
constexpr int kPairQueue = 4;               // entries of the shared-memory queue of claimed board pairs
constexpr int kMaxBars = 32;                 // mbarrier slots reserved in shared memory
constexpr long long kTimeoutCycles = 2000000000LL;  // ~1 s: a wedged pipeline reports instead of hanging

// Timing experiments (tools/tower_time.py, tools/tf_time.py) switch parts of the kernels off through TowerParams::debug.
// They exist only in the experiments build (make experiments -> libat_b200_exp.so, -DAT_EXPERIMENTS); in the product
// library the tests below are compile-time false and the environment is never read.
#ifdef AT_EXPERIMENTS
#define AT_DBG(bits) ((p.debug & (bits)) != 0)
// Timeline of one board pair in CTA 0 (experiments build only): every warp's lane 0 stamps clock64() at named points into
// g_timeline[warp][event]; tools/tower_timeline.py prints them relative to the pair's start (at_debug_read_timeline).
constexpr int kTlEvents = 256;
__device__ unsigned long long g_timeline[32 * kTlEvents];
#define AT_TL_DECL int tl_i = 0; bool tl_on = false;
#define AT_TL_PAIR(round) tl_on = (blockIdx.x == 0 && (round) == 3u); tl_i = 0;
#define AT_TL(tag)                                                                                              \
  do {                                                                                                          \
    if (tl_on && (threadIdx.x & 31) == 0 && tl_i < kTlEvents) {                                                 \
      g_timeline[(threadIdx.x >> 5) * kTlEvents + tl_i++] = ((unsigned long long)(tag) << 48) | (clock64() & 0xFFFFFFFFFFFFull); \
    }                                                                                                           \
  } while (0)
#else
#define AT_DBG(bits) false
#define AT_TL_DECL
#define AT_TL_PAIR(round)
#define AT_TL(tag)
#endif

struct TowerLayer {
  int cin, cout, taps, ks, n_slabs;
  int src, dst, res;  // buffer ids 0/1; dst = 2 -> global FC operand; res = -1 -> none
  int act;
  int bias_slot;
  uint32_t slab_bytes;
  // where each weight slab's first MMA reads its A operand, relative to the source image, >> 4: the slab covers channels
  // [c0, c0 + ks) of one tap, i.e. plane c0/8 of the image shifted by (dy*16 + dx) rows
  uint32_t aoff[kMaxSlabs];
};

struct TowerParams {
  const PackedState* states;
  int64_t batch;
  const uint8_t* wstream;  // all MMA layers' weight slabs, in execution order
  const float* bias;       // [n_layers][128] fp32
  const float* w0;         // layer 0: [taps][c0] fp32 (folded), taps = 9
  const float* b0;         // [c0]
  int c0, act0;
  int n_layers;
  TowerLayer layers[kMaxLayers];
  uint8_t* fc_a;           // output: [ceil(B/128)][n_kc][128][8] bf16
  int n_kc;                // L * C_last / 8
  int rows, cols;
  uint32_t* status;
  int* pair_counter;  // null: board pairs by blockIdx stride; else handed out by this counter (zeroed before the launch)
  uint32_t debug;  // experiments build only (AT_TOWER_DEBUG): 1 no weight copies, 2 no MMAs, 4 no epilogue math, 8 no layer 0, 16 no weight ring
};

constexpr size_t tower_smem_bytes() {
  return (size_t)kStages * kStageBytes + (size_t)kBoards * 2 * kBufBytes + (size_t)kMaxLayers * 128 * 4 + 9 * 128 * 4 + 128 * 4 +
         kMaxBars * 8 + 16 + kPairQueue * 8;
}

__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
  return pred != 0;
}

// Wait for an mbarrier phase. A wedged pipeline reports through `status` instead of hanging the GPU (after ~1 s).
// Mode 0 (shipped): plain try_wait in a loop. A waiting warp then re-issues ~9 instructions every ~65 cycles -- two thirds of
// all warp instructions tower_kernel executes are this loop (ncu source view, profiles/r02_wait_loop.txt) -- yet parking the
// warp instead (mode 1: try_wait with a suspend-time hint, no clock reads) was measured 1-2 % SLOWER on both tower kernels:
// what matters is how fast the waiter wakes, not the issue slots it burns meanwhile.
#ifndef AT_WAIT_MODE
#define AT_WAIT_MODE 0
#endif
constexpr uint32_t kWaitHintNs = 1u << 20;
constexpr uint32_t kWaitTries = 1u << 20;  // the hardware caps a single suspension well below the hint: bound by count, not by clock
__device__ __forceinline__ bool wait_bar(uint32_t bar, uint32_t parity, uint32_t* status, uint32_t code) {
#if AT_WAIT_MODE == 0
  if (mbar_try_wait(bar, parity)) return true;  // common case: no clock reads on the fast path
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > kTimeoutCycles) {
      atomicOr(status, code);
      return false;
    }
  }
  return true;
#else
  if (mbar_try_wait_hint(bar, parity, kWaitHintNs)) return true;
  for (uint32_t tries = 0; tries < kWaitTries; ++tries)
    if (mbar_try_wait_hint(bar, parity, kWaitHintNs)) return true;
  atomicOr(status, code);
  return false;
#endif
}

// ---- which board pair a CTA works on next ------------------------------------------------------------------------------
// Both tower kernels are persistent, one CTA per SM. With pairs assigned by blockIdx stride every CTA must get its SM at
// once: if anything else holds a few SMs (the NCCL kernel of a weight publication waiting for a slow rank), the CTAs left
// over run as a second wave and the launch takes twice as long. So pairs are handed out by an atomic counter instead: a CTA
// that starts late simply takes fewer. The weight producer, which runs ahead of everyone, claims a CTA's pairs one round
// in advance and publishes them in a 4-entry shared-memory queue (round number in the high word, pair index or -1 = "no
// more" in the low word); issuers and epilogue warps read entry `it` at their round `it` (the epilogue also entry it + 1,
// for the hand-over).
__device__ __forceinline__ void pairq_put(volatile unsigned long long* q, uint32_t it, int64_t pair) {
  q[it & (kPairQueue - 1)] = ((unsigned long long)it << 32) | (uint32_t)(int32_t)pair;
}
__device__ __forceinline__ bool pairq_get(const volatile unsigned long long* q, uint32_t it, int64_t* pair, uint32_t* status) {
  long long t0 = 0;
  for (;;) {
    const unsigned long long v = q[it & (kPairQueue - 1)];
    if ((uint32_t)(v >> 32) == it) {
      *pair = (int64_t)(int32_t)(uint32_t)v;
      return true;
    }
    if (t0 == 0) t0 = clock64();
    else if (clock64() - t0 > kTimeoutCycles) {
      atomicOr(status, 128u);
      return false;
    }
  }
}
// the producer's side: claim the next pair (or compute it, without a counter). The atomic's round trip to L2 is ~1 us, so
// the claim is issued where the round starts and its result only used (pair_resolve, pairq_put) a few slabs later.
__device__ __forceinline__ int64_t pair_claim_raw(int* counter, uint32_t it) {
  return counter ? (int64_t)atomicAdd(counter, 1) : (int64_t)blockIdx.x + (int64_t)it * gridDim.x;
}
__device__ __forceinline__ int64_t pair_resolve(int64_t raw, int64_t n_pairs) { return raw < n_pairs ? raw : -1; }
__device__ __forceinline__ int64_t pair_claim(int* counter, uint32_t it, int64_t n_pairs) {
  return pair_resolve(pair_claim_raw(counter, it), n_pairs);
}

// ---- 16-bit operand formats (FMT_BF16 / FMT_F16, at_prepjobs.cuh). Activations are rounded to the operand format once,
// where an epilogue hands them to the next MMA; fp16 conversions saturate at +-65504 instead of producing infinities.
template <int FMT>
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  uint32_t w;
  if constexpr (FMT == FMT_F16) asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(w) : "f"(hi), "f"(lo));
  else asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w) : "f"(hi), "f"(lo));
  return w;
}
// the same with the ReLU clamp built into the conversion (negative and NaN inputs give +0)
template <int FMT>
__device__ __forceinline__ uint32_t pack2_relu(float lo, float hi) {
  uint32_t w;
  if constexpr (FMT == FMT_F16) asm("cvt.rn.relu.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(w) : "f"(hi), "f"(lo));
  else asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(w) : "f"(hi), "f"(lo));
  return w;
}
template <int FMT>
__device__ __forceinline__ void unpack2(uint32_t w, float& lo, float& hi) {
  if constexpr (FMT == FMT_F16) {
    asm("{.reg .f16 l, h; mov.b32 {l, h}, %2; cvt.f32.f16 %0, l; cvt.f32.f16 %1, h;}" : "=f"(lo), "=f"(hi) : "r"(w));
  } else {
    lo = __uint_as_float(w << 16);
    hi = __uint_as_float(w & 0xFFFF0000u);
  }
}

template <int ACT>
__device__ __forceinline__ float act_fn(float x) {
  if constexpr (ACT == AT_ACT_RELU) return fmaxf(x, 0.0f);
  else return apply_act(x, ACT);
}
template <int ACT, int FMT>
__device__ __forceinline__ uint32_t pack2_act(float lo, float hi) {
  if constexpr (ACT == AT_ACT_RELU) return pack2_relu<FMT>(lo, hi);
  else return pack2<FMT>(act_fn<ACT>(lo), act_fn<ACT>(hi));
}
// eight fp32 values -> activation -> 16-bit operands, zeroed when `keep` is 0 (padding rows): one conversion per pair (the
// ReLU rides on it) and a single AND for the row mask.
template <int ACT, int FMT>
__device__ __forceinline__ uint4 pack8_act(const float (&v)[8], uint32_t keep) {
  return make_uint4(pack2_act<ACT, FMT>(v[0], v[1]) & keep, pack2_act<ACT, FMT>(v[2], v[3]) & keep,
                    pack2_act<ACT, FMT>(v[4], v[5]) & keep, pack2_act<ACT, FMT>(v[6], v[7]) & keep);
}
// v[0..7] += the eight 16-bit values of `w`
template <int FMT>
__device__ __forceinline__ void add8(float (&v)[8], const uint4& w) {
  float lo, hi;
  unpack2<FMT>(w.x, lo, hi); v[0] += lo; v[1] += hi;
  unpack2<FMT>(w.y, lo, hi); v[2] += lo; v[3] += hi;
  unpack2<FMT>(w.z, lo, hi); v[4] += lo; v[5] += hi;
  unpack2<FMT>(w.w, lo, hi); v[6] += lo; v[7] += hi;
}

// 32 bytes to global memory in one instruction (sm_100: STG.256); dst 32-byte aligned
__device__ __forceinline__ void stg256(void* dst, const uint4& lo, const uint4& hi) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst), "r"(lo.x), "r"(lo.y), "r"(lo.z), "r"(lo.w),
               "r"(hi.x), "r"(hi.y), "r"(hi.z), "r"(hi.w)
               : "memory");
}
__device__ __forceinline__ void sts128(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}

// Descriptor high words are constant: SBO = 128 B (rows of a core matrix 16 B apart), version 1, no swizzle.
constexpr uint32_t kDescHi = (128u >> 4) | (1u << 14);

// NM K-steps of one weight slab against one board's image (used by the fused tower + transformer kernel): A advances two
// 8-channel planes per step, B two planes of the slab.
template <int NM>
__device__ __forceinline__ void issue_slab1(uint32_t tmem_d, uint32_t a16, uint32_t b16, uint32_t a_lo_fixed, uint32_t b_lo_fixed,
                                            uint32_t b_step16, uint32_t idesc, bool first) {
#pragma unroll
  for (int j = 0; j < NM; ++j) {
    const uint32_t a_lo = a_lo_fixed | (a16 + (uint32_t)j * (2 * kPlaneBytes >> 4));
    const uint32_t b_lo = b_lo_fixed | (b16 + (uint32_t)j * b_step16);
    mma_bf16(tmem_d, ((uint64_t)kDescHi << 32) | a_lo, ((uint64_t)kDescHi << 32) | b_lo, idesc, !(first && j == 0));
  }
}

// The convolution tower: one CTA per SM, two boards per CTA, persistent over board pairs.
//
//   warp 0            weight producer: one thread streams the pair's slab sequence through a 4 x 16 KB ring (bulk copies
//                     global -> shared, completion on the stage's "full" mbarrier)
//   warp 1, warp 18   MMA issuers, one per board (converged warp, one elected lane): per layer wait for the board's input image, then per
//                     slab wait "full", issue the slab's K-steps (tcgen05.mma, M=128 positions x N=Cout x K=16) into the
//                     board's accumulator, commit to the stage's "empty" barrier (two commits, one per board, free a
//                     stage); after the last slab commit to the board's "accumulator complete" barrier
//   warps 2..17       epilogue, 8 per board (4 TMEM lane quarters x 2 channel shares): layer 0 on CUDA cores from the
//                     bitmasks, then per layer tcgen05.ld -> +bias (+residual) -> activation -> bf16 -> next image
//
// The two boards are independent pipelines that share only the weight ring: while one board's epilogue drains its
// accumulator the other board's MMAs can keep the tensor pipe busy.
//
// What was measured on the way here (profiles/r01_tower_*.txt, r01_mma_rate_probe.txt):
//  * two issuing threads on different accumulators saturate the pipe at 64 cycles per 128x128x16 MMA, also while bulk copies
//    and epilogue traffic hit shared memory; one thread alone reaches 85;
//  * a tcgen05.mma holds its issuing thread while the pipe is busy, so an issuer's per-slab time is its bookkeeping PLUS the
//    256 cycles of its four MMAs: the bookkeeping has to stay well under 256 cycles for the pipe to fill. Hence one plain
//    thread per board with a short loop (no votes, elections or warp syncs per slab), and the next slab's operand offset
//    fetched while the current one is issued;
//  * tried and dropped: CTA pairs with cta_group::2, an 8 x 8 KB ring, two issuing warps per board (no gain, and the
//    accumulation order -- hence the low bits -- then depends on timing), two accumulators per board with a channel-half
//    hand-off (overlap gain smaller than the added per-slab instructions).
template <int ACT, int FMT>
__global__ void __launch_bounds__(kTowerThreads, 1) tower_kernel(const __grid_constant__ TowerParams p) {
  constexpr int NS = kStages;
  constexpr uint32_t SB = kStageBytes;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* ring = smem;
  uint8_t* act = ring + kStages * kStageBytes;                        // [board][buf][kBufBytes]
  float* s_bias = reinterpret_cast<float*>(act + kBoards * 2 * kBufBytes);  // [kMaxLayers][128]
  float* s_w0 = s_bias + kMaxLayers * 128;                            // [9][c0]
  float* s_b0 = s_w0 + 9 * 128;                                       // [c0]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_b0 + 128);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + kMaxBars);
  volatile unsigned long long* pairq = reinterpret_cast<volatile unsigned long long*>(tmem_slot + 4);

  const int tid = threadIdx.x, warp = tid / 32, lane = tid % 32;
  const uint32_t ring_u = smem_u32(ring), act_u = smem_u32(act);
  const uint32_t bar_full = smem_u32(bars), bar_empty = smem_u32(bars + NS);
  const uint32_t bar_acc0 = smem_u32(bars + 2 * NS), bar_acc1 = smem_u32(bars + 2 * NS + 1);  // accumulator of board b complete
  const uint32_t bar_act0 = smem_u32(bars + 2 * NS + 2), bar_act1 = smem_u32(bars + 2 * NS + 3);  // input image of board b written
  const uint32_t bar_last0 = smem_u32(bars + 2 * NS + 4), bar_last1 = smem_u32(bars + 2 * NS + 5);  // a pair's last accumulator of board b
  const bool use_ring = !AT_DBG(16u);  // 16: timing experiment without the weight ring (issue loop + MMAs + epilogue only)

  // ---- one-time setup --------------------------------------------------------------------
  for (int i = tid; i < kBoards * 2 * kBufBytes / 16; i += kTowerThreads) reinterpret_cast<uint4*>(act)[i] = make_uint4(0, 0, 0, 0);
  for (int i = tid; i < p.n_layers * 128; i += kTowerThreads) s_bias[i] = p.bias[i];
  for (int i = tid; i < 9 * p.c0; i += kTowerThreads) s_w0[i] = p.w0[i];
  for (int i = tid; i < p.c0; i += kTowerThreads) s_b0[i] = p.b0[i];
  if (tid == 0) {
    for (int i = 0; i < kPairQueue; ++i) pairq[i] = ~0ull;
    for (int i = 0; i < NS; ++i) {
      mbar_init(bar_full + 8 * i, 1);
      mbar_init(bar_empty + 8 * i, kBoards);  // both boards' issuers are done with the slab
    }
    mbar_init(bar_acc0, 1);
    mbar_init(bar_acc1, 1);
    mbar_init(bar_act0, kEpiWarps / kBoards);  // one arrival per epilogue warp of the board
    mbar_init(bar_act1, kEpiWarps / kBoards);
    mbar_init(bar_last0, 1);
    mbar_init(bar_last1, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(smem_u32(tmem_slot), kTowerTmemCols);
    tmem_relinquish();
  }
  fence_proxy_async_smem();  // the zeroed halos are read by the tensor core (async proxy)
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int64_t n_pairs = (p.batch + kBoards - 1) / kBoards;
  // the pair of round `it`: from the producer's queue (without the ring -- experiments only -- nobody runs ahead: by stride)
  auto get_pair = [&](uint32_t it, int64_t* pair) -> bool {
    if (!use_ring) {
      *pair = pair_claim(nullptr, it, n_pairs);
      return true;
    }
    return pairq_get(pairq, it, pair, p.status);
  };

  if (warp == 0) {
    // ===== weight producer: one thread streams every layer's slabs through the ring, refilling each stage as soon as it
    // is released (a lane-per-stage variant refills in lock step -- all four lanes leave their waits together -- and was
    // measured slower) =====
    if (lane == 0 && use_ring) {
      uint32_t stage = 0, phase = 0;
      bool ok = true;
      AT_TL_DECL
      int64_t next = pair_claim(p.pair_counter, 0, n_pairs);
      pairq_put(pairq, 0, next);
      for (uint32_t it = 0; ok; ++it) {
        const int64_t pair = next;
        if (pair < 0) break;
        const int64_t raw = pair_claim_raw(p.pair_counter, it + 1);  // one round ahead: the epilogue prepares the next pair during this one
        AT_TL_PAIR(it)
        const uint8_t* src = p.wstream;
        for (int l = 0; l < p.n_layers && ok; ++l) {
          if (l == 1 || p.n_layers == 1) {
            next = pair_resolve(raw, n_pairs);
            pairq_put(pairq, it + 1, next);
          }
          AT_TL(100 + l);
          const uint32_t bytes = p.layers[l].slab_bytes;
          const int n_slabs = p.layers[l].n_slabs;
          for (int s = 0; s < n_slabs; ++s) {
            ok = wait_bar(bar_empty + 8 * stage, phase ^ 1, p.status, 1u);
            if (!ok) break;
            if (AT_DBG(1u)) {
              mbar_arrive(bar_full + 8 * stage);
            } else {
              mbar_arrive_expect_tx(bar_full + 8 * stage, bytes);
              bulk_g2s(ring_u + stage * SB, src, bytes, bar_full + 8 * stage);
            }
            src += bytes;
            stage = stage + 1 == NS ? 0 : stage + 1;
            phase ^= (stage == 0);
          }
        }
      }
    }
  } else if (warp == 1 || warp == kMmaWarpB) {
    // ===== MMA issuer of board `me`: the warp stays converged, one elected lane issues tcgen05.mma / commit =====
    // (Measured against a single-thread loop under one elect.sync: that one has fewer instructions per slab -- its
    // skeleton without MMAs runs 20 % faster -- but the full kernel is 1.5 % slower with it.)
    const int me = warp == 1 ? 0 : 1;
    const uint32_t my_acc = me == 0 ? bar_acc0 : bar_acc1, my_act = me == 0 ? bar_act0 : bar_act1;
    const uint32_t tmem_d0 = tmem_base + (uint32_t)me * 128;
    const uint32_t a_lo_fixed = (uint32_t)(kPlaneBytes >> 4) << 16;  // LBO = one 8-channel plane
    uint32_t par = 0;  // pair parity: which of the two image buffers is "buffer 0" and which accumulator is in use
    const bool no_mma = AT_DBG(2u);
    uint32_t stage = 0, phase = 0, lphase = 0;
    bool ok = true;
    AT_TL_DECL
    for (uint32_t it = 0; ok; ++it) {
      int64_t pair;
      ok = get_pair(it, &pair);
      if (!ok || pair < 0) break;
      AT_TL_PAIR(it)
      AT_TL(1);
      const uint32_t tmem_d = tmem_d0 + par * 256u;
      for (int l = 0; l < p.n_layers && ok; ++l) {
        const TowerLayer& L = p.layers[l];
        ok = __all_sync(0xffffffffu, wait_bar(my_act, lphase, p.status, 2u));  // this layer's input image is written
        if (!ok) break;
        AT_TL(10 + l);  // input image ready
        lphase ^= 1u;
        tc_fence_after();
        const uint32_t idesc = idesc_16(kM, (uint32_t)L.cout, FMT == FMT_BF16);
        // descriptor low words: [13:0] start address >> 4, [29:16] LBO >> 4; the address sums below stay inside 14 bits
        const uint32_t a_base = a_lo_fixed | ((act_u + (uint32_t)(me * 2 + ((uint32_t)L.src ^ par)) * kBufBytes) >> 4);
        const uint32_t b_base = ((uint32_t)L.cout << 16) | (ring_u >> 4);  // LBO = one 8-channel plane of the slab (Cout rows of 16 B)
        const uint32_t b_step = 2u * (uint32_t)L.cout;                      // one K-step = two planes
        const int n_slabs = L.n_slabs;
        const int steps = no_mma ? 0 : L.ks / 16;
        for (int s = 0; s < n_slabs; ++s) {
          const uint32_t a_lo = a_base + L.aoff[s];
          if (use_ring) {
            ok = __all_sync(0xffffffffu, wait_bar(bar_full + 8 * stage, phase, p.status, 4u));
            if (!ok) break;
          }
          tc_fence_after();
          const uint32_t b_lo = b_base + stage * (SB >> 4);
          if (elect_one()) {
            if (steps == 4) {
#pragma unroll
              for (int j = 0; j < 4; ++j)
                mma_bf16(tmem_d, ((uint64_t)kDescHi << 32) | (a_lo + (uint32_t)j * (2 * kPlaneBytes >> 4)),
                         ((uint64_t)kDescHi << 32) | (b_lo + (uint32_t)j * b_step), idesc, s > 0 || j > 0);
            } else {
              for (int j = 0; j < steps; ++j)
                mma_bf16(tmem_d, ((uint64_t)kDescHi << 32) | (a_lo + (uint32_t)j * (2 * kPlaneBytes >> 4)),
                         ((uint64_t)kDescHi << 32) | (b_lo + (uint32_t)j * b_step), idesc, s > 0 || j > 0);
            }
            if (use_ring) mma_commit(bar_empty + 8 * stage);  // the slab may be overwritten once these MMAs retire
          }
          __syncwarp();
          stage = stage + 1 == NS ? 0 : stage + 1;
          phase ^= (stage == 0);
        }
        if (ok && elect_one()) {
          mma_commit(my_acc);
          if (kPairDrain && l + 1 == p.n_layers) mma_commit(me == 0 ? bar_last0 : bar_last1);  // for the sibling board's epilogue warps
        }
        __syncwarp();
        AT_TL(30 + l);  // last slab issued, accumulator committed
      }
      if (AT_TOWER_OVERLAP) par ^= 1u;
    }
  } else {
    // ===== epilogue warps: 8 per board; warp%4 selects the TMEM lane quarter =====
    const int eb = (warp - 2) / (4 * kEpiSplit);   // board of the pair
    const int eh = ((warp - 2) / 4) % kEpiSplit;   // which share of the channels
    const int q = warp % 4;                        // TMEM lane quarter this warp may read
    const int m = q * 32 + lane;  // padded position = MMA row = TMEM lane
    const int y = m / kPadW, x = m % kPadW;
    const bool inside = (y < p.rows) && (x < p.cols);
    const uint32_t keep = inside ? 0xFFFFFFFFu : 0u;
    const int pos = y * p.cols + x;  // compact position index (token order h*W + w)
    const uint32_t my_act_u = act_u + (uint32_t)eb * 2 * kBufBytes + (uint32_t)(kHalo + m) * 16;  // this row, plane 0, buffer 0
    const uint32_t tmem_row0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)eb * 128;
    uint32_t acc_phase = 0, par = 0, last_phase = 0;
    bool ok = true;
    const uint32_t my_act = eb == 1 ? bar_act1 : bar_act0;
    const uint32_t my_acc = eb == 1 ? bar_acc1 : bar_acc0;
    auto arrive_act = [&]() {
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(my_act);
    };
    // ---- layer 0 on CUDA cores, straight from the bitmasks (extractor.py:33-46 + conv_body.0): this thread's position,
    // channels eh*8 + 16*i .. +7. Either kept in registers (c0 <= 16 * kL0Keep) or written to the image `dst_buf` at once.
    constexpr int kL0Keep = 2;
    auto layer0 = [&](int64_t board, bool valid, uint4 (&keepv)[kL0Keep], bool to_regs, uint32_t dst_buf) {
      if (AT_DBG(8u)) return;
      float g[9];
#pragma unroll
      for (int t = 0; t < 9; ++t) g[t] = 0.0f;
      if (valid) {
        const PackedState& st = p.states[board];
#pragma unroll
        for (int t = 0; t < 9; ++t) {
          const int yy = y + t / 3 - 1, xx = x + t % 3 - 1;
          if (yy >= 0 && yy < p.rows && xx >= 0 && xx < p.cols) {
            const int i = yy * p.cols + xx;
            const uint32_t bit = 1u << (i & 31);
            g[t] = (st.death[i >> 5] & bit) ? -1.0f : ((st.occ[i >> 5] & bit) ? 1.0f : 0.0f);
          }
        }
      }
      auto one = [&](int c8) {
        const float4 ba = *reinterpret_cast<const float4*>(s_b0 + c8), bb = *reinterpret_cast<const float4*>(s_b0 + c8 + 4);
        float v[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
#pragma unroll
        for (int t = 0; t < 9; ++t) {
          const float4 wa = *reinterpret_cast<const float4*>(s_w0 + t * p.c0 + c8);
          const float4 wb = *reinterpret_cast<const float4*>(s_w0 + t * p.c0 + c8 + 4);
          v[0] = fmaf(g[t], wa.x, v[0]); v[1] = fmaf(g[t], wa.y, v[1]); v[2] = fmaf(g[t], wa.z, v[2]); v[3] = fmaf(g[t], wa.w, v[3]);
          v[4] = fmaf(g[t], wb.x, v[4]); v[5] = fmaf(g[t], wb.y, v[5]); v[6] = fmaf(g[t], wb.z, v[6]); v[7] = fmaf(g[t], wb.w, v[7]);
        }
        return pack8_act<ACT, FMT>(v, keep);
      };
      if (to_regs) {
#pragma unroll
        for (int i = 0; i < kL0Keep; ++i) {
          const int c8 = eh * 8 + i * 8 * kEpiSplit;
          keepv[i] = c8 < p.c0 ? one(c8) : make_uint4(0, 0, 0, 0);
        }
      } else {
#pragma unroll 1
        for (int c8 = eh * 8; c8 < p.c0; c8 += 8 * kEpiSplit) sts128(my_act_u + dst_buf * kBufBytes + (uint32_t)(c8 / 8) * kPlaneBytes, one(c8));
      }
    };
    auto store_layer0 = [&](const uint4 (&keepv)[kL0Keep], uint32_t dst_buf) {
#pragma unroll
      for (int i = 0; i < kL0Keep; ++i) {
        const int c8 = eh * 8 + i * 8 * kEpiSplit;
        if (c8 < p.c0) sts128(my_act_u + dst_buf * kBufBytes + (uint32_t)(c8 / 8) * kPlaneBytes, keepv[i]);
      }
    };
    const bool l0_in_regs = p.c0 <= 8 * kEpiSplit * kL0Keep;
    uint4 l0v[kL0Keep];
    AT_TL_DECL
    bool first = true;
    for (uint32_t it = 0; ok; ++it) {
      int64_t pair, next_pair;
      ok = get_pair(it, &pair);
      if (!ok || pair < 0) break;
      next_pair = -1;
      AT_TL_PAIR(it)
      AT_TL(1);
      const int64_t board = pair * kBoards + eb;
      const bool valid = board < p.batch;
      const uint32_t tmem_row = tmem_row0 + par * 256u;
      if (!AT_TOWER_OVERLAP || first) {
        // The previous board's last epilogue may still be reading its residual from image 0 in the sibling warp (other
        // channel share, same rows): all eight epilogue warps of this board finish it before anyone writes the new image.
        if (!AT_DBG(64u)) asm volatile("bar.sync %0, %1;" ::"r"(1 + eb), "r"(kEpiWarps / kBoards * 32) : "memory");
        layer0(board, valid, l0v, false, par);
        arrive_act();
        first = false;
      }
      AT_TL(2);  // layer 0 written

      for (int l = 0; l < p.n_layers; ++l) {
        const TowerLayer& L = p.layers[l];
        // Overlap: while the last layer's MMAs run, work out the NEXT pair's layer 0 (registers); once they are done its
        // source image -- the pair's scratch buffer -- is free, so the next pair's first image goes there and its MMAs start
        // (on the other accumulator) while this pair's last accumulator is still being drained to HBM below.
        if (AT_TOWER_OVERLAP && l + 1 == p.n_layers) {  // published by the producer early in this round
          ok = get_pair(it + 1, &next_pair);
          if (!ok) break;
        }
        const bool hand_over = AT_TOWER_OVERLAP && l + 1 == p.n_layers && next_pair >= 0;
        const int64_t next_board = next_pair * kBoards + eb;
        if (hand_over && l0_in_regs) layer0(next_board, next_board < p.batch, l0v, true, 0);
        ok = wait_bar(my_acc, acc_phase, p.status, 8u);
        if (!ok) break;
        AT_TL(50 + l);  // accumulator seen complete
        acc_phase ^= 1;
        tc_fence_after();
        if (hand_over) {
          if (l0_in_regs) store_layer0(l0v, par ^ 1u);
          else layer0(next_board, next_board < p.batch, l0v, false, par ^ 1u);
          arrive_act();
        }
        const float* bias = s_bias + L.bias_slot * 128;
        const bool has_res = L.res >= 0;
        const uint32_t res_u = my_act_u + ((uint32_t)(has_res ? L.res : 0) ^ par) * kBufBytes;
        const bool to_smem = L.dst < 2;
        const uint32_t dst_u = my_act_u + ((uint32_t)(to_smem ? L.dst : 0) ^ par) * kBufBytes;
        uint8_t* gdst = p.fc_a + (((board >> 7) * p.n_kc + (int64_t)pos * (L.cout / 8)) * 128 + (board & 127)) * 16;
        const bool gstore = valid && inside;
        const int cout = AT_DBG(4u) ? 0 : L.cout;
        // the chunk loop, specialised on (residual, destination) so its body has no per-group branches
        auto drain = [&](auto HAS_RES, auto TO_SMEM) {
#pragma unroll 1
          for (int c0 = eh * 32; c0 < cout; c0 += 32 * kEpiSplit) {
            uint32_t r[32];
            tmem_ld32(tmem_row + (uint32_t)c0, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float4 ba = *reinterpret_cast<const float4*>(bias + c0 + i * 8), bb = *reinterpret_cast<const float4*>(bias + c0 + i * 8 + 4);
              float v[8] = {__uint_as_float(r[i * 8 + 0]) + ba.x, __uint_as_float(r[i * 8 + 1]) + ba.y, __uint_as_float(r[i * 8 + 2]) + ba.z,
                            __uint_as_float(r[i * 8 + 3]) + ba.w, __uint_as_float(r[i * 8 + 4]) + bb.x, __uint_as_float(r[i * 8 + 5]) + bb.y,
                            __uint_as_float(r[i * 8 + 6]) + bb.z, __uint_as_float(r[i * 8 + 7]) + bb.w};
              const int chunk = c0 / 8 + i;
              if constexpr (decltype(HAS_RES)::value) {
                add8<FMT>(v, lds128(res_u + (uint32_t)chunk * kPlaneBytes));
              }
              const uint4 out = pack8_act<ACT, FMT>(v, keep);
              if constexpr (decltype(TO_SMEM)::value) {
                sts128(dst_u + (uint32_t)chunk * kPlaneBytes, out);
              } else {
                if (gstore) *reinterpret_cast<uint4*>(gdst + (size_t)chunk * 2048) = out;
              }
            }
          }
        };
        // The LAST layer goes to HBM as the trunk FC's A operand, [k-chunk][128 boards][8 channels]: a board owns 16 bytes of
        // every 128-byte line, so a store instruction costs 32 line accesses whatever its width -- that, not bandwidth, is what
        // the drain takes. The pair's two boards sit next to each other there, so for this layer the threads change roles:
        // each takes 32 channels of BOTH boards (reading the sibling board's accumulator and residual image too) and writes
        // full 32-byte sectors with 256-bit stores -- half as many store instructions.
        auto drain_pair = [&](auto HAS_RES) {
          const int64_t board0 = pair * kBoards;
          const bool both = board0 + 1 < p.batch;
          uint8_t* g0 = p.fc_a + (((board0 >> 7) * p.n_kc + (int64_t)pos * (L.cout / 8)) * 128 + (board0 & 127)) * 16;
          const uint32_t tmem_q = tmem_base + ((uint32_t)(q * 32) << 16) + par * 256u;
          const uint32_t row_u = act_u + (uint32_t)(kHalo + m) * 16 + ((uint32_t)(has_res ? L.res : 0) ^ par) * kBufBytes;
#pragma unroll 1
          for (int c0 = (eb * kEpiSplit + eh) * 32; c0 < cout; c0 += 32 * kEpiSplit * kBoards) {
            uint4 out[kBoards][4];
#pragma unroll
            for (int b = 0; b < kBoards; ++b) {
              uint32_t r[32];
              tmem_ld32(tmem_q + (uint32_t)b * 128u + (uint32_t)c0, r);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const float4 ba = *reinterpret_cast<const float4*>(bias + c0 + i * 8), bb = *reinterpret_cast<const float4*>(bias + c0 + i * 8 + 4);
                float v[8] = {__uint_as_float(r[i * 8 + 0]) + ba.x, __uint_as_float(r[i * 8 + 1]) + ba.y, __uint_as_float(r[i * 8 + 2]) + ba.z,
                              __uint_as_float(r[i * 8 + 3]) + ba.w, __uint_as_float(r[i * 8 + 4]) + bb.x, __uint_as_float(r[i * 8 + 5]) + bb.y,
                              __uint_as_float(r[i * 8 + 6]) + bb.z, __uint_as_float(r[i * 8 + 7]) + bb.w};
                if constexpr (decltype(HAS_RES)::value)
                  add8<FMT>(v, lds128(row_u + (uint32_t)b * 2 * kBufBytes + (uint32_t)(c0 / 8 + i) * kPlaneBytes));
                out[b][i] = pack8_act<ACT, FMT>(v, keep);
              }
            }
            if (inside) {
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                uint8_t* dst = g0 + (size_t)(c0 / 8 + i) * 2048;
                if (both) stg256(dst, out[0][i], out[1][i]);
                else *reinterpret_cast<uint4*>(dst) = out[0][i];  // the batch's last, odd board
              }
            }
          }
        };
        using std::false_type;
        using std::true_type;
        if (kPairDrain && !to_smem) {
          // the sibling board's last accumulator: its own barrier, completed once per pair (the per-layer one could be a
          // phase further by the time a late warp looks at it)
          ok = wait_bar(eb == 1 ? bar_last0 : bar_last1, last_phase, p.status, 8u);
          if (!ok) break;
          last_phase ^= 1u;
          tc_fence_after();
          if (has_res) drain_pair(true_type{});
          else drain_pair(false_type{});
        } else if (has_res) {
          if (to_smem) drain(true_type{}, true_type{});
          else drain(true_type{}, false_type{});
        } else {
          if (to_smem) drain(false_type{}, true_type{});
          else drain(false_type{}, false_type{});
        }
        tc_fence_before();  // TMEM reads are done before the next layer's MMAs overwrite the accumulator
        if (l + 1 < p.n_layers) arrive_act();
        AT_TL(70 + l);  // drained, next image posted
      }
      if (AT_TOWER_OVERLAP) {
        // the sibling warp (other channel share, same rows) may still read this pair's residual image, which the next pair's
        // second layer overwrites: everyone finishes the drain first
        // (with the pair drain every epilogue warp reads both boards' images: all sixteen warps meet)
        if (!AT_DBG(64u)) {
          if (kPairDrain) asm volatile("bar.sync 3, %0;" ::"n"(kEpiWarps * 32) : "memory");
          else asm volatile("bar.sync %0, %1;" ::"r"(1 + eb), "r"(kEpiWarps / kBoards * 32) : "memory");
        }
        par ^= 1u;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, kTowerTmemCols);
}

// ---- trunk FC: split-K tcgen05 GEMM ---------------------------------------------------------
constexpr int kFcStages = 3;  // 96 KB: two CTAs per SM, one CTA's ring fill / accumulator drain under the other's stream
constexpr int kFcThreads = 192;

struct FcParams {
  const uint8_t* a;  // [Mtiles][n_kc][128][8] bf16
  const uint8_t* w;  // [n_kc][N][8] bf16
  float* partial;    // [splits][batch][N]
  int64_t batch;
  int n_kc, N, slabs_per_split, n_slabs;
  uint32_t* status;
  int fmt;  // operand format of a and w (FMT_BF16 / FMT_F16)
};

constexpr size_t fc_smem_bytes() { return (size_t)kFcStages * 2 * kStageBytes + (2 * kFcStages + 1) * 8 + 16; }

__global__ void __launch_bounds__(kFcThreads, 2) fc_kernel(const __grid_constant__ FcParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kFcStages * 2 * kStageBytes);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kFcStages + 1);
  const int tid = threadIdx.x, warp = tid / 32, lane = tid % 32;
  const uint32_t smem_u = smem_u32(smem);
  const uint32_t bar_full = smem_u32(bars), bar_empty = smem_u32(bars + kFcStages), bar_acc = smem_u32(bars + 2 * kFcStages);
  uint32_t cols = 32;
  while (cols < (uint32_t)p.N) cols <<= 1;

  if (tid == 0) {
    for (int i = 0; i < kFcStages; ++i) {
      mbar_init(bar_full + 8 * i, 1);
      mbar_init(bar_empty + 8 * i, 1);
    }
    mbar_init(bar_acc, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(smem_u32(tmem_slot), cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int mt = blockIdx.x, z = blockIdx.y;
  const int s_begin = z * p.slabs_per_split;
  const int s_end = min(p.n_slabs, s_begin + p.slabs_per_split);
  const uint32_t b_bytes = (uint32_t)p.N * 128;  // 8 chunks x N rows x 16 B

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int s = s_begin; s < s_end; ++s) {
        if (!wait_bar(bar_empty + 8 * stage, phase ^ 1, p.status, 16u)) break;
        mbar_arrive_expect_tx(bar_full + 8 * stage, kStageBytes + b_bytes);
        bulk_g2s(smem_u + stage * 2 * kStageBytes, p.a + ((size_t)mt * p.n_kc + (size_t)s * 8) * 2048, kStageBytes, bar_full + 8 * stage);
        bulk_g2s(smem_u + stage * 2 * kStageBytes + kStageBytes, p.w + (size_t)s * b_bytes, b_bytes, bar_full + 8 * stage);
        if (++stage == kFcStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t idesc = idesc_16(kM, (uint32_t)p.N, p.fmt == FMT_BF16);
      const uint32_t b_lbo = (uint32_t)p.N * 16;
      bool ok = true;
      for (int s = s_begin; s < s_end; ++s) {
        ok = wait_bar(bar_full + 8 * stage, phase, p.status, 32u);
        if (!ok) break;
        tc_fence_after();
        const uint32_t a_base = smem_u + stage * 2 * kStageBytes, b_base = a_base + kStageBytes;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint64_t da = smem_desc(a_base + (uint32_t)(2 * j) * 2048, 2048, 128, LAYOUT_NONE);
          const uint64_t db = smem_desc(b_base + (uint32_t)(2 * j) * b_lbo, b_lbo, 128, LAYOUT_NONE);
          mma_bf16(tmem_base, da, db, idesc, (s != s_begin) || j != 0);
        }
        mma_commit(bar_empty + 8 * stage);
        if (++stage == kFcStages) { stage = 0; phase ^= 1; }
      }
      mma_commit(bar_acc);
    }
  } else {
    const int q = warp % 4;
    const int m = q * 32 + lane;
    const int64_t board = (int64_t)mt * 128 + m;
    if (wait_bar(bar_acc, 0, p.status, 64u)) {
      tc_fence_after();
      float* dst = p.partial + ((int64_t)z * p.batch + board) * p.N;
      for (int c0 = 0; c0 < p.N; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, r);
        tmem_ld_wait();
        if (board < p.batch && s_end > s_begin) {
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (c0 + i * 4 < p.N)
              *reinterpret_cast<float4*>(dst + c0 + i * 4) = make_float4(__uint_as_float(r[i * 4]), __uint_as_float(r[i * 4 + 1]),
                                                                         __uint_as_float(r[i * 4 + 2]), __uint_as_float(r[i * 4 + 3]));
        } else if (board < p.batch) {
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (c0 + i * 4 < p.N) *reinterpret_cast<float4*>(dst + c0 + i * 4) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, cols);
}

#include "at_bf16_tf.inc"

constexpr size_t tower_tf_smem_bytes() {
  return (size_t)kTfStages * kStageBytes + (size_t)kBoards * 2 * kBufBytes + kMaskBytes + (size_t)kMaxLayers * 128 * 4 + 9 * 128 * 4 +
         128 * 4 + (size_t)kXchFloats * 4 + (size_t)kTfBars * 8 + 16 + kPairQueue * 8;
}

}  // namespace

// ---- host side ---------------------------------------------------------------------------------

struct Bf16Path {
  TowerParams tp{};           // template: pointers filled per call
  std::vector<int64_t> layer_wt;  // folded-set offset of each MMA layer's Wt
  std::vector<int> layer_k;
  std::vector<int64_t> bias_off;
  std::vector<int> bias_w;
  int64_t w0_off = 0, b0_off = 0;
  size_t wstream_bytes = 0;
  // per weight set
  uint8_t* wstream[2] = {nullptr, nullptr};
  uint8_t* wfc[2] = {nullptr, nullptr};
  float* bias[2] = {nullptr, nullptr};
  // trunk FC
  int fc_n = 0, fc_flat = 0, n_kc = 0;
  int64_t fc_wt = 0;
  // scratch
  int64_t max_batch = 0;
  uint8_t* fc_a = nullptr;
  float* partial = nullptr;
  int max_splits = 0;
  float* other = nullptr;
  float* grid = nullptr;
  uint32_t* status = nullptr;
  int sms = 148;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};  // tower begin/end, fc begin/end
  bool timed = false;
  bool timing = false;  // record the event pairs around the tower / FC kernels (at_engine_set_kernel_timing): off on the serving path
  // transformer extension (tower_tf_kernel)
  bool tf = false;
  TowerTfParams ttp{};
  float* fc_bias[2] = {nullptr, nullptr};  // trunk FC bias with the encoder's final LayerNorm folded in (fp32 [fc_n])
  int64_t fc_bias_off = 0;                  // the plain bias in the folded set
  size_t conv_stream_bytes = 0, pe_t_off = 0;
  int packed_fmt[2] = {-1, -1};             // operand format the 16-bit images of each weight set currently hold
};

int bf16_create(const Plan& p, int64_t max_batch, Bf16Path** out, std::string* reason) {
  *out = nullptr;
  const at_model_desc& d = p.d;
  auto no = [&](const std::string& why) {
    *reason = why;
    return AT_OK;
  };
  if (p.transformer) {
    if (d.t_dim != kTfD) return no("transformer on the tcgen05 path needs TRANSFORMER_DIM = 128");
    if (d.t_heads * kTfDh != kTfD) return no("transformer on the tcgen05 path needs head width 32");
    if (d.t_ff % 128 != 0 || d.t_ff > 1024) return no("TRANSFORMER_FC_DIM must be a multiple of 128, at most 1024");
    if (d.t_layers > kMaxTfLayers) return no("at most 4 transformer layers on the tcgen05 path");
  }
  if (d.grid_channels != 1) return no("GRID_INPUT_CHANNELS must be 1");
  if (d.env.rows * kPadW > kM || d.env.cols + 1 > kPadW) return no("board must fit a 16-wide, 128-position padded image (ROWS <= 8, COLS <= 15)");
  if (p.convs.empty()) return no("needs at least one conv_body layer");
  if (p.shared.empty()) return no("needs at least one shared FC layer");
  std::vector<const ConvLayer*> seq;
  for (const ConvLayer& c : p.convs) seq.push_back(&c);
  if (p.has_res_proj) seq.push_back(&p.res_proj);
  for (const ResBlock& r : p.res) {
    seq.push_back(&r.c1);
    seq.push_back(&r.c2);
  }
  // input_proj (model.py:179-186: a 1x1 convolution to TRANSFORMER_DIM when the tower ends on another width, no activation)
  // is one more tensor-core layer whose output becomes the residual stream
  ConvLayer proj;
  const bool with_proj = p.transformer && p.has_input_proj;
  if (with_proj) {
    proj.f = p.input_proj;
    proj.cin = p.cnn_ch; proj.cout = d.t_dim; proj.kh = proj.kw = 1;
    seq.push_back(&proj);
  }
  if ((int)seq.size() - 1 > kMaxLayers || seq.size() < 2) return no("needs 2..13 convolution layers");
  for (size_t i = 0; i < seq.size(); ++i) {
    const ConvLayer& c = *seq[i];
    const bool k3 = c.kh == 3 && c.kw == 3 && c.ph == 1 && c.pw == 1;
    const bool k1 = c.kh == 1 && c.kw == 1 && c.ph == 0 && c.pw == 0;
    if (c.sh != 1 || c.sw != 1 || !(k3 || (k1 && i > 0))) return no("convolutions must be 3x3 stride 1 pad 1 (or the 1x1 projection)");
    if (c.cout % 16 != 0 || c.cout > kMaxCh || (i > 0 && c.cin % 16 != 0)) return no("channel counts must be multiples of 16, at most 128");
    if (c.cout % 32 != 0) return no("channel counts must be multiples of 32");
    if (i > 0) {
      const int ks = c.cin >= 64 ? 64 : c.cin;
      if ((ks != 16 && ks != 32 && ks != 64) || c.cin % ks != 0 || c.kh * c.kw * c.cin / ks > kMaxSlabs)
        return no("input channel counts must be 16, 32 or a multiple of 64");
    }
  }
  const Affine& fc = p.shared[0];
  if (fc.N % 32 != 0 || fc.N > 128) return no("first shared FC width must be a multiple of 32, at most 128");
  if ((p.flat % 64) != 0) return no("flattened trunk size must be a multiple of 64");

  Bf16Path* b = new Bf16Path();
  TowerParams& tp = b->tp;
  tp.rows = d.env.rows;
  tp.cols = d.env.cols;
  tp.c0 = seq[0]->cout;
  tp.act0 = seq[0]->f.act;
  b->w0_off = seq[0]->f.wt;
  b->b0_off = seq[0]->f.bias;
  tp.n_layers = (int)seq.size() - 1;
  // buffer choreography: layer 0 writes buffer 0; plain layers ping-pong; a residual block's second conv
  // adds the block input (still sitting in the other buffer) and overwrites it in place.
  int cur = 0;
  size_t li = 1;
  int slot = 0;
  auto add = [&](const ConvLayer& c, int src, int dst, int res) {
    TowerLayer& L = tp.layers[slot];
    L.cin = c.cin; L.cout = c.cout; L.taps = c.kh * c.kw;
    L.ks = c.cin >= 64 ? 64 : c.cin;
    L.n_slabs = L.taps * c.cin / L.ks;
    L.slab_bytes = (uint32_t)L.ks * c.cout * 2;
    for (int sl = 0; sl < L.n_slabs && sl < kMaxSlabs; ++sl) {
      const int k0 = sl * L.ks;
      const int tap = k0 / c.cin, ch0 = k0 - tap * c.cin;
      const int shift = L.taps == 9 ? (tap / 3 - 1) * kPadW + (tap % 3 - 1) : 0;
      L.aoff[sl] = (uint32_t)((ch0 / 8) * kPlaneBytes + (kHalo + shift) * 16) >> 4;
    }
    L.src = src; L.dst = dst; L.res = res; L.act = c.f.act; L.bias_slot = slot;
    b->layer_wt.push_back(c.f.wt);
    b->layer_k.push_back(c.f.K);
    b->bias_off.push_back(c.f.bias);
    b->bias_w.push_back(c.cout);
    b->wstream_bytes += (size_t)c.f.K * c.cout * 2;
    ++slot;
  };
  const size_t n_plain = p.convs.size() + (p.has_res_proj ? 1 : 0);
  for (; li < n_plain; ++li) {
    add(*seq[li], cur, cur ^ 1, -1);
    cur ^= 1;
  }
  for (size_t r = 0; r < p.res.size(); ++r) {
    add(p.res[r].c1, cur, cur ^ 1, -1);
    add(p.res[r].c2, cur ^ 1, cur, cur);
  }
  if (with_proj) add(proj, cur, cur ^ 1, -1);
  if (n_plain == 1 && p.res.empty() && !with_proj) {
    delete b;
    return no("needs a tensor-core layer after the first convolution");
  }
  tp.layers[tp.n_layers - 1].dst = 2;
  const int c_last = tp.layers[tp.n_layers - 1].cout;
  b->n_kc = p.L * c_last / 8;
  tp.n_kc = b->n_kc;
  b->fc_n = fc.N;
  b->fc_flat = p.flat;
  b->fc_wt = fc.wt;
  b->max_batch = max_batch;
  b->conv_stream_bytes = b->wstream_bytes;
  if (p.transformer) {
    if (c_last != kTfD) {
      delete b;
      return no("last convolution must have 128 channels for the transformer path");
    }
    b->tf = true;
    TowerTfParams& t = b->ttp;
    t.t_layers = d.t_layers;
    t.heads = d.t_heads;
    t.ff_halves = d.t_ff / 128;
    tp.layers[tp.n_layers - 1].dst = 3;  // becomes the residual stream x in tensor memory
    int n = 0;
    auto slab = [&](uint32_t bytes) {
      if (n < kMaxRingSlabs) t.slab16[n] = (uint16_t)(bytes / 16);
      ++n;
    };
    for (int l = 0; l < tp.n_layers; ++l)
      for (int sl = 0; sl < tp.layers[l].n_slabs; ++sl) slab(tp.layers[l].slab_bytes);
    t.n_conv_slabs = n;
    // The encoder's slabs in the order the issuers consume them (tower_tf_kernel: qkv_gemm, deep_gemm, ring_slab). The operand
    // images are the same bytes whatever the cut -- K planes of 8 rows back to back, a GEMM's bias planes (two 8-wide K planes,
    // row 0 = the bias) right behind its weights, as bf16_add_prep_jobs writes them; no slab may exceed a 12 KB encoder stage.
    auto bias_bytes = [](int N) { return 2u * (uint32_t)N * 16u; };
    auto qkv_slabs = [&]() { slab(3 * kTfDh * 64 * 2); slab(3 * kTfDh * 64 * 2); slab(bias_bytes(3 * kTfDh)); };
    auto deep_slabs = [&](bool bias) { slab(128 * 48 * 2); slab(128 * 48 * 2); slab(128 * 32 * 2 + (bias ? bias_bytes(128) : 0u)); };
    for (int l = 0; l < d.t_layers; ++l) {
      qkv_slabs();                                                          // head 0
      for (int h = 0; h < d.t_heads; ++h) {
        if (h + 1 < d.t_heads) qkv_slabs();                                 // head h+1's projection is issued right behind head h's P V
        slab(kTfD * kTfDh * 2 + (h + 1 == d.t_heads ? bias_bytes(kTfD) : 0u));  // out_proj columns of head h; its bias with the last head
      }
      for (int f = 0; f < t.ff_halves; ++f) {
        if (f > 0) deep_slabs(f == 1);                                      // linear2, hidden half f-1 (+ its bias, once)
        deep_slabs(true);                                                   // linear1, hidden half f
      }
      deep_slabs(t.ff_halves == 1);
    }
    b->wstream_bytes = 0;
    for (int i = 0; i < n && i < kMaxRingSlabs; ++i) b->wstream_bytes += (size_t)t.slab16[i] * 16;
    b->pe_t_off = b->wstream_bytes;  // behind the slabs: the positional table, transposed for coalesced loads (PK_PE_T)
    b->wstream_bytes += (size_t)p.L * kTfD * sizeof(float);
    b->fc_bias_off = fc.bias;
    t.n_ring_slabs = n;
    if (n > kMaxRingSlabs) {
      delete b;
      return no("too many weight slabs per board for the fused tower + transformer kernel");
    }
  }

  cudaDeviceProp prop;
  int dev = 0;
  cudaGetDevice(&dev);
  cudaGetDeviceProperties(&prop, dev);
  b->sms = prop.multiProcessorCount;
  const int64_t m_tiles = (max_batch + 127) / 128;
  b->max_splits = 16;
  auto alloc = [&]() -> int {
    for (int i = 0; i < 2; ++i) {
      AT_CUDA_TRY(cudaMalloc(&b->wstream[i], b->wstream_bytes));
      AT_CUDA_TRY(cudaMalloc(&b->wfc[i], (size_t)p.flat * fc.N * 2));
      AT_CUDA_TRY(cudaMalloc(&b->bias[i], (size_t)kMaxLayers * 128 * sizeof(float)));
    }
    if (b->tf)
      for (int i = 0; i < 2; ++i) AT_CUDA_TRY(cudaMalloc(&b->fc_bias[i], (size_t)fc.N * sizeof(float)));
#define AT_TF_ATTR(ACT_, FMT_) \
  AT_CUDA_TRY(cudaFuncSetAttribute(tower_tf_kernel<ACT_, FMT_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tower_tf_smem_bytes()));
    AT_TF_ATTR(AT_ACT_RELU, FMT_BF16)
    AT_TF_ATTR(AT_ACT_RELU, FMT_F16)
    AT_TF_ATTR(AT_ACT_GELU, FMT_BF16)
    AT_TF_ATTR(AT_ACT_GELU, FMT_F16)
#undef AT_TF_ATTR
    AT_CUDA_TRY(cudaMalloc(&b->fc_a, (size_t)m_tiles * b->n_kc * 2048));
    AT_CUDA_TRY(cudaMemset(b->fc_a, 0, (size_t)m_tiles * b->n_kc * 2048));
    AT_CUDA_TRY(cudaMalloc(&b->partial, (size_t)(b->max_splits + 1) * max_batch * fc.N * sizeof(float)));
    AT_CUDA_TRY(cudaMalloc(&b->other, (size_t)max_batch * d.other_dim * sizeof(float)));
    AT_CUDA_TRY(cudaMalloc(&b->grid, (size_t)max_batch * d.env.rows * d.env.cols * sizeof(float)));
    AT_CUDA_TRY(cudaMalloc(&b->status, 4 * sizeof(uint32_t)));
    AT_CUDA_TRY(cudaMemset(b->status, 0, 4 * sizeof(uint32_t)));
    for (auto& e : b->ev) AT_CUDA_TRY(cudaEventCreate(&e));
#define AT_TOWER_ATTR(ACT_) \
  AT_CUDA_TRY(cudaFuncSetAttribute(tower_kernel<ACT_, FMT_BF16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tower_smem_bytes())); \
  AT_CUDA_TRY(cudaFuncSetAttribute(tower_kernel<ACT_, FMT_F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tower_smem_bytes()));
    AT_TOWER_ATTR(AT_ACT_RELU)
    AT_TOWER_ATTR(AT_ACT_GELU)
    AT_TOWER_ATTR(AT_ACT_SILU)
    AT_TOWER_ATTR(AT_ACT_TANH)
    AT_TOWER_ATTR(AT_ACT_SIGMOID)
#undef AT_TOWER_ATTR
    AT_CUDA_TRY(cudaFuncSetAttribute(fc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fc_smem_bytes()));
    return AT_OK;
  };
  const int rc = alloc();
  if (rc != AT_OK) {
    bf16_destroy(b);
    return rc;
  }
  *out = b;
  return AT_OK;
}

void bf16_destroy(Bf16Path* b) {
  if (!b) return;
  for (int i = 0; i < 2; ++i) {
    cudaFree(b->wstream[i]);
    cudaFree(b->wfc[i]);
    cudaFree(b->bias[i]);
    cudaFree(b->fc_bias[i]);
  }
  cudaFree(b->fc_a);
  cudaFree(b->partial);
  cudaFree(b->other);
  cudaFree(b->grid);
  cudaFree(b->status);
  for (auto& e : b->ev)
    if (e) cudaEventDestroy(e);
  delete b;
}

void bf16_prep_bases(const Bf16Path* b, int slot, PrepBases* bases) {
  bases->p[PB_WSTREAM] = b->wstream[slot];
  bases->p[PB_WFC] = b->wfc[slot];
  bases->p[PB_BIAS] = b->bias[slot];
  bases->p[PB_FCBIAS] = b->fc_bias[slot];
}

int bf16_packed_format(const Bf16Path* b, int slot) { return b->packed_fmt[slot]; }
void bf16_set_packed_format(Bf16Path* b, int slot, int fmt) { b->packed_fmt[slot] = fmt; }

// Phase-1 jobs of the weight-preparation table: every tower / encoder / FC matrix from the folded fp32 set into the
// core-matrix tiles the kernels stream ([K/8][N][8], in ring order) and the tower biases. For the encoder the LayerNorm
// weights are folded into the matrices that follow them and every bias becomes a 2-plane slab (see at_bf16_tf.inc); the
// order below is the order tower_tf_kernel's issuers consume the ring in.
void bf16_add_prep_jobs(const Bf16Path* b, const Plan& p, PrepList& list) {
  size_t off = 0;
  auto pack_b = [&](int64_t src, int K, int N, int base, size_t dst_bytes, int64_t row_scale, int period) {
    PrepJob j{};
    j.kind = PK_PACK_B; j.phase = 1;
    j.a[0] = src; j.a[1] = (int64_t)dst_bytes; j.a[2] = row_scale; j.i[0] = K; j.i[1] = N; j.i[2] = base; j.i[3] = period;
    list.add(j, (int64_t)K * N);
  };
  for (size_t l = 0; l < b->layer_wt.size(); ++l) {
    const int K = b->layer_k[l], N = b->bias_w[l];
    pack_b(b->layer_wt[l], K, N, PB_WSTREAM, off, -1, 1);
    off += (size_t)K * N * 2;
    PrepJob g{};
    g.kind = PK_GATHER_BIAS; g.phase = 1;
    g.a[0] = b->bias_off[l]; g.a[1] = (int64_t)l * 128; g.i[0] = N;
    list.add(g, 128);
  }
  if (!b->tf) {
    pack_b(b->fc_wt, b->fc_flat, b->fc_n, PB_WFC, 0, -1, 1);
    return;
  }
  // trunk FC over the encoder's output: final LayerNorm weight into the matrix rows (row k = token * 128 + channel),
  // its bias through the matrix into the FC bias
  pack_b(b->fc_wt, b->fc_flat, b->fc_n, PB_WFC, 0, p.lnf_w, kTfD);
  {
    PrepJob j{};
    j.kind = PK_FC_BIAS; j.phase = 1;
    j.a[0] = b->fc_bias_off; j.a[1] = b->fc_wt; j.a[2] = p.lnf_b; j.i[0] = b->fc_flat; j.i[1] = b->fc_n; j.i[2] = kTfD;
    list.add(j, 1);
  }
  const at_model_desc& d = p.d;
  size_t dst = b->conv_stream_bytes;
  struct Sel { int n_rows, k_count, k0, ld, seg_len, seg_base[3]; float seg0_scale; };
  auto pack = [&](int64_t src, const Sel& s, int64_t row_scale) {
    PrepJob j{};
    j.kind = PK_PACK_SEL; j.phase = 1;
    j.a[0] = src; j.a[1] = (int64_t)dst; j.a[2] = row_scale;
    j.i[0] = s.n_rows; j.i[1] = s.k_count; j.i[2] = s.k0; j.i[3] = s.ld; j.i[4] = s.seg_len;
    j.i[5] = s.seg_base[0]; j.i[6] = s.seg_base[1]; j.i[7] = s.seg_base[2];
    j.f[0] = s.seg0_scale;
    list.add(j, (int64_t)s.n_rows * s.k_count);
    dst += (size_t)s.n_rows * s.k_count * 2;
  };
  // bias slab of a GEMM with `n_rows` outputs: bias[col] (+ Wt[:, col] . beta when a LayerNorm precedes the GEMM)
  auto bias = [&](int64_t bias_src, int64_t wt, int64_t beta, const Sel& s) {
    PrepJob j{};
    j.kind = PK_BIAS_SLAB; j.phase = 1;
    j.a[0] = bias_src; j.a[1] = (int64_t)dst; j.a[2] = wt; j.a[3] = beta;
    j.i[0] = s.n_rows; j.i[1] = s.k_count; j.i[2] = s.ld; j.i[3] = s.seg_len;
    j.i[4] = s.seg_base[0]; j.i[5] = s.seg_base[1]; j.i[6] = s.seg_base[2];
    j.f[0] = s.seg0_scale;
    list.add(j, (int64_t)2 * s.n_rows * 8);
    dst += (size_t)2 * s.n_rows * 16;
  };
  const float qscale = 1.0f / sqrtf((float)kTfDh);
  const int halves = d.t_ff / 128;
  for (int l = 0; l < d.t_layers; ++l) {
    const EncoderLayer& e = p.enc[l];
    const Sel out_all{kTfD, kTfD, 0, kTfD, kTfD, {0, 0, 0}, 1.0f};
    auto out_proj = [&](int h) { pack(e.out.wt, Sel{kTfD, kTfDh, h * kTfDh, kTfD, kTfD, {0, 0, 0}, 1.0f}, -1); };
    auto qkv_proj = [&](int h) {
      const Sel qkv{3 * kTfDh, kTfD, 0, 3 * kTfD, kTfDh, {h * kTfDh, kTfD + h * kTfDh, 2 * kTfD + h * kTfDh}, qscale};
      pack(e.qkv.wt, qkv, e.ln1_w);           // two ring slabs (K halves), contiguous
      bias(e.qkv.bias, e.qkv.wt, e.ln1_b, qkv);
    };
    qkv_proj(0);
    for (int h = 0; h < d.t_heads; ++h) {
      if (h + 1 < d.t_heads) qkv_proj(h + 1);
      out_proj(h);
      if (h + 1 == d.t_heads) bias(e.out.bias, -1, -1, out_all);
    }
    auto linear2 = [&](int f) { pack(e.ff2.wt, Sel{kTfD, 128, f * 128, kTfD, kTfD, {0, 0, 0}, 1.0f}, -1); };
    for (int f = 0; f < halves; ++f) {
      if (f > 0) {
        linear2(f - 1);
        if (f == 1) bias(e.ff2.bias, -1, -1, out_all);
      }
      const Sel l1{128, kTfD, 0, d.t_ff, 128, {f * 128, 0, 0}, 1.0f};
      pack(e.ff1.wt, l1, e.ln2_w);
      bias(e.ff1.bias, e.ff1.wt, e.ln2_b, l1);
    }
    linear2(halves - 1);
    if (halves == 1) bias(e.ff2.bias, -1, -1, out_all);
    // this layer's norm1 belongs to its own q|k|v projection (handled above); the NEXT layer's norm1, or the encoder's
    // final norm, is applied to this layer's output and folded into whatever consumes it
  }
  PrepJob j{};
  j.kind = PK_PE_T; j.phase = 1;
  j.a[0] = p.pe; j.a[1] = (int64_t)b->pe_t_off;
  j.i[0] = p.L; j.i[1] = kTfD;
  list.add(j, (int64_t)p.L * kTfD);
}

int bf16_trunk_partials(Bf16Path* b, const Plan& p, const float* folded, int slot, const void* d_states, int64_t batch, int* n_splits,
                        cudaStream_t st, int64_t* launches, bool with_other) {
  const at_model_desc& d = p.d;
  const int fmt = b->packed_fmt[slot];
  if (fmt != FMT_BF16 && fmt != FMT_F16) return fail(AT_E_STATE, "tensor-core path: the weight set holds no packed operands yet");
  if (with_other) {
    // "other" features for the FC's last columns (the grid output of this launch is not used by the tower)
    AT_TRY(launch_features(d.env, 1, d.other_dim, d_states, batch, b->grid, b->other, st));
    ++*launches;
  }

  TowerParams tp = b->tp;
  tp.states = static_cast<const PackedState*>(d_states);
  tp.batch = batch;
  tp.wstream = b->wstream[slot];
  tp.bias = b->bias[slot];
  tp.w0 = folded + b->w0_off;
  tp.b0 = folded + b->b0_off;
  tp.fc_a = b->fc_a;
  tp.status = b->status;
#ifdef AT_EXPERIMENTS
  {
    const char* dbg = getenv("AT_TOWER_DEBUG");
    tp.debug = dbg ? (uint32_t)atoi(dbg) : 0u;
  }
#endif
  const int64_t n_pairs = (batch + kBoards - 1) / kBoards;
  const unsigned grid = (unsigned)(n_pairs < b->sms ? n_pairs : b->sms);
  // more pairs than CTAs: hand them out by counter (status word 2), so that CTAs which get their SM late take fewer
  tp.pair_counter = n_pairs > (int64_t)grid ? reinterpret_cast<int*>(b->status + 2) : nullptr;
  if (tp.pair_counter) AT_CUDA_TRY(cudaMemsetAsync(tp.pair_counter, 0, sizeof(int), st));
  if (b->timing) AT_CUDA_TRY(cudaEventRecord(b->ev[0], st));
  if (b->tf) {
    TowerTfParams t = b->ttp;
    t.conv = tp;
    t.pe = reinterpret_cast<const float*>(b->wstream[slot] + b->pe_t_off);
    t.n_pos = p.L;
#define AT_TF_LAUNCH(ACT_)                                                                                          \
  do {                                                                                                              \
    if (fmt == FMT_F16) tower_tf_kernel<ACT_, FMT_F16><<<grid, kTfThreads, tower_tf_smem_bytes(), st>>>(t);          \
    else tower_tf_kernel<ACT_, FMT_BF16><<<grid, kTfThreads, tower_tf_smem_bytes(), st>>>(t);                        \
  } while (0)
    if (tp.act0 == AT_ACT_RELU) AT_TF_LAUNCH(AT_ACT_RELU);
    else AT_TF_LAUNCH(AT_ACT_GELU);
#undef AT_TF_LAUNCH
  } else {
#define AT_TOWER_LAUNCH(ACT_)                                                                                       \
  do {                                                                                                              \
    if (fmt == FMT_F16) tower_kernel<ACT_, FMT_F16><<<grid, kTowerThreads, tower_smem_bytes(), st>>>(tp);            \
    else tower_kernel<ACT_, FMT_BF16><<<grid, kTowerThreads, tower_smem_bytes(), st>>>(tp);                          \
  } while (0)
    switch (tp.act0) {
      case AT_ACT_RELU: AT_TOWER_LAUNCH(AT_ACT_RELU); break;
      case AT_ACT_GELU: AT_TOWER_LAUNCH(AT_ACT_GELU); break;
      case AT_ACT_SILU: AT_TOWER_LAUNCH(AT_ACT_SILU); break;
      case AT_ACT_TANH: AT_TOWER_LAUNCH(AT_ACT_TANH); break;
      default: AT_TOWER_LAUNCH(AT_ACT_SIGMOID); break;
    }
#undef AT_TOWER_LAUNCH
  }
  AT_CUDA_TRY(cudaGetLastError());
  if (b->timing) AT_CUDA_TRY(cudaEventRecord(b->ev[1], st));
  ++*launches;

  // trunk FC: split-K. One CTA per SM (128 KB ring), so the launch runs in whole waves of `sms` CTAs: take the split count
  // that minimises waves x slabs per CTA (at batch 4096: 32 M-tiles x 9 splits = 288 CTAs = two waves of 27 slabs, where
  // 10 splits would spill 24 CTAs into a third wave).
  const int m_tiles = (int)((batch + 127) / 128);
  const int n_slabs = b->n_kc / 8;
  int splits = 1, per = n_slabs;
  {
    int64_t best = INT64_MAX;
    const int cap = b->max_splits < n_slabs ? b->max_splits : n_slabs;
    for (int sp = 1; sp <= cap; ++sp) {
      const int pr = (n_slabs + sp - 1) / sp;
      if ((n_slabs + pr - 1) / pr != sp) continue;  // same work as a smaller split count
      const int64_t waves = ((int64_t)m_tiles * sp + 2 * b->sms - 1) / (2 * b->sms);  // two CTAs per SM
      const int64_t cost = waves * (pr + 2);  // + ring fill / accumulator drain per CTA
      if (cost < best) {
        best = cost;
        splits = sp;
        per = pr;
      }
    }
  }
  FcParams fp{b->fc_a, b->wfc[slot], b->partial, batch, b->n_kc, b->fc_n, per, n_slabs, b->status, fmt};
  if (b->timing) AT_CUDA_TRY(cudaEventRecord(b->ev[2], st));
  fc_kernel<<<dim3((unsigned)m_tiles, (unsigned)splits), kFcThreads, fc_smem_bytes(), st>>>(fp);
  AT_CUDA_TRY(cudaGetLastError());
  if (b->timing) AT_CUDA_TRY(cudaEventRecord(b->ev[3], st));
  ++*launches;
  b->timed = b->timing;

  *n_splits = splits;
  return AT_OK;
}

const float* bf16_partial(const Bf16Path* b) { return b->partial; }
// trunk FC bias the head kernels must use: with the encoder's final LayerNorm folded in on the transformer path, else null
const float* bf16_fc_bias(const Bf16Path* b, int slot) { return b->tf ? b->fc_bias[slot] : nullptr; }
const float* bf16_other(const Bf16Path* b) { return b->other; }

// Unfused tail of the trunk FC: the "other" columns in fp32 as one more partial, then bias + BatchNorm (folded) + activation.
int bf16_finish_trunk(Bf16Path* b, const Plan& p, const float* folded, int slot, int splits, int64_t batch, float* d_trunk,
                      cudaStream_t st, int64_t* launches) {
  const at_model_desc& d = p.d;
  const Affine& fc = p.shared[0];
  GemmArgs o;
  o.A = b->other; o.lda = d.other_dim; o.M = batch; o.N = fc.N; o.K = d.other_dim;
  o.Wt = folded + fc.wt + (int64_t)p.flat * fc.N;
  o.C = b->partial + (int64_t)splits * batch * fc.N; o.ldc = fc.N;
  AT_TRY(launch_gemm(o, 1, st, launches));
  GemmArgs f;
  f.M = batch; f.N = fc.N; f.K = 1;
  f.partial = b->partial; f.bias = b->tf ? b->fc_bias[slot] : folded + fc.bias; f.act = fc.act;
  f.C = d_trunk; f.ldc = fc.N;
  AT_TRY(launch_splitk_finish(f, splits + 1, st, launches));
  return AT_OK;
}

const uint32_t* bf16_status_ptr(const Bf16Path* b) { return b->status; }
void bf16_set_timing(Bf16Path* b, bool on) { b->timing = on; if (!on) b->timed = false; }

int bf16_kernel_ms(Bf16Path* b, float* tower_ms, float* fc_ms) {
  if (!b || !b->timed) return fail(AT_E_STATE, "no timed tensor-core evaluation has run yet (at_engine_set_kernel_timing)");
  AT_CUDA_TRY(cudaEventSynchronize(b->ev[3]));
  AT_CUDA_TRY(cudaEventElapsedTime(tower_ms, b->ev[0], b->ev[1]));
  AT_CUDA_TRY(cudaEventElapsedTime(fc_ms, b->ev[2], b->ev[3]));
  return AT_OK;
}

// ---- small-batch evaluation whose scratch is NOT allocated: one static device array -----------------------------
//
// at_engine_evaluate_packed_static (include/at_b200.h) runs the same kernels as the bf16 path above, but the memory
// they use between themselves -- the "other" feature vectors, the tower's output tiles (the trunk FC's A operand) and
// the FC's split-K partial sums -- is carved out of the array below instead of the engine's cudaMalloc'ed buffers.
// Every GPU that loads this library carries the array (6 MiB) until the process exits, and every engine on that GPU
// shares it.
constexpr size_t kStaticScratchBytes = 6u << 20;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

namespace {
std::mutex g_static_mu;
unsigned char* g_static_addr[64] = {};  // per device ordinal: device address of g_static_scratch, resolved once
}  // namespace

// In host code `g_static_scratch` is only a handle for the runtime, not a device address: a kernel handed that name
// would fault (or, where the GPU can reach host memory, scribble over the wrong place). The address a kernel may use
// comes from cudaGetSymbolAddress, asked once per device with that device current.
static int static_scratch_base(unsigned char** out) {
  int dev = 0;
  AT_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return fail(AT_E_INVALID, "static scratch: device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_addr[dev]) {
    void* ptr = nullptr;
    AT_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_static_scratch));
    if (!ptr) return fail(AT_E_CUDA, "static scratch: cudaGetSymbolAddress returned a null address");
    g_static_addr[dev] = static_cast<unsigned char*>(ptr);
  }
  *out = g_static_addr[dev];
  return AT_OK;
}

int64_t bf16_static_max_batch() { return 128; }

namespace {
struct StaticCarve {
  size_t fc_a, partial, other, grid, total;  // byte sizes of the four regions, each rounded up to 256
};
StaticCarve static_carve(const Bf16Path* b, const Plan& p) {
  const size_t cap = (size_t)bf16_static_max_batch();
  auto up = [](size_t v) { return (v + 255) & ~size_t(255); };
  StaticCarve c;
  c.fc_a = up((size_t)b->n_kc * 2048);  // one 128-board tile of A-operand chunks
  c.partial = up((size_t)(b->max_splits + 1) * cap * b->fc_n * sizeof(float));
  c.other = up(cap * p.d.other_dim * sizeof(float));
  c.grid = up(cap * p.d.env.rows * p.d.env.cols * sizeof(float));
  c.total = c.fc_a + c.partial + c.other + c.grid;
  return c;
}
}  // namespace

// The two refusals of the static entry point, callable before anything is queued on the stream.
int bf16_static_check(const Bf16Path* b, const Plan& p, int64_t batch) {
  const int64_t cap = bf16_static_max_batch();
  if (batch > cap)
    return fail(AT_E_INVALID, "at_engine_evaluate_packed_static: batch " + std::to_string(batch) + " exceeds the static scratch limit of " +
                                  std::to_string(cap) + " states; use at_engine_evaluate_packed for larger batches");
  const size_t need = static_carve(b, p).total;
  if (need > kStaticScratchBytes)
    return fail(AT_E_UNSUPPORTED, "at_engine_evaluate_packed_static: this model needs " + std::to_string(need) +
                                      " bytes of scratch, the static array holds " + std::to_string(kStaticScratchBytes));
  return AT_OK;
}

int bf16_trunk_partials_static(Bf16Path* b, const Plan& p, const float* folded, int slot, const void* d_states, int64_t batch,
                               int* n_splits, const float** partial, const float** other, cudaStream_t st, int64_t* launches) {
  AT_TRY(bf16_static_check(b, p, batch));
  const StaticCarve c = static_carve(b, p);
  const size_t need_fc_a = c.fc_a, need_partial = c.partial, need_other = c.other;
  unsigned char* base = nullptr;
  AT_TRY(static_scratch_base(&base));
  // Nothing in the array has to start at zero: rows < batch of every region are fully rewritten by the kernels before
  // they are read, and rows >= batch (left over from earlier calls) only ever feed output rows that nobody reads.
  uint8_t* const keep_fc_a = b->fc_a;
  float* const keep_partial = b->partial;
  float* const keep_other = b->other;
  float* const keep_grid = b->grid;
  b->fc_a = base;
  b->partial = reinterpret_cast<float*>(base + need_fc_a);
  b->other = reinterpret_cast<float*>(base + need_fc_a + need_partial);
  b->grid = reinterpret_cast<float*>(base + need_fc_a + need_partial + need_other);
  *partial = b->partial;
  *other = b->other;
  const int rc = bf16_trunk_partials(b, p, folded, slot, d_states, batch, n_splits, st, launches);
  b->fc_a = keep_fc_a;
  b->partial = keep_partial;
  b->other = keep_other;
  b->grid = keep_grid;
  return rc;
}

#ifdef AT_EXPERIMENTS
int debug_read_timeline(unsigned long long* out, int n) {
  if (n > 32 * kTlEvents) n = 32 * kTlEvents;
  AT_CUDA_TRY(cudaDeviceSynchronize());
  AT_CUDA_TRY(cudaMemcpyFromSymbol(out, g_timeline, (size_t)n * sizeof(unsigned long long)));
  return n;
}
#endif

}  // namespace at
