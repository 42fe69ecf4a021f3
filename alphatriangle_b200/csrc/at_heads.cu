// Fused trunk-finish + policy / value heads + softmax epilogues (one kernel).
//
// For the common head shape -- one shared FC layer, one hidden layer per head
// (FC_DIMS_SHARED=[n], POLICY_HEAD_DIMS=[n], VALUE_HEAD_DIMS=[n], the reference's
// defaults: alphatriangle/config/model_config.py:41-52) -- this replaces nine small
// launches: split-K reduction of the trunk FC (+ the "other features" columns, + bias,
// BatchNorm folded, activation), both head MLPs (model.py:242-266), softmax over the
// actions, softmax over the value atoms dotted with the support, and the finite checks
// (network.py:157-162, 260-278). fp32 throughout. A block owns 32 boards; activations
// stay in shared memory; weights stream from L2 with float4 loads, 4 boards x 4 outputs
// of register tiling per thread.

#include <cuda_bf16.h>

#include "at_features.cuh"
#include "at_fp32.cuh"
#include "at_heads.cuh"
#include "at_prepjobs.cuh"

namespace at {

namespace {

constexpr int TB = 32;         // boards per block
constexpr int kThreads = 256;  // 8 warps: warp w owns boards 4w..4w+3 in the GEMM phases

// ReLU / identity inline, everything else (erf, tanh, exp based) behind one out-of-line call: this kernel applies the
// activation in several unrolled places and has to stay small enough for the instruction cache.
__device__ __noinline__ float act_slow(float x, int act) { return apply_act(x, act); }
__device__ __forceinline__ float act_small(float x, int act) {
  return act == AT_ACT_RELU ? fmaxf(x, 0.0f) : (act == AT_ACT_NONE ? x : act_slow(x, act));
}

constexpr int KC = 16;   // weight rows staged per chunk
constexpr int PW = 128;  // output columns per pass (32 lanes x float4)

// C_s[b][n] = act(bias[n] + sum_k A_s[b][k] * Wt[k][n]) for b < TB, n < N, in passes of 128 output columns.
// Weights go global -> registers -> shared in chunks of KC rows x 128 columns, double-buffered so the next chunk's
// loads are in flight while the current one is consumed. Warp w computes boards 4w..4w+3 (their A rows are warp-wide
// broadcast float4 loads), lane l the outputs 4l..4l+3 of the pass. Needs K % 4 == 0 and lda % 4 == 0. Kept out of
// line and free of big unrolls on purpose: the kernel calls it four times and must stay inside the instruction cache.
__device__ __noinline__ void block_gemm(const float* __restrict__ A_s, int lda, int K, const float* __restrict__ Wt,
                                        const float* __restrict__ bias, int N, float* __restrict__ C_s, int ldc, int act,
                                        float* __restrict__ Ws) {
  const int tid = threadIdx.x, warp = tid / 32, lane = tid % 32;
  const bool vec = (N & 3) == 0;
  const int n_chunks = (K + KC - 1) / KC;
  const int ld_k = tid / 32, ld_c = tid % 32;  // this thread stages rows ld_k and ld_k + 8, float4 column ld_c of every chunk
  const float* a_rows = A_s + (warp * 4) * lda;
  float4* Ws4 = reinterpret_cast<float4*>(Ws);

  for (int n_base = 0; n_base < N; n_base += PW) {
    const int n0 = n_base + lane * 4;       // this lane's four output columns
    const int ln = n_base + ld_c * 4;       // the four columns this thread stages
    float4 pre[2];
    auto fetch = [&](int c) {
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int k = c * KC + ld_k + r * 8;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (k < K && ln < N) {
          const float* src = Wt + (size_t)k * N + ln;
          if (vec) {
            v = __ldg(reinterpret_cast<const float4*>(src));
          } else {
            v.x = __ldg(src);
            v.y = ln + 1 < N ? __ldg(src + 1) : 0.f;
            v.z = ln + 2 < N ? __ldg(src + 2) : 0.f;
            v.w = ln + 3 < N ? __ldg(src + 3) : 0.f;
          }
        }
        pre[r] = v;
      }
    };
    auto stash = [&](int buf) {
      Ws4[(buf * KC + ld_k) * 32 + ld_c] = pre[0];
      Ws4[(buf * KC + ld_k + 8) * 32 + ld_c] = pre[1];
    };
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

    fetch(0);
    __syncthreads();  // every warp is done with whatever used Ws before
    stash(0);
    __syncthreads();
    for (int c = 0; c < n_chunks; ++c) {
      if (c + 1 < n_chunks) fetch(c + 1);
      const float4* W4 = Ws4 + (c & 1) * KC * 32 + lane;
      const int k_here = min(KC, K - c * KC);
#pragma unroll 1
      for (int kk = 0; kk < k_here; kk += 4) {
        float4 a[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const float4*>(a_rows + i * lda + c * KC + kk);
#pragma unroll
        for (int j4 = 0; j4 < 4; ++j4) {
          const float4 w = W4[(kk + j4) * 32];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float av = j4 == 0 ? a[i].x : (j4 == 1 ? a[i].y : (j4 == 2 ? a[i].z : a[i].w));
            acc[i][0] = fmaf(av, w.x, acc[i][0]);
            acc[i][1] = fmaf(av, w.y, acc[i][1]);
            acc[i][2] = fmaf(av, w.z, acc[i][2]);
            acc[i][3] = fmaf(av, w.w, acc[i][3]);
          }
        }
      }
      if (c + 1 < n_chunks) stash((c + 1) & 1);
      __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (n0 + j < N) {
        const float bv = bias[n0 + j];
#pragma unroll
        for (int i = 0; i < 4; ++i) C_s[(warp * 4 + i) * ldc + n0 + j] = act_small(acc[i][j] + bv, act);
      }
    }
  }
}

// Policy softmax, value-support softmax and expectation, finite flags: warp w finishes boards 4w..4w+3 of the block from
// the logits in shared memory (network.py:157-162, 260-278). Returns this warp's non-finite flag bits.
__device__ __forceinline__ uint32_t softmax_epilogue(const HeadsParams& p, const float* pl_s, int pl_pitch, const float* vl_s,
                                                     int vl_pitch, int64_t b0, int n_here) {
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  uint32_t bad = 0;
  for (int i = 0; i < 4; ++i) {
    const int b = warp * 4 + i;
    if (b >= n_here) break;
    {
      const float* lg = pl_s + b * pl_pitch;
      float mx = -INFINITY;
      bool fin = true;
      for (int a = lane; a < p.actions; a += 32) {
        const float v = lg[a];
        fin = fin && is_finite_f(v);
        mx = fmaxf(mx, v);
      }
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
      if (!__all_sync(0xffffffffu, fin)) bad |= AT_FLAG_NONFINITE_POLICY_LOGITS;
      float sum = 0.0f;
      for (int a = lane; a < p.actions; a += 32) sum += expf(lg[a] - mx);
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
      const float inv = 1.0f / sum;
      bool pfin = true;
      float* pr = p.probs + (b0 + b) * p.actions;
      for (int a = lane; a < p.actions; a += 32) {
        const float q = expf(lg[a] - mx) * inv;
        pfin = pfin && is_finite_f(q);
        pr[a] = q;
      }
      if (!__all_sync(0xffffffffu, pfin)) bad |= AT_FLAG_NONFINITE_POLICY_PROBS;
    }
    {
      const float* lg = vl_s + b * vl_pitch;
      float mx = -INFINITY;
      bool fin = true;
      for (int a = lane; a < p.atoms; a += 32) {
        const float v = lg[a];
        fin = fin && is_finite_f(v);
        mx = fmaxf(mx, v);
      }
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
      if (!__all_sync(0xffffffffu, fin)) bad |= AT_FLAG_NONFINITE_VALUE_LOGITS;
      float sum = 0.0f, dot = 0.0f;
      const float dz = (p.v_max - p.v_min) / (float)(p.atoms - 1);
      for (int a = lane; a < p.atoms; a += 32) {
        const float e = expf(lg[a] - mx);
        sum += e;
        dot = fmaf(e, support_atom(a, p.atoms, p.v_min, p.v_max, dz), dot);
      }
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, off);
        dot += __shfl_xor_sync(0xffffffffu, dot, off);
      }
      if (lane == 0) p.value[b0 + b] = dot / sum;
    }
  }
  return bad;
}

__global__ void __launch_bounds__(kThreads, 2) heads_fused_kernel(const HeadsParams p) {
  extern __shared__ __align__(16) float sm[];
  const int T = p.trunk_n, HP = p.pol_hidden, HV = p.val_hidden;
  float* t_s = sm;                    // [TB][T]
  float* hp_s = t_s + TB * T;         // [TB][HP]
  float* hv_s = hp_s + TB * HP;       // [TB][HV]
  float* pl_s = hv_s + TB * HV;       // [TB][actions]
  float* vl_s = pl_s + TB * p.actions;  // [TB][atoms]
  float* o_s = vl_s + TB * p.atoms;   // [TB][other_dim]
  float* w_s = sm + p.w_stage_off;    // [2][KC][128] weight staging (16-byte aligned)

  const int tid = threadIdx.x, warp = tid / 32, lane = tid % 32;
  const int64_t b0 = (int64_t)blockIdx.x * TB;
  const int n_here = (int)min((int64_t)TB, p.batch - b0);

  // ---- trunk: split-K partials + other-feature columns + bias, activation ----
  for (int i = tid; i < TB * p.other_dim; i += kThreads) {
    const int b = i / p.other_dim, k = i - b * p.other_dim;
    o_s[i] = b < n_here ? p.other[(b0 + b) * p.other_dim + k] : 0.0f;
  }
  __syncthreads();
  for (int i = tid; i < TB * T; i += kThreads) {
    const int b = i / T, n = i - b * T;
    float v = 0.0f;
    if (b < n_here) {
      v = p.trunk_bias[n];
      for (int z = 0; z < p.splits; ++z) v += p.partial[((int64_t)z * p.batch + b0 + b) * T + n];
      const float* wo = p.w_other + n;
      const float* ob = o_s + b * p.other_dim;
      for (int k = 0; k < p.other_dim; ++k) v = fmaf(ob[k], wo[(size_t)k * T], v);
      v = act_small(v, p.act);
    }
    t_s[i] = v;
  }
  __syncthreads();
  // ---- hidden layers of both heads ----
  block_gemm(t_s, T, T, p.wp1, p.bp1, HP, hp_s, HP, p.act, w_s);
  block_gemm(t_s, T, T, p.wv1, p.bv1, HV, hv_s, HV, p.act, w_s);
  // ---- logits ----
  block_gemm(hp_s, HP, HP, p.wp2, p.bp2, p.actions, pl_s, p.actions, AT_ACT_NONE, w_s);
  block_gemm(hv_s, HV, HV, p.wv2, p.bv2, p.atoms, vl_s, p.atoms, AT_ACT_NONE, w_s);
  __syncthreads();
  uint32_t bad = softmax_epilogue(p, pl_s, p.actions, vl_s, p.atoms, b0, n_here);
  if (blockIdx.x == 0 && tid == 0 && p.status && *p.status) bad |= 0x80000000u | ((*p.status & 0xFFu) << 8);
  if (bad && lane == 0) atomicOr(p.flags, bad);
}


// ------------------------------------------------------------------------------------------------------------------
// Tensor-core variant. Same block shape (32 boards, 8 warps), but the four GEMMs run on warp-level bf16 MMAs
// (mma.sync m16n8k16, fp32 accumulate): a warp keeps the whole 32 x K activation tile as A fragments in registers and
// walks its share of the 8-column output tiles; each tile's weights arrive as pre-packed B fragments, one coalesced
// 256-byte load per 16 x 8 block straight from L2 (the packed image is 170 KB, shared by every block), double-buffered
// against the MMAs. No weight staging through shared memory, no block-wide barriers inside a GEMM.
// ------------------------------------------------------------------------------------------------------------------

constexpr int kPadH = 8;  // bf16 padding of an activation row: pitch (K+8)*2 bytes keeps fragment loads / stores conflict-free

// FMT: operand format of activations and packed weights (FMT_BF16 / FMT_F16, at_prepjobs.cuh), fp32 accumulate either way
template <int FMT>
__device__ __forceinline__ void mma_16816(float (&d)[4], const uint32_t (&a)[4], const uint2 b) {
  if constexpr (FMT == FMT_F16)
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y));
  else
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y));
}

template <int FMT>
__device__ __forceinline__ uint32_t pack16x2(float lo, float hi) {
  uint32_t w;
  if constexpr (FMT == FMT_F16) asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(w) : "f"(hi), "f"(lo));
  else asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w) : "f"(hi), "f"(lo));
  return w;
}

// A fragments of a [32][K] bf16 tile (row pitch K + kPadH) for both 16-row halves: a[m][kt][0..3].
template <int KT>
__device__ __forceinline__ void load_a_frags(const __nv_bfloat16* A_s, int pitch, uint32_t (&a)[2][KT][4]) {
  const int lane = threadIdx.x % 32, g = lane >> 2, t = lane & 3;
#pragma unroll
  for (int m = 0; m < 2; ++m)
#pragma unroll
    for (int kt = 0; kt < KT; ++kt) {
      const __nv_bfloat16* r0 = A_s + (m * 16 + g) * pitch + kt * 16 + 2 * t;
      const __nv_bfloat16* r1 = r0 + 8 * pitch;
      a[m][kt][0] = *reinterpret_cast<const uint32_t*>(r0);
      a[m][kt][1] = *reinterpret_cast<const uint32_t*>(r1);
      a[m][kt][2] = *reinterpret_cast<const uint32_t*>(r0 + 8);
      a[m][kt][3] = *reinterpret_cast<const uint32_t*>(r1 + 8);
    }
}

// Tiles first, first + step, ... < n_tiles of one packed matrix: acc[m] = A[m] x B(tile); `done(tile, acc)` stores them.
template <int KT, int FMT, typename Done>
__device__ __forceinline__ void mma_tiles(const uint32_t (&a)[2][KT][4], const uint2* __restrict__ frags, int first, int step,
                                          int n_tiles, Done&& done) {
  const int lane = threadIdx.x % 32;
  uint2 b[KT], bn[KT];
  int tile = first;
  if (tile < n_tiles) {
#pragma unroll
    for (int kt = 0; kt < KT; ++kt) b[kt] = __ldg(frags + ((size_t)tile * KT + kt) * 32 + lane);
  }
  while (tile < n_tiles) {
    const int next = tile + step;
    if (next < n_tiles) {
#pragma unroll
      for (int kt = 0; kt < KT; ++kt) bn[kt] = __ldg(frags + ((size_t)next * KT + kt) * 32 + lane);
    }
    float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
    for (int kt = 0; kt < KT; ++kt) {
      mma_16816<FMT>(acc[0], a[0][kt], b[kt]);
      mma_16816<FMT>(acc[1], a[1][kt], b[kt]);
    }
    done(tile, acc);
#pragma unroll
    for (int kt = 0; kt < KT; ++kt) b[kt] = bn[kt];
    tile = next;
  }
}

struct FragLayout {
  int p1_tiles, v1_tiles, p2_tiles, v2_tiles;  // 8-column tiles per matrix
  size_t p1, v1, p2, v2, total;                // uint2 offsets
};

__host__ __device__ inline FragLayout frag_layout(const HeadsParams& p) {
  FragLayout L;
  const int kt1 = p.trunk_n / 16;
  L.p1_tiles = p.pol_hidden / 8; L.v1_tiles = p.val_hidden / 8;
  L.p2_tiles = (p.actions + 7) / 8; L.v2_tiles = (p.atoms + 7) / 8;
  L.p1 = 0;
  L.v1 = L.p1 + (size_t)L.p1_tiles * kt1 * 32;
  L.p2 = L.v1 + (size_t)L.v1_tiles * kt1 * 32;
  L.v2 = L.p2 + (size_t)L.p2_tiles * (p.pol_hidden / 16) * 32;
  L.total = L.v2 + (size_t)L.v2_tiles * (p.val_hidden / 16) * 32;
  return L;
}


// Softmax epilogues with the row held in registers: one pass over shared memory, one exp per element. A warp finishes
// BPW boards (warp w: boards BPW*w ...). Needs actions <= 32 * kActPerLane and atoms <= 64; the launcher checks.
constexpr int kActPerLane = 12;

template <int BPW>
__device__ __forceinline__ uint32_t softmax_epilogue_regs(const HeadsParams& p, const float* pl_s, int pl_pitch, const float* vl_s,
                                                          int vl_pitch, int64_t b0, int n_here) {
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  constexpr float kLog2e = 1.4426950408889634f;
  uint32_t bad = 0;
  const float dz = (p.v_max - p.v_min) / (float)(p.atoms - 1);
#pragma unroll 1
  for (int i = 0; i < BPW; ++i) {
    const int b = warp * BPW + i;
    if (b >= n_here) break;
    {
      const float* lg = pl_s + b * pl_pitch;
      float v[kActPerLane];
      float mx = -INFINITY;
      bool fin = true;
#pragma unroll
      for (int j = 0; j < kActPerLane; ++j) {
        const int a = lane + 32 * j;
        v[j] = a < p.actions ? lg[a] : -INFINITY;
        fin = fin && (a >= p.actions || is_finite_f(v[j]));
        mx = fmaxf(mx, v[j]);
      }
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
      if (!__all_sync(0xffffffffu, fin)) bad |= AT_FLAG_NONFINITE_POLICY_LOGITS;
      float sum = 0.0f;
      const float mxl = mx * kLog2e;
#pragma unroll
      for (int j = 0; j < kActPerLane; ++j) {
        v[j] = exp2f(fmaf(v[j], kLog2e, -mxl));  // padding lanes: exp2(-inf) = 0
        sum += v[j];
      }
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
      const float inv = 1.0f / sum;
      bool pfin = true;
      float* pr = p.probs + (b0 + b) * p.actions;
#pragma unroll
      for (int j = 0; j < kActPerLane; ++j) {
        const int a = lane + 32 * j;
        const float q = v[j] * inv;
        if (a < p.actions) {
          pfin = pfin && is_finite_f(q);
          pr[a] = q;
        }
      }
      if (!__all_sync(0xffffffffu, pfin)) bad |= AT_FLAG_NONFINITE_POLICY_PROBS;
    }
    {
      const float* lg = vl_s + b * vl_pitch;
      const int a0 = lane, a1 = lane + 32;
      const float l0 = a0 < p.atoms ? lg[a0] : -INFINITY, l1 = a1 < p.atoms ? lg[a1] : -INFINITY;
      const bool fin = (a0 >= p.atoms || is_finite_f(l0)) && (a1 >= p.atoms || is_finite_f(l1));
      float mx = fmaxf(l0, l1);
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
      if (!__all_sync(0xffffffffu, fin)) bad |= AT_FLAG_NONFINITE_VALUE_LOGITS;
      const float e0 = a0 < p.atoms ? expf(l0 - mx) : 0.0f, e1 = a1 < p.atoms ? expf(l1 - mx) : 0.0f;
      float sum = e0 + e1;
      float dot = fmaf(e1, support_atom(min(a1, p.atoms - 1), p.atoms, p.v_min, p.v_max, dz),
                       e0 * support_atom(min(a0, p.atoms - 1), p.atoms, p.v_min, p.v_max, dz));
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, off);
        dot += __shfl_xor_sync(0xffffffffu, dot, off);
      }
      if (lane == 0) p.value[b0 + b] = dot / sum;
    }
  }
  return bad;
}

constexpr int kMmaWarps = 16;  // one block per SM at 128 registers: twice the warps of the CUDA-core kernel to hide L2 latency
constexpr int kMmaThreads = kMmaWarps * 32;

template <int KT, int FMT>
__global__ void __launch_bounds__(kMmaThreads, 1) heads_mma_kernel(const HeadsParams p) {
  constexpr int K = KT * 16, PITCH = K + kPadH;
  extern __shared__ __align__(16) float sm[];
  const int AP = (p.actions + 7) & ~7, VP = (p.atoms + 7) & ~7;
  float* pl_s = sm;                     // [TB][AP] policy logits
  float* vl_s = pl_s + TB * AP;         // [TB][VP] value logits
  float* o_s = vl_s + TB * VP;          // [TB][other_dim]
  __nv_bfloat16* t_s = reinterpret_cast<__nv_bfloat16*>(o_s + ((TB * p.other_dim + 3) & ~3));  // [TB][PITCH] trunk
  __nv_bfloat16* hp_s = t_s + TB * PITCH;                                                      // policy hidden
  __nv_bfloat16* hv_s = hp_s + TB * PITCH;                                                     // value hidden

  const int tid = threadIdx.x, warp = tid / 32, lane = tid % 32, g = lane >> 2, t = lane & 3;
  const int64_t b0 = (int64_t)blockIdx.x * TB;
  const int n_here = (int)min((int64_t)TB, p.batch - b0);
  const FragLayout L = frag_layout(p);

  // ---- trunk: split-K partials + other-feature columns + bias, activation -> bf16 ----
  if (p.states) {
    // other features straight from the packed records: stage the block's 32 states (the trunk tile is still unused),
    // then one thread per board (extractor.py:48-139, shared with the feature kernel)
    PackedState* st_s = reinterpret_cast<PackedState*>(t_s);
    const uint4* src = reinterpret_cast<const uint4*>(p.states + b0);
    for (int i = tid; i < n_here * (AT_STATE_BYTES / 16); i += kMmaThreads) reinterpret_cast<uint4*>(st_s)[i] = __ldg(src + i);
    __syncthreads();
    if (tid < TB) {
      float* o = o_s + tid * p.other_dim;
      if (tid < n_here) {
        other_features_bits(st_s[tid], p.env_rows, p.env_cols, p.env_slots, p.other_dim, o);
      } else {
        for (int d = 0; d < p.other_dim; ++d) o[d] = 0.0f;
      }
    }
  } else {
    for (int i = tid; i < TB * p.other_dim; i += kMmaThreads) {
      const int b = i / p.other_dim;
      o_s[i] = b < n_here ? p.other[b0 * p.other_dim + i] : 0.0f;
    }
  }
  __syncthreads();
  {
    constexpr int K4 = K / 4;
    const size_t split_stride = (size_t)p.batch * K4;  // float4 units between two splits
    for (int i = tid; i < TB * K4; i += kMmaThreads) {
      const int b = i / K4, n = (i - b * K4) * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (b < n_here) {
        v = __ldg(reinterpret_cast<const float4*>(p.trunk_bias + n));
        const float4* pp = reinterpret_cast<const float4*>(p.partial) + (size_t)(b0 + b) * K4 + (n >> 2);
#pragma unroll 8
        for (int z = 0; z < p.splits; ++z) {  // independent L2 loads: keep many in flight
          const float4 q = __ldg(pp + z * split_stride);
          v.x += q.x; v.y += q.y; v.z += q.z; v.w += q.w;
        }
        const float4* wo = reinterpret_cast<const float4*>(p.w_other + n);
        const float* ob = o_s + b * p.other_dim;
#pragma unroll 10
        for (int k = 0; k < p.other_dim; ++k) {
          const float4 w = __ldg(wo + (size_t)k * K4);
          const float o = ob[k];
          v.x = fmaf(o, w.x, v.x); v.y = fmaf(o, w.y, v.y); v.z = fmaf(o, w.z, v.z); v.w = fmaf(o, w.w, v.w);
        }
        v.x = act_small(v.x, p.act); v.y = act_small(v.y, p.act); v.z = act_small(v.z, p.act); v.w = act_small(v.w, p.act);
      }
      *reinterpret_cast<uint2*>(t_s + b * PITCH + n) = make_uint2(pack16x2<FMT>(v.x, v.y), pack16x2<FMT>(v.z, v.w));
    }
  }
  __syncthreads();

  uint32_t a[2][KT][4];
  // ---- hidden layers of both heads: output tiles [policy hidden | value hidden], all from the trunk tile ----
  load_a_frags<KT>(t_s, PITCH, a);
  auto hidden_done = [&](__nv_bfloat16* H_s, const float* bias) {
    return [=](int tile, float (&acc)[2][4]) {
      const int n = tile * 8 + 2 * t;
      const float b0v = __ldg(bias + n), b1v = __ldg(bias + n + 1);
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        __nv_bfloat16* r0 = H_s + (m * 16 + g) * PITCH + n;
        *reinterpret_cast<uint32_t*>(r0) = pack16x2<FMT>(act_small(acc[m][0] + b0v, p.act), act_small(acc[m][1] + b1v, p.act));
        *reinterpret_cast<uint32_t*>(r0 + 8 * PITCH) = pack16x2<FMT>(act_small(acc[m][2] + b0v, p.act), act_small(acc[m][3] + b1v, p.act));
      }
    };
  };
  mma_tiles<KT, FMT>(a, p.frags + L.p1, warp, kMmaWarps, L.p1_tiles, hidden_done(hp_s, p.bp1));
  mma_tiles<KT, FMT>(a, p.frags + L.v1, warp, kMmaWarps, L.v1_tiles, hidden_done(hv_s, p.bv1));
  __syncthreads();

  // ---- logits (fp32, + bias) into shared memory ----
  auto logits_done = [&](float* out_s, int pitch, const float* bias, int N) {
    return [=](int tile, float (&acc)[2][4]) {
      const int n = tile * 8 + 2 * t;
      const float b0v = n < N ? __ldg(bias + n) : 0.0f, b1v = n + 1 < N ? __ldg(bias + n + 1) : 0.0f;
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        float* r0 = out_s + (m * 16 + g) * pitch + n;
        *reinterpret_cast<float2*>(r0) = make_float2(acc[m][0] + b0v, acc[m][1] + b1v);
        *reinterpret_cast<float2*>(r0 + 8 * pitch) = make_float2(acc[m][2] + b0v, acc[m][3] + b1v);
      }
    };
  };
  load_a_frags<KT>(hp_s, PITCH, a);
  mma_tiles<KT, FMT>(a, p.frags + L.p2, warp, kMmaWarps, L.p2_tiles, logits_done(pl_s, AP, p.bp2, p.actions));
  // value tiles go to the warps from the top down: those have the fewest policy tiles
  if (kMmaWarps - 1 - warp < L.v2_tiles) {
    load_a_frags<KT>(hv_s, PITCH, a);
    mma_tiles<KT, FMT>(a, p.frags + L.v2, kMmaWarps - 1 - warp, kMmaWarps, L.v2_tiles, logits_done(vl_s, VP, p.bv2, p.atoms));
  }
  __syncthreads();

  uint32_t bad = softmax_epilogue_regs<TB / kMmaWarps>(p, pl_s, AP, vl_s, VP, b0, n_here);
  if (blockIdx.x == 0 && tid == 0 && p.status && *p.status) bad |= 0x80000000u | ((*p.status & 0xFFu) << 8);
  if (bad && lane == 0) atomicOr(p.flags, bad);
}

}  // namespace

static size_t act_floats(const HeadsParams& p) {
  const size_t n = (size_t)TB * (p.trunk_n + p.pol_hidden + p.val_hidden + p.actions + p.atoms + p.other_dim);
  return (n + 3) & ~size_t(3);
}

size_t heads_fused_smem(const HeadsParams& p) { return (act_floats(p) + (size_t)2 * KC * PW) * sizeof(float); }

bool heads_fused_supported(const HeadsParams& p) {
  const bool dims4 = (p.trunk_n % 4 == 0) && (p.pol_hidden % 4 == 0) && (p.val_hidden % 4 == 0);
  return p.trunk_n > 0 && dims4 && heads_fused_smem(p) <= 200 * 1024;
}

int launch_heads_fused(const HeadsParams& p, cudaStream_t stream, int64_t* launches) {
  if (p.batch <= 0) return AT_OK;
  HeadsParams q = p;
  q.w_stage_off = (int)act_floats(p);
  const size_t smem = heads_fused_smem(p);
  AT_SMEM_OPT_IN_ONCE(heads_fused_kernel, 200 * 1024);
  heads_fused_kernel<<<(unsigned)((q.batch + TB - 1) / TB), kThreads, smem, stream>>>(q);
  AT_CUDA_TRY(cudaGetLastError());
  if (launches) ++*launches;
  return AT_OK;
}

// ---- tensor-core variant: host side ----

static int mma_kt(const HeadsParams& p) {
  const int K = p.trunk_n;
  if (p.pol_hidden != K || p.val_hidden != K) return 0;
  return (K == 128) ? 8 : (K == 64 ? 4 : 0);
}

static size_t heads_mma_smem(const HeadsParams& p) {
  const int AP = (p.actions + 7) & ~7, VP = (p.atoms + 7) & ~7, K = p.trunk_n;
  return ((size_t)TB * (AP + VP) + (((size_t)TB * p.other_dim + 3) & ~size_t(3))) * sizeof(float) + (size_t)3 * TB * (K + kPadH) * 2;
}

bool heads_mma_supported(const HeadsParams& p) {
  return mma_kt(p) != 0 && p.actions > 0 && p.actions <= 32 * kActPerLane && p.atoms > 1 && p.atoms <= 64 &&
         heads_mma_smem(p) <= 100 * 1024;
}

size_t heads_frag_count(const HeadsParams& p) { return frag_layout(p).total; }

// Phase-1 jobs of the weight-preparation table: the four head matrices (folded fp32 Wt at float offsets src[0..3] =
// policy hidden, value hidden, policy logits, value logits) into mma.sync B-fragment order.
void heads_add_prep_jobs(const HeadsParams& p, const int64_t src[4], PrepList& list) {
  const FragLayout L = frag_layout(p);
  auto pack = [&](int64_t from, int K, int N, int tiles, size_t off) {
    PrepJob j{};
    j.kind = PK_PACK_FRAG; j.phase = 1;
    j.a[0] = from; j.a[1] = (int64_t)off; j.i[0] = K; j.i[1] = N; j.i[2] = tiles;
    list.add(j, (int64_t)tiles * (K / 16) * 32 * 4);
  };
  pack(src[0], p.trunk_n, p.pol_hidden, L.p1_tiles, L.p1);
  pack(src[1], p.trunk_n, p.val_hidden, L.v1_tiles, L.v1);
  pack(src[2], p.pol_hidden, p.actions, L.p2_tiles, L.p2);
  pack(src[3], p.val_hidden, p.atoms, L.v2_tiles, L.v2);
}

int launch_heads_mma(const HeadsParams& p, cudaStream_t stream, int64_t* launches) {
  if (p.batch <= 0) return AT_OK;
  if (!p.frags) return fail(AT_E_INVALID, "launch_heads_mma: packed head weights missing");
  if (p.states && (p.env_cols > 32 || p.env_rows * p.env_cols > AT_MAX_CELLS || (size_t)TB * AT_STATE_BYTES > (size_t)TB * (p.trunk_n + kPadH) * 2))
    return fail(AT_E_INVALID, "launch_heads_mma: in-kernel other features need a board of at most 32 columns");
  if (!p.states && !p.other) return fail(AT_E_INVALID, "launch_heads_mma: neither other features nor packed states given");
  const size_t smem = heads_mma_smem(p);
  const unsigned grid = (unsigned)((p.batch + TB - 1) / TB);
#define AT_HEADS_LAUNCH(KT_, FMT_)                                                                                              \
  do {                                                                                                                          \
    AT_SMEM_OPT_IN_ONCE((heads_mma_kernel<KT_, FMT_>), 100 * 1024);                                                             \
    heads_mma_kernel<KT_, FMT_><<<grid, kMmaThreads, smem, stream>>>(p);                                                        \
  } while (0)
  if (mma_kt(p) == 8) {
    if (p.fmt == FMT_F16) AT_HEADS_LAUNCH(8, FMT_F16);
    else AT_HEADS_LAUNCH(8, FMT_BF16);
  } else {
    if (p.fmt == FMT_F16) AT_HEADS_LAUNCH(4, FMT_F16);
    else AT_HEADS_LAUNCH(4, FMT_BF16);
  }
#undef AT_HEADS_LAUNCH
  AT_CUDA_TRY(cudaGetLastError());
  if (launches) ++*launches;
  return AT_OK;
}

}  // namespace at
