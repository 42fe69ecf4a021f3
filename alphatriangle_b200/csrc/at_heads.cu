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

#include "at_fp32.cuh"
#include "at_heads.cuh"

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
  // ---- softmax epilogues: warp w finishes boards 4w..4w+3 ----
  uint32_t bad = 0;
  for (int i = 0; i < 4; ++i) {
    const int b = warp * 4 + i;
    if (b >= n_here) break;
    {
      const float* lg = pl_s + b * p.actions;
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
      const float* lg = vl_s + b * p.atoms;
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
  AT_CUDA_TRY(cudaFuncSetAttribute(heads_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  heads_fused_kernel<<<(unsigned)((q.batch + TB - 1) / TB), kThreads, smem, stream>>>(q);
  AT_CUDA_TRY(cudaGetLastError());
  if (launches) ++*launches;
  return AT_OK;
}

}  // namespace at
