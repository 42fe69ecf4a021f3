// fp32 CUDA-core path: the configuration-generic implementation of
// AlphaTriangleNet.forward (alphatriangle/nn/model.py:268-297) used for parity
// (max-abs 1e-4 against the reference's fp32 CPU path) and for every layer the
// tcgen05 path does not specialise. Activations are NHWC / token-major fp32.
//
//   gemm_kernel      C = epi(A * Wt)   A dense [M,K] or gathered on the fly from an
//                    image (implicit-GEMM convolution, model.py:15-38), Wt = [K,N]
//   splitk_finish    sum of split-K partials + the same epilogue
//   layernorm_kernel nn.LayerNorm over the last dim (one warp per row)
//   attention_kernel softmax(Q K^T / sqrt(dh)) V per (board, head), online softmax
//   heads_kernel     policy softmax, value softmax . support, finite flags
//                    (alphatriangle/nn/network.py:157-162,260-278), one warp per state

#include "at_fp32.cuh"

namespace at {

namespace {

constexpr int BM = 64, BN = 64, BK = 16;
constexpr int kGemmThreads = 256;

__device__ __forceinline__ void epilogue_store(const GemmArgs& g, int64_t m, int n, float v) {
  if (g.bias) v += g.bias[n];
  if (g.res) v += g.res[m * g.ldr + n];
  v = apply_act(v, g.act);
  if (g.post) v += g.post[(m % g.post_period) * (int64_t)g.N + n];
  const int64_t b = m / g.out_rows_per_batch, p = m - b * g.out_rows_per_batch;
  g.C[b * g.out_batch_stride + p * g.ldc + n] = v;
}

template <bool CONV, bool VEC>
__global__ void __launch_bounds__(kGemmThreads) gemm_kernel(const GemmArgs g) {
  __shared__ __align__(16) float As[BK][BM + 4];
  __shared__ __align__(16) float Bs[BK][BN];

  const int tid = threadIdx.x;
  const int tx = tid % 16, ty = tid / 16;
  const int64_t m0 = (int64_t)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;

  // split-K range (multiples of BK)
  const int k_steps_total = (g.K + BK - 1) / BK;
  const int steps_per = (k_steps_total + gridDim.z - 1) / gridDim.z;
  const int step_begin = blockIdx.z * steps_per;
  const int step_end = min(k_steps_total, step_begin + steps_per);

  // A loader: row a_m, four consecutive k starting at a_k
  const int a_m = tid / 4, a_k = (tid % 4) * 4;
  const int64_t am = m0 + a_m;
  const bool a_row_ok = am < g.M;
  int64_t a_base = 0;
  int oy = 0, ox = 0;
  if (a_row_ok) {
    if (CONV) {
      const int hw = g.out_h * g.out_w;
      const int64_t b = am / hw;
      const int p = (int)(am - b * hw);
      oy = p / g.out_w;
      ox = p - oy * g.out_w;
      a_base = b * g.in_sb;
    } else {
      a_base = am * g.lda;
    }
  }
  // B loader: row b_k, four consecutive n starting at b_n
  const int b_k = tid / 16, b_n = (tid % 16) * 4;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

  for (int step = step_begin; step < step_end; ++step) {
    const int k0 = step * BK;
    // ---- A tile ----
    float av[4] = {0.f, 0.f, 0.f, 0.f};
    if (a_row_ok) {
      const int k = k0 + a_k;
      if (CONV) {
        if (VEC) {
          if (k < g.K) {
            const int tap = k / g.cin, c = k - tap * g.cin;
            const int ky = tap / g.kw, kx = tap - ky * g.kw;
            const int iy = oy * g.sh - g.ph + ky, ix = ox * g.sw - g.pw + kx;
            if (iy >= 0 && iy < g.in_h && ix >= 0 && ix < g.in_w) {
              const float4 v = *reinterpret_cast<const float4*>(g.A + a_base + (int64_t)iy * g.in_sy + (int64_t)ix * g.in_sx + c);
              av[0] = v.x; av[1] = v.y; av[2] = v.z; av[3] = v.w;
            }
          }
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int kk = k + j;
            if (kk < g.K) {
              const int tap = kk / g.cin, c = kk - tap * g.cin;
              const int ky = tap / g.kw, kx = tap - ky * g.kw;
              const int iy = oy * g.sh - g.ph + ky, ix = ox * g.sw - g.pw + kx;
              if (iy >= 0 && iy < g.in_h && ix >= 0 && ix < g.in_w)
                av[j] = g.A[a_base + (int64_t)iy * g.in_sy + (int64_t)ix * g.in_sx + (int64_t)c * g.in_sc];
            }
          }
        }
      } else {
        if (VEC) {
          if (k < g.K) {
            const float4 v = *reinterpret_cast<const float4*>(g.A + a_base + k);
            av[0] = v.x; av[1] = v.y; av[2] = v.z; av[3] = v.w;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (k + j < g.K) av[j] = g.A[a_base + k + j];
        }
      }
    }
    // ---- B tile ----
    float bv[4] = {0.f, 0.f, 0.f, 0.f};
    {
      const int k = k0 + b_k, n = n0 + b_n;
      if (k < g.K) {
        if (VEC) {
          if (n < g.N) {
            const float4 v = *reinterpret_cast<const float4*>(g.Wt + (int64_t)k * g.N + n);
            bv[0] = v.x; bv[1] = v.y; bv[2] = v.z; bv[3] = v.w;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (n + j < g.N) bv[j] = g.Wt[(int64_t)k * g.N + n + j];
        }
      }
    }
    __syncthreads();  // previous step's reads are done
#pragma unroll
    for (int j = 0; j < 4; ++j) As[a_k + j][a_m] = av[j];
    *reinterpret_cast<float4*>(&Bs[b_k][b_n]) = make_float4(bv[0], bv[1], bv[2], bv[3]);
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 b4 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float a[4] = {a4.x, a4.y, a4.z, a4.w};
      const float b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
  }

  if (gridDim.z > 1) {
    float* part = g.partial + (int64_t)blockIdx.z * g.M * g.N;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t m = m0 + ty * 4 + i;
      if (m >= g.M) continue;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = n0 + tx * 4 + j;
        if (n < g.N) part[m * g.N + n] = acc[i][j];
      }
    }
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t m = m0 + ty * 4 + i;
      if (m >= g.M) continue;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = n0 + tx * 4 + j;
        if (n < g.N) epilogue_store(g, m, n, acc[i][j]);
      }
    }
  }
}

__global__ void __launch_bounds__(256) splitk_finish_kernel(const GemmArgs g, int splits) {
  const int64_t total = g.M * g.N;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    float v = 0.0f;
    for (int z = 0; z < splits; ++z) v += g.partial[(int64_t)z * total + i];
    const int64_t m = i / g.N;
    epilogue_store(g, m, (int)(i - m * g.N), v);
  }
}

// nn.LayerNorm(d, eps=1e-5): biased variance, y = (x - mean) / sqrt(var + eps) * w + b
__global__ void __launch_bounds__(256) layernorm_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                        const float* __restrict__ b, float* __restrict__ y, int64_t rows, int d,
                                                        float eps, int64_t out_rows_per_batch, int64_t out_batch_stride, int64_t ldy) {
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x / 32) + warp;
  if (row >= rows) return;
  const float* xr = x + row * d;
  float s = 0.0f;
  for (int i = lane; i < d; i += 32) s += xr[i];
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
  const float mean = s / (float)d;
  float v = 0.0f;
  for (int i = lane; i < d; i += 32) {
    const float t = xr[i] - mean;
    v = fmaf(t, t, v);
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  const float rstd = 1.0f / sqrtf(v / (float)d + eps);
  const int64_t bb = row / out_rows_per_batch, p = row - bb * out_rows_per_batch;
  float* yr = y + bb * out_batch_stride + p * ldy;
  for (int i = lane; i < d; i += 32) yr[i] = (xr[i] - mean) * rstd * w[i] + b[i];
}

// One block per (board, head); K and V of that head staged in shared memory, one
// thread per query row, single pass with a running max (torch MHA math path:
// scores scaled by 1/sqrt(dh), softmax over all L keys, no mask).
template <int DH>
__global__ void __launch_bounds__(128) attention_kernel(const float* __restrict__ qkv, float* __restrict__ out, int L, int d,
                                                        int heads) {
  extern __shared__ float sm[];
  float* Ks = sm;            // [L][DH]
  float* Vs = sm + L * DH;   // [L][DH]
  const int64_t b = blockIdx.x / heads;
  const int h = blockIdx.x % heads;
  const float* base = qkv + b * (int64_t)L * 3 * d;
  for (int i = threadIdx.x; i < L * DH; i += blockDim.x) {
    const int l = i / DH, j = i - l * DH;
    Ks[i] = base[(int64_t)l * 3 * d + d + h * DH + j];
    Vs[i] = base[(int64_t)l * 3 * d + 2 * d + h * DH + j];
  }
  __syncthreads();
  const float scale = 1.0f / sqrtf((float)DH);
  for (int q = threadIdx.x; q < L; q += blockDim.x) {
    float qv[DH], o[DH];
#pragma unroll
    for (int j = 0; j < DH; ++j) {
      qv[j] = base[(int64_t)q * 3 * d + h * DH + j] * scale;
      o[j] = 0.0f;
    }
    float mx = -INFINITY, den = 0.0f;
    for (int l = 0; l < L; ++l) {
      float sc = 0.0f;
#pragma unroll
      for (int j = 0; j < DH; ++j) sc = fmaf(qv[j], Ks[l * DH + j], sc);
      const float nmx = fmaxf(mx, sc);
      const float corr = expf(mx - nmx);  // exp(-inf) = 0 on the first key
      const float p = expf(sc - nmx);
      den = den * corr + p;
#pragma unroll
      for (int j = 0; j < DH; ++j) o[j] = fmaf(p, Vs[l * DH + j], o[j] * corr);
      mx = nmx;
    }
    const float inv = 1.0f / den;
    float* dst = out + (b * L + q) * (int64_t)d + h * DH;
#pragma unroll
    for (int j = 0; j < DH; ++j) dst[j] = o[j] * inv;
  }
}

// One warp per state: softmax over the action logits, softmax over the value
// atoms dotted with the support, and the reference's isfinite checks as flag bits.
__global__ void __launch_bounds__(256) heads_kernel(const float* __restrict__ policy_logits, const float* __restrict__ value_logits,
                                                    int64_t batch, int actions, int atoms, float v_min, float v_max,
                                                    float* __restrict__ probs, float* __restrict__ value, uint32_t* __restrict__ flags) {
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int64_t b = (int64_t)blockIdx.x * (blockDim.x / 32) + warp;
  if (b >= batch) return;
  uint32_t bad = 0;
  {
    const float* lg = policy_logits + b * actions;
    float mx = -INFINITY;
    bool fin = true;
    for (int i = lane; i < actions; i += 32) {
      const float v = lg[i];
      fin = fin && is_finite_f(v);
      mx = fmaxf(mx, v);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
    if (!__all_sync(0xffffffffu, fin)) bad |= AT_FLAG_NONFINITE_POLICY_LOGITS;
    float sum = 0.0f;
    for (int i = lane; i < actions; i += 32) sum += expf(lg[i] - mx);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
    const float inv = 1.0f / sum;
    bool pfin = true;
    float* pr = probs + b * actions;
    for (int i = lane; i < actions; i += 32) {
      const float p = expf(lg[i] - mx) * inv;
      pfin = pfin && is_finite_f(p);
      pr[i] = p;
    }
    if (!__all_sync(0xffffffffu, pfin)) bad |= AT_FLAG_NONFINITE_POLICY_PROBS;
  }
  {
    const float* lg = value_logits + b * atoms;
    float mx = -INFINITY;
    bool fin = true;
    for (int i = lane; i < atoms; i += 32) {
      const float v = lg[i];
      fin = fin && is_finite_f(v);
      mx = fmaxf(mx, v);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
    if (!__all_sync(0xffffffffu, fin)) bad |= AT_FLAG_NONFINITE_VALUE_LOGITS;
    float sum = 0.0f, dot = 0.0f;
    const float dz = (v_max - v_min) / (float)(atoms - 1);
    for (int i = lane; i < atoms; i += 32) {
      const float e = expf(lg[i] - mx);
      sum += e;
      dot = fmaf(e, support_atom(i, atoms, v_min, v_max, dz), dot);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      sum += __shfl_xor_sync(0xffffffffu, sum, off);
      dot += __shfl_xor_sync(0xffffffffu, dot, off);
    }
    if (lane == 0) value[b] = dot / sum;
  }
  if (bad && lane == 0) atomicOr(flags, bad);
}

__global__ void __launch_bounds__(256) finite_check_kernel(const float* __restrict__ x, int64_t rows, int64_t row_elems, int64_t sb,
                                                           uint32_t bit, uint32_t* __restrict__ flags) {
  bool fin = true;
  const int64_t total = rows * row_elems;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / row_elems;
    fin = fin && is_finite_f(x[r * sb + (i - r * row_elems)]);
  }
  if (!__all_sync(0xffffffffu, fin) && (threadIdx.x % 32) == 0) atomicOr(flags, bit);
}

__global__ void __launch_bounds__(256) copy_rows_kernel(const float* __restrict__ src, int64_t rows, int cols, float* __restrict__ dst,
                                                        int64_t ld_dst) {
  const int64_t total = rows * cols;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / cols;
    dst[r * ld_dst + (i - r * cols)] = src[i];
  }
}

// pipeline-timeout bits of the tcgen05 kernels -> bit 31 (+ detail in bits 8..15) of the caller's flag word
__global__ void merge_status_kernel(const uint32_t* __restrict__ status, uint32_t* __restrict__ flags) {
  if (threadIdx.x == 0 && *status) atomicOr(flags, 0x80000000u | ((*status & 0xFFu) << 8));
}

inline unsigned grid_for(int64_t total, int threads, int cap = 148 * 16) {
  int64_t blocks = (total + threads - 1) / threads;
  if (blocks < 1) blocks = 1;
  if (blocks > cap) blocks = cap;
  return (unsigned)blocks;
}

}  // namespace

int launch_gemm(const GemmArgs& g_in, int splits, cudaStream_t stream, int64_t* launches) {
  GemmArgs g = g_in;
  if (g.M <= 0 || g.N <= 0 || g.K <= 0) return fail(AT_E_INVALID, "gemm: empty problem");
  if (g.out_rows_per_batch <= 0) {
    g.out_rows_per_batch = g.M;
    g.out_batch_stride = 0;
  }
  if (g.post && g.post_period <= 0) return fail(AT_E_INVALID, "gemm: post table without a period");
  if (!g.post) g.post_period = 1;
  if (splits < 1) splits = 1;
  if (splits > 1 && !g.partial) return fail(AT_E_INVALID, "gemm: split-K without a partial buffer");
  const bool conv = g.conv != 0;
  const bool w_vec = (g.N % 4 == 0) && ((reinterpret_cast<uintptr_t>(g.Wt) & 15u) == 0);
  bool a_vec;
  if (conv)
    a_vec = g.in_sc == 1 && g.cin % 4 == 0 && g.in_sb % 4 == 0 && g.in_sy % 4 == 0 && g.in_sx % 4 == 0 &&
            (reinterpret_cast<uintptr_t>(g.A) & 15u) == 0;
  else
    a_vec = g.lda % 4 == 0 && g.K % 4 == 0 && (reinterpret_cast<uintptr_t>(g.A) & 15u) == 0;
  const bool vec = w_vec && a_vec;
  dim3 grid((unsigned)((g.M + BM - 1) / BM), (unsigned)((g.N + BN - 1) / BN), (unsigned)splits);
  if (conv) {
    if (vec) gemm_kernel<true, true><<<grid, kGemmThreads, 0, stream>>>(g);
    else gemm_kernel<true, false><<<grid, kGemmThreads, 0, stream>>>(g);
  } else {
    if (vec) gemm_kernel<false, true><<<grid, kGemmThreads, 0, stream>>>(g);
    else gemm_kernel<false, false><<<grid, kGemmThreads, 0, stream>>>(g);
  }
  AT_CUDA_TRY(cudaGetLastError());
  if (launches) ++*launches;
  if (splits > 1) {
    splitk_finish_kernel<<<grid_for(g.M * g.N, 256), 256, 0, stream>>>(g, splits);
    AT_CUDA_TRY(cudaGetLastError());
    if (launches) ++*launches;
  }
  return AT_OK;
}

int launch_splitk_finish(const GemmArgs& g_in, int splits, cudaStream_t stream, int64_t* launches) {
  GemmArgs g = g_in;
  if (g.M <= 0 || g.N <= 0 || splits < 1 || !g.partial || !g.C) return fail(AT_E_INVALID, "splitk_finish: bad arguments");
  if (g.out_rows_per_batch <= 0) {
    g.out_rows_per_batch = g.M;
    g.out_batch_stride = 0;
  }
  if (!g.post) g.post_period = 1;
  splitk_finish_kernel<<<grid_for(g.M * g.N, 256), 256, 0, stream>>>(g, splits);
  AT_CUDA_TRY(cudaGetLastError());
  if (launches) ++*launches;
  return AT_OK;
}

int launch_merge_status(const uint32_t* status, uint32_t* flags, cudaStream_t stream, int64_t* launches) {
  merge_status_kernel<<<1, 32, 0, stream>>>(status, flags);
  AT_CUDA_TRY(cudaGetLastError());
  if (launches) ++*launches;
  return AT_OK;
}

int launch_layernorm(const float* x, const float* w, const float* b, float* y, int64_t rows, int d, float eps,
                     int64_t out_rows_per_batch, int64_t out_batch_stride, int64_t ldy, cudaStream_t stream, int64_t* launches) {
  if (rows <= 0) return AT_OK;
  if (out_rows_per_batch <= 0) {
    out_rows_per_batch = rows;
    out_batch_stride = 0;
  }
  const int warps = 8;
  layernorm_kernel<<<(unsigned)((rows + warps - 1) / warps), warps * 32, 0, stream>>>(x, w, b, y, rows, d, eps, out_rows_per_batch,
                                                                                      out_batch_stride, ldy);
  AT_CUDA_TRY(cudaGetLastError());
  if (launches) ++*launches;
  return AT_OK;
}

int launch_attention(const float* qkv, float* out, int64_t boards, int L, int d, int heads, cudaStream_t stream, int64_t* launches) {
  if (boards <= 0) return AT_OK;
  const int dh = d / heads;
  const size_t smem = (size_t)2 * L * dh * sizeof(float);
  const unsigned grid = (unsigned)(boards * heads);
#define AT_ATTN_CASE(DH)                                                                                             \
  case DH:                                                                                                           \
    if (smem > 48 * 1024)                                                                                            \
      AT_CUDA_TRY(cudaFuncSetAttribute(attention_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    attention_kernel<DH><<<grid, 128, smem, stream>>>(qkv, out, L, d, heads);                                        \
    break;
  switch (dh) {
    AT_ATTN_CASE(4)
    AT_ATTN_CASE(8)
    AT_ATTN_CASE(16)
    AT_ATTN_CASE(32)
    AT_ATTN_CASE(64)
    default:
      return fail(AT_E_UNSUPPORTED, "attention: head dim must be one of 4, 8, 16, 32, 64 (got " + std::to_string(dh) + ")");
  }
#undef AT_ATTN_CASE
  AT_CUDA_TRY(cudaGetLastError());
  if (launches) ++*launches;
  return AT_OK;
}

int launch_heads(const float* policy_logits, const float* value_logits, int64_t batch, int actions, int atoms, float v_min, float v_max,
                 float* probs, float* value, uint32_t* flags, cudaStream_t stream, int64_t* launches) {
  if (batch <= 0) return AT_OK;
  const int warps = 8;
  heads_kernel<<<(unsigned)((batch + warps - 1) / warps), warps * 32, 0, stream>>>(policy_logits, value_logits, batch, actions, atoms,
                                                                                  v_min, v_max, probs, value, flags);
  AT_CUDA_TRY(cudaGetLastError());
  if (launches) ++*launches;
  return AT_OK;
}

int launch_finite_check(const float* x, int64_t rows, int64_t row_elems, int64_t row_stride, uint32_t bit, uint32_t* flags,
                        cudaStream_t stream, int64_t* launches) {
  if (rows <= 0 || row_elems <= 0) return AT_OK;
  finite_check_kernel<<<grid_for(rows * row_elems, 256), 256, 0, stream>>>(x, rows, row_elems, row_stride, bit, flags);
  AT_CUDA_TRY(cudaGetLastError());
  if (launches) ++*launches;
  return AT_OK;
}

int launch_copy_rows(const float* src, int64_t rows, int cols, float* dst, int64_t ld_dst, cudaStream_t stream, int64_t* launches) {
  if (rows <= 0 || cols <= 0) return AT_OK;
  copy_rows_kernel<<<grid_for(rows * cols, 256), 256, 0, stream>>>(src, rows, cols, dst, ld_dst);
  AT_CUDA_TRY(cudaGetLastError());
  if (launches) ++*launches;
  return AT_OK;
}

}  // namespace at
