// Launchers of the fp32 CUDA-core kernels (at_fp32.cu).
#pragma once

#include "at_common.cuh"

namespace at {

// C[b*out_batch_stride + p*ldc + n] = act(sum_k A[m,k] * Wt[k,n] + bias[n] + res[m*ldr+n]) + post[(m % post_period)*N + n]
// with m = b*out_rows_per_batch + p. A is either dense (lda) or gathered from an
// image with element strides (in_sb,in_sy,in_sx,in_sc): k = (ky*kw + kx)*cin + c.
struct GemmArgs {
  const float* A = nullptr;
  const float* Wt = nullptr;  // [K,N]
  const float* bias = nullptr;
  const float* res = nullptr;
  const float* post = nullptr;
  float* C = nullptr;
  float* partial = nullptr;  // split-K scratch [splits,M,N]
  int64_t M = 0;
  int N = 0, K = 0;
  int64_t lda = 0, ldc = 0, ldr = 0;
  int64_t out_rows_per_batch = 0, out_batch_stride = 0;
  int post_period = 0;
  int act = AT_ACT_NONE;
  // convolution gather
  int conv = 0;
  int cin = 0, kh = 0, kw = 0, sh = 1, sw = 1, ph = 0, pw = 0;
  int in_h = 0, in_w = 0, out_h = 0, out_w = 0;
  int64_t in_sb = 0, in_sy = 0, in_sx = 0, in_sc = 0;
};

// torch.linspace(v_min, v_max, atoms) element i, computed the way ATen's CPU kernel does
__host__ __device__ __forceinline__ float support_atom(int i, int atoms, float v_min, float v_max, float dz) {
  return i < atoms / 2 ? v_min + dz * (float)i : v_max - dz * (float)(atoms - 1 - i);
}

int launch_gemm(const GemmArgs& g, int splits, cudaStream_t stream, int64_t* launches);
int launch_splitk_finish(const GemmArgs& g, int splits, cudaStream_t stream, int64_t* launches);
int launch_merge_status(const uint32_t* status, uint32_t* flags, cudaStream_t stream, int64_t* launches);
int launch_layernorm(const float* x, const float* w, const float* b, float* y, int64_t rows, int d, float eps,
                     int64_t out_rows_per_batch, int64_t out_batch_stride, int64_t ldy, cudaStream_t stream, int64_t* launches);
int launch_attention(const float* qkv, float* out, int64_t boards, int L, int d, int heads, cudaStream_t stream, int64_t* launches);
int launch_heads(const float* policy_logits, const float* value_logits, int64_t batch, int actions, int atoms, float v_min, float v_max,
                 float* probs, float* value, uint32_t* flags, cudaStream_t stream, int64_t* launches);
int launch_finite_check(const float* x, int64_t rows, int64_t row_elems, int64_t row_stride, uint32_t bit, uint32_t* flags,
                        cudaStream_t stream, int64_t* launches);
int launch_copy_rows(const float* src, int64_t rows, int cols, float* dst, int64_t ld_dst, cudaStream_t stream, int64_t* launches);

}  // namespace at
