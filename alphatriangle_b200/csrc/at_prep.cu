// Weight preparation on the device: reference fp32 state_dict image -> folded set.
//
// Eval-mode BatchNorm (y = (x - mean) / sqrt(var + 1e-5) * gamma + beta, applied
// after a bias-free conv or a biased Linear: alphatriangle/nn/model.py:24-38,
// 232-266) is folded into the preceding affine map; weights are transposed to
// [K,N], permuted to the NHWC / tap-major K order the kernels use, and padded
// with zero rows up to Kpad.

#include "at_engine.cuh"

namespace at {

namespace {

constexpr float kBnEps = 1e-5f;  // nn.BatchNorm{1,2}d default, never overridden by the reference

__global__ void __launch_bounds__(256) fold_affine_kernel(const float* __restrict__ img, int64_t src_w, int64_t src_b, int64_t bn_g,
                                                          int64_t bn_b, int64_t bn_m, int64_t bn_v, int K, int Kpad, int N,
                                                          int perm_cols, int perm_inner, float* __restrict__ wt,
                                                          float* __restrict__ bias) {
  const int64_t total = (int64_t)Kpad * N;
  const int perm_outer = perm_inner > 0 ? perm_cols / perm_inner : 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int row = (int)(i / N), n = (int)(i - (int64_t)row * N);
    float v = 0.0f;
    if (row < K) {
      int k_ref = row;
      if (row < perm_cols) {
        const int inner = row / perm_outer, outer = row - inner * perm_outer;  // row = inner*outer_count + outer
        k_ref = outer * perm_inner + inner;
      }
      v = img[src_w + (int64_t)n * K + k_ref];
      if (bn_g >= 0) v *= img[bn_g + n] / sqrtf(img[bn_v + n] + kBnEps);
    }
    wt[i] = v;
  }
  for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < N; n += gridDim.x * blockDim.x) {
    float bv = src_b >= 0 ? img[src_b + n] : 0.0f;
    if (bn_g >= 0) {
      const float scale = img[bn_g + n] / sqrtf(img[bn_v + n] + kBnEps);
      bv = (bv - img[bn_m + n]) * scale + img[bn_b + n];
    }
    bias[n] = bv;
  }
}

int fold_one(const Affine& f, const float* img, float* folded, cudaStream_t stream) {
  const int64_t total = (int64_t)f.Kpad * f.N;
  int64_t blocks = (total + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  fold_affine_kernel<<<(unsigned)blocks, 256, 0, stream>>>(img, f.src_w, f.src_b, f.bn_g, f.bn_b, f.bn_m, f.bn_v, f.K, f.Kpad, f.N,
                                                          f.perm_cols, f.perm_inner, folded + f.wt, folded + f.bias);
  AT_CUDA_TRY(cudaGetLastError());
  return AT_OK;
}

int copy_vec(const float* img, int64_t src, float* folded, int64_t dst, int64_t n, cudaStream_t stream) {
  AT_CUDA_TRY(cudaMemcpyAsync(folded + dst, img + src, (size_t)n * sizeof(float), cudaMemcpyDeviceToDevice, stream));
  return AT_OK;
}

}  // namespace

int fold_weights(const Plan& p, const float* img, float* folded, cudaStream_t stream) {
  for (const ConvLayer& c : p.convs) AT_TRY(fold_one(c.f, img, folded, stream));
  if (p.has_res_proj) AT_TRY(fold_one(p.res_proj.f, img, folded, stream));
  for (const ResBlock& r : p.res) {
    AT_TRY(fold_one(r.c1.f, img, folded, stream));
    AT_TRY(fold_one(r.c2.f, img, folded, stream));
  }
  if (p.transformer) {
    const int dm = p.d.t_dim;
    if (p.has_input_proj) AT_TRY(fold_one(p.input_proj, img, folded, stream));
    AT_TRY(copy_vec(img, p.src_pe, folded, p.pe, (int64_t)p.L * dm, stream));  // rows 0..L-1 of pe[5000,1,d]
    for (const EncoderLayer& e : p.enc) {
      AT_TRY(fold_one(e.qkv, img, folded, stream));
      AT_TRY(fold_one(e.out, img, folded, stream));
      AT_TRY(fold_one(e.ff1, img, folded, stream));
      AT_TRY(fold_one(e.ff2, img, folded, stream));
      AT_TRY(copy_vec(img, e.src_ln1_w, folded, e.ln1_w, dm, stream));
      AT_TRY(copy_vec(img, e.src_ln1_b, folded, e.ln1_b, dm, stream));
      AT_TRY(copy_vec(img, e.src_ln2_w, folded, e.ln2_w, dm, stream));
      AT_TRY(copy_vec(img, e.src_ln2_b, folded, e.ln2_b, dm, stream));
    }
    AT_TRY(copy_vec(img, p.src_lnf_w, folded, p.lnf_w, dm, stream));
    AT_TRY(copy_vec(img, p.src_lnf_b, folded, p.lnf_b, dm, stream));
  }
  for (const Affine& f : p.shared) AT_TRY(fold_one(f, img, folded, stream));
  for (const Affine& f : p.policy) AT_TRY(fold_one(f, img, folded, stream));
  for (const Affine& f : p.value) AT_TRY(fold_one(f, img, folded, stream));
  return AT_OK;
}

}  // namespace at
