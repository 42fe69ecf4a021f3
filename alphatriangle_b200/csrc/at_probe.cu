// Bring-up probe: one CTA runs a chain of tcgen05.mma (M=128, bf16 -> fp32 in
// TMEM) over operand images the caller laid out byte-for-byte, with every
// shared-memory-descriptor field supplied by the caller. tests/ uses it to pin
// the descriptor conventions (core-matrix strides, row-shifted start addresses,
// 128-byte swizzle) that the convolution tower depends on.

#include "at_common.cuh"
#include "at_umma.cuh"

namespace at {

namespace {

using namespace umma;

struct ProbeArgs {
  const uint8_t* a_image;
  const uint8_t* b_image;
  uint32_t a_bytes, b_bytes;
  int n, k;
  uint32_t a_start, a_lbo, a_sbo, a_kstep;
  uint32_t b_start, b_lbo, b_sbo, b_kstep;
  uint32_t layout;
  float* out;
  uint32_t* status;
};

__global__ void __launch_bounds__(128, 1) probe_kernel(const ProbeArgs p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[2];
  __shared__ uint32_t tmem_base_s;

  const int tid = threadIdx.x, warp = tid / 32;
  const uint32_t a_smem = smem_u32(smem);
  const uint32_t b_smem = a_smem + ((p.a_bytes + 1023u) & ~1023u);
  const uint32_t bar_load = smem_u32(&bars[0]), bar_mma = smem_u32(&bars[1]);
  uint32_t cols = 32;
  while (cols < (uint32_t)p.n) cols <<= 1;

  if (tid == 0) {
    mbar_init(bar_load, 1);
    mbar_init(bar_mma, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(smem_u32(&tmem_base_s), cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;

  bool ok = true;
  if (tid == 0) {
    mbar_arrive_expect_tx(bar_load, p.a_bytes + p.b_bytes);
    bulk_g2s(a_smem, p.a_image, p.a_bytes, bar_load);
    bulk_g2s(b_smem, p.b_image, p.b_bytes, bar_load);
    ok = mbar_wait(bar_load, 0);
    if (ok) {
      tc_fence_after();
      const uint32_t idesc = idesc_bf16(128, (uint32_t)p.n);
      for (int j = 0; j < p.k / 16; ++j) {
        const uint64_t da = smem_desc(a_smem + p.a_start + j * p.a_kstep, p.a_lbo, p.a_sbo, p.layout);
        const uint64_t db = smem_desc(b_smem + p.b_start + j * p.b_kstep, p.b_lbo, p.b_sbo, p.layout);
        mma_bf16(tmem_base, da, db, idesc, j > 0);
      }
    }
    mma_commit(bar_mma);  // with no MMA outstanding this arrives immediately, so nobody hangs
    if (!ok) atomicOr(p.status, 1u);
  }
  __syncwarp();
  if (!mbar_wait(bar_mma, 0)) {
    atomicOr(p.status, 2u);
  } else {
    tc_fence_after();
    for (int c0 = 0; c0 < p.n; c0 += 32) {
      uint32_t r[32];
      tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, r);
      tmem_ld_wait();
      float* dst = p.out + (size_t)tid * p.n + c0;
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (c0 + j < p.n) dst[j] = __uint_as_float(r[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, cols);
}

}  // namespace

}  // namespace at

extern "C" int at_probe_umma(const void* d_a_image, int64_t a_bytes, const void* d_b_image, int64_t b_bytes, int n, int k,
                             uint32_t a_start_off, uint32_t a_lbo, uint32_t a_sbo, uint32_t a_kstep, uint32_t b_start_off,
                             uint32_t b_lbo, uint32_t b_sbo, uint32_t b_kstep, int layout_type, float* d_out, uint32_t* d_status,
                             void* stream) {
  using namespace at;
  if (!d_a_image || !d_b_image || !d_out || !d_status) return fail(AT_E_INVALID, "at_probe_umma: null pointer");
  if (n < 16 || n > 256 || n % 16 != 0) return fail(AT_E_INVALID, "at_probe_umma: n must be a multiple of 16 in [16,256]");
  if (k < 16 || k % 16 != 0) return fail(AT_E_INVALID, "at_probe_umma: k must be a positive multiple of 16");
  if (a_bytes <= 0 || b_bytes <= 0 || a_bytes % 16 || b_bytes % 16) return fail(AT_E_INVALID, "at_probe_umma: image sizes must be positive multiples of 16");
  const size_t smem = (((size_t)a_bytes + 1023) & ~(size_t)1023) + (size_t)b_bytes + 1024;
  if (smem > 200 * 1024) return fail(AT_E_INVALID, "at_probe_umma: images do not fit in shared memory");
  ProbeArgs p{static_cast<const uint8_t*>(d_a_image), static_cast<const uint8_t*>(d_b_image), (uint32_t)a_bytes, (uint32_t)b_bytes,
              n, k, a_start_off, a_lbo, a_sbo, a_kstep, b_start_off, b_lbo, b_sbo, b_kstep, (uint32_t)layout_type, d_out, d_status};
  AT_CUDA_TRY(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  probe_kernel<<<1, 128, smem, static_cast<cudaStream_t>(stream)>>>(p);
  AT_CUDA_TRY(cudaGetLastError());
  return AT_OK;
}
