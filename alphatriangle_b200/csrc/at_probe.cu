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

// A operand from tensor memory: every thread packs its row's bf16 pairs into TMEM columns, then the MMA chain reads A
// from there (B still through a shared-memory descriptor). Pins the TMEM A layout the attention P*V product relies on.
struct ProbeTsArgs {
  const uint16_t* a_rows;  // [128][k] bf16 bit patterns, row-major
  const uint8_t* b_image;
  uint32_t b_bytes;
  int n, k;
  int pack_mode;  // 0: element 2c in the low half of column c; 1: in the high half
  uint32_t a_col, a_cols_per_kstep;
  uint32_t b_start, b_lbo, b_sbo, b_kstep;
  float* out;
  uint32_t* status;
};

__global__ void __launch_bounds__(128, 1) probe_ts_kernel(const ProbeTsArgs p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[2];
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid / 32;
  const uint32_t b_smem = smem_u32(smem);
  const uint32_t bar_load = smem_u32(&bars[0]), bar_mma = smem_u32(&bars[1]);
  if (tid == 0) {
    mbar_init(bar_load, 1);
    mbar_init(bar_mma, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(smem_u32(&tmem_base_s), 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
  for (int k0 = 0; k0 < p.k; k0 += 16) {
    uint32_t r[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const uint32_t e0 = p.a_rows[(size_t)tid * p.k + k0 + 2 * c], e1 = p.a_rows[(size_t)tid * p.k + k0 + 2 * c + 1];
      r[c] = p.pack_mode == 0 ? (e0 | (e1 << 16)) : (e1 | (e0 << 16));
    }
    tmem_st8(lane_addr + p.a_col + (uint32_t)(k0 / 16) * p.a_cols_per_kstep, r);
  }
  tmem_st_wait();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (tid == 0) {
    mbar_arrive_expect_tx(bar_load, p.b_bytes);
    bulk_g2s(b_smem, p.b_image, p.b_bytes, bar_load);
    const bool ok = mbar_wait(bar_load, 0);
    if (ok) {
      tc_fence_after();
      const uint32_t idesc = idesc_bf16(128, (uint32_t)p.n);
      for (int j = 0; j < p.k / 16; ++j) {
        const uint64_t db = smem_desc(b_smem + p.b_start + j * p.b_kstep, p.b_lbo, p.b_sbo, LAYOUT_NONE);
        mma_bf16_ts(tmem_base, tmem_base + p.a_col + (uint32_t)j * p.a_cols_per_kstep, db, idesc, j > 0);
      }
    } else {
      atomicOr(p.status, 1u);
    }
    mma_commit(bar_mma);
  }
  __syncwarp();
  if (!mbar_wait(bar_mma, 0)) {
    atomicOr(p.status, 2u);
  } else {
    tc_fence_after();
    for (int c0 = 0; c0 < p.n; c0 += 32) {
      uint32_t r[32];
      tmem_ld32(lane_addr + (uint32_t)c0, r);
      tmem_ld_wait();
      float* dst = p.out + (size_t)tid * p.n + c0;
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (c0 + j < p.n) dst[j] = __uint_as_float(r[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 512);
}

}  // namespace

}  // namespace at

extern "C" int at_probe_umma_ts(const void* d_a_rows, const void* d_b_image, int64_t b_bytes, int n, int k, int pack_mode, uint32_t a_col,
                                uint32_t a_cols_per_kstep, uint32_t b_start_off, uint32_t b_lbo, uint32_t b_sbo, uint32_t b_kstep,
                                float* d_out, uint32_t* d_status, void* stream) {
  using namespace at;
  if (!d_a_rows || !d_b_image || !d_out || !d_status) return fail(AT_E_INVALID, "at_probe_umma_ts: null pointer");
  if (n < 16 || n > 256 || n % 16 != 0 || k < 16 || k % 16 != 0) return fail(AT_E_INVALID, "at_probe_umma_ts: bad n / k");
  if (b_bytes <= 0 || b_bytes % 16 || b_bytes > 200 * 1024) return fail(AT_E_INVALID, "at_probe_umma_ts: bad image size");
  if (a_col < (uint32_t)n || a_col + (uint32_t)(k / 16) * a_cols_per_kstep > 512) return fail(AT_E_INVALID, "at_probe_umma_ts: A columns out of range");
  ProbeTsArgs p{static_cast<const uint16_t*>(d_a_rows), static_cast<const uint8_t*>(d_b_image), (uint32_t)b_bytes, n, k, pack_mode,
                a_col, a_cols_per_kstep, b_start_off, b_lbo, b_sbo, b_kstep, d_out, d_status};
  const size_t smem = (size_t)b_bytes + 1024;
  AT_CUDA_TRY(cudaFuncSetAttribute(probe_ts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  probe_ts_kernel<<<1, 128, smem, static_cast<cudaStream_t>(stream)>>>(p);
  AT_CUDA_TRY(cudaGetLastError());
  return AT_OK;
}

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
