// Bring-up probe: one CTA runs a chain of tcgen05.mma (M=128, bf16 -> fp32 in
// TMEM) over operand images the caller laid out byte-for-byte, with every
// shared-memory-descriptor field supplied by the caller. tests/ uses it to pin
// the descriptor conventions (core-matrix strides, row-shifted start addresses,
// 128-byte swizzle) that the convolution tower depends on.

#include "../../include/at_b200_probe.h"
#include "at_common.cuh"
#include "at_umma.cuh"

// libat_b200_probe.so is linked on its own: it carries its own copy of the error plumbing at_common.cuh declares
namespace at {
namespace {
thread_local std::string g_probe_error;
}
void set_error(const std::string& msg) { g_probe_error = msg; }
int fail(int code, const std::string& msg) {
  g_probe_error = msg;
  return code;
}
}  // namespace at
extern "C" const char* at_probe_last_error(void) { return at::g_probe_error.c_str(); }

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

// ---- throughput probe -----------------------------------------------------------------------------------------
// How fast does one SM retire M=128 x N x K=16 bf16 MMAs with both operands in shared memory (the convolution tower's
// shape), and what do concurrent bulk copies into shared memory / tensor-memory reads by other warps cost? Every CTA of
// the grid runs the same loop so the chip is in its loaded clock / power state; out[cta] = cycles per MMA seen by issuer 0.
struct RateArgs {
  int n, mmas, issuers, copy, epi_warps;
  const uint8_t* src;  // >= 32 KB of readable global memory (bulk-copy source)
  float* out;
};

constexpr int kRateThreads = 32 * 20;
constexpr uint32_t kRateImage = 2 * 38912, kRateRing = 4 * 16384, kRateCopy = 2 * 16384, kRateScratch = 16 * 1024;  // one image per issuer

__global__ void __launch_bounds__(kRateThreads, 1) mma_rate_kernel(const RateArgs p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[6];
  __shared__ uint32_t tmem_base_s;
  __shared__ volatile int done_s;
  const int tid = threadIdx.x, warp = tid / 32, lane = tid % 32;
  const uint32_t ring_u = smem_u32(smem), img_u = ring_u + kRateRing, copy_u = img_u + ((kRateImage + 1023u) & ~1023u);
  const uint32_t scratch_u = copy_u + kRateCopy;
  for (uint32_t i = tid; i < (scratch_u + kRateScratch - ring_u) / 16; i += kRateThreads) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  if (tid == 0) {
    for (int i = 0; i < 6; ++i) mbar_init(smem_u32(&bars[i]), 1);
    fence_mbar_init();
    done_s = 0;
  }
  if (warp == 0) {
    tmem_alloc(smem_u32(&tmem_base_s), 512);
    tmem_relinquish();
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;

  if (p.copy >= 64 && p.copy < 80) {
    // tensor-memory read bandwidth (out[cta] = bytes per cycle over the whole SM): `epi_warps` warps (<= 16) issue back-to-back
    // tcgen05.ld of one shape, `mmas` rounds each, no other work. copy = 64: 32x32b.x32 (128 B per thread), 65: 32x32b.x16,
    // 66: 16x256b.x8 (128 B per thread, 16 lanes x 64 columns), 67: 16x128b.x8 (64 B per thread), 68: 32x32b.x64 (256 B per thread)
    const int ew = warp - 3;
    if (ew >= 0 && ew < p.epi_warps) {
      const uint32_t ta = tmem_base + ((uint32_t)((warp % 4) * 32) << 16) + (uint32_t)((ew / 4) % 4) * 64;
      uint32_t sink = 0;
      __syncwarp();
      const long long t0 = clock64();
      for (int i = 0; i < p.mmas; ++i) {
        uint32_t r[64];
        if (p.copy == 64) {
          asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                       : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                       : "r"(ta) : "memory");
          sink += r[0] + r[31];
        } else if (p.copy == 65) {
          asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                       : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                       : "r"(ta) : "memory");
          sink += r[0] + r[15];
        } else if (p.copy == 66) {
          asm volatile("tcgen05.ld.sync.aligned.16x256b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                       : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                       : "r"(ta) : "memory");
          sink += r[0] + r[31];
        } else if (p.copy == 67) {
          asm volatile("tcgen05.ld.sync.aligned.16x128b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                       : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                       : "r"(ta) : "memory");
          sink += r[0] + r[15];
        } else {
          asm volatile("tcgen05.ld.sync.aligned.32x32b.x64.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32,%33,%34,%35,%36,%37,%38,%39,%40,%41,%42,%43,%44,%45,%46,%47,%48,%49,%50,%51,%52,%53,%54,%55,%56,%57,%58,%59,%60,%61,%62,%63}, [%64];"
                       : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]), "=r"(r[32]), "=r"(r[33]), "=r"(r[34]), "=r"(r[35]), "=r"(r[36]), "=r"(r[37]), "=r"(r[38]), "=r"(r[39]), "=r"(r[40]), "=r"(r[41]), "=r"(r[42]), "=r"(r[43]), "=r"(r[44]), "=r"(r[45]), "=r"(r[46]), "=r"(r[47]), "=r"(r[48]), "=r"(r[49]), "=r"(r[50]), "=r"(r[51]), "=r"(r[52]), "=r"(r[53]), "=r"(r[54]), "=r"(r[55]), "=r"(r[56]), "=r"(r[57]), "=r"(r[58]), "=r"(r[59]), "=r"(r[60]), "=r"(r[61]), "=r"(r[62]), "=r"(r[63])
                       : "r"(ta) : "memory");
          sink += r[0] + r[63];
        }
        tmem_ld_wait();
      }
      const long long t1 = clock64();
      if (ew == 0 && lane == 0) {
        const float bytes_per_thread = p.copy == 64 || p.copy == 66 ? 128.0f : (p.copy == 68 ? 256.0f : 64.0f);
        p.out[blockIdx.x] = bytes_per_thread * 32.0f * (float)p.epi_warps * (float)p.mmas / (float)(t1 - t0);
      }
      if (sink == 0x12345678u) p.out[0] = 0.0f;
    }
  } else if (p.copy >= 10 && p.copy < 20) {
    // latency variants (out[cta] = cycles per round): 10 / 11 = mbarrier ping-pong between two warps (try_wait / test_wait),
    // 12 = tcgen05.commit -> mbarrier with an empty pipe, 13 = one MMA + commit -> mbarrier (issue-to-completion latency)
    const uint32_t b0 = smem_u32(&bars[0]), b1 = smem_u32(&bars[1]);
    const int rounds = p.mmas;
    if (lane == 0 && warp == 0) {
      const uint32_t idesc = idesc_bf16(128, (uint32_t)p.n);
      constexpr uint32_t kDescHi = (128u >> 4) | (1u << 14);
      const uint64_t da = ((uint64_t)kDescHi << 32) | ((2320u >> 4) << 16) | ((img_u + 17 * 16) >> 4);
      const uint64_t db = ((uint64_t)kDescHi << 32) | (((uint32_t)p.n) << 16) | (ring_u >> 4);
      const long long t0 = clock64();
      for (int i = 0; i < rounds; ++i) {
        const uint32_t ph = (uint32_t)i & 1u;
        if (p.copy == 10 || p.copy == 11) {
          mbar_arrive(b0);
          if (p.copy == 10) { while (!mbar_try_wait(b1, ph)) {} } else { while (!mbar_test_wait(b1, ph)) {} }
        } else {
          if (p.copy == 13) mma_bf16(tmem_base, da, db, idesc, false);
          mma_commit(b0);
          while (!mbar_test_wait(b0, ph)) {}
        }
      }
      p.out[blockIdx.x] = (float)(clock64() - t0) / (float)rounds;
    } else if (lane == 0 && warp == 1 && (p.copy == 10 || p.copy == 11)) {
      for (int i = 0; i < rounds; ++i) {
        const uint32_t ph = (uint32_t)i & 1u;
        if (p.copy == 10) { while (!mbar_try_wait(b0, ph)) {} } else { while (!mbar_test_wait(b0, ph)) {} }
        mbar_arrive(b1);
      }
    }
  } else if (warp < 2) {
    if (warp < p.issuers && lane == 0) {
      const uint32_t idesc = idesc_bf16(128, (uint32_t)p.n);
      const uint32_t tmem_d = tmem_base + (uint32_t)warp * 256;
      const uint32_t bar = smem_u32(&bars[warp]);
      constexpr uint32_t kDescHi = (128u >> 4) | (1u << 14);
      const uint32_t a_fixed = (2320u >> 4) << 16, b_fixed = ((uint32_t)p.n) << 16;  // LBO: A one plane, B n rows of 16 bytes
      const long long t0 = clock64();
      if (p.copy >= 32) {
        // the tower's per-slab pattern: copy = 32 + flags; per four MMAs: 1 = mbarrier try_wait (on a phase that is already
        // complete), 2 = tcgen05.commit to a barrier nobody waits on, 4 = tcgen05.fence::after_thread_sync
        const bool f_wait = p.copy & 1, f_commit = p.copy & 2, f_fence = p.copy & 4;
        const uint32_t bw = smem_u32(&bars[5]), bc = smem_u32(&bars[4]) ;
        const uint32_t a_base = a_fixed | ((img_u + (uint32_t)warp * 38912u + 17u * 16u) >> 4), b_base = b_fixed | (ring_u >> 4);
        for (int sl = 0; sl < p.mmas / 4; ++sl) {
          if (f_wait) { while (!mbar_try_wait(bw, 1)) {} }
          if (f_fence) tc_fence_after();
          const uint32_t a_lo = a_base + (uint32_t)(sl & 7), b_lo = b_base + (uint32_t)(sl & 3) * 1024u;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            mma_bf16(tmem_d, ((uint64_t)kDescHi << 32) | (a_lo + (uint32_t)j * 290u), ((uint64_t)kDescHi << 32) | (b_lo + (uint32_t)j * 256u), idesc,
                     sl > 0 || j > 0);
          if (f_commit) mma_commit(bc);
        }
      } else
      for (int i = 0; i < p.mmas; ++i) {
        const int tap = (i >> 2) % 9, j = i & 3;
        const uint32_t a16 = (img_u + (uint32_t)warp * 38912u + (uint32_t)((17 + (tap / 3 - 1) * 16 + (tap % 3 - 1)) * 16) + (uint32_t)j * 2 * 2320u) >> 4;
        const uint32_t b16 = (ring_u + (uint32_t)((i >> 2) & 3) * 16384u + (uint32_t)j * 2 * (uint32_t)p.n * 16u) >> 4;
        mma_bf16(tmem_d, ((uint64_t)kDescHi << 32) | a_fixed | a16, ((uint64_t)kDescHi << 32) | b_fixed | b16, idesc, i > 0);
      }
      mma_commit(bar);
      mbar_wait(bar, 0);
      const long long t1 = clock64();
      if (warp == 0) {
        p.out[blockIdx.x] = (float)(t1 - t0) / (float)p.mmas;
        done_s = 1;
      }
    }
  } else if (warp == 2) {
    if (lane == 0 && p.copy && p.copy < 10) {  // back-to-back 16 KB bulk copies, two in flight, for as long as the issuer runs
      uint32_t ph[2] = {0, 0};
      int i = 0;
      for (; !done_s; ++i) {
        const int b = i & 1;
        const uint32_t bar = smem_u32(&bars[2 + b]);
        if (i >= 2) {
          mbar_wait(bar, ph[b]);
          ph[b] ^= 1;
        }
        mbar_arrive_expect_tx(bar, 16384);
        bulk_g2s(copy_u + b * 16384, p.src + b * 16384, 16384, bar);
      }
      for (int b = 0; b < 2 && b < i; ++b) mbar_wait(smem_u32(&bars[2 + b]), ph[b]);  // nothing may land after the CTA is gone
    }
  } else if (warp - 3 < p.epi_warps) {
    // epilogue-like traffic: tensor-memory reads of this warp's lane quarter + 128-bit shared stores / loads
    const uint32_t trow = tmem_base + ((uint32_t)((warp % 4) * 32) << 16) + (uint32_t)(((warp - 3) / 4) % 4) * 32;
    const uint32_t my = scratch_u + (uint32_t)(tid % 512) * 16;
    uint32_t sink = 0;
    if (p.copy >= 32 && (p.copy & 8)) {
      // flag 8: the epilogue warps poll an mbarrier that never completes (as the tower's epilogue does while MMAs run)
      const uint32_t bw = smem_u32(&bars[3]);
      while (!done_s) {
        if (mbar_try_wait(bw, 0)) break;
        sink += (uint32_t)clock64();
      }
    } else
    while (!done_s) {
      uint32_t r[32];
      tmem_ld32(trow, r);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(my + (uint32_t)(i & 1) * 8192u), "r"(r[i * 8]), "r"(r[i * 8 + 1]), "r"(r[i * 8 + 2]), "r"(r[i * 8 + 3]) : "memory");
        sink += r[i * 8 + 4];
      }
    }
    if (sink == 0x12345678u) p.out[0] = 0.0f;
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

extern "C" int at_probe_mma_rate(int n, int mmas, int issuers, int copy, int epi_warps, const void* d_src, int grid, float* d_cycles,
                                 void* stream) {
  using namespace at;
  if (!d_src || !d_cycles) return fail(AT_E_INVALID, "at_probe_mma_rate: null pointer");
  if (n < 16 || n > 256 || n % 16 != 0 || mmas < 1 || issuers < 1 || issuers > 2 || epi_warps < 0 || epi_warps > 16 || grid < 1)
    return fail(AT_E_INVALID, "at_probe_mma_rate: bad argument");
  RateArgs p{n, mmas, issuers, copy, epi_warps, static_cast<const uint8_t*>(d_src), d_cycles};
  const size_t smem = kRateRing + ((kRateImage + 1023u) & ~1023u) + kRateCopy + kRateScratch + 1024;
  AT_CUDA_TRY(cudaFuncSetAttribute(mma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  mma_rate_kernel<<<grid, kRateThreads, smem, static_cast<cudaStream_t>(stream)>>>(p);
  AT_CUDA_TRY(cudaGetLastError());
  return AT_OK;
}

// ---- SM hog: `grid` CTAs that each take a whole SM (200 KB of dynamic shared memory) and spin for `cycles` -------------------
namespace at {
__global__ void hog_kernel(long long cycles, unsigned* d_sink) {
  extern __shared__ unsigned hog_smem[];
  const long long t0 = clock64();
  unsigned v = 0;
  while (clock64() - t0 < cycles) v += hog_smem[(v + threadIdx.x) & 1023];
  if (v == 0xFFFFFFFFu) *d_sink = v;  // keeps the loop alive
}
}  // namespace at

extern "C" int at_probe_hog(int grid, long long cycles, void* d_sink, void* stream) {
  using namespace at;
  if (grid < 1 || cycles < 0 || !d_sink) return fail(AT_E_INVALID, "at_probe_hog: bad argument");
  const size_t smem = 200 * 1024;
  AT_CUDA_TRY(cudaFuncSetAttribute(hog_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  hog_kernel<<<grid, 32, smem, static_cast<cudaStream_t>(stream)>>>(cycles, static_cast<unsigned*>(d_sink));
  AT_CUDA_TRY(cudaGetLastError());
  return AT_OK;
}
