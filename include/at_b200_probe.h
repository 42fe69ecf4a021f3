/*
 * at_b200_probe.h -- C ABI of the bring-up probes (libat_b200_probe.so). NOT part of the product library: these entry
 * points have no counterpart in the reference; they pin the tcgen05 shared-memory / tensor-memory descriptor conventions
 * the tower kernels rely on (tests/test_gpu_probe.py) and measure the tensor pipe's issue rate (tools/mma_rate.py).
 * Return 0 or a negative AT_E_* code (include/at_b200.h); at_probe_last_error() has the message.
 */
#ifndef AT_B200_PROBE_H
#define AT_B200_PROBE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* at_probe_last_error(void);

/* One 128xN tcgen05.mma chain over K (bf16, fp32 accumulate in TMEM) on operand
 * images the caller laid out, with the descriptor fields given explicitly.
 * Test-only: pins the shared-memory descriptor conventions the tower relies on. */
int at_probe_umma(const void* d_a_image, int64_t a_bytes, const void* d_b_image, int64_t b_bytes, int n, int k,
                  uint32_t a_start_off, uint32_t a_lbo, uint32_t a_sbo, uint32_t a_kstep, uint32_t b_start_off, uint32_t b_lbo,
                  uint32_t b_sbo, uint32_t b_kstep, int layout_type, float* d_out /* [128,n] */, uint32_t* d_status, void* stream);

/* Same chain with the A operand in tensor memory (rows = TMEM lanes, two bf16 per 32-bit column, written with
 * tcgen05.st from row-major bf16 bit patterns [128,k]). Test-only: pins the layout the attention P*V product uses. */
int at_probe_umma_ts(const void* d_a_rows, const void* d_b_image, int64_t b_bytes, int n, int k, int pack_mode, uint32_t a_col,
                     uint32_t a_cols_per_kstep, uint32_t b_start_off, uint32_t b_lbo, uint32_t b_sbo, uint32_t b_kstep, float* d_out,
                     uint32_t* d_status, void* stream);

/* Throughput probe: every CTA of `grid` issues `mmas` back-to-back 128 x n x 16 bf16 MMAs (operands in shared memory, the
 * tower's layout) from `issuers` (1 or 2) warps, optionally while another warp streams 16 KB bulk copies into shared
 * memory (`copy` != 0) and `epi_warps` warps read tensor memory / store to shared memory like the tower's epilogue.
 * d_cycles[cta] = SM cycles per MMA of issuer 0. No reference counterpart: measures the hardware ceiling DESIGN.md quotes. */
int at_probe_mma_rate(int n, int mmas, int issuers, int copy, int epi_warps, const void* d_src /* >= 32 KB */, int grid,
                      float* d_cycles /* [grid] */, void* stream);

/* SM hog: `grid` CTAs, each taking a whole SM (200 KB of shared memory) and spinning for `cycles` SM cycles -- a stand-in for
 * a collective's kernel that holds SMs while it waits for its peers (tools/contention.py: what a tower launch costs when it
 * does not get every SM at once). No reference counterpart. d_sink: any 4 writable bytes. */
int at_probe_hog(int grid, long long cycles, void* d_sink, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* AT_B200_PROBE_H */
