// Weight preparation as a job table: every fold / copy / tile / fragment-pack step of at_engine_load_weights is one
// entry of a table built once at engine creation, and a publication runs the whole table in TWO launches (phase 0:
// image -> folded fp32 set; phase 1: folded set -> tensor-core operand images) instead of ~180 small ones.
// Offsets in a job are relative to one of a handful of base pointers, so the same table serves both weight sets.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <vector>

#include "at_common.cuh"

namespace at {

enum PrepBase : int { PB_IMAGE = 0, PB_FOLDED, PB_WSTREAM, PB_WFC, PB_BIAS, PB_FCBIAS, PB_FRAGS, PB_COUNT };
enum PrepKind : int { PK_FOLD = 0, PK_COPY, PK_PACK_B, PK_PACK_SEL, PK_GATHER_BIAS, PK_PACK_FRAG, PK_BIAS_SLAB, PK_FC_BIAS, PK_PE_T };

struct PrepJob {
  int kind, phase, first_block, n_blocks;
  int64_t a[8];  // element offsets into the job's source / destination bases (meaning per kind, see prep_run_job)
  int i[10];
  float f[2];
};

struct PrepBases {
  void* p[PB_COUNT];
};

// 16-bit operand formats of the tensor-core path (tcgen05 kind::f16 takes either): same exponent range as fp32 with an
// 8-bit mantissa, or an 11-bit mantissa with a 5-bit exponent (max 65504).
enum OperandFormat : int { FMT_BF16 = 0, FMT_F16 = 1 };

__device__ __forceinline__ unsigned short to_operand16(float v, int fmt) {
  if (fmt == FMT_F16) {
    // saturate instead of overflowing to infinity: a weight beyond fp16's range clamps to +-65504
    const __half h = __float2half_rn(fminf(fmaxf(v, -65504.0f), 65504.0f));
    return *reinterpret_cast<const unsigned short*>(&h);
  }
  const __nv_bfloat16 b = __float2bfloat16_rn(v);
  return *reinterpret_cast<const unsigned short*>(&b);
}

class PrepList {
 public:
  void add(PrepJob j, int64_t elements) {
    int64_t blocks = (elements + 1023) / 1024;  // 256 threads x 4 elements, grid-stride inside the job beyond that
    if (blocks < 1) blocks = 1;
    if (blocks > 96) blocks = 96;
    j.n_blocks = (int)blocks;
    jobs_.push_back(j);
  }
  const std::vector<PrepJob>& jobs() const { return jobs_; }
  std::vector<PrepJob>& jobs() { return jobs_; }

 private:
  std::vector<PrepJob> jobs_;
};

// Uploaded table + per-phase launch geometry.
struct PrepTable {
  PrepJob* d_jobs = nullptr;
  int n_jobs = 0;
  int phase_first_job[2] = {0, 0}, phase_jobs[2] = {0, 0}, phase_blocks[2] = {0, 0};
};

int prep_upload(PrepList& list, PrepTable* table);  // sorts by phase, assigns block ranges, copies to the device
void prep_free(PrepTable* table);
// first_phase = 0: the whole table; 1: only the 16-bit operand images (re-pack of a folded set in another operand format)
int prep_run(const PrepTable& table, const PrepBases& bases, int fmt, cudaStream_t stream, int64_t* launches, int first_phase = 0);

}  // namespace at
