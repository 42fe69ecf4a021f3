// tcgen05 / TMEM bf16 path (at_bf16.cu): interface towards the engine.
#pragma once

#include <string>

#include "at_engine.cuh"

namespace at {

struct Bf16Path;  // opaque: packed bf16 weight sets + scratch of the tcgen05 kernels

// Creates the path if the plan matches what the kernels specialise; otherwise
// leaves *out null, explains why in *reason and still returns AT_OK.
int bf16_create(const Plan& plan, int64_t max_batch, Bf16Path** out, std::string* reason);
void bf16_destroy(Bf16Path* p);
// folded fp32 set -> 16-bit operand images: the packing steps as jobs of the weight-preparation table (phase 1), and the
// base pointers of weight set `slot` (0/1) they write to
class PrepList;
struct PrepBases;
void bf16_add_prep_jobs(const Bf16Path* p, const Plan& plan, PrepList& list);
void bf16_prep_bases(const Bf16Path* p, int slot, PrepBases* bases);
// operand format (FMT_BF16 / FMT_F16, at_prepjobs.cuh) the 16-bit images of weight set `slot` hold; -1: none yet
int bf16_packed_format(const Bf16Path* p, int slot);
void bf16_set_packed_format(Bf16Path* p, int slot, int fmt);
const float* bf16_fc_bias(const Bf16Path* p, int slot);
// PackedState[batch] -> split-K partial sums of the first shared FC layer over the tower output, fp32
// [n_splits][batch][shared[0].N] (bf16_partial), plus -- unless the consumer derives them itself from the packed
// states (with_other = false) -- the "other" feature vectors (bf16_other)
int bf16_trunk_partials(Bf16Path* p, const Plan& plan, const float* d_folded, int slot, const void* d_states, int64_t batch,
                        int* n_splits, cudaStream_t stream, int64_t* launches, bool with_other = true);
const float* bf16_partial(const Bf16Path* p);
const float* bf16_other(const Bf16Path* p);
// partials + other columns + bias, activation -> d_trunk fp32 [batch, shared[0].N] (unfused tail, also the test hook)
int bf16_finish_trunk(Bf16Path* p, const Plan& plan, const float* d_folded, int slot, int n_splits, int64_t batch, float* d_trunk,
                      cudaStream_t stream, int64_t* launches);
// Same as bf16_trunk_partials with every intermediate buffer inside the library's static device array
// (no allocation, batch <= bf16_static_max_batch()); *partial / *other point into that array.
int64_t bf16_static_max_batch();
int bf16_static_check(const Bf16Path* p, const Plan& plan, int64_t batch);  // batch / scratch-size refusals, nothing queued
int bf16_trunk_partials_static(Bf16Path* p, const Plan& plan, const float* d_folded, int slot, const void* d_states, int64_t batch,
                               int* n_splits, const float** partial, const float** other, cudaStream_t stream, int64_t* launches);
const uint32_t* bf16_status_ptr(const Bf16Path* p);  // device word: non-zero if a pipeline wait timed out
void bf16_set_timing(Bf16Path* p, bool on);  // event pairs around the tower / FC kernels of every evaluation from now on
int bf16_kernel_ms(Bf16Path* p, float* tower_ms, float* fc_ms);

}  // namespace at
