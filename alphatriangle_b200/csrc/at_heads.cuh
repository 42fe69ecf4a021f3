// Fused trunk-finish + heads + softmax kernel (at_heads.cu).
#pragma once

#include "at_common.cuh"

namespace at {

struct HeadsParams {
  // trunk FC: split-K partials [splits][batch][trunk_n], other features [batch][other_dim] against rows w_other [other_dim][trunk_n]
  const float* partial;
  int splits;
  const float* other;
  int other_dim;
  const float* w_other;
  const float* trunk_bias;
  int trunk_n;
  int act;
  // heads: Wt [K][N] fp32 with folded BatchNorm, bias [N]
  const float *wp1, *bp1, *wp2, *bp2;
  const float *wv1, *bv1, *wv2, *bv2;
  int pol_hidden, val_hidden, actions, atoms;
  float v_min, v_max;
  int64_t batch;
  float* probs;
  float* value;
  uint32_t* flags;
  int w_stage_off;         // filled by the launcher: float offset of the weight staging area in shared memory
  const uint32_t* status;  // optional: tcgen05 pipeline status word folded into the flags
  int fmt;                 // operand format of `frags` and of the activations fed to the MMAs (FMT_BF16 / FMT_F16)
  const uint2* frags;      // tensor-core variant: the four head matrices as bf16 mma.sync B fragments (heads_add_prep_jobs), else null
  // tensor-core variant, optional: derive the other-feature vectors from the packed states in the prologue instead of
  // reading `other` (same arithmetic as the feature kernel, at_features.cuh); needs env_cols <= 32
  const PackedState* states;
  int env_rows, env_cols, env_slots;
};

size_t heads_fused_smem(const HeadsParams& p);
bool heads_fused_supported(const HeadsParams& p);
int launch_heads_fused(const HeadsParams& p, cudaStream_t stream, int64_t* launches);

// Tensor-core variant (heads_mma_kernel): same inputs and outputs, the four head GEMMs on warp-level bf16 MMAs with the
// weights pre-packed into B-fragment order (one coalesced 256-byte load per 16x8 tile, no shared-memory staging).
bool heads_mma_supported(const HeadsParams& p);
size_t heads_frag_count(const HeadsParams& p);  // uint2 elements of the packed image
class PrepList;
// the fragment packing as jobs of the weight-preparation table; src[] = folded-set float offsets of wp1 / wv1 / wp2 / wv2
void heads_add_prep_jobs(const HeadsParams& p, const int64_t src[4], PrepList& list);
int launch_heads_mma(const HeadsParams& p, cudaStream_t stream, int64_t* launches);

}  // namespace at
