// Engine internals: weight-image layout, folded weight sets, layer plan.
#pragma once

#include <string>
#include <vector>

#include "at_common.cuh"

namespace at {

struct ImageEntry {
  std::string name;  // reference state_dict key
  int64_t offset;    // in floats
  int64_t numel;
};

// One affine layer after folding: Wt [Kpad,N] fp32 (rows >= K are zero) + bias [N].
struct Affine {
  int K = 0, Kpad = 0, N = 0;
  int act = AT_ACT_NONE;
  // source tensors in the weight image (-1 = absent)
  int64_t src_w = -1, src_b = -1, bn_g = -1, bn_b = -1, bn_m = -1, bn_v = -1;
  // column permutation applied while folding: reference column a*perm_inner + i
  // (for the first perm_cols columns) lands in row i*(perm_cols/perm_inner) + a
  int perm_cols = 0, perm_inner = 1;
  // destination in a folded weight set (floats)
  int64_t wt = -1, bias = -1;
};

struct ConvLayer {
  Affine f;  // K = kh*kw*cin with k = (ky*kw+kx)*cin + c
  int cin = 0, cout = 0, kh = 0, kw = 0, sh = 1, sw = 1, ph = 0, pw = 0;
  int in_h = 0, in_w = 0, out_h = 0, out_w = 0;
};

struct ResBlock {
  ConvLayer c1, c2;
};

struct EncoderLayer {
  Affine qkv, out, ff1, ff2;
  int64_t src_ln1_w = -1, src_ln1_b = -1, src_ln2_w = -1, src_ln2_b = -1;
  int64_t ln1_w = -1, ln1_b = -1, ln2_w = -1, ln2_b = -1;  // in the folded set
};

struct Plan {
  at_model_desc d{};
  std::vector<ImageEntry> image;
  int64_t image_floats = 0;

  std::vector<ConvLayer> convs;  // conv_body
  bool has_res_proj = false;
  ConvLayer res_proj;
  std::vector<ResBlock> res;
  int cnn_ch = 0, out_h = 0, out_w = 0;  // CNN output geometry

  bool transformer = false;
  bool has_input_proj = false;
  Affine input_proj;
  int64_t src_pe = -1, pe = -1;  // pe: folded [L,d]
  std::vector<EncoderLayer> enc;
  int64_t src_lnf_w = -1, src_lnf_b = -1, lnf_w = -1, lnf_b = -1;

  int L = 0;        // tokens = out_h*out_w
  int flat = 0;     // flattened trunk features
  int concat_ld = 0;  // padded row length of [flat | other]
  std::vector<Affine> shared, policy, value;  // last policy / value entry is the logits layer
  int action_dim = 0;

  int64_t folded_floats = 0;
  // scratch geometry (floats per board)
  int64_t unit_pq = 0, unit_r = 0, max_mlp = 0, max_n = 0;
};

int build_plan(const at_model_desc& d, Plan& plan);  // at_plan.cu

// at_prep.cu: the image -> folded fp32 weight set steps, appended to a weight-preparation job table (at_prepjobs.cuh)
class PrepList;
void fold_add_prep_jobs(const Plan& plan, PrepList& list);

}  // namespace at
