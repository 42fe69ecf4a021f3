// Host-side plan: from ModelConfig dimensions to (a) the weight-image table keyed
// by the reference's state_dict names and (b) the layer list the kernels run.
// The module tree being mirrored is built at alphatriangle/nn/model.py:116-266.

#include "at_engine.cuh"

namespace at {

namespace {

struct Builder {
  Plan& p;
  int64_t img = 0, fold = 0;

  int64_t add(const std::string& name, int64_t numel) {
    p.image.push_back({name, img, numel});
    const int64_t off = img;
    img += (numel + 3) & ~int64_t(3);  // every tensor starts 16-byte aligned: the image doubles as the trainer's parameter storage
    return off;
  }
  int64_t reserve(int64_t floats) {
    const int64_t off = fold;
    fold += (floats + 3) & ~int64_t(3);  // keep every folded tensor 16-byte aligned
    return off;
  }
  void bn(const std::string& prefix, int n, Affine& f) {
    f.bn_g = add(prefix + ".weight", n);
    f.bn_b = add(prefix + ".bias", n);
    f.bn_m = add(prefix + ".running_mean", n);
    f.bn_v = add(prefix + ".running_var", n);
  }
  void place(Affine& f) {
    f.Kpad = (f.K + 15) & ~15;
    f.wt = reserve((int64_t)f.Kpad * f.N);
    f.bias = reserve(f.N);
    if (f.N > p.max_n) p.max_n = f.N;
  }
};

int conv_out(int in, int k, int s, int pad) { return (in + 2 * pad - k) / s + 1; }

}  // namespace

int build_plan(const at_model_desc& d, Plan& p) {
  p = Plan{};
  p.d = d;
  Builder b{p};
  const bool use_bn = d.use_batch_norm != 0;

  if (d.env.rows <= 0 || d.env.cols <= 0 || d.env.rows * d.env.cols > AT_MAX_CELLS)
    return fail(AT_E_INVALID, "model: need 0 < ROWS*COLS <= 128");
  if (d.env.slots <= 0 || d.env.slots > AT_MAX_SLOTS) return fail(AT_E_INVALID, "model: need 1 <= NUM_SHAPE_SLOTS <= 4");
  if (d.grid_channels <= 0 || d.other_dim <= 0) return fail(AT_E_INVALID, "model: GRID_INPUT_CHANNELS and OTHER_NN_INPUT_FEATURES_DIM must be positive");
  if (d.n_conv < 0 || d.n_conv > AT_MAX_CONVS) return fail(AT_E_INVALID, "model: at most 8 CONV_FILTERS entries");
  if (d.n_shared < 0 || d.n_shared > AT_MAX_MLP || d.n_policy < 0 || d.n_policy > AT_MAX_MLP || d.n_value < 0 || d.n_value > AT_MAX_MLP)
    return fail(AT_E_INVALID, "model: at most 4 hidden layers per MLP");
  if (d.num_atoms < 2 || !(d.v_min < d.v_max)) return fail(AT_E_INVALID, "model: need NUM_VALUE_ATOMS > 1 and VALUE_MIN < VALUE_MAX");
  if (d.activation < AT_ACT_RELU || d.activation > AT_ACT_SIGMOID) return fail(AT_E_INVALID, "model: unknown activation");
  if (d.n_conv == 0 && d.n_res_blocks == 0) return fail(AT_E_UNSUPPORTED, "model: a network with no convolution at all is not supported");

  int ch = d.grid_channels, h = d.env.rows, w = d.env.cols;
  auto conv = [&](const std::string& wname, const std::string& bn_prefix, int cout, int kh, int kw, int sh, int sw, int ph, int pw,
                  int act, ConvLayer& c) -> int {
    if (cout <= 0 || kh <= 0 || kw <= 0 || sh <= 0 || sw <= 0 || ph < 0 || pw < 0) return fail(AT_E_INVALID, "model: bad convolution geometry");
    c.cin = ch; c.cout = cout; c.kh = kh; c.kw = kw; c.sh = sh; c.sw = sw; c.ph = ph; c.pw = pw;
    c.in_h = h; c.in_w = w;
    c.out_h = conv_out(h, kh, sh, ph);
    c.out_w = conv_out(w, kw, sw, pw);
    if (c.out_h <= 0 || c.out_w <= 0) return fail(AT_E_INVALID, "model: convolution output is empty");
    c.f.K = kh * kw * ch;
    c.f.N = cout;
    c.f.act = act;
    c.f.src_w = b.add(wname + ".weight", (int64_t)cout * ch * kh * kw);
    if (use_bn) b.bn(bn_prefix, cout, c.f);
    else c.f.src_b = b.add(wname + ".bias", cout);
    c.f.perm_cols = c.f.K;       // reference column c*(kh*kw) + tap -> row tap*cin + c
    c.f.perm_inner = kh * kw;
    b.place(c.f);
    ch = cout; h = c.out_h; w = c.out_w;
    const int64_t unit = (int64_t)c.out_h * c.out_w * cout;
    if (unit > p.unit_pq) p.unit_pq = unit;
    return AT_OK;
  };

  p.convs.resize(d.n_conv);
  for (int i = 0; i < d.n_conv; ++i) {
    const std::string base = "conv_body." + std::to_string(i);
    AT_TRY(conv(base + ".0", base + ".1", d.conv_filters[i], d.conv_kernel[i][0], d.conv_kernel[i][1], d.conv_stride[i][0],
                d.conv_stride[i][1], d.conv_pad[i][0], d.conv_pad[i][1], d.activation, p.convs[i]));
  }
  if (d.n_res_blocks > 0) {
    int idx = 0;
    if (d.res_filters <= 0) return fail(AT_E_INVALID, "model: RESIDUAL_BLOCK_FILTERS must be positive");
    if (ch != d.res_filters) {
      p.has_res_proj = true;
      AT_TRY(conv("res_body.0.0", "res_body.0.1", d.res_filters, 1, 1, 1, 1, 0, 0, d.activation, p.res_proj));
      idx = 1;
    }
    p.res.resize(d.n_res_blocks);
    for (int r = 0; r < d.n_res_blocks; ++r, ++idx) {
      const std::string base = "res_body." + std::to_string(idx);
      AT_TRY(conv(base + ".conv1.0", base + ".conv1.1", ch, 3, 3, 1, 1, 1, 1, d.activation, p.res[r].c1));
      // conv2 -> bn2 -> (+x) -> activation: the activation is applied after the residual add
      AT_TRY(conv(base + ".conv2", base + ".bn2", ch, 3, 3, 1, 1, 1, 1, d.activation, p.res[r].c2));
    }
  }
  p.cnn_ch = ch; p.out_h = h; p.out_w = w;
  p.L = h * w;

  int feat_ch = ch;
  p.transformer = d.use_transformer != 0;
  if (p.transformer) {
    const int dm = d.t_dim;
    if (dm <= 0 || d.t_heads <= 0 || d.t_layers <= 0 || d.t_ff <= 0 || dm % d.t_heads != 0)
      return fail(AT_E_INVALID, "model: bad transformer dimensions");
    if (d.activation != AT_ACT_RELU && d.activation != AT_ACT_GELU)
      return fail(AT_E_INVALID, "model: nn.TransformerEncoderLayer accepts only relu / gelu (model.py:188-195)");
    if (p.L > 5000) return fail(AT_E_INVALID, "model: sequence longer than PositionalEncoding max_len");
    if (ch != dm) {
      p.has_input_proj = true;
      Affine& f = p.input_proj;
      f.K = ch; f.N = dm; f.act = AT_ACT_NONE;
      f.src_w = b.add("input_proj.weight", (int64_t)dm * ch);
      f.src_b = b.add("input_proj.bias", dm);
      b.place(f);
    }
    p.src_pe = b.add("pos_encoder.pe", (int64_t)5000 * dm);
    p.pe = b.reserve((int64_t)p.L * dm);
    p.enc.resize(d.t_layers);
    for (int l = 0; l < d.t_layers; ++l) {
      const std::string base = "transformer_body.layers." + std::to_string(l);
      EncoderLayer& e = p.enc[l];
      e.qkv.K = dm; e.qkv.N = 3 * dm;
      e.qkv.src_w = b.add(base + ".self_attn.in_proj_weight", (int64_t)3 * dm * dm);
      e.qkv.src_b = b.add(base + ".self_attn.in_proj_bias", 3 * dm);
      e.out.K = dm; e.out.N = dm;
      e.out.src_w = b.add(base + ".self_attn.out_proj.weight", (int64_t)dm * dm);
      e.out.src_b = b.add(base + ".self_attn.out_proj.bias", dm);
      e.ff1.K = dm; e.ff1.N = d.t_ff; e.ff1.act = d.activation;
      e.ff1.src_w = b.add(base + ".linear1.weight", (int64_t)d.t_ff * dm);
      e.ff1.src_b = b.add(base + ".linear1.bias", d.t_ff);
      e.ff2.K = d.t_ff; e.ff2.N = dm;
      e.ff2.src_w = b.add(base + ".linear2.weight", (int64_t)dm * d.t_ff);
      e.ff2.src_b = b.add(base + ".linear2.bias", dm);
      e.src_ln1_w = b.add(base + ".norm1.weight", dm);
      e.src_ln1_b = b.add(base + ".norm1.bias", dm);
      e.src_ln2_w = b.add(base + ".norm2.weight", dm);
      e.src_ln2_b = b.add(base + ".norm2.bias", dm);
      b.place(e.qkv); b.place(e.out); b.place(e.ff1); b.place(e.ff2);
      e.ln1_w = b.reserve(dm); e.ln1_b = b.reserve(dm); e.ln2_w = b.reserve(dm); e.ln2_b = b.reserve(dm);
    }
    p.src_lnf_w = b.add("transformer_body.norm.weight", dm);
    p.src_lnf_b = b.add("transformer_body.norm.bias", dm);
    p.lnf_w = b.reserve(dm); p.lnf_b = b.reserve(dm);
    feat_ch = dm;
    const int64_t unit = (int64_t)p.L * dm;
    if (unit > p.unit_pq) p.unit_pq = unit;
    const int widest = 3 * dm > d.t_ff ? 3 * dm : d.t_ff;
    p.unit_r = (int64_t)p.L * widest;
  }
  p.flat = p.L * feat_ch;
  p.concat_ld = (p.flat + d.other_dim + 15) & ~15;
  p.action_dim = d.env.slots * d.env.rows * d.env.cols;

  auto mlp = [&](const std::string& name, int in_dim, int n_hidden, const int32_t* dims, int out_dim, bool first_is_trunk,
                 std::vector<Affine>& dst) -> int {
    int q = 0;
    for (int i = 0; i < n_hidden; ++i) {
      if (dims[i] <= 0) return fail(AT_E_INVALID, "model: MLP widths must be positive");
      Affine f;
      f.K = in_dim; f.N = dims[i]; f.act = d.activation;
      f.src_w = b.add(name + "." + std::to_string(q) + ".weight", (int64_t)dims[i] * in_dim);
      f.src_b = b.add(name + "." + std::to_string(q) + ".bias", dims[i]);
      if (use_bn) b.bn(name + "." + std::to_string(q + 1), dims[i], f);
      if (first_is_trunk && i == 0 && !p.transformer) {
        // reference flattens NCHW (column c*HW + p, model.py:290); activations here are NHWC (row p*C + c)
        f.perm_cols = p.flat;
        f.perm_inner = p.L;
      }
      b.place(f);
      if (first_is_trunk && i == 0) f.Kpad = p.concat_ld;  // same value; the GEMM runs over the padded row
      dst.push_back(f);
      in_dim = dims[i];
      if (dims[i] > p.max_mlp) p.max_mlp = dims[i];
      q += use_bn ? 3 : 2;
    }
    if (out_dim > 0) {
      Affine f;
      f.K = in_dim; f.N = out_dim; f.act = AT_ACT_NONE;
      f.src_w = b.add(name + "." + std::to_string(q) + ".weight", (int64_t)out_dim * in_dim);
      f.src_b = b.add(name + "." + std::to_string(q) + ".bias", out_dim);
      b.place(f);
      dst.push_back(f);
    }
    return AT_OK;
  };
  // shared_fc may be empty (FC_DIMS_SHARED = []): the heads then read the concat row directly
  AT_TRY(mlp("shared_fc", p.flat + d.other_dim, d.n_shared, d.shared_dims, 0, true, p.shared));
  const int trunk = d.n_shared > 0 ? d.shared_dims[d.n_shared - 1] : p.flat + d.other_dim;
  if (d.n_shared == 0 && !p.transformer) return fail(AT_E_UNSUPPORTED, "model: FC_DIMS_SHARED = [] without a transformer is not supported");
  AT_TRY(mlp("policy_head", trunk, d.n_policy, d.policy_dims, p.action_dim, false, p.policy));
  AT_TRY(mlp("value_head", trunk, d.n_value, d.value_dims, d.num_atoms, false, p.value));

  p.image_floats = b.img;
  p.folded_floats = b.fold;
  return AT_OK;
}

}  // namespace at
