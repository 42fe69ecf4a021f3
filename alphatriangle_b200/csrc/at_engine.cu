// Engine: owns the folded weight sets and scratch memory, runs the layer plan,
// and exports the C ABI declared in include/at_b200.h.

#include <cstring>
#include <new>

#include "at_bf16.cuh"
#include "at_engine.cuh"
#include "at_fp32.cuh"
#include "at_heads.cuh"
#include "at_prepjobs.cuh"

namespace at {

namespace {
thread_local std::string g_last_error;
}

void set_error(const std::string& msg) { g_last_error = msg; }
int fail(int code, const std::string& msg) {
  g_last_error = msg;
  return code;
}

}  // namespace at

using namespace at;

struct at_engine {
  Plan plan;
  int device = 0;
  int64_t max_batch = 0;
  int64_t chunk = 0;  // boards per pass through the fp32 scratch

  // double-buffered folded weights: evaluation reads `active`, at_engine_load_weights fills the other one
  float* folded[2] = {nullptr, nullptr};
  cudaEvent_t ready[2] = {nullptr, nullptr};     // folding of set i finished
  cudaEvent_t last_use[2] = {nullptr, nullptr};  // last evaluation that read set i
  int active = -1;
  int pending = -1;  // set filled by at_engine_publish_weights, promoted to `active` by the first evaluation that finds it ready
  int64_t version[2] = {-1, -1};

  // scratch (fp32 path)
  float *buf_p = nullptr, *buf_q = nullptr, *buf_r = nullptr, *concat = nullptr;
  float *mlp_a = nullptr, *mlp_b = nullptr, *trunk = nullptr, *pol_logits = nullptr, *val_logits = nullptr;
  float* partial = nullptr;
  int64_t partial_floats = 0;
  // staging for the packed / host entry points
  float *feat_grid = nullptr, *feat_other = nullptr;
  void* d_states = nullptr;
  float *d_policy = nullptr, *d_value = nullptr;
  uint32_t* d_flags = nullptr;

  // host entry point: results of slice i drain over PCIe on this stream while slice i+1 computes
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t slice_done[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t copies_done = nullptr;

  // pipelined host entry points (at_engine_submit_host / at_engine_wait_host): two batches in flight, each with its own
  // device staging, so the PCIe copies of one batch run under the kernels of the other
  struct HostSlot {
    void* d_states = nullptr;
    float *d_policy = nullptr, *d_value = nullptr;
    uint32_t* d_flags = nullptr;
    cudaEvent_t staged = nullptr, computed = nullptr, done = nullptr;
    bool busy = false;
    int64_t launches = 0;
  } slot[2];
  cudaStream_t h2d_stream = nullptr;

  Bf16Path* bf16 = nullptr;  // tcgen05 path state (null if the configuration is not covered)
  std::string bf16_reason;
  uint2* heads_frag[2] = {nullptr, nullptr};  // per weight set: head matrices as bf16 MMA fragments (null: shape not covered)
  PrepTable prep;  // every fold / pack step of a weight load, run as two launches
  int tc_fmt = FMT_BF16;  // operand format weight loads pack for: the one the last tensor-core evaluation asked for

  int64_t last_launches = 0;
};

namespace {

template <typename T>
int dev_alloc(T** p, int64_t count) {
  *p = nullptr;
  if (count <= 0) return AT_OK;
  AT_CUDA_TRY(cudaMalloc(reinterpret_cast<void**>(p), (size_t)count * sizeof(T)));
  return AT_OK;
}

struct DeviceGuard {
  int prev = -1;
  bool ok = false;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) == cudaSuccess && cudaSetDevice(dev) == cudaSuccess) ok = true;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

int pick_splits(int64_t M, int N, int K, int64_t partial_floats) {
  const int64_t tiles = ((M + 63) / 64) * ((N + 63) / 64);
  const int k_steps = (K + 15) / 16;
  if (tiles >= 148 || k_steps < 16) return 1;
  int64_t s = (2 * 148 + tiles - 1) / tiles;
  if (s > k_steps / 8) s = k_steps / 8;
  while (s > 1 && s * M * N > partial_floats) --s;
  return s < 1 ? 1 : (int)s;
}

int run_affine(at_engine* e, const Affine& f, const float* W, const float* A, int64_t lda, int64_t M, float* C, int64_t ldc,
               const float* res, int64_t ldr, cudaStream_t st, bool a_zero_padded = false) {
  GemmArgs g;
  g.A = A; g.lda = lda; g.M = M; g.N = f.N;
  g.K = a_zero_padded ? f.Kpad : f.K;  // run over the zero tail when the input row has one (keeps float4 loads)
  g.Wt = W + f.wt; g.bias = W + f.bias;
  g.C = C; g.ldc = ldc; g.res = res; g.ldr = ldr; g.act = f.act;
  g.partial = e->partial;
  return launch_gemm(g, pick_splits(M, f.N, g.K, e->partial_floats), st, &e->last_launches);
}

int run_conv(at_engine* e, const ConvLayer& c, const float* W, const float* in, int64_t sb, int64_t sy, int64_t sx, int64_t sc,
             int64_t boards, float* out, int64_t out_batch_stride, const float* res, const float* post, int post_period,
             cudaStream_t st) {
  GemmArgs g;
  g.conv = 1;
  g.A = in; g.in_sb = sb; g.in_sy = sy; g.in_sx = sx; g.in_sc = sc;
  g.cin = c.cin; g.kh = c.kh; g.kw = c.kw; g.sh = c.sh; g.sw = c.sw; g.ph = c.ph; g.pw = c.pw;
  g.in_h = c.in_h; g.in_w = c.in_w; g.out_h = c.out_h; g.out_w = c.out_w;
  g.M = boards * c.out_h * c.out_w; g.N = c.cout; g.K = c.f.K;
  g.Wt = W + c.f.wt; g.bias = W + c.f.bias;
  g.C = out; g.ldc = c.cout;
  g.out_rows_per_batch = (int64_t)c.out_h * c.out_w;
  g.out_batch_stride = out_batch_stride;
  g.res = res; g.ldr = c.cout;
  g.post = post; g.post_period = post_period;
  g.act = c.f.act;
  return launch_gemm(g, 1, st, &e->last_launches);
}

// shared_fc[first_shared:] -> policy / value heads -> softmax + expected value (model.py:294-297, network.py:269-278)
int run_heads(at_engine* e, const float* W, const float* x, int64_t ld, bool zero_padded, size_t first_shared, int64_t n, float* probs,
              float* value, uint32_t* flags, cudaStream_t st) {
  const Plan& p = e->plan;
  const at_model_desc& d = p.d;
  const float* trunk = x;
  int64_t trunk_ld = ld;
  {
    float* pingpong[2] = {e->mlp_a, e->mlp_b};
    int w2 = 0;
    for (size_t i = first_shared; i < p.shared.size(); ++i) {
      float* out = (i + 1 == p.shared.size()) ? e->trunk : pingpong[w2];
      AT_TRY(run_affine(e, p.shared[i], W, trunk, trunk_ld, n, out, p.shared[i].N, nullptr, 0, st, zero_padded && i == 0));
      trunk = out; trunk_ld = p.shared[i].N; w2 ^= 1;
    }
  }
  auto head = [&](const std::vector<Affine>& layers, float* logits) -> int {
    const float* hx = trunk;
    int64_t hld = trunk_ld;
    float* pingpong[2] = {e->mlp_a, e->mlp_b};
    int w2 = 0;
    for (size_t i = 0; i < layers.size(); ++i) {
      float* out = (i + 1 == layers.size()) ? logits : pingpong[w2];
      AT_TRY(run_affine(e, layers[i], W, hx, hld, n, out, layers[i].N, nullptr, 0, st));
      hx = out; hld = layers[i].N; w2 ^= 1;
    }
    return AT_OK;
  };
  AT_TRY(head(p.policy, e->pol_logits));
  AT_TRY(head(p.value, e->val_logits));
  return launch_heads(e->pol_logits, e->val_logits, n, p.action_dim, d.num_atoms, d.v_min, d.v_max, probs, value, flags, st,
                      &e->last_launches);
}

// AlphaTriangleNet.forward for `n` boards (n <= chunk) on the fp32 path.
int forward_fp32_chunk(at_engine* e, const float* W, const float* grid, const float* other, int64_t n, float* probs, float* value,
                       uint32_t* flags, cudaStream_t st) {
  const Plan& p = e->plan;
  const at_model_desc& d = p.d;
  const int R = d.env.rows, Cc = d.env.cols;

  // ---- CNN trunk: conv_body, res_body (model.py:275-276) -------------------------------
  struct Img { const float* ptr; int64_t sb, sy, sx, sc; };
  Img cur{grid, (int64_t)d.grid_channels * R * Cc, Cc, 1, (int64_t)R * Cc};  // NCHW input
  float* pp[2] = {e->buf_p, e->buf_q};
  int which = 0;
  const bool tf = p.transformer;
  // where the final trunk activation goes: straight into the concat row (conv-only) or buf (transformer)
  const int n_trunk_layers = (int)p.convs.size() + (p.has_res_proj ? 1 : 0) + 2 * (int)p.res.size();
  int layer_no = 0;
  const float* pe_post = (tf && !p.has_input_proj) ? W + p.pe : nullptr;

  auto nhwc = [](const float* ptr, const ConvLayer& c) { return Img{ptr, (int64_t)c.out_h * c.out_w * c.cout, (int64_t)c.out_w * c.cout, c.cout, 1}; };
  auto plain_conv = [&](const ConvLayer& c) -> int {
    ++layer_no;
    const bool last = layer_no == n_trunk_layers;
    float* out = pp[which];
    int64_t obs = (int64_t)c.out_h * c.out_w * c.cout;
    if (last && !tf) { out = e->concat; obs = p.concat_ld; }
    AT_TRY(run_conv(e, c, W, cur.ptr, cur.sb, cur.sy, cur.sx, cur.sc, n, out, obs, nullptr, last ? pe_post : nullptr, p.L, st));
    cur = nhwc(out, c);
    if (last && !tf) cur.sb = obs;
    which ^= 1;
    return AT_OK;
  };
  for (const ConvLayer& c : p.convs) AT_TRY(plain_conv(c));
  if (p.has_res_proj) AT_TRY(plain_conv(p.res_proj));
  for (const ResBlock& rb : p.res) {
    // t = act(bn1(conv1(x)));  x' = act(bn2(conv2(t)) + x)   (model.py:52-60)
    const float* x = cur.ptr;  // NHWC, lives in pp[which ^ 1] (or is the NCHW input when there is no conv_body)
    if (cur.sc != 1) return fail(AT_E_UNSUPPORTED, "model: residual blocks directly on the input grid are not supported");
    ++layer_no;
    float* t = pp[which];
    AT_TRY(run_conv(e, rb.c1, W, cur.ptr, cur.sb, cur.sy, cur.sx, cur.sc, n, t, (int64_t)p.L * rb.c1.cout, nullptr, nullptr, 1, st));
    ++layer_no;
    const bool last = layer_no == n_trunk_layers;
    float* out = const_cast<float*>(x);  // in place over x: element (m,n) is read as residual and written by the same thread
    int64_t obs = (int64_t)p.L * rb.c2.cout;
    if (last && !tf) { out = e->concat; obs = p.concat_ld; }
    const Img ti = nhwc(t, rb.c1);
    // residual rows are addressed densely (m*cout + n): x is a dense NHWC buffer here
    AT_TRY(run_conv(e, rb.c2, W, ti.ptr, ti.sb, ti.sy, ti.sx, ti.sc, n, out, obs, x, last ? pe_post : nullptr, p.L, st));
    cur = nhwc(out, rb.c2);
    if (last && !tf) cur.sb = obs;
  }

  // ---- transformer (model.py:278-288) ---------------------------------------------------
  if (tf) {
    const int dm = d.t_dim;
    const int64_t rows = n * p.L;
    float* X = const_cast<float*>(cur.ptr);
    float* Y = (X == e->buf_p) ? e->buf_q : e->buf_p;
    if (p.has_input_proj) {
      // 1x1 conv with bias, then x + pe[:L]; token order l = h*W + w (model.py:285)
      GemmArgs g;
      g.A = X; g.lda = p.cnn_ch; g.M = rows; g.N = dm; g.K = p.input_proj.K;
      g.Wt = W + p.input_proj.wt; g.bias = W + p.input_proj.bias;
      g.C = Y; g.ldc = dm; g.post = W + p.pe; g.post_period = p.L;
      AT_TRY(launch_gemm(g, 1, st, &e->last_launches));
      float* t = X; X = Y; Y = t;
    }
    float* Rb = e->buf_r;
    for (const EncoderLayer& l : p.enc) {
      // x = x + out_proj(MHA(norm1(x)));  x = x + linear2(act(linear1(norm2(x))))   (norm_first=True)
      AT_TRY(launch_layernorm(X, W + l.ln1_w, W + l.ln1_b, Y, rows, dm, 1e-5f, 0, 0, dm, st, &e->last_launches));
      AT_TRY(run_affine(e, l.qkv, W, Y, dm, rows, Rb, 3 * dm, nullptr, 0, st));
      AT_TRY(launch_attention(Rb, Y, n, p.L, dm, d.t_heads, st, &e->last_launches));
      AT_TRY(run_affine(e, l.out, W, Y, dm, rows, X, dm, X, dm, st));
      AT_TRY(launch_layernorm(X, W + l.ln2_w, W + l.ln2_b, Y, rows, dm, 1e-5f, 0, 0, dm, st, &e->last_launches));
      AT_TRY(run_affine(e, l.ff1, W, Y, dm, rows, Rb, d.t_ff, nullptr, 0, st));
      AT_TRY(run_affine(e, l.ff2, W, Rb, d.t_ff, rows, X, dm, X, dm, st));
    }
    // final encoder norm, written token-major into the concat row: index l*d + j (model.py:288)
    AT_TRY(launch_layernorm(X, W + p.lnf_w, W + p.lnf_b, e->concat, rows, dm, 1e-5f, p.L, p.concat_ld, dm, st, &e->last_launches));
  }

  // ---- cat(other) -> shared_fc -> heads (model.py:292-297) ---------------------------------
  AT_TRY(launch_copy_rows(other, n, d.other_dim, e->concat + p.flat, p.concat_ld, st, &e->last_launches));
  return run_heads(e, W, e->concat, p.concat_ld, /*zero_padded=*/true, 0, n, probs, value, flags, st);
}

int begin_eval(at_engine* e, cudaStream_t st, const float** W) {
  if (e->active < 0) return fail(AT_E_STATE, "evaluate called before at_engine_load_weights");
  // a published set takes over once its fold / repack has finished; until then evaluations keep reading the old one
  if (e->pending >= 0 && cudaEventQuery(e->ready[e->pending]) == cudaSuccess) {
    e->active = e->pending;
    e->pending = -1;
  }
  AT_CUDA_TRY(cudaStreamWaitEvent(st, e->ready[e->active], 0));
  *W = e->folded[e->active];
  e->last_launches = 0;
  return AT_OK;
}

int end_eval(at_engine* e, cudaStream_t st) {
  AT_CUDA_TRY(cudaEventRecord(e->last_use[e->active], st));
  return AT_OK;
}

bool is_tc(int precision) { return precision == AT_PREC_BF16 || precision == AT_PREC_F16; }
int tc_format(int precision) { return precision == AT_PREC_F16 ? FMT_F16 : FMT_BF16; }

int check_precision(at_engine* e, int precision) {
  if (precision == AT_PREC_FP32) return AT_OK;
  if (is_tc(precision)) {
    if (e->bf16) return AT_OK;
    return fail(AT_E_UNSUPPORTED, "tensor-core path unavailable for this configuration: " + e->bf16_reason);
  }
  return fail(AT_E_INVALID, "unknown precision");
}

PrepBases prep_bases(at_engine* e, const float* d_image, int set);

// The 16-bit operand images of the active weight set are kept in ONE format. An evaluation that wants the other one
// re-packs them from the folded fp32 set on its own stream (two formats are not meant to alternate call by call: an
// engine normally serves one precision). Later weight loads pack for the format asked for last.
int ensure_format(at_engine* e, int fmt, cudaStream_t st) {
  e->tc_fmt = fmt;
  if (bf16_packed_format(e->bf16, e->active) == fmt) return AT_OK;
  AT_CUDA_TRY(cudaStreamWaitEvent(st, e->last_use[e->active], 0));  // evaluations still reading the old images
  const PrepBases bases = prep_bases(e, nullptr, e->active);
  AT_TRY(prep_run(e->prep, bases, fmt, st, nullptr, /*first_phase=*/1));
  bf16_set_packed_format(e->bf16, e->active, fmt);
  AT_CUDA_TRY(cudaEventRecord(e->ready[e->active], st));
  return AT_OK;
}

PrepBases prep_bases(at_engine* e, const float* d_image, int set) {
  PrepBases bases{};
  bases.p[PB_IMAGE] = const_cast<float*>(d_image);
  bases.p[PB_FOLDED] = e->folded[set];
  if (e->bf16) bf16_prep_bases(e->bf16, set, &bases);
  bases.p[PB_FRAGS] = e->heads_frag[set];
  return bases;
}

}  // namespace

extern "C" {

int at_abi_version(void) { return AT_ABI_VERSION; }
const char* at_last_error(void) { return g_last_error.c_str(); }

int at_extract_features(const at_env_desc* env, int grid_channels, int other_dim, const void* d_states, int64_t batch, float* d_grid,
                        float* d_other, void* stream) {
  if (!env) return fail(AT_E_INVALID, "at_extract_features: null env");
  return launch_features(*env, grid_channels, other_dim, d_states, batch, d_grid, d_other, static_cast<cudaStream_t>(stream));
}

int64_t at_model_describe(const at_model_desc* desc, char* buf, int64_t cap) {
  if (!desc) return fail(AT_E_INVALID, "at_model_describe: null desc");
  Plan plan;
  const int rc = build_plan(*desc, plan);
  if (rc != AT_OK) return rc;
  std::string text;
  for (const ImageEntry& en : plan.image) text += en.name + " " + std::to_string(en.offset) + " " + std::to_string(en.numel) + "\n";
  if (buf && cap > 0) {
    const int64_t n = (int64_t)text.size() < cap - 1 ? (int64_t)text.size() : cap - 1;
    std::memcpy(buf, text.data(), (size_t)n);
    buf[n] = 0;
  }
  return (int64_t)text.size();
}

// The fused head kernels' view of weight set `slot` (W = its folded image). Only for the one-shared-layer / one-hidden-
// layer-per-head shape (the reference's defaults); the caller checks that first.
static bool heads_shape_fused(const Plan& p) { return p.shared.size() == 1 && p.policy.size() == 2 && p.value.size() == 2; }

static HeadsParams heads_params(const at_engine* e, const float* W, int slot) {
  const Plan& p = e->plan;
  const Affine& fc = p.shared[0];
  HeadsParams hp{};
  hp.other_dim = p.d.other_dim;
  hp.w_other = W + fc.wt + (int64_t)p.flat * fc.N; hp.trunk_bias = W + fc.bias; hp.trunk_n = fc.N; hp.act = fc.act;
  hp.wp1 = W + p.policy[0].wt; hp.bp1 = W + p.policy[0].bias; hp.wp2 = W + p.policy[1].wt; hp.bp2 = W + p.policy[1].bias;
  hp.wv1 = W + p.value[0].wt; hp.bv1 = W + p.value[0].bias; hp.wv2 = W + p.value[1].wt; hp.bv2 = W + p.value[1].bias;
  hp.pol_hidden = p.policy[0].N; hp.val_hidden = p.value[0].N; hp.actions = p.action_dim; hp.atoms = p.d.num_atoms;
  hp.v_min = p.d.v_min; hp.v_max = p.d.v_max;
  hp.frags = e->heads_frag[slot];
  if (e->bf16 && W) {
    hp.fmt = bf16_packed_format(e->bf16, slot);
    if (const float* fb = bf16_fc_bias(e->bf16, slot)) hp.trunk_bias = fb;  // encoder's final LayerNorm folded in
  }
  return hp;
}

// Head widths decide the kernel: heads_mma_kernel (warp-level MMAs on fragment-packed weights) when the trunk and both
// hidden layers are 64 or 128 wide, else heads_fused_kernel (same fusion on CUDA cores); deeper heads take the unfused tail.
static bool heads_use_mma(const HeadsParams& hp) { return hp.frags && heads_mma_supported(hp); }
static bool heads_take_states(const at_engine* e, const HeadsParams& hp) {
  return heads_use_mma(hp) && heads_fused_supported(hp) && e->plan.d.env.cols <= 32;
}

// Trunk finish + heads + softmax epilogues in one launch: tensor-core variant when the packed fragments exist.
static int run_heads_fused(at_engine* e, HeadsParams& hp, cudaStream_t st) {
  if (heads_use_mma(hp)) return launch_heads_mma(hp, st, &e->last_launches);
  return launch_heads_fused(hp, st, &e->last_launches);
}

int at_engine_create(const at_model_desc* desc, int device, int64_t max_batch, at_engine** out) {
  if (!desc || !out) return fail(AT_E_INVALID, "at_engine_create: null argument");
  *out = nullptr;
  if (max_batch <= 0) return fail(AT_E_INVALID, "at_engine_create: max_batch must be positive");
  int n_dev = 0;
  AT_CUDA_TRY(cudaGetDeviceCount(&n_dev));
  if (device < 0 || device >= n_dev) return fail(AT_E_INVALID, "at_engine_create: no such CUDA device");
  cudaDeviceProp prop;
  AT_CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) return fail(AT_E_UNSUPPORTED, std::string("at_engine_create: this library is built for sm_100a (B200); device is ") + prop.name);

  at_engine* e = new (std::nothrow) at_engine();
  if (!e) return fail(AT_E_INVALID, "at_engine_create: out of host memory");
  int rc = build_plan(*desc, e->plan);
  if (rc != AT_OK) { delete e; return rc; }
  e->device = device;
  e->max_batch = max_batch;
  e->chunk = max_batch < 2048 ? max_batch : 2048;
  DeviceGuard guard(device);
  if (!guard.ok) { delete e; return fail(AT_E_CUDA, "at_engine_create: cudaSetDevice failed"); }

  const Plan& p = e->plan;
  const int64_t S = e->chunk;
  auto alloc_all = [&]() -> int {
    for (int i = 0; i < 2; ++i) {
      AT_TRY(dev_alloc(&e->folded[i], p.folded_floats));
      AT_CUDA_TRY(cudaEventCreateWithFlags(&e->ready[i], cudaEventDisableTiming));
      AT_CUDA_TRY(cudaEventCreateWithFlags(&e->last_use[i], cudaEventDisableTiming));
    }
    AT_TRY(dev_alloc(&e->buf_p, S * p.unit_pq));
    AT_TRY(dev_alloc(&e->buf_q, S * p.unit_pq));
    AT_TRY(dev_alloc(&e->buf_r, S * p.unit_r));
    AT_TRY(dev_alloc(&e->concat, S * p.concat_ld));
    AT_CUDA_TRY(cudaMemset(e->concat, 0, (size_t)S * p.concat_ld * sizeof(float)));  // pad columns stay zero forever
    const int64_t mlp_w = p.max_mlp > 0 ? p.max_mlp : 1;
    const int64_t H = max_batch;  // the heads run on whole batches (the bf16 trunk is not sliced)
    AT_TRY(dev_alloc(&e->mlp_a, H * mlp_w));
    AT_TRY(dev_alloc(&e->mlp_b, H * mlp_w));
    AT_TRY(dev_alloc(&e->trunk, H * mlp_w));
    AT_TRY(dev_alloc(&e->pol_logits, H * p.action_dim));
    AT_TRY(dev_alloc(&e->val_logits, H * (int64_t)p.d.num_atoms));
    e->partial_floats = (int64_t)8192 * (p.max_n > 0 ? p.max_n : 1);
    AT_TRY(dev_alloc(&e->partial, e->partial_floats));
    AT_TRY(dev_alloc(&e->feat_grid, S * (int64_t)p.d.grid_channels * p.d.env.rows * p.d.env.cols));
    AT_TRY(dev_alloc(&e->feat_other, S * (int64_t)p.d.other_dim));
    AT_CUDA_TRY(cudaMalloc(&e->d_states, (size_t)max_batch * AT_STATE_BYTES));
    AT_TRY(dev_alloc(&e->d_policy, max_batch * (int64_t)p.action_dim));
    AT_TRY(dev_alloc(&e->d_value, max_batch));
    AT_TRY(dev_alloc(&e->d_flags, 4));
    AT_CUDA_TRY(cudaMemset(e->d_flags, 0, 4 * sizeof(uint32_t)));
    AT_CUDA_TRY(cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking));
    AT_CUDA_TRY(cudaStreamCreateWithFlags(&e->h2d_stream, cudaStreamNonBlocking));
    for (auto& sl : e->slot) {
      AT_CUDA_TRY(cudaMalloc(&sl.d_states, (size_t)max_batch * AT_STATE_BYTES));
      AT_TRY(dev_alloc(&sl.d_policy, max_batch * (int64_t)p.action_dim));
      AT_TRY(dev_alloc(&sl.d_value, max_batch));
      AT_TRY(dev_alloc(&sl.d_flags, 4));
      AT_CUDA_TRY(cudaEventCreateWithFlags(&sl.staged, cudaEventDisableTiming));
      AT_CUDA_TRY(cudaEventCreateWithFlags(&sl.computed, cudaEventDisableTiming));
      AT_CUDA_TRY(cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
    }
    for (auto& ev : e->slice_done) AT_CUDA_TRY(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    AT_CUDA_TRY(cudaEventCreateWithFlags(&e->copies_done, cudaEventDisableTiming));
    return AT_OK;
  };
  rc = alloc_all();
  if (rc != AT_OK) { at_engine_destroy(e); return rc; }
  rc = bf16_create(p, max_batch, &e->bf16, &e->bf16_reason);
  if (rc != AT_OK) { at_engine_destroy(e); return rc; }
  if (e->bf16 && heads_shape_fused(p)) {
    const HeadsParams shape = heads_params(e, nullptr, 0);
    if (heads_mma_supported(shape)) {
      for (int i = 0; i < 2 && rc == AT_OK; ++i) rc = dev_alloc(&e->heads_frag[i], (int64_t)heads_frag_count(shape));
      if (rc != AT_OK) { at_engine_destroy(e); return rc; }
    }
  }
  {
    PrepList jobs;
    fold_add_prep_jobs(p, jobs);
    if (e->bf16) bf16_add_prep_jobs(e->bf16, p, jobs);
    if (e->heads_frag[0]) {
      const int64_t src[4] = {p.policy[0].wt, p.value[0].wt, p.policy[1].wt, p.value[1].wt};
      heads_add_prep_jobs(heads_params(e, nullptr, 0), src, jobs);
    }
    rc = prep_upload(jobs, &e->prep);
    if (rc != AT_OK) { at_engine_destroy(e); return rc; }
  }
  AT_CUDA_TRY(cudaDeviceSynchronize());
  *out = e;
  return AT_OK;
}

int at_engine_destroy(at_engine* e) {
  if (!e) return AT_OK;
  DeviceGuard guard(e->device);
  cudaDeviceSynchronize();
  bf16_destroy(e->bf16);
  prep_free(&e->prep);
  for (int i = 0; i < 2; ++i) {
    cudaFree(e->folded[i]);
    cudaFree(e->heads_frag[i]);
    if (e->ready[i]) cudaEventDestroy(e->ready[i]);
    if (e->last_use[i]) cudaEventDestroy(e->last_use[i]);
  }
  float* bufs[] = {e->buf_p, e->buf_q, e->buf_r, e->concat, e->mlp_a, e->mlp_b, e->trunk, e->pol_logits, e->val_logits,
                   e->partial, e->feat_grid, e->feat_other, e->d_policy, e->d_value};
  for (float* b : bufs) cudaFree(b);
  cudaFree(e->d_states);
  cudaFree(e->d_flags);
  if (e->copy_stream) cudaStreamDestroy(e->copy_stream);
  if (e->h2d_stream) cudaStreamDestroy(e->h2d_stream);
  for (auto& sl : e->slot) {
    cudaFree(sl.d_states);
    cudaFree(sl.d_policy);
    cudaFree(sl.d_value);
    cudaFree(sl.d_flags);
    if (sl.staged) cudaEventDestroy(sl.staged);
    if (sl.computed) cudaEventDestroy(sl.computed);
    if (sl.done) cudaEventDestroy(sl.done);
  }
  for (auto& ev : e->slice_done)
    if (ev) cudaEventDestroy(ev);
  if (e->copies_done) cudaEventDestroy(e->copies_done);
  delete e;
  return AT_OK;
}

int64_t at_engine_weight_floats(const at_engine* e) { return e ? e->plan.image_floats : 0; }
int at_engine_weight_entries(const at_engine* e) { return e ? (int)e->plan.image.size() : 0; }
int at_engine_weight_entry(const at_engine* e, int index, const char** name, int64_t* offset, int64_t* numel) {
  if (!e || index < 0 || index >= (int)e->plan.image.size()) return fail(AT_E_INVALID, "at_engine_weight_entry: index out of range");
  const ImageEntry& en = e->plan.image[index];
  if (name) *name = en.name.c_str();
  if (offset) *offset = en.offset;
  if (numel) *numel = en.numel;
  return AT_OK;
}

static int fill_weight_set(at_engine* e, const float* d_image, int64_t n_floats, int64_t version, void* stream, bool deferred,
                           const char* who) {
  if (!e || !d_image) return fail(AT_E_INVALID, std::string(who) + ": null argument");
  if (n_floats != e->plan.image_floats)
    return fail(AT_E_INVALID, std::string(who) + ": image has " + std::to_string(n_floats) + " floats, model needs " +
                                  std::to_string(e->plan.image_floats));
  DeviceGuard guard(e->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int target = e->active < 0 ? 0 : 1 - e->active;
  // evaluations that still read the set about to be overwritten must drain first; so must an earlier publication into it
  AT_CUDA_TRY(cudaStreamWaitEvent(st, e->last_use[target], 0));
  if (e->pending == target) AT_CUDA_TRY(cudaStreamWaitEvent(st, e->ready[target], 0));
  AT_TRY(prep_run(e->prep, prep_bases(e, d_image, target), e->tc_fmt, st, nullptr));
  if (e->bf16) bf16_set_packed_format(e->bf16, target, e->tc_fmt);
  AT_CUDA_TRY(cudaEventRecord(e->ready[target], st));
  e->version[target] = version;
  if (deferred && e->active >= 0) {
    e->pending = target;
  } else {
    e->active = target;
    e->pending = -1;
  }
  return AT_OK;
}

int at_engine_load_weights(at_engine* e, const float* d_image, int64_t n_floats, int64_t version, void* stream) {
  return fill_weight_set(e, d_image, n_floats, version, stream, /*deferred=*/false, "at_engine_load_weights");
}

int at_engine_publish_weights(at_engine* e, const float* d_image, int64_t n_floats, int64_t version, void* stream) {
  return fill_weight_set(e, d_image, n_floats, version, stream, /*deferred=*/true, "at_engine_publish_weights");
}

int64_t at_engine_weight_version(const at_engine* e) { return (e && e->active >= 0) ? e->version[e->active] : -1; }

int at_engine_supports(const at_engine* e, int precision) {
  if (!e) return 0;
  if (precision == AT_PREC_FP32) return 1;
  if (is_tc(precision)) {
    if (e->bf16) return 1;
    set_error(e->bf16_reason);
    return 0;
  }
  return 0;
}

int at_engine_action_dim(const at_engine* e) { return e ? e->plan.action_dim : 0; }
int64_t at_engine_last_launches(const at_engine* e) { return e ? e->last_launches : 0; }

int at_engine_set_kernel_timing(at_engine* e, int on) {
  if (!e) return fail(AT_E_INVALID, "at_engine_set_kernel_timing: null engine");
  if (!e->bf16) return fail(AT_E_UNSUPPORTED, "at_engine_set_kernel_timing: the tensor-core path is not available");
  bf16_set_timing(e->bf16, on != 0);
  return AT_OK;
}

int at_engine_kernel_ms(const at_engine* e, float* tower_ms, float* fc_ms) {
  if (!e || !tower_ms || !fc_ms) return fail(AT_E_INVALID, "at_engine_kernel_ms: null argument");
  if (!e->bf16) return fail(AT_E_UNSUPPORTED, "at_engine_kernel_ms: the bf16 path is not available");
  DeviceGuard guard(e->device);
  return bf16_kernel_ms(e->bf16, tower_ms, fc_ms);
}

int at_engine_bf16_trunk(at_engine* e, const void* d_states, int64_t batch, float* d_trunk, void* stream) {
  if (!e || !d_states || !d_trunk) return fail(AT_E_INVALID, "at_engine_bf16_trunk: null argument");
  if (batch <= 0 || batch > e->max_batch) return fail(AT_E_INVALID, "at_engine_bf16_trunk: batch out of range");
  AT_TRY(check_precision(e, AT_PREC_BF16));
  DeviceGuard guard(e->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const float* W = nullptr;
  AT_TRY(begin_eval(e, st, &W));
  AT_TRY(ensure_format(e, e->tc_fmt, st));
  int splits = 0;
  AT_TRY(bf16_trunk_partials(e->bf16, e->plan, W, e->active, d_states, batch, &splits, st, &e->last_launches));
  AT_TRY(bf16_finish_trunk(e->bf16, e->plan, W, e->active, splits, batch, d_trunk, st, &e->last_launches));
  return end_eval(e, st);
}

int at_engine_evaluate_packed_static(at_engine* e, const void* d_states, int64_t batch, float* d_policy, float* d_value,
                                     uint32_t* d_flags, void* stream) {
  if (!e) return fail(AT_E_INVALID, "at_engine_evaluate_packed_static: null engine");
  if (batch < 0) return fail(AT_E_INVALID, "at_engine_evaluate_packed_static: negative batch");
  if (batch == 0) return AT_OK;
  if (!d_states || !d_policy || !d_value || !d_flags) return fail(AT_E_INVALID, "at_engine_evaluate_packed_static: null pointer");
  if ((reinterpret_cast<uintptr_t>(d_states) & 15u) != 0)
    return fail(AT_E_INVALID, "at_engine_evaluate_packed_static: d_states must be 16-byte aligned");
  AT_TRY(check_precision(e, AT_PREC_BF16));
  const Plan& p = e->plan;
  if (!(p.shared.size() == 1 && p.policy.size() == 2 && p.value.size() == 2))
    return fail(AT_E_UNSUPPORTED, "at_engine_evaluate_packed_static: needs one shared FC layer and one hidden layer per head");
  // every refusal happens before anything is queued: a failed call leaves no kernels behind and the outputs untouched
  if (!heads_fused_supported(heads_params(e, nullptr, 0)))
    return fail(AT_E_UNSUPPORTED, "at_engine_evaluate_packed_static: head widths outside what the fused head kernel covers");
  AT_TRY(bf16_static_check(e->bf16, p, batch));
  DeviceGuard guard(e->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const float* W = nullptr;
  AT_TRY(begin_eval(e, st, &W));
  AT_TRY(ensure_format(e, e->tc_fmt, st));
  int splits = 0;
  const float *partial = nullptr, *other = nullptr;
  int rc = bf16_trunk_partials_static(e->bf16, p, W, e->active, d_states, batch, &splits, &partial, &other, st, &e->last_launches);
  if (rc == AT_OK) {
    HeadsParams hp = heads_params(e, W, e->active);
    hp.partial = partial; hp.splits = splits; hp.other = other; hp.batch = batch;
    hp.probs = d_policy; hp.value = d_value; hp.flags = d_flags; hp.status = bf16_status_ptr(e->bf16);
    rc = run_heads_fused(e, hp, st);
  }
  const int rc_end = end_eval(e, st);  // also after a failed launch: whatever was queued still reads this weight set
  return rc != AT_OK ? rc : rc_end;
}

int at_engine_forward(at_engine* e, int precision, const float* d_grid, const float* d_other, int64_t batch, float* d_policy,
                      float* d_value, uint32_t* d_flags, void* stream) {
  if (!e) return fail(AT_E_INVALID, "at_engine_forward: null engine");
  if (batch < 0) return fail(AT_E_INVALID, "at_engine_forward: negative batch");
  if (batch == 0) return AT_OK;
  if (!d_grid || !d_other || !d_policy || !d_value || !d_flags) return fail(AT_E_INVALID, "at_engine_forward: null pointer");
  if (precision != AT_PREC_FP32)
    return fail(AT_E_UNSUPPORTED, "at_engine_forward takes prepared fp32 feature tensors and runs the fp32 path only; "
                                  "the bf16 path starts from packed states (at_engine_evaluate_packed)");
  DeviceGuard guard(e->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const float* W = nullptr;
  AT_TRY(begin_eval(e, st, &W));
  const Plan& p = e->plan;
  const int64_t grid_per = (int64_t)p.d.grid_channels * p.d.env.rows * p.d.env.cols;
  AT_TRY(launch_finite_check(d_grid, batch, grid_per, grid_per, AT_FLAG_NONFINITE_GRID, d_flags, st, &e->last_launches));
  AT_TRY(launch_finite_check(d_other, batch, p.d.other_dim, p.d.other_dim, AT_FLAG_NONFINITE_OTHER, d_flags, st, &e->last_launches));
  for (int64_t at0 = 0; at0 < batch; at0 += e->chunk) {
    const int64_t n = batch - at0 < e->chunk ? batch - at0 : e->chunk;
    AT_TRY(forward_fp32_chunk(e, W, d_grid + at0 * grid_per, d_other + at0 * p.d.other_dim, n, d_policy + at0 * p.action_dim,
                              d_value + at0, d_flags, st));
  }
  return end_eval(e, st);
}

int at_engine_evaluate_packed(at_engine* e, int precision, const void* d_states, int64_t batch, float* d_policy, float* d_value,
                              uint32_t* d_flags, void* stream) {
  if (!e) return fail(AT_E_INVALID, "at_engine_evaluate_packed: null engine");
  if (batch < 0) return fail(AT_E_INVALID, "at_engine_evaluate_packed: negative batch");
  if (batch == 0) return AT_OK;
  if (!d_states || !d_policy || !d_value || !d_flags) return fail(AT_E_INVALID, "at_engine_evaluate_packed: null pointer");
  if ((reinterpret_cast<uintptr_t>(d_states) & 15u) != 0) return fail(AT_E_INVALID, "at_engine_evaluate_packed: d_states must be 16-byte aligned");
  AT_TRY(check_precision(e, precision));
  if (is_tc(precision) && batch > e->max_batch)
    return fail(AT_E_INVALID, "at_engine_evaluate_packed: batch exceeds the engine's max_batch");
  DeviceGuard guard(e->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const float* W = nullptr;
  AT_TRY(begin_eval(e, st, &W));
  const Plan& p = e->plan;
  if (is_tc(precision)) {
    AT_TRY(ensure_format(e, tc_format(precision), st));
    int splits = 0;
    // the tensor-core head kernel derives the other-feature vectors itself: no separate feature launch on this path
    const bool heads_own_other = heads_shape_fused(p) && heads_take_states(e, heads_params(e, W, e->active));
    AT_TRY(bf16_trunk_partials(e->bf16, p, W, e->active, d_states, batch, &splits, st, &e->last_launches, !heads_own_other));
    if (heads_shape_fused(p)) {
      HeadsParams hp = heads_params(e, W, e->active);
      hp.partial = bf16_partial(e->bf16); hp.splits = splits; hp.other = bf16_other(e->bf16); hp.batch = batch;
      if (heads_own_other) {
        hp.states = static_cast<const PackedState*>(d_states);
        hp.env_rows = p.d.env.rows; hp.env_cols = p.d.env.cols; hp.env_slots = p.d.env.slots;
      }
      hp.probs = d_policy; hp.value = d_value; hp.flags = d_flags; hp.status = bf16_status_ptr(e->bf16);
      if (heads_fused_supported(hp)) {
        AT_TRY(run_heads_fused(e, hp, st));
        return end_eval(e, st);
      }
    }
    float* trunk1 = p.shared.size() == 1 ? e->trunk : e->mlp_b;
    AT_TRY(bf16_finish_trunk(e->bf16, p, W, e->active, splits, batch, trunk1, st, &e->last_launches));
    AT_TRY(run_heads(e, W, trunk1, p.shared[0].N, false, 1, batch, d_policy, d_value, d_flags, st));
    AT_TRY(launch_merge_status(bf16_status_ptr(e->bf16), d_flags, st, &e->last_launches));
    return end_eval(e, st);
  }
  const char* states = static_cast<const char*>(d_states);
  for (int64_t at0 = 0; at0 < batch; at0 += e->chunk) {
    const int64_t n = batch - at0 < e->chunk ? batch - at0 : e->chunk;
    AT_TRY(launch_features(p.d.env, p.d.grid_channels, p.d.other_dim, states + at0 * AT_STATE_BYTES, n, e->feat_grid, e->feat_other, st));
    ++e->last_launches;
    AT_TRY(forward_fp32_chunk(e, W, e->feat_grid, e->feat_other, n, d_policy + at0 * p.action_dim, d_value + at0, d_flags, st));
  }
  return end_eval(e, st);
}

// Device-visible address of a pinned (page-locked, mapped) host buffer, or null for pageable memory.
static void* pinned_device_ptr(const void* h) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, h) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return a.type == cudaMemoryTypeHost ? a.devicePointer : nullptr;
}

constexpr int64_t kZeroCopyMaxBatch = 256;

int at_engine_evaluate_host(at_engine* e, int precision, const void* h_states, int64_t batch, float* h_policy, float* h_value,
                            uint32_t* h_flags, void* stream) {
  if (!e) return fail(AT_E_INVALID, "at_engine_evaluate_host: null engine");
  if (batch < 0) return fail(AT_E_INVALID, "at_engine_evaluate_host: negative batch");
  if (h_flags) *h_flags = 0;
  if (batch == 0) return AT_OK;
  if (batch > e->max_batch) return fail(AT_E_INVALID, "at_engine_evaluate_host: batch exceeds the engine's max_batch");
  if (!h_states || !h_policy || !h_value || !h_flags) return fail(AT_E_INVALID, "at_engine_evaluate_host: null pointer");
  DeviceGuard guard(e->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (batch <= kZeroCopyMaxBatch) {
    // The leaf batches one search sends (<= 32 states by default, mcts_config.py:14) are latency-, not bandwidth-bound: with
    // pinned host buffers the kernels read the records and write probabilities / values straight through PCIe, which
    // removes three copy-engine round trips from the call. Only the 4-byte status word is still copied.
    void* ds = pinned_device_ptr(h_states);
    void* dp = pinned_device_ptr(h_policy);
    void* dv = pinned_device_ptr(h_value);
    if (ds && dp && dv && (reinterpret_cast<uintptr_t>(ds) & 15u) == 0) {
      AT_CUDA_TRY(cudaMemsetAsync(e->d_flags, 0, sizeof(uint32_t), st));
      AT_TRY(at_engine_evaluate_packed(e, precision, ds, batch, static_cast<float*>(dp), static_cast<float*>(dv), e->d_flags, stream));
      AT_CUDA_TRY(cudaMemcpyAsync(h_flags, e->d_flags, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
      AT_CUDA_TRY(cudaStreamSynchronize(st));
      return AT_OK;
    }
  }
  AT_CUDA_TRY(cudaMemcpyAsync(e->d_states, h_states, (size_t)batch * AT_STATE_BYTES, cudaMemcpyHostToDevice, st));
  AT_CUDA_TRY(cudaMemsetAsync(e->d_flags, 0, sizeof(uint32_t), st));
  // Slices: the device->host copy of slice i (5.9 MB per 4096 states) overlaps the kernels of slice i+1. Only worth it
  // when a slice still fills the GPU for several waves (measured: slicing a 4096 batch four ways costs 40%).
  const int A = e->plan.action_dim;
  const int n_slices = batch >= 16384 ? 4 : (batch >= 8192 ? 2 : 1);
  const int64_t per = ((batch + n_slices - 1) / n_slices + 1) & ~int64_t(1);  // even, so board pairs stay aligned
  int64_t launches = 0;
  int used = 0;
  for (int64_t at0 = 0; at0 < batch; at0 += per, ++used) {
    const int64_t n = batch - at0 < per ? batch - at0 : per;
    AT_TRY(at_engine_evaluate_packed(e, precision, static_cast<const char*>(e->d_states) + at0 * AT_STATE_BYTES, n,
                                     e->d_policy + at0 * A, e->d_value + at0, e->d_flags, stream));
    launches += e->last_launches;
    const bool last = at0 + per >= batch;
    cudaStream_t cs = last ? st : e->copy_stream;
    if (!last) {
      AT_CUDA_TRY(cudaEventRecord(e->slice_done[used], st));
      AT_CUDA_TRY(cudaStreamWaitEvent(cs, e->slice_done[used], 0));
    }
    AT_CUDA_TRY(cudaMemcpyAsync(h_policy + at0 * A, e->d_policy + at0 * A, (size_t)n * A * sizeof(float), cudaMemcpyDeviceToHost, cs));
    AT_CUDA_TRY(cudaMemcpyAsync(h_value + at0, e->d_value + at0, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, cs));
  }
  e->last_launches = launches;
  AT_CUDA_TRY(cudaMemcpyAsync(h_flags, e->d_flags, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
  AT_CUDA_TRY(cudaEventRecord(e->copies_done, e->copy_stream));
  AT_CUDA_TRY(cudaStreamSynchronize(st));
  AT_CUDA_TRY(cudaEventSynchronize(e->copies_done));
  return AT_OK;
}

int at_engine_submit_host(at_engine* e, int precision, const void* h_states, int64_t batch, float* h_policy, float* h_value,
                          uint32_t* h_flags, int slot, void* stream) {
  if (!e) return fail(AT_E_INVALID, "at_engine_submit_host: null engine");
  if (slot < 0 || slot > 1) return fail(AT_E_INVALID, "at_engine_submit_host: slot must be 0 or 1");
  if (batch <= 0 || batch > e->max_batch) return fail(AT_E_INVALID, "at_engine_submit_host: batch must be in 1..max_batch");
  if (!h_states || !h_policy || !h_value || !h_flags) return fail(AT_E_INVALID, "at_engine_submit_host: null pointer");
  at_engine::HostSlot& sl = e->slot[slot];
  if (sl.busy) return fail(AT_E_INVALID, "at_engine_submit_host: slot still in flight (call at_engine_wait_host first)");
  DeviceGuard guard(e->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // records go up on their own stream, so they do not queue behind the previous batch's kernels
  AT_CUDA_TRY(cudaMemcpyAsync(sl.d_states, h_states, (size_t)batch * AT_STATE_BYTES, cudaMemcpyHostToDevice, e->h2d_stream));
  AT_CUDA_TRY(cudaEventRecord(sl.staged, e->h2d_stream));
  AT_CUDA_TRY(cudaStreamWaitEvent(st, sl.staged, 0));
  AT_CUDA_TRY(cudaMemsetAsync(sl.d_flags, 0, sizeof(uint32_t), st));
  AT_TRY(at_engine_evaluate_packed(e, precision, sl.d_states, batch, sl.d_policy, sl.d_value, sl.d_flags, stream));
  sl.launches = e->last_launches;
  AT_CUDA_TRY(cudaEventRecord(sl.computed, st));
  // results drain on the copy stream while the compute stream is free for the other slot's kernels
  AT_CUDA_TRY(cudaStreamWaitEvent(e->copy_stream, sl.computed, 0));
  const int A = e->plan.action_dim;
  AT_CUDA_TRY(cudaMemcpyAsync(h_policy, sl.d_policy, (size_t)batch * A * sizeof(float), cudaMemcpyDeviceToHost, e->copy_stream));
  AT_CUDA_TRY(cudaMemcpyAsync(h_value, sl.d_value, (size_t)batch * sizeof(float), cudaMemcpyDeviceToHost, e->copy_stream));
  AT_CUDA_TRY(cudaMemcpyAsync(h_flags, sl.d_flags, sizeof(uint32_t), cudaMemcpyDeviceToHost, e->copy_stream));
  AT_CUDA_TRY(cudaEventRecord(sl.done, e->copy_stream));
  sl.busy = true;
  return AT_OK;
}

int at_engine_wait_host(at_engine* e, int slot) {
  if (!e) return fail(AT_E_INVALID, "at_engine_wait_host: null engine");
  if (slot < 0 || slot > 1) return fail(AT_E_INVALID, "at_engine_wait_host: slot must be 0 or 1");
  at_engine::HostSlot& sl = e->slot[slot];
  if (!sl.busy) return fail(AT_E_INVALID, "at_engine_wait_host: nothing was submitted on this slot");
  DeviceGuard guard(e->device);
  sl.busy = false;
  AT_CUDA_TRY(cudaEventSynchronize(sl.done));
  e->last_launches = sl.launches;
  return AT_OK;
}

}  // extern "C"

#ifdef AT_EXPERIMENTS
namespace at { int debug_read_timeline(unsigned long long* out, int n); }
extern "C" int at_debug_read_timeline(unsigned long long* out, int n) { return at::debug_read_timeline(out, n); }
#endif
