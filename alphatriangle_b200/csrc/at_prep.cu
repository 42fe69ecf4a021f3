// Weight preparation on the device: reference fp32 state_dict image -> folded set.
//
// Eval-mode BatchNorm (y = (x - mean) / sqrt(var + 1e-5) * gamma + beta, applied
// after a bias-free conv or a biased Linear: alphatriangle/nn/model.py:24-38,
// 232-266) is folded into the preceding affine map; weights are transposed to
// [K,N], permuted to the NHWC / tap-major K order the kernels use, and padded
// with zero rows up to Kpad.

#include <algorithm>

#include "at_engine.cuh"
#include "at_prepjobs.cuh"

namespace at {

namespace {

constexpr float kBnEps = 1e-5f;  // nn.BatchNorm{1,2}d default, never overridden by the reference

// ---- the job bodies: block `rank` of `nb` blocks of one job, 256 threads each --------------------------------------

// PK_FOLD   a: src_w, src_b, bn_g, bn_b, bn_m, bn_v (image, -1 = absent), wt, bias (folded)   i: K, Kpad, N, perm_cols, perm_inner
__device__ void job_fold(const PrepJob& j, const float* __restrict__ img, float* __restrict__ folded, int rank, int nb) {
  const int64_t src_w = j.a[0], src_b = j.a[1], bn_g = j.a[2], bn_b = j.a[3], bn_m = j.a[4], bn_v = j.a[5];
  float* __restrict__ wt = folded + j.a[6];
  float* __restrict__ bias = folded + j.a[7];
  const int K = j.i[0], Kpad = j.i[1], N = j.i[2], perm_cols = j.i[3], perm_inner = j.i[4];
  const int64_t total = (int64_t)Kpad * N;
  const int perm_outer = perm_inner > 0 ? perm_cols / perm_inner : 0;
  for (int64_t i = (int64_t)rank * blockDim.x + threadIdx.x; i < total; i += (int64_t)nb * blockDim.x) {
    const int row = (int)(i / N), n = (int)(i - (int64_t)row * N);
    float v = 0.0f;
    if (row < K) {
      int k_ref = row;
      if (row < perm_cols) {
        const int inner = row / perm_outer, outer = row - inner * perm_outer;  // row = inner*outer_count + outer
        k_ref = outer * perm_inner + inner;
      }
      v = img[src_w + (int64_t)n * K + k_ref];
      if (bn_g >= 0) v *= img[bn_g + n] / sqrtf(img[bn_v + n] + kBnEps);
    }
    wt[i] = v;
  }
  for (int n = rank * blockDim.x + threadIdx.x; n < N; n += nb * blockDim.x) {
    float bv = src_b >= 0 ? img[src_b + n] : 0.0f;
    if (bn_g >= 0) {
      const float scale = img[bn_g + n] / sqrtf(img[bn_v + n] + kBnEps);
      bv = (bv - img[bn_m + n]) * scale + img[bn_b + n];
    }
    bias[n] = bv;
  }
}

// PK_COPY   a: src offset, dst offset (floats)   i: n, src base, dst base   f: scale
__device__ void job_copy(const PrepJob& j, const PrepBases& b, int rank, int nb) {
  const float* __restrict__ src = static_cast<const float*>(b.p[j.i[1]]) + j.a[0];
  float* __restrict__ dst = static_cast<float*>(b.p[j.i[2]]) + j.a[1];
  const float scale = j.f[0];
  for (int i = rank * blockDim.x + threadIdx.x; i < j.i[0]; i += nb * blockDim.x) dst[i] = src[i] * scale;
}

// PK_PACK_B   fp32 Wt [K,N] (folded + a0) -> 16-bit core-matrix tiles [K/8][N][8] (base i2, BYTE offset a1); optional row
// scale: source row k is multiplied by folded[a2 + k % i3] (a LayerNorm weight folded into the matrix), a2 = -1: none
// i: K, N, dst base, scale period
__device__ void job_pack_b(const PrepJob& j, const PrepBases& b, int fmt, int rank, int nb) {
  const float* __restrict__ folded = static_cast<const float*>(b.p[PB_FOLDED]);
  const float* __restrict__ wt = folded + j.a[0];
  unsigned short* __restrict__ dst = reinterpret_cast<unsigned short*>(static_cast<unsigned char*>(b.p[j.i[2]]) + j.a[1]);
  const int N = j.i[1];
  const int64_t total = (int64_t)j.i[0] * N;
  for (int64_t i = (int64_t)rank * blockDim.x + threadIdx.x; i < total; i += (int64_t)nb * blockDim.x) {
    const int64_t kc = i / ((int64_t)N * 8);
    const int rem = (int)(i - kc * N * 8);
    const int n = rem / 8, e = rem % 8;
    const int64_t k = kc * 8 + e;
    float v = wt[k * N + n];
    if (j.a[2] >= 0) v *= folded[j.a[2] + k % j.i[3]];
    dst[i] = to_operand16(v, fmt);
  }
}

// PK_PACK_SEL   B-operand tiles [K/8][n_rows][8] from selected rows / columns of a folded Wt [K_src][ld]: tile row r reads
// column seg_base[r / seg_len] + r % seg_len, tile k reads source row k0 + k; rows of segment 0 are scaled (1/sqrt(dh)
// folded into the query projection); source row k is also scaled by folded[a2 + k] when a2 >= 0 (LayerNorm weight folded in).
// a: src (folded floats), dst (wstream BYTES), row scale   i: n_rows, k_count, k0, ld, seg_len, seg_base[3]   f: seg0_scale
__device__ void job_pack_sel(const PrepJob& j, const PrepBases& b, int fmt, int rank, int nb) {
  const float* __restrict__ wt = static_cast<const float*>(b.p[PB_FOLDED]) + j.a[0];
  unsigned short* __restrict__ dst = reinterpret_cast<unsigned short*>(static_cast<unsigned char*>(b.p[PB_WSTREAM]) + j.a[1]);
  const int n_rows = j.i[0], k_count = j.i[1], k0 = j.i[2], ld = j.i[3], seg_len = j.i[4];
  const int total = n_rows * k_count;
  for (int i = rank * blockDim.x + threadIdx.x; i < total; i += nb * blockDim.x) {
    const int kc = i / (n_rows * 8);
    const int rem = i - kc * n_rows * 8;
    const int r = rem / 8, e = rem % 8;
    const int seg = r / seg_len;
    const int col = j.i[5 + seg] + r % seg_len;
    const int k = k0 + kc * 8 + e;
    float v = wt[(size_t)k * ld + col];
    if (seg == 0) v *= j.f[0];
    if (j.a[2] >= 0) v *= static_cast<const float*>(b.p[PB_FOLDED])[j.a[2] + k];
    dst[i] = to_operand16(v, fmt);
  }
}

// PK_GATHER_BIAS   one tower layer's folded bias, zero-padded to 128   a: src (folded), dst (bias floats)   i: width
__device__ void job_gather_bias(const PrepJob& j, const PrepBases& b, int rank, int nb) {
  const float* __restrict__ src = static_cast<const float*>(b.p[PB_FOLDED]) + j.a[0];
  float* __restrict__ dst = static_cast<float*>(b.p[PB_BIAS]) + j.a[1];
  for (int i = rank * blockDim.x + threadIdx.x; i < 128; i += nb * blockDim.x) dst[i] = i < j.i[0] ? src[i] : 0.0f;
}

// PK_PACK_FRAG   Wt [K][N] fp32 -> [N tiles][K/16][32 lanes] uint2: lane 4g+t of tile jt, k-block kt holds W[16kt+2t, +1][8jt+g]
// and W[16kt+2t+8, +9][8jt+g] (the m16n8k16 "col" B fragment); columns >= N are zero.   a: src (folded), dst (uint2)   i: K, N, tiles
__device__ void job_pack_frag(const PrepJob& j, const PrepBases& b, int fmt, int rank, int nb) {
  const float* __restrict__ Wt = static_cast<const float*>(b.p[PB_FOLDED]) + j.a[0];
  uint2* __restrict__ out = static_cast<uint2*>(b.p[PB_FRAGS]) + j.a[1];
  const int K = j.i[0], N = j.i[1], KT = K / 16;
  const int total = j.i[2] * KT * 32;
  for (int idx = rank * blockDim.x + threadIdx.x; idx < total; idx += nb * blockDim.x) {
    const int lane = idx % 32, kt = (idx / 32) % KT, jt = idx / (32 * KT);
    const int g = lane >> 2, t = lane & 3, n = 8 * jt + g, k0 = 16 * kt + 2 * t;
    auto w = [&](int k) -> uint32_t { return (uint32_t)to_operand16(n < N ? Wt[(size_t)k * N + n] : 0.0f, fmt); };
    out[idx] = make_uint2(w(k0) | (w(k0 + 1) << 16), w(k0 + 8) | (w(k0 + 9) << 16));
  }
}

// PK_BIAS_SLAB   the bias K-step of one GEMM: a [2 planes][n_rows][8] tile whose element (0, r, 0) is
//   (bias[col(r)] + sum_k Wt[k][col(r)] * beta[k]) * (segment 0 ? seg0_scale : 1)
// -- the layer's bias plus what the preceding LayerNorm's beta contributes through the matrix (a3 = -1: no LayerNorm) -- and
// everything else zero. col(r) as in PK_PACK_SEL.
// a: bias (folded), dst (wstream BYTES), Wt (folded), beta (folded)   i: n_rows, K, ld, seg_len, seg_base[3]   f: seg0_scale
__device__ void job_bias_slab(const PrepJob& j, const PrepBases& b, int fmt, int rank, int nb) {
  const float* __restrict__ folded = static_cast<const float*>(b.p[PB_FOLDED]);
  unsigned short* __restrict__ dst = reinterpret_cast<unsigned short*>(static_cast<unsigned char*>(b.p[PB_WSTREAM]) + j.a[1]);
  const int n_rows = j.i[0], K = j.i[1], ld = j.i[2], seg_len = j.i[3];
  for (int i = rank * blockDim.x + threadIdx.x; i < 2 * n_rows * 8; i += nb * blockDim.x) {
    const int r = (i / 8) % n_rows;
    float v = 0.0f;
    if (i < n_rows * 8 && (i & 7) == 0) {
      const int seg = r / seg_len;
      const int col = j.i[4 + seg] + r % seg_len;
      v = folded[j.a[0] + col];
      if (j.a[3] >= 0) {
        const float* __restrict__ wt = folded + j.a[2];
        const float* __restrict__ beta = folded + j.a[3];
        float acc = 0.0f;
        for (int k = 0; k < K; ++k) acc = fmaf(wt[(size_t)k * ld + col], beta[k], acc);
        v += acc;
      }
      if (seg == 0) v *= j.f[0];
    }
    dst[i] = to_operand16(v, fmt);
  }
}

// PK_PE_T   positional table [P][D] fp32 (folded + a0) -> [D / 4][P][4] fp32 at BYTE offset a1 of the weight stream: the fused
// kernel's threads are one per position, so a warp's 16-byte loads of one column group are then 512 contiguous bytes
// instead of 32 different cache lines.  i: P, D
__device__ void job_pe_t(const PrepJob& j, const PrepBases& b, int rank, int nb) {
  const float* __restrict__ src = static_cast<const float*>(b.p[PB_FOLDED]) + j.a[0];
  float* __restrict__ dst = reinterpret_cast<float*>(static_cast<unsigned char*>(b.p[PB_WSTREAM]) + j.a[1]);
  const int P = j.i[0], D = j.i[1];
  for (int i = rank * blockDim.x + threadIdx.x; i < P * D; i += nb * blockDim.x) {
    const int cg = i / (P * 4), rem = i - cg * P * 4;
    dst[i] = src[(rem >> 2) * D + cg * 4 + (rem & 3)];
  }
}

// PK_FC_BIAS   trunk FC bias with the encoder's final LayerNorm folded in: out[n] = bias[n] + sum_k Wt[k][n] * beta[k % period]
// a: bias, Wt, beta (folded floats)   i: K, N, period.   One block; thread (n, part) sums a K range, parts combined in shared memory.
__device__ void job_fc_bias(const PrepJob& j, const PrepBases& b, int rank, int nb) {
  if (rank != 0) return;
  __shared__ float4 part[256];
  const float* __restrict__ folded = static_cast<const float*>(b.p[PB_FOLDED]);
  const float4* __restrict__ wt4 = reinterpret_cast<const float4*>(folded + j.a[1]);  // 16-byte aligned: reserve() keeps it so
  const float* __restrict__ beta = folded + j.a[2];
  float* __restrict__ out = static_cast<float*>(b.p[PB_FCBIAS]);
  const int K = j.i[0], N = j.i[1], period = j.i[2];
  const int n4 = N / 4;                          // N is a multiple of 32 on this path: thread = (4 columns, K part)
  const int parts = 256 / n4;
  const int g = threadIdx.x % n4, part_id = threadIdx.x / n4;
  // a chain of L2 loads otherwise: float4 columns, eight rows in flight
  float4 acc[8];
#pragma unroll
  for (int u = 0; u < 8; ++u) acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (part_id < parts) {
    int k = part_id;
    for (; k + 7 * parts < K; k += 8 * parts) {
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int kk = k + u * parts;
        const float4 w = __ldg(wt4 + (size_t)kk * n4 + g);
        const float bt = beta[kk % period];
        acc[u].x = fmaf(w.x, bt, acc[u].x); acc[u].y = fmaf(w.y, bt, acc[u].y);
        acc[u].z = fmaf(w.z, bt, acc[u].z); acc[u].w = fmaf(w.w, bt, acc[u].w);
      }
    }
    for (; k < K; k += parts) {
      const float4 w = __ldg(wt4 + (size_t)k * n4 + g);
      const float bt = beta[k % period];
      acc[0].x = fmaf(w.x, bt, acc[0].x); acc[0].y = fmaf(w.y, bt, acc[0].y);
      acc[0].z = fmaf(w.z, bt, acc[0].z); acc[0].w = fmaf(w.w, bt, acc[0].w);
    }
  }
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int u = 0; u < 8; ++u) { s.x += acc[u].x; s.y += acc[u].y; s.z += acc[u].z; s.w += acc[u].w; }
  part[threadIdx.x] = part_id < parts ? s : make_float4(0.f, 0.f, 0.f, 0.f);
  __syncthreads();
  if (threadIdx.x < N) {
    const int gg = threadIdx.x / 4, c = threadIdx.x % 4;
    float v = folded[j.a[0] + threadIdx.x];
    for (int q = 0; q < parts; ++q) {
      const float4 t = part[q * n4 + gg];
      v += c == 0 ? t.x : (c == 1 ? t.y : (c == 2 ? t.z : t.w));
    }
    out[threadIdx.x] = v;
  }
}

__global__ void __launch_bounds__(256) prep_kernel(const PrepJob* __restrict__ jobs, int first_job, int n_jobs, PrepBases bases, int fmt) {
  __shared__ PrepJob job;
  if (threadIdx.x == 0) {
    int lo = first_job, hi = first_job + n_jobs - 1;  // jobs of a phase are contiguous and ordered by first_block
    while (lo < hi) {
      const int mid = (lo + hi + 1) / 2;
      if (jobs[mid].first_block <= (int)blockIdx.x) lo = mid;
      else hi = mid - 1;
    }
    job = jobs[lo];
  }
  __syncthreads();
  const int rank = (int)blockIdx.x - job.first_block, nb = job.n_blocks;
  switch (job.kind) {
    case PK_FOLD: job_fold(job, static_cast<const float*>(bases.p[PB_IMAGE]), static_cast<float*>(bases.p[PB_FOLDED]), rank, nb); break;
    case PK_COPY: job_copy(job, bases, rank, nb); break;
    case PK_PACK_B: job_pack_b(job, bases, fmt, rank, nb); break;
    case PK_PACK_SEL: job_pack_sel(job, bases, fmt, rank, nb); break;
    case PK_GATHER_BIAS: job_gather_bias(job, bases, rank, nb); break;
    case PK_PACK_FRAG: job_pack_frag(job, bases, fmt, rank, nb); break;
    case PK_BIAS_SLAB: job_bias_slab(job, bases, fmt, rank, nb); break;
    case PK_FC_BIAS: job_fc_bias(job, bases, rank, nb); break;
    case PK_PE_T: job_pe_t(job, bases, rank, nb); break;
    default: break;
  }
}

void add_fold(PrepList& list, const Affine& f) {
  PrepJob j{};
  j.kind = PK_FOLD; j.phase = 0;
  j.a[0] = f.src_w; j.a[1] = f.src_b; j.a[2] = f.bn_g; j.a[3] = f.bn_b; j.a[4] = f.bn_m; j.a[5] = f.bn_v; j.a[6] = f.wt; j.a[7] = f.bias;
  j.i[0] = f.K; j.i[1] = f.Kpad; j.i[2] = f.N; j.i[3] = f.perm_cols; j.i[4] = f.perm_inner;
  list.add(j, (int64_t)f.Kpad * f.N);
}

void add_copy0(PrepList& list, int64_t src, int64_t dst, int64_t n) {
  PrepJob j{};
  j.kind = PK_COPY; j.phase = 0;
  j.a[0] = src; j.a[1] = dst; j.i[0] = (int)n; j.i[1] = PB_IMAGE; j.i[2] = PB_FOLDED; j.f[0] = 1.0f;
  list.add(j, n);
}

}  // namespace

// image -> folded fp32 set (phase 0 jobs): eval-mode BatchNorm folded into the preceding affine map, weights transposed to
// [K,N], permuted to the NHWC / tap-major K order the kernels use, zero rows up to Kpad.
void fold_add_prep_jobs(const Plan& p, PrepList& list) {
  for (const ConvLayer& c : p.convs) add_fold(list, c.f);
  if (p.has_res_proj) add_fold(list, p.res_proj.f);
  for (const ResBlock& r : p.res) {
    add_fold(list, r.c1.f);
    add_fold(list, r.c2.f);
  }
  if (p.transformer) {
    const int dm = p.d.t_dim;
    if (p.has_input_proj) add_fold(list, p.input_proj);
    add_copy0(list, p.src_pe, p.pe, (int64_t)p.L * dm);  // rows 0..L-1 of pe[5000,1,d]
    for (const EncoderLayer& e : p.enc) {
      add_fold(list, e.qkv);
      add_fold(list, e.out);
      add_fold(list, e.ff1);
      add_fold(list, e.ff2);
      add_copy0(list, e.src_ln1_w, e.ln1_w, dm);
      add_copy0(list, e.src_ln1_b, e.ln1_b, dm);
      add_copy0(list, e.src_ln2_w, e.ln2_w, dm);
      add_copy0(list, e.src_ln2_b, e.ln2_b, dm);
    }
    add_copy0(list, p.src_lnf_w, p.lnf_w, dm);
    add_copy0(list, p.src_lnf_b, p.lnf_b, dm);
  }
  for (const Affine& f : p.shared) add_fold(list, f);
  for (const Affine& f : p.policy) add_fold(list, f);
  for (const Affine& f : p.value) add_fold(list, f);
}

int prep_upload(PrepList& list, PrepTable* t) {
  std::vector<PrepJob>& jobs = list.jobs();
  std::stable_sort(jobs.begin(), jobs.end(), [](const PrepJob& x, const PrepJob& y) { return x.phase < y.phase; });
  t->n_jobs = (int)jobs.size();
  for (int ph = 0; ph < 2; ++ph) t->phase_first_job[ph] = t->phase_jobs[ph] = t->phase_blocks[ph] = 0;
  for (int k = 0; k < t->n_jobs; ++k) {
    PrepJob& j = jobs[k];
    if (j.phase < 0 || j.phase > 1) return fail(AT_E_INVALID, "weight preparation: bad job phase");
    if (t->phase_jobs[j.phase] == 0) t->phase_first_job[j.phase] = k;
    j.first_block = t->phase_blocks[j.phase];
    t->phase_blocks[j.phase] += j.n_blocks;
    ++t->phase_jobs[j.phase];
  }
  if (t->n_jobs == 0) return AT_OK;
  AT_CUDA_TRY(cudaMalloc(&t->d_jobs, sizeof(PrepJob) * jobs.size()));
  AT_CUDA_TRY(cudaMemcpy(t->d_jobs, jobs.data(), sizeof(PrepJob) * jobs.size(), cudaMemcpyHostToDevice));
  return AT_OK;
}

void prep_free(PrepTable* t) {
  if (t->d_jobs) cudaFree(t->d_jobs);
  t->d_jobs = nullptr;
  t->n_jobs = 0;
}

int prep_run(const PrepTable& t, const PrepBases& bases, int fmt, cudaStream_t stream, int64_t* launches, int first_phase) {
  for (int ph = first_phase; ph < 2; ++ph) {
    if (t.phase_jobs[ph] == 0) continue;
    prep_kernel<<<(unsigned)t.phase_blocks[ph], 256, 0, stream>>>(t.d_jobs, t.phase_first_job[ph], t.phase_jobs[ph], bases, fmt);
    AT_CUDA_TRY(cudaGetLastError());
    if (launches) ++*launches;
  }
  return AT_OK;
}

}  // namespace at
