// Feature extraction: PackedState[B] -> grid f32 [B,Cg,R,C] + other f32 [B,D].
//
// Restates, per state, alphatriangle/features/extractor.py:33-147 and
// alphatriangle/features/grid_features.py:7-42. The reference evaluates every
// "other" feature in float64 (Python floats / numpy float64 scalars) and rounds
// once when the value is stored into a float32 array, so the kernel does the same
// operations in the same order with explicitly rounded fp64 intrinsics (no FMA
// contraction) and one __double2float_rn per feature. Result: bit-exact.
//
// HBM-bound: 80 B in + (Cg*R*C + D)*4 B out per state (680 B at the defaults). Boards up to 32 columns take
// features_fast_kernel (bit-parallel rows, one thread per state); wider ones the generic kernel below it.
// One block = 16 states; the 16 lanes of a state share its columns (heights, holes),
// then its 30 float64 feature divisions, then its output rows (float4 stores).

#include "at_common.cuh"
#include "at_features.cuh"

namespace at {

namespace {

constexpr int kThreads = 256;
constexpr int kLanesPerState = 16;
constexpr int kStatesPerBlock = kThreads / kLanesPerState;
constexpr int kMaxOther = AT_MAX_SLOTS * 8 + 6;  // slots*7 + slots + 6

__device__ __forceinline__ bool bit_at(const uint32_t* m, int i) { return (m[i >> 5] >> (i & 31)) & 1u; }

__global__ void __launch_bounds__(kThreads) features_kernel(const PackedState* __restrict__ states, int64_t batch, int rows, int cols,
                                                           int slots, int grid_channels, int other_dim, int grid_vec4,
                                                           float* __restrict__ grid, float* __restrict__ other) {
  __shared__ PackedState s_state[kStatesPerBlock];
  __shared__ int s_height[kStatesPerBlock][AT_MAX_CELLS];
  __shared__ float s_other[kStatesPerBlock][kMaxOther];

  const int tid = threadIdx.x;
  const int64_t first = (int64_t)blockIdx.x * kStatesPerBlock;
  const int n_here = (int)min((int64_t)kStatesPerBlock, batch - first);

  // stage the packed records (16-byte coalesced loads)
  {
    const uint4* src = reinterpret_cast<const uint4*>(states + first);
    uint4* dst = reinterpret_cast<uint4*>(s_state);
    const int n_vec = n_here * (AT_STATE_BYTES / 16);
    for (int i = tid; i < n_vec; i += kThreads) dst[i] = __ldg(src + i);
  }
  __syncthreads();

  const int s = tid / kLanesPerState;
  const int lane = tid % kLanesPerState;
  const unsigned group_mask = 0xFFFFu << ((tid & 31) & ~15);
  const int cells = rows * cols;
  const int n_full = slots * 7 + slots + 6;

  if (s < n_here) {
    const PackedState& st = s_state[s];
    // grid_features.get_column_heights / count_holes (grid_features.py:7-33)
    int sum_h = 0, max_h = 0, holes = 0;
    for (int c = lane; c < cols; c += kLanesPerState) {
      int top = -1;
      for (int r = 0; r < rows; ++r) {
        const int i = r * cols + c;
        if (bit_at(st.occ, i) && !bit_at(st.death, i)) top = r;
      }
      const int h = top + 1;
      int empty_below = 0;
      for (int r = 0; r < h; ++r) {
        const int i = r * cols + c;
        if (!bit_at(st.occ, i) && !bit_at(st.death, i)) ++empty_below;
      }
      s_height[s][c] = h;
      sum_h += h;
      max_h = max(max_h, h);
      holes += empty_below;
    }
    __syncwarp(group_mask);
    // grid_features.get_bumpiness (grid_features.py:36-42): integer-valued, exact in fp64
    int bump = 0;
    for (int c = lane; c + 1 < cols; c += kLanesPerState) bump += abs(s_height[s][c] - s_height[s][c + 1]);
#pragma unroll
    for (int off = kLanesPerState / 2; off > 0; off >>= 1) {
      sum_h += __shfl_xor_sync(group_mask, sum_h, off, kLanesPerState);
      max_h = max(max_h, __shfl_xor_sync(group_mask, max_h, off, kLanesPerState));
      holes += __shfl_xor_sync(group_mask, holes, off, kLanesPerState);
      bump += __shfl_xor_sync(group_mask, bump, off, kLanesPerState);
    }
    int dead = 0;
    for (int w = 0; w < 4; ++w) {
      const int lo = w * 32;
      const uint32_t valid = cells >= lo + 32 ? 0xFFFFFFFFu : (cells > lo ? ((1u << (cells - lo)) - 1u) : 0u);
      dead += __popc(st.death[w] & valid);
    }
    const int playable = cells - dead;
    const double drows = (double)rows, dcols = (double)cols;

    // Every feature of the vector has the form clip((a / b) / c, lo, hi) in float64, rounded once to float32
    // (extractor.py:48-118; dividing by 1.0 is exact, and the unclipped ratios lie in [0, 1] by construction), so the
    // 16 lanes of a state share them out: lane l computes features l, l + 16, ...
    for (int d = lane; d < n_full; d += kLanesPerState) {
      double a = 0.0, b = 1.0, c = 1.0, lo = 0.0, hi = 1.0;
      if (d < slots * 7) {
        // _get_shape_features (extractor.py:48-85)
        const int i = d / 7, k = d - i * 7;
        const int8_t* sl = st.slot[i];
        const int n = (int)(uint8_t)sl[1];
        if (sl[0] != 0 && n > 0) {
          const int ups = (int)(uint8_t)sl[2];
          const int min_r = sl[4], min_c = sl[5], max_r = sl[6], max_c = sl[7];
          switch (k) {
            case 0: a = (double)n; b = 5.0; break;
            case 1: a = (double)ups; b = (double)n; break;
            case 2: a = (double)(n - ups); b = (double)n; break;
            case 3: a = (double)(max_r - min_r + 1); b = drows; break;
            case 4: a = __dadd_rn(__dmul_rn((double)(max_c - min_c + 1), 0.75), 0.25); b = dcols; break;
            case 5: a = (double)(min_r + max_r); b = 2.0; c = drows; break;
            default: a = (double)(min_c + max_c); b = 2.0; c = dcols; break;
          }
          if (k == 1 || k == 2) { lo = -1.0e300; hi = 1.0e300; }  // the reference does not clip these two
        }
      } else if (d < slots * 8) {
        // _get_shape_availability (extractor.py:87-90)
        a = st.slot[d - slots * 7][0] != 0 ? 1.0 : 0.0;
      } else {
        // _get_explicit_features (extractor.py:92-118)
        lo = -1.0e300; hi = 1.0e300;
        switch (d - slots * 8) {
          case 0: a = st.score; b = 100.0; lo = -5.0; hi = 5.0; break;
          case 1: a = (double)sum_h; b = dcols; c = drows; break;                    // np.mean(heights) / rows
          case 2: a = (double)max_h; b = drows; break;
          case 3: if (playable > 0) { a = (double)holes; b = (double)playable; } break;
          case 4: if (cols > 1) { a = (double)bump; b = (double)(cols - 1); c = drows; } break;
          default: a = (double)st.step; b = 1000.0; lo = 0.0; hi = 1.0; break;
        }
      }
      s_other[s][d] = store_f32(clip_d(__ddiv_rn(__ddiv_rn(a, b), c), lo, hi));
    }
  }
  __syncthreads();

  // other: concat [shape(7*slots), avail(slots), explicit(6)], zero-pad or truncate to D (extractor.py:120-139)
  {
    float* dst = other + first * other_dim;
    if (s < n_here)
      for (int d = lane; d < other_dim; d += kLanesPerState) dst[s * other_dim + d] = d < n_full ? s_other[s][d] : 0.0f;
  }
  // grid: channel 0 = +1 occupied&~death, -1 death, 0 otherwise; further channels zero (extractor.py:33-46)
  if (s < n_here) {
    const PackedState& st = s_state[s];
    const int per = grid_channels * cells;
    float* dst = grid + (first + s) * per;
    if (grid_vec4) {
      for (int v4 = lane; v4 * 4 < per; v4 += kLanesPerState) {
        float q[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int e = v4 * 4 + j;
          q[j] = e < cells ? (bit_at(st.death, e) ? -1.0f : (bit_at(st.occ, e) ? 1.0f : 0.0f)) : 0.0f;
        }
        *reinterpret_cast<float4*>(dst + v4 * 4) = make_float4(q[0], q[1], q[2], q[3]);
      }
    } else {
      for (int e = lane; e < per; e += kLanesPerState)
        dst[e] = e < cells ? (bit_at(st.death, e) ? -1.0f : (bit_at(st.occ, e) ? 1.0f : 0.0f)) : 0.0f;
    }
  }
}


// ---- fast path (COLS <= 32): one thread per state, whole board rows as bit masks -------------------------------
//
// heights / holes / bumpiness without a per-cell loop: with P_r = playable-occupied cells of row r as a bit mask and
// G_r = P_r | P_{r+1} | ... (columns whose height exceeds r):
//   sum(heights) = sum_r popc(G_r)        max(heights) = #{r : G_r != 0}
//   holes        = sum_r popc(empty_r & G_{r+1})
//   bumpiness    = sum_r popc((G_r ^ (G_r >> 1)) & low(cols-1))      (|h[c]-h[c+1]| = #rows where exactly one exceeds r)
//
// Divisions: every ratio except score/100 has small integer operands, and for integers |a|, b < 2^24
//   float32(float64(a) / float64(b)) == float32(a) / float32(b)     (IEEE single division)
// because a/b is at least 2^-49 (relative) away from every float32 rounding boundary it does not sit on, far more than
// the 2^-53 the float64 rounding can move it; the same bound covers the reference's two-step quotients (a/b)/c = a/(bc)
// and (0.75 w + 0.25)/cols = (3w+1)/(4 cols). tests/test_feature_identity.py checks the identity exhaustively over the
// operand ranges a PackedState can carry. Clipping to [0,1] commutes with the final rounding (bounds are exact).
constexpr int kFastThreads = 256;

__global__ void __launch_bounds__(kFastThreads) features_fast_kernel(const PackedState* __restrict__ states, int64_t batch, int rows,
                                                                      int cols, int slots, int grid_channels, int other_dim,
                                                                      int vec_ok, float* __restrict__ grid, float* __restrict__ other) {
  extern __shared__ __align__(16) uint8_t fsm[];
  PackedState* s_state = reinterpret_cast<PackedState*>(fsm);
  float* s_other = reinterpret_cast<float*>(fsm + kFastThreads * sizeof(PackedState));  // [state][other_dim]

  const int tid = threadIdx.x;
  const int64_t first = (int64_t)blockIdx.x * kFastThreads;
  const int n_here = (int)min((int64_t)kFastThreads, batch - first);
  {
    const uint4* src = reinterpret_cast<const uint4*>(states + first);
    uint4* dst = reinterpret_cast<uint4*>(s_state);
    for (int i = tid; i < n_here * (AT_STATE_BYTES / 16); i += kFastThreads) dst[i] = __ldg(src + i);
  }
  __syncthreads();

  const int cells = rows * cols;
  if (tid < n_here) other_features_bits(s_state[tid], rows, cols, slots, other_dim, s_other + tid * other_dim);
  __syncthreads();

  {  // other: the block's rows are one contiguous run
    float* dst = other + first * other_dim;
    const int total = n_here * other_dim;
    if (vec_ok && (total & 3) == 0) {
      const float4* s4 = reinterpret_cast<const float4*>(s_other);
      float4* d4 = reinterpret_cast<float4*>(dst);
      for (int i = tid; i < total / 4; i += kFastThreads) d4[i] = s4[i];
    } else {
      for (int i = tid; i < total; i += kFastThreads) dst[i] = s_other[i];
    }
  }
  {  // grid: one warp per state and pass, each lane four cells (extractor.py:33-46)
    const int per = grid_channels * cells;
    const int warp = tid / 32, lane = tid % 32;
    for (int ss = warp; ss < n_here; ss += kFastThreads / 32) {
      const PackedState& st = s_state[ss];
      float* dst = grid + (first + ss) * per;
      if (vec_ok && (per & 3) == 0) {
        for (int v4 = lane; v4 * 4 < per; v4 += 32) {
          const int e0 = v4 * 4;
          uint32_t on = 0, dn = 0;
          if (e0 < cells) {  // four cells never straddle a 32-bit word
            on = (st.occ[e0 >> 5] >> (e0 & 31)) & 0xFu;
            dn = (st.death[e0 >> 5] >> (e0 & 31)) & 0xFu;
            if (e0 + 4 > cells) {
              const uint32_t keep = (1u << (cells - e0)) - 1u;
              on &= keep;
              dn &= keep;
            }
          }
          float q[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) q[j] = (dn >> j) & 1u ? -1.0f : ((on >> j) & 1u ? 1.0f : 0.0f);
          *reinterpret_cast<float4*>(dst + e0) = make_float4(q[0], q[1], q[2], q[3]);
        }
      } else {
        for (int e = lane; e < per; e += 32)
          dst[e] = e < cells ? (bit_at(st.death, e) ? -1.0f : (bit_at(st.occ, e) ? 1.0f : 0.0f)) : 0.0f;
      }
    }
  }
}

}  // namespace

int launch_features(const at_env_desc& env, int grid_channels, int other_dim, const void* d_states, int64_t batch, float* d_grid,
                    float* d_other, cudaStream_t stream) {
  if (env.rows <= 0 || env.cols <= 0 || env.rows * env.cols > AT_MAX_CELLS)
    return fail(AT_E_INVALID, "at_extract_features: need 0 < rows*cols <= 128");
  if (env.slots <= 0 || env.slots > AT_MAX_SLOTS) return fail(AT_E_INVALID, "at_extract_features: need 1 <= slots <= 4");
  if (grid_channels <= 0 || other_dim <= 0) return fail(AT_E_INVALID, "at_extract_features: grid_channels and other_dim must be positive");
  if (batch < 0) return fail(AT_E_INVALID, "at_extract_features: negative batch");
  if (batch == 0) return AT_OK;
  if (!d_states || !d_grid || !d_other) return fail(AT_E_INVALID, "at_extract_features: null pointer");
  if ((reinterpret_cast<uintptr_t>(d_states) & 15u) != 0) return fail(AT_E_INVALID, "at_extract_features: d_states must be 16-byte aligned");
  const int64_t blocks = (batch + kStatesPerBlock - 1) / kStatesPerBlock;
  if (blocks > 0x7FFFFFFF) return fail(AT_E_INVALID, "at_extract_features: batch too large");
  // slot layout inside s_other: shape features at [i*7], availability at [slots*7+i], explicit at [slots*8]
  if (env.cols <= 32) {
    const int vec_ok = ((reinterpret_cast<uintptr_t>(d_grid) & 15u) == 0) && ((reinterpret_cast<uintptr_t>(d_other) & 15u) == 0) &&
                       ((kFastThreads * other_dim) % 4 == 0) && ((grid_channels * env.rows * env.cols) % 4 == 0);
    const size_t smem = (size_t)kFastThreads * (sizeof(PackedState) + (size_t)other_dim * sizeof(float));
    if (smem <= 200 * 1024) {
      if (smem > 48 * 1024) AT_SMEM_OPT_IN_ONCE(features_fast_kernel, 200 * 1024);
      const int64_t fblocks = (batch + kFastThreads - 1) / kFastThreads;
      features_fast_kernel<<<(unsigned)fblocks, kFastThreads, smem, stream>>>(static_cast<const PackedState*>(d_states), batch, env.rows,
                                                                              env.cols, env.slots, grid_channels, other_dim, vec_ok,
                                                                              d_grid, d_other);
      AT_CUDA_TRY(cudaGetLastError());
      return AT_OK;
    }
  }
  const int grid_vec4 = ((grid_channels * env.rows * env.cols) % 4 == 0) && ((reinterpret_cast<uintptr_t>(d_grid) & 15u) == 0);
  features_kernel<<<(unsigned)blocks, kThreads, 0, stream>>>(static_cast<const PackedState*>(d_states), batch, env.rows, env.cols,
                                                              env.slots, grid_channels, other_dim, grid_vec4, d_grid, d_other);
  AT_CUDA_TRY(cudaGetLastError());
  return AT_OK;
}

}  // namespace at
