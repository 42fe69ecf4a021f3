// Per-state "other features" arithmetic shared by the feature kernel (at_features.cu) and the head kernel's prologue
// (at_heads.cu), so both produce the same bits. Restates alphatriangle/features/extractor.py:48-139 and
// alphatriangle/features/grid_features.py:7-42 for boards up to 32 columns with whole rows as bit masks.
#pragma once

#include "at_common.cuh"

namespace at {

__device__ __forceinline__ double clip_d(double x, double lo, double hi) {
  // np.clip on a float64 scalar: NaN propagates, otherwise min(max(x, lo), hi)
  if (x != x) return x;
  return fmin(fmax(x, lo), hi);
}

__device__ __forceinline__ float store_f32(double v) {
  // value assigned into a float32 ndarray, then nan_to_num(nan=0, posinf=0, neginf=0)
  float f = __double2float_rn(v);
  return is_finite_f(f) ? f : 0.0f;
}

// bits [off, off + 32) of the 128-bit mask (lo, hi), masked to the board width; register-only (no indexed array)
__device__ __forceinline__ uint32_t row_bits(uint64_t lo, uint64_t hi, int off, uint32_t mask) {
  const uint64_t x = off < 64 ? lo : hi, y = off < 64 ? hi : 0ull;
  const int sh = off & 63;
  const uint64_t v = (x >> sh) | (sh ? (y << (64 - sh)) : 0ull);
  return (uint32_t)v & mask;
}
__device__ __forceinline__ float ratio01(int a, int b) {  // clip(a / b, 0, 1) as the reference's float32 sees it
  return fminf(fmaxf(__fdiv_rn((float)a, (float)b), 0.0f), 1.0f);
}

// One thread, one state: writes o[0 .. other_dim). cols <= 32.
__device__ __forceinline__ void other_features_bits(const PackedState& st, int rows, int cols, int slots, int other_dim, float* o) {
  const int cells = rows * cols;
  const int n_full = slots * 8 + 6;
  uint32_t po[4], em[4];
  int dead = 0;
#pragma unroll
  for (int w = 0; w < 4; ++w) {
    const int lo = w * 32;
    const uint32_t valid = cells >= lo + 32 ? 0xFFFFFFFFu : (cells > lo ? ((1u << (cells - lo)) - 1u) : 0u);
    po[w] = st.occ[w] & ~st.death[w] & valid;
    em[w] = ~st.occ[w] & ~st.death[w] & valid;
    dead += __popc(st.death[w] & valid);
  }
  const uint64_t po_lo = po[0] | ((uint64_t)po[1] << 32), po_hi = po[2] | ((uint64_t)po[3] << 32);
  const uint64_t em_lo = em[0] | ((uint64_t)em[1] << 32), em_hi = em[2] | ((uint64_t)em[3] << 32);
  const uint32_t cmask = cols >= 32 ? 0xFFFFFFFFu : ((1u << cols) - 1u);
  const uint32_t pmask = cmask >> 1;  // adjacent column pairs (c, c+1), c = 0 .. cols-2
  uint32_t G = 0;
  int sum_h = 0, max_h = 0, holes = 0, bump = 0;
  for (int r = rows - 1; r >= 0; --r) {
    holes += __popc(row_bits(em_lo, em_hi, r * cols, cmask) & G);
    G |= row_bits(po_lo, po_hi, r * cols, cmask);
    sum_h += __popc(G);
    max_h += G != 0u;
    bump += __popc((G ^ (G >> 1)) & pmask);
  }
  const int playable = cells - dead;
  auto put = [&](int d, float v) {
    if (d < other_dim) o[d] = v;
  };
  for (int i = 0; i < slots; ++i) {  // _get_shape_features / _get_shape_availability (extractor.py:48-90)
    const int8_t* sl = st.slot[i];
    const int n = (int)(uint8_t)sl[1];
    const bool present = sl[0] != 0;
    float f[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (present && n > 0) {
      const int ups = (int)(uint8_t)sl[2];
      const int min_r = sl[4], min_c = sl[5], max_r = sl[6], max_c = sl[7];
      f[0] = ratio01(n, 5);
      f[1] = __fdiv_rn((float)ups, (float)n);
      f[2] = __fdiv_rn((float)(n - ups), (float)n);
      f[3] = ratio01(max_r - min_r + 1, rows);
      f[4] = ratio01(3 * (max_c - min_c + 1) + 1, 4 * cols);
      f[5] = ratio01(min_r + max_r, 2 * rows);
      f[6] = ratio01(min_c + max_c, 2 * cols);
    }
#pragma unroll
    for (int k = 0; k < 7; ++k) put(i * 7 + k, f[k]);
    put(slots * 7 + i, present ? 1.0f : 0.0f);
  }
  {  // _get_explicit_features (extractor.py:92-118)
    const int e = slots * 8;
    put(e + 0, store_f32(clip_d(__ddiv_rn(st.score, 100.0), -5.0, 5.0)));
    put(e + 1, __fdiv_rn((float)sum_h, (float)(cols * rows)));
    put(e + 2, __fdiv_rn((float)max_h, (float)rows));
    put(e + 3, playable > 0 ? __fdiv_rn((float)holes, (float)playable) : 0.0f);
    put(e + 4, cols > 1 ? __fdiv_rn((float)bump, (float)((cols - 1) * rows)) : 0.0f);
    put(e + 5, st.step <= 0 ? 0.0f : (st.step >= 1000 ? 1.0f : __fdiv_rn((float)st.step, 1000.0f)));
  }
  for (int d = n_full; d < other_dim; ++d) o[d] = 0.0f;  // zero padding (extractor.py:127-139)
}

}  // namespace at
