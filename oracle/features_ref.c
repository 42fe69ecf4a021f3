/* Plain-C restatement of the reference's feature extraction (TEST INFRASTRUCTURE -- only tests/, __graft_entry__.smoke()
 * and bench.py's CPU-baseline legs may use anything under oracle/; the product path never does).
 *
 * Input is the PackedState record (alphatriangle_b200/packing.py) -- exactly what the reference's extractor reads from a
 * GameState, nothing derived. Follows, function by function:
 *   column heights      alphatriangle/features/grid_features.py:7-19
 *   holes               grid_features.py:22-33
 *   bumpiness           grid_features.py:36-42
 *   grid channel        features/extractor.py:33-46
 *   shape features      extractor.py:48-85
 *   shape availability  extractor.py:87-90
 *   explicit features   extractor.py:92-118
 *   concat / pad / cut  extractor.py:120-147
 * Arithmetic as the reference does it: Python floats / numpy float64 scalars, ONE rounding when a value is stored into
 * the float32 output (compile without -ffast-math; FP contraction cannot change a chain of single divisions).
 * Pinned by tests/test_oracle_c.py against the vectors the reference itself produced (tests/golden/).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define STATE_BYTES 80
#define MAX_CELLS 128
#define MAX_SLOTS 4

static int bit(const uint8_t* mask, int i) { return (mask[i >> 3] >> (i & 7)) & 1; }
static double clip(double x, double lo, double hi) { return x != x ? x : (x < lo ? lo : (x > hi ? hi : x)); }
static float store(double v) { /* assignment into a float32 array, then nan_to_num(nan = posinf = neginf = 0) */
  const float f = (float)v;
  return isfinite(f) ? f : 0.0f;
}

/* returns 0, or -1 for sizes outside the record format */
int at_oracle_features(const uint8_t* packed, int64_t n, int rows, int cols, int slots, int grid_channels, int other_dim,
                       float* grid /* [n, grid_channels, rows, cols] */, float* other /* [n, other_dim] */) {
  if (rows <= 0 || cols <= 0 || rows * cols > MAX_CELLS || slots <= 0 || slots > MAX_SLOTS || grid_channels <= 0 || other_dim <= 0) return -1;
  const int cells = rows * cols;
  for (int64_t s = 0; s < n; ++s) {
    const uint8_t* rec = packed + s * STATE_BYTES;
    const uint8_t *occ = rec, *death = rec + 16;
    const int8_t* slot = (const int8_t*)(rec + 32);
    double score;
    int32_t step;
    memcpy(&score, rec + 64, 8);
    memcpy(&step, rec + 72, 4);

    /* grid: +1 occupied and not death, -1 death, 0 otherwise; further channels zero */
    float* g = grid + s * (int64_t)grid_channels * cells;
    memset(g, 0, sizeof(float) * (size_t)grid_channels * cells);
    for (int i = 0; i < cells; ++i) g[i] = bit(death, i) ? -1.0f : (bit(occ, i) ? 1.0f : 0.0f);

    double full[MAX_SLOTS * 8 + 6];
    int n_full = 0;
    /* shape features, 7 per slot */
    for (int i = 0; i < slots; ++i) {
      const int8_t* sl = slot + 8 * i;
      const int present = sl[0] != 0, nt = (uint8_t)sl[1], ups = (uint8_t)sl[2];
      double f[7] = {0, 0, 0, 0, 0, 0, 0};
      if (present && nt > 0) {
        const int min_r = sl[4], min_c = sl[5], max_r = sl[6], max_c = sl[7];
        const int height = max_r - min_r + 1;
        const double width_eff = (max_c - min_c + 1) * 0.75 + 0.25;
        f[0] = clip(nt / 5.0, 0, 1);
        f[1] = (double)ups / nt;
        f[2] = (double)(nt - ups) / nt;
        f[3] = clip((double)height / rows, 0, 1);
        f[4] = clip(width_eff / cols, 0, 1);
        f[5] = clip(((min_r + max_r) / 2.0) / rows, 0, 1);
        f[6] = clip(((min_c + max_c) / 2.0) / cols, 0, 1);
      }
      /* the reference stores each into a float32 matrix first: round here, keep as double for the concat below */
      for (int k = 0; k < 7; ++k) full[n_full++] = (double)(float)f[k];
    }
    for (int i = 0; i < slots; ++i) full[n_full++] = slot[8 * i] != 0 ? 1.0 : 0.0;

    /* explicit features */
    int heights[MAX_CELLS];
    int sum_h = 0, max_h = 0, holes = 0, dead = 0;
    for (int c = 0; c < cols; ++c) {
      int max_r = -1;
      for (int r = 0; r < rows; ++r)
        if (bit(occ, r * cols + c) && !bit(death, r * cols + c)) max_r = r;
      heights[c] = max_r + 1;
      sum_h += heights[c];
      if (heights[c] > max_h) max_h = heights[c];
    }
    for (int c = 0; c < cols; ++c)
      for (int r = 0; r < heights[c]; ++r)
        if (!bit(occ, r * cols + c) && !bit(death, r * cols + c)) ++holes;
    double bump = 0.0;
    for (int c = 0; c + 1 < cols; ++c) bump += abs(heights[c] - heights[c + 1]);
    for (int i = 0; i < cells; ++i) dead += bit(death, i);
    const int playable = cells - dead;
    float e[6];
    e[0] = store(clip(score / 100.0, -5.0, 5.0));
    e[1] = store(((double)sum_h / cols) / rows);  /* np.mean(int32 heights) is float64: sum / cols, then / rows */
    e[2] = store((double)max_h / rows);
    e[3] = playable > 0 ? store((double)holes / playable) : 0.0f;
    e[4] = cols > 1 ? store((bump / (cols - 1)) / rows) : 0.0f;
    e[5] = store(clip(step / 1000.0, 0, 1));
    for (int k = 0; k < 6; ++k) full[n_full++] = e[k];

    float* o = other + s * (int64_t)other_dim;
    for (int d = 0; d < other_dim; ++d) {
      const float v = d < n_full ? (float)full[d] : 0.0f;
      o[d] = isfinite(v) ? v : 0.0f;
    }
  }
  return 0;
}
