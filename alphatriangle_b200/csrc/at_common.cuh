// Shared declarations of libat_b200 (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <string>

#include "../../include/at_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libat_b200 is written for sm_100a (B200) only"
#endif

namespace at {

// ---- error plumbing (host) ---------------------------------------------------
void set_error(const std::string& msg);
int fail(int code, const std::string& msg);

#define AT_CUDA_TRY(expr)                                                                          \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess)                                                                         \
      return ::at::fail(AT_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));            \
  } while (0)

// Opt a kernel in to more than 48 KB of dynamic shared memory ONCE per device and call site (the attribute is per device;
// setting it on every launch costs a driver call on the path trimcts drives with 32-leaf batches).
#define AT_SMEM_OPT_IN_ONCE(kernel, bytes)                                                                        \
  do {                                                                                                            \
    static std::atomic<unsigned long long> done_{0};                                                              \
    int dev_ = 0;                                                                                                 \
    AT_CUDA_TRY(cudaGetDevice(&dev_));                                                                            \
    const unsigned long long bit_ = 1ull << (dev_ & 63);                                                          \
    if (!(done_.load(std::memory_order_acquire) & bit_)) {                                                        \
      AT_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes)));        \
      done_.fetch_or(bit_, std::memory_order_release);                                                            \
    }                                                                                                             \
  } while (0)

#define AT_TRY(expr)              \
  do {                            \
    int _rc = (expr);             \
    if (_rc != AT_OK) return _rc; \
  } while (0)

// ---- PackedState (see alphatriangle_b200/packing.py) ------------------------------
struct __align__(16) PackedState {
  uint32_t occ[4];
  uint32_t death[4];
  int8_t slot[AT_MAX_SLOTS][8];  // present, n_tris, n_up, 0, min_r, min_c, max_r, max_c
  double score;
  int32_t step;
  uint32_t pad;
};
static_assert(sizeof(PackedState) == AT_STATE_BYTES, "PackedState must be 80 bytes");

// ---- activations (device) -----------------------------------------------------
__device__ __forceinline__ float apply_act(float x, int act) {
  switch (act) {
    case AT_ACT_RELU: return fmaxf(x, 0.0f);
    case AT_ACT_GELU: return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));  // nn.GELU(approximate='none')
    case AT_ACT_SILU: return x / (1.0f + expf(-x));
    case AT_ACT_TANH: return tanhf(x);
    case AT_ACT_SIGMOID: return 1.0f / (1.0f + expf(-x));
    default: return x;
  }
}

__device__ __forceinline__ bool is_finite_f(float x) { return fabsf(x) <= 3.402823466e+38f; }  // false for NaN and +-inf

// ---- kernel launchers (host), one per .cu file --------------------------------
// at_features.cu
int launch_features(const at_env_desc& env, int grid_channels, int other_dim, const void* d_states, int64_t batch, float* d_grid,
                    float* d_other, cudaStream_t stream);

}  // namespace at
