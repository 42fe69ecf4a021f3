// tcgen05 / TMEM bf16 path -- placeholder until the tower kernels land.
#include "at_bf16.cuh"

namespace at {

struct Bf16Path {};

int bf16_create(const Plan&, int64_t, Bf16Path** out, std::string* reason) {
  *out = nullptr;
  *reason = "tcgen05 tower not built yet";
  return AT_OK;
}
void bf16_destroy(Bf16Path* p) { delete p; }
int bf16_pack_weights(Bf16Path*, const Plan&, const float*, int, cudaStream_t) { return AT_OK; }
int bf16_evaluate(Bf16Path*, const Plan&, const float*, int, const void*, int64_t, float*, float*, uint32_t*, cudaStream_t, int64_t*) {
  return fail(AT_E_UNSUPPORTED, "bf16 path not built");
}

}  // namespace at
