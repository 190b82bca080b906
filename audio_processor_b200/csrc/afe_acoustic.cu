// afe_acoustic.cu -- C ABI entry point of the batched acoustic features (placeholder body).
#include "afe_ctx.h"
using namespace afe;
extern "C" {
size_t afe_acoustic_workspace_bytes(const int64_t*, int) { return 256; }
int afe_acoustic_batch(afe_ctx*, void*, const float* const*, const int64_t*, int, int, float*, int32_t*, void*, size_t) {
  return fail(AFE_ERR_UNSUPPORTED, "afe_acoustic_batch: not built yet");
}
}
