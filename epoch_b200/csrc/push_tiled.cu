// push_tiled.cu -- cell-sorted fast path for push + deposit (placeholder until the
// tiled kernel lands; the dispatcher falls back to the general kernel).
#include "push_common.cuh"
namespace eb {
int32_t launch_push_tiled(eb200_ctx* c, const SpeciesD& sp, const PushP& p, bool* handled) {
  (void)c; (void)sp; (void)p;
  *handled = false;
  return 0;
}
}  // namespace eb
