// scene.cu — scene upload / BVH / path-tracing state (north_star stages n3-n7).
#include "ptb_internal.h"

namespace ptb {
void scene_destroy(ptb_context *ctx) { (void)ctx; }
}  // namespace ptb
