// wavefront.cu — wavefront pipeline (raygen / extend / per-material shade queues).  Placeholder
// until the queue kernels land: the entry point reports an error instead of falling back.
#include "render_internal.hpp"
#include "scene.hpp"

namespace rtb {
struct WavefrontState {};
void destroy_wavefront_state(WavefrontState* w) { delete w; }
int launch_wavefront(DeviceScene*, const DCamera&, int, int, int, const rtb_render_opts&, float*, cudaStream_t, uint64_t*) {
  return fail(RTB_ERR_INVALID, "wavefront pipeline not built yet");
}
}  // namespace rtb
