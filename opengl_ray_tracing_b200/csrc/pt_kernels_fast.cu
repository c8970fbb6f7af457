// pt_kernels_fast.cu — the production configuration: FMA contraction on, SFU transcendental
// approximations in shading.  Ray generation and all ray/box/triangle tests still use the strict
// (un-fused, IEEE) helpers of pt_common.cuh, so the primary-hit triangle ids stay bit-exact.
#include "pt_shade.cuh"

namespace pt {
cudaError_t launch_fast(const RenderParams& p, int variant, cudaStream_t stream) {
    return launch_render<CfgFast>(p, variant, stream);
}
}  // namespace pt
