// pt_kernels_fast.cu — the production configuration: FMA contraction on, SFU transcendental
// approximations in shading.  Ray generation and all ray/box/triangle tests still use the strict
// (un-fused, IEEE) helpers of pt_common.cuh, so the primary-hit triangle ids stay bit-exact.
#include "pt_wavefront.cuh"

namespace pt {
cudaError_t launch_fast(const RenderParams& p, int variant, bool count, cudaStream_t stream) {
    if (count) return launch_render<CfgFastCount>(p, variant, stream);   // same arithmetic + traversal work counters
    return launch_render<CfgFast>(p, variant, stream);
}
cudaError_t launch_persistent_fast_v2(const RenderParams& p, int grid, cudaStream_t stream);   // pt_kernels_fast_v2.cu (-ftz=true)
cudaError_t launch_persistent_fast(const RenderParams& p, int variant, bool count, int grid, cudaStream_t stream) {
    if (count) return launch_persistent<CfgFastCount>(p, variant, grid, stream);
    if (variant == 0) return launch_persistent_variant<CfgFast, 0>(p, grid, stream);
    if (variant == 1) return launch_persistent_variant<CfgFast, 1>(p, grid, stream);
    return launch_persistent_fast_v2(p, grid, stream);    // the metric kernel lives in its own translation unit
}
// PT_ENV_CDF_SEARCH: variant 2 on the persistent pipeline only
cudaError_t launch_persistent_fast_cdf(const RenderParams& p, int grid, cudaStream_t stream) { return launch_persistent_variant<CfgFastCdf, 2>(p, grid, stream); }
cudaError_t launch_wavefront_fast(const RenderParams& p, int variant, bool count, int num_sms, cudaStream_t stream, WaveStats* ws) {
    if (count) return launch_wavefront<CfgFastCount>(p, variant, num_sms, stream, ws);
    return launch_wavefront<CfgFast>(p, variant, num_sms, stream, ws);
}
}  // namespace pt
