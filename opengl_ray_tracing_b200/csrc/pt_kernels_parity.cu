// pt_kernels_parity.cu — the two bit-exact configurations.  MUST be compiled with -fmad=false (no FMA
// contraction) and without --use_fast_math: under those flags every plain float operator in
// pt_shade.cuh is the same IEEE binary32 operation the CPU oracle performs (oracle/ptoracle.cpp built
// with -ffp-contract=off), and include/ptmath.h supplies the transcendental functions on both sides.
#include "pt_wavefront.cuh"

namespace pt {
// which: 1 = PT_MODE_PARITY_REF, 2 = PT_MODE_PARITY_WIDE, 3 = PT_MODE_PARITY_FAST
cudaError_t launch_parity(const RenderParams& p, int variant, int which, cudaStream_t stream) {
    if (which == 1) return launch_render<CfgParityRef>(p, variant, stream);
    if (which == 2) return launch_render<CfgParityWide>(p, variant, stream);
    return launch_render<CfgParityFast>(p, variant, stream);
}
cudaError_t launch_persistent_parity(const RenderParams& p, int variant, int which, int grid, cudaStream_t stream) {
    if (which == 1) return launch_persistent<CfgParityRef>(p, variant, grid, stream);
    if (which == 2) return launch_persistent<CfgParityWide>(p, variant, grid, stream);
    return launch_persistent<CfgParityFast>(p, variant, grid, stream);
}
cudaError_t launch_persistent_parity_cdf(const RenderParams& p, int grid, cudaStream_t stream) { return launch_persistent_variant<CfgParityFastCdf, 2>(p, grid, stream); }
// the wavefront pipeline under the parity arithmetic: which = 2 (reference topology, wide nodes) or 3 (fast topology)
cudaError_t launch_wavefront_parity(const RenderParams& p, int variant, int which, int num_sms, cudaStream_t stream, WaveStats* ws) {
    if (which == 2) return launch_wavefront<CfgParityWide>(p, variant, num_sms, stream, ws);
    return launch_wavefront<CfgParityFast>(p, variant, num_sms, stream, ws);
}
cudaError_t build_light_table_parity(const DevScene& sc, float4* tab, cudaStream_t stream) { return build_light_table<CfgParityWide>(sc, tab, stream); }
cudaError_t derive_materials_parity(float4* materials, int n, cudaStream_t stream) { return derive_materials<CfgParityWide>(materials, n, stream); }
}  // namespace pt
