// pt_kernels_parity.cu — the two bit-exact configurations.  MUST be compiled with -fmad=false (no FMA
// contraction) and without --use_fast_math: under those flags every plain float operator in
// pt_shade.cuh is the same IEEE binary32 operation the CPU oracle performs (oracle/ptoracle.cpp built
// with -ffp-contract=off), and include/ptmath.h supplies the transcendental functions on both sides.
#include "pt_shade.cuh"

namespace pt {
cudaError_t launch_parity(const RenderParams& p, int variant, bool reference_layout, cudaStream_t stream) {
    return reference_layout ? launch_render<CfgParityRef>(p, variant, stream)
                            : launch_render<CfgParityWide>(p, variant, stream);
}
}  // namespace pt
