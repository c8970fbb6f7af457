// pt_kernels_fast_v2.cu — the ONE kernel the headline metric runs: render_persistent<CfgFast, 2> (production arithmetic,
// variant 2 = Disney BRDF + Sobol + HDR importance sampling + MIS), in a translation unit of its own so that it can be
// compiled with -ftz=true on top of pt_kernels_fast.cu's flags.
//
// Flushing denormals shortens every approximate division, reciprocal and square root of the shading code (the non-ftz forms
// carry a scaling path for denormal operands): 3,560 instead of 3,696 SASS instructions, is-1M 21.97 vs 21.60 Grays/s in the
// same box (tools/ab.py, profiles/r02_raw/c31_ab_ftz.log) at an unchanged rRMSE against the oracle (1.094e-5, same pixels
// differing).  The SAME flag costs the other two variants' kernels 1.7 % (disney-1M) and 3.5 % (lambert-100k) — ptxas lays
// their loops out differently — so they stay in pt_kernels_fast.cu without it.
//
// What -ftz touches in the strict (_rn) helpers: a denormal operand or result of a ray/box/triangle test becomes zero.  The
// decisions built on them compare against 1e-5, 5e-4 or the sign of products of scene-scale numbers; a value below 1.2e-38
// there means an exact-zero geometry case, which the comparisons treat the same way.  Primary-hit ids stay bit-exact in every
// test (full 1080p id image, metric scene, 4K / 10 M triangles).  The parity configurations are compiled elsewhere, without it.
#include "pt_shade.cuh"

namespace pt {
cudaError_t launch_persistent_fast_v2(const RenderParams& p, int grid, cudaStream_t stream) {
    return launch_persistent_variant<CfgFast, 2>(p, grid, stream);
}
}  // namespace pt
