/* oracle/ref_pass3_wrap.cpp — TEST INFRASTRUCTURE (never linked into the product).
 *
 * Hosts the reference's display pass, shaders/pass3.fsh (toneMapping(c, 1.5), IS/shaders/pass3.fsh:14-24), on the CPU the
 * same way ref_shader_wrap.cpp hosts pass1.fsh: oracle/build_ref.py writes the lightly rewritten shader text to
 * oracle/_ref/pass3.inc and this file includes it as the body of a struct.  Around it: the full-screen quad
 * (vshader.vsh:1-9, pix = NDC of the fragment centre) and the texPass0 binding of display() (IS/main.cpp:706-707).
 * Output: the fragment colours as floats, row 0 = bottom; the window's 8-bit conversion is GL's, not the shader's, and is
 * applied by the caller (tests/golden/make_golden_pass3.py).
 */
#include "glsl_compat.h"

namespace glslc {
#define in
#define out
#define uniform
struct Pass3 {
#include REF_PASS3_INC
};
#undef in
#undef out
#undef uniform
}  // namespace glslc

using namespace glslc;

extern "C" int ref_pass3_render(const float* rgba, int width, int height, float* out_rgba) {
    for (int y = 0; y < height; y++) {
        for (int x = 0; x < width; x++) {
            Pass3 f;
            f.pix = vec3(((float)x + 0.5f) / (float)width * 2.0f - 1.0f, ((float)y + 0.5f) / (float)height * 2.0f - 1.0f, 0.0f);
            f.texPass0 = sampler2D{rgba, width, height, 4};
            f.main();
            float* o = out_rgba + 4 * ((long)y * width + x);
            o[0] = f.fragColor.x; o[1] = f.fragColor.y; o[2] = f.fragColor.z; o[3] = f.fragColor.w;
        }
    }
    return 0;
}
