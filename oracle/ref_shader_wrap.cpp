/* oracle/ref_shader_wrap.cpp — TEST INFRASTRUCTURE (never linked into the product).
 *
 * Hosts the reference's OWN fragment-shader text on the CPU.  oracle/build_ref.py writes the lightly
 * rewritten shader (see oracle/glsl_compat.h for the rewrite rules) to oracle/_ref/shader_<v>.inc and
 * compiles this file with -DREF_SHADER_INC="…" -DREF_VARIANT=<0|1|2>; the shader becomes the body of
 * `struct Fragment`, so its global `uint seed = …` initialiser (pass1.fsh:73-76) runs once per
 * fragment exactly like a GLSL global.
 *
 * What this file adds around the shader is the part of the reference that lives in main.cpp / GL:
 *   - the full-screen quad + vshader.vsh:1-9: `pix` = NDC of the fragment centre;
 *   - the uniforms and sampler bindings of display() (IS/main.cpp:677-701);
 *   - pass1 -> pass2 ping-pong (IS/main.cpp:704-705): lastFrame := pass1 output after every frame.
 * Output is the accumulated RGBA32F image, row 0 = bottom (GL window origin).
 */
#include "glsl_compat.h"
#include <vector>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace glslc {

#define in
#define out
#define uniform

struct Fragment {
#include REF_SHADER_INC

    /* GLSL argument lists are evaluated left to right; build_ref.py rewrites the one call site whose
       arguments have side effects, SampleHdr(rand(), rand()), to SampleHdr(vec2{rand(), rand()}). */
#if REF_VARIANT == 2
    vec3 SampleHdr(vec2 xi) { return SampleHdr(xi.x, xi.y); }
#endif
    vec4 gl_FragData[1];

    struct Uniforms {
        uint frameCounter; int nTriangles, nNodes, width, height, hdrResolution;
        samplerBuffer triangles, nodes;
        sampler2D lastFrame, hdrMap, hdrCache;
        vec3 eye; mat4 cameraRotate;
    };
    /* members are initialised in declaration order, so `seed` (declared after the uniforms and `pix`)
       sees them already set — the same order a GLSL implementation guarantees */
    Fragment(const Uniforms& u, vec3 pix_)
        : pix(pix_), frameCounter(u.frameCounter), nTriangles(u.nTriangles), nNodes(u.nNodes),
          width(u.width), height(u.height),
#if REF_VARIANT == 2
          hdrResolution(u.hdrResolution),
#endif
          triangles(u.triangles), nodes(u.nodes), lastFrame(u.lastFrame), hdrMap(u.hdrMap),
#if REF_VARIANT == 2
          hdrCache(u.hdrCache),
#endif
          eye(u.eye), cameraRotate(u.cameraRotate) {}
};

#undef in
#undef out
#undef uniform

} /* namespace glslc */

using namespace glslc;

extern "C" {

int ref_shader_variant(void) { return REF_VARIANT; }
int ref_shader_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* NDC coordinate of fragment centre k on an n-pixel axis (rasteriser interpolation of the quad's
   vPosition, vshader.vsh:7-8); this formula is the one definition shared with oracle/ptoracle.cpp and
   the CUDA kernels (DESIGN.md "pixel mapping") */
static inline float ndc_of(int k, int n) { return ((float)k + 0.5f) / (float)n * 2.0f - 1.0f; }

/* Renders frames first_frame .. first_frame+n_frames-1 into accum (W*H*4 floats, in/out, row 0 =
   bottom) over the pixel rectangle [x0,x1) x [y0,y1).  hdrCache may be NULL for variants 0/1. */
int ref_shader_render(const float* tris, int nTriangles, const float* nodes, int nNodes,
                      const float* hdr, const float* hdrCache, int hdrW, int hdrH,
                      int width, int height, const float* eye, const float* cameraRotate,
                      unsigned first_frame, unsigned n_frames, float* accum,
                      int x0, int y0, int x1, int y1)
{
    std::vector<float> last((size_t)width * height * 4);
    for (unsigned f = 0; f < n_frames; f++) {
        memcpy(last.data(), accum, last.size() * sizeof(float));      /* pass2: copy into lastFrame */
        Fragment::Uniforms u;
        u.frameCounter = first_frame + f;
        u.nTriangles = nTriangles; u.nNodes = nNodes; u.width = width; u.height = height;
        u.hdrResolution = hdrW;
        u.triangles.data = tris; u.nodes.data = nodes;
        u.lastFrame = sampler2D{last.data(), width, height, 4};
        u.hdrMap = sampler2D{hdr, hdrW, hdrH, 3};
        u.hdrCache = sampler2D{hdrCache, hdrW, hdrH, 3};
        u.eye = vec3(eye[0], eye[1], eye[2]);
        for (int c = 0; c < 4; c++)                                      /* column-major, glm value_ptr */
            u.cameraRotate.c[c] = vec4(cameraRotate[4*c], cameraRotate[4*c+1], cameraRotate[4*c+2], cameraRotate[4*c+3]);
        /* fragments are independent; runs of 16 pixels of a row per task: a mesh row costs ~100x a sky row, so whole rows
           (the round-1 schedule) left most threads idle behind the few that drew the mesh rows */
        const int rw = x1 - x0, runs_per_row = (rw + 15) / 16;
        const long n_runs = (long)runs_per_row * (y1 - y0);
        #pragma omp parallel for schedule(dynamic, 1)
        for (long run = 0; run < n_runs; run++) {
            const int y = y0 + (int)(run / runs_per_row), xa = x0 + (int)(run % runs_per_row) * 16, xb = xa + 16 < x1 ? xa + 16 : x1;
            for (int x = xa; x < xb; x++) {
                Fragment frag(u, vec3(ndc_of(x, width), ndc_of(y, height), 0.0f));
                frag.main();
                float* o = accum + 4 * ((size_t)y * width + x);
                o[0] = frag.gl_FragData[0].x; o[1] = frag.gl_FragData[0].y;
                o[2] = frag.gl_FragData[0].z; o[3] = frag.gl_FragData[0].w;
            }
        }
    }
    return 0;
}

/* Known-answer hooks: the shader's own rand() (pass1.fsh:73-89) and sobolVec2 / CranleyPattersonRotation
   (pass1.fsh:112-136) evaluated for pixel (px,py) of a width x height target at frame `frame`. */
static Fragment make_fragment(int px, int py, int width, int height, unsigned frame) {
    Fragment::Uniforms u;
    memset((void*)&u, 0, sizeof u);
    u.frameCounter = frame; u.width = width; u.height = height;
    return Fragment(u, vec3(ndc_of(px, width), ndc_of(py, height), 0.0f));
}
void ref_shader_rand_seq(int px, int py, int width, int height, unsigned frame, int n, float* out) {
    Fragment f = make_fragment(px, py, width, height, frame);
    for (int i = 0; i < n; i++) out[i] = f.rand();
}
#if REF_VARIANT == 2
void ref_shader_sobol_cp(int px, int py, int width, int height, unsigned i, unsigned b, float* out4) {
    Fragment f = make_fragment(px, py, width, height, 0);
    vec2 s = f.sobolVec2(i, b);
    vec2 r = f.CranleyPattersonRotation(s);
    out4[0] = s.x; out4[1] = s.y; out4[2] = r.x; out4[3] = r.y;
}
#endif

} /* extern "C" */
