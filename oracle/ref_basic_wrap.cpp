// oracle/ref_basic_wrap.cpp — TEST INFRASTRUCTURE.  The reference's CPU tracer (BasicRayTracingWithC++/main.cpp) with a
// REPRODUCIBLE random stream, as a library.
//
// build_ref.py generates oracle/_ref/basic_det.inc from the reference source where it lies, with these edits only:
//   * `const int SAMPLE/WIDTH/HEIGHT` become run-time ints, BRIGHTNESS (a function of SAMPLE, main.cpp:20) a macro
//   * randf() (main.cpp:211-214, an mt19937 seeded from random_device and shared by 50 threads) returns det_randf()
//   * main() becomes ref_basic_main(); det_seed(k, i, j) is inserted at the top of the per-pixel body so that every
//     (sample, pixel) path draws from its own counter-based stream; one OpenMP thread (the reference's 50 threads race on
//     the image); imshow(...) becomes det_export(image)
// Everything else — shapes, intersections, shoot(), pathTracing(), the scene literal, the per-pixel loop — is the
// reference's own code.  The CUDA kernel pt_basic_render consumes the same stream in the same order.
#include <stdint.h>
#include <string.h>

static uint32_t g_det_seed = 1;
static thread_local uint32_t t_det_state = 1;
static double* g_det_out = nullptr;

static inline uint32_t det_wang(uint32_t s) {          // the hash of the reference's shaders (pass1.fsh:77-84)
    s = (s ^ 61u) ^ (s >> 16);
    s *= 9u;
    s = s ^ (s >> 4);
    s *= 0x27d4eb2du;
    s = s ^ (s >> 15);
    return s;
}
static inline void det_seed(int k, int i, int j) {
    uint32_t s = det_wang(g_det_seed ^ ((uint32_t)k * 0x9e3779b1u));
    s = det_wang(s ^ ((uint32_t)i * 0x85ebca77u));
    s = det_wang(s ^ ((uint32_t)j * 0xc2b2ae3du));
    t_det_state = s;
}
static inline double det_randf() {                      // [0, 1): 32 bits / 2^32, exact in double
    t_det_state = det_wang(t_det_state);
    return (double)t_det_state * (1.0 / 4294967296.0);
}
static void det_export(const double* image);

#include "basic_det.inc"

static void det_export(const double* image) { memcpy(g_det_out, image, sizeof(double) * (size_t)WIDTH * HEIGHT * 3); }

// image: height*width*3 doubles, row 0 = top, the per-pixel SUM over the samples (what imshow receives, main.cpp:434)
extern "C" int ref_basic_render(int width, int height, int samples, unsigned seed, double* image) {
    WIDTH = width; HEIGHT = height; SAMPLE = samples;
    g_det_seed = seed; g_det_out = image;
    return ref_basic_main();
}
