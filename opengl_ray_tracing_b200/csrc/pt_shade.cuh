// pt_shade.cuh — per-sample shading: RNG, Sobol, Disney BRDF, HDR environment sampling, MIS, and the
// render kernel.  One source, three arithmetic configurations (pt_kernels_parity.cu / pt_kernels_fast.cu).
//
// Reference functions, by line ("P:" = ImportanceSampling_LowDiscrepancySequence/shaders/pass1.fsh,
// "D:" = DisneyBRDF/shaders/pass1.fsh, "O:" = OpenglRayTracing/shaders/pass1.fsh):
//   seed/wang_hash/rand P:73-89 · sobol/grayCode/sobolVec2 P:96-116 · CranleyPattersonRotation P:118-136
//   SampleHemisphere O:89-95 · toNormalHemisphere P:153-159 · getTangent P:161-172
//   SampleSphericalMap/sampleHdr P:175-189 · BRDF terms P:386-421 · BRDF_Evaluate_aniso P:423-482
//   SampleCosineHemisphere/SampleGTR2/SampleGTR1/SampleBRDF P:485-570 · SampleHdr P:573-585
//   BRDF_Evaluate P:587-636 · hdrColor/hdrPdf P:647-666 · BRDF_Pdf P:669-706 · misMixWeight P:708-711
//   pathTracing O:329-364, D:443-481 · pathTracingImportanceSampling P:761-841 · main P:844-871
//
// Structure: the three shaders call hitBVH from up to three places; here every closest-hit query of a
// path (primary and bounce rays) goes through ONE traversal site inside a loop, and shadow rays through a
// second, any-hit site.  What must survive a traversal is kept small: f_r, N·L, pdf, history, Lo.
#pragma once
#include <algorithm>

#include "pt_trace.cuh"
#include "../../include/ptmath.h"

namespace pt {

#define PT_PI 3.1415926f          // P:23 — the shader's own constant, not pi

static __constant__ uint32_t c_sobol_v[8 * 32] = {
#include "pt_sobol_v.inc"
};

// ---- arithmetic configurations ------------------------------------------------------------------
struct MathExact {                // include/ptmath.h: the oracle's definition of the GLSL transcendentals
    static __device__ __forceinline__ float sin(float a) { return pt_sinf(a); }
    static __device__ __forceinline__ float cos(float a) { return pt_cosf(a); }
    static __device__ __forceinline__ float asin(float a) { return pt_asinf(a); }
    static __device__ __forceinline__ float atan2(float y, float x) { return pt_atan2f(y, x); }
    static __device__ __forceinline__ float log(float a) { return pt_logf(a); }
    static __device__ __forceinline__ float pow(float a, float b) { return pt_powf(a, b); }
    static __device__ __forceinline__ float sin_dir(float a) { return pt_sinf(a); }
    static __device__ __forceinline__ float cos_dir(float a) { return pt_cosf(a); }
    static __device__ __forceinline__ float pow_dir(float a, float b) { return pt_powf(a, b); }
};
// Production arithmetic: SFU approximations where CUDA has them.  (`*_dir` = the calls that build a sampled direction;
// PT_DIR_TRIG_EXACT = 1 routes them to include/ptmath.h.  Measured on the bench's crop of the metric scene, 2 spp, rRMSE
// of PT_MODE_FAST against the oracle: 9.3e-6 with the exact functions, 1.09e-5 with the SFU's — no gain worth 4.8 % of
// the frame time.  What DID separate the production image from the oracle's, 1.14e-3, was the cancellation inside the
// lobe samplers: see SampleBRDF.)
#ifndef PT_DIR_TRIG_EXACT
#define PT_DIR_TRIG_EXACT 0
#endif
#ifndef PT_STRICT_SAMPLERS
#define PT_STRICT_SAMPLERS 1
#endif
struct MathFast {
    static __device__ __forceinline__ float sin(float a) { return __sinf(a); }
    static __device__ __forceinline__ float cos(float a) { return __cosf(a); }
    static __device__ __forceinline__ float asin(float a) { return asinf(a); }
    static __device__ __forceinline__ float atan2(float y, float x) { return atan2f(y, x); }
    static __device__ __forceinline__ float log(float a) { return __logf(a); }
    static __device__ __forceinline__ float pow(float a, float b) { return __powf(a, b); }
#if PT_DIR_TRIG_EXACT
    static __device__ __forceinline__ float sin_dir(float a) { return pt_sinf(a); }
    static __device__ __forceinline__ float cos_dir(float a) { return pt_cosf(a); }
    static __device__ __forceinline__ float pow_dir(float a, float b) { return pt_powf(a, b); }
#else
    static __device__ __forceinline__ float sin_dir(float a) { return __sinf(a); }
    static __device__ __forceinline__ float cos_dir(float a) { return __cosf(a); }
    static __device__ __forceinline__ float pow_dir(float a, float b) { return __powf(a, b); }
#endif
};
struct CfgParityRef  { typedef MathExact M; static const bool RAW = true;  static const bool FAST_TREE = false; static const bool COUNT = false; static const bool LIGHT_TAB = false; static const bool ENV_CDF = false; };
struct CfgParityWide { typedef MathExact M; static const bool RAW = false; static const bool FAST_TREE = false; static const bool COUNT = false; static const bool LIGHT_TAB = true;  static const bool ENV_CDF = false; };
struct CfgParityFast { typedef MathExact M; static const bool RAW = false; static const bool FAST_TREE = true;  static const bool COUNT = false; static const bool LIGHT_TAB = true;  static const bool ENV_CDF = false; };
struct CfgFast       { typedef MathFast  M; static const bool RAW = false; static const bool FAST_TREE = true;  static const bool COUNT = false; static const bool LIGHT_TAB = true;  static const bool ENV_CDF = false; };
struct CfgFastCount  { typedef MathFast  M; static const bool RAW = false; static const bool FAST_TREE = true;  static const bool COUNT = true;  static const bool LIGHT_TAB = true;  static const bool ENV_CDF = false; };
// PT_ENV_CDF_SEARCH: light samples by binary search of the marginal / conditional CDFs (no hdrCache table, no light table)
struct CfgParityFastCdf { typedef MathExact M; static const bool RAW = false; static const bool FAST_TREE = true; static const bool COUNT = false; static const bool LIGHT_TAB = false; static const bool ENV_CDF = true; };
struct CfgFastCdf       { typedef MathFast  M; static const bool RAW = false; static const bool FAST_TREE = true; static const bool COUNT = false; static const bool LIGHT_TAB = false; static const bool ENV_CDF = true; };

// counting configuration: histogram of traversal steps (inner nodes + triangles) per ray, log2 buckets
__device__ __forceinline__ uint32_t ray_steps(const Counter<false>&) { return 0; }
__device__ __forceinline__ uint32_t ray_steps(const Counter<true>& c) { return c.c.inner + c.c.tris; }
__device__ __forceinline__ void note_ray(const Counter<false>&, uint32_t, unsigned long long*) {}
__device__ __forceinline__ void note_ray(const Counter<true>& c, uint32_t before, unsigned long long* hist) {
    const uint32_t n = c.c.inner + c.c.tris - before;
    atomicAdd(hist + (n ? 32 - __clz(n) : 0), 1ull);
}

// `far_origin`: the ray starts far outside the scene (a primary ray of a distant eye, RenderParams.eye_far).  What
// hitTriangle accepts is then fuzzy on the scale of |o| 2^-23 — the rounding of P = o + d t — which the fixed padding of
// the production tree's boxes (scene radius 2^-17, pt_build.cpp) no longer covers, so such rays are traced on the
// reference topology with the reference's own box test (trace_wide: exact by construction).  Secondary rays start on
// the scene and never take this path.
template <class Cfg>
__device__ __forceinline__ TraceResult trace(const DevScene& sc, V3 o, V3 d, const bool anyhit, Counter<Cfg::COUNT>& cnt, const bool far_origin = false) {
    if (Cfg::RAW) return trace_ref(sc, o, d);         // the reference has no any-hit query: closest hit answers both
    if (Cfg::FAST_TREE) {
#if PT_FAR_EYE
        if (far_origin) return trace_wide_fallback(sc, o, d);
#endif
        const uint32_t steps0 = ray_steps(cnt);
        TraceResult tr = trace_fast<PT_STACK>(sc, o, d, anyhit, cnt);
        note_ray(cnt, steps0, sc.ray_histogram);
        if (tr.tri == -2) { cnt.tie(); tr = trace_wide_fallback(sc, o, d); }   // exact tie: let the reference order decide
        return tr;
    }
    return anyhit ? trace_wide<true, PT_STACK>(sc, o, d) : trace_wide<false, PT_STACK>(sc, o, d);
}

// ---- RNG and QMC -----------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t wang_hash(uint32_t& seed) {
    seed = (seed ^ 61u) ^ (seed >> 16);
    seed *= 9u;
    seed = seed ^ (seed >> 4);
    seed *= 0x27d4eb2du;
    seed = seed ^ (seed >> 15);
    return seed;
}
__device__ __forceinline__ float u2f(uint32_t h) { return s_mul(__uint2float_rn(h), 2.3283064365386963e-10f); }  // /2^32, exact
__device__ __forceinline__ float rnd(uint32_t& seed) { return u2f(wang_hash(seed)); }

__device__ __forceinline__ float sobol(uint32_t d, uint32_t i) {
    uint32_t result = 0, offset = d * 32;
    for (uint32_t j = 0; i != 0; i >>= 1, j++)
        if (i & 1u) result ^= c_sobol_v[j + offset];
    return s_mul(__uint2float_rn(result), 2.3283064365386963e-10f);          // 1.0f/float(0xFFFFFFFFU) == 2^-32
}
// The Sobol point of a frame is the same for every pixel (P:795-799: index = frameCounter + 1; the per-pixel part is the
// Cranley-Patterson shift), and the loop above runs once per set bit POSITION of the frame index — 13 rounds at frame
// 5000.  One tiny launch tabulates the 8 coordinates of each frame of a render launch with the very same function.
static __global__ void sobol_table_kernel(float* __restrict__ tab, uint32_t first_frame, uint32_t n_frames, uint32_t stride) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_frames * 8u) return;
    const uint32_t frame = first_frame + (i >> 3) * stride;
    tab[i] = sobol(i & 7u, (frame + 1u) ^ ((frame + 1u) >> 1));
}
// row = the frame's 8 tabulated coordinates, or null: compute them here
__device__ __forceinline__ void sobol_pair(const float* __restrict__ row, uint32_t bounce, uint32_t frame, float& u, float& v) {
    if (row != nullptr && bounce < 4u) {
        const float2 t = __ldg(reinterpret_cast<const float2*>(row) + bounce);
        u = t.x; v = t.y;
    } else {
        u = sobol(bounce * 2u, (frame + 1u) ^ ((frame + 1u) >> 1));
        v = sobol(bounce * 2u + 1u, (frame + 1u) ^ ((frame + 1u) >> 1));
    }
}
__device__ __forceinline__ void cranley_patterson(int px, int py, float& u, float& v) {
    uint32_t pseed = ((uint32_t)px * 1973u + (uint32_t)py * 9277u + (uint32_t)(114514 / 1919) * 26699u) | 1u;
    float a = rnd(pseed), b = rnd(pseed);
    u = s_add(u, a);
    if (u > 1.0f) u = s_sub(u, 1.0f);
    if (u < 0.0f) u = s_add(u, 1.0f);
    v = s_add(v, b);
    if (v > 1.0f) v = s_sub(v, 1.0f);
    if (v < 0.0f) v = s_add(v, 1.0f);
}

// ---- frames ---------------------------------------------------------------------------------------
__device__ __forceinline__ V3 toNormalHemisphere(V3 v, V3 N) {
    V3 helper = mk3(1, 0, 0);
    if (fabsf(N.x) > 0.999f) helper = mk3(0, 0, 1);
    V3 tangent = normalize(cross(N, helper));
    V3 bitangent = normalize(cross(N, tangent));
    return tangent * v.x + bitangent * v.y + N * v.z;
}
__device__ __forceinline__ void getTangent(V3 N, V3& tangent, V3& bitangent) {
    V3 helper = mk3(1, 0, 0);
    if (fabsf(N.x) > 0.999f) helper = mk3(0, 0, 1);
    bitangent = normalize(cross(N, helper));
    tangent = normalize(cross(N, bitangent));
}

// ---- environment map ----------------------------------------------------------------------------------
// Direction <-> texel mapping, evaluated in the oracle's arithmetic (strict ops + include/ptmath.h) in EVERY
// mode.  Reason: calculateHdrCache stores texel CORNERS (x/W, y/H; IS/main.cpp:632-633), so the direction
// SampleHdr builds from a cache entry maps back to a uv that sits exactly on a texel boundary; which of the
// two neighbouring texels hdrColor()/hdrPdf() then read (P:647-666) is decided by the last bits of
// cos/sin/atan/asin.  Leaving those to the SFU approximations flips about half of all light-sample lookups.
__device__ __forceinline__ void spherical_uv(V3 v, float& u, float& w) {      // SampleSphericalMap, P:175-181
    V3 n = s_normalize(v);
    u = s_div(pt_atan2f(n.z, n.x), s_mul(2.0f, PT_PI));
    w = s_div(pt_asinf(n.y), PT_PI);
    u = s_add(u, 0.5f);
    w = s_sub(1.0f, s_add(w, 0.5f));
}
__device__ __forceinline__ float4 tex_nearest(const float4* img, int w, int h, float u, float v) {
    return ldg4(img + (size_t)nearest_texel(v, h) * w + nearest_texel(u, w));
}
template <class M> __device__ __forceinline__ V3 sampleHdr(const DevScene& sc, V3 v) {      // clamped to 10, P:184-189
    float u, w;
    spherical_uv(v, u, w);
    float4 c = tex_nearest(sc.hdr4, sc.hdrW, sc.hdrH, u, w);
    return mk3(fminf(c.x, 10.0f), fminf(c.y, 10.0f), fminf(c.z, 10.0f));
}
// hdrColor (unclamped, P:647-651) and hdrPdf (P:655-666) look up the same uv: one atan2/asin for both
// the table's B channel (pdf of the texel): from the hdrCache texture, or (PT_ENV_CDF_SEARCH) from the pdf map itself
template <bool CDF> __device__ __forceinline__ float env_texel_pdf(const DevScene& sc, float u, float w) {
    if (CDF) return ldg1(sc.env_pdf + (size_t)nearest_texel(w, sc.hdrH) * sc.hdrW + nearest_texel(u, sc.hdrW));
    return tex_nearest(sc.cache4, sc.hdrW, sc.hdrH, u, w).z;
}
template <class M, bool CDF = false> __device__ __forceinline__ void hdr_color_pdf(const DevScene& sc, V3 L, V3& color, float& pdf_light) {
    float u, w;
    spherical_uv(L, u, w);
    float4 c = tex_nearest(sc.hdr4, sc.hdrW, sc.hdrH, u, w);
    color = mk3(c.x, c.y, c.z);
    float pdf = env_texel_pdf<CDF>(sc, u, w);
    float theta = PT_PI * (0.5f - w);
    float sin_theta = fmaxf(M::sin(theta), 1e-10f);
    float p_convert = (float)(sc.hdrW * sc.hdrW / 2) / (2.0f * PT_PI * PT_PI * sin_theta);
    pdf_light = pdf * p_convert;
}
// hdrPdf at an already computed uv, P:655-666
template <class M, bool CDF = false> __device__ __forceinline__ float hdr_pdf(const DevScene& sc, float u, float w) {
    float pdf = env_texel_pdf<CDF>(sc, u, w);
    float theta = PT_PI * (0.5f - w);
    float sin_theta = fmaxf(M::sin(theta), 1e-10f);
    float p_convert = (float)(sc.hdrW * sc.hdrW / 2) / (2.0f * PT_PI * PT_PI * sin_theta);
    return pdf * p_convert;
}
__device__ __forceinline__ V3 hdr_entry_direction(float4 c) {                                              // P:575-584
    float x = c.x, y = s_sub(1.0f, c.y);
    float phi = s_mul(s_mul(2.0f, PT_PI), s_sub(x, 0.5f));
    float theta = s_mul(PT_PI, s_sub(y, 0.5f));
    float ct = pt_cosf(theta);
    return mk3(s_mul(ct, pt_cosf(phi)), pt_sinf(theta), s_mul(ct, pt_sinf(phi)));
}
// std::lower_bound as libstdc++ writes it (first position whose element is not < val)
__device__ __forceinline__ int lower_bound_ldg(const float* __restrict__ a, int n, float val) {
    int first = 0, len = n;
    while (len > 0) {
        const int half = len >> 1, middle = first + half;
        if (ldg1(a + middle) < val) { first = middle + 1; len = len - half - 1; }
        else len = half;
    }
    return first;
}
// SampleHdr, P:573-585.  CDF = false: the reference's NEAREST lookup in the precomputed table.  CDF = true
// (PT_ENV_CDF_SEARCH): the two searches calculateHdrCache ran per table entry (M:621-636) run per sample on the continuous
// random numbers.  The table entry a lookup (xi_1, xi_2) reads is [row floor(xi_2 H)][column floor(xi_1 W)], which the
// reference computed from "xi_1" = row / H and "xi_2" = column / W (its transposed roles, M:616-634): so xi_2 searches the
// marginal CDF for the column x and xi_1 the column's conditional CDF for the row y — at the table's grid points the
// search returns the table's entries (tests/test_env_cdf.py).
template <class M, bool CDF = false> __device__ __forceinline__ V3 SampleHdr(const DevScene& sc, float xi_1, float xi_2) {   // P:573-585
    if (CDF) {
        int x = lower_bound_ldg(sc.env_cdf_x, sc.hdrW, xi_2);
        if (x >= sc.hdrW) x = sc.hdrW - 1;
        const int y = lower_bound_ldg(sc.env_cdf_y + (size_t)x * sc.hdrH, sc.hdrH, xi_1);
        return hdr_entry_direction(make_float4(s_div((float)x, (float)sc.hdrW), s_div((float)y, (float)sc.hdrH), 0.0f, 0.0f));
    }
    return hdr_entry_direction(tex_nearest(sc.cache4, sc.hdrW, sc.hdrH, xi_1, xi_2));
}

// ---- light table (Cfg::LIGHT_TAB) ------------------------------------------------------------------------
// Everything the light-sampling half of a bounce computes from its two random numbers is a function of the hdrCache TEXEL
// they select (P:573-585: NEAREST lookup, then a direction from the entry): the direction Ld (two sines, two cosines), and,
// if the shadow ray escapes, hdrColor(Ld) and hdrPdf(Ld) (P:647-666: normalize, atan, asin, two more texel fetches, a sine,
// a division).  There are only W x H such texels, so the upload evaluates those very functions once per texel, with the
// arithmetic of the mode the context renders in, and the render kernel reads the results back: bit for bit what the
// straight-line code computes, ~350 instructions and two scattered 16-byte fetches per bounce cheaper.
//   tab[2 i]     = (Ld.x, Ld.y, Ld.z, -)             read where the reference calls SampleHdr
//   tab[2 i + 1] = (hdrColor(Ld).rgb, hdrPdf(Ld))    read only when the shadow ray misses (same 32-byte sector)
template <class M>
__global__ void light_table_kernel(const DevScene sc, float4* __restrict__ tab) {
    const size_t n = (size_t)sc.hdrW * sc.hdrH;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const V3 Ld = hdr_entry_direction(ldg4(sc.cache4 + i));
        float u, w;
        spherical_uv(Ld, u, w);
        const float4 c = tex_nearest(sc.hdr4, sc.hdrW, sc.hdrH, u, w);
        tab[2 * i] = make_float4(Ld.x, Ld.y, Ld.z, 0.0f);
        tab[2 * i + 1] = make_float4(c.x, c.y, c.z, hdr_pdf<M>(sc, u, w));
    }
}
template <class Cfg>
cudaError_t build_light_table(const DevScene& sc, float4* tab, cudaStream_t stream) {
    light_table_kernel<typename Cfg::M><<<148 * 8, 256, 0, stream>>>(sc, tab);
    return cudaGetLastError();
}

// ---- Disney terms -------------------------------------------------------------------------------------
__device__ __forceinline__ float SchlickFresnel(float u) {
    float m = fminf(fmaxf(1.0f - u, 0.0f), 1.0f);
    float m2 = m * m;
    return m2 * m2 * m;
}
// What the Disney functions compute from the MATERIAL ALONE — the tinted specular colours of P:436-441, the lobe
// probabilities of P:547-553 / P:690-699, the two GTR widths with their squares and PI*log(a^2) of P:398-404, the anisotropic
// widths of P:456-458 — is the same for every evaluation of that material: ~145 instructions per bounce on the metric scene.
// MaterialD reads those values from the second half of the material's 128-byte record, where the upload evaluated the very
// same expressions in the context's arithmetic (material_derive_kernel); a plain Material (PT_MODE_PARITY_REF, whose
// materials live in the triangle records, and the wavefront stages) evaluates them in place.  Same values bit for bit.
struct MatSpec { V3 Cspec0, Csheen; };
struct MatLobes { float p_diffuse, p_specular, p_clearcoat; };
struct MatGtr { float alpha1, a2m1_1, pilog_1, a2_2, a2m1_2; };   // clearcoat GTR1: a, a^2 - 1, PI log(a^2); specular GTR2: a^2, a^2 - 1
struct MatAniso { float ax, ay; };
struct MaterialD : Material { const float4* ext; };               // ext = the record's derived half (4 float4)

template <class M> __device__ __forceinline__ MatSpec mat_spec(const Material& mt) {                       // P:436-441
    MatSpec r;
    V3 Cdlin = mt.baseColor;
    float Cdlum = 0.3f * Cdlin.x + 0.6f * Cdlin.y + 0.1f * Cdlin.z;
    V3 Ctint = (Cdlum > 0.0f) ? (Cdlin / Cdlum) : mk3(1, 1, 1);
    V3 Cspec = mix3(mk3(1, 1, 1), Ctint, mt.specularTint) * mt.specular;
    r.Cspec0 = mix3(Cspec * 0.08f, Cdlin, mt.metallic);
    r.Csheen = mix3(mk3(1, 1, 1), Ctint, mt.sheenTint);
    return r;
}
template <class M> __device__ __forceinline__ MatLobes mat_lobes(const Material& mt) {                     // P:547-553
    float r_diffuse = (1.0f - mt.metallic), r_specular = 1.0f, r_clearcoat = 0.25f * mt.clearcoat;
    float r_sum = r_diffuse + r_specular + r_clearcoat;
    MatLobes r;
    r.p_diffuse = r_diffuse / r_sum; r.p_specular = r_specular / r_sum; r.p_clearcoat = r_clearcoat / r_sum;
    return r;
}
template <class M> __device__ __forceinline__ MatGtr mat_gtr(const Material& mt) {
    MatGtr r;
    r.alpha1 = mixf(0.1f, 0.001f, mt.clearcoatGloss);
    float a2 = r.alpha1 * r.alpha1;
    r.a2m1_1 = a2 - 1.0f;
    r.pilog_1 = PT_PI * M::log(a2);
    float alpha2 = fmaxf(0.001f, sqr(mt.roughness));
    r.a2_2 = alpha2 * alpha2;
    r.a2m1_2 = r.a2_2 - 1.0f;
    return r;
}
template <class M> __device__ __forceinline__ MatAniso mat_aniso(const Material& mt) {                    // P:456-458
    float aspect = sqrtf(1.0f - mt.anisotropic * 0.9f);
    MatAniso r;
    r.ax = fmaxf(0.001f, sqr(mt.roughness) / aspect);
    r.ay = fmaxf(0.001f, sqr(mt.roughness) * aspect);
    return r;
}
// derived half of the record: (Cspec0, a2_2) (Csheen, alpha1) (p_diffuse, p_specular, p_clearcoat, a2m1_2) (a2m1_1, pilog_1, ax, ay)
template <class M> __device__ __forceinline__ MatSpec mat_spec(const MaterialD& mt) {
    const float4 a = ldg4(mt.ext), b = ldg4(mt.ext + 1);
    MatSpec r; r.Cspec0 = mk3(a.x, a.y, a.z); r.Csheen = mk3(b.x, b.y, b.z);
    return r;
}
template <class M> __device__ __forceinline__ MatLobes mat_lobes(const MaterialD& mt) {
    const float4 c = ldg4(mt.ext + 2);
    MatLobes r; r.p_diffuse = c.x; r.p_specular = c.y; r.p_clearcoat = c.z;
    return r;
}
template <class M> __device__ __forceinline__ MatGtr mat_gtr(const MaterialD& mt) {
    const float4 a = ldg4(mt.ext), b = ldg4(mt.ext + 1), c = ldg4(mt.ext + 2), d = ldg4(mt.ext + 3);
    MatGtr r; r.alpha1 = b.w; r.a2m1_1 = d.x; r.pilog_1 = d.y; r.a2_2 = a.w; r.a2m1_2 = c.w;
    return r;
}
template <class M> __device__ __forceinline__ MatAniso mat_aniso(const MaterialD& mt) {
    const float4 d = ldg4(mt.ext + 3);
    MatAniso r; r.ax = d.z; r.ay = d.w;
    return r;
}
// one thread per material: the derived half of each 128-byte record, in the arithmetic of the translation unit
template <class M>
__global__ void material_derive_kernel(float4* __restrict__ materials, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4* p = materials + 8 * (size_t)i;
    const float4 a = p[0], b = p[1], c = p[2], d = p[3];
    Material m;
    m.emissive = mk3(a.x, a.y, a.z); m.subsurface = a.w;
    m.baseColor = mk3(b.x, b.y, b.z); m.metallic = b.w;
    m.specular = c.x; m.specularTint = c.y; m.roughness = c.z; m.anisotropic = c.w;
    m.sheen = d.x; m.sheenTint = d.y; m.clearcoat = d.z; m.clearcoatGloss = d.w;
    const MatSpec sp = mat_spec<M>(m);
    const MatLobes lb = mat_lobes<M>(m);
    const MatGtr g = mat_gtr<M>(m);
    const MatAniso an = mat_aniso<M>(m);
    p[4] = make_float4(sp.Cspec0.x, sp.Cspec0.y, sp.Cspec0.z, g.a2_2);
    p[5] = make_float4(sp.Csheen.x, sp.Csheen.y, sp.Csheen.z, g.alpha1);
    p[6] = make_float4(lb.p_diffuse, lb.p_specular, lb.p_clearcoat, g.a2m1_2);
    p[7] = make_float4(g.a2m1_1, g.pilog_1, an.ax, an.ay);
}
template <class Cfg>
cudaError_t derive_materials(float4* materials, int n, cudaStream_t stream) {
    material_derive_kernel<typename Cfg::M><<<(n + 127) / 128, 128, 0, stream>>>(materials, n);
    return cudaGetLastError();
}

// GTR1 (P:398-404) and GTR2 (P:406-410) on the derived widths
__device__ __forceinline__ float GTR1(float NdotH, const MatGtr& g) {
    if (g.alpha1 >= 1.0f) return 1.0f / PT_PI;
    float t = 1.0f + g.a2m1_1 * NdotH * NdotH;
    return g.a2m1_1 / (g.pilog_1 * t);
}
__device__ __forceinline__ float GTR2(float NdotH, const MatGtr& g) {
    float t = 1.0f + g.a2m1_2 * NdotH * NdotH;
    return g.a2_2 / (PT_PI * t * t);
}
__device__ __forceinline__ float GTR2_aniso(float NdotH, float HdotX, float HdotY, float ax, float ay) {
    return 1.0f / (PT_PI * ax * ay * sqr(sqr(HdotX / ax) + sqr(HdotY / ay) + NdotH * NdotH));
}
__device__ __forceinline__ float smithG_GGX(float NdotV, float alphaG) {
    float a = alphaG * alphaG, b = NdotV * NdotV;
    return 1.0f / (NdotV + sqrtf(a + b - a * b));
}
__device__ __forceinline__ float smithG_GGX_aniso(float NdotV, float VdotX, float VdotY, float ax, float ay) {
    return 1.0f / (NdotV + sqrtf(sqr(VdotX * ax) + sqr(VdotY * ay) + sqr(NdotV)));
}

// Terms shared by both evaluate functions (P:432-448 == P:597-613)
struct DisneyCommon {
    V3 Cdlin, Cspec0, Csheen;
    float Fd, ss, FH, NdotH, LdotH;
};
template <class M, class MT>
__device__ __forceinline__ DisneyCommon disney_common(V3 V, V3 N, V3 L, float NdotL, float NdotV, const MT& mt) {
    DisneyCommon c;
    V3 H = normalize(L + V);
    c.NdotH = dot(N, H);
    c.LdotH = dot(L, H);
    c.Cdlin = mt.baseColor;
    const MatSpec sp = mat_spec<M>(mt);
    c.Cspec0 = sp.Cspec0;
    c.Csheen = sp.Csheen;
    float Fd90 = 0.5f + 2.0f * c.LdotH * c.LdotH * mt.roughness;
    float FL = SchlickFresnel(NdotL), FV = SchlickFresnel(NdotV);
    c.Fd = mixf(1.0f, Fd90, FL) * mixf(1.0f, Fd90, FV);
    float Fss90 = c.LdotH * c.LdotH * mt.roughness;
    float Fss = mixf(1.0f, Fss90, FL) * mixf(1.0f, Fss90, FV);
    c.ss = 1.25f * (Fss * (1.0f / (NdotL + NdotV) - 0.5f) + 0.5f);
    c.FH = SchlickFresnel(c.LdotH);
    return c;
}
template <class M, class MT>
__device__ __forceinline__ V3 disney_combine(const DisneyCommon& c, float Gs, float Ds, float NdotL, float NdotV, const MatGtr& g,
                                             const MT& mt) {
    V3 Fs = mix3(c.Cspec0, mk3(1, 1, 1), c.FH);
    V3 specular = (Fs * Gs) * Ds;
    float Dr = GTR1(c.NdotH, g);
    float Fr = mixf(0.04f, 1.0f, c.FH);
    float Gr = smithG_GGX(NdotL, 0.25f) * smithG_GGX(NdotV, 0.25f);
    float cc = 0.25f * Gr * Fr * Dr * mt.clearcoat;
    V3 Fsheen = c.Csheen * (c.FH * mt.sheen);
    V3 diffuse = c.Cdlin * ((1.0f / PT_PI) * mixf(c.Fd, c.ss, mt.subsurface)) + Fsheen;
    return diffuse * (1.0f - mt.metallic) + specular + mk3(cc, cc, cc);
}

// BRDF_Evaluate_aniso P:423-482 (the Disney variant's BRDF_Evaluate, D:381-440)
template <class M, class MT>
__device__ __forceinline__ V3 BRDF_Evaluate_aniso(V3 V, V3 N, V3 L, V3 X, V3 Y, const MT& mt) {
    float NdotL = dot(N, L), NdotV = dot(N, V);
    if (NdotL < 0.0f || NdotV < 0.0f) return mk3(0, 0, 0);
    DisneyCommon c = disney_common<M>(V, N, L, NdotL, NdotV, mt);
    V3 H = normalize(L + V);
    const MatAniso an = mat_aniso<M>(mt);
    float Ds = GTR2_aniso(c.NdotH, dot(H, X), dot(H, Y), an.ax, an.ay);
    float Gs = smithG_GGX_aniso(NdotL, dot(L, X), dot(L, Y), an.ax, an.ay);
    Gs *= smithG_GGX_aniso(NdotV, dot(V, X), dot(V, Y), an.ax, an.ay);
    return disney_combine<M>(c, Gs, Ds, NdotL, NdotV, mat_gtr<M>(mt), mt);
}
// isotropic BRDF_Evaluate of the IS shader, P:587-636 (smithG_GGX takes roughness, not alpha: P:620-621)
template <class M, class MT>
__device__ __forceinline__ V3 BRDF_Evaluate_iso(V3 V, V3 N, V3 L, const MT& mt) {
    float NdotL = dot(N, L), NdotV = dot(N, V);
    if (NdotL < 0.0f || NdotV < 0.0f) return mk3(0, 0, 0);
    DisneyCommon c = disney_common<M>(V, N, L, NdotL, NdotV, mt);
    const MatGtr g = mat_gtr<M>(mt);
    float Ds = GTR2(c.NdotH, g);
    float Gs = smithG_GGX(NdotL, mt.roughness);
    Gs *= smithG_GGX(NdotV, mt.roughness);
    return disney_combine<M>(c, Gs, Ds, NdotL, NdotV, g, mt);
}

// BRDF_Pdf, P:669-706
template <class M, class MT> __device__ __forceinline__ float BRDF_Pdf(V3 V, V3 N, V3 L, const MT& mt) {
    float NdotL = dot(N, L), NdotV = dot(N, V);
    if (NdotL < 0.0f || NdotV < 0.0f) return 0.0f;
    V3 H = normalize(L + V);
    float NdotH = dot(N, H), LdotH = dot(L, H);
    const MatGtr g = mat_gtr<M>(mt);
    float Ds = GTR2(NdotH, g);
    float Dr = GTR1(NdotH, g);
    float pdf_diffuse = NdotL / PT_PI;
    float pdf_specular = Ds * NdotH / (4.0f * LdotH);
    float pdf_clearcoat = Dr * NdotH / (4.0f * LdotH);
    const MatLobes lb = mat_lobes<M>(mt);
    float pdf = lb.p_diffuse * pdf_diffuse + lb.p_specular * pdf_specular + lb.p_clearcoat * pdf_clearcoat;
    return fmaxf(1e-10f, pdf);
}
__device__ __forceinline__ float misMixWeight(float a, float b) { float t = a * a; return t / (b * b + t); }

// SampleBRDF, P:539-570 with its three lobe samplers P:485-536
__device__ __forceinline__ V3 reflect3(V3 I, V3 N) { return I - N * (2.0f * dot(N, I)); }
template <class M, class MT>
__device__ __forceinline__ V3 SampleBRDF(float xi_1, float xi_2, float xi_3, V3 V, V3 N, const MT& mt) {
    const MatGtr g = mat_gtr<M>(mt);
    const MatLobes lb = mat_lobes<M>(mt);
    const float p_diffuse = lb.p_diffuse, p_specular = lb.p_specular;
    float rd = xi_3;
    // One toNormalHemisphere site for the three lobes (a warp on a glossy surface holds lanes of all of them): the branches
    // only produce the local vector; the same operations in the same order per lane as the reference's three functions.
    V3 v;
    bool half_vector = true;
    // The polar angle of each lobe goes through a cancellation: sqrt(1 - x^2 - y^2), 1 + (a^2 - 1) xi, 1 - cos^2.  On a
    // mirror-like surface (the metric scene's floor: alpha = 0.001, 1 - cos^2 ~ 1e-6) an ulp of cos_theta_h is 10 % of
    // sin_theta_h, i.e. the sampled direction moves by 1e-5 rad when the compiler fuses or approximates one operation —
    // enough to land on the neighbouring environment texel in ~1 % of the samples.  These few operations therefore run in
    // the oracle's arithmetic (un-fused IEEE mul/add/div/sqrt) in EVERY mode, like the ray/triangle tests do, and the
    // material-derived a^2 - 1 they consume comes from the exact-arithmetic upload kernel.  Measured (tools/fast_rrmse.py,
    // the bench's crop of the metric scene, 2 spp): rRMSE of PT_MODE_FAST against the oracle 1.14e-3 -> 1.1e-5, for no
    // measurable time (21.3 Grays/s either way); un-fusing or un-approximating the REST of the shading arithmetic buys
    // nothing more until all of it is exact (-fmad=false + IEEE div/sqrt: 4.4e-8, at 16.3 Grays/s).
    if (rd <= p_diffuse) {                                        // SampleCosineHemisphere
#if PT_STRICT_SAMPLERS
        float r = s_sqrt(xi_1);
        float theta = s_mul(s_mul(xi_2, 2.0f), PT_PI);
        float x = s_mul(r, M::cos_dir(theta)), y = s_mul(r, M::sin_dir(theta));
        v = mk3(x, y, s_sqrt(s_sub(s_sub(1.0f, s_mul(x, x)), s_mul(y, y))));
#else
        float r = sqrtf(xi_1);
        float theta = xi_2 * 2.0f * PT_PI;
        float x = r * M::cos_dir(theta), y = r * M::sin_dir(theta);
        v = mk3(x, y, sqrtf(1.0f - x * x - y * y));
#endif
        half_vector = false;
    } else {
        float cos_theta_h;
        if (p_diffuse < rd && rd <= p_diffuse + p_specular) {     // SampleGTR2
#if PT_STRICT_SAMPLERS
            cos_theta_h = s_sqrt(s_div(s_sub(1.0f, xi_2), s_add(1.0f, s_mul(g.a2m1_2, xi_2))));
#else
            cos_theta_h = sqrtf((1.0f - xi_2) / (1.0f + g.a2m1_2 * xi_2));
#endif
        } else if (p_diffuse + p_specular < rd) {                 // SampleGTR1
#if PT_STRICT_SAMPLERS
            const float a2 = s_mul(g.alpha1, g.alpha1);
            cos_theta_h = s_sqrt(s_div(s_sub(1.0f, M::pow_dir(a2, s_sub(1.0f, xi_2))), s_sub(1.0f, a2)));
#else
            const float a2 = g.alpha1 * g.alpha1;
            cos_theta_h = sqrtf((1.0f - M::pow_dir(a2, 1.0f - xi_2)) / (1.0f - a2));
#endif
        } else {
            return mk3(0, 1, 0);                                  // only reachable through NaNs
        }
#if PT_STRICT_SAMPLERS
        float phi_h = s_mul(s_mul(2.0f, PT_PI), xi_1);
        float sin_phi_h = M::sin_dir(phi_h), cos_phi_h = M::cos_dir(phi_h);
        float sin_theta_h = s_sqrt(fmaxf(0.0f, s_sub(1.0f, s_mul(cos_theta_h, cos_theta_h))));
        v = mk3(s_mul(sin_theta_h, cos_phi_h), s_mul(sin_theta_h, sin_phi_h), cos_theta_h);
#else
        float phi_h = 2.0f * PT_PI * xi_1;
        float sin_phi_h = M::sin_dir(phi_h), cos_phi_h = M::cos_dir(phi_h);
        float sin_theta_h = sqrtf(fmaxf(0.0f, 1.0f - cos_theta_h * cos_theta_h));
        v = mk3(sin_theta_h * cos_phi_h, sin_theta_h * sin_phi_h, cos_theta_h);
#endif
    }
    const V3 w = toNormalHemisphere(v, N);
    return half_vector ? reflect3(-V, w) : w;
}

// ray generation, P:846-850, in strict arithmetic (primary-hit ids must be bit-exact in every mode).
// pix = NDC of the fragment centre: ((k + 0.5) / n) * 2 - 1 (vshader.vsh:7-8 interpolated over the quad)
__device__ __forceinline__ void generate_ray(const RenderParams& p, int px, int py, uint32_t& seed, V3& o, V3& d) {
    float pixx = s_sub(s_mul(s_div(s_add((float)px, 0.5f), (float)p.width), 2.0f), 1.0f);
    float pixy = s_sub(s_mul(s_div(s_add((float)py, 0.5f), (float)p.height), 2.0f), 1.0f);
    float a1 = s_div(s_sub(rnd(seed), 0.5f), (float)p.width);
    float a2 = s_div(s_sub(rnd(seed), 0.5f), (float)p.height);
    float vx = s_add(pixx, a1), vy = s_add(pixy, a2), vz = -1.5f, vw = 0.0f;
    const float* m = p.cam;
    d = mk3(s_add(s_add(s_add(s_mul(m[0], vx), s_mul(m[4], vy)), s_mul(m[8], vz)), s_mul(m[12], vw)),
            s_add(s_add(s_add(s_mul(m[1], vx), s_mul(m[5], vy)), s_mul(m[9], vz)), s_mul(m[13], vw)),
            s_add(s_add(s_add(s_mul(m[2], vx), s_mul(m[6], vy)), s_mul(m[10], vz)), s_mul(m[14], vw)));
    d = s_normalize(d);
    o = mk3(p.eye[0], p.eye[1], p.eye[2]);
}

template <bool RAW> struct MatOf { typedef Material T; };
template <> struct MatOf<false> { typedef MaterialD T; };
__device__ __forceinline__ void set_ext(Material&, const DevScene&, int) {}
__device__ __forceinline__ void set_ext(MaterialD& m, const DevScene& sc, int mat_id) { m.ext = sc.materials + 8 * (size_t)mat_id + 4; }

// ---- one sample of one pixel: main() of the three shaders ------------------------------------------
// Returns this frame's colour BEFORE accumulation; primary_id = triangle of the primary hit (-1 miss).
// Written as a per-lane state machine around ONE traversal site: the ray in flight is the primary ray, a BRDF
// bounce ray (closest hit) or an environment shadow ray (any hit); whatever must survive the traversal is a few
// registers (f_r, N.L, pdf, history, Lo, and while a shadow ray flies: the pending bounce ray).  Lanes of a warp
// whose paths are in different phases still run the traversal loop together, and the kernel holds one copy of the
// traversal, one of the environment lookup and one of the BRDF evaluation (the megakernel's stall profile was
// dominated by instruction fetch, profiles/r01_megakernel_v2.md).  Every value is computed by the same operations
// in the same order as in the reference's straight-line code, so results are unchanged bit for bit.
#define PT_PH_PRIMARY 0
#define PT_PH_BOUNCE 1
#define PT_PH_SHADOW 2

template <class Cfg, int VARIANT>
__device__ __forceinline__ V3 shade_sample(const RenderParams& p, int px, int py, uint32_t frame, const float* __restrict__ sobol_row,
                                           int& primary_id, uint32_t& nrays, Counter<Cfg::COUNT>& cnt) {
    typedef typename Cfg::M M;
    const DevScene& sc = p.sc;
    uint32_t seed = ((uint32_t)px * 1973u + (uint32_t)py * 9277u + frame * 26699u) | 1u;            // P:73-76

    V3 o, d;
    generate_ray(p, px, py, seed, o, d);

    V3 Le0 = mk3(0, 0, 0), Lo = mk3(0, 0, 0), history = mk3(1, 1, 1);
    V3 f_r = mk3(0, 0, 0);
    float cosw = 0.0f, pdf = 1.0f;                    // of the bounce ray in flight, or pending behind a shadow ray
    V3 fl = mk3(0, 0, 0), nextL = mk3(0, 0, 0);       // shadow phase: BRDF value toward the light; the pending bounce ray
    float ndl = 0.0f, pdf_b = 0.0f;                   //               N.Ld and the BRDF's pdf of Ld
    bool next_ok = false;
    uint32_t light_idx = 0;                           // hdrCache texel of the light sample in flight (Cfg::LIGHT_TAB)
    int bounce = -1;
    int phase = PT_PH_PRIMARY;
    primary_id = -1;
    while (true) {
        const TraceResult tr = trace<Cfg>(sc, o, d, phase == PT_PH_SHADOW, cnt, p.eye_far && phase == PT_PH_PRIMARY);
        nrays++;
        if (VARIANT == 2 && Cfg::LIGHT_TAB && tr.tri < 0 && phase == PT_PH_SHADOW) {
            // ---- the shadow ray reached the environment: colour and pdf of its texel come from the light table (P:781-792) ----
            const float4 e = ldg4(sc.light_tab + 2 * (size_t)light_idx + 1);
            cnt.texel(2u);
            const float pdf_light = e.w;
            const float mis_weight = misMixWeight(pdf_light, pdf_b);
            Lo = Lo + (history * mis_weight) * mk3(e.x, e.y, e.z) * fl * ndl / pdf_light;
        } else if (tr.tri < 0) {
            // ---- the ray left the scene: one environment lookup serves all three phases ----
            float u, w;
            spherical_uv(d, u, w);
            const float4 c = tex_nearest(sc.hdr4, sc.hdrW, sc.hdrH, u, w);
            cnt.texel(VARIANT == 2 && phase != PT_PH_PRIMARY ? 2u : 1u);
            if (phase == PT_PH_PRIMARY) return mk3(fminf(c.x, 10.0f), fminf(c.y, 10.0f), fminf(c.z, 10.0f));   // P:857-859
            if (VARIANT == 2) {
                const V3 color = mk3(c.x, c.y, c.z);                                                  // hdrColor, P:647-651
                const float pdf_light = hdr_pdf<M, Cfg::ENV_CDF>(sc, u, w);                           // hdrPdf, P:655-666
                if (!Cfg::LIGHT_TAB && phase == PT_PH_SHADOW) {                                       // P:781-792
                    const float mis_weight = misMixWeight(pdf_light, pdf_b);
                    Lo = Lo + (history * mis_weight) * color * fl * ndl / pdf_light;
                } else {                                                                              // P:819-830
                    const float mis_weight = misMixWeight(pdf, pdf_light);
                    Lo = Lo + (history * mis_weight) * color * f_r * cosw / pdf;
                    break;
                }
            } else {                                                                                  // O:349-353, D:465-469
                const V3 sky = mk3(fminf(c.x, 10.0f), fminf(c.y, 10.0f), fminf(c.z, 10.0f));
                Lo = Lo + history * sky * f_r * cosw / pdf;
                break;
            }
        } else if (phase != PT_PH_SHADOW) {
            // ---- closest hit of the primary or a bounce ray ----
            Hit hit;
            int mat_id;
            finalize_hit<Cfg::RAW>(sc, o, d, tr, hit, mat_id);
            cnt.shaded();
            typename MatOf<Cfg::RAW>::T mat;
            load_material<Cfg::RAW>(sc, mat_id, mat);
            set_ext(mat, sc, mat_id);
            if (phase == PT_PH_PRIMARY) {
                primary_id = tr.tri;
                Le0 = mat.emissive;                                                                   // P:862
            } else {
                Lo = Lo + history * mat.emissive * f_r * cosw / pdf;                                  // P:833-834
                history = history * (f_r * cosw / pdf);                                               // P:837
            }
            bounce++;
            if (bounce >= p.max_bounce) break;
            const V3 V = -hit.viewDir, N = hit.normal;
            o = hit.hitPoint;
            if (VARIANT == 2) {
                // light sampling through the hdrCache, P:770-775: the shadow ray flies first, the BRDF ray waits
                const float e1 = rnd(seed);
                const float e2 = rnd(seed);
                V3 Ld;
                if (Cfg::LIGHT_TAB) {
                    light_idx = (uint32_t)nearest_texel(e2, sc.hdrH) * (uint32_t)sc.hdrW + (uint32_t)nearest_texel(e1, sc.hdrW);
                    const float4 r0 = ldg4(sc.light_tab + 2 * (size_t)PT_IDX(sc, light_idx, (size_t)sc.hdrW * sc.hdrH));
                    Ld = mk3(r0.x, r0.y, r0.z);
                } else {
                    Ld = SampleHdr<M, Cfg::ENV_CDF>(sc, e1, e2);
                }
                cnt.texel(1u);
                ndl = dot(N, Ld);
                const bool want_shadow = ndl > 0.0f;
                // BRDF sampling driven by Sobol + Cranley-Patterson, P:795-816
                float su, sv;
                sobol_pair(sobol_row, (uint32_t)bounce, frame, su, sv);
                cranley_patterson(px, py, su, sv);
                const float xi_3 = rnd(seed);
                const V3 L = SampleBRDF<M>(su, sv, xi_3, V, N, mat);
                cosw = dot(N, L);
                next_ok = cosw > 0.0f;
                // BRDF value and pdf toward the light (k = 0) and along the sampled direction (k = 1): one code copy
#pragma unroll 1
                for (int k = 0; k < 2; k++) {
                    if (k == 0 ? !want_shadow : !next_ok) continue;
                    const V3 Lk = k == 0 ? Ld : L;
                    const V3 f = BRDF_Evaluate_iso<M>(V, N, Lk, mat);
                    const float pf = BRDF_Pdf<M>(V, N, Lk, mat);
                    if (k == 0) { fl = f; pdf_b = pf; }
                    else { f_r = f; pdf = pf; }
                }
                if (next_ok && pdf <= 0.0f) next_ok = false;  // the reference traces this ray and then discards it (P:811-816)
                if (want_shadow) { phase = PT_PH_SHADOW; d = Ld; nextL = L; continue; }
                if (!next_ok) break;
                phase = PT_PH_BOUNCE; d = L;
            } else {
                // uniform hemisphere, O:334-346 / D:448-456
                const float z = rnd(seed);
                const float r = fmaxf(0.0f, sqrtf(1.0f - z * z));
                const float phi = 2.0f * PT_PI * rnd(seed);
                const V3 L = toNormalHemisphere(mk3(r * M::cos_dir(phi), r * M::sin_dir(phi), z), N);
                pdf = 1.0f / (2.0f * PT_PI);
                cosw = fmaxf(0.0f, dot(L, N));
                if (VARIANT == 1) {
                    V3 tangent, bitangent;
                    getTangent(N, tangent, bitangent);
                    f_r = BRDF_Evaluate_aniso<M>(V, N, L, tangent, bitangent, mat);
                } else {
                    f_r = mat.baseColor / PT_PI;
                }
                phase = PT_PH_BOUNCE; d = L;
            }
            continue;
        }
        // ---- a shadow ray has been resolved (occluded, or its light term added): release the pending bounce ray ----
        if (!next_ok) break;
        phase = PT_PH_BOUNCE; d = nextL;
    }
    return Le0 + Lo;
}

// adds a thread's traversal counters to the global ones (counting configuration only)
__device__ __forceinline__ void flush_counts(Counter<false>&, unsigned long long*) {}
__device__ __forceinline__ void flush_counts(Counter<true>& cnt, unsigned long long* out) {
    uint32_t a = cnt.c.inner, b = cnt.c.tris, c = cnt.c.cand, d = cnt.c.ties, e = cnt.c.shaded, f = cnt.c.texels;
    for (int o = 16; o > 0; o >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o); c += __shfl_xor_sync(0xffffffffu, c, o);
        d += __shfl_xor_sync(0xffffffffu, d, o);
        e += __shfl_xor_sync(0xffffffffu, e, o); f += __shfl_xor_sync(0xffffffffu, f, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(out, (unsigned long long)a); atomicAdd(out + 1, (unsigned long long)b); atomicAdd(out + 2, (unsigned long long)c);
        if (d) atomicAdd(out + 3, (unsigned long long)d);
        atomicAdd(out + 4, (unsigned long long)e); atomicAdd(out + 5, (unsigned long long)f);
    }
}

// ---- render kernel: one thread per pixel, a warp covers an 8x4 pixel tile, frames looped in registers ----

template <class Cfg, int VARIANT>
__global__ void __launch_bounds__(256) render_kernel(const __grid_constant__ RenderParams p) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const TileMap tm = make_tile_map(p.x0, p.y0, p.x1, p.y1, p.band_world, p.band_rank, p.band_diag);
    const uint32_t tile = blockIdx.x * 8u + (uint32_t)warp;          // one owned 8x4 tile per warp
    int tx = 0, strip = 0;
    if (tile < tm.count()) tm.locate(tile, tx, strip);
    const int px = tile < tm.count() ? p.x0 + tx * 8 + (lane & 7) : p.x1;
    const int py = p.y0 + strip * 4 + (lane >> 3);
    uint32_t nrays = 0;
    Counter<Cfg::COUNT> cnt;
    if (px < p.x1 && py < p.y1) {
        const size_t idx = (size_t)py * p.width + px;
        float4 acc = p.accum[idx];
        int primary = -1;
        const long long t_start = p.cost ? clock64() : 0;
        for (uint32_t f = 0; f < p.n_frames; f++) {
            const uint32_t frame = p.first_frame + f * p.frame_stride;
            V3 c = shade_sample<Cfg, VARIANT>(p, px, py, frame, p.sobol_tab ? p.sobol_tab + (size_t)f * 8 : nullptr, primary, nrays, cnt);
            if (p.accumulate_sum) {
                acc.x = s_add(acc.x, c.x); acc.y = s_add(acc.y, c.y); acc.z = s_add(acc.z, c.z);
            } else {
                // mix(lastColor, color, 1/(frameCounter+1)), P:868-869 — strict so the running mean is mode-independent
                float a = s_div(1.0f, (float)(frame + 1u));
                float b = s_sub(1.0f, a);
                acc.x = s_add(s_mul(acc.x, b), s_mul(c.x, a));
                acc.y = s_add(s_mul(acc.y, b), s_mul(c.y, a));
                acc.z = s_add(s_mul(acc.z, b), s_mul(c.z, a));
            }
        }
        acc.w = 1.0f;
        p.accum[idx] = acc;
        if (p.primary_ids) p.primary_ids[idx] = primary;
        if (p.cost) p.cost[idx] = (unsigned)(clock64() - t_start);
    }
    // one atomic per warp for the ray counter
    for (int o = 16; o > 0; o >>= 1) nrays += __shfl_xor_sync(0xffffffffu, nrays, o);
    if (lane == 0 && nrays) atomicAdd(p.ray_counter, (unsigned long long)nrays);
    flush_counts(cnt, p.work_counters);
}

// ---- persistent render kernel: (8x4 pixel tile, frame) work items, frame-major, claimed dynamically ----------
// The one-thread-per-pixel kernel above runs the n_frames of a pixel back to back in one thread, so a launch lasts
// at least as long as its most expensive pixel times n_frames (on is-1M: a mesh tile = 3.6 ms of a 6.1 ms launch,
// SMs 69 % busy, profiles/r01_megakernel_v2.md).  Here a resident warp repeatedly claims the next (tile, frame)
// item, computes that one sample for its 32 pixels, and folds it into the framebuffer IN FRAME ORDER: a lane that
// finishes frame f early waits for f-1 of its pixel (which was claimed before it, hence is running or done: no
// deadlock).  The running mean therefore sees the same sequence of operations as the reference's ping-pong
// (P:868-869) and results stay bit-identical.
//
// The frame a pixel is ready for travels in the .w of the pixel's own float4, so the "flag" and the data are ONE
// aligned 16-byte word — written by one st.relaxed.gpu.b128 and read by one ld.relaxed.gpu.b128, each ONE access in the
// PTX memory model and a single L2 transaction.  No fence, no release/acquire pair, no warp
// barrier and no separate progress array: the poll IS the accumulator read.  Within a launch w = PT_TAG | (next frame
// index of the launch); the launch's last frame stores w = 1 (the reference's alpha, P:871); the first frame of a
// launch waits for nothing (kernel boundary).  A NaN pattern cannot collide with the 1.0f / 0.0f a finished or cleared
// framebuffer holds.  (Measured against the protocol it replaced — lane 0 ld.acquire on a per-tile progress word,
// __syncwarp, __ldcg, __stcg, __syncwarp, lane 0 st.release — and against K consecutive frames per item, which
// PT_ITEM_FRAMES still selects, and an L2 prefetch of the accumulator line: DESIGN.md §5, §9.)
#define PT_TAG 0xffc00000u
#ifndef PT_ITEM_FRAMES
#define PT_ITEM_FRAMES 1           // consecutive frames of a tile computed by one claim and folded together
#endif
// ONE 128-bit access each: `.b128` (PTX ISA 8.3+) is a single scalar access in the PTX memory model — tag and data are
// read, and written, as one morally-strong relaxed operation (a `.v4.u32` access is, formally, four scalar accesses in
// unspecified order, which would allow "new tag, old rgb").  SASS: LDG.E.128.STRONG.GPU / STG.E.128.STRONG.GPU.
#if defined(PT_FOLD_V4) && PT_FOLD_V4       // the round-1 form (A/B only, tools/build_variant.py): same SASS opcodes, four scalar accesses in the PTX model
__device__ __forceinline__ uint4 ld_relaxed_v4(const float4* p) {
    uint4 v;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_v4(float4* p, float x, float y, float z, uint32_t w) {
    asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(__float_as_uint(x)), "r"(__float_as_uint(y)),
                 "r"(__float_as_uint(z)), "r"(w) : "memory");
}
#else
__device__ __forceinline__ uint4 ld_relaxed_v4(const float4* p) {
    unsigned long long lo, hi;
    asm volatile("{\n\t.reg .b128 t;\n\tld.relaxed.gpu.global.b128 t, [%2];\n\tmov.b128 {%0, %1}, t;\n\t}" : "=l"(lo), "=l"(hi) : "l"(p) : "memory");
    uint4 v;
    v.x = (uint32_t)lo; v.y = (uint32_t)(lo >> 32); v.z = (uint32_t)hi; v.w = (uint32_t)(hi >> 32);
    return v;
}
__device__ __forceinline__ void st_relaxed_v4(float4* p, float x, float y, float z, uint32_t w) {
    const unsigned long long lo = (unsigned long long)__float_as_uint(x) | ((unsigned long long)__float_as_uint(y) << 32);
    const unsigned long long hi = (unsigned long long)__float_as_uint(z) | ((unsigned long long)w << 32);
    asm volatile("{\n\t.reg .b128 t;\n\tmov.b128 t, {%1, %2};\n\tst.relaxed.gpu.global.b128 [%0], t;\n\t}" ::"l"(p), "l"(lo), "l"(hi) : "memory");
}
#endif

#define PT_PERSIST_CHUNK 1              // (tile, frame) items claimed per atomic
#ifndef PT_PERSIST_THREADS
#define PT_PERSIST_THREADS 256
#endif
#ifndef PT_PERSIST_MIN_CTAS
#define PT_PERSIST_MIN_CTAS 5           // measured on B200 (is-1M / lambert-100k / disney-1M, Grays/s): 3 CTAs 11.6/5.7/5.3, 4: 12.4/6.5/6.1, 6: 12.6/7.0/6.8, 8: 12.1/6.9/6.7
#endif
// the t0-th tile of a frame in visiting order (S interleaved passes over the tile rows, see render_persistent) -> (tx, strip)
__device__ __forceinline__ void persist_tile(const TileMap& tm, uint32_t S, uint32_t t0, int& tx, int& strip) {
    const uint32_t row0 = t0 / (uint32_t)tm.per_row, j0 = t0 - row0 * (uint32_t)tm.per_row;
    const uint32_t R = (uint32_t)tm.rows, q = R / S, rem = R - q * S;
    uint32_t a, k;
    if (row0 < rem * (q + 1)) { a = row0 / (q + 1); k = row0 - a * (q + 1); }
    else { const uint32_t i2 = row0 - rem * (q + 1); a = rem + i2 / q; k = i2 - (a - rem) * q; }
    const uint32_t tile = (a + k * S) * (uint32_t)tm.per_row + j0;
    tm.locate(tile, tx, strip);
}
// That mapping costs seven 32-bit integer divisions (~120 warp instructions) per work item, 5 % of ALL instructions the kernel
// executes on the metric scene (2,490 per item on average, ~500 for a sky tile), and it is the same for every frame: one
// small launch tabulates it (PT_TILE_TAB), the render kernel reads 4 bytes per item.
#ifndef PT_TILE_TAB
#define PT_TILE_TAB 1
#endif
static __global__ void tile_order_kernel(const TileMap tm, uint32_t S, uint32_t ntiles, uint32_t* __restrict__ tab) {
    const uint32_t t0 = blockIdx.x * blockDim.x + threadIdx.x;
    if (t0 >= ntiles) return;
    int tx, strip;
    persist_tile(tm, S, t0, tx, strip);
    tab[t0] = (uint32_t)tx | ((uint32_t)strip << 16);
}

template <class Cfg, int VARIANT>
__global__ void __launch_bounds__(PT_PERSIST_THREADS, PT_PERSIST_MIN_CTAS) render_persistent(const __grid_constant__ RenderParams p) {
    const int lane = threadIdx.x & 31;
    const TileMap tm = make_tile_map(p.x0, p.y0, p.x1, p.y1, p.band_world, p.band_rank, p.band_diag);
    const uint32_t ntiles = tm.count();                // owned 8x4 warp tiles
    const uint32_t ngroups = (p.n_frames + PT_ITEM_FRAMES - 1u) / PT_ITEM_FRAMES;   // an item = PT_ITEM_FRAMES consecutive frames of one tile
    const uint32_t nitems = ntiles * ngroups;
    // Order of the tile rows inside a frame: S passes over the rows a, a + S, a + 2S, ... (a = 0 .. S-1) instead of bottom to
    // top.  On the metric scene the lower half of the image is floor (three rays and two BRDF evaluations per sample: issue
    // bound) and the upper half sky (one trivial ray: bound by the framebuffer stream); bottom to top, the resident warps
    // are all of one kind at any time.  With one pass ~ 1.35 x the resident warps the two kinds overlap: +5.8 % at 1080p
    // (S = 8), +6.5 % at 4K (S = 32), the measured optimum at both; scenes without such a split (variants 0 and 1 here) are
    // neutral to -2.5 %, so the rule is applied to variant 2 only.  PT_SCHED_STRIDE overrides.  DESIGN.md §5.
    uint32_t S = p.sched_stride;
    if (S == 0u) S = VARIANT == 2 ? max(1u, (uint32_t)(0.74f * (float)ntiles / (float)(gridDim.x * (PT_PERSIST_THREADS / 32u)) + 0.5f)) : 1u;
    S = min(S, max(1u, (uint32_t)tm.rows));
    uint32_t* head = p.sched;                          // [0] = next item
    uint32_t nrays = 0;
    Counter<Cfg::COUNT> cnt;
    for (;;) {
        uint32_t first_item = 0;
        if (lane == 0) first_item = atomicAdd(head, (uint32_t)PT_PERSIST_CHUNK);
        first_item = __shfl_sync(0xffffffffu, first_item, 0);
        if (first_item >= nitems) break;
        // a claimed chunk is worked off in increasing item order; the item a warp may have to wait for (same tile, previous
        // frame = item - ntiles) always lies in an EARLIER chunk, so waits point strictly backwards: no deadlock
#pragma unroll 1
        for (uint32_t item = first_item; item < first_item + PT_PERSIST_CHUNK && item < nitems; item++) {
            const uint32_t f = item / ntiles, t0 = item - f * ntiles;
            int tx, strip;
            if (PT_TILE_TAB && p.tile_tab != nullptr) {
                const uint32_t packed = __ldg(p.tile_tab + t0);
                tx = (int)(packed & 0xffffu); strip = (int)(packed >> 16);
            } else {
                persist_tile(tm, S, t0, tx, strip);
            }
            const int px = p.x0 + tx * 8 + (lane & 7);
            const int py = p.y0 + strip * 4 + (lane >> 3);
            const bool valid = px < p.x1 && py < p.y1;
            const long long t_start = p.cost ? clock64() : 0;
            const uint32_t fa = f * PT_ITEM_FRAMES;         // first frame (index within the launch) of this item
            const uint32_t nk = min((uint32_t)PT_ITEM_FRAMES, p.n_frames - fa);
            V3 c[PT_ITEM_FRAMES];
            int primary = -1;
#pragma unroll 1
            for (uint32_t k2 = 0; k2 < nk; k2++) {
                const uint32_t frame = p.first_frame + (fa + k2) * p.frame_stride;
                V3 ck = mk3(0, 0, 0);
                if (valid) ck = shade_sample<Cfg, VARIANT>(p, px, py, frame, p.sobol_tab ? p.sobol_tab + (size_t)(fa + k2) * 8 : nullptr, primary, nrays, cnt);
                if (PT_ITEM_FRAMES == 1) c[0] = ck; else c[k2] = ck;
            }
            // ordered accumulation
            if (valid) {
                float4* const slot = p.accum + PT_IDX(p.sc, (size_t)py * p.width + px, (size_t)p.width * p.height);
                uint4 raw = ld_relaxed_v4(slot);
                if (fa != 0u) {
                    const uint32_t want = PT_TAG | (fa & 0x3fffffu);
                    while (raw.w != want) { __nanosleep(32); raw = ld_relaxed_v4(slot); }
                }
                if (p.cost) __threadfence();               // profiling hook only: the cost word rides behind the tag
                float ax = __uint_as_float(raw.x), ay = __uint_as_float(raw.y), az = __uint_as_float(raw.z);
#pragma unroll
                for (uint32_t k2 = 0; k2 < PT_ITEM_FRAMES; k2++) {
                    if (k2 >= nk) break;
                    if (p.accumulate_sum) {
                        ax = s_add(ax, c[k2].x); ay = s_add(ay, c[k2].y); az = s_add(az, c[k2].z);
                    } else {
                        // mix(lastColor, color, 1/(frameCounter+1)), P:868-869, in the reference's frame order
                        const float a = s_div(1.0f, (float)(p.first_frame + (fa + k2) * p.frame_stride + 1u));
                        const float b = s_sub(1.0f, a);
                        ax = s_add(s_mul(ax, b), s_mul(c[k2].x, a));
                        ay = s_add(s_mul(ay, b), s_mul(c[k2].y, a));
                        az = s_add(s_mul(az, b), s_mul(c[k2].z, a));
                    }
                }
                const size_t idx = (size_t)py * p.width + px;
                const bool last = fa + nk == p.n_frames;
                if (last && p.primary_ids) p.primary_ids[idx] = primary;
                if (p.cost) {
                    const unsigned dt = (unsigned)(clock64() - t_start);
                    __stcg(p.cost + idx, fa == 0 ? dt : __ldcg(p.cost + idx) + dt);
                    __threadfence();
                }
                st_relaxed_v4(slot, ax, ay, az, last ? 0x3f800000u : (PT_TAG | ((fa + nk) & 0x3fffffu)));
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) nrays += __shfl_xor_sync(0xffffffffu, nrays, o);
    if (lane == 0 && nrays) atomicAdd(p.ray_counter, (unsigned long long)nrays);
    flush_counts(cnt, p.work_counters);
}

template <class Cfg, int VARIANT>
cudaError_t launch_persistent_variant(const RenderParams& p, int num_sms, cudaStream_t stream) {
    const TileMap tm = make_tile_map(p.x0, p.y0, p.x1, p.y1, p.band_world, p.band_rank, p.band_diag);
    const size_t ntiles = tm.count();
    if (ntiles == 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(p.sched, 0, 64 * sizeof(uint32_t), stream);   // the work head (its own 256 bytes)
    if (e != cudaSuccess) return e;
    if (VARIANT == 2 && p.sobol_tab)
        sobol_table_kernel<<<(p.n_frames * 8 + 255) / 256, 256, 0, stream>>>(const_cast<float*>(p.sobol_tab), p.first_frame, p.n_frames, p.frame_stride);
    // one resident set of CTAs: SMs x (CTAs that fit per SM), never more warps than there are tiles
    void (*kern)(const RenderParams) = render_persistent<Cfg, VARIANT>;
    int per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, PT_PERSIST_THREADS, 0);
    if (e != cudaSuccess) return e;
    int grid = num_sms * (per_sm > 0 ? per_sm : 1);
    const int need = (int)((ntiles + PT_PERSIST_THREADS / 32 - 1) / (PT_PERSIST_THREADS / 32));
    if (grid > need) grid = need;
    if (PT_TILE_TAB && p.tile_tab != nullptr) {
        // the visiting order of a frame's tiles, tabulated with the kernel's own rule for S (which depends on this grid)
        uint32_t S = p.sched_stride;
        if (S == 0u) S = VARIANT == 2 ? std::max(1u, (uint32_t)(0.74f * (float)ntiles / (float)((uint32_t)grid * (PT_PERSIST_THREADS / 32u)) + 0.5f)) : 1u;
        S = std::min(S, std::max(1u, (uint32_t)tm.rows));
        tile_order_kernel<<<(unsigned)((ntiles + 255) / 256), 256, 0, stream>>>(tm, S, (uint32_t)ntiles, const_cast<uint32_t*>(p.tile_tab));
    }
    kern<<<grid, PT_PERSIST_THREADS, 0, stream>>>(p);
    return cudaGetLastError();
}
template <class Cfg>
cudaError_t launch_persistent(const RenderParams& p, int variant, int num_sms, cudaStream_t stream) {
    if (variant == 0) return launch_persistent_variant<Cfg, 0>(p, num_sms, stream);
    if (variant == 1) return launch_persistent_variant<Cfg, 1>(p, num_sms, stream);
    return launch_persistent_variant<Cfg, 2>(p, num_sms, stream);
}

template <class Cfg>
cudaError_t launch_render(const RenderParams& p, int variant, cudaStream_t stream) {
    const TileMap tm = make_tile_map(p.x0, p.y0, p.x1, p.y1, p.band_world, p.band_rank, p.band_diag);
    if (tm.count() == 0) return cudaSuccess;
    dim3 grid((tm.count() + 7) / 8), block(256);                                     // one warp per owned 8x4 tile
    if (variant == 2 && p.sobol_tab)
        sobol_table_kernel<<<(p.n_frames * 8 + 255) / 256, 256, 0, stream>>>(const_cast<float*>(p.sobol_tab), p.first_frame, p.n_frames, p.frame_stride);
    if (variant == 0) render_kernel<Cfg, 0><<<grid, block, 0, stream>>>(p);
    else if (variant == 1) render_kernel<Cfg, 1><<<grid, block, 0, stream>>>(p);
    else render_kernel<Cfg, 2><<<grid, block, 0, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace pt
