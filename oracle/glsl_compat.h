/* oracle/glsl_compat.h — TEST INFRASTRUCTURE (never linked into the product).
 *
 * A GLSL-on-C++ prelude.  oracle/build_ref.py takes the reference's fragment shader text where it lies
 * under /root/reference (e.g. ImportanceSampling_LowDiscrepancySequence/shaders/pass1.fsh), strips
 * comments / `#version`, adds `f` suffixes to float literals (GLSL literals are 32-bit), turns the
 * few multi-component swizzles into calls (.xyz -> .xyz()) and `inout T x` into `T& x`, and writes
 * the result to oracle/_ref/ (git-ignored).  oracle/ref_shader_wrap.cpp then includes that text as
 * the BODY OF A C++ STRUCT with this header in scope, so the reference's own shader source runs on
 * the host cores, one struct instance per fragment.  No reference source is stored in this repo.
 *
 * Semantics chosen where GLSL leaves them open (these are the definitions the oracle restatement
 * and the CUDA parity kernels reproduce):
 *   - all arithmetic is IEEE binary32, no contraction (build with -ffp-contract=off);
 *   - dot(a,b) = (ax*bx + ay*by) + az*bz ; cross as written below ; normalize(v) = v * (1/sqrt(dot(v,v)));
 *   - mix(x,y,a) = x*(1-a) + y*a ; reflect(I,N) = I - 2*dot(N,I)*N ; clamp = min(max(x,lo),hi);
 *   - min/max follow IEEE minNum/maxNum (fminf/fmaxf), which is what NVIDIA hardware's FMNMX does;
 *   - mat4*vec4 = ((c0*x + c1*y) + c2*z) + c3*w ;
 *   - sin/cos/atan/asin/log/pow: include/ptmath.h when REF_MATH_PTMATH is defined, else the C libm;
 *   - texelFetch on a GL_RGB32F buffer texture = 3 consecutive floats; texture2D with
 *     GL_NEAREST + GL_CLAMP_TO_EDGE (main.cpp:185-195) = texel floor(u*W), clamped to [0, W-1].
 */
#ifndef GLSL_COMPAT_H
#define GLSL_COMPAT_H

#include <math.h>
#include <stdint.h>
#include "ptmath.h"

namespace glslc {

typedef unsigned int uint;

struct vec2 {
    union { float x; float r; };
    union { float y; float g; };
    vec2() : x(0), y(0) {}
    explicit vec2(float s) : x(s), y(s) {}
    vec2(float a, float b) : x(a), y(b) {}
};
struct vec3 {
    union { float x; float r; };
    union { float y; float g; };
    union { float z; float b; };
    vec3() : x(0), y(0), z(0) {}
    explicit vec3(float s) : x(s), y(s), z(s) {}
    vec3(float a, float b_, float c) : x(a), y(b_), z(c) {}
    vec2 xy() const { return vec2(x, y); }
    vec2 rg() const { return vec2(x, y); }
};
struct vec4 {
    union { float x; float r; };
    union { float y; float g; };
    union { float z; float b; };
    union { float w; float a; };
    vec4() : x(0), y(0), z(0), w(0) {}
    vec4(float a_, float b_, float c, float d) : x(a_), y(b_), z(c), w(d) {}
    vec4(vec3 v, float d) : x(v.x), y(v.y), z(v.z), w(d) {}
    vec4(vec2 v, float c, float d) : x(v.x), y(v.y), z(c), w(d) {}
    vec3 xyz() const { return vec3(x, y, z); }
    vec3 rgb() const { return vec3(x, y, z); }
    vec2 rg() const { return vec2(x, y); }
    vec2 xy() const { return vec2(x, y); }
};
struct ivec3 {
    int x, y, z;
    explicit ivec3(vec3 v) : x((int)v.x), y((int)v.y), z((int)v.z) {}
};
struct mat4 {
    vec4 c[4];
};

/* ---- vec2 ---- */
inline vec2 operator+(vec2 a, vec2 b) { return vec2(a.x + b.x, a.y + b.y); }
inline vec2 operator+(vec2 a, float s) { return vec2(a.x + s, a.y + s); }
inline vec2 operator*(vec2 a, float s) { return vec2(a.x * s, a.y * s); }
inline vec2& operator/=(vec2& a, vec2 b) { a.x = a.x / b.x; a.y = a.y / b.y; return a; }
inline vec2& operator+=(vec2& a, float s) { a.x = a.x + s; a.y = a.y + s; return a; }

/* ---- vec3 ---- */
inline vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(vec3 a, vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator/(vec3 a, vec3 b) { return vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline vec3 operator*(vec3 a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, vec3 a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(vec3 a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline vec3 operator/(float s, vec3 a) { return vec3(s / a.x, s / a.y, s / a.z); }
inline vec3 operator+(vec3 a, float s) { return vec3(a.x + s, a.y + s, a.z + s); }
inline vec3& operator+=(vec3& a, vec3 b) { a = a + b; return a; }
inline vec3& operator*=(vec3& a, vec3 b) { a = a * b; return a; }

/* ---- vec4 / mat4 ---- */
inline vec4 operator*(vec4 a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline vec4 operator+(vec4 a, vec4 b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline vec4 operator*(const mat4& m, vec4 v) {
    return ((m.c[0] * v.x + m.c[1] * v.y) + m.c[2] * v.z) + m.c[3] * v.w;
}

/* ---- scalar built-ins ---- */
inline float min(float a, float b) { return fminf(a, b); }
inline float max(float a, float b) { return fmaxf(a, b); }
inline float abs(float a) { return fabsf(a); }
inline float sqrt(float a) { return sqrtf(a); }
inline float clamp(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
inline float mix(float x, float y, float a) { return x * (1.0f - a) + y * a; }
#ifdef REF_MATH_PTMATH
inline float sin(float a) { return pt_sinf(a); }
inline float cos(float a) { return pt_cosf(a); }
inline float asin(float a) { return pt_asinf(a); }
inline float atan(float y, float x) { return pt_atan2f(y, x); }
inline float log(float a) { return pt_logf(a); }
inline float pow(float a, float b) { return pt_powf(a, b); }
#else
inline float sin(float a) { return sinf(a); }
inline float cos(float a) { return cosf(a); }
inline float asin(float a) { return asinf(a); }
inline float atan(float y, float x) { return atan2f(y, x); }
inline float log(float a) { return logf(a); }
inline float pow(float a, float b) { return powf(a, b); }
#endif

/* ---- vector built-ins ---- */
inline vec3 min(vec3 a, vec3 b) { return vec3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
inline vec3 max(vec3 a, vec3 b) { return vec3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }
inline float dot(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline vec3 cross(vec3 a, vec3 b) {
    return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
inline vec3 normalize(vec3 v) { return v * (1.0f / sqrtf(dot(v, v))); }
inline vec3 reflect(vec3 I, vec3 N) { return I - (2.0f * dot(N, I)) * N; }
inline vec3 mix(vec3 x, vec3 y, float a) { return x * (1.0f - a) + y * a; }

/* ---- samplers ---- */
struct samplerBuffer {
    const float* data;   /* GL_RGB32F texels: 3 floats each */
};
struct sampler2D {
    const float* data;   /* row 0 = t=0 */
    int w, h, ch;        /* ch = floats per texel (3 for hdr maps, 4 for lastFrame) */
};
inline vec4 texelFetch(const samplerBuffer& s, int i) {
    const float* p = s.data + 3 * (long)i;
    return vec4(p[0], p[1], p[2], 1.0f);
}
inline int nearest_texel(float u, int n) {
    float f = floorf(u * (float)n);
    f = fminf(fmaxf(f, 0.0f), (float)(n - 1));
    return (int)f;
}
inline vec4 texture2D(const sampler2D& s, vec2 uv) {
    int i = nearest_texel(uv.x, s.w), j = nearest_texel(uv.y, s.h);
    const float* p = s.data + (long)s.ch * ((long)j * s.w + i);
    return vec4(p[0], p[1], p[2], 1.0f);
}

} /* namespace glslc */
#endif
