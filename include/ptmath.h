/* ptmath.h — portable single-precision transcendental functions.
 *
 * The reference shaders (ImportanceSampling_LowDiscrepancySequence/shaders/pass1.fsh) call the GLSL
 * built-ins sin/cos/atan/asin/log/pow.  GLSL does not define those bit-for-bit, so "the same result as
 * the reference" needs ONE concrete definition that the CPU oracle and the CUDA parity kernels can
 * both evaluate.  This header is that definition: every function is a fixed sequence of IEEE-754
 * binary32 +,-,*,/ and sqrt operations (Cody–Waite range reduction + the classic Cephes minimax
 * polynomials), so gcc (-ffp-contract=off) and nvcc (-fmad=false) produce identical bits.
 * Accuracy is about 1–2 ulp on the ranges the shaders use (checked in tests/test_ptmath.py).
 *
 * Rules for this file: no FMA, no double, no libm calls except sqrtf/floorf (both exact by IEEE),
 * only float<->int conversions that are exact or truncate.  Every multiply/add/subtract/divide goes
 * through the PTM_* macros: in CUDA device code they are the _rn intrinsics (never fused, immune to
 * -fmad / --use_fast_math), on the host they are the plain operators (compile with -ffp-contract=off).
 */
#ifndef PTMATH_H
#define PTMATH_H

#include <math.h>
#include <string.h>

#ifdef __CUDACC__
#define PTM_FN static __host__ __device__ __forceinline__
#else
#define PTM_FN static inline
#endif

#if defined(__CUDA_ARCH__)
#define PTM_MUL(a, b) __fmul_rn((a), (b))
#define PTM_ADD(a, b) __fadd_rn((a), (b))
#define PTM_SUB(a, b) __fsub_rn((a), (b))
#define PTM_DIV(a, b) __fdiv_rn((a), (b))
#define PTM_SQRT(a) __fsqrt_rn(a)
#else
#define PTM_MUL(a, b) ((a) * (b))
#define PTM_ADD(a, b) ((a) + (b))
#define PTM_SUB(a, b) ((a) - (b))
#define PTM_DIV(a, b) ((a) / (b))
#define PTM_SQRT(a) sqrtf(a)
#endif
/* Horner step: acc*x + c */
#define PTM_H(acc, x, c) PTM_ADD(PTM_MUL((acc), (x)), (c))

PTM_FN int ptm_float_as_int(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_int(f);
#else
    int i; memcpy(&i, &f, 4); return i;
#endif
}
PTM_FN float ptm_int_as_float(int i) {
#if defined(__CUDA_ARCH__)
    return __int_as_float(i);
#else
    float f; memcpy(&f, &i, 4); return f;
#endif
}

/* ---- octant reduction shared by sin and cos: x >= 0, returns r in [-pi/4, pi/4] and octant j ---- */
PTM_FN float ptm_reduce_pio4(float x, int* jout) {
    const float FOPI = 1.27323954473516f;            /* 4/pi */
    const float DP1 = 0.78515625f;                   /* pi/4 split in three parts; j*DP1, j*DP2 exact */
    const float DP2 = 2.4187564849853515625e-4f;
    const float DP3 = 3.77489497744594108e-8f;
    int j = (int)PTM_MUL(FOPI, x);                   /* truncation; |x| < 2^20 assumed */
    if (j & 1) j += 1;
    float y = (float)j;                              /* exact */
    *jout = j;
    return PTM_SUB(PTM_SUB(PTM_SUB(x, PTM_MUL(y, DP1)), PTM_MUL(y, DP2)), PTM_MUL(y, DP3));
}
PTM_FN float ptm_sin_poly(float r) {
    float z = PTM_MUL(r, r);
    float p = PTM_H(PTM_H(-1.9515295891E-4f, z, 8.3321608736E-3f), z, -1.6666654611E-1f);
    return PTM_ADD(PTM_MUL(PTM_MUL(p, z), r), r);
}
PTM_FN float ptm_cos_poly(float r) {
    float z = PTM_MUL(r, r);
    float p = PTM_H(PTM_H(2.443315711809948E-005f, z, -1.388731625493765E-003f), z, 4.166664568298827E-002f);
    return PTM_ADD(PTM_SUB(PTM_MUL(PTM_MUL(p, z), z), PTM_MUL(0.5f, z)), 1.0f);
}

PTM_FN float pt_sinf(float x) {
    int neg = 0, j;
    if (x < 0.0f) { neg = 1; x = -x; }
    float r = ptm_reduce_pio4(x, &j);
    j &= 7;
    if (j > 3) { neg = !neg; j -= 4; }
    float y = (j == 1 || j == 2) ? ptm_cos_poly(r) : ptm_sin_poly(r);
    return neg ? -y : y;
}

PTM_FN float pt_cosf(float x) {
    int neg = 0, j;
    if (x < 0.0f) x = -x;
    float r = ptm_reduce_pio4(x, &j);
    j &= 7;
    if (j > 3) { neg = !neg; j -= 4; }
    if (j > 1) neg = !neg;
    float y = (j == 1 || j == 2) ? ptm_sin_poly(r) : ptm_cos_poly(r);
    return neg ? -y : y;
}

/* atan on [0, inf) by two-point range reduction, then odd minimax polynomial */
PTM_FN float ptm_atan_pos(float x) {
    float y;
    if (x > 2.414213562373095f) { y = 1.5707963267948966f; x = -PTM_DIV(1.0f, x); }
    else if (x > 0.4142135623730950f) { y = 0.7853981633974483f; x = PTM_DIV(PTM_SUB(x, 1.0f), PTM_ADD(x, 1.0f)); }
    else y = 0.0f;
    float z = PTM_MUL(x, x);
    float p = PTM_H(PTM_H(PTM_H(8.05374449538e-2f, z, -1.38776856032E-1f), z, 1.99777106478E-1f), z, -3.33329491539E-1f);
    return PTM_ADD(y, PTM_ADD(PTM_MUL(PTM_MUL(p, z), x), x));
}

/* GLSL atan(y, x): angle of (x, y) in (-pi, pi] */
PTM_FN float pt_atan2f(float y, float x) {
    const float PI_F = 3.14159265358979323846f;
    const float PIO2_F = 1.5707963267948966f;
    if (x == 0.0f) {
        if (y > 0.0f) return PIO2_F;
        if (y < 0.0f) return -PIO2_F;
        return 0.0f;
    }
    if (y == 0.0f) return (x > 0.0f) ? 0.0f : PI_F;
    float q = PTM_DIV(y, x);
    float a = (q < 0.0f) ? -ptm_atan_pos(-q) : ptm_atan_pos(q);
    if (x < 0.0f) a = (y < 0.0f) ? PTM_SUB(a, PI_F) : PTM_ADD(a, PI_F);
    return a;
}

PTM_FN float pt_asinf(float x) {
    int neg = 0, flag = 0;
    float a = x, z;
    if (a < 0.0f) { neg = 1; a = -a; }
    if (a > 1.0f) a = 1.0f;                          /* shaders pass normalised components; clamp rounding spill */
    if (a < 1.0e-4f) return x;
    if (a > 0.5f) { z = PTM_MUL(0.5f, PTM_SUB(1.0f, a)); a = PTM_SQRT(z); flag = 1; }
    else z = PTM_MUL(a, a);
    float p = PTM_H(PTM_H(PTM_H(PTM_H(4.2163199048E-2f, z, 2.4181311049E-2f), z, 4.5470025998E-2f), z, 7.4953002686E-2f), z,
                    1.6666752422E-1f);
    z = PTM_ADD(PTM_MUL(PTM_MUL(p, z), a), a);
    if (flag) { z = PTM_ADD(z, z); z = PTM_SUB(1.5707963267948966f, z); }
    return neg ? -z : z;
}

/* natural log, x > 0 and normal (shaders call it on alpha^2 in [1e-6, 1e-2]) */
PTM_FN float pt_logf(float x) {
    int ix = ptm_float_as_int(x);
    int e = ((ix >> 23) & 0xff) - 126;               /* x = m * 2^e, m in [0.5, 1) */
    float m = ptm_int_as_float((ix & 0x007fffff) | 0x3f000000);
    if (m < 0.707106781186547524f) { e -= 1; m = PTM_SUB(PTM_ADD(m, m), 1.0f); }
    else m = PTM_SUB(m, 1.0f);
    float z = PTM_MUL(m, m);
    float p = PTM_H(PTM_H(PTM_H(PTM_H(7.0376836292E-2f, m, -1.1514610310E-1f), m, 1.1676998740E-1f), m, -1.2420140846E-1f), m,
                    1.4249322787E-1f);
    p = PTM_H(PTM_H(PTM_H(PTM_H(p, m, -1.6668057665E-1f), m, 2.0000714765E-1f), m, -2.4999993993E-1f), m, 3.3333331174E-1f);
    float y = PTM_MUL(PTM_MUL(p, m), z);
    float fe = (float)e;
    if (e) y = PTM_ADD(y, PTM_MUL(-2.12194440e-4f, fe));
    y = PTM_ADD(y, PTM_MUL(-0.5f, z));
    z = PTM_ADD(m, y);
    if (e) z = PTM_ADD(z, PTM_MUL(0.693359375f, fe));
    return z;
}

/* e^x for |x| < 80 */
PTM_FN float pt_expf(float x) {
    float n = floorf(PTM_ADD(PTM_MUL(1.44269504088896341f, x), 0.5f));
    x = PTM_SUB(x, PTM_MUL(n, 0.693359375f));
    x = PTM_SUB(x, PTM_MUL(n, -2.12194440e-4f));
    float z = PTM_MUL(x, x);
    float p = PTM_H(PTM_H(PTM_H(PTM_H(PTM_H(1.9875691500E-4f, x, 1.3981999507E-3f), x, 8.3334519073E-3f), x, 4.1665795894E-2f), x,
                          1.6666665459E-1f), x, 5.0000001201E-1f);
    z = PTM_ADD(PTM_ADD(PTM_MUL(p, z), x), 1.0f);
    int k = (int)n;                                  /* z in [~0.7, ~1.42]; scale by 2^k through the exponent field */
    return PTM_MUL(z, ptm_int_as_float((k + 127) << 23));
}

/* GLSL pow(x, y) for x > 0 (the shaders only call it as pow(alpha^2, 1 - xi)) */
PTM_FN float pt_powf(float x, float y) {
    return pt_expf(PTM_MUL(y, pt_logf(x)));
}

#endif /* PTMATH_H */
