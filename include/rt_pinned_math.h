/*
 * rt_pinned_math.h — pinned definitions of the WGSL builtins whose precision the WGSL
 * specification leaves to the implementation: sin, cos (compute_shader.wgsl:239-240,329) and
 * pow(x, 5.0) (compute_shader.wgsl:140).
 *
 * Why this exists.  The reference shader is a chaotic path tracer: a 1-ulp difference in a
 * scattered direction can flip a grazing hit several bounces later and change a whole pixel.
 * WGSL allows sin/cos an absolute error of 2^-11 and pow several ulp, so the reference is not
 * bit-reproducible between two GPUs either.  To make parity a bit-exact statement instead of a
 * statistical one, the CPU oracle and the CUDA kernels evaluate these three builtins with the
 * SAME sequence of IEEE-754 binary32 operations, written out below with one rounding per
 * operator (no FMA contraction: the oracle is built with -ffp-contract=off, the kernels with
 * -fmad=false; the kernels' FFMA use is confined to explicit fmaf() calls in the broad phase).
 * Every other operation of the shader (+ - * / sqrt, comparisons, integer hashing, conversions)
 * is exactly specified by IEEE-754 and needs no pinning.
 *
 * sin and cos are within 2 ulp (pow5: 3 ulp) of the correctly rounded result over the argument ranges
 * the shader produces (tests/test_oracle.py checks them against libm), far inside the WGSL
 * tolerance.  oracle/ can also be built against libm (-DRT_ORACLE_LIBM) to measure how far an
 * implementation with different builtin rounding lands from the pinned one.
 *
 * Polynomial coefficients: the classic single-precision minimax kernels on [-pi/4, pi/4]
 * (FreeBSD msun k_sinf.c / k_cosf.c coefficients, public domain), evaluated in binary32.
 */
#ifndef RT_PINNED_MATH_H_
#define RT_PINNED_MATH_H_

#if defined(__CUDACC__)
#define RT_HD __host__ __device__ __forceinline__
#else
#define RT_HD static inline
#endif

/* sin and cos of a finite angle 0 <= a < 8 (the shader only forms a = u * 2pi with u in [0,1]).
 * Cody-Waite reduction by k * pi/2 with a 3-term split whose first two terms have <= 12
 * significant bits (k <= 5, so k*hi and k*mid are exact), then the minimax kernels. */
RT_HD void rt_sincos(float a, float* s_out, float* c_out) {
    const float two_over_pi = 0.636619772f;
    const float pio2_hi  = 1.5703125f;               /* 0x3FC90000: 1.1001001 b            */
    const float pio2_mid = 4.8375129699707031e-4f;   /* 0x39FDA000: next 11 bits of pi/2   */
    const float pio2_lo  = 7.549790126404332e-08f;   /* 0x33A22169: pi/2 - hi - mid        */
    const float S1 = -0.166666666416265235595f, S2 = 0.0083333293858894631756f;
    const float S3 = -0.000198393348360966317347f, S4 = 0.0000027183114939898219064f;
    const float C0 = -0.499999997251031003120f, C1 = 0.0416666233237390631894f;
    const float C2 = -0.00138867637746099294692f, C3 = 0.0000243904487962774090654f;

    int k = (int)(a * two_over_pi + 0.5f);           /* a >= 0: truncation == floor */
    float kf = (float)k;
    float r = a - kf * pio2_hi;
    r = r - kf * pio2_mid;
    r = r - kf * pio2_lo;
    float z = r * r;
    float sp = S3 + z * S4;
    sp = S2 + z * sp;
    sp = S1 + z * sp;
    float sn = r + (r * z) * sp;
    float cp = C2 + z * C3;
    cp = C1 + z * cp;
    cp = C0 + z * cp;
    float cs = 1.0f + z * cp;
    switch (k & 3) {
        case 0:  *s_out = sn;  *c_out = cs;  break;
        case 1:  *s_out = cs;  *c_out = -sn; break;
        case 2:  *s_out = -sn; *c_out = -cs; break;
        default: *s_out = -cs; *c_out = sn;  break;
    }
}

/* pow(x, 5.0) for the Schlick term, x = 1 - cos_theta in [0, 2]. */
RT_HD float rt_pow5(float x) {
    float x2 = x * x;
    float x4 = x2 * x2;
    return x4 * x;
}

#endif /* RT_PINNED_MATH_H_ */
