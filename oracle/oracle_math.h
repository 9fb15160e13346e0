/*
 * oracle_math.h -- TEST INFRASTRUCTURE (CPU oracle), not product code.
 *
 * Pinned definitions of the transcendental / vector built-ins that the
 * reference's OpenCL program takes from an unspecified NVIDIA OpenCL driver
 * (powr, exp, sqrt, distance, normalize, round, float->int conversion).
 * Their true bits are unknowable ("parity unpinned"), so the
 * oracle DEFINES them here from IEEE-754 correctly rounded primitives only
 * (+ - * / sqrt fma), evaluated strictly in the written order.  The CUDA
 * product carries its own, separately written copy of the same definitions
 * (cl_multiview_stereo_b200/csrc/mvs_math.cuh); tests compare the two
 * bit-for-bit and both against libm within a few ULP.
 *
 * Compile with -ffp-contract=off (no implicit contraction) and -mfma (so
 * fmaf/fma are single hardware instructions; they are correctly rounded
 * either way).
 */
#ifndef ORACLE_MATH_H
#define ORACLE_MATH_H

#include <math.h>
#include <stdint.h>
#include <string.h>

static inline float o_as_float(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint32_t o_as_uint(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline double o_as_double(uint64_t u) { double f; memcpy(&f, &u, 8); return f; }

/* 2^k for k in [-126,127] as float */
static inline float o_pow2i(int k) { return o_as_float((uint32_t)(k + 127) << 23); }
static inline double o_pow2i_d(int k) { return o_as_double((uint64_t)(k + 1023) << 52); }

/* float -> int conversion: truncation toward zero, saturating, NaN -> 0.
 * (OpenCL leaves out-of-range conversions implementation-defined; this is the
 * pinned behaviour, equal to PTX cvt.rzi.s32.f32.) */
static inline int o_f2i(float x)
{
    if (x != x) return 0;
    if (x >= 2147483648.0f) return 2147483647;
    if (x <= -2147483648.0f) return (-2147483647 - 1);
    return (int)x;
}
static inline int o_d2i(double x)
{
    if (x != x) return 0;
    if (x >= 2147483648.0) return 2147483647;
    if (x <= -2147483648.0) return (-2147483647 - 1);
    return (int)x;
}

/* exp(x), float.  Cody-Waite reduction by ln2 (hi/lo), degree-5 Cephes
 * polynomial in Horner form with fmaf, scaling in two exact power-of-two steps
 * (the second may round once into the denormal range). */
static inline float o_expf(float x)
{
    if (x != x) return x;
    if (x > 88.72284f) return INFINITY;
    if (x < -104.0f) return 0.0f;
    float t = fmaf(x, 1.44269504f, 12582912.0f);
    float kf = t - 12582912.0f;
    int k = (int)kf;
    float r = fmaf(kf, -0.693145751953125f, x);
    r = fmaf(kf, -1.428606765330187e-06f, r);
    float p = 1.9875691500e-4f;
    p = fmaf(p, r, 1.3981999507e-3f);
    p = fmaf(p, r, 8.3334519073e-3f);
    p = fmaf(p, r, 4.1665795894e-2f);
    p = fmaf(p, r, 1.6666665459e-1f);
    p = fmaf(p, r, 5.0000001201e-1f);
    float z = r * r;
    float y = fmaf(p, z, r);
    y = y + 1.0f;
    int k1 = k >> 1;
    int k2 = k - k1;
    y = y * o_pow2i(k1);
    y = y * o_pow2i(k2);
    return y;
}

/* exp(x), double (only used for the one double-typed expression of the
 * reference, clcode.cl:1131).  Same scheme, Taylor degree 13. */
static inline double o_exp(double x)
{
    if (x != x) return x;
    if (x > 709.78) return (double)INFINITY;
    if (x < -745.2) return 0.0;
    double t = fma(x, 1.4426950408889634, 6755399441055744.0);
    double kd = t - 6755399441055744.0;
    int k = (int)kd;
    double r = fma(kd, -6.93147180369123816490e-01, x);
    r = fma(kd, -1.90821492927058770002e-10, r);
    double p = 1.0 / 6227020800.0;
    p = fma(p, r, 1.0 / 479001600.0);
    p = fma(p, r, 1.0 / 39916800.0);
    p = fma(p, r, 1.0 / 3628800.0);
    p = fma(p, r, 1.0 / 362880.0);
    p = fma(p, r, 1.0 / 40320.0);
    p = fma(p, r, 1.0 / 5040.0);
    p = fma(p, r, 1.0 / 720.0);
    p = fma(p, r, 1.0 / 120.0);
    p = fma(p, r, 1.0 / 24.0);
    p = fma(p, r, 1.0 / 6.0);
    p = fma(p, r, 0.5);
    double z = r * r;
    double y = fma(p, z, r);
    y = y + 1.0;
    int k1 = k >> 1;
    int k2 = k - k1;
    y = y * o_pow2i_d(k1);
    y = y * o_pow2i_d(k2);
    return y;
}

/* powr(x, 1.0f/3.0f) for x > 0: pinned as the real cube root, DIVISION-FREE (round 2; the round-1 definition --
 * two Halley steps with IEEE divisions -- made the colour conversion issue-bound on the GPU).  Bit-hack seed of the
 * INVERSE cube root z ~ x^(-1/3) (3.4 % off at worst), two Newton steps z <- z * (4/3 - (x/3) z^3) (multiplications
 * and one fma each), y = x * z^2, and one correction with the fma residual x - y^3 and 1 / (3 y^2) ~ z^2 / 3.
 * At most 1 ulp from the correctly rounded cube root over the whole positive normal range (exhaustively checked on
 * [0.008856, 1.2], the interval rgb2lab feeds it), exact on exact cubes; tests/test_cpu_abi_math.py. */
static inline float o_cbrtf_pos(float x)
{
    uint32_t i = o_as_uint(x);
    i = 0x54a23400u - i / 3u;
    float z = o_as_float(i);
    const float x3 = x * 0.33333334f;
    for (int it = 0; it < 2; ++it) {
        float z3 = z * z * z;
        float t = fmaf(-x3, z3, 1.3333334f);
        z = z * t;
    }
    float zz = z * z;
    float y = x * zz;
    float y2 = y * y;
    float e = fmaf(-y2, y, x);
    float c = zz * 0.33333334f;
    y = fmaf(e, c, y);
    return y;
}

/* round(): half away from zero (OpenCL C / C99 semantics). */
static inline float o_roundf(float x) { return roundf(x); }

/* distance(a,b) = sqrt(dx*dx + dy*dy + dz*dz), summed left to right */
static inline float o_dist3(float ax, float ay, float az, float bx, float by, float bz)
{
    float dx = ax - bx, dy = ay - by, dz = az - bz;
    return sqrtf(dx * dx + dy * dy + dz * dz);
}

#endif
