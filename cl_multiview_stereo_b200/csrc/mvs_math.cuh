// mvs_math.cuh -- device math of the B200 depth pipeline.
//
// The reference takes powr / exp / sqrt / distance / normalize / round and the
// float->int conversion from an unspecified OpenCL driver (clcode.cl:45-50,
// 1130-1131, 1158-1162, 1308-1309, 1702); their bits are driver-defined.  This
// library pins them: every function below is built from IEEE-754 correctly
// rounded primitives only (+ - * / sqrt fma) in a fixed order, so that results
// do not depend on compiler contraction (the translation units are compiled
// with -fmad=false; every fused operation is an explicit fmaf/fma).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mvs {

__device__ __forceinline__ float pow2i(int k) { return __int_as_float((k + 127) << 23); }
__device__ __forceinline__ double pow2i_d(int k) { return __longlong_as_double((long long)(k + 1023) << 52); }

// float -> int, truncation toward zero, saturating, NaN -> 0 (cvt.rzi.s32.f32)
__device__ __forceinline__ int f2i(float x) { return __float2int_rz(x); }
__device__ __forceinline__ int d2i(double x) { return __double2int_rz(x); }

// exp(x), binary32: Cody-Waite reduction by ln2 (hi/lo), degree-5 polynomial in
// Horner form, scaling by two exact powers of two.
// NONPOS = the caller guarantees x <= 0 or NaN (x = -(d*d)*w with w >= 0): the overflow select can never
// fire and a NaN argument already leaves y = NaN (every step propagates it), so both selects are dropped --
// same value for every admissible argument, four instructions fewer.
template <bool NONPOS = false>
__device__ __forceinline__ float expf_pinned(float x)
{
    // branch-free: the polynomial is always evaluated and the three special ranges are selected
    // afterwards (same values as returning early; lets the compiler interleave neighbouring calls)
    float t = fmaf(x, 1.44269504f, 12582912.0f);
    float kf = t - 12582912.0f;
    int k = f2i(kf);
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
    y = y * pow2i(k1);
    y = y * pow2i(k2);
    y = x < -104.0f ? 0.0f : y;
    if (!NONPOS) {
        y = x > 88.72284f ? __int_as_float(0x7f800000) : y;
        y = x != x ? x : y;
    }
    return y;
}

// expf_pinned restricted to |x| < 0.3465 (|x * log2(e)| < 1/2): there the reduction of expf_pinned gives k = 0 and
// r = x exactly, both scalings multiply by 1, and none of the range selects fires -- what is left is the polynomial.
// Same bits as expf_pinned(x) on that interval (tests/test_gpu_parity.py::test_small_exp...), 8 instructions
// instead of 22.  The caller guarantees the interval (NaN passes through: every step propagates it).
__device__ __forceinline__ float expf_pinned_small(float x)
{
    float p = 1.9875691500e-4f;
    p = fmaf(p, x, 1.3981999507e-3f);
    p = fmaf(p, x, 8.3334519073e-3f);
    p = fmaf(p, x, 4.1665795894e-2f);
    p = fmaf(p, x, 1.6666665459e-1f);
    p = fmaf(p, x, 5.0000001201e-1f);
    const float z = x * x;
    float y = fmaf(p, z, x);
    y = y + 1.0f;
    return y;
}

// exp(x), binary64 (only the one double-typed expression of the path,
// clcode.cl:1131, needs it)
__device__ __forceinline__ double exp_pinned(double x)
{
    if (x != x) return x;
    if (x > 709.78) return __longlong_as_double(0x7ff0000000000000LL);
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
    y = y * pow2i_d(k1);
    y = y * pow2i_d(k2);
    return y;
}

// powr(x, 1/3) for x > 0, pinned as the cube root without a division: integer seed of the inverse cube root,
// two Newton steps on it (multiplications + one fma), y = x z^2, one correction with the fma residual x - y^3
// scaled by z^2 / 3.  <= 1 ulp from the correctly rounded root, exact on exact cubes (about 19 instructions;
// the first definition spent three IEEE divisions here and kept k_cvt issue-bound).
__device__ __forceinline__ float cbrtf_pos(float x)
{
    const uint32_t i = 0x54a23400u - __float_as_uint(x) / 3u;
    float z = __uint_as_float(i);
    const float x3 = x * 0.33333334f;
#pragma unroll
    for (int it = 0; it < 2; ++it) {
        const float z3 = z * z * z;
        const float t = fmaf(-x3, z3, 1.3333334f);
        z = z * t;
    }
    const float zz = z * z;
    float y = x * zz;
    const float y2 = y * y;
    const float e = fmaf(-y2, y, x);
    const float c = zz * 0.33333334f;
    y = fmaf(e, c, y);
    return y;
}

// distance(a,b): sqrt of the left-to-right sum of squared differences
__device__ __forceinline__ float dist3(float ax, float ay, float az, float bx, float by, float bz)
{
    float dx = ax - bx, dy = ay - by, dz = az - bz;
    return sqrtf(dx * dx + dy * dy + dz * dz);
}

// round(): half away from zero
__device__ __forceinline__ float round_haz(float x) { return roundf(x); }

} // namespace mvs
