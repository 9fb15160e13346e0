/*
 * clshim.h -- TEST INFRASTRUCTURE.  A small OpenCL-C-on-C++ shim so that the reference's
 * own device source (/root/reference/clMVDE/clcode.cl) compiles with g++ and runs on the
 * host as the second oracle (oracle/_ref/libclref.so).  Nothing of the reference is copied
 * into this repository: the Makefile reads clcode.cl where it lies, rewrites OpenCL vector
 * literals "(float2)(a, b)" to constructor calls and includes the result.
 *
 * What the shim supplies: the vector types the kernels use (float2/3/4/8, int2/3, uchar3/8,
 * uint) with the .x/.sN members and the four swizzles that occur (.s12 .s345 .xy .xyz),
 * element-wise operators, work-item functions backed by thread-local ids, barrier() backed
 * by a per-work-group pthread barrier, and the math built-ins.
 *
 * Built-ins whose bits an OpenCL driver defines (exp, powr(x, 1/3), distance, normalize,
 * round, sqrt) are mapped to the pinned definitions of ../oracle_math.h -- the one and only
 * substitution; everything else is the reference's text executed as written (C++ evaluates
 * float expressions in float, literals keep their C types, no contraction: -ffp-contract=off).
 */
#ifndef CLSHIM_H
#define CLSHIM_H

#include <stdint.h>
#include <type_traits>
#include "../oracle_math.h"

namespace clref {

typedef unsigned int uint;
typedef unsigned char uchar;

/* ---- work-item state ------------------------------------------------------------------ */
struct WorkItem {
    int gid[3], lid[3], grp[3], gsz[3], lsz[3];
    void *group_barrier;
};
extern thread_local WorkItem wi;
inline int get_global_id(int d) { return wi.gid[d]; }
inline int get_local_id(int d) { return wi.lid[d]; }
inline int get_group_id(int d) { return wi.grp[d]; }
inline int get_global_size(int d) { return wi.gsz[d]; }
inline int get_local_size(int d) { return wi.lsz[d]; }
void barrier(int flags);
#define CLK_LOCAL_MEM_FENCE 1
#define CLK_GLOBAL_MEM_FENCE 2

/* ---- vector types ----------------------------------------------------------------------- */
struct float2;
struct float3;

template <int N, int A, int B> struct Swz2 {   /* two lanes of an N-float vector */
    float v[N];
    operator float2() const;
    Swz2 &operator=(const float2 &o);
};
template <int N, int A, int B, int C> struct Swz3 {
    float v[N];
    operator float3() const;
    Swz3 &operator=(const float3 &o);
};

struct float2 {
    union {
        struct { float x, y; };
        struct { float s0, s1; };
        float v[2];
    };
    float2() {}
    float2(float a, float b) : x(a), y(b) {}
    explicit float2(float a) : x(a), y(a) {}
};

struct float3 {   /* 16 bytes, 4th lane is padding like cl_float3 */
    union {
        struct { float x, y, z, w; };
        struct { float s0, s1, s2, s3; };
        float v[4];
        Swz2<4, 0, 1> xy;
        Swz3<4, 0, 1, 2> xyz;
    };
    float3() {}
    float3(float a, float b, float c) : x(a), y(b), z(c), w(0.0f) {}
    explicit float3(float a) : x(a), y(a), z(a), w(0.0f) {}
    float3 &operator=(float a) { x = a; y = a; z = a; return *this; }   /* dead kernel only (clcode.cl:194) */
};

struct float4 {
    union {
        struct { float x, y, z, w; };
        struct { float s0, s1, s2, s3; };
        float v[4];
        Swz2<4, 0, 1> xy;
        Swz3<4, 0, 1, 2> xyz;
    };
    float4() {}
    float4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
};

struct float8 {
    union {
        struct { float s0, s1, s2, s3, s4, s5, s6, s7; };
        float v[8];
        Swz2<8, 1, 2> s12;
        Swz3<8, 3, 4, 5> s345;
    };
    float8() {}
    float8(float a, float b, float c, float d, float e, float f, float g, float h)
        : s0(a), s1(b), s2(c), s3(d), s4(e), s5(f), s6(g), s7(h) {}
    /* (float8)(s0, float2, float3, s6, s7): the one mixed literal in the source (clcode.cl:666) */
    float8(float a, float2 b, float3 c, float g, float h) : s0(a), s1(b.x), s2(b.y), s3(c.x), s4(c.y), s5(c.z), s6(g), s7(h) {}
    float8(float a, float b, float c, float3 d, float g, float h) : s0(a), s1(b), s2(c), s3(d.x), s4(d.y), s5(d.z), s6(g), s7(h) {}
};

struct int2 {
    union {
        struct { int x, y; };
        struct { int s0, s1; };
    };
    int2() {}
    int2(int a, int b) : x(a), y(b) {}
};
struct int3 {
    union {
        struct { int x, y, z, w; };
        struct { int s0, s1, s2, s3; };
    };
    int3() {}
    int3(int a, int b, int c) : x(a), y(b), z(c), w(0) {}
};
struct uchar3 {
    union {
        struct { uchar x, y, z, w; };
        struct { uchar s0, s1, s2, s3; };
    };
};
struct uchar8 {
    union {
        struct { uchar s0, s1, s2, s3, s4, s5, s6, s7; };
        uchar v[8];
    };
    uchar8() {}
    uchar8(uchar a, uchar b, uchar c, uchar d, uchar e, uchar f, uchar g, uchar h)
        : s0(a), s1(b), s2(c), s3(d), s4(e), s5(f), s6(g), s7(h) {}
};

template <int N, int A, int B> Swz2<N, A, B>::operator float2() const { return float2(v[A], v[B]); }
template <int N, int A, int B> Swz2<N, A, B> &Swz2<N, A, B>::operator=(const float2 &o) { v[A] = o.x; v[B] = o.y; return *this; }
template <int N, int A, int B, int C> Swz3<N, A, B, C>::operator float3() const { return float3(v[A], v[B], v[C]); }
template <int N, int A, int B, int C> Swz3<N, A, B, C> &Swz3<N, A, B, C>::operator=(const float3 &o) { v[A] = o.x; v[B] = o.y; v[C] = o.z; return *this; }

/* ---- element-wise operators (only the forms the kernels use) --------------------------- */
inline float2 operator+(float2 a, float2 b) { return float2(a.x + b.x, a.y + b.y); }
inline float2 operator-(float2 a, float2 b) { return float2(a.x - b.x, a.y - b.y); }
inline float2 operator*(float2 a, float2 b) { return float2(a.x * b.x, a.y * b.y); }
inline float2 operator*(float2 a, float s) { return float2(a.x * s, a.y * s); }
inline float2 operator*(float s, float2 a) { return float2(s * a.x, s * a.y); }
inline float2 operator/(float2 a, float s) { return float2(a.x / s, a.y / s); }
inline float2 &operator+=(float2 &a, float2 b) { a.x += b.x; a.y += b.y; return a; }
inline float2 &operator/=(float2 &a, float s) { a.x /= s; a.y /= s; return a; }

inline float3 operator+(float3 a, float3 b) { return float3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline float3 operator-(float3 a, float3 b) { return float3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline float3 operator*(float3 a, float3 b) { return float3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline float3 operator*(float3 a, float s) { return float3(a.x * s, a.y * s, a.z * s); }
inline float3 operator*(float s, float3 a) { return float3(s * a.x, s * a.y, s * a.z); }
inline float3 operator/(float3 a, float s) { return float3(a.x / s, a.y / s, a.z / s); }
inline float3 &operator+=(float3 &a, float3 b) { a.x += b.x; a.y += b.y; a.z += b.z; return a; }
inline float3 &operator/=(float3 &a, float s) { a.x /= s; a.y /= s; a.z /= s; return a; }

inline float4 operator-(float4 a, float4 b) { return float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
inline float4 operator+(float4 a, float4 b) { return float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }

inline int2 operator+(int2 a, int2 b) { return int2(a.x + b.x, a.y + b.y); }
inline int2 operator-(int2 a, int2 b) { return int2(a.x - b.x, a.y - b.y); }
inline int2 operator*(int2 a, int s) { return int2(a.x * s, a.y * s); }
inline int2 operator*(int s, int2 a) { return int2(s * a.x, s * a.y); }
/* forms that only the reference's dead kernels use (they must still compile) */
inline int2 operator-(int2 a, int s) { return int2(a.x - s, a.y - s); }
inline int2 operator+(int2 a, int s) { return int2(a.x + s, a.y + s); }
inline int2 max(int2 a, int2 b) { return int2(a.x < b.x ? b.x : a.x, a.y < b.y ? b.y : a.y); }
inline int2 min(int2 a, int2 b) { return int2(b.x < a.x ? b.x : a.x, b.y < a.y ? b.y : a.y); }

/* ---- built-ins ---------------------------------------------------------------------------- */
/* pinned (driver-defined in OpenCL): see ../oracle_math.h */
inline float exp(float x) { return o_expf(x); }
inline double exp(double x) { return o_exp(x); }
inline float powr(float x, float y)
{
    /* the live source only calls powr(t, 1.0f / 3.0f) with t > 0.008856 (clcode.cl:45-49) */
    if (y == 1.0f / 3.0f && x > 0.0f) return o_cbrtf_pos(x);
    return ::powf(x, y);
}
inline double pow(double x, double y) { return ::pow(x, y); }
inline float round(float x) { return o_roundf(x); }
inline double round(double x) { return ::round(x); }
inline float sqrt(float x) { return ::sqrtf(x); }
inline double sqrt(double x) { return ::sqrt(x); }
inline float fabs(float x) { return ::fabsf(x); }
inline double fabs(double x) { return ::fabs(x); }
inline float sign(float x) { return x > 0.0f ? 1.0f : (x < 0.0f ? -1.0f : 0.0f); }

template <class A, class B> inline typename std::common_type<A, B>::type fmax(A a, B b)
{
    typedef typename std::common_type<A, B>::type T;
    T x = (T)a, y = (T)b;
    return x < y ? y : x;   /* no NaN inputs on the live path */
}
template <class A, class B> inline typename std::common_type<A, B>::type max(A a, B b)
{
    typedef typename std::common_type<A, B>::type T;
    T x = (T)a, y = (T)b;
    return x < y ? y : x;
}
template <class A, class B> inline typename std::common_type<A, B>::type min(A a, B b)
{
    typedef typename std::common_type<A, B>::type T;
    T x = (T)a, y = (T)b;
    return y < x ? y : x;
}

inline float dot(float2 a, float2 b) { return a.x * b.x + a.y * b.y; }
inline float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float dot(float4 a, float4 b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
/* distance / normalize: pinned as sqrt of the left-to-right sum of squares (oracle_math.h) */
inline float distance(float3 a, float3 b) { return o_dist3(a.x, a.y, a.z, b.x, b.y, b.z); }
inline float distance(float2 a, float2 b) { float dx = a.x - b.x, dy = a.y - b.y; return ::sqrtf(dx * dx + dy * dy); }
inline float4 normalize(float4 a)
{
    float len = ::sqrtf(a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w);
    if (len == 0.0f) return float4(0.0f, 0.0f, 0.0f, 0.0f);
    return float4(a.x / len, a.y / len, a.z / len, a.w / len);
}

} /* namespace clref */

/* address-space and function qualifiers of OpenCL C */
#define kernel
#define global
#define local
#define constant const

#endif
