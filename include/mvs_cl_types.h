/*
 * mvs_cl_types.h -- layout-identical stand-ins for the OpenCL host types the reference's
 * public headers use (cl_platform.h vector types behind header.h:24-33, and cl::Program in
 * the stage constructors, clSLIC.h:15, photo_consistency.h:8, depth_refinement.h:6,
 * pipeline.h:14).  No OpenCL is involved: cl::Program is a shared handle to a CUDA context
 * of the C ABI (include/mvs_capi.h).
 *
 * Sizes / alignments follow the Khronos definitions (3-component types are the 4-component
 * ones): cl_uchar3 4/4, cl_uchar8 8/8, cl_int2 8/8, cl_int3 16/16, cl_float2 8/8,
 * cl_float3 = cl_float4 16/16, cl_float8 32/32.
 */
#ifndef MVS_CL_TYPES_H
#define MVS_CL_TYPES_H

#include <stdint.h>
#include <memory>
#include <string>

#include "mvs_capi.h"

typedef int32_t cl_int;
typedef uint32_t cl_uint;
typedef uint8_t cl_uchar;
typedef float cl_float;
#define CL_SUCCESS 0

#if defined(__GNUC__)
#define MVS_ALIGNED(n) __attribute__((aligned(n)))
#else
#define MVS_ALIGNED(n) __declspec(align(n))
#endif

#if defined(__GNUC__)
#pragma GCC diagnostic push
#pragma GCC diagnostic ignored "-Wpedantic"
#endif

typedef union {
    cl_uchar MVS_ALIGNED(4) s[4];
    struct { cl_uchar x, y, z, w; };
    struct { cl_uchar s0, s1, s2, s3; };
} cl_uchar4;
typedef cl_uchar4 cl_uchar3;

typedef union {
    cl_uchar MVS_ALIGNED(8) s[8];
    struct { cl_uchar x, y, z, w; };
    struct { cl_uchar s0, s1, s2, s3, s4, s5, s6, s7; };
} cl_uchar8;

typedef union {
    cl_int MVS_ALIGNED(8) s[2];
    struct { cl_int x, y; };
    struct { cl_int s0, s1; };
} cl_int2;

typedef union {
    cl_int MVS_ALIGNED(16) s[4];
    struct { cl_int x, y, z, w; };
    struct { cl_int s0, s1, s2, s3; };
} cl_int4;
typedef cl_int4 cl_int3;

typedef union {
    cl_float MVS_ALIGNED(8) s[2];
    struct { cl_float x, y; };
    struct { cl_float s0, s1; };
} cl_float2;

typedef union {
    cl_float MVS_ALIGNED(16) s[4];
    struct { cl_float x, y, z, w; };
    struct { cl_float s0, s1, s2, s3; };
} cl_float4;
typedef cl_float4 cl_float3;

typedef union {
    cl_float MVS_ALIGNED(32) s[8];
    struct { cl_float x, y, z, w; };
    struct { cl_float s0, s1, s2, s3, s4, s5, s6, s7; };
} cl_float8;

#if defined(__GNUC__)
#pragma GCC diagnostic pop
#endif

static_assert(sizeof(cl_uchar3) == 4 && alignof(cl_uchar3) == 4, "vec3u layout");
static_assert(sizeof(cl_uchar8) == 8 && alignof(cl_uchar8) == 8, "vec8u layout");
static_assert(sizeof(cl_int2) == 8 && alignof(cl_int2) == 8, "vec2i layout");
static_assert(sizeof(cl_int3) == 16 && alignof(cl_int3) == 16, "vec3i layout");
static_assert(sizeof(cl_float2) == 8 && alignof(cl_float2) == 8, "vec2f layout");
static_assert(sizeof(cl_float3) == 16 && alignof(cl_float3) == 16, "vec3f layout");
static_assert(sizeof(cl_float8) == 32 && alignof(cl_float8) == 32, "vec8f layout");

namespace cl {

/* Stand-in for cl::Program: what the reference builds from clcode.cl at start-up
 * (pipeline.cpp:31-56) is here a CUDA context of the C ABI -- the kernels are compiled
 * ahead of time into libmvs_b200.so.  Copies share the context, like cl::Program copies
 * share the underlying cl_program. */
class Program {
public:
    Program() {}
    /* binds CUDA device `device`; throws std::runtime_error when there is none (no CPU fallback) */
    explicit Program(int device);
    mvs_ctx *ctx() const { return h_ ? h_->ctx : nullptr; }
    bool valid() const { return ctx() != nullptr; }

private:
    struct Holder {
        mvs_ctx *ctx = nullptr;
        ~Holder() { if (ctx) mvs_ctx_destroy(ctx); }
    };
    std::shared_ptr<Holder> h_;
};

} // namespace cl

#endif
