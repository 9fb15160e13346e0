/* TEST INFRASTRUCTURE.  Inert stand-in for <CL/cl.hpp> so that the reference's host sources
 * (clMVDE/photo_consistency.cpp, depth_refinement.cpp, file_handler.cpp) compile UNMODIFIED with g++.
 * Only their CPU twins (*_host methods) are ever executed; every OpenCL call site compiles against
 * classes that do nothing (no buffer is allocated, no kernel runs, reads leave the host array as it is).
 * Host vector types follow the Khronos cl_platform.h layouts (3-component = 4-component). */
#ifndef MVS_SHIM_CL_HPP
#define MVS_SHIM_CL_HPP
#include <stddef.h>
#include <stdint.h>
#include <string>
#include <vector>

typedef int32_t cl_int;
typedef uint32_t cl_uint;
typedef uint8_t cl_uchar;
typedef float cl_float;
typedef cl_uint cl_bool;
typedef uint64_t cl_mem_flags;

#define CL_SUCCESS 0
#define CL_TRUE 1
#define CL_FALSE 0
#define CL_MEM_READ_WRITE (1 << 0)
#define CL_MEM_WRITE_ONLY (1 << 1)
#define CL_MEM_READ_ONLY (1 << 2)
#define CL_MEM_USE_HOST_PTR (1 << 3)
#define CL_MEM_ALLOC_HOST_PTR (1 << 4)
#define CL_MEM_COPY_HOST_PTR (1 << 5)
#define CL_MEM_HOST_WRITE_ONLY (1 << 7)
#define CL_MEM_HOST_READ_ONLY (1 << 8)
#define CL_PROGRAM_CONTEXT 0x1161
#define CL_CONTEXT_DEVICES 0x1081
#define CL_DEVICE_TYPE_GPU (1 << 2)

#define MVS_SHIM_VEC(NAME, T, N, ALIGN, FIELDS)            \
    typedef union {                                         \
        T __attribute__((aligned(ALIGN))) s[N];             \
        struct { FIELDS };                                  \
    } NAME
#pragma GCC diagnostic push
#pragma GCC diagnostic ignored "-Wpedantic"
typedef union { cl_uchar __attribute__((aligned(4))) s[4]; struct { cl_uchar x, y, z, w; }; struct { cl_uchar s0, s1, s2, s3; }; } cl_uchar4;
typedef cl_uchar4 cl_uchar3;
typedef union { cl_uchar __attribute__((aligned(8))) s[8]; struct { cl_uchar x, y, z, w; }; struct { cl_uchar s0, s1, s2, s3, s4, s5, s6, s7; }; } cl_uchar8;
typedef union { cl_int __attribute__((aligned(8))) s[2]; struct { cl_int x, y; }; struct { cl_int s0, s1; }; } cl_int2;
typedef union { cl_int __attribute__((aligned(16))) s[4]; struct { cl_int x, y, z, w; }; struct { cl_int s0, s1, s2, s3; }; } cl_int4;
typedef cl_int4 cl_int3;
typedef union { cl_float __attribute__((aligned(8))) s[2]; struct { cl_float x, y; }; struct { cl_float s0, s1; }; } cl_float2;
typedef union { cl_float __attribute__((aligned(16))) s[4]; struct { cl_float x, y, z, w; }; struct { cl_float s0, s1, s2, s3; }; } cl_float4;
typedef cl_float4 cl_float3;
typedef union { cl_float __attribute__((aligned(32))) s[8]; struct { cl_float x, y, z, w; }; struct { cl_float s0, s1, s2, s3, s4, s5, s6, s7; }; } cl_float8;
#pragma GCC diagnostic pop
#undef MVS_SHIM_VEC

namespace cl {
struct Device {};
struct Platform {};
struct Context {
    template <int I> std::vector<Device> getInfo() const { return std::vector<Device>(1); }
};
struct Program {
    template <int I> Context getInfo() const { return Context(); }
};
struct Buffer {
    Buffer() {}
    Buffer(const Context &, cl_mem_flags, size_t, void * = nullptr, cl_int *err = nullptr) { if (err) *err = CL_SUCCESS; }
};
struct NDRange {
    NDRange() {}
    NDRange(size_t) {}
    NDRange(size_t, size_t) {}
    NDRange(size_t, size_t, size_t) {}
};
static const NDRange NullRange;
struct Kernel {
    Kernel(const Program &, const char *, cl_int *err = nullptr) { if (err) *err = CL_SUCCESS; }
    template <class T> cl_int setArg(cl_uint, const T &) { return CL_SUCCESS; }
};
struct CommandQueue {
    CommandQueue(const Context &, const Device &, cl_mem_flags = 0, cl_int *err = nullptr) { if (err) *err = CL_SUCCESS; }
    cl_int enqueueNDRangeKernel(const Kernel &, const NDRange &, const NDRange &, const NDRange & = NullRange) { return CL_SUCCESS; }
    cl_int enqueueReadBuffer(const Buffer &, cl_bool, size_t, size_t, void *) { return CL_SUCCESS; }
    cl_int enqueueWriteBuffer(const Buffer &, cl_bool, size_t, size_t, const void *) { return CL_SUCCESS; }
    cl_int finish() { return CL_SUCCESS; }
};
} // namespace cl
#endif
