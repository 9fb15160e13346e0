/*
 * header.h -- shared types of the drop-in host API.
 *
 * Same type names, field order and byte layouts as the reference's header.h (typedefs
 * header.h:24-33, system_settings :55-77, LOCAL_SIZE :37-38, helper prototypes :40-53), but
 * with no OpenCL and no OpenCV behind them: the cl_* PODs come from mvs_cl_types.h and the
 * images are plain 8-bit BGR rasters (struct Image below, the layout cv::imread returns).
 */
#ifndef HEADERS
#define HEADERS

#include "mvs_cl_types.h"

#include <algorithm>
#include <chrono>
#include <fstream>
#include <iostream>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string>
#include <vector>

using namespace std;

/* boundary data types; sizes in bytes as the device sees them */
typedef cl_float8 vec8f;   /* 32: superpixel record {id, cx, cy, L, a, b, count, disparity} */
typedef cl_float4 vec4f;   /* 16 */
typedef cl_float3 vec3f;   /* 16: Lab pixel, 4th lane is padding */
typedef cl_float2 vec2f;   /*  8: flatness pair */

typedef cl_int3 vec3i;     /* 16 */
typedef cl_int2 vec2i;     /*  8: image / map size */

typedef cl_uchar8 vec8u;   /*  8: ray extents {nw, w, sw, n, s, ne, e, se} */
typedef cl_uchar3 vec3u;   /*  4: input pixel {s0 = R, s1 = G, s2 = B, pad} */

enum CL_Error_TYPE { buff_alloc = 0, kernel_def = 1, kernel_exe = 2, read_mem = 3 };

#define LOCAL_SIZE 16
#define LOCAL_SIZE_UPDATE 16

/* 8-bit, 3-channel, BGR-interleaved raster: what the reference holds in a cv::Mat after
 * cv::imread (file_handler.cpp:47).  Named Mat so that loadImageIn/loadImageOut keep their
 * reference signatures. */
struct Mat {
    int rows = 0, cols = 0;
    std::vector<unsigned char> data;   /* rows * cols * 3, B G R */
    bool empty() const { return data.empty(); }
    unsigned char *ptr(int r, int c) { return data.data() + 3 * ((size_t)r * cols + c); }
    const unsigned char *ptr(int r, int c) const { return data.data() + 3 * ((size_t)r * cols + c); }
};

void printText(std::string txt);
void printError(cl_int err_no, std::string message);
void errorHandler(CL_Error_TYPE type, cl_int err, std::string msg);

vec2i makeVec2i(int x, int y);
vec3i makeVec3i(int x, int y, int z);

vec2f makeVec2f(float x, float y);
vec3f makeVec3f(float x, float y, float z);
vec8f makeVec8f(float s0, float s1, float s2, float s3, float s4, float s5, float s6, float s7);

float euDistance3D(vec3f d1, vec3f d2);
vec3f crossVec3f(vec3f a, vec3f b);
vec3f normalizeVec3f(vec3f a);

/* The settings block main() fills (clMVDE.cpp:14-36) and every stage reads.  Field order and types are
 * the reference's (header.h:55-77); the comments say where a value ends up on the device path. */
struct system_settings {
    /* ---- superpixels (clSLIC) ---- */
    int spixl_size;                  /* S: cell size in pixels; map is ceil(W/S) x ceil(H/S) (pipeline.cpp:18-19); S >= 6 */
    float slic_color_weight;         /* w in sqrt(dc*Nc + w*(ds*Ns)) (clcode.cl:425-435) */
    int array_width, array_height;   /* camera grid; view index = gy * array_width + gx */
    int no_iter;                     /* centre-update / association rounds after the first association */
    bool enforce_connectivity = false;   /* two passes of the 5x5 label vote (clSLIC.cpp:390-410) */
    bool edge_enable = false;            /* edge-aware centre placement (clSLIC.cpp:86, 186-233; semantics pinned in DESIGN.md section 3) */

    /* ---- disparity initialisation (clPhotoConsistency) ---- */
    int num_disp_levels;             /* written by pipeline::perform_depth_est, not read from main() */
    int neib_hor;                    /* neighbour views: |dx| <= neib_hor ...                         */
    int neib_ver;                    /* ... and |dy| <= neib_ver on the grid (pipeline.cpp:130-142)    */
    int min_disp;                    /* levels min_disp, min_disp + inc, ... <= max_disp as floats     */
    int max_disp;
    int inc;
    float bl_ratio;                  /* vertical / horizontal baseline: y shift = bl_ratio * d * dy (clcode.cl:1034) */

    /* ---- plane refinement (clDepthRefinement) ---- */
    int kernel_size;                 /* halved by the pipeline (pipeline.cpp:166); far-neighbour stride = kernel_size/2 / kernel_step * S */
    int kernel_step;                 /* K: far-neighbour rings of the smoothness cost, divided by (iter + 1) per iteration */
    float fuse;                      /* visibility threshold, halved on the way to the kernels (depth_refinement.cpp:339, 1328) */
    float gamma;                     /* colour-similarity bandwidth: kernels see 1 / (2 gamma^2) (pipeline.cpp:164, depth_refinement.cpp:335) */
    float alpha;                     /* disparity-difference bandwidth: kernels see 1 / (2 alpha^2) */
    int no_prop;                     /* propagate iterations (state ping-pong A -> B -> A ...) */
};

#endif // !HEADERS
