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

typedef cl_float8 vec8f;
typedef cl_float4 vec4f;
typedef cl_float3 vec3f;
typedef cl_float2 vec2f;

typedef cl_int3 vec3i;
typedef cl_int2 vec2i;

typedef cl_uchar8 vec8u;
typedef cl_uchar3 vec3u;

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

struct system_settings {
    int spixl_size;
    float slic_color_weight;
    int array_width, array_height;
    int no_iter;
    bool enforce_connectivity = false;
    bool edge_enable = false;

    int num_disp_levels;
    int neib_hor;
    int neib_ver;
    int min_disp;
    int max_disp;
    int inc;
    float bl_ratio;

    int kernel_size;
    int kernel_step;
    float fuse;
    float gamma;
    float alpha;
    int no_prop;
};

#endif // !HEADERS
