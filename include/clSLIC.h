/*
 * clSLIC.h -- superpixel segmentation stage, drop-in for the reference class of the same
 * name (clSLIC.h:10-57).  Public signatures are the reference's; the body runs the sm_100a
 * kernels behind mvs_slic_run (include/mvs_capi.h) instead of OpenCL.
 */
#pragma once
#include "header.h"

class clSLIC
{
public:
    /* clSLIC.h:15.  map_size must be ceil(img_size / spixl_size) (pipeline.cpp:18-19). */
    clSLIC(cl::Program program, system_settings *settings_, cl_int2 img_size, cl_int2 map_size);
    ~clSLIC();

    /* clSLIC.h:19, clSLIC.cpp:67-122: fills cvt_img (Lab), spixl_map (centres) and idx_img
     * (labels) for one image; pointers are caller-owned host memory. */
    void do_super_pixel_seg(vec3u *in_img, vec3f *cvt_img, vec8f *spixl_map, cl_uint *idx_img);
    /* clSLIC.h:20, clSLIC.cpp:447-478: paints superpixel borders red, from the labels of the
     * last do_super_pixel_seg call */
    void draw_segmentation_lines(cl_uchar3 *in_img, cl_uchar3 *out_img);

    /* additive: `count` images of the same size in one call (arrays are count x image) */
    void do_super_pixel_seg_batch(int count, vec3u *in_img, vec3f *cvt_img, vec8f *spixl_map, cl_uint *idx_img);

private:
    cl::Program program;
    system_settings *settings;
    vec2i img_size;
    vec2i map_size;
    cl_uint *last_idx;   /* borrowed: labels written by the last call */
};
