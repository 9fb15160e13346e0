/*
 * photo_consistency.h -- block-matching disparity initialisation, drop-in for the reference
 * class (photo_consistency.h:5-54).  The device path runs the sm_100a kernels behind
 * mvs_bm_run (include/mvs_capi.h).
 */
#pragma once
#include "header.h"

class clPhotoConsistency
{
public:
    /* photo_consistency.h:8 */
    clPhotoConsistency(cl::Program program, int view_count, int spixl_size, int num_disp_levels, vec2i img_size_,
                       vec2i map_size_);
    ~clPhotoConsistency();

    /* photo_consistency.h:10-19, .cpp:21-150: fills spixel_rep and spixel_map[].s7 for all views */
    void do_initial_depth_estimation(vec8f *spixel_map, vec8u *spixel_rep, vec3f *cvg_img, cl_uint *idx_img,
                                     int array_width, float bl_ratio, vector<vector<int> > &view_subset,
                                     vector<float> &disp_levels);

    /* photo_consistency.h:23-34.  The reference's host twin is a CPU debugging aid that has
     * diverged from its kernels; this library has no CPU path, so
     * the entry point runs the SAME device kernels for views [start_view, end_view) with the
     * flattened subset tables the caller passes. */
    void do_initial_depth_estimation_host(vec8f *spixel_map, vec8u *spixel_rep, vec3f *in_img, cl_uint *idx_img,
                                          int array_width, float bl_ratio, vector<float> &disp_levels,
                                          int *view_subset_mat, int *subset_num_per_cam, int start_view,
                                          int end_view);

private:
    cl::Program program;
    int spixel_size;
    int num_disp_levels;
    int view_count;
    vec2i map_size;
    vec2i img_size;
};
