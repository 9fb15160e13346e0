/*
 * depth_refinement.h -- per-superpixel plane refinement and fusion, drop-in for the
 * reference class (depth_refinement.h:3-81).  The device path runs the sm_100a kernels
 * behind mvs_refine_run (include/mvs_capi.h).
 */
#pragma once
#include "header.h"

class clDepthRefinement
{
public:
    /* depth_refinement.h:6-7, .cpp:5-71: borrows the caller's host arrays of all views */
    clDepthRefinement(cl::Program program, vec2i img_size, vec2i map_size, vec2i camera_array_size, vec3f *cvt_img,
                      vec8f *spixl_map, cl_uint *idx_img, vec8u *spixl_rep,
                      std::vector<std::vector<int> > view_subset_vec, int spixl_size, float bl_ratio);
    ~clDepthRefinement();
    /* depth_refinement.h:9, .cpp:91-118: flatness, init state, no_prop x propagate, fusion.
     * gamma/alpha/kernel_size are the values pipeline::refine_depth_map passes (2g^2, 2a^2,
     * kernel_size/2: pipeline.cpp:164-166). */
    void do_refinement(float gamma, float alpha, float fuse, int kernel_step, int kernel_size, int no_prop);

    /* additive accessors (the reference keeps these arrays private and only writes PNGs):
     * valid after do_refinement, owned by the object */
    const vec2f *flatness() const { return flatness_map.data(); }
    const float *state_a() const { return current_state.data(); }    /* [V][mh][mw][6] */
    const float *state_b() const { return current_state2.data(); }
    const float *disparity() const { return disp_img.data(); }       /* [V][H][W] */

    /* additive: the multi-view fusion the reference keeps commented out at the end of fusion()
     * (depth_refinement.cpp:1382-1451: project_to_reference_inv + remove_view_inconsistency per
     * reference view).  Call after do_refinement; fuse as given to do_refinement. */
    void multi_view_fusion(float fuse);
    const float *disparity_projected() const { return disp_proj.data(); }   /* [V][H][W] */
    const float *disparity_fused() const { return disp_fused.data(); }      /* [V][H][W] */

private:
    cl::Program program;
    vec2i img_size;
    vec2i map_size;
    int array_width, array_height;
    int view_count;
    int spixl_size;
    float bl_ratio;
    std::vector<int> subset_num;
    std::vector<int> view_subset;     /* [V][V], row stride V (depth_refinement.cpp:25-31) */
    std::vector<vec2f> flatness_map;
    std::vector<float> current_state, current_state2, disp_img, disp_proj, disp_fused;
    vec8f *spixl_map;
    cl_uint *idx_img;
    vec8u *spixl_rep;
};
