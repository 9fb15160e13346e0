/*
 * pipeline.h -- orchestrator, drop-in for the reference class (pipeline.h:6-36).
 * exe_pipeline() runs the whole per-view depth path (SLIC -> block matching -> refinement ->
 * fusion); the reference ships with perform_depth_est() commented out (pipeline.cpp:62-63),
 * here it is enabled.
 */
#pragma once
#include "header.h"
#include "clSLIC.h"
#include "photo_consistency.h"
#include "depth_refinement.h"

class pipeline
{
public:
    pipeline(std::string file_addr, system_settings *settings);   /* pipeline.cpp:7-28 */
    ~pipeline();
    void exe_pipeline();                                          /* pipeline.cpp:60-64 */

    // Variables (pipeline.h:13-17)
    cl::Program program;
    vec2i img_size, map_size;
    system_settings *settings;
    int view_num;

    /* additive: results of the last exe_pipeline (host arrays owned by the pipeline) */
    const vec3f *lab() const { return cvt_img; }
    const cl_uint *labels() const { return idx_img; }
    const vec8f *superpixels() const { return spixl_map; }
    const vec8u *extents() const { return spixl_rep; }
    const float *disparity() const { return disp_img.data(); }
    const float *state_a() const { return state_a_.data(); }
    /* additive: false = run the three stage classes one after another through host memory
     * exactly like the reference; true (default) = one device-resident frame (mvs_frame_run) */
    bool device_resident = true;
    /* additive: also run the multi-view fusion the reference keeps disabled
     * (depth_refinement.cpp:1382-1451); result in disparity_fused() */
    bool multi_view_fusion = false;
    const float *disparity_fused() const { return disp_fused_.data(); }
    /* per-stage device milliseconds of the last device-resident run (mvs_frame_stage_ms order) */
    float stage_ms[8];

private:
    void perform_segmentation();
    void perform_depth_est();
    void init_depth_map(std::vector<std::vector<int> > &view_subset, std::vector<float> &disp_levels,
                        int num_disp_levels);
    void refine_depth_map(std::vector<std::vector<int> > &view_subset);
    void compile_device_code();
    void run_device_resident();

    vector<Mat> img_array;
    vec3f *cvt_img;
    cl_uint *idx_img;
    vec8f *spixl_map;
    vec8u *spixl_rep;
    std::vector<float> disp_img, state_a_, disp_fused_;
};
