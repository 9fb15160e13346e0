// stages.cpp -- the three stage classes of the drop-in host API (include/clSLIC.h,
// photo_consistency.h, depth_refinement.h) over the C ABI.  Host-pointer semantics are the
// reference's: the caller owns every array, results are written in place
// (clSLIC.cpp:110-116, photo_consistency.cpp:110,144).  A failed device call prints the
// library's message and throws -- the reference prints and carries on, which
// this library never does.
#include "clSLIC.h"
#include "depth_refinement.h"
#include "photo_consistency.h"

#include <stdexcept>
#include <string.h>

static void must(int rc, const char *where)
{
    if (rc == 0) return;
    std::string msg = std::string(where) + ": " + mvs_last_error();
    std::cout << msg << std::endl;
    throw std::runtime_error(msg);
}

static mvs_ctx *need_ctx(const cl::Program &p, const char *where)
{
    if (!p.valid())
        throw std::runtime_error(std::string(where) + ": cl::Program holds no CUDA context (construct it with a device index)");
    return p.ctx();
}

// ---- clSLIC --------------------------------------------------------------------------------
clSLIC::clSLIC(cl::Program program_, system_settings *settings_, cl_int2 img_size_, cl_int2 map_size_)
    : program(program_), settings(settings_), img_size(img_size_), map_size(map_size_), last_idx(nullptr)
{
    int S = settings->spixl_size;
    int mw = (int)ceilf((float)img_size.x / (float)S), mh = (int)ceilf((float)img_size.y / (float)S);
    if (mw != map_size.x || mh != map_size.y)
        throw std::runtime_error("clSLIC: map_size must be ceil(img_size / spixl_size) (pipeline.cpp:18-19)");
}

clSLIC::~clSLIC() {}

void clSLIC::do_super_pixel_seg_batch(int count, vec3u *in_img, vec3f *cvt_img, vec8f *spixl_map, cl_uint *idx_img)
{
    must(mvs_slic_run_edge(need_ctx(program, "clSLIC"), img_size.x, img_size.y, settings->spixl_size,
                           settings->slic_color_weight, settings->no_iter, settings->enforce_connectivity ? 1 : 0,
                           settings->edge_enable ? 1 : 0, count, in_img, cvt_img, spixl_map, idx_img, nullptr),
         "clSLIC::do_super_pixel_seg");
    last_idx = idx_img;
}

void clSLIC::do_super_pixel_seg(vec3u *in_img, vec3f *cvt_img, vec8f *spixl_map, cl_uint *idx_img)
{
    do_super_pixel_seg_batch(1, in_img, cvt_img, spixl_map, idx_img);
}

// overlay (clSLIC.cpp:447-478) on the device: interior pixels whose 4-neighbourhood holds another label become
// {0,0,255}, the others take in_img; the 1-pixel frame of out_img is left untouched, as in the reference
void clSLIC::draw_segmentation_lines(cl_uchar3 *in_img, cl_uchar3 *out_img)
{
    if (!last_idx) throw std::runtime_error("draw_segmentation_lines: no segmentation has been computed yet");
    must(mvs_draw_segmentation_run(need_ctx(program, "clSLIC"), img_size.x, img_size.y, 1, last_idx, in_img, out_img),
         "clSLIC::draw_segmentation_lines");
}

// ---- clPhotoConsistency ----------------------------------------------------------------------
clPhotoConsistency::clPhotoConsistency(cl::Program program_, int view_count_, int spixl_size, int num_disp_levels_,
                                       vec2i img_size_, vec2i map_size_)
    : program(program_), spixel_size(spixl_size), num_disp_levels(num_disp_levels_), view_count(view_count_),
      map_size(map_size_), img_size(img_size_)
{
}

clPhotoConsistency::~clPhotoConsistency() {}

void clPhotoConsistency::do_initial_depth_estimation(vec8f *spixel_map, vec8u *spixel_rep, vec3f *cvg_img,
                                                     cl_uint *idx_img, int array_width, float bl_ratio,
                                                     vector<vector<int> > &view_subset, vector<float> &disp_levels)
{
    // flatten with row stride V (photo_consistency.cpp:39-47)
    std::vector<int> mat((size_t)view_count * view_count, 0), num(view_count, 0);
    for (int i = 0; i < view_count && i < (int)view_subset.size(); i++) {
        num[i] = (int)view_subset[i].size();
        for (int j = 0; j < num[i]; j++) mat[(size_t)i * view_count + j] = view_subset[i][j];
    }
    must(mvs_bm_run(need_ctx(program, "clPhotoConsistency"), view_count, img_size.x, img_size.y, spixel_size,
                    (float *)spixel_map, (uint8_t *)spixel_rep, (const float *)cvg_img, idx_img, array_width, bl_ratio,
                    mat.data(), num.data(), disp_levels.data(), (int)disp_levels.size()),
         "clPhotoConsistency::do_initial_depth_estimation");
}

void clPhotoConsistency::do_initial_depth_estimation_host(vec8f *spixel_map, vec8u *spixel_rep, vec3f *in_img,
                                                          cl_uint *idx_img, int array_width, float bl_ratio,
                                                          vector<float> &disp_levels, int *view_subset_mat,
                                                          int *subset_num_per_cam, int start_view, int end_view)
{
    if (start_view < 0 || end_view > view_count || start_view > end_view)
        throw std::runtime_error("do_initial_depth_estimation_host: bad view range");
    size_t M = (size_t)map_size.x * map_size.y;
    std::vector<vec8f> tmp_map(spixel_map, spixel_map + M * view_count);
    std::vector<vec8u> tmp_rep(M * view_count);
    must(mvs_bm_run(need_ctx(program, "clPhotoConsistency"), view_count, img_size.x, img_size.y, spixel_size,
                    (float *)tmp_map.data(), (uint8_t *)tmp_rep.data(), (const float *)in_img, idx_img, array_width,
                    bl_ratio, view_subset_mat, subset_num_per_cam, disp_levels.data(), (int)disp_levels.size()),
         "clPhotoConsistency::do_initial_depth_estimation_host");
    // only the requested views are written back
    for (size_t i = M * start_view; i < M * end_view; ++i) { spixel_map[i].s7 = tmp_map[i].s7; spixel_rep[i] = tmp_rep[i]; }
}

// ---- clDepthRefinement -------------------------------------------------------------------------
clDepthRefinement::clDepthRefinement(cl::Program program_, vec2i img_size_, vec2i map_size_, vec2i camera_array_size,
                                     vec3f * /*cvt_img: not read by the refinement kernels*/, vec8f *spixl_map_,
                                     cl_uint *idx_img_, vec8u *spixl_rep_,
                                     std::vector<std::vector<int> > view_subset_vec, int spixl_size_, float bl_ratio_)
    : program(program_), img_size(img_size_), map_size(map_size_), array_width(camera_array_size.x),
      array_height(camera_array_size.y), view_count(camera_array_size.x * camera_array_size.y),
      spixl_size(spixl_size_), bl_ratio(bl_ratio_), spixl_map(spixl_map_), idx_img(idx_img_), spixl_rep(spixl_rep_)
{
    subset_num.assign(view_count, 0);
    view_subset.assign((size_t)view_count * view_count, 0);
    for (int i = 0; i < view_count && i < (int)view_subset_vec.size(); i++) {   // depth_refinement.cpp:25-31
        subset_num[i] = (int)view_subset_vec[i].size();
        for (int j = 0; j < subset_num[i]; j++) view_subset[(size_t)i * view_count + j] = view_subset_vec[i][j];
    }
}

clDepthRefinement::~clDepthRefinement() {}

void clDepthRefinement::do_refinement(float gamma, float alpha, float fuse, int kernel_step, int kernel_size, int no_prop)
{
    size_t M = (size_t)map_size.x * map_size.y * view_count, N = (size_t)img_size.x * img_size.y * view_count;
    flatness_map.resize(M);
    current_state.resize(6 * M);
    current_state2.resize(6 * M);
    disp_img.resize(N);
    must(mvs_refine_run(need_ctx(program, "clDepthRefinement"), img_size.x, img_size.y, array_width, array_height,
                        spixl_size, bl_ratio, (const float *)spixl_map, idx_img, (const uint8_t *)spixl_rep,
                        view_subset.data(), subset_num.data(), gamma, alpha, fuse, kernel_step, kernel_size, no_prop,
                        (float *)flatness_map.data(), current_state.data(), current_state2.data(), disp_img.data()),
         "clDepthRefinement::do_refinement");
}

void clDepthRefinement::multi_view_fusion(float fuse)
{
    size_t N = (size_t)img_size.x * img_size.y * view_count;
    if (disp_img.size() != N) {
        printText("ERROR: multi_view_fusion called before do_refinement");
        throw std::runtime_error("clDepthRefinement::multi_view_fusion: no disparity images");
    }
    disp_proj.resize(N);
    disp_fused.resize(N);
    must(mvs_view_fusion_run(need_ctx(program, "clDepthRefinement"), view_count, img_size.x, img_size.y, array_width,
                             bl_ratio, fuse, disp_img.data(), disp_proj.data(), disp_fused.data()),
         "clDepthRefinement::multi_view_fusion");
}
