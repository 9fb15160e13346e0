// pipeline.cpp -- orchestrator of the drop-in host API (include/pipeline.h; reference
// pipeline.cpp:7-175).  Same construction (image list -> sizes -> device set-up -> host
// arrays) and the same stage order; perform_depth_est is enabled.  By default the whole
// frame stays resident on the device between stages (one mvs_frame) and the host arrays the
// reference hands from stage to stage are filled once at the end; device_resident = false
// walks the three stage classes through host memory exactly like the reference.
#include "pipeline.h"
#include "file_handler.h"

#include <stdexcept>
#include <string.h>

static void must(int rc, const char *where)
{
    if (rc == 0) return;
    std::string msg = std::string(where) + ": " + mvs_last_error();
    std::cout << msg << std::endl;
    throw std::runtime_error(msg);
}

pipeline::pipeline(std::string file_addr, system_settings *settings_)
    : settings(settings_), cvt_img(nullptr), idx_img(nullptr), spixl_map(nullptr), spixl_rep(nullptr)
{
    memset(stage_ms, 0, sizeof stage_ms);
    view_num = settings->array_height * settings->array_width;
    if (read_image_array(img_array, file_addr, view_num))
        throw std::runtime_error("pipeline: cannot read " + std::to_string(view_num) + " images listed in " + file_addr);
    img_size = makeVec2i(img_array[0].cols, img_array[0].rows);
    for (const Mat &m : img_array)
        if (m.cols != img_size.x || m.rows != img_size.y) throw std::runtime_error("pipeline: views differ in size");
    map_size.x = (int)ceil((float)img_size.x / (float)settings->spixl_size);   // pipeline.cpp:18-19
    map_size.y = (int)ceil((float)img_size.y / (float)settings->spixl_size);
    compile_device_code();
    size_t N = (size_t)img_size.x * img_size.y * view_num, M = (size_t)map_size.x * map_size.y * view_num;
    cvt_img = new vec3f[N];
    idx_img = new cl_uint[N];
    spixl_map = new vec8f[M];
    spixl_rep = new vec8u[M];
}

pipeline::~pipeline()
{
    delete[] cvt_img;
    delete[] idx_img;
    delete[] spixl_map;
    delete[] spixl_rep;
}

// the reference JIT-builds clcode.cl here (pipeline.cpp:31-56); the CUDA kernels are built
// ahead of time, so this only binds the device (MVS_DEVICE selects it, default 0)
void pipeline::compile_device_code()
{
    const char *dev = getenv("MVS_DEVICE");
    program = cl::Program(dev ? atoi(dev) : 0);
}

void pipeline::exe_pipeline()
{
    if (device_resident) { run_device_resident(); return; }
    perform_segmentation();
    perform_depth_est();
}

static mvs_params params_of(const system_settings &s, int W, int H)
{
    mvs_params p;
    p.width = W; p.height = H; p.spixl_size = s.spixl_size;
    p.array_width = s.array_width; p.array_height = s.array_height;
    p.slic_color_weight = s.slic_color_weight; p.no_iter = s.no_iter;
    p.enforce_connectivity = s.enforce_connectivity ? 1 : 0;
    p.neib_hor = s.neib_hor; p.neib_ver = s.neib_ver;
    p.min_disp = s.min_disp; p.max_disp = s.max_disp; p.inc = s.inc; p.bl_ratio = s.bl_ratio;
    p.kernel_size = s.kernel_size; p.kernel_step = s.kernel_step;
    p.fuse = s.fuse; p.gamma = s.gamma; p.alpha = s.alpha; p.no_prop = s.no_prop;
    p.edge_enable = s.edge_enable ? 1 : 0;
    return p;
}

void pipeline::run_device_resident()
{
    size_t n1 = (size_t)img_size.x * img_size.y, N = n1 * view_num, M = (size_t)map_size.x * map_size.y * view_num;
    std::vector<vec3u> in_img(N);
    for (int i = 0; i < view_num; i++) loadImageIn(img_array[i], in_img.data() + i * n1, img_size.y, img_size.x);
    mvs_params p = params_of(*settings, img_size.x, img_size.y);
    mvs_frame *f = nullptr;
    must(mvs_frame_create(program.ctx(), &p, 0, view_num, &f), "pipeline: frame");
    disp_img.resize(N);
    state_a_.resize(6 * M);
    if (multi_view_fusion) disp_fused_.resize(N);
    int rc = mvs_frame_set_timing(f, 1) ||   // pipeline::stage_ms reports the split of this run
             mvs_frame_upload_async(f, MVS_BUF_RGBX, 0, view_num, in_img.data()) || mvs_frame_run(f) ||
             (multi_view_fusion && (mvs_frame_view_fusion(f) ||
                                    mvs_frame_download_async(f, MVS_BUF_DISP_FUSED, 0, view_num, disp_fused_.data()))) ||
             mvs_frame_download_async(f, MVS_BUF_LAB, 0, view_num, cvt_img) ||
             mvs_frame_download_async(f, MVS_BUF_IDX, 0, view_num, idx_img) ||
             mvs_frame_download_async(f, MVS_BUF_MAP, 0, view_num, spixl_map) ||
             mvs_frame_download_async(f, MVS_BUF_REP, 0, view_num, spixl_rep) ||
             mvs_frame_download_async(f, MVS_BUF_STATE_A, 0, view_num, state_a_.data()) ||
             mvs_frame_download_async(f, MVS_BUF_DISP, 0, view_num, disp_img.data()) ||
             mvs_frame_stage_ms(f, stage_ms);
    std::string err = rc ? mvs_last_error() : "";
    mvs_frame_destroy(f);
    if (rc) { std::cout << "pipeline: " << err << std::endl; throw std::runtime_error("pipeline: " + err); }
}

void pipeline::perform_segmentation()
{
    size_t n1 = (size_t)img_size.x * img_size.y, m1 = (size_t)map_size.x * map_size.y;
    std::vector<vec3u> in_img(n1 * view_num);
    clSLIC slic_obj(program, settings, img_size, map_size);
    for (int i = 0; i < view_num; i++) {
        loadImageIn(img_array[i], in_img.data() + i * n1, img_size.y, img_size.x);
        auto t0 = std::chrono::high_resolution_clock::now();
        slic_obj.do_super_pixel_seg(in_img.data() + i * n1, cvt_img + i * n1, spixl_map + i * m1, idx_img + i * n1);
        auto t1 = std::chrono::high_resolution_clock::now();
        if (getenv("MVS_LOG"))
            std::cout << "Time of SLIC = " << std::chrono::duration_cast<std::chrono::milliseconds>(t1 - t0).count() << std::endl;
    }
}

void pipeline::perform_depth_est()
{
    // disparity levels and neighbour lists: pipeline.cpp:121-142
    std::vector<float> disp_levels;
    for (int i = 0; i <= (settings->max_disp - settings->min_disp) / settings->inc; i++)
        disp_levels.push_back((float)(settings->min_disp + i * settings->inc));
    const int aw = settings->array_width, ah = settings->array_height;
    vector<vector<int> > view_subset(view_num);
    for (int i = 0; i < view_num; i++)
        for (int x = i % aw - settings->neib_hor; x <= i % aw + settings->neib_hor; x++)
            for (int y = i / aw - settings->neib_ver; y <= i / aw + settings->neib_ver; y++) {
                int indx = y * aw + x;
                if (x >= 0 && x < aw && y >= 0 && y < ah && indx != i) view_subset[i].push_back(indx);
            }
    init_depth_map(view_subset, disp_levels, (int)disp_levels.size());
    refine_depth_map(view_subset);
}

void pipeline::init_depth_map(std::vector<std::vector<int> > &view_subset, std::vector<float> &disp_levels,
                              int num_disp_levels)
{
    clPhotoConsistency consistency_obj(program, view_num, settings->spixl_size, num_disp_levels, img_size, map_size);
    consistency_obj.do_initial_depth_estimation(spixl_map, spixl_rep, cvt_img, idx_img, settings->array_width,
                                                settings->bl_ratio, view_subset, disp_levels);
}

void pipeline::refine_depth_map(std::vector<std::vector<int> > &view_subset)
{
    float gamma = 2 * pow(settings->gamma, 2);   // pipeline.cpp:164-166
    float alpha = 2 * pow(settings->alpha, 2);
    int kernel_size = settings->kernel_size / 2;
    clDepthRefinement refinement_obj(program, img_size, map_size, makeVec2i(settings->array_width, settings->array_height),
                                     cvt_img, spixl_map, idx_img, spixl_rep, view_subset, settings->spixl_size,
                                     settings->bl_ratio);
    refinement_obj.do_refinement(gamma, alpha, settings->fuse, settings->kernel_step, kernel_size, settings->no_prop);
    size_t N = (size_t)img_size.x * img_size.y * view_num, M = (size_t)map_size.x * map_size.y * view_num;
    disp_img.assign(refinement_obj.disparity(), refinement_obj.disparity() + N);
    state_a_.assign(refinement_obj.state_a(), refinement_obj.state_a() + 6 * M);
    if (multi_view_fusion) {
        refinement_obj.multi_view_fusion(settings->fuse);
        disp_fused_.assign(refinement_obj.disparity_fused(), refinement_obj.disparity_fused() + N);
    }
}
