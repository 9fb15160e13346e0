// clMVDE.cpp -- entry point in the shape of the reference's (clMVDE.cpp:12-43): fill
// system_settings with the shipped literals, build a pipeline from an image list, run it.
// Additions: key=value overrides, and raw dumps of every stage output for parity checks.
//
//   clMVDE <list.txt> [out_dir] [key=value ...]        keys = system_settings field names,
//                                                      staged=1 walks the stage classes through host memory
#include "header.h"
#include "pipeline.h"

#include <map>
#include <string.h>

static void dump(const std::string &path, const void *p, size_t bytes)
{
    std::ofstream o(path, std::ios::binary);
    o.write((const char *)p, (std::streamsize)bytes);
}

int main(int argc, char **argv)
{
    system_settings *s = new system_settings;
    s->spixl_size = 8;
    s->slic_color_weight = 0.6f;
    s->array_height = 3;
    s->array_width = 3;
    s->no_iter = 5;
    s->enforce_connectivity = false;
    s->edge_enable = false;
    s->num_disp_levels = 30;
    s->neib_hor = 1;
    s->neib_ver = 1;
    s->min_disp = 30;
    s->max_disp = 60;
    s->inc = 1;
    s->bl_ratio = 1.03590f;
    s->kernel_size = 1080;
    s->kernel_step = 13;
    s->fuse = 1;
    s->gamma = 2;
    s->alpha = 6;
    s->no_prop = 5;

    std::string list = argc > 1 ? argv[1] : "data.txt";
    std::string out_dir;
    bool staged = false, view_fusion = false;
    for (int i = 2; i < argc; ++i) {
        const char *eq = strchr(argv[i], '=');
        if (!eq) { out_dir = argv[i]; continue; }
        std::string k(argv[i], eq - argv[i]);
        double v = atof(eq + 1);
        if (k == "spixl_size") s->spixl_size = (int)v;
        else if (k == "slic_color_weight") s->slic_color_weight = (float)v;
        else if (k == "array_width") s->array_width = (int)v;
        else if (k == "array_height") s->array_height = (int)v;
        else if (k == "no_iter") s->no_iter = (int)v;
        else if (k == "enforce_connectivity") s->enforce_connectivity = v != 0;
        else if (k == "neib_hor") s->neib_hor = (int)v;
        else if (k == "neib_ver") s->neib_ver = (int)v;
        else if (k == "min_disp") s->min_disp = (int)v;
        else if (k == "max_disp") s->max_disp = (int)v;
        else if (k == "inc") s->inc = (int)v;
        else if (k == "bl_ratio") s->bl_ratio = (float)v;
        else if (k == "kernel_size") s->kernel_size = (int)v;
        else if (k == "kernel_step") s->kernel_step = (int)v;
        else if (k == "fuse") s->fuse = (float)v;
        else if (k == "gamma") s->gamma = (float)v;
        else if (k == "alpha") s->alpha = (float)v;
        else if (k == "no_prop") s->no_prop = (int)v;
        else if (k == "staged") staged = v != 0;
        else if (k == "multi_view_fusion") view_fusion = v != 0;
        else { std::cout << "unknown setting " << k << std::endl; return 2; }
    }

    try {
        pipeline *model = new pipeline(list, s);
        model->device_resident = !staged;
        model->multi_view_fusion = view_fusion;
        auto t0 = std::chrono::high_resolution_clock::now();
        model->exe_pipeline();
        auto t1 = std::chrono::high_resolution_clock::now();
        std::cout << "views " << model->view_num << "  image " << model->img_size.x << "x" << model->img_size.y
                  << "  map " << model->map_size.x << "x" << model->map_size.y << "  wall ms "
                  << std::chrono::duration_cast<std::chrono::microseconds>(t1 - t0).count() / 1e3 << std::endl;
        if (!out_dir.empty()) {
            size_t N = (size_t)model->img_size.x * model->img_size.y * model->view_num;
            size_t M = (size_t)model->map_size.x * model->map_size.y * model->view_num;
            dump(out_dir + "/lab.bin", model->lab(), N * sizeof(vec3f));
            dump(out_dir + "/idx.bin", model->labels(), N * sizeof(cl_uint));
            dump(out_dir + "/map.bin", model->superpixels(), M * sizeof(vec8f));
            dump(out_dir + "/rep.bin", model->extents(), M * sizeof(vec8u));
            dump(out_dir + "/state_a.bin", model->state_a(), M * 6 * sizeof(float));
            dump(out_dir + "/disp.bin", model->disparity(), N * sizeof(float));
            if (view_fusion) dump(out_dir + "/disp_fused.bin", model->disparity_fused(), N * sizeof(float));
        }
        delete model;
    } catch (const std::exception &e) {
        std::cout << "error: " << e.what() << std::endl;
        return 1;
    }
    delete s;
    return 0;
}
