// clref.cpp -- TEST INFRASTRUCTURE.  Second oracle: the reference's OWN device source
// (/root/reference/clMVDE/clcode.cl), compiled as C++ over clshim.h and driven through an
// NDRange emulation that follows the reference's host launches (global size rounded up to a
// multiple of 16, 16x16 work-groups: clSLIC.cpp:134-148, 275-296, 325-364;
// photo_consistency.cpp:95-140; depth_refinement.cpp:132-147, 343-372, 743-783, 1349-1370).
// The kernel text is included from oracle/_ref/clcode_cxx.inc, which the Makefile generates
// from the reference tree; no reference source is stored in this repository.
#include "clshim.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

namespace clref {

thread_local WorkItem wi;

void barrier(int)
{
    if (wi.group_barrier) pthread_barrier_wait((pthread_barrier_t *)wi.group_barrier);
}

#include "../_ref/clcode_cxx.inc"

static inline int round_up16(int v) { return (v + 15) / 16 * 16; }

// run `body` once per work-item of a barrier-free kernel; (gx, gy, gz) is the global size
template <class F> static void ndrange(int gx, int gy, int gz, F body)
{
#pragma omp parallel for collapse(2) schedule(dynamic, 4)
    for (int z = 0; z < gz; ++z)
        for (int y = 0; y < gy; ++y) {
            WorkItem &w = wi;
            w.group_barrier = nullptr;
            w.gsz[0] = gx; w.gsz[1] = gy; w.gsz[2] = gz;
            w.lsz[0] = 16; w.lsz[1] = 16; w.lsz[2] = 1;
            w.gid[1] = y; w.gid[2] = z; w.lid[1] = y % 16; w.lid[2] = 0; w.grp[1] = y / 16; w.grp[2] = z;
            for (int x = 0; x < gx; ++x) {
                w.gid[0] = x; w.lid[0] = x % 16; w.grp[0] = x / 16;
                body();
            }
        }
}

// the same, restricted to work-items with y in [y0, y1) and z in [z0, z1) of a (gx, gy, gz) NDRange: a
// bounded SAMPLE of a launch, used only to time the reference's kernels on part of a frame
template <class F> static void ndrange_part(int gx, int gy, int gz, int y0, int y1, int z0, int z1, F body)
{
    // a sample is a few rows: parallel over the work-items of a row as well, so that every core has work
    for (int z = z0; z < z1; ++z) {
#pragma omp parallel for collapse(2) schedule(dynamic, 4)
        for (int y = y0; y < y1; ++y)
            for (int x = 0; x < gx; ++x) {
                WorkItem &w = wi;
                w.group_barrier = nullptr;
                w.gsz[0] = gx; w.gsz[1] = gy; w.gsz[2] = gz;
                w.lsz[0] = 16; w.lsz[1] = 16; w.lsz[2] = 1;
                w.gid[1] = y; w.gid[2] = z; w.lid[1] = y % 16; w.lid[2] = 0; w.grp[1] = y / 16; w.grp[2] = z;
                w.gid[0] = x; w.lid[0] = x % 16; w.grp[0] = x / 16;
                body();
            }
    }
}

} // namespace clref

using namespace clref;

// ---- update_cluster_center needs real work-groups: 256 threads and a barrier -----------------
namespace {
struct UpdJob {
    float3 *lab; clref::uint *idx; float8 *accum;
    int2 map_size, img_size;
    int S, cpl, G;
    float3 *color_local; float2 *location_local; float *count_local; bool *should_add;
    pthread_barrier_t bar;
    int tid;
};
struct UpdThread { UpdJob *job; int tid; };

void *upd_worker(void *p)
{
    UpdThread *t = (UpdThread *)p;
    UpdJob *j = t->job;
    WorkItem &w = wi;
    w.group_barrier = &j->bar;
    w.lsz[0] = 16; w.lsz[1] = 16; w.lsz[2] = 1;
    w.gsz[0] = j->map_size.x * 16; w.gsz[1] = j->map_size.y * 16; w.gsz[2] = j->G;
    w.lid[0] = t->tid % 16; w.lid[1] = t->tid / 16; w.lid[2] = 0;
    for (int g = 0; g < j->G; ++g)
        for (int cy = 0; cy < j->map_size.y; ++cy)
            for (int cx = 0; cx < j->map_size.x; ++cx) {
                w.grp[0] = cx; w.grp[1] = cy; w.grp[2] = g;
                w.gid[0] = cx * 16 + w.lid[0]; w.gid[1] = cy * 16 + w.lid[1]; w.gid[2] = g;
                update_cluster_center(j->lab, j->idx, j->accum, j->map_size, j->img_size, j->S, j->cpl,
                                      j->color_local, j->location_local, j->count_local, j->should_add);
            }
    return nullptr;
}
} // namespace

extern "C" {

int clref_ok(void) { return 1; }

// clSLIC::cvt_color_space (clSLIC.cpp:125-152)
void clref_cvt(const void *rgbx, void *lab, int W, int H)
{
    int2 sz(W, H);
    ndrange(round_up16(W), round_up16(H), 1, [&] { cvt((uchar3 *)rgbx, (float3 *)lab, sz); });
}

// clSLIC::init_cluster_centers (clSLIC.cpp:156-183)
void clref_init_centers(const void *lab, void *map, int W, int H, int mw, int mh, int S)
{
    int2 isz(W, H), msz(mw, mh);
    ndrange(round_up16(mw), round_up16(mh), 1, [&] { init_cluster_centers((float3 *)lab, (float8 *)map, isz, msz, S); });
}

// clSLIC::find_center_association (clSLIC.cpp:268-303)
void clref_assoc(const void *lab, const void *map, void *idx, int W, int H, int mw, int mh, int S, float max_xy_dist,
                 float max_color_dist, float weight)
{
    int2 isz(W, H), msz(mw, mh);
    ndrange(round_up16(W), round_up16(H), 1, [&] {
        find_center_association((float3 *)lab, (float8 *)map, (unsigned int *)idx, isz, msz, S, max_xy_dist,
                                max_color_dist, weight);
    });
}

// clSLIC::update_cluster_center (clSLIC.cpp:307-366): kernel 1 with 16x16 work-groups and four
// local arrays, then finalize_reduction_result
void clref_update_centers(const void *lab, const void *idx, void *map, int W, int H, int mw, int mh, int S)
{
    UpdJob j;
    j.lab = (float3 *)lab; j.idx = (clref::uint *)idx;
    j.map_size = int2(mw, mh); j.img_size = int2(W, H);
    j.S = S;
    float n = (float)(S * S * 9);
    j.G = (int)ceilf(n / (float)(16 * 16));   // clSLIC.cpp:22-23
    j.cpl = S * 3 / 16;                       // clSLIC.cpp:313
    std::vector<float8> accum((size_t)mw * mh * j.G);
    memset(accum.data(), 0, accum.size() * sizeof(float8));
    j.accum = accum.data();
    std::vector<float3> cl(256); std::vector<float2> ll(256); std::vector<float> cn(256);
    bool sa = false;
    j.color_local = cl.data(); j.location_local = ll.data(); j.count_local = cn.data(); j.should_add = &sa;
    pthread_barrier_init(&j.bar, nullptr, 256);
    std::vector<pthread_t> th(256);
    std::vector<UpdThread> ta(256);
    for (int t = 0; t < 256; ++t) { ta[t].job = &j; ta[t].tid = t; pthread_create(&th[t], nullptr, upd_worker, &ta[t]); }
    for (int t = 0; t < 256; ++t) pthread_join(th[t], nullptr);
    pthread_barrier_destroy(&j.bar);
    int2 msz(mw, mh);
    int G = j.G;
    ndrange(round_up16(mw), round_up16(mh), 1, [&] { finalize_reduction_result(accum.data(), (float8 *)map, msz, G); });
}

// apply_edge_alternative (clSLIC.cpp:215-232) from the reference source: deterministic for a given edge map
void clref_apply_edge(const void *lab, const float *edge, void *map, int W, int H, int mw, int mh)
{
    int2 isz(W, H), msz(mw, mh);
    ndrange(round_up16(mw), round_up16(mh), 1, [&] { apply_edge_alternative((float3 *)lab, (float *)edge, (float8 *)map, isz, msz); });
}

// The ARITHMETIC of edge_compute_alternative (clSLIC.cpp:194-211) from the reference source.  The kernel overwrites
// the Lab buffer it reads (clcode.cl:194), so every work-item is run on a private copy of the Lab SNAPSHOT and the
// value it stores at its own pixel is collected: what the kernel computes when no neighbour has been overwritten
// yet -- the semantics this repository pins (oracle.c orc_edge_compute).  Small images only (one copy per pixel).
// Its `float3 color[8]` is indexed up to 8 (clcode.cl:169-184): the ninth store lands outside the array; the
// kernel text is compiled with room for it (see prep.py).
void clref_edge_compute_snapshot(const void *lab, float *edge, int W, int H)
{
    int2 sz(W, H);
    const size_t n = (size_t)W * H;
#pragma omp parallel
    {
        std::vector<float3> copy(n);
#pragma omp for schedule(dynamic, 4)
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x) {
                memcpy(copy.data(), lab, n * sizeof(float3));
                WorkItem &w = wi;
                w.group_barrier = nullptr;
                w.gsz[0] = round_up16(W); w.gsz[1] = round_up16(H); w.gsz[2] = 1;
                w.lsz[0] = 16; w.lsz[1] = 16; w.lsz[2] = 1;
                w.gid[0] = x; w.gid[1] = y; w.gid[2] = 0; w.lid[0] = x % 16; w.lid[1] = y % 16; w.lid[2] = 0;
                w.grp[0] = x / 16; w.grp[1] = y / 16; w.grp[2] = 0;
                edge_compute_alternative(copy.data(), edge, sz);
                edge[(size_t)W * y + x] = copy[(size_t)W * y + x].x;   // the kernel broadcasts the scalar into the float3
            }
    }
}

// clSLIC::enforce_connectivity, one pass (clSLIC.cpp:390-410)
void clref_suppress(const void *in, void *out, int W, int H)
{
    int2 isz(W, H);
    ndrange(round_up16(W), round_up16(H), 1, [&] { supress_local_lable((clref::uint *)in, (clref::uint *)out, isz); });
}

// clPhotoConsistency: find_super_pixel_boundary over all views (photo_consistency.cpp:95-107)
void clref_boundary(const void *map, const void *idx, void *rep, int V, int W, int H, int mw, int mh, int S)
{
    int2 isz(W, H), msz(mw, mh);
    ndrange(round_up16(mw), round_up16(mh), V, [&] {
        find_super_pixel_boundary((float8 *)map, (clref::uint *)idx, (uchar8 *)rep, msz, isz, S);
    });
}

// clPhotoConsistency: initial_depth_estimation_v2, one launch per view (photo_consistency.cpp:114-140)
void clref_block_match(const void *lab, void *map, const void *rep, const void *idx, const float *levels, int nlev,
                       const int *view_subset, const int *subset_num, int array_width, int V, int W, int H, int mw,
                       int mh, int S, float bl_ratio)
{
    int2 isz(W, H), msz(mw, mh);
    for (int z = 0; z < V; ++z)
        ndrange(round_up16(mw), round_up16(mh), 1, [&] {
            initial_depth_estimation_v2((float3 *)lab, (float8 *)map, (uchar8 *)rep, (clref::uint *)idx, (float *)levels,
                                        (int *)view_subset, (int *)subset_num, array_width, msz, isz, bl_ratio, S, nlev,
                                        V, z);
        });
}

// timing sample: superpixel rows [y0, y1) of view z only
void clref_block_match_rows(const void *lab, void *map, const void *rep, const void *idx, const float *levels, int nlev,
                            const int *view_subset, const int *subset_num, int array_width, int V, int W, int H, int mw,
                            int mh, int S, float bl_ratio, int z, int y0, int y1)
{
    int2 isz(W, H), msz(mw, mh);
    ndrange_part(round_up16(mw), round_up16(mh), 1, y0, y1, 0, 1, [&] {
        initial_depth_estimation_v2((float3 *)lab, (float8 *)map, (uchar8 *)rep, (clref::uint *)idx, (float *)levels,
                                    (int *)view_subset, (int *)subset_num, array_width, msz, isz, bl_ratio, S, nlev, V, z);
    });
}

// clDepthRefinement::compute_flatness (depth_refinement.cpp:122-147); gamma = the kernel argument
void clref_flatness(const void *map, void *flat, int V, int mw, int mh, float gamma)
{
    int2 msz(mw, mh);
    ndrange(round_up16(mw), round_up16(mh), V, [&] { compute_flatness((float8 *)map, (float2 *)flat, msz, gamma); });
}

// clDepthRefinement::init_spixl_state (depth_refinement.cpp:343-372); kernel-side scalars
void clref_init_state(const void *map, const void *idx, const void *rep, const void *flat, float *state, float gamma,
                      float alpha, int no_kernel_steps, float kernel_step_size, float bl_ratio, float fuse, int mw, int mh,
                      const int *view_subset, const int *subset_num, int array_width, int W, int H, int V)
{
    int2 isz(W, H), msz(mw, mh);
    ndrange(round_up16(mw), round_up16(mh), V, [&] {
        init_current_state((float8 *)map, (clref::uint *)idx, (uchar8 *)rep, (float2 *)flat, state, gamma, alpha,
                           no_kernel_steps, kernel_step_size, bl_ratio, fuse, msz, (int *)view_subset, (int *)subset_num,
                           array_width, isz, V);
    });
}

// clDepthRefinement::propagate, one iteration (depth_refinement.cpp:743-783); the grid covers all
// views, which the kernel relies on (get_global_size(2) is its subset row stride)
void clref_propagate(float *state_out, const float *state_in, const void *map, const void *idx, const void *rep,
                     const void *flat, const int *view_subset, const int *subset_num, int iter, int mw, int mh, int W,
                     int H, float alpha, float gamma, float bl_ratio, float fuse, int no_kernel_steps,
                     float kernel_step_size, int array_width, int V)
{
    int2 isz(W, H), msz(mw, mh);
    ndrange(round_up16(mw), round_up16(mh), V, [&] {
        propagate(state_out, (float *)state_in, (float8 *)map, (clref::uint *)idx, (uchar8 *)rep, (float2 *)flat,
                  (int *)view_subset, (int *)subset_num, iter, msz, isz, alpha, gamma, bl_ratio, fuse, no_kernel_steps,
                  kernel_step_size, array_width, 1);
    });
}

// timing sample: superpixel rows [y0, y1) of view z only (the grid keeps its full size V in z)
void clref_propagate_rows(float *state_out, const float *state_in, const void *map, const void *idx, const void *rep,
                          const void *flat, const int *view_subset, const int *subset_num, int iter, int mw, int mh, int W,
                          int H, float alpha, float gamma, float bl_ratio, float fuse, int no_kernel_steps,
                          float kernel_step_size, int array_width, int V, int z, int y0, int y1)
{
    int2 isz(W, H), msz(mw, mh);
    ndrange_part(round_up16(mw), round_up16(mh), V, y0, y1, z, z + 1, [&] {
        propagate(state_out, (float *)state_in, (float8 *)map, (clref::uint *)idx, (uchar8 *)rep, (float2 *)flat,
                  (int *)view_subset, (int *)subset_num, iter, msz, isz, alpha, gamma, bl_ratio, fuse, no_kernel_steps,
                  kernel_step_size, array_width, 1);
    });
}

// clDepthRefinement::fusion as shipped = spixl_to_image (depth_refinement.cpp:1349-1370)
void clref_spixl_to_image(const void *map, const void *idx, const float *state, float *disp, int V, int W, int H, int mw,
                          int mh)
{
    int2 isz(W, H), msz(mw, mh);
    ndrange(round_up16(W), round_up16(H), V, [&] {
        spixl_to_image((float8 *)map, (clref::uint *)idx, (float *)state, disp, 0.0f, isz, msz);
    });
}

// multi-view fusion exactly as the host code the reference keeps commented out would run it
// (depth_refinement.cpp:1382-1451): per reference view, project_to_reference_inv then
// remove_view_inconsistency, 2-D NDRanges; disp_proj zero-filled first (the reference never
// initialises it; 0 is the kernels' own "no value").  view_num / view_neighbor arguments are unused
// by the kernel (clcode.cl:2037-2038).
void clref_view_fusion(const float *full, float *proj, float *out, int V, int W, int H, int array_width, float bl_ratio,
                       float fuse_)
{
    int2 isz(W, H);
    float fuse = 0.5 * fuse_;   // depth_refinement.cpp:1328
    memset(proj, 0, sizeof(float) * (size_t)W * H * V);
    for (int r = 0; r < V; ++r) {
        ndrange(round_up16(W), round_up16(H), 1, [&] {
            project_to_reference_inv((float *)full, proj, r, isz, bl_ratio, array_width, V);
        });
        ndrange(round_up16(W), round_up16(H), 1, [&] {
            remove_view_inconsistency(proj, (float *)full, nullptr, nullptr, r, bl_ratio, fuse, isz, array_width, V, out);
        });
    }
}

} // extern "C"
