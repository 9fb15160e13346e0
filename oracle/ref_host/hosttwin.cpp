// hosttwin.cpp -- TEST / BASELINE INFRASTRUCTURE.  Drives the reference's CPU twins -- the *_host methods of
// clPhotoConsistency and clDepthRefinement -- exactly as the reference's own (commented-out) call sites do
// (photo_consistency.cpp:153-164; depth_refinement.cpp:160-166, 395-399, 816-827), from the reference's
// sources compiled UNMODIFIED over the inert headers under shim/ (oracle/ref_host/Makefile reads them where
// they lie under /root/reference; nothing is copied).  Built with -fno-access-control: the twins are private.
//
// They are the CPU path north_star asks to see timed, not an oracle: they have diverged from the kernels
// (hard-coded array_width = 5 / bl_ratio = 0.625, no centre clamp, truncated steps,
// no far-neighbour and pair-plane stages in propagate_host ...).  All are serial except propagate_host
// (#pragma omp parallel for over views, depth_refinement.cpp:1310).
#include "stdafx.h"
#include "photo_consistency.h"
#include "depth_refinement.h"

#include <chrono>
#include <string.h>

static double now_s()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

extern "C" {

int hosttwin_ok(void) { return 1; }

// do_initial_depth_estimation_host for views [z0, z1): map (in/out: s7) and rep (out) in the API layouts
double hosttwin_block_match(float *map, unsigned char *rep, const float *lab, const unsigned *idx, int V, int W, int H,
                            int mw, int mh, int S, int array_width, float bl_ratio, const float *levels, int nlev,
                            const int *subset_mat, const int *subset_num, int z0, int z1)
{
    vec2i isz, msz;
    isz.x = W; isz.y = H; msz.x = mw; msz.y = mh;
    // heap object, never deleted: the reference declares ~clPhotoConsistency (photo_consistency.h:21) but defines it nowhere
    clPhotoConsistency *pc = new clPhotoConsistency(cl::Program(), V, S, nlev, isz, msz);
    // the twin loops `dl < no_disp` (photo_consistency.cpp:332), a member no constructor or method ever assigns
    // (photo_consistency.h:42): the harness gives it the level count, the only value it can have been meant to hold
    pc->no_disp = nlev;
    std::vector<float> lv(levels, levels + nlev);
    double t = now_s();
    pc->do_initial_depth_estimation_host((vec8f *)map, (vec8u *)rep, (vec3f *)lab, (cl_uint *)idx, array_width, bl_ratio, lv,
                                        (int *)subset_mat, (int *)subset_num, z0, z1);
    return now_s() - t;
}

// compute_flatness_host, init_spixl_state_host and no_prop x propagate_host with the scalar preparation of
// compute_flatness / init_spixl_state / propagate (depth_refinement.cpp:130, 335-339, 734-738, 768-769 as the
// commented call at :822-825 passes them).  gamma_, alpha_, kernel_size as pipeline::refine_depth_map passes them.
// times4 = seconds of {flatness, init state, all propagate iterations, 0}.
int hosttwin_refine(const float *map, const unsigned *idx, const unsigned char *rep, const float *lab, int V, int W, int H,
                    int mw, int mh, int aw, int ah, int S, float bl_ratio, const int *subset_mat, const int *subset_num,
                    float gamma_, float alpha_, float fuse_, int kernel_step, int kernel_size, int no_prop, float *flat,
                    float *state_a, float *state_b, double *times4)
{
    vec2i isz, msz, asz;
    isz.x = W; isz.y = H; msz.x = mw; msz.y = mh; asz.x = aw; asz.y = ah;
    std::vector<std::vector<int> > sub(V);
    for (int i = 0; i < V; ++i) sub[i].assign(subset_mat + (size_t)i * V, subset_mat + (size_t)i * V + subset_num[i]);
    clDepthRefinement *dr = new clDepthRefinement(cl::Program(), isz, msz, asz, (vec3f *)lab, (vec8f *)map, (cl_uint *)idx,
                                                  (vec8u *)rep, sub, S, bl_ratio);
    size_t M = (size_t)mw * mh * V;
    // the twins read these members (depth_refinement.cpp:694, 918, 1104): point them at the caller's arrays
    dr->flatness_map = (vec2f *)flat;
    dr->current_state = state_a;
    dr->current_state2 = state_b;
    float gamma_flat = (1.0 / (float)gamma_);                                          // :130
    double t0 = now_s();
    dr->compute_flatness_host((vec2f *)flat, gamma_flat);
    double t1 = now_s();
    float gamma = 1 / gamma_, alpha = 1 / alpha_;                                      // :335-336
    float sp_kernel_step = std::max(1, kernel_size / kernel_step * S);                 // :338
    float fuse = 0.5 * fuse_;                                                          // :339
    dr->init_spixl_state_host(state_a, 0, V, gamma, alpha, fuse, kernel_step, sp_kernel_step);
    double t2 = now_s();
    memcpy(state_b, state_a, sizeof(float) * 6 * M);
    for (int iter = 0; iter < no_prop; ++iter) {                                       // :822-825
        // H7: compute_smoothness / compute_consistency read the MEMBER current_state, whatever they are passed
        dr->current_state = iter % 2 == 0 ? state_a : state_b;
        if (iter % 2 == 0)
            dr->propagate_host(state_a, state_b, iter, 1, gamma, alpha, kernel_step / (iter + 1), sp_kernel_step / (iter + 1), fuse, 0, V);
        else
            dr->propagate_host(state_b, state_a, iter, 1, gamma, alpha, kernel_step / (iter + 1), sp_kernel_step / (iter + 1), fuse, 0, V);
    }
    double t3 = now_s();
    times4[0] = t1 - t0; times4[1] = t2 - t1; times4[2] = t3 - t2; times4[3] = 0;
    // the reference's destructor deletes arrays it does not own here; detach them and leak the small rest
    dr->flatness_map = nullptr; dr->current_state = nullptr; dr->current_state2 = nullptr;
    return 0;
}

} // extern "C"
