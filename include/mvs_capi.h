/*
 * mvs_capi.h -- C ABI of the B200-native per-view depth pipeline
 * (SLIC -> block matching -> plane refinement -> fusion).
 *
 * The reference (kianoosh-j/CL_MultiView_Stereo, directory clMVDE/) has no C
 * ABI: its stage classes call OpenCL directly.  This layer sits UNDER the kept
 * C++ classes (include/clSLIC.h, photo_consistency.h, depth_refinement.h,
 * pipeline.h) and is what a foreign-language binding (ctypes, cgo, JNI ...)
 * binds.  Each entry point cites the reference interface it replaces.
 *
 * Conventions
 *  - every function returns 0 on success, non-zero on failure;
 *    mvs_last_error() returns a thread-local message for the last failure.
 *  - there is NO CPU fallback: without a CUDA device every compute call fails.
 *  - host pointers may be pageable; "dev" pointers are CUDA device pointers.
 *  - layouts are the reference's:
 *      rgbx   uchar4 {s0,s1,s2,pad}           [V][H][W]      (vec3u, file_handler.cpp:6-15)
 *      lab    float4 {L,a,b,pad=0}            [V][H][W]      (vec3f)
 *      idx    uint32 local superpixel index   [V][H][W]      (cl_uint)
 *      map    float8 {id,cx,cy,L,a,b,n,disp}  [V][mh][mw]    (vec8f)
 *      rep    uchar8 {nw,w,sw,n,s,ne,e,se}    [V][mh][mw]    (vec8u)
 *      flat   float2                          [V][mh][mw]    (vec2f)
 *      state  float[6] {d,sm,cs,nx,ny,nz}     [V][mh][mw]
 *      disp   float                           [V][H][W]
 */
#ifndef MVS_CAPI_H
#define MVS_CAPI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mvs_ctx mvs_ctx;       /* one CUDA device + streams (+ NCCL communicator) */
typedef struct mvs_frame mvs_frame;   /* device-resident buffers of one V-view frame */

/* Parameters of a frame.  Mirrors system_settings (header.h:55-77) plus the
 * image geometry pipeline::pipeline derives (pipeline.cpp:15-19). */
typedef struct mvs_params {
    int width, height;           /* image size */
    int spixl_size;              /* S */
    int array_width, array_height;
    float slic_color_weight;
    int no_iter;
    int enforce_connectivity;
    int neib_hor, neib_ver;
    int min_disp, max_disp, inc;
    float bl_ratio;
    int kernel_size;             /* as in system_settings (halved internally, pipeline.cpp:166) */
    int kernel_step;
    float fuse, gamma, alpha;    /* as in system_settings (2*g^2, 2*a^2 applied internally, pipeline.cpp:164-165) */
    int no_prop;
    int edge_enable;             /* clSLIC::apply_edge_values before the first association (clSLIC.cpp:86, 186-233) */
} mvs_params;

/* selectors for mvs_frame_download / mvs_frame_upload / mvs_frame_dev_ptr */
enum mvs_buffer {
    MVS_BUF_RGBX = 0, MVS_BUF_LAB = 1, MVS_BUF_IDX = 2, MVS_BUF_MAP = 3, MVS_BUF_REP = 4,
    MVS_BUF_FLAT = 5, MVS_BUF_STATE_A = 6, MVS_BUF_STATE_B = 7, MVS_BUF_DISP = 8,
    MVS_BUF_DISP_PROJ = 9, MVS_BUF_DISP_FUSED = 10   /* multi-view fusion, allocated by mvs_frame_view_fusion */
};

const char *mvs_last_error(void);
int mvs_device_count(int *count);

/* replaces pipeline::compile_device_code (pipeline.cpp:31-56): binds a device, creates streams */
int mvs_ctx_create(int device, mvs_ctx **out);
int mvs_ctx_destroy(mvs_ctx *ctx);
int mvs_ctx_sync(mvs_ctx *ctx);
/* number of kernel launches issued through this context so far */
int mvs_ctx_launch_count(mvs_ctx *ctx, long long *count);
/* the CUDA stream (cudaStream_t) every call on this context is ordered on */
int mvs_ctx_stream(mvs_ctx *ctx, void **stream);
/* CUDA-event timing on that stream: record into slot 0..15, elapsed ms between two slots */
int mvs_ctx_event_record(mvs_ctx *ctx, int slot);
int mvs_ctx_event_elapsed(mvs_ctx *ctx, int slot_a, int slot_b, float *ms);
/* Orders the work queued on `ctx` from now on after the event `other` recorded into `slot` (same device).  Two
 * contexts that stream frames alternately use it to run their KERNELS one frame after the other (a frame's stages
 * keep the L2 to themselves) while the host<->device copies of one frame overlap the kernels of the other. */
int mvs_ctx_wait_event(mvs_ctx *ctx, mvs_ctx *other, int slot);
/* page-locked host memory for the asynchronous copies */
int mvs_host_alloc(void **p, size_t bytes);
int mvs_host_free(void *p);

/* multi-GPU: one process per GPU.  id is an opaque 128-byte NCCL unique id made by rank 0
 * and distributed by the caller (torch.distributed / MPI / files). */
int mvs_comm_unique_id(void *id128);
int mvs_comm_init(mvs_ctx *ctx, const void *id128, int nranks, int rank);
/* contiguous partition of V views over nranks (the plan mvs_frame_exchange assumes) */
int mvs_view_partition(int V, int nranks, int rank, int *lo, int *hi);
/* contiguous partition that balances per-view WEIGHTS instead of view counts (a view's work grows with its number
 * of neighbour views: pipeline.cpp:121-142 gives a corner view 3 and an interior view 8): minimises the heaviest
 * rank.  lo_hi: 2*nranks ints {lo_0,hi_0,lo_1,hi_1,...}.  Hand the table to mvs_frame_set_partition on every rank. */
int mvs_view_partition_weighted(int V, int nranks, const int *weight, int *lo_hi);
/* the halo plan mvs_frame_exchange follows (pure host code, no device needed): every transfer
 * {view, source rank, destination rank} in issue order, for the neighbour lists the reference
 * builds at pipeline.cpp:121-142.  triples may be NULL to query *count only. */
int mvs_exchange_plan(int V, int nranks, const int *view_subset, const int *subset_num, int *triples, int cap,
                      int *count);
/* the same for an explicit table of view ranges (mvs_view_partition_weighted) */
int mvs_exchange_plan_for(int V, int nranks, const int *lo_hi, const int *view_subset, const int *subset_num,
                          int *triples, int cap, int *count);

/* ---- frame: device-resident pipeline over V = array_width*array_height views ---- */
/* view_lo..view_hi-1 are the views this context computes (all views for single GPU);
 * buffers for every view are allocated because neighbour views are read (clcode.cl:1046,1314,1581). */
int mvs_frame_create(mvs_ctx *ctx, const mvs_params *p, int view_lo, int view_hi, mvs_frame **out);
int mvs_frame_destroy(mvs_frame *f);
/* The cameras of the array in another SLOT order: slot s of every per-view array, and of every view index of this
 * API (uploads, downloads, view ranges, the partition table), holds camera camera_of_slot[s] -- a permutation of
 * 0..V-1 in the row-major numbering pipeline.cpp:130-142 uses.  Neighbour lists keep the reference's order per camera,
 * so a camera's results are the same bits in any order.  Lets a sharded frame give a rank ANY set of cameras as one
 * contiguous slot range.  Call before the first upload; mvs_frame_view_fusion refuses a frame in non-identity order.
 * mvs_view_order_balanced proposes an order and the ranks' slot ranges {lo_0,hi_0,...} for nranks ranks, minimising the
 * heaviest rank under cost = sum over own cameras (w_base + w_neighbour * neighbour views) + w_halo * views read but
 * not owned.  Pure host code; no counterpart in the reference (single device). */
int mvs_frame_set_view_order(mvs_frame *f, const int *camera_of_slot);
int mvs_view_order_balanced(int array_width, int array_height, int neib_hor, int neib_ver, int nranks, int w_base,
                            int w_neighbour, int w_halo, int *camera_of_slot, int *lo_hi);
/* view ranges of all ranks when they are not mvs_view_partition's: same table on every rank, this rank's entry
 * must be the [view_lo,view_hi) the frame was created with.  Pure host call. */
int mvs_frame_set_partition(mvs_frame *f, int nranks, const int *lo_hi);
/* host <-> device copy of one buffer for views [v0,v1) */
int mvs_frame_upload(mvs_frame *f, int which, int v0, int v1, const void *host);
int mvs_frame_download(mvs_frame *f, int which, int v0, int v1, void *host);
/* same, asynchronous on the context stream (host memory should be page-locked) */
int mvs_frame_upload_async(mvs_frame *f, int which, int v0, int v1, const void *host);
int mvs_frame_download_async(mvs_frame *f, int which, int v0, int v1, void *host);
int mvs_frame_dev_ptr(mvs_frame *f, int which, void **dev);
/* loadImageIn (file_handler.cpp:6-15) on the device: packed BGR raster (cv::Mat order, 3 B/px) of views
 * [v0,v1) -> MVS_BUF_RGBX {s0=R,s1=G,s2=B,pad} */
int mvs_frame_upload_bgr(mvs_frame *f, int v0, int v1, const uint8_t *bgr);
/* 8-bit visualisation images quantised on the device: plot_full_image (depth_refinement.cpp:1476-1497) for
 * MVS_BUF_DISP / MVS_BUF_DISP_FUSED; img_translate_state (depth_refinement.cpp:1500-1560) for MVS_BUF_STATE_A/B
 * with element 0 (d), 1 (smoothness), 2 (consistency); img_translate (photo_consistency.cpp:414-436) for
 * MVS_BUF_MAP (block-matching disparity s7); img_translate_flatness (depth_refinement.cpp:299-321) for
 * MVS_BUF_FLAT (ceil(flatness.s0 * 255), min/max unused).  host: uint8[(v1-v0)*H*W] */
int mvs_frame_export_u8(mvs_frame *f, int which, int element, int min_disp, int max_disp, int v0, int v1,
                        uint8_t *host);
/* clSLIC::draw_segmentation_lines (clSLIC.h:20, clSLIC.cpp:447-478) from the resident image and labels of views
 * [v0,v1): interior pixels next to another label become {0,0,255}, the others take the input colour.  rgbx
 * (uint8[(v1-v0)*H*W*4]) is in/out like the reference's out_img: its one-pixel frame and pad bytes are kept. */
int mvs_frame_draw_segmentation(mvs_frame *f, int v0, int v1, uint8_t *rgbx);
int mvs_frame_sizes(mvs_frame *f, int *V, int *W, int *H, int *mw, int *mh, int *num_levels);

/* clSLIC::do_super_pixel_seg (clSLIC.cpp:67-122) for the frame's own views, all views per launch */
int mvs_frame_slic(mvs_frame *f);
/* clPhotoConsistency::do_initial_depth_estimation (photo_consistency.cpp:21-150):
 * find_super_pixel_boundary + initial_depth_estimation_v2, writes rep and map.s7 */
int mvs_frame_block_match(mvs_frame *f);
/* clDepthRefinement::do_refinement (depth_refinement.cpp:91-118): flatness, init state,
 * no_prop x propagate (ping-pong A/B), spixl_to_image from state A */
int mvs_frame_refine(mvs_frame *f);
/* the individual refinement stages (depth_refinement.cpp:122, 327, 724, 1318) */
int mvs_frame_flatness(mvs_frame *f);
int mvs_frame_init_state(mvs_frame *f);
int mvs_frame_propagate(mvs_frame *f, int iter);
/* one propagate iteration with work counters (diagnostic): stats12 = {superpixels, candidate planes, unique planes
 * evaluated, 0, 0 (two retired counters), (plane, view) items, items with samples inside the neighbour
 * image, items with a visible sample, samples inside, visible samples, 0, 0} */
int mvs_frame_propagate_stats(mvs_frame *f, int iter, unsigned long long *stats12);
int mvs_frame_fusion(mvs_frame *f);
/* multi-view fusion: project_to_reference_inv + remove_view_inconsistency per reference view
 * (clcode.cl:1995-2101) in the host sequence the reference keeps disabled (depth_refinement.cpp:1382-1451),
 * disp_proj zero-filled.  Reads MVS_BUF_DISP of ALL views, fills MVS_BUF_DISP_PROJ and MVS_BUF_DISP_FUSED.
 * Not part of mvs_frame_run (the shipped pipeline stops at spixl_to_image). */
int mvs_frame_view_fusion(mvs_frame *f);
/* pipeline::exe_pipeline with perform_depth_est enabled (pipeline.cpp:60-64, 108-151) */
int mvs_frame_run(mvs_frame *f);
/* multi-GPU halo exchange of the buffers neighbouring views read (no-op without a communicator) */
int mvs_frame_exchange(mvs_frame *f, int which);
/* per-stage device time (ms) of the last mvs_frame_run: slic, boundary, block_match, flatness,
 * init_state, propagate(total), fusion, exchange -- 8 floats */
int mvs_frame_stage_ms(mvs_frame *f, float *ms8);
/* diagnostic: the device's pinned math (csrc/mvs_math.cuh) on host arrays, for bit-for-bit comparison with the oracle's
 * copy.  which = 0 expf, 1 expf (arguments <= 0), 2 polynomial-only expf (|x| < 0.3465), 3 cube root (x > 0),
 * 4 round half away from zero, 5 float -> int (y holds the int's bits) */
int mvs_math_eval(mvs_ctx *ctx, int which, const float *x, float *y, size_t n);
/* view-sharded frame: the caller uploads the INPUT images of the halo views (the views this rank's own views read) as
 * well as its own, so that no image crosses between ranks; mvs_frame_halo_views lists them */
int mvs_frame_set_halo_inputs(mvs_frame *f, int local);
int mvs_frame_halo_views(mvs_frame *f, int *views, int cap, int *count);
/* per-stage CUDA events on/off (off by default; mvs_frame_stage_ms reads and resets them) */
int mvs_frame_set_timing(mvs_frame *f, int enable);

/* ---- class-level entry points with the reference's host-pointer semantics ---- */

/* clSLIC::do_super_pixel_seg(in_img, cvt_img, spixl_map, idx_img) (clSLIC.h:19, clSLIC.cpp:67-122)
 * for `count` images of the same size; outputs may be NULL to skip the copy back. */
int mvs_slic_run(mvs_ctx *ctx, int width, int height, int spixl_size, float color_weight, int no_iter,
                 int enforce_connectivity, int count, const void *rgbx, void *lab, void *map, uint32_t *idx);
/* the same with system_settings::edge_enable (clSLIC.cpp:86): edge-aware centre placement before the first
 * association; `edge` (may be NULL) receives the Sobel magnitude image [count][H][W] float */
int mvs_slic_run_edge(mvs_ctx *ctx, int width, int height, int spixl_size, float color_weight, int no_iter,
                      int enforce_connectivity, int edge_enable, int count, const void *rgbx, void *lab, void *map,
                      uint32_t *idx, float *edge);

/* clSLIC::draw_segmentation_lines (clSLIC.h:20, clSLIC.cpp:447-478) on the device: `count` images [count][H][W]
 * vec3u + their labels in; out_rgbx is in/out (interior pixels are rewritten, the one-pixel frame is kept). */
int mvs_draw_segmentation_run(mvs_ctx *ctx, int width, int height, int count, const uint32_t *idx, const void *in_rgbx,
                              void *out_rgbx);
/* clPhotoConsistency::do_initial_depth_estimation (photo_consistency.h:10-19, .cpp:21-150).
 * view_subset is the flattened int[V*V] (row stride V) and subset_num int[V] the reference
 * builds at photo_consistency.cpp:39-47.  Fills rep and map[].s7. */
int mvs_bm_run(mvs_ctx *ctx, int view_count, int width, int height, int spixl_size, float *map, uint8_t *rep,
               const float *lab, const uint32_t *idx, int array_width, float bl_ratio, const int *view_subset,
               const int *subset_num, const float *disp_levels, int num_levels);

/* clDepthRefinement ctor + do_refinement (depth_refinement.h:6-9, .cpp:5-118).  gamma/alpha/
 * kernel_size are the values pipeline::refine_depth_map passes (pipeline.cpp:164-166).
 * Outputs (any may be NULL): flat float2[V*M], state_a/state_b float[6*V*M], disp float[V*N]. */
int mvs_refine_run(mvs_ctx *ctx, int width, int height, int array_width, int array_height, int spixl_size,
                   float bl_ratio, const float *map, const uint32_t *idx, const uint8_t *rep,
                   const int *view_subset, const int *subset_num, float gamma, float alpha, float fuse,
                   int kernel_step, int kernel_size, int no_prop, float *flat, float *state_a, float *state_b,
                   float *disp);

/* the disabled tail of clDepthRefinement::fusion (depth_refinement.cpp:1382-1451) on host arrays:
 * disp float[V*N] in; disp_proj / disp_fused float[V*N] out (either may be NULL).  fuse is
 * system_settings::fuse (halved internally, depth_refinement.cpp:1328). */
/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The computation of
 * mvs_view_fusion_run's disp_fused, allocating nothing: disp and disp_fused are DEVICE pointers to
 * view_count*height*width floats each (input, output; caller-owned), and the scratch between the two kernels (the
 * chased disparities) is a file-scope device array of the library, MVS_VIEW_FUSION_STATIC_MAX_FLOATS floats = 16 MiB,
 * which every GPU that loads the library holds until the process exits.  view_count*height*width beyond that limit is
 * refused.  The work is queued on the context's stream and not waited for (mvs_ctx_sync).  Every call on a device
 * shares the one array, which keeps the last call's contents (the part a call uses is cleared first): calls that
 * overlap on the GPU -- two contexts of one device -- race; order them (mvs_ctx_wait_event) or keep to one context.
 * No cudaMalloc / cudaFree / allocating library call on this path.  fuse as in mvs_view_fusion_run. */
#define MVS_VIEW_FUSION_STATIC_MAX_FLOATS (4u << 20)
int mvs_view_fusion_static(mvs_ctx *ctx, int view_count, int width, int height, int array_width, float bl_ratio,
                           float fuse, const float *disp_dev, float *disp_fused_dev);

int mvs_view_fusion_run(mvs_ctx *ctx, int view_count, int width, int height, int array_width, float bl_ratio,
                        float fuse, const float *disp, float *disp_proj, float *disp_fused);

#ifdef __cplusplus
}
#endif
#endif
