// common.cuh -- shared declarations of the sm_100a depth-pipeline library.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace mvs {

// Host-side cache of the block matcher's tensor maps (TMA boxes over the frame's Lab stack), one per geometric class
// of a neighbour view: same camera row (0), same camera column (1), diagonal (2).  Rebuilt when the frame's Lab
// pointer, the disparity range or bl_ratio change.  state: 0 = not built, 1 = ready, -1 = this geometry does not fit.
struct BmTma {
    CUtensorMap map[3];
    int bw[3], bh[3];
    const void *lab;
    float dmin, dmax, bl_ratio;
    int nlev;
    int state;
};

// the same for the centre update of SLIC (S = 8): box loads of the Lab tile and the label tile of a CTA
struct SlicTma {
    CUtensorMap map_lab, map_idx;
    const void *lab, *idx;
    int state;
};

// Geometry + device pointers of one frame (all V views; view-major arrays).
struct FrameDev {
    int V, W, H, mw, mh, S, aw, ah;
    int N;              // W*H
    int M;              // mw*mh
    uchar4 *rgbx;       // [V][N]
    float4 *lab;        // [V][N]   {L,a,b,0}
    float4 *labT;       // [V][W][H]         Lab transposed            } level-aligned copies for the block matcher
    float4 *labSp;      // [V][W+H-1][H]     Lab sheared along x - y   } (launch_lab_shear; allocated on demand,
    float4 *labSm;      // [V][W+H-1][H]     Lab sheared along x + y   }  photo.cu)
    uint32_t *idx;      // [V][N]
    uint32_t *idx_tmp;  // [V][N]   connectivity ping-pong (allocated on demand)
    float4 *map;        // [V][M][2] float8 {id,cx,cy,L | a,b,n,disp}
    uint2 *rep;         // [V][M]   uchar8 packed little-endian {nw,w,sw,n | s,ne,e,se}
    float2 *flat;       // [V][M]
    float4 *plane[2];   // [V][M]   {d,nx,ny,nz}  internal SoA split of the 6-float state (A,B)
    float2 *cost[2];    // [V][M]   {sm,cs}
    float *state6[2];   // [V][M][6] API layout, filled on export
    float *disp;        // [V][N]
    uint2 *pixd;        // [V][N]   {label, disparity of the label's INPUT plane at the pixel}: what propagate's consistency
                        //          walk gathers per sample (rebuilt before every iteration; allocated on demand)
    float *disp_proj;   // [V][N]   multi-view fusion: chased disparity (allocated on demand)
    float *disp_fused;  // [V][N]   multi-view fusion: voted disparity (allocated on demand)
    float4 *spcol;      // [V][M]   {L, a, b, 0} per superpixel, packed from map after block matching (allocated on demand)
    float *edge;        // [V][N]   edge-aware SLIC: Sobel magnitude of Lab (edge_enable; allocated on demand)
    uint8_t *io_stage;  // [V][N][4] staging of packed BGR input / 8-bit and overlay output (allocated on demand)
    int2 *cam;          // [V]      (x, y) of every view's camera in the array: view z sits at (z % aw, z / aw) unless
                        //          mvs_frame_set_view_order placed the cameras in another slot order
    int *view_subset;   // [V][V]
    int *subset_num;    // [V]
    float *levels;      // [L]
    int num_levels;
    BmTma *bm_tma;      // host pointer (owned by the frame): tensor maps of the block matcher
    SlicTma *slic_tma;  // host pointer (owned by the frame): tensor maps of the centre update
    int unit_baseline;  // every neighbour view at most one camera apart in x and y
    int max_neighbours; // largest subset_num over the views
};

struct SlicArgs {
    float max_xy_dist, max_color_dist, weight;
};

struct RefineArgs {
    float gamma, alpha, fuse, bl_ratio;
    int no_kernel_steps;
    float kernel_step_size;
    int iter;
};

// launchers (all asynchronous on `st`); z0..z1-1 = views to compute.  Each returns the
// number of kernel launches it issued.
int launch_cvt(const FrameDev &f, int z0, int z1, cudaStream_t st);
int launch_init_centers(const FrameDev &f, int z0, int z1, cudaStream_t st);
int launch_apply_edge_values(const FrameDev &f, int z0, int z1, cudaStream_t st);
int launch_assoc(const FrameDev &f, const SlicArgs &a, int z0, int z1, cudaStream_t st);
int launch_update_centers(const FrameDev &f, int z0, int z1, cudaStream_t st);
int prepare_update_centers_tma(const FrameDev &f, const char **why);
int launch_connectivity(const FrameDev &f, int z0, int z1, cudaStream_t st);

int launch_boundary(const FrameDev &f, int z0, int z1, cudaStream_t st);
int launch_block_match(const FrameDev &f, float bl_ratio, int z0, int z1, cudaStream_t st);
int launch_lab_shear(const FrameDev &f, int z0, int z1, cudaStream_t st);
int prepare_block_match_tma(const FrameDev &f, const float *levels_host, int nlev, float bl_ratio, const char **why);

int launch_flatness(const FrameDev &f, float gamma_flat, int z0, int z1, cudaStream_t st);
int launch_init_state(const FrameDev &f, const RefineArgs &a, int z0, int z1, cudaStream_t st);
int launch_plane_from_map(const FrameDev &f, int z0, int z1, cudaStream_t st);
int launch_propagate(const FrameDev &f, const RefineArgs &a, int src, int dst, int z0, int z1, cudaStream_t st,
                     unsigned long long *stats = nullptr);
int launch_spixl_to_image(const FrameDev &f, int src, int z0, int z1, cudaStream_t st);
int launch_pixel_plane(const FrameDev &f, int src, int z0, int z1, cudaStream_t st);
int launch_spixl_colour(const FrameDev &f, int z0, int z1, cudaStream_t st);
int launch_view_fusion(const FrameDev &f, float bl_ratio, float fuse, int z0, int z1, cudaStream_t st);
// the no-allocation variant (fuse.cu): scratch in a file-scope device array of this many floats (16 MiB)
constexpr size_t VIEW_FUSION_STATIC_MAX_FLOATS = (size_t)4 << 20;
int launch_view_fusion_static(int device, cudaStream_t st, const float *disp, float *fused, int V, int W, int H, int aw,
                              float bl_ratio, float fuse, const char **why);
int launch_ingest_bgr(const FrameDev &f, const void *bgr_dev, int z0, int z1, cudaStream_t st);
int launch_quantise_disp(const FrameDev &f, const float *src, uint8_t *out_dev, float mn, float range, int z0, int z1,
                         cudaStream_t st);
int launch_quantise_state(const FrameDev &f, int which, int element, uint8_t *out_dev, float mn, float range, int z0,
                          int z1, cudaStream_t st);
int launch_quantise_spixl(const FrameDev &f, const float *src, int stride, int off, int mode, uint8_t *out_dev, float mn,
                          float range, int z0, int z1, cudaStream_t st);
int launch_segmentation_lines(const FrameDev &f, uchar4 *out_dev, int z0, int z1, cudaStream_t st);
int launch_math_eval(int which, const float *x_dev, float *y_dev, size_t n, cudaStream_t st);
int launch_export_state(const FrameDev &f, int which, int z0, int z1, cudaStream_t st);
int launch_import_state(const FrameDev &f, int which, int z0, int z1, cudaStream_t st);

} // namespace mvs
