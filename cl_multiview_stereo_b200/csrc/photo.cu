// photo.cu -- photo-consistency disparity initialisation (sm_100a).
//
// Replaces the OpenCL kernels driven by clPhotoConsistency::do_initial_depth_estimation
// (reference clMVDE/photo_consistency.cpp:21-150): find_super_pixel_boundary
// (clcode.cl:791-855) and initial_depth_estimation_v2 (clcode.cl:972-1069).  The reference
// launches the matcher once per view; here every view is covered by one launch.
#include "common.cuh"
#include "mvs_math.cuh"

namespace mvs {

// ---------------------------------------------------------------------------------------
// Superpixel extent along 8 compass rays.  Centre truncated and clamped away from the
// border by S (clcode.cl:806-819); for i = 1..S-1 the last own-label hit wins and is
// stored as i-1 (clcode.cl:826-851).  The reference reads the label before testing the
// bounds; bounds are tested first here (same truth value).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_boundary(const float4 *__restrict__ map, const uint32_t *__restrict__ idx,
                                                  uint2 *__restrict__ rep, int W, int H, int S, int N, int M, int z0)
{
    int sp = blockIdx.x * blockDim.x + threadIdx.x;
    int z = z0 + blockIdx.y;
    if (sp >= M) return;
    float4 m0 = map[2 * ((size_t)z * M + sp)];
    int cx = f2i(m0.y), cy = f2i(m0.z);
    if (cx < S) cx += (S - cx);
    if (cx + S > W) cx -= S;
    if (cy < S) cy += (S - cy);
    if (cy + S > H) cy -= S;
    const uint32_t *im = idx + (size_t)z * N;
    uint32_t dir[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const int rdx[8] = {-1, -1, -1, 0, 0, 1, 1, 1};
    const int rdy[8] = {-1, 0, 1, -1, 1, -1, 0, 1};
    for (int i = 1; i < S; i++) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            int x = cx + rdx[k] * i, y = cy + rdy[k] * i;
            if (x >= 0 && y >= 0 && x < W && y < H && im[(size_t)y * W + x] == (uint32_t)sp) dir[k] = (uint32_t)(i - 1);
        }
    }
    uint2 o;
    o.x = dir[0] | (dir[1] << 8) | (dir[2] << 16) | (dir[3] << 24);
    o.y = dir[4] | (dir[5] << 8) | (dir[6] << 16) | (dir[7] << 24);
    rep[(size_t)z * M + sp] = o;
}

int launch_boundary(const FrameDev &f, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    dim3 g((f.M + 255) / 256, z1 - z0);
    k_boundary<<<g, 256, 0, st>>>(f.map, f.idx, f.rep, f.W, f.H, f.S, f.N, f.M, z0);
    return 1;
}

// ---------------------------------------------------------------------------------------
// Block matching.  One WARP per superpixel, lanes over disparity levels (31 levels at the
// shipped settings fill a warp; more levels loop in chunks of 32).  The 25 reference
// samples of the superpixel (positions + Lab) are level- and view-invariant: they are
// staged once per warp in shared memory and read back as warp-wide broadcasts.  For a
// neighbouring view in the same camera row, consecutive levels project to consecutive
// pixels, so a warp-wide gather of the projected Lab is one contiguous 512-byte segment.
// Cost accumulation order is the reference's: per (level, view) a sequential float sum
// over i (x) outer / j (y) inner with the '+30, -30' sequence (clcode.cl:1030-1053);
// min over views and argmin over levels use strict '<' (first view / lowest level wins,
// clcode.cl:1054-1063); the cross-lane argmin breaks ties toward the lower level.
// ---------------------------------------------------------------------------------------
constexpr int BM_WARPS = 8;

__global__ void __launch_bounds__(BM_WARPS * 32)
k_block_match(const float4 *__restrict__ lab, float4 *__restrict__ map, const uint2 *__restrict__ rep,
              const float *__restrict__ levels, int disp_num, const int *__restrict__ view_subset,
              const int *__restrict__ subset_num, int aw, int V, int W, int H, int N, int M, float bl_ratio, int z0)
{
    __shared__ float4 s_ref[BM_WARPS][25];   // Lab of the reference samples
    __shared__ int2 s_xy[BM_WARPS][25];      // their pixel coordinates (x = -1 marks "outside")
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int sp = blockIdx.x * BM_WARPS + warp;
    int z = z0 + blockIdx.y;
    if (sp >= M) return;   // whole warp leaves together
    size_t mi = (size_t)z * M + sp;
    uint2 r = rep[mi];
    int d0 = r.x & 255, d1 = (r.x >> 8) & 255, d2 = (r.x >> 16) & 255, d3 = r.x >> 24;
    int d4 = r.y & 255, d5 = (r.y >> 8) & 255, d6 = (r.y >> 16) & 255, d7 = r.y >> 24;
    int bb_l = max(d0, max(d1, d2)), bb_r = max(d5, max(d6, d7));
    int bb_t = max(d0, max(d3, d5)), bb_b = max(d2, max(d4, d7));
    // fmax(1.0, 0.25*(float)(l+r)) is exact in float (multiples of 1/4 below 128)
    float stepx = fmaxf(1.0f, 0.25f * (float)(bb_l + bb_r));
    float stepy = fmaxf(1.0f, 0.25f * (float)(bb_t + bb_b));
    float4 m0 = map[2 * mi];
    float cxf = m0.y, cyf = m0.z;
    const float4 *ref_img = lab + (size_t)z * N;
    if (lane < 25) {
        int i = lane / 5 - 2, j = lane % 5 - 2;
        int xr = f2i(cxf + (float)i * stepx);
        int yr = f2i(cyf + (float)j * stepy);
        bool ok = xr >= 0 && yr >= 0 && xr < W && yr < H;
        s_xy[warp][lane] = make_int2(ok ? xr : -1, yr);
        s_ref[warp][lane] = ok ? ref_img[(size_t)yr * W + xr] : make_float4(0, 0, 0, 0);
    }
    __syncwarp();

    int refx = z % aw, refy = z / aw;
    int ns = subset_num[z];
    float cost_est = 1000000.0f, disp_est = 0.0f;
    for (int base = 0; base < disp_num; base += 32) {
        int dl = base + lane;
        bool active = dl < disp_num;
        float d = active ? levels[dl] : 0.0f;
        float bd = bl_ratio * d;
        float min_val = 1000000.0f;
        for (int n = 0; n < ns; n++) {
            int view = view_subset[V * z + n];
            float dx = (float)(view % aw - refx), dy = (float)(view / aw - refy);
            float sx = d * dx, sy = bd * dy;
            const float4 *pimg = lab + (size_t)view * N;
            float val = 0.0f;
#pragma unroll 5
            for (int s = 0; s < 25; ++s) {
                int2 xy = s_xy[warp][s];
                int xp = f2i((float)xy.x - sx);
                int yp = f2i((float)xy.y - sy);
                val += 30.0f;
                if (xy.x >= 0 && xp >= 0 && yp >= 0 && xp < W && yp < H) {
                    val -= 30.0f;
                    float4 cr = s_ref[warp][s];
                    float4 cp = pimg[(size_t)yp * W + xp];
                    val += fabsf(cr.x - cp.x) + fabsf(cr.y - cp.y) + fabsf(cr.z - cp.z);
                }
            }
            if (val < min_val) min_val = val;
        }
        // warp argmin over this chunk of levels, ties -> lower level index
        float bv = active ? min_val : __int_as_float(0x7f800000);
        int bi = dl;
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov < bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (bv < cost_est) { cost_est = bv; disp_est = levels[bi]; }
    }
    if (lane == 0) reinterpret_cast<float *>(map + 2 * mi + 1)[3] = disp_est;
}

int launch_block_match(const FrameDev &f, float bl_ratio, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    dim3 g((f.M + BM_WARPS - 1) / BM_WARPS, z1 - z0);
    k_block_match<<<g, BM_WARPS * 32, 0, st>>>(f.lab, f.map, f.rep, f.levels, f.num_levels, f.view_subset,
                                               f.subset_num, f.aw, f.V, f.W, f.H, f.N, f.M, bl_ratio, z0);
    return 1;
}

} // namespace mvs
