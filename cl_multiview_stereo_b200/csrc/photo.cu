// photo.cu -- photo-consistency disparity initialisation (sm_100a).
//
// Replaces the OpenCL kernels driven by clPhotoConsistency::do_initial_depth_estimation
// (reference clMVDE/photo_consistency.cpp:21-150): find_super_pixel_boundary
// (clcode.cl:791-855) and initial_depth_estimation_v2 (clcode.cl:972-1069).  The reference
// launches the matcher once per view; here every view is covered by one launch.
#include "common.cuh"
#include "mvs_math.cuh"
#include <algorithm>

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
// Block matching.  One WARP per superpixel.  The 25 reference samples of the superpixel
// (positions + Lab) are level- and view-invariant: staged once per warp in shared memory.
//
// The kernel is bound by the L1 tag stage (one cycle per distinct 128-byte line a warp-wide
// gather touches), so the lane mapping of the gather follows the geometry of the pair:
//  * neighbour in the same camera row (dy = 0): lanes = disparity levels.  Consecutive levels
//    project to consecutive pixels of ONE image row: a warp-wide gather is one contiguous
//    512-byte segment (4-5 lines).
//  * neighbour in another camera row: levels land on 32 different image rows (32 lines per
//    gather with lanes = levels).  Here the lanes run over (level, x-sample) instead, so that
//    the five x-samples of a sample row share a line (6-10 lines per gather); the per-sample
//    terms are staged in shared memory and the cost chain is then run with lanes = levels.
// Cost accumulation order is the reference's in both forms: per (level, view) a sequential
// float sum over i (x) outer / j (y) inner with the '+30, -30' sequence (clcode.cl:1030-1053);
// min over views and argmin over levels use strict '<' (first view / lowest level wins,
// clcode.cl:1054-1063); the cross-lane argmin breaks ties toward the lower level.
// ---------------------------------------------------------------------------------------
constexpr int BM_WARPS = 8;
constexpr float BM_OUTSIDE = -1.0f;   // staged term of a sample that projects outside (terms are >= 0 or NaN)

__global__ void __launch_bounds__(BM_WARPS * 32, 5)
k_block_match(const float4 *__restrict__ lab, float4 *__restrict__ map, const uint2 *__restrict__ rep,
              const float *__restrict__ levels, int disp_num, const int *__restrict__ view_subset,
              const int *__restrict__ subset_num, int aw, int V, int W, int H, int N, int M, float bl_ratio, int z0,
              int band_blocks)
{
    __shared__ float4 s_ref[BM_WARPS][25];      // Lab of the reference samples
    __shared__ float2 s_xy[BM_WARPS][25];       // their pixel coordinates as floats (exact below 2^24)
    __shared__ float s_t[BM_WARPS][32 * 25];    // per (level, sample) cost terms of the current pair
    const unsigned FULL = 0xffffffffu;
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // grid = (blocks of a band of superpixel rows, views, bands): all views work on the same image band at the
    // same time, so the neighbour views' Lab rows a band reads stay in L2 across the views that share them
    int sp = (blockIdx.z * band_blocks + blockIdx.x) * BM_WARPS + warp;
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
    bool ref_ok = false;
    if (lane < 25) {
        int i = lane / 5 - 2, j = lane % 5 - 2;
        int xr = f2i(cxf + (float)i * stepx);
        int yr = f2i(cyf + (float)j * stepy);
        ref_ok = xr >= 0 && yr >= 0 && xr < W && yr < H;
        s_xy[warp][lane] = ref_ok ? make_float2((float)xr, (float)yr) : make_float2(0.0f, 0.0f);
        s_ref[warp][lane] = ref_ok ? ref_img[(size_t)yr * W + xr] : make_float4(0, 0, 0, 0);
    }
    const unsigned vmask = __ballot_sync(FULL, ref_ok);   // bit s: reference sample s lies inside the image
    __syncwarp();

    int refx = z % aw, refy = z / aw;
    int ns = subset_num[z];
    float cost_est = 1000000.0f, disp_est = 0.0f;
    float *t = s_t[warp];
    for (int base = 0; base < disp_num; base += 32) {
        const int nlev = min(32, disp_num - base);
        int dl = base + lane;
        bool active = dl < disp_num;
        float d = active ? levels[dl] : 0.0f;
        float bd = bl_ratio * d;
        float min_val = 1000000.0f;
        for (int n = 0; n < ns; n++) {
            int view = view_subset[V * z + n];
            float dx = (float)(view % aw - refx), dy = (float)(view / aw - refy);
            const float4 *pimg = lab + (size_t)view * N;
            float val = 0.0f;
            if (dy == 0.0f) {
                float sx = d * dx, sy = bd * dy;
#pragma unroll
                for (int s = 0; s < 25; ++s) {
                    float2 xy = s_xy[warp][s];
                    int xp = f2i(xy.x - sx);
                    int yp = f2i(xy.y - sy);
                    val += 30.0f;
                    if (((vmask >> s) & 1u) && (unsigned)xp < (unsigned)W && (unsigned)yp < (unsigned)H) {
                        val -= 30.0f;
                        float4 cr = s_ref[warp][s];
                        float4 cp = pimg[yp * W + xp];
                        val += fabsf(cr.x - cp.x) + fabsf(cr.y - cp.y) + fabsf(cr.z - cp.z);
                    }
                }
            } else {
                // gather with lanes over (level, x-sample); j = sample row (uniform per pass)
#pragma unroll 1
                for (int j = 0; j < 5; ++j) {
#pragma unroll
                    for (int rr = 0; rr < 5; ++rr) {   // nlev * 5 <= 160 items per sample row
                        const int e = rr * 32 + lane;
                        const bool item = e < nlev * 5;
                        const int l = item ? e / 5 : 0, i = e - 5 * l;
                        const int s = item ? i * 5 + j : 0;
                        const float dd = __shfl_sync(FULL, d, l);
                        const float2 xy = s_xy[warp][s];
                        const int xp = f2i(xy.x - dd * dx);
                        const int yp = f2i(xy.y - (bl_ratio * dd) * dy);
                        float tv = BM_OUTSIDE;
                        if (item && ((vmask >> s) & 1u) && (unsigned)xp < (unsigned)W && (unsigned)yp < (unsigned)H) {
                            const float4 cr = s_ref[warp][s];
                            const float4 cp = pimg[yp * W + xp];
                            tv = fabsf(cr.x - cp.x) + fabsf(cr.y - cp.y) + fabsf(cr.z - cp.z);
                        }
                        if (item) t[l * 25 + s] = tv;
                    }
                }
                __syncwarp();
                if (active) {
#pragma unroll
                    for (int s = 0; s < 25; ++s) {
                        const float tv = t[lane * 25 + s];
                        val += 30.0f;
                        if (!(tv == BM_OUTSIDE)) { val -= 30.0f; val += tv; }
                    }
                }
                __syncwarp();
            }
            if (val < min_val) min_val = val;
        }
        // warp argmin over this chunk of levels, ties -> lower level index
        float bv = active ? min_val : __int_as_float(0x7f800000);
        int bi = dl;
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            float ov = __shfl_xor_sync(FULL, bv, o);
            int oi = __shfl_xor_sync(FULL, bi, o);
            if (ov < bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (bv < cost_est) { cost_est = bv; disp_est = levels[bi]; }
    }
    if (lane == 0) reinterpret_cast<float *>(map + 2 * mi + 1)[3] = disp_est;
}

int launch_block_match(const FrameDev &f, float bl_ratio, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    const int nblocks = (f.M + BM_WARPS - 1) / BM_WARPS;
    int band_blocks = std::max(1, (2 * f.mw + BM_WARPS - 1) / BM_WARPS);   // two superpixel rows per band
    int bands = (nblocks + band_blocks - 1) / band_blocks;
    if (bands > 65535) { band_blocks = (nblocks + 65534) / 65535; bands = (nblocks + band_blocks - 1) / band_blocks; }
    dim3 g(band_blocks, z1 - z0, bands);
    k_block_match<<<g, BM_WARPS * 32, 0, st>>>(f.lab, f.map, f.rep, f.levels, f.num_levels, f.view_subset,
                                               f.subset_num, f.aw, f.V, f.W, f.H, f.N, f.M, bl_ratio, z0, band_blocks);
    return 1;
}

} // namespace mvs
