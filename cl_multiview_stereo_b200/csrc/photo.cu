// photo.cu -- photo-consistency disparity initialisation (sm_100a).
//
// Replaces the OpenCL kernels driven by clPhotoConsistency::do_initial_depth_estimation
// (reference clMVDE/photo_consistency.cpp:21-150): find_super_pixel_boundary
// (clcode.cl:791-855) and initial_depth_estimation_v2 (clcode.cl:972-1069).  The reference
// launches the matcher once per view; here every view is covered by one launch.
#include "common.cuh"
#include "mvs_math.cuh"
#include "tma.cuh"
#include <algorithm>
#include <math.h>
#include <stdlib.h>
#include <string.h>

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
              const int *__restrict__ subset_num, const int2 *__restrict__ cam, int V, int W, int H, int N, int M, float bl_ratio, int z0,
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

    const int2 cz = cam[z];   // position of the view's camera in the array (slot order need not be row-major)
    int refx = cz.x, refy = cz.y;
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
            const int2 cv = cam[view];
            float dx = (float)(cv.x - refx), dy = (float)(cv.y - refy);
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

// ---------------------------------------------------------------------------------------
// Block matching on level-aligned copies of Lab.
//
// With lanes = disparity levels, the 32 gathers of a warp for one sample are one contiguous run of a neighbour in the
// SAME camera row (consecutive levels shift by one pixel along the image row: 4-5 cache lines), but 32 different image
// rows for every other neighbour -- the reason the kernel above regroups lanes and stages terms in shared memory for
// those pairs, and sits on the L1 tag stage (97 % busy).  Here the neighbour image is read from a copy whose memory
// order follows the direction consecutive levels move in:
//   same row     (dy = 0)      lab   [y][x]                row-major, as produced by cvt
//   same column  (dx = 0)      labT  [x][y]                transposed: levels walk down a column
//   diagonal, dx = dy          labSp [x - y + H - 1][y]    sheared + transposed: levels walk (-1, -1) per step
//   diagonal, dx = -dy         labSm [x + y][y]
// (bl_ratio close to 1: the row shift of level d is round-ish(bl * d), so a warp's gathers fall into two or three
// runs; any other bl_ratio is still exact, only less aligned).  Every pair then runs the same loop: pixel address =
// xp * KX + yp * KY + K0 with three per-neighbour constants, lanes = levels, the reference's '+30, -30' chain
// (clcode.cl:1030-1053) per lane, no shared-memory staging of terms.  k_lab_shear makes the three copies once per
// frame (tile through shared memory, runs along y stored coalesced).
// ---------------------------------------------------------------------------------------
constexpr int SH_T = 32;   // tile edge of the copy kernel

__global__ void __launch_bounds__(256) k_lab_shear(const float4 *__restrict__ lab, float4 *__restrict__ labT,
                                                   float4 *__restrict__ labSp, float4 *__restrict__ labSm, int W, int H,
                                                   int N, size_t sh_stride, int z0)
{
    __shared__ float4 tile[SH_T][SH_T + 1];
    const int z = z0 + blockIdx.z;
    const int x0 = blockIdx.x * SH_T, y0 = blockIdx.y * SH_T;
    const int tx = threadIdx.x, ty = threadIdx.y;   // 32 x 8
    const float4 *src = lab + (size_t)z * N;
    for (int yl = ty; yl < SH_T; yl += 8) {
        const int x = x0 + tx, y = y0 + yl;
        if (x < W && y < H) tile[yl][tx] = src[(size_t)y * W + x];
    }
    __syncthreads();
    float4 *dT = labT + (size_t)z * N, *dP = labSp + (size_t)z * sh_stride, *dM = labSm + (size_t)z * sh_stride;
    const int y = y0 + tx;          // lanes run along y in all three copies
    if (y < H) {
        for (int xl = ty; xl < SH_T; xl += 8) {
            const int x = x0 + xl;
            if (x < W) dT[(size_t)x * H + y] = tile[tx][xl];
        }
        for (int d = -(SH_T - 1) + ty; d <= SH_T - 1; d += 8) {          // diagonal x - y = const of the tile
            const int xl = d + tx;
            if (xl >= 0 && xl < SH_T && x0 + xl < W) dP[(size_t)(x0 - y0 + d + H - 1) * H + y] = tile[tx][xl];
        }
        for (int e = ty; e <= 2 * (SH_T - 1); e += 8) {                  // anti-diagonal x + y = const
            const int xl = e - tx;
            if (xl >= 0 && xl < SH_T && x0 + xl < W) dM[(size_t)(x0 + y0 + e) * H + y] = tile[tx][xl];
        }
    }
}

int launch_lab_shear(const FrameDev &f, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    dim3 b(32, 8), g((f.W + SH_T - 1) / SH_T, (f.H + SH_T - 1) / SH_T, z1 - z0);
    k_lab_shear<<<g, b, 0, st>>>(f.lab, f.labT, f.labSp, f.labSm, f.W, f.H, f.N, (size_t)(f.W + f.H - 1) * f.H, z0);
    return 1;
}

#ifndef BS_UNROLL
#define BS_UNROLL 5
#endif
#ifndef BS_MINB
#define BS_MINB 4
#endif
constexpr int BS_WARPS = 8;
constexpr int BS_UN = BS_UNROLL;

__global__ void __launch_bounds__(BS_WARPS * 32, BS_MINB)
k_block_match_sh(const float4 *__restrict__ lab, const float4 *__restrict__ labT, const float4 *__restrict__ labSp,
                 const float4 *__restrict__ labSm, float4 *__restrict__ map, const uint2 *__restrict__ rep,
                 const float *__restrict__ levels, int disp_num, const int *__restrict__ view_subset,
                 const int *__restrict__ subset_num, const int2 *__restrict__ cam, int V, int W, int H, int N, int M, int mw, int mh,
                 float bl_ratio, int z0, int cbw, int bands)
{
    __shared__ float4 s_ref[BS_WARPS][25];      // Lab of the reference samples
    __shared__ float2 s_xy[BS_WARPS][25];       // their pixel coordinates as floats (exact below 2^24)
    const unsigned FULL = 0xffffffffu;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // grid = (blocks of a map tile, views, map tiles).  A tile is cbw superpixel columns x 2 rows; tiles are numbered
    // band-fastest inside a column block, so that consecutive tiles (all views each) read almost the same rows of
    // the same columns of every copy: that set -- a quarter of the image width, tile rows +- the disparity range,
    // all views, four layouts -- fits the L2 (a full-width band does not: 42 % L2 hits, 6.3 GB of DRAM reads)
    const int cb = blockIdx.z / bands, band = blockIdx.z - cb * bands;
    const int t = blockIdx.x * BS_WARPS + warp;                 // superpixel inside the tile, row-major (2 rows)
    const int gx = cb * cbw + t % cbw, gy = 2 * band + t / cbw;
    const int z = z0 + blockIdx.y;
    if (t >= 2 * cbw || gx >= mw || gy >= mh) return;   // whole warp leaves together
    const int sp = gy * mw + gx;
    const size_t mi = (size_t)z * M + sp;
    const uint2 r = rep[mi];
    const int d0 = r.x & 255, d1 = (r.x >> 8) & 255, d2 = (r.x >> 16) & 255, d3 = r.x >> 24;
    const int d4 = r.y & 255, d5 = (r.y >> 8) & 255, d6 = (r.y >> 16) & 255, d7 = r.y >> 24;
    const int bb_l = max(d0, max(d1, d2)), bb_r = max(d5, max(d6, d7));
    const int bb_t = max(d0, max(d3, d5)), bb_b = max(d2, max(d4, d7));
    const float stepx = fmaxf(1.0f, 0.25f * (float)(bb_l + bb_r));
    const float stepy = fmaxf(1.0f, 0.25f * (float)(bb_t + bb_b));
    const float4 m0 = map[2 * mi];
    const float4 *ref_img = lab + (size_t)z * N;
    bool ref_ok = false;
    if (lane < 25) {
        const int i = lane / 5 - 2, j = lane % 5 - 2;
        const int xr = f2i(m0.y + (float)i * stepx);
        const int yr = f2i(m0.z + (float)j * stepy);
        ref_ok = xr >= 0 && yr >= 0 && xr < W && yr < H;
        s_xy[warp][lane] = ref_ok ? make_float2((float)xr, (float)yr) : make_float2(0.0f, 0.0f);
        s_ref[warp][lane] = ref_ok ? ref_img[(size_t)yr * W + xr] : make_float4(0, 0, 0, 0);
    }
    const unsigned vmask = __ballot_sync(FULL, ref_ok);
    __syncwarp();

    const int2 cz = cam[z];
    const int refx = cz.x, refy = cz.y;
    const int ns = subset_num[z];
    const size_t sh_stride = (size_t)(W + H - 1) * H;
    float cost_est = 1000000.0f, disp_est = 0.0f;
    for (int base = 0; base < disp_num; base += 32) {
        const int dl = base + lane;
        const bool active = dl < disp_num;
        const float d = active ? levels[dl] : 0.0f;
        const float bd = bl_ratio * d;
        float min_val = 1000000.0f;
        for (int n = 0; n < ns; n++) {
            const int view = view_subset[V * z + n];
            const int2 cv = cam[view];
            const int ix = cv.x - refx, iy = cv.y - refy;
            const float sx = d * (float)ix, sy = bd * (float)iy;
            // the copy whose memory order follows the level direction of this neighbour, and pixel -> element
            const float4 *img;
            int KX, KY, K0;
            if (iy == 0) { img = lab + (size_t)view * N; KX = 1; KY = W; K0 = 0; }
            else if (ix == 0) { img = labT + (size_t)view * N; KX = H; KY = 1; K0 = 0; }
            else if ((ix > 0) == (iy > 0)) { img = labSp + (size_t)view * sh_stride; KX = H; KY = 1 - H; K0 = (H - 1) * H; }
            else { img = labSm + (size_t)view * sh_stride; KX = H; KY = 1 + H; K0 = 0; }
            float val = 0.0f;
#pragma unroll BS_UN
            for (int s = 0; s < 25; ++s) {
                const float2 xy = s_xy[warp][s];
                const int xp = f2i(xy.x - sx);
                const int yp = f2i(xy.y - sy);
                val += 30.0f;
                if (((vmask >> s) & 1u) && (unsigned)xp < (unsigned)W && (unsigned)yp < (unsigned)H) {
                    val -= 30.0f;
                    const float4 cr = s_ref[warp][s];
                    const float4 cp = img[xp * KX + yp * KY + K0];
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
            const float ov = __shfl_xor_sync(FULL, bv, o);
            const int oi = __shfl_xor_sync(FULL, bi, o);
            if (ov < bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (bv < cost_est) { cost_est = bv; disp_est = levels[bi]; }
    }
    if (lane == 0) reinterpret_cast<float *>(map + 2 * mi + 1)[3] = disp_est;
}

// ---------------------------------------------------------------------------------------
// Block matching out of TMA-staged windows.
//
// The gather kernel above pushes every one of its 1.9e9 sixteen-byte samples per frame through the L1 tag stage
// (L1 97 % busy, 33 % hits).  Here a CTA owns a TILE of BT_W x BT_H adjacent superpixels of one view.  For one
// neighbour view at a time, everything the tile can sample there -- the bounding box of its reference samples,
// displaced by the whole disparity range along the camera offset -- is a dense box of that view's Lab image: ONE
// cp.async.bulk.tensor box load (TMA) brings it into shared memory, completion on an mbarrier, and the (level,
// sample) loops of clcode.cl:1030-1053 then read shared memory with lanes = disparity levels.  The box width is
// odd for a neighbour in the same camera column and even for a diagonal one, so that the 16-byte reads of the
// eight lanes of a quarter-warp -- one image row apart, or one row and one column -- fall on distinct bank groups.
// A sample that lands outside the staged box (a centre that has drifted unusually far) takes the global load.
// The cost chain ('+30, -30', x outer / y inner), the minimum over views and the argmin over levels are those of
// the kernel above; results are bit-identical.
// ---------------------------------------------------------------------------------------
constexpr int BT_W = 8, BT_H = 4;            // tile in superpixels
constexpr int BT_SP = BT_W * BT_H;
constexpr int BT_WARPS = 16;                 // two superpixels per warp
static_assert(BT_SP == 2 * BT_WARPS, "two superpixels per warp");

struct BmTmaArgs {
    const float4 *lab;
    float4 *map;
    const uint2 *rep;
    const float *levels;
    const int *view_subset, *subset_num;
    const int2 *cam;
    int disp_num, V, W, H, N, M, mw, mh, z0;
    float bl_ratio, dmin, dmax;
    int bw[3], bh[3];
};

__global__ void __launch_bounds__(BT_WARPS * 32, 1)
k_block_match_tma(const __grid_constant__ CUtensorMap tm_row, const __grid_constant__ CUtensorMap tm_col,
                  const __grid_constant__ CUtensorMap tm_diag, const BmTmaArgs a)
{
    extern __shared__ __align__(128) unsigned char bt_smem[];
    __shared__ float4 s_ref[BT_SP][25];
    __shared__ float2 s_xy[BT_SP][25];
    __shared__ int s_box[4];                 // xmin, xmax, ymin, ymax of the tile's reference samples
    __shared__ __align__(8) uint64_t s_bar;
    float4 *win = reinterpret_cast<float4 *>(bt_smem);
    const unsigned FULL = 0xffffffffu;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // grid = (tile columns, views, tile rows): the views of one band of rows run together (L2 reuse of the rows a
    // band reads from its neighbour views)
    const int z = a.z0 + blockIdx.y;
    if (threadIdx.x == 0) {
        s_box[0] = 0x7fffffff; s_box[1] = -0x7fffffff; s_box[2] = 0x7fffffff; s_box[3] = -0x7fffffff;
        mbar_init(&s_bar, 1);
        mbar_fence_init();
        tma_prefetch_desc(&tm_row); tma_prefetch_desc(&tm_col); tma_prefetch_desc(&tm_diag);
    }
    __syncthreads();

    // ---- reference samples of the warp's two superpixels ----
    const float4 *ref_img = a.lab + (size_t)z * a.N;
    unsigned vmask[2];
    bool live[2];
    size_t mi[2];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        const int t = 2 * warp + q;
        const int gx = blockIdx.x * BT_W + t % BT_W, gy = blockIdx.z * BT_H + t / BT_W;
        live[q] = gx < a.mw && gy < a.mh;
        mi[q] = (size_t)z * a.M + (live[q] ? gy * a.mw + gx : 0);
        bool ref_ok = false;
        int xr = 0, yr = 0;
        if (live[q]) {
            const uint2 r = a.rep[mi[q]];
            const int d0 = r.x & 255, d1 = (r.x >> 8) & 255, d2 = (r.x >> 16) & 255, d3 = r.x >> 24;
            const int d4 = r.y & 255, d5 = (r.y >> 8) & 255, d6 = (r.y >> 16) & 255, d7 = r.y >> 24;
            const int bb_l = max(d0, max(d1, d2)), bb_r = max(d5, max(d6, d7));
            const int bb_t = max(d0, max(d3, d5)), bb_b = max(d2, max(d4, d7));
            const float stepx = fmaxf(1.0f, 0.25f * (float)(bb_l + bb_r));
            const float stepy = fmaxf(1.0f, 0.25f * (float)(bb_t + bb_b));
            const float4 m0 = a.map[2 * mi[q]];
            if (lane < 25) {
                const int i = lane / 5 - 2, j = lane % 5 - 2;
                xr = f2i(m0.y + (float)i * stepx);
                yr = f2i(m0.z + (float)j * stepy);
                ref_ok = xr >= 0 && yr >= 0 && xr < a.W && yr < a.H;
                s_xy[t][lane] = ref_ok ? make_float2((float)xr, (float)yr) : make_float2(0.0f, 0.0f);
                s_ref[t][lane] = ref_ok ? ref_img[(size_t)yr * a.W + xr] : make_float4(0, 0, 0, 0);
            }
        }
        vmask[q] = __ballot_sync(FULL, ref_ok);
        const int lo_x = __reduce_min_sync(FULL, ref_ok ? xr : 0x7fffffff), hi_x = __reduce_max_sync(FULL, ref_ok ? xr : -0x7fffffff);
        const int lo_y = __reduce_min_sync(FULL, ref_ok ? yr : 0x7fffffff), hi_y = __reduce_max_sync(FULL, ref_ok ? yr : -0x7fffffff);
        if (lane == 0 && vmask[q]) {
            atomicMin(&s_box[0], lo_x); atomicMax(&s_box[1], hi_x);
            atomicMin(&s_box[2], lo_y); atomicMax(&s_box[3], hi_y);
        }
    }
    __syncthreads();
    const int bx_lo = s_box[0], by_lo = s_box[2];
    const bool any_ref = s_box[1] >= s_box[0];

    const int2 cz = a.cam[z];
    const int refx = cz.x, refy = cz.y;
    const int ns = a.subset_num[z];
    const bool active = lane < a.disp_num;
    const float d = active ? a.levels[lane] : 0.0f;
    const float bd = a.bl_ratio * d;
    float min_val[2] = {1000000.0f, 1000000.0f};

    for (int n = 0; n < ns; ++n) {
        const int view = a.view_subset[a.V * z + n];
        const int2 cv = a.cam[view];
        const int idx_x = cv.x - refx, idx_y = cv.y - refy;
        const float dx = (float)idx_x, dy = (float)idx_y;
        const int cls = idx_y == 0 ? 0 : (idx_x == 0 ? 1 : 2);
        const int BW = cls == 0 ? a.bw[0] : (cls == 1 ? a.bw[1] : a.bw[2]);
        const int BH = cls == 0 ? a.bh[0] : (cls == 1 ? a.bh[1] : a.bh[2]);
        // window origin: lowest pixel any (sample, level) of the tile can project to (one pixel of slack for the
        // truncations); samples beyond origin + box take the global load
        const float sx_hi = fmaxf(a.dmin * dx, a.dmax * dx), sy_hi = fmaxf(a.bl_ratio * a.dmin * dy, a.bl_ratio * a.dmax * dy);
        const int x0 = any_ref ? bx_lo - (int)ceilf(sx_hi) - 1 : 0, y0 = any_ref ? by_lo - (int)ceilf(sy_hi) - 1 : 0;
        __syncthreads();                                  // every warp is done with the previous window
        if (threadIdx.x == 0) {
            fence_proxy_async();
            mbar_arrive_expect_tx(&s_bar, (uint32_t)(BW * BH * 16));
            const CUtensorMap *tm = cls == 0 ? &tm_row : (cls == 1 ? &tm_col : &tm_diag);
            tma_load_2d(win, tm, 2 * x0, y0 + view * a.H, &s_bar);   // x in 8-byte elements; rows of all views stacked
        }
        mbar_wait(&s_bar, (uint32_t)(n & 1));
        const float4 *pimg = a.lab + (size_t)view * a.N;
        const float sx = d * dx, sy = bd * dy;
        float val[2] = {0.0f, 0.0f};
#pragma unroll 5
        for (int s = 0; s < 25; ++s) {
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const int t = 2 * warp + q;
                const float2 xy = s_xy[t][s];
                const int xp = f2i(xy.x - sx);
                const int yp = f2i(xy.y - sy);
                val[q] += 30.0f;
                if (((vmask[q] >> s) & 1u) && (unsigned)xp < (unsigned)a.W && (unsigned)yp < (unsigned)a.H) {
                    val[q] -= 30.0f;
                    const float4 cr = s_ref[t][s];
                    const int wx = xp - x0, wy = yp - y0;
                    float4 cp;
                    if ((unsigned)wx < (unsigned)BW && (unsigned)wy < (unsigned)BH) cp = win[wy * BW + wx];
                    else cp = pimg[yp * a.W + xp];
                    val[q] += fabsf(cr.x - cp.x) + fabsf(cr.y - cp.y) + fabsf(cr.z - cp.z);
                }
            }
        }
#pragma unroll
        for (int q = 0; q < 2; ++q)
            if (val[q] < min_val[q]) min_val[q] = val[q];
    }

    // warp argmin over the levels, ties -> lower level index (strict '<' of clcode.cl:1059-1063)
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        float bv = active ? min_val[q] : __int_as_float(0x7f800000);
        int bi = lane;
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            const float ov = __shfl_xor_sync(FULL, bv, o);
            const int oi = __shfl_xor_sync(FULL, bi, o);
            if (ov < bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        float disp_est = 0.0f;
        if (bv < 1000000.0f) disp_est = a.levels[bi];
        if (lane == 0 && live[q]) reinterpret_cast<float *>(a.map + 2 * mi[q] + 1)[3] = disp_est;
    }
}

// box sizes of the three neighbour classes for superpixel size S and the disparity range; false when the tile's
// windows do not fit a TMA box / shared memory (the gather kernel serves those geometries)
static bool bm_tma_boxes(int S, float dmin, float dmax, float bl_ratio, int bw[3], int bh[3], size_t *smem)
{
    if (!(dmax >= dmin) || !isfinite(dmin) || !isfinite(dmax) || !isfinite(bl_ratio)) return false;
    const int base_w = BT_W * S + 3 * S, base_h = BT_H * S + 3 * S;   // tile + sample reach (S) + drift (S / 2) on both sides
    const double rx = ceil((double)dmax - dmin) + 3, ry = ceil(fabs((double)bl_ratio) * ((double)dmax - dmin)) + 3;
    if (rx > 200 || ry > 200) return false;
    bw[0] = base_w + (int)rx;           bh[0] = base_h;
    bw[1] = base_w | 1;                 bh[1] = base_h + (int)ry;     // odd pitch: rows one apart -> distinct bank groups
    bw[2] = (base_w + (int)rx + 1) & ~1; bh[2] = base_h + (int)ry;    // even pitch: (row, column) one apart -> distinct
    size_t mx = 0;
    for (int c = 0; c < 3; ++c) {
        if (bw[c] > 128 || bh[c] > 256) return false;
        mx = std::max(mx, (size_t)bw[c] * bh[c] * 16);
    }
    *smem = mx;
    return mx <= 190 * 1024;
}

static int launch_block_match_tma(const FrameDev &f, float bl_ratio, int z0, int z1, cudaStream_t st)
{
    BmTma *t = f.bm_tma;
    if (!t || f.num_levels < 1 || f.num_levels > 32 || !f.levels) return 0;
    // level range (host copy of the levels is not kept here: the launcher is given them through the frame's cache)
    if (t->state == 0 || t->lab != (const void *)f.lab || t->bl_ratio != bl_ratio) return 0;
    if (t->state < 0) return 0;
    BmTmaArgs a;
    a.lab = f.lab; a.map = f.map; a.rep = f.rep; a.levels = f.levels; a.view_subset = f.view_subset; a.subset_num = f.subset_num;
    a.disp_num = f.num_levels; a.cam = f.cam; a.V = f.V; a.W = f.W; a.H = f.H; a.N = f.N; a.M = f.M; a.mw = f.mw; a.mh = f.mh;
    a.z0 = z0; a.bl_ratio = bl_ratio; a.dmin = t->dmin; a.dmax = t->dmax;
    size_t smem = 0;
    for (int c = 0; c < 3; ++c) { a.bw[c] = t->bw[c]; a.bh[c] = t->bh[c]; smem = std::max(smem, (size_t)t->bw[c] * t->bh[c] * 16); }
    cudaFuncSetAttribute(k_block_match_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    dim3 g((f.mw + BT_W - 1) / BT_W, z1 - z0, (f.mh + BT_H - 1) / BT_H);
    k_block_match_tma<<<g, BT_WARPS * 32, smem, st>>>(t->map[0], t->map[1], t->map[2], a);
    return 1;
}

// (re)build the tensor maps of the frame for this Lab buffer / level set / bl_ratio; host levels are passed in
int prepare_block_match_tma(const FrameDev &f, const float *levels_host, int nlev, float bl_ratio, const char **why)
{
    static const char *none = "";
    *why = none;
    BmTma *t = f.bm_tma;
    if (!t) return 0;
    float dmin = 0, dmax = 0;
    for (int i = 0; i < nlev; ++i) {
        dmin = i ? std::min(dmin, levels_host[i]) : levels_host[i];
        dmax = i ? std::max(dmax, levels_host[i]) : levels_host[i];
    }
    if (t->state != 0 && t->lab == (const void *)f.lab && t->dmin == dmin && t->dmax == dmax && t->bl_ratio == bl_ratio &&
        t->nlev == nlev)
        return 0;
    t->lab = f.lab; t->dmin = dmin; t->dmax = dmax; t->bl_ratio = bl_ratio; t->nlev = nlev;
    t->state = -1;
    size_t smem = 0;
    if (nlev < 1 || nlev > 32 || (f.W * 16) % 16 || !bm_tma_boxes(f.S, dmin, dmax, bl_ratio, t->bw, t->bh, &smem)) return 0;
    if ((double)f.V * f.H >= 2147483647.0) return 0;
    for (int c = 0; c < 3; ++c)
        if (tma_make_image_map(&t->map[c], f.lab, 16, f.W, f.H, f.V, t->bw[c], t->bh[c], why)) return 1;
    t->state = 1;
    return 0;
}

int launch_block_match(const FrameDev &f, float bl_ratio, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    // The TMA-window kernel is bit-identical but measured slower on the bench frames (6.8 ms against 3.8 ms per
    // 15-view 1080p frame: one resident CTA per SM, window fill and compute strictly in turn); it runs on request
    // (MVS_BM_MODE=tma) so that the comparison stays reproducible.  DESIGN.md section 4.1b has the numbers.
    static const bool want_tma = [] { const char *e = getenv("MVS_BM_MODE"); return e && !strcmp(e, "tma"); }();
    if (want_tma && launch_block_match_tma(f, bl_ratio, z0, z1, st)) return 1;
    const int nblocks = (f.M + BM_WARPS - 1) / BM_WARPS;
    // level-aligned copies of Lab present (made by launch_lab_shear for this frame): the lanes = levels kernel for all pairs
    static const bool want_gather = [] { const char *e = getenv("MVS_BM_MODE"); return e && !strcmp(e, "gather"); }();
    if (!want_gather && f.labT && f.labSp && f.labSm && (double)(f.W + f.H) * f.H < 2147483647.0) {
        // column blocks of about a quarter of the map width (multiples of 8 superpixels), bands of two map rows
        int cbw = std::max(8, ((f.mw + 3) / 4 + 7) / 8 * 8);
        const int bands = (f.mh + 1) / 2;
        int ncb = (f.mw + cbw - 1) / cbw;
        if ((long long)ncb * bands > 65535) { cbw = f.mw; ncb = 1; }
        if (bands <= 65535) {
            dim3 g((2 * cbw + BS_WARPS - 1) / BS_WARPS, z1 - z0, ncb * bands);
            k_block_match_sh<<<g, BS_WARPS * 32, 0, st>>>(f.lab, f.labT, f.labSp, f.labSm, f.map, f.rep, f.levels, f.num_levels,
                                                          f.view_subset, f.subset_num, f.cam, f.V, f.W, f.H, f.N, f.M, f.mw, f.mh,
                                                          bl_ratio, z0, cbw, bands);
            return 1;
        }
    }
    int band_blocks = std::max(1, (2 * f.mw + BM_WARPS - 1) / BM_WARPS);   // two superpixel rows per band
    int bands = (nblocks + band_blocks - 1) / band_blocks;
    if (bands > 65535) { band_blocks = (nblocks + 65534) / 65535; bands = (nblocks + band_blocks - 1) / band_blocks; }
    dim3 g(band_blocks, z1 - z0, bands);
    k_block_match<<<g, BM_WARPS * 32, 0, st>>>(f.lab, f.map, f.rep, f.levels, f.num_levels, f.view_subset,
                                               f.subset_num, f.cam, f.V, f.W, f.H, f.N, f.M, bl_ratio, z0, band_blocks);
    return 1;
}

} // namespace mvs
