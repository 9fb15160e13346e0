// slic.cu -- superpixel segmentation kernels (sm_100a).
//
// Replaces the OpenCL kernels driven by clSLIC::do_super_pixel_seg
// (reference clMVDE/clSLIC.cpp:67-122): cvt (clcode.cl:125-151), init_cluster_centers
// (:259-294), find_center_association (:447-520), update_cluster_center +
// finalize_reduction_result (:533-611, :719-773), supress_local_lable (:676-711).
// All views of a frame go through one launch (blockIdx.z / .y = view).
#include "common.cuh"
#include <algorithm>
#include "mvs_math.cuh"
#include "tma.cuh"
#include <stdlib.h>
#include <string.h>

namespace mvs {

// ---------------------------------------------------------------------------------------
// RGB -> CIELAB.  The loader stores s0=R,s1=G,s2=B (file_handler.cpp:11-13) while the
// conversion reads s0 as blue and s2 as red (clcode.cl:25-27); kept as is.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ float lab_f(float t)
{
    return t > 0.008856f ? cbrtf_pos(t) : (903.3f * t + 16.0f) / 116.0f;
}

__device__ __forceinline__ float4 rgb2lab(uchar4 p)
{
    float b = (float)p.x * 0.0039216f;
    float g = (float)p.y * 0.0039216f;
    float r = (float)p.z * 0.0039216f;
    float x = r * 0.412453f + g * 0.357580f + b * 0.180423f;
    float y = r * 0.212671f + g * 0.715160f + b * 0.072169f;
    float z = r * 0.019334f + g * 0.119193f + b * 0.950227f;
    float fx = lab_f(x / 0.950456f);
    float fy = lab_f(y / 1.0f);
    float fz = lab_f(z / 1.088754f);
    return make_float4(116.0f * fy - 16.0f, 500.0f * (fx - fy), 200.0f * (fy - fz), 0.0f);
}

__global__ void __launch_bounds__(256) k_cvt(const uchar4 *__restrict__ rgbx, float4 *__restrict__ lab, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) lab[i] = rgb2lab(rgbx[i]);
}

int launch_cvt(const FrameDev &f, int z0, int z1, cudaStream_t st)
{
    size_t n = (size_t)(z1 - z0) * f.N;
    if (!n) return 0;
    k_cvt<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(f.rgbx + (size_t)z0 * f.N, f.lab + (size_t)z0 * f.N, n);
    return 1;
}

// ---------------------------------------------------------------------------------------
// Cluster-centre initialisation: cell midpoint, '>' edge test (clcode.cl:273-277).  The
// reference can read one past the image when W mod S == S/2; the read is clamped here.
// s7 (left unwritten by the reference) is set to 0.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_init_centers(const float4 *__restrict__ lab, float4 *__restrict__ map,
                                                      int W, int H, int mw, int mh, int S, int N, int M, int z0)
{
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    int z = z0 + blockIdx.y;
    if (c >= M) return;
    int col = c % mw, row = c / mw;
    int cx = col * S + S / 2, cy = row * S + S / 2;
    if (cx > W) cx = (col * S + W) / 2;
    if (cy > H) cy = (row * S + H) / 2;
    int rx = min(cx, W - 1), ry = min(cy, H - 1);
    float4 p = lab[(size_t)z * N + (size_t)ry * W + rx];
    float4 *m = map + 2 * ((size_t)z * M + c);
    m[0] = make_float4((float)c, (float)cx, (float)cy, p.x);
    m[1] = make_float4(p.y, p.z, 0.0f, 0.0f);
}

int launch_init_centers(const FrameDev &f, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    dim3 g((f.M + 255) / 256, z1 - z0);
    k_init_centers<<<g, 256, 0, st>>>(f.lab, f.map, f.W, f.H, f.mw, f.mh, f.S, f.N, f.M, z0);
    return 1;
}

// ---------------------------------------------------------------------------------------
// Edge-aware centre placement (edge_enable; clSLIC::apply_edge_values, clSLIC.cpp:186-233).
// edge_compute_alternative (clcode.cl:161-195) as pinned in DESIGN.md section 3 and oracle.c: the Sobel magnitude
// of Lab is computed from the values cvt wrote (every window reads a snapshot: the kernel here reads `lab` and
// writes `edge`, two different buffers) and stored in the edge image the reference's parameter list names.  The
// 3x3 window is indexed as the reference indexes it: nine entries INCLUDING the centre, so entry 4 of the
// formulas is the centre pixel and entry 5 the right-hand neighbour (clcode.cl:172-192).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_edge_sobel(const float4 *__restrict__ lab, float *__restrict__ edge, int W, int H,
                                                    int N, int z0)
{
    const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
    const int z = z0 + blockIdx.z;
    if (x >= W || y >= H) return;
    const float4 *vlab = lab + (size_t)z * N;
    float4 c[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        const int cx = max(0, min(W - 1, x + k % 3 - 1)), cy = max(0, min(H - 1, y + k / 3 - 1));   // clcode.cl:178
        c[k] = vlab[(size_t)cy * W + cx];
    }
    const float cx_[9] = {c[0].x, c[1].x, c[2].x, c[3].x, c[4].x, c[5].x, c[6].x, c[7].x, c[8].x};
    const float cy_[9] = {c[0].y, c[1].y, c[2].y, c[3].y, c[4].y, c[5].y, c[6].y, c[7].y, c[8].y};
    const float cz_[9] = {c[0].z, c[1].z, c[2].z, c[3].z, c[4].z, c[5].z, c[6].z, c[7].z, c[8].z};
    float v[3];
    const float *ch[3] = {cx_, cy_, cz_};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float *q = ch[k];
        const float dx = -1 * q[0] + q[2] - 2 * q[3] + 2 * q[4] - q[5] + q[7];   // clcode.cl:190
        const float dy = -1 * q[0] - 2 * q[1] - q[2] + q[5] + 2 * q[6] + q[7];   // clcode.cl:191
        v[k] = dx * dx + dy * dy;
    }
    edge[(size_t)z * N + (size_t)y * W + x] = sqrtf(v[0] * 1.0f + v[1] * 1.0f + v[2] * 1.0f);   // dot(v, (1,1,1))
}

// apply_edge_alternative (clcode.cl:204-248): each centre moves to the 8-neighbour of its truncated position with the
// smallest edge value (strict '<'; scan order W, NW, N, NE, E, SE, S, SW) and takes that pixel's colour
__global__ void __launch_bounds__(256) k_apply_edge(const float4 *__restrict__ lab, const float *__restrict__ edge,
                                                    float4 *__restrict__ map, int W, int H, int N, int M, int z0)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    const int z = z0 + blockIdx.y;
    if (c >= M) return;
    float4 *m = map + 2 * ((size_t)z * M + c);
    const float4 m0 = m[0];
    const int cx = f2i(m0.y), cy = f2i(m0.z);
    const float *ve = edge + (size_t)z * N;
    // the reference reads edge_img at the centre without a bounds test; a centre is inside the image by construction
    // (init_cluster_centers), the clamp only guards the W mod S == S / 2 corner case (Appendix B S4)
    float edge_val = ve[(size_t)min(max(cy, 0), H - 1) * W + min(max(cx, 0), W - 1)];
    const int dx[8] = {-1, -1, 0, 1, 1, 1, 0, -1}, dy[8] = {0, -1, -1, -1, 0, 1, 1, 1};
    bool changed = false;
    int c2x = 0, c2y = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int nx = cx + dx[i], ny = cy + dy[i];
        if (nx >= 0 && ny >= 0 && nx < W && ny < H) {
            const float ne = ve[(size_t)ny * W + nx];
            if (ne < edge_val) { edge_val = ne; c2x = nx; c2y = ny; changed = true; }
        }
    }
    if (changed) {
        const float4 col = lab[(size_t)z * N + (size_t)c2y * W + c2x];
        const float4 m1 = m[1];
        m[0] = make_float4(m0.x, (float)c2x, (float)c2y, col.x);
        m[1] = make_float4(col.y, col.z, m1.z, m1.w);
    }
}

int launch_apply_edge_values(const FrameDev &f, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    dim3 b(32, 8), g((f.W + 31) / 32, (f.H + 7) / 8, z1 - z0);
    k_edge_sobel<<<g, b, 0, st>>>(f.lab, f.edge, f.W, f.H, f.N, z0);
    dim3 g2((f.M + 255) / 256, z1 - z0);
    k_apply_edge<<<g2, 256, 0, st>>>(f.lab, f.edge, f.map, f.W, f.H, f.N, f.M, z0);
    return 2;
}

// ---------------------------------------------------------------------------------------
// Pixel -> centre association.  Live rule of the reference: 2x2 candidate cells chosen
// by the half-cell phase, the outer loop variable (built from the x phase) offsets the
// cell ROW and the inner one (y phase) the cell COLUMN (clcode.cl:475-479); strict '<'
// so the first minimum wins; distance = sqrt(dc*Nc + w*(ds*Ns)) (clcode.cl:425-435).
//
// A CTA covers a TX x TY pixel tile; the (TX/S+2) x (TY/S+2) centres it can touch are
// staged once in shared memory as {cx,cy,L,a,b}.
// ---------------------------------------------------------------------------------------
constexpr int AS_TX = 64, AS_TY = 8;   // tile in pixels (thread block = 64x8/ (4 px per thread in x))
constexpr int AS_PX = 4;               // pixels per thread (one STG.128 of labels)

template <int S_T>
__global__ void __launch_bounds__(AS_TX / AS_PX * AS_TY)
k_assoc(const float4 *__restrict__ lab, const float4 *__restrict__ map, uint32_t *__restrict__ idx, int W, int H,
        int mw, int mh, int S_rt, int N, int M, int z0, float max_xy_dist, float max_color_dist, float weight)
{
    const int S = S_T > 0 ? S_T : S_rt;
    extern __shared__ float s_ctr[];   // [cells][5]
    int z = z0 + blockIdx.z;
    int tx0 = blockIdx.x * AS_TX, ty0 = blockIdx.y * AS_TY;
    // candidate cell range of the tile: cells (tx0/S - 1 .. (tx0+TX-1)/S + 1), same in y
    int ccx0 = tx0 / S - 1, ccy0 = ty0 / S - 1;
    int ncx = (tx0 + AS_TX - 1) / S + 1 - ccx0 + 1;
    int ncy = (ty0 + AS_TY - 1) / S + 1 - ccy0 + 1;
    const float4 *vmap = map + 2 * (size_t)z * M;
    int tid = threadIdx.y * blockDim.x + threadIdx.x;
    for (int k = tid; k < ncx * ncy; k += blockDim.x * blockDim.y) {
        int gx = ccx0 + k % ncx, gy = ccy0 + k / ncx;
        float4 a = make_float4(0, 0, 0, 0), b = make_float4(0, 0, 0, 0);
        if (gx >= 0 && gy >= 0 && gx < mw && gy < mh) {
            a = vmap[2 * (gy * mw + gx)];
            b = vmap[2 * (gy * mw + gx) + 1];
        }
        float *s = s_ctr + 5 * k;
        s[0] = a.y; s[1] = a.z; s[2] = a.w; s[3] = b.x; s[4] = b.y;
    }
    __syncthreads();

    int row = ty0 + threadIdx.y;
    int colb = tx0 + threadIdx.x * AS_PX;
    if (row >= H || colb >= W) return;
    const float4 *vlab = lab + (size_t)z * N + (size_t)row * W;
    uint32_t out[AS_PX];
    int cluster_y = row / S;
    int deltaY = (row + S / 2) / S - cluster_y;
    if (S_T > 0 && S_T % (2 * AS_PX) == 0) {
        // the AS_PX pixels of a thread lie in one half-cell (colb is a multiple of AS_PX, S/2 too): they share
        // the four candidate cells, which are read from shared memory once and kept in registers
        const int cluster_x = colb / S;
        const int deltaX = (colb + S / 2) / S - cluster_x;
        float cc[4][5], dy2[4];
        int cid[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int i = (k >> 1) - 1, j = (k & 1) - 1;   // i outer, j inner (clcode.cl:475-479)
            const int cix = cluster_x + j + deltaY, ciy = cluster_y + i + deltaX;
            const bool ok = cix >= 0 && ciy >= 0 && cix < mw && ciy < mh;
            cid[k] = ok ? ciy * mw + cix : -1;
            const float *c = s_ctr + 5 * (ok ? (ciy - ccy0) * ncx + (cix - ccx0) : 0);
#pragma unroll
            for (int q = 0; q < 5; ++q) cc[k][q] = c[q];
            const float dy = (float)row - cc[k][1];
            dy2[k] = dy * dy;
        }
        float4 px[AS_PX];
#pragma unroll
        for (int p = 0; p < AS_PX; ++p) px[p] = colb + p < W ? vlab[colb + p] : make_float4(0, 0, 0, 0);
#pragma unroll
        for (int p = 0; p < AS_PX; ++p) {
            const float fcol = (float)(colb + p);
            float min_dist = 999999.9999f;
            int min_id = -1;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float dL = px[p].x - cc[k][2], da = px[p].y - cc[k][3], db = px[p].z - cc[k][4];
                const float color_distance = dL * dL + da * da + db * db;
                const float dx = fcol - cc[k][0];
                const float space_distance = dx * dx + dy2[k];
                const float dist = sqrtf((color_distance * max_color_dist) + weight * (space_distance * max_xy_dist));
                if (cid[k] >= 0 && dist < min_dist) { min_dist = dist; min_id = cid[k]; }
            }
            out[p] = (uint32_t)min_id;
        }
    } else {
#pragma unroll
    for (int p = 0; p < AS_PX; ++p) {
        int col = colb + p;
        out[p] = 0;
        if (col >= W) continue;
        float4 px = vlab[col];
        int cluster_x = col / S;
        int deltaX = (col + S / 2) / S - cluster_x;
        float min_dist = 999999.9999f;
        int min_id = -1;
#pragma unroll
        for (int i = -1; i <= 0; i++)
#pragma unroll
            for (int j = -1; j <= 0; j++) {
                int cix = cluster_x + j + deltaY;   // column offset from the y phase
                int ciy = cluster_y + i + deltaX;   // row offset from the x phase
                if (cix >= 0 && ciy >= 0 && cix < mw && ciy < mh) {
                    const float *c = s_ctr + 5 * ((ciy - ccy0) * ncx + (cix - ccx0));
                    float dL = px.x - c[2], da = px.y - c[3], db = px.z - c[4];
                    float color_distance = dL * dL + da * da + db * db;
                    float dx = (float)col - c[0], dy = (float)row - c[1];
                    float space_distance = dx * dx + dy * dy;
                    float dist = sqrtf((color_distance * max_color_dist) + weight * (space_distance * max_xy_dist));
                    if (dist < min_dist) { min_dist = dist; min_id = ciy * mw + cix; }
                }
            }
        out[p] = (uint32_t)min_id;   // -1 -> 0xFFFFFFFF (only when no candidate cell exists)
    }
    }
    uint32_t *o = idx + (size_t)z * N + (size_t)row * W + colb;
    if (colb + AS_PX <= W && ((W & 3) == 0)) {
        *reinterpret_cast<uint4 *>(o) = make_uint4(out[0], out[1], out[2], out[3]);
    } else {
#pragma unroll
        for (int p = 0; p < AS_PX; ++p)
            if (colb + p < W) o[p] = out[p];
    }
}

int launch_assoc(const FrameDev &f, const SlicArgs &a, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    dim3 b(AS_TX / AS_PX, AS_TY);
    dim3 g((f.W + AS_TX - 1) / AS_TX, (f.H + AS_TY - 1) / AS_TY, z1 - z0);
    int ncx = (AS_TX - 1) / f.S + 4, ncy = (AS_TY - 1) / f.S + 4;
    size_t sm = (size_t)ncx * ncy * 5 * sizeof(float);
    if (f.S == 8)
        k_assoc<8><<<g, b, sm, st>>>(f.lab, f.map, f.idx, f.W, f.H, f.mw, f.mh, f.S, f.N, f.M, z0, a.max_xy_dist,
                                     a.max_color_dist, a.weight);
    else if (f.S == 16)
        k_assoc<16><<<g, b, sm, st>>>(f.lab, f.map, f.idx, f.W, f.H, f.mw, f.mh, f.S, f.N, f.M, z0, a.max_xy_dist,
                                      a.max_color_dist, a.weight);
    else
        k_assoc<0><<<g, b, sm, st>>>(f.lab, f.map, f.idx, f.W, f.H, f.mw, f.mh, f.S, f.N, f.M, z0, a.max_xy_dist,
                                     a.max_color_dist, a.weight);
    return 1;
}

// ---------------------------------------------------------------------------------------
// Centre update.  The reference runs one 16x16 work-group per (cluster, window tile g),
// reduces 256 local-memory slots with the tree i = 128,64,...,1 and then adds the G
// partials in g order (clcode.cl:584-596, 741-754).  The association of that float sum is
// reproduced exactly by ONE WARP per cluster: lane l owns the eight slots l + 32k
// (k = 0..7) of a window tile, folds them in registers in the order of tree steps 128, 64,
// 32 and finishes steps 16..1 with shuffles.  Position and count sums are exact integers
// (< 2^24), so they use integer warp reductions.  Window geometry, including the
// under-coverage at S = 8 (cluster_per_line = 3S/16 = 1), follows clSLIC.cpp:22-23, 313
// and clcode.cl:552-566.  finalize_reduction_result is fused: lane 0 divides and stores.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_update_centers(const float4 *__restrict__ lab,
                                                        const uint32_t *__restrict__ idx, float4 *__restrict__ map,
                                                        int W, int H, int mw, int mh, int S, int N, int M, int G,
                                                        int cpl, int z0)
{
    int lane = threadIdx.x & 31;
    int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    int z = z0 + blockIdx.y;
    if (c >= M) return;
    int gx = c % mw, gy = c / mw;
    const uint32_t *vidx = idx + (size_t)z * N;
    const float4 *vlab = lab + (size_t)z * N;
    int lx = lane & 15, lyh = lane >> 4;
    float c0 = 0.0f, c1 = 0.0f, c2 = 0.0f;
    int sx = 0, sy = 0, sn = 0;
    for (int g = 0; g < G; ++g) {
        int nbr_x = g % cpl, nbr_y = g / cpl;
        int pxo = nbr_x * 16 + lx;
        if (nbr_y * 16 >= S * 3) break;   // every slot of this and later tiles is empty: partial = +0
        int px = gx * S - S + pxo;
        bool colok = pxo < S * 3 && px >= 0 && px < W;
        float v0[8], v1[8], v2[8];
        int tx = 0, ty = 0, tn = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            int pyo = nbr_y * 16 + 2 * k + lyh;
            int py = gy * S - S + pyo;
            v0[k] = 0.0f; v1[k] = 0.0f; v2[k] = 0.0f;
            if (colok && pyo < S * 3 && py >= 0 && py < H) {
                size_t pi = (size_t)py * W + px;
                if (vidx[pi] == (uint32_t)c) {
                    float4 p = vlab[pi];
                    v0[k] = p.x; v1[k] = p.y; v2[k] = p.z;
                    tx += px; ty += py; tn += 1;
                }
            }
        }
        // tree steps 128, 64, 32 (slots differ by 4, 2, 1 in k)
#pragma unroll
        for (int h = 4; h >= 1; h >>= 1)
#pragma unroll
            for (int k = 0; k < h; ++k) { v0[k] += v0[k + h]; v1[k] += v1[k + h]; v2[k] += v2[k + h]; }
        float a0 = v0[0], a1 = v1[0], a2 = v2[0];
        // tree steps 16, 8, 4, 2, 1
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            a0 += __shfl_down_sync(0xffffffffu, a0, o);
            a1 += __shfl_down_sync(0xffffffffu, a1, o);
            a2 += __shfl_down_sync(0xffffffffu, a2, o);
        }
        sx += __reduce_add_sync(0xffffffffu, tx);
        sy += __reduce_add_sync(0xffffffffu, ty);
        sn += __reduce_add_sync(0xffffffffu, tn);
        c0 = c0 + a0; c1 = c1 + a1; c2 = c2 + a2;   // only lane 0 holds the tree root
    }
    if (lane == 0) {
        float4 *m = map + 2 * ((size_t)z * M + c);
        float keep7 = m[1].w;   // s7 is not touched by SLIC
        if (sn != 0) {
            float n = (float)sn;
            m[0] = make_float4((float)c, (float)sx / n, (float)sy / n, c0 / n);
            m[1] = make_float4(c1 / n, c2 / n, n, keep7);
        } else {
            m[0] = make_float4((float)c, 0.0f, 0.0f, 0.0f);
            m[1] = make_float4(0.0f, 0.0f, 0.0f, keep7);
        }
    }
}

// Tiled form of the same update: a CTA stages the pixels its TCX x TCY clusters can see
// (labels + Lab, once, coalesced) in shared memory and its 8 warps then run the per-cluster
// window reduction above out of shared memory.  The direct kernel tests every pixel from six
// clusters through L1/L2 (latency bound: 1.1 ms per launch on the 15-view frame); the tile
// reads each pixel from HBM/L2 once plus the halo.  Arithmetic and its order are unchanged.
// Pixels outside the image are staged with label 0xFFFFFFFF, which no cluster carries.
constexpr int UPD_THREADS = 512;
struct UpdTile { int tcx, tcy, xext, yext, rw, rh, pitch; };

// tile geometry for superpixel size S (window extent actually scanned: clSLIC.cpp:22-23, 313;
// clcode.cl:552-566); constexpr so that the S = 8 / 16 instantiations fold every index computation
__host__ __device__ constexpr UpdTile upd_tile(int S)
{
    UpdTile t{};
    const int G = (S * S * 9 + 255) / 256;   // ceil(9 S^2 / 256), exact in integers
    const int cpl = S * 3 / 16;
    t.xext = cpl * 16 < 3 * S ? cpl * 16 : 3 * S;
    const int gy = ((G + cpl - 1) / cpl) * 16;
    t.yext = gy < 3 * S ? gy : 3 * S;
    t.tcx = 64 / S > 1 ? 64 / S : 1;
    t.tcy = 32 / S > 1 ? 32 / S : 1;
    t.rw = (t.tcx - 1) * S + t.xext;
    t.rh = (t.tcy - 1) * S + t.yext;
    t.pitch = t.rw + ((16 - t.rw % 32) % 32 + 32) % 32;   // rows two apart in a half-warp pair land on disjoint banks
    return t;
}

// per-cluster window reduction out of the staged tile (s_lab / s_idx, row pitch t.pitch); the eight floats of a
// centre are written by eight lanes, one component each (five of them a division: one division's worth of
// instructions instead of five on a single lane; s7 is simply not stored)
template <int S_T, int NT = UPD_THREADS>
__device__ __forceinline__ void upd_reduce(const float4 *s_lab, const uint32_t *s_idx, float4 *__restrict__ map, int S, int mw,
                                           int mh, int M, int z, int gx0, int gy0, int x0, int y0)
{
    const UpdTile t = upd_tile(S);
    const int G = (S * S * 9 + 255) / 256, cpl = S * 3 / 16;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int lx = lane & 15, lyh = lane >> 4;
    for (int ci = warp; ci < t.tcx * t.tcy; ci += NT / 32) {
        const int lcx = ci % t.tcx, lcy = ci / t.tcx;
        const int gx = gx0 + lcx, gy = gy0 + lcy;
        if (gx >= mw || gy >= mh) continue;   // warp-uniform
        const int c = gy * mw + gx;
        float c0 = 0.0f, c1 = 0.0f, c2 = 0.0f;
        // position and count sums are exact integers: per lane over all window tiles, one warp
        // reduction at the end.  Per slot the lane adds 1 | k << 8 | nbr_y << 16 (count, row pair, tile row).
        int tx = 0, ty = 0, tn = 0;
        // S = 8: the two live window tiles are unrolled, so the second one (rows 16..23) tests only its four
        // in-window row pairs
#pragma unroll(S_T == 8 ? 3 : 1)
        for (int g = 0; g < G; ++g) {
            const int nbr_x = g % cpl, nbr_y = g / cpl;
            if (nbr_y * 16 >= S * 3) break;
            const int pxo = nbr_x * 16 + lx;
            const bool colok = pxo < S * 3;
            float v0[8], v1[8], v2[8];
            int acc = 0;
            const int sbase = (lcy * S + nbr_y * 16 + lyh) * t.pitch + lcx * S + pxo;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int pyo = nbr_y * 16 + 2 * k + lyh;
                v0[k] = 0.0f; v1[k] = 0.0f; v2[k] = 0.0f;
                if (colok && pyo < S * 3) {
                    const int si = sbase + 2 * k * t.pitch;
                    if (s_idx[si] == (uint32_t)c) {
                        const float4 p = s_lab[si];
                        v0[k] = p.x; v1[k] = p.y; v2[k] = p.z;
                        acc += 1 + (k << 8);
                    }
                }
            }
            {
                const int n = acc & 255, ks = acc >> 8;
                tn += n;
                tx += n * (x0 + lcx * S + pxo);
                ty += n * (y0 + lcy * S + nbr_y * 16 + lyh) + 2 * ks;
            }
#pragma unroll
            for (int h = 4; h >= 1; h >>= 1)
#pragma unroll
                for (int k = 0; k < h; ++k) { v0[k] += v0[k + h]; v1[k] += v1[k + h]; v2[k] += v2[k + h]; }
            float a0 = v0[0], a1 = v1[0], a2 = v2[0];
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) {
                a0 += __shfl_down_sync(0xffffffffu, a0, o);
                a1 += __shfl_down_sync(0xffffffffu, a1, o);
                a2 += __shfl_down_sync(0xffffffffu, a2, o);
            }
            c0 = c0 + a0; c1 = c1 + a1; c2 = c2 + a2;   // only lane 0 holds the tree root
        }
        const int sx = __reduce_add_sync(0xffffffffu, tx);
        const int sy = __reduce_add_sync(0xffffffffu, ty);
        const int sn = __reduce_add_sync(0xffffffffu, tn);
        // finalize_reduction_result (clcode.cl:719-773): lane L writes component L of the float8
        const float r0 = __shfl_sync(0xffffffffu, c0, 0), r1 = __shfl_sync(0xffffffffu, c1, 0), r2 = __shfl_sync(0xffffffffu, c2, 0);
        const float n = (float)sn;
        const float num = lane == 1 ? (float)sx : lane == 2 ? (float)sy : lane == 3 ? r0 : lane == 4 ? r1 : r2;
        float val = num / n;
        if (lane == 0) val = (float)c;
        if (lane == 6) val = n;
        if (sn == 0 && lane != 0) val = 0.0f;   // empty cluster: centre (0, 0), colour 0, count 0 (clcode.cl:731-734)
        if (lane < 7) reinterpret_cast<float *>(map + 2 * ((size_t)z * M + c))[lane] = val;
    }
}

template <int S_T>
__global__ void __launch_bounds__(UPD_THREADS) k_update_centers_tiled(const float4 *__restrict__ lab,
                                                              const uint32_t *__restrict__ idx,
                                                              float4 *__restrict__ map, int W, int H, int mw, int mh,
                                                              int S_rt, int N, int M, int z0)
{
    const int S = S_T > 0 ? S_T : S_rt;
    const UpdTile t = upd_tile(S);
    extern __shared__ __align__(16) unsigned char upd_smem[];
    float4 *s_lab = reinterpret_cast<float4 *>(upd_smem);
    uint32_t *s_idx = reinterpret_cast<uint32_t *>(s_lab + (size_t)t.rh * t.pitch);
    const int z = z0 + blockIdx.z;
    const int gx0 = blockIdx.x * t.tcx, gy0 = blockIdx.y * t.tcy;
    const int x0 = gx0 * S - S, y0 = gy0 * S - S;
    const uint32_t *vidx = idx + (size_t)z * N;
    const float4 *vlab = lab + (size_t)z * N;
    // staging: four elements per thread in flight (loads first, then the shared-memory stores)
    const int total = t.rw * t.rh;
    for (int e0 = threadIdx.x; e0 < total; e0 += 4 * UPD_THREADS) {
        uint32_t l[4];
        float4 c[4];
        int so[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int e = e0 + j * UPD_THREADS;
            const int ry = e / t.rw, rx = e - ry * t.rw;
            const int px = x0 + rx, py = y0 + ry;
            l[j] = 0xFFFFFFFFu;
            c[j] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            so[j] = e < total ? ry * t.pitch + rx : -1;
            if (e < total && py >= 0 && py < H && px >= 0 && px < W) {
                const size_t pi = (size_t)py * W + px;
                l[j] = __ldg(vidx + pi);
                c[j] = __ldg(vlab + pi);
            }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (so[j] >= 0) { s_idx[so[j]] = l[j]; s_lab[so[j]] = c[j]; }
    }
    __syncthreads();
    upd_reduce<S_T>(s_lab, s_idx, map, S, mw, mh, M, z, gx0, gy0, x0, y0);
}

// The same with the tile staged by TMA: one box load of the Lab tile and one of the label tile (cp.async.bulk.tensor,
// completion on an mbarrier) replace the per-thread index arithmetic, loads and shared-memory stores of the staging
// loop -- 28 % of the kernel's instructions and its long-scoreboard stalls (profiles/r02_v3_k_update_centers_tiled*).
// The box is t.pitch pixels wide (the bank-friendly row pitch of the reduction, a few columns more than the window
// needs).  TMA fills out-of-range box parts with zeros, and a box that leaves a view vertically reads the rows of the
// neighbouring view in the stack: label 0 is a real cluster and so are the neighbour's labels, therefore CTAs whose
// box leaves the image overwrite those label slots with 0xFFFFFFFF, which no cluster carries.
template <int S_T>
__global__ void __launch_bounds__(UPD_THREADS) k_update_centers_tma(const __grid_constant__ CUtensorMap tm_lab,
                                                                    const __grid_constant__ CUtensorMap tm_idx,
                                                                    float4 *__restrict__ map, int W, int H, int mw, int mh,
                                                                    int M, int z0)
{
    constexpr int S = S_T;
    constexpr UpdTile t = upd_tile(S);
    extern __shared__ __align__(128) unsigned char upd_smem[];
    __shared__ __align__(8) uint64_t s_bar;
    float4 *s_lab = reinterpret_cast<float4 *>(upd_smem);
    uint32_t *s_idx = reinterpret_cast<uint32_t *>(s_lab + (size_t)t.rh * t.pitch);
    const int z = z0 + blockIdx.z;
    const int gx0 = blockIdx.x * t.tcx, gy0 = blockIdx.y * t.tcy;
    const int x0 = gx0 * S - S, y0 = gy0 * S - S;
    if (threadIdx.x == 0) {
        mbar_init(&s_bar, 1);
        mbar_fence_init();
        mbar_arrive_expect_tx(&s_bar, (uint32_t)(t.rh * t.pitch * (sizeof(float4) + sizeof(uint32_t))));
        tma_load_2d(s_lab, &tm_lab, 2 * x0, y0 + z * H, &s_bar);   // Lab pixels are two 8-byte elements wide
        tma_load_2d(s_idx, &tm_idx, x0, y0 + z * H, &s_bar);
    }
    __syncthreads();           // the initialised barrier is visible to every thread before it waits on it
    mbar_wait(&s_bar, 0);
    if (x0 < 0 || y0 < 0 || x0 + t.pitch > W || y0 + t.rh > H) {   // CTA-uniform: the box leaves the image
        for (int e = threadIdx.x; e < t.rh * t.pitch; e += UPD_THREADS) {
            const int ry = e / t.pitch, rx = e - ry * t.pitch;
            const int px = x0 + rx, py = y0 + ry;
            if (px < 0 || px >= W || py < 0 || py >= H) s_idx[e] = 0xFFFFFFFFu;
        }
        __syncthreads();
    }
    upd_reduce<S_T>(s_lab, s_idx, map, S, mw, mh, M, z, gx0, gy0, x0, y0);
}

// The TMA-staged update as a PERSISTENT, double-buffered kernel: one CTA per SM walks the tiles of the frame (x
// fastest, so that the CTAs of a wave work on neighbouring tiles and share their halos in L2) with two tile buffers;
// the boxes of tile i + 1 are in flight while the warps reduce tile i.  MEASURED AND REJECTED (MVS_UPD_MODE=pipe):
// 0.31 ms per launch against 0.24 for the one-tile kernel above with two resident CTAs, with 512 or 1024 threads --
// the update is bound by the instructions of its pinned reduction tree (30 shuffles + 45 adds per cluster for three
// sums, issue 63 %), not by waiting for its tile; two independent CTAs overlap their phases better than one CTA
// whose 32 warps meet at a barrier per tile.
// A buffer is refilled only after every thread has left it: block barrier, then fence.proxy.async by the issuing thread
// (generic-proxy reads and the label patches before it, the TMA writes after it).
#ifndef UPP_THREADS
#define UPP_THREADS 1024
#endif
template <int S_T, int NT>
__global__ void __launch_bounds__(NT, 1) k_update_centers_pipe(const __grid_constant__ CUtensorMap tm_lab,
                                                               const __grid_constant__ CUtensorMap tm_idx,
                                                               float4 *__restrict__ map, int W, int H, int mw, int mh,
                                                               int M, int z0, int ntx, int nty, int ntiles)
{
    constexpr int S = S_T;
    constexpr UpdTile t = upd_tile(S);
    constexpr uint32_t LAB_BYTES = (uint32_t)(t.rh * t.pitch * sizeof(float4));
    constexpr uint32_t STAGE_BYTES = (uint32_t)(t.rh * t.pitch * (sizeof(float4) + sizeof(uint32_t)));
    static_assert(LAB_BYTES % 128 == 0 && STAGE_BYTES % 128 == 0, "TMA destinations are 128-byte aligned");
    extern __shared__ __align__(128) unsigned char upd_smem[];
    __shared__ __align__(8) uint64_t s_bar[2];
    auto tile_origin = [&](int tile, int &gx0, int &gy0, int &z) {
        const int bx = tile % ntx, r = tile / ntx;
        gx0 = bx * t.tcx; gy0 = (r % nty) * t.tcy; z = z0 + r / nty;
    };
    auto issue = [&](int tile, int stage) {   // one thread
        int gx0, gy0, z;
        tile_origin(tile, gx0, gy0, z);
        const int x0 = gx0 * S - S, y0 = gy0 * S - S;
        unsigned char *dst = upd_smem + (size_t)stage * STAGE_BYTES;
        mbar_arrive_expect_tx(&s_bar[stage], STAGE_BYTES);
        tma_load_2d(dst, &tm_lab, 2 * x0, y0 + z * H, &s_bar[stage]);   // Lab pixels are two 8-byte elements wide
        tma_load_2d(dst + LAB_BYTES, &tm_idx, x0, y0 + z * H, &s_bar[stage]);
    };
    if (threadIdx.x == 0) {
        mbar_init(&s_bar[0], 1);
        mbar_init(&s_bar[1], 1);
        mbar_fence_init();
        if ((int)blockIdx.x < ntiles) issue(blockIdx.x, 0);
    }
    __syncthreads();
    int i = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++i) {
        const int stage = i & 1, next = tile + gridDim.x;
        if (threadIdx.x == 0 && next < ntiles) {
            fence_proxy_async();
            issue(next, stage ^ 1);
        }
        int gx0, gy0, z;
        tile_origin(tile, gx0, gy0, z);
        const int x0 = gx0 * S - S, y0 = gy0 * S - S;
        const float4 *s_lab = reinterpret_cast<const float4 *>(upd_smem + (size_t)stage * STAGE_BYTES);
        uint32_t *s_idx = reinterpret_cast<uint32_t *>(upd_smem + (size_t)stage * STAGE_BYTES + LAB_BYTES);
        mbar_wait(&s_bar[stage], (uint32_t)((i >> 1) & 1));
        if (x0 < 0 || y0 < 0 || x0 + t.pitch > W || y0 + t.rh > H) {   // CTA-uniform: the box leaves the image
            for (int e = threadIdx.x; e < t.rh * t.pitch; e += NT) {
                const int ry = e / t.pitch, rx = e - ry * t.pitch;
                const int px = x0 + rx, py = y0 + ry;
                if (px < 0 || px >= W || py < 0 || py >= H) s_idx[e] = 0xFFFFFFFFu;
            }
            __syncthreads();
        }
        upd_reduce<S_T, NT>(s_lab, s_idx, map, S, mw, mh, M, z, gx0, gy0, x0, y0);
        __syncthreads();   // every thread has left this buffer: the next iteration may refill it
    }
}

// (re)build the tensor maps of the centre update for this frame's Lab / label buffers (S = 8 tile: 80 x 48 box)
int prepare_update_centers_tma(const FrameDev &f, const char **why)
{
    static const char *none = "";
    *why = none;
    SlicTma *tm = f.slic_tma;
    if (!tm || f.S != 8) return 0;
    if (tm->state != 0 && tm->lab == (const void *)f.lab && tm->idx == (const void *)f.idx) return 0;
    tm->lab = f.lab; tm->idx = f.idx;
    tm->state = -1;
    const UpdTile t = upd_tile(8);
    if (f.W % 4 || (double)f.V * f.H >= 2147483647.0) return 0;   // label rows must be 16-byte multiples for a tensor map
    if (tma_make_image_map(&tm->map_lab, f.lab, 16, f.W, f.H, f.V, t.pitch, t.rh, why)) return 1;
    if (tma_make_image_map(&tm->map_idx, f.idx, 4, f.W, f.H, f.V, t.pitch, t.rh, why)) return 1;
    tm->state = 1;
    return 0;
}

int launch_update_centers(const FrameDev &f, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    const int S = f.S;
    int G = (int)ceilf((float)(S * S * 9) / 256.0f);   // clSLIC.cpp:22-23
    int cpl = S * 3 / 16;
    const UpdTile t = upd_tile(S);
    size_t smem = (size_t)t.rh * t.pitch * (sizeof(float4) + sizeof(uint32_t));
    if (G == (S * S * 9 + 255) / 256 && smem <= 200 * 1024) {
        dim3 g((f.mw + t.tcx - 1) / t.tcx, (f.mh + t.tcy - 1) / t.tcy, z1 - z0);
        // S = 8 (the shipped setting): tile staged by TMA when the frame's tensor maps could be built
        // (MVS_UPD_MODE=loads forces the load-staged kernel for A/B measurement)
        // (MVS_UPD_MODE=pipe: the persistent double-buffered kernel, a measured rejection)
        static const bool force_loads = [] { const char *e = getenv("MVS_UPD_MODE"); return e && !strcmp(e, "loads"); }();
        static const bool force_tile = [] { const char *e = getenv("MVS_UPD_MODE"); return !(e && !strcmp(e, "pipe")); }();
        SlicTma *tm = f.slic_tma;
        const bool maps_ok = S == 8 && !force_loads && tm && tm->state == 1 && tm->lab == (const void *)f.lab && tm->idx == (const void *)f.idx;
        if (maps_ok && !force_tile && 2 * smem + 1024 <= 227 * 1024) {
            static int sms[64] = {0};
            int dev = 0;
            cudaGetDevice(&dev);
            if (dev >= 0 && dev < 64 && !sms[dev]) cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev);
            const int nsm = dev >= 0 && dev < 64 && sms[dev] > 0 ? sms[dev] : 148;
            const int ntiles = (int)(g.x * g.y * g.z);
            auto kern = k_update_centers_pipe<8, UPP_THREADS>;
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * smem));
            kern<<<std::min(nsm, ntiles), UPP_THREADS, 2 * smem, st>>>(tm->map_lab, tm->map_idx, f.map, f.W, f.H, f.mw, f.mh,
                                                                       f.M, z0, (int)g.x, (int)g.y, ntiles);
            return 1;
        }
        if (maps_ok) {
            cudaFuncSetAttribute(k_update_centers_tma<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            k_update_centers_tma<8><<<g, UPD_THREADS, smem, st>>>(tm->map_lab, tm->map_idx, f.map, f.W, f.H, f.mw, f.mh, f.M, z0);
            return 1;
        }
#define UPD_LAUNCH(ST)                                                                                              \
    do {                                                                                                            \
        cudaFuncSetAttribute(k_update_centers_tiled<ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); \
        k_update_centers_tiled<ST><<<g, UPD_THREADS, smem, st>>>(f.lab, f.idx, f.map, f.W, f.H, f.mw, f.mh, S, f.N, f.M, z0); \
    } while (0)
        if (S == 8) UPD_LAUNCH(8);
        else if (S == 16) UPD_LAUNCH(16);
        else UPD_LAUNCH(0);
#undef UPD_LAUNCH
        return 1;
    }
    dim3 g((f.M + 7) / 8, z1 - z0);
    k_update_centers<<<g, 256, 0, st>>>(f.lab, f.idx, f.map, f.W, f.H, f.mw, f.mh, f.S, f.N, f.M, G, cpl, z0);
    return 1;
}

// ---------------------------------------------------------------------------------------
// Connectivity: 5x5 vote, last differing label wins, threshold >= 16, 2-pixel frame copied
// (clcode.cl:676-711); two passes idx -> tmp -> idx (clSLIC.cpp:390-410).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_suppress(const uint32_t *__restrict__ in, uint32_t *__restrict__ out, int W,
                                                  int H, int N, int z0)
{
    int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
    int z = z0 + blockIdx.z;
    if (x >= W || y >= H) return;
    const uint32_t *vin = in + (size_t)z * N;
    int id = W * y + x;
    uint32_t cl = vin[id];
    uint32_t res = cl;
    if (!(x <= 1 || y <= 1 || x >= W - 2 || y >= H - 2)) {
        int diff_count = 0;
        uint32_t diff_lable = 0xFFFFFFFFu;
#pragma unroll
        for (int j = -2; j <= 2; j++)
#pragma unroll
            for (int i = -2; i <= 2; i++) {
                uint32_t nl = vin[(y + j) * W + (x + i)];
                if (nl != cl) { diff_lable = nl; diff_count++; }
            }
        if (diff_count >= 16) res = diff_lable;
    }
    out[(size_t)z * N + id] = res;
}

int launch_connectivity(const FrameDev &f, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    dim3 b(32, 8), g((f.W + 31) / 32, (f.H + 7) / 8, z1 - z0);
    k_suppress<<<g, b, 0, st>>>(f.idx, f.idx_tmp, f.W, f.H, f.N, z0);
    k_suppress<<<g, b, 0, st>>>(f.idx_tmp, f.idx, f.W, f.H, f.N, z0);
    return 2;
}

} // namespace mvs
