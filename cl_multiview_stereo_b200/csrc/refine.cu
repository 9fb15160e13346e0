// refine.cu -- per-superpixel plane refinement and fusion (sm_100a).
//
// Replaces the OpenCL kernels driven by clDepthRefinement::do_refinement (reference
// clMVDE/depth_refinement.cpp:91-118): compute_flatness (clcode.cl:1076-1132),
// init_current_state (:1362-1404, with init_smoothness :1136-1254 and
// initialize_consistency :1260-1357), propagate (:1727-1900, with update :1635-1673,
// spatialRefinement :1687-1723, compute_smoothness :1407-1525, compute_consistency
// :1528-1631) and spixl_to_image (:1906-1931).
//
// Internal state layout: the reference's 6-float record {d,sm,cs,nx,ny,nz} is kept as
// two arrays, plane = float4{d,nx,ny,nz} (what neighbours and other views gather: one
// 16-byte load) and cost = float2{sm,cs} (read only by the owner).  k_export_state /
// k_import_state convert to and from the 6-float API layout.
#include "common.cuh"
#include "mvs_math.cuh"

namespace mvs {

struct RCtx {
    const float4 *map;
    const uint32_t *idx;
    const uint2 *rep;
    const float2 *flat;
    const int *view_subset, *subset_num;
    const int2 *cam;
    int V, W, H, mw, mh, N, M;
    float gamma, alpha, bl_ratio, fuse;
};

__device__ __forceinline__ RCtx make_ctx(const FrameDev &f, const RefineArgs &a)
{
    RCtx c;
    c.map = f.map; c.idx = f.idx; c.rep = f.rep; c.flat = f.flat;
    c.view_subset = f.view_subset; c.subset_num = f.subset_num;
    c.V = f.V; c.W = f.W; c.H = f.H; c.mw = f.mw; c.mh = f.mh; c.N = f.N; c.M = f.M; c.cam = f.cam;
    c.gamma = a.gamma; c.alpha = a.alpha; c.bl_ratio = a.bl_ratio; c.fuse = a.fuse;
    return c;
}

// ---------------------------------------------------------------------------------------
// Flatness: 1 + sum of squared colour differences to the W, E, S, N neighbours;
// out = (exp(-fl*g), 1 - exp(-0.25*fl*g)) -- the second one is a double expression in the
// reference (unsuffixed literals, clcode.cl:1131) narrowed to float on the store.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_flatness(const float4 *__restrict__ map, float2 *__restrict__ flat, int mw,
                                                  int mh, int M, float gamma, int z0)
{
    int sp = blockIdx.x * blockDim.x + threadIdx.x;
    int z = z0 + blockIdx.y;
    if (sp >= M) return;
    int x = sp % mw, y = sp / mw;
    const float4 *vm = map + 2 * (size_t)z * M;
    float4 a = vm[2 * sp], b = vm[2 * sp + 1];
    float c0x = a.w, c0y = b.x, c0z = b.y;
    float fl = 1.0f;
    const int off[4] = {-1, 1, mw, -mw};
    const bool ok[4] = {x - 1 >= 0, x + 1 < mw, y + 1 < mh, y - 1 >= 0};
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (ok[k]) {
            float4 a1 = vm[2 * (sp + off[k])], b1 = vm[2 * (sp + off[k]) + 1];
            float dx = a1.w - c0x, dy = b1.x - c0y, dz = b1.y - c0z;
            fl += dx * dx + dy * dy + dz * dz;
        }
    float2 o;
    o.x = expf_pinned(-fl * gamma);
    o.y = (float)(1.0 - exp_pinned(-0.25 * (double)fl * (double)gamma));
    flat[(size_t)z * M + sp] = o;
}

int launch_flatness(const FrameDev &f, float gamma_flat, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    dim3 g((f.M + 255) / 256, z1 - z0);
    k_flatness<<<g, 256, 0, st>>>(f.map, f.flat, f.mw, f.mh, f.M, gamma_flat, z0);
    return 1;
}

// ---------------------------------------------------------------------------------------
// helpers shared by init state and propagate
// ---------------------------------------------------------------------------------------
struct Samples { int s[9]; };

__device__ __forceinline__ Samples rep_samples(uint2 r)
{
    Samples o;
    o.s[0] = r.x & 255; o.s[1] = (r.x >> 8) & 255; o.s[2] = (r.x >> 16) & 255; o.s[3] = r.x >> 24;
    o.s[4] = 0;
    o.s[5] = r.y & 255; o.s[6] = (r.y >> 8) & 255; o.s[7] = (r.y >> 16) & 255; o.s[8] = r.y >> 24;
    return o;
}

// tail of one neighbour view (clcode.cl:1339-1347, 1614-1622); '+= 0.5*fl.y' is double
__device__ __forceinline__ void cs_view_tail(float num, float vws, float ows, float visibility, float visible,
                                             float fl_y, float &consistency, int &view_counter)
{
    if (num > 0.0f) {
        view_counter++;
        if (vws > 0.0f) consistency += (vws / num) * (visibility / vws) * (visible / vws);
        if (ows > 0.0f) consistency = (float)((double)consistency + 0.5 * (double)fl_y);
    }
}

__device__ __forceinline__ float cs_finish(float consistency, int view_counter)
{
    const float margin = (float)0.01;
    if (view_counter > 0) return fmaxf(margin, consistency / (float)view_counter);
    return margin;
}

// 'visible += when_visible * exp(-diff*diff*alpha)' without evaluating exp for occluded
// samples: 0 * exp(t) is +0 for every non-NaN diff, and NaN for a NaN diff.
__device__ __forceinline__ void visible_term(float diff, float fuse, float alpha, float &visible, float &vws,
                                             float &ows)
{
    if (fabsf(diff) < fuse) {
        visible += expf_pinned(-diff * diff * alpha);
        vws += 1.0f;
    } else {
        if (diff != diff) visible += diff;
        ows += 1.0f;
    }
}

// ---------------------------------------------------------------------------------------
// init_current_state
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void sm_term_init(const float4 *vm, int q, float cx, float cy, float cz, float disp,
                                             float g, float alpha, float &smoothness, float &weight_norm)
{
    float4 a = vm[2 * q], b = vm[2 * q + 1];
    float diff = dist3(cx, cy, cz, a.w, b.x, b.y);
    float similarity = expf_pinned(-diff * diff * g);
    diff = disp - b.w;
    smoothness += similarity * expf_pinned(-diff * diff * alpha);
    weight_norm += similarity;
}

__global__ void __launch_bounds__(128) k_init_state(FrameDev f, RefineArgs ra, int z0)
{
    RCtx c = make_ctx(f, ra);
    int sp = blockIdx.x * blockDim.x + threadIdx.x;
    int z = z0 + blockIdx.y;
    if (sp >= c.M) return;
    int x = sp % c.mw, y = sp / c.mw;
    size_t id = (size_t)z * c.M + sp;
    const float4 *vm = c.map + 2 * (size_t)z * c.M;
    float4 ma = vm[2 * sp], mb = vm[2 * sp + 1];
    float cx = ma.w, cy = mb.x, cz = mb.y, d = mb.w;
    float2 fl = c.flat[id];

    // ---- init_smoothness (clcode.cl:1136-1254)
    float smoothness = 0.0f, weight_norm = 0.0f;
    for (int i = -1; i <= 1; i++)
        for (int j = -1; j <= 1; j++) {
            int qx = x + i, qy = y + j;
            if (qx >= 0 && qy >= 0 && qx < c.mw && qy < c.mh && (i != 0 || j != 0)) {
                float4 a = vm[2 * (qy * c.mw + qx)], b = vm[2 * (qy * c.mw + qx) + 1];
                float diff = dist3(a.w, b.x, b.y, cx, cy, cz);
                float similarity = expf_pinned(-diff * diff * c.gamma);
                diff = d - b.w;
                smoothness += similarity * expf_pinned(-diff * diff * c.alpha);
                weight_norm += similarity;
            }
        }
    int step_size = max(1, d2i((double)(fl.x * ra.kernel_step_size) + 0.5));
    for (int i = 1; i <= ra.no_kernel_steps; i++) {
        float gamma_i = c.gamma * (float)(1 + i);
        int step = i * step_size;
        if (x > step) sm_term_init(vm, c.mw * y + x - (step + 1), cx, cy, cz, d, gamma_i, c.alpha, smoothness, weight_norm);
        if (x < c.mw - step - 1) sm_term_init(vm, c.mw * y + x + (step + 1), cx, cy, cz, d, gamma_i, c.alpha, smoothness, weight_norm);
        if (y > step) sm_term_init(vm, c.mw * (y - step - 1) + x, cx, cy, cz, d, gamma_i, c.alpha, smoothness, weight_norm);
        if (y < c.mh - step - 1) sm_term_init(vm, c.mw * (y + step + 1) + x, cx, cy, cz, d, gamma_i, c.alpha, smoothness, weight_norm);
    }
    float sm = weight_norm > 0.0f ? smoothness / weight_norm : (float)0.000001;

    // ---- initialize_consistency (clcode.cl:1260-1357)
    float consistency = 0.0f;
    int view_counter = 0;
    const int2 camz = c.cam[z];
    int camx = camz.x, camy = camz.y;
    Samples ss = rep_samples(c.rep[id]);
    int icx = f2i(ma.y), icy = f2i(ma.z);
    int ns = c.subset_num[z];
    for (int n = 0; n < ns; n++) {
        int view = c.view_subset[z * c.V + n];
        const int2 cv = c.cam[view];
        float rx = round_haz(d * (float)(cv.x - camx));
        float ry = round_haz(c.bl_ratio * d * (float)(cv.y - camy));
        const uint32_t *vidx = c.idx + (size_t)view * c.N;
        const float4 *pm = c.map + 2 * (size_t)view * c.M;
        float vws = 0.0f, ows = 0.0f, num = 0.0f, visibility = 0.0f, visible = 0.0f;
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            int i = k / 3 - 1, j = k % 3 - 1;
            int xr = icx + ss.s[k] * i, yr = icy + ss.s[k] * j;
            int xp = f2i((float)xr - rx), yp = f2i((float)yr - ry);
            if (xp >= 0 && yp >= 0 && xp < c.W && yp < c.H) {
                uint32_t q = vidx[(size_t)yp * c.W + xp];
                float4 a = pm[2 * (size_t)q], b = pm[2 * (size_t)q + 1];
                visible_term(b.w - d, c.fuse, c.alpha, visible, vws, ows);
                float diff = dist3(a.w, b.x, b.y, cx, cy, cz);
                visibility += expf_pinned(-diff * diff * c.gamma);
                num += 1.0f;
            }
        }
        cs_view_tail(num, vws, ows, visibility, visible, fl.y, consistency, view_counter);
    }
    float cs = cs_finish(consistency, view_counter);
    f.plane[0][id] = make_float4(d, 0.0f, 0.0f, 1.0f);
    f.cost[0][id] = make_float2(sm, cs);
}

// The plane part of init_current_state alone: {d, n} = {s7, (0, 0, 1)} (clcode.cl:1398-1403).  On a view-sharded frame
// a rank fills the INPUT planes of its halo views for iteration 0 with this from the centres it already holds,
// instead of receiving them from their owners (one exchange less per frame; same values by construction).
__global__ void __launch_bounds__(256) k_plane_from_map(const float4 *__restrict__ map, float4 *__restrict__ plane, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) plane[i] = make_float4(map[2 * i + 1].w, 0.0f, 0.0f, 1.0f);
}

int launch_plane_from_map(const FrameDev &f, int z0, int z1, cudaStream_t st)
{
    size_t n = (size_t)(z1 - z0) * f.M, o = (size_t)z0 * f.M;
    if (!n) return 0;
    k_plane_from_map<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(f.map + 2 * o, f.plane[0] + o, n);
    return 1;
}

int launch_init_state(const FrameDev &f, const RefineArgs &a, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    dim3 g((f.M + 127) / 128, z1 - z0);
    k_init_state<<<g, 128, 0, st>>>(f, a, z0);
    return 1;
}

// ---------------------------------------------------------------------------------------
// spixl_to_image (clcode.cl:1906-1931): evaluate the own superpixel's plane at each pixel
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_spixl_to_image(const float4 *__restrict__ map,
                                                        const uint32_t *__restrict__ idx,
                                                        const float4 *__restrict__ plane, float *__restrict__ disp,
                                                        int W, int H, int N, int M, int z0)
{
    int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
    int z = z0 + blockIdx.z;
    if (x >= W || y >= H) return;
    size_t pi = (size_t)z * N + (size_t)y * W + x;
    uint32_t q = idx[pi];
    float4 a = map[2 * ((size_t)z * M + q)];
    float4 pl = plane[(size_t)z * M + q];
    disp[pi] = (pl.y * (a.y - (float)x) + pl.z * (a.z - (float)y) + pl.w * pl.x) / pl.w;
}

// the same with four pixels per thread (W % 4 == 0): 128-bit label load and disparity store; the four pixels
// mostly share a superpixel, so their centre / plane gathers hit the same lines
__global__ void __launch_bounds__(256) k_spixl_to_image4(const float4 *__restrict__ map,
                                                         const uint4 *__restrict__ idx4,
                                                         const float4 *__restrict__ plane, float4 *__restrict__ disp4,
                                                         int W, int H, int N, int M, int z0)
{
    const int xq = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
    const int z = z0 + blockIdx.z;
    const int x = 4 * xq;
    if (x >= W || y >= H) return;
    const size_t pq = ((size_t)z * N + (size_t)y * W + x) / 4;
    const uint4 q4 = idx4[pq];
    const uint32_t q[4] = {q4.x, q4.y, q4.z, q4.w};
    float o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float4 a = map[2 * ((size_t)z * M + q[k])];
        const float4 pl = plane[(size_t)z * M + q[k]];
        o[k] = (pl.y * (a.y - (float)(x + k)) + pl.z * (a.z - (float)y) + pl.w * pl.x) / pl.w;
    }
    disp4[pq] = make_float4(o[0], o[1], o[2], o[3]);
}

int launch_spixl_to_image(const FrameDev &f, int src, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    dim3 b(32, 8);
    if (f.W % 4 == 0) {
        dim3 g((f.W / 4 + 31) / 32, (f.H + 7) / 8, z1 - z0);
        k_spixl_to_image4<<<g, b, 0, st>>>(f.map, (const uint4 *)f.idx, f.plane[src], (float4 *)f.disp, f.W, f.H, f.N, f.M, z0);
    } else {
        dim3 g((f.W + 31) / 32, (f.H + 7) / 8, z1 - z0);
        k_spixl_to_image<<<g, b, 0, st>>>(f.map, f.idx, f.plane[src], f.disp, f.W, f.H, f.N, f.M, z0);
    }
    return 1;
}

// ---------------------------------------------------------------------------------------
// Per-pixel record of a state: {label, d} with d = the plane of the pixel's superpixel evaluated at
// the pixel.  compute_consistency (clcode.cl:1581-1600) looks up, for every reprojected sample, the
// label at the projected pixel, that superpixel's centre and its INPUT plane, and evaluates
//   d_intrp_proj = (n.x*(c.x - xp) + n.y*(c.y - yp) + n.z*d) / n.z
// -- a pure function of (view, pixel) for a fixed input state, the same for every candidate plane of
// every superpixel that lands on the pixel.  It is therefore evaluated ONCE per pixel and iteration
// here (same expression, same operand order, same bits) and k_propagate gathers one 8-byte record per
// sample instead of label -> (32-byte centre record, 16-byte plane) -> 8 flops and an IEEE division.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_pixel_plane(const float4 *__restrict__ map, const uint32_t *__restrict__ idx,
                                                     const float4 *__restrict__ plane, uint2 *__restrict__ pixd,
                                                     int W, int H, int N, int M, int z0)
{
    int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
    int z = z0 + blockIdx.z;
    if (x >= W || y >= H) return;
    size_t pi = (size_t)z * N + (size_t)y * W + x;
    uint32_t q = idx[pi];
    float4 a = map[2 * ((size_t)z * M + q)];
    float4 pl = plane[(size_t)z * M + q];
    float d = (pl.y * (a.y - (float)x) + pl.z * (a.z - (float)y) + pl.w * pl.x) / pl.w;
    pixd[pi] = make_uint2(q, __float_as_uint(d));
}

__global__ void __launch_bounds__(256) k_pixel_plane4(const float4 *__restrict__ map, const uint4 *__restrict__ idx4,
                                                      const float4 *__restrict__ plane, uint4 *__restrict__ pixd4,
                                                      int W, int H, int N, int M, int z0)
{
    const int xq = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
    const int z = z0 + blockIdx.z;
    const int x = 4 * xq;
    if (x >= W || y >= H) return;
    const size_t pq = ((size_t)z * N + (size_t)y * W + x) / 4;
    const uint4 q4 = idx4[pq];
    const uint32_t q[4] = {q4.x, q4.y, q4.z, q4.w};
    uint32_t o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float4 a = map[2 * ((size_t)z * M + q[k])];
        const float4 pl = plane[(size_t)z * M + q[k]];
        o[k] = __float_as_uint((pl.y * (a.y - (float)(x + k)) + pl.z * (a.z - (float)y) + pl.w * pl.x) / pl.w);
    }
    pixd4[2 * pq] = make_uint4(q[0], o[0], q[1], o[1]);
    pixd4[2 * pq + 1] = make_uint4(q[2], o[2], q[3], o[3]);
}

// {L, a, b, 0} of every superpixel as one 16-byte record (the colours are fixed once SLIC is done)
__global__ void __launch_bounds__(256) k_spixl_colour(const float4 *__restrict__ map, float4 *__restrict__ spcol, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = map[2 * i], b = map[2 * i + 1];
    spcol[i] = make_float4(a.w, b.x, b.y, 0.0f);
}

int launch_spixl_colour(const FrameDev &f, int z0, int z1, cudaStream_t st)
{
    size_t n = (size_t)(z1 - z0) * f.M, o = (size_t)z0 * f.M;
    if (!n) return 0;
    k_spixl_colour<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(f.map + 2 * o, f.spcol + o, n);
    return 1;
}

int launch_pixel_plane(const FrameDev &f, int src, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    dim3 b(32, 8);
    if (f.W % 4 == 0) {
        dim3 g((f.W / 4 + 31) / 32, (f.H + 7) / 8, z1 - z0);
        k_pixel_plane4<<<g, b, 0, st>>>(f.map, (const uint4 *)f.idx, f.plane[src], (uint4 *)f.pixd, f.W, f.H, f.N, f.M, z0);
    } else {
        dim3 g((f.W + 31) / 32, (f.H + 7) / 8, z1 - z0);
        k_pixel_plane<<<g, b, 0, st>>>(f.map, f.idx, f.plane[src], f.pixd, f.W, f.H, f.N, f.M, z0);
    }
    return 1;
}

// ---------------------------------------------------------------------------------------
// 6-float API layout <-> internal plane/cost split
// ---------------------------------------------------------------------------------------
__global__ void k_export_state(const float4 *__restrict__ plane, const float2 *__restrict__ cost,
                               float *__restrict__ st6, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 p = plane[i];
    float2 c = cost[i];
    float *o = st6 + 6 * i;
    o[0] = p.x; o[1] = c.x; o[2] = c.y; o[3] = p.y; o[4] = p.z; o[5] = p.w;
}

__global__ void k_import_state(float4 *__restrict__ plane, float2 *__restrict__ cost, const float *__restrict__ st6,
                               size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *o = st6 + 6 * i;
    plane[i] = make_float4(o[0], o[3], o[4], o[5]);
    cost[i] = make_float2(o[1], o[2]);
}

int launch_export_state(const FrameDev &f, int which, int z0, int z1, cudaStream_t st)
{
    size_t n = (size_t)(z1 - z0) * f.M, o = (size_t)z0 * f.M;
    if (!n) return 0;
    k_export_state<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(f.plane[which] + o, f.cost[which] + o,
                                                                 f.state6[which] + 6 * o, n);
    return 1;
}

int launch_import_state(const FrameDev &f, int which, int z0, int z1, cudaStream_t st)
{
    size_t n = (size_t)(z1 - z0) * f.M, o = (size_t)z0 * f.M;
    if (!n) return 0;
    k_import_state<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(f.plane[which] + o, f.cost[which] + o,
                                                                 f.state6[which] + 6 * o, n);
    return 1;
}

} // namespace mvs
