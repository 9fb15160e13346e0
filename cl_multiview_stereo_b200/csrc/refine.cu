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
    int V, W, H, mw, mh, N, M, aw;
    float gamma, alpha, bl_ratio, fuse;
};

__device__ __forceinline__ RCtx make_ctx(const FrameDev &f, const RefineArgs &a)
{
    RCtx c;
    c.map = f.map; c.idx = f.idx; c.rep = f.rep; c.flat = f.flat;
    c.view_subset = f.view_subset; c.subset_num = f.subset_num;
    c.V = f.V; c.W = f.W; c.H = f.H; c.mw = f.mw; c.mh = f.mh; c.N = f.N; c.M = f.M; c.aw = f.aw;
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
    int camx = z % c.aw, camy = z / c.aw;
    Samples ss = rep_samples(c.rep[id]);
    int icx = f2i(ma.y), icy = f2i(ma.z);
    int ns = c.subset_num[z];
    for (int n = 0; n < ns; n++) {
        int view = c.view_subset[z * c.V + n];
        float rx = round_haz(d * (float)(view % c.aw - camx));
        float ry = round_haz(c.bl_ratio * d * (float)(view / c.aw - camy));
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

int launch_init_state(const FrameDev &f, const RefineArgs &a, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    dim3 g((f.M + 127) / 128, z1 - z0);
    k_init_state<<<g, 128, 0, st>>>(f, a, z0);
    return 1;
}

// ---------------------------------------------------------------------------------------
// propagate
// ---------------------------------------------------------------------------------------
struct Plane { float d, sm, cs, nx, ny, nz; };

struct Self {            // per-superpixel invariants of one propagate call
    int x, y, z;
    float cx, cy;        // centre
    float L, a, b;       // colour
    float2 fl;
    Samples ss;
    int icx, icy;
    int step_size;       // flatness-scaled far-neighbour stride of compute_smoothness
};

__device__ __forceinline__ void sm_term(const float4 *vm, const float4 *vp, int q, const Self &s, float d, float nx,
                                        float ny, float nz, float g, float alpha, float &smoothness,
                                        float &weight_norm)
{
    float4 a = vm[2 * q], b = vm[2 * q + 1];
    float diff = dist3(s.L, s.a, s.b, a.w, b.x, b.y);
    float similarity = expf_pinned(-diff * diff * g);
    float d_intrp = (nx * (s.cx - a.y) + ny * (s.cy - a.z) + nz * d) / nz;
    diff = d_intrp - vp[q].x;
    smoothness += similarity * expf_pinned(-diff * diff * alpha);
    weight_norm += similarity;
}

// compute_smoothness (clcode.cl:1407-1525)
__device__ float compute_smoothness(const RCtx &c, const float4 *__restrict__ plane_in, const Self &s, float d,
                                    float nx, float ny, float nz, int no_kernel_steps)
{
    float smoothness = 0.0f, weight_norm = 0.0f;
    const float4 *vm = c.map + 2 * (size_t)s.z * c.M;
    const float4 *vp = plane_in + (size_t)s.z * c.M;
    for (int i = -1; i <= 1; i++)
        for (int j = -1; j <= 1; j++) {
            int qx = s.x + i, qy = s.y + j;
            if (qx >= 0 && qy >= 0 && qx < c.mw && qy < c.mh && (i != 0 || j != 0))
                sm_term(vm, vp, qy * c.mw + qx, s, d, nx, ny, nz, c.gamma, c.alpha, smoothness, weight_norm);
        }
    for (int i = 1; i <= no_kernel_steps; i++) {
        float gamma_i = c.gamma * (float)(1 + i);
        int step = i * s.step_size;
        if (s.x > step) sm_term(vm, vp, c.mw * s.y + s.x - (step + 1), s, d, nx, ny, nz, gamma_i, c.alpha, smoothness, weight_norm);
        if (s.x < c.mw - step - 1) sm_term(vm, vp, c.mw * s.y + s.x + (step + 1), s, d, nx, ny, nz, gamma_i, c.alpha, smoothness, weight_norm);
        if (s.y > step) sm_term(vm, vp, c.mw * (s.y - (step + 1)) + s.x, s, d, nx, ny, nz, gamma_i, c.alpha, smoothness, weight_norm);
        if (s.y < c.mh - step - 1) sm_term(vm, vp, c.mw * (s.y + (step + 1)) + s.x, s, d, nx, ny, nz, gamma_i, c.alpha, smoothness, weight_norm);
    }
    return weight_norm > 0.0f ? smoothness / weight_norm : (float)0.000001;
}

// compute_consistency (clcode.cl:1528-1631).  The per-sample plane interpolation does not
// depend on the neighbour view, so it is evaluated once per sample; the subset row stride
// is V (the reference takes it from get_global_size(2), which equals V for its launch).
__device__ float compute_consistency(const RCtx &c, const float4 *__restrict__ plane_in, const Self &s, float d,
                                     float nx, float ny, float nz)
{
    float di[9], bdi[9];
    int sx[9], sy[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        int i = k / 3 - 1, j = k % 3 - 1;
        sx[k] = s.icx + s.ss.s[k] * i;
        sy[k] = s.icy + s.ss.s[k] * j;
        di[k] = (nx * (s.cx - (float)sx[k]) + ny * (s.cy - (float)sy[k]) + nz * d) / nz;
        bdi[k] = c.bl_ratio * di[k];
    }
    float consistency = 0.0f;
    int view_counter = 0;
    int camx = s.z % c.aw, camy = s.z / c.aw;
    int ns = c.subset_num[s.z];
    for (int n = 0; n < ns; n++) {
        int view = c.view_subset[c.V * s.z + n];
        float dvx = (float)(view % c.aw - camx), dvy = (float)(view / c.aw - camy);
        const uint32_t *vidx = c.idx + (size_t)view * c.N;
        const float4 *pm = c.map + 2 * (size_t)view * c.M;
        const float4 *pp = plane_in + (size_t)view * c.M;
        float vws = 0.0f, ows = 0.0f, num = 0.0f, visibility = 0.0f, visible = 0.0f;
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            int xp = f2i((float)sx[k] - round_haz(di[k] * dvx));
            int yp = f2i((float)sy[k] - round_haz(bdi[k] * dvy));
            if (xp >= 0 && yp >= 0 && xp < c.W && yp < c.H) {
                uint32_t q = vidx[(size_t)yp * c.W + xp];
                float4 a = pm[2 * (size_t)q], b = pm[2 * (size_t)q + 1];
                float4 pl = pp[q];
                float d_proj = (pl.y * (a.y - (float)xp) + pl.z * (a.z - (float)yp) + pl.w * pl.x) / pl.w;
                visible_term(d_proj - di[k], c.fuse, c.alpha, visible, vws, ows);
                float diff = dist3(a.w, b.x, b.y, s.L, s.a, s.b);
                visibility += expf_pinned(-diff * diff * c.gamma);
                num += 1.0f;
            }
        }
        cs_view_tail(num, vws, ows, visibility, visible, s.fl.y, consistency, view_counter);
    }
    return cs_finish(consistency, view_counter);
}

// update (clcode.cl:1635-1673): try the plane of superpixel q (same view) extrapolated to
// the own centre
__device__ __forceinline__ void try_neighbour_plane(const RCtx &c, const float4 *__restrict__ plane_in,
                                                    const Self &s, int q, int iter, int no_kernel_steps, Plane &p)
{
    const float4 *vm = c.map + 2 * (size_t)s.z * c.M;
    float4 pl = plane_in[(size_t)s.z * c.M + q];
    float4 a = vm[2 * q], b = vm[2 * q + 1];
    float d_intrp = (pl.y * (a.y - s.cx) + pl.z * (a.z - s.cy) + pl.w * pl.x) / pl.w;
    float sm1 = compute_smoothness(c, plane_in, s, d_intrp, pl.y, pl.z, pl.w, no_kernel_steps);
    float cs1 = compute_consistency(c, plane_in, s, d_intrp, pl.y, pl.z, pl.w);
    float diff = dist3(s.L, s.a, s.b, a.w, b.x, b.y);
    float similarity = expf_pinned(-diff * diff * c.gamma);
    if ((iter < 4 && sm1 * similarity > p.sm) || cs1 * sm1 > p.sm * p.cs) {
        p.d = d_intrp; p.sm = sm1; p.cs = cs1; p.nx = pl.y; p.ny = pl.z; p.nz = pl.w;
    }
}

// spatialRefinement (clcode.cl:1687-1723): plane through the own point and two
// neighbouring centres; normalize pinned as v / sqrt(x*x+y*y+z*z+0*0), zero stays zero.
__device__ __forceinline__ void try_pair_plane(const RCtx &c, const float4 *__restrict__ plane_in, const Self &s,
                                               int q1, int q2, int iter, int no_kernel_steps, Plane &p)
{
    const float4 *vm = c.map + 2 * (size_t)s.z * c.M;
    const float4 *vp = plane_in + (size_t)s.z * c.M;
    float4 m1 = vm[2 * q1], m2 = vm[2 * q2];
    float ax = m1.y - s.cx, ay = m1.z - s.cy, az = vp[q1].x - p.d;
    float bx = m2.y - s.cx, by = m2.z - s.cy, bz = vp[q2].x - p.d;
    float crx = ay * bz - az * by;
    float cry = bx * az - ax * bz;
    float crz = ax * by - ay * bx;
    float len = sqrtf(crx * crx + cry * cry + crz * crz + 0.0f * 0.0f);
    float nx = 0.0f, ny = 0.0f, nz = 0.0f;
    if (len != 0.0f) { nx = crx / len; ny = cry / len; nz = crz / len; }
    float sm1 = compute_smoothness(c, plane_in, s, p.d, nx, ny, nz, no_kernel_steps);
    float cs1 = compute_consistency(c, plane_in, s, p.d, nx, ny, nz);
    if ((iter < 4 && sm1 > p.sm) || sm1 * cs1 > p.sm * p.cs) {
        p.sm = sm1; p.cs = cs1; p.nx = nx; p.ny = ny; p.nz = nz;
    }
}

__global__ void __launch_bounds__(128) k_propagate(FrameDev f, RefineArgs ra, int src, int dst, int z0)
{
    RCtx c = make_ctx(f, ra);
    const float4 *__restrict__ plane_in = f.plane[src];
    int sp = blockIdx.x * blockDim.x + threadIdx.x;
    int z = z0 + blockIdx.y;
    if (sp >= c.M) return;
    Self s;
    s.x = sp % c.mw; s.y = sp / c.mw; s.z = z;
    size_t id = (size_t)z * c.M + sp;
    float4 ma = c.map[2 * id], mb = c.map[2 * id + 1];
    s.cx = ma.y; s.cy = ma.z; s.L = ma.w; s.a = mb.x; s.b = mb.y;
    s.fl = c.flat[id];
    s.ss = rep_samples(c.rep[id]);
    s.icx = f2i(s.cx); s.icy = f2i(s.cy);
    s.step_size = max(1, d2i((double)(s.fl.x * ra.kernel_step_size) + 0.5));
    float4 pl = plane_in[id];
    float2 co = f.cost[src][id];
    Plane p = {pl.x, co.x, co.y, pl.y, pl.z, pl.w};
    const int K = ra.no_kernel_steps, iter = ra.iter;

    // 1. planes of the 8 immediate neighbours, x offset outer / y offset inner
    for (int i = -1; i <= 1; i++)
        for (int j = -1; j <= 1; j++) {
            int qx = s.x + i, qy = s.y + j;
            if (qx >= 0 && qy >= 0 && qx < c.mw && qy < c.mh && !(i == 0 && j == 0))
                try_neighbour_plane(c, plane_in, s, qy * c.mw + qx, iter, K, p);
        }
    // 2. far neighbours UP, DOWN, LEFT, RIGHT at multiples of (int)kernel_step_size
    int stride = f2i(ra.kernel_step_size);
    for (int i = 1; i <= K; i++) {
        int offset = i * stride;
        if (s.y > offset) try_neighbour_plane(c, plane_in, s, c.mw * (s.y - (offset + 1)) + s.x, iter, K, p);
        if (s.y < c.mh - offset - 1) try_neighbour_plane(c, plane_in, s, c.mw * (s.y + offset + 1) + s.x, iter, K, p);
        if (s.x > offset) try_neighbour_plane(c, plane_in, s, c.mw * s.y + s.x - offset - 1, iter, K, p);
        if (s.x < c.mw - offset - 1) try_neighbour_plane(c, plane_in, s, c.mw * s.y + s.x + offset + 1, iter, K, p);
    }
    // 3. planes through consecutive neighbour pairs, ring W, NW, N, NE, E, SE, S, SW
    const int nbx[8] = {-1, -1, 0, 1, 1, 1, 0, -1};
    const int nby[8] = {0, -1, -1, -1, 0, 1, 1, 1};
    for (int i = 0; i < 8; i++) {
        int j = (i + 1) & 7;
        int x1 = s.x + nbx[i], y1 = s.y + nby[i], x2 = s.x + nbx[j], y2 = s.y + nby[j];
        if (x1 > -1 && y1 > -1 && x1 < c.mw && y1 < c.mh && x2 > -1 && y2 > -1 && x2 < c.mw && y2 < c.mh)
            try_pair_plane(c, plane_in, s, y1 * c.mw + x1, y2 * c.mw + x2, iter, K, p);
    }
    f.plane[dst][id] = make_float4(p.d, p.nx, p.ny, p.nz);
    f.cost[dst][id] = make_float2(p.sm, p.cs);
}

int launch_propagate(const FrameDev &f, const RefineArgs &a, int src, int dst, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    dim3 g((f.M + 127) / 128, z1 - z0);
    k_propagate<<<g, 128, 0, st>>>(f, a, src, dst, z0);
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

int launch_spixl_to_image(const FrameDev &f, int src, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    dim3 b(32, 8), g((f.W + 31) / 32, (f.H + 7) / 8, z1 - z0);
    k_spixl_to_image<<<g, b, 0, st>>>(f.map, f.idx, f.plane[src], f.disp, f.W, f.H, f.N, f.M, z0);
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
