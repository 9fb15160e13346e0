// fuse.cu -- multi-view fusion of the per-view disparity images (sm_100a).
//
// Replaces the OpenCL kernels project_to_reference_inv (reference clMVDE/clcode.cl:1995-2034) and
// remove_view_inconsistency (clcode.cl:2037-2101) together with the host sequence the reference
// keeps commented out in clDepthRefinement::fusion (depth_refinement.cpp:1382-1451): for every
// reference view r in turn, slice r of disp_proj is written, then slice r of the result is voted from
// ALL slices of disp_proj.  The reference never initialises disp_proj; the semantics pinned here (and
// in both oracles) are a zero-filled buffer, 0 being the kernels' own "no value" (clcode.cl:2056,
// 2062).  Under that schedule view r sees exactly the slices 0..r, which makes every reference view
// independent: the 2V launches of the reference become two, with the view in the grid.
#include "common.cuh"
#include "mvs_math.cuh"

namespace mvs {

// Chase the largest disparity seen along the reprojection of a pixel through every other view, in view
// order; each step reprojects with the running value (clcode.cl:2013-2031).
__global__ void __launch_bounds__(256) k_project_inv(const float *__restrict__ full, float *__restrict__ proj, int W,
                                                     int H, int N, int V, int aw, float bl_ratio, int z0)
{
    extern __shared__ float2 pi_dv[];   // [V] camera offset (ref - view) of every view, as floats
    const int pix = blockIdx.x * blockDim.x + threadIdx.x;
    const int ref = z0 + blockIdx.y;
    const int rx = ref % aw, ry = ref / aw;
    for (int i = threadIdx.x; i < V; i += blockDim.x) pi_dv[i] = make_float2((float)(rx - i % aw), (float)(ry - i / aw));
    __syncthreads();
    if (pix >= N) return;
    const int y = pix / W, x = pix - y * W;
    const float fx = (float)x, fy = (float)y;
    float m = full[(size_t)N * ref + pix];
    for (int i = 0; i < V; ++i) {
        if (i == ref) continue;
        const float2 dv = pi_dv[i];
        const int px = f2i(fx - round_haz(m * dv.x));
        const int py = f2i(fy - round_haz(bl_ratio * m * dv.y));
        if (px >= 0 && py >= 0 && px < W && py < H) {
            const float cur = full[(size_t)N * i + (size_t)W * py + px];
            if (m < cur) m = cur;
        }
    }
    proj[(size_t)N * ref + pix] = m;
}

// Vote (clcode.cl:2052-2097).  The candidates of a pixel (slices 0..ref of disp_proj) are staged in
// shared memory, one column per thread.  A candidate whose value already occurred at a lower slice is
// skipped: its stability is the same number, and an accepted value can only have raised d_est to at
// least that value, so the second visit never changes d_est (NaN never compares equal and is never
// skipped).
constexpr int RV_THREADS = 128;

__global__ void __launch_bounds__(RV_THREADS) k_remove_inconsistency(const float *__restrict__ proj,
                                                                     const float *__restrict__ full,
                                                                     float *__restrict__ out, int W, int H, int N, int V,
                                                                     int aw, float bl_ratio, float fuse, int z0)
{
    extern __shared__ float rv_s[];   // [V][RV_THREADS] candidates, then [V] float2 camera offsets (view - ref)
    float2 *rv_dv = reinterpret_cast<float2 *>(rv_s + (size_t)V * RV_THREADS);
    const int t = threadIdx.x;
    const int pix = blockIdx.x * RV_THREADS + t;
    const int ref = z0 + blockIdx.y;
    const int rx = ref % aw, ry = ref / aw;
    for (int i = t; i < V; i += RV_THREADS) rv_dv[i] = make_float2((float)(i % aw - rx), (float)(i / aw - ry));
    __syncthreads();
    if (pix >= N) return;
    const int y = pix / W, x = pix - y * W;
    const float fx = (float)x, fy = (float)y;
    const int nc = ref + 1;   // slices above ref are still zero when view ref is voted
    for (int i = 0; i < nc; ++i) rv_s[i * RV_THREADS + t] = proj[(size_t)N * i + pix];
    float d_est = 0.0f;
    for (int i = 0; i < nc; ++i) {
        const float d = rv_s[i * RV_THREADS + t];
        if (d != 0) {
            bool seen = false;
            for (int k = 0; k < i; ++k) seen |= rv_s[k * RV_THREADS + t] == d;
            if (seen) continue;
            float stability = 0.0f;
            for (int j = 0; j < nc; ++j) {
                const float dc = rv_s[j * RV_THREADS + t];
                if (dc != 0) {
                    const float diff = dc - d;
                    if (fabsf(diff) > fuse) stability -= 1;
                    if (fabsf(diff) <= fuse) stability += 1;
                }
            }
            const float bd = bl_ratio * d;
            for (int j = 0; j < V; ++j) {
                const float2 dv = rv_dv[j];
                const int px = f2i(fx - round_haz(d * dv.x));
                const int py = f2i(fy - round_haz(bd * dv.y));
                if (px >= 0 && py >= 0 && px < W && py < H) {
                    const float dc = full[(size_t)N * j + (size_t)W * py + px];
                    const float diff = dc - d;
                    if (fabsf(diff) > fuse) stability -= 1;
                    if (fabsf(diff) < fuse) stability += 1;
                }
            }
            if (stability >= 0 && (d_est == 0 || d_est < d)) d_est = d;
        }
    }
    out[(size_t)N * ref + pix] = d_est;
}

// views z0..z1-1 as reference views; `full` must hold ALL V views.  Voting view r reads the chased
// slices 0..r, so a rank that owns [z0, z1) chases 0..z1-1 itself (cheap) rather than receiving them.
int launch_view_fusion(const FrameDev &f, float bl_ratio, float fuse, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    if (!f.disp || !f.disp_proj || !f.disp_fused) return -1;
    const size_t smem = sizeof(float) * (size_t)f.V * RV_THREADS + sizeof(float2) * (size_t)f.V;
    if (smem > 200 * 1024) return -1;
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(k_remove_inconsistency, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    dim3 g1((f.N + 255) / 256, z1);
    k_project_inv<<<g1, 256, sizeof(float2) * (size_t)f.V, st>>>(f.disp, f.disp_proj, f.W, f.H, f.N, f.V, f.aw, bl_ratio, 0);
    dim3 g2((f.N + RV_THREADS - 1) / RV_THREADS, z1 - z0);
    k_remove_inconsistency<<<g2, RV_THREADS, smem, st>>>(f.disp_proj, f.disp, f.disp_fused, f.W, f.H, f.N, f.V, f.aw,
                                                         bl_ratio, fuse, z0);
    return 2;
}

// ---------------------------------------------------------------------------------------
// The same two kernels for callers that cannot allocate (mvs_view_fusion_static, capi.cu): the chased disparities
// -- the scratch between the two launches -- live in a file-scope device array instead of a cudaMalloc'd buffer.
// 4 Mi floats = 16 MiB, held by every GPU that loads the library until the process exits.
// In host code `g_fusion_scratch` is only a stand-in for the symbol, not a device address: the kernels get the
// address cudaGetSymbolAddress returns, asked once per device with that device current.
// ---------------------------------------------------------------------------------------
static __device__ __align__(256) float g_fusion_scratch[VIEW_FUSION_STATIC_MAX_FLOATS];

static float *fusion_scratch_of(int device, const char **why)
{
    static float *ptr[64] = {nullptr};   // written once per device; callers of one device are ordered by its context
    if (device < 0 || device >= 64) { *why = "device number out of range"; return nullptr; }
    if (!ptr[device]) {
        float *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_fusion_scratch);   // the caller made `device` current
        if (e != cudaSuccess) { *why = cudaGetErrorString(e); return nullptr; }
        ptr[device] = p;
    }
    return ptr[device];
}

// disp (all V views) -> fused, both device pointers of V*W*H floats; nothing is allocated.  0, or -1 with *why.
int launch_view_fusion_static(int device, cudaStream_t st, const float *disp, float *fused, int V, int W, int H, int aw,
                              float bl_ratio, float fuse, const char **why)
{
    const size_t total = (size_t)V * W * H;
    if (total > VIEW_FUSION_STATIC_MAX_FLOATS) { *why = "more pixels than the static scratch holds"; return -1; }
    const size_t smem = sizeof(float) * (size_t)V * RV_THREADS + sizeof(float2) * (size_t)V;
    if (smem > 200 * 1024) { *why = "too many views for the vote's shared-memory staging"; return -1; }
    float *proj = fusion_scratch_of(device, why);
    if (!proj) return -1;
    cudaError_t e = cudaSuccess;
    if (smem > 48 * 1024)
        e = cudaFuncSetAttribute(k_remove_inconsistency, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    // The array keeps what the last call left.  The vote of view r reads slices 0..r, all of which the chase below
    // writes in full, so no stale value can be read; the part in use is cleared all the same, so that the scratch
    // starts every call exactly as the zero-filled buffer of the allocating path does.
    if (e == cudaSuccess) e = cudaMemsetAsync(proj, 0, sizeof(float) * total, st);
    if (e != cudaSuccess) { *why = cudaGetErrorString(e); return -1; }
    const int N = W * H;
    dim3 g1((N + 255) / 256, V);
    k_project_inv<<<g1, 256, sizeof(float2) * (size_t)V, st>>>(disp, proj, W, H, N, V, aw, bl_ratio, 0);
    dim3 g2((N + RV_THREADS - 1) / RV_THREADS, V);
    k_remove_inconsistency<<<g2, RV_THREADS, smem, st>>>(proj, disp, fused, W, H, N, V, aw, bl_ratio, fuse, 0);
    e = cudaGetLastError();
    if (e != cudaSuccess) { *why = cudaGetErrorString(e); return -1; }
    return 2;
}

} // namespace mvs
