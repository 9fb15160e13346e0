// capi.cu -- extern "C" layer of the B200 depth pipeline (see include/mvs_capi.h).
//
// Owns contexts (device + stream + optional NCCL communicator) and frames (device-resident
// buffers of all views).  Host-side scalar preparation follows the reference:
// clSLIC.cpp:15-23 (SLIC normalisers), pipeline.cpp:121-142 (levels, neighbour lists),
// pipeline.cpp:164-166 + depth_refinement.cpp:130, 335-339, 734-738, 768-769 (refinement
// parameters, per-iteration arguments), depth_refinement.cpp:744-753, 1352 (ping-pong and
// the buffer fusion reads).
#include "../../include/mvs_capi.h"
#include "common.cuh"
#include <nccl.h>   // types and prototypes only: the library is bound at run time (below)
#include <dlfcn.h>
#include <math.h>
#include <stdio.h>
#include <stdarg.h>
#include <string.h>
#include <algorithm>
#include <string>
#include <vector>

using namespace mvs;

static thread_local std::string g_err;

static int fail(const char *fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return 1;
}

#define CK(call)                                                                                            \
    do {                                                                                                    \
        cudaError_t e_ = (call);                                                                            \
        if (e_ != cudaSuccess) return fail("%s:%d CUDA error %d (%s) in %s", __FILE__, __LINE__, (int)e_,  \
                                           cudaGetErrorString(e_), #call);                                  \
    } while (0)
// NCCL is bound at run time, not at link time: a host process may already carry its own NCCL
// (PyTorch bundles one), and two different NCCL builds in one address space do not mix.  The
// copy that is already loaded wins (RTLD_NOLOAD); otherwise the system library is opened.
// Single-GPU use never touches NCCL.
struct NcclApi {
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclSend) Send = nullptr;
    decltype(&ncclRecv) Recv = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    void *handle = nullptr;
    bool ok = false;
};
static NcclApi g_nccl;

static int nccl_bind()
{
    if (g_nccl.ok) return 0;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
    if (!h) return fail("NCCL is not available: %s", dlerror());
    g_nccl.handle = h;
#define BIND(name)                                                                  \
    g_nccl.name = (decltype(g_nccl.name))dlsym(h, "nccl" #name);                    \
    if (!g_nccl.name) return fail("libnccl lacks nccl" #name)
    BIND(GetUniqueId); BIND(CommInitRank); BIND(CommDestroy); BIND(GetErrorString);
    BIND(GroupStart); BIND(GroupEnd); BIND(Send); BIND(Recv); BIND(AllGather);
#undef BIND
    g_nccl.ok = true;
    return 0;
}

#define NK(call)                                                                                            \
    do {                                                                                                    \
        ncclResult_t r_ = (g_nccl.call);                                                                    \
        if (r_ != ncclSuccess) return fail("%s:%d NCCL error %d (%s) in nccl%s", __FILE__, __LINE__, (int)r_, \
                                           g_nccl.GetErrorString(r_), #call);                               \
    } while (0)

struct mvs_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    long long launches = 0;
    ncclComm_t comm = nullptr;
    int nranks = 1, rank = 0;
    cudaEvent_t user_ev[16] = {nullptr};
};

enum { ST_SLIC = 0, ST_BOUNDARY, ST_BM, ST_FLAT, ST_INIT, ST_PROP, ST_FUSION, ST_EXCH, ST_COUNT };

struct mvs_frame {
    mvs_ctx *ctx = nullptr;
    mvs_params p{};
    FrameDev d{};
    int lo = 0, hi = 0;
    std::vector<int> part;         // {lo, hi} of every rank (mvs_frame_set_partition); empty = mvs_view_partition
    std::vector<int> h_subset, h_subset_num;
    std::vector<int> h_cam_of_slot;   // camera (row-major number in the array) held by every view slot; identity by default
    bool identity_order = true;
    std::vector<float> h_levels;
    // refinement scalars (set by prepare_refine)
    float gamma_flat = 0, gamma = 0, alpha = 0, fuse = 0, kss = 0;
    int kernel_step = 0, no_prop = 0;
    float bl_ratio = 0;
    float stage_ms[ST_COUNT] = {0};
    std::vector<cudaEvent_t> ev;   // pairs (start, stop) tagged with a stage
    std::vector<int> ev_stage;
    size_t ev_used = 0;
    bool timing = false;           // per-stage CUDA events are opt-in (mvs_frame_set_timing)
    std::string timer_err;         // first failure of an event call inside a StageTimer
    bool no_shear = false;            // the level-aligned Lab copies could not be allocated
    bool halo_inputs_local = false;   // the caller uploads the halo views' input images (mvs_frame_set_halo_inputs)
    bool halo_plane0 = false;      // plane[0] of the halo views already holds the initial state (mvs_frame_init_state)
};

extern "C" const char *mvs_last_error(void) { return g_err.c_str(); }

extern "C" int mvs_device_count(int *count)
{
    CK(cudaGetDeviceCount(count));
    return 0;
}

extern "C" int mvs_ctx_create(int device, mvs_ctx **out)
{
    if (!out) return fail("mvs_ctx_create: null out");
    int n = 0;
    CK(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n) return fail("mvs_ctx_create: device %d out of range (%d CUDA devices)", device, n);
    CK(cudaSetDevice(device));
    mvs_ctx *c = new mvs_ctx;
    c->device = device;
    cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) { delete c; return fail("cudaStreamCreate: %s", cudaGetErrorString(e)); }
    *out = c;
    return 0;
}

extern "C" int mvs_ctx_destroy(mvs_ctx *c)
{
    if (!c) return 0;
    cudaSetDevice(c->device);
    if (c->comm && g_nccl.ok) g_nccl.CommDestroy(c->comm);
    for (cudaEvent_t e : c->user_ev) if (e) cudaEventDestroy(e);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return 0;
}

extern "C" int mvs_ctx_sync(mvs_ctx *c)
{
    if (!c) return fail("null ctx");
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int mvs_ctx_launch_count(mvs_ctx *c, long long *count)
{
    if (!c || !count) return fail("null argument");
    *count = c->launches;
    return 0;
}

extern "C" int mvs_ctx_stream(mvs_ctx *c, void **stream)
{
    if (!c || !stream) return fail("null argument");
    *stream = (void *)c->stream;
    return 0;
}

// CUDA-event timing on the context stream (what every launch of this context is ordered on)
extern "C" int mvs_ctx_event_record(mvs_ctx *c, int slot)
{
    if (!c || slot < 0 || slot >= 16) return fail("bad event slot");
    CK(cudaSetDevice(c->device));
    if (!c->user_ev[slot]) CK(cudaEventCreate(&c->user_ev[slot]));
    CK(cudaEventRecord(c->user_ev[slot], c->stream));
    return 0;
}

extern "C" int mvs_ctx_wait_event(mvs_ctx *c, mvs_ctx *other, int slot)
{
    if (!c || !other || slot < 0 || slot >= 16 || !other->user_ev[slot]) return fail("bad event slot");
    if (c->device != other->device) return fail("mvs_ctx_wait_event: contexts on devices %d and %d", c->device, other->device);
    CK(cudaSetDevice(c->device));
    CK(cudaStreamWaitEvent(c->stream, other->user_ev[slot], 0));
    return 0;
}

extern "C" int mvs_ctx_event_elapsed(mvs_ctx *c, int a, int b, float *ms)
{
    if (!c || !ms || a < 0 || b < 0 || a >= 16 || b >= 16 || !c->user_ev[a] || !c->user_ev[b]) return fail("bad event slot");
    CK(cudaSetDevice(c->device));
    CK(cudaEventSynchronize(c->user_ev[b]));
    CK(cudaEventElapsedTime(ms, c->user_ev[a], c->user_ev[b]));
    return 0;
}

extern "C" int mvs_comm_unique_id(void *id128)
{
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
    if (nccl_bind()) return 1;
    NK(GetUniqueId((ncclUniqueId *)id128));
    return 0;
}

extern "C" int mvs_comm_init(mvs_ctx *c, const void *id128, int nranks, int rank)
{
    if (!c) return fail("null ctx");
    CK(cudaSetDevice(c->device));
    ncclUniqueId id;
    memcpy(&id, id128, sizeof id);
    if (nccl_bind()) return 1;
    NK(CommInitRank(&c->comm, nranks, id, rank));
    c->nranks = nranks;
    c->rank = rank;
    return 0;
}

// the device's pinned math on host arrays (diagnostic): which = 0 expf, 1 expf for non-positive arguments, 2 the
// polynomial-only expf for |x| < 0.3465, 3 cube root, 4 round half away from zero, 5 float -> int (bits of the int)
extern "C" int mvs_math_eval(mvs_ctx *c, int which, const float *x, float *y, size_t n)
{
    if (!c || !x || !y) return fail("null argument");
    if (which < 0 || which > 5) return fail("mvs_math_eval: unknown function %d", which);
    CK(cudaSetDevice(c->device));
    float *dx = nullptr, *dy = nullptr;
    CK(cudaMalloc(&dx, sizeof(float) * std::max<size_t>(n, 1)));
    cudaError_t e = cudaMalloc(&dy, sizeof(float) * std::max<size_t>(n, 1));
    if (e == cudaSuccess) e = cudaMemcpyAsync(dx, x, sizeof(float) * n, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) { c->launches += launch_math_eval(which, dx, dy, n, c->stream); e = cudaGetLastError(); }
    if (e == cudaSuccess) e = cudaMemcpyAsync(y, dy, sizeof(float) * n, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(dx); cudaFree(dy);
    if (e != cudaSuccess) return fail("mvs_math_eval: %s", cudaGetErrorString(e));
    return 0;
}

extern "C" int mvs_host_alloc(void **p, size_t bytes)
{
    CK(cudaHostAlloc(p, bytes, cudaHostAllocDefault));
    return 0;
}

extern "C" int mvs_host_free(void *p)
{
    CK(cudaFreeHost(p));
    return 0;
}

// contiguous partition of V views over n ranks (first V%n ranks get one more)
extern "C" int mvs_view_partition(int V, int nranks, int rank, int *lo, int *hi)
{
    if (nranks <= 0 || rank < 0 || rank >= nranks) return fail("bad rank %d/%d", rank, nranks);
    int base = V / nranks, rem = V % nranks;
    int l = rank * base + std::min(rank, rem);
    *lo = l;
    *hi = l + base + (rank < rem ? 1 : 0);
    return 0;
}

// Contiguous partition of V weighted views over n ranks that minimises the heaviest rank; among those, the one
// with the smallest sum of squared rank weights (so no rank is left empty while another holds two views).
// A view's work grows with its number of neighbour views (3 at a corner of the camera grid, 8 inside), so equal
// view counts leave the ranks that own the interior views with up to twice the work of the others.
// lo_hi: 2*nranks ints {lo_0, hi_0, lo_1, hi_1, ...}.  Pure host code, the same table on every rank.
extern "C" int mvs_view_partition_weighted(int V, int nranks, const int *weight, int *lo_hi)
{
    if (V < 0 || nranks <= 0 || !lo_hi || (V > 0 && !weight)) return fail("mvs_view_partition_weighted: bad argument");
    std::vector<long long> pre(V + 1, 0);
    for (int v = 0; v < V; ++v) {
        if (weight[v] <= 0) return fail("weight[%d] = %d must be positive", v, weight[v]);
        pre[v + 1] = pre[v] + weight[v];
    }
    struct Cost { long long mx; double sq; };
    auto better = [](const Cost &a, const Cost &b) { return a.mx < b.mx || (a.mx == b.mx && a.sq < b.sq); };
    const Cost none{-1, 0.0};
    // best[k][i]: the first i views over the first k ranks
    std::vector<std::vector<Cost>> best(nranks + 1, std::vector<Cost>(V + 1, none));
    std::vector<std::vector<int>> cut(nranks + 1, std::vector<int>(V + 1, 0));
    best[0][0] = Cost{0, 0.0};
    for (int k = 1; k <= nranks; ++k)
        for (int i = 0; i <= V; ++i)
            for (int j = 0; j <= i; ++j) {
                if (best[k - 1][j].mx < 0) continue;
                const long long w = pre[i] - pre[j];
                const Cost c{std::max(best[k - 1][j].mx, w), best[k - 1][j].sq + (double)w * (double)w};
                if (best[k][i].mx < 0 || better(c, best[k][i])) { best[k][i] = c; cut[k][i] = j; }
            }
    for (int k = nranks, i = V; k >= 1; --k) {
        lo_hi[2 * (k - 1)] = cut[k][i];
        lo_hi[2 * (k - 1) + 1] = i;
        i = cut[k][i];
    }
    return 0;
}

// A slot order of the cameras (mvs_frame_set_view_order) and the contiguous slot ranges of nranks ranks such that the
// heaviest rank is as light as a greedy assignment plus pairwise improvement makes it.  Cost of a rank = sum over its
// cameras of (w_base + w_neighbour * neighbour views) + w_halo * (views it reads but does not own: it segments those
// itself when the halo inputs are local).  A contiguous run of row-major views cannot separate the interior cameras
// of a small array (5x3 over 8 ranks: views 6 and 7, eight neighbours each, share a rank); any SET of cameras can.
// Pure host code, the same answer on every rank.  camera_of_slot: V ints; lo_hi: 2*nranks ints.
extern "C" int mvs_view_order_balanced(int aw, int ah, int neib_hor, int neib_ver, int nranks, int w_base, int w_neighbour,
                                       int w_halo, int *camera_of_slot, int *lo_hi)
{
    if (aw <= 0 || ah <= 0 || nranks <= 0 || neib_hor < 0 || neib_ver < 0 || !camera_of_slot || !lo_hi)
        return fail("mvs_view_order_balanced: bad argument");
    if (w_base < 0 || w_neighbour < 0 || w_halo < 0) return fail("mvs_view_order_balanced: negative weight");
    const int V = aw * ah;
    std::vector<std::vector<int>> nb(V);
    for (int c = 0; c < V; ++c)
        for (int x = c % aw - neib_hor; x <= c % aw + neib_hor; ++x)
            for (int y = c / aw - neib_ver; y <= c / aw + neib_ver; ++y)
                if (x >= 0 && x < aw && y >= 0 && y < ah && y * aw + x != c) nb[c].push_back(y * aw + x);
    std::vector<int> owner(V, -1);
    auto rank_cost = [&](int r) {
        long long own = 0;
        std::vector<char> read(V, 0);
        for (int c = 0; c < V; ++c)
            if (owner[c] == r) { own += w_base + (long long)w_neighbour * (long long)nb[c].size(); for (int q : nb[c]) read[q] = 1; }
        long long halo = 0;
        for (int c = 0; c < V; ++c) if (read[c] && owner[c] != r) halo += w_halo;
        return own + halo;
    };
    struct Cost { long long mx; double sq; };
    auto total = [&]() {
        Cost t{0, 0.0};
        for (int r = 0; r < nranks; ++r) { const long long c = rank_cost(r); t.mx = std::max(t.mx, c); t.sq += (double)c * (double)c; }
        return t;
    };
    auto better = [](const Cost &a, const Cost &b) { return a.mx < b.mx || (a.mx == b.mx && a.sq < b.sq); };
    // improvement: move one camera or swap two while that helps
    auto improve = [&]() {
        Cost cur = total();
        for (int pass = 0; pass < 4 * V; ++pass) {
            bool improved = false;
            for (int a = 0; a < V && !improved; ++a) {
                const int ra = owner[a];
                for (int r = 0; r < nranks && !improved; ++r) {
                    if (r == ra) continue;
                    owner[a] = r;
                    const Cost t = total();
                    if (better(t, cur)) { cur = t; improved = true; } else owner[a] = ra;
                }
                for (int b = a + 1; b < V && !improved; ++b) {
                    if (owner[b] == ra) continue;
                    std::swap(owner[a], owner[b]);
                    const Cost t = total();
                    if (better(t, cur)) { cur = t; improved = true; } else std::swap(owner[a], owner[b]);
                }
            }
            if (!improved) break;
        }
        return cur;
    };
    // three starts, each improved, the best kept: greedy (heaviest camera first, to the rank that ends up best), the
    // even row-major ranges, the weighted row-major ranges
    std::vector<int> best_owner;
    Cost best_cost{-1, 0.0};
    for (int start = 0; start < 3; ++start) {
        if (start == 0) {
            std::fill(owner.begin(), owner.end(), -1);
            std::vector<int> order(V);
            for (int c = 0; c < V; ++c) order[c] = c;
            std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return nb[a].size() > nb[b].size(); });
            for (int c : order) {
                int best_r = 0;
                Cost bc{-1, 0.0};
                for (int r = 0; r < nranks; ++r) {
                    owner[c] = r;
                    const Cost t = total();
                    if (bc.mx < 0 || better(t, bc)) { bc = t; best_r = r; }
                }
                owner[c] = best_r;
            }
        } else {
            std::vector<int> lh(2 * nranks), w(V);
            for (int c = 0; c < V; ++c) w[c] = std::max(1, w_base + w_neighbour * (int)nb[c].size());
            if (start == 1) for (int r = 0; r < nranks; ++r) mvs_view_partition(V, nranks, r, &lh[2 * r], &lh[2 * r + 1]);
            else if (mvs_view_partition_weighted(V, nranks, w.data(), lh.data())) return 1;
            for (int r = 0; r < nranks; ++r) for (int c = lh[2 * r]; c < lh[2 * r + 1]; ++c) owner[c] = r;
        }
        const Cost c = improve();
        if (best_cost.mx < 0 || better(c, best_cost)) { best_cost = c; best_owner = owner; }
    }
    owner = best_owner;
    int s = 0;
    for (int r = 0; r < nranks; ++r) {
        lo_hi[2 * r] = s;
        for (int c = 0; c < V; ++c) if (owner[c] == r) camera_of_slot[s++] = c;
        lo_hi[2 * r + 1] = s;
    }
    return 0;
}

// ---------------------------------------------------------------------------------------
// frame
// ---------------------------------------------------------------------------------------
static size_t elem_bytes(const mvs_frame *f, int which, bool *per_pixel)
{
    switch (which) {
    case MVS_BUF_RGBX: *per_pixel = true; return 4;
    case MVS_BUF_LAB: *per_pixel = true; return 16;
    case MVS_BUF_IDX: *per_pixel = true; return 4;
    case MVS_BUF_DISP:
    case MVS_BUF_DISP_PROJ:
    case MVS_BUF_DISP_FUSED: *per_pixel = true; return 4;
    case MVS_BUF_MAP: *per_pixel = false; return 32;
    case MVS_BUF_REP: *per_pixel = false; return 8;
    case MVS_BUF_FLAT: *per_pixel = false; return 8;
    case MVS_BUF_STATE_A:
    case MVS_BUF_STATE_B: *per_pixel = false; return 24;
    }
    *per_pixel = false;
    return 0;
}

static void *buf_ptr(mvs_frame *f, int which)
{
    switch (which) {
    case MVS_BUF_RGBX: return f->d.rgbx;
    case MVS_BUF_LAB: return f->d.lab;
    case MVS_BUF_IDX: return f->d.idx;
    case MVS_BUF_DISP: return f->d.disp;
    case MVS_BUF_DISP_PROJ: return f->d.disp_proj;
    case MVS_BUF_DISP_FUSED: return f->d.disp_fused;
    case MVS_BUF_MAP: return f->d.map;
    case MVS_BUF_REP: return f->d.rep;
    case MVS_BUF_FLAT: return f->d.flat;
    case MVS_BUF_STATE_A: return f->d.state6[0];
    case MVS_BUF_STATE_B: return f->d.state6[1];
    }
    return nullptr;
}

static int set_tables(mvs_frame *f, const int *subset, const int *subset_num, const float *levels, int nlev)
{
    int V = f->d.V;
    f->h_subset.assign(subset, subset + (size_t)V * V);
    f->h_subset_num.assign(subset_num, subset_num + V);
    for (int z = 0; z < V; ++z) {
        if (subset_num[z] < 0 || subset_num[z] > V) return fail("subset_num[%d]=%d out of range", z, subset_num[z]);
        for (int k = 0; k < subset_num[z]; ++k)
            if (subset[z * V + k] < 0 || subset[z * V + k] >= V) return fail("view_subset[%d][%d] out of range", z, k);
    }
    f->d.unit_baseline = 1;
    f->d.max_neighbours = 1;
    for (int z = 0; z < V; ++z) f->d.max_neighbours = std::max(f->d.max_neighbours, subset_num[z]);
    for (int z = 0; z < V; ++z)
        for (int k = 0; k < subset_num[z]; ++k) {
            const int aw = f->d.aw, cv = f->h_cam_of_slot[subset[z * V + k]], cz = f->h_cam_of_slot[z];
            if (abs(cv % aw - cz % aw) > 1 || abs(cv / aw - cz / aw) > 1) f->d.unit_baseline = 0;
        }
    cudaStream_t st = f->ctx->stream;
    CK(cudaMemcpyAsync(f->d.view_subset, f->h_subset.data(), sizeof(int) * V * V, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(f->d.subset_num, f->h_subset_num.data(), sizeof(int) * V, cudaMemcpyHostToDevice, st));
    if (levels) {
        if (f->d.levels) CK(cudaFree(f->d.levels));
        f->h_levels.assign(levels, levels + nlev);
        CK(cudaMalloc(&f->d.levels, sizeof(float) * std::max(nlev, 1)));
        CK(cudaMemcpyAsync(f->d.levels, f->h_levels.data(), sizeof(float) * nlev, cudaMemcpyHostToDevice, st));
        f->d.num_levels = nlev;
    }
    CK(cudaStreamSynchronize(st));
    return 0;
}

// refinement scalars from the values pipeline::refine_depth_map passes (gamma_ = 2g^2 ...)
static void prepare_refine(mvs_frame *f, float gamma_, float alpha_, float fuse_, int kernel_step, int kernel_size,
                           int no_prop, float bl_ratio)
{
    f->gamma_flat = (float)(1.0 / (double)gamma_);                                  // depth_refinement.cpp:130
    f->gamma = 1 / gamma_;                                                           // :335
    f->alpha = 1 / alpha_;                                                           // :336
    f->kss = (float)std::max(1, kernel_size / kernel_step * f->d.S);                 // :338 (int math)
    f->fuse = (float)(0.5 * fuse_);                                                  // :339
    f->kernel_step = kernel_step;
    f->no_prop = no_prop;
    f->bl_ratio = bl_ratio;
}

extern "C" int mvs_frame_destroy(mvs_frame *f)
{
    if (!f) return 0;
    cudaError_t first = cudaSetDevice(f->ctx->device);
    auto keep = [&](cudaError_t e) { if (first == cudaSuccess && e != cudaSuccess) first = e; };
    keep(cudaStreamSynchronize(f->ctx->stream));
    FrameDev &d = f->d;
    void *bufs[] = {d.rgbx, d.lab, d.idx, d.idx_tmp, d.map, d.rep, d.flat, d.disp, d.pixd, d.disp_proj, d.disp_fused,
                    d.io_stage, d.view_subset, d.subset_num, d.levels, d.plane[0], d.plane[1], d.cost[0], d.cost[1],
                    d.state6[0], d.state6[1], d.edge, d.spcol, d.labT, d.labSp, d.labSm, d.cam};
    for (void *p : bufs) if (p) keep(cudaFree(p));
    for (cudaEvent_t e : f->ev) keep(cudaEventDestroy(e));
    delete d.bm_tma;
    delete d.slic_tma;
    delete f;
    if (first != cudaSuccess) return fail("mvs_frame_destroy: %s", cudaGetErrorString(first));
    return 0;
}

// camera positions of the view slots -> device
static int upload_cameras(mvs_frame *f)
{
    std::vector<int2> cam(f->d.V);
    f->identity_order = true;
    for (int v = 0; v < f->d.V; ++v) {
        const int c = f->h_cam_of_slot[v];
        cam[v] = make_int2(c % f->d.aw, c / f->d.aw);
        if (c != v) f->identity_order = false;
    }
    CK(cudaMemcpyAsync(f->d.cam, cam.data(), sizeof(int2) * cam.size(), cudaMemcpyHostToDevice, f->ctx->stream));
    CK(cudaStreamSynchronize(f->ctx->stream));
    return 0;
}

static int frame_alloc(mvs_ctx *ctx, int W, int H, int S, int aw, int ah, bool need_rgbx, bool need_lab, bool need_disp,
                       bool need_tmp, mvs_frame **out)
{
    if (!ctx || !out) return fail("null argument");
    if (W <= 0 || H <= 0 || aw <= 0 || ah <= 0) return fail("bad geometry %dx%d array %dx%d", W, H, aw, ah);
    if (S < 6) return fail("spixl_size %d < 6 is not supported (cluster_per_line = 3S/16 would be 0, clSLIC.cpp:313)", S);
    if (W < S || H < S) return fail("image %dx%d smaller than one superpixel (%d)", W, H, S);
    // The centre update sums pixel positions as exact integers; the reference sums them as floats through its tree
    // (clcode.cl:584-596, 741-754).  The two agree while every partial sum is an integer below 2^24, i.e. while
    // 9 S^2 max(W, H) < 2^24 (every combination of the named workloads: S = 8 / 16 up to 4K); beyond that the
    // reference rounds progressively and this library would differ in the last bit, so it refuses instead.
    if (9.0 * S * S * (double)std::max(W, H) >= 16777216.0)
        return fail("spixl_size %d with a %dx%d image: position sums of the centre update leave the exact range "
                    "(9 S^2 max(W,H) >= 2^24)", S, W, H);
    CK(cudaSetDevice(ctx->device));
    mvs_frame *f = new mvs_frame;
    f->ctx = ctx;
    FrameDev &d = f->d;
    d.V = aw * ah; d.W = W; d.H = H; d.S = S; d.aw = aw; d.ah = ah;
    d.mw = (int)ceilf((float)W / (float)S);   // pipeline.cpp:18-19
    d.mh = (int)ceilf((float)H / (float)S);
    d.N = W * H; d.M = d.mw * d.mh;
    d.bm_tma = new BmTma();
    d.bm_tma->state = 0;
    d.slic_tma = new SlicTma();
    d.slic_tma->state = 0;
    if ((double)d.M >= 16777216.0) { delete d.bm_tma; delete d.slic_tma; delete f; return fail("more than 2^24 superpixels per view"); }
    size_t VN = (size_t)d.V * d.N, VM = (size_t)d.V * d.M;
    cudaError_t e = cudaSuccess;
    size_t asked = 0;
    auto A = [&](void **p, size_t bytes) {
        if (e != cudaSuccess) return;
        asked += bytes;
        e = cudaMalloc(p, bytes);
        if (e == cudaSuccess) e = cudaMemsetAsync(*p, 0, bytes, ctx->stream);
    };
    if (need_rgbx) A((void **)&d.rgbx, VN * 4);
    if (need_lab) A((void **)&d.lab, VN * 16);
    A((void **)&d.idx, VN * 4);
    if (need_tmp) A((void **)&d.idx_tmp, VN * 4);
    A((void **)&d.map, VM * 32);
    A((void **)&d.rep, VM * 8);
    A((void **)&d.flat, VM * 8);
    for (int k = 0; k < 2; ++k) { A((void **)&d.plane[k], VM * 16); A((void **)&d.cost[k], VM * 8); A((void **)&d.state6[k], VM * 24); }
    if (need_disp) A((void **)&d.disp, VN * 4);
    A((void **)&d.view_subset, sizeof(int) * d.V * d.V);
    A((void **)&d.subset_num, sizeof(int) * d.V);
    A((void **)&d.cam, sizeof(int2) * d.V);
    if (e != cudaSuccess) {
        std::string why = cudaGetErrorString(e);
        cudaGetLastError();
        mvs_frame_destroy(f);
        return fail("frame allocation failed after %.1f MB (%d views of %dx%d): %s", asked / 1e6, aw * ah, W, H, why.c_str());
    }
    f->lo = 0; f->hi = d.V;
    f->h_cam_of_slot.resize(d.V);
    for (int v = 0; v < d.V; ++v) f->h_cam_of_slot[v] = v;
    if (upload_cameras(f)) { mvs_frame_destroy(f); return 1; }
    *out = f;
    return 0;
}

// neighbour lists of every view slot, pipeline.cpp:130-142: cameras x outer, y inner, self excluded -- in the
// reference's order whatever slot the cameras sit in (the sums over neighbour views run in list order)
static int neighbour_tables(mvs_frame *f, const float *levels, int nlev)
{
    const mvs_params *p = &f->p;
    const int V = f->d.V, aw = p->array_width, ah = p->array_height;
    std::vector<int> slot_of(V), sub((size_t)V * V, 0), num(V, 0);
    for (int s = 0; s < V; ++s) slot_of[f->h_cam_of_slot[s]] = s;
    for (int s = 0; s < V; s++) {
        const int i = f->h_cam_of_slot[s];
        for (int x = i % aw - p->neib_hor; x <= i % aw + p->neib_hor; x++)
            for (int y = i / aw - p->neib_ver; y <= i / aw + p->neib_ver; y++) {
                int indx = y * aw + x;
                if (x >= 0 && x < aw && y >= 0 && y < ah && indx != i) sub[(size_t)s * V + num[s]++] = slot_of[indx];
            }
    }
    return set_tables(f, sub.data(), num.data(), levels, nlev);
}

// Places the cameras of the array in another slot order: slot s of every per-view array (and of every view index of
// this API: uploads, downloads, view ranges, the partition table) holds camera camera_of_slot[s], a permutation of
// 0 .. V-1 in the row-major numbering of the array.  Lets a sharded frame give every rank ANY set of cameras as one
// contiguous slot range (sets balanced by work, tiles of the camera grid) instead of a run of row-major neighbours.
// Call it before the first upload.  The results of a camera are the same bits in any order; multi-view fusion, whose
// semantics depend on the order of the slices, wants the identity order and refuses any other.
extern "C" int mvs_frame_set_view_order(mvs_frame *f, const int *camera_of_slot)
{
    if (!f || !camera_of_slot) return fail("null argument");
    if (f->p.array_width * f->p.array_height != f->d.V) return fail("mvs_frame_set_view_order: not a frame of mvs_frame_create");
    const int V = f->d.V;
    std::vector<char> seen(V, 0);
    for (int s = 0; s < V; ++s) {
        const int c = camera_of_slot[s];
        if (c < 0 || c >= V || seen[c]) return fail("mvs_frame_set_view_order: not a permutation of the %d cameras (slot %d)", V, s);
        seen[c] = 1;
    }
    CK(cudaSetDevice(f->ctx->device));
    f->h_cam_of_slot.assign(camera_of_slot, camera_of_slot + V);
    if (upload_cameras(f)) return 1;
    f->halo_plane0 = false;
    return neighbour_tables(f, nullptr, 0);
}

extern "C" int mvs_frame_create(mvs_ctx *ctx, const mvs_params *p, int view_lo, int view_hi, mvs_frame **out)
{
    if (!p) return fail("null params");
    int V = p->array_width * p->array_height;
    if (view_lo < 0 || view_hi > V || view_lo > view_hi) return fail("bad view range [%d,%d) of %d", view_lo, view_hi, V);
    if (p->inc <= 0 || p->max_disp < p->min_disp) return fail("bad disparity range");
    if (p->kernel_step <= 0) return fail("kernel_step must be positive");
    mvs_frame *f = nullptr;
    if (frame_alloc(ctx, p->width, p->height, p->spixl_size, p->array_width, p->array_height, true, true, true,
                    p->enforce_connectivity != 0, &f))
        return 1;
    f->p = *p;
    f->lo = view_lo; f->hi = view_hi;
    // disparity levels, pipeline.cpp:121-124
    std::vector<float> lev;
    for (int i = 0; i <= (p->max_disp - p->min_disp) / p->inc; i++) lev.push_back((float)(p->min_disp + i * p->inc));
    if (neighbour_tables(f, lev.data(), (int)lev.size())) { mvs_frame_destroy(f); return 1; }
    float g = (float)(2 * pow((double)p->gamma, 2)), a = (float)(2 * pow((double)p->alpha, 2));   // pipeline.cpp:164-165
    prepare_refine(f, g, a, p->fuse, p->kernel_step, p->kernel_size / 2, p->no_prop, p->bl_ratio);
    *out = f;
    return 0;
}

extern "C" int mvs_frame_sizes(mvs_frame *f, int *V, int *W, int *H, int *mw, int *mh, int *num_levels)
{
    if (!f) return fail("null frame");
    if (V) *V = f->d.V;
    if (W) *W = f->d.W;
    if (H) *H = f->d.H;
    if (mw) *mw = f->d.mw;
    if (mh) *mh = f->d.mh;
    if (num_levels) *num_levels = f->d.num_levels;
    return 0;
}

extern "C" int mvs_frame_dev_ptr(mvs_frame *f, int which, void **dev)
{
    if (!f || !dev) return fail("null argument");
    void *p = buf_ptr(f, which);
    if (!p) return fail("buffer %d not allocated", which);
    *dev = p;
    return 0;
}

static int xfer(mvs_frame *f, int which, int v0, int v1, void *host, bool up, bool sync)
{
    if (!f || !host) return fail("null argument");
    if (v0 < 0 || v1 > f->d.V || v0 > v1) return fail("bad view range [%d,%d)", v0, v1);
    bool pp;
    size_t eb = elem_bytes(f, which, &pp);
    char *dev = (char *)buf_ptr(f, which);
    if (!eb || !dev) return fail("buffer %d not available", which);
    CK(cudaSetDevice(f->ctx->device));
    size_t per_view = eb * (pp ? (size_t)f->d.N : (size_t)f->d.M);
    cudaStream_t st = f->ctx->stream;
    bool state = which == MVS_BUF_STATE_A || which == MVS_BUF_STATE_B;
    int w = which == MVS_BUF_STATE_B ? 1 : 0;
    if (!up && state) f->ctx->launches += launch_export_state(f->d, w, v0, v1, st);
    if (up) CK(cudaMemcpyAsync(dev + per_view * v0, host, per_view * (v1 - v0), cudaMemcpyHostToDevice, st));
    else CK(cudaMemcpyAsync(host, dev + per_view * v0, per_view * (v1 - v0), cudaMemcpyDeviceToHost, st));
    if (up && state) { f->ctx->launches += launch_import_state(f->d, w, v0, v1, st); f->halo_plane0 = false; }
    if (up && which == MVS_BUF_MAP) f->halo_plane0 = false;
    if (sync) CK(cudaStreamSynchronize(st));
    return 0;
}

extern "C" int mvs_frame_upload(mvs_frame *f, int which, int v0, int v1, const void *host)
{
    return xfer(f, which, v0, v1, (void *)host, true, true);
}
extern "C" int mvs_frame_download(mvs_frame *f, int which, int v0, int v1, void *host)
{
    return xfer(f, which, v0, v1, host, false, true);
}
// asynchronous variants on the context stream (host memory should be pinned)
extern "C" int mvs_frame_upload_async(mvs_frame *f, int which, int v0, int v1, const void *host)
{
    return xfer(f, which, v0, v1, (void *)host, true, false);
}
extern "C" int mvs_frame_download_async(mvs_frame *f, int which, int v0, int v1, void *host)
{
    return xfer(f, which, v0, v1, host, false, false);
}

// ---- stage timing -----------------------------------------------------------------------
// Off by default (mvs_frame_set_timing).  Every timed stage call takes one (start, stop) event pair from the
// frame's pool; mvs_frame_stage_ms() sums them per stage and hands the pairs back.  The pool is bounded: a caller
// that loops over mvs_frame_run without ever reading the times stops being timed after MAX_TIMED_PAIRS pairs
// instead of growing without bound.  A failing event call is remembered and reported by the stage that made it.
static const size_t MAX_TIMED_PAIRS = 1024;

struct StageTimer {
    mvs_frame *f;
    bool on;
    StageTimer(mvs_frame *f_, int stage) : f(f_), on(f_->timing)
    {
        if (!on) return;
        if (f->ev_used + 2 > f->ev.size()) {
            if (f->ev.size() >= 2 * MAX_TIMED_PAIRS) { on = false; return; }
            cudaEvent_t a = nullptr, b = nullptr;
            cudaError_t e = cudaEventCreate(&a);
            if (e == cudaSuccess) e = cudaEventCreate(&b);
            if (e != cudaSuccess) {
                if (a) cudaEventDestroy(a);
                note(e, "cudaEventCreate");
                on = false;
                return;
            }
            f->ev.push_back(a); f->ev.push_back(b);
            f->ev_stage.push_back(stage); f->ev_stage.push_back(stage);
        }
        f->ev_stage[f->ev_used] = stage;
        cudaError_t e = cudaEventRecord(f->ev[f->ev_used], f->ctx->stream);
        if (e != cudaSuccess) { note(e, "cudaEventRecord"); on = false; }
    }
    ~StageTimer()
    {
        if (!on) return;
        cudaError_t e = cudaEventRecord(f->ev[f->ev_used + 1], f->ctx->stream);
        if (e != cudaSuccess) { note(e, "cudaEventRecord"); return; }   // the pair is not counted
        f->ev_used += 2;
    }
    void note(cudaError_t e, const char *what)
    {
        if (f->timer_err.empty()) f->timer_err = std::string(what) + ": " + cudaGetErrorString(e);
    }
};

extern "C" int mvs_frame_set_timing(mvs_frame *f, int enable)
{
    if (!f) return fail("null frame");
    f->timing = enable != 0;
    return 0;
}

extern "C" int mvs_frame_stage_ms(mvs_frame *f, float *ms8)
{
    if (!f || !ms8) return fail("null argument");
    CK(cudaSetDevice(f->ctx->device));
    CK(cudaStreamSynchronize(f->ctx->stream));
    for (int k = 0; k < ST_COUNT; ++k) f->stage_ms[k] = 0;
    for (size_t i = 0; i + 1 < f->ev_used; i += 2) {
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, f->ev[i], f->ev[i + 1]));
        f->stage_ms[f->ev_stage[i]] += ms;
    }
    f->ev_used = 0;
    memcpy(ms8, f->stage_ms, sizeof(float) * ST_COUNT);
    return 0;
}

static int check_launch(const char *what)
{
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail("%s: kernel launch failed: %s", what, cudaGetErrorString(e));
    return 0;
}

// after a timed stage: a failed event call inside its StageTimer is this stage's error
static int check_stage(mvs_frame *f, const char *what)
{
    if (check_launch(what)) return 1;
    if (!f->timer_err.empty()) {
        std::string m = f->timer_err;
        f->timer_err.clear();
        return fail("%s: stage timing failed: %s", what, m.c_str());
    }
    return 0;
}

// ---- multi-GPU halo exchange ---------------------------------------------------------------
// Rank r needs view v when v is a neighbour (view_subset) of one of r's own views.  Every
// rank derives the same global plan; transfers are grouped ncclSend/ncclRecv over NVLink.
// The plan is a pure host function of (V, nranks, neighbour lists) so that it can be tested
// without a GPU: transfers (view, src rank, dst rank) in issue order.
static void build_exchange_plan(int V, int n, const int *part, const int *subset, const int *subset_num,
                                std::vector<int> &triples)
{
    std::vector<int> lo(n), hi(n), owner(V);
    for (int r = 0; r < n; ++r) {
        if (part) { lo[r] = part[2 * r]; hi[r] = part[2 * r + 1]; }
        else mvs_view_partition(V, n, r, &lo[r], &hi[r]);
        for (int v = lo[r]; v < hi[r]; ++v) owner[v] = r;
    }
    for (int r = 0; r < n; ++r) {
        std::vector<char> need(V, 0);
        for (int u = lo[r]; u < hi[r]; ++u)
            for (int k = 0; k < subset_num[u]; ++k) need[subset[(size_t)u * V + k]] = 1;
        for (int v = 0; v < V; ++v)
            if (need[v] && owner[v] != r) { triples.push_back(v); triples.push_back(owner[v]); triples.push_back(r); }
    }
}

static int check_partition(int V, int nranks, const int *lo_hi)
{
    int at = 0;
    for (int r = 0; r < nranks; ++r) {
        if (lo_hi[2 * r] != at || lo_hi[2 * r + 1] < at) return fail("partition: rank %d holds [%d,%d), expected to start at %d", r, lo_hi[2 * r], lo_hi[2 * r + 1], at);
        at = lo_hi[2 * r + 1];
    }
    if (at != V) return fail("partition covers %d of %d views", at, V);
    return 0;
}

static int exchange_plan(int V, int nranks, const int *part, const int *view_subset, const int *subset_num, int *triples,
                         int cap, int *count);

extern "C" int mvs_exchange_plan(int V, int nranks, const int *view_subset, const int *subset_num, int *triples, int cap,
                                 int *count)
{
    return exchange_plan(V, nranks, nullptr, view_subset, subset_num, triples, cap, count);
}

extern "C" int mvs_exchange_plan_for(int V, int nranks, const int *lo_hi, const int *view_subset, const int *subset_num,
                                     int *triples, int cap, int *count)
{
    if (!lo_hi || nranks <= 0) return fail("mvs_exchange_plan_for: bad argument");
    if (check_partition(V, nranks, lo_hi)) return 1;
    return exchange_plan(V, nranks, lo_hi, view_subset, subset_num, triples, cap, count);
}

static int exchange_plan(int V, int nranks, const int *part, const int *view_subset, const int *subset_num, int *triples,
                         int cap, int *count)
{
    if (V <= 0 || nranks <= 0 || !view_subset || !subset_num || !count) return fail("mvs_exchange_plan: bad argument");
    for (int z = 0; z < V; ++z) {
        if (subset_num[z] < 0 || subset_num[z] > V) return fail("subset_num[%d]=%d out of range", z, subset_num[z]);
        for (int k = 0; k < subset_num[z]; ++k)
            if (view_subset[z * V + k] < 0 || view_subset[z * V + k] >= V) return fail("view_subset[%d][%d] out of range", z, k);
    }
    std::vector<int> t;
    build_exchange_plan(V, nranks, part, view_subset, subset_num, t);
    *count = (int)t.size() / 3;
    if (triples) {
        if (cap < *count) return fail("mvs_exchange_plan: %d transfers, room for %d", *count, cap);
        memcpy(triples, t.data(), sizeof(int) * t.size());
    }
    return 0;
}

// A frame takes part in the halo exchanges when its context has a communicator of more than one rank AND it holds only
// part of the views.  A frame that holds every view reads nothing it does not have (and in any partition that gives
// one rank every view, all other ranks are empty and post nothing either).
static bool sharded(const mvs_frame *f)
{
    return f->ctx->comm && f->ctx->nranks > 1 && !(f->lo == 0 && f->hi == f->d.V);
}

// `count` buffers go out in ONE group (one launch, one synchronisation with the neighbour ranks); transfers between
// a pair of ranks are matched in issue order, which is the same plan order on both sides
static int exchange_raw_multi(mvs_frame *f, int count, char *const *base, const size_t *per_view)
{
    mvs_ctx *c = f->ctx;
    if (!sharded(f)) return 0;
    std::vector<int> t;
    build_exchange_plan(f->d.V, c->nranks, f->part.empty() ? nullptr : f->part.data(), f->h_subset.data(),
                        f->h_subset_num.data(), t);
    NK(GroupStart());
    ncclResult_t bad = ncclSuccess;
    for (int b = 0; b < count && bad == ncclSuccess; ++b)
        for (size_t i = 0; i + 2 < t.size() && bad == ncclSuccess; i += 3) {
            const int v = t[i], src = t[i + 1], dst = t[i + 2];
            if (src == c->rank) bad = g_nccl.Send(base[b] + per_view[b] * v, per_view[b], ncclChar, dst, c->comm, c->stream);
            if (dst == c->rank && bad == ncclSuccess)
                bad = g_nccl.Recv(base[b] + per_view[b] * v, per_view[b], ncclChar, src, c->comm, c->stream);
        }
    ncclResult_t end = g_nccl.GroupEnd();   // the group is closed on every path
    if (bad != ncclSuccess) return fail("halo exchange: NCCL error %d (%s)", (int)bad, g_nccl.GetErrorString(bad));
    if (end != ncclSuccess) return fail("halo exchange: NCCL error %d (%s) in ncclGroupEnd", (int)end, g_nccl.GetErrorString(end));
    return 0;
}

static int exchange_raw(mvs_frame *f, char *base, size_t per_view) { return exchange_raw_multi(f, 1, &base, &per_view); }

// views this rank reads but does not own
static void halo_views(mvs_frame *f, std::vector<int> &out)
{
    int V = f->d.V;
    std::vector<char> need(V, 0);
    for (int u = f->lo; u < f->hi; ++u)
        for (int k = 0; k < f->h_subset_num[u]; ++k) need[f->h_subset[(size_t)u * V + k]] = 1;
    for (int v = 0; v < V; ++v)
        if (need[v] && (v < f->lo || v >= f->hi)) out.push_back(v);
}

// The view ranges of ALL ranks when they are not mvs_view_partition's (e.g. mvs_view_partition_weighted): every
// rank passes the same table; this rank's entry must be the range the frame was created with.
extern "C" int mvs_frame_set_partition(mvs_frame *f, int nranks, const int *lo_hi)
{
    if (!f || !lo_hi) return fail("null argument");
    mvs_ctx *c = f->ctx;
    const int n = c->comm ? c->nranks : nranks, rank = c->comm ? c->rank : -1;
    if (nranks != n) return fail("partition for %d ranks, communicator has %d", nranks, n);
    if (check_partition(f->d.V, nranks, lo_hi)) return 1;
    if (rank >= 0 && (lo_hi[2 * rank] != f->lo || lo_hi[2 * rank + 1] != f->hi))
        return fail("partition gives rank %d views [%d,%d), the frame was created for [%d,%d)", rank, lo_hi[2 * rank],
                    lo_hi[2 * rank + 1], f->lo, f->hi);
    f->part.assign(lo_hi, lo_hi + 2 * nranks);
    return 0;
}

extern "C" int mvs_frame_exchange(mvs_frame *f, int which)
{
    if (!f) return fail("null frame");
    CK(cudaSetDevice(f->ctx->device));
    StageTimer t(f, ST_EXCH);
    bool pp;
    size_t eb;
    char *base;
    if (which == MVS_BUF_STATE_A || which == MVS_BUF_STATE_B) {   // internal plane array, 16 B / superpixel
        base = (char *)f->d.plane[which == MVS_BUF_STATE_B];
        eb = 16; pp = false;
    } else {
        eb = elem_bytes(f, which, &pp);
        base = (char *)buf_ptr(f, which);
    }
    if (!eb || !base) return fail("buffer %d not available", which);
    return exchange_raw(f, base, eb * (pp ? (size_t)f->d.N : (size_t)f->d.M));
}

// ---- stages --------------------------------------------------------------------------------
extern "C" int mvs_frame_slic(mvs_frame *f)
{
    if (!f) return fail("null frame");
    if (!f->d.rgbx || !f->d.lab) return fail("frame has no image buffers");
    CK(cudaSetDevice(f->ctx->device));
    StageTimer t(f, ST_SLIC);
    cudaStream_t st = f->ctx->stream;
    int S = f->d.S;
    SlicArgs a;
    a.max_xy_dist = 1.0f / (1.4242f * S);        // clSLIC.cpp:15-18
    a.max_color_dist = 15.0f / (1.7321f * 128);
    a.max_xy_dist *= a.max_xy_dist;
    a.max_color_dist *= a.max_color_dist;
    a.weight = f->p.slic_color_weight;
    long long &L = f->ctx->launches;
    {
        const char *why = "";
        if (prepare_update_centers_tma(f->d, &why)) return fail("slic: tensor map: %s", why);
    }
    L += launch_cvt(f->d, f->lo, f->hi, st);
    L += launch_init_centers(f->d, f->lo, f->hi, st);
    if (f->p.edge_enable) {   // clSLIC.cpp:86
        if (!f->d.edge) CK(cudaMalloc(&f->d.edge, sizeof(float) * (size_t)f->d.V * f->d.N));
        L += launch_apply_edge_values(f->d, f->lo, f->hi, st);
    }
    L += launch_assoc(f->d, a, f->lo, f->hi, st);
    for (int i = 0; i < f->p.no_iter; ++i) {
        L += launch_update_centers(f->d, f->lo, f->hi, st);
        L += launch_assoc(f->d, a, f->lo, f->hi, st);
    }
    if (f->p.enforce_connectivity) {
        if (!f->d.idx_tmp) return fail("connectivity buffer not allocated");
        L += launch_connectivity(f->d, f->lo, f->hi, st);
    }
    return check_stage(f, "slic");
}

// Sharded frame, halo data flow.  Block matching of a rank's own views reads only the Lab of their neighbour views;
// labels and centres of the neighbours are first read by the refinement.  So:
//   * the INPUT images of the halo views arrive first -- uploaded by the caller like the rank's own images
//     (mvs_frame_set_halo_inputs; no transfer between ranks at all) or, failing that, sent by their owners in one
//     NCCL group before anything is computed -- and the rank converts them to Lab itself (4 B/px instead of 16 B/px,
//     the same bits by construction);
//   * ONE group after block matching carries labels and centres (initial disparity included) of the halo views.
static int halo_inputs(mvs_frame *f)
{
    if (!sharded(f)) return 0;
    CK(cudaSetDevice(f->ctx->device));
    StageTimer t(f, ST_EXCH);
    if (!f->halo_inputs_local)
        if (exchange_raw(f, (char *)f->d.rgbx, sizeof(uchar4) * (size_t)f->d.N)) return 1;
    std::vector<int> halo;
    halo_views(f, halo);
    for (int v : halo) f->ctx->launches += launch_cvt(f->d, v, v + 1, f->ctx->stream);
    return check_launch("halo cvt");
}

static int exchange_after_block_match(mvs_frame *f)
{
    if (!sharded(f)) return 0;
    CK(cudaSetDevice(f->ctx->device));
    StageTimer t(f, ST_EXCH);
    char *base[2] = {(char *)f->d.idx, (char *)f->d.map};
    const size_t per_view[2] = {sizeof(uint32_t) * (size_t)f->d.N, 32 * (size_t)f->d.M};
    return exchange_raw_multi(f, 2, base, per_view);
}

// The caller uploads the input images of the halo views itself (every view this rank's views read: the camera images
// are host data, a rank can ingest its neighbours' as well as its own), so no image crosses between ranks.
extern "C" int mvs_frame_set_halo_inputs(mvs_frame *f, int local)
{
    if (!f) return fail("null frame");
    f->halo_inputs_local = local != 0;
    return 0;
}

// views this rank reads but does not own: count, and the view numbers when `views` has room
extern "C" int mvs_frame_halo_views(mvs_frame *f, int *views, int cap, int *count)
{
    if (!f || !count) return fail("null argument");
    std::vector<int> halo;
    halo_views(f, halo);
    *count = (int)halo.size();
    if (views) {
        if (cap < *count) return fail("mvs_frame_halo_views: %d views, room for %d", *count, cap);
        for (int i = 0; i < *count; ++i) views[i] = halo[i];
    }
    return 0;
}

static void read_view_runs(mvs_frame *f, std::vector<std::pair<int, int>> &runs);

extern "C" int mvs_frame_block_match(mvs_frame *f)
{
    if (!f) return fail("null frame");
    if (!f->d.lab || !f->d.levels) return fail("frame has no Lab / level buffers");
    CK(cudaSetDevice(f->ctx->device));
    cudaStream_t st = f->ctx->stream;
    {
        StageTimer t(f, ST_BOUNDARY);
        f->ctx->launches += launch_boundary(f->d, f->lo, f->hi, st);
    }
    {
        const char *why = "";
        if (prepare_block_match_tma(f->d, f->h_levels.data(), (int)f->h_levels.size(), f->bl_ratio, &why))
            return fail("block_match: tensor map: %s", why);
        // level-aligned copies of Lab (transposed, two sheared) for every view this rank's matcher reads, unless the
        // caller forces another kernel or the memory is not there (then the gather kernel runs)
        static const bool other = [] { const char *e = getenv("MVS_BM_MODE"); return e && (!strcmp(e, "gather") || !strcmp(e, "tma")); }();
        FrameDev &d = f->d;
        if (!other && !f->no_shear) {
            const size_t nT = sizeof(float4) * (size_t)d.V * d.N, nS = sizeof(float4) * (size_t)d.V * (d.W + d.H - 1) * d.H;
            if (!d.labT) {
                cudaError_t e = cudaMalloc(&d.labT, nT);
                if (e == cudaSuccess) e = cudaMalloc(&d.labSp, nS);
                if (e == cudaSuccess) e = cudaMalloc(&d.labSm, nS);
                if (e != cudaSuccess) {   // not enough memory for the copies: stay with the gather kernel
                    cudaGetLastError();
                    cudaFree(d.labT); cudaFree(d.labSp); cudaFree(d.labSm);
                    d.labT = d.labSp = d.labSm = nullptr;
                    f->no_shear = true;
                }
            }
        }
        StageTimer t(f, ST_BM);
        if (d.labT && !other) {
            std::vector<std::pair<int, int>> runs;
            read_view_runs(f, runs);
            for (auto &r : runs) f->ctx->launches += launch_lab_shear(d, r.first, r.second, st);
        }
        f->ctx->launches += launch_block_match(f->d, f->bl_ratio, f->lo, f->hi, st);
    }
    return check_stage(f, "block_match");
}

extern "C" int mvs_frame_flatness(mvs_frame *f)
{
    if (!f) return fail("null frame");
    CK(cudaSetDevice(f->ctx->device));
    StageTimer t(f, ST_FLAT);
    f->ctx->launches += launch_flatness(f->d, f->gamma_flat, f->lo, f->hi, f->ctx->stream);
    return check_stage(f, "flatness");
}

extern "C" int mvs_frame_init_state(mvs_frame *f)
{
    if (!f) return fail("null frame");
    CK(cudaSetDevice(f->ctx->device));
    StageTimer t(f, ST_INIT);
    RefineArgs a{f->gamma, f->alpha, f->fuse, f->bl_ratio, f->kernel_step, f->kss, 0};
    f->ctx->launches += launch_init_state(f->d, a, f->lo, f->hi, f->ctx->stream);
    if (sharded(f)) {
        // the initial planes of the halo views follow from the centres this rank holds since the exchange after block
        // matching: iteration 0 needs no plane exchange
        std::vector<int> halo;
        halo_views(f, halo);
        for (int v : halo) f->ctx->launches += launch_plane_from_map(f->d, v, v + 1, f->ctx->stream);
        f->halo_plane0 = true;
    }
    return check_stage(f, "init_state");
}

// contiguous runs of the views this rank's propagate reads: its own range plus the halo views
static void read_view_runs(mvs_frame *f, std::vector<std::pair<int, int>> &runs)
{
    std::vector<int> halo;
    halo_views(f, halo);
    std::vector<char> need(f->d.V, 0);
    for (int v = f->lo; v < f->hi; ++v) need[v] = 1;
    for (int v : halo) need[v] = 1;
    for (int v = 0; v < f->d.V;) {
        if (!need[v]) { ++v; continue; }
        int e = v;
        while (e < f->d.V && need[e]) ++e;
        runs.push_back({v, e});
        v = e;
    }
}

static int propagate_impl(mvs_frame *f, int iter, unsigned long long *stats_dev)
{
    if (!f) return fail("null frame");
    if (iter < 0) return fail("negative iteration");
    CK(cudaSetDevice(f->ctx->device));
    int src = iter & 1, dst = src ^ 1;   // even: A -> B, odd: B -> A (depth_refinement.cpp:744-753)
    const bool have_halo = iter == 0 && f->halo_plane0;
    f->halo_plane0 = false;
    if (sharded(f) && !have_halo)
        if (mvs_frame_exchange(f, src ? MVS_BUF_STATE_B : MVS_BUF_STATE_A)) return 1;
    if (!f->d.pixd) CK(cudaMalloc(&f->d.pixd, sizeof(uint2) * (size_t)f->d.V * f->d.N));
    StageTimer t(f, ST_PROP);
    // per-pixel {label, input-plane disparity} of every view this rank's superpixels project into
    std::vector<std::pair<int, int>> runs;
    read_view_runs(f, runs);
    if (iter == 0 || !f->d.spcol) {   // superpixel colours as 16-byte records (fixed from SLIC on)
        if (!f->d.spcol) CK(cudaMalloc(&f->d.spcol, sizeof(float4) * (size_t)f->d.V * f->d.M));
        for (auto &r : runs) f->ctx->launches += launch_spixl_colour(f->d, r.first, r.second, f->ctx->stream);
    }
    for (auto &r : runs) f->ctx->launches += launch_pixel_plane(f->d, src, r.first, r.second, f->ctx->stream);
    // per-iteration arguments, depth_refinement.cpp:768-769 (int / int, float / int)
    RefineArgs a{f->gamma, f->alpha, f->fuse, f->bl_ratio, f->kernel_step / (iter + 1), f->kss / (float)(iter + 1), iter};
    int nl = launch_propagate(f->d, a, src, dst, f->lo, f->hi, f->ctx->stream, stats_dev);
    if (nl == -3) return fail("propagate: %d views of %d pixels exceed the 2^31 pixels one frame may hold", f->d.V, f->d.N);
    if (nl == -2) return fail("propagate: kernel_step %d gives more than 255 candidate planes per superpixel", a.no_kernel_steps);
    if (nl < 0) return fail("propagate: kernel_step %d with %d neighbour views needs more shared memory than one SM has",
                            a.no_kernel_steps, f->d.max_neighbours);
    f->ctx->launches += nl;
    return check_stage(f, "propagate");
}

extern "C" int mvs_frame_propagate(mvs_frame *f, int iter) { return propagate_impl(f, iter, nullptr); }

// One propagate iteration with work counters (a diagnostic, not part of mvs_frame_run): stats12 =
// {superpixels, candidate planes, unique planes evaluated, 0, 0 (two retired counters), (plane, view)
// items, items with samples inside the neighbour image, items with a visible sample, samples inside, samples visible}.
// How many of a superpixel's candidates are bit-identical duplicates depends on the scene; the bench
// line reports it so that a frame rate can be read against the work it stands for.
extern "C" int mvs_frame_propagate_stats(mvs_frame *f, int iter, unsigned long long *stats12)
{
    if (!f || !stats12) return fail("null argument");
    CK(cudaSetDevice(f->ctx->device));
    unsigned long long *dev = nullptr;
    CK(cudaMalloc(&dev, 12 * sizeof(unsigned long long)));
    cudaError_t e = cudaMemsetAsync(dev, 0, 12 * sizeof(unsigned long long), f->ctx->stream);
    int rc = e != cudaSuccess ? fail("cudaMemsetAsync: %s", cudaGetErrorString(e)) : propagate_impl(f, iter, dev);
    if (!rc) {
        e = cudaMemcpyAsync(stats12, dev, 12 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, f->ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(f->ctx->stream);
        if (e != cudaSuccess) rc = fail("propagate_stats: %s", cudaGetErrorString(e));
    }
    cudaFree(dev);
    return rc;
}

extern "C" int mvs_frame_fusion(mvs_frame *f)
{
    if (!f) return fail("null frame");
    if (!f->d.disp) return fail("frame has no disparity buffer");
    CK(cudaSetDevice(f->ctx->device));
    StageTimer t(f, ST_FUSION);
    // fusion always reads buffer A (depth_refinement.cpp:1352)
    f->ctx->launches += launch_spixl_to_image(f->d, 0, f->lo, f->hi, f->ctx->stream);
    return check_stage(f, "fusion");
}

// Multi-view fusion of the disparity images (clcode.cl:1995-2101; host sequence
// depth_refinement.cpp:1382-1451, disabled in the reference -- not part of mvs_frame_run).  Every view
// reads every other view's disparity image: on a sharded frame each rank first receives the images it
// does not own (4 B/px per view, grouped ncclSend/ncclRecv), then votes its own views.
extern "C" int mvs_frame_view_fusion(mvs_frame *f)
{
    if (!f) return fail("null frame");
    if (!f->d.disp) return fail("frame has no disparity buffer");
    if (!f->identity_order)
        return fail("view fusion: the slices are chased and voted in camera order (row-major), this frame holds the "
                    "cameras in another slot order (mvs_frame_set_view_order)");
    CK(cudaSetDevice(f->ctx->device));
    FrameDev &d = f->d;
    const size_t bytes = sizeof(float) * (size_t)d.V * d.N;
    if (!d.disp_proj) { CK(cudaMalloc(&d.disp_proj, bytes)); }
    if (!d.disp_fused) { CK(cudaMalloc(&d.disp_fused, bytes)); }
    cudaStream_t st = f->ctx->stream;
    mvs_ctx *c = f->ctx;
    if (sharded(f)) {
        // Transfers pair up by the partition table alone: rank s sends its own images to rank r exactly when both
        // hold views (a rank without views votes nothing and needs no images; a rank that holds every view is not
        // sharded, and then every other rank is empty and posts nothing).
        StageTimer t(f, ST_EXCH);
        std::vector<int> lo(c->nranks), hi(c->nranks);
        for (int r = 0; r < c->nranks; ++r) {
            if (f->part.empty()) mvs_view_partition(d.V, c->nranks, r, &lo[r], &hi[r]);
            else { lo[r] = f->part[2 * r]; hi[r] = f->part[2 * r + 1]; }
        }
        if (lo[c->rank] != f->lo || hi[c->rank] != f->hi)
            return fail("view fusion: frame holds views [%d,%d), the partition gives rank %d [%d,%d)", f->lo, f->hi, c->rank,
                        lo[c->rank], hi[c->rank]);
        const size_t per_view = sizeof(float) * (size_t)d.N;
        const bool mine = hi[c->rank] > lo[c->rank];
        NK(GroupStart());
        ncclResult_t bad = ncclSuccess;
        for (int r = 0; r < c->nranks && bad == ncclSuccess && mine; ++r) {
            if (r == c->rank || hi[r] <= lo[r]) continue;
            bad = g_nccl.Send((char *)d.disp + per_view * lo[c->rank], per_view * (hi[c->rank] - lo[c->rank]), ncclChar, r, c->comm, st);
            if (bad == ncclSuccess)
                bad = g_nccl.Recv((char *)d.disp + per_view * lo[r], per_view * (hi[r] - lo[r]), ncclChar, r, c->comm, st);
        }
        ncclResult_t end = g_nccl.GroupEnd();   // the group is closed on every path
        if (bad != ncclSuccess) return fail("view fusion exchange: NCCL error %d (%s)", (int)bad, g_nccl.GetErrorString(bad));
        if (end != ncclSuccess) return fail("view fusion exchange: NCCL error %d (%s) in ncclGroupEnd", (int)end, g_nccl.GetErrorString(end));
    } else if (f->lo != 0 || f->hi != d.V) {
        return fail("view fusion reads every view: frame holds views [%d,%d) of %d and has no communicator", f->lo, f->hi, d.V);
    }
    StageTimer t(f, ST_FUSION);
    CK(cudaMemsetAsync(d.disp_proj, 0, bytes, st));
    CK(cudaMemsetAsync(d.disp_fused, 0, bytes, st));
    int nl = launch_view_fusion(d, f->bl_ratio, f->fuse, f->lo, f->hi, st);
    if (nl < 0) return fail("view fusion: %d views need more shared memory than one SM has", d.V);
    f->ctx->launches += nl;
    return check_stage(f, "view_fusion");
}

extern "C" int mvs_frame_refine(mvs_frame *f)
{
    if (mvs_frame_flatness(f) || mvs_frame_init_state(f)) return 1;
    for (int i = 0; i < f->no_prop; ++i)
        if (mvs_frame_propagate(f, i)) return 1;
    return mvs_frame_fusion(f);
}

extern "C" int mvs_frame_run(mvs_frame *f)
{
    if (halo_inputs(f)) return 1;
    if (mvs_frame_slic(f)) return 1;
    if (mvs_frame_block_match(f)) return 1;
    if (exchange_after_block_match(f)) return 1;
    return mvs_frame_refine(f);
}

// ---- device-side ingest / export -------------------------------------
static int need_io_stage(mvs_frame *f)
{
    if (!f->d.io_stage) CK(cudaMalloc(&f->d.io_stage, (size_t)f->d.V * f->d.N * 4));
    return 0;
}

// loadImageIn (file_handler.cpp:6-15) for views [v0, v1): `bgr` is the packed 3-byte BGR raster of
// those views (cv::Mat order); 3 B/px cross the bus and the repack to vec3u {R,G,B,pad} runs on the device
extern "C" int mvs_frame_upload_bgr(mvs_frame *f, int v0, int v1, const uint8_t *bgr)
{
    if (!f || !bgr) return fail("null argument");
    if (!f->d.rgbx) return fail("frame has no image buffers");
    if (v0 < 0 || v1 > f->d.V || v0 > v1) return fail("bad view range [%d,%d)", v0, v1);
    CK(cudaSetDevice(f->ctx->device));
    if (need_io_stage(f)) return 1;
    cudaStream_t st = f->ctx->stream;
    CK(cudaMemcpyAsync(f->d.io_stage, bgr, (size_t)(v1 - v0) * f->d.N * 3, cudaMemcpyHostToDevice, st));
    f->ctx->launches += launch_ingest_bgr(f->d, f->d.io_stage, v0, v1, st);
    if (check_launch("ingest_bgr")) return 1;
    CK(cudaStreamSynchronize(st));
    return 0;
}

// 8-bit visualisation images, quantised on the device (1 B/px crosses the bus):
//   which = MVS_BUF_DISP / MVS_BUF_DISP_FUSED: plot_full_image (depth_refinement.cpp:1476-1497)
//   which = MVS_BUF_STATE_A / _B with element 0 (d), 1 (smoothness), 2 (consistency):
//           img_translate_state (depth_refinement.cpp:1500-1560)
//   which = MVS_BUF_MAP: img_translate (photo_consistency.cpp:414-436), the block-matching disparity s7
//   which = MVS_BUF_FLAT: img_translate_flatness (depth_refinement.cpp:299-321), ceil(flatness.s0 * 255)
extern "C" int mvs_frame_export_u8(mvs_frame *f, int which, int element, int min_disp, int max_disp, int v0, int v1,
                                   uint8_t *host)
{
    if (!f || !host) return fail("null argument");
    if (v0 < 0 || v1 > f->d.V || v0 > v1) return fail("bad view range [%d,%d)", v0, v1);
    CK(cudaSetDevice(f->ctx->device));
    if (need_io_stage(f)) return 1;
    cudaStream_t st = f->ctx->stream;
    const float mn = (float)min_disp, range = (float)(max_disp - min_disp);
    if (which == MVS_BUF_DISP || which == MVS_BUF_DISP_FUSED) {
        const float *src = (const float *)buf_ptr(f, which);
        if (!src) return fail("buffer %d not available", which);
        f->ctx->launches += launch_quantise_disp(f->d, src, f->d.io_stage, mn, range, v0, v1, st);
    } else if (which == MVS_BUF_STATE_A || which == MVS_BUF_STATE_B) {
        if (element < 0 || element > 2) return fail("state element %d: only 0 (d), 1 (smoothness), 2 (consistency)", element);
        f->ctx->launches += launch_quantise_state(f->d, which == MVS_BUF_STATE_B, element, f->d.io_stage, mn, range, v0, v1, st);
    } else if (which == MVS_BUF_MAP || which == MVS_BUF_FLAT) {
        if (!f->d.idx || !f->d.map) return fail("frame has no label / superpixel buffers");
        if (which == MVS_BUF_FLAT && !f->d.flat) return fail("frame has no flatness buffer");
        f->ctx->launches += which == MVS_BUF_MAP
                                ? launch_quantise_spixl(f->d, (const float *)f->d.map, 8, 7, 0, f->d.io_stage, mn, range, v0, v1, st)
                                : launch_quantise_spixl(f->d, (const float *)f->d.flat, 2, 0, 1, f->d.io_stage, mn, range, v0, v1, st);
    } else {
        return fail("mvs_frame_export_u8: buffer %d has no 8-bit image form", which);
    }
    if (check_launch("export_u8")) return 1;
    CK(cudaMemcpyAsync(host, f->d.io_stage, (size_t)(v1 - v0) * f->d.N, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return 0;
}

// clSLIC::draw_segmentation_lines (clSLIC.cpp:447-478) for views [v0, v1) from the frame's resident image and
// labels.  `rgbx` is in/out like the reference's out_img: its one-pixel frame and pad bytes come back unchanged.
extern "C" int mvs_frame_draw_segmentation(mvs_frame *f, int v0, int v1, uint8_t *rgbx)
{
    if (!f || !rgbx) return fail("null argument");
    if (!f->d.rgbx || !f->d.idx) return fail("frame has no image / label buffers");
    if (v0 < 0 || v1 > f->d.V || v0 > v1) return fail("bad view range [%d,%d)", v0, v1);
    CK(cudaSetDevice(f->ctx->device));
    if (need_io_stage(f)) return 1;
    cudaStream_t st = f->ctx->stream;
    const size_t bytes = (size_t)(v1 - v0) * f->d.N * 4;
    CK(cudaMemcpyAsync(f->d.io_stage, rgbx, bytes, cudaMemcpyHostToDevice, st));
    f->ctx->launches += launch_segmentation_lines(f->d, (uchar4 *)f->d.io_stage, v0, v1, st);
    if (check_launch("draw_segmentation")) return 1;
    CK(cudaMemcpyAsync(rgbx, f->d.io_stage, bytes, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return 0;
}

// ---------------------------------------------------------------------------------------
// class-level entry points (host pointers, reference semantics)
// ---------------------------------------------------------------------------------------
extern "C" int mvs_slic_run_edge(mvs_ctx *ctx, int width, int height, int spixl_size, float color_weight, int no_iter,
                                 int enforce_connectivity, int edge_enable, int count, const void *rgbx, void *lab,
                                 void *map, uint32_t *idx, float *edge)
{
    if (!rgbx) return fail("null input image");
    if (count <= 0) return fail("count must be positive");
    if (edge && !edge_enable) return fail("an edge image was asked for with edge_enable off");
    mvs_frame *f = nullptr;
    // `count` independent images = a count x 1 "array" with no neighbour relations
    if (frame_alloc(ctx, width, height, spixl_size, count, 1, true, true, false, enforce_connectivity != 0, &f)) return 1;
    f->p.slic_color_weight = color_weight;
    f->p.no_iter = no_iter;
    f->p.enforce_connectivity = enforce_connectivity;
    f->p.edge_enable = edge_enable;
    f->timing = false;
    int rc = mvs_frame_upload_async(f, MVS_BUF_RGBX, 0, count, rgbx) || mvs_frame_slic(f) ||
             (lab && mvs_frame_download_async(f, MVS_BUF_LAB, 0, count, lab)) ||
             (map && mvs_frame_download_async(f, MVS_BUF_MAP, 0, count, map)) ||
             (idx && mvs_frame_download_async(f, MVS_BUF_IDX, 0, count, idx));
    if (!rc && edge) {
        cudaError_t e = cudaMemcpyAsync(edge, f->d.edge, sizeof(float) * (size_t)count * f->d.N, cudaMemcpyDeviceToHost, ctx->stream);
        if (e != cudaSuccess) rc = fail("edge image download: %s", cudaGetErrorString(e));
    }
    rc = rc || mvs_ctx_sync(ctx);
    mvs_frame_destroy(f);
    return rc;
}

extern "C" int mvs_slic_run(mvs_ctx *ctx, int width, int height, int spixl_size, float color_weight, int no_iter,
                            int enforce_connectivity, int count, const void *rgbx, void *lab, void *map, uint32_t *idx)
{
    return mvs_slic_run_edge(ctx, width, height, spixl_size, color_weight, no_iter, enforce_connectivity, 0, count, rgbx, lab,
                             map, idx, nullptr);
}

// clSLIC::draw_segmentation_lines (clSLIC.cpp:447-478) with host arrays: `count` images + their labels in, the
// overlay written into out_rgbx (in/out: its one-pixel frame and pad bytes are kept, as in the reference)
extern "C" int mvs_draw_segmentation_run(mvs_ctx *ctx, int width, int height, int count, const uint32_t *idx,
                                         const void *in_rgbx, void *out_rgbx)
{
    if (!idx || !in_rgbx || !out_rgbx) return fail("null argument");
    if (count <= 0) return fail("count must be positive");
    mvs_frame *f = nullptr;
    // only the per-pixel image / label buffers matter here; S = 8 sizes the (unused) superpixel arrays
    if (frame_alloc(ctx, width, height, width < 8 || height < 8 ? 6 : 8, count, 1, true, false, false, false, &f)) return 1;
    int rc = mvs_frame_upload_async(f, MVS_BUF_RGBX, 0, count, in_rgbx) || mvs_frame_upload_async(f, MVS_BUF_IDX, 0, count, idx) ||
             mvs_frame_draw_segmentation(f, 0, count, (uint8_t *)out_rgbx);
    mvs_frame_destroy(f);
    return rc;
}

extern "C" int mvs_bm_run(mvs_ctx *ctx, int view_count, int width, int height, int spixl_size, float *map, uint8_t *rep,
                          const float *lab, const uint32_t *idx, int array_width, float bl_ratio, const int *view_subset,
                          const int *subset_num, const float *disp_levels, int num_levels)
{
    if (!map || !rep || !lab || !idx || !view_subset || !subset_num || !disp_levels) return fail("null argument");
    if (array_width <= 0 || view_count <= 0) return fail("bad view geometry");
    mvs_frame *f = nullptr;
    // the kernels only use array_width (view -> grid position); rows follow from the view count
    int ah = (view_count + array_width - 1) / array_width;
    if (ah * array_width != view_count) return fail("view_count %d is not a multiple of array_width %d", view_count, array_width);
    if (frame_alloc(ctx, width, height, spixl_size, array_width, ah, false, true, false, false, &f)) return 1;
    f->timing = false;
    f->bl_ratio = bl_ratio;
    int V = view_count;
    int rc = set_tables(f, view_subset, subset_num, disp_levels, num_levels) ||
             mvs_frame_upload_async(f, MVS_BUF_LAB, 0, V, lab) || mvs_frame_upload_async(f, MVS_BUF_IDX, 0, V, idx) ||
             mvs_frame_upload_async(f, MVS_BUF_MAP, 0, V, map) || mvs_frame_block_match(f) ||
             mvs_frame_download_async(f, MVS_BUF_REP, 0, V, rep) || mvs_frame_download_async(f, MVS_BUF_MAP, 0, V, map) ||
             mvs_ctx_sync(ctx);
    mvs_frame_destroy(f);
    return rc;
}

extern "C" int mvs_refine_run(mvs_ctx *ctx, int width, int height, int array_width, int array_height, int spixl_size,
                              float bl_ratio, const float *map, const uint32_t *idx, const uint8_t *rep,
                              const int *view_subset, const int *subset_num, float gamma, float alpha, float fuse,
                              int kernel_step, int kernel_size, int no_prop, float *flat, float *state_a, float *state_b,
                              float *disp)
{
    if (!map || !idx || !rep || !view_subset || !subset_num) return fail("null argument");
    if (kernel_step <= 0) return fail("kernel_step must be positive");
    mvs_frame *f = nullptr;
    if (frame_alloc(ctx, width, height, spixl_size, array_width, array_height, false, false, true, false, &f)) return 1;
    f->timing = false;
    int V = f->d.V;
    prepare_refine(f, gamma, alpha, fuse, kernel_step, kernel_size, no_prop, bl_ratio);
    int rc = set_tables(f, view_subset, subset_num, nullptr, 0) || mvs_frame_upload_async(f, MVS_BUF_MAP, 0, V, map) ||
             mvs_frame_upload_async(f, MVS_BUF_IDX, 0, V, idx) || mvs_frame_upload_async(f, MVS_BUF_REP, 0, V, rep) ||
             mvs_frame_refine(f) || (flat && mvs_frame_download_async(f, MVS_BUF_FLAT, 0, V, flat)) ||
             (state_a && mvs_frame_download_async(f, MVS_BUF_STATE_A, 0, V, state_a)) ||
             (state_b && mvs_frame_download_async(f, MVS_BUF_STATE_B, 0, V, state_b)) ||
             (disp && mvs_frame_download_async(f, MVS_BUF_DISP, 0, V, disp)) || mvs_ctx_sync(ctx);
    mvs_frame_destroy(f);
    return rc;
}

// The vote of mvs_view_fusion_run on device buffers, allocating nothing (include/mvs_capi.h): the scratch between the
// two kernels is the file-scope device array of fuse.cu.
static_assert(MVS_VIEW_FUSION_STATIC_MAX_FLOATS == VIEW_FUSION_STATIC_MAX_FLOATS, "header and kernel file agree on the limit");

extern "C" int mvs_view_fusion_static(mvs_ctx *ctx, int view_count, int width, int height, int array_width,
                                      float bl_ratio, float fuse, const float *disp_dev, float *disp_fused_dev)
{
    if (!ctx || !disp_dev || !disp_fused_dev) return fail("mvs_view_fusion_static: null argument");
    if (view_count <= 0 || width <= 0 || height <= 0 || array_width <= 0) return fail("mvs_view_fusion_static: bad geometry");
    if (view_count % array_width) return fail("view_count %d is not a multiple of array_width %d", view_count, array_width);
    const double total = (double)view_count * width * height;
    if (total > (double)MVS_VIEW_FUSION_STATIC_MAX_FLOATS)
        return fail("mvs_view_fusion_static: %d views of %dx%d are %.0f disparities, the static scratch holds %u "
                    "(use mvs_view_fusion_run or mvs_frame_view_fusion, which allocate)", view_count, width, height, total,
                    (unsigned)MVS_VIEW_FUSION_STATIC_MAX_FLOATS);
    CK(cudaSetDevice(ctx->device));
    for (const void *p : {(const void *)disp_dev, (const void *)disp_fused_dev}) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, p) != cudaSuccess || (at.type != cudaMemoryTypeDevice && at.type != cudaMemoryTypeManaged)) {
            cudaGetLastError();
            return fail("mvs_view_fusion_static: %p is not a device pointer", p);
        }
        if (at.type == cudaMemoryTypeDevice && at.device != ctx->device)
            return fail("mvs_view_fusion_static: %p lives on device %d, the context on device %d", p, at.device, ctx->device);
    }
    const char *why = "";
    const int nl = launch_view_fusion_static(ctx->device, ctx->stream, disp_dev, disp_fused_dev, view_count, width, height,
                                             array_width, bl_ratio, (float)(0.5 * fuse), &why);   // depth_refinement.cpp:1328
    if (nl < 0) return fail("mvs_view_fusion_static: %s", why);
    ctx->launches += nl;
    return 0;
}

extern "C" int mvs_view_fusion_run(mvs_ctx *ctx, int view_count, int width, int height, int array_width, float bl_ratio,
                                   float fuse, const float *disp, float *disp_proj, float *disp_fused)
{
    if (!disp) return fail("null disparity input");
    if (array_width <= 0 || view_count <= 0) return fail("bad view geometry");
    int ah = (view_count + array_width - 1) / array_width;
    if (ah * array_width != view_count) return fail("view_count %d is not a multiple of array_width %d", view_count, array_width);
    mvs_frame *f = nullptr;
    // only the per-pixel disparity buffers matter here; S = 8 sizes the (unused) superpixel arrays
    if (frame_alloc(ctx, width, height, 8, array_width, ah, false, false, true, false, &f)) return 1;
    f->timing = false;
    f->bl_ratio = bl_ratio;
    f->fuse = (float)(0.5 * fuse);   // depth_refinement.cpp:1328
    int V = view_count;
    int rc = mvs_frame_upload_async(f, MVS_BUF_DISP, 0, V, disp) || mvs_frame_view_fusion(f) ||
             (disp_proj && mvs_frame_download_async(f, MVS_BUF_DISP_PROJ, 0, V, disp_proj)) ||
             (disp_fused && mvs_frame_download_async(f, MVS_BUF_DISP_FUSED, 0, V, disp_fused)) || mvs_ctx_sync(ctx);
    mvs_frame_destroy(f);
    return rc;
}
