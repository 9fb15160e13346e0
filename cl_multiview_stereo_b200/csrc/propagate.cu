// propagate.cu -- plane propagation, the dominant kernel of the depth path (sm_100a).
//
// Replaces the OpenCL kernel `propagate` (reference clMVDE/clcode.cl:1727-1900) with its
// helpers update (:1635-1673), spatialRefinement (:1687-1723), compute_smoothness
// (:1407-1525), compute_consistency (:1528-1631) and cross_product_test (:1676-1685).
//
// What the reference does per superpixel: a sequential greedy walk over 16+ candidate planes
// (8 neighbours, up to 4K far neighbours, 8 planes through neighbour pairs); every candidate
// costs ~70 reprojected samples (9 samples x up to 8 neighbour views, dependent gathers) plus
// 8..60 smoothness terms, and is then accepted or not against the running best.
//
// What this kernel does instead (same results, bit for bit):
//  * EIGHT lanes (an "octet") share one superpixel; a CTA is 4 warps = a 4x4 tile of superpixels.
//  * The two costs of a candidate are a pure function of its plane (d, n) for the fixed input
//    state, and the candidate planes of the first two groups do not depend on the running best.
//    So the octet first GENERATES all neighbour / far-neighbour planes in parallel (one lane per
//    candidate), removes bit-identical duplicates (neighbouring superpixels often carry the same
//    plane), EVALUATES the unique planes in batches, and only then replays the reference's
//    sequential acceptance chain from the table of costs.  The pair planes keep d (clcode.cl:1716),
//    so all eight depend only on the d left by that chain: one more generate / dedup / evaluate /
//    replay round.
//  * Inside a batch the work is spread so that every float sum of the reference stays a
//    sequential chain in one lane, in the reference's order: a lane owns one (plane, neighbour
//    view) pair and walks its 9 samples in order; smoothness products are computed across the
//    octet, staged in shared memory and added in order by the lane that owns the plane.
//  * Candidate-invariant factors (colour similarity of a smoothness term, centre offsets, the
//    neighbour's disparity, the weight norm) are computed once per superpixel instead of once
//    per candidate -- same operations, same operands, same bits.
#include "common.cuh"
#include "mvs_math.cuh"
#include <stdlib.h>
#include <string.h>
#include <algorithm>

namespace mvs {

namespace {

constexpr int PR_WARPS = 4;                 // warps per CTA (tile rows)
constexpr int PR_OCT = 4;                   // octets per warp (tile columns)
constexpr int PR_SPB = PR_WARPS * PR_OCT;   // superpixels per CTA
constexpr int PR_UB = 4;                    // unique planes evaluated per batch
// build-time knob (MVS_NVCC_DEFS): resident CTAs the register allocation aims for.  Where the colour of the superpixel
// under a projected sample comes from was a knob too; the two alternatives to the 16-byte colour record (a
// per-superpixel cache of similarities in shared memory, a per-pixel colour image) were measured slower and are gone
// (DESIGN 4.1).
#ifndef PR_MINB
#define PR_MINB 6
#endif

struct PArgs {
    const float4 *map;        // [V][M][2]
    const uint32_t *idx;      // [V][N]
    const uint2 *rep;         // [V][M]
    const float2 *flat;       // [V][M]
    const int *view_subset, *subset_num;
    const uint2 *pixd;        // [V][N] {label, input-plane disparity at the pixel} (k_pixel_plane)
    const float4 *spcol;      // [V][M] {L, a, b, 0} of every superpixel: half the bytes and half the cache lines of the
                              //        32-byte centre records for the colour gather of the consistency walk
    const float4 *plane_in;   // [V][M] {d,nx,ny,nz}
    const float2 *cost_in;    // [V][M] {sm,cs}
    float4 *plane_out;
    float2 *cost_out;
    const int2 *cam;          // [V] position of every view's camera in the array
    int V, W, H, mw, mh, N, M;
    float gamma, alpha, bl_ratio, fuse;
    int K;                    // no_kernel_steps of this iteration
    float kernel_step_size;
    int iter, z0;
    int tmax;                 // smoothness terms per superpixel at most (8 + 4K)
    int ncmax;                // candidate slots per superpixel at most (valid neighbours + far + 8 pairs)
    int nsmax;                // neighbour views per view at most (8 on a grid with neib_hor = neib_ver = 1)
    int hp;                   // planes whose smoothness products are staged at a time (phases G, H): PR_UB when their
                              // products fit the storage of the sample table, else 2
    unsigned long long *stats;   // optional work counters (mvs_frame_propagate_stats)
};

// shared-memory record of one superpixel (dynamic part sized by tmax / ncmax)
struct SpLayout {
    int meta, tab, cp, cc, cw, own, uniq, sres, stage, total;
};

// CTA layout: float4 viewt[nsmax] (one table: every superpixel of a CTA belongs to the same view), then one
// record of L.total bytes per superpixel
__host__ __device__ inline int sp_hp(int tmax) { return 4 * PR_UB * (tmax + 1) <= 16 * PR_UB * 9 ? PR_UB : 2; }

__host__ __device__ inline SpLayout sp_layout(int tmax, int ncmax, int nsmax)
{
    SpLayout L;
    int o = 9 * 16;                           // samp[9]
    L.meta = o;  o += 16;                     // {L, a, b} of the own centre (read by whichever lane samples for it)
    L.tab = o;   o += 16 * tmax;              // float4 tab[tmax]
    L.cp = o;    o += 16 * ncmax;             // float4 candidate planes
    L.cc = o;    o += 8 * ncmax;              // float2 costs of the unique ones
    L.sres = o;  o += 8 * PR_UB * nsmax;      // float2 per (plane, view): {term, flags}
    const int prod_bytes = 4 * sp_hp(tmax) * (tmax + 1), sdi_bytes = 16 * PR_UB * 9;
    L.stage = o; o += ((prod_bytes > sdi_bytes ? prod_bytes : sdi_bytes) + 15) & ~15;   // sdi (E,F) then prod (G,H)
    L.cw = o;    o += 4 * ncmax;              // float acceptance weight (colour similarity to the source)
    L.own = o;   o += ncmax;                  // u8: slot whose costs this candidate takes
    L.uniq = o;  o += ncmax;                  // u8: slots of the unique planes, in order
    L.total = (o + 15) & ~15;
    return L;
}

// in-range test + linear index of the map cell (x, y)
__device__ __forceinline__ bool cell(const PArgs &a, int x, int y, int &q)
{
    q = y * a.mw + x;
    return x >= 0 && y >= 0 && x < a.mw && y < a.mh;
}

// near neighbour number s = 0..7 in the reference's loop order: x offset outer, y offset
// inner, centre skipped (clcode.cl:1413-1416, 1768-1770)
__device__ __forceinline__ void near_offset(int s, int &i, int &j)
{
    int t = s < 4 ? s : s + 1;
    i = t / 3 - 1;
    j = t % 3 - 1;
}

// the 32-byte superpixel record {id,cx,cy,L | a,b,n,disp} as ONE 256-bit gather (LDG.E.256, sm_100):
// one L1 request per lane instead of two
struct Rec8 { float id, cx, cy, L, a, b, n, disp; };
__device__ __forceinline__ Rec8 ld_rec8(const float4 *p)
{
    Rec8 r;
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r.id), "=f"(r.cx), "=f"(r.cy), "=f"(r.L), "=f"(r.a), "=f"(r.b), "=f"(r.n), "=f"(r.disp)
                 : "l"(p));
    return r;
}

__device__ __forceinline__ bool same_bits(const float4 &p, const float4 &q)
{
    return __float_as_int(p.x) == __float_as_int(q.x) && __float_as_int(p.y) == __float_as_int(q.y) &&
           __float_as_int(p.z) == __float_as_int(q.z) && __float_as_int(p.w) == __float_as_int(q.w);
}

} // namespace

// UNIT: every neighbour view is at most one camera apart in x and y (neib_hor, neib_ver <= 1).
// Then round(d * dv) == dv * round(d) exactly (dv in {-1, 0, 1}; holds for inf / NaN too), and
// the two rounds of a sample are taken once per (plane, sample) instead of once per view.
// NP: gamma >= 0 and alpha >= 0 (checked by the launcher), so every exp argument -(d*d)*w is <= 0 or NaN
// PIXD: the consistency walk gathers the per-pixel {label, plane disparity} record of k_pixel_plane and keeps the
// colour similarities of the superpixels it meets in a small per-superpixel cache (false: the walk of the first
// round of this kernel, label -> centre record + plane, kept for A/B measurement: MVS_PROP_MODE=legacy)
// SV: fuse^2 * alpha < 0.3465, so the exp of a VISIBLE sample (|diff| < fuse) has its argument in the interval where
// expf_pinned reduces to its polynomial (expf_pinned_small: same bits, a third of the instructions)
template <bool UNIT, int MINB, bool NP, bool PIXD, bool STATS, bool SV>
__global__ void __launch_bounds__(PR_WARPS * 32, MINB) k_propagate(PArgs a)
{
    extern __shared__ __align__(16) unsigned char pr_smem[];
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int l8 = lane & 7, obase = lane & 24;
    const unsigned lower8 = (1u << l8) - 1u;
    // grid = (column tiles, views, row tiles): the views of one band of rows run together, so the label rows
    // and records a band gathers from its neighbour views are still in L2 when those views reach the band
    const int x = blockIdx.x * PR_OCT + (lane >> 3), y = blockIdx.z * PR_WARPS + warp, z = a.z0 + blockIdx.y;
    const bool live = x < a.mw && y < a.mh;

    const SpLayout L = sp_layout(a.tmax, a.ncmax, a.nsmax);
    float4 *viewt = reinterpret_cast<float4 *>(pr_smem);         // {view * N (int bits), dvx, dvy, view * M (int bits)}
    unsigned char *sp_smem = pr_smem + 16 * a.nsmax;
    unsigned char *sbase = sp_smem + (size_t)(warp * PR_OCT + (lane >> 3)) * L.total;
    float4 *samp = reinterpret_cast<float4 *>(sbase);            // {sx, sy (integers held as float), cx - sx, cy - sy}
    float4 *tab = reinterpret_cast<float4 *>(sbase + L.tab);     // {cx-cxq, cy-cyq, d_q, similarity}
    float4 *cp = reinterpret_cast<float4 *>(sbase + L.cp);       // candidate planes {d, nx, ny, nz}
    float2 *cc = reinterpret_cast<float2 *>(sbase + L.cc);       // {smoothness, consistency}
    float *cw = reinterpret_cast<float *>(sbase + L.cw);
    unsigned char *own = sbase + L.own, *uniq = sbase + L.uniq;
    float2 *sres = reinterpret_cast<float2 *>(sbase + L.sres);
    float4 *sdi = reinterpret_cast<float4 *>(sbase + L.stage);   // per (plane, sample): {d_intrp, bl*d_intrp, round, round}
    float *prod = reinterpret_cast<float *>(sbase + L.stage);    // per (plane, term), row stride tmax + 1
    const int pstride = a.tmax + 1;

    const size_t zM = (size_t)z * a.M;
    const float4 *vm = a.map + 2 * zM;
    const float4 *vp = a.plane_in + zM;
    const int sp = live ? y * a.mw + x : 0;

    // ---- own record (the octet's lanes read the same addresses: one transaction) ----
    const float4 ma = vm[2 * sp], mb = vm[2 * sp + 1];
    const float cx = ma.y, cy = ma.z, cL = ma.w, ca = mb.x, cb = mb.y;
    const float2 fl = a.flat[zM + sp];
    if (l8 == 0) *reinterpret_cast<float4 *>(sbase + L.meta) = make_float4(cL, ca, cb, 0.0f);
    const int ns = live ? a.subset_num[z] : 0;
    const int2 cz = a.cam[z];
    const int camx = cz.x, camy = cz.y;
    const double half_fly = 0.5 * (double)fl.y;   // clcode.cl:1621 is a double expression
    const int nsw = a.subset_num[z];   // warp-uniform (ns is 0 for octets outside the map)

    // sample table: sample k = (i+1)*3 + (j+1) sits s_k pixels along ray (i, j) from the
    // truncated centre (clcode.cl:1547-1566); rep bytes are {nw,w,sw,n,s,ne,e,se}
    {
        const uint2 r = a.rep[zM + sp];
        const unsigned long long r64 = ((unsigned long long)r.y << 32) | r.x;
        const int icx = f2i(cx), icy = f2i(cy);
        for (int k = l8; k < 9; k += 8) {
            int s = k == 4 ? 0 : (int)((r64 >> (8 * (k < 4 ? k : k - 1))) & 255);
            int sx = icx + s * (k / 3 - 1), sy = icy + s * (k % 3 - 1);
            // |sx|, |sy| < 2^24: the float copies are exact
            samp[k] = make_float4((float)sx, (float)sy, cx - (float)sx, cy - (float)sy);
        }
    }
    // the CTA's view table: the first warp fills it, the block barrier publishes it (no lane has left the kernel:
    // octets outside the map stay as idle members of their warp)
    if (warp == 0)
        for (int n = lane; n < nsw; n += 32) {
            int view = a.view_subset[a.V * z + n];
            const int2 cv = a.cam[view];
            viewt[n] = make_float4(__int_as_float(view * a.N), (float)(cv.x - camx), (float)(cv.y - camy),
                                   __int_as_float(view * a.M));
        }
    __syncthreads();

    // ---- smoothness term table (candidate-invariant part of compute_smoothness) ----
    // slot order = the reference's accumulation order: 8 near neighbours, then for
    // i = 1..K the far neighbours x-, x+, y-, y+ at i * step_size (clcode.cl:1413-1518);
    // only in-range slots are kept, compacted in order.
    int T = 0;
    {
        const int step_size = max(1, d2i((double)(fl.x * a.kernel_step_size) + 0.5));   // clcode.cl:1437
        const int nslots = 8 + 4 * a.K;
        for (int s0 = 0; s0 < nslots; s0 += 8) {
            const int s = s0 + l8;
            bool ok = false;
            int q = 0;
            float g = a.gamma;
            if (live && s < nslots) {
                if (s < 8) {
                    int i, j;
                    near_offset(s, i, j);
                    ok = cell(a, x + i, y + j, q);
                } else {
                    const int ii = 1 + ((s - 8) >> 2), dir = (s - 8) & 3;
                    const int step = ii * step_size;
                    g = a.gamma * (float)(1 + ii);
                    if (dir == 0) { ok = x > step; q = a.mw * y + x - (step + 1); }
                    else if (dir == 1) { ok = x < a.mw - step - 1; q = a.mw * y + x + (step + 1); }
                    else if (dir == 2) { ok = y > step; q = a.mw * (y - (step + 1)) + x; }
                    else { ok = y < a.mh - step - 1; q = a.mw * (y + (step + 1)) + x; }
                }
            }
            const unsigned m8 = (__ballot_sync(FULL, ok) >> obase) & 255u;
            if (ok) {
                const float4 qa = vm[2 * q], qb = vm[2 * q + 1];
                const float diff = dist3(cL, ca, cb, qa.w, qb.x, qb.y);
                const float sim = expf_pinned<NP>(-diff * diff * g);
                tab[T + __popc(m8 & lower8)] = make_float4(cx - qa.y, cy - qa.z, vp[q].x, sim);
            }
            T += __popc(m8);
        }
    }

    // ---- candidate planes of update(): neighbours, then far neighbours (clcode.cl:1768-1850) ----
    // slot s: 0..7 near neighbours (x offset outer); 8 + 4(i-1) + {0,1,2,3} = UP, DOWN, LEFT, RIGHT
    // at i * (int)kernel_step_size.  One lane per slot; valid ones are compacted in order.
    int NU = 0;
    {
        const int stride = f2i(a.kernel_step_size);
        const bool far_possible = stride < max(a.mw, a.mh);
        const int nslots = far_possible ? 8 + 4 * a.K : 8;
        for (int s0 = 0; s0 < nslots; s0 += 8) {
            const int s = s0 + l8;
            bool ok = false;
            int q1 = 0;
            if (live && s < nslots) {
                if (s < 8) {
                    int i, j;
                    near_offset(s, i, j);
                    ok = cell(a, x + i, y + j, q1);
                } else {
                    const int off = (1 + ((s - 8) >> 2)) * stride, dir = (s - 8) & 3;
                    if (dir == 0) { ok = y > off; q1 = a.mw * (y - (off + 1)) + x; }
                    else if (dir == 1) { ok = y < a.mh - off - 1; q1 = a.mw * (y + off + 1) + x; }
                    else if (dir == 2) { ok = x > off; q1 = a.mw * y + x - off - 1; }
                    else { ok = x < a.mw - off - 1; q1 = a.mw * y + x + off + 1; }
                }
            }
            const unsigned m8 = (__ballot_sync(FULL, ok) >> obase) & 255u;
            if (ok) {   // neighbour's plane extrapolated to the own centre (clcode.cl:1643-1651)
                const float4 pl = vp[q1];
                const float4 qa = vm[2 * q1], qb = vm[2 * q1 + 1];
                const float cd = (pl.y * (qa.y - cx) + pl.z * (qa.z - cy) + pl.w * pl.x) / pl.w;
                const float diff = dist3(cL, ca, cb, qa.w, qb.x, qb.y);
                const int pos = NU + __popc(m8 & lower8);
                cp[pos] = make_float4(cd, pl.y, pl.z, pl.w);
                cw[pos] = expf_pinned<NP>(-diff * diff * a.gamma);
            }
            NU += __popc(m8);
        }
    }
    __syncwarp();

    float weight_norm = 0.0f;
    for (int t = 0; t < T; ++t) weight_norm += tab[t].w;

    // ---- evaluation of the unique planes uniq[u0 .. u1) in batches of PR_UB ----
    // (u0, u1 differ between the octets of a warp; all barriers are warp-wide)
    auto evaluate = [&](int u0, int u1) {
        for (int b0 = u0; __any_sync(FULL, b0 < u1); b0 += PR_UB) {
            const int nb = min(max(u1 - b0, 0), PR_UB);
            // E: plane disparity at the 9 samples (clcode.cl:1571-1573), once per (plane, sample)
            for (int e = l8; e < nb * 9; e += 8) {
                const int u = e / 9, k = e - 9 * u;
                const float4 pl = cp[uniq[b0 + u]];
                const float4 sv = samp[k];
                const float di = (pl.y * sv.z + pl.z * sv.w + pl.w * pl.x) / pl.w;
                const float bdi = a.bl_ratio * di;
                sdi[e] = make_float4(di, bdi, round_haz(di), round_haz(bdi));
            }
            __syncwarp();
            // F: compute_consistency, one lane per (plane, neighbour view), samples in order
            // The (plane, view) items of the four octets are pooled over the 32 lanes of the warp: an octet with
            // few unique planes lends its lanes to the others (all octets of a CTA belong to one view: same ns).
            const int nb0 = __shfl_sync(FULL, nb, 0), nb1 = __shfl_sync(FULL, nb, 8);
            const int nb2 = __shfl_sync(FULL, nb, 16), nb3 = __shfl_sync(FULL, nb, 24);
            // Items are ordered VIEW-major over the whole warp: neighbouring lanes walk the planes of the four
            // (adjacent) superpixels in the SAME neighbour view, so their label / record / plane gathers fall into
            // the same few cache lines (15.2 -> 14.7 ms per frame when done per octet, further when done per warp).
            const int q1 = nb0 + nb1, q2 = q1 + nb2, nbt = q2 + nb3;
            const unsigned nbt_inv = 16777216u / (unsigned)max(nbt, 1) + 1u;   // it / nbt == (it * nbt_inv) >> 24, it * nbt < 2^24
            for (int it = lane; it < nbt * nsw; it += 32) {
                const int n = (int)(((unsigned long long)(unsigned)it * nbt_inv) >> 24), r = it - n * nbt;
                const int oc = (r >= nb0) + (r >= q1) + (r >= q2);
                const int u = r - (oc == 0 ? 0 : oc == 1 ? nb0 : oc == 2 ? q1 : q2);
                const unsigned char *ob = sp_smem + (size_t)(warp * PR_OCT + oc) * L.total;
                const float4 *samp = reinterpret_cast<const float4 *>(ob);
                const float4 *sdi = reinterpret_cast<const float4 *>(ob + L.stage);
                float2 *sres = reinterpret_cast<float2 *>(const_cast<unsigned char *>(ob) + L.sres);
                const float4 own = *reinterpret_cast<const float4 *>(ob + L.meta);
                const float cL = own.x, ca = own.y, cb = own.z;
                const float4 vt = viewt[n];
                const uint32_t nbase = (uint32_t)__float_as_int(vt.x), mbase = (uint32_t)__float_as_int(vt.w);
                float num = 0.0f, vws = 0.0f, ows = 0.0f, visibility = 0.0f, visible = 0.0f;
                if (PIXD) {
                    // Per sample ONE 8-byte gather: the label at the projected pixel and the disparity of that
                    // superpixel's input plane there (k_pixel_plane) -- no centre / plane gathers, no division.
                    // The colour of that superpixel (for the similarity to the own colour) is a second, dependent
                    // gather of its 16-byte colour record.
                    int n_vis = 0;
#pragma unroll 1
                    for (int k0 = 0; k0 < 9; k0 += 3) {
                        int pix[3];
                        float dik[3];
                        bool inb[3];
#pragma unroll
                        for (int j = 0; j < 3; ++j) {
                            const float4 sd = sdi[u * 9 + k0 + j];
                            const float4 sv = samp[k0 + j];
                            int xp, yp;
                            if (UNIT) {
                                xp = f2i(sv.x - sd.z * vt.y);
                                yp = f2i(sv.y - sd.w * vt.z);
                            } else {
                                xp = f2i(sv.x - round_haz(sd.x * vt.y));   // clcode.cl:1576-1577
                                yp = f2i(sv.y - round_haz(sd.y * vt.z));
                            }
                            inb[j] = (unsigned)xp < (unsigned)a.W && (unsigned)yp < (unsigned)a.H;
                            pix[j] = inb[j] ? yp * a.W + xp : 0;
                            dik[j] = sd.x;
                        }
                        uint2 pr[3];
#pragma unroll
                        for (int j = 0; j < 3; ++j) pr[j] = __ldg(a.pixd + nbase + (uint32_t)pix[j]);
                        float4 col[3];
#pragma unroll
                        for (int j = 0; j < 3; ++j) col[j] = __ldg(a.spcol + mbase + pr[j].x);
#pragma unroll
                        for (int j = 0; j < 3; ++j)
                            if (inb[j]) {
                                const float diff = __uint_as_float(pr[j].y) - dik[j];
                                if (fabsf(diff) < a.fuse) {
                                    const float ea = -diff * diff * a.alpha;
                                    visible += SV ? expf_pinned_small(ea) : expf_pinned<NP>(ea);
                                    vws += 1.0f;
                                    if (STATS) n_vis++;
                                }
                                else { if (diff != diff) visible += diff; ows += 1.0f; }
                                const float cdiff = dist3(col[j].x, col[j].y, col[j].z, cL, ca, cb);
                                const float sim = expf_pinned<NP>(-cdiff * cdiff * a.gamma);
                                visibility += sim;
                                num += 1.0f;
                            }
                    }
                    if (STATS) {   // (plane, view) items: all / with samples / with a visible sample; samples: inside / visible
                        atomicAdd(a.stats + 5, 1ull);
                        if (num > 0.0f) atomicAdd(a.stats + 6, 1ull);
                        if (vws > 0.0f) atomicAdd(a.stats + 7, 1ull);
                        atomicAdd(a.stats + 8, (unsigned long long)num);
                        atomicAdd(a.stats + 9, (unsigned long long)n_vis);
                    }
                } else {
                    // three samples at a time: their label gathers, then their centre / plane gathers, are
                    // issued together (a sample that projects outside reads pixel 0 and is masked afterwards)
#pragma unroll 1
                    for (int k0 = 0; k0 < 9; k0 += 3) {
                        int pix[3];
                        float fxp[3], fyp[3], dik[3];
                        bool inb[3];
#pragma unroll
                        for (int j = 0; j < 3; ++j) {
                            const float4 sd = sdi[u * 9 + k0 + j];
                            const float4 sv = samp[k0 + j];
                            int xp, yp;
                            if (UNIT) {
                                xp = f2i(sv.x - sd.z * vt.y);
                                yp = f2i(sv.y - sd.w * vt.z);
                            } else {
                                xp = f2i(sv.x - round_haz(sd.x * vt.y));   // clcode.cl:1576-1577
                                yp = f2i(sv.y - round_haz(sd.y * vt.z));
                            }
                            inb[j] = (unsigned)xp < (unsigned)a.W && (unsigned)yp < (unsigned)a.H;
                            pix[j] = inb[j] ? yp * a.W + xp : 0;
                            fxp[j] = (float)xp; fyp[j] = (float)yp; dik[j] = sd.x;
                        }
                        uint32_t vq[3];
#pragma unroll
                        for (int j = 0; j < 3; ++j) vq[j] = mbase + a.idx[nbase + (uint32_t)pix[j]];
                        float qcx[3], qcy[3], qL[3], qA[3], qB[3];
                        float4 pl[3];
#pragma unroll
                        for (int j = 0; j < 3; ++j) {
                            // one 256-bit gather instead of a 128-bit + a 64-bit one: 17.8 -> 16.1 ms per frame
                            const Rec8 r = ld_rec8(a.map + 2 * vq[j]);
                            qcx[j] = r.cx; qcy[j] = r.cy; qL[j] = r.L; qA[j] = r.a; qB[j] = r.b;
                            pl[j] = a.plane_in[vq[j]];
                        }
#pragma unroll
                        for (int j = 0; j < 3; ++j)
                            if (inb[j]) {
                                const float d_proj = (pl[j].y * (qcx[j] - fxp[j]) + pl[j].z * (qcy[j] - fyp[j]) + pl[j].w * pl[j].x) / pl[j].w;
                                const float diff = d_proj - dik[j];
                                // visible += when_visible * exp(..): 0 * exp is +0 unless diff is NaN
                                if (fabsf(diff) < a.fuse) { visible += expf_pinned<NP>(-diff * diff * a.alpha); vws += 1.0f; }
                                else { if (diff != diff) visible += diff; ows += 1.0f; }
                                const float cdiff = dist3(qL[j], qA[j], qB[j], cL, ca, cb);
                                visibility += expf_pinned<NP>(-cdiff * cdiff * a.gamma);
                                num += 1.0f;
                            }
                    }
                }
                float term = 0.0f;
                int flags = 0;   // bit0: view has samples, bit1: some visible, bit2: some occluded
                if (num > 0.0f) {
                    flags = 1;
                    if (vws > 0.0f) { flags |= 2; term = (vws / num) * (visibility / vws) * (visible / vws); }
                    if (ows > 0.0f) flags |= 4;
                }
                sres[u * a.nsmax + n] = make_float2(term, __int_as_float(flags));
            }
            __syncwarp();   // sdi is dead from here on: prod reuses its storage
            // G + H, a.hp planes at a time: their products share the storage sdi had.  With many smoothness terms
            // (iteration 0: 8 + 4 * 13) all PR_UB planes at once would make `prod` the largest array of a superpixel's
            // record and cost a resident CTA: two at a time then; from iteration 1 on all four fit.
            // G: compute_smoothness products, terms spread over the octet (pooling these over the warp like the
            //    consistency items costs more in index arithmetic than it saves: 15.2 -> 15.8 ms)
            // H: ordered sums, one lane per plane: views in subset order, then the smoothness terms
            for (int h0 = 0; __any_sync(FULL, h0 < nb); h0 += a.hp) {
                const int h1 = min(h0 + a.hp, nb);
                for (int u = h0; u < h1; ++u) {
                    const float4 pl = cp[uniq[b0 + u]];
                    for (int t = l8; t < T; t += 8) {
                        const float4 e = tab[t];
                        const float d_intrp = (pl.y * e.x + pl.z * e.y + pl.w * pl.x) / pl.w;   // clcode.cl:1426-1428
                        const float diff = d_intrp - e.z;
                        prod[(u - h0) * pstride + t] = e.w * expf_pinned<NP>(-diff * diff * a.alpha);
                    }
                }
                __syncwarp();
                if (l8 >= h0 && l8 < h1) {
                    float consistency = 0.0f;
                    int view_counter = 0;
                    for (int n = 0; n < ns; ++n) {   // clcode.cl:1614-1622
                        const float2 rn = sres[l8 * a.nsmax + n];
                        const int fn = __float_as_int(rn.y);
                        if (fn & 1) {
                            view_counter++;
                            if (fn & 2) consistency += rn.x;
                            if (fn & 4) consistency = (float)((double)consistency + half_fly);
                        }
                    }
                    const float margin = (float)0.01;
                    const float cs1 = view_counter > 0 ? fmaxf(margin, consistency / (float)view_counter) : margin;
                    float smoothness = 0.0f;
                    const float *pr = prod + (l8 - h0) * pstride;
                    for (int t = 0; t < T; ++t) smoothness += pr[t];
                    const float sm1 = weight_norm > 0.0f ? smoothness / weight_norm : (float)0.000001;
                    cc[uniq[b0 + l8]] = make_float2(sm1, cs1);
                }
                __syncwarp();
            }
        }
    };

    // ---- duplicates among slots [c_lo, c_hi): own[c] = first slot with the same plane bits among the
    // unique planes found so far and the earlier slots of the range; new unique planes are appended ----
    int UQ = 0;
    auto dedup = [&](int c_lo, int c_hi) {
        for (int c0 = c_lo; __any_sync(FULL, c0 < c_hi); c0 += 8) {
            const int c = c0 + l8;
            const bool valid = c < c_hi;
            int o = c;
            if (valid) {
                const float4 P = cp[c];
                bool hit = false;
                for (int k = 0; k < UQ && !hit; ++k) {
                    const int e = uniq[k];
                    if (same_bits(P, cp[e])) { o = e; hit = true; }
                }
                for (int e = c0; e < c && !hit; ++e)
                    if (same_bits(P, cp[e])) { o = e; hit = true; }
                own[c] = (unsigned char)o;
            }
            const bool isu = valid && o == c;
            const unsigned m8 = (__ballot_sync(FULL, isu) >> obase) & 255u;
            if (isu) uniq[UQ + __popc(m8 & lower8)] = (unsigned char)c;
            UQ += __popc(m8);
            __syncwarp();
        }
    };

    // ---- current plane, carried greedily through the candidates (clcode.cl:1781-1784) ----
    float pd, pnx, pny, pnz, psm, pcs;
    {
        const float4 pl = vp[sp];
        const float2 co = a.cost_in[zM + sp];
        pd = pl.x; pnx = pl.y; pny = pl.z; pnz = pl.w; psm = co.x; pcs = co.y;
    }

    // ---- two rounds of dedup / evaluate / replay.  Round 0: the update() candidates generated above,
    // accepted by clcode.cl:1663-1669.  Round 1: the spatialRefinement() candidates, the plane through
    // the own point and two consecutive neighbour centres of the ring W,NW,N,NE,E,SE,S,SW
    // (clcode.cl:1852-1877, 1687-1723).  That plane keeps d, so all eight follow from the d round 0
    // left; lane r builds pair r.  Its acceptance test (clcode.cl:1713-1718) is the update() test with
    // a weight of exactly 1 (x * 1 == x, products commute), so both rounds share the replay code. ----
    int c_lo = 0, c_hi = NU;
#pragma unroll 1
    for (int round = 0; round < 2; ++round) {
        if (round == 1) {
            const int r = l8, r2 = (r + 1) & 7;
            const int ox1 = (r == 0 || r == 1 || r == 7) ? -1 : ((r >= 3 && r <= 5) ? 1 : 0);
            const int oy1 = (r >= 1 && r <= 3) ? -1 : ((r >= 5) ? 1 : 0);
            const int ox2 = (r2 == 0 || r2 == 1 || r2 == 7) ? -1 : ((r2 >= 3 && r2 <= 5) ? 1 : 0);
            const int oy2 = (r2 >= 1 && r2 <= 3) ? -1 : ((r2 >= 5) ? 1 : 0);
            int q1, q2;
            const bool ok1 = cell(a, x + ox1, y + oy1, q1), ok2 = cell(a, x + ox2, y + oy2, q2);
            const bool ok = live && ok1 && ok2;
            const unsigned pm8 = (__ballot_sync(FULL, ok) >> obase) & 255u;
            if (ok) {
                const float4 m1 = vm[2 * q1], m2 = vm[2 * q2];
                const float ax = m1.y - cx, ay = m1.z - cy, az = vp[q1].x - pd;
                const float bx = m2.y - cx, by = m2.z - cy, bz = vp[q2].x - pd;
                const float crx = ay * bz - az * by;
                const float cry = bx * az - ax * bz;
                const float crz = ax * by - ay * bx;
                const float len = sqrtf(crx * crx + cry * cry + crz * crz + 0.0f * 0.0f);
                float cnx = 0.0f, cny = 0.0f, cnz = 0.0f;
                if (len != 0.0f) { cnx = crx / len; cny = cry / len; cnz = crz / len; }
                const int pos = NU + __popc(pm8 & lower8);
                cp[pos] = make_float4(pd, cnx, cny, cnz);
                cw[pos] = 1.0f;
            }
            c_lo = NU;
            c_hi = NU + __popc(pm8);
            __syncwarp();
        }
        const int u_lo = UQ;
        dedup(c_lo, c_hi);
        evaluate(u_lo, UQ);
        for (int c = c_lo; c < c_hi; ++c) {
            const float2 co = cc[own[c]];
            const float wsim = cw[c];
            if ((a.iter < 4 && co.x * wsim > psm) || co.y * co.x > psm * pcs) {
                const float4 pl = cp[c];
                pd = pl.x; pnx = pl.y; pny = pl.z; pnz = pl.w; psm = co.x; pcs = co.y;
            }
        }
    }

    if (live && l8 == 0) {
        a.plane_out[zM + sp] = make_float4(pd, pnx, pny, pnz);
        a.cost_out[zM + sp] = make_float2(psm, pcs);
        if (STATS) {   // candidates / unique planes evaluated (both rounds)
            atomicAdd(a.stats + 0, 1ull);
            atomicAdd(a.stats + 1, (unsigned long long)c_hi);
            atomicAdd(a.stats + 2, (unsigned long long)UQ);
        }
    }
}

int launch_propagate(const FrameDev &f, const RefineArgs &r, int src, int dst, int z0, int z1, cudaStream_t st,
                     unsigned long long *stats)
{
    if (z1 <= z0) return 0;
    PArgs a;
    a.map = f.map; a.idx = f.idx; a.rep = f.rep; a.flat = f.flat;
    a.view_subset = f.view_subset; a.subset_num = f.subset_num;
    a.pixd = f.pixd; a.spcol = f.spcol;
    a.plane_in = f.plane[src]; a.cost_in = f.cost[src]; a.plane_out = f.plane[dst]; a.cost_out = f.cost[dst];
    a.V = f.V; a.W = f.W; a.H = f.H; a.mw = f.mw; a.mh = f.mh; a.N = f.N; a.M = f.M; a.cam = f.cam;
    a.gamma = r.gamma; a.alpha = r.alpha; a.bl_ratio = r.bl_ratio; a.fuse = r.fuse;
    a.K = r.no_kernel_steps < 0 ? 0 : r.no_kernel_steps;
    a.kernel_step_size = r.kernel_step_size; a.iter = r.iter; a.z0 = z0;
    a.tmax = 8 + 4 * a.K;
    a.hp = sp_hp(a.tmax);
    a.stats = stats;
    // candidate slots a superpixel can fill: 8 neighbours, the far neighbours whose offset fits the map
    // (same float -> int conversion as the kernel: truncation, saturating), 8 pairs
    long long stride = (long long)r.kernel_step_size;
    if (r.kernel_step_size != r.kernel_step_size) stride = 0;
    if (r.kernel_step_size >= 2147483647.0f) stride = 2147483647LL;
    if (r.kernel_step_size <= -2147483648.0f) stride = -2147483648LL;
    int nc = 16;
    if (stride < (long long)(f.mw > f.mh ? f.mw : f.mh))
        for (int i = 1; i <= a.K; ++i) {
            long long off = (long long)i * stride;
            if (off != (long long)(int)off) { nc += 4; continue; }   // int overflow in the kernel: be generous
            if (off < f.mh - 1) nc += 2;
            if (off < f.mw - 1) nc += 2;
        }
    a.ncmax = nc;
    if (a.ncmax > 255) return -2;   // slot numbers are kept in bytes
    a.nsmax = std::max(1, f.max_neighbours);
    const SpLayout L = sp_layout(a.tmax, a.ncmax, a.nsmax);
    size_t smem = 16 * (size_t)a.nsmax + (size_t)PR_SPB * L.total;
    if ((double)f.V * f.N >= 2147483648.0) return -3;   // 32-bit gather offsets
    if (smem > 227 * 1024) return -1;
    // every neighbour at most one camera apart?  (host copy of the lists lives with the frame)
    const bool unit = f.unit_baseline != 0;
    dim3 g((f.mw + PR_OCT - 1) / PR_OCT, z1 - z0, (f.mh + PR_WARPS - 1) / PR_WARPS);
    const bool np = a.gamma >= 0.0f && a.alpha >= 0.0f;
    // the per-pixel record path needs f.pixd (k_pixel_plane of the SAME input state) and view slots below 255
    static const bool legacy = [] { const char *e = getenv("MVS_PROP_MODE"); return e && !strcmp(e, "legacy"); }();
    const bool pixd = !legacy && f.pixd && f.spcol;
    // visible samples: |diff| < fuse and alpha >= 0, so the exp argument lies in (-fuse^2 alpha, 0]
    const bool sv = np && a.fuse >= 0.0f && (double)a.fuse * a.fuse * a.alpha < 0.34;
#define PR_LAUNCH(U, B, P, X, S, SVV)                                                                            \
    do {                                                                                                         \
        auto kern = k_propagate<U, B, P, X, S, SVV>;                                                             \
        if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        kern<<<g, PR_WARPS * 32, smem, st>>>(a);                                                                 \
    } while (0)
#define PR_PICK(U, P)                                                                                            \
    do {                                                                                                         \
        if (!pixd) PR_LAUNCH(U, 5, P, false, false, false);                                                      \
        else if (stats) PR_LAUNCH(U, PR_MINB, P, true, true, false);                                             \
        else if (sv && P) PR_LAUNCH(U, PR_MINB, P, true, false, P);                                              \
        else PR_LAUNCH(U, PR_MINB, P, true, false, false);                                                       \
    } while (0)
    if (!unit) { if (np) PR_PICK(false, true); else PR_PICK(false, false); }
    else if (np) PR_PICK(true, true);
    else PR_PICK(true, false);
#undef PR_PICK
#undef PR_LAUNCH
    return 1;
}

} // namespace mvs
