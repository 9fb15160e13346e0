// propagate.cu -- plane propagation, the dominant kernel of the depth path (sm_100a).
//
// Replaces the OpenCL kernel `propagate` (reference clMVDE/clcode.cl:1727-1900) with its
// helpers update (:1635-1673), spatialRefinement (:1687-1723), compute_smoothness
// (:1407-1525), compute_consistency (:1528-1631) and cross_product_test (:1676-1685).
//
// Mapping.  The reference runs one work-item per superpixel and evaluates its 16+ candidate
// planes one after another; each evaluation is ~70 reprojected samples (9 samples x up to 8
// neighbour views, three dependent gathers each) plus 8..60 smoothness terms.  Here EIGHT
// lanes (an "octet") share one superpixel: the (view, sample) items and the smoothness terms
// of a candidate are spread over the octet, so eight gather chains are in flight per
// superpixel, registers per thread stay low and there is ONE copy of the evaluation code
// (the one-thread-per-superpixel form inlines it at every call site: 18k SASS instructions,
// stalled on instruction fetch).  A CTA is 4 warps = a 4x4 tile of superpixels.
//
// Bit-exactness.  Every float sum of the reference is a sequential chain in a fixed order;
// the octet computes the addends in parallel, stages them in shared memory and adds them in
// the reference's order (samples k = 0..8 inside a view, views in subset order, smoothness
// terms near-then-far).  Candidate-invariant factors (colour similarity of a smoothness
// term, centre offsets, the neighbour's disparity, the weight norm) are computed once per
// superpixel instead of once per candidate -- same operations, same operands, same bits.
#include "common.cuh"
#include "mvs_math.cuh"

namespace mvs {

namespace {

constexpr int PR_WARPS = 4;                 // warps per CTA (tile rows)
constexpr int PR_OCT = 4;                   // octets per warp (tile columns)
constexpr int PR_SPB = PR_WARPS * PR_OCT;   // superpixels per CTA
constexpr int PR_ITEMS = 72;                // 9 samples x 8 neighbour views at most

// shared-memory record of one superpixel (dynamic: tab/prod hold tmax entries)
constexpr int PR_MEMO = 4;                  // evaluated planes remembered per superpixel

struct SpShared {
    float4 samp[9];          // {sx, sy (integers held as float), cx - sx, cy - sy}
    float4 viewt[8];         // {view * N (int bits), dvx, dvy, view * M (int bits)}
    float2 item[PR_ITEMS];   // per (view, sample): {visible term or flag, visibility term or -1}
    float2 di[10];           // per sample of the current candidate: {d_intrp, bl_ratio * d_intrp}
    float4 memo_plane[PR_MEMO];   // {d, nx, ny, nz} of the last evaluated candidates
    float2 memo_cost[PR_MEMO];    // their {smoothness, consistency}
};
static_assert(sizeof(SpShared) % 16 == 0, "SpShared must keep float4 alignment");

__host__ __device__ inline size_t sp_shared_bytes(int tmax)
{
    return sizeof(SpShared) + (size_t)tmax * (sizeof(float4) + sizeof(float));
}

struct PArgs {
    const float4 *map;        // [V][M][2]
    const uint32_t *idx;      // [V][N]
    const uint2 *rep;         // [V][M]
    const float2 *flat;       // [V][M]
    const int *view_subset, *subset_num;
    const float4 *plane_in;   // [V][M] {d,nx,ny,nz}
    const float2 *cost_in;    // [V][M] {sm,cs}
    float4 *plane_out;
    float2 *cost_out;
    int V, W, H, mw, mh, N, M, aw;
    float gamma, alpha, bl_ratio, fuse;
    int K;                    // no_kernel_steps of this iteration
    float kernel_step_size;
    int iter, z0, tmax;
};

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

// acceptance test shared by update() and spatialRefinement() (clcode.cl:1663-1669,
// 1713-1718): update() weighs sm1 by the colour similarity to the source neighbour,
// spatialRefinement() does not (wsim = 1, exact) and keeps d (its cd IS the current d)
__device__ __forceinline__ void accept(int iter, bool, float sm1, float cs1, float wsim, float cd, float cnx,
                                       float cny, float cnz, float &pd, float &pnx, float &pny, float &pnz,
                                       float &psm, float &pcs)
{
    if ((iter < 4 && sm1 * wsim > psm) || cs1 * sm1 > psm * pcs) {
        pd = cd; psm = sm1; pcs = cs1; pnx = cnx; pny = cny; pnz = cnz;
    }
}

} // namespace

__global__ void __launch_bounds__(PR_WARPS * 32) k_propagate(PArgs a)
{
    extern __shared__ __align__(16) unsigned char pr_smem[];
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int l8 = lane & 7, obase = lane & 24;
    const int x = blockIdx.x * PR_OCT + (lane >> 3), y = blockIdx.y * PR_WARPS + warp, z = a.z0 + blockIdx.z;
    const bool live = x < a.mw && y < a.mh;

    unsigned char *sbase = pr_smem + (size_t)(warp * PR_OCT + (lane >> 3)) * sp_shared_bytes(a.tmax);
    SpShared *sh = reinterpret_cast<SpShared *>(sbase);
    float4 *tab = reinterpret_cast<float4 *>(sbase + sizeof(SpShared));   // {cx-cxq, cy-cyq, d_q, similarity}
    float *prod = reinterpret_cast<float *>(tab + a.tmax);

    const size_t zM = (size_t)z * a.M;
    const float4 *vm = a.map + 2 * zM;
    const float4 *vp = a.plane_in + zM;
    const int sp = live ? y * a.mw + x : 0;

    // ---- own record (the octet's lanes read the same addresses: one transaction) ----
    const float4 ma = vm[2 * sp], mb = vm[2 * sp + 1];
    const float cx = ma.y, cy = ma.z, cL = ma.w, ca = mb.x, cb = mb.y;
    const float2 fl = a.flat[zM + sp];
    const int ns = a.subset_num[z];
    const int camx = z % a.aw, camy = z / a.aw;
    const double half_fly = 0.5 * (double)fl.y;   // clcode.cl:1621 is a double expression

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
            sh->samp[k] = make_float4((float)sx, (float)sy, cx - (float)sx, cy - (float)sy);
        }
        if (l8 < ns) {
            int view = a.view_subset[a.V * z + l8];
            sh->viewt[l8] = make_float4(__int_as_float(view * a.N), (float)(view % a.aw - camx),
                                        (float)(view / a.aw - camy), __int_as_float(view * a.M));
        }
    }

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
                const float sim = expf_pinned(-diff * diff * g);
                tab[T + __popc(m8 & ((1u << l8) - 1u))] = make_float4(cx - qa.y, cy - qa.z, vp[q].x, sim);
            }
            T += __popc(m8);
        }
    }
    __syncwarp();
    float weight_norm = 0.0f;
    for (int t = 0; t < T; ++t) weight_norm += tab[t].w;

    // ---- current plane, carried greedily through the candidates (clcode.cl:1781-1784) ----
    float pd, pnx, pny, pnz, psm, pcs;
    {
        const float4 pl = vp[sp];
        const float2 co = a.cost_in[zM + sp];
        pd = pl.x; pnx = pl.y; pny = pl.z; pnz = pl.w; psm = co.x; pcs = co.y;
    }

    // candidate numbering (clcode.cl:1768-1877): 0..7 near neighbours (x offset outer);
    // 8..8+4K-1 far neighbours UP, DOWN, LEFT, RIGHT at i * (int)kernel_step_size;
    // then the 8 planes through consecutive neighbour pairs, ring W,NW,N,NE,E,SE,S,SW.
    const int stride = f2i(a.kernel_step_size);
    const int far_end = 8 + 4 * a.K, ncand = far_end + 8;
    const bool far_possible = stride < max(a.mw, a.mh);
    int ci = live ? 0 : ncand;
    int nmemo = 0;   // planes evaluated so far by this octet (memo holds the last PR_MEMO)

    for (;;) {
        // ---- next candidate of this octet that still needs an evaluation ----
        // A candidate's costs are a pure function of its plane (cd, cn) for the fixed input
        // state, and neighbouring superpixels often carry bit-identical planes: planes already
        // evaluated for this superpixel take their costs from the memo and go straight to the
        // acceptance test.
        bool act = false, pair = false;
        float cd = 0.0f, cnx = 0.0f, cny = 0.0f, cnz = 0.0f, wsim = 1.0f;
        while (ci < ncand && !act) {
            const int c = ci++;
            int q1 = 0, q2 = 0;
            bool ok;
            pair = false;
            if (c < 8) {
                int i, j;
                near_offset(c, i, j);
                ok = cell(a, x + i, y + j, q1);
            } else if (c < far_end) {
                if (!far_possible) { ci = far_end; continue; }
                const int off = (1 + ((c - 8) >> 2)) * stride, dir = (c - 8) & 3;
                if (dir == 0) { ok = y > off; q1 = a.mw * (y - (off + 1)) + x; }
                else if (dir == 1) { ok = y < a.mh - off - 1; q1 = a.mw * (y + off + 1) + x; }
                else if (dir == 2) { ok = x > off; q1 = a.mw * y + x - off - 1; }
                else { ok = x < a.mw - off - 1; q1 = a.mw * y + x + off + 1; }
            } else {
                const int r = c - far_end, r2 = (r + 1) & 7;
                // ring position r -> offset: W,NW,N,NE,E,SE,S,SW
                const int ox1 = (r == 0 || r == 1 || r == 7) ? -1 : ((r >= 3 && r <= 5) ? 1 : 0);
                const int oy1 = (r >= 1 && r <= 3) ? -1 : ((r >= 5) ? 1 : 0);
                const int ox2 = (r2 == 0 || r2 == 1 || r2 == 7) ? -1 : ((r2 >= 3 && r2 <= 5) ? 1 : 0);
                const int oy2 = (r2 >= 1 && r2 <= 3) ? -1 : ((r2 >= 5) ? 1 : 0);
                const bool ok1 = cell(a, x + ox1, y + oy1, q1), ok2 = cell(a, x + ox2, y + oy2, q2);
                ok = ok1 && ok2;
                pair = true;
            }
            if (!ok) continue;
            // the candidate plane (cd, cn) and its acceptance weight
            if (!pair) {   // update(): neighbour's plane extrapolated to the own centre (clcode.cl:1643-1651)
                const float4 pl = vp[q1];
                const float4 qa = vm[2 * q1], qb = vm[2 * q1 + 1];
                cd = (pl.y * (qa.y - cx) + pl.z * (qa.z - cy) + pl.w * pl.x) / pl.w;
                cnx = pl.y; cny = pl.z; cnz = pl.w;
                const float diff = dist3(cL, ca, cb, qa.w, qb.x, qb.y);
                wsim = expf_pinned(-diff * diff * a.gamma);
            } else {       // spatialRefinement(): plane through the own point and two neighbour centres
                const float4 m1 = vm[2 * q1], m2 = vm[2 * q2];
                const float ax = m1.y - cx, ay = m1.z - cy, az = vp[q1].x - pd;
                const float bx = m2.y - cx, by = m2.z - cy, bz = vp[q2].x - pd;
                const float crx = ay * bz - az * by;
                const float cry = bx * az - ax * bz;
                const float crz = ax * by - ay * bx;
                const float len = sqrtf(crx * crx + cry * cry + crz * crz + 0.0f * 0.0f);
                cnx = 0.0f; cny = 0.0f; cnz = 0.0f;
                if (len != 0.0f) { cnx = crx / len; cny = cry / len; cnz = crz / len; }
                cd = pd;
                wsim = 1.0f;
            }
            bool hit = false;
            for (int e = 0; e < min(nmemo, PR_MEMO) && !hit; ++e) {
                const float4 mp = sh->memo_plane[e];
                if (__float_as_int(mp.x) == __float_as_int(cd) && __float_as_int(mp.y) == __float_as_int(cnx) &&
                    __float_as_int(mp.z) == __float_as_int(cny) && __float_as_int(mp.w) == __float_as_int(cnz)) {
                    const float2 mc = sh->memo_cost[e];
                    accept(a.iter, pair, mc.x, mc.y, wsim, cd, cnx, cny, cnz, pd, pnx, pny, pnz, psm, pcs);
                    hit = true;
                }
            }
            act = !hit;
        }
        if (!__any_sync(FULL, act)) break;

        // ---- per-sample plane disparity of the candidate (clcode.cl:1571-1573), once per sample ----
        if (act)
            for (int k = l8; k < 9; k += 8) {
                const float4 sv = sh->samp[k];
                const float di = (cnx * sv.z + cny * sv.w + cnz * cd) / cnz;
                sh->di[k] = make_float2(di, a.bl_ratio * di);
            }
        __syncwarp();

        if (act) {
            // ---- compute_consistency items: (neighbour view n, sample k) spread over the octet ----
            const int nitems = 9 * ns;
            for (int it = l8; it < nitems; it += 8) {
                const int n = it / 9, k = it - 9 * n;
                const float4 sv = sh->samp[k], vt = sh->viewt[n];
                const float2 dd = sh->di[k];
                const int xp = f2i(sv.x - round_haz(dd.x * vt.y));   // clcode.cl:1576-1577
                const int yp = f2i(sv.y - round_haz(dd.y * vt.z));
                float2 out = make_float2(0.0f, -1.0f);              // -1: sample projects outside
                if (xp >= 0 && yp >= 0 && xp < a.W && yp < a.H) {
                    const uint32_t q = a.idx[(uint32_t)__float_as_int(vt.x) + (uint32_t)(yp * a.W + xp)];
                    const uint32_t vq = (uint32_t)__float_as_int(vt.w) + q;
                    const float4 qa = a.map[2 * vq], qb = a.map[2 * vq + 1];
                    const float4 pl = a.plane_in[vq];
                    const float d_proj = (pl.y * (qa.y - (float)xp) + pl.z * (qa.z - (float)yp) + pl.w * pl.x) / pl.w;
                    const float diff = d_proj - dd.x;
                    // visible: exp term; occluded: -1 (adds +0 in the reference); NaN stays NaN
                    out.x = fabsf(diff) < a.fuse ? expf_pinned(-diff * diff * a.alpha) : (diff != diff ? diff : -1.0f);
                    const float cdiff = dist3(qa.w, qb.x, qb.y, cL, ca, cb);
                    out.y = expf_pinned(-cdiff * cdiff * a.gamma);
                }
                sh->item[it] = out;
            }
            // ---- compute_smoothness terms ----
            for (int t = l8; t < T; t += 8) {
                const float4 e = tab[t];
                const float d_intrp = (cnx * e.x + cny * e.y + cnz * cd) / cnz;   // clcode.cl:1426-1428
                const float diff = d_intrp - e.z;
                prod[t] = e.w * expf_pinned(-diff * diff * a.alpha);
            }
        }
        __syncwarp();

        // ---- ordered sums: samples within a view (lane n), then views, then smoothness ----
        float term = 0.0f;
        int flags = 0;   // bit0: view has samples, bit1: some visible, bit2: some occluded
        if (act && l8 < ns) {
            float num = 0.0f, vws = 0.0f, ows = 0.0f, visibility = 0.0f, visible = 0.0f;
            const float2 *iv = sh->item + 9 * l8;
#pragma unroll
            for (int k = 0; k < 9; ++k) {
                const float2 v = iv[k];
                if (!(v.y == -1.0f)) {
                    if (v.x >= 0.0f) { visible += v.x; vws += 1.0f; }
                    else { if (v.x != v.x) visible += v.x; ows += 1.0f; }
                    visibility += v.y;
                    num += 1.0f;
                }
            }
            if (num > 0.0f) {
                flags = 1;
                if (vws > 0.0f) { flags |= 2; term = (vws / num) * (visibility / vws) * (visible / vws); }
                if (ows > 0.0f) flags |= 4;
            }
        }
        float consistency = 0.0f;
        int view_counter = 0;
        for (int n = 0; n < ns; ++n) {   // clcode.cl:1614-1622, views in subset order
            const float tn = __shfl_sync(FULL, term, obase + n);
            const int fn = __shfl_sync(FULL, flags, obase + n);
            if (fn & 1) {
                view_counter++;
                if (fn & 2) consistency += tn;
                if (fn & 4) consistency = (float)((double)consistency + half_fly);
            }
        }
        if (act) {
            const float margin = (float)0.01;
            const float cs1 = view_counter > 0 ? fmaxf(margin, consistency / (float)view_counter) : margin;
            float smoothness = 0.0f;
            int t = 0;
            for (; t + 4 <= T; t += 4) {
                const float4 p4 = *reinterpret_cast<const float4 *>(prod + t);
                smoothness += p4.x; smoothness += p4.y; smoothness += p4.z; smoothness += p4.w;
            }
            for (; t < T; ++t) smoothness += prod[t];
            const float sm1 = weight_norm > 0.0f ? smoothness / weight_norm : (float)0.000001;
            if (l8 == 0) {
                sh->memo_plane[nmemo % PR_MEMO] = make_float4(cd, cnx, cny, cnz);
                sh->memo_cost[nmemo % PR_MEMO] = make_float2(sm1, cs1);
            }
            nmemo++;
            accept(a.iter, pair, sm1, cs1, wsim, cd, cnx, cny, cnz, pd, pnx, pny, pnz, psm, pcs);
        }
        __syncwarp();   // staged terms consumed, memo visible to the octet
    }

    if (live && l8 == 0) {
        a.plane_out[zM + sp] = make_float4(pd, pnx, pny, pnz);
        a.cost_out[zM + sp] = make_float2(psm, pcs);
    }
}


int launch_propagate(const FrameDev &f, const RefineArgs &r, int src, int dst, int z0, int z1, cudaStream_t st)
{
    if (z1 <= z0) return 0;
    PArgs a;
    a.map = f.map; a.idx = f.idx; a.rep = f.rep; a.flat = f.flat;
    a.view_subset = f.view_subset; a.subset_num = f.subset_num;
    a.plane_in = f.plane[src]; a.cost_in = f.cost[src]; a.plane_out = f.plane[dst]; a.cost_out = f.cost[dst];
    a.V = f.V; a.W = f.W; a.H = f.H; a.mw = f.mw; a.mh = f.mh; a.N = f.N; a.M = f.M; a.aw = f.aw;
    a.gamma = r.gamma; a.alpha = r.alpha; a.bl_ratio = r.bl_ratio; a.fuse = r.fuse;
    a.K = r.no_kernel_steps < 0 ? 0 : r.no_kernel_steps;
    a.kernel_step_size = r.kernel_step_size; a.iter = r.iter; a.z0 = z0;
    a.tmax = 8 + 4 * a.K;
    size_t smem = PR_SPB * sp_shared_bytes(a.tmax);
    if (smem > 227 * 1024 || (double)f.V * f.N >= 2147483648.0) return -1;   // kernel_step too large for the staged term table
    if (smem > 48 * 1024) cudaFuncSetAttribute(k_propagate, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    dim3 g((f.mw + PR_OCT - 1) / PR_OCT, (f.mh + PR_WARPS - 1) / PR_WARPS, z1 - z0);
    k_propagate<<<g, PR_WARPS * 32, smem, st>>>(a);
    return 1;
}

} // namespace mvs
