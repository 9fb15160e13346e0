/*
 * oracle.c -- TEST INFRASTRUCTURE, not product code.
 *
 * CPU restatement of the LIVE kernels of the reference's OpenCL program
 * (/root/reference/clMVDE/clcode.cl) and of the host-side scalar preparation
 * that feeds them.  Every function cites the reference lines it follows.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.  The product (the CUDA
 * library under cl_multiview_stereo_b200/csrc) never does.
 *
 * PARITY UNPINNED by the reference itself: the reference ships no golden
 * vectors / KATs and its OpenCL built-ins (powr, exp) have driver-defined
 * bits.  The oracle is pinned by definition: strict
 * left-to-right IEEE binary32 evaluation, no contraction, the math of
 * oracle_math.h, OpenCL C conversion semantics (float->int truncates; usual
 * C promotions for unsuffixed literals, i.e. double sub-expressions where the
 * kernel text has them).
 *
 * Layouts: rgbx uchar4 {s0,s1,s2,pad}; lab float4
 * {L,a,b,pad}; spixl_map float8 {id,cx,cy,L,a,b,count,disp}; idx uint32 local
 * superpixel index; rep uchar8 {nw,w,sw,n,s,ne,e,se}; flatness float2;
 * state 6 floats {d,sm,cs,nx,ny,nz}; all view-major.
 *
 * OpenMP pragmas only parallelise independent work-items; they never change
 * any arithmetic order.
 */
#include "oracle_math.h"
#include <stdlib.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* host threads the oracle's OpenMP loops use (reported as cpu_baseline.cores) */
int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

#define LOCAL_SIZE_UPDATE 16 /* header.h:38 */

typedef struct { int x, y; } int2_t;

/* ------------------------------------------------------------------------ */
/* SLIC                                                                      */
/* ------------------------------------------------------------------------ */

/* rgb2lab, clcode.cl:21-59.  s0 is read as blue, s2 as red (Appendix B S1). */
static void rgb2lab(const uint8_t *pix, float *out)
{
    float _b = (float)pix[0] * 0.0039216f;
    float _g = (float)pix[1] * 0.0039216f;
    float _r = (float)pix[2] * 0.0039216f;

    float x = _r * 0.412453f + _g * 0.357580f + _b * 0.180423f;
    float y = _r * 0.212671f + _g * 0.715160f + _b * 0.072169f;
    float z = _r * 0.019334f + _g * 0.119193f + _b * 0.950227f;

    float epsilon = 0.008856f;
    float kappa = 903.3f;
    float Xr = 0.950456f, Yr = 1.0f, Zr = 1.088754f;

    float xr = x / Xr;
    float yr = y / Yr;
    float zr = z / Zr;

    float fx, fy, fz;
    if (xr > epsilon) fx = o_cbrtf_pos(xr); else fx = (kappa * xr + 16.0f) / 116.0f;
    if (yr > epsilon) fy = o_cbrtf_pos(yr); else fy = (kappa * yr + 16.0f) / 116.0f;
    if (zr > epsilon) fz = o_cbrtf_pos(zr); else fz = (kappa * zr + 16.0f) / 116.0f;

    out[0] = 116.0f * fy - 16.0f;
    out[1] = 500.0f * (fx - fy);
    out[2] = 200.0f * (fy - fz);
    out[3] = 0.0f; /* pad lane: never written by the reference (S10); oracle writes 0 */
}

/* kernel cvt, clcode.cl:125-151 */
void orc_cvt(const uint8_t *rgbx, float *lab, int W, int H)
{
#pragma omp parallel for schedule(static)
    for (int row = 0; row < H; ++row)
        for (int col = 0; col < W; ++col) {
            int tid = W * row + col;
            rgb2lab(rgbx + 4 * (size_t)tid, lab + 4 * (size_t)tid);
        }
}

/* kernel init_cluster_centers, clcode.cl:259-294.  s7 is left unwritten by
 * the reference; the oracle writes 0 (S10).  The '>' edge test (S4) can index
 * one past the image when W mod S == S/2; the oracle clamps the read. */
void orc_init_centers(const float *lab, float *map, int W, int H, int mw, int mh, int S)
{
    for (int row = 0; row < mh; ++row)
        for (int col = 0; col < mw; ++col) {
            int c = row * mw + col;
            int cx = col * S + S / 2;
            int cy = row * S + S / 2;
            if (cx > W) cx = (col * S + W) / 2;
            if (cy > H) cy = (row * S + H) / 2;
            int rx = cx < W ? cx : W - 1;
            int ry = cy < H ? cy : H - 1;
            float *m = map + 8 * (size_t)c;
            const float *p = lab + 4 * ((size_t)ry * W + rx);
            m[0] = (float)c;
            m[1] = (float)cx;
            m[2] = (float)cy;
            m[3] = p[0];
            m[4] = p[1];
            m[5] = p[2];
            m[6] = 0.0f;
            m[7] = 0.0f;
        }
}

/* slic_distance_function, clcode.cl:422-438 */
static float slic_distance(const float *pixel, int y, int x, const float *cl, float weight,
                           float space_normalizer, float color_normalizer)
{
    float color_distance = (pixel[0] - cl[3]) * (pixel[0] - cl[3])
                         + (pixel[1] - cl[4]) * (pixel[1] - cl[4])
                         + (pixel[2] - cl[5]) * (pixel[2] - cl[5]);
    float space_distance = ((float)x - cl[1]) * ((float)x - cl[1])
                         + ((float)y - cl[2]) * ((float)y - cl[2]);
    float distance = (color_distance * color_normalizer) + weight * (space_distance * space_normalizer);
    return sqrtf(distance);
}

/* kernel find_center_association, clcode.cl:447-520 (live 2x2 rule; note the
 * outer loop variable built from deltaX is added to cluster_y, S7). */
void orc_assoc(const float *lab, const float *map, uint32_t *idx, int W, int H, int mw, int mh,
               int S, float max_xy_dist, float max_color_dist, float weight)
{
#pragma omp parallel for schedule(static)
    for (int row = 0; row < H; ++row)
        for (int col = 0; col < W; ++col) {
            int pixel_id = row * W + col;
            int cluster_x = col / S;
            int cluster_y = row / S;
            int deltaX = (col + S / 2) / S - cluster_x;
            int deltaY = (row + S / 2) / S - cluster_y;
            float min_dist = 999999.9999f;
            float min_id = -1;
            for (int i = -1 + deltaX; i <= deltaX; i++)
                for (int j = -1 + deltaY; j <= deltaY; j++) {
                    int cix = cluster_x + j;
                    int ciy = cluster_y + i;
                    if (cix >= 0 && ciy >= 0 && cix < mw && ciy < mh) {
                        int ci = ciy * mw + cix;
                        float dist = slic_distance(lab + 4 * (size_t)pixel_id, row, col,
                                                   map + 8 * (size_t)ci, weight, max_xy_dist, max_color_dist);
                        if (dist < min_dist) { min_dist = dist; min_id = (float)ci; }
                    }
                }
            /* (unsigned)(-1.0f) is undefined in C; the pinned value is 0xFFFFFFFF
             * (only reachable when no candidate cell exists, i.e. never for W,H >= S). */
            idx[pixel_id] = min_id < 0 ? 0xFFFFFFFFu : (uint32_t)min_id;
        }
}

/* kernels update_cluster_center + finalize_reduction_result,
 * clcode.cl:533-611, 719-773, host clSLIC.cpp:307-366 (G, cluster_per_line).
 * Reproduces the 256-slot local-memory tree (S6) and the window
 * under-coverage at S=8 (S5). */
void orc_update_centers(const float *lab, const uint32_t *idx, float *map, int W, int H, int mw, int mh, int S)
{
    float num_pixl_to_search = (float)(S * S * 9);
    int G = (int)ceilf(num_pixl_to_search / (float)(LOCAL_SIZE_UPDATE * LOCAL_SIZE_UPDATE)); /* clSLIC.cpp:22-23 */
    int cpl = S * 3 / LOCAL_SIZE_UPDATE; /* clSLIC.cpp:313 */
    const int L = LOCAL_SIZE_UPDATE;
#pragma omp parallel for schedule(static)
    for (int c = 0; c < mw * mh; ++c) {
        int gx = c % mw, gy = c / mw;
        float col3[256][3], loc[256][2], cnt[256];
        float xy0 = 0.0f, xy1 = 0.0f, c0 = 0.0f, c1 = 0.0f, c2 = 0.0f, n = 0.0f;
        for (int g = 0; g < G; ++g) {
            /* cpl == 0 (S < 6) would divide by zero in the reference; the oracle
             * requires S >= 6. */
            int nbr_x = g % cpl, nbr_y = g / cpl;
            for (int ly = 0; ly < L; ++ly)
                for (int lx = 0; lx < L; ++lx) {
                    int li = ly * L + lx;
                    col3[li][0] = col3[li][1] = col3[li][2] = 0.0f;
                    loc[li][0] = loc[li][1] = 0.0f;
                    cnt[li] = 0.0f;
                    int pxo = nbr_x * L + lx, pyo = nbr_y * L + ly;
                    if (pyo < S * 3 && pxo < S * 3) {
                        int px = gx * S - S + pxo, py = gy * S - S + pyo;
                        if (py >= 0 && px >= 0 && px < W && py < H) {
                            size_t pi = (size_t)py * W + px;
                            if (idx[pi] == (uint32_t)c) {
                                col3[li][0] = lab[4 * pi + 0];
                                col3[li][1] = lab[4 * pi + 1];
                                col3[li][2] = lab[4 * pi + 2];
                                loc[li][0] = (float)px;
                                loc[li][1] = (float)py;
                                cnt[li] = 1.0f;
                            }
                        }
                    }
                }
            for (int i = L * L / 2; i != 0; i /= 2)
                for (int li = 0; li < i; ++li) {
                    col3[li][0] += col3[li + i][0];
                    col3[li][1] += col3[li + i][1];
                    col3[li][2] += col3[li + i][2];
                    loc[li][0] += loc[li + i][0];
                    loc[li][1] += loc[li + i][1];
                    cnt[li] += cnt[li + i];
                }
            /* finalize: partials added sequentially in g order (clcode.cl:741-754) */
            xy0 += loc[0][0]; xy1 += loc[0][1];
            c0 = c0 + col3[0][0]; c1 = c1 + col3[0][1]; c2 = c2 + col3[0][2];
            n += cnt[0];
        }
        float *m = map + 8 * (size_t)c;
        m[0] = (float)c;
        m[1] = 0.0f; m[2] = 0.0f; m[3] = 0.0f; m[4] = 0.0f; m[5] = 0.0f; m[6] = 0.0f;
        if (n != 0) {
            m[1] = xy0 / n; m[2] = xy1 / n;
            m[3] = c0 / n; m[4] = c1 / n; m[5] = c2 / n;
            m[6] = n;
        }
        /* m[7] untouched (SLIC never writes s7) */
    }
}

/* kernel supress_local_lable, clcode.cl:676-711 (one pass) */
void orc_suppress(const uint32_t *in, uint32_t *out, int W, int H)
{
#pragma omp parallel for schedule(static)
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            int id = W * y + x;
            int clable = (int)in[id];
            if (x <= 1 || y <= 1 || x >= W - 2 || y >= H - 2) { out[id] = (uint32_t)clable; continue; }
            int diff_count = 0, diff_lable = -1;
            for (int j = -2; j <= 2; j++)
                for (int i = -2; i <= 2; i++) {
                    int nlable = (int)in[(y + j) * W + (x + i)];
                    if (nlable != clable) { diff_lable = nlable; diff_count++; }
                }
            out[id] = (uint32_t)(diff_count >= 16 ? diff_lable : clable);
        }
}

/* clSLIC::clSLIC scalar preparation (clSLIC.cpp:15-18) */
void orc_slic_params(int S, float *max_xy_dist, float *max_color_dist)
{
    float a = 1.0f / (1.4242f * S);
    float b = 15.0f / (1.7321f * 128);
    a *= a; b *= b;
    *max_xy_dist = a; *max_color_dist = b;
}

static inline int imax(int a, int b) { return a > b ? a : b; }
static inline int imin(int a, int b) { return a < b ? a : b; }

/* Edge-aware centre placement (edge_enable, clSLIC.cpp:186-233).
 *
 * The reference's first kernel, edge_compute_alternative (clcode.cl:161-195), has no defined result as written:
 * it stores the edge magnitude INTO cvt_img (the Lab buffer its neighbouring work-items are still reading the 3x3
 * window from -- a data race whose outcome depends on work-group scheduling), never writes the edge_img parameter,
 * and apply_edge_alternative (clcode.cl:204-248) then reads that never-written buffer.  Pinned semantics, stated
 * once in DESIGN.md section 3: the Sobel magnitude is computed from a SNAPSHOT of Lab (every window reads the
 * values cvt produced) and stored in edge_img, the buffer the kernel's own parameter list names for it; Lab is
 * left intact.  Everything else is the kernel's text, quirks included: the 3x3 window is stored with its centre
 * (nine entries into color[], so color[4] is the CENTRE pixel and color[5], not color[4], the right-hand
 * neighbour), and the Sobel formulas index it as if the centre had been skipped (clcode.cl:172-192). */
void orc_edge_compute(const float *lab, float *edge, int W, int H)
{
#pragma omp parallel for schedule(static)
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            float c[9][3];
            int i = 0;
            for (int yo = -1; yo <= 1; yo++)
                for (int xo = -1; xo <= 1; xo++) {
                    int cx = imax(0, imin(W - 1, x + xo)), cy = imax(0, imin(H - 1, y + yo));   /* clcode.cl:178 */
                    const float *p = lab + 4 * ((size_t)W * cy + cx);
                    c[i][0] = p[0]; c[i][1] = p[1]; c[i][2] = p[2];
                    i++;
                }
            float sum = 0.0f;   /* dot(v, (1,1,1)) = v.x*1 + v.y*1 + v.z*1, left to right */
            for (int k = 0; k < 3; ++k) {
                float dx = -1 * c[0][k] + c[2][k] - 2 * c[3][k] + 2 * c[4][k] - c[5][k] + c[7][k];   /* clcode.cl:190 */
                float dy = -1 * c[0][k] - 2 * c[1][k] - c[2][k] + c[5][k] + 2 * c[6][k] + c[7][k];   /* clcode.cl:191 */
                float v = dx * dx + dy * dy;
                sum = k == 0 ? v * 1.0f : sum + v * 1.0f;
            }
            edge[(size_t)W * y + x] = sqrtf(sum);
        }
}

/* apply_edge_alternative, clcode.cl:204-248: every centre moves to the 8-neighbour of its truncated position with
 * the smallest edge value (strict '<', scan order W, NW, N, NE, E, SE, S, SW) and takes that pixel's colour */
void orc_apply_edge(const float *lab, const float *edge, float *map, int W, int H, int mw, int mh)
{
    static const int dxy[8][2] = {{-1, 0}, {-1, -1}, {0, -1}, {1, -1}, {1, 0}, {1, 1}, {0, 1}, {-1, 1}};
    for (int y = 0; y < mh; ++y)
        for (int x = 0; x < mw; ++x) {
            float *sp = map + 8 * ((size_t)mw * y + x);
            int cx = o_f2i(sp[1]), cy = o_f2i(sp[2]);
            float edge_val = edge[(size_t)W * cy + cx];
            int changed = 0, c2x = 0, c2y = 0;
            for (int i = 0; i < 8; ++i) {
                int nx = cx + dxy[i][0], ny = cy + dxy[i][1];
                if (nx >= 0 && ny >= 0 && nx < W && ny < H) {
                    float ne = edge[(size_t)W * ny + nx];
                    if (ne < edge_val) { edge_val = ne; c2x = nx; c2y = ny; changed = 1; }
                }
            }
            if (changed) {
                const float *col = lab + 4 * ((size_t)W * c2y + c2x);
                sp[1] = (float)c2x; sp[2] = (float)c2y;
                sp[3] = col[0]; sp[4] = col[1]; sp[5] = col[2];
            }
        }
}

/* clSLIC::do_super_pixel_seg, clSLIC.cpp:67-122; edge_enable: apply_edge_values between the centre initialisation
 * and the first association (clSLIC.cpp:86) */
void orc_slic_edge(const uint8_t *rgbx, float *lab, float *map, uint32_t *idx, int W, int H, int mw, int mh,
                   int S, float weight, int no_iter, int enforce_connectivity, int edge_enable)
{
    float mxy, mcol;
    orc_slic_params(S, &mxy, &mcol);
    orc_cvt(rgbx, lab, W, H);
    orc_init_centers(lab, map, W, H, mw, mh, S);
    if (edge_enable) {
        float *edge = (float *)malloc(sizeof(float) * (size_t)W * H);
        orc_edge_compute(lab, edge, W, H);
        orc_apply_edge(lab, edge, map, W, H, mw, mh);
        free(edge);
    }
    orc_assoc(lab, map, idx, W, H, mw, mh, S, mxy, mcol, weight);
    for (int i = 0; i < no_iter; ++i) {
        orc_update_centers(lab, idx, map, W, H, mw, mh, S);
        orc_assoc(lab, map, idx, W, H, mw, mh, S, mxy, mcol, weight);
    }
    if (enforce_connectivity) { /* clSLIC.cpp:373-411: idx->tmp, tmp->idx */
        uint32_t *tmp = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)W * H);
        orc_suppress(idx, tmp, W, H);
        orc_suppress(tmp, idx, W, H);
        free(tmp);
    }
}

void orc_slic(const uint8_t *rgbx, float *lab, float *map, uint32_t *idx, int W, int H, int mw, int mh,
              int S, float weight, int no_iter, int enforce_connectivity)
{
    orc_slic_edge(rgbx, lab, map, idx, W, H, mw, mh, S, weight, no_iter, enforce_connectivity, 0);
}

/* ------------------------------------------------------------------------ */
/* Block matching (photo consistency)                                        */
/* ------------------------------------------------------------------------ */

/* kernel find_super_pixel_boundary, clcode.cl:791-855.  The reference reads
 * idx_img before testing bounds in each && chain (B1); the oracle tests the
 * bounds first -- same truth value, no out-of-bounds read. */
void orc_boundary(const float *map, const uint32_t *idx, uint8_t *rep, int V, int W, int H, int mw, int mh, int S)
{
#pragma omp parallel for schedule(static) collapse(2)
    for (int z = 0; z < V; ++z)
        for (int ty = 0; ty < mh; ++ty)
            for (int tx = 0; tx < mw; ++tx) {
                size_t mi = (size_t)z * mw * mh + (size_t)ty * mw + tx;
                int cx = o_f2i(map[8 * mi + 1]);
                int cy = o_f2i(map[8 * mi + 2]);
                if (cx < S) cx += (S - cx);
                if (cx + S > W) cx -= S;
                if (cy < S) cy += (S - cy);
                if (cy + S > H) cy -= S;
                uint32_t sp = (uint32_t)(ty * mw + tx);
                const uint32_t *im = idx + (size_t)z * W * H;
                uint8_t dir[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#define LBL(xx, yy) im[(size_t)(yy) * W + (xx)]
#define INB(xx, yy) ((xx) >= 0 && (yy) >= 0 && (xx) < W && (yy) < H)
                for (int i = 1; i < S; i++) {
                    if (cx - i >= 0 && cy - i >= 0 && INB(cx - i, cy - i) && sp == LBL(cx - i, cy - i)) dir[0] = (uint8_t)(i - 1);
                    if (cx - i >= 0 && INB(cx - i, cy) && sp == LBL(cx - i, cy)) dir[1] = (uint8_t)(i - 1);
                    if (cx - i >= 0 && cy + i < H && INB(cx - i, cy + i) && sp == LBL(cx - i, cy + i)) dir[2] = (uint8_t)(i - 1);
                    if (cy - i >= 0 && INB(cx, cy - i) && sp == LBL(cx, cy - i)) dir[3] = (uint8_t)(i - 1);
                    if (cy + i < H && INB(cx, cy + i) && sp == LBL(cx, cy + i)) dir[4] = (uint8_t)(i - 1);
                    if (cx + i < W && cy - i >= 0 && INB(cx + i, cy - i) && sp == LBL(cx + i, cy - i)) dir[5] = (uint8_t)(i - 1);
                    if (cx + i < W && INB(cx + i, cy) && sp == LBL(cx + i, cy)) dir[6] = (uint8_t)(i - 1);
                    if (cx + i < W && cy + i < H && INB(cx + i, cy + i) && sp == LBL(cx + i, cy + i)) dir[7] = (uint8_t)(i - 1);
                }
#undef LBL
#undef INB
                for (int k = 0; k < 8; ++k) rep[8 * mi + k] = dir[k];
            }
}


/* kernel initial_depth_estimation_v2, clcode.cl:972-1069, for views z0..z1-1
 * (the reference launches it once per view, photo_consistency.cpp:133-140).
 * Writes only spixl_map[].s7. */
void orc_block_match_rows(const float *lab, float *map, const uint8_t *rep, const float *disp_level, int disp_num,
                          const int *view_subset, const int *subset_num, int array_width, int V, int W, int H,
                          int mw, int mh, float bl_ratio, int z0, int z1, int y0, int y1)
{
#pragma omp parallel for schedule(dynamic, 64) collapse(2)
    for (int z = z0; z < z1; ++z)
        for (int sp = y0 * mw; sp < y1 * mw; ++sp) {
            size_t idx = (size_t)mw * mh * z + sp;
            const uint8_t *dir = rep + 8 * idx;
            int bb_l = imax(dir[0], imax(dir[1], dir[2]));
            int bb_r = imax(dir[5], imax(dir[6], dir[7]));
            int bb_t = imax(dir[0], imax(dir[3], dir[5]));
            int bb_b = imax(dir[2], imax(dir[4], dir[7]));
            /* double expression narrowed to float (clcode.cl:1006-1007); exact */
            float stepx = (float)fmax(1.0, 0.25 * (float)(bb_l + bb_r));
            float stepy = (float)fmax(1.0, 0.25 * (float)(bb_t + bb_b));
            float cxf = map[8 * idx + 1], cyf = map[8 * idx + 2];
            int refx = z % array_width, refy = z / array_width;
            float cost_est = 1000000.0f, disp_est = 0.0f;
            for (int dl = 0; dl < disp_num; dl++) {
                float d = disp_level[dl];
                float min_val = 1000000.0f;
                for (int n = 0; n < subset_num[z]; n++) {
                    int view = view_subset[V * z + n];
                    int vx = view % array_width, vy = view / array_width;
                    float val = 0.0f;
                    for (int i = -2; i <= 2; i++)
                        for (int j = -2; j <= 2; j++) {
                            int xr = o_f2i(cxf + (float)i * stepx);
                            int yr = o_f2i(cyf + (float)j * stepy);
                            int xp = o_f2i((float)xr - d * (float)(vx - refx));
                            int yp = o_f2i((float)yr - bl_ratio * d * (float)(vy - refy));
                            val += 30;
                            if (xr >= 0 && yr >= 0 && xp >= 0 && yp >= 0 && xr < W && yr < H && xp < W && yp < H) {
                                val -= 30;
                                const float *cr = lab + 4 * ((size_t)W * H * z + (size_t)W * yr + xr);
                                const float *cp = lab + 4 * ((size_t)W * H * view + (size_t)W * yp + xp);
                                val += fabsf(cr[0] - cp[0]) + fabsf(cr[1] - cp[1]) + fabsf(cr[2] - cp[2]);
                            }
                        }
                    if (val < min_val) min_val = val;
                }
                if (min_val < cost_est) { cost_est = min_val; disp_est = d; }
            }
            map[8 * idx + 7] = disp_est;
        }
}

void orc_block_match(const float *lab, float *map, const uint8_t *rep, const float *disp_level, int disp_num,
                     const int *view_subset, const int *subset_num, int array_width, int V, int W, int H,
                     int mw, int mh, float bl_ratio, int z0, int z1)
{
    orc_block_match_rows(lab, map, rep, disp_level, disp_num, view_subset, subset_num, array_width, V, W, H, mw, mh,
                         bl_ratio, z0, z1, 0, mh);
}

/* pipeline::perform_depth_est, pipeline.cpp:121-142: disparity levels and the
 * neighbour-view lists (x outer, y inner, self excluded), flattened with row
 * stride V as photo_consistency.cpp:39-47 does.  Returns the level count. */
int orc_make_levels(int min_disp, int max_disp, int inc, float *levels)
{
    int n = 0;
    for (int i = 0; i <= (max_disp - min_disp) / inc; i++) levels[n++] = (float)(min_disp + i * inc);
    return n;
}
void orc_make_subsets(int aw, int ah, int neib_hor, int neib_ver, int *subset_mat, int *subset_num)
{
    int V = aw * ah;
    for (int i = 0; i < V; i++) {
        int n = 0;
        for (int k = 0; k < V; ++k) subset_mat[i * V + k] = 0;
        for (int x = i % aw - neib_hor; x <= i % aw + neib_hor; x++)
            for (int y = i / aw - neib_ver; y <= i / aw + neib_ver; y++) {
                int indx = y * aw + x;
                if (x >= 0 && x < aw && y >= 0 && y < ah && indx != i) subset_mat[i * V + n++] = indx;
            }
        subset_num[i] = n;
    }
}

/* ------------------------------------------------------------------------ */
/* Depth refinement                                                          */
/* ------------------------------------------------------------------------ */

/* kernel compute_flatness, clcode.cl:1076-1132.  gamma here is the value the
 * host passes, (float)(1.0/(float)gamma_) (depth_refinement.cpp:130).
 * Neighbour order W, E, S(y+1), N(y-1) (R2).  Line 1131 is a double
 * expression (unsuffixed -0.25 and 1), narrowed to float on the store (R1). */
void orc_flatness(const float *map, float *flat, int V, int mw, int mh, float gamma)
{
#pragma omp parallel for schedule(static) collapse(2)
    for (int z = 0; z < V; ++z)
        for (int y = 0; y < mh; ++y)
            for (int x = 0; x < mw; ++x) {
                size_t idx = (size_t)mw * mh * z + (size_t)mw * y + x;
                const float *m0 = map + 8 * idx;
                float c0x = m0[3], c0y = m0[4], c0z = m0[5];
                float fl = 1.0f;
                const long off[4] = {-1, 1, mw, -(long)mw};
                const int ok[4] = {x - 1 >= 0, x + 1 < mw, y + 1 < mh, y - 1 >= 0};
                for (int k = 0; k < 4; ++k)
                    if (ok[k]) {
                        const float *m1 = map + 8 * (idx + off[k]);
                        float diff = (m1[3] - c0x) * (m1[3] - c0x) + (m1[4] - c0y) * (m1[4] - c0y)
                                   + (m1[5] - c0z) * (m1[5] - c0z);
                        fl += diff;
                    }
                flat[2 * idx + 0] = o_expf(-fl * gamma);
                flat[2 * idx + 1] = (float)(1.0 - o_exp(-0.25 * (double)fl * (double)gamma));
            }
}

typedef struct {
    const float *map;      /* float8[V*M] */
    const uint32_t *idx;   /* uint[V*N] */
    const uint8_t *rep;    /* uchar8[V*M] */
    const float *flat;     /* float2[V*M] */
    const int *view_subset, *subset_num;
    int V, W, H, mw, mh, array_width;
    float gamma, alpha, bl_ratio, fuse;
} rf_ctx;

/* the far-neighbour term shared by init_smoothness (clcode.cl:1176-1245) */
static inline void sm_term_init(const float *sp_check, float cx, float cy, float cz, float disp,
                                float g, float alpha, float *smoothness, float *weight_norm)
{
    float diff = o_dist3(cx, cy, cz, sp_check[3], sp_check[4], sp_check[5]);
    float similarity = o_expf(-diff * diff * g);
    diff = disp - sp_check[7];
    *smoothness += similarity * o_expf(-diff * diff * alpha);
    *weight_norm += similarity;
}

/* init_smoothness, clcode.cl:1136-1254 */
static float init_smoothness(const rf_ctx *c, const float *sp_ref, const float *fl, int px, int py, int pz,
                             int no_kernel_steps, float kernel_step_size)
{
    float smoothness = 0.0f, weight_norm = 0.0f;
    const float *vm = c->map + 8 * (size_t)c->mw * c->mh * pz;
    float cx = sp_ref[3], cy = sp_ref[4], cz = sp_ref[5], disp = sp_ref[7];
    for (int i = -1; i <= 1; i++)
        for (int j = -1; j <= 1; j++) {
            int qx = px + i, qy = py + j;
            if (qx >= 0 && qy >= 0 && qx < c->mw && qy < c->mh && (i != 0 || j != 0)) {
                const float *sp_check = vm + 8 * ((size_t)c->mw * qy + qx);
                /* distance(color_check, color) */
                float diff = o_dist3(sp_check[3], sp_check[4], sp_check[5], cx, cy, cz);
                float similarity = o_expf(-diff * diff * c->gamma);
                diff = disp - sp_check[7];
                smoothness += similarity * o_expf(-diff * diff * c->alpha);
                weight_norm += similarity;
            }
        }
    int step_size = imax(1, o_d2i((double)(fl[0] * kernel_step_size) + 0.5));
    for (int i = 1; i <= no_kernel_steps; i++) {
        float gamma_i = c->gamma * (float)(1 + i);
        int step = i * step_size;
        if (px > step)
            sm_term_init(vm + 8 * ((size_t)c->mw * py + px - (step + 1)), cx, cy, cz, disp, gamma_i, c->alpha, &smoothness, &weight_norm);
        if (px < c->mw - step - 1)
            sm_term_init(vm + 8 * ((size_t)c->mw * py + px + (step + 1)), cx, cy, cz, disp, gamma_i, c->alpha, &smoothness, &weight_norm);
        if (py > step)
            sm_term_init(vm + 8 * ((size_t)c->mw * (py - step - 1) + px), cx, cy, cz, disp, gamma_i, c->alpha, &smoothness, &weight_norm);
        if (py < c->mh - step - 1)
            sm_term_init(vm + 8 * ((size_t)c->mw * (py + step + 1) + px), cx, cy, cz, disp, gamma_i, c->alpha, &smoothness, &weight_norm);
    }
    if (weight_norm > 0) return smoothness / weight_norm;
    return (float)0.000001;
}

static inline void rep_samples(const uint8_t *rep, int *s)
{
    s[0] = rep[0]; s[1] = rep[1]; s[2] = rep[2]; s[3] = rep[3]; s[4] = 0;
    s[5] = rep[4]; s[6] = rep[5]; s[7] = rep[6]; s[8] = rep[7];
}

/* the tail shared by initialize_consistency / compute_consistency
 * (clcode.cl:1339-1347, 1614-1622); line 1346/1621 is a double expression */
static inline void cs_view_tail(float num, float visib_weight_sum, float occl_weight_sum, float visibility,
                                float visible, float fl_y, float *consistency, int *view_counter)
{
    if (num > 0) {
        (*view_counter)++;
        if (visib_weight_sum > 0)
            *consistency += (visib_weight_sum / num) * (visibility / visib_weight_sum) * (visible / visib_weight_sum);
        if (occl_weight_sum > 0)
            *consistency = (float)((double)*consistency + 0.5 * (double)fl_y);
    }
}

/* initialize_consistency, clcode.cl:1260-1357 */
static float initialize_consistency(const rf_ctx *c, int px, int py, int pz, const float *sp, const float *fl)
{
    float consistency = 0.0f;
    int view_counter = 0;
    int camx = pz % c->array_width, camy = pz / c->array_width;
    size_t M = (size_t)c->mw * c->mh, N = (size_t)c->W * c->H;
    int sp_samples[9];
    rep_samples(c->rep + 8 * (M * pz + (size_t)c->mw * py + px), sp_samples);
    float colx = sp[3], coly = sp[4], colz = sp[5];
    float d = sp[7];
    int icx = o_f2i(sp[1]), icy = o_f2i(sp[2]);
    for (int n = 0; n < c->subset_num[pz]; n++) {
        int view = c->view_subset[pz * c->V + n];
        int vx = view % c->array_width, vy = view / c->array_width;
        float visib_weight_sum = 0.0f, occl_weight_sum = 0.0f, num = 0.0f, visibility = 0.0f, visible = 0.0f;
        for (int i = -1; i <= 1; i++)
            for (int j = -1; j <= 1; j++) {
                int s = sp_samples[(i + 1) * 3 + j + 1];
                int xr = icx + s * i, yr = icy + s * j;
                int xp = o_f2i((float)xr - o_roundf(d * (float)(vx - camx)));
                int yp = o_f2i((float)yr - o_roundf(c->bl_ratio * d * (float)(vy - camy)));
                if (xp >= 0 && yp >= 0 && xp < c->W && yp < c->H) {
                    uint32_t idx_proj = c->idx[N * view + (size_t)c->W * yp + xp];
                    uint32_t cpx = idx_proj % (uint32_t)c->mw, cpy = idx_proj / (uint32_t)c->mw;
                    const float *sp_proj = c->map + 8 * (M * view + (size_t)c->mw * cpy + cpx);
                    float diff = sp_proj[7] - d;
                    float when_visible = 0.0f;
                    if (fabsf(diff) < c->fuse) when_visible = 1.0f;
                    visible += when_visible * o_expf(-diff * diff * c->alpha);
                    visib_weight_sum += when_visible;
                    occl_weight_sum += (1 - when_visible);
                    diff = o_dist3(sp_proj[3], sp_proj[4], sp_proj[5], colx, coly, colz);
                    visibility += o_expf(-diff * diff * c->gamma);
                    num += 1;
                }
            }
        cs_view_tail(num, visib_weight_sum, occl_weight_sum, visibility, visible, fl[1], &consistency, &view_counter);
    }
    float margin = (float)0.01;
    if (view_counter > 0) return fmaxf(margin, consistency / (float)view_counter);
    return margin;
}

/* host scalar preparation shared by init_spixl_state and propagate
 * (depth_refinement.cpp:335-339, 734-738; Appendix C) */
void orc_refine_params(float gamma_, float alpha_, float fuse_, int kernel_steps, int kernel_size, int S,
                       float *gamma, float *alpha, float *fuse, float *sp_kernel_step)
{
    *gamma = 1 / gamma_;
    *alpha = 1 / alpha_;
    *sp_kernel_step = (float)imax(1, kernel_size / kernel_steps * S);
    *fuse = (float)(0.5 * fuse_);
}

/* kernel init_current_state, clcode.cl:1362-1404.  gamma/alpha/fuse are the
 * kernel-side values (orc_refine_params). */
void orc_init_state(const float *map, const uint32_t *idx, const uint8_t *rep, const float *flat, float *state,
                    float gamma, float alpha, int no_kernel_steps, float kernel_step_size, float bl_ratio, float fuse,
                    int mw, int mh, const int *view_subset, const int *subset_num, int array_width, int W, int H, int V)
{
    rf_ctx c = {map, idx, rep, flat, view_subset, subset_num, V, W, H, mw, mh, array_width, gamma, alpha, bl_ratio, fuse};
#pragma omp parallel for schedule(dynamic, 16) collapse(2)
    for (int z = 0; z < V; ++z)
        for (int y = 0; y < mh; ++y)
            for (int x = 0; x < mw; ++x) {
                size_t i = (size_t)mw * mh * z + (size_t)mw * y + x;
                const float *sp = map + 8 * i;
                const float *fl = flat + 2 * i;
                float sm = init_smoothness(&c, sp, fl, x, y, z, no_kernel_steps, kernel_step_size);
                float cs = initialize_consistency(&c, x, y, z, sp, fl);
                float *st = state + 6 * i;
                st[0] = sp[7]; st[1] = sm; st[2] = cs; st[3] = 0.0f; st[4] = 0.0f; st[5] = 1.0f;
            }
}

typedef struct { float x, y, z; } f3;

/* one term of compute_smoothness (clcode.cl:1419-1432, 1444-1518) */
static inline void sm_term(const float *sp_check, float st_d, float d, f3 n, float cx, float cy, f3 color,
                           float g, float alpha, float *smoothness, float *weight_norm)
{
    float diff = o_dist3(color.x, color.y, color.z, sp_check[3], sp_check[4], sp_check[5]);
    float similarity = o_expf(-diff * diff * g);
    float d_intrp = (n.x * (cx - sp_check[1]) + n.y * (cy - sp_check[2]) + n.z * d) / n.z;
    diff = d_intrp - st_d;
    *smoothness += similarity * o_expf(-diff * diff * alpha);
    *weight_norm += similarity;
}

/* compute_smoothness, clcode.cl:1407-1525 */
static float compute_smoothness(const rf_ctx *c, const float *state, float d, f3 n, float cx, float cy, f3 color,
                                int x, int y, int z, const float *fl, int no_kernel_steps, float kernel_step_size)
{
    float smoothness = 0.0f, weight_norm = 0.0f;
    size_t M = (size_t)c->mw * c->mh;
    const float *vm = c->map + 8 * M * z;
    const float *vs = state + 6 * M * z;
    for (int i = -1; i <= 1; i++)
        for (int j = -1; j <= 1; j++) {
            int qx = x + i, qy = y + j;
            if (qx >= 0 && qy >= 0 && qx < c->mw && qy < c->mh && (i != 0 || j != 0)) {
                size_t q = (size_t)c->mw * qy + qx;
                sm_term(vm + 8 * q, vs[6 * q], d, n, cx, cy, color, c->gamma, c->alpha, &smoothness, &weight_norm);
            }
        }
    int step_size = imax(1, o_d2i((double)(fl[0] * kernel_step_size) + 0.5));
    for (int i = 1; i <= no_kernel_steps; i++) {
        float gamma_i = c->gamma * (float)(1 + i);
        int step = i * step_size;
        size_t q;
        if (x > step) {
            q = (size_t)c->mw * y + (x - (step + 1));
            sm_term(vm + 8 * q, vs[6 * q], d, n, cx, cy, color, gamma_i, c->alpha, &smoothness, &weight_norm);
        }
        if (x < c->mw - step - 1) {
            q = (size_t)c->mw * y + (x + (step + 1));
            sm_term(vm + 8 * q, vs[6 * q], d, n, cx, cy, color, gamma_i, c->alpha, &smoothness, &weight_norm);
        }
        if (y > step) {
            q = (size_t)c->mw * (y - (step + 1)) + x;
            sm_term(vm + 8 * q, vs[6 * q], d, n, cx, cy, color, gamma_i, c->alpha, &smoothness, &weight_norm);
        }
        if (y < c->mh - step - 1) {
            q = (size_t)c->mw * (y + (step + 1)) + x;
            sm_term(vm + 8 * q, vs[6 * q], d, n, cx, cy, color, gamma_i, c->alpha, &smoothness, &weight_norm);
        }
    }
    if (weight_norm > 0) return smoothness / weight_norm;
    return (float)0.000001;
}

/* compute_consistency, clcode.cl:1528-1631.  The reference takes the subset
 * row stride from get_global_size(2) (R10); here it is V explicitly. */
static float compute_consistency(const rf_ctx *c, const float *state, const uint8_t *sp_rep, float d, f3 n,
                                 float cx, float cy, f3 color, int z, const float *fl)
{
    float consistency = 0.0f;
    int view_counter = 0;
    int camx = z % c->array_width, camy = z / c->array_width;
    size_t M = (size_t)c->mw * c->mh, N = (size_t)c->W * c->H;
    int sp_samples[9];
    rep_samples(sp_rep, sp_samples);
    int icx = o_f2i(cx), icy = o_f2i(cy);
    for (int k = 0; k < c->subset_num[z]; k++) {
        int view = c->view_subset[c->V * z + k];
        float visib_weight_sum = 0.0f, occl_weight_sum = 0.0f, num = 0.0f, visibility = 0.0f, visible = 0.0f;
        int vx = view % c->array_width, vy = view / c->array_width;
        for (int i = -1; i <= 1; i++)
            for (int j = -1; j <= 1; j++) {
                int s = sp_samples[(i + 1) * 3 + j + 1];
                int xx = icx + s * i, xy = icy + s * j;
                float d_intrp = (n.x * (cx - (float)xx) + n.y * (cy - (float)xy) + n.z * d) / n.z;
                int xp = o_f2i((float)xx - o_roundf(d_intrp * (float)(vx - camx)));
                int yp = o_f2i((float)xy - o_roundf(c->bl_ratio * d_intrp * (float)(vy - camy)));
                if (xp >= 0 && yp >= 0 && xp < c->W && yp < c->H) {
                    uint32_t idx_proj = c->idx[N * view + (size_t)c->W * yp + xp];
                    uint32_t cpx = idx_proj % (uint32_t)c->mw, cpy = idx_proj / (uint32_t)c->mw;
                    size_t q = M * view + (size_t)c->mw * cpy + cpx;
                    const float *sp_proj = c->map + 8 * q;
                    const float *st = state + 6 * q;
                    float d_proj = st[0];
                    float npx = st[3], npy = st[4], npz = st[5];
                    float d_intrp_proj = (npx * (sp_proj[1] - (float)xp) + npy * (sp_proj[2] - (float)yp) + npz * d_proj) / npz;
                    float diff = d_intrp_proj - d_intrp;
                    float when_visible = 0.0f;
                    if (fabsf(diff) < c->fuse) when_visible = 1.0f;
                    visible += when_visible * o_expf(-diff * diff * c->alpha);
                    visib_weight_sum += when_visible;
                    occl_weight_sum += (1 - when_visible);
                    diff = o_dist3(sp_proj[3], sp_proj[4], sp_proj[5], color.x, color.y, color.z);
                    visibility += o_expf(-diff * diff * c->gamma);
                    num += 1;
                }
            }
        cs_view_tail(num, visib_weight_sum, occl_weight_sum, visibility, visible, fl[1], &consistency, &view_counter);
    }
    float margin = (float)0.01;
    if (view_counter > 0) return fmaxf(margin, consistency / (float)view_counter);
    return margin;
}

typedef struct { float d, sm, cs; f3 n; } plane_t;

/* update, clcode.cl:1635-1673 */
static plane_t update_plane(const rf_ctx *c, const float *state, const uint8_t *sp_rep, int iter, int no_kernel_steps,
                            float kernel_step_size, int x, int y, int z, f3 color, float cx, float cy, plane_t p0,
                            size_t idx_check, const float *fl)
{
    const float *st = state + 6 * idx_check;
    f3 n1 = {st[3], st[4], st[5]};
    float d1 = st[0];
    const float *sp_check = c->map + 8 * idx_check;
    float d_intrp = (n1.x * (sp_check[1] - cx) + n1.y * (sp_check[2] - cy) + n1.z * d1) / n1.z;
    float sm1 = compute_smoothness(c, state, d_intrp, n1, cx, cy, color, x, y, z, fl, no_kernel_steps, kernel_step_size);
    float cs1 = compute_consistency(c, state, sp_rep, d_intrp, n1, cx, cy, color, z, fl);
    float diff = o_dist3(color.x, color.y, color.z, sp_check[3], sp_check[4], sp_check[5]);
    float similarity = o_expf(-diff * diff * c->gamma);
    plane_t r = p0;
    if ((iter < 4 && sm1 * similarity > p0.sm) || cs1 * sm1 > p0.sm * p0.cs) {
        r.d = d_intrp; r.sm = sm1; r.cs = cs1; r.n = n1;
    }
    return r;
}

/* spatialRefinement + cross_product_test + normalize, clcode.cl:1676-1723.
 * normalize is pinned as v / sqrt(x*x+y*y+z*z+w*w) with w = 0, and 0 for a
 * zero vector (R9). */
static plane_t spatial_refinement(const rf_ctx *c, const float *state, const uint8_t *sp_rep, int n1x, int n1y,
                                  int n2x, int n2y, int x, int y, int z, float cx, float cy, f3 color, plane_t p0,
                                  const float *fl, int no_kernel_steps, float kernel_step_size, int iter)
{
    size_t M = (size_t)c->mw * c->mh;
    size_t q1 = M * z + (size_t)c->mw * n1y + n1x, q2 = M * z + (size_t)c->mw * n2y + n2x;
    float c1x = c->map[8 * q1 + 1], c1y = c->map[8 * q1 + 2], c1z = state[6 * q1];
    float c2x = c->map[8 * q2 + 1], c2y = c->map[8 * q2 + 2], c2z = state[6 * q2];
    float ax = c1x - cx, ay = c1y - cy, az = c1z - p0.d;
    float bx = c2x - cx, by = c2y - cy, bz = c2z - p0.d;
    float crx = ay * bz - az * by;
    float cry = bx * az - ax * bz;
    float crz = ax * by - ay * bx;
    float len = sqrtf(crx * crx + cry * cry + crz * crz + 0.0f * 0.0f);
    f3 n1;
    if (len == 0.0f) { n1.x = 0.0f; n1.y = 0.0f; n1.z = 0.0f; }
    else { n1.x = crx / len; n1.y = cry / len; n1.z = crz / len; }
    float sm1 = compute_smoothness(c, state, p0.d, n1, cx, cy, color, x, y, z, fl, no_kernel_steps, kernel_step_size);
    float cs1 = compute_consistency(c, state, sp_rep, p0.d, n1, cx, cy, color, z, fl);
    plane_t r = p0;
    if ((iter < 4 && sm1 > p0.sm) || sm1 * cs1 > p0.sm * p0.cs) {
        r.sm = sm1; r.cs = cs1; r.n = n1;
    }
    return r;
}

/* kernel propagate, clcode.cl:1727-1900, for views z0..z1-1 (all views' state_in
 * must be present: Jacobi update).  no_kernel_steps and kernel_step_size are
 * the per-iteration kernel arguments (depth_refinement.cpp:768-769). */
void orc_propagate_rows(float *state_out, const float *state_in, const float *map, const uint32_t *idx,
                        const uint8_t *rep, const float *flat, const int *view_subset, const int *subset_num, int iter,
                        int mw, int mh, int W, int H, float alpha, float gamma, float bl_ratio, float fuse,
                        int no_kernel_steps, float kernel_step_size, int array_width, int V, int z0, int z1, int y0,
                        int y1)
{
    rf_ctx c = {map, idx, rep, flat, view_subset, subset_num, V, W, H, mw, mh, array_width, gamma, alpha, bl_ratio, fuse};
    size_t M = (size_t)mw * mh;
#pragma omp parallel for schedule(dynamic, 8) collapse(3)
    for (int z = z0; z < z1; ++z)
        for (int y = y0; y < y1; ++y)
            for (int x = 0; x < mw; ++x) {
                size_t id = M * z + (size_t)mw * y + x;
                const float *sp = map + 8 * id;
                float cx = sp[1], cy = sp[2];
                f3 color = {sp[3], sp[4], sp[5]};
                const float *st = state_in + 6 * id;
                plane_t p = {st[0], st[1], st[2], {st[3], st[4], st[5]}};
                const float *fl = flat + 2 * id;
                const uint8_t *sp_rep = rep + 8 * id;
                /* 1. immediate neighbours, x offset outer (R12) */
                for (int i = -1; i <= 1; i++)
                    for (int j = -1; j <= 1; j++) {
                        int qx = x + i, qy = y + j;
                        if (qx >= 0 && qy >= 0 && qx < mw && qy < mh && !(i == 0 && j == 0))
                            p = update_plane(&c, state_in, sp_rep, iter, no_kernel_steps, kernel_step_size, x, y, z,
                                             color, cx, cy, p, M * z + (size_t)mw * qy + qx, fl);
                    }
                /* 2. far neighbours UP, DOWN, LEFT, RIGHT (R8, R12) */
                int step_size = o_f2i(kernel_step_size);
                for (int i = 1; i <= no_kernel_steps; i++) {
                    int offset = i * step_size;
                    if (y > offset)
                        p = update_plane(&c, state_in, sp_rep, iter, no_kernel_steps, kernel_step_size, x, y, z, color,
                                         cx, cy, p, M * z + (size_t)mw * (y - (offset + 1)) + x, fl);
                    if (y < mh - offset - 1)
                        p = update_plane(&c, state_in, sp_rep, iter, no_kernel_steps, kernel_step_size, x, y, z, color,
                                         cx, cy, p, M * z + (size_t)mw * (y + offset + 1) + x, fl);
                    if (x > offset)
                        p = update_plane(&c, state_in, sp_rep, iter, no_kernel_steps, kernel_step_size, x, y, z, color,
                                         cx, cy, p, M * z + (size_t)mw * y + (x - offset - 1), fl);
                    if (x < mw - offset - 1)
                        p = update_plane(&c, state_in, sp_rep, iter, no_kernel_steps, kernel_step_size, x, y, z, color,
                                         cx, cy, p, M * z + (size_t)mw * y + (x + offset + 1), fl);
                }
                /* 3. planes through consecutive neighbour pairs, ring W,NW,N,NE,E,SE,S,SW */
                const int nbx[8] = {x - 1, x - 1, x, x + 1, x + 1, x + 1, x, x - 1};
                const int nby[8] = {y, y - 1, y - 1, y - 1, y, y + 1, y + 1, y + 1};
                for (int i = 0; i < 8; i++) {
                    int j = (i + 1) % 8;
                    if (nbx[i] > -1 && nby[i] > -1 && nbx[i] < mw && nby[i] < mh && nbx[j] > -1 && nby[j] > -1 &&
                        nbx[j] < mw && nby[j] < mh)
                        p = spatial_refinement(&c, state_in, sp_rep, nbx[i], nby[i], nbx[j], nby[j], x, y, z, cx, cy,
                                               color, p, fl, no_kernel_steps, kernel_step_size, iter);
                }
                float *so = state_out + 6 * id;
                so[0] = p.d; so[1] = p.sm; so[2] = p.cs; so[3] = p.n.x; so[4] = p.n.y; so[5] = p.n.z;
            }
}

void orc_propagate(float *state_out, const float *state_in, const float *map, const uint32_t *idx, const uint8_t *rep,
                   const float *flat, const int *view_subset, const int *subset_num, int iter, int mw, int mh, int W,
                   int H, float alpha, float gamma, float bl_ratio, float fuse, int no_kernel_steps,
                   float kernel_step_size, int array_width, int V, int z0, int z1)
{
    orc_propagate_rows(state_out, state_in, map, idx, rep, flat, view_subset, subset_num, iter, mw, mh, W, H, alpha,
                       gamma, bl_ratio, fuse, no_kernel_steps, kernel_step_size, array_width, V, z0, z1, 0, mh);
}

/* kernel spixl_to_image, clcode.cl:1906-1931 */
void orc_spixl_to_image(const float *map, const uint32_t *idx, const float *state, float *disp, int V, int W, int H,
                        int mw, int mh)
{
    size_t M = (size_t)mw * mh, N = (size_t)W * H;
#pragma omp parallel for schedule(static) collapse(2)
    for (int z = 0; z < V; ++z)
        for (int py = 0; py < H; ++py)
            for (int px = 0; px < W; ++px) {
                int id = (int)idx[N * z + (size_t)W * py + px];
                int sx = id % mw, sy = id / mw;
                size_t q = M * z + (size_t)mw * sy + sx;
                const float *sp = map + 8 * q;
                const float *st = state + 6 * q;
                float nx = st[3], ny = st[4], nz = st[5];
                disp[N * z + (size_t)W * py + px] = (nx * (sp[1] - (float)px) + ny * (sp[2] - (float)py) + nz * st[0]) / nz;
            }
}

/* clDepthRefinement::do_refinement, depth_refinement.cpp:91-118 with the
 * scalar preparation of :130, :335-339, :734-738, :768-769 and the ping-pong
 * of :744-753; fusion reads buffer A (:1352, R7).  gamma_/alpha_ are the
 * values pipeline::refine_depth_map passes (2*g^2, 2*a^2, pipeline.cpp:164-166).
 * state_a/state_b: float[6*M*V] each; flat: float2[M*V]; disp: float[N*V]. */
void orc_refine(const float *map, const uint32_t *idx, const uint8_t *rep, const int *view_subset,
                const int *subset_num, int V, int W, int H, int mw, int mh, int array_width, int S, float bl_ratio,
                float gamma_, float alpha_, float fuse_, int kernel_step, int kernel_size, int no_prop,
                float *flat, float *state_a, float *state_b, float *disp)
{
    float gflat = (float)(1.0 / (double)gamma_);
    orc_flatness(map, flat, V, mw, mh, gflat);
    float gamma, alpha, fuse, kss;
    orc_refine_params(gamma_, alpha_, fuse_, kernel_step, kernel_size, S, &gamma, &alpha, &fuse, &kss);
    orc_init_state(map, idx, rep, flat, state_a, gamma, alpha, kernel_step, kss, bl_ratio, fuse, mw, mh,
                   view_subset, subset_num, array_width, W, H, V);
    for (int i = 0; i < no_prop; ++i) {
        float *out = (i % 2 == 0) ? state_b : state_a;
        const float *in = (i % 2 == 0) ? state_a : state_b;
        orc_propagate(out, in, map, idx, rep, flat, view_subset, subset_num, i, mw, mh, W, H, alpha, gamma, bl_ratio,
                      fuse, kernel_step / (i + 1), kss / (float)(i + 1), array_width, V, 0, V);
    }
    orc_spixl_to_image(map, idx, state_a, disp, V, W, H, mw, mh);
}

/* ------------------------------------------------------------------------ */
/* Multi-view fusion: kernels project_to_reference_inv (clcode.cl:1995-2034) */
/* and remove_view_inconsistency (clcode.cl:2037-2101) in the host sequence   */
/* the reference keeps commented out (depth_refinement.cpp:1382-1451): for    */
/* every reference view r in turn, slice r of disp_proj is written and then   */
/* slice r of the output is voted from ALL slices of disp_proj.  The          */
/* reference never initialises disp_proj; the semantics pinned here are a     */
/* zero-filled buffer (0 = "no value", clcode.cl:2056,2062), so view r sees   */
/* the slices 0..r.  fuse is the kernel argument (host passes 0.5 * fuse_,    */
/* depth_refinement.cpp:1328).                                                */
/* ------------------------------------------------------------------------ */
void orc_project_to_reference_inv(const float *full, float *proj, int ref, int V, int W, int H, int array_width,
                                  float bl_ratio)
{
    size_t N = (size_t)W * H;
    int rx = ref % array_width, ry = ref / array_width;
#pragma omp parallel for schedule(static)
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            size_t tid = N * ref + (size_t)W * y + x;
            float min_disp = full[tid];
            for (int i = 0; i < V; ++i) {
                if (i == ref) continue;
                int cx = i % array_width, cy = i / array_width;
                int px = o_f2i((float)x - o_roundf(min_disp * (float)(rx - cx)));
                int py = o_f2i((float)y - o_roundf(bl_ratio * min_disp * (float)(ry - cy)));
                if (px >= 0 && py >= 0 && px < W && py < H) {
                    float cur = full[N * i + (size_t)W * py + px];
                    if (min_disp < cur) min_disp = cur;
                }
            }
            proj[tid] = min_disp;
        }
}

void orc_remove_view_inconsistency(const float *proj, const float *full, float *out, int ref, int V, int W, int H,
                                   int array_width, float bl_ratio, float fuse)
{
    size_t N = (size_t)W * H;
    int rx = ref % array_width, ry = ref / array_width;
#pragma omp parallel for schedule(static)
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            size_t pix = (size_t)W * y + x;
            float d_est = 0.0f;
            for (int i = 0; i < V; ++i) {
                float d = proj[N * i + pix];
                if (d != 0) {
                    float stability = 0.0f;
                    for (int j = 0; j < V; ++j) {
                        float dc = proj[N * j + pix];
                        if (dc != 0) {
                            float diff = dc - d;
                            if (fabsf(diff) > fuse) stability -= 1;
                            if (fabsf(diff) <= fuse) stability += 1;
                        }
                    }
                    for (int j = 0; j < V; ++j) {
                        int cx = j % array_width, cy = j / array_width;
                        int px = o_f2i((float)x - o_roundf(d * (float)(cx - rx)));
                        int py = o_f2i((float)y - o_roundf(bl_ratio * d * (float)(cy - ry)));
                        if (px >= 0 && py >= 0 && px < W && py < H) {
                            float dc = full[N * j + (size_t)W * py + px];
                            float diff = dc - d;
                            if (fabsf(diff) > fuse) stability -= 1;
                            if (fabsf(diff) < fuse) stability += 1;
                        }
                    }
                    if (stability >= 0 && (d_est == 0 || d_est < d)) d_est = d;
                }
            }
            out[N * ref + pix] = d_est;
        }
}

/* the whole sequence; proj and out are float[N*V], proj is zero-filled here */
void orc_view_fusion(const float *full, float *proj, float *out, int V, int W, int H, int array_width, float bl_ratio,
                     float fuse_)
{
    float fuse = (float)(0.5 * (double)fuse_);   /* depth_refinement.cpp:1328: float fuse = 0.5 * fuse_ */
    memset(proj, 0, sizeof(float) * (size_t)W * H * V);
    for (int r = 0; r < V; ++r) {
        orc_project_to_reference_inv(full, proj, r, V, W, H, array_width, bl_ratio);
        orc_remove_view_inconsistency(proj, full, out, r, V, W, H, array_width, bl_ratio, fuse);
    }
}

/* ------------------------------------------------------------------------ */
/* Host-side byte work of the reference (no OpenCL involved, restated from    */
/* the C++ sources): loadImageIn (file_handler.cpp:6-15), plot_full_image     */
/* (depth_refinement.cpp:1476-1497), img_translate_state (:1500-1560).        */
/* The (unsigned char) cast of an out-of-range float is undefined in C; the   */
/* value pinned is what the reference's x86 build does: cvttss2si, low byte   */
/* (NaN and |x| >= 2^31 give 0x80000000 -> 0).                                */
/* ------------------------------------------------------------------------ */
void orc_load_image_in(const uint8_t *bgr, uint8_t *rgbx, int W, int H)
{
    for (size_t p = 0; p < (size_t)W * H; ++p) {
        rgbx[4 * p + 0] = bgr[3 * p + 2];   /* s0 = red   */
        rgbx[4 * p + 1] = bgr[3 * p + 1];   /* s1 = green */
        rgbx[4 * p + 2] = bgr[3 * p + 0];   /* s2 = blue  */
        rgbx[4 * p + 3] = 0;
    }
}

static uint8_t o_to_u8(float x)
{
    float fl = floorf(x);
    int v = (fl >= -2147483648.0f && fl < 2147483648.0f) ? (int)fl : (int)0x80000000u;
    return (uint8_t)(v & 255);
}

void orc_plot_full_image(const float *disp, uint8_t *out, size_t n, int min_disp, int max_disp)
{
    for (size_t p = 0; p < n; ++p)
        out[p] = o_to_u8(((disp[p] - (float)min_disp) / (float)(max_disp - min_disp)) * 255);
}

void orc_img_translate_state(const float *state, const uint32_t *idx, uint8_t *out, int V, int W, int H, int mw,
                             int mh, int element, int min_disp, int max_disp)
{
    size_t N = (size_t)W * H, M = (size_t)mw * mh;
    for (int k = 0; k < V; ++k)
        for (size_t p = 0; p < N; ++p) {
            const float *st = state + 6 * (M * k + idx[N * k + p]);
            float v;
            if (element == 0) v = ((st[0] - (float)min_disp) / (float)(max_disp - min_disp)) * 255;
            else v = st[element] * 255;
            out[N * k + p] = o_to_u8(v);
        }
}

/* clPhotoConsistency::img_translate (photo_consistency.cpp:414-436): the block-matching disparity
 * (spixl_map .s7) of each pixel's superpixel, floor-quantised.  map: float[V][M][8] */
void orc_img_translate(const float *map, const uint32_t *idx, uint8_t *out, int V, int W, int H, int mw, int mh,
                       float min_disp, float max_disp)
{
    size_t N = (size_t)W * H, M = (size_t)mw * mh;
    for (int k = 0; k < V; ++k)
        for (size_t p = 0; p < N; ++p) {
            float d = map[8 * (M * k + idx[N * k + p]) + 7];
            out[N * k + p] = o_to_u8(((d - min_disp) / (max_disp - min_disp)) * 255);
        }
}

/* clDepthRefinement::img_translate_flatness (depth_refinement.cpp:299-321): (unsigned char)ceil(fl.s0 * 255).
 * flat: float[V][M][2] */
void orc_img_translate_flatness(const float *flat, const uint32_t *idx, uint8_t *out, int V, int W, int H, int mw, int mh)
{
    size_t N = (size_t)W * H, M = (size_t)mw * mh;
    for (int k = 0; k < V; ++k)
        for (size_t p = 0; p < N; ++p) {
            float d = flat[2 * (M * k + idx[N * k + p])];
            out[N * k + p] = o_to_u8(ceilf(d * 255));
        }
}

/* clSLIC::draw_segmentation_lines (clSLIC.cpp:447-478), one image.  in/out: uchar[H][W][4]; only bytes 0..2 of
 * interior pixels are written */
void orc_draw_segmentation_lines(const uint32_t *idx, const uint8_t *in, uint8_t *out, int W, int H)
{
    for (int i = 1; i < H - 1; i++)
        for (int j = 1; j < W - 1; j++) {
            size_t p = (size_t)i * W + j;
            uint32_t l = idx[p];
            if (l != idx[p + 1] || l != idx[p - 1] || l != idx[p - W] || l != idx[p + W]) {
                out[4 * p] = 0; out[4 * p + 1] = 0; out[4 * p + 2] = 255;
            } else {
                out[4 * p] = in[4 * p]; out[4 * p + 1] = in[4 * p + 1]; out[4 * p + 2] = in[4 * p + 2];
            }
        }
}

/* ------------------------------------------------------------------------ */
/* vector wrappers of the pinned math (tests compare them with libm and with
 * the CUDA copy in csrc/mvs_math.cuh)                                       */
/* ------------------------------------------------------------------------ */
void orc_math_expf(const float *x, float *y, int n) { for (int i = 0; i < n; ++i) y[i] = o_expf(x[i]); }
void orc_math_exp(const double *x, double *y, int n) { for (int i = 0; i < n; ++i) y[i] = o_exp(x[i]); }
void orc_math_cbrtf(const float *x, float *y, int n) { for (int i = 0; i < n; ++i) y[i] = o_cbrtf_pos(x[i]); }
void orc_math_roundf(const float *x, float *y, int n) { for (int i = 0; i < n; ++i) y[i] = o_roundf(x[i]); }
void orc_math_f2i(const float *x, int *y, int n) { for (int i = 0; i < n; ++i) y[i] = o_f2i(x[i]); }
