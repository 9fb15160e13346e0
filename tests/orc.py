"""ctypes binding of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY: imported
by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ODIR = os.path.join(ROOT, "oracle")
_lib = None

f32, u32, u8, i32 = np.float32, np.uint32, np.uint8, np.int32


def load():
    global _lib
    if _lib is None:
        so = os.path.join(ODIR, "liboracle.so")
        src = [os.path.join(ODIR, f) for f in ("oracle.c", "oracle_math.h")]
        if not os.path.exists(so) or any(os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(so) for s in src):
            subprocess.check_call(["make", "-C", ODIR, "-s"])
        _lib = C.CDLL(so)
        for name in dir(Oracle):
            pass
    return Oracle(_lib)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _cf(x):
    return C.c_float(float(x))


class Oracle:
    def __init__(self, lib):
        self.lib = lib

    # ---- SLIC
    def slic(self, rgbx, S=8, weight=0.6, no_iter=5, connectivity=False):
        rgbx = np.ascontiguousarray(rgbx, u8)
        H, W, _ = rgbx.shape
        mw, mh = -(-W // S), -(-H // S)
        lab = np.zeros((H, W, 4), f32)
        mp = np.zeros((mh, mw, 8), f32)
        idx = np.zeros((H, W), u32)
        self.lib.orc_slic(_p(rgbx), _p(lab), _p(mp), _p(idx), W, H, mw, mh, S, _cf(weight), no_iter, int(connectivity))
        return lab, mp, idx

    def slic_edge(self, rgbx, S=8, weight=0.6, no_iter=5, connectivity=False, edge=True):
        """do_super_pixel_seg with edge_enable (pinned semantics, oracle.c orc_edge_compute)"""
        rgbx = np.ascontiguousarray(rgbx, u8)
        H, W, _ = rgbx.shape
        mw, mh = -(-W // S), -(-H // S)
        lab = np.zeros((H, W, 4), f32)
        mp = np.zeros((mh, mw, 8), f32)
        idx = np.zeros((H, W), u32)
        self.lib.orc_slic_edge(_p(rgbx), _p(lab), _p(mp), _p(idx), W, H, mw, mh, S, _cf(weight), no_iter, int(connectivity),
                               int(edge))
        return lab, mp, idx

    def edge_compute(self, lab):
        H, W, _ = lab.shape
        edge = np.zeros((H, W), f32)
        self.lib.orc_edge_compute(_p(np.ascontiguousarray(lab, f32)), _p(edge), W, H)
        return edge

    def apply_edge(self, lab, edge, mp):
        H, W, _ = lab.shape
        mh, mw, _ = mp.shape
        mp = np.array(mp, f32, order="C", copy=True)
        self.lib.orc_apply_edge(_p(np.ascontiguousarray(lab, f32)), _p(np.ascontiguousarray(edge, f32)), _p(mp), W, H, mw, mh)
        return mp

    def init_centers(self, lab, S):
        H, W, _ = lab.shape
        mw, mh = -(-W // S), -(-H // S)
        mp = np.zeros((mh, mw, 8), f32)
        self.lib.orc_init_centers(_p(np.ascontiguousarray(lab, f32)), _p(mp), W, H, mw, mh, S)
        return mp

    def cvt(self, rgbx):
        rgbx = np.ascontiguousarray(rgbx, u8)
        H, W, _ = rgbx.shape
        lab = np.zeros((H, W, 4), f32)
        self.lib.orc_cvt(_p(rgbx), _p(lab), W, H)
        return lab

    def slic_params(self, S):
        a, b = C.c_float(), C.c_float()
        self.lib.orc_slic_params(S, C.byref(a), C.byref(b))
        return a.value, b.value

    def assoc(self, lab, mp, S, weight):
        H, W, _ = lab.shape
        mh, mw, _ = mp.shape
        mxy, mcol = self.slic_params(S)
        idx = np.zeros((H, W), u32)
        self.lib.orc_assoc(_p(lab), _p(mp), _p(idx), W, H, mw, mh, S, _cf(mxy), _cf(mcol), _cf(weight))
        return idx

    def update_centers(self, lab, idx, mp, S):
        H, W, _ = lab.shape
        mh, mw, _ = mp.shape
        mp = mp.copy()
        self.lib.orc_update_centers(_p(lab), _p(idx), _p(mp), W, H, mw, mh, S)
        return mp

    def suppress(self, idx):
        H, W = idx.shape
        out = np.zeros_like(idx)
        self.lib.orc_suppress(_p(np.ascontiguousarray(idx, u32)), _p(out), W, H)
        return out

    # ---- block matching
    def make_levels(self, mn, mx, inc):
        buf = np.zeros(((mx - mn) // inc + 2,), f32)
        n = self.lib.orc_make_levels(mn, mx, inc, _p(buf))
        return buf[:n].copy()

    def make_subsets(self, aw, ah, nh=1, nv=1):
        V = aw * ah
        mat = np.zeros((V, V), i32)
        num = np.zeros((V,), i32)
        self.lib.orc_make_subsets(aw, ah, nh, nv, _p(mat), _p(num))
        return mat, num

    def boundary(self, mp, idx, S):
        V, H, W = idx.shape
        mh, mw = mp.shape[1:3]
        rep = np.zeros((V, mh, mw, 8), u8)
        self.lib.orc_boundary(_p(mp), _p(idx), _p(rep), V, W, H, mw, mh, S)
        return rep

    def block_match(self, lab, mp, rep, levels, subset, subset_num, aw, bl_ratio, z0=0, z1=None):
        V, H, W, _ = lab.shape
        mh, mw = mp.shape[1:3]
        mp = mp.copy()
        z1 = V if z1 is None else z1
        self.lib.orc_block_match(_p(lab), _p(mp), _p(rep), _p(levels), len(levels), _p(subset), _p(subset_num), aw, V,
                                 W, H, mw, mh, _cf(bl_ratio), z0, z1)
        return mp

    # ---- refinement
    def refine_params(self, gamma_, alpha_, fuse_, kernel_steps, kernel_size, S):
        g, a, f, k = C.c_float(), C.c_float(), C.c_float(), C.c_float()
        self.lib.orc_refine_params(_cf(gamma_), _cf(alpha_), _cf(fuse_), kernel_steps, kernel_size, S, C.byref(g),
                                   C.byref(a), C.byref(f), C.byref(k))
        return g.value, a.value, f.value, k.value

    def flatness(self, mp, gamma_flat):
        V, mh, mw, _ = mp.shape
        flat = np.zeros((V, mh, mw, 2), f32)
        self.lib.orc_flatness(_p(mp), _p(flat), V, mw, mh, _cf(gamma_flat))
        return flat

    def refine(self, mp, idx, rep, subset, subset_num, aw, S, bl_ratio, gamma_, alpha_, fuse_, kernel_step, kernel_size,
               no_prop):
        V, H, W = idx.shape
        mh, mw = mp.shape[1:3]
        flat = np.zeros((V, mh, mw, 2), f32)
        sa = np.zeros((V, mh, mw, 6), f32)
        sb = np.zeros((V, mh, mw, 6), f32)
        disp = np.zeros((V, H, W), f32)
        self.lib.orc_refine(_p(mp), _p(idx), _p(rep), _p(subset), _p(subset_num), V, W, H, mw, mh, aw, S, _cf(bl_ratio),
                            _cf(gamma_), _cf(alpha_), _cf(fuse_), kernel_step, kernel_size, no_prop, _p(flat), _p(sa),
                            _p(sb), _p(disp))
        return flat, sa, sb, disp

    def init_state(self, mp, idx, rep, flat, subset, subset_num, aw, bl_ratio, gamma, alpha, fuse, ksteps, kss):
        V, H, W = idx.shape
        mh, mw = mp.shape[1:3]
        st = np.zeros((V, mh, mw, 6), f32)
        self.lib.orc_init_state(_p(mp), _p(idx), _p(rep), _p(flat), _p(st), _cf(gamma), _cf(alpha), ksteps, _cf(kss),
                                _cf(bl_ratio), _cf(fuse), mw, mh, _p(subset), _p(subset_num), aw, W, H, V)
        return st

    def propagate(self, st_in, mp, idx, rep, flat, subset, subset_num, it, aw, bl_ratio, gamma, alpha, fuse, ksteps,
                  kss, z0=0, z1=None):
        V, H, W = idx.shape
        mh, mw = mp.shape[1:3]
        out = np.zeros_like(st_in)
        z1 = V if z1 is None else z1
        self.lib.orc_propagate(_p(out), _p(st_in), _p(mp), _p(idx), _p(rep), _p(flat), _p(subset), _p(subset_num), it,
                               mw, mh, W, H, _cf(alpha), _cf(gamma), _cf(bl_ratio), _cf(fuse), ksteps, _cf(kss), aw, V,
                               z0, z1)
        return out

    def num_threads(self):
        return int(self.lib.orc_num_threads())

    def block_match_rows(self, lab, mp, rep, levels, subset, subset_num, aw, bl_ratio, z, y0, y1):
        """timing helper: rows [y0,y1) of view z, result written into a copy"""
        V, H, W, _ = lab.shape
        mh, mw = mp.shape[1:3]
        mp = mp.copy()
        self.lib.orc_block_match_rows(_p(lab), _p(mp), _p(rep), _p(levels), len(levels), _p(subset), _p(subset_num), aw,
                                      V, W, H, mw, mh, _cf(bl_ratio), z, z + 1, y0, y1)
        return mp

    def propagate_rows(self, st_in, out, mp, idx, rep, flat, subset, subset_num, it, aw, bl_ratio, gamma, alpha, fuse,
                       ksteps, kss, z, y0, y1):
        V, H, W = idx.shape
        mh, mw = mp.shape[1:3]
        self.lib.orc_propagate_rows(_p(out), _p(st_in), _p(mp), _p(idx), _p(rep), _p(flat), _p(subset), _p(subset_num),
                                    it, mw, mh, W, H, _cf(alpha), _cf(gamma), _cf(bl_ratio), _cf(fuse), ksteps, _cf(kss),
                                    aw, V, z, z + 1, y0, y1)
        return out

    def spixl_to_image(self, mp, idx, st):
        V, H, W = idx.shape
        mh, mw = mp.shape[1:3]
        disp = np.zeros((V, H, W), f32)
        self.lib.orc_spixl_to_image(_p(mp), _p(idx), _p(st), _p(disp), V, W, H, mw, mh)
        return disp

    def view_fusion(self, disp, aw, bl_ratio, fuse_):
        """multi-view fusion (project_to_reference_inv + remove_view_inconsistency per reference view);
        returns (disp_proj, fused)"""
        disp = np.ascontiguousarray(disp, f32)
        V, H, W = disp.shape
        proj = np.zeros_like(disp)
        out = np.zeros_like(disp)
        self.lib.orc_view_fusion(_p(disp), _p(proj), _p(out), V, W, H, aw, _cf(bl_ratio), _cf(fuse_))
        return proj, out

    def project_inv(self, disp, proj, ref, aw, bl_ratio):
        """kernel project_to_reference_inv for one reference view, in place into proj[ref]"""
        V, H, W = disp.shape
        self.lib.orc_project_to_reference_inv(_p(disp), _p(proj), int(ref), V, W, H, aw, _cf(bl_ratio))

    def remove_inconsistency(self, proj, disp, out, ref, aw, bl_ratio, fuse_kernel):
        """kernel remove_view_inconsistency for one reference view, in place into out[ref]; fuse_kernel is the
        kernel argument (0.5 * system_settings::fuse)"""
        V, H, W = disp.shape
        self.lib.orc_remove_view_inconsistency(_p(proj), _p(disp), _p(out), int(ref), V, W, H, aw, _cf(bl_ratio),
                                               _cf(fuse_kernel))

    # ---- host-side byte work (file_handler.cpp:6-15, depth_refinement.cpp:1476-1560)
    def load_image_in(self, bgr):
        bgr = np.ascontiguousarray(bgr, u8)
        H, W, _ = bgr.shape
        out = np.zeros((H, W, 4), u8)
        self.lib.orc_load_image_in(_p(bgr), _p(out), W, H)
        return out

    def plot_full_image(self, disp, min_disp, max_disp):
        disp = np.ascontiguousarray(disp, f32)
        out = np.zeros(disp.shape, u8)
        self.lib.orc_plot_full_image(_p(disp), _p(out), C.c_size_t(disp.size), int(min_disp), int(max_disp))
        return out

    def img_translate_state(self, state, idx, element, min_disp, max_disp):
        V, H, W = idx.shape
        mh, mw = state.shape[1:3]
        out = np.zeros((V, H, W), u8)
        self.lib.orc_img_translate_state(_p(np.ascontiguousarray(state, f32)), _p(idx), _p(out), V, W, H, mw, mh,
                                         int(element), int(min_disp), int(max_disp))
        return out

    def img_translate(self, mp, idx, min_disp, max_disp):
        V, H, W = idx.shape
        mh, mw = mp.shape[1:3]
        out = np.zeros((V, H, W), u8)
        self.lib.orc_img_translate(_p(np.ascontiguousarray(mp, f32)), _p(idx), _p(out), V, W, H, mw, mh,
                                   C.c_float(min_disp), C.c_float(max_disp))
        return out

    def img_translate_flatness(self, flat, idx):
        V, H, W = idx.shape
        mh, mw = flat.shape[1:3]
        out = np.zeros((V, H, W), u8)
        self.lib.orc_img_translate_flatness(_p(np.ascontiguousarray(flat, f32)), _p(idx), _p(out), V, W, H, mw, mh)
        return out

    def draw_segmentation_lines(self, idx, img, out):
        """idx [H,W] u32, img / out [H,W,4] u8; out is updated in place and returned"""
        H, W = idx.shape
        self.lib.orc_draw_segmentation_lines(_p(np.ascontiguousarray(idx, np.uint32)), _p(np.ascontiguousarray(img, u8)),
                                             _p(out), W, H)
        return out

    # ---- whole pipeline (pipeline::exe_pipeline with perform_depth_est enabled)
    def pipeline(self, rgbx, s):
        """rgbx [V,H,W,4]; s = SystemSettings-like.  Returns dict of every stage output."""
        V, H, W, _ = rgbx.shape
        labs, maps, idxs = [], [], []
        for v in range(V):
            l, m, i = self.slic(rgbx[v], s.spixl_size, s.slic_color_weight, s.no_iter, s.enforce_connectivity)
            labs.append(l); maps.append(m); idxs.append(i)
        lab, mp, idx = np.stack(labs), np.stack(maps), np.stack(idxs)
        out = dict(lab=lab, idx=idx, map_slic=mp.copy())
        levels = self.make_levels(s.min_disp, s.max_disp, s.inc)
        sub, num = self.make_subsets(s.array_width, s.array_height, s.neib_hor, s.neib_ver)
        rep = self.boundary(mp, idx, s.spixl_size)
        mp = self.block_match(lab, mp, rep, levels, sub, num, s.array_width, s.bl_ratio)
        out.update(rep=rep, map=mp)
        g = np.float32(2 * float(s.gamma) ** 2)
        a = np.float32(2 * float(s.alpha) ** 2)
        flat, sa, sb, disp = self.refine(mp, idx, rep, sub, num, s.array_width, s.spixl_size, s.bl_ratio, g, a, s.fuse,
                                         s.kernel_step, s.kernel_size // 2, s.no_prop)
        out.update(flat=flat, state_a=sa, state_b=sb, disp=disp)
        return out
