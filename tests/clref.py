"""ctypes binding of oracle/_ref/libclref.so: the reference's own clcode.cl compiled as C++
(oracle/ref/).  TEST INFRASTRUCTURE ONLY.  Same method names and array conventions as
tests/orc.py, so that a test can run one call against both oracles.  The library exists only
where /root/reference was available at build time (and travels to the GPU box as a built file)."""
import ctypes as C
import os
import subprocess

import numpy as np

from .orc import ODIR, _cf, _p, f32, i32, u8, u32

SO = os.path.join(ODIR, "_ref", "libclref.so")
REF_SRC = "/root/reference/clMVDE/clcode.cl"
_lib = None


def available():
    return os.path.exists(SO) or os.path.exists(REF_SRC)


def load():
    """returns a ClRef, or None when neither the built library nor the reference tree exists"""
    global _lib
    if _lib is None:
        if os.path.exists(REF_SRC):
            subprocess.check_call(["make", "-C", os.path.join(ODIR, "ref"), "-s"])
        if not os.path.exists(SO):
            return None
        _lib = C.CDLL(SO)
    return ClRef(_lib)


class ClRef:
    def __init__(self, lib):
        self.lib = lib

    # ---- SLIC (clSLIC.cpp:67-122)
    def cvt(self, rgbx):
        rgbx = np.ascontiguousarray(rgbx, u8)
        H, W, _ = rgbx.shape
        lab = np.zeros((H, W, 4), f32)
        self.lib.clref_cvt(_p(rgbx), _p(lab), W, H)
        return lab

    def init_centers(self, lab, S):
        H, W, _ = lab.shape
        mw, mh = -(-W // S), -(-H // S)
        mp = np.zeros((mh, mw, 8), f32)
        self.lib.clref_init_centers(_p(lab), _p(mp), W, H, mw, mh, S)
        return mp

    def assoc(self, lab, mp, S, weight, mxy, mcol):
        H, W, _ = lab.shape
        mh, mw, _ = mp.shape
        idx = np.zeros((H, W), u32)
        self.lib.clref_assoc(_p(lab), _p(mp), _p(idx), W, H, mw, mh, S, _cf(mxy), _cf(mcol), _cf(weight))
        return idx

    def update_centers(self, lab, idx, mp, S):
        H, W, _ = lab.shape
        mh, mw, _ = mp.shape
        mp = mp.copy()
        self.lib.clref_update_centers(_p(lab), _p(idx), _p(mp), W, H, mw, mh, S)
        return mp

    def edge_compute_snapshot(self, lab):
        """the arithmetic of edge_compute_alternative, every work-item on its own copy of the Lab snapshot"""
        H, W, _ = lab.shape
        edge = np.zeros((H, W), f32)
        self.lib.clref_edge_compute_snapshot(_p(np.ascontiguousarray(lab, f32)), _p(edge), W, H)
        return edge

    def apply_edge(self, lab, edge, mp):
        H, W, _ = lab.shape
        mh, mw, _ = mp.shape
        mp = np.array(mp, f32, order="C", copy=True)
        self.lib.clref_apply_edge(_p(np.ascontiguousarray(lab, f32)), _p(np.ascontiguousarray(edge, f32)), _p(mp), W, H, mw, mh)
        return mp

    def suppress(self, idx):
        H, W = idx.shape
        out = np.zeros_like(idx)
        self.lib.clref_suppress(_p(np.ascontiguousarray(idx, u32)), _p(out), W, H)
        return out

    def slic(self, rgbx, S, weight, no_iter, connectivity, mxy, mcol):
        """clSLIC::do_super_pixel_seg: cvt, init, assoc, no_iter x (update, assoc), [2 x suppress]"""
        lab = self.cvt(rgbx)
        mp = self.init_centers(lab, S)
        idx = self.assoc(lab, mp, S, weight, mxy, mcol)
        for _ in range(no_iter):
            mp = self.update_centers(lab, idx, mp, S)
            idx = self.assoc(lab, mp, S, weight, mxy, mcol)
        if connectivity:
            idx = self.suppress(self.suppress(idx))
        return lab, mp, idx

    # ---- block matching (photo_consistency.cpp:21-150)
    def boundary(self, mp, idx, S):
        V, H, W = idx.shape
        mh, mw = mp.shape[1:3]
        rep = np.zeros((V, mh, mw, 8), u8)
        self.lib.clref_boundary(_p(mp), _p(idx), _p(rep), V, W, H, mw, mh, S)
        return rep

    def block_match(self, lab, mp, rep, idx, levels, subset, subset_num, aw, bl_ratio, S):
        V, H, W, _ = lab.shape
        mh, mw = mp.shape[1:3]
        mp = mp.copy()
        levels = np.ascontiguousarray(levels, f32)
        self.lib.clref_block_match(_p(lab), _p(mp), _p(rep), _p(idx), _p(levels), len(levels), _p(subset),
                                   _p(subset_num), aw, V, W, H, mw, mh, S, _cf(bl_ratio))
        return mp

    def block_match_rows(self, lab, mp, rep, idx, levels, subset, subset_num, aw, bl_ratio, S, z, y0, y1):
        """timing helper: superpixel rows [y0,y1) of view z, result written into a copy"""
        V, H, W, _ = lab.shape
        mh, mw = mp.shape[1:3]
        mp = mp.copy()
        levels = np.ascontiguousarray(levels, f32)
        self.lib.clref_block_match_rows(_p(lab), _p(mp), _p(rep), _p(idx), _p(levels), len(levels), _p(subset),
                                        _p(subset_num), aw, V, W, H, mw, mh, S, _cf(bl_ratio), z, y0, y1)
        return mp

    def propagate_rows(self, st_in, out, mp, idx, rep, flat, subset, subset_num, it, aw, bl_ratio, gamma, alpha, fuse,
                       ksteps, kss, z, y0, y1):
        """timing helper: superpixel rows [y0,y1) of view z into `out`"""
        V, H, W = idx.shape
        mh, mw = mp.shape[1:3]
        self.lib.clref_propagate_rows(_p(out), _p(st_in), _p(mp), _p(idx), _p(rep), _p(flat), _p(subset), _p(subset_num),
                                      it, mw, mh, W, H, _cf(alpha), _cf(gamma), _cf(bl_ratio), _cf(fuse), ksteps, _cf(kss),
                                      aw, V, z, y0, y1)
        return out

    # ---- refinement (depth_refinement.cpp:91-118)
    def flatness(self, mp, gamma_flat):
        V, mh, mw, _ = mp.shape
        flat = np.zeros((V, mh, mw, 2), f32)
        self.lib.clref_flatness(_p(mp), _p(flat), V, mw, mh, _cf(gamma_flat))
        return flat

    def init_state(self, mp, idx, rep, flat, subset, subset_num, aw, bl_ratio, gamma, alpha, fuse, ksteps, kss):
        V, H, W = idx.shape
        mh, mw = mp.shape[1:3]
        st = np.zeros((V, mh, mw, 6), f32)
        self.lib.clref_init_state(_p(mp), _p(idx), _p(rep), _p(flat), _p(st), _cf(gamma), _cf(alpha), ksteps, _cf(kss),
                                  _cf(bl_ratio), _cf(fuse), mw, mh, _p(subset), _p(subset_num), aw, W, H, V)
        return st

    def propagate(self, st_in, mp, idx, rep, flat, subset, subset_num, it, aw, bl_ratio, gamma, alpha, fuse, ksteps, kss):
        V, H, W = idx.shape
        mh, mw = mp.shape[1:3]
        out = np.zeros_like(st_in)
        self.lib.clref_propagate(_p(out), _p(st_in), _p(mp), _p(idx), _p(rep), _p(flat), _p(subset), _p(subset_num), it,
                                 mw, mh, W, H, _cf(alpha), _cf(gamma), _cf(bl_ratio), _cf(fuse), ksteps, _cf(kss), aw, V)
        return out

    def spixl_to_image(self, mp, idx, st):
        V, H, W = idx.shape
        mh, mw = mp.shape[1:3]
        disp = np.zeros((V, H, W), f32)
        self.lib.clref_spixl_to_image(_p(mp), _p(idx), _p(st), _p(disp), V, W, H, mw, mh)
        return disp

    def view_fusion(self, disp, aw, bl_ratio, fuse_):
        disp = np.ascontiguousarray(disp, f32)
        V, H, W = disp.shape
        proj = np.zeros_like(disp)
        out = np.zeros_like(disp)
        self.lib.clref_view_fusion(_p(disp), _p(proj), _p(out), V, W, H, aw, _cf(bl_ratio), _cf(fuse_))
        return proj, out
