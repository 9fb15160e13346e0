"""ctypes binding of oracle/_ref/libhosttwin.so: the reference's CPU twins (do_initial_depth_estimation_host,
compute_flatness_host, init_spixl_state_host, propagate_host) compiled UNMODIFIED from the reference's host
sources over inert OpenCL / OpenCV headers (oracle/ref_host/).  TEST / BASELINE INFRASTRUCTURE ONLY: these are
the CPU path north_star asks to see timed; they have diverged from the kernels and are
not an oracle.  The library exists only where /root/reference was available at build time."""
import ctypes as C
import os
import subprocess

import numpy as np

from .orc import ODIR, _cf, _p, f32, i32, u8, u32

SO = os.path.join(ODIR, "_ref", "libhosttwin.so")
REF_SRC = "/root/reference/clMVDE/depth_refinement.cpp"
_lib = None


def available():
    return os.path.exists(SO) or os.path.exists(REF_SRC)


def load():
    global _lib
    if _lib is None:
        if os.path.exists(REF_SRC):
            subprocess.check_call(["make", "-C", os.path.join(ODIR, "ref_host"), "-s"])
        if not os.path.exists(SO):
            return None
        _lib = C.CDLL(SO)
        _lib.hosttwin_block_match.restype = C.c_double
    return HostTwin(_lib)


class HostTwin:
    def __init__(self, lib):
        self.lib = lib

    def block_match(self, lab, mp, idx, levels, subset, subset_num, aw, bl_ratio, S, z0=0, z1=None):
        """do_initial_depth_estimation_host (photo_consistency.cpp:385-410) -> (map with s7, rep, seconds)"""
        V, H, W, _ = lab.shape
        mh, mw = mp.shape[1:3]
        mp = np.array(mp, f32, order="C", copy=True)
        rep = np.zeros((V, mh, mw, 8), u8)
        levels = np.ascontiguousarray(levels, f32)
        z1 = V if z1 is None else z1
        sec = self.lib.hosttwin_block_match(_p(mp), _p(rep), _p(np.ascontiguousarray(lab, f32)), _p(np.ascontiguousarray(idx, u32)),
                                            V, W, H, mw, mh, S, aw, _cf(bl_ratio), _p(levels), len(levels),
                                            _p(np.ascontiguousarray(subset, i32)), _p(np.ascontiguousarray(subset_num, i32)), z0, z1)
        return mp, rep, float(sec)

    def refine(self, mp, idx, rep, lab, aw, ah, S, bl_ratio, subset, subset_num, gamma_, alpha_, fuse_, kernel_step,
               kernel_size, no_prop):
        """compute_flatness_host + init_spixl_state_host + no_prop x propagate_host -> (flat, state_a, state_b,
        {stage: seconds}); gamma_/alpha_/kernel_size as pipeline::refine_depth_map passes them"""
        V, H, W = idx.shape
        mh, mw = mp.shape[1:3]
        flat = np.zeros((V, mh, mw, 2), f32)
        sa = np.zeros((V, mh, mw, 6), f32)
        sb = np.zeros((V, mh, mw, 6), f32)
        t = (C.c_double * 4)()
        lab = np.ascontiguousarray(lab, f32) if lab is not None else np.zeros((1,), f32)
        self.lib.hosttwin_refine(_p(np.ascontiguousarray(mp, f32)), _p(np.ascontiguousarray(idx, u32)),
                                 _p(np.ascontiguousarray(rep, u8)), _p(lab), V, W, H, mw, mh, aw, ah, S, _cf(bl_ratio),
                                 _p(np.ascontiguousarray(subset, i32)), _p(np.ascontiguousarray(subset_num, i32)),
                                 _cf(gamma_), _cf(alpha_), _cf(fuse_), kernel_step, kernel_size, no_prop, _p(flat), _p(sa),
                                 _p(sb), t)
        return flat, sa, sb, {"flatness": t[0], "init_state": t[1], "propagate": t[2]}
