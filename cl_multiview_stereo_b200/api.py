"""Python host-side mirror of the reference's stage interface, over the C ABI.

The reference's host language is C++ (the drop-in classes are in include/ and host/); this
module exists for the tests and bench.py, which drive the same C entry points through ctypes.
Names follow the reference: Context ~ pipeline::compile_device_code (pipeline.cpp:31-56),
Frame.slic ~ clSLIC::do_super_pixel_seg (clSLIC.cpp:67), Frame.block_match ~
clPhotoConsistency::do_initial_depth_estimation (photo_consistency.cpp:21), Frame.refine ~
clDepthRefinement::do_refinement (depth_refinement.cpp:91).
"""
import ctypes as C
import math
from dataclasses import dataclass

import numpy as np

from ._native import MvsError, Params, check, lib

BUF_RGBX, BUF_LAB, BUF_IDX, BUF_MAP, BUF_REP, BUF_FLAT, BUF_STATE_A, BUF_STATE_B, BUF_DISP = range(9)
BUF_DISP_PROJ, BUF_DISP_FUSED = 9, 10   # multi-view fusion (allocated by Frame.view_fusion)
STAGES = ("slic", "boundary", "block_match", "flatness", "init_state", "propagate", "fusion", "exchange")


@dataclass
class SystemSettings:
    """system_settings (reference header.h:55-77) with the literals of clMVDE.cpp:14-36."""
    spixl_size: int = 8
    slic_color_weight: float = 0.6
    array_width: int = 3
    array_height: int = 3
    no_iter: int = 5
    enforce_connectivity: bool = False
    edge_enable: bool = False
    num_disp_levels: int = 30
    neib_hor: int = 1
    neib_ver: int = 1
    min_disp: int = 30
    max_disp: int = 60
    inc: int = 1
    bl_ratio: float = 1.03590
    kernel_size: int = 1080
    kernel_step: int = 13
    fuse: float = 1.0
    gamma: float = 2.0
    alpha: float = 6.0
    no_prop: int = 5

    def params(self, width, height):
        return Params(width, height, self.spixl_size, self.array_width, self.array_height,
                      self.slic_color_weight, self.no_iter, int(self.enforce_connectivity),
                      self.neib_hor, self.neib_ver, self.min_disp, self.max_disp, self.inc, self.bl_ratio,
                      self.kernel_size, self.kernel_step, self.fuse, self.gamma, self.alpha, self.no_prop,
                      int(self.edge_enable))


def map_size(width, height, S):
    """pipeline.cpp:18-19"""
    return int(math.ceil(np.float32(width) / np.float32(S))), int(math.ceil(np.float32(height) / np.float32(S)))


def make_levels(min_disp, max_disp, inc):
    """pipeline.cpp:121-124"""
    return np.array([min_disp + i * inc for i in range((max_disp - min_disp) // inc + 1)], dtype=np.float32)


def make_subsets(aw, ah, neib_hor=1, neib_ver=1):
    """pipeline.cpp:133-142 flattened as photo_consistency.cpp:39-47 (row stride V)."""
    V = aw * ah
    mat = np.zeros((V, V), dtype=np.int32)
    num = np.zeros(V, dtype=np.int32)
    for i in range(V):
        for x in range(i % aw - neib_hor, i % aw + neib_hor + 1):
            for y in range(i // aw - neib_ver, i // aw + neib_ver + 1):
                indx = y * aw + x
                if 0 <= x < aw and 0 <= y < ah and indx != i:
                    mat[i, num[i]] = indx
                    num[i] += 1
    return mat, num


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class Context:
    def __init__(self, device=0):
        self._h = C.c_void_p()
        check(lib().mvs_ctx_create(int(device), C.byref(self._h)))
        self.device = device

    def close(self):
        if self._h:
            lib().mvs_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sync(self):
        check(lib().mvs_ctx_sync(self._h))

    @property
    def launches(self):
        n = C.c_longlong()
        check(lib().mvs_ctx_launch_count(self._h, C.byref(n)))
        return n.value

    @property
    def stream(self):
        s = C.c_void_p()
        check(lib().mvs_ctx_stream(self._h, C.byref(s)))
        return s.value or 0

    def event_record(self, slot):
        check(lib().mvs_ctx_event_record(self._h, int(slot)))

    def wait_event(self, other, slot):
        """work queued on this context from now on runs after `other`'s event `slot`"""
        check(lib().mvs_ctx_wait_event(self._h, other._h, int(slot)))

    def event_elapsed_ms(self, a, b):
        ms = C.c_float()
        check(lib().mvs_ctx_event_elapsed(self._h, int(a), int(b), C.byref(ms)))
        return ms.value

    @staticmethod
    def unique_id():
        buf = (C.c_char * 128)()
        check(lib().mvs_comm_unique_id(buf))
        return bytes(buf)

    def comm_init(self, uid, nranks, rank):
        buf = (C.c_char * 128).from_buffer_copy(uid)
        check(lib().mvs_comm_init(self._h, buf, int(nranks), int(rank)))


def device_count():
    n = C.c_int()
    check(lib().mvs_device_count(C.byref(n)))
    return n.value


def math_eval(ctx, which, x):
    """the device's pinned math on an array (diagnostic): 0 expf, 1 expf for x <= 0, 2 polynomial-only expf for
    |x| < 0.3465, 3 cube root, 4 round, 5 float -> int (returned as int32)"""
    x = np.ascontiguousarray(x, np.float32)
    y = np.empty_like(x)
    check(lib().mvs_math_eval(ctx._h, int(which), _ptr(x), _ptr(y), C.c_size_t(x.size)))
    return y.view(np.int32) if which == 5 else y


def view_partition(V, nranks, rank):
    lo, hi = C.c_int(), C.c_int()
    check(lib().mvs_view_partition(int(V), int(nranks), int(rank), C.byref(lo), C.byref(hi)))
    return lo.value, hi.value


def view_partition_weighted(weights, nranks):
    """contiguous view ranges [(lo, hi)] * nranks that minimise the heaviest rank's summed weight"""
    w = np.ascontiguousarray(weights, np.int32)
    out = np.zeros((int(nranks), 2), np.int32)
    check(lib().mvs_view_partition_weighted(len(w), int(nranks), _ptr(w), _ptr(out)))
    return [(int(a), int(b)) for a, b in out]


def view_order_balanced(aw, ah, nranks, neib_hor=1, neib_ver=1, w_base=19, w_neighbour=20, w_halo=4):
    """(camera_of_slot, [(lo, hi)] per rank): a slot order of the cameras whose contiguous slot ranges balance the
    ranks' work (mvs_view_order_balanced).  Default weights: measured per-view times in units of 10 us at 1080p (own
    view: segmentation etc. 0.19 ms + 0.2 ms per neighbour view; halo view: Lab conversion and receiving its labels)."""
    V = aw * ah
    order = (C.c_int * V)()
    lo_hi = (C.c_int * (2 * nranks))()
    check(lib().mvs_view_order_balanced(aw, ah, neib_hor, neib_ver, nranks, w_base, w_neighbour, w_halo, order, lo_hi))
    return list(order), [(lo_hi[2 * r], lo_hi[2 * r + 1]) for r in range(nranks)]


def view_weights(subset_num, base=33, per_neighbour=10):
    """work of a view relative to the others: a constant part (SLIC, boundary, fusion, the smoothness and bookkeeping
    share of propagate) plus a part per neighbour view (block matching, state initialisation, the consistency share of
    propagate).  33 : 10 is the split measured on the 15-view 1080p frame"""
    return [int(base + per_neighbour * int(k)) for k in subset_num]


def exchange_plan(V, nranks, subset, subset_num, parts=None):
    """the halo transfers mvs_frame_exchange issues: list of (view, source rank, destination rank); parts = the
    view ranges of all ranks when they are not view_partition's"""
    subset = np.ascontiguousarray(subset, np.int32)
    subset_num = np.ascontiguousarray(subset_num, np.int32)
    n = C.c_int()
    if parts is None:
        def call(t, cap):
            return lib().mvs_exchange_plan(int(V), int(nranks), _ptr(subset), _ptr(subset_num), t, cap, C.byref(n))
    else:
        table = np.ascontiguousarray(parts, np.int32).reshape(-1)
        if table.size != 2 * int(nranks):
            raise MvsError(f"exchange_plan: {table.size // 2} view ranges for {nranks} ranks")

        def call(t, cap):
            return lib().mvs_exchange_plan_for(int(V), int(nranks), _ptr(table), _ptr(subset), _ptr(subset_num), t, cap,
                                               C.byref(n))
    check(call(None, 0))
    t = np.zeros((max(n.value, 1), 3), np.int32)
    check(call(_ptr(t), n.value))
    return [tuple(int(x) for x in row) for row in t[:n.value]]


class PinnedBuffer:
    """page-locked host memory (cudaHostAlloc) exposed as a numpy array"""

    def __init__(self, shape, dtype):
        self.shape = tuple(shape)
        self.dtype = np.dtype(dtype)
        self.nbytes = int(np.prod(self.shape)) * self.dtype.itemsize
        self._p = C.c_void_p()
        check(lib().mvs_host_alloc(C.byref(self._p), C.c_size_t(max(self.nbytes, 1))))
        raw = (C.c_char * max(self.nbytes, 1)).from_address(self._p.value)
        self.array = np.frombuffer(raw, dtype=self.dtype, count=int(np.prod(self.shape))).reshape(self.shape)

    def free(self):
        if self._p:
            self.array = None
            lib().mvs_host_free(self._p)
            self._p = C.c_void_p()


class Frame:
    """Device-resident buffers of one V-view frame + the stage calls."""

    _SHAPES = {
        BUF_RGBX: ("N", 4, np.uint8), BUF_LAB: ("N", 4, np.float32), BUF_IDX: ("N", 1, np.uint32),
        BUF_MAP: ("M", 8, np.float32), BUF_REP: ("M", 8, np.uint8), BUF_FLAT: ("M", 2, np.float32),
        BUF_STATE_A: ("M", 6, np.float32), BUF_STATE_B: ("M", 6, np.float32), BUF_DISP: ("N", 1, np.float32),
        BUF_DISP_PROJ: ("N", 1, np.float32), BUF_DISP_FUSED: ("N", 1, np.float32),
    }

    def __init__(self, ctx, settings, width, height, view_lo=0, view_hi=None):
        self.ctx = ctx
        self.settings = settings
        self.V = settings.array_width * settings.array_height
        self.W, self.H = width, height
        self.mw, self.mh = map_size(width, height, settings.spixl_size)
        self.lo = view_lo
        self.hi = self.V if view_hi is None else view_hi
        p = settings.params(width, height)
        self._h = C.c_void_p()
        check(lib().mvs_frame_create(ctx._h, C.byref(p), self.lo, self.hi, C.byref(self._h)))

    def close(self):
        if self._h:
            lib().mvs_frame_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def shape(self, which, nviews):
        kind, k, dt = self._SHAPES[which]
        if kind == "N":
            s = (nviews, self.H, self.W) + ((k,) if k > 1 else ())
        else:
            s = (nviews, self.mh, self.mw) + ((k,) if k > 1 else ())
        return s, dt

    def upload(self, which, arr, v0=0, v1=None, sync=True):
        v1 = self.V if v1 is None else v1
        s, dt = self.shape(which, v1 - v0)
        a = np.ascontiguousarray(arr, dtype=dt)
        if a.size != int(np.prod(s)):
            raise MvsError(f"upload: expected {s}, got {arr.shape}")
        fn = lib().mvs_frame_upload if sync else lib().mvs_frame_upload_async
        check(fn(self._h, which, v0, v1, _ptr(a)))

    def download(self, which, v0=0, v1=None, out=None, sync=True):
        v1 = self.V if v1 is None else v1
        s, dt = self.shape(which, v1 - v0)
        if out is None:
            out = np.empty(s, dtype=dt)
        fn = lib().mvs_frame_download if sync else lib().mvs_frame_download_async
        check(fn(self._h, which, v0, v1, _ptr(out)))
        return out

    def upload_bgr(self, bgr, v0=0, v1=None):
        """loadImageIn on the device: packed BGR rasters [v1-v0, H, W, 3] (cv::Mat order) -> BUF_RGBX"""
        v1 = self.V if v1 is None else v1
        a = np.ascontiguousarray(bgr, dtype=np.uint8)
        if a.size != (v1 - v0) * self.H * self.W * 3:
            raise MvsError(f"upload_bgr: expected {(v1 - v0, self.H, self.W, 3)}, got {a.shape}")
        check(lib().mvs_frame_upload_bgr(self._h, v0, v1, _ptr(a)))

    def export_u8(self, which, min_disp, max_disp, element=0, v0=0, v1=None):
        """8-bit visualisation image quantised on the device: plot_full_image (BUF_DISP, BUF_DISP_FUSED),
        img_translate_state (BUF_STATE_A/B + element), img_translate (BUF_MAP), img_translate_flatness (BUF_FLAT)"""
        v1 = self.V if v1 is None else v1
        out = np.empty((v1 - v0, self.H, self.W), np.uint8)
        check(lib().mvs_frame_export_u8(self._h, which, int(element), int(min_disp), int(max_disp), v0, v1, _ptr(out)))
        return out

    def set_view_order(self, camera_of_slot):
        """slot s of every per-view array holds camera camera_of_slot[s] (mvs_frame_set_view_order)"""
        arr = (C.c_int * len(camera_of_slot))(*[int(c) for c in camera_of_slot])
        check(lib().mvs_frame_set_view_order(self._h, arr))

    def set_partition(self, parts):
        """view ranges [(lo, hi)] of ALL ranks when they are not view_partition's (same table on every rank)"""
        table = np.ascontiguousarray(parts, np.int32).reshape(-1)
        check(lib().mvs_frame_set_partition(self._h, table.size // 2, _ptr(table)))

    def draw_segmentation(self, out, v0=0, v1=None):
        """clSLIC::draw_segmentation_lines on the device; `out` uint8 [v1-v0, H, W, 4] is updated in place
        (its one-pixel frame and pad bytes are kept, as in the reference)"""
        v1 = self.V if v1 is None else v1
        if out.dtype != np.uint8 or not out.flags.c_contiguous or out.size != (v1 - v0) * self.H * self.W * 4:
            raise MvsError(f"draw_segmentation: expected contiguous uint8 {(v1 - v0, self.H, self.W, 4)}")
        check(lib().mvs_frame_draw_segmentation(self._h, v0, v1, _ptr(out)))
        return out

    def dev_ptr(self, which):
        p = C.c_void_p()
        check(lib().mvs_frame_dev_ptr(self._h, which, C.byref(p)))
        return p.value

    def slic(self):
        check(lib().mvs_frame_slic(self._h))

    def block_match(self):
        check(lib().mvs_frame_block_match(self._h))

    def flatness(self):
        check(lib().mvs_frame_flatness(self._h))

    def init_state(self):
        check(lib().mvs_frame_init_state(self._h))

    def propagate(self, it):
        check(lib().mvs_frame_propagate(self._h, int(it)))

    def propagate_stats(self, it):
        """one propagate iteration with work counters (diagnostic)"""
        st = (C.c_ulonglong * 12)()
        check(lib().mvs_frame_propagate_stats(self._h, int(it), st))
        return dict(zip(("superpixels", "candidates", "unique_planes", "retired_3", "retired_4", "items",
                         "items_inside", "items_visible", "samples_inside", "samples_visible"), [int(x) for x in st]))

    def fusion(self):
        check(lib().mvs_frame_fusion(self._h))

    def view_fusion(self):
        """multi-view fusion of BUF_DISP (all views) -> BUF_DISP_PROJ, BUF_DISP_FUSED; the stage the
        reference keeps disabled (depth_refinement.cpp:1382-1451), not part of run()"""
        check(lib().mvs_frame_view_fusion(self._h))

    def refine(self):
        check(lib().mvs_frame_refine(self._h))

    def run(self):
        check(lib().mvs_frame_run(self._h))

    def exchange(self, which):
        check(lib().mvs_frame_exchange(self._h, which))

    def halo_views(self):
        """views this rank's own views read but the rank does not own (needs the communicator to be set up)"""
        n = C.c_int()
        check(lib().mvs_frame_halo_views(self._h, None, 0, C.byref(n)))
        v = np.zeros(max(n.value, 1), np.int32)
        check(lib().mvs_frame_halo_views(self._h, _ptr(v), n.value, C.byref(n)))
        return [int(x) for x in v[:n.value]]

    def set_halo_inputs(self, local=True):
        """the caller uploads the halo views' input images itself (no image exchange between ranks)"""
        check(lib().mvs_frame_set_halo_inputs(self._h, int(bool(local))))

    def set_timing(self, on):
        check(lib().mvs_frame_set_timing(self._h, int(bool(on))))

    def stage_ms(self):
        ms = (C.c_float * 8)()
        check(lib().mvs_frame_stage_ms(self._h, ms))
        return dict(zip(STAGES, [float(x) for x in ms]))


# ---- class-level entry points with host arrays (the reference's per-stage semantics) ----
def slic_run(ctx, rgbx, S=8, weight=0.6, no_iter=5, enforce_connectivity=False, edge_enable=False, want_edge=False):
    """clSLIC::do_super_pixel_seg for a batch rgbx[count,H,W,4] -> (lab, map, idx) [+ edge image with want_edge]"""
    rgbx = np.ascontiguousarray(rgbx, dtype=np.uint8)
    if rgbx.ndim == 3:
        rgbx = rgbx[None]
    cnt, H, W, _ = rgbx.shape
    mw, mh = map_size(W, H, S)
    lab = np.empty((cnt, H, W, 4), np.float32)
    mp = np.empty((cnt, mh, mw, 8), np.float32)
    idx = np.empty((cnt, H, W), np.uint32)
    edge = np.empty((cnt, H, W), np.float32) if want_edge else None
    check(lib().mvs_slic_run_edge(ctx._h, W, H, S, C.c_float(weight), no_iter, int(enforce_connectivity), int(edge_enable),
                                  cnt, _ptr(rgbx), _ptr(lab), _ptr(mp), _ptr(idx), _ptr(edge) if want_edge else None))
    if want_edge:
        return lab, mp, idx, edge
    return lab, mp, idx


def bm_run(ctx, mp, lab, idx, S, array_width, bl_ratio, subset_mat, subset_num, levels):
    """clPhotoConsistency::do_initial_depth_estimation -> (map with s7, rep)"""
    mp = np.array(mp, dtype=np.float32, order="C", copy=True)
    lab = np.ascontiguousarray(lab, np.float32)
    idx = np.ascontiguousarray(idx, np.uint32)
    V, H, W = idx.shape
    rep = np.empty(mp.shape[:3] + (8,), np.uint8)
    sm = np.ascontiguousarray(subset_mat, np.int32)
    sn = np.ascontiguousarray(subset_num, np.int32)
    lv = np.ascontiguousarray(levels, np.float32)
    check(lib().mvs_bm_run(ctx._h, V, W, H, S, _ptr(mp), _ptr(rep), _ptr(lab), _ptr(idx), array_width,
                           C.c_float(bl_ratio), _ptr(sm), _ptr(sn), _ptr(lv), len(lv)))
    return mp, rep


def refine_run(ctx, mp, idx, rep, S, aw, ah, bl_ratio, subset_mat, subset_num, gamma, alpha, fuse, kernel_step,
               kernel_size, no_prop):
    """clDepthRefinement ctor + do_refinement -> (flat, state_a, state_b, disp).
    gamma/alpha/kernel_size as pipeline::refine_depth_map passes them (pipeline.cpp:164-166)."""
    mp = np.ascontiguousarray(mp, np.float32)
    idx = np.ascontiguousarray(idx, np.uint32)
    rep = np.ascontiguousarray(rep, np.uint8)
    V, H, W = idx.shape
    mh, mw = mp.shape[1:3]
    flat = np.empty((V, mh, mw, 2), np.float32)
    sa = np.empty((V, mh, mw, 6), np.float32)
    sb = np.empty((V, mh, mw, 6), np.float32)
    disp = np.empty((V, H, W), np.float32)
    sm = np.ascontiguousarray(subset_mat, np.int32)
    sn = np.ascontiguousarray(subset_num, np.int32)
    check(lib().mvs_refine_run(ctx._h, W, H, aw, ah, S, C.c_float(bl_ratio), _ptr(mp), _ptr(idx), _ptr(rep),
                               _ptr(sm), _ptr(sn), C.c_float(gamma), C.c_float(alpha), C.c_float(fuse),
                               kernel_step, kernel_size, no_prop, _ptr(flat), _ptr(sa), _ptr(sb), _ptr(disp)))
    return flat, sa, sb, disp


VIEW_FUSION_STATIC_MAX_FLOATS = 4 << 20


def view_fusion_static(ctx, disp_dev, fused_dev, view_count, width, height, aw, bl_ratio, fuse):
    """mvs_view_fusion_static: the vote of view_fusion_run on DEVICE pointers (ints), allocating nothing; queued on the
    context's stream (ctx.sync() before reading the output).  At most VIEW_FUSION_STATIC_MAX_FLOATS disparities."""
    check(lib().mvs_view_fusion_static(ctx._h, int(view_count), int(width), int(height), int(aw), C.c_float(bl_ratio),
                                       C.c_float(fuse), C.c_void_p(disp_dev), C.c_void_p(fused_dev)))


def view_fusion_run(ctx, disp, aw, bl_ratio, fuse):
    """the disabled tail of clDepthRefinement::fusion (depth_refinement.cpp:1382-1451) on a stack of
    per-view disparity images [V,H,W] -> (disp_proj, disp_fused).  fuse as in system_settings."""
    disp = np.ascontiguousarray(disp, np.float32)
    V, H, W = disp.shape
    proj = np.empty_like(disp)
    out = np.empty_like(disp)
    check(lib().mvs_view_fusion_run(ctx._h, V, W, H, aw, C.c_float(bl_ratio), C.c_float(fuse), _ptr(disp), _ptr(proj),
                                    _ptr(out)))
    return proj, out
