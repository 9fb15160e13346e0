"""ctypes binding of the C ABI (include/mvs_capi.h).  Fails loudly when the CUDA library is
missing: there is no CPU fallback on the product path."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmvs_b200.so")


class MvsError(RuntimeError):
    pass


class Params(C.Structure):
    """mvs_params (include/mvs_capi.h) -- mirrors system_settings, reference header.h:55-77."""
    _fields_ = [
        ("width", C.c_int), ("height", C.c_int), ("spixl_size", C.c_int),
        ("array_width", C.c_int), ("array_height", C.c_int),
        ("slic_color_weight", C.c_float), ("no_iter", C.c_int), ("enforce_connectivity", C.c_int),
        ("neib_hor", C.c_int), ("neib_ver", C.c_int),
        ("min_disp", C.c_int), ("max_disp", C.c_int), ("inc", C.c_int), ("bl_ratio", C.c_float),
        ("kernel_size", C.c_int), ("kernel_step", C.c_int),
        ("fuse", C.c_float), ("gamma", C.c_float), ("alpha", C.c_float), ("no_prop", C.c_int),
        ("edge_enable", C.c_int),
    ]


# every symbol include/mvs_capi.h declares (tests check the library exports all of them)
SYMBOLS = [
    "mvs_last_error", "mvs_device_count", "mvs_ctx_create", "mvs_ctx_destroy", "mvs_ctx_sync",
    "mvs_ctx_launch_count", "mvs_ctx_stream", "mvs_ctx_event_record", "mvs_ctx_event_elapsed", "mvs_ctx_wait_event", "mvs_comm_unique_id", "mvs_comm_init", "mvs_host_alloc",
    "mvs_host_free", "mvs_view_partition", "mvs_exchange_plan", "mvs_frame_create", "mvs_frame_destroy", "mvs_frame_upload",
    "mvs_frame_download", "mvs_frame_upload_async", "mvs_frame_download_async", "mvs_frame_dev_ptr",
    "mvs_frame_sizes", "mvs_frame_slic", "mvs_frame_block_match", "mvs_frame_refine", "mvs_frame_flatness",
    "mvs_frame_init_state", "mvs_frame_propagate", "mvs_frame_fusion", "mvs_frame_run", "mvs_frame_exchange",
    "mvs_frame_stage_ms", "mvs_frame_set_timing", "mvs_slic_run", "mvs_bm_run", "mvs_refine_run",
    "mvs_frame_view_fusion", "mvs_view_fusion_run", "mvs_frame_upload_bgr", "mvs_frame_export_u8",
    "mvs_frame_draw_segmentation", "mvs_view_partition_weighted", "mvs_exchange_plan_for", "mvs_frame_set_partition",
    "mvs_frame_propagate_stats", "mvs_draw_segmentation_run", "mvs_slic_run_edge", "mvs_frame_set_halo_inputs", "mvs_frame_halo_views", "mvs_math_eval", "mvs_view_fusion_static", "mvs_frame_set_view_order", "mvs_view_order_balanced",
]

_lib = None


def _check_fresh():
    """refuse a library that was not built from the sources beside it (the .so is git-ignored and travels to the
    GPU box prebuilt: this is what ties it to the commit).  MVS_ALLOW_STALE=1 skips the check."""
    if os.environ.get("MVS_ALLOW_STALE"):
        return
    from . import build as b
    stamp = LIB_PATH + ".srchash"
    if not os.path.exists(stamp):
        raise MvsError(f"{stamp} is missing: rebuild with `python -m cl_multiview_stereo_b200.build`")
    have = open(stamp).read().split()
    cu, deps = b.native_sources()
    want = b._source_hash(cu + deps)
    if not have or have[0] != want:
        raise MvsError(f"{LIB_PATH} was built from other sources (hash {have[0][:12] if have else '?'}, "
                       f"tree {want[:12]}): rebuild with `python -m cl_multiview_stereo_b200.build`")


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MvsError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a).  There is no CPU fallback.")
        _check_fresh()
        _lib = C.CDLL(LIB_PATH)
        _lib.mvs_last_error.restype = C.c_char_p
        for s in SYMBOLS:
            if s != "mvs_last_error":
                getattr(_lib, s).restype = C.c_int
    return _lib


def check(rc):
    if rc != 0:
        raise MvsError(lib().mvs_last_error().decode("utf-8", "replace"))
