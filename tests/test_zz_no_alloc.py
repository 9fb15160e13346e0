"""mvs_view_fusion_static: the no-allocation variant of the multi-view fusion vote (scratch in a file-scope device
array of the library) against the allocating entry points, through the C ABI.  In a file of its own that pytest
collects last; passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest

import cl_multiview_stereo_b200 as mvs
from cl_multiview_stereo_b200 import synth

pytestmark = pytest.mark.gpu


def _settings(aw, ah):
    return mvs.SystemSettings(array_width=aw, array_height=ah, min_disp=0, max_disp=8, inc=1, bl_ratio=1.0, kernel_size=128)


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_static_scratch_fusion_equals_the_allocating_entry_points_twice_in_a_row(ctx):
    """two calls in a row on DIFFERENT frames (the second must not see what the first left in the shared scratch), then
    the first again; bit-exact against mvs_frame_view_fusion on the resident frame and mvs_view_fusion_run on host arrays"""
    aw, ah, W, H = 3, 3, 96, 64
    V = aw * ah
    s = _settings(aw, ah)
    frames, want = [], []
    out = mvs.Frame(ctx, s, W, H)          # its disparity buffer is the caller-owned OUTPUT of the new entry point
    try:
        for seed in (21, 22):
            rgbx, _ = synth.make_frame(seed, W, H, aw, ah, bl_ratio=1.0, n_patches=3, dmin=1, dmax=6)
            f = mvs.Frame(ctx, s, W, H)
            frames.append(f)
            f.upload(mvs.BUF_RGBX, rgbx)
            f.run()
            f.view_fusion()                                   # the existing entry point (allocates disp_proj / disp_fused)
            fused = f.download(mvs.BUF_DISP_FUSED)
            _, fused_run = mvs.view_fusion_run(ctx, f.download(mvs.BUF_DISP), aw, s.bl_ratio, s.fuse)
            assert np.array_equal(_bits(fused), _bits(fused_run))
            assert np.count_nonzero(fused) > 0                # not vacuous
            want.append(fused)
        assert not np.array_equal(_bits(want[0]), _bits(want[1]))
        launches = ctx.launches
        for k in (0, 1, 0):
            out.upload(mvs.BUF_DISP, np.full((V, H, W), -7.0, np.float32))      # poison the output
            mvs.view_fusion_static(ctx, frames[k].dev_ptr(mvs.BUF_DISP), out.dev_ptr(mvs.BUF_DISP), V, W, H, aw,
                                   s.bl_ratio, s.fuse)
            ctx.sync()
            got = out.download(mvs.BUF_DISP)
            assert np.array_equal(_bits(got), _bits(want[k])), f"call on frame {k}"
        assert ctx.launches - launches == 6                   # two kernels per call, counted as ours
    finally:
        for f in frames:
            f.close()
        out.close()


def test_static_scratch_fusion_refuses_what_does_not_fit(ctx):
    """beyond MVS_VIEW_FUSION_STATIC_MAX_FLOATS disparities the entry point refuses (before touching any pointer); host
    pointers are refused too"""
    s = _settings(3, 3)
    f = mvs.Frame(ctx, s, 96, 64)
    try:
        p = f.dev_ptr(mvs.BUF_DISP)
        with pytest.raises(mvs.MvsError, match="static scratch holds"):
            mvs.view_fusion_static(ctx, p, p, 15, 1920, 1080, 5, 1.0, 1.0)       # 31 M disparities > 4 Mi
        assert 9 * 683 * 683 > mvs.VIEW_FUSION_STATIC_MAX_FLOATS
        with pytest.raises(mvs.MvsError, match="static scratch holds"):
            mvs.view_fusion_static(ctx, p, p, 9, 683, 683, 3, 1.0, 1.0)          # just beyond the limit
        host = np.zeros((9, 64, 96), np.float32)
        with pytest.raises(mvs.MvsError, match="not a device pointer"):
            mvs.view_fusion_static(ctx, host.ctypes.data, p, 9, 96, 64, 3, 1.0, 1.0)
        with pytest.raises(mvs.MvsError, match="multiple of array_width"):
            mvs.view_fusion_static(ctx, p, p, 9, 96, 64, 4, 1.0, 1.0)
    finally:
        f.close()
