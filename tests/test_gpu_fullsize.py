"""Full-size parity: the bench workload itself (15 views, 1920x1080, shipped literals,
clMVDE.cpp:14-36) run on the GPU stage by stage, each stage checked bit for bit against the CPU
oracle where the oracle finishes in seconds: one whole view of SLIC, the cheap stages on every view,
and bands of superpixel rows of block matching and of every propagate iteration for a corner, an edge
and an interior view (each oracle call takes the GPU's own input of that stage, so a band is an exact
check of that stage).  Plus size-independent properties on the whole frame."""
import numpy as np
import pytest

import cl_multiview_stereo_b200 as mvs
from tests.test_gpu_parity import assert_bits_equal

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def full_frame(ctx):
    import bench
    s, W, H, rgbx = bench.make_input("15view_1080p")
    f = mvs.Frame(ctx, s, W, H)
    f.upload(mvs.BUF_RGBX, rgbx)
    yield s, W, H, rgbx, f
    f.close()


def test_fullsize_stagewise_bands(ctx, oracle, full_frame):
    s, W, H, rgbx, f = full_frame
    V, aw = s.array_width * s.array_height, s.array_width
    S = s.spixl_size
    mw, mh = mvs.map_size(W, H, S)
    M = mw * mh

    # ---- SLIC: one whole view against the oracle, all views through invariants
    f.slic()
    idx = f.download(mvs.BUF_IDX)
    lab = f.download(mvs.BUF_LAB)
    mp = f.download(mvs.BUF_MAP)
    assert idx.max() < M
    z = 7
    olab, omap, oidx = oracle.slic(rgbx[z], S, s.slic_color_weight, s.no_iter, s.enforce_connectivity)
    assert np.array_equal(idx[z], oidx)
    assert_bits_equal(lab[z], olab, "lab of view 7")
    assert_bits_equal(mp[z], omap, "centres of view 7")
    # the count lane of every centre is the number of pixels carrying that label inside the window the
    # reference scans; summed over a view it can only lose pixels (Appendix B S5), never invent them
    assert (mp[..., 6].reshape(V, -1).sum(1) <= W * H).all()

    # ---- boundary (all views) and block matching (row bands)
    f.block_match()
    rep = f.download(mvs.BUF_REP)
    assert np.array_equal(rep, oracle.boundary(mp, idx, S))
    mp_bm = f.download(mvs.BUF_MAP)
    assert_bits_equal(mp_bm[..., :7], mp[..., :7], "block matching must only touch s7")
    levels = oracle.make_levels(s.min_disp, s.max_disp, s.inc)
    assert np.isin(mp_bm[..., 7], levels).all()
    sub, num = oracle.make_subsets(aw, s.array_height, s.neib_hor, s.neib_ver)
    for z, y0 in ((0, 0), (2, 66), (7, 131)):   # corner view / top band, edge view / middle, interior / bottom
        want = oracle.block_match_rows(lab, mp, rep, levels, sub, num, aw, s.bl_ratio, z, y0, y0 + 4)
        assert np.array_equal(mp_bm[z, y0:y0 + 4, :, 7], want[z, y0:y0 + 4, :, 7]), (z, y0)

    # ---- refinement
    gam, alp, fus, kss = oracle.refine_params(2 * s.gamma ** 2, 2 * s.alpha ** 2, s.fuse, s.kernel_step,
                                              s.kernel_size // 2, S)
    f.flatness()
    flat = f.download(mvs.BUF_FLAT)
    assert_bits_equal(flat, oracle.flatness(mp_bm, np.float32(1.0 / float(np.float32(2 * s.gamma ** 2)))), "flatness")
    f.init_state()
    st = f.download(mvs.BUF_STATE_A)
    assert_bits_equal(st, oracle.init_state(mp_bm, idx, rep, flat, sub, num, aw, s.bl_ratio, gam, alp, fus,
                                            s.kernel_step, kss), "init state")
    states = {0: st}
    for it in range(s.no_prop):
        f.propagate(it)
        got = f.download(mvs.BUF_STATE_B if it % 2 == 0 else mvs.BUF_STATE_A)
        want = np.zeros_like(got)
        for z, y0 in ((0, 0), (9, 60), (7, 131)):
            oracle.propagate_rows(st, want, mp_bm, idx, rep, flat, sub, num, it, aw, s.bl_ratio, gam, alp, fus,
                                  s.kernel_step // (it + 1), np.float32(kss) / np.float32(it + 1), z, y0, y0 + 4)
            assert_bits_equal(got[z, y0:y0 + 4], want[z, y0:y0 + 4], f"propagate {it}, view {z}, rows {y0}..{y0 + 3}")
        # plane normals stay unit length or become the zero vector (normalize(0) = 0, clcode.cl:1702)
        n2 = (got[..., 3:6].astype(np.float64) ** 2).sum(-1)
        fin = np.isfinite(n2)
        assert (np.abs(n2[fin] - 1.0) < 1e-4).mean() > 0.99
        st = got
        states[it + 1] = got
    f.fusion()
    disp = f.download(mvs.BUF_DISP)
    # fusion reads buffer A: with no_prop = 5 that is iteration 3's output (Appendix B R7)
    last_a = states[s.no_prop if s.no_prop % 2 == 0 else s.no_prop - 1]
    assert_bits_equal(disp, oracle.spixl_to_image(mp_bm, idx, last_a), "disparity image")


def test_fullsize_run_equals_stagewise_and_is_deterministic(ctx, full_frame):
    s, W, H, rgbx, f = full_frame
    f.run()
    a = f.download(mvs.BUF_DISP)
    sa = f.download(mvs.BUF_STATE_A)
    f.run()
    assert np.array_equal(a.view(np.uint32), f.download(mvs.BUF_DISP).view(np.uint32))
    assert np.array_equal(sa.view(np.uint32), f.download(mvs.BUF_STATE_A).view(np.uint32))
