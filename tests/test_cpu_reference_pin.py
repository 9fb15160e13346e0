"""Pins the CPU oracle (oracle/oracle.c) to the reference itself.

  * live: the reference's own device source, clcode.cl compiled as C++ (oracle/ref ->
    oracle/_ref/libclref.so), is run beside the restatement on seeded inputs -- every stage must
    agree bit for bit.  Needs /root/reference at build time; skipped where it never existed.
  * golden: tests/golden/*.npz hold outputs that libclref produced here
    (tests/golden/make_golden.py); the restatement must reproduce them bit for bit anywhere.
"""
import glob
import json
import os

import numpy as np
import pytest

import cl_multiview_stereo_b200 as mvs
from cl_multiview_stereo_b200 import synth
from tests import clref

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def assert_same(got, want, what):
    got, want = np.asarray(got), np.asarray(want)
    assert got.shape == want.shape, f"{what}: shape {got.shape} vs {want.shape}"
    if got.dtype == np.float32:
        neq = (bits(got) != bits(want)) & ~(np.isnan(got) & np.isnan(want))
    else:
        neq = got != want
    if neq.any():
        k = tuple(np.argwhere(neq)[0])
        raise AssertionError(f"{what}: {int(neq.sum())} of {neq.size} differ, first at {k}: {got[k]!r} vs {want[k]!r}")


@pytest.fixture(scope="module")
def ref():
    r = clref.load()
    if r is None:
        pytest.skip("oracle/_ref/libclref.so absent and /root/reference not present to build it")
    return r


@pytest.mark.parametrize("W,H,S,conn,iters", [(64, 48, 8, False, 2), (45, 37, 8, True, 1), (64, 48, 16, False, 1),
                                              (50, 40, 10, False, 1)])
def test_slic_restatement_equals_reference_source(oracle, ref, W, H, S, conn, iters):
    img = synth.voronoi_image(W + S, W, H, sites=18)
    mxy, mcol = oracle.slic_params(S)
    lab_r, mp_r, idx_r = ref.slic(img, S, 0.6, iters, conn, mxy, mcol)
    lab_o, mp_o, idx_o = oracle.slic(img, S, 0.6, iters, conn)
    assert_same(lab_o, lab_r, "lab")
    assert_same(idx_o, idx_r, "labels")
    assert_same(mp_o, mp_r, "spixl_map")


@pytest.mark.parametrize("aw,ah,W,H,bl,levels,ksize,nh,nv", [
    (3, 2, 96, 64, 1.0, (0, 16, 1), 128, 1, 1), (2, 2, 77, 53, 0.625, (2, 40, 2), 64, 1, 1),
    (5, 3, 64, 48, 1.0359, (0, 12, 1), 1080, 1, 1),
    (5, 3, 64, 48, 1.0, (0, 10, 1), 128, 2, 1), (4, 4, 56, 40, 1.0, (0, 10, 1), 128, 2, 2)])   # > 8 neighbours, wide baselines
def test_depth_restatement_equals_reference_source(oracle, ref, aw, ah, W, H, bl, levels, ksize, nh, nv):
    """block matching, flatness, init state, every propagate iteration, fusion"""
    s = mvs.SystemSettings(array_width=aw, array_height=ah, min_disp=levels[0], max_disp=levels[1], inc=levels[2],
                           bl_ratio=bl, kernel_size=ksize, neib_hor=nh, neib_ver=nv)
    rgbx, _ = synth.make_frame(aw * 10 + ah, W, H, aw, ah, bl_ratio=bl, n_patches=4, dmin=levels[0] + 2,
                               dmax=min(levels[1] - 2, 14))
    o = oracle.pipeline(rgbx, s)          # SLIC by the restatement (checked against the source above)
    S = s.spixl_size
    lv = oracle.make_levels(s.min_disp, s.max_disp, s.inc)
    sub, num = oracle.make_subsets(aw, ah, nh, nv)
    assert_same(o["rep"], ref.boundary(o["map_slic"], o["idx"], S), "spixl_rep")
    assert_same(o["map"], ref.block_match(o["lab"], o["map_slic"], o["rep"], o["idx"], lv, sub, num, aw, bl, S), "initD")
    g_, a_ = np.float32(2 * s.gamma ** 2), np.float32(2 * s.alpha ** 2)
    gam, alp, fus, kss = oracle.refine_params(g_, a_, s.fuse, s.kernel_step, s.kernel_size // 2, S)
    assert_same(o["flat"], ref.flatness(o["map"], np.float32(1.0 / float(g_))), "flatness")
    st_o = oracle.init_state(o["map"], o["idx"], o["rep"], o["flat"], sub, num, aw, bl, gam, alp, fus, s.kernel_step, kss)
    st_r = ref.init_state(o["map"], o["idx"], o["rep"], o["flat"], sub, num, aw, bl, gam, alp, fus, s.kernel_step, kss)
    assert_same(st_o, st_r, "init state")
    for it in range(s.no_prop):
        args = (o["map"], o["idx"], o["rep"], o["flat"], sub, num, it, aw, bl, gam, alp, fus, s.kernel_step // (it + 1),
                np.float32(kss) / np.float32(it + 1))
        nxt_o, nxt_r = oracle.propagate(st_o, *args), ref.propagate(st_o, *args)
        assert_same(nxt_o, nxt_r, f"state after propagate {it}")
        st_o = nxt_o
    assert_same(o["disp"], ref.spixl_to_image(o["map"], o["idx"], o["state_a"]), "disparity")


def golden_files(prefix):
    return sorted(glob.glob(os.path.join(GOLDEN, prefix + "*.npz")))


def settings_from(g):
    d = json.loads(str(g["settings"]))
    return mvs.SystemSettings(**d)


@pytest.mark.parametrize("path", golden_files("synth_") + golden_files("bar_"), ids=os.path.basename)
def test_restatement_reproduces_pipeline_golden(oracle, path):
    g = np.load(path)
    s = settings_from(g)
    o = oracle.pipeline(g["rgbx"], s)
    for k in ("idx", "map_slic", "rep", "map", "flat", "state_a", "state_b", "disp"):
        assert_same(o[k], g[k], f"{os.path.basename(path)}:{k}")


@pytest.mark.parametrize("path", golden_files("slic_"), ids=os.path.basename)
def test_restatement_reproduces_slic_golden(oracle, path):
    g = np.load(path)
    lab, mp, idx = oracle.slic(g["rgbx"], int(g["S"]), float(g["weight"]), int(g["no_iter"]), bool(g["connectivity"]))
    assert_same(lab, g["lab"], "lab")
    assert_same(idx, g["idx"], "labels")
    assert_same(mp, g["map"], "spixl_map")


def test_golden_set_is_present():
    names = {os.path.basename(p) for p in glob.glob(os.path.join(GOLDEN, "*.npz"))}
    assert {"synth_3x2.npz", "bar_5x3_crop.npz", "slic_dance_s8_conn.npz", "slic_horse_s16.npz",
            "slic_horse_s8.npz", "fusion_synth_3x2.npz", "fusion_bar_5x3_crop.npz"} <= names


def _fusion_input(seed, aw, ah, W, H, bl):
    """per-view disparity images of a synthetic scene: ground truth + blocky errors + holes (exact 0)"""
    rgbx, gt = synth.make_frame(seed, W, H, aw, ah, bl_ratio=bl, n_patches=5, dmin=2, dmax=9)
    rng = np.random.default_rng(seed)
    d = gt.astype(np.float32)
    noise = rng.normal(0, 0.4, size=d.shape).astype(np.float32)
    d = d + noise
    wrong = rng.random(d.shape) < 0.05
    d[wrong] = rng.uniform(0, 12, size=int(wrong.sum())).astype(np.float32)
    d[rng.random(d.shape) < 0.02] = 0.0
    return np.ascontiguousarray(d, np.float32)


@pytest.mark.parametrize("aw,ah,W,H,bl,fuse", [(3, 2, 64, 48, 1.0, 1.0), (2, 2, 45, 37, 0.625, 2.0), (5, 3, 48, 32, 1.0359, 1.0),
                                               (1, 1, 20, 17, 1.0, 1.0)])
def test_view_fusion_restatement_equals_reference_source(oracle, ref, aw, ah, W, H, bl, fuse):
    d = _fusion_input(aw * 100 + W, aw, ah, W, H, bl)
    proj_r, out_r = ref.view_fusion(d, aw, bl, fuse)
    proj_o, out_o = oracle.view_fusion(d, aw, bl, fuse)
    assert_same(proj_o, proj_r, "disp_proj")
    assert_same(out_o, out_r, "fused disparity")
    # what the stage is for: a fused value is 0 or one of the candidates at that pixel
    assert ((out_o == 0) | (out_o[:, None] == proj_o[None]).any(1)).all()


@pytest.mark.parametrize("path", golden_files("fusion_"), ids=os.path.basename)
def test_restatement_reproduces_view_fusion_golden(oracle, path):
    g = np.load(path)
    proj, fused = oracle.view_fusion(g["disp"], int(g["array_width"]), float(g["bl_ratio"]), float(g["fuse"]))
    assert_same(proj, g["proj"], "disp_proj")
    assert_same(fused, g["fused"], "fused disparity")


def test_reference_host_twins_build_and_run():
    """the reference's CPU twins (*_host methods) compile unmodified over the inert headers of oracle/ref_host and
    run on a small frame; where a twin computes what the kernel computes (flatness: the same terms, N before S) it
    agrees to the tolerance of the reference's own compare_host_to_device_flatness (1e-4, depth_refinement.cpp:209)"""
    from tests import hosttwin, orc
    if not hosttwin.available():
        pytest.skip("no /root/reference and no prebuilt oracle/_ref/libhosttwin.so")
    h = hosttwin.load()
    o = orc.load()
    import cl_multiview_stereo_b200 as mvs
    from cl_multiview_stereo_b200 import synth
    s = mvs.SystemSettings(array_width=5, array_height=3, min_disp=0, max_disp=16, inc=1, bl_ratio=0.625, kernel_size=128)
    rgbx, _ = synth.make_frame(7, 96, 64, 5, 3, bl_ratio=0.625, n_patches=4, dmin=4, dmax=14)
    ref = o.pipeline(rgbx, s)
    sub, num = o.make_subsets(5, 3)
    levels = o.make_levels(0, 16, 1)
    mp, rep, sec = h.block_match(ref["lab"], ref["map_slic"], ref["idx"], levels, sub, num, 5, 0.625, 8)
    assert np.isin(mp[..., 7], levels).all()
    # array_width = 5 and bl_ratio = 0.625 are what the twin hard-codes (photo_consistency.cpp:325-326): on this
    # geometry it mostly lands on the kernel's level (it has no centre clamp and truncates its steps, H3-H5)
    assert (mp[..., 7] == ref["map"][..., 7]).mean() > 0.5
    flat, sa, sb, t = h.refine(ref["map"], ref["idx"], ref["rep"], ref["lab"], 5, 3, 8, 0.625, sub, num, np.float32(8), np.float32(72),
                               1.0, 13, 64, 3)
    assert np.abs(flat - ref["flat"]).max() < 1e-4
    assert np.isfinite(sa[..., :3]).all() and (sa[..., 5] != 0).any()


@pytest.mark.parametrize("W,H,S", [(61, 43, 8), (96, 64, 8), (75, 50, 16)])
def test_edge_aware_slic_pinned_to_reference_source(W, H, S):
    """edge_enable (clSLIC.cpp:186-233).  The reference's first kernel overwrites the buffer it reads and leaves the
    edge image unwritten (clcode.cl:176-194), so its result is undefined as written; the pinned semantics (DESIGN.md
    section 3: Sobel of a Lab snapshot into the edge image) are held against the reference's own source where it IS
    defined: the arithmetic of edge_compute_alternative per work-item on a private snapshot, and apply_edge_alternative
    on a given edge map."""
    from tests import orc
    r = clref.load()
    if r is None:
        pytest.skip("oracle/_ref/libclref.so not built")
    o = orc.load()
    img = synth.voronoi_image(40 + W, W, H, sites=24)
    lab = o.cvt(img)
    edge = o.edge_compute(lab)
    assert_same(edge, r.edge_compute_snapshot(lab), "edge magnitude")
    mp = o.init_centers(lab, S)
    moved = o.apply_edge(lab, edge, mp)
    assert_same(moved[..., :7], r.apply_edge(lab, edge, mp)[..., :7], "centres after apply_edge_alternative")
    assert (moved[..., 1:3] != mp[..., 1:3]).any(), "the test image should move some centre"
    # the whole segmentation with the flag: stages as clSLIC::do_super_pixel_seg orders them (clSLIC.cpp:83-96)
    lab2, mp2, idx2 = o.slic_edge(img, S, 0.6, 3, False, True)
    mxy, mcol = o.slic_params(S)
    ix = r.assoc(lab, moved, S, 0.6, mxy, mcol)
    m = moved
    for _ in range(3):
        m = o.update_centers(lab, ix, m, S)
        ix = r.assoc(lab, m, S, 0.6, mxy, mcol)
    assert np.array_equal(idx2, ix)
    assert_same(mp2[..., :7], m[..., :7], "centres of the edge-aware segmentation")
