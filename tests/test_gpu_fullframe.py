"""Whole-frame parity at full size: EVERY stage, EVERY view, EVERY superpixel of a 1920x1080 frame
against the CPU oracle, on the bench's synthetic 15-view frame and on the reference's own data (the
bundled 15-view set c0f1..c14f1 under the shipped literals and under the legacy 15-view settings, and
the 3x3 Beer-Garden set the shipped main is configured for).

Each oracle stage is fed the GPU's own output of the stage before it, so a difference is pinned to
the stage that made it; the bar is identical bits (labels, rays, initial disparities: integers /
levels; Lab, centres, flatness, state, disparity image: float bit patterns, NaN = NaN).  The oracle
needs about ten seconds per frame on the box's cores.

Plus: the propagate kernel on a state in which no two superpixels share a plane (no duplicate
candidates to skip), parity bands at the sizes of the other two workloads (4x4 1080p, 5x5 4K), and the
view-sharded frame against the single-GPU frame on two devices.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

import cl_multiview_stereo_b200 as mvs
from cl_multiview_stereo_b200 import datasets, synth
from tests.test_gpu_parity import assert_bits_equal

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def legacy_settings(aw, ah):
    """settings B of SURVEY 8(d): the 15-view run the reference's results/ folder came from (clMVDE.cpp:26-29
    comments, photo_consistency.cpp:325-326): levels 10..100 step 2, bl_ratio 0.625"""
    return mvs.SystemSettings(array_width=aw, array_height=ah, min_disp=10, max_disp=100, inc=2, bl_ratio=0.625)


def check_whole_frame(ctx, oracle, rgbx, s, what):
    """stage by stage, all views; returns the unique-plane statistics of the propagate iterations"""
    V, H, W, _ = rgbx.shape
    S, aw = s.spixl_size, s.array_width
    f = mvs.Frame(ctx, s, W, H)
    try:
        f.upload(mvs.BUF_RGBX, rgbx)
        # ---- SLIC, every view
        f.slic()
        lab, idx, mp = f.download(mvs.BUF_LAB), f.download(mvs.BUF_IDX), f.download(mvs.BUF_MAP)
        for z in range(V):
            olab, omap, oidx = oracle.slic(rgbx[z], S, s.slic_color_weight, s.no_iter, s.enforce_connectivity)
            assert np.array_equal(idx[z], oidx), f"{what}: labels of view {z} differ at {(idx[z] != oidx).sum()} pixels"
            assert_bits_equal(lab[z], olab, f"{what}: Lab of view {z}")
            assert_bits_equal(mp[z], omap, f"{what}: centres of view {z}")
        # ---- boundary + block matching, every superpixel
        f.block_match()
        rep, mp_bm = f.download(mvs.BUF_REP), f.download(mvs.BUF_MAP)
        assert np.array_equal(rep, oracle.boundary(mp, idx, S)), f"{what}: rays"
        levels = oracle.make_levels(s.min_disp, s.max_disp, s.inc)
        sub, num = oracle.make_subsets(aw, s.array_height, s.neib_hor, s.neib_ver)
        want = oracle.block_match(lab, mp, rep, levels, sub, num, aw, s.bl_ratio)
        assert_bits_equal(mp_bm, want, f"{what}: centres + initial disparities")
        # ---- refinement
        g_, a_ = np.float32(2 * s.gamma ** 2), np.float32(2 * s.alpha ** 2)
        gam, alp, fus, kss = oracle.refine_params(g_, a_, s.fuse, s.kernel_step, s.kernel_size // 2, S)
        f.flatness()
        flat = f.download(mvs.BUF_FLAT)
        assert_bits_equal(flat, oracle.flatness(mp_bm, np.float32(1.0 / float(g_))), f"{what}: flatness")
        f.init_state()
        st = f.download(mvs.BUF_STATE_A)
        assert_bits_equal(st, oracle.init_state(mp_bm, idx, rep, flat, sub, num, aw, s.bl_ratio, gam, alp, fus,
                                                s.kernel_step, kss), f"{what}: initial state")
        states = {0: st}
        stats = []
        for it in range(s.no_prop):
            stats.append(f.propagate_stats(it))
            got = f.download(mvs.BUF_STATE_B if it % 2 == 0 else mvs.BUF_STATE_A)
            want = oracle.propagate(st, mp_bm, idx, rep, flat, sub, num, it, aw, s.bl_ratio, gam, alp, fus,
                                    s.kernel_step // (it + 1), np.float32(kss) / np.float32(it + 1))
            assert_bits_equal(got, want, f"{what}: state after propagate {it}")
            st = got
            states[it + 1] = got
        f.fusion()
        disp = f.download(mvs.BUF_DISP)
        last_a = states[s.no_prop if s.no_prop % 2 == 0 else s.no_prop - 1]   # fusion reads buffer A (R7)
        assert_bits_equal(disp, oracle.spixl_to_image(mp_bm, idx, last_a), f"{what}: disparity image")
        # the one-call pipeline gives the same frame
        f.upload(mvs.BUF_RGBX, rgbx)
        f.run()
        assert_bits_equal(f.download(mvs.BUF_DISP), disp, f"{what}: mvs_frame_run vs the staged calls")
        return stats
    finally:
        f.close()


def test_whole_frame_synthetic_15view_1080p(ctx, oracle):
    import bench
    s, W, H, rgbx = bench.make_input("15view_1080p")
    stats = check_whole_frame(ctx, oracle, rgbx, s, "synthetic 15-view 1080p")
    assert all(st["superpixels"] == 15 * 240 * 135 for st in stats)


def need(name):
    if not datasets.available(name):
        pytest.skip(f"bundled set {name} not under tests/data/_ref (run tests/data/fetch_bundled.py where /root/reference exists)")


def test_whole_frame_bundled_15view_shipped_literals(ctx, oracle):
    need("bar15")
    rgbx, aw, ah = datasets.load_set("bar15")
    check_whole_frame(ctx, oracle, rgbx, mvs.SystemSettings(array_width=aw, array_height=ah), "bundled c0f1..c14f1, settings A")


def test_whole_frame_bundled_15view_legacy_settings(ctx, oracle):
    need("bar15")
    rgbx, aw, ah = datasets.load_set("bar15")
    check_whole_frame(ctx, oracle, rgbx, legacy_settings(aw, ah), "bundled c0f1..c14f1, settings B")


def test_whole_frame_beer_garden_3x3(ctx, oracle):
    need("beer9")
    rgbx, aw, ah = datasets.load_set("beer9")
    check_whole_frame(ctx, oracle, rgbx, mvs.SystemSettings(array_width=aw, array_height=ah), "Beer-Garden 3x3")


def test_propagate_without_duplicate_planes(ctx, oracle):
    """every superpixel starts from its own random plane: no candidate of any superpixel is a bit-identical
    duplicate of another (beyond the pair planes of round 1), the case the dedup stage cannot help with"""
    s = mvs.SystemSettings(array_width=3, array_height=3, min_disp=0, max_disp=16, inc=1, bl_ratio=1.0, kernel_size=128)
    W, H = 256, 160
    rgbx, _ = synth.make_frame(21, W, H, 3, 3, bl_ratio=1.0, n_patches=6, dmin=4, dmax=14)
    ref = oracle.pipeline(rgbx, s)
    rng = np.random.default_rng(5)
    V, mh, mw = ref["state_a"].shape[:3]
    st = ref["state_a"].copy()
    n = rng.normal(size=(V, mh, mw, 3)).astype(np.float32)
    n[..., 2] = np.abs(n[..., 2]) + 0.5
    n /= np.linalg.norm(n, axis=-1, keepdims=True).astype(np.float32)
    st[..., 0] = rng.uniform(2, 15, size=(V, mh, mw)).astype(np.float32)
    st[..., 3:6] = n
    sub, num = oracle.make_subsets(3, 3)
    gam, alp, fus, kss = oracle.refine_params(2 * s.gamma ** 2, 2 * s.alpha ** 2, s.fuse, s.kernel_step,
                                              s.kernel_size // 2, s.spixl_size)
    f = mvs.Frame(ctx, s, W, H)
    try:
        f.upload(mvs.BUF_MAP, ref["map"]); f.upload(mvs.BUF_IDX, ref["idx"]); f.upload(mvs.BUF_REP, ref["rep"])
        f.flatness()
        f.upload(mvs.BUF_STATE_A, st)
        for it in range(3):
            stats = f.propagate_stats(it)
            got = f.download(mvs.BUF_STATE_B if it % 2 == 0 else mvs.BUF_STATE_A)
            want = oracle.propagate(st, ref["map"], ref["idx"], ref["rep"], ref["flat"], sub, num, it, 3, s.bl_ratio,
                                    gam, alp, fus, s.kernel_step // (it + 1), np.float32(kss) / np.float32(it + 1))
            assert_bits_equal(got, want, f"propagate {it} from unique planes")
            if it == 0:   # every candidate of round 0 is unique; round 1 can only repeat a degenerate pair plane
                assert stats["unique_planes"] >= 0.95 * stats["candidates"], stats
            st = got
    finally:
        f.close()


@pytest.mark.parametrize("workload,views", [("4x4_1080p", (0, 5, 15)), ("5x5_4k", (0, 12, 24))])
def test_other_workloads_parity_bands(ctx, oracle, workload, views):
    """the other two named workloads at their full size: SLIC of one view, and four-row bands of block matching
    and of every propagate iteration for a corner, an interior and the last view, each oracle call on the GPU's
    own input of that stage"""
    import bench
    s, W, H, rgbx = bench.make_input(workload)
    V, aw, S = s.array_width * s.array_height, s.array_width, s.spixl_size
    mw, mh = mvs.map_size(W, H, S)
    f = mvs.Frame(ctx, s, W, H)
    try:
        f.upload(mvs.BUF_RGBX, rgbx)
        f.slic()
        z = views[1]
        idx, lab, mp = f.download(mvs.BUF_IDX), f.download(mvs.BUF_LAB), f.download(mvs.BUF_MAP)
        olab, omap, oidx = oracle.slic(rgbx[z], S, s.slic_color_weight, s.no_iter, s.enforce_connectivity)
        assert np.array_equal(idx[z], oidx)
        assert_bits_equal(mp[z], omap, f"{workload}: centres of view {z}")
        f.block_match()
        rep, mp_bm = f.download(mvs.BUF_REP), f.download(mvs.BUF_MAP)
        assert np.array_equal(rep, oracle.boundary(mp, idx, S))
        levels = oracle.make_levels(s.min_disp, s.max_disp, s.inc)
        sub, num = oracle.make_subsets(aw, s.array_height, s.neib_hor, s.neib_ver)
        bands = list(zip(views, (0, mh // 2, mh - 4)))
        for z, y0 in bands:
            want = oracle.block_match_rows(lab, mp, rep, levels, sub, num, aw, s.bl_ratio, z, y0, y0 + 4)
            assert np.array_equal(mp_bm[z, y0:y0 + 4, :, 7], want[z, y0:y0 + 4, :, 7]), (workload, z, y0)
        del lab
        gam, alp, fus, kss = oracle.refine_params(2 * s.gamma ** 2, 2 * s.alpha ** 2, s.fuse, s.kernel_step,
                                                  s.kernel_size // 2, S)
        f.flatness(); f.init_state()
        flat, st = f.download(mvs.BUF_FLAT), f.download(mvs.BUF_STATE_A)
        for it in range(s.no_prop):
            f.propagate(it)
            got = f.download(mvs.BUF_STATE_B if it % 2 == 0 else mvs.BUF_STATE_A)
            want = np.zeros_like(got)
            for z, y0 in bands:
                oracle.propagate_rows(st, want, mp_bm, idx, rep, flat, sub, num, it, aw, s.bl_ratio, gam, alp, fus,
                                      s.kernel_step // (it + 1), np.float32(kss) / np.float32(it + 1), z, y0, y0 + 4)
                assert_bits_equal(got[z, y0:y0 + 4], want[z, y0:y0 + 4], f"{workload}: propagate {it}, view {z}, rows {y0}..")
            st = got
    finally:
        f.close()


def test_view_sharded_frame_equals_single_gpu_on_two_devices():
    """two ranks, one GPU each, NCCL halo exchange: every buffer of the sharded frame equals the single-GPU frame
    bit for bit (tools/sharded_check.py asserts it per rank and prints one line per rank)"""
    ndev = mvs.device_count()
    if ndev < 2:
        pytest.skip(f"needs two CUDA devices, this box has {ndev}")
    port = 29500 + os.getpid() % 400
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tools", "sharded_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert out.stdout.count("SHARDED_OK") == 2, out.stdout[-2000:]
