"""CPU tests (no GPU): the C-ABI library loads and exports every symbol include/mvs_capi.h
declares (no compute calls), the host-side helpers agree with the oracle's restatement of
pipeline.cpp:121-142, and the oracle's pinned math sits within a few ULP of libm."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import cl_multiview_stereo_b200 as mvs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "mvs_capi.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(mvs_[a-z0-9_]+)\s*\(", txt)))


def test_header_declares_what_python_binds():
    assert set(header_symbols()) == set(mvs.SYMBOLS)


def test_library_exports_every_declared_symbol():
    lib = mvs.lib()
    for s in header_symbols():
        assert hasattr(lib, s), f"libmvs_b200.so does not export {s}"


def test_params_struct_layout():
    # mvs_params: 21 four-byte fields (edge_enable last), no padding
    assert C.sizeof(mvs.Params) == 84


def test_no_cpu_fallback_without_device():
    """without a CUDA device every compute entry point fails loudly"""
    lib = mvs.lib()
    n = C.c_int(-1)
    rc = lib.mvs_device_count(C.byref(n))
    if rc == 0 and n.value > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(mvs.MvsError):
        mvs.Context(0)


def test_view_partition_covers_all_views():
    for V in (1, 9, 15, 16, 25):
        for n in (1, 2, 4, 8):
            spans = [mvs.view_partition(V, n, r) for r in range(n)]
            assert spans[0][0] == 0 and spans[-1][1] == V
            for a, b in zip(spans, spans[1:]):
                assert a[1] == b[0]
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def test_levels_and_subsets_match_oracle(oracle):
    # pipeline.cpp:121-124 and :130-142 (x outer, y inner, self excluded)
    assert np.array_equal(mvs.make_levels(30, 60, 1), oracle.make_levels(30, 60, 1))
    assert len(mvs.make_levels(10, 100, 2)) == 46
    for aw, ah in ((5, 3), (3, 3), (4, 4), (1, 1), (5, 5)):
        sub, num = mvs.make_subsets(aw, ah)
        osub, onum = oracle.make_subsets(aw, ah)
        assert np.array_equal(num, onum)
        for z in range(aw * ah):
            assert np.array_equal(sub[z, :num[z]], osub[z, :onum[z]])
    sub, num = mvs.make_subsets(5, 3)
    assert list(sub[6, :num[6]]) == [0, 5, 10, 1, 11, 2, 7, 12]


def ulp_diff(a, b):
    a = np.asarray(a, np.float32).view(np.int32).astype(np.int64)
    b = np.asarray(b, np.float32).view(np.int32).astype(np.int64)
    a = np.where(a < 0, -(a & 0x7FFFFFFF), a)
    b = np.where(b < 0, -(b & 0x7FFFFFFF), b)
    return np.abs(a - b)


def test_pinned_expf_close_to_libm(oracle):
    rng = np.random.default_rng(1)
    x = np.concatenate([rng.uniform(-87, 88, 200000), rng.uniform(-2, 2, 100000), [0.0, -0.0, 1.0, -1.0]]).astype(np.float32)
    y = np.empty_like(x)
    oracle.lib.orc_math_expf(x.ctypes.data_as(C.c_void_p), y.ctypes.data_as(C.c_void_p), len(x))
    want = np.exp(x.astype(np.float64)).astype(np.float32)
    assert ulp_diff(y, want).max() <= 2
    # edges: overflow, underflow, NaN
    e = np.array([89.0, 1000.0, -104.5, -1000.0, np.nan, -100.0], np.float32)
    r = np.empty_like(e)
    oracle.lib.orc_math_expf(e.ctypes.data_as(C.c_void_p), r.ctypes.data_as(C.c_void_p), len(e))
    assert np.isinf(r[0]) and np.isinf(r[1]) and r[2] == 0 and r[3] == 0 and np.isnan(r[4])
    assert r[5] > 0 and ulp_diff(r[5:6], np.float32(np.exp(-100.0))).max() <= 1   # denormal result


def test_pinned_exp_double_close_to_libm(oracle):
    rng = np.random.default_rng(2)
    x = rng.uniform(-50, 50, 100000)
    y = np.empty_like(x)
    oracle.lib.orc_math_exp(x.ctypes.data_as(C.c_void_p), y.ctypes.data_as(C.c_void_p), len(x))
    assert np.max(np.abs(y / np.exp(x) - 1)) < 5e-16


def test_pinned_cbrt_close_to_libm(oracle):
    """the division-free cube root: <= 1 ulp over the whole positive normal range (sampled) and over EVERY float
    of the interval rgb2lab can feed it, exact on exact cubes"""
    rng = np.random.default_rng(3)
    lo, hi = np.float32(0.008856).view(np.uint32), np.float32(1.2).view(np.uint32)
    every = np.arange(int(lo), int(hi) + 1, dtype=np.uint32).view(np.float32)
    wide = np.exp(rng.uniform(np.log(1e-37), np.log(1e37), 400000)).astype(np.float32)
    x = np.concatenate([every, wide, np.array([0.008857, 1.0, 0.125, 8.0], np.float32)])
    y = np.empty_like(x)
    oracle.lib.orc_math_cbrtf(x.ctypes.data_as(C.c_void_p), y.ctypes.data_as(C.c_void_p), len(x))
    want = np.cbrt(x.astype(np.float64)).astype(np.float32)
    assert ulp_diff(y, want).max() <= 1
    assert y[-3] == 1.0 and y[-2] == 0.5 and y[-1] == 2.0


def test_round_and_truncation_semantics(oracle):
    x = np.array([0.5, -0.5, 1.5, 2.5, -2.5, 0.49999997, -0.9, 0.9, 3e9, -3e9, np.nan], np.float32)
    r = np.empty_like(x)
    oracle.lib.orc_math_roundf(x.ctypes.data_as(C.c_void_p), r.ctypes.data_as(C.c_void_p), len(x))
    assert list(r[:8]) == [1.0, -1.0, 2.0, 3.0, -3.0, 0.0, -1.0, 1.0]      # half away from zero (R4)
    i = np.empty(len(x), np.int32)
    oracle.lib.orc_math_f2i(x.ctypes.data_as(C.c_void_p), i.ctypes.data_as(C.c_void_p), len(x))
    assert list(i[:8]) == [0, 0, 1, 2, -2, 0, 0, 0]                          # toward zero (B5)
    assert i[8] == 2147483647 and i[9] == -2147483648 and i[10] == 0          # cvt.rzi saturation


def test_host_byte_work_restatement(oracle):
    """loadImageIn / plot_full_image restated in oracle.c against plain numpy (file_handler.cpp:6-15,
    depth_refinement.cpp:1487)"""
    rng = np.random.default_rng(5)
    bgr = rng.integers(0, 256, size=(9, 7, 3), dtype=np.uint8)
    out = oracle.load_image_in(bgr)
    assert np.array_equal(out[..., :3], bgr[..., ::-1]) and (out[..., 3] == 0).all()
    d = rng.uniform(30, 60, size=(4, 5)).astype(np.float32)
    want = np.floor((d - np.float32(30)) / np.float32(30) * np.float32(255)).astype(np.uint8)
    assert np.array_equal(oracle.plot_full_image(d, 30, 60), want)
    hostile = np.array([np.nan, np.inf, -np.inf, 1e20, -1.0, 61.0], np.float32)
    assert list(oracle.plot_full_image(hostile, 30, 60)) == [0, 0, 0, 0, (-264) & 255, 263 & 255]


def test_visualisation_restatements(oracle):
    """img_translate, img_translate_flatness and draw_segmentation_lines restated in oracle.c against plain numpy
    (photo_consistency.cpp:414-436, depth_refinement.cpp:299-321, clSLIC.cpp:447-478)"""
    rng = np.random.default_rng(11)
    V, H, W, mh, mw = 2, 9, 11, 3, 4
    idx = rng.integers(0, mh * mw, size=(V, H, W), dtype=np.uint32)
    mp = rng.uniform(30, 60, size=(V, mh, mw, 8)).astype(np.float32)
    flat = rng.uniform(0, 1, size=(V, mh, mw, 2)).astype(np.float32)
    flat[0, 0, 0, 0], flat[0, 0, 1, 0] = 1.0, 0.0
    d = np.stack([mp[v].reshape(-1, 8)[idx[v], 7] for v in range(V)])
    want = np.floor((d - np.float32(30)) / np.float32(30) * np.float32(255)).astype(np.uint8)
    assert np.array_equal(oracle.img_translate(mp, idx, 30, 60), want)
    fl = np.stack([flat[v].reshape(-1, 2)[idx[v], 0] for v in range(V)])
    want = np.ceil(fl * np.float32(255)).astype(np.int32).astype(np.uint8)
    assert np.array_equal(oracle.img_translate_flatness(flat, idx), want)
    # overlay: the frame and the pad byte of `out` are never written
    lab = (np.arange(H)[:, None] // 4 * 3 + np.arange(W)[None, :] // 4).astype(np.uint32)
    img = rng.integers(0, 256, size=(H, W, 4), dtype=np.uint8)
    out0 = rng.integers(0, 256, size=(H, W, 4), dtype=np.uint8)
    out = oracle.draw_segmentation_lines(lab, img, out0.copy())
    edge = np.zeros((H, W), bool)
    c = lab[1:-1, 1:-1]
    edge[1:-1, 1:-1] = (c != lab[1:-1, 2:]) | (c != lab[1:-1, :-2]) | (c != lab[:-2, 1:-1]) | (c != lab[2:, 1:-1])
    want = out0.copy()
    want[1:-1, 1:-1, :3] = img[1:-1, 1:-1, :3]
    want[edge, :3] = (0, 0, 255)
    assert np.array_equal(out, want) and edge.any() and not edge.all()
    tiny = out0[:2, :2].copy()
    assert np.array_equal(oracle.draw_segmentation_lines(lab[:2, :2], img[:2, :2].copy(), tiny.copy()), tiny)


def test_weighted_view_partition_minimises_the_heaviest_rank():
    """mvs_view_partition_weighted against brute force over every contiguous split"""
    import itertools
    rng = np.random.default_rng(2)
    cases = [(mvs.view_weights(mvs.make_subsets(5, 3)[1], b, p), n) for n in (1, 2, 4, 8) for b, p in ((33, 10), (5, 6))]
    cases += [(list(rng.integers(1, 60, size=V)), n) for V in (1, 3, 7, 10) for n in (1, 2, 3, 5)]
    for w, n in cases:
        V = len(w)
        parts = mvs.view_partition_weighted(w, n)
        assert len(parts) == n and parts[0][0] == 0 and parts[-1][1] == V
        assert all(a[1] == b[0] for a, b in zip(parts, parts[1:])) and all(lo <= hi for lo, hi in parts)
        got = max(sum(w[lo:hi]) for lo, hi in parts)
        best = min(max(sum(w[a:b]) for a, b in zip((0,) + cuts, cuts + (V,)))
                   for cuts in itertools.combinations_with_replacement(range(V + 1), n - 1))
        assert got == best, (w, n, parts)
        if V >= n:
            assert all(hi > lo for lo, hi in parts), "a rank left without a view although there are enough"
    # the 15-view grid on 8 GPUs, work counted in neighbour views only: the heaviest rank carries 13 instead of the
    # even split's 16
    num = mvs.make_subsets(5, 3)[1]
    load = lambda parts: max(sum(int(k) for k in num[lo:hi]) for lo, hi in parts)    # noqa: E731
    assert load(mvs.view_partition_weighted([int(k) for k in num], 8)) == 13
    assert load([mvs.view_partition(15, 8, r) for r in range(8)]) == 16
    with pytest.raises(mvs.MvsError):
        mvs.view_partition_weighted([3, 0, 2], 2)
    with pytest.raises(mvs.MvsError):
        mvs.exchange_plan(4, 2, *mvs.make_subsets(2, 2), parts=[(0, 3), (2, 4)])


def test_balanced_view_order_is_a_permutation_and_beats_contiguous_ranges():
    """mvs_view_order_balanced: a permutation of the cameras with contiguous slot ranges; under its own cost model the
    heaviest rank is never heavier than with the row-major even or weighted ranges, and strictly lighter on the
    15-view array over 8 ranks (where a contiguous run cannot separate the interior cameras)"""
    wb, wn, wh = 19, 20, 16

    def cost(sets, sub, num):
        worst = 0
        for cams in sets:
            own = sum(wb + wn * int(num[c]) for c in cams)
            read = {int(sub[c, k]) for c in cams for k in range(num[c])} - set(cams)
            worst = max(worst, own + wh * len(read))
        return worst

    for aw, ah, nh, nv in ((5, 3, 1, 1), (4, 4, 1, 1), (5, 5, 1, 1), (3, 3, 2, 1), (7, 1, 1, 1), (2, 2, 1, 1)):
        V = aw * ah
        sub, num = mvs.make_subsets(aw, ah, nh, nv)
        for n in (1, 2, 3, 4, 8):
            order, parts = mvs.view_order_balanced(aw, ah, n, nh, nv, wb, wn, wh)
            assert sorted(order) == list(range(V))
            assert parts[0][0] == 0 and parts[-1][1] == V and all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
            assert (order, parts) == mvs.view_order_balanced(aw, ah, n, nh, nv, wb, wn, wh)      # deterministic
            mine = cost([order[lo:hi] for lo, hi in parts], sub, num)
            even = cost([list(range(*mvs.view_partition(V, n, r))) for r in range(n)], sub, num)
            wtd = cost([list(range(lo, hi)) for lo, hi in mvs.view_partition_weighted(mvs.view_weights(num, wb, wn), n)], sub, num)
            assert mine <= min(even, wtd), (aw, ah, n, mine, even, wtd)
            if (aw, ah, n) == (5, 3, 8):
                assert mine < 0.9 * min(even, wtd)
    with pytest.raises(mvs.MvsError):
        mvs.view_order_balanced(5, 3, 0)
