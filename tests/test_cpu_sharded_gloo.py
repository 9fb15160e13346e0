"""N > 1 host logic on CPU: the view partition and the halo-exchange plan of the C ABI
(mvs_view_partition, mvs_exchange_plan -- what mvs_frame_exchange issues as grouped
ncclSend/ncclRecv) drive a world_size-2 `gloo` run in which each rank computes only its own
views with the CPU oracle, receives only the planned halo views, and must reproduce the
single-process result bit for bit.  Views that are neither owned nor planned stay poisoned, so a
missing transfer shows up as a mismatch (dependencies reach the grid neighbours only).
"""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import cl_multiview_stereo_b200 as mvs  # noqa: E402
from cl_multiview_stereo_b200 import synth  # noqa: E402


@pytest.mark.parametrize("aw,ah,nh,nv", [(5, 3, 1, 1), (4, 4, 1, 1), (5, 5, 1, 1), (3, 3, 2, 1), (2, 4, 1, 1), (7, 1, 1, 1)])
@pytest.mark.parametrize("n", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("weighted", [False, True])
def test_exchange_plan_delivers_every_needed_view_once(aw, ah, nh, nv, n, weighted):
    V = aw * ah
    sub, num = mvs.make_subsets(aw, ah, nh, nv)
    if weighted:
        parts = mvs.view_partition_weighted(mvs.view_weights(num), n)
        plan = mvs.exchange_plan(V, n, sub, num, parts)
    else:
        plan = mvs.exchange_plan(V, n, sub, num)
        parts = [mvs.view_partition(V, n, r) for r in range(n)]
    owner = {v: r for r, (lo, hi) in enumerate(parts) for v in range(lo, hi)}
    assert sorted(owner) == list(range(V))
    got = {}
    for v, src, dst in plan:
        assert src == owner[v] and dst != src
        assert (v, dst) not in got, "view sent twice to the same rank"
        got[(v, dst)] = src
    for r, (lo, hi) in enumerate(parts):
        need = {int(sub[u, k]) for u in range(lo, hi) for k in range(num[u])}
        need = {v for v in need if owner[v] != r}
        assert need == {v for (v, dst) in got if dst == r}
    if n == 1:
        assert plan == []


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


AW, AH, W, H = 2, 4, 64, 48


def _settings():
    return mvs.SystemSettings(array_width=AW, array_height=AH, min_disp=0, max_disp=10, inc=1, bl_ratio=1.0,
                              kernel_size=64, no_prop=3, no_iter=2)


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    from tests import orc
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    o = orc.load()
    s = _settings()
    V = AW * AH
    rgbx_all, _ = synth.make_frame(5, W, H, AW, AH, 1.0, 4, 2.0, 8.0)
    lo, hi = mvs.view_partition(V, world, rank)
    sub, num = mvs.make_subsets(AW, AH, s.neib_hor, s.neib_ver)
    plan = mvs.exchange_plan(V, world, sub, num)
    mw, mh = mvs.map_size(W, H, s.spixl_size)

    def exchange(arr):
        reqs = []
        bufs = []
        for v, src, dst in plan:
            if src == rank:
                reqs.append(dist.isend(torch.from_numpy(np.ascontiguousarray(arr[v]).view(np.uint8).reshape(-1).copy()), dst, tag=v))
            if dst == rank:
                t = torch.empty(arr[v].nbytes, dtype=torch.uint8)
                bufs.append((v, t))
                reqs.append(dist.irecv(t, src, tag=v))
        for r in reqs:
            r.wait()
        for v, t in bufs:
            arr[v] = t.numpy().view(arr.dtype).reshape(arr[v].shape)

    # this rank's frame: own views real, everything else poisoned.  The sequence below is mvs_frame_run's on a
    # sharded frame (capi.cu halo_inputs / exchange_after_block_match / mvs_frame_init_state / propagate_impl).
    rgbx = np.full_like(rgbx_all, 0)
    rgbx[lo:hi] = rgbx_all[lo:hi]
    lab = np.full((V, H, W, 4), np.nan, np.float32)
    mp = np.full((V, mh, mw, 8), np.nan, np.float32)
    idx = np.zeros((V, H, W), np.uint32)
    halo = sorted({v for v, src, dst in plan if dst == rank})
    # (1) the halo views' INPUT images arrive first (sent by their owners here; the caller may upload them itself
    #     instead, mvs_frame_set_halo_inputs) and their Lab is recomputed locally
    exchange(rgbx)
    for v in halo:
        lab[v] = o.cvt(rgbx[v])
    for v in range(lo, hi):
        lab[v], mp[v], idx[v] = o.slic(rgbx[v], s.spixl_size, s.slic_color_weight, s.no_iter, s.enforce_connectivity)
    # (2) block matching of the own views reads only Lab of the neighbours: no labels / centres needed yet
    levels = o.make_levels(s.min_disp, s.max_disp, s.inc)
    mp_clean = np.where(np.isnan(mp), np.float32(0), mp)   # boundary walks every view; poisoned centres are not used below
    rep = o.boundary(mp_clean, idx, s.spixl_size)
    own_map = o.block_match(lab, mp, rep, levels, sub, num, AW, s.bl_ratio, lo, hi)
    mp[lo:hi] = own_map[lo:hi]
    # (3) ONE group after block matching: labels and centres (initial disparity included) of the halo views
    exchange(idx); exchange(mp)
    g = np.float32(2 * float(s.gamma) ** 2)
    a = np.float32(2 * float(s.alpha) ** 2)
    gam, alp, fus, kss = o.refine_params(g, a, s.fuse, s.kernel_step, s.kernel_size // 2, s.spixl_size)
    flat = o.flatness(mp, np.float32(1.0 / float(g)))
    st = o.init_state(mp, idx, rep, flat, sub, num, AW, s.bl_ratio, gam, alp, fus, s.kernel_step, kss)
    poison = np.ones(V, bool)
    poison[lo:hi] = False
    st[poison] = np.nan
    # (4) iteration 0 needs no plane exchange: the halo views' initial planes are {s7, (0, 0, 1)} of their centres
    for v in halo:
        st[v, ..., 0] = mp[v, ..., 7]
        st[v, ..., 3:6] = (0.0, 0.0, 1.0)
    states = []
    for it in range(s.no_prop):
        if it > 0:
            exchange(st)
        nxt = o.propagate(st, mp, idx, rep, flat, sub, num, it, AW, s.bl_ratio, gam, alp, fus, s.kernel_step // (it + 1),
                          np.float32(kss) / np.float32(it + 1), lo, hi)
        nxt[poison] = np.nan
        states.append(nxt[lo:hi].copy())
        st = nxt
    # multi-view fusion on the sharded frame (mvs_frame_view_fusion): own disparity images, every rank sends its
    # own to every other rank, chases the slices 0..hi-1 itself and votes its own views in view order
    st_a = states[s.no_prop - 1] if s.no_prop % 2 == 0 else states[s.no_prop - 2]   # fusion reads buffer A
    st_full = np.full((V, mh, mw, 6), np.nan, np.float32)
    st_full[lo:hi] = st_a
    disp = np.zeros((V, H, W), np.float32)
    disp[lo:hi] = o.spixl_to_image(mp, idx, st_full)[lo:hi]
    reqs, bufs = [], []
    for r in range(world):
        if r == rank:
            continue
        rlo, rhi = mvs.view_partition(V, world, r)
        if hi > lo:
            reqs.append(dist.isend(torch.from_numpy(disp[lo:hi].reshape(-1).copy()), r, tag=1000 + rank))
        if rhi > rlo:
            t = torch.empty((rhi - rlo) * H * W, dtype=torch.float32)
            bufs.append((rlo, rhi, t))
            reqs.append(dist.irecv(t, r, tag=1000 + r))
    for q in reqs:
        q.wait()
    for rlo, rhi, t in bufs:
        disp[rlo:rhi] = t.numpy().reshape(rhi - rlo, H, W)
    proj, fused = np.zeros_like(disp), np.zeros_like(disp)
    for r in range(hi):
        o.project_inv(disp, proj, r, AW, s.bl_ratio)
        if r >= lo:
            o.remove_inconsistency(proj, disp, fused, r, AW, s.bl_ratio, 0.5 * s.fuse)
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), lo=lo, hi=hi, idx=idx[lo:hi], map=mp[lo:hi],
             states=np.stack(states), halo=np.array(halo), disp=disp[lo:hi], fused=fused[lo:hi])
    dist.barrier()
    dist.destroy_process_group()


def test_view_sharded_over_gloo_equals_single_process(oracle, tmp_path):
    import torch.multiprocessing as tmp_mp
    world = 2
    port = _free_port()
    tmp_mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)

    s = _settings()
    V = AW * AH
    rgbx, _ = synth.make_frame(5, W, H, AW, AH, 1.0, 4, 2.0, 8.0)
    ref = oracle.pipeline(rgbx, s)
    sub, num = mvs.make_subsets(AW, AH, s.neib_hor, s.neib_ver)
    g = np.float32(2 * float(s.gamma) ** 2)
    a = np.float32(2 * float(s.alpha) ** 2)
    gam, alp, fus, kss = oracle.refine_params(g, a, s.fuse, s.kernel_step, s.kernel_size // 2, s.spixl_size)
    st = oracle.init_state(ref["map"], ref["idx"], ref["rep"], ref["flat"], sub, num, AW, s.bl_ratio, gam, alp, fus,
                           s.kernel_step, kss)
    full_states = []
    for it in range(s.no_prop):
        st = oracle.propagate(st, ref["map"], ref["idx"], ref["rep"], ref["flat"], sub, num, it, AW, s.bl_ratio, gam, alp,
                              fus, s.kernel_step // (it + 1), np.float32(kss) / np.float32(it + 1))
        full_states.append(st)
    proj_full, fused_full = oracle.view_fusion(ref["disp"], AW, s.bl_ratio, s.fuse)
    seen = 0
    for r in range(world):
        d = np.load(tmp_path / f"rank{r}.npz")
        lo, hi = int(d["lo"]), int(d["hi"])
        assert (lo, hi) == mvs.view_partition(V, world, r)
        # 2 x 4 grid over two ranks: the far camera row is never shipped
        assert len(d["halo"]) == AW
        assert np.array_equal(d["idx"], ref["idx"][lo:hi])
        assert np.array_equal(d["map"].view(np.uint32), ref["map"][lo:hi].view(np.uint32))
        for it in range(s.no_prop):
            assert np.array_equal(d["states"][it].view(np.uint32), full_states[it][lo:hi].view(np.uint32)), \
                f"rank {r}: state after iteration {it} differs from the single-process run"
        assert np.array_equal(d["disp"].view(np.uint32), ref["disp"][lo:hi].view(np.uint32))
        assert np.array_equal(d["fused"].view(np.uint32), fused_full[lo:hi].view(np.uint32)), \
            f"rank {r}: sharded multi-view fusion differs from the single-process run"
        seen += hi - lo
    assert seen == V
