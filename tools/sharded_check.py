#!/usr/bin/env python
"""One frame sharded by camera view over the ranks of a torchrun launch (one GPU per rank, NCCL halo
exchange through the C ABI) against the same frame computed whole on this rank's GPU: every buffer a
rank owns must hold the same bits.  Prints SHARDED_OK <rank> per rank; exits non-zero on a difference.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 tools/sharded_check.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import cl_multiview_stereo_b200 as mvs
    from cl_multiview_stereo_b200 import synth
    s = mvs.SystemSettings(array_width=4, array_height=3, min_disp=0, max_disp=16, inc=1, bl_ratio=1.0, kernel_size=128)
    W, H = 256, 160
    V = s.array_width * s.array_height
    rgbx, _ = synth.make_frame(31, W, H, s.array_width, s.array_height, bl_ratio=1.0, n_patches=6, dmin=4, dmax=14)
    ctx = mvs.Context(local)
    whole = mvs.Frame(ctx, s, W, H)
    whole.upload(mvs.BUF_RGBX, rgbx)
    whole.run()
    whole.view_fusion()
    uid = [mvs.Context.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    ctx.comm_init(uid[0], world, rank)
    bad = []
    # (halo images sent by their owners | uploaded by this rank itself) x (row-major view ranges | camera sets in a
    # balanced slot order, which has no multi-view fusion)
    border, btable = mvs.view_order_balanced(s.array_width, s.array_height, world, w_halo=40)
    if border == list(range(V)):       # make sure a non-trivial order is exercised
        border = border[::-1]
    for local_halo, order, table in ((False, None, None), (True, None, None), (False, border, btable), (True, border, btable)):
      lo, hi = table[rank] if table else mvs.view_partition(V, world, rank)
      cams = order if order else list(range(V))          # slot -> camera
      part = mvs.Frame(ctx, s, W, H, lo, hi)
      if order:
        part.set_view_order(order)
        part.set_partition(table)
      if hi > lo:
        part.upload(mvs.BUF_RGBX, np.ascontiguousarray(rgbx[cams[lo:hi]]), lo, hi)
      if local_halo:
        for v in part.halo_views():
            part.upload(mvs.BUF_RGBX, rgbx[cams[v]:cams[v] + 1], v, v + 1)
        part.set_halo_inputs(True)
      part.run()
      bufs = [("labels", mvs.BUF_IDX), ("centres", mvs.BUF_MAP), ("rays", mvs.BUF_REP), ("flatness", mvs.BUF_FLAT),
              ("state A", mvs.BUF_STATE_A), ("state B", mvs.BUF_STATE_B), ("disparity", mvs.BUF_DISP)]
      if not order:
        part.view_fusion()
        bufs.append(("fused disparity", mvs.BUF_DISP_FUSED))
      for name, which in bufs:
        if hi > lo:
            a = part.download(which, lo, hi)
            b = np.stack([whole.download(which, c, c + 1)[0] for c in cams[lo:hi]])
            if not np.array_equal(np.ascontiguousarray(a).view(np.uint8), np.ascontiguousarray(b).view(np.uint8)):
                bad.append(f"{name} (halo inputs {'local' if local_halo else 'exchanged'}, {'balanced order' if order else 'row-major'})")
      part.close()
    lo, hi = mvs.view_partition(V, world, rank)
    flag = torch.tensor([len(bad)], device=f"cuda:{local}")
    dist.all_reduce(flag)
    whole.close(); ctx.close()
    dist.barrier()
    dist.destroy_process_group()
    if bad:
        print(f"rank {rank}: sharded frame differs from the single-GPU frame in: {bad}", flush=True)
    if int(flag.item()):
        sys.exit(1)
    print(f"SHARDED_OK {rank} views [{lo},{hi}) of {V} on {world} ranks", flush=True)


if __name__ == "__main__":
    main()
