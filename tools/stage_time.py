#!/usr/bin/env python
"""Per-stage CUDA-event times of one workload + a checksum of every output (kernel experiments:
all variants of a kernel must print the same checksums).  python tools/stage_time.py [workload] [steps]"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import cl_multiview_stereo_b200 as mvs  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "15view_1080p"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    s, W, H, rgbx = bench.make_input(name)
    ctx = mvs.Context(0)
    f = mvs.Frame(ctx, s, W, H)
    f.upload(mvs.BUF_RGBX, rgbx)
    for _ in range(2):
        f.run()
    ctx.sync()
    ctx.event_record(0)
    for _ in range(steps):
        f.run()
    ctx.event_record(1)
    ctx.sync()
    total = ctx.event_elapsed_ms(0, 1) / steps
    f.set_timing(True)          # the per-stage split in a pass of its own, the total above carries no stage events
    for _ in range(steps):
        f.run()
    st = {k: round(v / steps, 4) for k, v in f.stage_ms().items()}
    f.set_timing(False)
    sums = {}
    for nm, which in (("idx", mvs.BUF_IDX), ("map", mvs.BUF_MAP), ("state_a", mvs.BUF_STATE_A),
                      ("state_b", mvs.BUF_STATE_B), ("disp", mvs.BUF_DISP)):
        sums[nm] = hashlib.sha1(np.ascontiguousarray(f.download(which)).tobytes()).hexdigest()[:12]
    out = {"workload": name, "ms_per_frame": round(total, 4), "stage_ms": st, "sha1": sums,
           "defs": os.environ.get("MVS_NVCC_DEFS", ""), "mode": " ".join(x for x in (os.environ.get("MVS_PROP_MODE", ""), os.environ.get("MVS_BM_MODE", ""), os.environ.get("MVS_UPD_MODE", "")) if x)}
    if os.environ.get("MVS_STATS"):
        f.slic(); f.block_match(); f.flatness(); f.init_state()
        out["propagate_stats"] = [f.propagate_stats(i) for i in range(s.no_prop)]
    print(json.dumps(out))
    f.close(); ctx.close()


if __name__ == "__main__":
    main()
