#!/usr/bin/env python
"""Device time of the multi-view fusion stage (mvs_frame_view_fusion) on a bench workload."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import cl_multiview_stereo_b200 as mvs  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "15view_1080p"
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    s, W, H, rgbx = bench.make_input(name)
    ctx = mvs.Context(0)
    f = mvs.Frame(ctx, s, W, H)
    f.set_timing(False)
    f.upload(mvs.BUF_RGBX, rgbx)
    f.run()
    f.view_fusion()
    ctx.sync()
    tot = 0.0
    for _ in range(reps):
        ctx.event_record(0)
        f.view_fusion()
        ctx.event_record(1)
        tot += ctx.event_elapsed_ms(0, 1)
    import numpy as np
    d, o = f.download(mvs.BUF_DISP), f.download(mvs.BUF_DISP_FUSED)
    print("%s: view fusion %.3f ms per frame; fused==0 on %.2f%% of pixels; |fused-disp|<=0.5 on %.2f%%" % (
        name, tot / reps, 100 * float((o == 0).mean()), 100 * float((np.abs(o - d) <= 0.5).mean())))


if __name__ == "__main__":
    main()
