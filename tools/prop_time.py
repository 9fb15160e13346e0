#!/usr/bin/env python
"""Per-iteration device time of the propagate kernel on the bench workload (CUDA events on the
context stream).  Usage: MVS_PROP_MINB=5 python tools/prop_time.py [workload] [reps]"""
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
    ctx.sync()
    tot = [0.0] * s.no_prop
    for _ in range(reps):
        f.flatness(); f.init_state()
        for it in range(s.no_prop):
            ctx.event_record(0)
            f.propagate(it)
            ctx.event_record(1)
            tot[it] += ctx.event_elapsed_ms(0, 1)
    per = [t / reps for t in tot]
    print("MINB=%s per-iteration ms: %s  total %.3f" % (os.environ.get("MVS_PROP_MINB", "default"),
                                                        " ".join("%.3f" % p for p in per), sum(per)))


if __name__ == "__main__":
    main()
