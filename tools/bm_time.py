#!/usr/bin/env python
"""Device time of the block-matching kernel on a bench workload (CUDA events on the context stream)."""
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
    f.upload(mvs.BUF_RGBX, rgbx)
    f.set_timing(True)
    f.slic()
    f.block_match()
    ctx.sync(); f.stage_ms()
    for _ in range(reps):
        f.block_match()
    ms = f.stage_ms()
    print("block_match %.3f ms boundary %.3f ms" % (ms["block_match"] / reps,
                                                                   ms["boundary"] / reps))


if __name__ == "__main__":
    main()
