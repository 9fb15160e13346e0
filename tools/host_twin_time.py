#!/usr/bin/env python
"""Times the reference's CPU twins (the *_host methods of clPhotoConsistency / clDepthRefinement, compiled
unmodified from the reference sources: oracle/ref_host/ -> oracle/_ref/libhosttwin.so) on a whole frame of a
bench workload, on this machine's host cores.  Their inputs (Lab, labels, centres) come from the CPU oracle's
SLIC, since the reference has no CPU SLIC.  All twins are serial except propagate_host (OpenMP over views).
Prints one JSON line.   python tools/host_twin_time.py [workload] [bm_views]

bm_views limits do_initial_depth_estimation_host to the first views (it is serial: about 1.5 s per 1080p view);
the line reports both the measured views and -- labelled as such -- views * mean for the frame."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("OMP_NUM_THREADS", str(os.cpu_count() or 1))
import bench  # noqa: E402
from tests import hosttwin, orc  # noqa: E402


# the twins print progress lines to stdout ("Image 3 Is Done."): keep stdout for the one JSON line
sys.stdout.flush()
REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "15view_1080p"
    s, W, H, rgbx = bench.make_input(name)
    V = s.array_width * s.array_height
    bm_views = int(sys.argv[2]) if len(sys.argv) > 2 else V
    o = orc.load()
    h = hosttwin.load()
    if h is None:
        os.write(REAL_STDOUT, (json.dumps({"unavailable": "oracle/_ref/libhosttwin.so was never built (no /root/reference at build time)"}) + "\n").encode())
        return
    S = s.spixl_size
    labs, maps, idxs = [], [], []
    for v in range(V):
        l, m, i = o.slic(rgbx[v], S, s.slic_color_weight, s.no_iter, s.enforce_connectivity)
        labs.append(l); maps.append(m); idxs.append(i)
    lab, mp, idx = np.stack(labs), np.stack(maps), np.stack(idxs)
    levels = o.make_levels(s.min_disp, s.max_disp, s.inc)
    sub, num = o.make_subsets(s.array_width, s.array_height, s.neib_hor, s.neib_ver)
    mp_bm, rep, bm_s = h.block_match(lab, mp, idx, levels, sub, num, s.array_width, s.bl_ratio, S, 0, bm_views)
    if bm_views < V:   # the remaining views still need an initial disparity for the refinement twins
        rep_o = o.boundary(mp, idx, S)
        mp_o = o.block_match(lab, mp, rep_o, levels, sub, num, s.array_width, s.bl_ratio)
        mp_bm[bm_views:], rep[bm_views:] = mp_o[bm_views:], rep_o[bm_views:]
    g_, a_ = np.float32(2 * float(s.gamma) ** 2), np.float32(2 * float(s.alpha) ** 2)
    t = time.perf_counter()
    flat, sa, sb, T = h.refine(mp_bm, idx, rep, lab, s.array_width, s.array_height, S, s.bl_ratio, sub, num, g_, a_, s.fuse,
                               s.kernel_step, s.kernel_size // 2, s.no_prop)
    wall = time.perf_counter() - t
    # agreement with the kernels where the twins compute the same thing (they are not an oracle)
    ref_flat = o.flatness(mp_bm, np.float32(1.0 / float(g_)))
    out = {"workload": name, "cores": os.cpu_count(), "omp_threads": int(os.environ["OMP_NUM_THREADS"]),
           "twins": "reference host sources compiled unmodified over inert OpenCL/OpenCV headers (oracle/ref_host)",
           "block_match_host_s": bm_s, "block_match_views_timed": bm_views,
           "block_match_host_s_frame_estimate": bm_s * V / max(bm_views, 1),
           "flatness_host_s": T["flatness"], "init_state_host_s": T["init_state"], "propagate_host_s": T["propagate"],
           "refine_wall_s": wall,
           "note": "do_initial_depth_estimation_host / compute_flatness_host / init_spixl_state_host are serial as written; "
                   "propagate_host is '#pragma omp parallel for' over views (<= V threads busy) and skips the far-neighbour "
                   "and pair-plane stages of the kernel (SURVEY Appendix B H8)",
           "flatness_max_abs_diff_vs_kernel": float(np.abs(flat - ref_flat).max()),
           "state_finite": bool(np.isfinite(sa[..., :3]).all())}
    os.write(REAL_STDOUT, (json.dumps(out) + "\n").encode())


if __name__ == "__main__":
    main()
