#!/usr/bin/env python
"""Generates the golden vectors under tests/golden/ by running the REFERENCE'S OWN device source
(/root/reference/clMVDE/clcode.cl compiled as C++, oracle/ref/ -> oracle/_ref/libclref.so)
through the reference's host sequence (clSLIC.cpp:67-122, photo_consistency.cpp:21-150,
depth_refinement.cpp:91-118, pipeline.cpp:108-175).  Runs only where /root/reference exists
(this container); the .npz files it writes are committed and are what travels.

  python tests/golden/make_golden.py            (from the repository root)

Cases
  synth_3x2      synthetic piecewise-planar 3x2 array, 96x64, small disparity range
  bar_5x3_crop   160x96 crop of the bundled 15-view set Images/c0f1..c14f1.png (view order of
                 backup/data-bar.txt), shipped literals of clMVDE.cpp:14-36
  fusion_*       multi-view fusion (project_to_reference_inv + remove_view_inconsistency in the host
                 sequence of depth_refinement.cpp:1382-1451, disp_proj zero-filled) on the disparity
                 images of the two pipeline cases above
  slic_*         SLIC only: crops of Images/dance.jpg / horse.jpg, odd sizes, S=8 with
                 connectivity, S=16
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import clref, orc  # noqa: E402
import cl_multiview_stereo_b200 as mvs  # noqa: E402
from cl_multiview_stereo_b200 import synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
REF_IMG = "/root/reference/Images"


def load_rgbx(path, x0, y0, w, h):
    """cv2 decodes to BGR; loadImageIn (file_handler.cpp:6-15) stores s0=R, s1=G, s2=B"""
    import cv2
    bgr = cv2.imread(path, cv2.IMREAD_COLOR)
    assert bgr is not None, path
    crop = bgr[y0:y0 + h, x0:x0 + w]
    out = np.zeros((h, w, 4), np.uint8)
    out[..., 0], out[..., 1], out[..., 2] = crop[..., 2], crop[..., 1], crop[..., 0]
    return out


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def run_pipeline(r, o, rgbx, s):
    """the reference's stage sequence, every kernel from libclref"""
    V, H, W, _ = rgbx.shape
    S = s.spixl_size
    mxy, mcol = o.slic_params(S)                        # clSLIC.cpp:15-18 (host scalars)
    labs, maps, idxs = [], [], []
    for v in range(V):
        l, m, i = r.slic(rgbx[v], S, s.slic_color_weight, s.no_iter, s.enforce_connectivity, mxy, mcol)
        labs.append(l); maps.append(m); idxs.append(i)
    lab, mp, idx = np.stack(labs), np.stack(maps), np.stack(idxs)
    out = dict(idx=idx, map_slic=mp.copy(), lab_sha256=np.array(sha(lab)))
    levels = o.make_levels(s.min_disp, s.max_disp, s.inc)                 # pipeline.cpp:121-124
    sub, num = o.make_subsets(s.array_width, s.array_height, s.neib_hor, s.neib_ver)   # :130-142
    rep = r.boundary(mp, idx, S)
    mp = r.block_match(lab, mp, rep, idx, levels, sub, num, s.array_width, s.bl_ratio, S)
    out.update(rep=rep, map=mp)
    g_ = np.float32(2 * float(s.gamma) ** 2)                              # pipeline.cpp:164-166
    a_ = np.float32(2 * float(s.alpha) ** 2)
    gam, alp, fus, kss = o.refine_params(g_, a_, s.fuse, s.kernel_step, s.kernel_size // 2, S)
    flat = r.flatness(mp, np.float32(1.0 / float(g_)))                    # depth_refinement.cpp:130
    st = [r.init_state(mp, idx, rep, flat, sub, num, s.array_width, s.bl_ratio, gam, alp, fus, s.kernel_step, kss), None]
    st[1] = np.zeros_like(st[0])
    for it in range(s.no_prop):                                           # ping-pong, :744-753
        src, dst = it % 2, 1 - it % 2
        st[dst] = r.propagate(st[src], mp, idx, rep, flat, sub, num, it, s.array_width, s.bl_ratio, gam, alp, fus,
                              s.kernel_step // (it + 1), np.float32(kss) / np.float32(it + 1))
    disp = r.spixl_to_image(mp, idx, st[0])                               # fusion reads buffer A, :1352
    out.update(flat=flat, state_a=st[0], state_b=st[1], disp=disp)
    return out


def settings_dict(s):
    return {k: getattr(s, k) for k in s.__dataclass_fields__}


def make_fusion(r):
    """from the committed pipeline goldens, so that it can be regenerated on its own"""
    import json
    for src in ("synth_3x2", "bar_5x3_crop"):
        g = np.load(os.path.join(OUT, src + ".npz"))
        st = json.loads(str(g["settings"]))
        disp = g["disp"]
        assert np.isfinite(disp).all()   # NaN would reproject through an undefined float->int conversion
        proj, fused = r.view_fusion(disp, st["array_width"], st["bl_ratio"], st["fuse"])
        np.savez_compressed(os.path.join(OUT, "fusion_" + src + ".npz"), disp=disp, array_width=st["array_width"],
                            bl_ratio=st["bl_ratio"], fuse=st["fuse"], proj=proj, fused=fused)
        print("fusion_" + src, "done", flush=True)


def main():
    r, o = clref.load(), orc.load()
    assert r is not None, "needs /root/reference (libclref.so)"
    import json
    if len(sys.argv) > 1 and sys.argv[1] == "fusion":
        make_fusion(r)
        return

    # 1. synthetic 3x2
    s = mvs.SystemSettings(array_width=3, array_height=2, min_disp=0, max_disp=16, inc=1, bl_ratio=1.0, kernel_size=128)
    rgbx, _ = synth.make_frame(7, 96, 64, 3, 2, bl_ratio=1.0, n_patches=4, dmin=4, dmax=14)
    np.savez_compressed(os.path.join(OUT, "synth_3x2.npz"), rgbx=rgbx, settings=json.dumps(settings_dict(s)),
                        **run_pipeline(r, o, rgbx, s))
    print("synth_3x2 done", flush=True)

    # 2. bundled 15-view set, crop, shipped literals (array 5x3)
    s = mvs.SystemSettings(array_width=5, array_height=3)
    x0, y0, w, h = 880, 500, 160, 96
    rgbx = np.stack([load_rgbx(f"{REF_IMG}/c{k}f1.png", x0, y0, w, h) for k in range(15)])
    np.savez_compressed(os.path.join(OUT, "bar_5x3_crop.npz"), rgbx=rgbx, settings=json.dumps(settings_dict(s)),
                        crop=np.array([x0, y0, w, h]), **run_pipeline(r, o, rgbx, s))
    print("bar_5x3_crop done", flush=True)

    # 3. SLIC only, odd sizes
    for name, path, (x0, y0, w, h), S, conn in [
        ("slic_dance_s8_conn", f"{REF_IMG}/dance.jpg", (300, 200, 141, 94), 8, True),
        ("slic_horse_s16", f"{REF_IMG}/horse.jpg", (400, 250, 125, 75), 16, False),
        ("slic_horse_s8", f"{REF_IMG}/horse.jpg", (500, 300, 37, 29), 8, False),
    ]:
        img = load_rgbx(path, x0, y0, w, h)
        mxy, mcol = o.slic_params(S)
        lab, mp, idx = r.slic(img, S, 0.6, 5, conn, mxy, mcol)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), rgbx=img, S=S, connectivity=conn, weight=0.6, no_iter=5,
                            lab=lab, map=mp, idx=idx)
        print(name, "done", flush=True)
    make_fusion(r)


if __name__ == "__main__":
    main()
