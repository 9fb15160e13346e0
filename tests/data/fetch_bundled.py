#!/usr/bin/env python
"""TEST / BENCH DATA.  Copies the reference's bundled camera-array images -- the 15-view set
Images/c0f1..c14f1.png (view order: clMVDE/backup/data-bar.txt:1-15), the 3x3 Beer-Garden set
(clMVDE/data.txt:1-9) and dance.jpg / horse.jpg -- from /root/reference/Images, where they lie, into
tests/data/_ref/ (git-ignored: the files never enter the history; NOT gpurun-ignored: they travel to
the GPU box like the built libraries, where /root/reference does not exist).

  python tests/data/fetch_bundled.py        (run by __graft_entry__.build() when /root/reference exists)
"""
import os
import shutil
import sys

SRC = "/root/reference/Images"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def fetch(verbose=False):
    if not os.path.isdir(SRC):
        return False
    names = [f"c{k}f1.png" for k in range(15)] + [f"Beer-Garden/img{k}.png" for k in range(9)] + ["dance.jpg", "horse.jpg"]
    for n in names:
        s, d = os.path.join(SRC, n), os.path.join(DST, n)
        if not os.path.exists(s):
            raise FileNotFoundError(s)
        if os.path.exists(d) and os.path.getsize(d) == os.path.getsize(s):
            continue
        os.makedirs(os.path.dirname(d), exist_ok=True)
        shutil.copyfile(s, d)
        if verbose:
            print("copied", n)
    return True


if __name__ == "__main__":
    ok = fetch(verbose=True)
    print("bundled images:", "ready" if ok else "no /root/reference here", file=sys.stderr)
