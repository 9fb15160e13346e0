"""The reference's bundled camera-array data as frames of the pipeline's input format.

read_image_array (file_handler.cpp:30-57) reads one image path per line of a list file and
loadImageIn (file_handler.cpp:6-15) repacks cv::Mat BGR to vec3u {s0=R, s1=G, s2=B, pad}.  The list
files are clMVDE/backup/data-bar.txt:1-15 (the 15-view set c0f1..c14f1, a 5 x 3 array,
photo_consistency.cpp:325) and clMVDE/data.txt:1-9 (Beer-Garden, the 3 x 3 array clMVDE.cpp:17-18
is configured for).  The images live under tests/data/_ref/ (copied there from the reference tree by
tests/data/fetch_bundled.py; git-ignored) and are decoded here with OpenCV, the reference's decoder.
"""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DATA = os.path.join(ROOT, "tests", "data", "_ref")

SETS = {
    # name: (file names in list-file order, array width, array height)
    "bar15": ([f"c{k}f1.png" for k in range(15)], 5, 3),
    "beer9": ([f"Beer-Garden/img{k}.png" for k in range(9)], 3, 3),
}


def available(name="bar15"):
    names = SETS[name][0]
    return all(os.path.exists(os.path.join(DATA, n)) for n in names)


def load_bgr(path):
    import cv2
    bgr = cv2.imread(path, cv2.IMREAD_COLOR)
    if bgr is None:
        raise FileNotFoundError(path)
    return bgr


def bgr_to_rgbx(bgr):
    """loadImageIn, file_handler.cpp:6-15"""
    out = np.zeros(bgr.shape[:2] + (4,), np.uint8)
    out[..., 0], out[..., 1], out[..., 2] = bgr[..., 2], bgr[..., 1], bgr[..., 0]
    return out


def load_set(name, crop=None):
    """-> (rgbx uint8 [V, H, W, 4], array width, array height); crop = (x0, y0, w, h) of every view"""
    names, aw, ah = SETS[name]
    views = []
    for n in names:
        bgr = load_bgr(os.path.join(DATA, n))
        if crop is not None:
            x0, y0, w, h = crop
            bgr = bgr[y0:y0 + h, x0:x0 + w]
        views.append(bgr_to_rgbx(bgr))
    return np.stack(views), aw, ah


def load_bgr_set(name):
    """the packed BGR rasters themselves [V, H, W, 3] (input of the device-side loadImageIn)"""
    names, aw, ah = SETS[name]
    return np.stack([load_bgr(os.path.join(DATA, n)) for n in names]), aw, ah
