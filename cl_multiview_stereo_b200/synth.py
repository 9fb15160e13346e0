"""Synthetic camera-array frames: a piecewise-planar scene
rendered into every view of an aw x ah camera grid with the reference's reprojection convention
x_v = x_ref - d*(cx_v - cx_ref), y_v = y_ref - bl_ratio*d*(cy_v - cy_ref) (clcode.cl:1034),
plus ground-truth disparity per view.  Pure numpy, deterministic for a given seed.
"""
import numpy as np


def _texture(rng, H, W, pad):
    """random blocky + fine-noise RGB texture, uint8 [H+2pad, W+2pad, 3]"""
    h, w = H + 2 * pad, W + 2 * pad
    out = np.zeros((h, w, 3), np.float32)
    for cell, amp in ((32, 60.0), (8, 40.0), (3, 25.0)):
        ch, cw = (h + cell - 1) // cell + 1, (w + cell - 1) // cell + 1
        g = rng.uniform(-amp, amp, size=(ch, cw, 3)).astype(np.float32)
        out += np.repeat(np.repeat(g, cell, axis=0), cell, axis=1)[:h, :w]
    out += rng.uniform(-6, 6, size=(h, w, 3)).astype(np.float32)
    base = rng.uniform(70, 190, size=(1, 1, 3)).astype(np.float32)
    return np.clip(out + base, 0, 255).astype(np.uint8)


def make_scene(seed, W, H, n_patches=12, dmin=32.0, dmax=58.0):
    """planar layers in the reference (grid-origin) view: d(x,y) = a*x + b*y + c inside a rectangle"""
    rng = np.random.default_rng(seed)
    pad = 96
    layers = []
    # slanted background covering everything
    a, b = rng.uniform(-1.5, 1.5) / W, rng.uniform(-1.5, 1.5) / H
    c0 = dmin + 2.0
    layers.append(dict(a=a, b=b, c=c0, rect=(-10 * W, -10 * H, 10 * W, 10 * H), tex=_texture(rng, H, W, pad)))
    ds = np.sort(rng.uniform(dmin + 4.0, dmax, size=n_patches))
    for k in range(n_patches):
        pw, ph = rng.uniform(0.12, 0.35) * W, rng.uniform(0.12, 0.35) * H
        x0, y0 = rng.uniform(-0.05 * W, W - 0.5 * pw), rng.uniform(-0.05 * H, H - 0.5 * ph)
        a, b = rng.uniform(-4.0, 4.0) / W, rng.uniform(-4.0, 4.0) / H
        c = ds[k] - a * (x0 + pw / 2) - b * (y0 + ph / 2)
        layers.append(dict(a=a, b=b, c=c, rect=(x0, y0, x0 + pw, y0 + ph), tex=_texture(rng, H, W, pad)))
    return dict(W=W, H=H, pad=pad, layers=layers)


def render_view(scene, gx, gy, bl_ratio=1.0):
    """view at grid position (gx, gy) -> (rgbx uint8 [H,W,4] with s0=R,s1=G,s2=B, gt disparity float32 [H,W])"""
    W, H, pad = scene["W"], scene["H"], scene["pad"]
    rgb = np.zeros((H, W, 3), np.uint8)
    gt = np.zeros((H, W), np.float32)
    for L in scene["layers"]:   # painter's order: later layers are nearer
        a, b, c = np.float32(L["a"]), np.float32(L["b"]), np.float32(L["c"])
        den = np.float32(1.0) - a * np.float32(gx) - b * np.float32(bl_ratio * gy)
        x0, y0, x1, y1 = L["rect"]
        # only the part of the view the layer can reach: its disparity is planar, so the extremes over the
        # frame sit at the corners (a conservative box; the mask below decides pixel by pixel)
        dc = [(float(a) * cx + float(b) * cy + float(c)) / float(den) for cx in (0, W - 1) for cy in (0, H - 1)]
        dlo, dhi = min(dc), max(dc)
        sx, sy = [dlo * gx, dhi * gx], [dlo * bl_ratio * gy, dhi * bl_ratio * gy]
        bx0 = int(max(0, np.floor(x0 - max(sx)) - 2)); bx1 = int(min(W, np.ceil(x1 - min(sx)) + 3))
        by0 = int(max(0, np.floor(y0 - max(sy)) - 2)); by1 = int(min(H, np.ceil(y1 - min(sy)) + 3))
        if bx1 <= bx0 or by1 <= by0:
            continue
        ys, xs = np.mgrid[by0:by1, bx0:bx1].astype(np.float32)
        d = (a * xs + b * ys + c) / den
        xr = xs + d * np.float32(gx)
        yr = ys + d * np.float32(bl_ratio * gy)
        m = (xr >= x0) & (xr < x1) & (yr >= y0) & (yr < y1)
        xi = np.clip(np.rint(xr).astype(np.int32) + pad, 0, W + 2 * pad - 1)
        yi = np.clip(np.rint(yr).astype(np.int32) + pad, 0, H + 2 * pad - 1)
        col = L["tex"][yi, xi]
        rgb[by0:by1, bx0:bx1][m] = col[m]
        gt[by0:by1, bx0:bx1][m] = d[m]
    rgbx = np.zeros((H, W, 4), np.uint8)
    rgbx[..., :3] = rgb
    return rgbx, gt


def make_frame(seed, W, H, aw, ah, bl_ratio=1.0, n_patches=12, dmin=32.0, dmax=58.0):
    """all views of an aw x ah array -> rgbx [V,H,W,4], gt [V,H,W]; view index = gy*aw + gx"""
    scene = make_scene(seed, W, H, n_patches, dmin, dmax)
    V = aw * ah
    rgbx = np.zeros((V, H, W, 4), np.uint8)
    gt = np.zeros((V, H, W), np.float32)
    for v in range(V):
        rgbx[v], gt[v] = render_view(scene, v % aw, v // aw, bl_ratio)
    return rgbx, gt


def voronoi_image(seed, W, H, sites=4096):
    """SLIC-only workload (config C3): blocky mosaic + noise, uint8 [H,W,4]"""
    rng = np.random.default_rng(seed)
    cell = max(8, int(np.sqrt(W * H / sites)))
    ch, cw = (H + cell - 1) // cell + 1, (W + cell - 1) // cell + 1
    g = rng.integers(0, 256, size=(ch, cw, 3), dtype=np.int32)
    img = np.repeat(np.repeat(g, cell, axis=0), cell, axis=1)
    oy, ox = rng.integers(0, cell, 2)
    img = img[oy:oy + H, ox:ox + W] + rng.integers(-8, 9, size=(H, W, 3))
    out = np.zeros((H, W, 4), np.uint8)
    out[..., :3] = np.clip(img, 0, 255)
    return out
