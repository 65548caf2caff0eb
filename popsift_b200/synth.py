"""Seeded synthetic grayscale test images.

Blobs (scale-space extrema at many sigmas) + multi-scale value noise + dark/bright
rectangles (edges exercise the edge-rejection test), clamped and rounded to u8.
The generator is shared by the parity tests, the golden-fixture script and bench.py,
so that the reference build and the new build always see byte-identical inputs
(the bytes are written to a file that both read).
"""
import numpy as np


def _value_noise(rng, w, h, cell, amp):
    gw, gh = w // cell + 2, h // cell + 2
    g = rng.uniform(-1.0, 1.0, size=(gh, gw)).astype(np.float32)
    ys = np.arange(h, dtype=np.float32) / cell
    xs = np.arange(w, dtype=np.float32) / cell
    y0 = ys.astype(np.int32); x0 = xs.astype(np.int32)
    fy = (ys - y0)[:, None]; fx = (xs - x0)[None, :]
    a = g[y0][:, x0]; b = g[y0][:, x0 + 1]
    c = g[y0 + 1][:, x0]; d = g[y0 + 1][:, x0 + 1]
    return amp * ((a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy)


def gen(w, h, seed):
    """Return an (h, w) uint8 image, deterministic in (w, h, seed)."""
    rng = np.random.default_rng(seed)
    img = np.full((h, w), 128.0, dtype=np.float32)
    nblob = max(4, (w * h) // 600)
    cx = rng.uniform(0, w, nblob); cy = rng.uniform(0, h, nblob)
    sg = np.exp(rng.uniform(np.log(1.5), np.log(24.0), nblob))
    am = rng.uniform(20.0, 90.0, nblob) * rng.choice([-1.0, 1.0], nblob)
    for i in range(nblob):
        r = int(np.ceil(4.0 * sg[i]))
        x0 = max(0, int(cx[i]) - r); x1 = min(w, int(cx[i]) + r + 1)
        y0 = max(0, int(cy[i]) - r); y1 = min(h, int(cy[i]) + r + 1)
        if x0 >= x1 or y0 >= y1:
            continue
        gx = np.exp(-0.5 * ((np.arange(x0, x1) - cx[i]) / sg[i]) ** 2).astype(np.float32)
        gy = np.exp(-0.5 * ((np.arange(y0, y1) - cy[i]) / sg[i]) ** 2).astype(np.float32)
        img[y0:y1, x0:x1] += np.float32(am[i]) * gy[:, None] * gx[None, :]
    for cell, amp in ((64, 24.0), (16, 12.0), (4, 6.0)):
        img += _value_noise(rng, w, h, cell, amp)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    for k in range(12):
        rcx = rng.uniform(0, w); rcy = rng.uniform(0, h)
        hw = rng.uniform(0.02, 0.15) * w; hh = rng.uniform(0.02, 0.15) * h
        ang = 0.0 if k % 2 == 0 else rng.uniform(0, np.pi)
        val = rng.uniform(30.0, 70.0) * (1.0 if rng.uniform() < 0.5 else -1.0)
        ca, sa = np.cos(ang), np.sin(ang)
        # bounding box of the rotated rectangle keeps this cheap on large images
        ext = int(np.ceil(np.hypot(hw, hh))) + 1
        x0 = max(0, int(rcx) - ext); x1 = min(w, int(rcx) + ext + 1)
        y0 = max(0, int(rcy) - ext); y1 = min(h, int(rcy) + ext + 1)
        if x0 >= x1 or y0 >= y1:
            continue
        dx = xx[y0:y1, x0:x1] - rcx; dy = yy[y0:y1, x0:x1] - rcy
        u = dx * ca + dy * sa; v = -dx * sa + dy * ca
        img[y0:y1, x0:x1] += np.where((np.abs(u) <= hw) & (np.abs(v) <= hh), np.float32(val), np.float32(0))
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def gen_tiled(w, h, seed, tile=2048):
    """A giant image (config 5: 16384 x 16384) as a mosaic of independent gen() tiles: the same kind of content at a
    cost linear in the area (gen() itself draws every blob in Python and holds two float coordinate planes)."""
    img = np.empty((h, w), np.uint8)
    k = 0
    for y in range(0, h, tile):
        for x in range(0, w, tile):
            th, tw = min(tile, h - y), min(tile, w - x)
            img[y:y + th, x:x + tw] = gen(tw, th, seed * 1000 + k)
            k += 1
    return img


def constant(w, h, value=128):
    return np.full((h, w), value, dtype=np.uint8)


def single_blob(w, h, cx, cy, sigma, amp=100.0, base=60.0):
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    img = base + amp * np.exp(-0.5 * (((xx - cx) / sigma) ** 2 + ((yy - cy) / sigma) ** 2))
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def checkerboard(w, h, cell=8, lo=40, hi=215):
    yy, xx = np.mgrid[0:h, 0:w]
    return np.where(((xx // cell) + (yy // cell)) % 2 == 0, lo, hi).astype(np.uint8)
