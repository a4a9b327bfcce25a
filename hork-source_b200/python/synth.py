"""Deterministic synthetic inputs for the BASELINE.json configs.

Everything derives from a counter-based hash (splitmix64 of (seed, x, y, channel, octave)),
so the images are reproducible bit-for-bit anywhere numpy runs.  Test / bench glue only.
"""
import numpy as np

MASK = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & MASK
    z = x
    z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & MASK
    z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & MASK
    return z ^ (z >> np.uint64(31))


def hash01(seed, x, y, c=0, o=0):
    """uniform [0,1) float64 from integer lattice coordinates (arrays)."""
    with np.errstate(over="ignore"):
        k = (np.uint64(seed) * np.uint64(0x100000001B3)
             + x.astype(np.uint64) * np.uint64(0x9E3779B1)
             + y.astype(np.uint64) * np.uint64(0x85EBCA77)
             + np.uint64(c) * np.uint64(0xC2B2AE3D)
             + np.uint64(o) * np.uint64(0x27D4EB2F165667C5)) & MASK
        return (splitmix64(k) >> np.uint64(11)).astype(np.float64) * (1.0 / (1 << 53))


def value_noise(w, h, seed, c, octaves, base_cell=64, wrap=True):
    """fractal value noise in [0,1], periodic when wrap (so mip wrap filtering is seamless).
    The hash is evaluated on the cell lattice only and gathered per pixel."""
    ys, xs = np.arange(h), np.arange(w)
    acc = np.zeros((h, w), np.float64)
    amp, tot = 1.0, 0.0
    for o in range(octaves):
        cell = max(base_cell >> o, 1)
        gx, gy = xs // cell, ys // cell
        fx, fy = (xs % cell) / cell, (ys % cell) / cell
        fx, fy = fx * fx * (3 - 2 * fx), fy * fy * (3 - 2 * fy)
        nx, ny = max(w // cell, 1), max(h // cell, 1)
        x1, y1 = ((gx + 1) % nx, (gy + 1) % ny) if wrap else (gx + 1, gy + 1)
        ly, lx = np.mgrid[0:int(max(gy.max(), y1.max())) + 1, 0:int(max(gx.max(), x1.max())) + 1]
        lat = hash01(seed, lx, ly, c, o)
        fx, fy = fx[None, :], fy[:, None]
        v00, v10 = lat[np.ix_(gy, gx)], lat[np.ix_(gy, x1)]
        v01, v11 = lat[np.ix_(y1, gx)], lat[np.ix_(y1, x1)]
        acc += amp * ((v00 * (1 - fx) + v10 * fx) * (1 - fy) + (v01 * (1 - fx) + v11 * fx) * fy)
        tot += amp
        amp *= 0.5
    return acc / tot


def albedo_rgba8(w, h, seed, alpha="opaque"):
    """Config 1/3: correlated luma + chroma noise, ~8% hard-edged Voronoi-ish cells, ~5% flat cells.
    alpha: 'opaque' (255) | 'mask' (smooth mask, ~35% of blocks mixed) | 'noise'."""
    luma = value_noise(w, h, seed, 0, 5)
    img = np.empty((h, w, 4), np.float64)
    for c in range(3):
        img[..., c] = np.clip(0.65 * luma + 0.35 * value_noise(w, h, seed, 1 + c, 5, 32) + 0.04 * (hash01(seed, *np.mgrid[0:h, 0:w][::-1], 7 + c) - 0.5), 0, 1)
    yy, xx = np.mgrid[0:h, 0:w]
    ncy, ncx = (h + 15) // 16, (w + 15) // 16
    lcy, lcx = np.mgrid[0:ncy, 0:ncx]
    iy, ix = np.arange(h) // 16, np.arange(w) // 16

    def cell(ch):  # hash of the 16x16 cell each pixel lies in
        return hash01(seed, lcx, lcy, ch)[np.ix_(iy, ix)]
    sel = cell(11)
    flat = sel < 0.05
    edge = (sel >= 0.05) & (sel < 0.13)
    half = ((xx % 16) + (yy % 16) * (cell(15) - 0.5) * 2) > 8
    for c in range(3):
        cellcol = cell(12 + c)
        img[..., c] = np.where(flat, cellcol, img[..., c])
        img[..., c] = np.where(edge & half, 1.0 - cellcol, img[..., c])
    if alpha == "opaque":
        img[..., 3] = 1.0
    elif alpha == "mask":
        m = value_noise(w, h, seed, 20, 3, 128)
        img[..., 3] = np.clip((m - 0.42) * 6.0, 0, 1)
    else:
        img[..., 3] = hash01(seed, xx, yy, 21)
    return np.round(img * 255.0).astype(np.uint8)


def roughness_r8(w, h, seed):
    """Config 2: 4-octave noise mapped to [16, 240]."""
    n = value_noise(w, h, seed, 0, 4)
    return np.round(16 + n * 224).astype(np.uint8)


def normals_rg8(w, h, seed):
    """Config 2: 6-octave height field -> central-difference tangent normals -> round(sat(n*0.5+0.5)*255) (RG)."""
    hgt = value_noise(w, h, seed, 0, 6) * 24.0
    dx = (np.roll(hgt, -1, 1) - np.roll(hgt, 1, 1)) * 0.5
    dy = (np.roll(hgt, -1, 0) - np.roll(hgt, 1, 0)) * 0.5
    inv = 1.0 / np.sqrt(dx * dx + dy * dy + 1.0)
    nx, ny = -dx * inv, -dy * inv
    out = np.empty((h, w, 2), np.uint8)
    out[..., 0] = np.floor(np.clip(nx * 0.5 + 0.5, 0, 1) * 255.0 + 0.5).astype(np.uint8)
    out[..., 1] = np.floor(np.clip(ny * 0.5 + 0.5, 0, 1) * 255.0 + 0.5).astype(np.uint8)
    return out


def sky_rgba32f(w, h, seed, face=0):
    """Config 4: vertical gradient 0.05..4.0, cloud noise, a sun disc peaking at 5e4 on face 0, alpha 1."""
    yy, xx = np.mgrid[0:h, 0:w]
    t = yy / max(h - 1, 1)
    base = 0.05 + (4.0 - 0.05) * (1 - t) ** 2
    cloud = value_noise(w, h, seed + face, 0, 5, 128)
    img = np.empty((h, w, 4), np.float32)
    img[..., 0] = base * (0.55 + 0.9 * cloud)
    img[..., 1] = base * (0.70 + 0.8 * cloud)
    img[..., 2] = base * (1.00 + 0.6 * cloud)
    if face == 0:
        d2 = ((xx - 0.7 * w) ** 2 + (yy - 0.3 * h) ** 2) / (0.03 * w) ** 2
        sun = 5e4 * np.exp(-d2 * 3.0)
        img[..., 0] += sun
        img[..., 1] += sun * 0.9
        img[..., 2] += sun * 0.7
    img[..., 3] = 1.0
    return img


def random_blocks(n, bpp, seed):
    """n packed 4x4 blocks of uniform random bytes."""
    idx = np.arange(n * 16 * bpp, dtype=np.uint64)
    with np.errstate(over="ignore"):
        v = splitmix64(idx + np.uint64(seed) * np.uint64(0x9E3779B97F4A7C15))
    return (v >> np.uint64(56)).astype(np.uint8)


def mip_dims(w, h, mips):
    return [(max(w >> i, 4), max(h >> i, 4)) for i in range(mips)]
