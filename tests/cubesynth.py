"""TEST INFRASTRUCTURE: a synthetic HDR sky as six RGBA32F cube faces (face frames as in DESIGN.md 8d)."""
import numpy as np


def face_dirs(n):
    c = (np.arange(n, dtype=np.float32) + 0.5) * (2.0 / n) - 1.0
    u, v = np.meshgrid(c, c)  # u along x (columns), v along y (rows)
    one = np.ones_like(u)
    d = [np.stack(a, -1) for a in ((one, -v, -u), (-one, -v, u), (u, one, v), (u, -one, -v), (u, -v, one), (-u, -v, -one))]
    d = np.stack(d, 0).astype(np.float32)
    return d / np.linalg.norm(d, axis=-1, keepdims=True)


def sky_cube(n, seed=0, sun=40.0):
    d = face_dirs(n)
    rng = np.random.default_rng(seed)
    sun_dir = np.array([0.4, 0.7, 0.6], np.float32)
    sun_dir /= np.linalg.norm(sun_dir)
    up = np.clip(d[..., 1] * 0.5 + 0.5, 0, 1)
    lobe = np.exp(-80.0 * (1.0 - d @ sun_dir)).astype(np.float32) * sun
    img = np.zeros((6, n, n, 4), np.float32)
    img[..., 0] = 0.15 + 0.35 * up + lobe
    img[..., 1] = 0.25 + 0.45 * up + 0.9 * lobe
    img[..., 2] = 0.35 + 0.60 * up + 0.6 * lobe
    img[..., :3] += rng.random((6, n, n, 3), dtype=np.float32) * 0.02
    img[..., 3] = 1.0
    return img
