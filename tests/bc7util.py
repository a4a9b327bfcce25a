"""TEST INFRASTRUCTURE: BC7 quality comparison helpers -- CUDA encoder vs the reference's bc7enc (oracle/_ref):
PSNR, the encoder's own perceptual error metric per block, mode histograms."""
import numpy as np

import hkbcn
import refshim

BC7 = 66


def to_blocks(img, w, h, ch=4):
    """(h, w, ch) -> (nblocks, 16, ch) in block raster order"""
    return img.reshape(h // 4, 4, w // 4, 4, ch).transpose(0, 2, 1, 3, 4).reshape(-1, 16, ch)


def perceptual_err(dec, src, alpha):
    """bc7enc.cpp:505-535 compute_color_distance_rgb[a] with weights 512/103/18/128, per pixel"""
    d = dec.astype(np.int64)
    s = src.astype(np.int64)

    def ycc(c):
        l = c[..., 0] * 109 + c[..., 1] * 366 + c[..., 2] * 37
        return l, (c[..., 0] << 9) - l, (c[..., 2] << 9) - l
    l1, cr1, cb1 = ycc(d)
    l2, cr2, cb2 = ycc(s)
    e = 512 * ((l1 - l2) >> 8) ** 2 + 103 * ((cr1 - cr2) >> 8) ** 2 + 18 * ((cb1 - cb2) >> 8) ** 2
    if alpha:
        e = e + 128 * (d[..., 3] - s[..., 3]) ** 2
    return e


def modes(blocks):
    b0 = blocks.reshape(-1, 16)[:, 0].astype(np.int32)
    m = np.full(b0.shape, 8)
    for k in range(7, -1, -1):
        m[(b0 & (1 << k)) != 0] = k
    return np.bincount(m, minlength=9)


def compare(img, w, h, name):
    src = img.reshape(h, w, 4)
    got = hkbcn.compress(BC7, img, w, h)
    ref = refshim.compress(BC7, img, w, h)
    dg = refshim.decompress(BC7, got, w, h).reshape(h, w, 4)
    dr = refshim.decompress(BC7, ref, w, h).reshape(h, w, 4)
    sb = to_blocks(src, w, h)
    gb, rb = to_blocks(dg, w, h), to_blocks(dr, w, h)
    has_alpha = (sb[..., 3] < 255).any(axis=1)
    out = {}
    for cls, sel in (("opaque", ~has_alpha), ("alpha", has_alpha)):
        if not sel.any():
            continue
        eg = np.where(sel[:, None], perceptual_err(gb, sb, False) + (128 * (gb[..., 3].astype(np.int64) - sb[..., 3]) ** 2 if cls == "alpha" else 0), 0).sum(1)[sel]
        er = np.where(sel[:, None], perceptual_err(rb, sb, False) + (128 * (rb[..., 3].astype(np.int64) - sb[..., 3]) ** 2 if cls == "alpha" else 0), 0).sum(1)[sel]
        nc = 4 if cls == "alpha" else 3
        mg = ((gb[sel][..., :nc].astype(np.float64) - sb[sel][..., :nc]) ** 2).mean()
        mr = ((rb[sel][..., :nc].astype(np.float64) - sb[sel][..., :nc]) ** 2).mean()
        pg = 10 * np.log10(255 ** 2 / max(mg, 1e-12))
        pr = 10 * np.log10(255 ** 2 / max(mr, 1e-12))
        out[cls] = dict(blocks=int(sel.sum()), psnr_gpu=pg, psnr_ref=pr, perr_gpu=float(eg.mean()), perr_ref=float(er.mean()),
                        worse=float((eg > er).mean()), better=float((eg < er).mean()), same=float((eg == er).mean()),
                        identical=float((got.reshape(-1, 16)[sel] == ref.reshape(-1, 16)[sel]).all(1).mean()))
        print(f"{name:>14} {cls:>6}: n={sel.sum():6d} PSNR gpu {pg:7.3f} ref {pr:7.3f} (d {pg - pr:+.3f} dB) | perceptual err/block gpu {eg.mean():10.1f} "
              f"ref {er.mean():10.1f} (ratio {eg.mean() / max(er.mean(), 1e-9):.4f}) | worse {100 * (eg > er).mean():5.1f}% better {100 * (eg < er).mean():5.1f}% "
              f"same {100 * (eg == er).mean():5.1f}% identical-bytes {100 * (got.reshape(-1, 16)[sel] == ref.reshape(-1, 16)[sel]).all(1).mean():5.1f}%")
    print(f"{'':>14} modes gpu {modes(got).tolist()} ref {modes(ref).tolist()}")
    return out
