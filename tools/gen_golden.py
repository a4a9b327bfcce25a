#!/usr/bin/env python3
"""Generate tests/golden/bcn_golden.npz by running the reference itself (oracle/_ref/libhkref.so,
the unmodified rgbcx / bcdec / bc7decomp / bc7enc / bc6h_enc sources) on seeded inputs.
Run in the build container (needs /root/reference to have built oracle/_ref); the fixture travels."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "hork-source_b200", "python"))
import refshim  # noqa: E402
import synth  # noqa: E402
from test_gpu_parity import adversarial_r8, adversarial_rgba8, force_modes_bc6h, force_modes_bc7  # noqa: E402

BC1, BC2, BC3, BC4, BC5, BC6U, BC6S, BC7 = 54, 56, 58, 60, 62, 64, 65, 66
out = {}
rgba = np.concatenate([adversarial_rgba8()[: 64 * 700], synth.albedo_rgba8(64, 32, 0xB2000001, "mask").reshape(8, 4, 16, 4, 4).transpose(0, 2, 1, 3, 4).ravel()])
r8 = adversarial_r8()[: 16 * 900]
rg8 = synth.random_blocks(300, 2, 11)
out["rgba"], out["r8"], out["rg8"] = rgba, r8, rg8
for fmt, key in ((BC1, "bc1"), (BC2, "bc2"), (BC3, "bc3")):
    out[key + "_enc"] = refshim.compress(fmt, rgba, 4, rgba.size // 16, 1)
out["bc4_enc"] = refshim.compress(BC4, r8, 4, r8.size // 4, 1)
out["bc5_enc"] = refshim.compress(BC5, rg8, 4, rg8.size // 8, 1)
for fmt, key, blk in ((BC1, "bc1", 8), (BC2, "bc2", 16), (BC3, "bc3", 16), (BC4, "bc4", 8), (BC5, "bc5", 16), (BC7, "bc7", 16),
                      (BC6U, "bc6u", 16), (BC6S, "bc6s", 16)):
    n = 360
    b = synth.random_blocks(n, 1, 50 + fmt)[: n * blk]
    if fmt == BC7:
        b = force_modes_bc7(b)
    if fmt in (BC6U, BC6S):
        b = force_modes_bc6h(b)
    out[key + "_blocks"] = b
    out[key + "_dec"] = refshim.decompress(fmt, b, 4, 4 * n, nthreads=1)
# BC7 / BC6H encoder goldens: the reference encoder's own output blocks for seeded inputs
img7 = synth.albedo_rgba8(64, 64, 0xB2000005, "mask")
out["bc7_src"] = img7
out["bc7_ref_enc"] = refshim.compress(BC7, img7, 64, 64, 0)
hdr = synth.sky_rgba32f(32, 32, 0xB2000006)
out["bc6h_src"] = hdr
out["bc6h_ref_enc"] = refshim.compress(BC6U, hdr, 32, 32, 0)  # as shipped (moves with stack leftovers: PSNR oracle only)
out["bc6h_ref0_enc"] = refshim.compress(BC6U, hdr, 32, 32, 0, lib=refshim.ref0())  # zero-init build: the bit-exact oracle
os.makedirs(os.path.join(ROOT, "tests", "golden"), exist_ok=True)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "bcn_golden.npz"), **out)
print("wrote", {k: v.shape for k, v in out.items()})
