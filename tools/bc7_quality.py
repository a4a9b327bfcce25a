#!/usr/bin/env python3
"""BC7 quality comparison, CUDA encoder vs the reference's bc7enc (oracle/_ref): PSNR, the encoder's own
perceptual error metric per block, mode histograms.  Test / analysis glue only (runs on the GPU box)."""
import os
import sys


ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hork-source_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import hkbcn  # noqa: E402
import synth  # noqa: E402
from bc7util import compare  # noqa: E402


if __name__ == "__main__":
    hkbcn.init()
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    compare(synth.albedo_rgba8(n, n, 0xB2000004), n, n, "albedo-B")
    compare(synth.albedo_rgba8(n, n, 0xB2000005, "mask"), n, n, "albedo-B'")
    compare(synth.albedo_rgba8(n, n, 0xB2000007, "noise"), n, n, "noise-alpha")
    rb = synth.random_blocks(4096, 4, 99).reshape(-1, 4).copy()
    compare(rb.reshape(-1), 4, 4 * 4096, "random")
    rb[:, 3] = 255
    compare(rb.reshape(-1), 4, 4 * 4096, "random-opaque")
