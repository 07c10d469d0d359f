#!/usr/bin/env python
"""Diff-image tool for parity debugging.

    python tools/diff_images.py a.npy b.npy out.png [--scale 32]
    python tools/diff_images.py a.png b.png out.png

Inputs are rgba32f accumulations saved with numpy ((h, w, 4) float32: what Renderer.read_accum() and the oracle return)
or PNG files written by gpu_ray_tracing_b200.write_png.  Writes a picture of the per-pixel RGBA8 difference (x scale,
pixels over the 1/255 tolerance in red) and prints the mismatch report."""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_ray_tracing_b200 as rt  # noqa: E402


def load(path):
    return rt.read_png(path) if path.lower().endswith(".png") else np.load(path)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("a")
    ap.add_argument("b")
    ap.add_argument("out")
    ap.add_argument("--scale", type=int, default=32)
    args = ap.parse_args()
    pic, report = rt.diff_image(load(args.a), load(args.b), args.scale)
    rt.write_png(args.out, pic)
    print(json.dumps(report))
