"""Generates tests/golden/render_800_rows200_207.npz: the CPU oracle's pixels for rows 200..207 of the 800x800 lego
frame (64 coarse + 128 fine samples, deterministic sampling) - 6 400 rays through the object at the headline
resolution, including the thin structures on which single-pass 16-bit coarse densities move fine samples (row 207,
column 394 is the worst pixel of the all-fp16 frame).  About a minute of CPU time:
    python tests/golden/make_band800.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle as O  # noqa: E402

DATA = os.path.join(ROOT, "data", "lego_rust")


def main():
    cam = O.camera_from_json(os.path.join(DATA, "tf_reference_samples.json"))
    oc, of = O.Network(os.path.join(DATA, "coarse")), O.Network(os.path.join(DATA, "fine"))
    W = H = 800
    rb, re_ = 200 * W, 208 * W
    rgb = O.render(oc, of, cam, W, H, 64, 128, ray_begin=rb, ray_end=re_)
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "render_800_rows200_207.npz"),
                        W=W, H=H, ray_begin=rb, ray_end=re_, rgb=rgb)
    print("rows 200..207:", rgb.shape, "mean", rgb.mean(axis=0), "non-white pixels", int((rgb.min(axis=1) < 0.999).sum()))


if __name__ == "__main__":
    main()
