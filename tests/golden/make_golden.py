"""Generates tests/golden/render_8x8_stages.npz: the CPU oracle's per-stage outputs for an 8x8 lego render
(64 coarse + 128 fine samples, deterministic sampling) and a Philox-jittered 8x8 frame, plus the oracle's
forward_batch on 256 seeded samples.  The Rust reference cannot be run here (no cargo), so these freeze the oracle
(itself pinned by lego_rust/tf_reference_samples.json) against drift; regenerate only when the oracle is
deliberately changed:   python tests/golden/make_golden.py
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
    # rays through the object: rows 28..35 x cols 28..35 of a 64x64 frame, as 8 row segments
    W = H = 64
    segs = [(r * W + 28, r * W + 36) for r in range(28, 36)]
    rgb, st = [], {}
    for b, e in segs:
        img, s = O.render(oc, of, cam, W, H, 64, 128, ray_begin=b, ray_end=e, stages=True)
        rgb.append(img)
        for k, v in s.items():
            st.setdefault(k, []).append(v)
    out = {"W": W, "H": H, "segments": np.array(segs, np.int64), "rgb": np.concatenate(rgb)}
    out.update({k: np.concatenate(v) for k, v in st.items()})
    out["rgb_philox_seed7"] = np.concatenate(
        [O.render(oc, of, cam, W, H, 64, 128, mode=O.SAMPLING_PHILOX, seed=7, ray_begin=b, ray_end=e) for b, e in segs])
    rng = np.random.default_rng(123)
    dirs = O.ray_dirs(cam, 800, 800)[rng.integers(0, 640000, 256)]
    t = rng.uniform(2.0, 6.0, 256).astype(np.float32)
    pts = (np.array(cam.pos, np.float32)[:, None] + dirs.T * t[None, :]).astype(np.float32)
    out["fwd_points"], out["fwd_dirs"] = pts, dirs
    for name, net in (("coarse", oc), ("fine", of)):
        c, s = net.forward_batch(pts, dirs)
        out[f"fwd_{name}_rgb"], out[f"fwd_{name}_sigma"] = c, s
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "render_8x8_stages.npz"), **out)
    print({k: getattr(v, "shape", v) for k, v in out.items()})


if __name__ == "__main__":
    main()
