"""Generates tests/golden/frame_800.npz: the CPU oracle's WHOLE 800x800 lego frame (64 coarse + 128 fine samples,
deterministic sampling, tf_reference camera) - the reference render the north star's image criterion names
(render_image, src/lib.rs:474-565).  About 7 core-hours of CPU work, so the frame is rendered in bands of 8 rows with
a resumable scratch file (tests/golden/_frame_800_partial.npy, git-ignored):
    python tests/golden/make_frame800.py [threads] [det | philox SEED]
`philox SEED` writes frame_800_philox<SEED>.npz instead: the same frame with the Philox jitter of the product's default
sampling mode (the reproducible stand-in for the reference's OS-seeded thread_rng, lib.rs:243,332).

Stored as: `white` = bit mask of pixels that are exactly (1,1,1) in f32 (background: zero density along the whole
ray), `rgb` = the f32 values of all other pixels in pixel-index order; `sha256` of the f32 frame bytes.
"""
import hashlib
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import pyoracle as O  # noqa: E402

DATA = os.path.join(ROOT, "data", "lego_rust")
W = H = 800
BAND = 8


def pack(frame):
    white = np.all(frame == np.float32(1.0), axis=1)
    return dict(W=W, H=H, nc=64, nf=128, white=np.packbits(white), rgb=frame[~white],
                sha256=hashlib.sha256(frame.tobytes()).hexdigest())


def unpack(z):
    n = int(z["W"]) * int(z["H"])
    white = np.unpackbits(z["white"])[:n].astype(bool)
    frame = np.ones((n, 3), np.float32)
    frame[~white] = z["rgb"]
    return frame


def main():
    threads = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    philox = len(sys.argv) > 3 and sys.argv[2] == "philox"
    seed = int(sys.argv[3]) if philox else 0
    tag = f"_philox{seed}" if philox else ""
    kw = dict(mode=O.SAMPLING_PHILOX, seed=seed) if philox else {}
    cam = O.camera_from_json(os.path.join(DATA, "tf_reference_samples.json"))
    oc, of = O.Network(os.path.join(DATA, "coarse")), O.Network(os.path.join(DATA, "fine"))
    part = os.path.join(HERE, f"_frame_800{tag}_partial.npy")
    done_f = os.path.join(HERE, f"_frame_800{tag}_done.npy")
    frame = np.load(part) if os.path.exists(part) else np.zeros((W * H, 3), np.float32)
    done = np.load(done_f) if os.path.exists(done_f) else np.zeros(H // BAND, bool)
    t0 = time.time()
    for b in range(H // BAND):
        if done[b]:
            continue
        rb, re_ = b * BAND * W, (b + 1) * BAND * W
        frame[rb:re_] = O.render(oc, of, cam, W, H, 64, 128, ray_begin=rb, ray_end=re_, threads=threads, **kw)
        done[b] = True
        np.save(part, frame)
        np.save(done_f, done)
        print(f"band {b + 1}/{H // BAND} done, {time.time() - t0:.0f} s", flush=True)
    out = pack(frame)
    assert np.array_equal(unpack(out), frame)
    if philox:
        out["seed"] = seed
    np.savez_compressed(os.path.join(HERE, f"frame_800{tag}.npz"), **out)
    print(f"frame_800{tag}.npz written; sha256", out["sha256"], "non-white pixels", len(out["rgb"]))


if __name__ == "__main__":
    main()
