"""NERF_OPT_SKIP_EMPTY_RAYS on the 800x800 reference pose: rays skipped, frame time and the distance from the oracle's
frame of the reference pipeline (tests/golden/frame_800.npz).  (An experimental threshold on the coarse opacity was
measured with an earlier version of this script - profiles/r02_skip_rays_thresholds.json - and removed: up to 3 % of
opacity it skips 36.2 % of the rays instead of 35.4 %.)
    python scripts/skip_rays_check.py            (GPU box)
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nerf_rs_b200 as nb  # noqa: E402


def main():
    z = np.load(os.path.join(ROOT, "tests", "golden", "frame_800.npz"))
    n = 640000
    white = np.unpackbits(z["white"])[:n].astype(bool)
    want = np.ones((n, 3), np.float32)
    want[~white] = z["rgb"]
    r = nb.Renderer(nb.DATA_ROOT, device=0)
    cam = nb.camera_from_json(os.path.join(nb.DATA_ROOT, "tf_reference_samples.json"))
    rows = []

    def run(label):
        for _ in range(3):
            got = r.render_image(cam, 800, 800, 64, 128, precision="mixed")
        ms = []
        for _ in range(5):
            got = r.render_image(cam, 800, 800, 64, 128, precision="mixed")
            t = r.timing()
            ms.append(t["total_ms"])
        d = np.abs(got - want).max(axis=1)
        rows.append({"mode": label, "frame_ms": float(np.median(ms)), "fine_ms": t["mlp_fine_ms"],
                     "rays_skipped": int(t["rays_skipped"]), "skipped_fraction": t["rays_skipped"] / n,
                     "max_x255": float(d.max() * 255), "pixels_over_2_255": int((d > 2 / 255).sum()),
                     "pixels_over_1_255": int((d > 1 / 255).sum()),
                     "psnr_db": float(10 * np.log10(1.0 / max(float(np.mean((got.astype(np.float64) - want) ** 2)), 1e-20)))})
        print(json.dumps(rows[-1]), flush=True)
        return got

    full = run("off (the reference's pipeline)")
    r.set_option(nb.OPT_SKIP_EMPTY_RAYS, 1)
    got = run("skip rays without any coarse weight")
    ch = (got != full).any(axis=1)
    assert (got[ch] == 1.0).all()
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "skip_rays_check.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
