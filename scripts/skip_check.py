"""Developer diagnostic on a GPU box: the fine pass with and without the empty-sample colour skip (NERF_B200_SKIP_EMPTY):
frames must be bit-identical (a sample with sigma = 0 has weight exactly 0), timing per stage.
    python scripts/skip_check.py [res] [precision]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nerf_rs_b200 as nb  # noqa: E402

CAM = os.path.join(nb.DATA_ROOT, "tf_reference_samples.json")


def run(skip, cam, W, H, prec, reps=3):
    os.environ["NERF_B200_SKIP_EMPTY"] = "1" if skip else "0"
    r = nb.Renderer(nb.DATA_ROOT)
    best, img = None, None
    for _ in range(reps):
        img = r.render_image(cam, W, H, 64, 128, precision=prec)
        t = r.timing()
        if best is None or t["total_ms"] < best["total_ms"]:
            best = t
    ph = r.render_image(cam, W, H, 64, 128, precision=prec, mode=nb.SAMPLING_PHILOX, seed=4)
    r.close()
    return img, ph, best


def main():
    res = int(sys.argv[1]) if len(sys.argv) > 1 else 800
    prec = sys.argv[2] if len(sys.argv) > 2 else "mixed"
    cams = [nb.camera_from_json(CAM)] + nb.orbit_cameras(CAM, 8)[2::3]
    for ci, cam in enumerate(cams):
        W = H = res if ci == 0 else res // 2
        a, pa, ta = run(False, cam, W, H, prec)
        b, pb, tb = run(True, cam, W, H, prec)
        print(f"pose {ci} {W}x{H} {prec}: identical det {np.array_equal(a, b)} philox {np.array_equal(pa, pb)} finite "
              f"{bool(np.isfinite(b).all())} | fine {ta['mlp_fine_ms']:.2f} -> {tb['mlp_fine_ms']:.2f} ms, total "
              f"{ta['total_ms']:.2f} -> {tb['total_ms']:.2f} ms, white pixels {float(np.all(b == 1.0, axis=1).mean()):.3f}",
              flush=True)


if __name__ == "__main__":
    main()
