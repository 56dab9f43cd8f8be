"""Developer diagnostic on a GPU box: how good is the fine pass's empty-slot path at best?  Renders the 800x800 frame
with the camera turned away from the object (every ray empty: every slot skips its colour branch) and the reference
pose, with NERF_B200_SKIP_EMPTY on and off.  Skipped MMA work is 17.8 % of a quad.
    python scripts/empty_frame_check.py"""
import copy
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nerf_rs_b200 as nb  # noqa: E402

CAM = os.path.join(nb.DATA_ROOT, "tf_reference_samples.json")


def run(skip, cam, prec="fp16"):
    os.environ["NERF_B200_SKIP_EMPTY"] = "1" if skip else "0"
    r = nb.Renderer(nb.DATA_ROOT)
    best = None
    for _ in range(4):
        img = r.render_image(cam, 800, 800, 64, 128, precision=prec)
        t = r.timing()
        if best is None or t["mlp_fine_ms"] < best["mlp_fine_ms"]:
            best = t
    r.close()
    return img, best


cam = nb.camera_from_json(CAM)
away = copy.copy(cam)
away.forward[:] = [-x for x in cam.forward[:]]
for name, c in (("reference pose", cam), ("looking away", away)):
    a, ta = run(False, c)
    b, tb = run(True, c)
    print(f"{name}: white {float(np.all(b == 1.0, axis=1).mean()):.3f} identical {np.array_equal(a, b)} | fine "
          f"{ta['mlp_fine_ms']:.2f} -> {tb['mlp_fine_ms']:.2f} ms ({100 * (1 - tb['mlp_fine_ms'] / ta['mlp_fine_ms']):.1f} % saved), "
          f"coarse {ta['mlp_coarse_ms']:.2f} ms", flush=True)
