"""Generates tests/golden/render_800_orbit_bands.npz: the CPU oracle's pixels for rows 396..403 of the 800x800 frame
(64 coarse + 128 fine samples, deterministic sampling) from two OTHER cameras - poses 2 and 5 of the 8-pose orbit of
nerf_rs_b200.orbit_poses (same radius and elevation as the tf_reference camera, azimuth +90 and +225 degrees) - so that
the image bound is held away from the reference pose too.  About four minutes of CPU time on two cores:
    python tests/golden/make_band_orbit.py [threads]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle as O  # noqa: E402
import nerf_rs_b200 as nb  # noqa: E402  (orbit_poses is host-side numpy; the CUDA library is not loaded)

DATA = os.path.join(ROOT, "data", "lego_rust")
POSES, N_VIEWS = (2, 5), 8


def main():
    threads = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    oc, of = O.Network(os.path.join(DATA, "coarse")), O.Network(os.path.join(DATA, "fine"))
    poses = nb.orbit_poses(os.path.join(DATA, "tf_reference_samples.json"), N_VIEWS)
    W = H = 800
    rb, re_ = 396 * W, 404 * W
    out = dict(W=W, H=H, ray_begin=rb, ray_end=re_, poses=np.array(POSES), n_views=N_VIEWS)
    for k in POSES:
        p = poses[k]
        cam = O.camera_from_values(p["near"], p["far"], p["camera_origin"], p["camera_forward"], p["camera_up"], p["hwf"])
        rgb = O.render(oc, of, cam, W, H, 64, 128, ray_begin=rb, ray_end=re_, threads=threads)
        out[f"rgb_{k}"] = rgb
        print(f"pose {k}: mean {rgb.mean(axis=0)}, non-white pixels {int((rgb.min(axis=1) < 0.999).sum())}", flush=True)
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "render_800_orbit_bands.npz"), **out)


if __name__ == "__main__":
    main()
