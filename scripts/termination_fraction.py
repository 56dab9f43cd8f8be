"""How much MLP work would exact early ray termination save?  compute_weights (lib.rs:250-283) zeroes every weight
after the transmittance drops below 1e-4, so samples behind that point cannot change the pixel.  Counts them."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nerf_rs_b200 as nb

r = nb.Renderer(nb.DATA_ROOT)
cam = nb.camera_from_json(nb.DATA_ROOT + "/tf_reference_samples.json")
W = H = 800
tot_c = dead_c = tot_f = dead_f = 0
rays_term = 0
for row0 in range(0, 800, 100):
    rb, re_ = row0 * W, (row0 + 100) * W
    _, st = r.render_image(cam, W, H, 64, 128, precision="mixed", ray_begin=rb, ray_end=re_, stages=True)
    wf = r.compute_weights(st["sigma_f"], st["t_f"], cam.far)
    wc = st["w_c"]
    for w, sg, name in ((wc, st["sigma_c"], "c"), (wf, st["sigma_f"], "f")):
        n = w.shape[1]
        # index of the last sample that can carry weight: everything after the early exit is zero by construction;
        # use the transmittance itself
        t = st["t_c"] if name == "c" else st["t_f"]
        d = np.diff(np.concatenate([t, np.full((t.shape[0], 1), cam.far, np.float32)], axis=1), axis=1).clip(min=0)
        T = np.cumprod(np.exp(-sg * d), axis=1)
        first_dead = np.where((T < 1e-4).any(axis=1), (T < 1e-4).argmax(axis=1) + 1, n)
        dead = (n - first_dead).sum()
        if name == "c":
            tot_c += w.size; dead_c += dead
        else:
            tot_f += w.size; dead_f += dead; rays_term += ((T < 1e-4).any(axis=1)).sum()
print(f"coarse samples behind T<1e-4: {dead_c / tot_c:.3f}; fine samples behind T<1e-4: {dead_f / tot_f:.3f}; "
      f"rays that terminate: {rays_term / (W * H):.3f}")
