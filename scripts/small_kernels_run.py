"""Developer helper for ncu: two mixed-mode 800x800 renders (deterministic, then Philox) and nothing else, so that a
capture of the sampling / compositing kernels has a short, clean command line.
    python scripts/small_kernels_run.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nerf_rs_b200 as nb  # noqa: E402

r = nb.Renderer(nb.DATA_ROOT)
cam = nb.camera_from_json(os.path.join(nb.DATA_ROOT, "tf_reference_samples.json"))
for kw in (dict(mode=nb.SAMPLING_DET), dict(mode=nb.SAMPLING_PHILOX, seed=1)):
    img = r.render_image(cam, 800, 800, 64, 128, precision="mixed", **kw)
    print(kw, float(img.mean()), r.timing())
r.close()
