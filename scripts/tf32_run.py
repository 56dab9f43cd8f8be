"""Developer helper for ncu: one 400x400 render in the tf32 mode (two launches of mlp_tc_tf32_kernel).
    python scripts/tf32_run.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nerf_rs_b200 as nb  # noqa: E402

r = nb.Renderer(nb.DATA_ROOT)
cam = nb.camera_from_json(os.path.join(nb.DATA_ROOT, "tf_reference_samples.json"))
img = r.render_image(cam, 400, 400, 64, 128, precision="tf32")
print(float(img.mean()), r.timing())
r.close()
