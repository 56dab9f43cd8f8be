"""Developer tool: fine-pass MLP time per sample at several frame sizes, each after an idle gap, to separate the
kernel's cycle count from the power-capped clock (small frames finish before the cap pulls the SM clock down)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nerf_rs_b200 as nb
r = nb.Renderer(nb.DATA_ROOT)
cam = nb.camera_from_json(os.path.join(nb.DATA_ROOT, "tf_reference_samples.json"))
for side in (800, 256, 400, 800):
    best = None
    for rep in range(2):
        time.sleep(1.0)
        r.render_image(cam, side, side)
        ns = r.timing()["mlp_fine_ms"] * 1e6 / (side * side * 192)
        best = ns if best is None else min(best, ns)
    print(f"{side}x{side}: fine pass {best * 1e3:.3f} ps per sample ({192 * side * side / 1e6:.1f} M samples, {best * side * side * 192 / 1e6:.2f} ms)")
