import os, sys
sys.path.insert(0, "/root/repo")
import nerf_rs_b200 as nb
cam = nb.camera_from_json(os.path.join(nb.DATA_ROOT, "tf_reference_samples.json"))
r = nb.Renderer(nb.DATA_ROOT, device=0)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    r.render_image(cam, 800, 800, 64, 128, precision="mixed")
print(r.timing())
