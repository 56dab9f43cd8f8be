import os, sys
import numpy as np
sys.path.insert(0, "/root/repo")
import nerf_rs_b200 as nb
cam = nb.camera_from_json(os.path.join(nb.DATA_ROOT, "tf_reference_samples.json"))
def make(env):
    for k in ("NERF_B200_TWO_PHASE", "NERF_B200_FINE_SUB_RAYS"): os.environ.pop(k, None)
    os.environ.update(env)
    return nb.Renderer(nb.DATA_ROOT, device=0)
rs = [("fused", make({"NERF_B200_TWO_PHASE": "0"}))]
for g in (40960, 81920, 163840, 327680):
    rs.append((f"two-phase {g}", make({"NERF_B200_TWO_PHASE": "1", "NERF_B200_FINE_SUB_RAYS": str(g)})))
for rep in range(2):
    for name, r in rs:
        for _ in range(3): r.render_image(cam, 800, 800, 64, 128, precision="mixed")
        ms, fine = [], []
        for _ in range(5):
            r.render_image(cam, 800, 800, 64, 128, precision="mixed"); t = r.timing()
            ms.append(t["total_ms"]); fine.append(t["mlp_fine_ms"])
        print(f"{name}: frame {np.median(ms):.2f} ms, fine {np.median(fine):.2f} ms", flush=True)
