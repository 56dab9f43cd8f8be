"""Image parity of the tensor-core modes against the order-faithful f32 kernel over an orbit of poses (not only the
reference camera): max per-pixel difference, pixels above 2/255 and PSNR per mode."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nerf_rs_b200 as nb
from oracle import pyoracle as O  # developer script: the oracle arbitrates single pixels

n_views = int(sys.argv[1]) if len(sys.argv) > 1 else 8
res = int(sys.argv[2]) if len(sys.argv) > 2 else 400
r = nb.Renderer(nb.DATA_ROOT)
cam_json = os.path.join(nb.DATA_ROOT, "tf_reference_samples.json")
cams = nb.orbit_cameras(cam_json, n_views)
poses = nb.orbit_poses(cam_json, n_views)
onets = (O.Network(nb.DATA_ROOT + "/coarse"), O.Network(nb.DATA_ROOT + "/fine"))
worst = {}
for v, cam in enumerate(cams):
    ref = r.render_image(cam, res, res, 64, 128, precision="fp32")
    line = [f"view {v}:"]
    frames = {}
    for m in ("fp16x3", "mixed", "fp16", "bf16"):
        frames[m] = r.render_image(cam, res, res, 64, 128, precision=m)
        d = np.abs(frames[m] - ref)
        mx, over = float(d.max()) * 255, int((d.max(axis=1) > 2 / 255).sum())
        psnr = 10 * np.log10(1 / max(float(np.mean(d.astype(np.float64) ** 2)), 1e-20))
        line.append(f"{m} {mx:.2f}/255 over={over} {psnr:.1f}dB |")
        w = worst.setdefault(m, [0.0, 0, 1e9])
        w[0] = max(w[0], mx); w[1] += over; w[2] = min(w[2], psnr)
    print(" ".join(line), flush=True)
    # pixels where even the split-operand mode leaves the bound: what does the CPU oracle say there?
    dx = np.abs(frames["fp16x3"] - ref).max(axis=1)
    for px in np.nonzero(dx > 2 / 255)[0][:4]:
        px = int(px)
        o = O.render(onets[0], onets[1], O.camera_from_json(poses[v]), res, res, 64, 128, ray_begin=px, ray_end=px + 1)[0]
        print(f"    pixel {px} (row {px // res}, col {px % res}): oracle {o} | f32 kernel {ref[px]} | fp16x3 {frames['fp16x3'][px]} | "
              f"mixed {frames['mixed'][px]} -> |f32 - oracle| {np.abs(ref[px] - o).max() * 255:.2f}/255, "
              f"|fp16x3 - oracle| {np.abs(frames['fp16x3'][px] - o).max() * 255:.2f}/255", flush=True)
print("worst over", n_views, "views at", res, "x", res, ":", {k: (round(a, 2), b, round(c, 1)) for k, (a, b, c) in worst.items()})
