import sys, os, numpy as np
sys.path.insert(0, ".")
import nerf_rs_b200 as nb
r = nb.Renderer(nb.DATA_ROOT)
cam = nb.camera_from_json(nb.DATA_ROOT + "/tf_reference_samples.json")
W = H = 800
tot = 0; pos = 0; pos16 = 0; both = 0; only32 = 0; mx_only32 = 0.0
for row0 in range(0, 800, 100):
    rb, re_ = row0 * W, (row0 + 100) * W
    _, s32 = r.render_image(cam, W, H, 64, 128, precision="fp16x3", ray_begin=rb, ray_end=re_, stages=True)
    _, s16 = r.render_image(cam, W, H, 64, 128, precision="fp16", ray_begin=rb, ray_end=re_, stages=True)
    a = s32["sigma_c"].max(axis=1); b = s16["sigma_c"].max(axis=1)
    tot += len(a); pos += (a > 0).sum(); pos16 += (b > 0).sum(); both += ((a > 0) & (b > 0)).sum()
    m = (a > 0) & (b == 0); only32 += m.sum()
    if m.any(): mx_only32 = max(mx_only32, float(a[m].max()))
print(f"rays {tot}: max sigma_c > 0 in x3 {pos/tot:.3f}, in fp16 {pos16/tot:.3f}, both {both/tot:.3f}, x3-only {only32} rays (max sigma there {mx_only32:.4f})")
