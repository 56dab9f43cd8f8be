"""Developer diagnostic: where do the few large per-pixel differences of the 16-bit modes come from?
Renders the 800x800 frame in fp32 and fp16, takes the worst pixels, re-renders them with stage dumps in both modes and
recomposites mixtures (fp32 sample positions with fp16 network outputs and vice versa, via the stage entry points)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nerf_rs_b200 as nb  # noqa: E402

CAM = os.path.join(nb.DATA_ROOT, "tf_reference_samples.json")


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "fp16"
    W = H = 800
    r = nb.Renderer(nb.DATA_ROOT)
    cam = nb.camera_from_json(CAM)
    a = r.render_image(cam, W, H, 64, 128, precision="fp32")
    b = r.render_image(cam, W, H, 64, 128, precision=mode)
    d = np.abs(a - b).max(axis=1)
    order = np.argsort(-d)[:12]
    print(f"{mode} vs fp32: pixels > 2/255: {(d > 2 / 255).sum()}, > 1/255: {(d > 1 / 255).sum()}, max {d.max() * 255:.2f}")
    for px in order[:8]:
        px = int(px)
        ra, sa = r.render_image(cam, W, H, 64, 128, precision="fp32", ray_begin=px, ray_end=px + 1, stages=True)
        rb, sb = r.render_image(cam, W, H, 64, 128, precision=mode, ray_begin=px, ray_end=px + 1, stages=True)
        dt = np.abs(sa["t_f"] - sb["t_f"]).max()
        dsc = np.abs(sa["sigma_c"] - sb["sigma_c"]).max()
        dwc = np.abs(sa["w_c"] - sb["w_c"]).max()
        # fine network of `mode` evaluated at the fp32 pass's sample positions
        o = np.array(cam.origin[:], np.float32)
        dirs = np.repeat(sa["dirs"], 192, axis=0)
        pts = (o[:, None] + dirs.T * sa["t_f"].reshape(-1)[None, :]).astype(np.float32)
        rgb_m, sig_m = r.fine.forward_batch(pts, dirs, precision=mode)
        mix = r.integrate_rays(rgb_m.reshape(1, 192, 3), sig_m.reshape(1, 192), sa["t_f"], cam.far) if hasattr(r, "integrate_rays") else None
        print(f"px {px} (row {px // W}, col {px % W}): diff*255 {d[px] * 255:.2f} | fp32 {ra[0]} {mode} {rb[0]} | "
              f"max|dt_f| {dt:.3e} max|dsigma_c| {dsc:.3e} (max sigma_c {sa['sigma_c'].max():.1f}) max|dw_c| {dwc:.3e} | "
              f"max sigma_f {sa['sigma_f'].max():.1f} | {mode} net at fp32 positions -> {None if mix is None else mix[0]}")
        k = int(np.argmax(np.abs(sa["t_f"] - sb["t_f"])))
        print("    t_f fp32 around max dt:", sa["t_f"][0, max(0, k - 3):k + 4], "\n    t_f", mode, "           :", sb["t_f"][0, max(0, k - 3):k + 4])
        print("    w_c fp32 nonzero:", np.count_nonzero(sa["w_c"]), mode, np.count_nonzero(sb["w_c"]))
    r.close()


if __name__ == "__main__":
    main()
