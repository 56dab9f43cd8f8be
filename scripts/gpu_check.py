"""Developer diagnostic for a GPU box: per-mode forward_batch error statistics and a timed 800x800 render.
Prints everything it can even when a mode is wrong, so one gpurun call gives the full picture."""
import os
import sys
import time
import traceback

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nerf_rs_b200 as nb  # noqa: E402
from oracle import pyoracle as O  # noqa: E402

CAM = os.path.join(nb.DATA_ROOT, "tf_reference_samples.json")


def main():
    modes = sys.argv[1].split(",") if len(sys.argv) > 1 else ["fp32", "fp16", "bf16"]
    big = int(sys.argv[2]) if len(sys.argv) > 2 else 800
    r = nb.Renderer(nb.DATA_ROOT)
    cam = nb.camera_from_json(CAM)
    ocam = O.camera_from_json(CAM)
    oc, of = O.Network(nb.DATA_ROOT + "/coarse"), O.Network(nb.DATA_ROOT + "/fine")
    rng = np.random.default_rng(0)
    n = 1000
    dirs = O.ray_dirs(ocam, 800, 800)[rng.integers(0, 640000, n)]
    t = rng.uniform(2, 6, n).astype(np.float32)
    pts = (np.array(ocam.pos, np.float32)[:, None] + dirs.T * t[None]).astype(np.float32)
    ref = {0: oc.forward_batch(pts, dirs), 1: of.forward_batch(pts, dirs)}
    for m in modes:
        for net_id, net in ((0, r.coarse), (1, r.fine)):
            try:
                rgb, sig = net.forward_batch(pts, dirs, precision=m)
                rr, sr = ref[net_id]
                rel = np.abs(sig - sr) / np.maximum(1, np.abs(sr))
                print(f"[fwd] {m} net{net_id}: sigma rel max {rel.max():.3e} med {np.median(rel):.3e} "
                      f"rgb abs max {np.abs(rgb - rr).max():.3e} finite {np.isfinite(rgb).all() and np.isfinite(sig).all()} "
                      f"| sig[:4] {sig[:4]} ref {sr[:4]}", flush=True)
            except Exception:
                traceback.print_exc()
    want = O.render(oc, of, ocam, 32, 32, 64, 128)
    for m in modes:
        try:
            got = r.render_image(cam, 32, 32, 64, 128, precision=m)
            mse = float(np.mean((got - want) ** 2))
            print(f"[render32] {m}: max px diff*255 {np.abs(got - want).max() * 255:.3f} "
                  f"PSNR {10 * np.log10(1 / max(mse, 1e-20)):.1f} dB", flush=True)
        except Exception:
            traceback.print_exc()
    for m in modes:
        if m == "fp32" and big > 256:
            continue
        try:
            out = np.empty((big * big, 3), np.float32)
            for it in range(3):
                t0 = time.time()
                r.render_image(cam, big, big, 64, 128, precision=m, out=out)
                dt = time.time() - t0
                tm = r.timing()
                fl = big * big * (64 * 982528 + 192 * 1186816)
                print(f"[render{big}] {m} it{it}: host {dt * 1e3:.1f} ms, device {tm['total_ms']:.2f} ms "
                      f"(coarse {tm['mlp_coarse_ms']:.2f} fine {tm['mlp_fine_ms']:.2f} sampling {tm['sampling_ms']:.2f} "
                      f"composite {tm['composite_ms']:.2f}) -> {big * big / tm['total_ms'] / 1e3:.3f} Mrays/s, "
                      f"{fl / tm['total_ms'] / 1e9:.1f} TFLOP/s", flush=True)
            nb.save_ppm(os.path.join(ROOT, "gpurun_out", f"lego_{big}_{m}.ppm"), big, big, out)
        except Exception:
            traceback.print_exc()
    r.close()


if __name__ == "__main__":
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    main()
