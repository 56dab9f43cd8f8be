"""Developer diagnostic for the fp16x3 mode on a GPU box: forward_batch errors (golden JSON + random, ragged sizes),
a small render against the oracle, and timed 800x800 renders of fp16x3 / fp16 / fp32."""
import json
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
    big = int(sys.argv[1]) if len(sys.argv) > 1 else 800
    r = nb.Renderer(nb.DATA_ROOT)
    cam = nb.camera_from_json(CAM)
    ocam = O.camera_from_json(CAM)
    oc, of = O.Network(nb.DATA_ROOT + "/coarse"), O.Network(nb.DATA_ROOT + "/fine")
    golden = json.load(open(CAM))
    o = np.array(golden["camera_origin"], np.float32)
    z = np.array(golden["z_vals"], np.float32)
    P, D = [], []
    for ex in golden["examples"]:
        rd = np.array(ex["ray_d"], np.float32)
        P.append((o[:, None] + rd[:, None] * z[None, :]).astype(np.float32))
        D.append(np.tile(np.array(ex["viewdir_unit"], np.float32), (len(z), 1)))
    gp, gd = np.concatenate(P, axis=1), np.concatenate(D, axis=0)
    for net, key in ((r.coarse, "coarse"), (r.fine, "fine")):
        ws = np.concatenate([ex[key + "_sigma"] for ex in golden["examples"]]).astype(np.float32)
        wc = np.concatenate([ex[key + "_rgb"] for ex in golden["examples"]]).astype(np.float32)
        for m in ("fp16x3", "fp32"):
            try:
                rgb, sig = net.forward_batch(gp, gd, precision=m)
                print(f"[golden] {m} {key}: sigma abs max {np.abs(sig - ws).max():.3e} rgb abs max {np.abs(rgb - wc).max():.3e}", flush=True)
            except Exception:
                traceback.print_exc()
    rng = np.random.default_rng(0)
    for n in (1, 127, 128, 129, 255, 257, 1000, 4173, 100000):
        dirs = O.ray_dirs(ocam, 800, 800)[rng.integers(0, 640000, n)]
        t = rng.uniform(2, 6, n).astype(np.float32)
        pts = (np.array(ocam.pos, np.float32)[:, None] + dirs.T * t[None]).astype(np.float32)
        for net_id, (net, onet) in enumerate(((r.coarse, oc), (r.fine, of))):
            try:
                rr, sr = onet.forward_batch(pts, dirs)
                rgb, sig = net.forward_batch(pts, dirs, precision="fp16x3")
                r32, s32 = net.forward_batch(pts, dirs, precision="fp32")
                print(f"[fwd] n={n} net{net_id}: x3 sigma abs max {np.abs(sig - sr).max():.3e} rel max "
                      f"{(np.abs(sig - sr) / np.maximum(1, np.abs(sr))).max():.3e} rgb abs max {np.abs(rgb - rr).max():.3e} | "
                      f"fp32 sigma abs max {np.abs(s32 - sr).max():.3e} rgb {np.abs(r32 - rr).max():.3e} | max sigma {sr.max():.1f}",
                      flush=True)
            except Exception:
                traceback.print_exc()
    for sz in (32, 64):
        want = O.render(oc, of, ocam, sz, sz, 64, 128)
        for m in ("fp16x3", "fp32", "fp16"):
            try:
                got = r.render_image(cam, sz, sz, 64, 128, precision=m)
                mse = float(np.mean((got - want) ** 2))
                print(f"[render{sz}] {m}: max px diff*255 {np.abs(got - want).max() * 255:.4f} "
                      f"PSNR {10 * np.log10(1 / max(mse, 1e-20)):.1f} dB", flush=True)
            except Exception:
                traceback.print_exc()
    frames = {}
    for m in ("fp16x3", "mixed", "fp16", "fp32"):
        try:
            out = np.empty((big * big, 3), np.float32)
            for it in range(2 if m != "fp32" else 1):
                t0 = time.time()
                r.render_image(cam, big, big, 64, 128, precision=m, out=out)
                dt = time.time() - t0
                tm = r.timing()
                fl = big * big * (64 * 982528 + 192 * 1186816)
                print(f"[render{big}] {m} it{it}: host {dt * 1e3:.1f} ms, device {tm['total_ms']:.2f} ms "
                      f"(coarse {tm['mlp_coarse_ms']:.2f} fine {tm['mlp_fine_ms']:.2f}) -> "
                      f"{big * big / tm['total_ms'] / 1e3:.3f} Mrays/s, {fl / tm['total_ms'] / 1e9:.1f} algorithmic TFLOP/s", flush=True)
            frames[m] = out.copy()
        except Exception:
            traceback.print_exc()
    if "fp32" in frames:
        for m in ("fp16x3", "mixed", "fp16"):
            if m in frames:
                d = np.abs(frames[m] - frames["fp32"])
                mse = float(np.mean(d ** 2))
                print(f"[frame{big}] {m} vs fp32 GPU frame: max px diff*255 {d.max() * 255:.4f}, pixels > 2/255: "
                      f"{int((d.max(axis=1) > 2 / 255).sum())}, PSNR {10 * np.log10(1 / max(mse, 1e-20)):.1f} dB", flush=True)
    r.close()


if __name__ == "__main__":
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    main()
