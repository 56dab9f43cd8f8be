"""Developer diagnostic for the two-pass coarse densities of the mixed mode, on a GPU box:
is the frame bit-identical to the one with the split-operand kernel over EVERY coarse sample, how many samples are
re-evaluated, what does it cost, and how much slack do the margins have?
    python scripts/select_check.py [res]
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nerf_rs_b200 as nb  # noqa: E402

CAM = os.path.join(nb.DATA_ROOT, "tf_reference_samples.json")


def render(env, cam, W, H, prec="mixed", stages=True, reps=1):
    for k in ("NERF_B200_COARSE_SELECT", "NERF_B200_SELECT_MARGIN", "NERF_B200_SELECT_TAU", "NERF_B200_COARSE_STAGES"):
        os.environ.pop(k, None)
    os.environ.update(env)
    r = nb.Renderer(nb.DATA_ROOT)
    n = W * H
    rgb = np.empty((n, 3), np.float32)
    out = dict(rgb=rgb)
    smp = nb.Sampling(mode=nb.SAMPLING_DET, seed=0)
    if stages:
        out.update(sigma_c=np.zeros((n, 64), np.float32), w_c=np.zeros((n, 64), np.float32),
                   t_f=np.zeros((n, 192), np.float32))
        st = nb.Stages(**{k: nb._p(v) for k, v in out.items() if k != "rgb"})
        nb._check(nb.lib().nerf_render_stages(r._h, C.byref(cam), W, H, 64, 128, nb.PRECISIONS[prec], C.byref(smp), 0, n,
                                              nb._p(rgb), C.byref(st)))
    best = None
    for _ in range(reps):
        nb._check(nb.lib().nerf_render(r._h, C.byref(cam), W, H, 64, 128, nb.PRECISIONS[prec], C.byref(smp), 0, n,
                                       nb._p(rgb)))
        t = r.timing()
        if best is None or t["total_ms"] < best["total_ms"]:
            best = t
    out["timing"] = best
    r.close()
    return out


def compare(tag, a, b):
    same_t = np.array_equal(a["t_f"], b["t_f"])
    same_w = np.array_equal(a["w_c"], b["w_c"])
    d = np.abs(a["rgb"] - b["rgb"]).max()
    rays = int((a["t_f"] != b["t_f"]).any(axis=1).sum())
    t = b["timing"]
    frac = t["coarse_refined"] / max(t["coarse_samples"], 1)
    print(f"[{tag}] t_f identical {same_t} (rays differing {rays}), w_c identical {same_w}, max |d rgb| {d:.3e}, "
          f"sigma_c entries differing {int((a['sigma_c'] != b['sigma_c']).sum())} | refined {frac:.4f} of the coarse samples, "
          f"coarse {t['mlp_coarse_ms']:.2f} ms (refine part {t['coarse_refine_ms']:.2f}), fine {t['mlp_fine_ms']:.2f}, "
          f"total {t['total_ms']:.2f} ms", flush=True)


def main():
    res = int(sys.argv[1]) if len(sys.argv) > 1 else 800
    cams = [nb.camera_from_json(CAM)] + nb.orbit_cameras(CAM, 8)[1::3]
    for ci, cam in enumerate(cams):
        W = H = res if ci == 0 else res // 2
        full = render({"NERF_B200_COARSE_SELECT": "0"}, cam, W, H, reps=2 if ci == 0 else 1)
        t = full["timing"]
        print(f"pose {ci} {W}x{H}: every sample split-operand: coarse {t['mlp_coarse_ms']:.2f} ms, fine "
              f"{t['mlp_fine_ms']:.2f}, total {t['total_ms']:.2f}", flush=True)
        compare(f"pose {ci} default", full, render({}, cam, W, H, reps=2 if ci == 0 else 1))
        if ci == 0:
            for stages in ("1", "2", "4", "8"):
                compare(f"stages {stages}", full, render({"NERF_B200_COARSE_STAGES": stages}, cam, W, H, reps=3))
            for margin in ("0.5", "0.25", "0.1", "0.03", "0.0"):
                compare(f"margin {margin}", full, render({"NERF_B200_SELECT_MARGIN": margin}, cam, W, H))
            for tau in ("10.5", "9.5", "9.0", "8.0"):
                compare(f"tau {tau}", full, render({"NERF_B200_SELECT_TAU": tau}, cam, W, H))


if __name__ == "__main__":
    main()
