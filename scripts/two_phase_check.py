"""Two-phase fine pass against the fused fine pass: frames must be bit-identical; timing of both.
    python scripts/two_phase_check.py        (GPU box)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nerf_rs_b200 as nb  # noqa: E402


def make(env):
    keys = ("NERF_B200_TWO_PHASE", "NERF_B200_FINE_SUB_RAYS")
    for k in keys:
        os.environ.pop(k, None)
    os.environ.update(env)
    try:
        return nb.Renderer(nb.DATA_ROOT, device=0)
    finally:
        for k in keys:
            os.environ.pop(k, None)


def main():
    cam = nb.camera_from_json(os.path.join(nb.DATA_ROOT, "tf_reference_samples.json"))
    fused = make({"NERF_B200_TWO_PHASE": "0"})
    two = make({"NERF_B200_TWO_PHASE": "1"})
    two_small = make({"NERF_B200_TWO_PHASE": "1", "NERF_B200_FINE_SUB_RAYS": "700"})
    ok = True
    for prec in ("mixed", "fp16", "bf16"):
        for kw in (dict(mode=nb.SAMPLING_DET), dict(mode=nb.SAMPLING_PHILOX, seed=3)):
            a = fused.render_image(cam, 64, 48, 64, 128, precision=prec, **kw)
            for name, r in (("two-phase", two), ("two-phase, 700-ray groups", two_small)):
                b = r.render_image(cam, 64, 48, 64, 128, precision=prec, **kw)
                same = bool((a == b).all())
                ok &= same
                print(f"{prec} mode {kw['mode']} {name}: identical {same}, max diff {np.abs(a - b).max():.3e}, "
                      f"finite {bool(np.isfinite(b).all())}", flush=True)
    if len(sys.argv) > 1 and sys.argv[1] == "small":
        return 0 if ok else 1
    res = {}
    for name, r in (("fused", fused), ("two-phase", two), ("fused again", fused), ("two-phase again", two)):
        for _ in range(3):
            img = r.render_image(cam, 800, 800, 64, 128, precision="mixed")
        ms, fine = [], []
        for _ in range(5):
            img = r.render_image(cam, 800, 800, 64, 128, precision="mixed")
            t = r.timing()
            ms.append(t["total_ms"]); fine.append(t["mlp_fine_ms"])
        res[name] = img
        print(f"{name}: frame {np.median(ms):.2f} ms, fine pass {np.median(fine):.2f} ms, launches {t['kernel_launches']}", flush=True)
    same = bool((res["fused"] == res["two-phase"]).all())
    print("800x800 frames identical:", same, "max diff", float(np.abs(res["fused"] - res["two-phase"]).max()))
    return 0 if ok and same else 1


if __name__ == "__main__":
    sys.exit(main())
