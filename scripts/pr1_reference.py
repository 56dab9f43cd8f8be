"""BASELINE.json configs[0], the reference's own run (`cargo run --release` -> main.rs -> render_cli_image,
src/lib.rs:647-677): 256x256 (lib.rs:657-658), 64 coarse + 128 fine samples (lib.rs:603-612), `render_image` alone
inside the timed region (lib.rs:668-670), output.ppm written afterwards (lib.rs:676) - executed by the CPU oracle,
because the Rust crate cannot be built here (no cargo / rustc).

  python scripts/pr1_reference.py [threads]      (about 20 minutes on 8 cores)

Rows recorded in profiles/r02_pr1_reference.json:
  R256-det     deterministic sampling (jitter 0.5, u_k = (k + 0.5) / 128): the frame the GPU path is compared with;
               its PPM is committed as tests/golden/pr1_output_det_256.ppm
  R256-philox  Philox seed 0, the reproducible stand-in for the reference's OS-seeded thread_rng (lib.rs:243,332)
  literal      16 blocks of 8x8 pixels through the object with the reference's literal GEMM loop order
               (in -> col -> out, src/network.rs:134-143), scaled to the frame and labelled extrapolated: what the
               unmodified Rust loop costs; the other rows use the bit-identical cache-friendly order
"""
import hashlib
import json
import os
import platform
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle as O  # noqa: E402

DATA = os.path.join(ROOT, "data", "lego_rust")
W = H = 256
NC, NF = 64, 128


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return platform.processor()


def main():
    threads = int(sys.argv[1]) if len(sys.argv) > 1 else len(os.sched_getaffinity(0))
    cam = O.camera_from_json(os.path.join(DATA, "tf_reference_samples.json"))
    oc, of = O.Network(os.path.join(DATA, "coarse")), O.Network(os.path.join(DATA, "fine"))
    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    rec = {"config": "BASELINE.json configs[0]: lego_rust coarse+fine render at the reference's default 256x256, 64 + 128 "
                     "samples/ray, written to output.ppm",
           "executed_by": "oracle/oracle.cpp (CPU restatement of src/lib.rs + src/network.rs; no Rust toolchain here)",
           "cpu": cpu_model(), "threads": threads, "rows": {}}
    for name, kw in (("R256-det", dict(mode=O.SAMPLING_DET)), ("R256-philox", dict(mode=O.SAMPLING_PHILOX, seed=0))):
        t0 = time.perf_counter()
        rgb = O.render(oc, of, cam, W, H, NC, NF, threads=threads, **kw)  # the reference's timed region
        dt = time.perf_counter() - t0
        path = os.path.join(out_dir, f"pr1_output_{name.split('-')[1]}_256.ppm")
        O.save_ppm(path, W, H, rgb)
        ppm = open(path, "rb").read()
        rec["rows"][name] = {"seconds": dt, "rays_per_s": W * H / dt, "samples_per_s": W * H * (2 * NC + NF) / dt,
                             "ppm_sha256": hashlib.sha256(ppm).hexdigest(), "ppm_bytes": len(ppm),
                             "f32_frame_sha256": hashlib.sha256(rgb.tobytes()).hexdigest(),
                             "mean_rgb": [float(x) for x in rgb.mean(axis=0)]}
        print(name, rec["rows"][name], flush=True)
        if name == "R256-det":
            with open(os.path.join(ROOT, "tests", "golden", "pr1_output_det_256.ppm"), "wb") as f:
                f.write(ppm)
    # the reference's literal loop order on 16 blocks (rows 128..135, columns 64..191: through the object)
    rays = 0
    t0 = time.perf_counter()
    parts = []
    for row in range(128, 136):
        rb = row * W + 64
        parts.append(O.render(oc, of, cam, W, H, NC, NF, ray_begin=rb, ray_end=rb + 128, gemm_order=O.GEMM_LITERAL,
                              threads=threads))
        rays += 128
    dt = time.perf_counter() - t0
    t1 = time.perf_counter()
    same = [O.render(oc, of, cam, W, H, NC, NF, ray_begin=row * W + 64, ray_end=row * W + 192, threads=threads)
            for row in range(128, 136)]
    dt_fast = time.perf_counter() - t1
    rec["rows"]["literal"] = {
        "rays": rays, "blocks": rays // 64, "seconds": dt, "rays_per_s": rays / dt,
        "frame_seconds_extrapolated": dt * W * H / rays,
        "same_rays_reordered_loop_seconds": dt_fast,
        "bit_identical_to_reordered_loop": bool(all(np.array_equal(a, b) for a, b in zip(parts, same))),
        "note": "src/network.rs:134-143 loop order (in -> col -> out); per-ray cost is constant, so the frame time is rays-scaled"}
    print("literal", rec["rows"]["literal"], flush=True)
    with open(os.path.join(ROOT, "profiles", "r02_pr1_reference.json"), "w") as f:
        json.dump(rec, f, indent=1)
        f.write("\n")


if __name__ == "__main__":
    main()
