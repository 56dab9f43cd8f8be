"""Developer tool: run one fine-network forward on a few tiles with NERF_B200_TRACE set and print the CTA-0 timeline."""
import os, sys, struct
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
path = os.path.join(ROOT, "gpurun_out", "trace.bin")
os.makedirs(os.path.dirname(path), exist_ok=True)
if os.path.exists(path):
    os.unlink(path)
import nerf_rs_b200 as nb
r = nb.Renderer(nb.DATA_ROOT)
n = 148 * 256 * 3  # 3 tile pairs per CTA
rng = np.random.default_rng(0)
pts = rng.uniform(-1, 1, (3, n)).astype(np.float32)
dirs = rng.normal(size=(n, 3)).astype(np.float32)
dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
r.fine.forward_batch(pts, dirs, precision="fp16")  # warm
os.environ["NERF_B200_TRACE"] = path
r.fine.forward_batch(pts, dirs, precision="fp16")
del os.environ["NERF_B200_TRACE"]
raw = np.fromfile(path, np.uint64).reshape(7, 2048, 2)
names = ["mma", "tma", "epi0", "epi1", "epi2", "epi3", "relay"]
t0 = min(int(raw[k, 0, 1]) for k in range(7) if raw[k, 0, 1])
for k in range(7):
    ev = [(int(t) - t0, int(tag)) for tag, t in raw[k] if t]
    print(names[k], len(ev))
    print(" ".join(f"{tag:03x}@{t}" for t, tag in ev[:int(sys.argv[1]) if len(sys.argv) > 1 else 120]))
