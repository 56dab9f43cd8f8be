"""Committed golden fixtures (tests/golden/render_8x8_stages.npz, made by tests/golden/make_golden.py):
the CPU test pins the oracle against them (drift guard), the GPU test compares the CUDA path with them directly."""
import os

import numpy as np
import pytest

from oracle import pyoracle as O

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "render_8x8_stages.npz"))
W, H = int(G["W"]), int(G["H"])
SEGS = [tuple(int(x) for x in s) for s in G["segments"]]


def test_oracle_reproduces_golden(onets, ocam):
    rgb = np.concatenate([O.render(onets[0], onets[1], ocam, W, H, 64, 128, ray_begin=b, ray_end=e) for b, e in SEGS])
    # same binary, same libm: bit-exact; a different glibc may move sinf/expf by an ulp
    assert np.abs(rgb - G["rgb"]).max() <= 2e-6
    c, s = onets[1].forward_batch(G["fwd_points"], G["fwd_dirs"])
    assert np.abs(c - G["fwd_fine_rgb"]).max() <= 1e-6
    assert (np.abs(s - G["fwd_fine_sigma"]) <= 1e-5 * np.maximum(1, np.abs(G["fwd_fine_sigma"]))).all()
    ph = np.concatenate([O.render(onets[0], onets[1], ocam, W, H, 64, 128, mode=O.SAMPLING_PHILOX, seed=7, ray_begin=b,
                                  ray_end=e) for b, e in SEGS])
    assert np.abs(ph - G["rgb_philox_seed7"]).max() <= 2e-6


@pytest.mark.gpu
def test_gpu_matches_golden_stages(renderer, cam):
    got, st = [], {}
    for b, e in SEGS:
        img, s = renderer.render_image(cam, W, H, 64, 128, precision="fp32", stages=True, ray_begin=b, ray_end=e)
        got.append(img)
        for k, v in s.items():
            st.setdefault(k, []).append(v)
    got = np.concatenate(got)
    st = {k: np.concatenate(v) for k, v in st.items()}
    assert np.abs(st["dirs"] - G["dirs"]).max() <= 1.2e-7
    assert np.abs(st["t_c"] - G["t_c"]).max() == 0.0
    rel = np.abs(st["sigma_c"] - G["sigma_c"]) / np.maximum(1, np.abs(G["sigma_c"]))
    assert rel.max() < 5e-4
    assert np.abs(st["w_c"] - G["w_c"]).max() < 5e-4
    assert np.abs(got - G["rgb"]).max() <= 2.0 / 255.0  # north-star image bound (fp32 mode is far inside it)
    assert np.abs(got - G["rgb"]).max() < 2e-3
    ph = np.concatenate([renderer.render_image(cam, W, H, 64, 128, precision="fp32", mode=1, seed=7, ray_begin=b,
                                               ray_end=e) for b, e in SEGS])
    assert np.abs(ph - G["rgb_philox_seed7"]).max() < 2e-3


@pytest.mark.gpu
@pytest.mark.parametrize("prec,stol,ctol", [("fp32", 2e-4, 2e-5), ("fp16", 2e-2, 6e-3), ("bf16", 3e-1, 6e-2)])
def test_gpu_forward_batch_matches_golden(renderer, prec, stol, ctol):
    for net, name in ((renderer.coarse, "coarse"), (renderer.fine, "fine")):
        c, s = net.forward_batch(G["fwd_points"], G["fwd_dirs"], precision=prec)
        ws = G[f"fwd_{name}_sigma"]
        assert (np.abs(s - ws) <= stol * np.maximum(1, np.abs(ws))).all()
        assert np.abs(c - G[f"fwd_{name}_rgb"]).max() <= ctol
