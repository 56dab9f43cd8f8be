"""NERF_OPT_SKIP_EMPTY_RAYS (include/nerf_b200.h): the opt-in mode in which rays without any
coarse weight get the background and skip the fine pass.  It is NOT the reference's image, so it has its own oracle
definition (oracle/pyoracle.py::render_skip_empty_rays) and its own tests:

  - against that definition evaluated by the CPU oracle, on rows that cross the edge of the object;
  - every ray that is not skipped is bit-identical to the same ray without the option, every skipped ray is exactly
    the background, and the skipped rays are exactly those whose coarse weights (the library's own stage dump) are all
    zero - every precision mode, every sampling mode, chunked, striped, 8-bit and depth / opacity outputs;
  - the whole 800x800 frame against the oracle's frame of the REFERENCE pipeline: what the option costs in accuracy.
"""
import os

import numpy as np
import pytest

from oracle import pyoracle as O
import nerf_rs_b200 as nb

from .conftest import DATA
from .test_gpu_round2 import _frame800, _psnr

pytestmark = pytest.mark.gpu


@pytest.fixture()
def fast(renderer):
    """The session's renderer with the option set for the duration of one test."""
    assert renderer.get_option(nb.OPT_SKIP_EMPTY_RAYS) == 0  # off unless asked for
    renderer.set_option(nb.OPT_SKIP_EMPTY_RAYS, 1)
    yield renderer
    renderer.set_option(nb.OPT_SKIP_EMPTY_RAYS, 0)


def test_option_api(renderer):
    assert renderer.get_option(nb.OPT_SKIP_EMPTY_RAYS) == 0
    renderer.set_option(nb.OPT_SKIP_EMPTY_RAYS, 5)
    assert renderer.get_option(nb.OPT_SKIP_EMPTY_RAYS) == 1
    renderer.set_option(nb.OPT_SKIP_EMPTY_RAYS, 0)
    assert renderer.get_option(nb.OPT_SKIP_EMPTY_RAYS) == 0
    with pytest.raises(nb.NerfError):
        renderer.set_option(77, 1)
    with pytest.raises(nb.NerfError):
        renderer.get_option(0)
    os.environ["NERF_B200_SKIP_EMPTY_RAYS"] = "1"
    try:
        r = nb.Renderer(DATA, device=0)
    finally:
        del os.environ["NERF_B200_SKIP_EMPTY_RAYS"]
    try:
        assert r.get_option(nb.OPT_SKIP_EMPTY_RAYS) == 1
    finally:
        r.close()


@pytest.mark.parametrize("kw", [dict(mode=O.SAMPLING_DET), dict(mode=O.SAMPLING_PHILOX, seed=4)])
def test_against_the_oracle_definition(fast, cam, onets, ocam, kw):
    """Row 12 of a 64x64 frame (rays without any coarse weight, then the coarse network's haze, then the object) and
    half of row 40, order-faithful f32 and the default mixed mode, at the reference's 64 + 128 samples."""
    W = H = 64
    seen = 0
    for b, e in ((12 * W, 13 * W), (40 * W, 40 * W + 32)):
        want, skipped = O.render_skip_empty_rays(onets[0], onets[1], ocam, W, H, 64, 128, ray_begin=b, ray_end=e, **kw)
        seen += int(skipped.sum())
        for prec, tol in (("fp32", 1.0), ("mixed", 2.0)):
            got = fast.render_image(cam, W, H, 64, 128, precision=prec, ray_begin=b, ray_end=e, **kw)
            t = fast.timing()
            d = np.abs(got - want).max(axis=1)
            print(f"rays [{b},{e}) {prec}: oracle skips {int(skipped.sum())}, library {t['rays_skipped']} of {t['rays']}; "
                  f"max {d.max() * 255:.3f}/255")
            assert t["rays"] == e - b
            # the two sides agree on which rays are skipped, up to rays that graze a density of ~1e-7
            assert abs(int(t["rays_skipped"]) - int(skipped.sum())) <= 2
            assert int((d > tol / 255.0).sum()) <= 2
    assert seen >= 20
    # a range in which no ray has any coarse weight: nothing reaches the fine pass (the list of length 0)
    want, skipped = O.render_skip_empty_rays(onets[0], onets[1], ocam, W, H, 64, 128, ray_begin=12 * W, ray_end=12 * W + 8, **kw)
    assert skipped.all() and (want == 1.0).all()
    got = fast.render_image(cam, W, H, 64, 128, precision="mixed", ray_begin=12 * W, ray_end=12 * W + 8, **kw)
    t = fast.timing()
    assert (got == 1.0).all() and t["rays_skipped"] == t["rays"] == 8 and t["mlp_fine_ms"] == 0.0


@pytest.mark.parametrize("prec", ["mixed", "fp32", "fp16x3", "fp16", "tf32", "bf16"])
def test_kept_rays_are_bit_identical_and_skipped_rays_are_background(renderer, cam, prec):
    W, H, nc, nf = 64, 48, 64, 128
    n = W * H
    rng = np.random.default_rng(5)
    jit = rng.uniform(0, 1, (n, nc)).astype(np.float32)
    u = rng.uniform(0, 1, (n, nf)).astype(np.float32)
    modes = [dict(mode=nb.SAMPLING_DET), dict(mode=nb.SAMPLING_PHILOX, seed=11)]
    if prec in ("mixed", "fp32"):
        modes.append(dict(mode=nb.SAMPLING_ARRAYS, jitter=jit, u=u))
    for kw in modes:
        full, st = renderer.render_image(cam, W, H, nc, nf, precision=prec, stages=True, **kw)
        empty = np.all(st["w_c"] == 0.0, axis=1)
        assert 0.1 < empty.mean() < 0.9
        renderer.set_option(nb.OPT_SKIP_EMPTY_RAYS, 1)
        try:
            got = renderer.render_image(cam, W, H, nc, nf, precision=prec, **kw)
            t = renderer.timing()
            # stage dumps are indexed by ray: they ignore the option
            again, st2 = renderer.render_image(cam, W, H, nc, nf, precision=prec, stages=True, **kw)
        finally:
            renderer.set_option(nb.OPT_SKIP_EMPTY_RAYS, 0)
        assert t["rays"] == n and t["rays_skipped"] == int(empty.sum()), (prec, kw["mode"])
        assert (got[~empty] == full[~empty]).all(), (prec, kw["mode"])
        assert (got[empty] == 1.0).all(), (prec, kw["mode"])
        assert (again == full).all() and (st2["sigma_f"] == st["sigma_f"]).all()
        # what the option changes: rays on which only the fine network finds density
        print(f"{prec} mode {kw['mode']}: {int(empty.sum())} of {n} rays skipped, "
              f"{int((full[empty] != 1.0).any(axis=1).sum())} of them not white in the full render, "
              f"max {np.abs(full[empty] - 1.0).max() * 255:.3f}/255")


def test_other_entry_points_and_chunks(renderer, cam):
    """Depth / opacity of skipped rays are 0; the 8-bit, striped, ranged and chunked paths take the same route."""
    import torch
    from nerf_rs_b200 import shard
    W, H, nc, nf = 40, 32, 64, 128
    n = W * H
    os.environ["NERF_B200_CHUNK_RAYS"] = "150"
    os.environ["NERF_B200_SKIP_EMPTY_RAYS"] = "1"
    try:
        small = nb.Renderer(DATA, device=0)
    finally:
        del os.environ["NERF_B200_CHUNK_RAYS"], os.environ["NERF_B200_SKIP_EMPTY_RAYS"]
    try:
        for kw in (dict(mode=nb.SAMPLING_DET), dict(mode=nb.SAMPLING_PHILOX, seed=3)):
            full, st = renderer.render_image(cam, W, H, nc, nf, precision="mixed", stages=True, **kw)
            empty = np.all(st["w_c"] == 0.0, axis=1)
            want = full.copy()
            want[empty] = 1.0
            rf, df, af = renderer.render_aux(cam, W, H, nc, nf, precision="mixed", **kw)
            renderer.set_option(nb.OPT_SKIP_EMPTY_RAYS, 1)
            try:
                rgb, depth, acc = renderer.render_aux(cam, W, H, nc, nf, precision="mixed", **kw)
                q = renderer.render_rgb8(cam, W, H, nc, nf, precision="mixed", **kw)
                part = renderer.render_image(cam, W, H, nc, nf, precision="mixed", ray_begin=77, ray_end=1001, **kw)
                host = np.full((n, 3), -7.0, np.float32)
                dev = torch.full((n, 3), -7.0, device="cuda:0")
                for p in range(3):
                    renderer.render_striped(cam, W, H, host, 53, 3, p, precision="mixed", **kw)
                    renderer.render_striped_device(cam, W, H, dev.data_ptr(), 53, 3, p, precision="mixed", **kw)
                torch.cuda.synchronize()
            finally:
                renderer.set_option(nb.OPT_SKIP_EMPTY_RAYS, 0)
            assert (rgb == want).all()
            assert (depth[empty] == 0).all() and (acc[empty] == 0).all()
            assert (depth[~empty] == df[~empty]).all() and (acc[~empty] == af[~empty]).all()
            assert (q == nb.quantize_rgb8(want)).all()
            assert (part == want[77:1001]).all()
            assert (host == want).all() and (dev.cpu().numpy() == want).all()
            # chunks of 150 rays (some all background, some mixed), option from the environment
            assert (small.render_image(cam, W, H, nc, nf, precision="mixed", **kw) == want).all()
            t = small.timing()
            assert t["rays"] == n and t["rays_skipped"] == int(empty.sum())
            r2, d2, a2 = small.render_aux(cam, W, H, nc, nf, precision="mixed", **kw)
            assert (r2 == want).all() and (d2 == depth).all() and (a2 == acc).all()
    finally:
        small.close()


def test_full_800_frame_what_the_option_costs(renderer, cam):
    """The 800x800 frame with the option against the oracle's frame of the reference pipeline."""
    want = _frame800()
    full = renderer.render_image(cam, 800, 800, 64, 128, precision="mixed")
    t0 = renderer.timing()
    renderer.set_option(nb.OPT_SKIP_EMPTY_RAYS, 1)
    try:
        renderer.render_image(cam, 800, 800, 64, 128, precision="mixed")  # warm the new workspaces
        got = renderer.render_image(cam, 800, 800, 64, 128, precision="mixed")
        t1 = renderer.timing()
    finally:
        renderer.set_option(nb.OPT_SKIP_EMPTY_RAYS, 0)
    changed = (got != full).any(axis=1)
    assert (got[changed] == 1.0).all()  # the only change is a ray turned into background
    d = np.abs(got - want).max(axis=1)
    bad = int((d > 2.0 / 255.0).sum())
    print(f"800x800 skip-empty-rays vs the oracle's reference frame: {t1['rays_skipped']} of {t1['rays']} rays skipped, "
          f"{int(changed.sum())} pixels differ from the full render (max {np.abs(got - full).max() * 255:.3f}/255); "
          f"vs oracle max {d.max() * 255:.3f}/255, pixels > 2/255: {bad}, PSNR {_psnr(got, want):.2f} dB; "
          f"frame {t0['total_ms']:.1f} ms -> {t1['total_ms']:.1f} ms (fine pass {t0['mlp_fine_ms']:.1f} -> "
          f"{t1['mlp_fine_ms']:.1f} ms)")
    assert t1["rays"] == 640000 and 0.25 < t1["rays_skipped"] / 640000 < 0.5  # measured: 35.4 %
    assert _psnr(got, want) >= 70.0 and bad <= 60  # measured: 79.9 dB, 29 pixels
    assert t1["total_ms"] < 0.85 * t0["total_ms"]  # measured: 124 -> 92 ms
