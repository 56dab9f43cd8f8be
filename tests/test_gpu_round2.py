"""GPU tests added in round 2 (all through the C ABI, against the CPU oracle or against the library's own
slower-but-simpler path where the claim is bit-identity):

  - the WHOLE 800x800 frame of every mode against the oracle's frame (tests/golden/frame_800.npz) - the north star's
    image criterion "against the reference render" (render_image, src/lib.rs:474-565) on all 640 000 pixels;
  - the product default (Philox jitter + mixed precision) at the reference's sample counts against the oracle;
  - the mixed mode's two-pass coarse densities == the split-operand kernel over every coarse sample, bit for bit;
  - renders cut into chunks == the single-chunk render, bit for bit, in every sampling mode;
  - two calls on two streams == the same calls one after the other.
"""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import pyoracle as O
import nerf_rs_b200 as nb

from .conftest import CAM_JSON as CAM, DATA

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _psnr(a, b):
    return 10 * np.log10(1.0 / max(float(np.mean((a.astype(np.float64) - b) ** 2)), 1e-20))


def _frame800(name="frame_800.npz"):
    """tests/golden/frame_800*.npz (made by tests/golden/make_frame800.py): background mask + the other pixels."""
    z = np.load(os.path.join(HERE, "golden", name))
    n = int(z["W"]) * int(z["H"])
    white = np.unpackbits(z["white"])[:n].astype(bool)
    frame = np.ones((n, 3), np.float32)
    frame[~white] = z["rgb"]
    return frame


# Pixels of the 800x800 frame that may differ from the oracle by more than 2/255, per mode.  The reference's pipeline is
# discontinuous in its own rounding (a fine sample falls on one side or the other of a sliver of density, DESIGN.md
# 4.3), so even the order-faithful f32 kernel - same sums in the same order, sinf/expf within 2 ulp of glibc - may
# land a pixel elsewhere; the bounds below are what was measured plus slack, and the single-pass modes are recorded, not
# hidden.
FRAME_BOUNDS = {  # mode: (pixels > 2/255 allowed, PSNR floor in dB)
    "fp32": (2, 100.0), "fp16x3": (2, 95.0), "mixed": (2, 80.0), "fp16": (300, 60.0), "tf32": (300, 60.0),
    "bf16": (4000, 50.0)}  # measured: 0 / 0 / 0 / 100 / (tf32: as fp16) / 1 673 pixels; 119 / 108 / 90 / 71 / 59 dB


@pytest.mark.parametrize("prec", ["fp32", "fp16x3", "mixed", "fp16", "tf32", "bf16"])
def test_full_800_frame_against_the_oracle(renderer, cam, prec):
    want = _frame800()
    got = renderer.render_image(cam, 800, 800, 64, 128, precision=prec)
    assert got.shape == want.shape and np.isfinite(got).all()
    d = np.abs(got - want).max(axis=1)
    bad = int((d > 2.0 / 255.0).sum())
    psnr = _psnr(got, want)
    print(f"800x800 frame vs oracle, {prec}: max {d.max() * 255:.3f}/255, pixels > 2/255: {bad} of {d.size}, "
          f"> 1/255: {int((d > 1.0 / 255.0).sum())}, PSNR {psnr:.2f} dB")
    allowed, floor = FRAME_BOUNDS[prec]
    assert psnr >= 45.0 and psnr >= floor, (prec, psnr)
    assert bad <= allowed, (prec, bad, float(d.max()) * 255)
    # background (exactly 1.0 = zero density along the whole ray) where the oracle has it, up to rays that graze a
    # density of ~1e-7 on one side only (0.4 % of the pixels, all of them within 1e-3 of white on both sides)
    if prec in ("fp32", "fp16x3", "mixed"):
        white_o = np.all(want == 1.0, axis=1)
        white_g = np.all(got == 1.0, axis=1)
        flip = white_o != white_g
        assert flip.mean() < 1e-2 and (not flip.any() or d[flip].max() < 1e-3)


@pytest.mark.parametrize("prec,tol", [("mixed", 2.0), ("fp32", 1.0)])
def test_philox_default_sampling_at_reference_counts(renderer, cam, onets, ocam, prec, tol):
    """The CLI's and the mirrors' default: Philox jitter, 64 + 128 samples, mixed precision.  Jittered coarse samples
    and random u make the merge unsorted-by-construction paths run (bitonic sort of the fine samples)."""
    W = H = 64
    want = O.render(onets[0], onets[1], ocam, W, H, 64, 128, mode=O.SAMPLING_PHILOX, seed=5)
    got = renderer.render_image(cam, W, H, 64, 128, precision=prec, mode=nb.SAMPLING_PHILOX, seed=5)
    d = np.abs(got - want).max(axis=1)
    print(f"philox 64x64 {prec}: max {d.max() * 255:.3f}/255, PSNR {_psnr(got, want):.1f} dB")
    assert d.max() * 255 <= tol and _psnr(got, want) >= 60.0
    other = renderer.render_image(cam, W, H, 64, 128, precision=prec, mode=nb.SAMPLING_PHILOX, seed=6)
    assert np.abs(other - got).max() > 1e-4  # the seed matters


def _fresh(env):
    keys = ("NERF_B200_COARSE_SELECT", "NERF_B200_CHUNK_RAYS", "NERF_B200_SELECT_MARGIN", "NERF_B200_SELECT_TAU",
            "NERF_B200_SKIP_EMPTY", "NERF_B200_COARSE_STAGES", "NERF_B200_TWO_PHASE", "NERF_B200_FINE_SUB_RAYS")
    saved = {k: os.environ.pop(k, None) for k in keys}
    os.environ.update(env)
    try:
        return nb.Renderer(DATA, device=0)
    finally:
        for k in keys:
            os.environ.pop(k, None)
            if saved[k] is not None:
                os.environ[k] = saved[k]


def test_two_pass_coarse_equals_split_operands_everywhere(renderer, cam):
    """mixed mode: fp16 over every coarse sample + split operands over the samples that can reach a coarse weight must
    give the frame the split-operand kernel over EVERY coarse sample gives - same fine sample positions, same pixels."""
    full = _fresh({"NERF_B200_COARSE_SELECT": "0"})
    try:
        W = H = 200
        for mode, seed in ((nb.SAMPLING_DET, 0), (nb.SAMPLING_PHILOX, 3)):
            a, sa = full.render_image(cam, W, H, 64, 128, precision="mixed", mode=mode, seed=seed, stages=True)
            b, sb = renderer.render_image(cam, W, H, 64, 128, precision="mixed", mode=mode, seed=seed, stages=True)
            assert (sa["w_c"] == sb["w_c"]).all() and (sa["t_f"] == sb["t_f"]).all() and (a == b).all()
            t = renderer.timing()
            assert t["coarse_samples"] == W * H * 64 and 0 < t["coarse_refined"] < 0.5 * t["coarse_samples"]
            # the densities that were not re-evaluated are zero or lie behind an opaque stretch of the ray
            diff = sa["sigma_c"] != sb["sigma_c"]
            assert (sb["w_c"][diff] == 0).all()
        # ray ranges of the two-pass path equal the same rows of its full frame (what multi-GPU sharding relies on)
        det = renderer.render_image(cam, W, H, 64, 128, precision="mixed")
        part = renderer.render_image(cam, W, H, 64, 128, precision="mixed", ray_begin=12345, ray_end=23456)
        assert (part == det[12345:23456]).all()
    finally:
        full.close()


def test_chunked_render_equals_single_chunk(renderer, cam):
    """Frames above chunk_rays rays (every 1600x1600 orbit frame) are rendered in chunks; offsets of the jitter / u
    arrays, Philox pixel indices and output rows must not depend on the chunking (chunk of 40 rays, not a divisor)."""
    small = _fresh({"NERF_B200_CHUNK_RAYS": "40"})
    try:
        W, H, nc, nf = 24, 16, 16, 24
        n = W * H
        rng = np.random.default_rng(8)
        jit = rng.uniform(0, 1, (n, nc)).astype(np.float32)
        u = rng.uniform(0, 1, (n, nf)).astype(np.float32)
        for prec in ("fp32", "mixed"):
            for kw in (dict(mode=nb.SAMPLING_DET), dict(mode=nb.SAMPLING_PHILOX, seed=9),
                       dict(mode=nb.SAMPLING_ARRAYS, jitter=jit, u=u)):
                a = renderer.render_image(cam, W, H, nc, nf, precision=prec, **kw)
                b = small.render_image(cam, W, H, nc, nf, precision=prec, **kw)
                assert (a == b).all(), (prec, kw["mode"])
                a = renderer.render_image(cam, W, H, nc, nf, precision=prec, ray_begin=31, ray_end=333, **(
                    kw if kw["mode"] != nb.SAMPLING_ARRAYS else dict(mode=nb.SAMPLING_ARRAYS, jitter=jit[31:333], u=u[31:333])))
                b = small.render_image(cam, W, H, nc, nf, precision=prec, ray_begin=31, ray_end=333, **(
                    kw if kw["mode"] != nb.SAMPLING_ARRAYS else dict(mode=nb.SAMPLING_ARRAYS, jitter=jit[31:333], u=u[31:333])))
                assert (a == b).all(), (prec, kw["mode"], "range")
        # the reference's sample counts: chunks together with the staged coarse pass (per-chunk ray lists and counters)
        # and the fixed-size importance kernel
        for kw in (dict(mode=nb.SAMPLING_DET), dict(mode=nb.SAMPLING_PHILOX, seed=9)):
            a = renderer.render_image(cam, W, H, 64, 128, precision="mixed", **kw)
            b = small.render_image(cam, W, H, 64, 128, precision="mixed", **kw)
            assert (a == b).all(), ("64+128", kw["mode"])
        ra, da, aa = renderer.render_aux(cam, W, H, nc, nf, precision="fp32")
        rb, db, ab = small.render_aux(cam, W, H, nc, nf, precision="fp32")
        assert (ra == rb).all() and (da == db).all() and (aa == ab).all()
        assert (small.render_rgb8(cam, W, H, nc, nf, precision="fp16") == renderer.render_rgb8(cam, W, H, nc, nf, precision="fp16")).all()
    finally:
        small.close()


def test_two_streams_on_one_context_do_not_overlap(renderer, cam):
    """A context has one set of workspaces: a *_device call on a second stream must wait (on the device) for the call
    before it.  Two different renders enqueued back to back on two streams, then compared with the serial results."""
    import torch
    W, H = 96, 64
    n = W * H
    want_a = renderer.render_image(cam, W, H, 64, 128, precision="mixed")
    want_b = renderer.render_image(cam, W, H, 64, 128, precision="mixed", mode=nb.SAMPLING_PHILOX, seed=2)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    out_a = torch.zeros(n, 3, device="cuda:0")
    out_b = torch.zeros(n, 3, device="cuda:0")
    torch.cuda.synchronize()
    for _ in range(3):
        renderer.render_device(cam, W, H, out_a.data_ptr(), 64, 128, precision="mixed", stream=s1.cuda_stream)
        renderer.render_device(cam, W, H, out_b.data_ptr(), 64, 128, precision="mixed", mode=nb.SAMPLING_PHILOX, seed=2,
                               stream=s2.cuda_stream)
    torch.cuda.synchronize()
    assert (out_a.cpu().numpy() == want_a).all() and (out_b.cpu().numpy() == want_b).all()


@pytest.mark.parametrize("prec", ["mixed", "fp32"])
def test_pr1_reference_output_ppm(renderer, cam, prec):
    """BASELINE.json configs[0]: the reference's own run (render_cli_image, lib.rs:647-677: 256x256, 64 + 128 samples,
    output.ppm) as the CPU oracle wrote it with deterministic sampling (scripts/pr1_reference.py ->
    tests/golden/pr1_output_det_256.ppm).  The GPU frame through the device-side save_ppm quantiser: same header, every
    byte within 1 of the oracle's, almost all identical."""
    want = open(os.path.join(HERE, "golden", "pr1_output_det_256.ppm"), "rb").read()
    head = b"P6\n256 256\n255\n"
    assert want.startswith(head) and len(want) == len(head) + 256 * 256 * 3
    got = renderer.render_rgb8(cam, 256, 256, 64, 128, precision=prec).tobytes()
    a = np.frombuffer(want[len(head):], np.uint8).astype(np.int16)
    b = np.frombuffer(got, np.uint8).astype(np.int16)
    d = np.abs(a - b)
    print(f"PR1 256x256 output.ppm, {prec}: bytes differing {int((d > 0).sum())} of {d.size}, max {int(d.max())}")
    assert d.max() <= 2 and (d > 1).sum() <= 3 and (d > 0).mean() < 0.02


@pytest.mark.parametrize("prec", ["mixed", "fp16", "bf16"])
def test_empty_sample_colour_skip_is_invisible(renderer, cam, prec):
    """The fine pass does not evaluate the colour branch where a whole 256-sample slot has no density (a sample with
    sigma = 0 has weight exactly 0 in integrate_ray, lib.rs:176-195): frames, depth and opacity must be bit-identical to
    the pass that evaluates every colour, and a caller who asks for the rgb stage still gets every sample's colour."""
    every = _fresh({"NERF_B200_SKIP_EMPTY": "0"})
    try:
        W, H = 160, 120
        for kw in (dict(mode=nb.SAMPLING_DET), dict(mode=nb.SAMPLING_PHILOX, seed=12)):
            a = every.render_image(cam, W, H, 64, 128, precision=prec, **kw)
            b = renderer.render_image(cam, W, H, 64, 128, precision=prec, **kw)
            assert (a == b).all() and np.isfinite(b).all()
        assert 0.2 < float(np.all(b == 1.0, axis=1).mean()) < 0.95  # the frame has background to skip and an object not to
        ra, da, aa = every.render_aux(cam, W, H, 64, 128, precision=prec)
        rb, db, ab = renderer.render_aux(cam, W, H, 64, 128, precision=prec)
        assert (ra == rb).all() and (da == db).all() and (aa == ab).all()
        part = renderer.render_image(cam, W, H, 64, 128, precision=prec, ray_begin=777, ray_end=9999)
        assert (part == renderer.render_image(cam, W, H, 64, 128, precision=prec)[777:9999]).all()
        # the stage dump disables the skip: colours of empty samples are the network's, as in the other context
        _, sa = every.render_image(cam, 48, 32, 64, 128, precision=prec, stages=True)
        _, sb = renderer.render_image(cam, 48, 32, 64, 128, precision=prec, stages=True)
        assert (sa["rgb_f"] == sb["rgb_f"]).all() and (sa["sigma_f"] == sb["sigma_f"]).all()
        assert (sb["rgb_f"][sb["sigma_f"] == 0] > 0).all()  # sigmoid outputs, not the skip's zeros
    finally:
        every.close()


def test_striped_views_assemble_the_frame_bit_for_bit(renderer, cam):
    """nerf_render_striped(_device): the frame dealt out in stripes to N renderers and rendered part by part equals the
    whole-frame render bit for bit - host frame and device frame, deterministic and Philox sampling, stripes that do
    not divide the frame - and each part only touches its own pixels."""
    import torch
    from nerf_rs_b200 import shard
    W, H = 64, 40
    n = W * H
    for kw in (dict(mode=nb.SAMPLING_DET), dict(mode=nb.SAMPLING_PHILOX, seed=21)):
        want = renderer.render_image(cam, W, H, 64, 128, precision="mixed", **kw)
        for stripe, parts in ((W, 4), (3 * W + 5, 3), (n, 2), (17, 5)):
            host = np.full((n, 3), -7.0, np.float32)
            dev = torch.full((n, 3), -7.0, device="cuda:0")
            for part in range(parts):
                before = host.copy()
                renderer.render_striped(cam, W, H, host, stripe, parts, part, precision="mixed", **kw)
                renderer.render_striped_device(cam, W, H, dev.data_ptr(), stripe, parts, part, precision="mixed", **kw)
                mine = shard.view_pixels(part, parts, stripe, n)
                other = np.setdiff1d(np.arange(n), mine)
                assert (host[other] == before[other]).all() and (host[mine] == want[mine]).all(), (stripe, parts, part)
            torch.cuda.synchronize()
            assert (host == want).all() and (dev.cpu().numpy() == want).all(), (stripe, parts)
    with pytest.raises(nb.NerfError):
        renderer.render_striped(cam, W, H, np.zeros((n, 3), np.float32), 0, 2, 0)
    with pytest.raises(nb.NerfError):
        renderer.render_striped(cam, W, H, np.zeros((n, 3), np.float32), W, 2, 2)


def test_staged_coarse_pass_is_invisible(renderer, cam):
    """mixed mode: the fp16 coarse pass runs front to back in slices of the ray and later slices skip the rays that have
    gone opaque (optical depth beyond tau_cut: every weight behind is exactly 0, lib.rs:276-279).  Fine sample positions,
    coarse weights and pixels must not depend on the number of slices - one launch over every sample included."""
    W = H = 200
    want = {}
    for stages in ("1", "2", "4", "8"):
        r = _fresh({"NERF_B200_COARSE_STAGES": stages})
        try:
            for mode, seed in ((nb.SAMPLING_DET, 0), (nb.SAMPLING_PHILOX, 17)):
                img, st = r.render_image(cam, W, H, 64, 128, precision="mixed", mode=mode, seed=seed, stages=True)
                assert np.isfinite(img).all() and np.isfinite(st["sigma_c"]).all() and (st["sigma_c"] >= 0).all()
                if stages == "1":
                    want[mode] = (img, st["w_c"], st["t_f"])
                else:
                    assert (img == want[mode][0]).all() and (st["w_c"] == want[mode][1]).all(), stages
                    assert (st["t_f"] == want[mode][2]).all(), stages
            # sample counts the slices do not divide fall back to one launch and still render
            odd = r.render_image(cam, 32, 24, 20, 24, precision="mixed")
            assert (odd == renderer.render_image(cam, 32, 24, 20, 24, precision="mixed")).all()
        finally:
            r.close()
    assert (renderer.render_image(cam, W, H, 64, 128, precision="mixed") == want[nb.SAMPLING_DET][0]).all()


def test_fixed_size_importance_kernel_equals_the_generic_one(renderer, cam):
    """64 + 128 samples (the reference's counts) run importance_fixed_kernel: unrolled folds, verified rank hints,
    slot-ordered stores.  NERF_B200_GENERIC_IMPORTANCE=1 forces the generic kernel on the same context: fine sample
    positions must agree bit for bit - sorted input (fast merge) and unsorted input (sort + rank merge) alike - and the
    oracle must agree with both."""
    def both(fn):
        os.environ.pop("NERF_B200_GENERIC_IMPORTANCE", None)
        fixed = fn()
        os.environ["NERF_B200_GENERIC_IMPORTANCE"] = "1"
        try:
            return fixed, fn()
        finally:
            os.environ.pop("NERF_B200_GENERIC_IMPORTANCE", None)

    W = H = 160
    for kw in (dict(mode=nb.SAMPLING_DET), dict(mode=nb.SAMPLING_PHILOX, seed=23)):
        (a, sa), (b, sb) = both(lambda: renderer.render_image(cam, W, H, 64, 128, precision="mixed", stages=True, **kw))
        assert (sa["t_f"] == sb["t_f"]).all() and (sa["w_c"] == sb["w_c"]).all() and (a == b).all(), kw
        assert (np.diff(sa["t_f"], axis=1) >= 0).all()
    # the stage entry point: regular u (fast merge), random u (fine samples unsorted), duplicates and ties, empty
    # weights (flat cdf), and coarse samples that arrive unsorted
    rng = np.random.default_rng(41)
    n, cnt, rays = 64, 128, 300
    t = np.sort(rng.uniform(2, 6, (rays, n)).astype(np.float32), axis=1)
    w = (rng.uniform(0, 1, (rays, n)) ** 6).astype(np.float32)
    w[:7] = 0
    w[7:14, 10:] = 0
    grid = np.tile(((np.arange(cnt) + 0.5) / cnt).astype(np.float32), (rays, 1))
    dup = np.repeat(rng.uniform(0, 1, (rays, cnt // 4)).astype(np.float32), 4, axis=1)
    shuffled = t.copy()
    shuffled[::3] = shuffled[::3, ::-1]
    cases = [(t, grid), (t, rng.uniform(0, 1, (rays, cnt)).astype(np.float32)), (t, np.sort(dup, axis=1)), (t, dup),
             (shuffled, grid)]
    for tt, u in cases:
        got, gen = both(lambda: renderer.sample_importance(tt, w, u))
        assert (got == gen).all()
        want = np.stack([np.sort(np.concatenate([tt[i], O.sample_importance(tt[i], w[i], u[i])]), kind="stable")
                         for i in range(0, rays, 5)])
        assert np.abs(got[::5] - want).max() < 2e-6 and (np.diff(got, axis=1) >= 0).all()


@pytest.mark.parametrize("prec", ["mixed", "fp32"])
def test_full_800_frame_with_philox_jitter_against_the_oracle(renderer, cam, prec):
    """The product default - Philox jitter seed 0 (CLI, C++ / Python / Rust mirrors), mixed precision - on all 640 000
    pixels of the headline frame against the oracle with the same counter-based stream (frame_800_philox0.npz).  Jittered
    samples take the unsorted paths of the sampler (sort + rank merge) at full size."""
    want = _frame800("frame_800_philox0.npz")
    got = renderer.render_image(cam, 800, 800, 64, 128, precision=prec, mode=nb.SAMPLING_PHILOX, seed=0)
    d = np.abs(got - want).max(axis=1)
    bad = int((d > 2.0 / 255.0).sum())
    psnr = _psnr(got, want)
    print(f"800x800 Philox frame vs oracle, {prec}: max {d.max() * 255:.3f}/255, pixels > 2/255: {bad} of {d.size}, "
          f"> 1/255: {int((d > 1.0 / 255.0).sum())}, PSNR {psnr:.2f} dB")
    assert np.isfinite(got).all() and bad <= 2 and psnr >= (95.0 if prec == "fp32" else 80.0), (prec, bad, psnr)


def test_stratified_samples_bit_exact_for_every_count_and_mode(renderer, cam):
    """K1 has a kernel for sample counts that are a multiple of four (four samples per thread, one Philox block, one
    128-bit store) and the per-sample kernel for the others: both against the oracle's stratified_samples
    (src/lib.rs:233-248) and Philox stream, bit for bit, on a ray range that does not start at pixel 0."""
    W, H = 16, 8
    b, e = 5, 117
    n = e - b
    rng = np.random.default_rng(77)
    for nc in (64, 12, 4, 10, 7, 3):
        jit = rng.uniform(0, 1, (n, nc)).astype(np.float32)
        u = rng.uniform(0, 1, (n, 4)).astype(np.float32)
        for kw, jitter_of in ((dict(mode=nb.SAMPLING_DET), lambda r: np.full(nc, 0.5, np.float32)),
                              (dict(mode=nb.SAMPLING_PHILOX, seed=(5 << 32) | 9),
                               lambda r: O.philox_uniforms((5 << 32) | 9, b + r, 0, nc)),
                              (dict(mode=nb.SAMPLING_ARRAYS, jitter=jit, u=u), lambda r: jit[r])):
            _, st = renderer.render_image(cam, W, H, nc, 4, precision="fp16", ray_begin=b, ray_end=e, stages=True, **kw)
            want = np.stack([O.stratified(cam.near, cam.far, nc, jitter_of(r)) for r in range(n)])
            assert (st["t_c"] == want).all(), (nc, kw["mode"])
            assert np.abs(st["dirs"] - O.ray_dirs(O.camera_from_json(CAM), W, H, ray_begin=b, ray_end=e)).max() <= 1.2e-7


@pytest.mark.parametrize("prec", ["mixed", "fp16", "bf16"])
def test_two_phase_fine_pass_is_invisible(renderer, cam, prec):
    """The fine pass runs as two launches per group of rays (density for every sample; the colour branch over the
    samples that have density): frames, depth and opacity bit-identical to the fused launch
    (NERF_B200_TWO_PHASE=0), with groups of 700 rays (ragged last group, groups without any density) as with one group,
    and together with the opt-in skipping of rays."""
    fused = _fresh({"NERF_B200_TWO_PHASE": "0"})
    groups = _fresh({"NERF_B200_TWO_PHASE": "1", "NERF_B200_FINE_SUB_RAYS": "700"})
    try:
        W, H = 160, 120
        for kw in (dict(mode=nb.SAMPLING_DET), dict(mode=nb.SAMPLING_PHILOX, seed=12)):
            a = fused.render_image(cam, W, H, 64, 128, precision=prec, **kw)
            n_fused = fused.timing()["kernel_launches"]
            b = renderer.render_image(cam, W, H, 64, 128, precision=prec, **kw)
            n_two = renderer.timing()["kernel_launches"]
            assert n_two == n_fused + 1  # one group: the fused launch became two
            c = groups.render_image(cam, W, H, 64, 128, precision=prec, **kw)
            assert groups.timing()["kernel_launches"] > n_two  # 28 groups of two launches
            assert (a == b).all() and (a == c).all() and np.isfinite(b).all()
        ra, da, aa = fused.render_aux(cam, W, H, 64, 128, precision=prec)
        for r in (renderer, groups):
            rb, db, ab = r.render_aux(cam, W, H, 64, 128, precision=prec)
            assert (ra == rb).all() and (da == db).all() and (aa == ab).all()
        # other sample counts (192 is the only one with kernels of its own elsewhere in the pipeline)
        for nc, nf in ((16, 24), (64, 0), (32, 96)):
            assert (fused.render_image(cam, 64, 48, nc, nf, precision=prec) ==
                    groups.render_image(cam, 64, 48, nc, nf, precision=prec)).all(), (nc, nf)
        # with rays skipped (opt-in): the list of rays and the list of samples compose
        for r in (fused, groups):
            r.set_option(nb.OPT_SKIP_EMPTY_RAYS, 1)
        assert (fused.render_image(cam, W, H, 64, 128, precision=prec) ==
                groups.render_image(cam, W, H, 64, 128, precision=prec)).all()
    finally:
        fused.close()
        groups.close()
