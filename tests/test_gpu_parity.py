"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the same inputs.

Tolerances.  fp32 mode: the GEMMs are order-faithful, only sinf/cosf/expf differ from glibc by <= 2 ulp, so
sigma/rgb agree to ~1e-4 relative.  fp16x3 mode (split fp16 operands, three tensor-core MMAs per product) carries
~22-bit operands: it is held to the north star's ABSOLUTE bounds (sigma and rgb <= 1e-3 on the golden samples, frame
<= 2/255 and >= 45 dB) and measured at ~2e-6 relative on sigma.  fp16/bf16 modes round the 256-wide GEMM operands to 11/8 significant
bits; an absolute sigma bound is meaningless at sigma ~ 200, so sigma is checked as
|d| <= tol * max(1, |sigma|) and rgb as an absolute bound, with tol from the operand precision."""
import numpy as np
import pytest

from oracle import pyoracle as O
import nerf_rs_b200 as nb

pytestmark = pytest.mark.gpu

SIG_TOL = {"fp32": 2e-4, "fp16": 4e-2, "bf16": 3e-1, "tf32": 4e-2, "fp16x3": 3e-4}  # max: single samples at a ReLU knee
SIG_P99 = {"fp32": 1e-4, "fp16": 5e-3, "bf16": 4e-2, "tf32": 5e-3, "fp16x3": 1e-4}  # 99th percentile of the same error
RGB_TOL = {"fp32": 2e-5, "fp16": 6e-3, "bf16": 6e-2, "tf32": 6e-3, "fp16x3": 1e-4}
ROUND = {"fp16": O.ROUND_FP16, "bf16": O.ROUND_BF16, "tf32": O.ROUND_TF32}


def _golden_inputs(golden):
    o = np.array(golden["camera_origin"], np.float32)
    z = np.array(golden["z_vals"], np.float32)
    P, D = [], []
    for ex in golden["examples"]:
        rd = np.array(ex["ray_d"], np.float32)
        P.append((o[:, None] + rd[:, None] * z[None, :]).astype(np.float32))
        D.append(np.tile(np.array(ex["viewdir_unit"], np.float32), (len(z), 1)))
    return np.concatenate(P, axis=1), np.concatenate(D, axis=0)


def _random_samples(ocam, n, seed):
    rng = np.random.default_rng(seed)
    dirs = O.ray_dirs(ocam, 800, 800)[rng.integers(0, 640000, n)]
    t = rng.uniform(2.0, 6.0, n).astype(np.float32)
    pts = (np.array(ocam.pos, np.float32)[:, None] + dirs.T * t[None, :]).astype(np.float32)
    return pts, dirs


def _check(prec, rgb, sig, rgb_ref, sig_ref, scale=1.0):
    ds = np.abs(sig - sig_ref) / np.maximum(1.0, np.abs(sig_ref))
    dc = np.abs(rgb - rgb_ref)
    assert np.isfinite(rgb).all() and np.isfinite(sig).all()
    assert ds.max() <= SIG_TOL[prec] * scale, (prec, "sigma", float(ds.max()))
    if ds.size >= 1000:  # isolated samples near the ReLU knee carry the max; the bulk is an order tighter
        assert np.percentile(ds, 99) <= SIG_P99[prec] * scale, (prec, "sigma p99", float(np.percentile(ds, 99)))
    assert dc.max() <= RGB_TOL[prec] * scale, (prec, "rgb", float(dc.max()))


@pytest.mark.parametrize("prec", ["fp32", "fp16", "bf16", "tf32", "fp16x3"])
def test_forward_batch_golden_json(renderer, golden, prec):
    """The reference's own unit test (lib.rs:753-916) against the GPU path: 1e-2 abs is its tolerance; sigma in the
    16-bit modes is held to the relative bound above instead (sigma reaches 213 here)."""
    pts, dirs = _golden_inputs(golden)
    for net, key in ((renderer.coarse, "coarse"), (renderer.fine, "fine")):
        rgb, sig = net.forward_batch(pts, dirs, precision=prec)
        want_s = np.concatenate([ex[key + "_sigma"] for ex in golden["examples"]]).astype(np.float32)
        want_c = np.concatenate([ex[key + "_rgb"] for ex in golden["examples"]]).astype(np.float32)
        assert np.abs(rgb - want_c).max() < 1e-2
        if prec == "fp32":
            assert np.abs(sig - want_s).max() < 1e-2
        if prec == "fp16x3":  # north star, "tf32 mode": <= 1e-3 abs on sigma and rgb, met on the tensor cores
            assert np.abs(sig - want_s).max() <= 1e-3 and np.abs(rgb - want_c).max() <= 1e-3
        _check(prec, rgb, sig, want_c, want_s)


@pytest.mark.parametrize("prec", ["fp32", "fp16", "bf16", "tf32", "fp16x3"])
@pytest.mark.parametrize("n", [1, 127, 128, 129, 255, 257, 1000, 4096 + 77])
def test_forward_batch_random_vs_oracle(renderer, onets, ocam, prec, n):
    """Ragged batch sizes around the 64 / 128-sample tiles, both networks."""
    pts, dirs = _random_samples(ocam, n, seed=n)
    for net, onet in ((renderer.coarse, onets[0]), (renderer.fine, onets[1])):
        rgb, sig = net.forward_batch(pts, dirs, precision=prec)
        rgb_ref, sig_ref = onet.forward_batch(pts, dirs)
        _check(prec, rgb, sig, rgb_ref, sig_ref)


@pytest.mark.parametrize("prec", ["fp16", "bf16", "tf32"])
def test_tensor_core_matches_rounding_emulation(renderer, onets, ocam, prec):
    """Against the oracle run with the same operand rounding the kernel applies, the only difference left is
    the f32 accumulation order inside the tensor core: an order of magnitude tighter than against exact f32."""
    pts, dirs = _random_samples(ocam, 2048, seed=5)
    rgb, sig = renderer.fine.forward_batch(pts, dirs, precision=prec)
    rgb_e, sig_e = onets[1].forward_batch(pts, dirs, round_mode=ROUND[prec])
    rgb_x, sig_x = onets[1].forward_batch(pts, dirs)
    err_emul = np.abs(sig - sig_e) / np.maximum(1, np.abs(sig_e))
    err_exact = np.abs(sig - sig_x) / np.maximum(1, np.abs(sig_x))
    assert np.median(err_emul) <= np.median(err_exact) + 1e-7
    _check(prec, rgb, sig, rgb_e, sig_e, scale=0.5)


def test_forward_batch_empty_and_errors(renderer):
    import nerf_rs_b200 as nb
    rgb, sig = renderer.coarse.forward_batch(np.zeros((3, 0), np.float32), np.zeros((0, 3), np.float32))
    assert rgb.shape == (0, 3) and sig.shape == (0,)
    with pytest.raises(ValueError):
        renderer.coarse.forward_batch(np.zeros((3, 4), np.float32), np.zeros((3, 3), np.float32))
    with pytest.raises(nb.NerfError):
        renderer.coarse.forward_batch(np.zeros((3, 4), np.float32), np.zeros((4, 3), np.float32), precision=9)
    # NERF_PREC_TF32 is a mode of its own (tcgen05 kind::tf32, mlp_tc_tf32.cu), not an alias of fp16 as in round 1: the
    # two agree to operand rounding but not bit for bit (other K-blocking, other accumulation order)
    rng = np.random.default_rng(5)
    pts = rng.uniform(-1.2, 1.2, (3, 600)).astype(np.float32)
    dirs = rng.normal(size=(600, 3)).astype(np.float32)
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    ct, st = renderer.fine.forward_batch(pts, dirs, precision="tf32")
    ch, sh = renderer.fine.forward_batch(pts, dirs, precision="fp16")
    assert np.isfinite(ct).all() and np.isfinite(st).all() and st.max() > 1.0
    assert np.abs(ct - ch).max() < 6e-3 and (np.abs(st - sh) <= 4e-2 * np.maximum(1, np.abs(sh))).all()
    assert not (np.array_equal(ct, ch) and np.array_equal(st, sh))


# ----------------------------------------------------------------------------------------- stage-wise parity
def test_ray_dirs(renderer, cam, ocam):
    for W, H in ((8, 8), (800, 800), (256, 64)):
        got = renderer.ray_dirs(cam, W, H)
        want = O.ray_dirs(ocam, W, H)
        assert np.abs(got - want).max() <= 1.2e-7  # IEEE div/sqrt, no FMA: bit-exact up to 1 ulp of the final div


def test_compute_weights(renderer):
    rng = np.random.default_rng(1)
    for n in (1, 5, 64, 192, 200):
        t = np.sort(rng.uniform(2, 6, (300, n)).astype(np.float32), axis=1)
        sig = (rng.uniform(0, 1, (300, n)) ** 8 * 300).astype(np.float32)
        got = renderer.compute_weights(sig, t, 6.0)
        want = np.stack([O.compute_weights(sig[i], t[i], 6.0) for i in range(300)])
        assert np.abs(got - want).max() < 2e-6
        # early-exit mask (lib.rs:276-279): past the T < 1e-4 cut-off both sides are exactly 0.  Weights below
        # 1e-6 are excluded: 1 - expf(-tiny) is 0 or 6e-8 depending on the last ulp of expf.
        big = np.maximum(np.abs(want), np.abs(got)) > 1e-6
        assert ((want == 0) == (got == 0))[big].all()
        assert (got[want == 0] < 1e-6).all()


def test_sample_importance_and_sort(renderer):
    rng = np.random.default_rng(2)
    n, cnt, rays = 64, 128, 200
    t = np.sort(rng.uniform(2, 6, (rays, n)).astype(np.float32), axis=1)
    w = (rng.uniform(0, 1, (rays, n)) ** 6).astype(np.float32)
    w[:5] = 0
    for u in (np.tile(((np.arange(cnt) + 0.5) / cnt).astype(np.float32), (rays, 1)),
              rng.uniform(0, 1, (rays, cnt)).astype(np.float32)):
        got = renderer.sample_importance(t, w, u)
        want = np.stack([np.sort(np.concatenate([t[i], O.sample_importance(t[i], w[i], u[i])]), kind="stable")
                         for i in range(rays)])
        assert (np.diff(got, axis=1) >= 0).all()
        assert np.abs(got - want).max() < 2e-6


def test_integrate_rays(renderer):
    rng = np.random.default_rng(3)
    n, rays = 192, 300
    t = np.sort(rng.uniform(2, 6, (rays, n)).astype(np.float32), axis=1)
    sig = (rng.uniform(0, 1, (rays, n)) ** 8 * 200).astype(np.float32)
    rgb = rng.uniform(0, 1, (rays, n, 3)).astype(np.float32)
    got = renderer.integrate_rays(rgb, sig, t, 6.0)
    want = np.stack([O.integrate_ray(rgb[i], sig[i], t[i], 6.0) for i in range(rays)])
    assert np.abs(got - want).max() < 3e-6


# ----------------------------------------------------------------------------------------- end to end
def _psnr(a, b):
    return 10 * np.log10(1.0 / max(float(np.mean((a - b) ** 2)), 1e-20))


def test_render_fp32_parity_with_stages(renderer, cam, onets, ocam):
    W = H = 24
    want, wst = O.render(onets[0], onets[1], ocam, W, H, 64, 128, stages=True)
    got, gst = renderer.render_image(cam, W, H, 64, 128, precision="fp32", stages=True)
    assert np.abs(gst["dirs"] - wst["dirs"]).max() <= 1.2e-7
    assert np.abs(gst["t_c"] - wst["t_c"]).max() == 0.0  # no transcendental involved: bit-exact
    rel = np.abs(gst["sigma_c"] - wst["sigma_c"]) / np.maximum(1, np.abs(wst["sigma_c"]))
    assert rel.max() < 5e-4
    assert np.abs(gst["w_c"] - wst["w_c"]).max() < 5e-4
    assert np.abs(got - want).max() <= 2.0 / 255.0  # north-star bound; fp32 is far inside it
    assert np.abs(got - want).max() < 2e-3 and _psnr(got, want) > 70.0


@pytest.mark.parametrize("prec,psnr_min", [("fp16", 50.0), ("bf16", 45.0), ("tf32", 50.0)])
def test_render_tensor_core_psnr(renderer, cam, onets, ocam, prec, psnr_min):
    """North star: PSNR >= 45 dB against the reference render with identical non-jittered samples."""
    W = H = 32
    want = O.render(onets[0], onets[1], ocam, W, H, 64, 128)
    got = renderer.render_image(cam, W, H, 64, 128, precision=prec)
    assert np.isfinite(got).all()
    assert _psnr(got, want) >= psnr_min, (prec, _psnr(got, want), float(np.abs(got - want).max()) * 255)


def test_render_fp16x3_and_mixed_meet_the_image_bound(renderer, cam, onets, ocam):
    """North star: max per-pixel abs diff <= 2/255 and PSNR >= 45 dB against the reference render with identical,
    non-jittered sample positions - on the tensor cores.  Small frame against the CPU oracle; an 800-pixel-wide band
    through the object (rows 380..419 of the 800x800 frame) against the order-faithful f32 kernel, which the test
    above ties to the oracle."""
    W = H = 32
    want = O.render(onets[0], onets[1], ocam, W, H, 64, 128)
    got = renderer.render_image(cam, W, H, 64, 128, precision="fp16x3")
    assert np.abs(got - want).max() <= 2.0 / 255.0 and _psnr(got, want) >= 45.0
    assert np.abs(got - want).max() < 1e-3 and _psnr(got, want) > 90.0  # what it actually achieves
    W = H = 800
    rb, re_ = 380 * W, 420 * W
    f32 = renderer.render_image(cam, W, H, 64, 128, precision="fp32", ray_begin=rb, ray_end=re_)
    x3 = renderer.render_image(cam, W, H, 64, 128, precision="fp16x3", ray_begin=rb, ray_end=re_)
    assert f32.std() > 0.05  # the band crosses the object
    assert np.abs(x3 - f32).max() <= 2.0 / 255.0 and _psnr(x3, f32) >= 45.0
    assert _psnr(x3, f32) > 90.0
    # "mixed": only the coarse network (which places the fine samples) runs fp16x3, the fine network single-pass fp16
    mx = renderer.render_image(cam, W, H, 64, 128, precision="mixed", ray_begin=rb, ray_end=re_)
    assert np.abs(mx - f32).max() <= 2.0 / 255.0 and _psnr(mx, f32) >= 80.0
    # and that is where the 16-bit modes lose it: same fine network, single-pass coarse densities
    fp16 = renderer.render_image(cam, W, H, 64, 128, precision="fp16", ray_begin=rb, ray_end=re_)
    assert _psnr(fp16, f32) >= 50.0 and _psnr(mx, f32) > _psnr(fp16, f32) + 10.0


def test_render_modes_and_ranges(renderer, cam, onets, ocam):
    W, H, nc, nf = 16, 8, 16, 24
    n = W * H
    # Philox: same counter-based stream on both sides
    want = O.render(onets[0], onets[1], ocam, W, H, nc, nf, mode=O.SAMPLING_PHILOX, seed=11)
    got = renderer.render_image(cam, W, H, nc, nf, precision="fp32", mode=1, seed=11)
    assert np.abs(got - want).max() < 2e-3
    # caller-supplied uniforms
    rng = np.random.default_rng(4)
    jit = rng.uniform(0, 1, (n, nc)).astype(np.float32)
    u = rng.uniform(0, 1, (n, nf)).astype(np.float32)
    want = O.render(onets[0], onets[1], ocam, W, H, nc, nf, mode=O.SAMPLING_ARRAYS, jitter=jit, u=u)
    got = renderer.render_image(cam, W, H, nc, nf, precision="fp32", mode=2, jitter=jit, u=u)
    assert np.abs(got - want).max() < 2e-3
    # a ray range equals the same rows of the full frame, bit for bit (what multi-GPU sharding relies on)
    for prec in ("fp32", "fp16", "fp16x3"):
        full = renderer.render_image(cam, W, H, nc, nf, precision=prec)
        part = renderer.render_image(cam, W, H, nc, nf, precision=prec, ray_begin=37, ray_end=101)
        assert (part == full[37:101]).all()
    # coarse-only render (n_fine = 0)
    want = O.render(onets[0], onets[1], ocam, W, H, nc, 0)
    got = renderer.render_image(cam, W, H, nc, 0, precision="fp32")
    assert np.abs(got - want).max() < 2e-3


def test_render_argument_errors(renderer, cam):
    import nerf_rs_b200 as nb
    for W, H, nc in ((12, 8, 4), (8, 12, 4), (8, 8, 0)):
        with pytest.raises(nb.NerfError) as e:
            renderer.render_image(cam, W, H, nc, 4)
        assert e.value.code == -1
    assert renderer.render_image(cam, 8, 8, 4, 4, ray_begin=5, ray_end=5).shape == (0, 3)


def test_cli_writes_the_same_ppm(renderer, cam, tmp_path):
    """nerf_cli (main.rs / render_cli_image, lib.rs:647-677) == the library path, byte for byte, and within the
    north-star image bound of the oracle's PPM quantisation."""
    import os
    import subprocess
    import nerf_rs_b200 as nb
    cli = os.path.join(os.path.dirname(nb.LIB_PATH), "nerf_cli")
    out = tmp_path / "cli.ppm"
    r = subprocess.run([cli, "--weights", nb.DATA_ROOT, "--width", "32", "--height", "24", "--det", "--precision",
                        "fp32", "--out", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "Rendering completed in" in r.stdout and "Starting image rendering..." in r.stdout
    img = renderer.render_image(cam, 32, 24, 64, 128, precision="fp32")
    want = b"P6\n32 24\n255\n" + nb.quantize_rgb8(img).tobytes()
    assert out.read_bytes() == want
    # reference defaults: 256x256, 64 + 128 samples, output.ppm in the working directory (lib.rs:603-612,657-658,676)
    r = subprocess.run([cli, "--weights", nb.DATA_ROOT], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    data = (tmp_path / "output.ppm").read_bytes()
    assert data.startswith(b"P6\n256 256\n255\n") and len(data) == 15 + 256 * 256 * 3
    assert "Rendering with 64 coarse samples and 128 fine samples per ray" in r.stdout
    # --cache: first run writes the pre-swizzled weight cache, second run starts from it; same bytes as from the directory
    cache = tmp_path / "lego.nrfw"
    outs = []
    for k in range(2):
        o = tmp_path / f"cache{k}.ppm"
        r = subprocess.run([cli, "--weights", nb.DATA_ROOT, "--cache", str(cache), "--width", "32", "--height", "24", "--det",
                            "--out", str(o)], capture_output=True, text=True)
        assert r.returncode == 0 and cache.exists(), r.stderr
        outs.append(o.read_bytes())
    mixed = renderer.render_image(cam, 32, 24, 64, 128)  # the CLI's default precision is the library's: mixed
    assert outs[0] == outs[1] == b"P6\n32 24\n255\n" + nb.quantize_rgb8(mixed).tobytes()
    # --gpus N: one context and one host thread per device, contiguous pixel-index ranges (rayon's role, lib.rs:533-550)
    import torch
    n = min(torch.cuda.device_count(), 4)
    o = tmp_path / "multi.ppm"
    r = subprocess.run([cli, "--weights", nb.DATA_ROOT, "--gpus", str(n), "--width", "32", "--height", "24", "--det", "--out",
                        str(o)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert o.read_bytes() == outs[0] and r.stdout.count("device time") == n
    r = subprocess.run([cli, "--weights", nb.DATA_ROOT, "--gpus", str(torch.cuda.device_count() + 1)], capture_output=True,
                       text=True, cwd=tmp_path)
    assert r.returncode == 1 and "device ordinal out of range" in r.stderr


# ---------------------------------------------------------------- 8-bit output on the device
def test_quantiser_on_device_is_bit_exact(renderer):
    """save_ppm's (clamp(v,0,1)*255+0.5) as u8 (lib.rs:574-576) on the GPU == the oracle, byte for byte, including
    every rounding boundary k/255 +- 1 ulp, values outside [0,1], infinities and NaN (Rust: NaN as u8 == 0)."""
    rng = np.random.default_rng(7)
    edges = (np.arange(0, 256, dtype=np.float64)[:, None] + np.array([-0.5, 0.0, 0.5])[None, :]) / 255.0
    edges = edges.astype(np.float32).ravel()
    near = np.concatenate([np.nextafter(edges, np.float32(-1)), edges, np.nextafter(edges, np.float32(2))])
    special = np.array([np.nan, -np.nan, np.inf, -np.inf, -0.0, 0.0, 1.0, 1.0000001, -1e-30, 1e30, 0.5, 255.0], np.float32)
    v = np.concatenate([near, special, rng.uniform(-0.2, 1.2, 30000).astype(np.float32)])
    v = np.resize(v, (v.size // 3 + 1) * 3).astype(np.float32)
    want = O.quantize(v).reshape(-1, 3)
    got3 = renderer.quantize_rgb8_gpu(v, channels=3)
    assert (got3 == want).all()
    got4 = renderer.quantize_rgb8_gpu(v, channels=4)  # pixels_to_rgba (lib.rs:582-592)
    assert (got4[:, :3] == want).all() and (got4[:, 3] == 255).all()
    assert renderer.quantize_rgb8_gpu(np.zeros((0, 3), np.float32)).shape == (0, 3)


@pytest.mark.parametrize("channels", [3, 4])
def test_render_rgb8_equals_quantised_float_frame(renderer, cam, channels):
    W, H = 48, 32
    img = renderer.render_image(cam, W, H, 64, 128, precision="fp16")
    got = renderer.render_rgb8(cam, W, H, 64, 128, precision="fp16", channels=channels)
    assert got.shape == (W * H, channels)
    assert (got[:, :3] == O.quantize(img).reshape(-1, 3)).all()
    if channels == 4:
        assert (got[:, 3] == 255).all()
    part = renderer.render_rgb8(cam, W, H, 64, 128, precision="fp16", channels=channels, ray_begin=100, ray_end=777)
    assert (part == got[100:777]).all()
    with pytest.raises(nb.NerfError):
        renderer.render_rgb8(cam, W, H, 64, 128, channels=2)


def test_render_many_equals_single_views(renderer, cam):
    """A batch of poses (orbit sweep) returns, view by view, the bytes of the single-view call; Philox view v uses
    seed + v."""
    import copy
    W, H = 40, 24
    cams = []
    for k in range(5):
        c = copy.copy(cam)
        a = 0.3 * k
        o = np.array(cam.origin[:], np.float32)
        rot = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]], np.float32)
        c.origin[:] = (rot @ o).tolist()
        c.forward[:] = (rot @ np.array(cam.forward[:], np.float32)).tolist()
        c.up[:] = (rot @ np.array(cam.up[:], np.float32)).tolist()
        cams.append(c)
    many = renderer.render_many(cams, W, H, 16, 32, precision="fp16")
    assert many.shape == (5, W * H, 3)
    for k, c in enumerate(cams):
        assert (many[k] == renderer.render_rgb8(c, W, H, 16, 32, precision="fp16")).all()
    assert len({many[k].tobytes() for k in range(5)}) == 5  # the poses really differ
    ph = renderer.render_many(cams[:3], W, H, 16, 32, precision="fp16", mode=nb.SAMPLING_PHILOX, seed=11, channels=4)
    for k in range(3):
        one = renderer.render_rgb8(cams[k], W, H, 16, 32, precision="fp16", mode=nb.SAMPLING_PHILOX, seed=11 + k, channels=4)
        assert (ph[k] == one).all()
    assert renderer.render_many([], W, H, 16, 32).shape == (0, W * H, 3)


# ---------------------------------------------------------------- depth / accumulated opacity
def test_render_aux_depth_and_opacity(renderer, cam):
    """depth = sum w_i t_i and acc = sum w_i (oracle.cpp::oracle_integrate_aux; the reference has neither) on the
    GPU's own fine samples; rgb is unchanged by asking for them, and 1 - acc is integrate_ray's background term."""
    W, H, nc, nf = 24, 16, 16, 32
    rgb0, st = renderer.render_image(cam, W, H, nc, nf, precision="fp32", stages=True)
    rgb, depth, acc = renderer.render_aux(cam, W, H, nc, nf, precision="fp32")
    assert (rgb == rgb0).all()
    want = np.stack([O.integrate_aux(st["sigma_f"][r], st["t_f"][r], cam.far) for r in range(W * H)])
    assert np.abs(depth - want[:, 0]).max() <= 2e-5 * cam.far
    assert np.abs(acc - want[:, 1]).max() <= 2e-6
    assert (acc >= 0).all() and (acc <= 1 + 1e-6).all() and (depth >= 0).all() and (depth <= cam.far * (1 + 1e-6)).all()
    hit = acc > 0.5
    assert hit.any() and (~hit).any()  # the frame contains both object and background
    assert (depth[hit] / acc[hit] > cam.near).all() and (depth[hit] / acc[hit] < cam.far).all()
    colour = np.einsum("rnc,rn->rc", st["rgb_f"].astype(np.float64),
                       np.stack([O.compute_weights(st["sigma_f"][r], st["t_f"][r], cam.far) for r in range(W * H)]))
    assert np.abs(rgb - (colour + (1.0 - acc[:, None]))).max() < 1e-5
    part = renderer.render_aux(cam, W, H, nc, nf, precision="fp32", ray_begin=50, ray_end=200)
    assert (part[1] == depth[50:200]).all() and (part[2] == acc[50:200]).all()


def test_context_from_weight_cache_is_bit_identical(renderer, cam, tmp_path):
    """a context built from the pre-swizzled cache file renders and evaluates exactly what the
    context built from the weight directory does, in every precision mode."""
    path = str(tmp_path / "lego.nrfw")
    nb.build_weight_cache(nb.DATA_ROOT, path)
    r2 = nb.Renderer(cache=path, device=0)
    try:
        rng = np.random.default_rng(3)
        pts = rng.uniform(-1.5, 1.5, (3, 777)).astype(np.float32)
        dirs = rng.normal(size=(777, 3)).astype(np.float32)
        dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
        for prec in ("fp32", "bf16", "fp16", "tf32", "fp16x3"):
            for a, b in ((renderer.coarse, r2.coarse), (renderer.fine, r2.fine)):
                ra, sa = a.forward_batch(pts, dirs, precision=prec)
                rb, sb = b.forward_batch(pts, dirs, precision=prec)
                assert (ra == rb).all() and (sa == sb).all(), prec
        assert (renderer.render_image(cam, 16, 16, 64, 128) == r2.render_image(cam, 16, 16, 64, 128)).all()
    finally:
        r2.close()


_PEER_CHILD = r"""
import sys, ctypes as C
sys.path.insert(0, sys.argv[1])
import nerf_rs_b200 as nb
handle, W, H, rb, re_ = bytes.fromhex(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]), int(sys.argv[6])
r = nb.Renderer(nb.DATA_ROOT, device=0)
cam = nb.camera_from_json(nb.DATA_ROOT + "/tf_reference_samples.json")
p = C.c_void_p()
assert nb.lib().nerf_shared_frame_open(r._h, handle, C.byref(p)) == 0, nb.lib().nerf_last_error()
r.render_device(cam, W, H, p.value + 12 * rb, 64, 128, precision="mixed", ray_begin=rb, ray_end=re_)
import torch
torch.cuda.synchronize()
assert nb.lib().nerf_shared_frame_close(r._h, p, 0) == 0
r.close()
print("child ok")
"""


def test_shared_frame_second_process_stores_into_the_owner_buffer(renderer, cam):
    """the frame buffer lives with rank 0, every other rank (one process per GPU) maps it through CUDA
    IPC and its compositing kernel stores its ray range there.  Here the second process shares the one GPU of the test
    box; with one process per GPU the same mapping goes over NVLink.  The assembled frame is bit-identical to the
    single-process frame."""
    import ctypes as C
    import subprocess
    import sys
    import torch
    from nerf_rs_b200 import shard
    from tests.conftest import ROOT
    W, H = 48, 40
    n = W * H
    sf = shard.SharedFrame(renderer, n, 0, 1, how="ipc")
    try:
        assert len(sf.handle) == 64 and sf.ptr
        frame = sf.tensor()
        frame.zero_()
        torch.cuda.synchronize()
        half = shard.ray_range(1, 2, n)[0]
        child = subprocess.run([sys.executable, "-c", _PEER_CHILD, ROOT, sf.handle.hex(), str(W), str(H), str(half), str(n)],
                               capture_output=True, text=True, timeout=300)
        assert child.returncode == 0 and "child ok" in child.stdout, child.stderr[-2000:]
        renderer.render_device(cam, W, H, sf.row_ptr(0), 64, 128, precision="mixed", ray_begin=0, ray_end=half)
        sf.complete(torch.cuda.current_stream())
        torch.cuda.synchronize()
        got = frame.cpu().numpy()
        want = renderer.render_image(cam, W, H, 64, 128, precision="mixed")
        assert (got == want).all()
    finally:
        sf.close()


def _random_scene(root, seed):
    """Two random networks in the reference's weight-directory format (lego_rust/README.md:23-36): shapes.txt +
    <name>.bin, little-endian f32.  He-style scales keep activations O(1); biases are non-zero."""
    import os
    rng = np.random.default_rng(seed)
    shapes = [l.split() for l in open(os.path.join(nb.DATA_ROOT, "coarse", "shapes.txt")) if l.strip()]
    nets = []
    for net in ("coarse", "fine"):
        os.makedirs(os.path.join(root, net))
        d = {}
        for name, *dims in shapes:
            dims = [int(x) for x in dims]
            if name.endswith("_kernel"):
                a = rng.normal(0.0, np.sqrt(2.0 / dims[0]), dims)
            else:
                a = rng.normal(0.0, 0.1, dims)
            if name == "alpha_bias":
                a = a + 1.0  # keep a good share of the densities above the ReLU knee
            d[name] = a.astype(np.float32)
            d[name].tofile(os.path.join(root, net, name + ".bin"))
        with open(os.path.join(root, net, "shapes.txt"), "w") as f:
            f.writelines(" ".join(s) + "\n" for s in shapes)
        nets.append(d)
    return nets


def test_other_weights_from_tensors_directory_and_cache(cam, ocam, tmp_path):
    """Network::new (network.rs:181-195) through nerf_create_from_tensors, on weights that are not the lego ones: the
    context from caller-held tensors, the one from the same tensors written as a weight directory and the one from
    that directory's cache file are the same bit for bit, and match the oracle loading the directory."""
    root = str(tmp_path / "scene")
    coarse, fine = _random_scene(root, seed=7)
    onet_c, onet_f = O.Network(root + "/coarse"), O.Network(root + "/fine")
    assert (onet_f.tensor("dense5_kernel") == fine["dense5_kernel"].ravel()).all()
    nb.build_weight_cache(root, root + "/scene.nrfw")
    rs = [nb.Renderer(tensors=(coarse, fine), device=0), nb.Renderer(root, device=0),
          nb.Renderer(cache=root + "/scene.nrfw", device=0)]
    try:
        pts, dirs = _random_samples(ocam, 1500, seed=21)
        for which, onet in ((0, onet_c), (1, onet_f)):
            rgb_ref, sig_ref = onet.forward_batch(pts, dirs)
            assert sig_ref.max() > 0.5 and (sig_ref > 0).mean() > 0.2  # the test has something to compare
            for prec in ("fp32", "fp16x3", "fp16", "bf16"):
                outs = [(r.coarse, r.fine)[which].forward_batch(pts, dirs, precision=prec) for r in rs]
                for rgb, sig in outs[1:]:
                    assert (rgb == outs[0][0]).all() and (sig == outs[0][1]).all(), prec
                _check(prec, outs[0][0], outs[0][1], rgb_ref, sig_ref)
        want = O.render(onet_c, onet_f, ocam, 16, 16, 64, 128)
        for prec, tol in (("fp32", 2e-3), ("fp16x3", 2e-3)):
            got = rs[0].render_image(cam, 16, 16, 64, 128, precision=prec)
            assert np.abs(got - want).max() < tol, (prec, float(np.abs(got - want).max()))
    finally:
        for r in rs:
            r.close()


def test_forward_batch_million_samples(renderer, onets, ocam):
    """SURVEY.md T3 at full size: 2^20 random in-volume samples through the fine network (1.2e12 FLOP on the oracle,
    spread over the host cores in 2 048-sample slices), every mode against the same oracle values."""
    import os
    from concurrent.futures import ThreadPoolExecutor
    n = 1 << 20
    pts, dirs = _random_samples(ocam, n, seed=99)

    def oracle_slice(k):
        s = slice(k * 2048, (k + 1) * 2048)
        return onets[1].forward_batch(np.ascontiguousarray(pts[:, s]), dirs[s])

    with ThreadPoolExecutor(os.cpu_count() or 8) as ex:
        outs = list(ex.map(oracle_slice, range(n // 2048)))
    rgb_ref = np.concatenate([o[0] for o in outs])
    sig_ref = np.concatenate([o[1] for o in outs])
    assert sig_ref.max() > 300.0  # the sample set reaches the dense parts of the scene
    for prec in ("fp32", "fp16x3", "fp16", "bf16"):
        rgb, sig = renderer.fine.forward_batch(pts, dirs, precision=prec)
        # the single-pass modes' maximum sits on isolated samples at a ReLU knee and grows with the sample count
        # (fp16: 0.12 here against 0.04 over 4 000 samples); their bulk (99th percentile) is held to the usual bound
        wide = prec in ("fp16", "bf16")
        _check(prec, rgb, sig, rgb_ref, sig_ref, scale=4.0 if wide else 1.0)
        ds = np.abs(sig - sig_ref) / np.maximum(1.0, np.abs(sig_ref))
        assert np.percentile(ds, 99) <= SIG_P99[prec], (prec, float(np.percentile(ds, 99)))
        if prec == "fp16x3":  # absolute, at densities up to ~1 000: the mode's f32-class claim
            assert np.abs(sig - sig_ref).max() <= 4e-3 and np.abs(rgb - rgb_ref).max() <= 1e-4


@pytest.mark.parametrize("prec,ok", [("fp32", True), ("fp16x3", True), ("fp16", False), ("default", True)])
def test_reference_unit_test_through_the_cpp_mirror(renderer, cam, tmp_path, prec, ok):
    """The reference crate's own unit test (src/lib.rs:753-916), restated in C++ against include/nerf_rs.hpp (the
    compiled-language mirror of the crate's API over the C ABI; tests/cpp/reference_examples.cpp), with the reference's
    1e-2 absolute tolerance on every sigma and rgb channel.  It holds in the f32 and the split-operand modes; the
    single-pass fp16 mode misses it on sigma (0.14 at sigma = 213), as documented.  The same binary then renders
    through the mirror's render_image / save_ppm and must write the bytes of the library path."""
    import os
    import subprocess
    exe = os.path.join(os.path.dirname(nb.LIB_PATH), "reference_examples")
    out = tmp_path / "mirror.ppm"
    r = subprocess.run([exe, nb.DATA_ROOT, prec, str(out)], capture_output=True, text=True, timeout=300)
    assert r.returncode == (0 if ok else 1), r.stdout[-2000:] + r.stderr[-500:]
    if ok:
        assert "ok: 0 assertion(s) failed" in r.stdout
    else:
        assert "sigma" in r.stdout and " rgb" not in r.stdout  # only densities miss the bound
    # "default": nothing set on the mirror's Network - forward_batch then runs fp16x3 (and passes the reference's test),
    # render_image runs mixed
    img = renderer.render_image(cam, 32, 24, 64, 128, precision="mixed" if prec == "default" else prec)
    assert out.read_bytes() == b"P6\n32 24\n255\n" + nb.quantize_rgb8(img).tobytes()


_VMM_CHILD = r"""
import sys, ctypes as C
sys.path.insert(0, sys.argv[1])
import nerf_rs_b200 as nb
fd, size, W, H, rb, re_ = (int(x) for x in sys.argv[2:8])
r = nb.Renderer(nb.DATA_ROOT, device=0)
cam = nb.camera_from_json(nb.DATA_ROOT + "/tf_reference_samples.json")
p = C.c_void_p()
assert nb.lib().nerf_shared_frame_open_fd(r._h, fd, size, C.byref(p)) == 0, nb.lib().nerf_last_error()
r.render_device(cam, W, H, p.value + 12 * rb, 64, 128, precision="mixed", ray_begin=rb, ray_end=re_)
import torch
torch.cuda.synchronize()
assert nb.lib().nerf_shared_frame_close_fd(r._h, p) == 0
r.close()
print("child ok")
"""


def test_shared_frame_per_allocation_mapping(renderer, cam):
    """The same assembly with the frame buffer shared through the CUDA virtual-memory API (nerf_shared_frame_*_fd, what
    shard.SharedFrame and bench.py use by default): rank 0 exports a file descriptor, the second process maps it and
    stores its ray range; no device-wide peer access is involved."""
    import ctypes as C
    import os
    import subprocess
    import sys
    import torch
    from nerf_rs_b200 import shard
    from tests.conftest import ROOT
    W, H = 48, 40
    n = W * H
    L = nb.lib()
    ptr, size, fd = C.c_void_p(), C.c_uint64(), C.c_int(-1)
    assert L.nerf_shared_frame_create_fd(renderer._h, n * 12, C.byref(ptr), C.byref(size), C.byref(fd)) == 0, L.nerf_last_error()
    try:
        assert size.value >= n * 12 and fd.value >= 0
        half = shard.ray_range(1, 2, n)[0]
        child = subprocess.run([sys.executable, "-c", _VMM_CHILD, ROOT, str(fd.value), str(size.value), str(W), str(H),
                                str(half), str(n)], capture_output=True, text=True, timeout=300, pass_fds=(fd.value,))
        assert child.returncode == 0 and "child ok" in child.stdout, child.stderr[-2000:]
        renderer.render_device(cam, W, H, ptr.value, 64, 128, precision="mixed", ray_begin=0, ray_end=half)
        torch.cuda.synchronize()
        class _Raw:
            __cuda_array_interface__ = {"shape": (n, 3), "typestr": "<f4", "data": (ptr.value, False), "version": 2}
        got = torch.as_tensor(_Raw(), device="cuda:0").cpu().numpy()
        want = renderer.render_image(cam, W, H, 64, 128, precision="mixed")
        assert (got == want).all()
        # the helper the bench uses: single rank, creates and maps without a socket
        sf = shard.SharedFrame(renderer, n, 0, 1)
        t = sf.tensor()
        t.zero_()
        assert t.shape == (n, 3) and float(t.sum()) == 0.0
        sf.close()
    finally:
        os.close(fd.value)
        assert L.nerf_shared_frame_close_fd(renderer._h, ptr) == 0
