"""CPU suite: pins the oracle against the reference's only golden (tf_reference_samples.json, mirrored by the
reference's own unit test src/lib.rs:753-916) and checks the host-side boundary logic.  No GPU needed."""
import ctypes as C
import os
import struct

import numpy as np
import pytest

from oracle import pyoracle as O

from .conftest import CAM_JSON, DATA


def _example_inputs(golden, ex):
    """The reference test's inputs (lib.rs:849-860): p = ray_o + ray_d * z (un-normalised ray_d), dir = viewdir_unit."""
    o = np.array(golden["camera_origin"], np.float32)
    rd = np.array(ex["ray_d"], np.float32)
    vd = np.array(ex["viewdir_unit"], np.float32)
    z = np.array(golden["z_vals"], np.float32)
    pts = (o[:, None] + rd[:, None] * z[None, :]).astype(np.float32)
    return pts, np.tile(vd, (len(z), 1))


def test_loader_sizes(onets):
    # shapes.txt: 12 kernels + 12 biases, 595 844 params = 2 383 376 B per net (lego_rust/README.md)
    names = ["dense%d" % i for i in range(8)] + ["bottleneck", "viewdirs", "rgb", "alpha"]
    for net in onets:
        total = sum(net.tensor(n + "_kernel").size + net.tensor(n + "_bias").size for n in names)
        assert total == 595844
    w = onets[0].tensor("dense5_kernel")
    raw = np.fromfile(os.path.join(DATA, "coarse", "dense5_kernel.bin"), "<f4")
    assert w.size == 319 * 256 and (w == raw).all()  # byte-exact, little-endian f32 (lib.rs:35-42)


def test_oracle_matches_tf_golden(golden, onets):
    """The reference's own assertion is |d| < 1e-2 (lib.rs:732-742); the f32 restatement is far inside it."""
    worst_s = worst_c = 0.0
    for ex in golden["examples"]:
        pts, dirs = _example_inputs(golden, ex)
        for net, key in zip(onets, ("coarse", "fine")):
            rgb, sig = net.forward_batch(pts, dirs)
            worst_s = max(worst_s, np.abs(sig - np.array(ex[key + "_sigma"])).max())
            worst_c = max(worst_c, np.abs(rgb - np.array(ex[key + "_rgb"])).max())
    assert worst_s < 1e-4 and worst_c < 1e-6, (worst_s, worst_c)


def test_literal_loop_order_is_bit_identical(golden, onets):
    pts, dirs = _example_inputs(golden, golden["examples"][0])
    a = onets[1].forward_batch(pts, dirs, gemm_order=O.GEMM_REORDERED)
    b = onets[1].forward_batch(pts, dirs, gemm_order=O.GEMM_LITERAL)
    assert (a[0] == b[0]).all() and (a[1] == b[1]).all()


def test_empty_batch(onets):
    rgb, sig = onets[0].forward_batch(np.zeros((3, 0), np.float32), np.zeros((0, 3), np.float32))
    assert rgb.shape == (0, 3) and sig.shape == (0,)


def test_ray_dirs_match_json_ray_d(golden, ocam):
    """JSON pixel = [row, col] of the 400x400 TF image; un-normalised dirs with the pixel-corner offset equal ray_d."""
    for ex in golden["examples"]:
        i, j = ex["pixel"]
        d = O.ray_dirs(ocam, 400, 400, offset=0.0, normalize=False, ray_begin=i * 400 + j, ray_end=i * 400 + j + 1)
        assert np.abs(d[0] - np.array(ex["ray_d"])).max() < 5e-7
    d = O.ray_dirs(ocam, 8, 8)
    assert np.allclose(np.linalg.norm(d, axis=1), 1.0, atol=1e-6)


def test_stratified_det(ocam):
    t = O.stratified(2.0, 6.0, 64, np.full(64, 0.5, np.float32))
    assert np.allclose(t, 2.0 + (np.arange(64) + 0.5) * 0.0625, atol=1e-6)
    assert (np.diff(t) > 0).all()


def test_compute_weights_closed_form():
    n, far = 64, 6.0
    t = np.linspace(2.0, 6.0, n, endpoint=False).astype(np.float32)
    w = O.compute_weights(np.full(n, 0.7, np.float32), t, far)
    a = 1 - np.exp(-0.7 * 0.0625)
    assert np.allclose(w, (1 - a) ** np.arange(n) * a, rtol=2e-5)
    # early exit (lib.rs:276-279): everything after T < 1e-4 is exactly zero, the crossing sample keeps its weight
    w = O.compute_weights(np.full(n, 100.0, np.float32), t, far)
    assert w[0] > 0.99 and w[1] > 0 and (w[2:] == 0).all()
    # negative delta clamps to 0 (lib.rs:267-269)
    w = O.compute_weights(np.ones(3, np.float32), np.array([3, 2, 4], np.float32), far)
    assert w[0] == 0.0


def test_sample_importance_semantics():
    n = 64
    t = (2.0 + (np.arange(n) + 0.5) * 0.0625).astype(np.float32)
    w = np.zeros(n, np.float32)
    w[20] = 1.0
    u = ((np.arange(128) + 0.5) / 128).astype(np.float32)
    s = O.sample_importance(t, w, u)
    assert len(s) == 128 and (np.diff(s) >= 0).all()
    inside = (s >= 0.5 * (t[19] + t[20])) & (s <= 0.5 * (t[20] + t[21]))
    assert inside.mean() > 0.99  # nearly all mass in bin 19 of the 62 bins
    assert len(O.sample_importance(t[:2], w[:2], u)) == 0  # n < 3 (lib.rs:295-297)
    assert len(O.sample_importance(t, w, u[:0])) == 0      # count == 0
    # u at the top of the range falls in the last bin through the forced cdf[last] = 1 (lib.rs:326-328)
    s = O.sample_importance(t, np.ones(n, np.float32), np.array([0.0, 0.99999994], np.float32))
    assert abs(s[0] - 0.5 * (t[0] + t[1])) < 1e-6 and s[1] <= 0.5 * (t[-2] + t[-1]) + 1e-6


def test_integrate_ray_white_background():
    t = np.linspace(2, 6, 8, endpoint=False).astype(np.float32)
    out = O.integrate_ray(np.zeros((8, 3), np.float32), np.zeros(8, np.float32), t, 6.0)
    assert (out == 1.0).all()  # empty space -> white (lib.rs:194)
    out = O.integrate_ray(np.full((8, 3), 0.25, np.float32), np.full(8, 1e4, np.float32), t, 6.0)
    assert np.allclose(out, 0.25, atol=1e-6)
    assert (O.integrate_ray(np.zeros((0, 3)), np.zeros(0), np.zeros(0), 6.0) == 0).all()  # lib.rs:178-180


def test_render_arg_checks(onets, ocam):
    with pytest.raises(ValueError, match="width must be multiple"):
        O.render(onets[0], onets[1], ocam, 12, 8, 4, 4)
    with pytest.raises(ValueError, match="height must be multiple"):
        O.render(onets[0], onets[1], ocam, 8, 12, 4, 4)
    with pytest.raises(ValueError, match="greater than 0"):
        O.render(onets[0], onets[1], ocam, 8, 8, 0, 4)


def test_render_range_equals_full_and_stages(onets, ocam):
    full, st = O.render(onets[0], onets[1], ocam, 8, 8, 8, 8, stages=True)
    part = O.render(onets[0], onets[1], ocam, 8, 8, 8, 8, ray_begin=13, ray_end=41)
    assert (part == full[13:41]).all()
    assert (np.diff(st["t_f"], axis=1) >= 0).all() and st["t_f"].shape == (64, 16)
    ph = O.render(onets[0], onets[1], ocam, 8, 8, 8, 8, mode=O.SAMPLING_PHILOX, seed=7)
    ph2 = O.render(onets[0], onets[1], ocam, 8, 8, 8, 8, mode=O.SAMPLING_PHILOX, seed=7, ray_begin=8, ray_end=16)
    assert (ph2 == ph[8:16]).all() and not (ph == full).all()


def test_philox_known_answer():
    # Philox4x32-10 known-answer vector (Random123 kat_vectors): counter = key = 0
    u = O.philox_uniforms(0, 0, 0, 4)
    want = np.array([0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8], np.uint64)
    assert (u == ((want >> 8).astype(np.float32) / np.float32(16777216.0))).all()


def test_ppm_byte_exact(tmp_path):
    W, H = 16, 8
    g = np.linspace(-0.2, 1.2, W * H * 3, dtype=np.float32).reshape(H * W, 3)
    g[3, 1] = np.nan
    p = tmp_path / "o.ppm"
    O.save_ppm(str(p), W, H, g)
    raw = p.read_bytes()
    head = b"P6\n16 8\n255\n"
    assert raw.startswith(head) and len(raw) == len(head) + W * H * 3
    want = np.where(np.isnan(g), 0, np.floor(np.clip(g, 0, 1) * np.float32(255.0) + np.float32(0.5))).astype(np.uint8)
    assert (np.frombuffer(raw[len(head):], np.uint8) == want.ravel()).all()


# --------------------------------------------------------------------------- the product's host side (no GPU)
def test_abi_exports_every_declared_symbol():
    import re
    import nerf_rs_b200 as nb
    L = nb.lib()
    hdr = open(os.path.join(os.path.dirname(DATA), "..", "include", "nerf_b200.h")).read()
    declared = set(re.findall(r"\b(nerf_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(nb.EXPORTS)
    for name in declared:
        assert hasattr(L, name), name
    assert b"sm_100a" in L.nerf_version()


def test_host_camera_and_ppm_match_oracle(tmp_path, ocam):
    import nerf_rs_b200 as nb
    cam = nb.camera_from_json(CAM_JSON)
    for a, b in ((cam.origin, ocam.pos), (cam.forward, ocam.dir), (cam.up, ocam.up)):
        assert list(a) == list(b)
    assert (cam.alpha_w, cam.alpha_h, cam.near, cam.far) == (ocam.alpha_w, ocam.alpha_h, ocam.near, ocam.far)
    g = np.random.default_rng(0).uniform(-0.5, 1.5, (64, 3)).astype(np.float32)
    assert (nb.quantize_rgb8(g) == O.quantize(g)).all()
    nb.save_ppm(str(tmp_path / "a.ppm"), 8, 8, g)
    O.save_ppm(str(tmp_path / "b.ppm"), 8, 8, g)
    assert (tmp_path / "a.ppm").read_bytes() == (tmp_path / "b.ppm").read_bytes()
    with pytest.raises(nb.NerfError):
        nb.camera_from_json(str(tmp_path / "missing.json"))


def test_no_cpu_fallback():
    """Without a GPU the product must fail loudly, never compute on the CPU."""
    import torch
    import nerf_rs_b200 as nb
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(nb.NerfError) as e:
        nb.Renderer(DATA)
    assert e.value.code == -4
    h = C.c_void_p()
    assert nb.lib().nerf_create(b"/nonexistent", 0, C.byref(h)) == -2  # NERF_E_IO before touching CUDA


def test_cli_builds_and_fails_loudly_without_gpu(tmp_path):
    """nerf_cli mirrors main.rs / render_cli_image (src/lib.rs:647-677); without a GPU it must fail, not fall back."""
    import subprocess
    import nerf_rs_b200 as nb
    cli = os.path.join(os.path.dirname(nb.LIB_PATH), "nerf_cli")
    assert os.path.exists(cli), "run __graft_entry__.build()"
    assert subprocess.run([cli, "--help"], capture_output=True).returncode == 0
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("GPU present: covered by the gpu test")
    except ImportError:
        pass
    r = subprocess.run([cli, "--weights", DATA, "--out", str(tmp_path / "o.ppm")], capture_output=True, text=True)
    assert r.returncode != 0 and "no CPU fallback" in r.stderr and not (tmp_path / "o.ppm").exists()


def test_cpp_mirror_builds_and_panics_without_gpu():
    """include/nerf_rs.hpp (the reference crate's API over the C ABI) compiles into the reference's unit test
    (tests/cpp/reference_examples.cpp); without a GPU load_network_from_dir panics - exit code 2, no fallback."""
    import subprocess
    import nerf_rs_b200 as nb
    from tests.conftest import _have_gpu
    exe = os.path.join(os.path.dirname(nb.LIB_PATH), "reference_examples")
    assert os.path.exists(exe)
    if not _have_gpu():
        r = subprocess.run([exe, DATA], capture_output=True, text=True)
        assert r.returncode == 2 and "no CPU fallback" in r.stderr
    r = subprocess.run([exe, "/nonexistent"], capture_output=True, text=True)
    assert r.returncode == 2 and "panic:" in r.stderr


def test_integrate_aux_closed_form():
    """Constant sigma on a uniform grid: T_i = exp(-sigma * i * d), acc = 1 - exp(-sigma * L) up to the T < 1e-4
    cut-off of compute_weights (lib.rs:276-279), depth = sum w_i t_i by direct summation."""
    n, sigma, near, far = 64, 0.7, 2.0, 6.0
    t = (near + (far - near) * np.arange(n) / n).astype(np.float32)
    s = np.full(n, sigma, np.float32)
    depth, acc = O.integrate_aux(s, t, far)
    w = O.compute_weights(s, t, far).astype(np.float64)
    assert abs(acc - w.sum()) < 1e-6 and abs(depth - (w * t).sum()) < 1e-5
    assert abs(acc - (1.0 - np.exp(-sigma * (far - near)))) < 2e-4
    assert tuple(O.integrate_aux(np.zeros(0, np.float32), np.zeros(0, np.float32), far)) == (0.0, 0.0)
    d0, a0 = O.integrate_aux(np.zeros(n, np.float32), t, far)  # empty space
    assert d0 == 0.0 and a0 == 0.0


def test_orbit_cameras_pose0_is_the_json_camera_and_views_shard():
    """The throughput sweep's synthetic poses (BASELINE.json configs[4]): pose 0 is the reference camera, every pose
    keeps its radius / elevation, looks at the origin with an orthonormal basis, and the contiguous view split used by
    `bench.py --orbit` covers all views exactly once."""
    import nerf_rs_b200 as nb
    from nerf_rs_b200 import shard
    from tests.conftest import CAM_JSON
    c0 = nb.camera_from_json(CAM_JSON)
    cams = nb.orbit_cameras(CAM_JSON, 200)
    assert len(cams) == 200
    for f in ("origin", "forward", "up"):
        assert np.abs(np.array(getattr(c0, f)[:]) - np.array(getattr(cams[0], f)[:])).max() < 2e-7
    assert cams[0].alpha_w == c0.alpha_w and cams[0].alpha_h == c0.alpha_h
    r0, z0 = np.linalg.norm(c0.origin[:]), c0.origin[2]
    for c in cams[::17]:
        o, f, u = (np.array(x[:], np.float64) for x in (c.origin, c.forward, c.up))
        assert abs(np.linalg.norm(o) - r0) < 1e-5 and abs(o[2] - z0) < 1e-6
        assert np.abs(f + o / np.linalg.norm(o)).max() < 1e-6 and abs(f @ u) < 1e-6 and abs(np.linalg.norm(u) - 1) < 1e-6
        assert (c.near, c.far) == (c0.near, c0.far)
    a = np.arctan2(cams[50].origin[1], cams[50].origin[0]) - np.arctan2(c0.origin[1], c0.origin[0])
    assert abs((a % (2 * np.pi)) - np.pi / 2) < 1e-5  # a quarter turn after 50 of 200 views
    for world in (1, 2, 3, 8):
        got = [v for r in range(world) for v in range(*shard.ray_range(r, world, 200))]
        assert got == list(range(200))


def test_weight_cache_file_is_deterministic_and_checked(tmp_path):
    """nerf_build_weight_cache (host only): same bytes every time, holds the 2 x 24 f32 tensors bit for bit, and a
    loader refuses a truncated / corrupted / foreign file before touching the GPU (checksum first, then device)."""
    import struct
    import nerf_rs_b200 as nb
    a, b = str(tmp_path / "a.nrfw"), str(tmp_path / "b.nrfw")
    nb.build_weight_cache(DATA, a)
    nb.build_weight_cache(DATA, b)
    raw = open(a, "rb").read()
    assert raw == open(b, "rb").read()
    version, n_nets = struct.unpack_from("<II", raw, 8)
    assert raw[:8] == b"NRFB200W" and n_nets == 2
    assert version >> 24 == 2 and version & 0xffffff != 0  # format revision | hash of the chunk-stream layout tag
    # two nets x (f32 tensors + bf16 + fp16 + split-fp16 + tf32 streams): 2 x (2 383 376 + 1 196 032 x 4 + 2 375 680) B
    # + names / dims
    assert 2 * (2383376 + 4 * 1196032 + 2375680) < len(raw) < 2 * (2383376 + 4 * 1196032 + 2375680) + 4096
    w = np.fromfile(os.path.join(DATA, "fine", "dense5_kernel.bin"), np.float32)
    assert raw.count(w.tobytes()) == 1  # stored verbatim
    with pytest.raises(nb.NerfError) as e:
        nb.build_weight_cache(str(tmp_path / "nowhere"), a)
    assert e.value.code == nb.E_IO
    import ctypes as C
    h = C.c_void_p()
    L = nb.lib()
    bad = bytearray(raw)
    bad[len(bad) // 2] ^= 0x40
    for name, blob in (("flip", bytes(bad)), ("trunc", raw[:-100]), ("magic", b"XXXXXXXX" + raw[8:]), ("empty", b"")):
        path = str(tmp_path / name)
        open(path, "wb").write(blob)
        assert L.nerf_create_from_cache(os.fsencode(path), 0, C.byref(h)) == nb.E_IO, name
        assert b"weight cache" in L.nerf_last_error()
    assert L.nerf_create_from_cache(os.fsencode(str(tmp_path / "missing")), 0, C.byref(h)) == nb.E_IO
    # a file written for another chunk-stream layout (an older kernel generation): same size, valid checksum, other
    # layout hash in the version word -> refused, never rendered
    def fnv1a64(b):
        hsh = 1469598103934665603  # weight_cache.cu::fnv1a
        for x in np.frombuffer(b, np.uint8).tolist():
            hsh = ((hsh ^ x) * 0x100000001b3) & 0xffffffffffffffff
        return hsh
    small = bytearray(raw[:8] + struct.pack("<II", version ^ 0x5a5a, 2))
    small += struct.pack("<Q", fnv1a64(bytes(small)))
    path = str(tmp_path / "other_layout")
    open(path, "wb").write(bytes(small))
    assert L.nerf_create_from_cache(os.fsencode(path), 0, C.byref(h)) == nb.E_UNSUPPORTED
    assert b"unsupported version" in L.nerf_last_error()
    # an intact file gets as far as the device: no GPU here, no CPU fallback
    from tests.conftest import _have_gpu
    if not _have_gpu():
        assert L.nerf_create_from_cache(os.fsencode(a), 0, C.byref(h)) == nb.E_CUDA


def test_shared_frame_entry_points_reject_bad_arguments():
    """The multi-GPU frame-buffer entry points (both forms) return NERF_E_ARG on null / empty arguments before touching
    CUDA, so the check runs without a GPU; closing nothing is a no-op."""
    import ctypes as C
    import nerf_rs_b200 as nb
    L = nb.lib()
    p, size, fd, buf = C.c_void_p(), C.c_uint64(), C.c_int(-1), C.create_string_buffer(64)
    assert L.nerf_shared_frame_create(None, 1024, C.byref(p), buf) == nb.E_ARG
    assert L.nerf_shared_frame_open(None, buf, C.byref(p)) == nb.E_ARG
    assert L.nerf_shared_frame_close(None, None, 1) == 0
    assert L.nerf_shared_frame_create_fd(None, 1024, C.byref(p), C.byref(size), C.byref(fd)) == nb.E_ARG
    assert L.nerf_shared_frame_open_fd(None, 3, 1024, C.byref(p)) == nb.E_ARG
    assert L.nerf_shared_frame_close_fd(None, None) == 0
    assert b"bad argument" in L.nerf_last_error()


def test_rust_shim_binds_only_exported_symbols():
    """rust/ cannot be compiled here (no cargo / rustc); what can be checked is that every `extern "C"` function its
    ffi.rs declares is a symbol libnerf_b200.so exports and include/nerf_b200.h declares, and that the crate has the
    reference's file layout (Cargo.toml, build.rs, src/{lib,main,network,vec3,weights}.rs)."""
    import re
    import ctypes as C
    import nerf_rs_b200 as nb
    root = os.path.join(os.path.dirname(DATA.rstrip("/")), "..", "rust")
    for rel in ("Cargo.toml", "build.rs", "README.md", "src/lib.rs", "src/main.rs", "src/network.rs", "src/vec3.rs",
                "src/weights.rs", "src/ffi.rs"):
        assert os.path.exists(os.path.join(root, rel)), rel
    ffi = open(os.path.join(root, "src", "ffi.rs")).read()
    names = set(re.findall(r"pub fn (nerf_\w+)\s*\(", ffi))
    assert {"nerf_create", "nerf_forward_batch", "nerf_render", "nerf_save_ppm", "nerf_create_from_tensors"} <= names
    header = open(os.path.join(root, "..", "include", "nerf_b200.h")).read()
    L = C.CDLL(nb.LIB_PATH)
    for n in names:
        assert n in nb.EXPORTS and re.search(r"\b%s\s*\(" % n, header), n
        getattr(L, n)
    cargo = open(os.path.join(root, "Cargo.toml")).read()
    assert 'build = "build.rs"' in cargo and 'crate-type = ["rlib", "cdylib"]' in cargo and 'name = "nerf-rs"' in cargo
    # precision constants of the shim == the header's enum
    for name, val in re.findall(r"pub const PREC_(\w+): c_int = (\d+);", ffi):
        assert re.search(r"NERF_PREC_%s = %s\b" % (name, val), header), name
