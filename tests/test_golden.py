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


B800 = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "render_800_rows200_207.npz"))


def test_oracle_reproduces_800_band_sample(onets, ocam):
    """Drift guard for the headline-resolution fixture (tests/golden/make_band800.py): 64 of its 6 400 rays."""
    rb = int(B800["ray_begin"]) + 394 - 32  # around row 200, column 394
    rgb = O.render(onets[0], onets[1], ocam, 800, 800, 64, 128, ray_begin=rb, ray_end=rb + 64)
    off = rb - int(B800["ray_begin"])
    assert np.abs(rgb - B800["rgb"][off:off + 64]).max() <= 2e-6


@pytest.mark.gpu
def test_gpu_800_band_against_the_oracle(renderer, cam):
    """North star at the headline resolution, against the CPU oracle itself: rows 200..207 of the 800x800 frame
    (6 400 rays through the object, thin structures included), identical deterministic sample positions.  The f32
    kernel, the split-operand mode and the default mixed mode stay within 2/255 and far above 45 dB; the all-fp16 mode
    keeps the PSNR but not the per-pixel bound on this band (row 207 holds its worst pixel), which is why it is not
    the default."""
    rb, re_ = int(B800["ray_begin"]), int(B800["ray_end"])
    want = B800["rgb"]

    def err(prec):
        got = renderer.render_image(cam, 800, 800, 64, 128, precision=prec, ray_begin=rb, ray_end=re_)
        d = np.abs(got - want)
        return float(d.max()) * 255.0, 10 * np.log10(1.0 / max(float(np.mean(d.astype(np.float64) ** 2)), 1e-20))

    for prec, psnr_min in (("fp32", 90.0), ("fp16x3", 80.0), ("mixed", 75.0)):
        mx, psnr = err(prec)
        print(f"800-band vs oracle, {prec}: max {mx:.3f}/255, PSNR {psnr:.1f} dB")
        assert mx <= 2.0 and psnr >= psnr_min, (prec, mx, psnr)
    mx, psnr = err("fp16")
    print(f"800-band vs oracle, fp16: max {mx:.3f}/255, PSNR {psnr:.1f} dB")
    assert psnr >= 45.0 and mx > 2.0, ("fp16", mx, psnr)  # documented behaviour, see DESIGN.md 4.3


def _frame800(name="frame_800.npz"):
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name))
    n = int(z["W"]) * int(z["H"])
    white = np.unpackbits(z["white"])[:n].astype(bool)
    frame = np.ones((n, 3), np.float32)
    frame[~white] = z["rgb"]
    return z, frame


def test_oracle_reproduces_full_frame_fixture(onets, ocam):
    """Drift guard for tests/golden/frame_800.npz (the oracle's whole 800x800 frame, tests/golden/make_frame800.py):
    the stored hash, its rows 200..207 against the older band fixture, and three 64-ray strips re-rendered now."""
    import hashlib
    z, frame = _frame800()
    assert hashlib.sha256(frame.tobytes()).hexdigest() == str(z["sha256"])
    assert np.array_equal(frame[int(B800["ray_begin"]):int(B800["ray_end"])], B800["rgb"])
    assert 0.70 < float(np.all(frame == 1.0, axis=1).mean()) < 0.75  # background: zero density along the whole ray
    for row, col in ((120, 400), (430, 300), (700, 520)):
        rb = row * 800 + col
        rgb = O.render(onets[0], onets[1], ocam, 800, 800, 64, 128, ray_begin=rb, ray_end=rb + 64)
        assert np.abs(rgb - frame[rb:rb + 64]).max() <= 2e-6


ORBIT = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "render_800_orbit_bands.npz"))


def _orbit_cams(make):
    """The fixture's cameras through `make(near, far, origin, forward, up, hwf)`; poses as nerf_rs_b200.orbit_poses."""
    import nerf_rs_b200 as nb
    from .conftest import CAM_JSON
    poses = nb.orbit_poses(CAM_JSON, int(ORBIT["n_views"]))
    return {int(k): make(poses[k]["near"], poses[k]["far"], poses[k]["camera_origin"], poses[k]["camera_forward"],
                         poses[k]["camera_up"], poses[k]["hwf"]) for k in ORBIT["poses"]}


def test_oracle_reproduces_orbit_bands(onets):
    """Drift guard for tests/golden/render_800_orbit_bands.npz (make_band_orbit.py): 64 rays of each pose."""
    rb = int(ORBIT["ray_begin"]) + 2 * 800 + 380
    for k, cam in _orbit_cams(O.camera_from_values).items():
        rgb = O.render(onets[0], onets[1], cam, 800, 800, 64, 128, ray_begin=rb, ray_end=rb + 64)
        off = rb - int(ORBIT["ray_begin"])
        assert np.abs(rgb - ORBIT[f"rgb_{k}"][off:off + 64]).max() <= 2e-6, k


@pytest.mark.gpu
def test_gpu_orbit_bands_against_the_oracle(renderer):
    """The image bound away from the reference pose: rows 396..403 of the 800x800 frame from two other cameras of the
    orbit (6 400 rays each, through the object), against the CPU oracle with identical deterministic sample positions.
    The reference's pipeline is discontinuous in its own rounding on thin structures, so a pixel or two
    of a band may leave 2/255 even in the order-faithful f32 mode; what is held is at most 2 such pixels per band and
    the PSNR."""
    import nerf_rs_b200 as nb
    rb, re_ = int(ORBIT["ray_begin"]), int(ORBIT["ray_end"])
    for k, cam in _orbit_cams(nb.camera_from_values).items():
        want = ORBIT[f"rgb_{k}"]
        assert want.std() > 0.05  # the band crosses the object
        for prec, psnr_min in (("fp32", 85.0), ("fp16x3", 80.0), ("mixed", 70.0)):
            got = renderer.render_image(cam, 800, 800, 64, 128, precision=prec, ray_begin=rb, ray_end=re_)
            d = np.abs(got - want).max(axis=1)
            psnr = 10 * np.log10(1.0 / max(float(np.mean((got.astype(np.float64) - want) ** 2)), 1e-20))
            print(f"orbit pose {k} band vs oracle, {prec}: max {d.max() * 255:.3f}/255, pixels > 2/255: "
                  f"{int((d > 2 / 255).sum())}, PSNR {psnr:.1f} dB")
            assert int((d > 2.0 / 255.0).sum()) <= 2 and psnr >= psnr_min, (k, prec, float(d.max()) * 255, psnr)


def test_oracle_reproduces_philox_frame_fixture(onets, ocam):
    """Drift guard for tests/golden/frame_800_philox0.npz (make_frame800.py 6 philox 0): the oracle's whole 800x800
    frame with the Philox jitter of the product's default sampling mode, seed 0."""
    import hashlib
    z, frame = _frame800("frame_800_philox0.npz")
    assert int(z["seed"]) == 0 and hashlib.sha256(frame.tobytes()).hexdigest() == str(z["sha256"])
    for row, col in ((150, 380), (450, 310)):
        rb = row * 800 + col
        rgb = O.render(onets[0], onets[1], ocam, 800, 800, 64, 128, mode=O.SAMPLING_PHILOX, seed=0, ray_begin=rb,
                       ray_end=rb + 64)
        assert np.abs(rgb - frame[rb:rb + 64]).max() <= 2e-6
