"""nerf_rs_b200 — B200 (sm_100a) replacement for the hot path of elisabeth96/nerf-rs.

The product is libnerf_b200.so (CUDA kernels + C ABI, include/nerf_b200.h).  This package is the thin
ctypes binding used by the tests, the bench and Python callers; it mirrors the reference crate's surface
(load_network_from_dir / Network.forward_batch, src/network.rs:197; render_image, src/lib.rs:474;
camera_from_samples, src/lib.rs:614; save_ppm, src/lib.rs:567).

There is no CPU fallback: importing works anywhere (so the symbol table can be checked), but creating a
Renderer without an sm_100 GPU, or with the shared library missing, raises.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnerf_b200.so")
DATA_ROOT = os.path.join(os.path.dirname(_HERE), "data", "lego_rust")

PREC_FP32, PREC_BF16, PREC_FP16, PREC_TF32, PREC_FP16X3, PREC_MIXED = 0, 1, 2, 3, 4, 5
PRECISIONS = {"fp32": PREC_FP32, "bf16": PREC_BF16, "fp16": PREC_FP16, "tf32": PREC_TF32, "fp16x3": PREC_FP16X3,
              "mixed": PREC_MIXED}
NET_COARSE, NET_FINE = 0, 1
SAMPLING_DET, SAMPLING_PHILOX, SAMPLING_ARRAYS = 0, 1, 2

_f32p = C.POINTER(C.c_float)

# every symbol include/nerf_b200.h declares
EXPORTS = ["nerf_last_error", "nerf_version", "nerf_create", "nerf_create_from_tensors", "nerf_destroy",
           "nerf_forward_batch", "nerf_forward_batch_device", "nerf_render", "nerf_render_device",
           "nerf_render_stages", "nerf_ray_dirs", "nerf_compute_weights", "nerf_sample_importance",
           "nerf_integrate_rays", "nerf_get_timing", "nerf_camera_from_json", "nerf_camera_from_values",
           "nerf_save_ppm", "nerf_quantize_rgb8", "nerf_render_rgb8", "nerf_render_rgb8_device", "nerf_render_many",
           "nerf_quantize_rgb8_gpu", "nerf_render_aux", "nerf_render_aux_device", "nerf_build_weight_cache",
           "nerf_create_from_cache", "nerf_shared_frame_create", "nerf_shared_frame_open", "nerf_shared_frame_close",
           "nerf_shared_frame_create_fd", "nerf_shared_frame_open_fd", "nerf_shared_frame_close_fd",
           "nerf_render_striped", "nerf_render_striped_device", "nerf_integrate_rays_noalloc", "nerf_set_option",
           "nerf_get_option"]
OPT_SKIP_EMPTY_RAYS = 1  # nerf_option (include/nerf_b200.h)
NOALLOC_MAX_BYTES = 4 << 20  # NERF_NOALLOC_MAX_BYTES (include/nerf_b200.h)


E_ARG, E_IO, E_SHAPE, E_CUDA, E_UNSUPPORTED = -1, -2, -3, -4, -5  # nerf_status


class NerfError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"nerf_b200 error {code}: {msg}")
        self.code = code


class Camera(C.Structure):
    """nerf_camera == struct Camera after camera_from_samples (src/lib.rs:197-211,614-645)."""
    _fields_ = [("origin", C.c_float * 3), ("forward", C.c_float * 3), ("up", C.c_float * 3),
                ("alpha_w", C.c_float), ("alpha_h", C.c_float), ("near", C.c_float), ("far", C.c_float)]


class Sampling(C.Structure):
    _fields_ = [("mode", C.c_int32), ("seed", C.c_uint64), ("jitter", C.c_void_p), ("u", C.c_void_p)]


class Stages(C.Structure):
    _fields_ = [(n, _f32p) for n in ("dirs", "t_c", "sigma_c", "w_c", "t_f", "sigma_f", "rgb_f")]


class Timing(C.Structure):
    _fields_ = [("total_ms", C.c_float), ("mlp_coarse_ms", C.c_float), ("mlp_fine_ms", C.c_float),
                ("sampling_ms", C.c_float), ("composite_ms", C.c_float), ("kernel_launches", C.c_int32),
                ("coarse_refine_ms", C.c_float), ("coarse_samples", C.c_int64), ("coarse_refined", C.c_int64),
                ("rays", C.c_int64), ("rays_skipped", C.c_int64)]


class Tensor(C.Structure):
    _fields_ = [("name", C.c_char_p), ("ndim", C.c_int32), ("dims", C.c_int64 * 2), ("data", _f32p)]


def build(verbose=False):
    """Compile libnerf_b200.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    out = None if verbose else subprocess.DEVNULL
    subprocess.run(["make", "-C", os.path.join(_HERE, "csrc")], check=True, stdout=out)


_LIB = None


def lib():
    """Load libnerf_b200.so; fails loudly when it has not been built."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64 = C.c_void_p, C.c_int, C.c_int64
    L.nerf_last_error.restype = C.c_char_p
    L.nerf_version.restype = C.c_char_p
    L.nerf_create.argtypes = [C.c_char_p, i32, C.POINTER(vp)]
    L.nerf_shared_frame_create.argtypes = [vp, C.c_uint64, C.POINTER(vp), C.c_char_p]
    L.nerf_shared_frame_open.argtypes = [vp, C.c_char_p, C.POINTER(vp)]
    L.nerf_shared_frame_close.argtypes = [vp, vp, i32]
    L.nerf_shared_frame_create_fd.argtypes = [vp, C.c_uint64, C.POINTER(vp), C.POINTER(C.c_uint64), C.POINTER(i32)]
    L.nerf_shared_frame_open_fd.argtypes = [vp, i32, C.c_uint64, C.POINTER(vp)]
    L.nerf_shared_frame_close_fd.argtypes = [vp, vp]
    L.nerf_build_weight_cache.argtypes = [C.c_char_p, C.c_char_p]
    L.nerf_create_from_cache.argtypes = [C.c_char_p, i32, C.POINTER(vp)]
    L.nerf_create_from_tensors.argtypes = [C.POINTER(Tensor), i32, C.POINTER(Tensor), i32, i32, C.POINTER(vp)]
    L.nerf_destroy.argtypes = [vp]
    L.nerf_destroy.restype = None
    L.nerf_forward_batch.argtypes = [vp, i32, _f32p, _f32p, i64, i32, _f32p, _f32p]
    L.nerf_forward_batch_device.argtypes = [vp, i32, vp, vp, i64, i32, vp, vp, vp]
    rargs = [vp, C.POINTER(Camera), i32, i32, i32, i32, i32, C.POINTER(Sampling), i64, i64]
    L.nerf_render.argtypes = rargs + [_f32p]
    L.nerf_render_device.argtypes = rargs + [vp, vp]
    L.nerf_render_stages.argtypes = rargs + [_f32p, C.POINTER(Stages)]
    sargs = rargs[:8] + [i64, i32, i32]
    L.nerf_render_striped.argtypes = sargs + [_f32p]
    L.nerf_render_striped_device.argtypes = sargs + [vp, vp]
    u8p = C.POINTER(C.c_uint8)
    L.nerf_render_rgb8.argtypes = rargs + [i32, u8p]
    L.nerf_render_rgb8_device.argtypes = rargs + [i32, vp, vp]
    L.nerf_render_many.argtypes = [vp, C.POINTER(Camera), i32, i32, i32, i32, i32, i32, C.POINTER(Sampling), i32, u8p]
    L.nerf_quantize_rgb8_gpu.argtypes = [vp, _f32p, i64, i32, u8p]
    L.nerf_render_aux.argtypes = rargs + [_f32p, _f32p, _f32p]
    L.nerf_render_aux_device.argtypes = rargs + [vp, vp, vp, vp]
    L.nerf_ray_dirs.argtypes = [vp, C.POINTER(Camera), i32, i32, i64, i64, _f32p]
    L.nerf_compute_weights.argtypes = [vp, _f32p, _f32p, i64, i32, C.c_float, _f32p]
    L.nerf_sample_importance.argtypes = [vp, _f32p, _f32p, i64, i32, _f32p, i32, _f32p]
    L.nerf_integrate_rays.argtypes = [vp, _f32p, _f32p, _f32p, i64, i32, C.c_float, _f32p]
    L.nerf_integrate_rays_noalloc.argtypes = [vp, _f32p, _f32p, _f32p, i64, i32, C.c_float, _f32p]
    L.nerf_get_timing.argtypes = [vp, C.POINTER(Timing)]
    L.nerf_set_option.argtypes = [vp, i32, i32]
    L.nerf_get_option.argtypes = [vp, i32, C.POINTER(i32)]
    L.nerf_camera_from_json.argtypes = [C.c_char_p, C.POINTER(Camera)]
    d3 = C.POINTER(C.c_double)
    L.nerf_camera_from_values.argtypes = [C.c_double, C.c_double, d3, d3, d3, d3, C.POINTER(Camera)]
    L.nerf_save_ppm.argtypes = [C.c_char_p, i32, i32, _f32p]
    L.nerf_quantize_rgb8.argtypes = [_f32p, i64, C.POINTER(C.c_uint8)]
    _LIB = L
    return L


def _check(rc):
    if rc != 0:
        raise NerfError(rc, lib().nerf_last_error().decode())


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return a.ctypes.data_as(_f32p)


def camera_from_json(path):
    """camera_from_samples (src/lib.rs:614-645) from a JSON file with the reference's keys."""
    cam = Camera()
    _check(lib().nerf_camera_from_json(os.fsencode(path), C.byref(cam)))
    return cam


def camera_from_values(near, far, origin, forward, up, hwf):
    cam = Camera()
    d3 = C.c_double * 3
    _check(lib().nerf_camera_from_values(float(near), float(far), d3(*origin), d3(*forward), d3(*up), d3(*hwf),
                                         C.byref(cam)))
    return cam


def build_weight_cache(weights_root, cache_path):
    """Writes the pre-swizzled weight cache of a scene (host only, no GPU): the f32 tensors of
    load_network_from_dir (src/lib.rs:108-174) plus the tensor-core chunk streams, one checksummed file."""
    _check(lib().nerf_build_weight_cache(os.fsencode(weights_root), os.fsencode(cache_path)))


def orbit_poses(json_path, n_views):
    """Synthetic poses of the throughput sweep as dicts with the reference's JSON keys: the
    JSON camera's radius and elevation, azimuth theta_k = theta_0 + 2 pi k / n_views about +z (pose 0 is the JSON
    camera), forward = -pos / |pos| (the JSON camera looks exactly at the origin), up = +z orthogonalised against
    forward, same near / far / hwf."""
    import json
    with open(json_path) as f:
        d = json.load(f)
    o = np.array(d["camera_origin"], np.float64)
    radius, rho = float(np.linalg.norm(o)), float(np.hypot(o[0], o[1]))
    theta0 = float(np.arctan2(o[1], o[0]))
    poses = []
    for k in range(n_views):
        th = theta0 + 2.0 * np.pi * k / n_views
        pos = np.array([rho * np.cos(th), rho * np.sin(th), o[2]])
        fwd = -pos / radius
        up = np.array([0.0, 0.0, 1.0]) - fwd[2] * fwd
        up /= np.linalg.norm(up)
        poses.append({"near": d["near"], "far": d["far"], "camera_origin": pos.tolist(), "camera_forward": fwd.tolist(),
                      "camera_up": up.tolist(), "hwf": d["hwf"]})
    return poses


def orbit_cameras(json_path, n_views):
    """orbit_poses through camera_from_samples (src/lib.rs:614-645), like the reference's own camera."""
    return [camera_from_values(p["near"], p["far"], p["camera_origin"], p["camera_forward"], p["camera_up"], p["hwf"])
            for p in orbit_poses(json_path, n_views)]


def save_ppm(path, W, H, rgb):
    """save_ppm (src/lib.rs:567-580), byte-exact."""
    rgb = _f32(rgb)
    if rgb.size != W * H * 3:
        raise ValueError("pixels.len() != width * height")  # lib.rs:569
    _check(lib().nerf_save_ppm(os.fsencode(path), W, H, _p(rgb)))


def quantize_rgb8(rgb):
    rgb = _f32(rgb)
    out = np.empty(rgb.shape, np.uint8)
    _check(lib().nerf_quantize_rgb8(_p(rgb), rgb.size, out.ctypes.data_as(C.POINTER(C.c_uint8))))
    return out


from .renderer import Network, Renderer  # noqa: E402,F401
