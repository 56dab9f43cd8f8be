"""ctypes binding of liboracle.so — TEST INFRASTRUCTURE ONLY.

The oracle is the CPU restatement of the reference's algorithm (oracle/oracle.cpp).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module; the
product package (nerf_rs_b200) never does.
"""
import ctypes as C
import json
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

ROUND_NONE, ROUND_BF16, ROUND_FP16, ROUND_TF32 = 0, 1, 2, 3
GEMM_REORDERED, GEMM_LITERAL = 0, 1
SAMPLING_DET, SAMPLING_PHILOX, SAMPLING_ARRAYS = 0, 1, 2

_f32p = C.POINTER(C.c_float)


class Camera(C.Structure):
    _fields_ = [("pos", C.c_float * 3), ("dir", C.c_float * 3), ("up", C.c_float * 3),
                ("alpha_w", C.c_float), ("alpha_h", C.c_float), ("near", C.c_float), ("far", C.c_float)]


class Sampling(C.Structure):
    _fields_ = [("mode", C.c_int), ("seed", C.c_uint64), ("jitter", _f32p), ("u", _f32p)]


class Stages(C.Structure):
    _fields_ = [(n, _f32p) for n in ("dirs", "t_c", "sigma_c", "w_c", "t_f", "sigma_f", "rgb_f")]


def build():
    """Compile liboracle.so with oracle/Makefile (idempotent)."""
    subprocess.run(["make", "-C", _HERE], check=True, stdout=subprocess.DEVNULL)


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    path = os.path.join(_HERE, "liboracle.so")
    if not os.path.exists(path):
        build()
    L = C.CDLL(path)
    L.oracle_last_error.restype = C.c_char_p
    L.oracle_load_network.restype = C.c_void_p
    L.oracle_load_network.argtypes = [C.c_char_p]
    L.oracle_free_network.argtypes = [C.c_void_p]
    L.oracle_get_tensor.restype = C.c_int64
    L.oracle_get_tensor.argtypes = [C.c_void_p, C.c_char_p, _f32p, C.c_int64]
    L.oracle_forward_batch.argtypes = [C.c_void_p, _f32p, _f32p, C.c_int64, C.c_int, C.c_int, _f32p, _f32p]
    L.oracle_camera_from_values.argtypes = [C.c_double, C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                            C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(Camera)]
    L.oracle_ray_dirs.argtypes = [C.POINTER(Camera), C.c_int, C.c_int, C.c_float, C.c_int, C.c_int64, C.c_int64, _f32p]
    L.oracle_stratified.argtypes = [C.c_float, C.c_float, C.c_int, _f32p, _f32p]
    L.oracle_compute_weights.argtypes = [_f32p, _f32p, C.c_int, C.c_float, _f32p]
    L.oracle_sample_importance.argtypes = [_f32p, _f32p, C.c_int, _f32p, C.c_int, _f32p]
    L.oracle_integrate_ray.argtypes = [_f32p, _f32p, _f32p, C.c_int, C.c_float, _f32p]
    L.oracle_integrate_aux.argtypes = [_f32p, _f32p, C.c_int, C.c_float, _f32p]
    L.oracle_philox_uniforms.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_int, _f32p]
    L.oracle_render.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(Camera), C.c_int, C.c_int, C.c_int, C.c_int,
                                C.POINTER(Sampling), C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_int, _f32p,
                                C.POINTER(Stages)]
    L.oracle_save_ppm.argtypes = [C.c_char_p, C.c_int, C.c_int, _f32p]
    L.oracle_quantize.argtypes = [_f32p, C.c_int64, C.POINTER(C.c_uint8)]
    L.oracle_num_threads.restype = C.c_int
    _LIB = L
    return L


def _p(a):
    return a.ctypes.data_as(_f32p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def last_error():
    return lib().oracle_last_error().decode()


class Network:
    """load_network_from_dir (src/lib.rs:108-174) + Network::forward_batch (src/network.rs:197-237)."""

    def __init__(self, directory):
        self.h = lib().oracle_load_network(os.fsencode(directory))
        if not self.h:
            raise RuntimeError(last_error())

    def __del__(self):
        if getattr(self, "h", None):
            lib().oracle_free_network(self.h)
            self.h = None

    def tensor(self, name):
        n = lib().oracle_get_tensor(self.h, name.encode(), None, 0)
        if n < 0:
            raise KeyError(name)
        out = np.empty(n, np.float32)
        lib().oracle_get_tensor(self.h, name.encode(), _p(out), n)
        return out

    def forward_batch(self, points, dirs, gemm_order=GEMM_REORDERED, round_mode=ROUND_NONE):
        """points: [3,B] feature-major; dirs: [B,3].  Returns (rgb [B,3], sigma [B])."""
        points = _f32(points)
        dirs = _f32(dirs)
        B = points.shape[1]
        assert points.shape == (3, B) and dirs.shape == (B, 3)
        rgb = np.empty((B, 3), np.float32)
        sigma = np.empty(B, np.float32)
        rc = lib().oracle_forward_batch(self.h, _p(points), _p(dirs), B, gemm_order, round_mode, _p(rgb), _p(sigma))
        if rc:
            raise RuntimeError(last_error())
        return rgb, sigma


def camera_from_json(path_or_dict):
    """camera_from_samples (src/lib.rs:614-645): keys near, far, camera_origin, camera_forward, camera_up, hwf."""
    d = path_or_dict
    if not isinstance(d, dict):
        with open(d) as f:
            d = json.load(f)
    return camera_from_values(d["near"], d["far"], d["camera_origin"], d["camera_forward"], d["camera_up"], d["hwf"])


def camera_from_values(near, far, origin, forward, up, hwf):
    cam = Camera()
    d3 = C.c_double * 3
    lib().oracle_camera_from_values(float(near), float(far), d3(*origin), d3(*forward), d3(*up), d3(*hwf), C.byref(cam))
    return cam


def ray_dirs(cam, nx, ny, offset=0.5, normalize=True, ray_begin=0, ray_end=None):
    ray_end = nx * ny if ray_end is None else ray_end
    out = np.empty((ray_end - ray_begin, 3), np.float32)
    lib().oracle_ray_dirs(C.byref(cam), nx, ny, offset, int(normalize), ray_begin, ray_end, _p(out))
    return out


def stratified(near, far, count, jitter):
    jitter = _f32(jitter)
    out = np.empty(count, np.float32)
    lib().oracle_stratified(near, far, count, _p(jitter), _p(out))
    return out


def compute_weights(sigma, t, far):
    sigma, t = _f32(sigma), _f32(t)
    w = np.empty(len(t), np.float32)
    lib().oracle_compute_weights(_p(sigma), _p(t), len(t), far, _p(w))
    return w


def sample_importance(t, w, u):
    t, w, u = _f32(t), _f32(w), _f32(u)
    out = np.empty(len(u), np.float32)
    n = lib().oracle_sample_importance(_p(t), _p(w), len(t), _p(u), len(u), _p(out))
    return out[:n]


def integrate_aux(sigma, t, far):
    """(depth, acc) = (sum w_i t_i, sum w_i) over compute_weights' weights."""
    sigma, t = _f32(sigma), _f32(t)
    out = np.empty(2, np.float32)
    lib().oracle_integrate_aux(_p(sigma), _p(t), len(t), far, _p(out))
    return out


def integrate_ray(rgb, sigma, t, far):
    rgb, sigma, t = _f32(rgb), _f32(sigma), _f32(t)
    out = np.empty(3, np.float32)
    lib().oracle_integrate_ray(_p(rgb), _p(sigma), _p(t), len(t), far, _p(out))
    return out


def philox_uniforms(seed, pixel_index, stream, count):
    out = np.empty(count, np.float32)
    lib().oracle_philox_uniforms(seed, pixel_index, stream, count, _p(out))
    return out


def render(coarse, fine, cam, W, H, nc, nf, mode=SAMPLING_DET, seed=0, jitter=None, u=None, ray_begin=0,
           ray_end=None, gemm_order=GEMM_REORDERED, round_mode=ROUND_NONE, threads=0, stages=False):
    """render_image (src/lib.rs:474-565) over pixel indices [ray_begin, ray_end).

    Returns rgb [n,3] (and a dict of per-stage arrays when stages=True)."""
    ray_end = W * H if ray_end is None else ray_end
    n = ray_end - ray_begin
    rgb = np.empty((n, 3), np.float32)
    smp = Sampling(mode=mode, seed=seed)
    keep = []
    if mode == SAMPLING_ARRAYS:
        jitter, u = _f32(jitter), _f32(u)
        keep += [jitter, u]
        smp.jitter, smp.u = _p(jitter), _p(u)
    st = None
    dump = {}
    if stages:
        nm = nc + nf
        dump = dict(dirs=np.zeros((n, 3), np.float32), t_c=np.zeros((n, nc), np.float32),
                    sigma_c=np.zeros((n, nc), np.float32), w_c=np.zeros((n, nc), np.float32),
                    t_f=np.zeros((n, nm), np.float32), sigma_f=np.zeros((n, nm), np.float32),
                    rgb_f=np.zeros((n, nm, 3), np.float32))
        st = Stages(**{k: _p(v) for k, v in dump.items()})
    rc = lib().oracle_render(coarse.h, fine.h, C.byref(cam), W, H, nc, nf, C.byref(smp), ray_begin, ray_end,
                             gemm_order, round_mode, threads, _p(rgb), C.byref(st) if st is not None else None)
    if rc:
        raise ValueError(last_error())
    return (rgb, dump) if stages else rgb


def render_skip_empty_rays(coarse, fine, cam, W, H, nc, nf, **kw):
    """Definition of the product's opt-in NERF_OPT_SKIP_EMPTY_RAYS (include/nerf_b200.h; the reference has no such
    mode): render_image as above, except that a ray ALL of whose coarse weights (compute_weights, src/lib.rs:250-283)
    are exactly 0 takes integrate_ray's value for a ray without weight, 0 + (1 - 0) * white (src/lib.rs:194), instead
    of the fine pass's.  Returns (rgb [n,3], skipped [n] bool)."""
    rgb, st = render(coarse, fine, cam, W, H, nc, nf, stages=True, **kw)
    skipped = np.all(st["w_c"] == 0.0, axis=1)
    rgb = rgb.copy()
    rgb[skipped] = 1.0
    return rgb, skipped


def save_ppm(path, W, H, rgb):
    rgb = _f32(rgb)
    assert rgb.size == W * H * 3
    if lib().oracle_save_ppm(os.fsencode(path), W, H, _p(rgb)):
        raise OSError(last_error())


def quantize(v):
    v = _f32(v)
    out = np.empty(v.size, np.uint8)
    lib().oracle_quantize(_p(v), v.size, out.ctypes.data_as(C.POINTER(C.c_uint8)))
    return out.reshape(v.shape)


def num_threads():
    return lib().oracle_num_threads()
