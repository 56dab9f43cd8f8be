"""Renderer / Network: the reference crate's call surface over the C ABI (host-side mirror).

  Renderer(root)                       ~ load_network_from_dir x2          (src/lib.rs:108-174, 651-652)
  Renderer.coarse / .fine -> Network   ~ struct Network                    (src/network.rs:172-178)
  Network.forward_batch(points, dirs)  ~ Network::forward_batch            (src/network.rs:197-237)
  Renderer.render_image(cam, W, H, ..) ~ render_image                      (src/lib.rs:474-565)
"""
import ctypes as C

import numpy as np

from . import (Camera, DATA_ROOT, PRECISIONS, SAMPLING_ARRAYS, SAMPLING_DET, Sampling, Stages, Tensor, Timing, _check, _f32,
               _f32p, _p, lib)


def _prec(p):
    return PRECISIONS[p] if isinstance(p, str) else int(p)


class Network:
    """One of the two networks of a Renderer; forward_batch keeps the reference's argument layout."""

    def __init__(self, renderer, net_id):
        self._r = renderer
        self.net_id = net_id

    def forward_batch(self, points, view_dirs, precision="fp16x3"):
        """points: [3, B] feature-major (the reference's Matrix 3xB); view_dirs: [B, 3].

        Returns (rgb [B, 3], sigma [B]).  An empty batch returns empty arrays (network.rs:199-201).  The default
        precision is fp16x3: the one tensor-core mode that passes the reference's own unit test (1e-2 abs on sigma up
        to 213, src/lib.rs:732-742) with margin; render_image defaults to "mixed", where the image bound applies."""
        points = _f32(points)
        view_dirs = _f32(view_dirs)
        B = points.shape[1] if points.ndim == 2 else 0
        if points.ndim != 2 or points.shape[0] != 3:
            raise ValueError("points must be [3, B]")  # debug_assert_eq!(points.rows(), 3) (network.rs:264)
        if view_dirs.shape != (B, 3):
            raise ValueError("batch != view_dirs.len()")  # network.rs:202
        rgb = np.empty((B, 3), np.float32)
        sigma = np.empty(B, np.float32)
        _check(lib().nerf_forward_batch(self._r._h, self.net_id, _p(points), _p(view_dirs), B, _prec(precision),
                                        _p(rgb), _p(sigma)))
        return rgb, sigma

    def forward_batch_device(self, d_points, d_dirs, B, d_rgb, d_sigma, precision="fp16x3", stream=0):
        """Same on device pointers (ints), enqueued on `stream` without synchronising."""
        _check(lib().nerf_forward_batch_device(self._r._h, self.net_id, d_points, d_dirs, B, _prec(precision), d_rgb,
                                               d_sigma, stream))


class Renderer:
    def __init__(self, weights_root=DATA_ROOT, device=0, tensors=None, cache=None):
        """weights_root holds coarse/ and fine/ (shapes.txt + *.bin).  `tensors` = (coarse_dict, fine_dict) of
        name -> ndarray builds the context from memory instead (Network::new, network.rs:181-195); `cache` = path of a
        file written by build_weight_cache."""
        h = C.c_void_p()
        if cache is not None:
            import os
            _check(lib().nerf_create_from_cache(os.fsencode(cache), device, C.byref(h)))
        elif tensors is not None:
            keep = []

            def pack(d):
                arr = (Tensor * len(d))()
                for i, (name, a) in enumerate(d.items()):
                    a = _f32(a)
                    keep.append(a)
                    arr[i].name = name.encode()
                    arr[i].ndim = a.ndim
                    for k in range(a.ndim):
                        arr[i].dims[k] = a.shape[k]
                    arr[i].data = _p(a)
                return arr

            ca, fa = pack(tensors[0]), pack(tensors[1])
            _check(lib().nerf_create_from_tensors(ca, len(tensors[0]), fa, len(tensors[1]), device, C.byref(h)))
        else:
            import os
            _check(lib().nerf_create(os.fsencode(weights_root), device, C.byref(h)))
        self._h = h
        self.device = device
        self.coarse = Network(self, 0)
        self.fine = Network(self, 1)

    def close(self):
        if getattr(self, "_h", None):
            lib().nerf_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ render_image (lib.rs:474-565)
    def _sampling(self, mode, seed, jitter, u, keep):
        s = Sampling(mode=mode, seed=seed)
        if mode == SAMPLING_ARRAYS:
            if isinstance(jitter, int):
                s.jitter, s.u = jitter, u  # device pointers
            else:
                jitter, u = _f32(jitter), _f32(u)
                keep += [jitter, u]
                s.jitter, s.u = jitter.ctypes.data, u.ctypes.data
        return s

    def render_image(self, cam, W, H, n_coarse=64, n_fine=128, precision="mixed", mode=SAMPLING_DET, seed=0,
                     jitter=None, u=None, ray_begin=0, ray_end=None, stages=False, out=None):
        """Returns rgb [n, 3] for pixel indices [ray_begin, ray_end) (and the stage dict when stages=True)."""
        ray_end = W * H if ray_end is None else ray_end
        n = max(ray_end - ray_begin, 0)
        rgb = out if out is not None else np.empty((n, 3), np.float32)
        keep = []
        smp = self._sampling(mode, seed, jitter, u, keep)
        if not stages:
            _check(lib().nerf_render(self._h, C.byref(cam), W, H, n_coarse, n_fine, _prec(precision), C.byref(smp),
                                     ray_begin, ray_end, _p(rgb)))
            return rgb
        nm = n_coarse + n_fine
        dump = dict(dirs=np.zeros((n, 3), np.float32), t_c=np.zeros((n, n_coarse), np.float32),
                    sigma_c=np.zeros((n, n_coarse), np.float32), w_c=np.zeros((n, n_coarse), np.float32),
                    t_f=np.zeros((n, nm), np.float32), sigma_f=np.zeros((n, nm), np.float32),
                    rgb_f=np.zeros((n, nm, 3), np.float32))
        st = Stages(**{k: _p(v) for k, v in dump.items()})
        _check(lib().nerf_render_stages(self._h, C.byref(cam), W, H, n_coarse, n_fine, _prec(precision), C.byref(smp),
                                        ray_begin, ray_end, _p(rgb), C.byref(st)))
        return rgb, dump

    def render_device(self, cam, W, H, d_rgb, n_coarse=64, n_fine=128, precision="mixed", mode=SAMPLING_DET, seed=0,
                      d_jitter=0, d_u=0, ray_begin=0, ray_end=None, stream=0):
        """render_image into a device buffer (int pointer), enqueued on `stream` (int cudaStream_t)."""
        ray_end = W * H if ray_end is None else ray_end
        smp = Sampling(mode=mode, seed=seed, jitter=d_jitter or None, u=d_u or None)
        _check(lib().nerf_render_device(self._h, C.byref(cam), W, H, n_coarse, n_fine, _prec(precision), C.byref(smp),
                                        ray_begin, ray_end, d_rgb, stream))

    def render_striped(self, cam, W, H, frame, stripe_rays, parts, part, n_coarse=64, n_fine=128, precision="mixed",
                       mode=SAMPLING_DET, seed=0):
        """Renders part `part` of the frame dealt out in stripes of `stripe_rays` pixel indices to `parts` renderers
        (stripe k belongs to part k % parts) into the whole-frame host array `frame` [W*H, 3]; other pixels are left."""
        assert frame.dtype == np.float32 and frame.size == W * H * 3 and frame.flags["C_CONTIGUOUS"]
        smp = Sampling(mode=mode, seed=seed)
        _check(lib().nerf_render_striped(self._h, C.byref(cam), W, H, n_coarse, n_fine, _prec(precision), C.byref(smp),
                                         stripe_rays, parts, part, _p(frame)))
        return frame

    def render_striped_device(self, cam, W, H, d_frame, stripe_rays, parts, part, n_coarse=64, n_fine=128,
                              precision="mixed", mode=SAMPLING_DET, seed=0, stream=0):
        """The same into a whole-frame device buffer (int pointer), enqueued on `stream`."""
        smp = Sampling(mode=mode, seed=seed)
        _check(lib().nerf_render_striped_device(self._h, C.byref(cam), W, H, n_coarse, n_fine, _prec(precision),
                                                C.byref(smp), stripe_rays, parts, part, d_frame, stream))

    # ------------------------------------------------------------------ 8-bit output (save_ppm / pixels_to_rgba, lib.rs:567-592)
    def render_rgb8(self, cam, W, H, n_coarse=64, n_fine=128, precision="mixed", mode=SAMPLING_DET, seed=0,
                    jitter=None, u=None, ray_begin=0, ray_end=None, channels=3):
        """render_image + the save_ppm quantiser on the device: uint8 [n, channels] (3 = PPM body, 4 = RGBA)."""
        ray_end = W * H if ray_end is None else ray_end
        out = np.empty((max(ray_end - ray_begin, 0), channels), np.uint8)
        keep = []
        smp = self._sampling(mode, seed, jitter, u, keep)
        _check(lib().nerf_render_rgb8(self._h, C.byref(cam), W, H, n_coarse, n_fine, _prec(precision), C.byref(smp),
                                      ray_begin, ray_end, channels, out.ctypes.data_as(C.POINTER(C.c_uint8))))
        return out

    def render_many(self, cams, W, H, n_coarse=64, n_fine=128, precision="mixed", mode=SAMPLING_DET, seed=0, channels=3):
        """A batch of poses: uint8 [n_views, H * W, channels]; copies overlap the next view's render."""
        arr = (Camera * len(cams))(*cams)
        out = np.empty((len(cams), W * H, channels), np.uint8)
        smp = Sampling(mode=mode, seed=seed)
        _check(lib().nerf_render_many(self._h, arr, len(cams), W, H, n_coarse, n_fine, _prec(precision), C.byref(smp),
                                      channels, out.ctypes.data_as(C.POINTER(C.c_uint8))))
        return out

    def render_aux(self, cam, W, H, n_coarse=64, n_fine=128, precision="mixed", mode=SAMPLING_DET, seed=0,
                   jitter=None, u=None, ray_begin=0, ray_end=None):
        """render_image plus depth (sum w_i t_i) and accumulated opacity (sum w_i): (rgb [n,3], depth [n], acc [n])."""
        ray_end = W * H if ray_end is None else ray_end
        n = max(ray_end - ray_begin, 0)
        rgb, depth, acc = np.empty((n, 3), np.float32), np.empty(n, np.float32), np.empty(n, np.float32)
        keep = []
        smp = self._sampling(mode, seed, jitter, u, keep)
        _check(lib().nerf_render_aux(self._h, C.byref(cam), W, H, n_coarse, n_fine, _prec(precision), C.byref(smp),
                                     ray_begin, ray_end, _p(rgb), _p(depth), _p(acc)))
        return rgb, depth, acc

    def quantize_rgb8_gpu(self, rgb, channels=3):
        rgb = _f32(rgb).reshape(-1, 3)
        out = np.empty((rgb.shape[0], channels), np.uint8)
        _check(lib().nerf_quantize_rgb8_gpu(self._h, _p(rgb), rgb.shape[0], channels,
                                            out.ctypes.data_as(C.POINTER(C.c_uint8))))
        return out

    # ------------------------------------------------------------------ stage entry points
    def ray_dirs(self, cam, W, H, ray_begin=0, ray_end=None):
        ray_end = W * H if ray_end is None else ray_end
        out = np.empty((ray_end - ray_begin, 3), np.float32)
        _check(lib().nerf_ray_dirs(self._h, C.byref(cam), W, H, ray_begin, ray_end, _p(out)))
        return out

    def compute_weights(self, sigma, t, far):
        sigma, t = _f32(sigma), _f32(t)
        w = np.empty_like(t)
        _check(lib().nerf_compute_weights(self._h, _p(sigma), _p(t), t.shape[0], t.shape[1], far, _p(w)))
        return w

    def sample_importance(self, t, w, u):
        t, w, u = _f32(t), _f32(w), _f32(u)
        out = np.empty((t.shape[0], t.shape[1] + u.shape[1]), np.float32)
        _check(lib().nerf_sample_importance(self._h, _p(t), _p(w), t.shape[0], t.shape[1], _p(u), u.shape[1], _p(out)))
        return out

    def integrate_rays(self, rgb, sigma, t, far):
        rgb, sigma, t = _f32(rgb), _f32(sigma), _f32(t)
        out = np.empty((t.shape[0], 3), np.float32)
        _check(lib().nerf_integrate_rays(self._h, _p(rgb), _p(sigma), _p(t), t.shape[0], t.shape[1], far, _p(out)))
        return out

    def set_option(self, option, value):
        """nerf_set_option: e.g. (nerf_rs_b200.OPT_SKIP_EMPTY_RAYS, 1) - opt-in, not the reference's image."""
        _check(lib().nerf_set_option(self._h, int(option), int(value)))

    def get_option(self, option):
        v = C.c_int32(0)
        _check(lib().nerf_get_option(self._h, int(option), C.byref(v)))
        return v.value

    def timing(self):
        t = Timing()
        _check(lib().nerf_get_timing(self._h, C.byref(t)))
        return {k: getattr(t, k) for k, _ in Timing._fields_}
