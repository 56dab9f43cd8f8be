"""nerf_integrate_rays_noalloc (include/nerf_b200.h): the allocation-free variant of nerf_integrate_rays
(no counterpart in the reference).  Its device scratch is a file-scope static array of
NERF_NOALLOC_MAX_BYTES, so it never calls cudaMalloc; here it is called through the C ABI and compared with the existing
entry point (same kernel: the tolerance of test_integrate_rays, 3e-6, and in fact bit for bit) and with the CPU oracle,
on two calls in a row that share the array - the second smaller than the first, so that stale contents would show.

This file is named to be collected LAST: a failure here must not hide other tests.
It passes alone in a fresh process and after the other GPU tests (it only needs the session's renderer fixture).
"""
import ctypes as C

import numpy as np
import pytest

from oracle import pyoracle as O
import nerf_rs_b200 as nb

pytestmark = pytest.mark.gpu

TOL = 3e-6  # tests/test_gpu_parity.py::test_integrate_rays


def _call(fn, renderer, rgb, sigma, t, far):
    rgb, sigma, t = (np.ascontiguousarray(a, np.float32) for a in (rgb, sigma, t))
    out = np.full((t.shape[0], 3), np.nan, np.float32)
    rc = fn(renderer._h, nb._p(rgb), nb._p(sigma), nb._p(t), t.shape[0], t.shape[1], far, nb._p(out))
    return rc, out


def _rays(rng, rays, n):
    t = np.sort(rng.uniform(2, 6, (rays, n)).astype(np.float32), axis=1)
    sig = (rng.uniform(0, 1, (rays, n)) ** 8 * 200).astype(np.float32)
    rgb = rng.uniform(0, 1, (rays, n, 3)).astype(np.float32)
    return rgb, sig, t


def test_integrate_rays_noalloc_matches_the_existing_entry_point(renderer):
    L = nb.lib()
    rng = np.random.default_rng(31)
    # two calls in a row on the shared scratch array: 1 000 rays x 192 samples (3.85 MB of the 4 MiB), then a smaller
    # call with other sample counts whose segments land at other offsets inside what the first call left behind
    for rays, n in ((1000, 192), (37, 64), (300, 5)):
        rgb, sig, t = _rays(rng, rays, n)
        rc0, want = _call(L.nerf_integrate_rays, renderer, rgb, sig, t, 6.0)
        rc1, got = _call(L.nerf_integrate_rays_noalloc, renderer, rgb, sig, t, 6.0)
        assert rc0 == 0 and rc1 == 0, L.nerf_last_error()
        assert np.isfinite(got).all()
        assert np.abs(got - want).max() <= TOL
        assert (got == want).all()  # same kernel on the same values
        ref = np.stack([O.integrate_ray(rgb[i], sig[i], t[i], 6.0) for i in range(rays)])
        assert np.abs(got - ref).max() < TOL
    # the same inputs again give the same bytes (nothing accumulates in the array between calls)
    rc2, again = _call(L.nerf_integrate_rays_noalloc, renderer, rgb, sig, t, 6.0)
    assert rc2 == 0 and (again == got).all()
    # empty calls are no-ops, as in the existing entry point
    e = np.zeros((0, 7), np.float32)
    assert _call(L.nerf_integrate_rays_noalloc, renderer, np.zeros((0, 7, 3), np.float32), e, e, 6.0)[0] == 0


def test_integrate_rays_noalloc_rejects_what_does_not_fit(renderer):
    L = nb.lib()
    rng = np.random.default_rng(32)
    # 1 100 rays x 192 samples need 4.24 MB > NERF_NOALLOC_MAX_BYTES: refused, nothing written, clear message
    rgb, sig, t = _rays(rng, 1100, 192)
    assert 1100 * (192 * 20 + 12) > nb.NOALLOC_MAX_BYTES
    rc, out = _call(L.nerf_integrate_rays_noalloc, renderer, rgb, sig, t, 6.0)
    assert rc == nb.E_UNSUPPORTED and np.isnan(out).all()
    msg = L.nerf_last_error().decode()
    assert "NERF_NOALLOC_MAX_BYTES" in msg and str(nb.NOALLOC_MAX_BYTES) in msg and "1100 rays" in msg
    # sizes whose product would wrap 64 bits are refused too, before any arithmetic on them
    f = C.c_float(0.0)
    p = C.cast(C.pointer(f), C.POINTER(C.c_float))
    assert L.nerf_integrate_rays_noalloc(renderer._h, p, p, p, 2 ** 62, 2 ** 30, 6.0, p) == nb.E_UNSUPPORTED
    assert L.nerf_integrate_rays_noalloc(None, p, p, p, 1, 1, 6.0, p) == nb.E_ARG
    # the existing entry point still takes the large call, and the variant works right after its refusals
    rc, want = _call(L.nerf_integrate_rays, renderer, rgb, sig, t, 6.0)
    assert rc == 0
    rc, got = _call(L.nerf_integrate_rays_noalloc, renderer, rgb[:1000], sig[:1000], t[:1000], 6.0)
    assert rc == 0 and (got == want[:1000]).all()
