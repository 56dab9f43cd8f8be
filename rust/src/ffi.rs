//! ffi.rs — the C ABI of libnerf_b200.so (include/nerf_b200.h), declared by hand: one `extern` per entry point the
//! crate binds.  Struct layouts mirror the header's (`#[repr(C)]`).
use std::ffi::CStr;
use std::os::raw::{c_char, c_int, c_void};

#[repr(C)]
pub struct NerfCtx {
    _opaque: [u8; 0],
}

/// nerf_camera: struct Camera after camera_from_samples (lib.rs:197-211,614-645).
#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct NerfCamera {
    pub origin: [f32; 3],
    pub forward: [f32; 3],
    pub up: [f32; 3],
    pub alpha_w: f32,
    pub alpha_h: f32,
    pub near: f32,
    pub far: f32,
}

/// nerf_sampling: mode 0 = deterministic, 1 = Philox(seed, pixel, stream), 2 = caller arrays.
#[repr(C)]
pub struct NerfSampling {
    pub mode: i32,
    pub seed: u64,
    pub jitter: *const f32,
    pub u: *const f32,
}

#[repr(C)]
pub struct NerfTensor {
    pub name: *const c_char,
    pub ndim: i32,
    pub dims: [i64; 2],
    pub data: *const f32,
}

// nerf_precision
pub const PREC_FP32: c_int = 0;
pub const PREC_BF16: c_int = 1;
pub const PREC_FP16: c_int = 2;
pub const PREC_TF32: c_int = 3;
pub const PREC_FP16X3: c_int = 4;
pub const PREC_MIXED: c_int = 5;
pub const OPT_SKIP_EMPTY_RAYS: c_int = 1;
pub const NET_COARSE: c_int = 0;
pub const NET_FINE: c_int = 1;

extern "C" {
    pub fn nerf_last_error() -> *const c_char;
    pub fn nerf_create(root: *const c_char, device: c_int, out: *mut *mut NerfCtx) -> c_int;
    pub fn nerf_create_from_tensors(
        coarse: *const NerfTensor,
        n_coarse: c_int,
        fine: *const NerfTensor,
        n_fine: c_int,
        device: c_int,
        out: *mut *mut NerfCtx,
    ) -> c_int;
    pub fn nerf_create_from_cache(cache: *const c_char, device: c_int, out: *mut *mut NerfCtx) -> c_int;
    pub fn nerf_build_weight_cache(root: *const c_char, cache: *const c_char) -> c_int;
    pub fn nerf_destroy(ctx: *mut NerfCtx);
    pub fn nerf_forward_batch(
        ctx: *mut NerfCtx,
        net: c_int,
        points: *const f32,
        dirs: *const f32,
        batch: i64,
        precision: c_int,
        rgb: *mut f32,
        sigma: *mut f32,
    ) -> c_int;
    pub fn nerf_render(
        ctx: *mut NerfCtx,
        cam: *const NerfCamera,
        w: c_int,
        h: c_int,
        n_coarse: c_int,
        n_fine: c_int,
        precision: c_int,
        smp: *const NerfSampling,
        ray_begin: i64,
        ray_end: i64,
        rgb: *mut f32,
    ) -> c_int;
    pub fn nerf_render_rgb8(
        ctx: *mut NerfCtx,
        cam: *const NerfCamera,
        w: c_int,
        h: c_int,
        n_coarse: c_int,
        n_fine: c_int,
        precision: c_int,
        smp: *const NerfSampling,
        ray_begin: i64,
        ray_end: i64,
        channels: c_int,
        out: *mut u8,
    ) -> c_int;
    pub fn nerf_render_device(
        ctx: *mut NerfCtx,
        cam: *const NerfCamera,
        w: c_int,
        h: c_int,
        n_coarse: c_int,
        n_fine: c_int,
        precision: c_int,
        smp: *const NerfSampling,
        ray_begin: i64,
        ray_end: i64,
        d_rgb: *mut f32,
        stream: *mut c_void,
    ) -> c_int;
    /// Part `part` of the frame dealt out in stripes of `stripe_rays` pixel indices to `parts` renderers (stripe k
    /// belongs to part k % parts) into the WHOLE host frame: what several GPUs driven from one process call, one
    /// context and one thread each (rayon's role in the reference, lib.rs:533-550).
    pub fn nerf_render_striped(
        ctx: *mut NerfCtx,
        cam: *const NerfCamera,
        w: c_int,
        h: c_int,
        n_coarse: c_int,
        n_fine: c_int,
        precision: c_int,
        smp: *const NerfSampling,
        stripe_rays: i64,
        parts: c_int,
        part: c_int,
        frame: *mut f32,
    ) -> c_int;
    /// integrate_ray (lib.rs:176-195) over host arrays: rgb [n_rays][n][3], sigma and t [n_rays][n], out [n_rays][3].
    pub fn nerf_integrate_rays(
        ctx: *mut NerfCtx,
        rgb: *const f32,
        sigma: *const f32,
        t: *const f32,
        n_rays: i64,
        n: c_int,
        far: f32,
        out: *mut f32,
    ) -> c_int;
    /// The same without any device allocation (static 4 MiB device scratch; larger calls return NERF_E_UNSUPPORTED).
    /// No counterpart in the reference.
    pub fn nerf_integrate_rays_noalloc(
        ctx: *mut NerfCtx,
        rgb: *const f32,
        sigma: *const f32,
        t: *const f32,
        n_rays: i64,
        n: c_int,
        far: f32,
        out: *mut f32,
    ) -> c_int;
    /// Context options (nerf_option).  OPT_SKIP_EMPTY_RAYS is opt-in and NOT the reference's image: rays without any
    /// coarse weight get the background and skip the fine pass (include/nerf_b200.h).
    pub fn nerf_set_option(ctx: *mut NerfCtx, option: c_int, value: c_int) -> c_int;
    pub fn nerf_get_option(ctx: *mut NerfCtx, option: c_int, value: *mut c_int) -> c_int;
    pub fn nerf_camera_from_json(path: *const c_char, out: *mut NerfCamera) -> c_int;
    pub fn nerf_save_ppm(path: *const c_char, w: c_int, h: c_int, rgb: *const f32) -> c_int;
}

/// The reference panics on every error (lib.rs:36,118,483-501); so does the shim, with the library's message.
pub fn check(rc: c_int, what: &str) {
    if rc != 0 {
        let msg = unsafe { CStr::from_ptr(nerf_last_error()) }.to_string_lossy().into_owned();
        panic!("{what}: {msg}");
    }
}
