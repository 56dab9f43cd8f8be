//! lib.rs — the nerf-rs crate's surface (src/lib.rs of elisabeth96/nerf-rs) over libnerf_b200.so.
//!
//! Same names and argument meaning as the reference: `load_network_from_dir` (lib.rs:108), `Camera` /
//! `camera_from_samples` (lib.rs:197,614), `render_image` (lib.rs:474), `save_ppm` (lib.rs:567),
//! `default_sample_counts` (lib.rs:603), `render_cli_image` (lib.rs:648), and the unit test of lib.rs:728-917.
//! What the reference does per 8x8 pixel block on the host (ray generation, stratified sampling, coarse network,
//! compute_weights, sample_importance, merge + sort, fine network, integrate_ray: lib.rs:176-472) is one call,
//! `nerf_render`, into sm_100a kernels.  There is no CPU arithmetic left here and no fallback: without a B200 every
//! entry point panics with the library's message.
//!
//! NOT compiled in this repository's build image (it has no Rust toolchain); rust/README.md says how to build it.
pub mod ffi;
pub mod network;
pub mod vec3;
pub mod weights;

use network::{Matrix, Network};
use serde_json::Value;
use std::ffi::CString;
use std::fs;
use std::path::Path;
use std::sync::{Arc, Mutex, OnceLock};
use std::time::Instant;
use vec3::Vec3;

/// lib.rs:108-174.  `dir` is <root>/coarse or <root>/fine, as the reference calls it (lib.rs:651-652).
pub fn load_network_from_dir(dir: &Path) -> Network {
    let leaf = dir.file_name().and_then(|s| s.to_str()).unwrap_or("");
    let net_id = match leaf {
        "coarse" => ffi::NET_COARSE,
        "fine" => ffi::NET_FINE,
        _ => panic!("load_network_from_dir: expected <root>/coarse or <root>/fine, got {}", dir.display()),
    };
    let root = dir.parent().unwrap_or_else(|| Path::new("."));
    Network::from_context(weights::open_scene(root), net_id)
}

pub fn load_tf_samples(path: &Path) -> Result<Value, Box<dyn std::error::Error>> {
    Ok(serde_json::from_str(&fs::read_to_string(path)?)?) // lib.rs:94-99
}

/// lib.rs:197-211.
pub struct Camera {
    pub nx: usize,
    pub ny: usize,
    pub alpha_width: f32,
    pub alpha_height: f32,
    pub pos: Vec3,
    pub dir: Vec3,
    pub up: Vec3,
    pub near: f32,
    pub far: f32,
    pub samples_per_ray: usize,
}

fn get_vec3(node: &Value, key: &str) -> Vec3 {
    let a = node[key].as_array().unwrap_or_else(|| panic!("{key}"));
    Vec3::new(a[0].as_f64().unwrap() as f32, a[1].as_f64().unwrap() as f32, a[2].as_f64().unwrap() as f32)
}

pub fn default_sample_counts() -> (usize, usize) {
    (64, 128) // lib.rs:608-611 (native target)
}

/// lib.rs:614-645: near, far, camera_origin, camera_forward, camera_up, hwf of the samples file.
pub fn camera_from_samples(samples: &Value, width: usize, height: usize, coarse_samples_per_ray: usize) -> Camera {
    let hwf = samples["hwf"].as_array().expect("hwf");
    let (hh, hw, focal) = (hwf[0].as_f64().unwrap() as f32, hwf[1].as_f64().unwrap() as f32, hwf[2].as_f64().unwrap() as f32);
    Camera {
        nx: width,
        ny: height,
        alpha_width: ((0.5 * hw) / focal).atan(),
        alpha_height: ((0.5 * hh) / focal).atan(),
        pos: get_vec3(samples, "camera_origin"),
        dir: get_vec3(samples, "camera_forward").normalize(),
        up: get_vec3(samples, "camera_up").normalize(),
        near: samples["near"].as_f64().unwrap() as f32,
        far: samples["far"].as_f64().unwrap() as f32,
        samples_per_ray: coarse_samples_per_ray,
    }
}

/// How the jitter of stratified_samples / sample_importance (lib.rs:243,332) is drawn.  The reference uses an
/// OS-seeded thread_rng, which nothing can reproduce; the library offers Philox keyed by (seed, pixel, stream) - the
/// default, seed 0 - or no jitter at all (what the parity tests compare).
#[derive(Clone, Copy)]
pub enum Jitter {
    Deterministic,
    Philox(u64),
}

static JITTER: Mutex<Jitter> = Mutex::new(Jitter::Philox(0));
pub fn set_jitter(j: Jitter) {
    *JITTER.lock().unwrap() = j;
}

/// lib.rs:474-565.
pub fn render_image(coarse: &Network, fine: &Network, camera: &Camera, fine_samples_per_ray: usize) -> Vec<Vec3> {
    // the asserts of lib.rs:483-501 (samples > 0, width and height multiples of 8) come back as NERF_E_ARG -> panic
    let ctx = match (&coarse.host, &fine.host) {
        (Some(c), Some(f)) => pair_context(c, f),
        _ => {
            assert!(
                Arc::ptr_eq(&coarse.ctx, &fine.ctx) && coarse.net_id == ffi::NET_COARSE && fine.net_id == ffi::NET_FINE,
                "render_image: coarse and fine must come from the same weight root"
            );
            coarse.ctx.clone()
        }
    };
    let cam = ffi::NerfCamera {
        origin: camera.pos.to_array(),
        forward: camera.dir.to_array(),
        up: camera.up.to_array(),
        alpha_w: camera.alpha_width,
        alpha_h: camera.alpha_height,
        near: camera.near,
        far: camera.far,
    };
    let smp = match *JITTER.lock().unwrap() {
        Jitter::Deterministic => ffi::NerfSampling { mode: 0, seed: 0, jitter: std::ptr::null(), u: std::ptr::null() },
        Jitter::Philox(seed) => ffi::NerfSampling { mode: 1, seed, jitter: std::ptr::null(), u: std::ptr::null() },
    };
    let n = camera.nx * camera.ny;
    let mut image = vec![Vec3::new(0.0, 0.0, 0.0); n];
    ffi::check(
        unsafe {
            ffi::nerf_render(
                ctx.raw,
                &cam,
                camera.nx as i32,
                camera.ny as i32,
                camera.samples_per_ray as i32,
                fine_samples_per_ray as i32,
                fine.precision.unwrap_or(ffi::PREC_MIXED),
                &smp,
                0,
                n as i64,
                image.as_mut_ptr() as *mut f32,
            )
        },
        "render_image",
    );
    image
}

/// Two networks built with `Network::new`: the pair's own context, built once and kept.
fn pair_context(coarse: &Arc<network::HostNet>, fine: &Arc<network::HostNet>) -> Arc<network::Context> {
    type Key = (usize, usize);
    static PAIRS: OnceLock<Mutex<Vec<(Key, Arc<network::Context>)>>> = OnceLock::new();
    let key = (Arc::as_ptr(coarse) as usize, Arc::as_ptr(fine) as usize);
    let mut pairs = PAIRS.get_or_init(|| Mutex::new(Vec::new())).lock().unwrap();
    if let Some((_, ctx)) = pairs.iter().find(|(k, _)| *k == key) {
        return ctx.clone();
    }
    let ctx = network::context_from_hosts(coarse, fine, 0);
    pairs.push((key, ctx.clone()));
    ctx
}

/// lib.rs:567-580: "P6\n{w} {h}\n255\n" + (clamp(v, 0, 1) * 255 + 0.5) as u8, byte-identical to the reference's.
pub fn save_ppm(path: &Path, width: usize, height: usize, pixels: &[Vec3]) -> std::io::Result<()> {
    assert_eq!(pixels.len(), width * height); // lib.rs:569
    let c = CString::new(path.to_str().expect("utf-8 path")).unwrap();
    let rc = unsafe { ffi::nerf_save_ppm(c.as_ptr(), width as i32, height as i32, pixels.as_ptr() as *const f32) };
    if rc != 0 {
        return Err(std::io::Error::new(std::io::ErrorKind::Other, "nerf_save_ppm failed"));
    }
    Ok(())
}

/// lib.rs:647-677, line for line in behaviour: lego_rust next to the crate, 256 x 256, output.ppm in the working
/// directory, the same three progress lines, only `render_image` inside the timed region.
pub fn render_cli_image() {
    let root = Path::new(env!("NERF_RS_LEGO_ROOT")); // the reference bakes in CARGO_MANIFEST_DIR/lego_rust (lib.rs:649-650)
    let coarse_network = load_network_from_dir(&root.join("coarse"));
    let fine_network = load_network_from_dir(&root.join("fine"));
    let samples = load_tf_samples(&root.join("tf_reference_samples.json")).unwrap();
    let (coarse_samples_per_ray, fine_samples_per_ray) = default_sample_counts();
    let (width, height) = (256, 256);
    println!("Rendering with {} coarse samples and {} fine samples per ray", coarse_samples_per_ray, fine_samples_per_ray);
    let cam = camera_from_samples(&samples, width, height, coarse_samples_per_ray);
    println!("Starting image rendering...");
    let start = Instant::now();
    let image = render_image(&coarse_network, &fine_network, &cam, fine_samples_per_ray);
    println!("Rendering completed in {:.2} seconds", start.elapsed().as_secs_f64());
    save_ppm(Path::new("output.ppm"), cam.nx, cam.ny, &image).unwrap();
}

#[cfg(test)]
mod tests {
    //! The reference's own unit test (lib.rs:728-917), same shape: both networks from their directories, three example
    //! rays x five sample positions, `forward_batch` on the coarse and on the fine network, every sigma and every rgb
    //! channel within 1e-2.  The reference spells the expected numbers as literals (lib.rs:759-847); they are the
    //! `examples` of lego_rust/tf_reference_samples.json, which this test reads instead.
    use super::*;

    fn assert_close(label: &str, expected: f32, actual: f32) {
        let diff = (expected - actual).abs();
        assert!(diff < 1e-2, "{label} mismatch: expected {expected} got {actual} (diff {diff})");
    }

    fn floats(v: &Value) -> Vec<f32> {
        v.as_array().unwrap().iter().map(|x| x.as_f64().unwrap() as f32).collect()
    }

    fn check_network(name: &str, net: &Network, points: &Matrix, dirs: &[Vec3], ex: &Value, idx: usize) {
        let (colors, sigmas) = net.forward_batch(points, dirs);
        for (s, &want) in floats(&ex[format!("{name}_sigma")]).iter().enumerate() {
            assert_close(&format!("{name}[{idx}].sigma[{s}]"), want, sigmas[s]);
        }
        for (s, color) in colors.iter().enumerate() {
            let want = floats(&ex[format!("{name}_rgb")][s]);
            assert_close(&format!("{name}[{idx}].color[{s}].r"), want[0], color.x);
            assert_close(&format!("{name}[{idx}].color[{s}].g"), want[1], color.y);
            assert_close(&format!("{name}[{idx}].color[{s}].b"), want[2], color.z);
        }
    }

    #[test]
    fn coarse_and_fine_match_reference_examples() {
        let root = Path::new(env!("NERF_RS_LEGO_ROOT"));
        let coarse = load_network_from_dir(&root.join("coarse"));
        let fine = load_network_from_dir(&root.join("fine"));
        let samples = load_tf_samples(&root.join("tf_reference_samples.json")).unwrap();
        let origin = get_vec3(&samples, "camera_origin");
        let sample_positions = floats(&samples["z_vals"]); // 2, 3, 4, 5, 6 (lib.rs:762)
        for (idx, ex) in samples["examples"].as_array().unwrap().iter().enumerate() {
            let ray_dir = get_vec3(ex, "ray_d"); // un-normalised, as TensorFlow NeRF steps along it (lib.rs:855)
            let view_dir = get_vec3(ex, "viewdir_unit");
            let mut points = Matrix::zeros(3, sample_positions.len() as i32);
            let mut view_dirs = Vec::with_capacity(sample_positions.len());
            for (col, &t) in sample_positions.iter().enumerate() {
                let p = origin + ray_dir * t;
                points.set(0, col, p.x);
                points.set(1, col, p.y);
                points.set(2, col, p.z);
                view_dirs.push(view_dir);
            }
            check_network("coarse", &coarse, &points, &view_dirs, ex, idx);
            check_network("fine", &fine, &points, &view_dirs, ex, idx);
        }
    }

    #[test]
    fn network_new_matches_load_network_from_dir() {
        let root = Path::new(env!("NERF_RS_LEGO_ROOT"));
        let build = |dir: &str| {
            let mut l = weights::read_layers(&root.join(dir));
            let alpha = l.pop().unwrap();
            let rgb = l.pop().unwrap();
            let viewdirs = l.pop().unwrap();
            let bottleneck = l.pop().unwrap();
            Network::new(l, bottleneck, viewdirs, rgb, alpha)
        };
        let (coarse, fine) = (build("coarse"), build("fine"));
        let loaded = load_network_from_dir(&root.join("fine"));
        let points = Matrix::new(vec![0.1, -0.2, 0.3, 0.0, 0.4, -0.1], 3, 2);
        let dirs = [Vec3::new(0.6, 0.0, -0.8); 2];
        assert_eq!(fine.forward_batch(&points, &dirs).1, loaded.forward_batch(&points, &dirs).1);
        set_jitter(Jitter::Deterministic);
        let samples = load_tf_samples(&root.join("tf_reference_samples.json")).unwrap();
        let cam = camera_from_samples(&samples, 32, 24, 64);
        let a = render_image(&coarse, &fine, &cam, 128);
        let b = render_image(&load_network_from_dir(&root.join("coarse")), &loaded, &cam, 128);
        assert!(a == b);
    }
}
