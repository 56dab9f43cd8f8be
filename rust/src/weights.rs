//! weights.rs — where the networks' tensors come from.  The reference's file of this name is the wasm build's embedded
//! table (src/weights.rs:1-100: include_bytes! of every lego_rust/*.bin); the native loaders sit in lib.rs:34-92.
//! Here both routes end in a GPU context:
//!   * `open_scene(root)`            -> nerf_create (the library reads <root>/coarse and <root>/fine itself), or
//!   * `open_scene_cached(root)`     -> nerf_build_weight_cache once + nerf_create_from_cache (one sequential read of the
//!                                      pre-swizzled tensor-core tiles instead of 48 files and the tile builders), and
//!   * `read_layers(dir)`            -> caller-held `Layer`s for `Network::new` (load_shapes + load_tensor restated).
use crate::ffi;
use crate::network::{Context, Layer, Matrix};
use std::collections::HashMap;
use std::ffi::CString;
use std::fs;
use std::path::Path;
use std::sync::{Arc, Mutex, OnceLock, Weak};

fn cstr(p: &Path) -> CString {
    CString::new(p.to_str().expect("utf-8 path")).unwrap()
}

/// One context per weight root: `load_network_from_dir(root/coarse)` and `(root/fine)` share it.
pub(crate) fn open_scene(root: &Path) -> Arc<Context> {
    static SCENES: OnceLock<Mutex<HashMap<String, Weak<Context>>>> = OnceLock::new();
    let mut scenes = SCENES.get_or_init(|| Mutex::new(HashMap::new())).lock().unwrap();
    let key = root.to_string_lossy().into_owned();
    if let Some(ctx) = scenes.get(&key).and_then(Weak::upgrade) {
        return ctx;
    }
    let mut raw = std::ptr::null_mut();
    let cache = root.join("weights.nrfw");
    if cache.exists() {
        ffi::check(unsafe { ffi::nerf_create_from_cache(cstr(&cache).as_ptr(), 0, &mut raw) }, "nerf_create_from_cache");
    } else {
        ffi::check(unsafe { ffi::nerf_create(cstr(root).as_ptr(), 0, &mut raw) }, "load_network_from_dir");
    }
    let ctx = Arc::new(Context { raw });
    scenes.insert(key, Arc::downgrade(&ctx));
    ctx
}

/// Writes <root>/weights.nrfw (host only, no GPU); `open_scene` picks it up from then on.
pub fn build_weight_cache(root: &Path) {
    let out = root.join("weights.nrfw");
    ffi::check(unsafe { ffi::nerf_build_weight_cache(cstr(root).as_ptr(), cstr(&out).as_ptr()) }, "nerf_build_weight_cache");
}

/// shapes.txt: one `name dim0 [dim1]` line per tensor (lib.rs:63-74).
pub fn load_shapes(path: &Path) -> Vec<(String, Vec<usize>)> {
    let text = fs::read_to_string(path).unwrap_or_else(|e| panic!("read {}: {e}", path.display()));
    text.lines()
        .filter(|l| !l.trim().is_empty())
        .map(|l| {
            let mut it = l.split_whitespace();
            let name = it.next().unwrap().to_string();
            (name, it.map(|d| d.parse().expect("dimension")).collect())
        })
        .collect()
}

/// <name>.bin: little-endian f32, row-major (lib.rs:35-42).
pub fn load_tensor(path: &Path, dims: &[usize]) -> Vec<f32> {
    let bytes = fs::read(path).unwrap_or_else(|e| panic!("read {}: {e}", path.display()));
    let count: usize = dims.iter().product();
    assert_eq!(bytes.len(), count * 4, "{}: size does not match shapes.txt", path.display());
    bytes.chunks_exact(4).map(|b| f32::from_le_bytes([b[0], b[1], b[2], b[3]])).collect()
}

/// The twelve layers of <dir> in the order `Network::new` takes them: dense0..7, bottleneck, viewdirs, rgb, alpha.
pub fn read_layers(dir: &Path) -> Vec<Layer> {
    let shapes: HashMap<String, Vec<usize>> = load_shapes(&dir.join("shapes.txt")).into_iter().collect();
    let layer = |stem: &str| {
        let kd = shapes.get(&format!("{stem}_kernel")).unwrap_or_else(|| panic!("missing tensor {stem}_kernel"));
        let bd = shapes.get(&format!("{stem}_bias")).unwrap_or_else(|| panic!("missing tensor {stem}_bias"));
        let w = load_tensor(&dir.join(format!("{stem}_kernel.bin")), kd);
        let b = load_tensor(&dir.join(format!("{stem}_bias.bin")), bd);
        Layer::new(Matrix::new(w, kd[0] as i32, kd[1] as i32), b)
    };
    let mut out: Vec<Layer> = (0..8).map(|i| layer(&format!("dense{i}"))).collect();
    for stem in ["bottleneck", "viewdirs", "rgb", "alpha"] {
        out.push(layer(stem));
    }
    out
}
