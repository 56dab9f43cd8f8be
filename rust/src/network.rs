//! network.rs — `Matrix`, `Layer`, `Network` with the reference's signatures (src/network.rs:3-60,62-78,172-237).
//! `Network::forward_batch` no longer runs positional encodings and a chain of `Layer::forward_matrix` on the host:
//! it is one call into the fused tcgen05 kernel behind `nerf_forward_batch`.
use crate::ffi;
use crate::vec3::Vec3;
use std::ffi::CString;
use std::os::raw::c_int;
use std::sync::Arc;

#[derive(Clone)]
pub struct Matrix {
    elements: Vec<f32>,
    rows: i32,
    cols: i32,
}

impl Matrix {
    pub fn new(elements: Vec<f32>, rows: i32, cols: i32) -> Self {
        debug_assert_eq!(elements.len(), (rows * cols) as usize);
        Matrix { elements, rows, cols }
    }
    pub fn zeros(rows: i32, cols: i32) -> Self {
        Matrix { elements: vec![0.0; (rows * cols) as usize], rows, cols }
    }
    pub fn rows(&self) -> usize {
        self.rows as usize
    }
    pub fn cols(&self) -> usize {
        self.cols as usize
    }
    pub fn data(&self) -> &[f32] {
        &self.elements
    }
    pub fn data_mut(&mut self) -> &mut [f32] {
        &mut self.elements
    }
    pub fn get(&self, row: usize, col: usize) -> f32 {
        self.elements[row * self.cols() + col]
    }
    pub fn set(&mut self, row: usize, col: usize, value: f32) {
        let c = self.cols();
        self.elements[row * c + col] = value;
    }
    pub fn row_slice(&self, row: usize) -> &[f32] {
        let c = self.cols();
        &self.elements[row * c..(row + 1) * c]
    }
}

/// A dense layer as the caller holds it: weight is [in x out] row-major (the `_kernel.bin` layout), bias has `out`
/// entries (network.rs:62-78).  Host data only.
pub struct Layer {
    pub(crate) weight: Matrix,
    pub(crate) bias: Vec<f32>,
}

impl Layer {
    pub fn new(weight: Matrix, bias: Vec<f32>) -> Self {
        debug_assert_eq!(weight.cols(), bias.len());
        Layer { weight, bias }
    }
}

/// Owner of one `nerf_ctx` (device weights of the coarse AND the fine network of a scene, workspaces, streams).
pub(crate) struct Context {
    pub(crate) raw: *mut ffi::NerfCtx,
}
// nerf_b200.h: a context is internally serialised and may be used from any host thread
unsafe impl Send for Context {}
unsafe impl Sync for Context {}
impl Drop for Context {
    fn drop(&mut self) {
        unsafe { ffi::nerf_destroy(self.raw) }
    }
}

/// The tensors of one network under the names of shapes.txt, kept on the host for `Network::new`.
pub(crate) struct HostNet {
    pub(crate) layers: Vec<Layer>, // dense0..7, bottleneck, viewdirs, rgb, alpha
}

const STEMS: [&str; 12] = [
    "dense0", "dense1", "dense2", "dense3", "dense4", "dense5", "dense6", "dense7", "bottleneck", "viewdirs", "rgb",
    "alpha",
];

impl HostNet {
    /// (names kept alive, tensor table) for nerf_create_from_tensors
    fn table(&self) -> (Vec<CString>, Vec<ffi::NerfTensor>) {
        let mut names = Vec::new();
        let mut t = Vec::new();
        for (layer, stem) in self.layers.iter().zip(STEMS.iter()) {
            names.push(CString::new(format!("{stem}_kernel")).unwrap());
            t.push(ffi::NerfTensor {
                name: names.last().unwrap().as_ptr(),
                ndim: 2,
                dims: [layer.weight.rows() as i64, layer.weight.cols() as i64],
                data: layer.weight.data().as_ptr(),
            });
            names.push(CString::new(format!("{stem}_bias")).unwrap());
            t.push(ffi::NerfTensor {
                name: names.last().unwrap().as_ptr(),
                ndim: 1,
                dims: [layer.bias.len() as i64, 0],
                data: layer.bias.as_ptr(),
            });
        }
        (names, t)
    }
}

pub(crate) fn context_from_hosts(coarse: &HostNet, fine: &HostNet, device: c_int) -> Arc<Context> {
    let (_kc, tc) = coarse.table();
    let (_kf, tf) = fine.table();
    let mut raw = std::ptr::null_mut();
    ffi::check(
        unsafe { ffi::nerf_create_from_tensors(tc.as_ptr(), tc.len() as c_int, tf.as_ptr(), tf.len() as c_int, device, &mut raw) },
        "Network::new",
    );
    Arc::new(Context { raw })
}

pub struct Network {
    pub(crate) ctx: Arc<Context>,
    pub(crate) net_id: c_int,
    /// `None`: forward_batch runs FP16X3 (passes the crate's unit test with margin), render_image runs MIXED
    pub(crate) precision: Option<c_int>,
    pub(crate) host: Option<Arc<HostNet>>,
}

impl Network {
    /// network.rs:181-195.  A GPU context always holds the two networks of a scene, so a network built on its own
    /// sits in both slots of its context; `render_image` builds the coarse + fine context of a pair on first use.
    pub fn new(layers: Vec<Layer>, bottleneck: Layer, viewdirs: Layer, rgb: Layer, alpha: Layer) -> Self {
        assert_eq!(layers.len(), 8, "the trunk has eight layers (network.rs:206-213)");
        let mut all = layers;
        all.extend([bottleneck, viewdirs, rgb, alpha]);
        let host = Arc::new(HostNet { layers: all });
        let ctx = context_from_hosts(&host, &host, 0);
        Network { ctx, net_id: ffi::NET_FINE, precision: None, host: Some(host) }
    }

    pub(crate) fn from_context(ctx: Arc<Context>, net_id: c_int) -> Self {
        Network { ctx, net_id, precision: None, host: None }
    }

    /// Arithmetic of the 256-wide layers (ffi::PREC_*).
    pub fn set_precision(&mut self, precision: c_int) {
        self.precision = Some(precision);
    }

    /// network.rs:197-237: `points` is 3 x B (a column per sample), `view_dirs` has B unit vectors.
    pub fn forward_batch(&self, points: &Matrix, view_dirs: &[Vec3]) -> (Vec<Vec3>, Vec<f32>) {
        let batch = points.cols();
        if batch == 0 {
            return (Vec::new(), Vec::new()); // network.rs:199-201
        }
        debug_assert_eq!(points.rows(), 3);
        debug_assert_eq!(batch, view_dirs.len()); // network.rs:202
        let mut rgb = vec![Vec3::new(0.0, 0.0, 0.0); batch];
        let mut sigma = vec![0.0f32; batch];
        ffi::check(
            unsafe {
                ffi::nerf_forward_batch(
                    self.ctx.raw,
                    self.net_id,
                    points.data().as_ptr(), // Matrix 3 x B row-major == [3][B] feature-major
                    view_dirs.as_ptr() as *const f32,
                    batch as i64,
                    self.precision.unwrap_or(ffi::PREC_FP16X3),
                    rgb.as_mut_ptr() as *mut f32,
                    sigma.as_mut_ptr(),
                )
            },
            "forward_batch",
        );
        (rgb, sigma)
    }
}
