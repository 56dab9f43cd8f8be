/*
 * nerf_b200.h — C ABI of libnerf_b200.so: the B200 (sm_100a) replacement for the hot path of
 * elisabeth96/nerf-rs (per-sample NeRF MLP inference + ray marching + compositing over the coarse and
 * fine lego_rust networks).
 *
 * The reference is a pure-Rust crate without any FFI surface; the seam is its Rust API.  Each entry point
 * below names the reference function it stands in for (file:line into the reference crate).  The Rust
 * shim a maintainer would add is shown in INTEGRATION.md.
 *
 * Conventions
 *   - Every function returns NERF_OK (0) or a negative nerf_status; nothing unwinds across the ABI
 *     (the reference panics: lib.rs:36,118,483-501).  nerf_last_error() gives the message (thread-local).
 *   - The caller owns every buffer passed in or out.  A nerf_ctx owns device weights, workspaces and
 *     streams; it is internally serialised and may be used from any host thread.
 *   - The library is CUDA-only: there is no CPU fallback.  nerf_create fails with NERF_E_CUDA when no
 *     sm_100 device is present.
 *   - "host" entry points take host pointers and include H2D/D2H copies; "*_device" entry points take
 *     device pointers on the context's device and enqueue on `stream` (a cudaStream_t passed as void*),
 *     returning without synchronising.  A context has ONE set of workspaces: calls on different streams are ordered
 *     on the device in the order they were enqueued (the next call's stream waits for an event the previous call
 *     recorded), so they never overlap; use one context per stream for concurrency.
 *   - Device memory of a context: the weights (15 MB in all operand formats) plus workspaces that grow with the
 *     largest call, about 4.7 KB per ray of a render chunk at 64 + 128 samples (1 Mi rays at most: NERF_B200_CHUNK_RAYS) - 3.0 GB
 *     for an 800x800 frame - and, for renders in the 16-bit tensor-core modes, the hand-off buffer of the two-phase fine pass:
 *     528 B per fine sample of a group of rays (40 960 rays x 192 samples = 4.2 GB; NERF_B200_FINE_SUB_RAYS sets the
 *     group, NERF_B200_TWO_PHASE=0 renders with one fused launch and no such buffer).
 */
#ifndef NERF_B200_H
#define NERF_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  NERF_OK = 0,
  NERF_E_ARG = -1,   /* bad argument: W%8, H%8, n_coarse==0 (lib.rs:483-501), null pointers, ranges */
  NERF_E_IO = -2,    /* weight / camera / ppm file could not be read or written (lib.rs:36,65,569) */
  NERF_E_SHAPE = -3, /* missing tensor or dims mismatch (lib.rs:118,127) */
  NERF_E_CUDA = -4,  /* CUDA runtime error, or no sm_100 device */
  NERF_E_UNSUPPORTED = -5
} nerf_status;

/* Arithmetic of the 256-wide dense layers.  Heads (sigma, rgb), encodings, sampling and compositing are
 * always f32. */
typedef enum {
  NERF_PREC_FP32 = 0, /* CUDA-core path; per-element sum order identical to network.rs:134-143 */
  NERF_PREC_BF16 = 1, /* tcgen05 kind::f16, bf16 operands, f32 accumulate in TMEM */
  NERF_PREC_FP16 = 2, /* tcgen05 kind::f16, fp16 operands: the 10-bit operand mantissa of tf32 at twice its rate */
  NERF_PREC_TF32 = 3, /* tcgen05 kind::tf32, operands in 32-bit containers (8-bit exponent, 10-bit mantissa), f32
                       * accumulate in TMEM (mlp_tc_tf32.cu).  On these networks it computes what FP16 computes at
                       * less than half the speed; it is the north star's "tf32 mode", kept for completeness.  (Round 1
                       * silently ran the FP16 path under this name.) */
  NERF_PREC_FP16X3 = 4 /* tcgen05 kind::f16, every operand split x = hi + lo (two fp16), three MMAs per product
                        * (hi.hi + lo.hi + hi.lo), f32 accumulate: ~22-bit operands, f32-class results at a third
                        * of the FP16 rate.  The mode that meets the reference-accuracy bounds on the tensor cores. */
  ,
  NERF_PREC_MIXED = 5  /* coarse network at FP16X3 accuracy, fine network FP16.  The coarse densities only place the
                        * fine samples (lib.rs:409-420), and where the coarse weights are almost zero the inverse CDF
                        * turns a 1e-4 weight error into a visible shift of a sample; on thin structures that is the
                        * whole difference between an FP16 frame and the f32 one.  In a render the coarse pass is two
                        * passes: FP16 over every sample, then FP16X3 over the samples whose density can reach a coarse
                        * weight at all (pre-activation above -0.5 and transmittance in front of it above 6e-6; the
                        * others are exactly zero or behind the reference's own T < 1e-4 cut-off, lib.rs:276-279) - the
                        * frame is bit-identical to the one with FP16X3 over every coarse sample.  nerf_forward_batch
                        * has no rays, so there MIXED means FP16X3 for the coarse network and FP16 for the fine one. */
} nerf_precision;

typedef enum { NERF_NET_COARSE = 0, NERF_NET_FINE = 1 } nerf_net_id;

/* The reference draws jitter from an OS-seeded thread_rng (lib.rs:243,332,375,407), which cannot be
 * reproduced; these are the reproducible substitutes. */
typedef enum {
  NERF_SAMPLING_DET = 0,    /* jitter = 0.5, u_k = (k + 0.5) / n_fine */
  NERF_SAMPLING_PHILOX = 1, /* Philox4x32-10 keyed by (seed, pixel_index, stream) */
  NERF_SAMPLING_ARRAYS = 2  /* caller supplies the uniforms */
} nerf_sampling_mode;

typedef struct nerf_ctx nerf_ctx;

/* One named tensor, names and dims as in lego_rust/<net>/shapes.txt (weights.rs:11-95). */
typedef struct {
  const char* name;  /* e.g. "dense5_kernel" */
  int32_t ndim;      /* 2 for kernels [in,out], 1 for biases */
  int64_t dims[2];
  const float* data; /* host, row-major f32 */
} nerf_tensor;

/* struct Camera after camera_from_samples (lib.rs:197-211,614-645). */
typedef struct {
  float origin[3];
  float forward[3]; /* normalised */
  float up[3];      /* normalised */
  float alpha_w;    /* atan(0.5*W_ref/focal) */
  float alpha_h;
  float near_;
  float far_;
} nerf_camera;

typedef struct {
  int32_t mode; /* nerf_sampling_mode */
  uint64_t seed;
  const float* jitter; /* ARRAYS: [ray_end-ray_begin][n_coarse]; host ptr for host calls, device ptr for *_device */
  const float* u;      /* ARRAYS: [ray_end-ray_begin][n_fine] */
} nerf_sampling;

/* Optional per-stage outputs of nerf_render_stages (host pointers, any may be NULL), indexed by
 * pixel_index - ray_begin. */
typedef struct {
  float* dirs;    /* [n][3]        normalised ray directions      (lib.rs:214-230,371) */
  float* t_c;     /* [n][nc]       stratified samples             (lib.rs:233-248) */
  float* sigma_c; /* [n][nc]       coarse density                 (lib.rs:404) */
  float* w_c;     /* [n][nc]       coarse weights                 (lib.rs:250-283) */
  float* t_f;     /* [n][nc+nf]    merged + sorted samples        (lib.rs:416-420) */
  float* sigma_f; /* [n][nc+nf]    fine density                   (lib.rs:445) */
  float* rgb_f;   /* [n][nc+nf][3] fine colour */
} nerf_stages;

/* Timing of the last nerf_render / nerf_forward_batch on a context, from CUDA events (ms). */
typedef struct {
  float total_ms;      /* first kernel to last kernel, device time */
  float mlp_coarse_ms; /* fused MLP kernel, coarse pass */
  float mlp_fine_ms;   /* fused MLP kernel, fine pass */
  float sampling_ms;   /* ray generation + stratified + importance sampling kernels */
  float composite_ms;  /* compositing kernel */
  int32_t kernel_launches;
  /* mixed mode only (0 otherwise): the part of mlp_coarse_ms spent selecting and re-evaluating coarse samples with
   * split operands, how many coarse samples the render had, and how many of them were re-evaluated */
  float coarse_refine_ms;
  int64_t coarse_samples;
  int64_t coarse_refined;
  /* rays the render covered, and how many of them NERF_OPT_SKIP_EMPTY_RAYS finished without a fine pass (0 unless set) */
  int64_t rays;
  int64_t rays_skipped;
} nerf_timing;

/* Options of a context (nerf_set_option).  All default to the reference's behaviour. */
typedef enum {
  /* Opt-in, NOT the reference's image (the reference has no counterpart: lib.rs:353-472 runs
   * the fine network on every ray).  A ray ALL of whose coarse weights (compute_weights, lib.rs:250-283) are exactly 0
   * gets integrate_ray's value for a ray without weight - rgb (1,1,1), depth 0, opacity 0 (lib.rs:194) - and its fine
   * pass is not run; every other ray is computed exactly as without the option, bit for bit.  The image differs from
   * the reference's only where the fine network finds density on a ray the coarse network sees as empty.  On the lego
   * frame 35 % of the rays are skipped and the frame takes 92 ms instead of 124 (29 pixels of 640 000 move by more than
   * 2/255 from the oracle's frame, 79.9 dB).  The other background rays cannot be told from the coarse pass: the coarse
   * network puts a haze of more than 3 % opacity on them, which only the fine network clears (a threshold on the
   * coarse opacity was measured and skips next to nothing more up to 30 000 ppm, profiles/r02_skip_rays_thresholds.json, and was not kept).  A render in this mode synchronises its
   * stream once per chunk of rays (the size of the fine pass comes back to the host), so the *_device entry points are
   * not fully asynchronous while it is set; nerf_render_stages ignores it. */
  NERF_OPT_SKIP_EMPTY_RAYS = 1
} nerf_option;

const char* nerf_last_error(void);
const char* nerf_version(void);

/* load_network_from_dir x2 (lib.rs:108-174, called at lib.rs:651-652): reads <root>/coarse and <root>/fine
 * (shapes.txt + <name>.bin, little-endian f32), uploads them and builds the pre-swizzled bf16 / fp16
 * operand tiles.  `device` is a CUDA ordinal. */
int nerf_create(const char* weights_root, int device, nerf_ctx** out);
/* Same from caller-held tensors (Network::new, network.rs:181-195). */
int nerf_create_from_tensors(const nerf_tensor* coarse, int n_coarse, const nerf_tensor* fine, int n_fine, int device,
                             nerf_ctx** out);
/* The pre-swizzled weight cache (one file per scene): the 2 x 24 f32 tensors of the weight directory plus the
 * tensor-core chunk streams of every operand format, with a checksum.  nerf_build_weight_cache is host-only (no GPU):
 * it runs load_network_from_dir (lib.rs:108-174) on <root>/coarse and <root>/fine and the tile builders, and writes
 * the file; nerf_create_from_cache builds the same context nerf_create would, from one sequential read.  A missing,
 * truncated or corrupted file is NERF_E_IO, a file of another version NERF_E_UNSUPPORTED. */
int nerf_build_weight_cache(const char* weights_root, const char* cache_path);
int nerf_create_from_cache(const char* cache_path, int device, nerf_ctx** out);
void nerf_destroy(nerf_ctx*);

/* Network::forward_batch (network.rs:197-237).  points: [3][B] feature-major (the reference's Matrix 3xB),
 * dirs: [B][3], rgb: [B][3], sigma: [B].  B == 0 is a no-op (network.rs:199-201). */
int nerf_forward_batch(nerf_ctx*, int net, const float* points, const float* dirs, int64_t B, int precision,
                       float* rgb, float* sigma);
int nerf_forward_batch_device(nerf_ctx*, int net, const float* d_points, const float* d_dirs, int64_t B,
                              int precision, float* d_rgb, float* d_sigma, void* stream);

/* render_image (lib.rs:474-565) restricted to pixel indices [ray_begin, ray_end) (pixel_index = i*W + j,
 * lib.rs:371); pass 0 and W*H for the whole frame.  rgb: [ray_end-ray_begin][3] f32. */
int nerf_render(nerf_ctx*, const nerf_camera*, int W, int H, int n_coarse, int n_fine, int precision,
                const nerf_sampling*, int64_t ray_begin, int64_t ray_end, float* rgb);
int nerf_render_device(nerf_ctx*, const nerf_camera*, int W, int H, int n_coarse, int n_fine, int precision,
                       const nerf_sampling*, int64_t ray_begin, int64_t ray_end, float* d_rgb, void* stream);
/* nerf_render plus the intermediate stages of render_block (lib.rs:353-472). */
int nerf_render_stages(nerf_ctx*, const nerf_camera*, int W, int H, int n_coarse, int n_fine, int precision,
                       const nerf_sampling*, int64_t ray_begin, int64_t ray_end, float* rgb, const nerf_stages*);

/* The step after the path, on the device: render rays [ray_begin, ray_end) and return 8-bit pixels quantised as
 * save_ppm does (lib.rs:574-576): channels = 3 gives the PPM body bytes, channels = 4 the RGBA bytes of
 * pixels_to_rgba (lib.rs:582-592, alpha 255).  3 or 4 B per ray cross PCIe instead of 12. */
int nerf_render_rgb8(nerf_ctx*, const nerf_camera*, int W, int H, int n_coarse, int n_fine, int precision,
                     const nerf_sampling*, int64_t ray_begin, int64_t ray_end, int channels, uint8_t* out);
int nerf_render_rgb8_device(nerf_ctx*, const nerf_camera*, int W, int H, int n_coarse, int n_fine, int precision,
                            const nerf_sampling*, int64_t ray_begin, int64_t ray_end, int channels, uint8_t* d_out,
                            void* stream);
/* Multi-pose batch (an orbit sweep): n_views whole frames of the same size, out = [n_views][H*W][channels] on the
 * host.  The device-to-host copy of view v (pinned double buffer, own stream) overlaps the render of view v + 1.
 * Sampling: deterministic or Philox (view v uses seed + v); caller-supplied arrays are not accepted here. */
int nerf_render_many(nerf_ctx*, const nerf_camera* cams, int n_views, int W, int H, int n_coarse, int n_fine,
                     int precision, const nerf_sampling*, int channels, uint8_t* out);
/* The quantiser alone on the device (host buffers): rgb [n_pixels][3] f32 -> out [n_pixels][channels]. */
int nerf_quantize_rgb8_gpu(nerf_ctx*, const float* rgb, int64_t n_pixels, int channels, uint8_t* out);

/* nerf_render plus two per-ray outputs the reference does not have (its integrate_ray, lib.rs:176-195, returns the
 * colour only): depth = sum_i w_i * t_i and accumulated opacity acc = sum_i w_i over the weights of compute_weights
 * (lib.rs:250-283), so that the background term of integrate_ray is 1 - acc.  depth / acc: [ray_end-ray_begin] f32,
 * either may be null. */
int nerf_render_aux(nerf_ctx*, const nerf_camera*, int W, int H, int n_coarse, int n_fine, int precision,
                    const nerf_sampling*, int64_t ray_begin, int64_t ray_end, float* rgb, float* depth, float* acc);
int nerf_render_aux_device(nerf_ctx*, const nerf_camera*, int W, int H, int n_coarse, int n_fine, int precision,
                           const nerf_sampling*, int64_t ray_begin, int64_t ray_end, float* d_rgb, float* d_depth,
                           float* d_acc, void* stream);

/* Multi-GPU frame assembly without a collective (one process per GPU; the reference's counterpart is rayon's collect
 * + scatter into the image, lib.rs:533-557).  Rank 0 creates the frame buffer in its HBM and publishes the 64-byte
 * CUDA IPC handle (over any side channel); every other rank opens it and passes `ptr + 12 * ray_begin` as d_rgb to
 * nerf_render_device / nerf_render_aux_device for its own [ray_begin, ray_end): the compositing kernel then stores
 * its pixels straight into rank 0's buffer over NVLink - the gather is fused into the producing kernel.  The frame
 * is complete on rank 0 once every rank's stream has been synchronised (a host barrier).  `owner` = 1 frees the
 * allocation, 0 unmaps it. */
#define NERF_IPC_HANDLE_BYTES 64
int nerf_shared_frame_create(nerf_ctx*, uint64_t bytes, void** d_ptr, unsigned char handle[NERF_IPC_HANDLE_BYTES]);
int nerf_shared_frame_open(nerf_ctx*, const unsigned char handle[NERF_IPC_HANDLE_BYTES], void** d_ptr);
int nerf_shared_frame_close(nerf_ctx*, void* d_ptr, int owner);

/* The same frame buffer shared per allocation through the CUDA virtual-memory API instead: only this buffer is mapped
 * on the peers.  (The IPC form above turns on device-wide peer access, after which every allocation of the process is
 * mapped for the peer - measured to cost the MLP kernels 1-3 %.)  create returns the mapped size and a POSIX file
 * descriptor to hand to the other ranks (SCM_RIGHTS over a Unix socket; the caller closes its copy), open maps the
 * received descriptor, close_fd unmaps and drops this process's reference (the memory is freed with the last one). */
int nerf_shared_frame_create_fd(nerf_ctx*, uint64_t bytes, void** d_ptr, uint64_t* mapped_bytes, int* fd);
int nerf_shared_frame_open_fd(nerf_ctx*, int fd, uint64_t mapped_bytes, void** d_ptr);
int nerf_shared_frame_close_fd(nerf_ctx*, void* d_ptr);

/* Stage entry points (host buffers), one per reference function, used by the stage-wise parity tests. */
int nerf_ray_dirs(nerf_ctx*, const nerf_camera*, int W, int H, int64_t ray_begin, int64_t ray_end,
                  float* dirs /*[n][3]*/);                                   /* lib.rs:214-230 + :371 */
int nerf_compute_weights(nerf_ctx*, const float* sigma, const float* t, int64_t n_rays, int n, float far_,
                         float* w);                                          /* lib.rs:250-283 */
int nerf_sample_importance(nerf_ctx*, const float* t, const float* w, int64_t n_rays, int n, const float* u,
                           int count, float* t_merged /*[n_rays][n+count] sorted*/); /* lib.rs:289-351,416-420 */
int nerf_integrate_rays(nerf_ctx*, const float* rgb, const float* sigma, const float* t, int64_t n_rays, int n,
                        float far_, float* out /*[n_rays][3]*/);             /* lib.rs:176-195 */

/* Striped views: the frame's pixel indices are dealt out in stripes of `stripe_rays` to `parts` renderers in turn
 * (stripe k of the frame belongs to part k % parts); this call renders part `part`.  The cost of a ray depends on what
 * it hits - empty rows are cheaper than rows through the object since the mixed mode refines fewer coarse samples and
 * the fine pass skips the colour of empty tiles there - so interleaved stripes, not contiguous ranges, are what balances
 * several GPUs (rayon's work stealing over 8x8 blocks plays this role in the reference, lib.rs:533-550).  Pixels are
 * bit-identical to the same pixels of a whole-frame render.  `d_frame` / `frame` is the WHOLE frame [W*H][3]; only the
 * part's pixels are written.  Deterministic or Philox sampling. */
int nerf_render_striped_device(nerf_ctx*, const nerf_camera*, int W, int H, int n_coarse, int n_fine, int precision,
                               const nerf_sampling*, int64_t stripe_rays, int parts, int part, float* d_frame,
                               void* stream);
int nerf_render_striped(nerf_ctx*, const nerf_camera*, int W, int H, int n_coarse, int n_fine, int precision,
                        const nerf_sampling*, int64_t stripe_rays, int parts, int part, float* frame);

/* No counterpart in the reference: added for callers that may not allocate
 * device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * nerf_integrate_rays with its device scratch (the staged rgb / sigma / t inputs and the output) in ONE file-scope static
 * device array instead of cudaMalloc'ed workspaces: the call performs no cudaMalloc, no cudaFree and calls no allocating
 * library routine.  Same arguments (host pointers), same kernel, same result as nerf_integrate_rays.
 *   - Limit: the four segments, each rounded up to 256 B - 12*n_rays*n + 4*n_rays*n + 4*n_rays*n + 12*n_rays bytes -
 *     must fit NERF_NOALLOC_MAX_BYTES (1 088 rays per call at 64 + 128 samples per ray).  A larger call is refused with
 *     NERF_E_UNSUPPORTED and a message naming the sizes; nothing is truncated.
 *   - Every GPU that loads the library holds the array until the process exits.
 *   - Every call on a device shares the array and finds the last call's contents in it; each call overwrites what it
 *     reads, so nothing needs clearing.  Calls that overlap on the GPU (two contexts on one device) race: serialise them. */
#define NERF_NOALLOC_MAX_BYTES (4u << 20)
int nerf_integrate_rays_noalloc(nerf_ctx*, const float* rgb, const float* sigma, const float* t, int64_t n_rays, int n,
                                float far_, float* out /*[n_rays][3]*/);

int nerf_get_timing(nerf_ctx*, nerf_timing* out);

/* Context options (nerf_option).  Unknown options: NERF_E_ARG.  Also settable at creation through the environment
 * (NERF_B200_SKIP_EMPTY_RAYS=1). */
int nerf_set_option(nerf_ctx*, int option, int value);
int nerf_get_option(nerf_ctx*, int option, int* value);

/* Host-side helpers of the boundary (no GPU needed). */
/* camera_from_samples (lib.rs:614-645) over the JSON keys near, far, camera_origin, camera_forward,
 * camera_up, hwf (lib.rs:620-629). */
int nerf_camera_from_json(const char* path, nerf_camera* out);
int nerf_camera_from_values(double near_, double far_, const double origin[3], const double forward[3],
                            const double up[3], const double hwf[3], nerf_camera* out);
/* save_ppm (lib.rs:567-580): "P6\n{w} {h}\n255\n" + (clamp(v,0,1)*255+0.5) as u8. */
int nerf_save_ppm(const char* path, int W, int H, const float* rgb);
/* The same quantiser on a buffer. */
int nerf_quantize_rgb8(const float* rgb, int64_t n, uint8_t* out);

#ifdef __cplusplus
}
#endif
#endif /* NERF_B200_H */
