/*
 * oracle.h — TEST INFRASTRUCTURE ONLY.  CPU restatement of the nerf-rs hot path.
 *
 * This library is the checker for the CUDA product in nerf_rs_b200/.  It is a
 * line-faithful C++ restatement of the reference crate's CPU algorithm
 * (elisabeth96/nerf-rs, src/lib.rs + src/network.rs + src/vec3.rs); every
 * function cites the reference file:line it follows.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load it.  The product library (libnerf_b200.so) never links or calls it.
 *
 * Parity pin: oracle_forward_batch reproduces every (sigma, rgb) vector of
 * lego_rust/tf_reference_samples.json (3 rays x 5 samples x {coarse,fine}); ray
 * generation is pinned indirectly by the same file's ray_d (pixel-corner
 * offset).  Sampling / compositing / PPM are NOT pinned by any reference golden
 * ("parity unpinned" for those stages: the Rust crate cannot be built here and
 * holds no test for them); they are pinned by line-faithful restatement only.
 */
#ifndef NERF_ORACLE_H
#define NERF_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oracle_net oracle_net;

/* Camera after camera_from_samples (src/lib.rs:614-645). */
typedef struct {
  float pos[3];
  float dir[3]; /* normalised camera_forward */
  float up[3];  /* normalised camera_up */
  float alpha_w, alpha_h;
  float near_, far_;
} oracle_camera;

/* Operand rounding emulation for the dense layers (0 = exact f32 = the reference). */
enum { ORACLE_ROUND_NONE = 0, ORACLE_ROUND_BF16 = 1, ORACLE_ROUND_FP16 = 2, ORACLE_ROUND_TF32 = 3 };
/* GEMM loop order: 0 = col->out->in (bit-identical, fast); 1 = the reference's literal in->col->out. */
enum { ORACLE_GEMM_REORDERED = 0, ORACLE_GEMM_LITERAL = 1 };
/* Sampling modes (the reference uses an OS-seeded thread_rng, which is irreproducible). */
enum { ORACLE_SAMPLING_DET = 0, ORACLE_SAMPLING_PHILOX = 1, ORACLE_SAMPLING_ARRAYS = 2 };

typedef struct {
  int mode;
  uint64_t seed;
  const float* jitter; /* ARRAYS: [n_rays][n_coarse], indexed by pixel_index - ray_begin */
  const float* u;      /* ARRAYS: [n_rays][n_fine] */
} oracle_sampling;

/* Optional per-stage dumps; any pointer may be NULL.  Indexed by (pixel_index - ray_begin). */
typedef struct {
  float* dirs;    /* [n][3] normalised ray dirs */
  float* t_c;     /* [n][nc] */
  float* sigma_c; /* [n][nc] */
  float* w_c;     /* [n][nc] */
  float* t_f;     /* [n][nc+nf] merged + sorted */
  float* sigma_f; /* [n][nc+nf] */
  float* rgb_f;   /* [n][nc+nf][3] */
} oracle_stages;

const char* oracle_last_error(void);

oracle_net* oracle_load_network(const char* dir);
void oracle_free_network(oracle_net*);
/* Copies tensor `name` (e.g. "dense5_kernel") into out (n floats); returns element count or -1. */
int64_t oracle_get_tensor(const oracle_net*, const char* name, float* out, int64_t n);

int oracle_forward_batch(const oracle_net*, const float* points /*[3][B]*/, const float* dirs /*[B][3]*/,
                         int64_t B, int gemm_order, int round_mode, float* rgb /*[B][3]*/, float* sigma /*[B]*/);

void oracle_camera_from_values(double near_, double far_, const double origin[3], const double forward[3],
                               const double up[3], const double hwf[3], oracle_camera* out);
/* offset = 0.5 is the reference; 0.0 reproduces the JSON's ray_d.  normalize != 0 applies lib.rs:371. */
void oracle_ray_dirs(const oracle_camera*, int nx, int ny, float offset, int normalize, int64_t ray_begin,
                     int64_t ray_end, float* out /*[n][3]*/);
void oracle_stratified(float near_, float far_, int count, const float* jitter, float* t_out);
void oracle_compute_weights(const float* sigma, const float* t, int n, float far_, float* w_out);
/* returns number of samples written (0 or count) */
int oracle_sample_importance(const float* t, const float* w, int n, const float* u, int count, float* out);
/* depth = sum w_i t_i, acc = sum w_i over compute_weights' weights (no reference counterpart; see oracle.cpp) */
void oracle_integrate_aux(const float* sigma, const float* t, int n, float far_, float out[2]);
void oracle_integrate_ray(const float* rgb /*[n][3]*/, const float* sigma, const float* t, int n, float far_,
                          float out[3]);
/* Uniforms shared bit-for-bit with the CUDA path: Philox4x32-10, key=(seed lo,hi),
 * counter=(pixel_index, block, stream, 0); float = (x>>8) * 2^-24. */
void oracle_philox_uniforms(uint64_t seed, uint32_t pixel_index, uint32_t stream, int count, float* out);

/* render_image (src/lib.rs:474-565) over pixel_index in [ray_begin, ray_end); returns 0 / negative. */
int oracle_render(const oracle_net* coarse, const oracle_net* fine, const oracle_camera* cam, int W, int H,
                  int n_coarse, int n_fine, const oracle_sampling* sampling, int64_t ray_begin, int64_t ray_end,
                  int gemm_order, int round_mode, int n_threads, float* rgb_out /*[n][3]*/,
                  const oracle_stages* stages);

int oracle_save_ppm(const char* path, int W, int H, const float* rgb /*[H*W][3]*/);
/* quantiser of save_ppm alone: (clamp(v,0,1)*255+0.5) as u8 */
void oracle_quantize(const float* v, int64_t n, uint8_t* out);

int oracle_num_threads(void);

#ifdef __cplusplus
}
#endif
#endif
