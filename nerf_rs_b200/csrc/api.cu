// api.cu — the C ABI of libnerf_b200.so (include/nerf_b200.h): context, weight upload, workspaces and the
// kernel schedule of render_block / render_image (src/lib.rs:353-565).  CUDA-only: no CPU fallback.
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

#include <cuda.h>  // types only: the driver entry points are fetched with cudaGetDriverEntryPoint, libcuda is not linked
#include <unistd.h>

#include <map>

#include "common.cuh"
#include "host_util.h"
#include "mlp_tc.h"
#include "weight_cache.h"

namespace nerf {
namespace {

#define NERF_CUDA(expr)                                                                       \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess) {                                                                  \
      set_error(std::string(#expr) + ": " + cudaGetErrorString(_e));                          \
      return NERF_E_CUDA;                                                                     \
    }                                                                                         \
  } while (0)

// Grow-only device buffer.
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes) {
    if (bytes <= cap) return NERF_OK;
    void* q = nullptr;
    if (cudaMalloc(&q, bytes) != cudaSuccess) {
      // not enough room for old + new: the contents are scratch, so give the old block back and try once more
      cudaGetLastError();
      if (p) cudaFree(p);
      p = nullptr;
      cap = 0;
      NERF_CUDA(cudaMalloc(&q, bytes));
    } else if (p) {
      cudaFree(p);
    }
    p = q;
    cap = bytes;
    return NERF_OK;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  template <class T>
  T* as() const { return reinterpret_cast<T*>(p); }
};

struct DeviceNet {
  float* blob = nullptr;  // every f32 tensor of the network, file layout, 16-byte aligned segments
  NetF32 f32{};
  TcWeights tc[2];        // [0] bf16, [1] fp16 pre-swizzled operand tiles (mlp_tc.cu)
  TcWeights tcx3;         // fp16 hi / lo split tiles of the fp16x3 mode (mlp_tc_x3.cu)
  TcWeights tctf32;       // tf32 tiles (mlp_tc_tf32.cu)
};

enum StageKind { ST_SAMPLING = 0, ST_MLP_COARSE, ST_MLP_FINE, ST_COMPOSITE, ST_COARSE_REFINE, ST_COUNT };

}  // namespace
}  // namespace nerf

using namespace nerf;

struct nerf_ctx {
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;  // used by the host entry points
  std::recursive_mutex mu;
  DeviceNet nets[2];
  DevBuf dirs, tc, sigc, wc, tf, sigf, rgbf, out, in_a, in_b, jit, uni;
  // two-pass coarse densities of the mixed mode: ids of the samples the split-operand kernel re-evaluates
  DevBuf sel_list, sel_count, alive[2];
  int coarse_stages = 4;        // NERF_B200_COARSE_STAGES: slices of the front-to-back fp16 coarse pass (1 = one launch)
  bool coarse_select = true;    // NERF_B200_COARSE_SELECT=0: split-operand kernel over every coarse sample
  bool skip_empty = true;       // NERF_B200_SKIP_EMPTY=0: the fine pass evaluates the colour of empty samples too
  // NERF_B200_SELECT_MARGIN: bound on the fp16 pass's pre-activation error near zero.  Measured (scripts/select_check.py,
  // 800x800 reference pose): the frame stays bit-identical down to 0.03, so 0.5 keeps a factor 16.
  float select_margin = 0.5f;
  // nerf_set_option(NERF_OPT_SKIP_EMPTY_RAYS) / NERF_B200_SKIP_EMPTY_RAYS=1 (opt-in, NOT the reference's image; DESIGN.md
  // 4.6): rays without any coarse weight get the background and skip the fine pass.  The number of surviving rays comes
  // back through a pinned word, so a render in this mode synchronises its stream once per chunk.
  bool skip_empty_rays = false;
  DevBuf ray_list, ray_count, dirs_alive;
  uint32_t* h_ray_count = nullptr;
  int64_t rays_total = 0, rays_skipped = 0;  // since timing_start
  // Two-phase fine pass (NERF_B200_TWO_PHASE=0/1): density for every fine sample first, the colour
  // branch afterwards over the samples that have density, in groups of fine_sub_rays rays (NERF_B200_FINE_SUB_RAYS) so
  // that the rows handed from one kernel to the other (512 B per sample with density) fit a fixed buffer even if every
  // sample has density.  Needs skip_empty (samples without density get rgb = 0) and a 16-bit tensor-core mode.
  bool two_phase_fine = true;
  int64_t fine_sub_rays = 40960;
  DevBuf h8, h8_meta, h8_count;
  float select_tau = 12.0f;     // NERF_B200_SELECT_TAU: optical depth behind which no coarse weight survives (T < 6e-6)
  // 8-bit output path: two device frames + two pinned host frames so that the copy of one view overlaps the next render
  DevBuf q8[2], depth, acc;
  uint8_t* pinned[2] = {nullptr, nullptr};
  size_t pinned_cap = 0;
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_q[2] = {nullptr, nullptr}, ev_c[2] = {nullptr, nullptr};
  // timing
  struct Span { cudaEvent_t a, b; int kind; };
  std::vector<Span> spans;
  size_t spans_used = 0;
  cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
  cudaEvent_t ev_busy = nullptr;  // end of the last enqueued call, on whichever stream it ran
  bool busy_recorded = false;
  bool have_timing = false;
  int launches = 0;
  int64_t refine_total = 0;     // coarse samples that went through the two-pass path since timing_start
  bool refine_pending = false;  // sel_count holds the last chunk's list length
  int64_t chunk_rays = 1 << 20;
  struct VmmFrame { CUmemGenericAllocationHandle handle; size_t size; };
  std::map<void*, VmmFrame> vmm_frames;  // nerf_shared_frame_*_fd mappings owned by this context
};

namespace nerf {
namespace {

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
  }
  ~DeviceGuard() {
    int cur = -1;
    cudaGetDevice(&cur);
    if (prev >= 0 && cur != prev) cudaSetDevice(prev);
  }
};

// `images`: the chunk streams of a weight cache file (one per operand format), or null to build them from `t`.
int upload_network(const TensorMap& t, DeviceNet* net, const std::vector<uint8_t>* images = nullptr) {
  static const char* const names[12] = {"dense0", "dense1", "dense2", "dense3", "dense4", "dense5",
                                        "dense6", "dense7", "bottleneck", "viewdirs", "rgb", "alpha"};
  size_t total = 0;
  auto pad4 = [](size_t n) { return (n + 3) & ~(size_t)3; };
  for (const char* n : names) {
    total += pad4(t.at(std::string(n) + "_kernel").data.size());
    total += pad4(t.at(std::string(n) + "_bias").data.size());
  }
  std::vector<float> host(total, 0.0f);
  const float* dptr[24];
  NERF_CUDA(cudaMalloc(&net->blob, total * sizeof(float)));
  size_t off = 0;
  for (int i = 0; i < 12; ++i) {
    const auto& k = t.at(std::string(names[i]) + "_kernel").data;
    const auto& b = t.at(std::string(names[i]) + "_bias").data;
    std::memcpy(host.data() + off, k.data(), k.size() * 4);
    dptr[2 * i] = net->blob + off;
    off += pad4(k.size());
    std::memcpy(host.data() + off, b.data(), b.size() * 4);
    dptr[2 * i + 1] = net->blob + off;
    off += pad4(b.size());
  }
  NERF_CUDA(cudaMemcpy(net->blob, host.data(), total * sizeof(float), cudaMemcpyHostToDevice));
  for (int i = 0; i < 8; ++i) {
    net->f32.w[i] = dptr[2 * i];
    net->f32.b[i] = dptr[2 * i + 1];
  }
  net->f32.w_bott = dptr[16];  net->f32.b_bott = dptr[17];
  net->f32.w_view = dptr[18];  net->f32.b_view = dptr[19];
  net->f32.w_rgb = dptr[20];   net->f32.b_rgb = dptr[21];
  net->f32.w_alpha = dptr[22]; net->f32.b_alpha = dptr[23];
  // Pre-swizzled 16-bit operand tiles for the tensor-core kernel, one set per operand format.
  for (int fmt = 0; fmt < kNumTcFormats; ++fmt) {
    TcWeights* w = fmt == 3 ? &net->tctf32 : (fmt == 2 ? &net->tcx3 : &net->tc[fmt]);
    int rc = images ? upload_tc_weights(images[fmt].data(), images[fmt].size(), fmt, t, w) : build_tc_weights(t, fmt, w);
    if (rc != NERF_OK) return rc;
  }
  return NERF_OK;
}

int create_common(const TensorMap& coarse, const TensorMap& fine, int device, nerf_ctx** out,
                  const CachedWeights* cache = nullptr) {
  int rc = validate_network(coarse, "coarse");
  if (rc != NERF_OK) return rc;
  rc = validate_network(fine, "fine");
  if (rc != NERF_OK) return rc;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
    set_error("no CUDA device: libnerf_b200 has no CPU fallback");
    return NERF_E_CUDA;
  }
  if (device < 0 || device >= count) {
    set_error("device ordinal out of range");
    return NERF_E_ARG;
  }
  cudaDeviceProp prop;
  NERF_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    set_error(std::string("device is sm_") + std::to_string(prop.major * 10 + prop.minor) +
              "; libnerf_b200 is built for sm_100a only");
    return NERF_E_CUDA;
  }
  DeviceGuard g(device);
  nerf_ctx* ctx = new (std::nothrow) nerf_ctx();
  if (!ctx) {
    set_error("out of host memory");
    return NERF_E_ARG;
  }
  ctx->device = device;
  ctx->sm_count = prop.multiProcessorCount;
  if (const char* e = std::getenv("NERF_B200_CHUNK_RAYS")) {
    const long long v = std::atoll(e);
    if (v > 0) ctx->chunk_rays = v;
  }
  if (const char* e = std::getenv("NERF_B200_COARSE_SELECT")) ctx->coarse_select = std::atoi(e) != 0;
  if (const char* e = std::getenv("NERF_B200_SKIP_EMPTY")) ctx->skip_empty = std::atoi(e) != 0;
  if (const char* e = std::getenv("NERF_B200_SKIP_EMPTY_RAYS")) ctx->skip_empty_rays = std::atoi(e) != 0;
  if (const char* e = std::getenv("NERF_B200_TWO_PHASE")) ctx->two_phase_fine = std::atoi(e) != 0;
  if (const char* e = std::getenv("NERF_B200_FINE_SUB_RAYS")) ctx->fine_sub_rays = std::max<int64_t>(1, std::atoll(e));
  if (const char* e = std::getenv("NERF_B200_COARSE_STAGES")) ctx->coarse_stages = std::max(1, std::atoi(e));
  if (const char* e = std::getenv("NERF_B200_SELECT_MARGIN")) ctx->select_margin = (float)std::atof(e);
  if (const char* e = std::getenv("NERF_B200_SELECT_TAU")) ctx->select_tau = (float)std::atof(e);
  cudaError_t e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = init_mlp_tc();  // function attributes, once per device
  if (e == cudaSuccess) e = init_mlp_tc_x3();
  if (e == cudaSuccess) e = init_mlp_tc_tf32();
  if (e == cudaSuccess) e = init_sampling();
  if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev_begin);
  if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev_end);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->ev_busy, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking);
  for (int i = 0; i < 2 && e == cudaSuccess; ++i) {
    e = cudaEventCreateWithFlags(&ctx->ev_q[i], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->ev_c[i], cudaEventDisableTiming);
  }
  if (e != cudaSuccess) {
    set_error(std::string("stream/event create: ") + cudaGetErrorString(e));
    delete ctx;
    return NERF_E_CUDA;
  }
  rc = ctx->sel_count.reserve(256);
  if (rc == NERF_OK && cudaMemset(ctx->sel_count.p, 0, 256) != cudaSuccess) rc = NERF_E_CUDA;
  if (rc == NERF_OK) rc = upload_network(coarse, &ctx->nets[0], cache ? cache->images[0] : nullptr);
  if (rc == NERF_OK) rc = upload_network(fine, &ctx->nets[1], cache ? cache->images[1] : nullptr);
  if (rc != NERF_OK) {
    nerf_destroy(ctx);
    return rc;
  }
  *out = ctx;
  return NERF_OK;
}

// get_ray_dir's pixel-independent part (src/lib.rs:216-226), evaluated once on the host.
RayBasis make_basis(const nerf_camera& cam, int W, int H) {
  RayBasis rb{};
  normalize3(cam.forward, rb.f);  // f = dir.normalize()
  float tmp[3];
  cross3(rb.f, cam.up, tmp);
  normalize3(tmp, rb.r);  // r = f x up, normalised
  cross3(rb.r, rb.f, tmp);
  normalize3(tmp, rb.u);  // u = r x f, normalised
  rb.sx = std::tan(cam.alpha_w);
  rb.sy = std::tan(cam.alpha_h);
  for (int i = 0; i < 3; ++i) rb.origin[i] = cam.origin[i];
  rb.near_ = cam.near_;
  rb.far_ = cam.far_;
  rb.nx = W;
  rb.ny = H;
  return rb;
}

void span_begin(nerf_ctx* c, int kind, cudaStream_t s) {
  if (c->spans_used == c->spans.size()) {
    nerf_ctx::Span sp{};
    cudaEventCreate(&sp.a);
    cudaEventCreate(&sp.b);
    c->spans.push_back(sp);
  }
  c->spans[c->spans_used].kind = kind;
  cudaEventRecord(c->spans[c->spans_used].a, s);
}
void span_end(nerf_ctx* c, cudaStream_t s) {
  cudaEventRecord(c->spans[c->spans_used].b, s);
  ++c->spans_used;
}

int run_mlp(nerf_ctx* c, int net, const SampleSrc& src, const SampleDst& dst, int precision, cudaStream_t s) {
  if (src.B == 0) return NERF_OK;
  cudaError_t e;
  switch (precision) {
    case NERF_PREC_FP32:
      e = launch_mlp_fp32(c->nets[net].f32, src, dst, s);
      break;
    case NERF_PREC_BF16:
      e = launch_mlp_tc(c->nets[net].tc[0], src, dst, c->sm_count, s);
      break;
    case NERF_PREC_TF32:
      e = launch_mlp_tc_tf32(c->nets[net].tctf32, src, dst, c->sm_count, s);
      break;
    case NERF_PREC_FP16:
      e = launch_mlp_tc(c->nets[net].tc[1], src, dst, c->sm_count, s);
      break;
    case NERF_PREC_FP16X3:
      e = launch_mlp_tc_x3(c->nets[net].tcx3, src, dst, c->sm_count, s);
      break;
    case NERF_PREC_MIXED:  // coarse network at f32-class accuracy (it places the fine samples), fine network in fp16
      e = net == NERF_NET_COARSE ? launch_mlp_tc_x3(c->nets[net].tcx3, src, dst, c->sm_count, s)
                                 : launch_mlp_tc(c->nets[net].tc[1], src, dst, c->sm_count, s);
      break;
    default:
      set_error("unknown precision");
      return NERF_E_ARG;
  }
  ++c->launches;
  if (e != cudaSuccess) {
    set_error(std::string("mlp launch: ") + cudaGetErrorString(e));
    return NERF_E_CUDA;
  }
  return NERF_OK;
}

int check_render_args(const nerf_ctx* c, const nerf_camera* cam, int W, int H, int nc, int nf, const nerf_sampling* smp,
                      int64_t rb, int64_t re) {
  if (!c || !cam) { set_error("null context or camera"); return NERF_E_ARG; }
  if (nc <= 0) { set_error("coarse samples per ray must be greater than 0"); return NERF_E_ARG; }       // lib.rs:483-486
  if (W <= 0 || W % 8 != 0) { set_error("image width must be multiple of block size"); return NERF_E_ARG; }   // lib.rs:492-496
  if (H <= 0 || H % 8 != 0) { set_error("image height must be multiple of block size"); return NERF_E_ARG; }  // lib.rs:497-501
  if (nf < 0) { set_error("negative fine sample count"); return NERF_E_ARG; }
  if (nf > 0 && nc < 3) {
    // sample_importance returns no samples for n < 3 (lib.rs:295-297), which makes rays ragged.
    set_error("n_fine > 0 needs n_coarse >= 3 (ragged rays are not supported)");
    return NERF_E_UNSUPPORTED;
  }
  if ((size_t)(6 * (size_t)nc + 2 * (size_t)nf) * 2 * 8 * sizeof(float) > 200 * 1024) {  // pow2 padding <= 2x
    set_error("n_coarse/n_fine too large for the importance sampler's shared memory");
    return NERF_E_UNSUPPORTED;
  }
  if (rb < 0 || re < rb || re > (int64_t)W * H) { set_error("bad ray range"); return NERF_E_ARG; }
  if (smp && smp->mode == NERF_SAMPLING_ARRAYS && (!smp->jitter || (nf > 0 && !smp->u))) {
    set_error("NERF_SAMPLING_ARRAYS needs jitter and u");
    return NERF_E_ARG;
  }
  if (smp && (smp->mode < 0 || smp->mode > 2)) { set_error("unknown sampling mode"); return NERF_E_ARG; }
  return NERF_OK;
}

struct StagePtrs {  // device destinations for stage dumps (may be null)
  float *dirs = nullptr, *t_c = nullptr, *sigma_c = nullptr, *w_c = nullptr, *t_f = nullptr, *sigma_f = nullptr,
        *rgb_f = nullptr;
};

// The kernel schedule of render_block (lib.rs:353-472) for rays [rb, re): K1 -> MLP(coarse) -> K3 ->
// MLP(fine) -> K4.  d_jit/d_u are device pointers indexed from ray rb.
int render_range(nerf_ctx* c, const nerf_camera& cam, int W, int H, int nc, int nf, int precision, int mode,
                 uint64_t seed, const float* d_jit, const float* d_u, int64_t rb, int64_t re, float* d_rgb,
                 bool want_wc, cudaStream_t s, float* d_depth = nullptr, float* d_acc = nullptr,
                 const RayMap* view = nullptr, bool compact_out = false) {
  // view == nullptr: rays are the pixel indices [rb, re) and d_rgb / d_depth / d_acc point at pixel rb's element.
  // view != nullptr: rays are the view's indices [rb, re) (RayMap::pixel names their pixels) and the outputs are
  // whole-frame arrays indexed by pixel, or (compact_out) arrays indexed by view index.
  const RayBasis basis = make_basis(cam, W, H);
  const int nm = nc + nf;
  for (int64_t b0 = rb; b0 < re; b0 += c->chunk_rays) {
    const int64_t n = std::min<int64_t>(c->chunk_rays, re - b0);
    const int64_t off = b0 - rb;
    int rc;
    if ((rc = c->dirs.reserve(n * 3 * 4)) || (rc = c->tc.reserve(n * nc * 4)) || (rc = c->sigc.reserve(n * nc * 4)) ||
        (rc = c->tf.reserve(n * nm * 4)) || (rc = c->sigf.reserve(n * nm * 4)) || (rc = c->rgbf.reserve(n * nm * 12)) ||
        (want_wc && (rc = c->wc.reserve(n * nc * 4))))
      return rc;
    span_begin(c, ST_SAMPLING, s);
    const RayMap map = view ? RayMap{b0, view->stripe, view->parts, view->part} : RayMap{b0, 0, 1, 0};
    const RayMap omap = (view && !compact_out) ? map : RayMap{off, 0, 1, 0};
    NERF_CUDA(launch_raygen_stratified(basis, nc, mode, seed, d_jit ? d_jit + off * nc : nullptr, map, n,
                                       c->dirs.as<float>(), c->tc.as<float>(), s));
    span_end(c, s);
    ++c->launches;

    SampleSrc src{};
    src.mode = 1;
    src.S = nc;
    src.B = n * nc;
    src.dirs = c->dirs.as<float>();
    src.t = c->tc.as<float>();
    for (int i = 0; i < 3; ++i) src.origin[i] = cam.origin[i];
    SampleDst dst{c->sigc.as<float>(), nullptr, 0.0f};  // coarse colour is discarded by the reference (lib.rs:404)
    if (precision == NERF_PREC_MIXED && c->coarse_select) {
      // Two-pass coarse densities: fp16 pre-activations for every sample, then the split-operand kernel over the
      // compacted list of samples whose density can reach a coarse weight (typically 8-10 % of them).
      if ((rc = c->sel_list.reserve(n * nc * 4)) || (rc = c->sel_count.reserve(256))) return rc;
      span_begin(c, ST_MLP_COARSE, s);
      dst.sigma_floor = -INFINITY;
      // The fp16 pass itself runs front to back in `stages` slices of the ray: a ray whose optical depth over the slices
      // done so far is beyond tau_cut cannot give weight to anything behind it (the same argument as for the refinement
      // list), so the later slices only cover the rays that still have light.
      const int stages =
          (c->coarse_stages > 1 && nc % c->coarse_stages == 0 && nc / c->coarse_stages >= 8) ? c->coarse_stages : 1;
      if (stages == 1) {
        rc = run_mlp(c, NERF_NET_COARSE, src, dst, NERF_PREC_FP16, s);
      } else {
        if ((rc = c->alive[0].reserve(n * 4)) || (rc = c->alive[1].reserve(n * 4))) return rc;
        // samples never evaluated keep 0xFEFEFEFE = -1.7e38: a hugely negative pre-activation, density 0
        NERF_CUDA(cudaMemsetAsync(c->sigc.p, 0xFE, n * nc * 4, s));
        uint32_t* cnt = c->sel_count.as<uint32_t>() + 8;  // one counter per stage, behind the refinement's two
        NERF_CUDA(cudaMemsetAsync(cnt, 0, 4 * 16, s));
        const int per = nc / stages;
        SampleSrc part = src;
        part.sub_n = per;
        part.B = n * per;
        for (int k = 0; k < stages && rc == NERF_OK; ++k) {
          part.sub_s0 = k * per;
          if (k > 0) {
            NERF_CUDA(launch_coarse_alive(c->sigc.as<float>(), c->tc.as<float>(), n, nc, k * per, cam.far_, c->select_tau,
                                          k > 1 ? c->alive[(k - 1) & 1].as<int32_t>() : nullptr,
                                          k > 1 ? cnt + (k - 1) : nullptr, c->alive[k & 1].as<int32_t>(), cnt + k, s));
            ++c->launches;
            part.index = c->alive[k & 1].as<int32_t>();
            part.count = cnt + k;
          }
          rc = run_mlp(c, NERF_NET_COARSE, part, dst, NERF_PREC_FP16, s);
        }
      }
      span_end(c, s);
      if (rc) return rc;
      span_begin(c, ST_COARSE_REFINE, s);
      NERF_CUDA(cudaMemsetAsync(c->sel_count.p, 0, 4, s));
      NERF_CUDA(launch_coarse_select(c->sigc.as<float>(), c->tc.as<float>(), n, nc, cam.far_, c->select_margin,
                                     c->select_tau, c->sel_list.as<int32_t>(), c->sel_count.as<uint32_t>(), s));
      ++c->launches;
      SampleSrc ref = src;
      ref.index = c->sel_list.as<int32_t>();
      ref.count = c->sel_count.as<uint32_t>();
      dst.sigma_floor = 0.0f;
      rc = run_mlp(c, NERF_NET_COARSE, ref, dst, NERF_PREC_FP16X3, s);
      span_end(c, s);
      if (rc) return rc;
      c->refine_total += n * nc;
      c->refine_pending = true;
    } else {
      span_begin(c, ST_MLP_COARSE, s);
      rc = run_mlp(c, NERF_NET_COARSE, src, dst, precision, s);
      span_end(c, s);
      if (rc) return rc;
    }

    // Skip-empty-rays mode: the rays the coarse pass gives no weight are finished here (background); the fine pass runs
    // over the compacted list of the others.  Not taken for stage dumps, whose arrays are indexed by ray.
    int64_t nr = n;                     // rays of the fine pass
    const int32_t* ray_ids = nullptr;   // their ray numbers within the chunk (null: all, in order)
    if (c->skip_empty_rays && !want_wc) {
      if ((rc = c->ray_list.reserve(n * 4)) || (rc = c->ray_count.reserve(256)) || (rc = c->dirs_alive.reserve(n * 3 * 4)))
        return rc;
      if (c->h_ray_count == nullptr) NERF_CUDA(cudaMallocHost(reinterpret_cast<void**>(&c->h_ray_count), 64));
      span_begin(c, ST_SAMPLING, s);
      NERF_CUDA(cudaMemsetAsync(c->ray_count.p, 0, 4, s));
      NERF_CUDA(launch_ray_pick(c->sigc.as<float>(), c->tc.as<float>(), c->dirs.as<float>(), n, nc, cam.far_, omap,
                                c->ray_list.as<int32_t>(), c->ray_count.as<uint32_t>(), c->dirs_alive.as<float>(), d_rgb,
                                d_depth, d_acc, s));
      NERF_CUDA(cudaMemcpyAsync(c->h_ray_count, c->ray_count.p, 4, cudaMemcpyDeviceToHost, s));
      span_end(c, s);
      ++c->launches;
      NERF_CUDA(cudaStreamSynchronize(s));
      nr = (int64_t)*c->h_ray_count;
      ray_ids = c->ray_list.as<int32_t>();
      src.dirs = c->dirs_alive.as<float>();
      c->rays_skipped += n - nr;
    }
    c->rays_total += n;
    if (nr == 0) continue;

    span_begin(c, ST_SAMPLING, s);
    NERF_CUDA(launch_importance(c->tc.as<float>(), c->sigc.as<float>(), nullptr, nc, nf, cam.far_, mode, seed,
                                d_u ? d_u + off * nf : nullptr, map, nr, c->tf.as<float>(),
                                want_wc ? c->wc.as<float>() : nullptr, s, ray_ids));
    span_end(c, s);
    ++c->launches;

    src.S = nm;
    src.B = nr * nm;
    src.t = c->tf.as<float>();
    dst = SampleDst{c->sigf.as<float>(), c->rgbf.as<float>(), 0.0f, c->skip_empty ? 1 : 0};
    const bool two_phase = c->two_phase_fine && c->skip_empty &&
                           (precision == NERF_PREC_MIXED || precision == NERF_PREC_FP16 || precision == NERF_PREC_BF16);
    span_begin(c, ST_MLP_FINE, s);
    if (two_phase) {
      const TcWeights& wt = c->nets[NERF_NET_FINE].tc[precision == NERF_PREC_BF16 ? 0 : 1];
      const int64_t sub_max = std::min<int64_t>(nr, c->fine_sub_rays);
      // rows are stored as whole [128 x 256] tile images; a CTA pair of the colour kernel reads two tiles at a time
      if ((rc = c->h8.reserve(((sub_max * nm + 255) / 256) * 256 * 512)) || (rc = c->h8_meta.reserve(sub_max * nm * 16)) ||
          (rc = c->h8_count.reserve(256)))
        return rc;
      for (int64_t r0 = 0; r0 < nr && rc == NERF_OK; r0 += c->fine_sub_rays) {
        const int64_t ns = std::min<int64_t>(c->fine_sub_rays, nr - r0);
        SampleSrc s2 = src;
        s2.dirs = src.dirs + r0 * 3;
        s2.t = src.t + r0 * nm;
        s2.B = ns * nm;
        SampleDst d2 = dst;
        d2.sigma = dst.sigma + r0 * nm;
        d2.rgb = dst.rgb + r0 * nm * 3;
        d2.skip_empty_rgb = 0;
        d2.h8 = c->h8.as<uint16_t>();
        d2.h8_meta = c->h8_meta.as<float4>();
        d2.h8_count = c->h8_count.as<uint32_t>();
        NERF_CUDA(cudaMemsetAsync(d2.h8_count, 0, 4, s));
        rc = run_mlp(c, NERF_NET_FINE, s2, d2, precision, s);
        if (rc == NERF_OK) {
          NERF_CUDA(launch_mlp_colour(wt, d2.h8, d2.h8_meta, d2.h8_count, d2.rgb, c->sm_count, s));
          ++c->launches;
        }
      }
    } else {
      rc = run_mlp(c, NERF_NET_FINE, src, dst, precision, s);
    }
    span_end(c, s);
    if (rc) return rc;

    span_begin(c, ST_COMPOSITE, s);
    NERF_CUDA(launch_composite(c->rgbf.as<float>(), c->sigf.as<float>(), c->tf.as<float>(), nr, nm, cam.far_, omap, d_rgb,
                               d_depth, d_acc, s, ray_ids));
    span_end(c, s);
    ++c->launches;
  }
  return NERF_OK;
}

// Every call that enqueues work on the context's workspaces goes through timing_start / timing_stop.  The workspaces
// are one set per context, so a call on another stream must not start before the previous call's kernels have finished
// with them: timing_stop records ev_busy on the call's stream, timing_start makes the next call's stream wait for it
// (a no-op on the same stream).  Calls are serialised on the device, never on the host.
void timing_start(nerf_ctx* c, cudaStream_t s) {
  if (c->busy_recorded) cudaStreamWaitEvent(s, c->ev_busy, 0);
  c->spans_used = 0;
  c->launches = 0;
  c->refine_total = 0;
  c->refine_pending = false;
  c->rays_total = c->rays_skipped = 0;
  if (c->sel_count.p) cudaMemsetAsync(c->sel_count.p, 0, 8, s);  // [0] list length of the chunk, [1] running total
  c->have_timing = true;
  cudaEventRecord(c->ev_begin, s);
}
void timing_stop(nerf_ctx* c, cudaStream_t s) {
  cudaEventRecord(c->ev_end, s);
  cudaEventRecord(c->ev_busy, s);
  c->busy_recorded = true;
}

}  // namespace
}  // namespace nerf

// ---- per-allocation sharing through the CUDA virtual-memory API (driver entry points, resolved at run time)
namespace nerf {
namespace {
struct VmmApi {
  decltype(&cuMemGetAllocationGranularity) granularity = nullptr;
  decltype(&cuMemCreate) create = nullptr;
  decltype(&cuMemRelease) release = nullptr;
  decltype(&cuMemAddressReserve) reserve = nullptr;
  decltype(&cuMemAddressFree) addr_free = nullptr;
  decltype(&cuMemMap) map = nullptr;
  decltype(&cuMemUnmap) unmap = nullptr;
  decltype(&cuMemSetAccess) set_access = nullptr;
  decltype(&cuMemExportToShareableHandle) export_handle = nullptr;
  decltype(&cuMemImportFromShareableHandle) import_handle = nullptr;
  bool ok = false;
};
const VmmApi& vmm_api() {
  static const VmmApi api = [] {
    VmmApi a;
    bool ok = true;
    auto get = [&](const char* name, void** fn) {
      cudaDriverEntryPointQueryResult st;
      if (cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &st) != cudaSuccess || st != cudaDriverEntryPointSuccess ||
          *fn == nullptr)
        ok = false;
    };
    get("cuMemGetAllocationGranularity", (void**)&a.granularity);
    get("cuMemCreate", (void**)&a.create);
    get("cuMemRelease", (void**)&a.release);
    get("cuMemAddressReserve", (void**)&a.reserve);
    get("cuMemAddressFree", (void**)&a.addr_free);
    get("cuMemMap", (void**)&a.map);
    get("cuMemUnmap", (void**)&a.unmap);
    get("cuMemSetAccess", (void**)&a.set_access);
    get("cuMemExportToShareableHandle", (void**)&a.export_handle);
    get("cuMemImportFromShareableHandle", (void**)&a.import_handle);
    a.ok = ok;
    return a;
  }();
  return api;
}
CUmemAllocationProp frame_prop(int owner_device) {
  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = owner_device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}
#define NERF_CU(expr)                                                                \
  do {                                                                               \
    CUresult _r = (expr);                                                            \
    if (_r != CUDA_SUCCESS) {                                                        \
      set_error(std::string(#expr) + ": CUresult " + std::to_string((int)_r));       \
      return NERF_E_CUDA;                                                            \
    }                                                                                \
  } while (0)

// Reserves an address range, maps `handle` there and gives this context's device read/write access.
int vmm_map(nerf_ctx* c, CUmemGenericAllocationHandle handle, size_t size, void** out) {
  const VmmApi& a = vmm_api();
  CUdeviceptr va = 0;
  NERF_CU(a.reserve(&va, size, 0, 0, 0));
  CUresult r = a.map(va, size, 0, handle, 0);
  if (r == CUDA_SUCCESS) {
    CUmemAccessDesc acc{};
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = c->device;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    r = a.set_access(va, size, &acc, 1);
    if (r != CUDA_SUCCESS) a.unmap(va, size);
  }
  if (r != CUDA_SUCCESS) {
    a.addr_free(va, size);
    set_error("map shared frame: CUresult " + std::to_string((int)r));
    return NERF_E_CUDA;
  }
  *out = reinterpret_cast<void*>(va);
  c->vmm_frames[*out] = nerf_ctx::VmmFrame{handle, size};
  return NERF_OK;
}
}  // namespace
}  // namespace nerf

// ============================================================================ extern "C" entry points
extern "C" {

int nerf_create(const char* weights_root, int device, nerf_ctx** out) {
  if (!weights_root || !out) {
    set_error("nerf_create: null argument");
    return NERF_E_ARG;
  }
  TensorMap coarse, fine;
  int rc = load_tensor_dir(std::string(weights_root) + "/coarse", &coarse);  // lib.rs:651
  if (rc != NERF_OK) return rc;
  rc = load_tensor_dir(std::string(weights_root) + "/fine", &fine);  // lib.rs:652
  if (rc != NERF_OK) return rc;
  return create_common(coarse, fine, device, out);
}

int nerf_build_weight_cache(const char* weights_root, const char* cache_path) {
  if (!weights_root || !cache_path) {
    set_error("nerf_build_weight_cache: null argument");
    return NERF_E_ARG;
  }
  TensorMap nets[2];
  int rc = load_tensor_dir(std::string(weights_root) + "/coarse", &nets[0]);
  if (rc == NERF_OK) rc = load_tensor_dir(std::string(weights_root) + "/fine", &nets[1]);
  std::vector<uint8_t> blob;
  if (rc == NERF_OK) rc = serialize_weight_cache(nets, &blob);
  if (rc == NERF_OK) rc = write_file(cache_path, blob);
  return rc;
}

int nerf_create_from_cache(const char* cache_path, int device, nerf_ctx** out) {
  if (!cache_path || !out) {
    set_error("nerf_create_from_cache: null argument");
    return NERF_E_ARG;
  }
  std::vector<uint8_t> blob;
  int rc = read_file(cache_path, &blob);
  if (rc != NERF_OK) return rc;
  CachedWeights cw;
  rc = parse_weight_cache(blob.data(), blob.size(), &cw);
  if (rc != NERF_OK) return rc;
  for (int n = 0; n < 2; ++n) {  // a stream of the wrong length would make the kernel's bulk copies read past the end
    for (int fmt = 0; fmt < kNumTcFormats; ++fmt) {
      if (cw.images[n][fmt].size() != tc_image_bytes(fmt)) {
        set_error("weight cache: chunk stream of unexpected size");
        return NERF_E_SHAPE;
      }
    }
  }
  return create_common(cw.nets[0], cw.nets[1], device, out, &cw);
}

int nerf_create_from_tensors(const nerf_tensor* coarse, int n_coarse, const nerf_tensor* fine, int n_fine, int device,
                             nerf_ctx** out) {
  if (!coarse || !fine || !out) {
    set_error("nerf_create_from_tensors: null argument");
    return NERF_E_ARG;
  }
  auto to_map = [](const nerf_tensor* ts, int n, TensorMap* m) -> int {
    for (int i = 0; i < n; ++i) {
      const nerf_tensor& t = ts[i];
      if (!t.name || !t.data || t.ndim < 1 || t.ndim > 2) {
        set_error("nerf_create_from_tensors: malformed tensor entry");
        return NERF_E_SHAPE;
      }
      HostTensor h;
      int64_t cnt = 1;
      for (int d = 0; d < t.ndim; ++d) {
        h.dims.push_back(t.dims[d]);
        cnt *= t.dims[d];
      }
      h.data.assign(t.data, t.data + cnt);
      (*m)[t.name] = std::move(h);
    }
    return NERF_OK;
  };
  TensorMap c, f;
  int rc = to_map(coarse, n_coarse, &c);
  if (rc != NERF_OK) return rc;
  rc = to_map(fine, n_fine, &f);
  if (rc != NERF_OK) return rc;
  return create_common(c, f, device, out);
}

void nerf_destroy(nerf_ctx* c) {
  if (!c) return;
  {
    DeviceGuard g(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (!c->vmm_frames.empty()) {  // shared frames the caller never closed: unmap and drop this context's references
      cudaDeviceSynchronize();
      const VmmApi& a = vmm_api();
      for (auto& kv : c->vmm_frames) {
        const CUdeviceptr va = reinterpret_cast<CUdeviceptr>(kv.first);
        if (a.ok) {
          a.unmap(va, kv.second.size);
          a.addr_free(va, kv.second.size);
          a.release(kv.second.handle);
        }
      }
      c->vmm_frames.clear();
    }
    for (DevBuf* b : {&c->dirs, &c->tc, &c->sigc, &c->wc, &c->tf, &c->sigf, &c->rgbf, &c->out, &c->in_a, &c->in_b,
                      &c->jit, &c->uni, &c->q8[0], &c->q8[1], &c->depth, &c->acc, &c->sel_list, &c->sel_count, &c->alive[0], &c->alive[1],
                      &c->ray_list, &c->ray_count, &c->dirs_alive, &c->h8, &c->h8_meta, &c->h8_count})
      b->release();
    if (c->h_ray_count) cudaFreeHost(c->h_ray_count);
    if (c->copy_stream) cudaStreamSynchronize(c->copy_stream);
    for (int i = 0; i < 2; ++i) {
      if (c->pinned[i]) cudaFreeHost(c->pinned[i]);
      if (c->ev_q[i]) cudaEventDestroy(c->ev_q[i]);
      if (c->ev_c[i]) cudaEventDestroy(c->ev_c[i]);
    }
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    for (auto& n : c->nets) {
      if (n.blob) cudaFree(n.blob);
      for (auto& t : n.tc) free_tc_weights(&t);
      free_tc_weights(&n.tcx3);
      free_tc_weights(&n.tctf32);
    }
    for (auto& sp : c->spans) {
      cudaEventDestroy(sp.a);
      cudaEventDestroy(sp.b);
    }
    if (c->ev_begin) cudaEventDestroy(c->ev_begin);
    if (c->ev_end) cudaEventDestroy(c->ev_end);
    if (c->ev_busy) cudaEventDestroy(c->ev_busy);
    if (c->stream) cudaStreamDestroy(c->stream);
  }
  delete c;
}

int nerf_forward_batch_device(nerf_ctx* c, int net, const float* d_points, const float* d_dirs, int64_t B,
                              int precision, float* d_rgb, float* d_sigma, void* stream) {
  if (!c || net < 0 || net > 1 || B < 0) {
    set_error("nerf_forward_batch: bad argument");
    return NERF_E_ARG;
  }
  if (B == 0) return NERF_OK;  // network.rs:199-201
  if (!d_points || !d_dirs || !d_rgb || !d_sigma) {
    set_error("nerf_forward_batch: null buffer");
    return NERF_E_ARG;
  }
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  cudaStream_t s = (cudaStream_t)stream;
  SampleSrc src{};
  src.mode = 0;
  src.B = B;
  src.points = d_points;
  src.dirs = d_dirs;
  SampleDst dst{d_sigma, d_rgb};
  timing_start(c, s);
  span_begin(c, net == 0 ? ST_MLP_COARSE : ST_MLP_FINE, s);
  int rc = run_mlp(c, net, src, dst, precision, s);
  span_end(c, s);
  timing_stop(c, s);
  return rc;
}

int nerf_forward_batch(nerf_ctx* c, int net, const float* points, const float* dirs, int64_t B, int precision,
                       float* rgb, float* sigma) {
  if (!c || net < 0 || net > 1 || B < 0) {
    set_error("nerf_forward_batch: bad argument");
    return NERF_E_ARG;
  }
  if (B == 0) return NERF_OK;
  if (!points || !dirs || !rgb || !sigma) {
    set_error("nerf_forward_batch: null buffer");
    return NERF_E_ARG;
  }
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  int rc;
  if ((rc = c->in_a.reserve(B * 12)) || (rc = c->in_b.reserve(B * 12)) || (rc = c->rgbf.reserve(B * 12)) ||
      (rc = c->sigf.reserve(B * 4)))
    return rc;
  NERF_CUDA(cudaMemcpyAsync(c->in_a.p, points, B * 12, cudaMemcpyHostToDevice, c->stream));
  NERF_CUDA(cudaMemcpyAsync(c->in_b.p, dirs, B * 12, cudaMemcpyHostToDevice, c->stream));
  rc = nerf_forward_batch_device(c, net, c->in_a.as<float>(), c->in_b.as<float>(), B, precision,
                                 c->rgbf.as<float>(), c->sigf.as<float>(), c->stream);
  if (rc != NERF_OK) return rc;
  NERF_CUDA(cudaMemcpyAsync(rgb, c->rgbf.p, B * 12, cudaMemcpyDeviceToHost, c->stream));
  NERF_CUDA(cudaMemcpyAsync(sigma, c->sigf.p, B * 4, cudaMemcpyDeviceToHost, c->stream));
  NERF_CUDA(cudaStreamSynchronize(c->stream));
  return NERF_OK;
}

int nerf_render_device(nerf_ctx* c, const nerf_camera* cam, int W, int H, int nc, int nf, int precision,
                       const nerf_sampling* smp, int64_t rb, int64_t re, float* d_rgb, void* stream) {
  int rc = check_render_args(c, cam, W, H, nc, nf, smp, rb, re);
  if (rc != NERF_OK) return rc;
  if (re == rb) return NERF_OK;
  if (!d_rgb) {
    set_error("nerf_render: null output");
    return NERF_E_ARG;
  }
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  cudaStream_t s = (cudaStream_t)stream;
  const int mode = smp ? smp->mode : NERF_SAMPLING_DET;
  timing_start(c, s);
  rc = render_range(c, *cam, W, H, nc, nf, precision, mode, smp ? smp->seed : 0, smp ? smp->jitter : nullptr,
                    smp ? smp->u : nullptr, rb, re, d_rgb, false, s);
  timing_stop(c, s);
  return rc;
}

static int render_host(nerf_ctx* c, const nerf_camera* cam, int W, int H, int nc, int nf, int precision,
                       const nerf_sampling* smp, int64_t rb, int64_t re, float* rgb, const nerf_stages* st) {
  int rc = check_render_args(c, cam, W, H, nc, nf, smp, rb, re);
  if (rc != NERF_OK) return rc;
  const int64_t n = re - rb;
  if (n == 0) return NERF_OK;
  if (!rgb) {
    set_error("nerf_render: null output");
    return NERF_E_ARG;
  }
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  cudaStream_t s = c->stream;
  const int mode = smp ? smp->mode : NERF_SAMPLING_DET;
  const float *d_jit = nullptr, *d_u = nullptr;
  if (mode == NERF_SAMPLING_ARRAYS) {
    if ((rc = c->jit.reserve(n * nc * 4)) || (rc = c->uni.reserve(std::max<int64_t>(n * nf, 1) * 4))) return rc;
    NERF_CUDA(cudaMemcpyAsync(c->jit.p, smp->jitter, n * nc * 4, cudaMemcpyHostToDevice, s));
    if (nf > 0) NERF_CUDA(cudaMemcpyAsync(c->uni.p, smp->u, n * nf * 4, cudaMemcpyHostToDevice, s));
    d_jit = c->jit.as<float>();
    d_u = c->uni.as<float>();
  }
  if ((rc = c->out.reserve(n * 12))) return rc;
  const bool want_stages = st != nullptr;
  const int64_t saved_chunk = c->chunk_rays;
  const bool saved_skip = c->skip_empty;
  if (want_stages) c->chunk_rays = n;  // stage dumps read the workspaces of a single chunk
  if (want_stages && st->rgb_f) c->skip_empty = false;  // a caller who asks for every sample's colour gets it
  const bool saved_skip_rays = c->skip_empty_rays;
  if (want_stages) c->skip_empty_rays = false;  // stage arrays are indexed by ray: every ray takes the whole pipeline
  timing_start(c, s);
  rc = render_range(c, *cam, W, H, nc, nf, precision, mode, smp ? smp->seed : 0, d_jit, d_u, rb, re,
                    c->out.as<float>(), want_stages && st->w_c, s);
  timing_stop(c, s);
  c->chunk_rays = saved_chunk;
  c->skip_empty = saved_skip;
  c->skip_empty_rays = saved_skip_rays;
  if (rc != NERF_OK) return rc;
  NERF_CUDA(cudaMemcpyAsync(rgb, c->out.p, n * 12, cudaMemcpyDeviceToHost, s));
  if (want_stages) {
    const int nm = nc + nf;
    auto d2h = [&](float* dst, const DevBuf& b, size_t bytes) -> cudaError_t {
      return dst ? cudaMemcpyAsync(dst, b.p, bytes, cudaMemcpyDeviceToHost, s) : cudaSuccess;
    };
    NERF_CUDA(d2h(st->dirs, c->dirs, n * 12));
    NERF_CUDA(d2h(st->t_c, c->tc, n * nc * 4));
    NERF_CUDA(d2h(st->sigma_c, c->sigc, n * nc * 4));
    NERF_CUDA(d2h(st->w_c, c->wc, n * nc * 4));
    NERF_CUDA(d2h(st->t_f, c->tf, n * nm * 4));
    NERF_CUDA(d2h(st->sigma_f, c->sigf, n * nm * 4));
    NERF_CUDA(d2h(st->rgb_f, c->rgbf, n * nm * 12));
  }
  NERF_CUDA(cudaStreamSynchronize(s));
  return NERF_OK;
}

int nerf_render(nerf_ctx* c, const nerf_camera* cam, int W, int H, int nc, int nf, int precision,
                const nerf_sampling* smp, int64_t rb, int64_t re, float* rgb) {
  return render_host(c, cam, W, H, nc, nf, precision, smp, rb, re, rgb, nullptr);
}

int nerf_render_stages(nerf_ctx* c, const nerf_camera* cam, int W, int H, int nc, int nf, int precision,
                       const nerf_sampling* smp, int64_t rb, int64_t re, float* rgb, const nerf_stages* st) {
  static const nerf_stages none{};
  return render_host(c, cam, W, H, nc, nf, precision, smp, rb, re, rgb, st ? st : &none);
}

// ---------------------------------------------------------------------------- striped views (multi-GPU load balance)
namespace nerf {
namespace {
int check_stripes(int W, int H, int64_t stripe, int parts, int part) {
  if (stripe <= 0 || parts <= 0 || part < 0 || part >= parts || stripe > (int64_t)W * H) {
    set_error("striped view: need stripe_rays > 0 and 0 <= part < parts");
    return NERF_E_ARG;
  }
  return NERF_OK;
}
}  // namespace
}  // namespace nerf

int nerf_render_striped_device(nerf_ctx* c, const nerf_camera* cam, int W, int H, int nc, int nf, int precision,
                               const nerf_sampling* smp, int64_t stripe_rays, int parts, int part, float* d_frame,
                               void* stream) {
  int rc = check_render_args(c, cam, W, H, nc, nf, smp, 0, (int64_t)W * H);
  if (rc == NERF_OK) rc = check_stripes(W, H, stripe_rays, parts, part);
  if (rc != NERF_OK) return rc;
  if (smp && smp->mode == NERF_SAMPLING_ARRAYS) {
    set_error("striped views take deterministic or Philox sampling (caller arrays are indexed by contiguous ranges)");
    return NERF_E_UNSUPPORTED;
  }
  const int64_t n = view_size((int64_t)W * H, stripe_rays, parts, part);
  if (n == 0) return NERF_OK;
  if (!d_frame) {
    set_error("nerf_render_striped: null output");
    return NERF_E_ARG;
  }
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  cudaStream_t s = (cudaStream_t)stream;
  const RayMap view{0, stripe_rays, parts, part};
  timing_start(c, s);
  rc = render_range(c, *cam, W, H, nc, nf, precision, smp ? smp->mode : NERF_SAMPLING_DET, smp ? smp->seed : 0, nullptr,
                    nullptr, 0, n, d_frame, false, s, nullptr, nullptr, &view);
  timing_stop(c, s);
  return rc;
}

int nerf_render_striped(nerf_ctx* c, const nerf_camera* cam, int W, int H, int nc, int nf, int precision,
                        const nerf_sampling* smp, int64_t stripe_rays, int parts, int part, float* frame) {
  int rc = check_render_args(c, cam, W, H, nc, nf, smp, 0, (int64_t)W * H);
  if (rc == NERF_OK) rc = check_stripes(W, H, stripe_rays, parts, part);
  if (rc != NERF_OK) return rc;
  const int64_t total = (int64_t)W * H, n = view_size(total, stripe_rays, parts, part);
  if (n == 0) return NERF_OK;
  if (!frame) {
    set_error("nerf_render_striped: null output");
    return NERF_E_ARG;
  }
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  // the view is rendered compactly (view index order) and its stripes are copied to their places in the caller's frame
  if ((rc = c->out.reserve(n * 12))) return rc;
  cudaStream_t s = c->stream;
  {
    // a compact view is a contiguous "frame" of its own for the compositing kernel: render with the view's pixel map
    // for ray generation and Philox, but store by view index
    const RayMap view{0, stripe_rays, parts, part};
    timing_start(c, s);
    rc = render_range(c, *cam, W, H, nc, nf, precision, smp ? smp->mode : NERF_SAMPLING_DET, smp ? smp->seed : 0, nullptr,
                      nullptr, 0, n, c->out.as<float>(), false, s, nullptr, nullptr, &view, /*compact_out=*/true);
    timing_stop(c, s);
    if (rc != NERF_OK) return rc;
  }
  const size_t block = (size_t)stripe_rays * 12;
  const int64_t full = n / stripe_rays, tail = n - full * stripe_rays;
  if (full > 0)
    NERF_CUDA(cudaMemcpy2DAsync(reinterpret_cast<char*>(frame) + (size_t)part * block, block * parts, c->out.p, block, block,
                                (size_t)full, cudaMemcpyDeviceToHost, s));
  if (tail > 0)
    NERF_CUDA(cudaMemcpyAsync(reinterpret_cast<char*>(frame) + ((size_t)full * parts + part) * block,
                              reinterpret_cast<const char*>(c->out.p) + (size_t)full * block, (size_t)tail * 12,
                              cudaMemcpyDeviceToHost, s));
  NERF_CUDA(cudaStreamSynchronize(s));
  return NERF_OK;
}

// ---------------------------------------------------------------------------- 8-bit output
namespace nerf {
namespace {
int check_channels(int channels) {
  if (channels != 3 && channels != 4) {
    set_error("channels must be 3 (PPM body) or 4 (RGBA)");
    return NERF_E_ARG;
  }
  return NERF_OK;
}
int reserve_pinned(nerf_ctx* c, size_t bytes) {
  if (bytes <= c->pinned_cap) return NERF_OK;
  for (int i = 0; i < 2; ++i) {
    if (c->pinned[i]) cudaFreeHost(c->pinned[i]);
    c->pinned[i] = nullptr;
  }
  c->pinned_cap = 0;
  for (int i = 0; i < 2; ++i) NERF_CUDA(cudaMallocHost(reinterpret_cast<void**>(&c->pinned[i]), bytes));
  c->pinned_cap = bytes;
  return NERF_OK;
}
}  // namespace
}  // namespace nerf

int nerf_render_rgb8_device(nerf_ctx* c, const nerf_camera* cam, int W, int H, int nc, int nf, int precision,
                            const nerf_sampling* smp, int64_t rb, int64_t re, int channels, uint8_t* d_out,
                            void* stream) {
  int rc = check_render_args(c, cam, W, H, nc, nf, smp, rb, re);
  if (rc == NERF_OK) rc = check_channels(channels);
  if (rc != NERF_OK) return rc;
  if (re == rb) return NERF_OK;
  if (!d_out) {
    set_error("nerf_render_rgb8: null output");
    return NERF_E_ARG;
  }
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  cudaStream_t s = (cudaStream_t)stream;
  if ((rc = c->out.reserve((re - rb) * 12))) return rc;
  const int mode = smp ? smp->mode : NERF_SAMPLING_DET;
  timing_start(c, s);
  rc = render_range(c, *cam, W, H, nc, nf, precision, mode, smp ? smp->seed : 0, smp ? smp->jitter : nullptr,
                    smp ? smp->u : nullptr, rb, re, c->out.as<float>(), false, s);
  if (rc == NERF_OK) {
    NERF_CUDA(launch_quantize(c->out.as<float>(), re - rb, channels, d_out, s));
    ++c->launches;
  }
  timing_stop(c, s);
  return rc;
}

int nerf_render_rgb8(nerf_ctx* c, const nerf_camera* cam, int W, int H, int nc, int nf, int precision,
                     const nerf_sampling* smp, int64_t rb, int64_t re, int channels, uint8_t* out) {
  int rc = check_render_args(c, cam, W, H, nc, nf, smp, rb, re);
  if (rc == NERF_OK) rc = check_channels(channels);
  if (rc != NERF_OK) return rc;
  const int64_t n = re - rb;
  if (n == 0) return NERF_OK;
  if (!out) {
    set_error("nerf_render_rgb8: null output");
    return NERF_E_ARG;
  }
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  cudaStream_t s = c->stream;
  nerf_sampling dev = smp ? *smp : nerf_sampling{};
  if (!smp) dev.mode = NERF_SAMPLING_DET;
  if (dev.mode == NERF_SAMPLING_ARRAYS) {
    if ((rc = c->jit.reserve(n * nc * 4)) || (rc = c->uni.reserve(std::max<int64_t>(n * nf, 1) * 4))) return rc;
    NERF_CUDA(cudaMemcpyAsync(c->jit.p, smp->jitter, n * nc * 4, cudaMemcpyHostToDevice, s));
    if (nf > 0) NERF_CUDA(cudaMemcpyAsync(c->uni.p, smp->u, n * nf * 4, cudaMemcpyHostToDevice, s));
    dev.jitter = c->jit.as<float>();
    dev.u = c->uni.as<float>();
  }
  const size_t bytes = (size_t)n * channels;
  if ((rc = c->q8[0].reserve(bytes))) return rc;
  rc = nerf_render_rgb8_device(c, cam, W, H, nc, nf, precision, &dev, rb, re, channels, c->q8[0].as<uint8_t>(), s);
  if (rc != NERF_OK) return rc;
  // straight into the caller's buffer, as nerf_render does (a pinned or registered buffer is written by DMA; a pageable
  // one goes through the driver's own staging - no second copy here)
  NERF_CUDA(cudaMemcpyAsync(out, c->q8[0].p, bytes, cudaMemcpyDeviceToHost, s));
  NERF_CUDA(cudaStreamSynchronize(s));
  return NERF_OK;
}

int nerf_render_many(nerf_ctx* c, const nerf_camera* cams, int n_views, int W, int H, int nc, int nf, int precision,
                     const nerf_sampling* smp, int channels, uint8_t* out) {
  if (!c || !cams || n_views < 0) {
    set_error("nerf_render_many: null context or cameras");
    return NERF_E_ARG;
  }
  if (smp && smp->mode == NERF_SAMPLING_ARRAYS) {
    set_error("nerf_render_many: caller-supplied sample arrays are not supported");
    return NERF_E_UNSUPPORTED;
  }
  const int64_t n = (int64_t)W * H;
  int rc = check_channels(channels);
  for (int v = 0; v < n_views && rc == NERF_OK; ++v) rc = check_render_args(c, &cams[v], W, H, nc, nf, smp, 0, n);
  if (rc != NERF_OK) return rc;
  if (n_views == 0) return NERF_OK;
  if (!out) {
    set_error("nerf_render_many: null output");
    return NERF_E_ARG;
  }
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  const size_t bytes = (size_t)n * channels;
  if ((rc = c->q8[0].reserve(bytes)) || (rc = c->q8[1].reserve(bytes)) || (rc = reserve_pinned(c, bytes))) return rc;
  cudaStream_t s = c->stream;
  auto drain = [&](int v) -> int {  // view v's bytes: pinned -> caller
    NERF_CUDA(cudaEventSynchronize(c->ev_c[v & 1]));
    std::memcpy(out + (size_t)v * bytes, c->pinned[v & 1], bytes);
    return NERF_OK;
  };
  for (int v = 0; v < n_views; ++v) {
    nerf_sampling sv = smp ? *smp : nerf_sampling{};
    if (!smp) sv.mode = NERF_SAMPLING_DET;
    sv.seed += (uint64_t)v;
    // buffer v & 1 was last used by view v - 2, drained during iteration v - 1
    rc = nerf_render_rgb8_device(c, &cams[v], W, H, nc, nf, precision, &sv, 0, n, channels, c->q8[v & 1].as<uint8_t>(), s);
    if (rc != NERF_OK) return rc;
    NERF_CUDA(cudaEventRecord(c->ev_q[v & 1], s));
    NERF_CUDA(cudaStreamWaitEvent(c->copy_stream, c->ev_q[v & 1], 0));
    NERF_CUDA(cudaMemcpyAsync(c->pinned[v & 1], c->q8[v & 1].p, bytes, cudaMemcpyDeviceToHost, c->copy_stream));
    NERF_CUDA(cudaEventRecord(c->ev_c[v & 1], c->copy_stream));
    if (v > 0 && (rc = drain(v - 1))) return rc;  // overlaps view v's render
  }
  return drain(n_views - 1);
}

int nerf_quantize_rgb8_gpu(nerf_ctx* c, const float* rgb, int64_t n_pixels, int channels, uint8_t* out) {
  if (!c || n_pixels < 0 || (n_pixels > 0 && (!rgb || !out))) {
    set_error("nerf_quantize_rgb8_gpu: bad argument");
    return NERF_E_ARG;
  }
  int rc = check_channels(channels);
  if (rc != NERF_OK) return rc;
  if (n_pixels == 0) return NERF_OK;
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  cudaStream_t s = c->stream;
  if ((rc = c->out.reserve(n_pixels * 12)) || (rc = c->q8[0].reserve(n_pixels * channels))) return rc;
  NERF_CUDA(cudaMemcpyAsync(c->out.p, rgb, n_pixels * 12, cudaMemcpyHostToDevice, s));
  NERF_CUDA(launch_quantize(c->out.as<float>(), n_pixels, channels, c->q8[0].as<uint8_t>(), s));
  NERF_CUDA(cudaMemcpyAsync(out, c->q8[0].p, n_pixels * channels, cudaMemcpyDeviceToHost, s));
  NERF_CUDA(cudaStreamSynchronize(s));
  return NERF_OK;
}


// ---------------------------------------------------------------------------- depth / opacity
int nerf_render_aux_device(nerf_ctx* c, const nerf_camera* cam, int W, int H, int nc, int nf, int precision,
                           const nerf_sampling* smp, int64_t rb, int64_t re, float* d_rgb, float* d_depth,
                           float* d_acc, void* stream) {
  int rc = check_render_args(c, cam, W, H, nc, nf, smp, rb, re);
  if (rc != NERF_OK) return rc;
  if (re == rb) return NERF_OK;
  if (!d_rgb) {
    set_error("nerf_render_aux: null rgb output");
    return NERF_E_ARG;
  }
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  cudaStream_t s = (cudaStream_t)stream;
  const int mode = smp ? smp->mode : NERF_SAMPLING_DET;
  timing_start(c, s);
  rc = render_range(c, *cam, W, H, nc, nf, precision, mode, smp ? smp->seed : 0, smp ? smp->jitter : nullptr,
                    smp ? smp->u : nullptr, rb, re, d_rgb, false, s, d_depth, d_acc);
  timing_stop(c, s);
  return rc;
}

int nerf_render_aux(nerf_ctx* c, const nerf_camera* cam, int W, int H, int nc, int nf, int precision,
                    const nerf_sampling* smp, int64_t rb, int64_t re, float* rgb, float* depth, float* acc) {
  int rc = check_render_args(c, cam, W, H, nc, nf, smp, rb, re);
  if (rc != NERF_OK) return rc;
  const int64_t n = re - rb;
  if (n == 0) return NERF_OK;
  if (!rgb) {
    set_error("nerf_render_aux: null rgb output");
    return NERF_E_ARG;
  }
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  cudaStream_t s = c->stream;
  nerf_sampling dev = smp ? *smp : nerf_sampling{};
  if (!smp) dev.mode = NERF_SAMPLING_DET;
  if (dev.mode == NERF_SAMPLING_ARRAYS) {
    if ((rc = c->jit.reserve(n * nc * 4)) || (rc = c->uni.reserve(std::max<int64_t>(n * nf, 1) * 4))) return rc;
    NERF_CUDA(cudaMemcpyAsync(c->jit.p, smp->jitter, n * nc * 4, cudaMemcpyHostToDevice, s));
    if (nf > 0) NERF_CUDA(cudaMemcpyAsync(c->uni.p, smp->u, n * nf * 4, cudaMemcpyHostToDevice, s));
    dev.jitter = c->jit.as<float>();
    dev.u = c->uni.as<float>();
  }
  if ((rc = c->out.reserve(n * 12)) || (rc = c->depth.reserve(n * 4)) || (rc = c->acc.reserve(n * 4))) return rc;
  rc = nerf_render_aux_device(c, cam, W, H, nc, nf, precision, &dev, rb, re, c->out.as<float>(),
                              depth ? c->depth.as<float>() : nullptr, acc ? c->acc.as<float>() : nullptr, s);
  if (rc != NERF_OK) return rc;
  NERF_CUDA(cudaMemcpyAsync(rgb, c->out.p, n * 12, cudaMemcpyDeviceToHost, s));
  if (depth) NERF_CUDA(cudaMemcpyAsync(depth, c->depth.p, n * 4, cudaMemcpyDeviceToHost, s));
  if (acc) NERF_CUDA(cudaMemcpyAsync(acc, c->acc.p, n * 4, cudaMemcpyDeviceToHost, s));
  NERF_CUDA(cudaStreamSynchronize(s));
  return NERF_OK;
}

int nerf_ray_dirs(nerf_ctx* c, const nerf_camera* cam, int W, int H, int64_t rb, int64_t re, float* dirs) {
  if (!c || !cam || W <= 0 || H <= 0 || rb < 0 || re < rb || re > (int64_t)W * H || (!dirs && re > rb)) {
    set_error("nerf_ray_dirs: bad argument");
    return NERF_E_ARG;
  }
  const int64_t n = re - rb;
  if (n == 0) return NERF_OK;
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  int rc;
  if ((rc = c->dirs.reserve(n * 12))) return rc;
  NERF_CUDA(launch_ray_dirs(make_basis(*cam, W, H), rb, n, c->dirs.as<float>(), c->stream));
  NERF_CUDA(cudaMemcpyAsync(dirs, c->dirs.p, n * 12, cudaMemcpyDeviceToHost, c->stream));
  NERF_CUDA(cudaStreamSynchronize(c->stream));
  return NERF_OK;
}

int nerf_compute_weights(nerf_ctx* c, const float* sigma, const float* t, int64_t n_rays, int n, float far_,
                         float* w) {
  if (!c || n_rays < 0 || n < 0 || ((!sigma || !t || !w) && n_rays * n > 0)) {
    set_error("nerf_compute_weights: bad argument");
    return NERF_E_ARG;
  }
  const int64_t cnt = n_rays * n;
  if (cnt == 0) return NERF_OK;
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  int rc;
  if ((rc = c->sigc.reserve(cnt * 4)) || (rc = c->tc.reserve(cnt * 4)) || (rc = c->wc.reserve(cnt * 4))) return rc;
  NERF_CUDA(cudaMemcpyAsync(c->sigc.p, sigma, cnt * 4, cudaMemcpyHostToDevice, c->stream));
  NERF_CUDA(cudaMemcpyAsync(c->tc.p, t, cnt * 4, cudaMemcpyHostToDevice, c->stream));
  NERF_CUDA(launch_weights(c->sigc.as<float>(), c->tc.as<float>(), n_rays, n, far_, c->wc.as<float>(), c->stream));
  NERF_CUDA(cudaMemcpyAsync(w, c->wc.p, cnt * 4, cudaMemcpyDeviceToHost, c->stream));
  NERF_CUDA(cudaStreamSynchronize(c->stream));
  return NERF_OK;
}

// sample_importance (lib.rs:289-351) on caller-supplied weights, followed by render_block's merge + sort
// (lib.rs:416-420).
int nerf_sample_importance(nerf_ctx* c, const float* t, const float* w, int64_t n_rays, int n, const float* u,
                           int count, float* t_merged) {
  if (!c || n_rays < 0 || n < 3 || count < 0 || ((!t || !w || !t_merged) && n_rays > 0) || (count > 0 && !u)) {
    set_error("nerf_sample_importance: bad argument (n must be >= 3)");
    return NERF_E_ARG;
  }
  if (n_rays == 0) return NERF_OK;
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  const int nm = n + count;
  int rc;
  if ((rc = c->wc.reserve(n_rays * n * 4)) || (rc = c->tc.reserve(n_rays * n * 4)) ||
      (rc = c->uni.reserve(std::max<int64_t>(n_rays * count, 1) * 4)) || (rc = c->tf.reserve(n_rays * nm * 4)))
    return rc;
  NERF_CUDA(cudaMemcpyAsync(c->wc.p, w, n_rays * n * 4, cudaMemcpyHostToDevice, c->stream));
  NERF_CUDA(cudaMemcpyAsync(c->tc.p, t, n_rays * n * 4, cudaMemcpyHostToDevice, c->stream));
  if (count > 0) NERF_CUDA(cudaMemcpyAsync(c->uni.p, u, n_rays * count * 4, cudaMemcpyHostToDevice, c->stream));
  NERF_CUDA(launch_importance(c->tc.as<float>(), nullptr, c->wc.as<float>(), n, count, 0.0f, NERF_SAMPLING_ARRAYS, 0,
                              c->uni.as<float>(), RayMap{0, 0, 1, 0}, n_rays, c->tf.as<float>(), nullptr, c->stream));
  NERF_CUDA(cudaMemcpyAsync(t_merged, c->tf.p, n_rays * nm * 4, cudaMemcpyDeviceToHost, c->stream));
  NERF_CUDA(cudaStreamSynchronize(c->stream));
  return NERF_OK;
}

int nerf_integrate_rays(nerf_ctx* c, const float* rgb, const float* sigma, const float* t, int64_t n_rays, int n,
                        float far_, float* out) {
  if (!c || n_rays < 0 || n < 0 || ((!rgb || !sigma || !t) && n_rays * n > 0) || (!out && n_rays > 0)) {
    set_error("nerf_integrate_rays: bad argument");
    return NERF_E_ARG;
  }
  if (n_rays == 0) return NERF_OK;
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  const int64_t cnt = std::max<int64_t>(n_rays * n, 1);
  int rc;
  if ((rc = c->rgbf.reserve(cnt * 12)) || (rc = c->sigf.reserve(cnt * 4)) || (rc = c->tf.reserve(cnt * 4)) ||
      (rc = c->out.reserve(n_rays * 12)))
    return rc;
  if (n > 0) {
    NERF_CUDA(cudaMemcpyAsync(c->rgbf.p, rgb, n_rays * n * 12, cudaMemcpyHostToDevice, c->stream));
    NERF_CUDA(cudaMemcpyAsync(c->sigf.p, sigma, n_rays * n * 4, cudaMemcpyHostToDevice, c->stream));
    NERF_CUDA(cudaMemcpyAsync(c->tf.p, t, n_rays * n * 4, cudaMemcpyHostToDevice, c->stream));
  }
  NERF_CUDA(launch_composite(c->rgbf.as<float>(), c->sigf.as<float>(), c->tf.as<float>(), n_rays, n, far_,
                             RayMap{0, 0, 1, 0}, c->out.as<float>(), nullptr, nullptr, c->stream));
  NERF_CUDA(cudaMemcpyAsync(out, c->out.p, n_rays * 12, cudaMemcpyDeviceToHost, c->stream));
  NERF_CUDA(cudaStreamSynchronize(c->stream));
  return NERF_OK;
}

// ---------------------------------------------------------------------------------------- shared frame (multi-GPU)
int nerf_shared_frame_create(nerf_ctx* c, uint64_t bytes, void** d_ptr, unsigned char handle[NERF_IPC_HANDLE_BYTES]) {
  if (!c || !d_ptr || !handle || bytes == 0) {
    set_error("nerf_shared_frame_create: bad argument");
    return NERF_E_ARG;
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == NERF_IPC_HANDLE_BYTES, "CUDA IPC handle size");
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  void* p = nullptr;
  NERF_CUDA(cudaMalloc(&p, bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    set_error(std::string("cudaIpcGetMemHandle: ") + cudaGetErrorString(e));
    return NERF_E_CUDA;
  }
  std::memcpy(handle, &h, sizeof(h));
  *d_ptr = p;
  return NERF_OK;
}

int nerf_shared_frame_open(nerf_ctx* c, const unsigned char handle[NERF_IPC_HANDLE_BYTES], void** d_ptr) {
  if (!c || !d_ptr || !handle) {
    set_error("nerf_shared_frame_open: bad argument");
    return NERF_E_ARG;
  }
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle, sizeof(h));
  void* p = nullptr;
  // maps the owner's allocation into this process and enables peer access between the two devices (NVLink)
  NERF_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *d_ptr = p;
  return NERF_OK;
}

int nerf_shared_frame_close(nerf_ctx* c, void* d_ptr, int owner) {
  if (!c || !d_ptr) return NERF_OK;
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  NERF_CUDA(cudaDeviceSynchronize());
  if (owner) NERF_CUDA(cudaFree(d_ptr));
  else NERF_CUDA(cudaIpcCloseMemHandle(d_ptr));
  return NERF_OK;
}

int nerf_shared_frame_create_fd(nerf_ctx* c, uint64_t bytes, void** d_ptr, uint64_t* mapped_bytes, int* fd) {
  if (!c || !d_ptr || !mapped_bytes || !fd || bytes == 0) {
    set_error("nerf_shared_frame_create_fd: bad argument");
    return NERF_E_ARG;
  }
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  NERF_CUDA(cudaFree(nullptr));  // make sure the primary context exists
  const VmmApi& a = vmm_api();
  if (!a.ok) {
    set_error("CUDA virtual-memory entry points unavailable");
    return NERF_E_UNSUPPORTED;
  }
  const CUmemAllocationProp prop = frame_prop(c->device);
  size_t gran = 0;
  NERF_CU(a.granularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  const size_t size = (bytes + gran - 1) / gran * gran;
  CUmemGenericAllocationHandle h;
  NERF_CU(a.create(&h, size, &prop, 0));
  int out_fd = -1;
  CUresult r = a.export_handle(&out_fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) {
    a.release(h);
    set_error("cuMemExportToShareableHandle: CUresult " + std::to_string((int)r));
    return NERF_E_CUDA;
  }
  int rc = vmm_map(c, h, size, d_ptr);
  if (rc != NERF_OK) {
    ::close(out_fd);
    a.release(h);
    return rc;
  }
  *mapped_bytes = size;
  *fd = out_fd;
  return NERF_OK;
}

int nerf_shared_frame_open_fd(nerf_ctx* c, int fd, uint64_t mapped_bytes, void** d_ptr) {
  if (!c || !d_ptr || fd < 0 || mapped_bytes == 0) {
    set_error("nerf_shared_frame_open_fd: bad argument");
    return NERF_E_ARG;
  }
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  NERF_CUDA(cudaFree(nullptr));
  const VmmApi& a = vmm_api();
  if (!a.ok) {
    set_error("CUDA virtual-memory entry points unavailable");
    return NERF_E_UNSUPPORTED;
  }
  CUmemGenericAllocationHandle h;
  NERF_CU(a.import_handle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  int rc = vmm_map(c, h, (size_t)mapped_bytes, d_ptr);
  if (rc != NERF_OK) a.release(h);
  return rc;
}

int nerf_shared_frame_close_fd(nerf_ctx* c, void* d_ptr) {
  if (!c || !d_ptr) return NERF_OK;
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  auto it = c->vmm_frames.find(d_ptr);
  if (it == c->vmm_frames.end()) {
    set_error("nerf_shared_frame_close_fd: not a frame of this context");
    return NERF_E_ARG;
  }
  NERF_CUDA(cudaDeviceSynchronize());
  const VmmApi& a = vmm_api();
  const CUdeviceptr va = reinterpret_cast<CUdeviceptr>(d_ptr);
  NERF_CU(a.unmap(va, it->second.size));
  NERF_CU(a.addr_free(va, it->second.size));
  NERF_CU(a.release(it->second.handle));  // the memory goes away with the last mapping / handle, in whichever process
  c->vmm_frames.erase(it);
  return NERF_OK;
}

int nerf_set_option(nerf_ctx* c, int option, int value) {
  if (!c) {
    set_error("nerf_set_option: null context");
    return NERF_E_ARG;
  }
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  switch (option) {
    case NERF_OPT_SKIP_EMPTY_RAYS:
      c->skip_empty_rays = value != 0;
      return NERF_OK;
    default:
      set_error("nerf_set_option: unknown option " + std::to_string(option));
      return NERF_E_ARG;
  }
}

int nerf_get_option(nerf_ctx* c, int option, int* value) {
  if (!c || !value) {
    set_error("nerf_get_option: null argument");
    return NERF_E_ARG;
  }
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  switch (option) {
    case NERF_OPT_SKIP_EMPTY_RAYS:
      *value = c->skip_empty_rays ? 1 : 0;
      return NERF_OK;
    default:
      set_error("nerf_get_option: unknown option " + std::to_string(option));
      return NERF_E_ARG;
  }
}

int nerf_get_timing(nerf_ctx* c, nerf_timing* out) {
  if (!c || !out) {
    set_error("nerf_get_timing: null argument");
    return NERF_E_ARG;
  }
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);
  std::memset(out, 0, sizeof(*out));
  if (!c->have_timing) return NERF_OK;
  NERF_CUDA(cudaEventSynchronize(c->ev_end));
  NERF_CUDA(cudaEventElapsedTime(&out->total_ms, c->ev_begin, c->ev_end));
  for (size_t i = 0; i < c->spans_used; ++i) {
    float ms = 0.0f;
    NERF_CUDA(cudaEventElapsedTime(&ms, c->spans[i].a, c->spans[i].b));
    switch (c->spans[i].kind) {
      case ST_SAMPLING: out->sampling_ms += ms; break;
      case ST_MLP_COARSE: out->mlp_coarse_ms += ms; break;
      case ST_MLP_FINE: out->mlp_fine_ms += ms; break;
      case ST_COMPOSITE: out->composite_ms += ms; break;
      case ST_COARSE_REFINE:
        out->mlp_coarse_ms += ms;
        out->coarse_refine_ms += ms;
        break;
    }
  }
  out->kernel_launches = c->launches;
  out->coarse_samples = c->refine_total;
  out->rays = c->rays_total;
  out->rays_skipped = c->rays_skipped;
  if (c->refine_pending) {  // ev_end has completed, so the counters are final
    uint32_t cnt[2] = {0, 0};
    NERF_CUDA(cudaMemcpy(cnt, c->sel_count.p, sizeof(cnt), cudaMemcpyDeviceToHost));
    out->coarse_refined = cnt[1];
  }
  return NERF_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------- allocation-free compositing
// nerf_integrate_rays_noalloc: integrate_ray (lib.rs:176-195) for callers that may not allocate device memory (code
// under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  nerf_integrate_rays stages its
// host inputs and its output in four cudaMalloc'ed workspaces (DevBuf::reserve); here the same compositing kernel gets
// that scratch from ONE file-scope device array of fixed size, so the call performs no cudaMalloc / cudaFree and calls
// no allocating library routine; the reference has no counterpart.
//
//   - Limit: the staged inputs + output of a call must fit NERF_NOALLOC_MAX_BYTES (4 MiB; each of the four segments is
//     rounded up to 256 B).  At the reference's 64 + 128 samples per ray that is 1 088 rays per call.  Larger calls are
//     refused with NERF_E_UNSUPPORTED and a message that names the sizes - they are never truncated.
//   - The array lives in every GPU that loads the library, from load to exit, whether the entry point is used or not.
//   - Every call on a device shares the array and finds the previous call's contents in it.  Nothing has to start at
//     zero: the three input segments are overwritten by this call's host-to-device copies before the kernel reads them,
//     and the kernel writes all 3 floats of every ray of the output segment before it is copied back; bytes beyond this
//     call's sizes are never read.  Calls that overlap on the GPU (two contexts, or two processes' worth of streams, on
//     one device) RACE on the array: serialise them.  Calls on one context cannot overlap - the call holds the context's
//     mutex and returns only after its stream has drained.
//   - `g_noalloc_scratch` below is, in host code, a host-side stand-in for the symbol, NOT a device address: handing it
//     to a kernel faults (or, where the GPU can reach host memory, silently uses the wrong memory).  The device address
//     is fetched once per device with cudaGetSymbolAddress, with that device current, and that is what the kernel gets.
namespace nerf {
namespace {

constexpr size_t kNoAllocMaxBytes = NERF_NOALLOC_MAX_BYTES;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocMaxBytes];

constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mu;
unsigned char* g_noalloc_ptr[kNoAllocMaxDevices] = {};  // device address of g_noalloc_scratch, per device ordinal

// Device address of the scratch array on `device`, which must be the current device.
int noalloc_scratch(int device, unsigned char** out) {
  if (device < 0 || device >= kNoAllocMaxDevices) {
    set_error("nerf_integrate_rays_noalloc: device ordinal beyond the scratch table");
    return NERF_E_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (g_noalloc_ptr[device] == nullptr) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || p == nullptr) {
      set_error(std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e));
      return NERF_E_CUDA;
    }
    g_noalloc_ptr[device] = static_cast<unsigned char*>(p);
  }
  *out = g_noalloc_ptr[device];
  return NERF_OK;
}

inline size_t round256(size_t b) { return (b + 255) & ~(size_t)255; }

}  // namespace
}  // namespace nerf

int nerf_integrate_rays_noalloc(nerf_ctx* c, const float* rgb, const float* sigma, const float* t, int64_t n_rays, int n,
                                float far_, float* out) {
  if (!c || n_rays < 0 || n < 0 || ((!rgb || !sigma || !t) && n_rays * n > 0) || (!out && n_rays > 0)) {
    set_error("nerf_integrate_rays_noalloc: bad argument");
    return NERF_E_ARG;
  }
  if (n_rays == 0) return NERF_OK;
  // sizes first, in 64-bit arithmetic that cannot wrap: n_rays and n are bounded before they are multiplied
  if ((uint64_t)n_rays > kNoAllocMaxBytes || (uint64_t)n > kNoAllocMaxBytes) {
    set_error("nerf_integrate_rays_noalloc: request exceeds NERF_NOALLOC_MAX_BYTES (" + std::to_string(kNoAllocMaxBytes) +
              " B of static device scratch); use nerf_integrate_rays");
    return NERF_E_UNSUPPORTED;
  }
  const size_t samples = (size_t)n_rays * (size_t)n;
  const size_t b_rgb = round256(samples * 12), b_sig = round256(samples * 4), b_t = round256(samples * 4),
               b_out = round256((size_t)n_rays * 12);
  const size_t need = b_rgb + b_sig + b_t + b_out;
  if (need > kNoAllocMaxBytes) {
    set_error("nerf_integrate_rays_noalloc: " + std::to_string(n_rays) + " rays x " + std::to_string(n) + " samples need " +
              std::to_string(need) + " B of scratch, the static array holds NERF_NOALLOC_MAX_BYTES = " +
              std::to_string(kNoAllocMaxBytes) + " B; split the call or use nerf_integrate_rays");
    return NERF_E_UNSUPPORTED;
  }
  std::lock_guard<std::recursive_mutex> lk(c->mu);
  DeviceGuard g(c->device);  // the symbol's address is per device: resolve it with the context's device current
  unsigned char* base = nullptr;
  int rc = noalloc_scratch(c->device, &base);
  if (rc != NERF_OK) return rc;
  float* d_rgb = reinterpret_cast<float*>(base);
  float* d_sig = reinterpret_cast<float*>(base + b_rgb);
  float* d_t = reinterpret_cast<float*>(base + b_rgb + b_sig);
  float* d_out = reinterpret_cast<float*>(base + b_rgb + b_sig + b_t);
  cudaStream_t s = c->stream;
  if (c->busy_recorded) NERF_CUDA(cudaStreamWaitEvent(s, c->ev_busy, 0));  // order after a *_device call on another stream
  if (n > 0) {
    NERF_CUDA(cudaMemcpyAsync(d_rgb, rgb, samples * 12, cudaMemcpyHostToDevice, s));
    NERF_CUDA(cudaMemcpyAsync(d_sig, sigma, samples * 4, cudaMemcpyHostToDevice, s));
    NERF_CUDA(cudaMemcpyAsync(d_t, t, samples * 4, cudaMemcpyHostToDevice, s));
  }
  NERF_CUDA(launch_composite(d_rgb, d_sig, d_t, n_rays, n, far_, RayMap{0, 0, 1, 0}, d_out, nullptr, nullptr, s));
  NERF_CUDA(cudaMemcpyAsync(out, d_out, (size_t)n_rays * 12, cudaMemcpyDeviceToHost, s));
  NERF_CUDA(cudaStreamSynchronize(s));
  return NERF_OK;
}
