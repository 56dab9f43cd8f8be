#!/usr/bin/env python
"""bench.py — headline benchmark: lego 800x800, 64 coarse + 128 fine samples/ray, rays/s.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--precision mixed|fp16|bf16|fp16x3|fp32] [--impl reference]
                  [--gather peer|nccl] [--orbit VIEWS --res R] [--no-cpu]

A step = one full 800x800 frame (640 000 rays, 163.84 M MLP samples) through the whole hot path
(ray gen -> stratified -> coarse MLP -> importance resampling -> fine MLP -> compositing).
`value` times it with the frame buffer resident in HBM (nerf_render_device); `e2e` times the public host API
(pinned host output, D2H inside the timed region).  With N > 1 (torchrun) the frame is split into N contiguous
pixel-index ranges, one per rank, and assembled on rank 0 inside the timed region (strong scaling): by default every
rank's compositing kernel stores its rows straight into rank 0's frame buffer over NVLink (`--gather peer`, buffer
shared through the CUDA virtual-memory API), `--gather nccl` uses one group of NCCL point-to-point transfers instead.  The default precision is `mixed`
(coarse network split-fp16 x3, fine network fp16: the fastest mode whose frame stays within 2/255 of the f32 frame);
the all-fp16 frame rate is reported beside it.  `--orbit V --res R` runs the multi-pose sweep (BASELINE.json
configs[4]) instead.
`--impl reference` times the reference's CPU algorithm (the oracle restatement; the Rust crate cannot be built
here) on the host cores on a bounded ray sample of the same frame.
"""
import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W = H = 800
NC, NF = 64, 128
FLOP_COARSE = 982528    # sigma-only sample: dense0-7 + alpha
FLOP_FINE = 1186816     # full-network sample
WORKLOAD = "lego 800x800, 64 coarse + 128 fine samples/ray, deterministic sampling, tf_reference camera"


PRECISION_NOTE = {
    "mixed": "mixed: coarse densities at fp16x3 accuracy (fp16 pass over every coarse sample, then split-fp16 operands - 3 "
             "tensor-core MMAs per product - over the ~8 % of samples whose density can reach a coarse weight; bit-identical "
             "to fp16x3 over every coarse sample), fine network fp16; f32 accumulate, heads / sampling / compositing f32. "
             "See `parity` for the frame against the CPU oracle",
    "fp16": "fp16: both networks single-pass fp16 operands, f32 accumulate; PSNR ~71 dB, ~100 of 640 000 pixels (thin "
            "structures) above 2/255",
    "fp16x3": "fp16x3: both networks split-fp16 operands, 3 tensor-core MMAs per product",
    "tf32": "tf32: both networks tcgen05 kind::tf32 (32-bit containers, 10-bit mantissa), f32 accumulate; the accuracy of "
            "fp16 at less than half its speed",
}
# dram__bytes_read.sum + dram__bytes_write.sum of the fine-pass launch, 800x800 c64+f128: a CONSTANT copied from one
# `ncu --set full` capture (file below), not measured by this run
# the fine pass of the reference frame: 16 x mlp_tc_kernel<1,0,0,1> (0.55 GB read, 10.95 GB written: densities, zero colours
# and the 512-byte rows of the 16 % of the samples that have density) + 16 x mlp_colour_kernel (10.17 GB read, 0.27 GB
# written).  The fused single launch it replaced: 0.518 + 1.915 GB (profiles/r02_frame_ncu_full_summary.csv)
NCU_FINE_DRAM_GB = 0.549 + 10.949 + 10.40 + 0.269
NCU_FINE_DRAM_SOURCE = "profiles/r02_two_phase_ncu_full_summary.csv"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops_sustained", 1404.7), d.get("bf16_tflops", 1653.3), d.get("hbm_gbs", 6538.0), "measured"
    return 1400.0, 1590.0, 6650.0, "fallback"


# ------------------------------------------------------------------------------------------ CPU arm
def cpu_sample(threads=0, budget_s=15.0):
    """Times the oracle (reference CPU algorithm, reordered bit-identical GEMM loop) on a middle row band."""
    from oracle import pyoracle as O
    import nerf_rs_b200 as nb
    cam = O.camera_from_json(os.path.join(nb.DATA_ROOT, "tf_reference_samples.json"))
    oc, of = O.Network(nb.DATA_ROOT + "/coarse"), O.Network(nb.DATA_ROOT + "/fine")
    # torchrun exports OMP_NUM_THREADS=1 to every rank: ask the scheduler how many cores this process may use instead of
    # taking OpenMP's default, so that the CPU arm is the same at every N
    cores = len(os.sched_getaffinity(0)) if threads <= 0 else threads
    begin = (H // 2) * W
    n0 = 8 * cores  # calibration: one 8-ray strip per thread
    t0 = time.perf_counter()
    O.render(oc, of, cam, W, H, NC, NF, ray_begin=begin, ray_end=begin + n0, threads=cores)
    per_ray = (time.perf_counter() - t0) / n0
    n = int(max(64, min(W * 8, budget_s / per_ray)) // 8 * 8)
    return O, (oc, of, cam), cores, begin, n


def run_cpu_step(ctx, cores, begin, n, keep=None):
    O, (oc, of, cam) = ctx
    t0 = time.perf_counter()
    img = O.render(oc, of, cam, W, H, NC, NF, ray_begin=begin, ray_end=begin + n, threads=cores)
    dt = time.perf_counter() - t0
    if keep is not None:
        keep.append(img)
    return dt


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    O, ctx, cores, begin, n = cpu_sample(budget_s=10.0)
    for _ in range(min(args.warmup, 1)):
        run_cpu_step((O, ctx), cores, begin, n)
    times = [run_cpu_step((O, ctx), cores, begin, n) for _ in range(args.steps)]
    dt = float(np.mean(times))
    val = n / dt
    sample = f"{n} rays (pixel indices {begin}..{begin + n}) of the 800x800 frame per step, 256 MLP samples/ray"
    print(json.dumps({
        "impl": "reference", "metric": "rays/s", "value": val, "unit": "rays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample},
        "samples_per_s": val * (NC + NC + NF),
        "cpu_baseline": {"value": val, "unit": "rays/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------------ GPU arm
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.path = f"/tmp/bench_clocks_{os.getpid()}.csv"
        try:
            self.f = open(self.path, "w")
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            c = [x.strip() for x in line.split(",")]
            if len(c) < 7:
                continue
            try:
                sm.append(float(c[0]))
                mx.append(float(c[1]))
            except ValueError:
                continue
            for nm, v in zip(names, c[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.path)
        if sm:
            busy = [x for x in sm if x > 0.5 * max(sm)] or sm
            out = {"sm_mhz": float(np.median(busy)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                   "samples": len(sm)}
        return out


def gpu_arm(args):
    import torch
    import torch.distributed as dist
    import nerf_rs_b200 as nb

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: libnerf_b200 has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    r = nb.Renderer(nb.DATA_ROOT, device=local)
    cam = nb.camera_from_json(os.path.join(nb.DATA_ROOT, "tf_reference_samples.json"))
    from nerf_rs_b200 import shard
    n_rays = W * H
    rb, re_ = shard.ray_range(rank, world, n_rays)
    n_local = re_ - rb
    use_peer = world > 1 and args.gather == "peer"
    # N > 1, default: the frame is dealt out in one-row stripes (stripe k -> rank k % N).  A ray's cost depends on what
    # it hits (fewer refined coarse samples and no colour branch in empty rows), so contiguous ranges leave the ranks
    # with the object 20 % behind; the NCCL-gather variant keeps contiguous ranges (it sends one block per rank).
    striped = use_peer and not args.contiguous
    sf = None
    if use_peer:
        # the frame lives in rank 0's HBM; every rank maps it (VMM handle passed as a file descriptor) and its compositing
        # kernel stores there
        try:
            sf = shard.SharedFrame(r, n_rays, rank, world, dist)
            ok = 1
        except Exception as e:  # no peer mapping on this box: all ranks fall back to the NCCL gather together
            print(f"rank {rank}: shared frame unavailable ({e}); using the NCCL gather", file=sys.stderr)
            sf, ok = None, 0
        flag = torch.tensor([ok], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0:
            if sf is not None and rank != 0:
                sf.close()
            dist.barrier()
            if sf is not None and rank == 0:
                sf.close()
            sf, use_peer = None, False
    if striped and sf is None:
        striped = False
    if striped:
        n_local = len(shard.view_pixels(rank, world, W, n_rays))
    if use_peer:
        frame = sf.tensor() if rank == 0 else None
        mine = None
        mine_ptr = sf.row_ptr(rb)
    else:
        frame = torch.empty((n_rays, 3), dtype=torch.float32, device=dev)  # rank 0 assembles the frame here
        mine = frame[rb:re_] if world == 1 else torch.empty((n_local, 3), dtype=torch.float32, device=dev)
        mine_ptr = mine.data_ptr()
    # The e2e leg goes through the public host API (nerf_render) on EVERY rank: rank k renders its ray range straight into
    # its rows of one host frame.  With N > 1 that frame is a POSIX shared-memory file mapped by every rank and
    # page-locked (cudaHostRegister) in each, so the device-to-host copies are pinned copies as at N = 1.
    shm_path = None
    if world == 1:
        pinned = torch.empty((n_rays, 3), dtype=torch.float32).pin_memory()
        host_frame = pinned.numpy()
    else:
        shm_path = f"/dev/shm/nerf_b200_bench_{os.environ.get('MASTER_PORT', '0')}"
        if rank == 0:
            with open(shm_path, "wb") as f:
                f.truncate(n_rays * 12)
        dist.barrier()
        host_frame = np.memmap(shm_path, dtype=np.float32, mode="r+", shape=(n_rays, 3))
        err = torch.cuda.cudart().cudaHostRegister(host_frame.ctypes.data, n_rays * 12, 0)
        if int(err) != 0:
            print(f"rank {rank}: cudaHostRegister failed ({err}); the e2e copies run from pageable memory", file=sys.stderr)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream()

    def gather():
        if not use_peer:  # peer mode: the pixels are already in rank 0's buffer when the kernel retires
            shard.gather_frame(mine, frame if rank == 0 else None, rank, world, dist)

    def step_device(prec=None):
        if striped:
            r.render_striped_device(cam, W, H, sf.row_ptr(0), W, world, rank, NC, NF, precision=prec or args.precision,
                                    stream=stream.cuda_stream)
        else:
            r.render_device(cam, W, H, mine_ptr, NC, NF, precision=prec or args.precision, ray_begin=rb, ray_end=re_,
                            stream=stream.cuda_stream)
        gather()

    def step_e2e():
        # nerf_render with host pointers: kernels + the D2H copy of this rank's rows, synchronous
        if striped:
            r.render_striped(cam, W, H, host_frame, W, world, rank, NC, NF, precision=args.precision)
        else:
            r.render_image(cam, W, H, NC, NF, precision=args.precision, ray_begin=rb, ray_end=re_, out=host_frame[rb:re_])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    extra = {"refine_ms": 0.0, "refined_frac": 0.0, "launches": 0}

    def timed(fn, k, use_events):
        """K steps, L2 flushed (untimed) before each one; returns summed step time in ms (max over ranks)."""
        tot = 0.0
        extra["refine_ms"] = 0.0
        mlp_c = mlp_f = smp = comp = 0.0
        for _ in range(k):
            flush.zero_()
            barrier()
            if use_events:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                fn()
                e1.record(stream)
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1)
                tm = r.timing()  # CUDA events recorded by the library on the same stream around each kernel
                mlp_c += tm["mlp_coarse_ms"]; mlp_f += tm["mlp_fine_ms"]; smp += tm["sampling_ms"]; comp += tm["composite_ms"]
                extra["refine_ms"] += tm["coarse_refine_ms"] / k
                extra["refined_frac"] = tm["coarse_refined"] / max(tm["coarse_samples"], 1)
                extra["launches"] = tm["kernel_launches"]
            else:
                t0 = time.perf_counter()
                fn()
                torch.cuda.synchronize()
                ms = (time.perf_counter() - t0) * 1e3
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            tot += float(t.item())
        return tot, (mlp_c / k, mlp_f / k, smp / k, comp / k)

    for _ in range(args.warmup):
        step_device()
    barrier()
    clocks = ClockSampler(local) if rank == 0 else None
    total_ms, (mc, mf, sm_ms, cp_ms) = timed(step_device, args.steps, True)
    clk = clocks.stop() if clocks else {}
    head = dict(extra)  # the headline loop's refinement share and launch count (later legs overwrite `extra`)
    stages_all = [[mc, mf, sm_ms, cp_ms]]
    if world > 1:  # every rank's mean stage times, for the record (which rank bounds the frame, and in which kernel)
        stages_all = [None] * world
        dist.all_gather_object(stages_all, [mc, mf, sm_ms, cp_ms])
    fast_ms = None
    if args.precision == "mixed":  # the same frame with both networks in single-pass fp16, reported beside the headline
        step_device("fp16")
        fast_ms, _ = timed(lambda: step_device("fp16"), min(args.steps, 5), True)
        fast_ms /= min(args.steps, 5)
    # Opt-in leg, reported beside the headline and never as it: NERF_OPT_SKIP_EMPTY_RAYS (include/nerf_b200.h) gives
    # the rays without any coarse weight the background and skips their fine pass - NOT the reference's image.
    opt_in = None
    if world == 1 and args.precision == "mixed":
        r.set_option(nb.OPT_SKIP_EMPTY_RAYS, 1)
        try:
            step_device()
            k = min(args.steps, 5)
            skip_ms, (s_mc, s_mf, s_sm, s_cp) = timed(step_device, k, True)
            tm = r.timing()
            torch.cuda.synchronize()
            opt_in = {"ms_per_frame": skip_ms / k, "rays_per_s": n_rays / (skip_ms / k * 1e-3),
                      "rays_skipped_fraction": tm["rays_skipped"] / max(tm["rays"], 1),
                      "stage_ms": {"mlp_coarse": s_mc, "mlp_fine": s_mf, "sampling": s_sm, "composite": s_cp},
                      "frame": frame.cpu().numpy().copy(),
                      "note": "opt-in nerf_set_option(NERF_OPT_SKIP_EMPTY_RAYS): rays all of whose coarse weights are "
                              "exactly 0 get the background without a fine pass; every other ray is bit-identical to "
                              "the headline frame. Not the reference's pipeline, so never the headline value"}
        finally:
            r.set_option(nb.OPT_SKIP_EMPTY_RAYS, 0)
        extra.update(head)
    step_e2e()
    e2e_ms, _ = timed(step_e2e, args.steps, False)
    # ---- parity, at every N: the frame the ranks assembled (through the device path and through the host API) against
    # the frame rank 0 renders alone, bit for bit, and against the CPU oracle's whole frame (tests/golden/frame_800.npz)
    step_device()
    if use_peer:
        sf.complete(stream)
    barrier()
    parity = None
    if rank == 0:
        dev_frame = frame.cpu().numpy()
        host_copy = np.array(host_frame)  # what the e2e leg's last step left in host memory
        solo = r.render_image(cam, W, H, NC, NF, precision=args.precision) if world > 1 else dev_frame
        parity = {"bit_identical_to_1gpu": bool(np.array_equal(dev_frame, solo) and np.array_equal(host_copy, solo)),
                  "assembled_by": "device path (" + ("peer stores" if use_peer else "nccl gather" if world > 1 else "one rank") +
                                  ") and host API (nerf_render per rank into one shared pinned frame), both compared"}
        gpath = os.path.join(ROOT, "tests", "golden", "frame_800.npz")
        if os.path.exists(gpath):
            z = np.load(gpath)
            white = np.unpackbits(z["white"])[:n_rays].astype(bool)
            want = np.ones((n_rays, 3), np.float32)
            want[~white] = z["rgb"]
            d = np.abs(host_copy - want).max(axis=1)
            parity.update({
                "vs": "CPU oracle (the reference's algorithm restated), whole 800x800 frame, identical deterministic sample "
                      "positions (tests/golden/frame_800.npz)",
                "pixels": int(d.size), "max_abs_diff_x255": float(d.max() * 255.0),
                "pixels_over_2_255": int((d > 2.0 / 255.0).sum()), "pixels_over_1_255": int((d > 1.0 / 255.0).sum()),
                "psnr_db": float(10.0 * np.log10(1.0 / max(float(np.mean((host_copy.astype(np.float64) - want) ** 2)), 1e-20)))})
    if rank == 0 and opt_in is not None:
        f = opt_in.pop("frame")
        changed = (f != dev_frame).any(axis=1)
        opt_in["pixels_differing_from_headline_frame"] = int(changed.sum())
        opt_in["all_of_them_background"] = bool((f[changed] == 1.0).all())
        if os.path.exists(gpath):
            d = np.abs(f - want).max(axis=1)
            opt_in["vs_oracle_reference_frame"] = {
                "max_abs_diff_x255": float(d.max() * 255.0), "pixels_over_2_255": int((d > 2.0 / 255.0).sum()),
                "psnr_db": float(10.0 * np.log10(1.0 / max(float(np.mean((f.astype(np.float64) - want) ** 2)), 1e-20)))}
    barrier()
    rgb8_ms = many_ms = None
    if world == 1:
        # the same frame returned as 8-bit PPM bytes (3 B/ray over PCIe instead of 12), and a
        # four-pose sweep whose device-to-host copies overlap the next pose's render
        def step_rgb8():
            r.render_rgb8(cam, W, H, NC, NF, precision=args.precision)
        step_rgb8()
        rgb8_ms, _ = timed(step_rgb8, max(args.steps // 2, 1), False)
        cams4 = [cam] * 4
        def step_many():
            r.render_many(cams4, W, H, NC, NF, precision=args.precision)
        step_many()  # the first call allocates the 8-bit device frames and the pinned double buffer
        many_ms, _ = timed(step_many, 1, False)

    if rank == 0:
        import ctypes
        h2d_bytes = ctypes.sizeof(nb.Camera) + ctypes.sizeof(nb.Sampling)  # the only per-step input: the pose
        ms = total_ms / args.steps
        value = n_rays / (ms * 1e-3)
        e2e_val = n_rays / (e2e_ms / args.steps * 1e-3)
        sustained, burst, hbm, kind = load_peaks()
        fine_flop = n_local * (NC + NF) * FLOP_FINE  # rank 0's fine-pass launch
        ach = fine_flop / (mf * 1e-3) / 1e12
        frame_tflops = n_rays * (NC * FLOP_COARSE + (NC + NF) * FLOP_FINE) / (ms * 1e-3) / 1e12
        fine_kernel = {"fp16x3": "mlp_tc_x3_kernel (fine pass; 3 MMAs per product, credited once)",
                       "fp32": "mlp_fp32_kernel (fine pass; CUDA cores)",
                       "tf32": "mlp_tc_tf32_kernel (fine pass)"}.get(
            args.precision, "fine pass = per group of 40 960 rays one mlp_tc_kernel<store h8> launch (trunk + density head, "
                            "every sample) and one mlp_colour_kernel launch (colour branch, the samples with density); "
                            "credited the reference's FLOP for every sample, stage time = all launches of the pass")
        coarse_kernel = {"fp16x3": "mlp_tc_x3_kernel (coarse pass, sigma only; 3 MMAs per product, credited once)",
                         "fp32": "mlp_fp32_kernel (coarse pass)",
                         "mixed": "mlp_tc_kernel fp16 over every coarse sample + coarse_select_kernel + mlp_tc_x3_kernel over "
                                  "the selected samples (coarse pass, sigma only)"}.get(args.precision, "mlp_tc_kernel (coarse pass)")
        c_ach = n_local * NC * FLOP_COARSE / (mc * 1e-3) / 1e12
        dominant_is_fine = mf >= mc
        line = {
            "metric": "rays/s", "value": value, "unit": "rays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "fp16" if args.precision == "mixed" else args.precision, "data": "synthetic",
            "config": {"workload": WORKLOAD, "precision": PRECISION_NOTE.get(args.precision, args.precision), "parallelism": (f"one-row stripes dealt to {world} ranks (stripe k -> rank k % {world})" if striped else f"ray-range x{world}"),
                       "l2": "flushed (256 MiB memset, untimed) before every step; per-frame intermediates ~3 GB >> 126 MB L2",
                       "timing": "value: per-rank CUDA events on the launching stream around the step (render + gather), every "
                                 "step started from a barrier, max over ranks; e2e: host wall clock from the same barrier to the "
                                 "return of nerf_render on each rank, max over ranks",
                       "gather": ("none" if world == 1 else
                                  "peer stores: every rank's compositing kernel writes its rows into rank 0's frame buffer over "
                                  "NVLink (the buffer is shared per allocation through the CUDA virtual-memory API), inside the timed region; "
                                  "no collective on the data path" if use_peer else
                                  "NCCL point-to-point group: every rank sends its rgb range to rank 0, inside the timed region")},
            "samples_per_s": value * (NC + NC + NF),
            "tflops_frame": frame_tflops,
            "stage_ms_rank0": {"mlp_coarse": mc, "mlp_fine": mf, "sampling": sm_ms, "composite": cp_ms,
                               "mlp_coarse_refine_part": head["refine_ms"]},
            "coarse_refined_fraction": head["refined_frac"],
            "stage_ms_by_rank": [[round(x, 3) for x in st] for st in stages_all],  # [coarse, fine, sampling, composite]
            "gpu_launches": head["launches"] * args.steps,  # counted by the library (nerf_timing.kernel_launches) per step
            "parity": parity,
            "clocks": clk,
            "e2e": {"value": e2e_val, "unit": "rays/s", "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": n_rays * 12},
            "fp16_both_networks": None if fast_ms is None else {
                "value": n_rays / (fast_ms * 1e-3), "unit": "rays/s", "ms_per_step": fast_ms,
                "note": PRECISION_NOTE["fp16"] + "; same timing method as `value`"},
            "opt_in_skip_empty_rays": opt_in,
            "e2e_rgb8": None if rgb8_ms is None else {
                "value": n_rays / (rgb8_ms / max(args.steps // 2, 1) * 1e-3), "unit": "rays/s",
                "d2h_bytes_per_step": n_rays * 3, "api": "nerf_render_rgb8 (device-side save_ppm quantiser)"},
            "e2e_many": None if many_ms is None else {
                "value": 4 * n_rays / (many_ms * 1e-3), "unit": "rays/s", "views": 4,
                "api": "nerf_render_many (pinned double buffer, copies overlap the next view)"},
            # the kernel with the largest stage time on rank 0; `frame` is the whole step against the same peak
            "roofline": {"bound": "tensor", "kernel": fine_kernel if dominant_is_fine else coarse_kernel,
                         "achieved": ach if dominant_is_fine else c_ach, "peak": sustained, "unit": "TFLOP/s",
                         "frac": (ach if dominant_is_fine else c_ach) / sustained, "peak_kind": f"bf16 sustained, {kind}",
                         "frac_of_burst": (ach if dominant_is_fine else c_ach) / burst,
                         "flop_per_sample": FLOP_FINE if dominant_is_fine else FLOP_COARSE,
                         "stage_ms": mf if dominant_is_fine else mc,
                         "frame": {"achieved": frame_tflops, "frac": frame_tflops / sustained / world,
                                   "frac_of_burst": frame_tflops / burst / world,
                                   "note": "algorithmic FLOP of the whole step / step time / (peak x n_gpus)"},
                         "traffic": (NCU_FINE_DRAM_GB * n_local / n_rays
                                     if dominant_is_fine and args.precision in ("fp16", "bf16", "mixed") else None),
                         "traffic_unit": "GB per launch",
                         "traffic_source": f"constant from one ncu --set full capture ({NCU_FINE_DRAM_SOURCE}, dram read + "
                                           "write summed over the fine pass's launches), scaled to this rank's share of the rays; not "
                                           "measured by this run"},
            "roofline_other": {"bound": "tensor", "kernel": coarse_kernel if dominant_is_fine else fine_kernel,
                               "achieved": c_ach if dominant_is_fine else ach, "peak": sustained, "unit": "TFLOP/s",
                               "frac": (c_ach if dominant_is_fine else ach) / sustained,
                               "stage_ms": mc if dominant_is_fine else mf,
                               "note": "credited FLOP (one product per MAC); the split-operand kernel executes three"},
        }
        if world == 1 and not args.no_cpu:
            O, ctx, cores, begin, n = cpu_sample()
            band = []
            dt = run_cpu_step((O, ctx), cores, begin, n, keep=band)
            # BASELINE.json's metric ends "PSNR vs ref": the rows the CPU arm just rendered against the same rows of the
            # frame the e2e leg left in host memory (identical deterministic sample positions on both sides)
            d = np.abs(host_copy[begin:begin + n] - band[0])
            line["parity_live_band"] = {
                "vs": "CPU oracle run just now on this box, same rays, same sample positions",
                "rays": n, "max_abs_diff_x255": float(d.max() * 255.0),
                "pixels_over_2_255": int((d.max(axis=1) > 2.0 / 255.0).sum()),
                "psnr_db": float(10.0 * np.log10(1.0 / max(float(np.mean(d.astype(np.float64) ** 2)), 1e-20)))}
            line["cpu_baseline"] = {"value": n / dt, "unit": "rays/s", "cores": cores, "kind": "port",
                                    "sample": f"{n} rays of the same frame (pixel indices {begin}..{begin + n}), "
                                              f"{dt:.1f} s, oracle reordered bit-identical GEMM loop"}
        print(json.dumps(line))
    if sf is not None:
        frame = None
        dist.barrier()
        if rank != 0:
            sf.close()  # unmap before the owner frees
        dist.barrier()
        if rank == 0:
            sf.close()
    if shm_path is not None:
        torch.cuda.cudart().cudaHostUnregister(host_frame.ctypes.data)
        dist.barrier()
        if rank == 0:
            os.unlink(shm_path)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    r.close()


def orbit_arm(args):
    """BASELINE.json configs[4]: a synthetic orbit of `--orbit` poses at `--res`^2, views sharded across the ranks
    (contiguous blocks, no data-path collective), each rank through nerf_render_many: 8-bit frames to pinned host
    memory, the copy of view v overlapping the render of view v + 1.  Host wall clock around the whole sweep (the call
    ends with every frame in host memory), max over ranks."""
    import torch
    import torch.distributed as dist
    import nerf_rs_b200 as nb
    from nerf_rs_b200 import shard

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: libnerf_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    r = nb.Renderer(nb.DATA_ROOT, device=local)
    cams = nb.orbit_cameras(os.path.join(nb.DATA_ROOT, "tf_reference_samples.json"), args.orbit)
    vb, ve = shard.ray_range(rank, world, args.orbit)  # the same contiguous split, over views
    mine = cams[vb:ve]
    R = args.res
    r.render_many(mine[:1], R, R, NC, NF, precision=args.precision)  # warm-up: workspaces, pinned buffers, clocks
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = ClockSampler(local) if rank == 0 else None
    t0 = time.perf_counter()
    frames = r.render_many(mine, R, R, NC, NF, precision=args.precision)
    torch.cuda.synchronize()
    t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    clk = clocks.stop() if clocks else {}
    checksum = int(frames.astype(np.uint64).sum()) if len(mine) else 0
    parity = None
    if rank == 0 and len(mine) and not args.no_cpu:
        # pose 0, eight rows through the middle of the frame: the CPU oracle at this resolution against the 8-bit frame
        from oracle import pyoracle as O
        p0 = nb.orbit_poses(os.path.join(nb.DATA_ROOT, "tf_reference_samples.json"), args.orbit)[0]
        ocam = O.camera_from_values(p0["near"], p0["far"], p0["camera_origin"], p0["camera_forward"], p0["camera_up"], p0["hwf"])
        oc, of = O.Network(nb.DATA_ROOT + "/coarse"), O.Network(nb.DATA_ROOT + "/fine")
        begin = (R // 2) * R
        band = O.render(oc, of, ocam, R, R, NC, NF, ray_begin=begin, ray_end=begin + 8 * R,
                        threads=len(os.sched_getaffinity(0)))
        want = O.quantize(band).reshape(-1, 3).astype(np.int16)
        got = frames[0][begin:begin + 8 * R].astype(np.int16)
        d = np.abs(got - want).max(axis=1)
        parity = {"vs": "CPU oracle, pose 0, rows %d..%d of the %dx%d frame, 8-bit values" % (R // 2, R // 2 + 7, R, R),
                  "rays": int(d.size), "max_abs_diff_of_255": int(d.max()), "pixels_over_2_255": int((d > 2).sum()),
                  "psnr_db": float(10 * np.log10(255.0 ** 2 / max(float(np.mean((got - want).astype(np.float64) ** 2)), 1e-12)))}
    if rank == 0:
        dt = float(t.item())
        n_rays = args.orbit * R * R
        print(json.dumps({
            "metric": "rays/s", "value": n_rays / dt, "unit": "rays/s", "n_gpus": world, "steps": 1, "warmup": 1,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "fp16" if args.precision == "mixed" else args.precision, "data": "synthetic",
            "config": {"workload": f"orbit of {args.orbit} lego poses at {R}x{R}, 64 coarse + 128 fine samples/ray, "
                                   "deterministic sampling, 8-bit frames returned to host memory",
                       "precision": PRECISION_NOTE.get(args.precision, args.precision),
                       "parallelism": f"views x{world} ({ve - vb} views on rank 0)"},
            "samples_per_s": n_rays / dt * (NC + NC + NF), "views_per_s": args.orbit / dt,
            "tflops": n_rays * (NC * FLOP_COARSE + (NC + NF) * FLOP_FINE) / dt / 1e12,
            "d2h_bytes": n_rays * 3, "clocks": clk, "rank0_checksum": checksum, "parity": parity}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    r.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--precision", default="mixed", choices=["fp16", "bf16", "tf32", "mixed", "fp16x3", "fp32"])
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"],
                    help="N > 1: peer = compositing kernels store into rank 0's frame over NVLink; nccl = send/recv gather")
    ap.add_argument("--contiguous", action="store_true",
                    help="N > 1 with --gather peer: contiguous pixel-index ranges instead of interleaved one-row stripes")
    ap.add_argument("--orbit", type=int, default=0, help="run the multi-pose sweep with this many views instead")
    ap.add_argument("--res", type=int, default=1600, help="orbit sweep: frame width = height")
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm(args)
    elif args.orbit > 0:
        orbit_arm(args)
    else:
        gpu_arm(args)


if __name__ == "__main__":
    main()
