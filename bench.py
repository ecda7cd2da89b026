#!/usr/bin/env python
"""bench.py -- Mrays/s of the 4K free-form lens render (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl ours|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One step = one whole frame of the workload (default: cfg4, makeEllipsoid(181,91,(1,4,2)) -> 98 826 patches,
3840x2160, 16 rays/pixel = 132 710 400 primary rays), image rows dealt to the N ranks in round-robin stripes
of 8 rows, tables replicated, one NCCL all-gather of the rendered stripes per step when N > 1 (strong
scaling: the frame is fixed).  `value` = primary rays of the frame / device time per step, max over ranks.
`e2e` goes through the host-buffer C-ABI call (bz_render_host, N = 1) or render + gather + device->pinned-host
copy of the frame (N > 1), every step.  `--impl reference` times the reference's own CPU code
(oracle/_ref/libbzref.so: reference sources + the one documented hoist) on the host cores.
Prints ONE JSON line.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "cuda-bezier-triangle-raytracer_b200"))

FLOP_STAGE_A, FLOP_STAGE_B, FLOP_SNELL = 33, 1750, 45          # algorithmic FP32 ops as the reference writes them
BYTES_PER_PIXEL_OUT, BYTES_PER_PIXEL_ACCUM = 4, 8               # RGBA8 written; u32 accumulator zeroed + read back by the resolve


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--workload", default=None)
    p.add_argument("--cpu-seconds", type=float, default=15.0, help="target wall time of one CPU-baseline sample")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--pipeline", default="auto", choices=["auto", "persistent", "wavefront"], help="how the lens tiles are traced (bz_lens_set_pipeline)")
    p.add_argument("--mode", default="exact", choices=["exact", "fast"],
                   help="exact: the parity build (-fmad=false, bit-exact with the reference); fast: BZ_MODE_FAST (FMA, MUFU reciprocals; north_star's tolerance)")
    return p.parse_args()


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md recipe)"""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.rows, self.proc, self.thread = [], None, None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(device), "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1 + 0.2 and len(r) >= 7] or [r for _, r in self.rows if len(r) >= 7]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[0]) for r in rows)
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "power_w_max": max(float(r[2]) for r in rows),
                "samples": len(rows), "reasons": reasons}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def reference_sample(name, seconds, threads):
    """Times the compiled reference (oracle/_ref) on a pixel lattice of the workload: full spp per sampled pixel,
    lattice step chosen from a calibration run so that the sample takes about `seconds` on `threads` threads.
    -> dict(value Mrays/s, rays, seconds, step, calls)"""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import bzref
    from bezier_b200 import workloads
    if not bzref.available():
        return None
    spec, mesh, cam_kw, scene_kw = workloads.setup(name, bzref.RefMesh)
    lens = bzref.RefLens(spec["ri"], bzref.RefBezierMesh(mesh))
    cam, scene = bzref.make_camera(spec["width"], spec["height"], **cam_kw), bzref.make_scene(**scene_kw)

    def run(step):
        t0 = time.perf_counter()
        _, stats, rays = lens.render(cam, scene, pixel_step=step, threads=threads)
        return time.perf_counter() - t0, rays, stats

    step = max(spec["width"], spec["height"]) // 16           # coarse lattice first, refined until the sample fills the budget
    dt, rays, stats = run(step)
    while dt < 0.6 * seconds and step > 1:
        step = max(1, min(step - 1, int(step * (max(dt, 1e-3) / seconds) ** 0.5)))
        dt, rays, stats = run(step)
    return dict(value=rays / dt / 1e6, rays=int(rays), seconds=dt, step=step, calls=int(stats.refract_calls),
                sample="every %d-th pixel in x and y, all %d samples of each: %d primary rays, %d refract() calls, %.1f s on %d threads"
                       % (step, spec["strata"] ** 2, rays, stats.refract_calls, dt, threads))


def main_reference(args, name):
    """--impl reference: the reference's own CPU implementation on the host cores, rank 0 only"""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from bezier_b200 import workloads
    spec = workloads.WORKLOADS[name]
    threads = host_cores()
    samples = []
    for _ in range(args.warmup + args.steps):
        r = reference_sample(name, args.cpu_seconds, threads)
        if r is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libbzref.so is not built (needs /root/reference at build time)"}))
            return
        samples.append(r)
    timed = samples[args.warmup:]
    rays = sum(s["rays"] for s in timed)
    seconds = sum(s["seconds"] for s in timed)
    value = rays / seconds / 1e6
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": seconds / len(timed) * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": name, "patches": None, "image": "%dx%d" % (spec["width"], spec["height"]), "spp": spec["strata"] ** 2,
                   "note": "each step is a bounded pixel-lattice sample of the frame, rate = sampled primary rays / wall time"},
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": threads, "kind": "reference", "sample": timed[-1]["sample"], "cpu": cpu_model()},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def main_ours(args, name):
    import numpy as np
    import torch
    import torch.distributed as dist
    import bezier_b200 as bz
    from bezier_b200 import sharding, workloads

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available() or bz.device_count() == 0:
        raise SystemExit("bench.py: no CUDA device -- the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL may print its version banner on stdout when the communicator is created: keep stdout for the ONE JSON line
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)
    assert world == args.gpus or world == 1, "--gpus must match the torchrun world size"

    # ---- setup (untimed): host preprocessing in the product's own C++ (bz::Mesh / bz::BezierMesh), tables to HBM ----
    t_setup = time.perf_counter()
    spec, mesh, cam_kw, scene_kw = workloads.setup(name, bz.Mesh)
    bezier = bz.BezierMesh(mesh)
    lens = bz.BezierLens(spec["ri"], bezier, device=local)
    if args.mode == "fast":
        lens.set_mode(bz.MODE_FAST)
    lens.set_pipeline({"auto": bz.PIPELINE_AUTO, "persistent": bz.PIPELINE_PERSISTENT, "wavefront": bz.PIPELINE_WAVEFRONT}[args.pipeline])
    setup_s = time.perf_counter() - t_setup
    cam, scene = bz.make_camera(spec["width"], spec["height"], **cam_kw), bz.make_scene(**scene_kw)
    width, height, spp = spec["width"], spec["height"], spec["strata"] ** 2
    stripe = workloads.STRIPE_ROWS
    rows_local = bz.striped_rows(height, stripe, rank, world)
    band_rows = max(bz.striped_rows(height, stripe, r, world) for r in range(world))
    frame_rays = width * height * spp

    dev = torch.device("cuda", local)
    band = torch.zeros((band_rows, width, 4), dtype=torch.uint8, device=dev)          # this rank's stripes (padded to the largest band)
    bands = torch.zeros((world, band_rows, width, 4), dtype=torch.uint8, device=dev) if world > 1 else None
    image = torch.zeros((height, width, 4), dtype=torch.uint8, device=dev) if (world > 1 and rank == 0) else None
    stats = torch.zeros(len(bz.STAT_NAMES), dtype=torch.int64, device=dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)               # > 126 MB L2
    host_image = torch.empty((height, width, 4), dtype=torch.uint8).pin_memory() if rank == 0 else None
    kernel_ms = []

    def step(record):
        flush.fill_(1)                                                                  # L2 flush between iterations
        lens.render_striped_device(cam, scene, stripe, rank, world, band, stats=stats)
        if world > 1:
            sharding.gather_bands(band, bands)                                          # NCCL all-gather over NVLink
            if rank == 0:
                bz.assemble_stripes_device(bands, band_rows, width, height, stripe, world, image)
        if record:
            kernel_ms.append(lens.last_kernel_ms())                                     # events on the launching stream

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step(False)
    barrier()
    stats.zero_()
    sampler = ClockSampler(local) if rank == 0 else None
    begin, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    wall0 = time.time()
    begin.record()
    for _ in range(args.steps):
        step(True)
    end.record()
    barrier()
    wall1 = time.time()
    elapsed = torch.tensor([begin.elapsed_time(end)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(elapsed, op=dist.ReduceOp.MAX)
    ms_per_step = float(elapsed.item()) / args.steps
    clocks = sampler.stop(wall0, wall1) if sampler else None

    # ---- work counters of the timed region, summed over ranks ----
    work = stats.clone()
    kern = torch.tensor([sum(kernel_ms) / max(len(kernel_ms), 1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(work, op=dist.ReduceOp.SUM)
        dist.all_reduce(kern, op=dist.ReduceOp.MAX)
    work = dict(zip(bz.STAT_NAMES, [int(x) // args.steps for x in work.tolist()]))      # per step (frame)
    own = dict(zip(bz.STAT_NAMES, [int(x) // args.steps for x in stats.tolist()]))      # this rank, per launch

    # ---- e2e: host buffers in the timed region, every step ----
    e2e_steps = max(2, min(args.steps, 5))

    def e2e_step():
        if world == 1:
            lens.render_host_into(cam, scene, host_image.data_ptr())                   # C ABI: bz_render_host, D2H inside
        else:
            step(False)
            if rank == 0:
                host_image.copy_(image, non_blocking=True)
            barrier()

    for _ in range(max(1, min(args.warmup, 3))):                                       # the call's device scratch is allocated on first use
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    e2e_t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_s = float(e2e_t.item())

    if rank == 0:
        peak_tf, _ = bz.fp32_peak_probe(local)
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except OSError:
            pass
        traffic, traffic_source = None, None                   # DRAM bytes per launch of the same kernel from the committed ncu --set full capture
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                t = json.load(f)
            if t.get("workload") == name and world == 1:
                traffic, traffic_source = t["dram_bytes_read"] + t["dram_bytes_write"], t["source"]
        except (OSError, KeyError, ValueError):
            pass
        flop = FLOP_STAGE_A * own["stage_a"] + FLOP_STAGE_B * own["stage_b"] + FLOP_SNELL * own["snell"]     # rank 0's launch
        k_ms = float(kern.item())
        own_pixels = rows_local * width
        hbm_bytes = own_pixels * (BYTES_PER_PIXEL_OUT + BYTES_PER_PIXEL_ACCUM)
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        achieved_tf = flop / (k_ms * 1e-3) / 1e12
        roofline = {
            "bound": "fp32_alu", "kernel": "bz_warp_trace_kernel<RENDER> (packet walk; kernel_ms also holds bz_sky_kernel, 0.5 of 18.6 ms on cfg4)", "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s",
            "frac": achieved_tf / peak_tf, "traffic": traffic, "traffic_unit": "bytes per launch", "traffic_source": traffic_source,
            "frac_of_no_fma_ceiling": achieved_tf / (peak_tf / 2.0),
            "no_fma_note": "the parity build has no FMA contraction (bit-exact with the reference's x86 build): one flop per lane and instruction, half the probe's peak",
            "peak_source": "dependent-FMA probe kernel in this process (2 flop/lane/instr); nominal 148 SM x 128 lanes x 2 x %.0f MHz = %.1f"
                           % (peaks.get("sm_max_mhz", 1965.0), 148 * 128 * 2 * peaks.get("sm_max_mhz", 1965.0) * 1e6 / 1e12),
            "flop_model": "33 x stage_a + 1750 x stage_b + 45 x snell, executed counts (not brute-force-equivalent)",
            "kernel_ms": k_ms, "stage_a": own["stage_a"], "stage_b": own["stage_b"], "snell": own["snell"], "node_visits": own["node_visits"],
            "hbm": {"algorithmic_bytes": hbm_bytes, "achieved_gbs": hbm_bytes / (k_ms * 1e-3) / 1e9, "peak_gbs": hbm_peak,
                    "frac": hbm_bytes / (k_ms * 1e-3) / 1e9 / hbm_peak, "peak_source": "MEASURED_PEAKS.json" if "hbm_gbs" in peaks else "fallback"},
        }
        launches_per_step = lens.last_kernel_launches() + (1 if world > 1 else 0)       # sky + trace + resolve (+ assemble on rank 0)
        line = {
            "metric": "Mrays/s", "value": frame_rays / (ms_per_step * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": name, "patches": lens.patch_count, "image": "%dx%d" % (width, height), "spp": spp, "primary_rays_per_step": frame_rays,
                       "refractive_index": spec["ri"], "max_crossings": spec["max_crossings"], "parallelism": "row stripes of %d x %d ranks" % (stripe, world),
                       "l2": "flushed between steps (256 MiB fill)", "mode": "culled (exact candidate-set-preserving BVH)" + (", BZ_MODE_FAST arithmetic" if args.mode == "fast" else ""), "fmad": args.mode == "fast", "pipeline": args.pipeline,
                       "setup_s": setup_s},
            "e2e": {"value": frame_rays / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": int(bz.load().bz_sizeof_camera() + bz.load().bz_sizeof_scene()),
                    "d2h_bytes_per_step": width * height * 4, "ms_per_step": e2e_s * 1e3,
                    "call": "bz_render_host into pinned host memory" if world == 1 else "bz_render_striped + NCCL all-gather + bz_assemble_stripes + D2H to pinned host"},
            "gpu_launches": launches_per_step * args.steps,
            "clocks": clocks, "roofline": roofline, "work_per_step": work,
            "refract_calls_per_s": work["refract_calls"] / (ms_per_step * 1e-3),
        }
        if not args.no_cpu_baseline and world == 1:
            threads = host_cores()
            ref = reference_sample(name, args.cpu_seconds, threads)
            if ref is not None:
                line["cpu_baseline"] = {"value": ref["value"], "unit": "Mrays/s", "cores": threads, "kind": "reference", "sample": ref["sample"], "cpu": cpu_model()}
            else:
                line["cpu_baseline"] = {"value": None, "unit": "Mrays/s", "cores": threads, "kind": "reference", "sample": "oracle/_ref not built"}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main_illumination(args, name):
    """secondary workload: the illumination mode (emit -> group by hemisphere patch -> bounce loop -> target bins), 1 GPU"""
    import torch
    import bezier_b200 as bz
    from bezier_b200 import workloads
    if int(os.environ.get("RANK", "0")) != 0:
        return
    if not torch.cuda.is_available() or bz.device_count() == 0:
        raise SystemExit("bench.py: no CUDA device -- the product has no CPU path")
    spec, mesh, em_kw, tg_kw = workloads.setup_illumination(name, bz.Mesh)
    lens = bz.BezierLens(spec["ri"], bz.BezierMesh(mesh), device=0)
    emitter, target = bz.make_emitter(**em_kw), bz.make_target(**tg_kw)
    dev = torch.device("cuda", 0)
    bins = torch.zeros((target.bins_v, target.bins_u), dtype=torch.int32, device=dev)
    stats = torch.zeros(len(bz.STAT_NAMES), dtype=torch.int64, device=dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)

    def step():
        flush.fill_(1)
        bins.zero_()
        lens.illuminate_device(emitter, target, bins, spec["max_crossings"], stats=stats)

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    stats.zero_()
    sampler = ClockSampler(0)
    begin, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    wall0 = time.time()
    begin.record()
    for _ in range(args.steps):
        step()
    end.record()
    torch.cuda.synchronize()
    wall1 = time.time()
    ms = begin.elapsed_time(end) / args.steps
    clocks = sampler.stop(wall0, wall1)
    work = dict(zip(bz.STAT_NAMES, [int(x) // args.steps for x in stats.tolist()]))
    launches = lens.last_kernel_launches()             # emit, pass, count, scan, scatter, trace
    host_bins = torch.zeros((target.bins_v, target.bins_u), dtype=torch.int32).pin_memory().numpy().view("uint32")
    lens.illuminate(emitter, target, spec["max_crossings"], bins=host_bins)              # warm-up: the call's device scratch is allocated on first use
    t0 = time.perf_counter()
    for _ in range(3):
        host_bins[:] = 0
        lens.illuminate(emitter, target, spec["max_crossings"], bins=host_bins)          # C ABI: bz_illuminate_host
    e2e_s = (time.perf_counter() - t0) / 3
    rays = emitter.rays
    print(json.dumps({
        "metric": "Mrays/s", "value": rays / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32 (f64 in the emitter's acos/sin/cos, as the reference)", "data": "synthetic",
        "config": {"workload": name, "patches": lens.patch_count, "emitted_rays_per_step": rays, "belts": emitter.belts, "bins": "%dx%d" % (target.bins_u, target.bins_v),
                   "max_crossings": spec["max_crossings"], "l2": "flushed between steps (256 MiB fill)"},
        "e2e": {"value": rays / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": int(host_bins.nbytes + 140), "d2h_bytes_per_step": int(host_bins.nbytes),
                "ms_per_step": e2e_s * 1e3, "call": "bz_illuminate_host"},
        "gpu_launches": launches * args.steps, "clocks": clocks, "work_per_step": work, "landed": int(bins.sum().item()),
    }))


def main():
    args = parse_args()
    from bezier_b200 import workloads
    name = args.workload or workloads.HEADLINE
    if name in workloads.ILLUMINATION:
        main_illumination(args, name)
        return
    if args.impl == "reference":
        args.cpu_seconds = min(args.cpu_seconds, 150.0 / max(args.steps + args.warmup, 1))
        main_reference(args, name)
    else:
        main_ours(args, name)


if __name__ == "__main__":
    main()
