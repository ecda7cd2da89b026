#!/usr/bin/env python
"""bench.py -- Mrays/s of the 4K free-form lens render (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl ours|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One step = one whole frame of the workload (default: cfg4, makeEllipsoid(181,91,(1,4,2)) -> 98 826 patches,
3840x2160, 16 rays/pixel = 132 710 400 primary rays), image rows dealt to the N ranks in round-robin stripes
of 8 rows, tables replicated, the rendered stripes gathered on rank 0 by bz_gather_rows (C ABI: every rank stores its
rows into the root's IPC-mapped frame over NVLink; --gather nccl: send / recv + assembly) when N > 1 (strong scaling: the frame is fixed).
`value` = primary rays of the frame / device time per step (CUDA events around every step, the L2 flush between
steps outside them), max over ranks.  `e2e` goes through the host-buffer C-ABI call (bz_render_host, N = 1) or
render + gather + device->pinned-host copy of the frame (N > 1, the way home of frame k overlapping the tracing of
frame k + 1), every step; `e2e_moving_camera` is the same with a camera that moves every step (the frame plan is
rebuilt on the host each time); `frame_crc32` is the CRC-32 of the frame the e2e loop left in host memory -- it must
be the same at every N.  `fast` times BZ_MODE_FAST beside the parity build and counts the pixels that differ.
`--impl reference` times the reference's own CPU code (oracle/_ref/libbzref.so: reference sources + the one
documented hoist) on the host cores.
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
    p.add_argument("--gather", default="peer", choices=["peer", "nccl"], help="transport of bz_gather_rows at N > 1 (bz_comm_set_gather)")
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
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(device), "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits", "-lms", "50"],
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
        rows = [r for t, r in self.rows if t0 <= t <= t1 + 0.1 and len(r) >= 7]
        window = "timed region"
        if not rows:                                            # a timed region shorter than the sampling period: the samples since the warm-up
            rows, window = [r for _, r in self.rows if len(r) >= 7], "warm-up + timed region"

        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[0]) for r in rows)
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "power_w_max": max(float(r[2]) for r in rows),
                "samples": len(rows), "window": window, "reasons": reasons}


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
    bezier = bzref.RefBezierMesh(mesh)
    lens = bzref.RefLens(spec["ri"], bezier)
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
    return dict(value=rays / dt / 1e6, rays=int(rays), seconds=dt, step=step, calls=int(stats.refract_calls), patches=int(bezier.size()),
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
        "config": {"workload": name, "patches": timed[-1]["patches"], "image": "%dx%d" % (spec["width"], spec["height"]), "spp": spec["strata"] ** 2,
                   "note": "each step is a bounded pixel-lattice sample of the frame, rate = sampled primary rays / wall time"},
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": threads, "kind": "reference", "sample": timed[-1]["sample"], "cpu": cpu_model()},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def main_ours(args, name):
    import zlib
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
        # NCCL may print its version banner on stdout when a communicator is created: keep stdout for the ONE JSON line
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
    try:
        if world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))      # bootstrap + the scalar reductions of this script only
            dist.barrier()
        comm = sharding.make_comm(bz, local) if world > 1 else None                     # the image gather itself: bz_gather_rows (C ABI)
        if comm is not None:
            comm.set_gather(bz.GATHER_NCCL if args.gather == "nccl" else bz.GATHER_PEER)
        torch.cuda.synchronize()
    finally:
        if world > 1:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)
    assert world == args.gpus or world == 1, "--gpus must match the torchrun world size"

    # ---- setup (untimed): host preprocessing in the product's own C++ (bz::Mesh / bz::BezierMesh), tables to HBM ----
    t_setup = time.perf_counter()
    spec, mesh, cam_kw, scene_kw = workloads.setup(name, bz.Mesh)
    bezier = bz.BezierMesh(mesh)
    lens = bz.BezierLens(spec["ri"], bezier, device=local)
    exact_mode, fast_mode = bz.MODE_CULLED, bz.MODE_FAST
    lens.set_mode(fast_mode if args.mode == "fast" else exact_mode)
    lens.set_pipeline({"auto": bz.PIPELINE_AUTO, "persistent": bz.PIPELINE_PERSISTENT, "wavefront": bz.PIPELINE_WAVEFRONT}[args.pipeline])
    setup_s = time.perf_counter() - t_setup
    cam, scene = bz.make_camera(spec["width"], spec["height"], **cam_kw), bz.make_scene(**scene_kw)
    width, height, spp = spec["width"], spec["height"], spec["strata"] ** 2
    stripe = workloads.STRIPE_ROWS
    rows_local = bz.striped_rows(height, stripe, rank, world)
    band_rows = max(bz.striped_rows(height, stripe, r, world) for r in range(world))
    frame_rays = width * height * spp

    dev = torch.device("cuda", local)
    bands = [torch.zeros((band_rows, width, 4), dtype=torch.uint8, device=dev) for _ in range(2)]       # this rank's stripes, double-buffered for the pipelined e2e loop
    images = [torch.zeros((height, width, 4), dtype=torch.uint8, device=dev) for _ in range(2)] if (world > 1 and rank == 0) else [None, None]
    stats = torch.zeros(len(bz.STAT_NAMES), dtype=torch.int64, device=dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)               # > 126 MB L2
    hosts = [torch.empty((height, width, 4), dtype=torch.uint8).pin_memory() for _ in range(2)] if rank == 0 else [None, None]
    kernel_ms = []

    rendered_events, split = [], {}

    def render_and_gather(buf, record):
        lens.render_striped_device(cam, scene, stripe, rank, world, bands[buf], stats=stats)
        if record:
            kernel_ms.append(lens.last_kernel_ms())                                     # events on the launching stream around the trace kernels
            rendered_events.append(torch.cuda.Event(enable_timing=True))
            rendered_events[-1].record()
        if world > 1:
            comm.gather_rows(bands[buf], width, height, stripe, rgba=images[buf])        # every rank's rows to rank 0 over NVLink (peer stores, or NCCL send / recv + assembly)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_steps(steps, record):
        """K steps between two barriers; every step is bracketed by its own event pair AFTER the L2 flush, so that the flush
        (256 MiB fill, stated as flush_ms) is between the timed iterations and not inside them -> (ms per step, wall0, wall1)"""
        pairs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        barrier()
        wall0 = time.time()
        for begin, end in pairs:
            flush.fill_(1)
            begin.record()
            render_and_gather(0, record)
            end.record()
        barrier()
        wall1 = time.time()
        total = torch.tensor([sum(b.elapsed_time(e) for b, e in pairs)], dtype=torch.float64, device=dev)
        if record and len(rendered_events) == steps:
            split["render_ms"] = sum(b.elapsed_time(m) for (b, _), m in zip(pairs, rendered_events)) / steps      # this rank: memset + sky + trace + resolve
            split["gather_ms"] = sum(m.elapsed_time(e) for (_, e), m in zip(pairs, rendered_events)) / steps      # this rank: wait for the slowest sender + transfer + assembly
        if world > 1:
            dist.all_reduce(total, op=dist.ReduceOp.MAX)
        return float(total.item()) / steps, wall0, wall1

    sampler = ClockSampler(local) if rank == 0 else None                                # started before the warm-up: nvidia-smi needs a moment to come up
    for _ in range(args.warmup):
        flush.fill_(1)
        render_and_gather(0, False)
    barrier()
    stats.zero_()
    ms_per_step, wall0, wall1 = timed_steps(args.steps, True)
    clocks = sampler.stop(wall0, wall1) if sampler else None
    fb, fe = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    fb.record()
    for _ in range(4):
        flush.fill_(1)
    fe.record()
    torch.cuda.synchronize()
    flush_ms = fb.elapsed_time(fe) / 4

    # ---- work counters of the timed region, summed over ranks ----
    work = stats.clone()
    kern = torch.tensor([sum(kernel_ms) / max(len(kernel_ms), 1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(work, op=dist.ReduceOp.SUM)
        dist.all_reduce(kern, op=dist.ReduceOp.MAX)
    work = dict(zip(bz.STAT_NAMES, [int(x) // args.steps for x in work.tolist()]))      # per step (frame)
    own = dict(zip(bz.STAT_NAMES, [int(x) // args.steps for x in stats.tolist()]))      # this rank, per launch
    launches_per_step = lens.last_kernel_launches() + (1 if (world > 1 and rank == 0) else 0)       # (sky +) trace + resolve (+ the stripe assembly on rank 0)

    # ---- e2e: host buffers in the timed region, every step ----
    e2e_steps = max(2, min(args.steps, 8))
    side = torch.cuda.Stream(device=dev)
    moving = {"plan_ms": []}

    def e2e_loop(steps, move_camera):
        """N = 1: bz_render_host (C ABI) into pinned host memory, every step.  N > 1: render on the compute stream; gather to rank 0
        + assembly + device->pinned-host copy on a side stream, double-buffered, so frame k's way home overlaps frame k + 1's
        tracing; every frame is in host memory before the clock stops.  move_camera: the eye moves by one ulp per step, so every
        step plans its frame anew (bz_plan_tiles + upload, bz_last_plan_ms)."""
        rendered = [torch.cuda.Event() for _ in range(2)]
        home = [torch.cuda.Event() for _ in range(2)]
        eye0 = float(cam.eye[0])
        barrier()
        t0 = time.perf_counter()
        for k in range(steps):
            buf = k % 2
            if move_camera:
                cam.eye[0] = float(np.nextafter(np.float32(cam.eye[0]), np.float32(1e30)))
            if world == 1:
                lens.render_host_into(cam, scene, hosts[buf].data_ptr())
            else:
                if k >= 2:
                    torch.cuda.current_stream().wait_event(home[buf])                  # the band buffer is free again
                lens.render_striped_device(cam, scene, stripe, rank, world, bands[buf])
                rendered[buf].record()
                with torch.cuda.stream(side):
                    side.wait_event(rendered[buf])
                    comm.gather_rows(bands[buf], width, height, stripe, rgba=images[buf], stream=side.cuda_stream)
                    if rank == 0:
                        hosts[buf].copy_(images[buf], non_blocking=True)
                    home[buf].record()
            if move_camera:
                moving["plan_ms"].append(lens.last_plan_ms())
        barrier()
        seconds = (time.perf_counter() - t0) / steps
        cam.eye[0] = eye0
        t = torch.tensor([seconds], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), (steps - 1) % 2

    e2e_loop(max(2, min(args.warmup, 3)), False)                                       # warm-up: scratch of the calls is allocated on first use
    e2e_s, last_buf = e2e_loop(e2e_steps, False)
    crc = zlib.crc32(hosts[last_buf].numpy().tobytes()) if rank == 0 else 0             # the frame the e2e loop left in host memory
    e2e_moving_s, _ = e2e_loop(e2e_steps, True)
    plan_ms = float(np.median(moving["plan_ms"])) if moving["plan_ms"] else 0.0

    # ---- the fast build next to the parity build (N = 1): same frame, BZ_MODE_FAST ----
    fast = None
    if world == 1 and args.mode == "exact":
        exact_frame = torch.zeros((height, width, 4), dtype=torch.uint8, device=dev)
        fast_frame = torch.zeros_like(exact_frame)
        lens.render_device(cam, scene, exact_frame)
        lens.set_mode(fast_mode)
        fast_stats = torch.zeros(len(bz.STAT_NAMES), dtype=torch.int64, device=dev)
        fast_ms, fast_kernel = [], []
        for k in range(2 + min(args.steps, 8)):
            flush.fill_(1)
            b, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            b.record()
            lens.render_device(cam, scene, fast_frame, stats=fast_stats if k == 2 else None)
            e.record()
            torch.cuda.synchronize()
            if k >= 2:
                fast_ms.append(b.elapsed_time(e))
                fast_kernel.append(lens.last_kernel_ms())
        lens.set_mode(exact_mode)
        diff = (exact_frame[..., 0].to(torch.int16) - fast_frame[..., 0].to(torch.int16)).abs()
        fw = dict(zip(bz.STAT_NAMES, [int(x) for x in fast_stats.tolist()]))
        fast = {"mode": "BZ_MODE_FAST: FMA contraction, MUFU reciprocal / rsqrt, FP32 cos2; ray generation and shading unchanged",
                "value": frame_rays / (sum(fast_ms) / len(fast_ms) * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": sum(fast_ms) / len(fast_ms),
                "kernel_ms": sum(fast_kernel) / len(fast_kernel),
                "flop": FLOP_STAGE_A * fw["stage_a"] + FLOP_STAGE_B * fw["stage_b"] + FLOP_SNELL * fw["snell"],
                "pixels_identical_pct": 100.0 * float((diff == 0).sum().item()) / diff.numel(), "pixels_differing": int((diff != 0).sum().item()),
                "max_abs_diff": int(diff.max().item()), "one_sample_flip": (230 - 25) / spp,
                "note": "the path amplifies rounding: the reference itself rebuilt with FMA contraction (-march=haswell) differs from the reference as much (DESIGN.md); "
                        "north_star's 1e-5 tolerance is met only by the bit-exact parity build, which stays the headline"}

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
            for entry in (t if isinstance(t, list) else [t]):
                if entry.get("workload") == name and world == 1 and args.mode == "exact" and entry.get("pipeline") == ("wavefront" if launches_per_step > 4 else "persistent"):
                    traffic, traffic_source = entry["dram_bytes_read"] + entry["dram_bytes_write"], entry["source"]
        except (OSError, KeyError, ValueError):
            pass
        flop = FLOP_STAGE_A * own["stage_a"] + FLOP_STAGE_B * own["stage_b"] + FLOP_SNELL * own["snell"]     # rank 0's launch
        k_ms = float(kern.item())
        own_pixels = rows_local * width
        wavefront = launches_per_step > 4          # (sky + trace + resolve (+ assembly) otherwise)
        hbm_bytes = own_pixels * (BYTES_PER_PIXEL_OUT + BYTES_PER_PIXEL_ACCUM)
        if wavefront:
            # the lists of the wavefront pipeline: per evaluated candidate 8 B entry + 32 B ray read, 8 B key + 32 B hit record written;
            # per traced ray slot and crossing 36 B ray state + chain head written by the walk, 36 B read and 32 B rewritten by the resolve
            hbm_bytes += 80 * own["stage_b"] + 104 * (32 * own["packet_walks"] + own["lane_walk_calls"])
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        achieved_tf = flop / (k_ms * 1e-3) / 1e12
        sm_max = (clocks or {}).get("sm_max_mhz") or peaks.get("sm_max_mhz", 1965.0)
        nominal_tf = lens.info()["sm_count"] * 128 * 2 * sm_max * 1e6 / 1e12
        roofline = {
            "bound": "fp32_alu",
            "kernel": ("bz_wave_walk_kernel + bz_wave_eval_kernel + bz_wave_resolve_kernel (wavefront pipeline)" if wavefront else "bz_warp_trace_kernel<RENDER> (persistent kernel)")
                      + "; kernel_ms = the trace kernels + bz_sky_kernel of one step, CUDA events on the launching stream",
            "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved_tf / peak_tf,
            "peak_nominal": nominal_tf, "frac_of_nominal": achieved_tf / nominal_tf,
            "traffic": traffic, "traffic_unit": "DRAM bytes of the trace kernels of one step (ncu, summed over their launches)", "traffic_source": traffic_source,
            "frac_of_no_fma_ceiling": achieved_tf / (peak_tf / 2.0),
            "no_fma_note": "the parity build has no FMA contraction (bit-exact with the reference's x86 build): one flop per lane and instruction, half the probe's peak",
            "peak_source": "FMA probe kernel in this process (16 independent chains per thread, 2 flop/lane/instr); peak_nominal = SMs x 128 lanes x 2 x %.0f MHz" % sm_max,
            "flop_model": "33 x stage_a + 1750 x stage_b + 45 x snell, executed counts (not brute-force-equivalent)",
            "kernel_ms": k_ms, "stage_a": own["stage_a"], "stage_b": own["stage_b"], "snell": own["snell"], "node_visits": own["node_visits"],
            "hbm": {"algorithmic_bytes": hbm_bytes, "achieved_gbs": hbm_bytes / (k_ms * 1e-3) / 1e9, "peak_gbs": hbm_peak,
                    "frac": hbm_bytes / (k_ms * 1e-3) / 1e9 / hbm_peak, "peak_source": "MEASURED_PEAKS.json" if "hbm_gbs" in peaks else "fallback"},
        }
        if fast is not None:
            fast["roofline_frac"] = fast.pop("flop") / (fast["kernel_ms"] * 1e-3) / 1e12 / peak_tf
            fast["roofline_frac_of_nominal"] = fast["roofline_frac"] * peak_tf / nominal_tf
        line = {
            "metric": "Mrays/s", "value": frame_rays / (ms_per_step * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": name, "patches": lens.patch_count, "image": "%dx%d" % (width, height), "spp": spp, "primary_rays_per_step": frame_rays,
                       "refractive_index": spec["ri"], "max_crossings": spec["max_crossings"], "parallelism": "row stripes of %d x %d ranks" % (stripe, world),
                       "l2": "flushed between steps (256 MiB fill, outside the per-step event pairs; flush_ms says what it costs)",
                       "mode": "culled (exact candidate-set-preserving BVH)" + (", BZ_MODE_FAST arithmetic" if args.mode == "fast" else ""), "fmad": args.mode == "fast",
                       "pipeline": args.pipeline, "setup_s": setup_s},
            "step_split_rank0": split, "gather": (None if comm is None else ("peer stores into the root's IPC-mapped buffer, generation flags over NVLink (no NCCL call per frame)" if comm.gather_mode() == bz.GATHER_PEER else "ncclSend / ncclRecv + assembly pass")), "flush_ms": flush_ms, "wall_ms_per_step_with_flush": (wall1 - wall0) / args.steps * 1e3,
            "e2e": {"value": frame_rays / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": int(bz.load().bz_sizeof_camera() + bz.load().bz_sizeof_scene()),
                    "d2h_bytes_per_step": width * height * 4, "ms_per_step": e2e_s * 1e3,
                    "call": "bz_render_host into pinned host memory" if world == 1 else
                            "bz_render_striped on the compute stream; bz_gather_rows (to rank 0 over NVLink) + D2H to pinned host on a side stream, double-buffered"},
            "e2e_moving_camera": {"value": frame_rays / e2e_moving_s / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_moving_s * 1e3, "plan_ms": plan_ms,
                                  "note": "the eye moves by one ulp every step: every frame is planned anew (bz_plan_tiles on the host + upload of the tile list)"},
            "frame_crc32": "%08x" % crc,
            "gpu_launches": launches_per_step * args.steps,
            "clocks": clocks, "roofline": roofline, "work_per_step": work,
            "refract_calls_per_s": work["refract_calls"] / (ms_per_step * 1e-3),
        }
        if fast is not None:
            line["fast"] = fast
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
        comm.close()                       # every rank: bz_comm_destroy is collective (ncclCommDestroy waits for the other ranks)
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
