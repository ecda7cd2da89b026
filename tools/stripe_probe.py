#!/usr/bin/env python
"""tools/stripe_probe.py -- device time of ONE rank's share of the cfg4 frame (stripes of rank r of R, as bench.py deals them) on one GPU,
for both pipelines and a list of chunk sizes: how the pipeline choice behaves at the per-rank sizes of the sharded render (tuning aid).
usage: stripe_probe.py R [chunk_log2 ...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cuda-bezier-triangle-raytracer_b200"))
import torch
import bezier_b200 as bz
from bezier_b200 import workloads

ranks = int(sys.argv[1])
spec, mesh, cam_kw, scene_kw = workloads.setup(workloads.HEADLINE, bz.Mesh)
bezier = bz.BezierMesh(mesh)
cam, scene = bz.make_camera(spec["width"], spec["height"], **cam_kw), bz.make_scene(**scene_kw)
stripe = workloads.STRIPE_ROWS
band_rows = max(bz.striped_rows(cam.height, stripe, r, ranks) for r in range(ranks))
band = torch.zeros((band_rows, cam.width, 4), dtype=torch.uint8, device="cuda")
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
for arg in ["persistent", "auto"] + sys.argv[2:]:
    lens = bz.BezierLens(spec["ri"], bezier).set_mode(bz.MODE_CULLED)
    if arg == "persistent":
        lens.set_pipeline(bz.PIPELINE_PERSISTENT)
    elif arg == "auto":
        lens.set_pipeline(bz.PIPELINE_AUTO)
    else:
        lens.set_pipeline(bz.PIPELINE_WAVEFRONT).set_wave_chunk(1 << int(arg))
    worst = []
    for rank in (range(ranks) if os.environ.get("BZ_PROBE_ALL") else (0, ranks // 2, ranks - 1)):
        ms = []
        for _ in range(5):
            flush.fill_(1)
            lens.render_striped_device(cam, scene, stripe, rank, ranks, band)
            ms.append(lens.last_kernel_ms())
        worst.append(min(ms[1:]))
    print("R=%d" % ranks, arg, "kernel ms of ranks %s:" % ("0 .. R-1" if os.environ.get("BZ_PROBE_ALL") else "0 / R/2 / R-1"), " ".join("%.3f" % m for m in worst), "launches", lens.last_kernel_launches(), flush=True)
    lens.close()
