#!/usr/bin/env python
"""tools/tune_wave.py -- device time of one cfg4 frame through the wavefront pipeline for a list of chunk sizes (tuning aid).
usage: tune_wave.py [mode exact|fast] chunk_log2 ...   (BZ_B200_LIB selects the library build)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cuda-bezier-triangle-raytracer_b200"))
import torch
import bezier_b200 as bz
from bezier_b200 import workloads

mode = sys.argv[1]
name = os.environ.get("BZ_TUNE_WORKLOAD", workloads.HEADLINE)
spec, mesh, cam_kw, scene_kw = workloads.setup(name, bz.Mesh)
bezier = bz.BezierMesh(mesh)
cam, scene = bz.make_camera(spec["width"], spec["height"], **cam_kw), bz.make_scene(**scene_kw)
rgba = torch.zeros((cam.height, cam.width, 4), dtype=torch.uint8, device="cuda")
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
for arg in sys.argv[2:]:
    lens = bz.BezierLens(spec["ri"], bezier).set_mode(bz.MODE_FAST if mode == "fast" else bz.MODE_CULLED)
    if arg == "persistent":
        lens.set_pipeline(bz.PIPELINE_PERSISTENT)
    else:
        lens.set_pipeline(bz.PIPELINE_WAVEFRONT).set_wave_chunk(1 << int(arg))
    ms = []
    for _ in range(5):
        flush.fill_(1)
        lens.render_device(cam, scene, rgba)
        ms.append(lens.last_kernel_ms())
    print(os.path.basename(os.environ.get("BZ_B200_LIB", "default")), mode, arg, "min %.3f ms" % min(ms[1:]), " ".join("%.2f" % m for m in ms), "launches", lens.last_kernel_launches(), flush=True)
    lens.close()
