#!/usr/bin/env python
"""tools/fast_report.py -- what BZ_MODE_FAST costs in parity and what it buys in speed.

For every BASELINE.json configuration: the frame rendered in the parity build (bit-exact with the reference, tests/) and in
the fast build on the same GPU -> pixels identical, histogram of |difference| in LSB; a random ray batch through
bz_intersect_batch and two bz_refract_batch crossings in both builds -> status / patch agreement, largest relative
error of points and directions; device time per frame in both builds.  Prints one JSON object per configuration.
Needs a CUDA device."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cuda-bezier-triangle-raytracer_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def rel(a, b, scale):
    return np.abs(a.astype(np.float64) - b.astype(np.float64)).max(axis=-1) / scale


def main():
    import torch
    import bezier_b200 as bz
    from bezier_b200 import workloads
    import common
    names = sys.argv[1:] or ["cfg1_sphere_640x480", "cfg2_ellipsoid64x32_1080p", "cfg3_robot_1080p", "cfg4_freeform181x91_4k_16spp"]
    for name in names:
        spec, mesh, cam_kw, scene_kw = workloads.setup(name, bz.Mesh)
        bezier = bz.BezierMesh(mesh)
        cam, scene = bz.make_camera(spec["width"], spec["height"], **cam_kw), bz.make_scene(**scene_kw)
        out = {"workload": name, "patches": bezier.size()}
        frames, times = {}, {}
        for label, mode in (("exact", bz.MODE_CULLED), ("fast", bz.MODE_FAST)):
            lens = bz.BezierLens(spec["ri"], bezier).set_mode(mode)
            rgba = torch.zeros((cam.height, cam.width, 4), dtype=torch.uint8, device="cuda")
            ms = []
            for _ in range(4):
                lens.render_device(cam, scene, rgba)
                ms.append(lens.last_kernel_ms())
            frames[label] = rgba.cpu().numpy()
            times[label] = min(ms[1:])
            lens.close()
        d = np.abs(frames["exact"][..., 0].astype(np.int32) - frames["fast"][..., 0].astype(np.int32))
        pixels = d.size
        rays = cam.width * cam.height * spec["strata"] ** 2
        out["render"] = {"pixels": int(pixels), "identical_pct": 100.0 * float((d == 0).sum()) / pixels, "within_1_lsb_pct": 100.0 * float((d <= 1).sum()) / pixels,
                         "differing": int((d != 0).sum()), "max_lsb": int(d.max()),
                         "one_sample_flip_lsb": (230 - 25) / spec["strata"] ** 2,
                         "kernel_ms_exact": times["exact"], "kernel_ms_fast": times["fast"],
                         "mrays_exact": rays / times["exact"] / 1e3, "mrays_fast": rays / times["fast"] / 1e3}
        # ray batches: random rays aimed into the bounding box, half of them starting inside it
        tri = mesh.triangles().reshape(-1, 3)
        lo, hi = tri.min(0), tri.max(0)
        size = float(np.linalg.norm(hi - lo))
        n = 400000
        s, dvec = common.random_rays(lo, hi, n, seed=5)
        exact = bz.BezierLens(spec["ri"], bezier).set_mode(bz.MODE_CULLED)
        fast = bz.BezierLens(spec["ri"], bezier).set_mode(bz.MODE_FAST)
        a, b = exact.intersect(s, dvec), fast.intersect(s, dvec)
        both = (a["what"] == 4) & (b["what"] == 4)
        same_patch = both & (a["patch"] == b["patch"])
        out["intersect"] = {"rays": n, "hits_exact": int((a["what"] == 4).sum()), "what_differs": int((a["what"] != b["what"]).sum()),
                            "patch_differs_among_common_hits": int((both & ~same_patch).sum()),
                            "max_rel_point": float(rel(a["point"][same_patch], b["point"][same_patch], size).max()),
                            "max_abs_normal": float(rel(a["normal"][same_patch], b["normal"][same_patch], 1.0).max()),
                            "max_rel_distance": float((np.abs(a["distance"][same_patch].astype(np.float64) - b["distance"][same_patch]) / size).max()),
                            "points_beyond_1e-5": int((rel(a["point"][same_patch], b["point"][same_patch], size) > 1e-5).sum()),
                            "normals_beyond_1e-5": int((rel(a["normal"][same_patch], b["normal"][same_patch], 1.0) > 1e-5).sum())}
        s2, d2 = common.random_rays(lo, hi, n, seed=9, inside_fraction=0.0)
        expected = np.full(n, 1, np.uint32)
        crossings = []
        for _ in range(2):
            es, ed, est, ep = exact.refract(s2, d2, expected)
            fs, fd, fst, fp = fast.refract(s2, d2, expected)
            ok = (est == fst) & (est != 0) & (ep == fp)
            crossings.append({"refracted_exact": int((est != 0).sum()), "status_differs": int((est != fst).sum()),
                              "patch_differs": int(((est == fst) & (est != 0) & (ep != fp)).sum()),
                              "max_rel_start": float(rel(es[ok], fs[ok], size).max()), "max_abs_dir": float(rel(ed[ok], fd[ok], 1.0).max()),
                              "dirs_beyond_1e-5": int((rel(ed[ok], fd[ok], 1.0) > 1e-5).sum())})
            s2, d2 = es, ed                                    # both builds continue from the exact build's rays
            expected = np.where(est == 0, 0, 3 - expected).astype(np.uint32)
        out["refract"] = crossings
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
