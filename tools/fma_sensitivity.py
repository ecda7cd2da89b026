#!/usr/bin/env python
"""tools/fma_sensitivity.py -- how much does the REFERENCE ALGORITHM itself move when only the rounding changes?

Builds the plain-C oracle (oracle/bz_oracle.c, bit-identical to the compiled reference, tests/test_golden.py) a second time with
FMA contraction allowed (-mfma -ffp-contract=fast: what `-march=haswell` would do to the reference's own build; division and
square root stay IEEE) and compares BezierMesh::intersect of the two builds on the same random rays.  CPU only.
The point: the four fixed iterations of the root search amplify rounding differences, so a share of the hits moves by
more than north_star's 1e-5 whatever the implementation -- BZ_MODE_FAST differs from the reference as much as this build does."""
import importlib.util
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests"), os.path.join(ROOT, "cuda-bezier-triangle-raytracer_b200")]
import bzoracle      # noqa: E402
import common        # noqa: E402


def main():
    out = os.path.join(ROOT, "scratch")
    os.makedirs(out, exist_ok=True)
    lib = os.path.join(out, "libbzoracle_fma.so")
    subprocess.check_call(["gcc", "-O2", "-std=c11", "-fPIC", "-mfma", "-ffp-contract=fast", "-pthread", "-I" + os.path.join(ROOT, "include"), "-shared",
                           os.path.join(ROOT, "oracle", "bz_oracle.c"), os.path.join(ROOT, "oracle", "bz_oracle_illum.c"), "-o", lib, "-lm"])
    spec = importlib.util.spec_from_file_location("bzoracle_fma", os.path.join(ROOT, "oracle", "bzoracle.py"))
    fma = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(fma)
    fma.LIB = lib
    threads = len(os.sched_getaffinity(0))
    print("%-16s %8s %8s | %9s %9s | %12s %12s %12s | %10s %10s %10s" % ("mesh", "rays", "hits", "what diff", "patch dif", "normal>1e-5", "normal>1e-4", "point>1e-5", "median dn", "p90 dn", "p99 dn"))
    for name, rays in (("sphere32x16", 100000), ("ellipsoid64x32", 100000), ("robot", 100000), ("ellipsoid181x91", 30000)):
        mesh, bezier, patches = common.ref_config(name)
        lo, hi = common.bbox(mesh)
        size = float(np.linalg.norm(hi - lo))
        s, d = common.random_rays(lo, hi, rays, seed=5)
        a = bzoracle.intersect(patches, s, d, threads=threads)
        b = fma.intersect(patches, s, d, threads=threads)
        both = (a["what"] == 4) & (b["what"] == 4) & (a["patch"] == b["patch"])
        dn = np.abs(a["normal"][both] - b["normal"][both]).max(1)
        dp = np.abs(a["point"][both] - b["point"][both]).max(1) / size
        dn = dn[np.isfinite(dn)]
        dp = dp[np.isfinite(dp)]
        hits = int((a["what"] == 4).sum())
        print("%-16s %8d %8d | %9d %9d | %6d %4.1f%% %6d %4.1f%% %6d %4.1f%% | %10.2e %10.2e %10.2e"
              % (name, rays, hits, int((a["what"] != b["what"]).sum()), int(((a["what"] == 4) & (b["what"] == 4) & (a["patch"] != b["patch"])).sum()),
                 int((dn > 1e-5).sum()), 100.0 * (dn > 1e-5).mean(), int((dn > 1e-4).sum()), 100.0 * (dn > 1e-4).mean(), int((dp > 1e-5).sum()), 100.0 * (dp > 1e-5).mean(),
                 np.median(dn), np.percentile(dn, 90), np.percentile(dn, 99)))


if __name__ == "__main__":
    main()
