"""Generates tests/golden/*.npz from the REAL reference (oracle/_ref/libbzref.so = reference sources compiled from
/root/reference + the one documented hoist).  Run in the build container, where /root/reference is mounted:
    make -C oracle && python tests/golden/make_golden.py
The vectors pin (a) the bz_patch tables of BezierMesh(Mesh), (b) BezierMesh::intersect and (c) two successive
BezierLens::refract crossings on seeded rays, so that the plain-C oracle, the host preprocessing and the CUDA path can
be checked on machines without the reference (the GPU box); golden_hemisphere.npz pins (d) UniformHemisphere."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import bzref      # noqa: E402
import common     # noqa: E402

CASES = {"sphere16x8": 1.3, "robot": 1.5, "ellipsoid15x5": 1.3}


def main():
    assert bzref.available() and bzref.load().bzref_is_hoisted() == 1
    for name, ri in CASES.items():
        mesh, bezier, patches = common.ref_config(name)
        lo, hi = common.bbox(mesh)
        s, d = common.random_rays(lo, hi, 1500, seed=101)
        hit = bezier.intersect(s, d, threads=8)
        lens = bzref.RefLens(ri, bezier)
        s0, d0 = common.random_rays(lo, hi, 1500, seed=202, inside_fraction=0.0)
        e0 = np.full(s0.shape[0], 1, np.uint32)
        s1, d1, st1 = lens.refract(s0, d0, e0, threads=8)
        e1 = np.where(st1 == 0, 0, 2).astype(np.uint32)
        s2, d2, st2 = lens.refract(s1, d1, e1, threads=8)
        out = os.path.join(HERE, "golden_%s.npz" % name)
        np.savez_compressed(out, ri=np.float32(ri), triangles=mesh.triangles(), patches=patches,
                            i_start=s, i_dir=d, **{"i_" + k: v for k, v in hit.items()},
                            r_start0=s0, r_dir0=d0, r_expected0=e0, r_start1=s1, r_dir1=d1, r_status1=st1,
                            r_expected1=e1, r_start2=s2, r_dir2=d2, r_status2=st2)
        print(name, patches.shape, "hits", int((hit["what"] == 4).sum()), "entered", int((st1 == 1).sum()), "exited", int((st2 == 2).sum()),
              os.path.getsize(out), "bytes")


def hemisphere():
    """UniformHemisphere (reference/hostUtil.cpp:3-29): belt tables and getRandom() on the draws of the C ABI's generator"""
    seed, rays = 424242, 2048
    out = {"seed": np.uint64(seed), "rays": np.uint32(rays)}
    draw1, draw2 = bzref.abi_draws(seed, 0, rays)
    for belts in (1, 2, 7, 16, 64, 128):
        count, width, first = bzref.hemisphere(belts)
        direction, patch = bzref.hemisphere_random(belts, draw1, draw2)
        out.update({"count_%d" % belts: np.uint32(count), "width_%d" % belts: width, "first_%d" % belts: first,
                    "direction_%d" % belts: direction, "patch_%d" % belts: patch})
    path = os.path.join(HERE, "golden_hemisphere.npz")
    np.savez_compressed(path, **out)
    print("hemisphere", os.path.getsize(path), "bytes")


if __name__ == "__main__":
    hemisphere()
    main()
