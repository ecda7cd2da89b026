"""CPU checks of the product's device arithmetic (csrc/bz_math.cuh compiled for the host by tests/host_shim.cpp)
and of the cull hierarchy (host/bz_bvh.cpp) against the oracle: bit-exact, brute and culled."""
import numpy as np
import pytest

import common


@pytest.mark.parametrize("name", ["sphere16x8", "ellipsoid64x32", "robot"])
def test_device_math_and_cull_match_oracle(oracle_built, shim, name):
    import bzoracle
    mesh, bezier, patches = common.ref_config(name)
    lo, hi = common.bbox(mesh)
    s, d = common.random_rays(lo, hi, 3000, seed=5)
    ref = bzoracle.intersect(patches, s, d, threads=8)
    brute, work_brute = common.shim_intersect(shim, patches, s, d, 0)
    culled, work_culled = common.shim_intersect(shim, patches, s, d, 1)
    assert int((ref["what"] == 4).sum()) > 300
    common.assert_same_hits(ref, brute, name + "/brute")
    common.assert_same_hits(ref, culled, name + "/culled")
    assert work_culled[0] < work_brute[0] / 4      # the hierarchy really culls


def test_far_origin_rays_stay_exact(oracle_built, shim):
    """ray origins far from the lens: the per-ray slack of the box test must keep the cull conservative"""
    import bzoracle
    mesh, bezier, patches = common.ref_config("ellipsoid64x32")
    rng = np.random.default_rng(9)
    n = 3000
    tgt = (rng.random((n, 3)) - 0.5) * np.array([2.0, 8.0, 4.0])
    o = rng.normal(size=(n, 3))
    o = o / np.linalg.norm(o, axis=1, keepdims=True) * 2000.0
    d = (tgt - o)
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    o = o.astype(np.float32)
    ref = bzoracle.intersect(patches, o, d, threads=8)
    culled, _ = common.shim_intersect(shim, patches, o, d, 1)
    common.assert_same_hits(ref, culled, "far")


def test_wide_patches_stay_exact(oracle_built, shim):
    """ill-conditioned patches (kappa > 16) sit in the hierarchy with grown boxes: rays aimed at them, from
    origins inside and outside the origin limit, must still see exactly the reference's candidates"""
    import ctypes as C
    import bzoracle
    mesh, bezier, patches = common.ref_config("ellipsoid181x91")
    wide = np.zeros(4096, np.uint32)
    limit = C.c_float(0.0)
    shim.shim_bvh_wide.restype = C.c_uint32
    count = shim.shim_bvh_wide(patches.ctypes.data_as(C.c_void_p), C.c_uint32(patches.shape[0]), wide.ctypes.data_as(C.c_void_p),
                               C.c_uint32(wide.size), C.byref(limit))
    assert 0 < count <= wide.size and limit.value > 0
    wide = wide[:count]
    rng = np.random.default_rng(21)
    n = 4000
    pick = wide[rng.integers(0, count, n)]
    cp = patches[pick][:, 20:50].reshape(n, 10, 3)
    w = rng.dirichlet((1.0, 1.0, 1.0), n).astype(np.float32) * 1.4 - 0.133     # a little beyond the flat triangle too
    target = (cp[:, 0] * w[:, :1] + cp[:, 1] * w[:, 1:2] + cp[:, 2] * w[:, 2:3]).astype(np.float64)
    o = rng.normal(size=(n, 3))
    o /= np.linalg.norm(o, axis=1, keepdims=True)
    radius = np.where(np.arange(n) % 2 == 0, 12.0, limit.value * 1.5)          # inside / outside the origin limit
    o = target + o * radius[:, None]
    d = target - o
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    o = o.astype(np.float32)
    ref = bzoracle.intersect(patches, o, d, threads=8)
    culled, _ = common.shim_intersect(shim, patches, o, d, 1)
    assert int((ref["what"] == 4).sum()) > 500
    assert int(np.isin(ref["patch"], wide).sum()) > 50
    common.assert_same_hits(ref, culled, "wide")
