"""Golden vectors (tests/golden/*.npz, made by tests/golden/make_golden.py from the compiled reference) against
  - the plain-C oracle (CPU)                         -> pins the oracle where the reference is not mounted
  - the product's host preprocessing (CPU)           -> the bz_patch tables
  - the CUDA path through the C ABI (`-m gpu`)       -> BezierMesh::intersect and two BezierLens::refract crossings
All comparisons are bit-exact (NaN matches NaN)."""
import os

import numpy as np
import pytest

import common

GOLDEN = os.path.join(common.ROOT, "tests", "golden")
CASES = ["sphere16x8", "robot", "ellipsoid15x5"]
HIT_KEYS = ("what", "patch", "distance", "point", "normal", "cos", "bary")


def load(name):
    return np.load(os.path.join(GOLDEN, "golden_%s.npz" % name))


def oracle():
    import bzoracle
    bzoracle.build()
    return bzoracle


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_golden(name):
    g = load(name)
    o = oracle()
    hit = o.intersect(g["patches"], g["i_start"], g["i_dir"], threads=4)
    for k in HIT_KEYS:
        assert common.same_bits(hit[k], g["i_" + k]), k
    ri = float(g["ri"])
    s1, d1, st1, _, _ = o.refract(g["patches"], ri, g["r_start0"], g["r_dir0"], g["r_expected0"], threads=4)
    assert np.array_equal(st1, g["r_status1"]) and common.same_bits(s1, g["r_start1"]) and common.same_bits(d1, g["r_dir1"])
    s2, d2, st2, _, _ = o.refract(g["patches"], ri, g["r_start1"], g["r_dir1"], g["r_expected1"], threads=4)
    assert np.array_equal(st2, g["r_status2"]) and common.same_bits(s2, g["r_start2"]) and common.same_bits(d2, g["r_dir2"])
    assert int((st1 == 1).sum()) > 300 and int((st2 == 2).sum()) > 100


def test_survey_anchor_rays():
    """SURVEY.md section 4 sanity anchors of the patched reference on makeUnitSphere(16, 8), n = 1.3"""
    g = load("sphere16x8")
    o = oracle()
    s = np.array([[-10, 0.1, 0.2], [-10, 0.45, -0.35], [-10, -0.3, 0.6]], np.float32)
    d = np.tile(np.array([[1, 0, 0]], np.float32), (3, 1))
    s1, d1, st1, p1, _ = o.refract(g["patches"], 1.3, s, d, 1)
    assert list(st1) == [1, 1, 1] and list(p1) == [350, 455, 160]
    s2, d2, st2, p2, _ = o.refract(g["patches"], 1.3, s1, d1, 2)
    assert list(st2) == [2, 2, 2] and list(p2) == [303, 393, 215]
    assert np.allclose(d2[0], [0.99463236, -0.04163020, -0.09472825], rtol=0, atol=2e-6)
    assert np.allclose(s1[0], [-0.97332907, 0.10000009, 0.20000002], rtol=0, atol=2e-6)


@pytest.mark.parametrize("name", CASES)
def test_host_tables_match_golden(name):
    import bezier_b200 as bz
    g = load(name)
    m = bz.Mesh()
    if name == "sphere16x8":
        m.make_unit_sphere(16, 8)
    elif name == "robot":
        m.read_stl(common.ROBOT_STL)
    else:
        m.make_ellipsoid(15, 5, (1.0, 4.0, 2.0))
    m.standardize_vertices().standardize_normals()
    assert np.array_equal(m.triangles().view(np.uint32), g["triangles"].view(np.uint32))
    assert np.array_equal(bz.BezierMesh(m).patches().view(np.uint32), g["patches"].view(np.uint32))


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("mode", ["culled", "brute"])
def test_cuda_path_matches_golden(name, mode):
    import bezier_b200 as bz
    assert bz.device_count() > 0, "no CUDA device: the product has no CPU path"
    g = load(name)
    lens = bz.BezierLens(float(g["ri"]), g["patches"]).set_mode(bz.MODE_BRUTE if mode == "brute" else bz.MODE_CULLED)
    hit = lens.intersect(g["i_start"], g["i_dir"])
    for k in HIT_KEYS:
        assert common.same_bits(hit[k], g["i_" + k]), k
    s1, d1, st1, _ = lens.refract(g["r_start0"], g["r_dir0"], g["r_expected0"])
    assert np.array_equal(st1, g["r_status1"]) and common.same_bits(s1, g["r_start1"]) and common.same_bits(d1, g["r_dir1"])
    s2, d2, st2, _ = lens.refract(g["r_start1"], g["r_dir1"], g["r_expected1"])
    assert np.array_equal(st2, g["r_status2"]) and common.same_bits(s2, g["r_start2"]) and common.same_bits(d2, g["r_dir2"])


def test_reference_as_written_is_broken_on_this_path(oracle_built):
    """compiled exactly as written, the reference reads an uninitialised mPoint in the root search
    (3dGeomUtil.h:284-290 vs bezierTriangle.cpp:158-159) and loses almost every ray; the hoisted build is the oracle."""
    import bzref
    if not bzref.available(as_written=True):
        pytest.skip("libbzref_aswritten.so not built")
    ys, zs = np.meshgrid(np.linspace(-1.2, 1.2, 80, dtype=np.float32), np.linspace(-1.2, 1.2, 60, dtype=np.float32))
    s = np.stack([np.full(ys.size, -10.0, np.float32), ys.ravel(), zs.ravel()], axis=1)
    d = np.tile(np.array([[1, 0, 0]], np.float32), (s.shape[0], 1))
    entered = {}
    for as_written in (False, True):
        m = bzref.RefMesh(as_written=as_written).make_unit_sphere(16, 8).standardize_vertices().standardize_normals()
        lens = bzref.RefLens(1.3, bzref.RefBezierMesh(m))
        entered[as_written] = int((lens.refract(s, d, 1, threads=4)[2] == 1).sum())
    assert entered[False] > 2000                      # ~53 % of the grid enters the sphere
    assert entered[True] < entered[False] // 4        # as written: a small fraction (what the garbage happens to allow)
