"""CPU tests of the product's host preprocessing (bz::Mesh / bz::BezierMesh behind the C ABI) against the compiled
reference (oracle/_ref): every table the hot path reads must be BIT-identical, so that any later mismatch is
attributable to the kernels.  No GPU needed: these entry points never touch CUDA."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import common

ROOT = common.ROOT


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def build_both(kind, *args):
    import bezier_b200 as bz
    import bzref
    out = []
    for cls in (bzref.RefMesh, bz.Mesh):
        m = cls()
        getattr(m, kind)(*args)
        m.standardize_vertices()
        snapped = m.triangles().copy()
        m.standardize_normals()
        out.append((m, snapped))
    return out


CASES = [("make_unit_sphere", (16, 8)), ("make_unit_sphere", (32, 16)), ("make_ellipsoid", (64, 32, (1.0, 4.0, 2.0))),
         ("make_ellipsoid", (7, 3, (1.0, 4.0, 2.0))), ("make_ellipsoid", (15, 5, (1.0, 4.0, 2.0))), ("read_stl", (common.ROBOT_STL,))]


@pytest.mark.parametrize("kind,args", CASES)
def test_tables_bit_identical_to_reference(oracle_built, kind, args):
    import bezier_b200 as bz
    import bzref
    (ref, ref_snapped), (got, got_snapped) = build_both(kind, *args)
    assert np.array_equal(bits(ref_snapped), bits(got_snapped)), "standardizeVertices"
    assert np.array_equal(bits(ref.triangles()), bits(got.triangles())), "winding after standardizeNormals"
    for a, b in zip(ref.neighbours(), got.neighbours()):
        assert np.array_equal(a, b), "getFace2neighbours"
    assert np.array_equal(bits(ref.corner_normals()), bits(got.corner_normals())), "getVertex2averageNormals"
    want, have = bzref.RefBezierMesh(ref).patches(), bz.BezierMesh(got).patches()
    assert want.shape == have.shape and want.shape[0] == 3 * ref.size()
    assert np.array_equal(bits(want), bits(have)), "bz_patch table differs in %d words" % int((bits(want) != bits(have)).sum())


def test_large_freeform_lens_tables(oracle_built):
    """cfg 4/5 mesh: 98 826 patches"""
    import bezier_b200 as bz
    import bzref
    (ref, _), (got, _) = build_both("make_ellipsoid", 181, 91, (1.0, 4.0, 2.0))
    want, have = bzref.RefBezierMesh(ref).patches(), bz.BezierMesh(got).patches()
    assert want.shape == (98826, 64)
    assert np.array_equal(bits(want), bits(have))


def test_transforms_splits_tessellation(oracle_built):
    import bezier_b200 as bz
    import bzref
    meshes = []
    for cls in (bzref.RefMesh, bz.Mesh):
        m = cls().make_ellipsoid(9, 4, (1.0, 2.0, 0.5))
        m.translate(10.0, 0.25, -3.0)
        m.scale(1.5)
        m.transform([[0.0, -1.0, 0.0], [1.0, 0.0, 0.0], [0.0, 0.0, 1.0]], [0.5, 0.0, 0.0])
        m.standardize_vertices().standardize_normals()
        meshes.append(m)
    ref, got = meshes
    assert np.array_equal(bits(ref.triangles()), bits(got.triangles()))
    rb, gb = bzref.RefBezierMesh(ref), bz.BezierMesh(got)
    assert np.array_equal(bits(rb.patches()), bits(gb.patches()))
    assert np.array_equal(bits(rb.tessellate(3).triangles()), bits(gb.interpolate(3).triangles()))      # BezierMesh::interpolate(divisor)
    rs, gs = rb.split_thick(), gb.split_thick()                                                          # splitThickBezierTriangles
    assert rs.size() == gs.size() and rs.size() > ref.size()
    assert np.array_equal(bits(rs.triangles()), bits(gs.triangles()))
    # the refined mesh goes through the same pipeline again (test.cpp:435-440)
    for m in (rs, gs):
        m.standardize_vertices().standardize_normals()
    assert np.array_equal(bits(bzref.RefBezierMesh(rs).patches()), bits(bz.BezierMesh(gs).patches()))


def test_stl_round_trip_and_ascii(oracle_built, tmp_path):
    import bezier_b200 as bz
    import bzref
    path = str(tmp_path / "sphere.stl")
    bz.Mesh().make_unit_sphere(8, 4).write_stl(path)
    ref_path = str(tmp_path / "sphere_ref.stl")
    bzref.RefMesh().make_unit_sphere(8, 4).write_stl(ref_path)
    assert open(path, "rb").read() == open(ref_path, "rb").read()                     # writeMesh text is identical
    assert np.array_equal(bits(bz.Mesh().read_stl(path).triangles()), bits(bzref.RefMesh().read_stl(path).triangles()))   # ASCII reader
    assert bz.Mesh().read_stl(common.ROBOT_STL).size() == 150                         # binary reader, SURVEY.md section 2 row 14


def test_error_behaviour(oracle_built, tmp_path):
    import bezier_b200 as bz
    import bzref
    # the reference throws "Vertex on edge detected." for makeEllipsoid(182, 91)
    with pytest.raises(RuntimeError, match="Vertex on edge detected"):
        bzref.RefMesh().make_ellipsoid(182, 91, (1.0, 4.0, 2.0)).standardize_vertices().standardize_normals()
    with pytest.raises(bz.BezierError, match="Vertex on edge detected") as e:
        bz.Mesh().make_ellipsoid(182, 91, (1.0, 4.0, 2.0)).standardize_vertices().standardize_normals()
    assert e.value.code == bz.ERR_PREPROCESS
    # BezierMesh of a mesh that skipped standardizeNormals: std::out_of_range in the reference (bezierMesh.cpp:17-18)
    with pytest.raises(bz.BezierError) as e:
        bz.BezierMesh(bz.Mesh().make_unit_sphere(8, 4))
    assert e.value.code == bz.ERR_PREPROCESS
    with pytest.raises(bz.BezierError) as e:
        bz.Mesh().read_stl(str(tmp_path / "missing.stl"))
    assert e.value.code == bz.ERR_PREPROCESS


def test_hash_set_order_replay_matches_libstdcxx(tmp_path):
    """bz::hashSetIterationOrder must replay std::unordered_set<uint32_t> exactly (seed-face tie break of standardizeNormals)"""
    src = tmp_path / "order.cpp"
    src.write_text(r'''
#include <cstdio>
#include <cstdlib>
#include <random>
#include <unordered_set>
#include <vector>
#include "bz_mesh.h"
int main() {
  std::mt19937 rng(12345);
  for(int trial = 0; trial < 400; ++trial) {
    size_t n = 1 + rng() % (trial < 300 ? 40 : 700);
    uint32_t range = (trial % 3 == 0) ? 64u : 200000u;
    std::vector<uint32_t> in;
    std::unordered_set<uint32_t> set;
    for(size_t i = 0; i < n; ++i) { uint32_t v = rng() % range; in.push_back(v); set.insert(v); }
    std::vector<uint32_t> want(set.begin(), set.end());
    if(want != bz::hashSetIterationOrder(in)) { std::printf("mismatch in trial %d (n = %zu)\n", trial, n); return 1; }
  }
  std::printf("ok\n");
  return 0;
}
''')
    exe = str(tmp_path / "order")
    host = os.path.join(ROOT, "cuda-bezier-triangle-raytracer_b200", "host")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I" + host, "-I" + os.path.join(ROOT, "include"), str(src), os.path.join(host, "bz_mesh.cpp"), "-o", exe])
    assert subprocess.check_output([exe]).decode().strip() == "ok"


def test_c_abi_exports_every_declared_symbol():
    """the shared library loads without a GPU and exports exactly what include/bz_capi.h declares"""
    import bezier_b200 as bz
    header = open(os.path.join(ROOT, "include", "bz_capi.h")).read()
    declared = set(re.findall(r"\b(bz_[a-z0-9_]+)\s*\(", header))
    lib = bz.load()
    for name in sorted(declared):
        assert hasattr(lib, name), name
    assert declared == set(bz.SIGNATURES), declared ^ set(bz.SIGNATURES)
    assert lib.bz_sizeof_patch() == 256 and lib.bz_sizeof_stats() == 96
    assert lib.bz_sizeof_camera() == C.sizeof(bz.Camera) and lib.bz_sizeof_scene() == C.sizeof(bz.Scene)
    assert bz.striped_rows(2160, 16, 0, 8) + bz.striped_rows(2160, 16, 7, 8) == 272 + 256


def test_no_cpu_fallback_without_device():
    """computing entry points fail loudly when there is no CUDA device"""
    import bezier_b200 as bz
    if bz.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(bz.BezierError) as e:
        bz.BezierLens(1.3, np.zeros((0, 64), np.float32))
    assert e.value.code == bz.ERR_NO_DEVICE


def test_lens_create_rejects_bad_neighbour_indices():
    """the follow hop indexes the device tables with the neighbour indices (bezierMesh.cpp:213-217): a record whose neighbour lies
    outside the table is refused before anything is uploaded (checked without a device)"""
    import bezier_b200 as bz
    mesh = bz.Mesh().make_unit_sphere(8, 4)
    mesh.standardize_vertices().standardize_normals()
    table = bz.BezierMesh(mesh).patches().copy()
    table.view(np.uint32)[5, 17] = table.shape[0]            # words 16..18 hold the neighbours
    with pytest.raises(bz.BezierError) as e:
        bz.BezierLens(1.3, table)
    assert e.value.code == bz.ERR_INVALID and "neighbour" in str(e.value)


def test_comm_argument_checks():
    import ctypes as C
    import bezier_b200 as bz
    lib = bz.load()
    handle = C.c_void_p()
    ident = (C.c_uint8 * 128)()
    assert lib.bz_comm_create(ident, 3, 2, 0, C.byref(handle)) == bz.ERR_INVALID      # rank >= ranks
    assert lib.bz_comm_create(None, 0, 1, 0, C.byref(handle)) == bz.ERR_INVALID
    assert lib.bz_gather_rows(None, None, 4, 4, 8, 0, None, None) == bz.ERR_INVALID
    assert lib.bz_comm_set_gather(None, 0) == bz.ERR_INVALID
    assert lib.bz_comm_gather_mode(None, None) == bz.ERR_INVALID
    assert lib.bz_comm_destroy(None) == bz.OK
