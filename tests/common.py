"""Test helpers: configs of BASELINE.json built through the compiled reference (oracle/_ref), ray sets, comparisons."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ROBOT_STL = os.path.join(ROOT, "cuda-bezier-triangle-raytracer_b200", "data", "robot.stl")

_cache = {}


def ref_config(name):
    """(RefMesh, RefBezierMesh, patches[n,64]) of a named config, through the REAL reference (the checker)."""
    import bzref
    if name in _cache:
        return _cache[name]
    m = bzref.RefMesh()
    if name.startswith("sphere"):
        s, b = name[len("sphere"):].split("x")
        m.make_unit_sphere(int(s), int(b))
    elif name.startswith("ellipsoid"):
        s, b = name[len("ellipsoid"):].split("x")
        m.make_ellipsoid(int(s), int(b), (1.0, 4.0, 2.0))
    elif name == "robot":
        m.read_stl(ROBOT_STL)
    else:
        raise KeyError(name)
    m.standardize_vertices().standardize_normals()
    b = bzref.RefBezierMesh(m)
    _cache[name] = (m, b, b.patches())
    return _cache[name]


def bbox(mesh):
    tri = mesh.triangles().reshape(-1, 3)
    return tri.min(0), tri.max(0)


def random_rays(lo, hi, n, seed, inside_fraction=0.5):
    """unit-direction rays aimed into the bounding box; the last part starts inside the box"""
    rng = np.random.default_rng(seed)
    c = (lo + hi) / 2
    r = float(np.linalg.norm(hi - lo)) / 2
    o = rng.normal(size=(n, 3))
    o = o / np.linalg.norm(o, axis=1, keepdims=True) * r * 3 + c
    tgt = c + (rng.random((n, 3)) - 0.5) * (hi - lo)
    k = int(n * (1.0 - inside_fraction))
    o[k:] = tgt[k:] + (rng.random((n - k, 3)) - 0.5) * 1e-3
    tgt[k:] = c + (rng.random((n - k, 3)) - 0.5) * (hi - lo) * 2
    d = tgt - o
    d = d.astype(np.float32)
    d = d / np.sqrt((d * d).sum(1, keepdims=True, dtype=np.float32))
    return o.astype(np.float32), d.astype(np.float32)


def load_shim():
    out = os.path.join(ROOT, "tests", "_build")
    os.makedirs(out, exist_ok=True)
    so = os.path.join(out, "libbzshim.so")
    srcs = [os.path.join(ROOT, "tests", "host_shim.cpp"), os.path.join(ROOT, "cuda-bezier-triangle-raytracer_b200", "host", "bz_bvh.cpp")]
    deps = srcs + [os.path.join(ROOT, "cuda-bezier-triangle-raytracer_b200", "csrc", "bz_math.cuh"), os.path.join(ROOT, "include", "bz_capi.h")]
    if not os.path.exists(so) or any(os.path.getmtime(p) > os.path.getmtime(so) for p in deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-I" + os.path.join(ROOT, "include")] + srcs + ["-o", so])
    return C.CDLL(so)


def shim_intersect(lib, patches, start, direction, mode):
    n = start.shape[0]
    what, patch = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
    dist, cos = np.zeros(n, np.float32), np.zeros(n, np.float32)
    pt, nr, ba = (np.zeros((n, 3), np.float32) for _ in range(3))
    vis = np.zeros(2, np.uint64)
    f = lambda a: a.ctypes.data_as(C.c_void_p)
    patches = np.ascontiguousarray(patches, np.float32)
    lib.shim_intersect_batch(f(patches), C.c_uint32(patches.shape[0]), f(start), f(direction), C.c_uint64(n), C.c_int(mode), f(what), f(patch),
                             f(dist), f(pt), f(nr), f(cos), f(ba), f(vis))
    return dict(what=what, patch=patch, distance=dist, point=pt, normal=nr, cos=cos, bary=ba), vis


def same_bits(a, b):
    """element-wise equality where NaN matches NaN: the reference accepts a hit whose 4th-iteration surface point is
    NaN (`NaN > 0.01` is false at bezierTriangle.cpp:165), and x86 and the GPU give NaNs different payload bits"""
    a, b = np.asarray(a), np.asarray(b)
    if a.shape != b.shape:
        return False
    if a.dtype.kind == "f":
        return bool(np.all((a == b) | (np.isnan(a) & np.isnan(b))))
    return bool(np.array_equal(a, b))


def assert_same_hits(a, b, label):
    for k in ("what", "patch", "distance", "point", "normal", "cos", "bary"):
        assert same_bits(a[k], b[k]), "%s: %s differs on %d rays" % (label, k, int((a[k] != b[k]).sum()))
