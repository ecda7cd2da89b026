"""TEST INFRASTRUCTURE ONLY -- ctypes binding of oracle/libbzoracle.so (oracle/bz_oracle.c), the
plain-C CPU restatement of BezierTriangle::intersect / BezierMesh::intersect / BezierLens::refract
working on a bz_patch table.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg
may import this module; the product never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libbzoracle.so")
LIMIT_THIS, LIMIT_NONE = 0, 1
_lib = None


def build():
    subprocess.check_call(["make", "-C", HERE, "libbzoracle.so"], stdout=subprocess.DEVNULL)


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB):
        build()
    lib = C.CDLL(LIB)
    vp, f, u32, u64 = C.c_void_p, C.c_float, C.c_uint32, C.c_uint64
    fp, up, wp = C.POINTER(C.c_float), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)
    lib.bzo_refract_batch.restype = None
    lib.bzo_refract_batch.argtypes = [vp, u32, f, fp, fp, up, u64, fp, fp, up, up, C.c_int, wp]
    lib.bzo_intersect_batch.restype = None
    lib.bzo_intersect_batch.argtypes = [vp, u32, fp, fp, u64, up, up, fp, fp, fp, fp, fp, C.c_int, wp]
    lib.bzo_intersect_batch_runner_up.restype = None
    lib.bzo_intersect_batch_runner_up.argtypes = [vp, u32, fp, fp, u64, up, up, fp, fp, C.c_int]
    lib.bzo_patch_intersect.restype = u32
    lib.bzo_patch_intersect.argtypes = [vp, fp, fp, u32, fp, fp, fp, fp, fp]
    lib.bzo_interpolate.restype = None
    lib.bzo_interpolate.argtypes = [vp, f, f, f, fp]
    lib.bzo_hemisphere.restype = u32
    lib.bzo_hemisphere.argtypes = [u32, fp, up]
    lib.bzo_emit.restype = None
    lib.bzo_emit.argtypes = [vp, u64, u64, fp, fp, up]
    lib.bzo_land.restype = None
    lib.bzo_land.argtypes = [vp, fp, fp, C.POINTER(C.c_uint8), u64, up, up]
    _lib = lib
    return lib


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _up(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint32))


def _f32(a, shape):
    return np.ascontiguousarray(a, dtype=np.float32).reshape(shape)


def _table(patches):
    patches = np.ascontiguousarray(patches, dtype=np.float32)
    assert patches.ndim == 2 and patches.shape[1] == 64
    return patches


def refract(patches, ri, start, direction, expected, threads=1):
    """BezierLens::refract over a batch -> (out_start, out_dir, status, patch, work[3])"""
    lib = load()
    patches = _table(patches)
    s, d = _f32(start, (-1, 3)), _f32(direction, (-1, 3))
    n = s.shape[0]
    e = np.ascontiguousarray(np.broadcast_to(np.asarray(expected, np.uint32), (n,)))
    os_, od = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
    st, pa = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
    work = np.zeros(3, np.uint64)
    lib.bzo_refract_batch(patches.ctypes.data, patches.shape[0], ri, _fp(s), _fp(d), _up(e), n, _fp(os_), _fp(od), _up(st),
                          _up(pa), threads, work.ctypes.data_as(C.POINTER(C.c_uint64)))
    return os_, od, st, pa, work


def intersect(patches, start, direction, threads=1):
    """BezierMesh::intersect over a batch -> dict(what, patch, distance, point, normal, cos, bary, work)"""
    lib = load()
    patches = _table(patches)
    s, d = _f32(start, (-1, 3)), _f32(direction, (-1, 3))
    n = s.shape[0]
    what, patch = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
    dist, cos = np.zeros(n, np.float32), np.zeros(n, np.float32)
    point, normal, bary = (np.zeros((n, 3), np.float32) for _ in range(3))
    work = np.zeros(3, np.uint64)
    lib.bzo_intersect_batch(patches.ctypes.data, patches.shape[0], _fp(s), _fp(d), n, _up(what), _up(patch), _fp(dist),
                            _fp(point), _fp(normal), _fp(cos), _fp(bary), threads, work.ctypes.data_as(C.POINTER(C.c_uint64)))
    return dict(what=what, patch=patch, distance=dist, point=point, normal=normal, cos=cos, bary=bary, work=work)


def edge_ambiguous(patches, start, direction, size, threads=1):
    """rays whose two best accepted hits lie within 1e-3 * size of each other (patch-edge duplicates, SURVEY.md section 4 iii)
    -> boolean mask; the bit-exact tests compare these rays like all others and REPORT how many there were"""
    lib = load()
    patches = _table(patches)
    s, d = _f32(start, (-1, 3)), _f32(direction, (-1, 3))
    n = s.shape[0]
    what, patch = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
    dist, runner = np.zeros(n, np.float32), np.zeros(n, np.float32)
    lib.bzo_intersect_batch_runner_up(patches.ctypes.data, patches.shape[0], _fp(s), _fp(d), n, _up(what), _up(patch), _fp(dist), _fp(runner), threads)
    return (what == 4) & (runner < 3e38) & (np.abs(runner.astype(np.float64) - dist) < 1e-3 * size)


def patch_intersect(patches, index, start, direction, limit):
    lib = load()
    patches = _table(patches)
    s, d = _f32(start, (3,)), _f32(direction, (3,))
    dist, cos = np.zeros(1, np.float32), np.zeros(1, np.float32)
    point, normal, bary = (np.zeros(3, np.float32) for _ in range(3))
    what = lib.bzo_patch_intersect(patches.ctypes.data + 256 * int(index), _fp(s), _fp(d), limit, _fp(dist), _fp(point),
                                   _fp(normal), _fp(cos), _fp(bary))
    return dict(what=int(what), distance=dist[0], point=point, normal=normal, cos=cos[0], bary=bary)


def interpolate(patches, index, b0, b1, b2):
    lib = load()
    patches = _table(patches)
    out = np.zeros(3, np.float32)
    lib.bzo_interpolate(patches.ctypes.data + 256 * int(index), b0, b1, b2, _fp(out))
    return out


# ---- illumination mode (oracle/bz_oracle_illum.c); emitter / target are the ctypes structures of the product binding or
# any object with the same memory layout (include/bz_capi.h: bz_emitter, bz_target) ----
LAND_LOST, LAND_MISSED = 0xffffffff, 0xfffffffe


def hemisphere(belts):
    """UniformHemisphere belt table -> (patch count, width[belts], first[belts])"""
    lib = load()
    width, first = np.zeros(belts, np.float32), np.zeros(belts, np.uint32)
    return int(lib.bzo_hemisphere(belts, _fp(width), _up(first))), width, first


def emit(emitter, first, count):
    """UniformHemisphere::getRandom with the C ABI's counter-based generator -> (start, dir, patch)"""
    lib = load()
    s, d, p = np.zeros((count, 3), np.float32), np.zeros((count, 3), np.float32), np.zeros(count, np.uint32)
    lib.bzo_emit(C.addressof(emitter), first, count, _fp(s), _fp(d), _up(p))
    return s, d, p


def trace_to_target(patches, ri, start, direction, target, max_crossings, threads=1):
    """bounce loop of bz_render / bz_trace_to_target around refract(), then the landing on the target -> (landing, bins, counts)"""
    lib = load()
    s, d = _f32(start, (-1, 3)).copy(), _f32(direction, (-1, 3)).copy()
    n = s.shape[0]
    expected = np.full(n, 1, np.uint32)
    alive = np.ones(n, bool)
    counts = dict(refract_calls=0, entered=0, exited=0)
    for _ in range(max_crossings):
        idx = np.nonzero(alive)[0]
        if idx.size == 0:
            break
        os_, od, st, _, _ = refract(patches, ri, s[idx], d[idx], expected[idx], threads)
        counts["refract_calls"] += int(idx.size)
        counts["entered"] += int((st == 1).sum())
        counts["exited"] += int((st == 2).sum())
        ok = st != 0
        s[idx[ok]], d[idx[ok]] = os_[ok], od[ok]
        expected[idx[ok]] = 3 - expected[idx[ok]]
        alive[idx[~ok]] = False
    in_air = np.ascontiguousarray((expected == 1).astype(np.uint8))
    landing = np.zeros(n, np.uint32)
    bins = np.zeros(target.bins_v * target.bins_u, np.uint32)
    lib.bzo_land(C.addressof(target), _fp(s), _fp(d), in_air.ctypes.data_as(C.POINTER(C.c_uint8)), n, _up(landing), _up(bins))
    counts["lost"] = int((landing == LAND_LOST).sum())
    return landing, bins.reshape(target.bins_v, target.bins_u), counts
