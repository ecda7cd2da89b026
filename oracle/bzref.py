"""TEST INFRASTRUCTURE ONLY -- ctypes binding of oracle/_ref/libbzref*.so.

libbzref.so is the REAL reference (reference/mesh.cpp, bezierTriangle.cpp, bezierMesh.cpp,
bezierLens.cpp compiled from /root/reference by oracle/build_ref.sh against oracle/eigen_shim,
with the one documented mPoint hoist).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module; the product never does.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

NO_PATCH = 0xFFFFFFFF
STATUS_NONE, STATUS_INSIDE, STATUS_OUTSIDE = 0, 1, 2
WHAT_NONE, WHAT_INTERSECT = 3, 4
PATCH_WORDS = 64  # sizeof(bz_patch) / 4, include/bz_capi.h


class Camera(C.Structure):
    _fields_ = [("eye", C.c_float * 3), ("forward", C.c_float * 3), ("right", C.c_float * 3), ("up", C.c_float * 3),
                ("half_w", C.c_float), ("half_h", C.c_float),
                ("width", C.c_uint32), ("height", C.c_uint32), ("strata", C.c_uint32), ("ortho", C.c_uint32)]


class Scene(C.Structure):
    _fields_ = [("bg_plane", C.c_float * 4), ("bg_axis_a", C.c_float * 3), ("bg_axis_b", C.c_float * 3),
                ("cell", C.c_float), ("value_dark", C.c_uint32), ("value_bright", C.c_uint32),
                ("value_lost", C.c_uint32), ("max_crossings", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in
                ("primary_rays", "refract_calls", "stage_a", "stage_b", "snell", "entered", "exited", "lost", "node_visits",
                 "packet_nodes", "packet_walks", "lane_walk_calls")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _up(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint32))


def lib_path(as_written=False):
    return os.path.join(HERE, "_ref", "libbzref_aswritten.so" if as_written else "libbzref.so")


def available(as_written=False):
    return os.path.exists(lib_path(as_written))


_libs = {}


def load(as_written=False):
    key = bool(as_written)
    if key in _libs:
        return _libs[key]
    lib = C.CDLL(lib_path(as_written))
    vp, f, u32, i32, u64 = C.c_void_p, C.c_float, C.c_uint32, C.c_int32, C.c_uint64
    fp, up = C.POINTER(C.c_float), C.POINTER(C.c_uint32)
    sig = {
        "bzref_last_error": (C.c_char_p, []),
        "bzref_is_hoisted": (C.c_int, []),
        "bzref_mesh_new": (vp, []),
        "bzref_mesh_free": (None, [vp]),
        "bzref_mesh_clone": (vp, [vp]),
        "bzref_mesh_make_ellipsoid": (None, [vp, i32, i32, f, f, f]),
        "bzref_mesh_make_unit_sphere": (None, [vp, i32, i32]),
        "bzref_mesh_read_stl": (C.c_int, [vp, C.c_char_p]),
        "bzref_mesh_write_stl": (None, [vp, C.c_char_p]),
        "bzref_mesh_set_triangles": (None, [vp, fp, u32]),
        "bzref_mesh_size": (u32, [vp]),
        "bzref_mesh_get_triangles": (None, [vp, fp]),
        "bzref_mesh_translate": (None, [vp, f, f, f]),
        "bzref_mesh_scale": (None, [vp, f]),
        "bzref_mesh_transform": (None, [vp, fp, fp]),
        "bzref_mesh_standardize_vertices": (None, [vp]),
        "bzref_mesh_standardize_normals": (C.c_int, [vp]),
        "bzref_mesh_get_neighbours": (None, [vp, up, C.POINTER(C.c_uint8)]),
        "bzref_mesh_neighbour_count": (u32, [vp]),
        "bzref_mesh_get_corner_normals": (C.c_int, [vp, fp]),
        "bzref_bezier_new": (vp, [vp]),
        "bzref_bezier_free": (None, [vp]),
        "bzref_bezier_size": (u32, [vp]),
        "bzref_bezier_dump": (None, [vp, vp]),
        "bzref_sizeof_bezier_triangle": (u32, []),
        "bzref_bezier_interpolate": (None, [vp, u32, f, f, f, fp]),
        "bzref_bezier_split_thick": (vp, [vp]),
        "bzref_bezier_tessellate": (vp, [vp, i32]),
        "bzref_patch_intersect": (u32, [vp, u32, fp, fp, u32, fp, fp, fp, fp, fp]),
        "bzref_intersect_batch": (u64, [vp, fp, fp, u64, up, up, fp, fp, fp, fp, fp, C.c_int]),
        "bzref_lens_new": (vp, [f, vp]),
        "bzref_lens_free": (None, [vp]),
        "bzref_refract_batch": (None, [vp, fp, fp, up, u64, fp, fp, up, C.c_int]),
        "bzref_generate_rays": (None, [C.POINTER(Camera), u32, u32, fp, fp]),
        "bzref_render": (u64, [vp, C.POINTER(Camera), C.POINTER(Scene), u32, u32, u32, C.POINTER(C.c_uint8),
                               C.POINTER(Stats), C.c_int]),
        "bzref_hemisphere": (u32, [u32, fp, up]),
        "bzref_hemisphere_random": (u64, [u32, up, up, u64, fp, up]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _libs[key] = lib
    return lib


def _f32(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float32)
    if shape is not None:
        a = a.reshape(shape)
    return a


class RefMesh:
    """reference Mesh (reference/mesh.h:18-133)"""

    def __init__(self, as_written=False, handle=None):
        self.lib = load(as_written)
        self.as_written = as_written
        self.h = handle if handle is not None else self.lib.bzref_mesh_new()

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.bzref_mesh_free(self.h)
            self.h = None

    def clone(self):
        return RefMesh(self.as_written, self.lib.bzref_mesh_clone(self.h))

    def make_ellipsoid(self, sectors, belts, size):
        self.lib.bzref_mesh_make_ellipsoid(self.h, sectors, belts, *[float(x) for x in size])
        return self

    def make_unit_sphere(self, sectors, belts):
        self.lib.bzref_mesh_make_unit_sphere(self.h, sectors, belts)
        return self

    def read_stl(self, path):
        if self.lib.bzref_mesh_read_stl(self.h, path.encode()) != 0:
            raise RuntimeError(self.lib.bzref_last_error().decode())
        return self

    def write_stl(self, path):
        self.lib.bzref_mesh_write_stl(self.h, path.encode())

    def set_triangles(self, tris):
        tris = _f32(tris, (-1, 9))
        self.lib.bzref_mesh_set_triangles(self.h, _fp(tris), tris.shape[0])
        return self

    def size(self):
        return int(self.lib.bzref_mesh_size(self.h))

    def triangles(self):
        out = np.zeros((self.size(), 3, 3), np.float32)
        self.lib.bzref_mesh_get_triangles(self.h, _fp(out))
        return out

    def translate(self, x, y, z):
        self.lib.bzref_mesh_translate(self.h, x, y, z)
        return self

    def scale(self, s):
        self.lib.bzref_mesh_scale(self.h, s)
        return self

    def transform(self, m, d):
        m, d = _f32(m, (3, 3)), _f32(d, (3,))
        self.lib.bzref_mesh_transform(self.h, _fp(m), _fp(d))
        return self

    def standardize_vertices(self):
        self.lib.bzref_mesh_standardize_vertices(self.h)
        return self

    def standardize_normals(self):
        if self.lib.bzref_mesh_standardize_normals(self.h) != 0:
            raise RuntimeError(self.lib.bzref_last_error().decode())
        return self

    def neighbours(self):
        n = int(self.lib.bzref_mesh_neighbour_count(self.h))
        fellows = np.zeros((n, 3), np.uint32)
        starts = np.zeros((n, 3), np.uint8)
        self.lib.bzref_mesh_get_neighbours(self.h, _up(fellows), starts.ctypes.data_as(C.POINTER(C.c_uint8)))
        return fellows, starts

    def corner_normals(self):
        out = np.zeros((self.size(), 3, 3), np.float32)
        if self.lib.bzref_mesh_get_corner_normals(self.h, _fp(out)) != 0:
            raise RuntimeError(self.lib.bzref_last_error().decode())
        return out


class RefBezierMesh:
    """reference BezierMesh (reference/bezierMesh.h:10-45)"""

    def __init__(self, mesh):
        self.lib = mesh.lib
        self.as_written = mesh.as_written
        self.h = self.lib.bzref_bezier_new(mesh.h)
        if not self.h:
            raise RuntimeError(self.lib.bzref_last_error().decode())

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.bzref_bezier_free(self.h)
            self.h = None

    def size(self):
        return int(self.lib.bzref_bezier_size(self.h))

    def patches(self):
        """(n, 64) float32 view of bz_patch records; columns 16:19 hold the uint32 neighbours."""
        out = np.zeros((self.size(), PATCH_WORDS), np.float32)
        self.lib.bzref_bezier_dump(self.h, out.ctypes.data_as(C.c_void_p))
        return out

    def interpolate(self, patch, b0, b1, b2):
        out = np.zeros(3, np.float32)
        self.lib.bzref_bezier_interpolate(self.h, patch, b0, b1, b2, _fp(out))
        return out

    def split_thick(self):
        return RefMesh(self.as_written, self.lib.bzref_bezier_split_thick(self.h))

    def tessellate(self, divisor):
        return RefMesh(self.as_written, self.lib.bzref_bezier_tessellate(self.h, divisor))

    def patch_intersect(self, patch, start, direction, limit):
        s, d = _f32(start, (3,)), _f32(direction, (3,))
        dist, cos = np.zeros(1, np.float32), np.zeros(1, np.float32)
        point, normal, bary = (np.zeros(3, np.float32) for _ in range(3))
        what = self.lib.bzref_patch_intersect(self.h, patch, _fp(s), _fp(d), limit, _fp(dist), _fp(point), _fp(normal),
                                              _fp(cos), _fp(bary))
        return dict(what=int(what), distance=dist[0], point=point, normal=normal, cos=cos[0], bary=bary)

    def intersect(self, start, direction, threads=1):
        s, d = _f32(start, (-1, 3)), _f32(direction, (-1, 3))
        n = s.shape[0]
        what, patch = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        dist, cos = np.zeros(n, np.float32), np.zeros(n, np.float32)
        point, normal, bary = (np.zeros((n, 3), np.float32) for _ in range(3))
        bad = self.lib.bzref_intersect_batch(self.h, _fp(s), _fp(d), n, _up(what), _up(patch), _fp(dist), _fp(point),
                                             _fp(normal), _fp(cos), _fp(bary), threads)
        if bad != 0:
            raise RuntimeError("index-recovering loop disagrees with BezierMesh::intersect on %d rays" % bad)
        return dict(what=what, patch=patch, distance=dist, point=point, normal=normal, cos=cos, bary=bary)


class RefLens:
    """reference BezierLens (reference/bezierLens.h:14-28)"""

    def __init__(self, ri, bezier):
        self.lib = bezier.lib
        self.bezier = bezier
        self.h = self.lib.bzref_lens_new(ri, bezier.h)

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.bzref_lens_free(self.h)
            self.h = None

    def refract(self, start, direction, expected, threads=1):
        s, d = _f32(start, (-1, 3)), _f32(direction, (-1, 3))
        n = s.shape[0]
        e = np.ascontiguousarray(np.broadcast_to(np.asarray(expected, np.uint32), (n,)))
        os_, od = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
        st = np.zeros(n, np.uint32)
        self.lib.bzref_refract_batch(self.h, _fp(s), _fp(d), _up(e), n, _fp(os_), _fp(od), _up(st), threads)
        return os_, od, st

    def render(self, camera, scene, row0=0, row1=None, pixel_step=1, threads=1):
        row1 = camera.height if row1 is None else row1
        rgba = np.zeros((row1 - row0, camera.width, 4), np.uint8)
        stats = Stats()
        rays = self.lib.bzref_render(self.h, C.byref(camera), C.byref(scene), row0, row1, pixel_step,
                                     rgba.ctypes.data_as(C.POINTER(C.c_uint8)), C.byref(stats), threads)
        return rgba, stats, int(rays)


def generate_rays(camera, row0=0, row1=None, as_written=False):
    lib = load(as_written)
    row1 = camera.height if row1 is None else row1
    n = (row1 - row0) * camera.width * camera.strata * camera.strata
    s, d = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
    lib.bzref_generate_rays(C.byref(camera), row0, row1, _fp(s), _fp(d))
    return s, d


def make_camera(width, height, strata=1, eye=(-12.0, 0.0, 0.0), half_w=0.5, half_h=None, ortho=False,
                forward=(1.0, 0.0, 0.0), right=(0.0, 1.0, 0.0), up=(0.0, 0.0, 1.0)):
    cam = Camera()
    cam.eye[:] = eye
    cam.forward[:] = forward
    cam.right[:] = right
    cam.up[:] = up
    cam.half_w = half_w
    cam.half_h = half_w * height / width if half_h is None else half_h
    cam.width, cam.height, cam.strata, cam.ortho = width, height, strata, 1 if ortho else 0
    return cam


def make_scene(bg_x=12.0, cell=1.0, dark=25, bright=230, lost=0, max_crossings=2):
    sc = Scene()
    sc.bg_plane[:] = (1.0, 0.0, 0.0, bg_x)
    sc.bg_axis_a[:] = (0.0, 1.0, 0.0)
    sc.bg_axis_b[:] = (0.0, 0.0, 1.0)
    sc.cell = cell
    sc.value_dark, sc.value_bright, sc.value_lost, sc.max_crossings = dark, bright, lost, max_crossings
    return sc


def hemisphere(belts):
    """belt table of the REAL UniformHemisphere constructor (reference/hostUtil.cpp:3-14) -> (patch count, width[belts], first[belts])"""
    width, first = np.zeros(belts, np.float32), np.zeros(belts, np.uint32)
    count = load().bzref_hemisphere(belts, _fp(width), _up(first))
    return int(count), width, first


def hemisphere_random(belts, draw1, draw2):
    """the REAL UniformHemisphere::getRandom (reference/hostUtil.cpp:16-29) with its engine set so that the two draws of call i
    are the 24-bit integers draw1[i], draw2[i] (u = draw * 2^-24) -> (direction[n,3], patch[n])"""
    d1, d2 = np.ascontiguousarray(draw1, np.uint32), np.ascontiguousarray(draw2, np.uint32)
    assert d1.shape == d2.shape and d1.ndim == 1 and int(max(d1.max(initial=0), d2.max(initial=0))) < (1 << 24)
    direction, patch = np.zeros((d1.size, 3), np.float32), np.zeros(d1.size, np.uint32)
    wrong = load().bzref_hemisphere_random(belts, _up(d1), _up(d2), d1.size, _fp(direction), _up(patch))
    assert wrong == 0, "the reference engine did not deliver the chosen draws on %d calls" % wrong
    return direction, patch


def abi_draws(seed, first, count):
    """the C ABI's counter-based generator (include/bz_capi.h, bz_emitter): 24-bit draws (draw1, draw2) of rays [first, first + count)"""
    ray = np.arange(first, first + count, dtype=np.uint64)
    out = []
    with np.errstate(over="ignore"):
        for draw in (0, 1):
            z = np.uint64(seed) + (np.uint64(2) * ray + np.uint64(draw + 1)) * np.uint64(0x9E3779B97F4A7C15)
            z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
            z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
            z = z ^ (z >> np.uint64(31))
            out.append((z >> np.uint64(40)).astype(np.uint32))
    return out[0], out[1]
