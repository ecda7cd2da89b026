"""bezier_b200 -- Python face of libbz_b200.so (include/bz_capi.h), the B200-native
Bezier-triangle intersect/refract path.

The classes mirror the reference's host API for this path (reference/mesh.h, bezierMesh.h,
bezierLens.h): Mesh -> BezierMesh -> BezierLens.refract.  All arithmetic of the hot path runs in
the hand-written sm_100a kernels behind the C ABI; this module only moves pointers.  There is no
CPU path: without the built library or without a CUDA device every computing call raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BZ_B200_LIB") or os.path.join(os.path.dirname(_HERE), "libbz_b200.so")   # override: debugging builds only

OK, ERR_INVALID, ERR_CUDA, ERR_NO_DEVICE, ERR_PREPROCESS = 0, -1, -2, -3, -4
STATUS_NONE, STATUS_INSIDE, STATUS_OUTSIDE = 0, 1, 2
WHAT_NONE, WHAT_INTERSECT = 3, 4
NO_PATCH = 0xFFFFFFFF
MODE_BRUTE, MODE_CULLED, MODE_LANES, MODE_FAST = 0, 1, 2, 3
PIPELINE_AUTO, PIPELINE_PERSISTENT, PIPELINE_WAVEFRONT = 0, 1, 2
ERR_OVERFLOW = -5
GATHER_PEER, GATHER_NCCL = 0, 1
PATCH_WORDS = 64


class BezierError(RuntimeError):
    def __init__(self, code, text):
        super().__init__("libbz_b200 error %d: %s" % (code, text))
        self.code = code


class Camera(C.Structure):
    """bz_camera (include/bz_capi.h)"""
    _fields_ = [("eye", C.c_float * 3), ("forward", C.c_float * 3), ("right", C.c_float * 3), ("up", C.c_float * 3),
                ("half_w", C.c_float), ("half_h", C.c_float),
                ("width", C.c_uint32), ("height", C.c_uint32), ("strata", C.c_uint32), ("ortho", C.c_uint32)]


class Scene(C.Structure):
    """bz_scene (include/bz_capi.h)"""
    _fields_ = [("bg_plane", C.c_float * 4), ("bg_axis_a", C.c_float * 3), ("bg_axis_b", C.c_float * 3),
                ("cell", C.c_float), ("value_dark", C.c_uint32), ("value_bright", C.c_uint32),
                ("value_lost", C.c_uint32), ("max_crossings", C.c_uint32)]


STAT_NAMES = ("primary_rays", "refract_calls", "stage_a", "stage_b", "snell", "entered", "exited", "lost", "node_visits",
              "packet_nodes", "packet_walks", "lane_walk_calls")


class Stats(C.Structure):
    """bz_stats (include/bz_capi.h)"""
    _fields_ = [(n, C.c_uint64) for n in STAT_NAMES]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n in STAT_NAMES}


class Emitter(C.Structure):
    """bz_emitter (include/bz_capi.h)"""
    _fields_ = [("origin", C.c_float * 3), ("edge_u", C.c_float * 3), ("edge_v", C.c_float * 3), ("normal", C.c_float * 3),
                ("tangent_a", C.c_float * 3), ("tangent_b", C.c_float * 3), ("points_u", C.c_uint32), ("points_v", C.c_uint32),
                ("rays_per_point", C.c_uint32), ("belts", C.c_uint32), ("seed", C.c_uint64)]

    @property
    def rays(self):
        return self.points_u * self.points_v * self.rays_per_point


class Target(C.Structure):
    """bz_target (include/bz_capi.h)"""
    _fields_ = [("origin", C.c_float * 3), ("edge_u", C.c_float * 3), ("edge_v", C.c_float * 3), ("bins_u", C.c_uint32), ("bins_v", C.c_uint32)]


LAND_LOST, LAND_MISSED = 0xffffffff, 0xfffffffe


def make_emitter(origin, edge_u, edge_v, normal=(1.0, 0.0, 0.0), tangent_a=(0.0, 1.0, 0.0), tangent_b=(0.0, 0.0, 1.0),
                 points=(4, 4), rays_per_point=1024, belts=16, seed=1):
    em = Emitter()
    em.origin[:], em.edge_u[:], em.edge_v[:] = origin, edge_u, edge_v
    em.normal[:], em.tangent_a[:], em.tangent_b[:] = normal, tangent_a, tangent_b
    em.points_u, em.points_v, em.rays_per_point, em.belts, em.seed = points[0], points[1], rays_per_point, belts, seed
    return em


def make_target(origin, edge_u, edge_v, bins=(64, 64)):
    tg = Target()
    tg.origin[:], tg.edge_u[:], tg.edge_v[:] = origin, edge_u, edge_v
    tg.bins_u, tg.bins_v = bins
    return tg


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


_lib = None

_vp, _f, _u32, _i32, _u64, _int = C.c_void_p, C.c_float, C.c_uint32, C.c_int32, C.c_uint64, C.c_int

# every symbol include/bz_capi.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "bz_last_error": (C.c_char_p, []),
    "bz_device_count": (_int, [C.POINTER(_int)]),
    "bz_sizeof_patch": (_u32, []),
    "bz_sizeof_camera": (_u32, []),
    "bz_sizeof_scene": (_u32, []),
    "bz_sizeof_stats": (_u32, []),
    "bz_lens_create": (_int, [_vp, _u32, _f, _int, C.POINTER(_vp)]),
    "bz_lens_destroy": (_int, [_vp]),
    "bz_lens_set_mode": (_int, [_vp, _u32]),
    "bz_lens_patch_count": (_int, [_vp, C.POINTER(_u32)]),
    "bz_lens_set_pipeline": (_int, [_vp, _u32]),
    "bz_lens_check": (_int, [_vp]),
    "bz_lens_set_wave_chunk": (_int, [_vp, _u32]),
    "bz_lens_release_stream": (_int, [_vp, _vp]),
    "bz_last_plan_ms": (_int, [_vp, C.POINTER(_f)]),
    "bz_lens_info": (_int, [_vp, C.POINTER(_u32 * 4)]),
    "bz_refract_batch": (_int, [_vp, _vp, _vp, _vp, _u64, _vp, _vp, _vp, _vp, _vp]),
    "bz_refract_batch_host": (_int, [_vp, _vp, _vp, _vp, _u64, _vp, _vp, _vp, _vp]),
    "bz_intersect_batch": (_int, [_vp, _vp, _vp, _u64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "bz_intersect_batch_host": (_int, [_vp, _vp, _vp, _u64, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "bz_generate_rays": (_int, [C.POINTER(Camera), _u32, _u32, _vp, _vp, _vp]),
    "bz_render": (_int, [_vp, C.POINTER(Camera), C.POINTER(Scene), _u32, _u32, _vp, _vp, _vp]),
    "bz_render_host": (_int, [_vp, C.POINTER(Camera), C.POINTER(Scene), _u32, _u32, _vp, C.POINTER(Stats)]),
    "bz_render_static": (_int, [_vp, C.POINTER(Camera), C.POINTER(Scene), _u32, _u32, _vp, _vp, _vp]),
    "bz_striped_rows": (_u32, [_u32, _u32, _u32, _u32]),
    "bz_render_striped": (_int, [_vp, C.POINTER(Camera), C.POINTER(Scene), _u32, _u32, _u32, _vp, _vp, _vp]),
    "bz_assemble_stripes": (_int, [_vp, _u32, _u32, _u32, _u32, _u32, _vp, _vp]),
    "bz_comm_unique_id": (_int, [_vp]),
    "bz_comm_create": (_int, [_vp, _u32, _u32, _int, C.POINTER(_vp)]),
    "bz_comm_destroy": (_int, [_vp]),
    "bz_gather_rows": (_int, [_vp, _vp, _u32, _u32, _u32, _u32, _vp, _vp]),
    "bz_comm_set_gather": (_int, [_vp, _u32]),
    "bz_comm_gather_mode": (_int, [_vp, C.POINTER(_u32)]),
    "bz_fp32_peak_probe": (_int, [_int, C.POINTER(_f), C.POINTER(_f)]),
    "bz_mesh_create": (_int, [C.POINTER(_vp)]),
    "bz_mesh_destroy": (_int, [_vp]),
    "bz_mesh_clone": (_int, [_vp, C.POINTER(_vp)]),
    "bz_mesh_make_ellipsoid": (_int, [_vp, _i32, _i32, C.POINTER(_f * 3)]),
    "bz_mesh_make_unit_sphere": (_int, [_vp, _i32, _i32]),
    "bz_mesh_read_stl": (_int, [_vp, C.c_char_p]),
    "bz_mesh_write_stl": (_int, [_vp, C.c_char_p]),
    "bz_mesh_set_triangles": (_int, [_vp, _vp, _u32]),
    "bz_mesh_size": (_int, [_vp, C.POINTER(_u32)]),
    "bz_mesh_get_triangles": (_int, [_vp, _vp]),
    "bz_mesh_transform": (_int, [_vp, _vp, _vp]),
    "bz_mesh_split_triangles": (_int, [_vp, _i32]),
    "bz_mesh_split_triangles_by_side": (_int, [_vp, _f]),
    "bz_mesh_standardize_vertices": (_int, [_vp]),
    "bz_mesh_standardize_normals": (_int, [_vp]),
    "bz_mesh_neighbour_count": (_int, [_vp, C.POINTER(_u32)]),
    "bz_mesh_get_face2neighbours": (_int, [_vp, _vp, _vp]),
    "bz_mesh_get_corner_normals": (_int, [_vp, _vp]),
    "bz_bezier_create": (_int, [_vp, C.POINTER(_vp)]),
    "bz_bezier_destroy": (_int, [_vp]),
    "bz_bezier_size": (_int, [_vp, C.POINTER(_u32)]),
    "bz_bezier_get_patches": (_int, [_vp, _vp]),
    "bz_bezier_interpolate": (_int, [_vp, _i32, C.POINTER(_vp)]),
    "bz_bezier_split_thick": (_int, [_vp, C.POINTER(_vp)]),
    "bz_lens_create_from_bezier": (_int, [_vp, _f, _int, C.POINTER(_vp)]),
    "bz_last_kernel_ms": (_int, [_vp, C.POINTER(_f)]),
    "bz_last_kernel_launches": (_int, [_vp, C.POINTER(_u32)]),
    "bz_sizeof_emitter": (_u32, []),
    "bz_sizeof_target": (_u32, []),
    "bz_hemisphere_patches": (_int, [_u32, C.POINTER(_u32), _vp, _vp]),
    "bz_emit_rays": (_int, [C.POINTER(Emitter), _u64, _u64, _vp, _vp, _vp, _vp]),
    "bz_trace_to_target": (_int, [_vp, _vp, _vp, _u64, C.POINTER(Target), _u32, _vp, _vp, _vp, _vp]),
    "bz_illuminate": (_int, [_vp, C.POINTER(Emitter), C.POINTER(Target), _u32, _vp, _vp, _vp]),
    "bz_illuminate_host": (_int, [_vp, C.POINTER(Emitter), C.POINTER(Target), _u32, _vp, C.POINTER(Stats)]),
}


def load():
    """Load libbz_b200.so; raises if it has not been built (there is no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("libbz_b200.so is not built: run `make -C %s` (or __graft_entry__.build())" % os.path.dirname(LIB_PATH))
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def _check(rc):
    if rc != OK:
        raise BezierError(rc, load().bz_last_error().decode())


def device_count():
    n = _int(0)
    _check(load().bz_device_count(C.byref(n)))
    return n.value


def fp32_peak_probe(device=0):
    """(TFLOP/s, ms) of the dependent-FMA probe kernel on `device`."""
    tf, ms = _f(0.0), _f(0.0)
    _check(load().bz_fp32_peak_probe(device, C.byref(tf), C.byref(ms)))
    return tf.value, ms.value


def striped_rows(height, stripe_rows, rank, ranks):
    return int(load().bz_striped_rows(height, stripe_rows, rank, ranks))


def _np(a, dtype, shape):
    return np.ascontiguousarray(a, dtype=dtype).reshape(shape)


def _ptr(t):
    """device pointer of a torch tensor (or None)"""
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream_handle(stream):
    if stream is None:
        import torch
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)
    return C.c_void_p(int(stream))


class Mesh:
    """Mesh -- reference/mesh.h:18-133 (host preprocessing, C++ class bz::Mesh behind the C ABI)"""

    def __init__(self, handle=None):
        self.lib = load()
        if handle is None:
            handle = _vp()
            _check(self.lib.bz_mesh_create(C.byref(handle)))
        self.h = handle

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.bz_mesh_destroy(self.h)
            self.h = None

    def clone(self):
        out = _vp()
        _check(self.lib.bz_mesh_clone(self.h, C.byref(out)))
        return Mesh(out)

    def make_ellipsoid(self, sectors, belts, size):
        _check(self.lib.bz_mesh_make_ellipsoid(self.h, sectors, belts, C.byref((_f * 3)(*[float(x) for x in size]))))
        return self

    def make_unit_sphere(self, sectors, belts):
        _check(self.lib.bz_mesh_make_unit_sphere(self.h, sectors, belts))
        return self

    def read_stl(self, path):
        _check(self.lib.bz_mesh_read_stl(self.h, os.fsencode(path)))
        return self

    def write_stl(self, path):
        _check(self.lib.bz_mesh_write_stl(self.h, os.fsencode(path)))

    def set_triangles(self, triangles):
        t = _np(triangles, np.float32, (-1, 9))
        _check(self.lib.bz_mesh_set_triangles(self.h, t.ctypes.data_as(_vp), t.shape[0]))
        return self

    def size(self):
        n = _u32(0)
        _check(self.lib.bz_mesh_size(self.h, C.byref(n)))
        return n.value

    def triangles(self):
        out = np.zeros((self.size(), 3, 3), np.float32)
        _check(self.lib.bz_mesh_get_triangles(self.h, out.ctypes.data_as(_vp)))
        return out

    def transform(self, matrix, displacement):
        m, d = _np(matrix, np.float32, (3, 3)), _np(displacement, np.float32, (3,))
        _check(self.lib.bz_mesh_transform(self.h, m.ctypes.data_as(_vp), d.ctypes.data_as(_vp)))
        return self

    def translate(self, x, y, z):
        return self.transform(np.eye(3, dtype=np.float32), (x, y, z))

    def scale(self, factor):
        return self.transform(np.eye(3, dtype=np.float32) * np.float32(factor), (0.0, 0.0, 0.0))

    def split_triangles(self, divisor):
        _check(self.lib.bz_mesh_split_triangles(self.h, int(divisor)))
        return self

    def split_triangles_by_side(self, max_side):
        _check(self.lib.bz_mesh_split_triangles_by_side(self.h, float(max_side)))
        return self

    def standardize_vertices(self):
        _check(self.lib.bz_mesh_standardize_vertices(self.h))
        return self

    def standardize_normals(self):
        _check(self.lib.bz_mesh_standardize_normals(self.h))
        return self

    def neighbours(self):
        """getFace2neighbours -> (fellow[n,3] u32, starts[n,3] u8)"""
        n = _u32(0)
        _check(self.lib.bz_mesh_neighbour_count(self.h, C.byref(n)))
        fellow, starts = np.zeros((n.value, 3), np.uint32), np.zeros((n.value, 3), np.uint8)
        _check(self.lib.bz_mesh_get_face2neighbours(self.h, fellow.ctypes.data_as(_vp), starts.ctypes.data_as(_vp)))
        return fellow, starts

    def corner_normals(self):
        """getVertex2averageNormals looked up per corner -> [n,3,3]"""
        out = np.zeros((self.size(), 3, 3), np.float32)
        _check(self.lib.bz_mesh_get_corner_normals(self.h, out.ctypes.data_as(_vp)))
        return out


class BezierMesh:
    """BezierMesh(mesh) -- reference/bezierMesh.h:10-45 (C++ class bz::BezierMesh behind the C ABI)"""

    def __init__(self, mesh):
        self.lib = load()
        handle = _vp()
        _check(self.lib.bz_bezier_create(mesh.h, C.byref(handle)))
        self.h = handle

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.bz_bezier_destroy(self.h)
            self.h = None

    def size(self):
        n = _u32(0)
        _check(self.lib.bz_bezier_size(self.h, C.byref(n)))
        return n.value

    def patches(self):
        """(n, 64) float32 view of the bz_patch table; columns 16:19 hold the uint32 neighbours"""
        out = np.zeros((self.size(), PATCH_WORDS), np.float32)
        _check(self.lib.bz_bezier_get_patches(self.h, out.ctypes.data_as(_vp)))
        return out

    def interpolate(self, divisor):
        out = _vp()
        _check(self.lib.bz_bezier_interpolate(self.h, int(divisor), C.byref(out)))
        return Mesh(out)

    def split_thick(self):
        out = _vp()
        _check(self.lib.bz_bezier_split_thick(self.h, C.byref(out)))
        return Mesh(out)


def hemisphere_patches(belts):
    """UniformHemisphere belt table (reference/hostUtil.cpp:3-14) -> (patch count, width[belts], first[belts])"""
    n = _u32(0)
    width, first = np.zeros(belts, np.float32), np.zeros(belts, np.uint32)
    _check(load().bz_hemisphere_patches(belts, C.byref(n), width.ctypes.data_as(_vp), first.ctypes.data_as(_vp)))
    return n.value, width, first


def emit_rays(emitter, first=0, count=None, device=0):
    """bz_emit_rays on `device` -> host (start, dir, patch)"""
    import torch
    count = emitter.rays - first if count is None else count
    dev = torch.device("cuda", device)
    with torch.cuda.device(dev):
        s = torch.zeros((count, 3), dtype=torch.float32, device=dev)
        d = torch.zeros_like(s)
        p = torch.zeros(count, dtype=torch.int32, device=dev)
        _check(load().bz_emit_rays(C.byref(emitter), first, count, _ptr(s), _ptr(d), _ptr(p), _stream_handle(None)))
        torch.cuda.synchronize()
    return s.cpu().numpy(), d.cpu().numpy(), p.cpu().numpy().view(np.uint32)


class BezierLens:
    """BezierLens(ri, bezier_mesh) -- reference/bezierLens.h:14-28.

    `patches` is either a BezierMesh of this package or an (n, 64) float32 array of bz_patch
    records.  The lens owns the SoA tables and the cull hierarchy on `device`."""

    def __init__(self, ri, patches, device=0):
        self.lib = load()
        self.device = device
        self.ri = float(ri)
        handle = _vp()
        if isinstance(patches, BezierMesh):
            self.patch_count = patches.size()
            _check(self.lib.bz_lens_create_from_bezier(patches.h, self.ri, device, C.byref(handle)))
        else:
            table = np.ascontiguousarray(patches, dtype=np.float32)
            if table.size == 0:
                table = table.reshape(0, PATCH_WORDS)
            assert table.ndim == 2 and table.shape[1] == PATCH_WORDS
            self.patch_count = table.shape[0]
            _check(self.lib.bz_lens_create(table.ctypes.data_as(_vp), table.shape[0], self.ri, device, C.byref(handle)))
        self.h = handle

    def close(self):
        if getattr(self, "h", None):
            self.lib.bz_lens_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_mode(self, mode):
        _check(self.lib.bz_lens_set_mode(self.h, mode))
        return self

    def set_pipeline(self, pipeline):
        """PIPELINE_AUTO / PIPELINE_PERSISTENT / PIPELINE_WAVEFRONT for the render calls (same bits either way)"""
        _check(self.lib.bz_lens_set_pipeline(self.h, pipeline))
        return self

    def set_wave_chunk(self, ray_slots):
        _check(self.lib.bz_lens_set_wave_chunk(self.h, ray_slots))
        return self

    def check(self):
        """synchronise and raise if a device-side work list overflowed since the last check"""
        _check(self.lib.bz_lens_check(self.h))

    def info(self):
        out = (_u32 * 4)()
        _check(self.lib.bz_lens_info(self.h, C.byref(out)))
        return dict(nodes=out[0], levels=out[1], always=out[2], sm_count=out[3])

    def last_kernel_ms(self):
        ms = _f(0.0)
        _check(self.lib.bz_last_kernel_ms(self.h, C.byref(ms)))
        return ms.value

    def last_plan_ms(self):
        ms = _f(0.0)
        _check(self.lib.bz_last_plan_ms(self.h, C.byref(ms)))
        return ms.value

    def release_stream(self, stream):
        _check(self.lib.bz_lens_release_stream(self.h, C.c_void_p(int(stream))))

    def last_kernel_launches(self):
        count = _u32(0)
        _check(self.lib.bz_last_kernel_launches(self.h, C.byref(count)))
        return count.value

    # ---- host-buffer entry points (numpy in, numpy out; copies inside the call) ----

    def refract(self, start, direction, expected):
        """BezierLens::refract over a batch of host rays -> (start, dir, status, patch)"""
        s, d = _np(start, np.float32, (-1, 3)), _np(direction, np.float32, (-1, 3))
        n = s.shape[0]
        e = np.ascontiguousarray(np.broadcast_to(np.asarray(expected, np.uint32), (n,)))
        os_, od = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
        st, pa = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        p = lambda a: a.ctypes.data_as(_vp)
        _check(self.lib.bz_refract_batch_host(self.h, p(s), p(d), p(e), n, p(os_), p(od), p(st), p(pa)))
        return os_, od, st, pa

    def intersect(self, start, direction):
        """BezierMesh::intersect over a batch of host rays -> dict"""
        s, d = _np(start, np.float32, (-1, 3)), _np(direction, np.float32, (-1, 3))
        n = s.shape[0]
        what, patch = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        dist, cos = np.zeros(n, np.float32), np.zeros(n, np.float32)
        point, normal, bary = (np.zeros((n, 3), np.float32) for _ in range(3))
        p = lambda a: a.ctypes.data_as(_vp)
        _check(self.lib.bz_intersect_batch_host(self.h, p(s), p(d), n, p(what), p(patch), p(dist), p(point), p(normal), p(cos), p(bary)))
        return dict(what=what, patch=patch, distance=dist, point=point, normal=normal, cos=cos, bary=bary)

    def trace_to_target(self, start, direction, target, max_crossings=2):
        """bz_trace_to_target over host rays -> (landing per ray, bins (bins_v, bins_u), Stats)"""
        import torch
        dev = torch.device("cuda", self.device)
        with torch.cuda.device(dev):
            s = torch.from_numpy(_np(start, np.float32, (-1, 3))).to(dev)
            d = torch.from_numpy(_np(direction, np.float32, (-1, 3))).to(dev)
            n = s.shape[0]
            bins = torch.zeros((target.bins_v, target.bins_u), dtype=torch.int32, device=dev)
            landing = torch.zeros(n, dtype=torch.int32, device=dev)
            stats = torch.zeros(len(STAT_NAMES), dtype=torch.int64, device=dev)
            _check(self.lib.bz_trace_to_target(self.h, _ptr(s), _ptr(d), n, C.byref(target), max_crossings, _ptr(bins), _ptr(landing), _ptr(stats),
                                               _stream_handle(None)))
            torch.cuda.synchronize()
        out = Stats()
        for name, value in zip(STAT_NAMES, stats.tolist()):
            setattr(out, name, value)
        return landing.cpu().numpy().view(np.uint32), bins.cpu().numpy().view(np.uint32), out

    def illuminate(self, emitter, target, max_crossings=2, bins=None):
        """bz_illuminate_host: the whole simulation -> (bins (bins_v, bins_u) uint32, Stats); `bins` is accumulated into when given"""
        out = np.zeros((target.bins_v, target.bins_u), np.uint32) if bins is None else bins
        stats = Stats()
        _check(self.lib.bz_illuminate_host(self.h, C.byref(emitter), C.byref(target), max_crossings, out.ctypes.data_as(_vp), C.byref(stats)))
        return out, stats

    def illuminate_device(self, emitter, target, bins, max_crossings=2, stats=None, stream=None):
        _check(self.lib.bz_illuminate(self.h, C.byref(emitter), C.byref(target), max_crossings, _ptr(bins), _ptr(stats), _stream_handle(stream)))

    def render(self, camera, scene, row0=0, row1=None):
        """host image of rows [row0,row1): (rows, width, 4) uint8, Stats"""
        row1 = camera.height if row1 is None else row1
        rgba = np.zeros((row1 - row0, camera.width, 4), np.uint8)
        stats = Stats()
        _check(self.lib.bz_render_host(self.h, C.byref(camera), C.byref(scene), row0, row1, rgba.ctypes.data_as(_vp), C.byref(stats)))
        return rgba, stats

    def render_host_into(self, camera, scene, host_ptr, row0=0, row1=None, stats=None):
        """bz_render_host into caller-owned host memory (e.g. a pinned buffer); `stats` is a Stats to accumulate into"""
        row1 = camera.height if row1 is None else row1
        _check(self.lib.bz_render_host(self.h, C.byref(camera), C.byref(scene), row0, row1, C.c_void_p(int(host_ptr)),
                                       None if stats is None else C.byref(stats)))

    # ---- device-buffer entry points (torch CUDA tensors, asynchronous on the stream) ----

    def refract_device(self, start, direction, expected, out_start, out_dir, out_status, out_patch=None, stream=None):
        n = start.shape[0]
        _check(self.lib.bz_refract_batch(self.h, _ptr(start), _ptr(direction), _ptr(expected), n, _ptr(out_start), _ptr(out_dir),
                                         _ptr(out_status), _ptr(out_patch), _stream_handle(stream)))

    def render_device(self, camera, scene, rgba, row0=0, row1=None, stats=None, stream=None):
        row1 = camera.height if row1 is None else row1
        _check(self.lib.bz_render(self.h, C.byref(camera), C.byref(scene), row0, row1, _ptr(rgba), _ptr(stats), _stream_handle(stream)))

    def render_static_device(self, camera, scene, rgba, row0=0, row1=None, stats=None, stream=None):
        """bz_render_static: the render without any allocation (static device scratch, at most STATIC_MAX_PIXELS pixels per call)"""
        row1 = camera.height if row1 is None else row1
        _check(self.lib.bz_render_static(self.h, C.byref(camera), C.byref(scene), row0, row1, _ptr(rgba), _ptr(stats), _stream_handle(stream)))

    def render_striped_device(self, camera, scene, stripe_rows, rank, ranks, rgba, stats=None, stream=None):
        _check(self.lib.bz_render_striped(self.h, C.byref(camera), C.byref(scene), stripe_rows, rank, ranks, _ptr(rgba), _ptr(stats),
                                          _stream_handle(stream)))


class Comm:
    """bz_comm: the NCCL communicator of the image gather (include/bz_capi.h).  `share_id(id_bytes or None) -> id_bytes` ships the
    128-byte id from rank 0 to every rank (torch.distributed broadcast in sharding.make_comm, MPI, a file ...)."""

    def __init__(self, rank, ranks, device, share_id):
        self.lib = load()
        self.rank, self.ranks, self.device = rank, ranks, device
        ident = None
        if rank == 0:
            buf = (C.c_uint8 * 128)()
            _check(self.lib.bz_comm_unique_id(C.byref(buf)))
            ident = bytes(buf)
        ident = share_id(ident)
        assert len(ident) == 128
        handle = _vp()
        _check(self.lib.bz_comm_create((C.c_uint8 * 128).from_buffer_copy(ident), rank, ranks, device, C.byref(handle)))
        self.h = handle

    def gather_rows(self, band, width, height, stripe_rows, rgba=None, root=0, stream=None):
        """every rank: its stripe band (device tensor); root: `rgba` [height, width, 4] gets the assembled frame"""
        _check(self.lib.bz_gather_rows(self.h, _ptr(band), width, height, stripe_rows, root, _ptr(rgba), _stream_handle(stream)))

    def set_gather(self, mode):
        """GATHER_PEER (default: NVLink stores into the root's buffer) or GATHER_NCCL (send / recv + assembly); same on every rank"""
        _check(self.lib.bz_comm_set_gather(self.h, mode))
        return self

    def gather_mode(self):
        out = _u32(0)
        _check(self.lib.bz_comm_gather_mode(self.h, C.byref(out)))
        return int(out.value)

    def close(self):
        if getattr(self, "h", None):
            self.lib.bz_comm_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def assemble_stripes_device(bands, band_rows, width, height, stripe_rows, ranks, rgba, stream=None):
    """gathered stripe buffers [ranks, band_rows, width, 4] -> image [height, width, 4] (device tensors)"""
    _check(load().bz_assemble_stripes(_ptr(bands), band_rows, width, height, stripe_rows, ranks, _ptr(rgba), _stream_handle(stream)))


def generate_rays_device(camera, start, direction, row0=0, row1=None, stream=None):
    row1 = camera.height if row1 is None else row1
    _check(load().bz_generate_rays(C.byref(camera), row0, row1, _ptr(start), _ptr(direction), _stream_handle(stream)))
