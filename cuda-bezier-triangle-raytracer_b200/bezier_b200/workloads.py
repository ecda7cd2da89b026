"""The five configurations of BASELINE.json as concrete, deterministic workloads.

Only plain numbers are produced here, so the same definition drives the CUDA path (bezier_b200) and,
in tests / bench baselines, the checker (oracle/bzref.py) -- this module imports neither.

Framing (the reference has no camera, DESIGN.md "Renderer definitions"): view along +x, pinhole,
s = max(z half-extent, y half-extent * height / width) of the mesh bounding box, eye 6.5 s in front of
the box centre plane, vertical half-angle so that the lens fills 60 % of the limiting image axis,
checker background 6 s behind the box with cells of 0.5 s.
"""
import os

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ROBOT_STL = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "data", "robot.stl")      # BASELINE.json config 3 names this file (reference/robot.stl)

WORKLOADS = {
    # name: mesh recipe, image, strata (spp = strata^2), refractive index, max surface crossings
    "cfg1_sphere_640x480": dict(mesh=("unit_sphere", 32, 16), width=640, height=480, strata=1, ri=1.3, max_crossings=2),
    "cfg2_ellipsoid64x32_1080p": dict(mesh=("ellipsoid", 64, 32, (1.0, 4.0, 2.0)), width=1920, height=1080, strata=1, ri=1.3, max_crossings=2),
    "cfg3_robot_1080p": dict(mesh=("stl", ROBOT_STL), width=1920, height=1080, strata=1, ri=1.5, max_crossings=8),
    "cfg4_freeform181x91_4k_16spp": dict(mesh=("ellipsoid", 181, 91, (1.0, 4.0, 2.0)), width=3840, height=2160, strata=4, ri=1.3, max_crossings=2),
}
WORKLOADS["cfg5_freeform181x91_4k_16spp_sharded"] = WORKLOADS["cfg4_freeform181x91_4k_16spp"]
# illumination mode: an LED-sized emitter 1.5 lens depths in front of the lens, a target behind it
ILLUMINATION = {
    "illum_ellipsoid64x32_4m": dict(mesh=("ellipsoid", 64, 32, (1.0, 4.0, 2.0)), points=(8, 8), rays_per_point=65536, belts=16, bins=(256, 128),
                                    ri=1.3, max_crossings=2),
}
HEADLINE = "cfg4_freeform181x91_4k_16spp"
STRIPE_ROWS = 8       # one tile row of the 4K frame; 270 stripes: a stripe and its mirror image never fall on the same rank at N = 8


def build_mesh(mesh_class, recipe):
    """recipe -> standardized mesh of `mesh_class` (bezier_b200.Mesh or, for the checker, bzref.RefMesh)"""
    m = mesh_class()
    kind = recipe[0]
    if kind == "unit_sphere":
        m.make_unit_sphere(recipe[1], recipe[2])
    elif kind == "ellipsoid":
        m.make_ellipsoid(recipe[1], recipe[2], recipe[3])
    elif kind == "stl":
        m.read_stl(recipe[1])
    else:
        raise KeyError(kind)
    m.standardize_vertices()
    m.standardize_normals()
    return m


def frame(lo, hi, width, height, strata=1, max_crossings=2):
    """bounding box -> (camera kwargs, scene kwargs) for make_camera / make_scene of either binding"""
    cy, cz = (float(lo[1]) + float(hi[1])) / 2.0, (float(lo[2]) + float(hi[2])) / 2.0
    ey, ez = (float(hi[1]) - float(lo[1])) / 2.0, (float(hi[2]) - float(lo[2])) / 2.0
    s = max(ez, ey * height / width)
    distance = 6.5 * s
    centre_x = (float(lo[0]) + float(hi[0])) / 2.0
    half_h = s / 0.6 / distance
    camera = dict(strata=strata, eye=(centre_x - distance, cy, cz), half_w=half_h * width / height, half_h=half_h)
    scene = dict(bg_x=float(hi[0]) + 6.0 * s, cell=0.5 * s, max_crossings=max_crossings)
    return camera, scene


def setup(name, mesh_class):
    """-> (spec, mesh, camera kwargs, scene kwargs)"""
    spec = WORKLOADS[name]
    mesh = build_mesh(mesh_class, spec["mesh"])
    tri = mesh.triangles().reshape(-1, 3)
    camera, scene = frame(tri.min(0), tri.max(0), spec["width"], spec["height"], spec["strata"], spec["max_crossings"])
    return spec, mesh, camera, scene


def setup_illumination(name, mesh_class):
    """-> (spec, mesh, emitter kwargs, target kwargs) for bezier_b200.make_emitter / make_target"""
    spec = ILLUMINATION[name]
    mesh = build_mesh(mesh_class, spec["mesh"])
    tri = mesh.triangles().reshape(-1, 3)
    lo, hi = tri.min(0), tri.max(0)
    cy, cz = (float(lo[1]) + float(hi[1])) / 2.0, (float(lo[2]) + float(hi[2])) / 2.0
    ey, ez = (float(hi[1]) - float(lo[1])) / 2.0, (float(hi[2]) - float(lo[2])) / 2.0
    depth = float(hi[0]) - float(lo[0])
    side = 0.05 * min(ey, ez)
    emitter = dict(origin=(float(lo[0]) - 1.5 * depth, cy - side / 2.0, cz - side / 2.0), edge_u=(0.0, side, 0.0), edge_v=(0.0, 0.0, side),
                   points=spec["points"], rays_per_point=spec["rays_per_point"], belts=spec["belts"], seed=2026)
    target = dict(origin=(float(hi[0]) + 4.0 * depth, cy - 3.0 * ey, cz - 3.0 * ez), edge_u=(0.0, 6.0 * ey, 0.0), edge_v=(0.0, 0.0, 6.0 * ez), bins=spec["bins"])
    return spec, mesh, emitter, target
