"""`-m gpu`: the parity tests at SURVEY.md section 4's sample sizes and on the paths the small tests do not reach.

  * >= 10^6 rays per configuration (cfg1 sphere 32x16, cfg3 robot.stl; >= 2*10^5 for cfg2 ellipsoid 64x32) through
    bz_intersect_batch and two successive bz_refract_batch crossings, against the compiled reference (oracle/_ref)
  * the 98 826-patch lens of the headline with INCOHERENT rays: random origins (half of them inside the body), origins
    beyond the hierarchy's origin limit, rays aimed at the ill-conditioned (*wide*) patches -- packet walk and per-lane walk
  * a lens refined by splitThickBezierTriangles -> re-standardise (reference/test.cpp:435-440), through the kernels
  * the two render pipelines (persistent kernel, wavefront kernels) on whole frames
Every comparison is BIT-EXACT (NaN matches NaN).  Rays whose two best accepted hits lie within 1e-3 of the shape size of
each other (patch-edge duplicates, SURVEY.md section 4 iii) are compared like all others; their number is printed.
"""
import ctypes as C
import time

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu

THREADS = 16


@pytest.fixture(scope="module")
def bz():
    import bezier_b200
    assert bezier_b200.device_count() > 0, "no CUDA device: the product has no CPU path"
    return bezier_b200


def _size(mesh):
    lo, hi = common.bbox(mesh)
    return lo, hi, float(np.linalg.norm(hi - lo))


@pytest.mark.parametrize("name,rays,ri", [("sphere32x16", 1000000, 1.3), ("robot", 1000000, 1.5), ("ellipsoid64x32", 200000, 1.3)],
                         ids=["cfg1_sphere32x16-1000000rays", "cfg3_robot-1000000rays", "cfg2_ellipsoid64x32-200000rays"])
def test_intersect_and_two_crossings_at_survey_sample_size(bz, oracle_built, name, rays, ri):
    import bzoracle
    import bzref
    mesh, bezier, patches = common.ref_config(name)
    lo, hi, size = _size(mesh)
    lens = bz.BezierLens(ri, patches)
    t0 = time.time()
    # BezierMesh::intersect: random rays aimed into the bounding box, half of them starting inside it
    s, d = common.random_rays(lo, hi, rays, seed=31)
    ref = bezier.intersect(s, d, threads=THREADS)
    got = lens.intersect(s, d)
    common.assert_same_hits(ref, got, name)
    hits = int((ref["what"] == 4).sum())
    ambiguous = int(bzoracle.edge_ambiguous(patches, s[: rays // 4], d[: rays // 4], size, threads=THREADS).sum())
    # BezierLens::refract, entry then exit
    ref_lens = bzref.RefLens(ri, bezier)
    s, d = common.random_rays(lo, hi, rays, seed=32, inside_fraction=0.0)
    expected = np.full(rays, 1, np.uint32)
    crossed = []
    for crossing in range(2):
        rs, rd, rst = ref_lens.refract(s, d, expected, threads=THREADS)
        gs, gd, gst, _ = lens.refract(s, d, expected)
        assert np.array_equal(rst, gst), "crossing %d: status differs on %d rays" % (crossing, int((rst != gst).sum()))
        assert common.same_bits(rs, gs) and common.same_bits(rd, gd), "crossing %d" % crossing
        crossed.append(int((rst != 0).sum()))
        s, d = rs, rd
        expected = np.where(rst == 0, 0, 3 - expected).astype(np.uint32)
    print("\n%s: %d rays, %d hits, %d of the first %d rays edge-ambiguous (compared like the rest), refracted %d then %d, %.0f s"
          % (name, rays, hits, ambiguous, rays // 4, crossed[0], crossed[1], time.time() - t0))
    assert hits > rays // 4 and crossed[0] > rays // 4 and crossed[1] > rays // 20


_freeform = {}


def _freeform_rays(shim):
    """incoherent rays for the 98 826-patch lens + the reference's answers (computed once for both walk modes)"""
    if _freeform:
        return _freeform
    import bzoracle
    mesh, bezier, patches = common.ref_config("ellipsoid181x91")
    lo, hi, size = _size(mesh)
    rng = np.random.default_rng(77)
    sets = []
    # (1) random rays into the bounding box, half of them starting inside the body
    sets.append(common.random_rays(lo, hi, 100000, seed=41))
    # (2) origins far beyond the hierarchy's origin limit
    n = 20000
    target = (rng.random((n, 3)) - 0.5) * (hi - lo) + (lo + hi) / 2
    o = rng.normal(size=(n, 3))
    o = o / np.linalg.norm(o, axis=1, keepdims=True) * rng.choice([300.0, 3000.0, 30000.0], size=(n, 1))
    d = target - o
    sets.append((o.astype(np.float32), (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)))
    # (3) rays aimed at the ill-conditioned patches, from inside and outside the origin limit (the GPU twin of
    #     tests/test_cpu_shim.py::test_wide_patches_stay_exact)
    wide = np.zeros(4096, np.uint32)
    limit = C.c_float(0.0)
    shim.shim_bvh_wide.restype = C.c_uint32
    count = shim.shim_bvh_wide(patches.ctypes.data_as(C.c_void_p), C.c_uint32(patches.shape[0]), wide.ctypes.data_as(C.c_void_p), C.c_uint32(wide.size), C.byref(limit))
    assert 0 < count <= wide.size and limit.value > 0
    wide = wide[:count]
    n = 10000
    pick = wide[rng.integers(0, count, n)]
    cp = patches[pick][:, 20:50].reshape(n, 10, 3)
    w = rng.dirichlet((1.0, 1.0, 1.0), n).astype(np.float32) * 1.4 - 0.133
    target = (cp[:, 0] * w[:, :1] + cp[:, 1] * w[:, 1:2] + cp[:, 2] * w[:, 2:3]).astype(np.float64)
    o = rng.normal(size=(n, 3))
    o /= np.linalg.norm(o, axis=1, keepdims=True)
    o = target + o * np.where(np.arange(n) % 2 == 0, 12.0, limit.value * 1.5)[:, None]
    d = target - o
    sets.append((o.astype(np.float32), (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)))
    s = np.concatenate([a for a, _ in sets])
    d = np.concatenate([b for _, b in sets])
    t0 = time.time()
    ref = bzoracle.intersect(patches, s, d, threads=THREADS)
    # the plain-C oracle is itself pinned to the compiled reference (tests/test_golden.py); spot-check it here on this lens
    probe = np.arange(0, s.shape[0], 97)
    real = bezier.intersect(s[probe], d[probe], threads=THREADS)
    for key in ("what", "patch", "distance", "point", "normal", "cos", "bary"):
        assert common.same_bits(real[key], ref[key][probe]), key
    _freeform.update(patches=patches, s=s, d=d, ref=ref, wide=wide, size=size, seconds=time.time() - t0, counts=[a.shape[0] for a, _ in sets])
    return _freeform


@pytest.mark.parametrize("mode", ["culled", "lanes"])
def test_freeform_lens_incoherent_rays(bz, oracle_built, shim, mode):
    f = _freeform_rays(shim)
    lens = bz.BezierLens(1.3, f["patches"]).set_mode(bz.MODE_CULLED if mode == "culled" else bz.MODE_LANES)
    got = lens.intersect(f["s"], f["d"])
    common.assert_same_hits(f["ref"], got, "freeform/" + mode)
    hits = f["ref"]["what"] == 4
    on_wide = int(np.isin(f["ref"]["patch"][hits], f["wide"]).sum())
    print("\nfreeform lens, %s: %s rays (random / far origins / aimed at the %d wide patches), %d hits, %d of them on wide patches, reference %.0f s"
          % (mode, f["counts"], f["wide"].size, int(hits.sum()), on_wide, f["seconds"]))
    assert int(hits.sum()) > 50000 and on_wide > 100
    # two refract crossings of the random set through both walks agree with each other bit for bit (the reference side of
    # refract is covered on this lens by the lattice of tests/test_gpu_full_size.py)
    if mode == "lanes":
        other = bz.BezierLens(1.3, f["patches"])
        n = f["counts"][0]
        expected = np.full(n, 1, np.uint32)
        a, b = lens.refract(f["s"][:n], f["d"][:n], expected), other.refract(f["s"][:n], f["d"][:n], expected)
        for x, y in zip(a, b):
            assert common.same_bits(x, y)


def test_many_wide_patches_far_origin_packet_seed_list(bz, oracle_built):
    """more than 224 records outside the hierarchy for one packet: an off-centre, strongly stretched mesh has hundreds of
    ill-conditioned patches, and a packet of rays from beyond the origin limit seeds its candidate list with all of them
    (7 per entry, more than 32 entries -- the list is filled by a strided loop, csrc/bz_trace.cu).  Packet walk == per-lane
    walk == oracle."""
    import bzoracle
    import bzref
    m = bzref.RefMesh().make_ellipsoid(40, 20, (1.0, 400.0, 0.02))
    m.translate(3.0, 500.0, 1.0)
    m.standardize_vertices().standardize_normals()
    patches = bzref.RefBezierMesh(m).patches()
    lo, hi, size = _size(m)
    info = {}
    lenses = {}
    for mode in (bz.MODE_CULLED, bz.MODE_LANES):
        lenses[mode] = bz.BezierLens(1.3, patches).set_mode(mode)
    # a coherent fan from far away (packets stay packets) and random far rays
    rng = np.random.default_rng(5)
    n = 32 * 400
    centre = (lo + hi) / 2
    eye = centre + np.array([-1.0, 0.3, 0.2]) * size * 400.0
    target = centre + (rng.random((n, 3)) - 0.5) * (hi - lo) * 0.05 + np.repeat((rng.random((n // 32, 3)) - 0.5) * (hi - lo), 32, axis=0)
    d = target - eye
    s = np.tile(eye, (n, 1)).astype(np.float32)
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    ref = bzoracle.intersect(patches, s, d, threads=THREADS)
    for mode, lens in lenses.items():
        common.assert_same_hits(ref, lens.intersect(s, d), "wide-seed/%d" % mode)
    print("\nstretched off-centre ellipsoid: %d patches, %d far rays, %d hits" % (patches.shape[0], n, int((ref["what"] == 4).sum())))


def test_split_thick_lens_through_the_kernels(bz, oracle_built):
    """splitThickBezierTriangles -> standardize -> BezierMesh, twice (reference/test.cpp:435-440): the refined table goes through
    intersect, two refract crossings and a render, against the compiled reference built the same way"""
    import bzref
    ref_mesh = bzref.RefMesh().make_ellipsoid(15, 5, (1.0, 4.0, 2.0))
    own_mesh = bz.Mesh().make_ellipsoid(15, 5, (1.0, 4.0, 2.0))
    for m in (ref_mesh, own_mesh):
        m.standardize_vertices().standardize_normals()
    ref_bezier, own_bezier = bzref.RefBezierMesh(ref_mesh), bz.BezierMesh(own_mesh)
    before = ref_bezier.size()
    for _ in range(2):
        ref_mesh, own_mesh = ref_bezier.split_thick(), own_bezier.split_thick()
        for m in (ref_mesh, own_mesh):
            m.standardize_vertices().standardize_normals()
        ref_bezier, own_bezier = bzref.RefBezierMesh(ref_mesh), bz.BezierMesh(own_mesh)
    assert ref_bezier.size() == own_bezier.size() > before
    assert np.array_equal(ref_bezier.patches().view(np.uint32), own_bezier.patches().view(np.uint32))
    lens = bz.BezierLens(1.3, own_bezier)                    # the product's own host preprocessing feeds the kernels
    lo, hi, size = _size(ref_mesh)
    s, d = common.random_rays(lo, hi, 60000, seed=51)
    common.assert_same_hits(ref_bezier.intersect(s, d, threads=THREADS), lens.intersect(s, d), "split-thick")
    ref_lens = bzref.RefLens(1.3, ref_bezier)
    s, d = common.random_rays(lo, hi, 60000, seed=52, inside_fraction=0.0)
    expected = np.full(s.shape[0], 1, np.uint32)
    for crossing in range(2):
        rs, rd, rst = ref_lens.refract(s, d, expected, threads=THREADS)
        gs, gd, gst, _ = lens.refract(s, d, expected)
        assert np.array_equal(rst, gst) and common.same_bits(rs, gs) and common.same_bits(rd, gd)
        s, d = rs, rd
        expected = np.where(rst == 0, 0, 3 - expected).astype(np.uint32)
    w, h = 200, 150
    half = float(max(hi[1] - lo[1], hi[2] - lo[2])) * 0.5 / 0.6 / 12.0
    eye = (float(lo[0]) - 12.0, float(lo[1] + hi[1]) / 2, float(lo[2] + hi[2]) / 2)
    rcam, rscene = bzref.make_camera(w, h, strata=2, eye=eye, half_w=half), bzref.make_scene(bg_x=float(hi[0]) + 12.0)
    want, want_stats, _ = ref_lens.render(rcam, rscene, threads=THREADS)
    cam, scene = bz.make_camera(w, h, strata=2, eye=eye, half_w=half), bz.make_scene(bg_x=float(hi[0]) + 12.0)
    for pipeline in (bz.PIPELINE_PERSISTENT, bz.PIPELINE_WAVEFRONT):
        rgba, stats = lens.set_pipeline(pipeline).render(cam, scene)
        assert np.array_equal(rgba, want), "%d pixels differ" % int((rgba != want).any(-1).sum())
        assert stats.entered == want_stats.entered and stats.exited == want_stats.exited and stats.lost == want_stats.lost
    print("\nsplit-thick lens: %d -> %d patches, %d rays entered in the render" % (before, ref_bezier.size(), want_stats.entered))
    assert want_stats.entered > 1000


@pytest.mark.parametrize("name", ["cfg2_ellipsoid64x32_1080p", "cfg3_robot_1080p", "cfg4_freeform181x91_4k_16spp"])
def test_pipelines_give_the_same_frame(bz, name):
    """whole frames at full size: the wavefront kernels (several chunk sizes) and the persistent kernel write the same bytes
    and count the same work"""
    from bezier_b200 import workloads
    spec, mesh, cam_kw, scene_kw = workloads.setup(name, bz.Mesh)
    lens = bz.BezierLens(spec["ri"], bz.BezierMesh(mesh))
    cam, scene = bz.make_camera(spec["width"], spec["height"], **cam_kw), bz.make_scene(**scene_kw)
    want, want_stats = lens.set_pipeline(bz.PIPELINE_PERSISTENT).render(cam, scene)
    for chunk in (1 << 22, 1 << 18, 3000):
        rgba, stats = lens.set_pipeline(bz.PIPELINE_WAVEFRONT).set_wave_chunk(chunk).render(cam, scene)
        assert np.array_equal(rgba, want), "chunk %d: %d pixels differ" % (chunk, int((rgba != want).any(-1).sum()))
        assert stats.as_dict() == want_stats.as_dict(), chunk
        lens.check()


@pytest.mark.parametrize("strata,width,height", [(3, 333, 217), (5, 150, 101), (7, 64, 40), (1, 1001, 600)])
def test_wavefront_ragged_tiles_and_parts(bz, strata, width, height):
    """samples per pixel that make the packets of a tile ragged (9 spp: 18 packets = parts of 8 + 8 + 2; 25 spp: 13 packets, the last
    half full; 49 spp: one pixel per ... tile sides 4 and 2) and frame sizes that cut the border tiles: the walk kernel's hand-out in
    parts of tiles gives the persistent kernel's frame and counts"""
    from bezier_b200 import workloads
    spec, mesh, cam_kw, scene_kw = workloads.setup("cfg2_ellipsoid64x32_1080p", bz.Mesh)
    lens = bz.BezierLens(spec["ri"], bz.BezierMesh(mesh))
    cam, scene = bz.make_camera(width, height, **dict(cam_kw, strata=strata)), bz.make_scene(**scene_kw)
    want, want_stats = lens.set_pipeline(bz.PIPELINE_PERSISTENT).render(cam, scene)
    assert want_stats.entered > 1000
    for chunk in (1 << 26, 50000, 2000):
        rgba, stats = lens.set_pipeline(bz.PIPELINE_WAVEFRONT).set_wave_chunk(chunk).render(cam, scene)
        assert np.array_equal(rgba, want), "chunk %d: %d pixels differ" % (chunk, int((rgba != want).any(-1).sum()))
        assert stats.as_dict() == want_stats.as_dict(), chunk
        lens.check()


def test_one_lens_many_streams_and_threads(bz, oracle_built):
    """BezierLens::refract is const and re-entrant (reference/bezierLens.h:27): ONE lens handle serves two cameras rendered
    concurrently on two streams (both pipelines) while host threads refract and render through the *_host calls; every result
    matches the compiled reference"""
    import threading
    import torch
    import bzref
    mesh, bezier, patches = common.ref_config("sphere32x16")
    lo, hi, size = _size(mesh)
    ref_lens = bzref.RefLens(1.3, bezier)
    w, h = 192, 144
    cams = [dict(eye=(-12.0, 0.0, 0.0), half_w=0.14), dict(eye=(-9.0, 2.0, -1.5), forward=(0.9, -0.2, 0.15), right=(0.2, 1.0, 0.0), up=(-0.15, 0.0, 1.0), half_w=0.2)]
    want = []
    for kw in cams:
        rgba, _, _ = ref_lens.render(bzref.make_camera(w, h, strata=2, **kw), bzref.make_scene(bg_x=14.0), threads=THREADS)
        want.append(rgba)
    s, d = common.random_rays(lo, hi, 50000, seed=61, inside_fraction=0.0)
    expected = np.full(s.shape[0], 1, np.uint32)
    want_rays = ref_lens.refract(s, d, expected, threads=THREADS)
    lens = bz.BezierLens(1.3, patches)
    scene = bz.make_scene(bg_x=14.0)
    for pipeline in (bz.PIPELINE_PERSISTENT, bz.PIPELINE_WAVEFRONT):
        lens.set_pipeline(pipeline)
        streams = [torch.cuda.Stream(), torch.cuda.Stream()]
        frames = [torch.zeros((h, w, 4), dtype=torch.uint8, device="cuda") for _ in cams]
        errors = []

        def host_worker(kind):
            try:
                for _ in range(4):
                    if kind == "refract":
                        gs, gd, gst, _ = lens.refract(s, d, expected)
                        assert np.array_equal(gst, want_rays[2]) and common.same_bits(gs, want_rays[0]) and common.same_bits(gd, want_rays[1])
                    else:
                        rgba, _ = lens.render(bz.make_camera(w, h, strata=2, **cams[1]), scene)
                        assert np.array_equal(rgba, want[1])
            except Exception as e:      # noqa: BLE001 -- reported by the main thread
                errors.append(e)

        threads = [threading.Thread(target=host_worker, args=(kind,)) for kind in ("refract", "render", "refract")]
        for t in threads:
            t.start()
        for _ in range(6):                           # frames of the two cameras interleaved, no synchronisation in between
            for k, kw in enumerate(cams):
                lens.render_device(bz.make_camera(w, h, strata=2, **kw), scene, frames[k], stream=streams[k].cuda_stream)
        for t in threads:
            t.join()
        torch.cuda.synchronize()
        assert not errors, errors
        for k in range(2):
            assert np.array_equal(frames[k].cpu().numpy(), want[k]), "pipeline %d camera %d" % (pipeline, k)
        for st in streams:
            lens.release_stream(st.cuda_stream)
        lens.check()
