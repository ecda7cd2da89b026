"""`-m gpu` parity tests: the CUDA path through the C ABI against the oracle.

Tolerance: the parity kernels are built with -fmad=false and IEEE div/sqrt, so every comparison with the
oracle (oracle/_ref = the compiled reference + the documented hoist, and oracle/bz_oracle.c) is
BIT-EXACT -- stricter than north_star's 1e-5 relative / 1 LSB.  The render cases run through both
pipelines (persistent kernel, wavefront kernels).  The fast build (BZ_MODE_FAST) has its own file and
its own, looser, written tolerances: tests/test_gpu_fast.py.
"""
import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu

CONFIGS = ["sphere16x8", "sphere32x16", "ellipsoid64x32", "robot"]
RI = {"robot": 1.5}


@pytest.fixture(scope="module")
def bz():
    import bezier_b200
    assert bezier_b200.device_count() > 0, "no CUDA device: the product has no CPU path"
    return bezier_b200


@pytest.mark.parametrize("name", CONFIGS)
@pytest.mark.parametrize("mode", ["brute", "culled", "lanes"])
def test_intersect_batch_bit_exact(bz, oracle_built, name, mode):
    import bzoracle
    mesh, bezier, patches = common.ref_config(name)
    lo, hi = common.bbox(mesh)
    s, d = common.random_rays(lo, hi, 20000, seed=7)
    ref = bzoracle.intersect(patches, s, d, threads=8)
    lens = bz.BezierLens(RI.get(name, 1.3), patches).set_mode({"brute": bz.MODE_BRUTE, "culled": bz.MODE_CULLED, "lanes": bz.MODE_LANES}[mode])
    got = lens.intersect(s, d)
    assert int((ref["what"] == 4).sum()) > 1000
    common.assert_same_hits(ref, got, "%s/%s" % (name, mode))


@pytest.mark.parametrize("name", CONFIGS)
def test_refract_two_crossings_bit_exact_vs_reference(bz, oracle_built, name):
    """entry + exit refraction through the REAL reference classes (oracle/_ref)"""
    import bzref
    mesh, bezier, patches = common.ref_config(name)
    ri = RI.get(name, 1.3)
    ref_lens = bzref.RefLens(ri, bezier)
    lens = bz.BezierLens(ri, patches)
    lo, hi = common.bbox(mesh)
    s, d = common.random_rays(lo, hi, 8000, seed=11, inside_fraction=0.0)
    expected = np.full(s.shape[0], 1, np.uint32)
    for crossing in range(2):
        rs, rd, rst = ref_lens.refract(s, d, expected, threads=8)
        gs, gd, gst, gpatch = lens.refract(s, d, expected)
        assert np.array_equal(rst, gst), "status differs on %d rays" % int((rst != gst).sum())
        assert common.same_bits(rs, gs) and common.same_bits(rd, gd)
        assert int((rst != 0).sum()) > 500
        s, d = rs, rd
        expected = np.where(rst == 0, 0, 3 - expected).astype(np.uint32)   # dead rays expect cNone: they stay dead


def test_refract_edge_cases(bz, oracle_built):
    mesh, bezier, patches = common.ref_config("sphere16x8")
    lens = bz.BezierLens(1.3, patches)
    # empty batch
    gs, gd, gst, gp = lens.refract(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32), np.zeros(0, np.uint32))
    assert gs.shape == (0, 3) and gst.shape == (0,)
    # a single ray, a ragged count (not a multiple of the pool size), rays that miss, wrong expected side
    s = np.array([[-10, 0.1, 0.2], [-10, 5.0, 5.0], [-10, 0.1, 0.2]], np.float32)
    d = np.array([[1, 0, 0], [1, 0, 0], [1, 0, 0]], np.float32)
    gs, gd, gst, gp = lens.refract(s, d, np.array([1, 1, 2], np.uint32))
    assert list(gst) == [1, 0, 0]
    assert gp[0] == 350 and gp[1] == bz.NO_PATCH and gp[2] == 350      # SURVEY.md section 4 anchor: entry patch 350
    assert np.array_equal(gs[1], s[1]) and np.array_equal(gd[2], d[2])   # cNone leaves the ray as it came
    # empty lens: every ray passes untouched
    empty = bz.BezierLens(1.3, np.zeros((0, 64), np.float32))
    gs, gd, gst, gp = empty.refract(s, d, 1)
    assert list(gst) == [0, 0, 0]


@pytest.mark.parametrize("name,w,h,strata,crossings", [("sphere16x8", 160, 120, 1, 2), ("sphere16x8", 64, 48, 4, 2), ("robot", 96, 64, 2, 8),
                                                      ("sphere16x8", 131, 77, 3, 2), ("ellipsoid64x32", 67, 45, 1, 2),   # ragged edge tiles, spp not a power of two
                                                      ("sphere16x8", 50, 37, 5, 2), ("sphere16x8", 9, 5, 12, 2)])         # tile rays not a multiple of 32; a 2x2-pixel tile
def test_render_pixel_exact_vs_reference(bz, oracle_built, name, w, h, strata, crossings):
    import bzref
    mesh, bezier, patches = common.ref_config(name)
    ri = RI.get(name, 1.3)
    lo, hi = common.bbox(mesh)
    half = float(max(hi[1] - lo[1], hi[2] - lo[2])) * 0.5 / 0.6 / (12.0 + float(lo[0]))
    eye = (float(lo[0]) - 12.0, float(lo[1] + hi[1]) / 2, float(lo[2] + hi[2]) / 2)
    cam_args = dict(strata=strata, eye=eye, half_w=half)
    rcam, rscene = bzref.make_camera(w, h, **cam_args), bzref.make_scene(bg_x=float(hi[0]) + 12.0, max_crossings=crossings)
    ref_rgba, ref_stats, rays = bzref.RefLens(ri, bezier).render(rcam, rscene, threads=8)
    cam, scene = bz.make_camera(w, h, **cam_args), bz.make_scene(bg_x=float(hi[0]) + 12.0, max_crossings=crossings)
    counts = {}
    for mode, pipeline in ((bz.MODE_CULLED, bz.PIPELINE_PERSISTENT), (bz.MODE_CULLED, bz.PIPELINE_WAVEFRONT), (bz.MODE_LANES, bz.PIPELINE_AUTO), (bz.MODE_BRUTE, bz.PIPELINE_AUTO)):
        rgba, stats = bz.BezierLens(ri, patches).set_mode(mode).set_pipeline(pipeline).render(cam, scene)
        assert np.array_equal(rgba, ref_rgba), "mode %d pipeline %d: %d of %d pixels differ" % (mode, pipeline, int((rgba != ref_rgba).any(-1).sum()), w * h)
        got, want = stats.as_dict(), ref_stats.as_dict()
        for key in ("primary_rays", "refract_calls", "entered", "exited", "lost"):
            assert got[key] == want[key], key
        if mode == bz.MODE_CULLED:      # camera rays go through the warp-cooperative walk (a coarse image of a fine mesh falls back in part)
            assert stats.packet_walks > 0 and stats.lane_walk_calls < stats.refract_calls
            counts[pipeline] = got
    # the two pipelines do the same work: same stage-A tests, stage-B evaluations, Snell blocks, walks
    for key in ("stage_a", "stage_b", "snell", "packet_walks", "lane_walk_calls", "packet_nodes", "node_visits"):
        assert counts[bz.PIPELINE_PERSISTENT][key] == counts[bz.PIPELINE_WAVEFRONT][key], key
    stats = bz.BezierLens(ri, patches).set_mode(bz.MODE_LANES).render(cam, scene)[1]
    assert stats.packet_walks == 0 and stats.lane_walk_calls == stats.refract_calls
    assert ref_stats.entered > 100


@pytest.mark.parametrize("label,cam_args,plans", [
    ("oblique pinhole, skewed axes", dict(eye=(-9.0, 3.0, -2.0), forward=(0.9, -0.3, 0.2), right=(0.25, 1.0, 0.05), up=(-0.1, 0.1, 1.1), half_w=0.35), True),
    ("lens in a corner of the frame", dict(eye=(-12.0, 2.5, 1.5), half_w=0.4), True),
    ("lens off the frame", dict(eye=(-12.0, 0.0, 0.0), forward=(1.0, 1.5, 0.0), right=(-1.5, 1.0, 0.0), half_w=0.2), True),
    ("orthographic, oblique", dict(eye=(-8.0, 0.5, 0.25), forward=(1.0, -0.05, 0.1), right=(0.05, 1.0, 0.0), up=(-0.1, 0.0, 1.0), half_w=2.5, ortho=True), True),
    ("eye inside the body", dict(eye=(0.0, 0.1, 0.0), half_w=0.8), False),
    ("lens behind the eye", dict(eye=(6.0, 0.0, 0.0), half_w=0.5), True),
])
def test_render_sky_tiles_pixel_exact(bz, oracle_built, label, cam_args, plans):
    """tiles that cannot meet the lens are shaded by bz_sky_kernel: same pixels and counters as the reference and as the
    arm without a tile plan (BZ_MODE_LANES), for cameras that exercise the projection of the hierarchy's boxes"""
    import bzref
    mesh, bezier, patches = common.ref_config("ellipsoid64x32")
    w, h, strata = 192, 128, 2
    rcam, rscene = bzref.make_camera(w, h, strata=strata, **cam_args), bzref.make_scene(bg_x=30.0, max_crossings=2)
    ref_rgba, ref_stats, _ = bzref.RefLens(1.3, bezier).render(rcam, rscene, threads=8)
    cam, scene = bz.make_camera(w, h, strata=strata, **cam_args), bz.make_scene(bg_x=30.0, max_crossings=2)
    for mode, pipeline in ((bz.MODE_CULLED, bz.PIPELINE_PERSISTENT), (bz.MODE_CULLED, bz.PIPELINE_WAVEFRONT), (bz.MODE_LANES, bz.PIPELINE_AUTO)):
        lens = bz.BezierLens(1.3, patches).set_mode(mode).set_pipeline(pipeline)
        for again in range(2):                                  # the second frame reuses the kept tile plan
            rgba, stats = lens.render(cam, scene)
            assert np.array_equal(rgba, ref_rgba), "%s: %d of %d pixels differ" % (label, int((rgba != ref_rgba).any(-1).sum()), w * h)
            got, want = stats.as_dict(), ref_stats.as_dict()
            for key in ("primary_rays", "refract_calls", "entered", "exited", "lost"):
                assert got[key] == want[key], (label, key)
        launches = lens.last_kernel_launches()
        # one persistent kernel, or walk + eval + resolve per crossing (2) and chunk (two, one per lane, unless the lens is on one tile)
        traces = (6, 12) if pipeline == bz.PIPELINE_WAVEFRONT else (1,)
        if mode == bz.MODE_CULLED and plans:
            if ref_stats.entered > 0:
                assert launches - 2 in traces, label                                     # sky + trace + resolve
            else:
                assert launches == 2, label                                              # no trace when nothing can meet the lens
        else:
            assert launches - 1 in traces, label                                         # no plan: every tile is traced
        # a band of the frame and a moved camera make a new plan
        if mode == bz.MODE_CULLED:
            band, _ = lens.render(cam, scene, 37, 101)
            assert np.array_equal(band, ref_rgba[37:101]), label


def test_tile_plan_against_the_arm_without_one_random_cameras(bz, oracle_built):
    """60 random cameras (pinhole and orthographic, any orientation, skewed axes, near and far, lens partly or wholly off the
    frame): the planned frame (sky tiles + lens tiles) equals the frame of BZ_MODE_LANES, which traces every tile"""
    mesh, bezier, patches = common.ref_config("sphere32x16")
    planned, plain = bz.BezierLens(1.3, patches), bz.BezierLens(1.3, patches).set_mode(bz.MODE_LANES)
    rng = np.random.default_rng(2024)
    scene = bz.make_scene(bg_x=40.0, max_crossings=2)
    with_sky = 0
    for case in range(60):
        eye = rng.normal(size=3)
        eye = eye / np.linalg.norm(eye) * rng.uniform(1.3, 30.0)
        aim = rng.uniform(-1.5, 1.5, size=3) * (0.0 if case % 3 == 0 else 1.0)           # every third camera looks at the centre
        forward = aim - eye
        forward /= np.linalg.norm(forward)
        helper = np.array([0.0, 0.0, 1.0]) if abs(forward[2]) < 0.9 else np.array([0.0, 1.0, 0.0])
        right = np.cross(forward, helper)
        right /= np.linalg.norm(right)
        up = np.cross(right, forward)
        right = right * rng.uniform(0.5, 2.0) + forward * rng.uniform(-0.2, 0.2)            # skewed, unnormalised axes
        up = up * rng.uniform(0.5, 2.0) + right * rng.uniform(-0.2, 0.2)
        ortho = case % 4 == 3
        half_w = rng.uniform(0.5, 3.0) if ortho else rng.uniform(0.05, 1.5)
        cam = bz.make_camera(160, 96, strata=1 + case % 3, eye=tuple(eye), forward=tuple(forward), right=tuple(right), up=tuple(up), half_w=float(half_w), ortho=ortho)
        a, sa = planned.render(cam, scene)
        with_sky += planned.last_kernel_launches() == 3 or (planned.last_kernel_launches() == 2 and sa.packet_walks == 0 and sa.lane_walk_calls == 0)
        b, sb = plain.render(cam, scene)
        assert np.array_equal(a, b), "camera %d: %d pixels differ" % (case, int((a != b).any(-1).sum()))
        for key in ("primary_rays", "refract_calls", "entered", "exited", "lost"):
            assert getattr(sa, key) == getattr(sb, key), (case, key)
    assert with_sky >= 30          # most of these cameras get a plan


def test_render_bands_and_stripes_compose(bz, oracle_built):
    """row bands and round-robin stripes reproduce the full frame byte for byte (multi-GPU sharding, one device)"""
    import torch
    mesh, bezier, patches = common.ref_config("sphere16x8")
    lens = bz.BezierLens(1.3, patches)
    cam, scene = bz.make_camera(96, 70, strata=2, eye=(-12.0, 0.0, 0.0), half_w=0.14), bz.make_scene()
    full, _ = lens.render(cam, scene)
    top, _ = lens.render(cam, scene, 0, 33)
    bottom, _ = lens.render(cam, scene, 33, 70)
    assert np.array_equal(np.concatenate([top, bottom]), full)
    ranks, stripe = 4, 8
    reserve = max(bz.striped_rows(cam.height, stripe, r, ranks) for r in range(ranks))
    bands = torch.zeros((ranks, reserve, cam.width, 4), dtype=torch.uint8, device="cuda")       # what the all-gather leaves on a rank
    image = np.zeros_like(full)
    for rank in range(ranks):
        rows = bz.striped_rows(cam.height, stripe, rank, ranks)
        lens.render_striped_device(cam, scene, stripe, rank, ranks, bands[rank])
        torch.cuda.synchronize()
        local = bands[rank].cpu().numpy()
        r = 0
        for first in range(rank * stripe, cam.height, ranks * stripe):
            k = min(stripe, cam.height - first)
            image[first:first + k] = local[r:r + k]
            r += k
        assert r == rows
    assert np.array_equal(image, full)
    assembled = torch.zeros((cam.height, cam.width, 4), dtype=torch.uint8, device="cuda")
    bz.assemble_stripes_device(bands, reserve, cam.width, cam.height, stripe, ranks, assembled)  # bz_assemble_stripes
    torch.cuda.synchronize()
    assert np.array_equal(assembled.cpu().numpy(), full)


def test_generate_rays_bit_exact(bz, oracle_built):
    import bzref
    import torch
    for ortho in (False, True):
        args = dict(strata=3, eye=(-12.0, 0.25, -0.5), half_w=0.2, ortho=ortho)
        cam, rcam = bz.make_camera(50, 30, **args), bzref.make_camera(50, 30, **args)
        rs, rd = bzref.generate_rays(rcam, 5, 22)
        s = torch.zeros(rs.shape, dtype=torch.float32, device="cuda")
        d = torch.zeros_like(s)
        bz.generate_rays_device(cam, s, d, 5, 22)
        torch.cuda.synchronize()
        assert np.array_equal(s.cpu().numpy(), rs) and np.array_equal(d.cpu().numpy(), rd)


def test_queue_overflow_falls_back_inline(bz, oracle_built):
    """many rays through the same few patches: the candidate queues overflow and the inline path must give the same bits"""
    import bzoracle
    mesh, bezier, patches = common.ref_config("sphere16x8")
    rng = np.random.default_rng(3)
    n = 50000
    s = np.zeros((n, 3), np.float32)
    s[:, 0] = -10
    s[:, 1:] = (rng.random((n, 2)) * 1.6 - 0.8).astype(np.float32)
    d = np.tile(np.array([[1, 0, 0]], np.float32), (n, 1))
    ref = bzoracle.intersect(patches, s, d, threads=8)
    got = bz.BezierLens(1.3, patches).intersect(s, d)
    common.assert_same_hits(ref, got, "dense")


def test_packet_walk_equals_lane_walk_on_coherent_batches(bz, oracle_built):
    """parallel and fan-shaped ray grids (coherent packets: the warp-cooperative walk) against the per-lane walk and the oracle,
    including rays that start inside the lens and a far origin beyond the hierarchy's origin limit"""
    import bzoracle
    mesh, bezier, patches = common.ref_config("ellipsoid64x32")
    lo, hi = common.bbox(mesh)
    ys, zs = np.meshgrid(np.linspace(lo[1] * 1.1, hi[1] * 1.1, 256, dtype=np.float32), np.linspace(lo[2] * 1.1, hi[2] * 1.1, 160, dtype=np.float32))
    n = ys.size
    cases = []
    for x0 in (-10.0, 0.0, -2000.0):                                        # outside, inside the body, beyond origin_limit
        s = np.stack([np.full(n, x0, np.float32), ys.ravel(), zs.ravel()], axis=1)
        d = np.tile(np.array([[1, 0, 0]], np.float32), (n, 1))
        cases.append((s, d))
    eye = np.array([-9.0, 0.3, -0.2], np.float32)                            # a fan from one point
    target = np.stack([np.zeros(n, np.float32), ys.ravel(), zs.ravel()], axis=1)
    d = target - eye
    d = (d / np.sqrt((d * d).sum(1, keepdims=True))).astype(np.float32)
    cases.append((np.tile(eye, (n, 1)), d))
    packet, lanes = bz.BezierLens(1.3, patches), bz.BezierLens(1.3, patches).set_mode(bz.MODE_LANES)
    for s, d in cases:
        ref = bzoracle.intersect(patches, s, d, threads=8)
        common.assert_same_hits(ref, packet.intersect(s, d), "packet walk")
        common.assert_same_hits(ref, lanes.intersect(s, d), "lane walk")
    assert int((ref["what"] == 4).sum()) > 1000
