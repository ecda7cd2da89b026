"""`-m gpu` test of bz_render_static (include/bz_capi.h): the render that allocates nothing -- its scratch is a static device array.

Checked through the C ABI (ctypes on libbz_b200.so) against the existing entry point, bz_render, on the same lens.
Tolerance: the one the other render tests use -- BIT-EXACT (the same kernel in the parity build computes both frames).
Two calls in a row with different cameras and bands share the static array: the second must not see the first's accumulators.
The file is named to be collected last.
"""
import ctypes as C

import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def bz():
    import bezier_b200
    assert bezier_b200.device_count() > 0, "no CUDA device: the product has no CPU path"
    return bezier_b200


def _lens(bz, name):
    from bezier_b200 import workloads
    spec, mesh, cam_kw, scene_kw = workloads.setup(name, bz.Mesh)
    return spec, bz.BezierLens(spec["ri"], bz.BezierMesh(mesh)), cam_kw, scene_kw


def _render(lib, entry, lens, cam, scene, row0, row1, torch):
    """one call of `entry` (bz_render or bz_render_static) through the C ABI -> (frame, counters)"""
    rgba = torch.full((row1 - row0, cam.width, 4), 77, dtype=torch.uint8, device="cuda")
    stats = torch.zeros(32, dtype=torch.int64, device="cuda")
    rc = getattr(lib, entry)(lens.h, C.byref(cam), C.byref(scene), row0, row1, C.c_void_p(rgba.data_ptr()), C.c_void_p(stats.data_ptr()), None)
    assert rc == 0, lib.bz_last_error().decode()
    torch.cuda.synchronize()
    return rgba.cpu(), stats.cpu().tolist()


def test_render_static_matches_render_on_two_calls_in_a_row(bz):
    import torch
    lib = bz.load()
    spec, lens, cam_kw, scene_kw = _lens(bz, "cfg1_sphere_640x480")
    scene = bz.make_scene(**scene_kw)
    work = [bz.STAT_NAMES.index(n) for n in ("primary_rays", "refract_calls", "stage_a", "stage_b", "snell", "entered", "exited", "lost")]
    # first call: the whole 640 x 480 frame, 1 ray per pixel; second call: a band of a 4-rays-per-pixel frame of another shape, which
    # reuses the front of the static array (its accumulators still hold the first frame's sums unless the call clears them)
    calls = [(bz.make_camera(spec["width"], spec["height"], **cam_kw), 0, spec["height"]),
             (bz.make_camera(400, 300, **dict(cam_kw, strata=2)), 60, 260)]
    for mode in (bz.MODE_CULLED, bz.MODE_BRUTE):
        lens.set_mode(mode)
        for cam, row0, row1 in calls:
            want, want_stats = _render(lib, "bz_render", lens, cam, scene, row0, row1, torch)
            got, got_stats = _render(lib, "bz_render_static", lens, cam, scene, row0, row1, torch)
            assert int((want[..., 0] != want[0, 0, 0]).sum()) > 1000, "the frame shows the lens"
            assert torch.equal(want, got), "%d pixels differ" % int((want != got).any(dim=-1).sum())
            assert [want_stats[k] for k in work] == [got_stats[k] for k in work]
    lens.close()


def test_render_static_rejects_bands_beyond_its_scratch(bz):
    import torch
    lib = bz.load()
    spec, lens, cam_kw, scene_kw = _lens(bz, "cfg1_sphere_640x480")
    scene = bz.make_scene(**scene_kw)
    limit = 2097152                                   # BZ_STATIC_MAX_PIXELS
    cam = bz.make_camera(2048, 1025, **cam_kw)        # one row more than the array holds
    rgba = torch.zeros((1025, 2048, 4), dtype=torch.uint8, device="cuda")
    rc = lib.bz_render_static(lens.h, C.byref(cam), C.byref(scene), 0, 1025, C.c_void_p(rgba.data_ptr()), None, None)
    assert rc == bz.ERR_INVALID and b"BZ_STATIC_MAX_PIXELS" in lib.bz_last_error()
    assert 2048 * 1024 == limit
    # ... and the same frame in two bands is fine, and equals bz_render's
    want, _ = _render(lib, "bz_render", lens, cam, scene, 0, 1025, torch)
    top, _ = _render(lib, "bz_render_static", lens, cam, scene, 0, 1024, torch)
    bottom, _ = _render(lib, "bz_render_static", lens, cam, scene, 1024, 1025, torch)
    assert torch.equal(want, torch.cat([top, bottom]))
    lens.close()
