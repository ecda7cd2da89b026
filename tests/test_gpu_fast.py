"""`-m gpu`: BZ_MODE_FAST (the kernels rebuilt with FMA contraction, MUFU reciprocal / reciprocal square root, FP32 cos2) against the
compiled reference.

This build is NOT bit-exact, and it cannot meet north_star's "1e-5 relative on every ray" either -- no build that rounds differently
from the reference can: the reference's four fixed root-search iterations amplify rounding (tools/fma_sensitivity.py: the reference's own
algorithm compiled with FMA contraction on the CPU moves 2.5 % ... 48 % of the hit normals by more than 1e-5).  What
the fast build is held to, with the tolerances written here:
  * rendered frames: >= 98 % of the pixels identical to the reference's (north_star: >= 90 %); the others differ by whole flipped samples
    (a checker cell 25 <-> 230 or a lost ray, divided by the samples per pixel), the mean absolute difference stays below 0.25 LSB;
  * ray batches: status and hit patch agree on >= 98.5 % of the rays; on the agreeing hits the MEDIAN deviation of points (relative to
    the shape size) and of unit directions is <= 1e-5 and 90 % are within 1e-4;
  * ray generation and shading are the same bits as in the parity build (a frame without a lens is identical).
The parity build stays the default and the headline; tests/test_gpu_parity.py holds it to bit-exactness."""
import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def bz():
    import bezier_b200
    assert bezier_b200.device_count() > 0, "no CUDA device: the product has no CPU path"
    return bezier_b200


@pytest.mark.parametrize("name,ri,crossings", [("sphere32x16", 1.3, 2), ("ellipsoid64x32", 1.3, 2), ("robot", 1.5, 8)])
@pytest.mark.parametrize("pipeline", ["persistent", "wavefront"])
def test_fast_render_within_tolerance(bz, oracle_built, name, ri, crossings, pipeline):
    import bzref
    mesh, bezier, patches = common.ref_config(name)
    lo, hi = common.bbox(mesh)
    w, h, strata = 320, 240, 2
    half = float(max(hi[1] - lo[1], hi[2] - lo[2])) * 0.5 / 0.6 / 12.0
    eye = (float(lo[0]) - 12.0, float(lo[1] + hi[1]) / 2, float(lo[2] + hi[2]) / 2)
    rcam, rscene = bzref.make_camera(w, h, strata=strata, eye=eye, half_w=half), bzref.make_scene(bg_x=float(hi[0]) + 12.0, max_crossings=crossings)
    want, want_stats, _ = bzref.RefLens(ri, bezier).render(rcam, rscene, threads=16)
    cam, scene = bz.make_camera(w, h, strata=strata, eye=eye, half_w=half), bz.make_scene(bg_x=float(hi[0]) + 12.0, max_crossings=crossings)
    lens = bz.BezierLens(ri, patches).set_mode(bz.MODE_FAST).set_pipeline(bz.PIPELINE_WAVEFRONT if pipeline == "wavefront" else bz.PIPELINE_PERSISTENT)
    got, stats = lens.render(cam, scene)
    diff = np.abs(got[..., 0].astype(np.int32) - want[..., 0].astype(np.int32))
    identical = float((diff == 0).mean())
    sample = (230 - 25) / strata ** 2
    print("\n%s / %s: %.3f %% of the pixels identical, %d differ, largest difference %d (one flipped sample = %.1f)"
          % (name, pipeline, 100.0 * identical, int((diff != 0).sum()), int(diff.max()), sample))
    assert identical >= 0.98
    assert float(diff.mean()) < 0.25      # no drift: the frame differs by a few flipped samples (checker 25 <-> 230, or a lost ray: 0), not by a bias
    assert stats.primary_rays == want_stats.primary_rays
    assert abs(int(stats.entered) - int(want_stats.entered)) <= max(20, want_stats.entered // 200)


def test_fast_build_leaves_ray_generation_and_shading_alone(bz, oracle_built):
    """a frame that looks past the lens: only K1 and K3 run, and they are the parity build's bits"""
    mesh, bezier, patches = common.ref_config("sphere16x8")
    cam = bz.make_camera(200, 150, strata=3, eye=(-12.0, 6.0, 0.0), forward=(1.0, 0.3, 0.1), right=(-0.3, 1.0, 0.0), up=(-0.1, 0.0, 1.0), half_w=0.2)
    scene = bz.make_scene(bg_x=14.0)
    exact, stats = bz.BezierLens(1.3, patches).render(cam, scene)
    fast, _ = bz.BezierLens(1.3, patches).set_mode(bz.MODE_FAST).render(cam, scene)
    assert stats.entered == 0 and np.array_equal(exact, fast)


@pytest.mark.parametrize("name,ri", [("sphere32x16", 1.3), ("ellipsoid64x32", 1.3), ("robot", 1.5)])
def test_fast_batches_within_tolerance(bz, oracle_built, name, ri):
    import bzoracle
    mesh, bezier, patches = common.ref_config(name)
    lo, hi = common.bbox(mesh)
    size = float(np.linalg.norm(hi - lo))
    s, d = common.random_rays(lo, hi, 100000, seed=71)
    ref = bzoracle.intersect(patches, s, d, threads=16)
    lens = bz.BezierLens(ri, patches).set_mode(bz.MODE_FAST)
    got = lens.intersect(s, d)
    agree = (ref["what"] == got["what"]) & ((ref["what"] != 4) | (ref["patch"] == got["patch"]))
    hits = agree & (ref["what"] == 4)
    dp = np.abs(ref["point"][hits].astype(np.float64) - got["point"][hits]).max(1) / size
    dn = np.abs(ref["normal"][hits].astype(np.float64) - got["normal"][hits]).max(1)
    dp, dn = dp[np.isfinite(dp)], dn[np.isfinite(dn)]
    print("\n%s: status / patch agree on %.2f %% of %d rays; points: median %.1e, p90 %.1e; normals: median %.1e, p90 %.1e"
          % (name, 100.0 * agree.mean(), s.shape[0], np.median(dp), np.percentile(dp, 90), np.median(dn), np.percentile(dn, 90)))
    assert agree.mean() >= 0.985
    assert np.median(dp) <= 1e-5 and np.median(dn) <= 1e-5
    assert np.percentile(dp, 90) <= 1e-4 and np.percentile(dn, 90) <= 1e-4
    # two crossings through refract(): statuses agree on the same share
    s, d = common.random_rays(lo, hi, 100000, seed=72, inside_fraction=0.0)
    expected = np.full(s.shape[0], 1, np.uint32)
    rs, rd, rst, _, _ = bzoracle.refract(patches, ri, s, d, expected, threads=16)
    gs, gd, gst, _ = lens.refract(s, d, expected)
    same = rst == gst
    assert same.mean() >= 0.985
    moved = same & (rst != 0)
    dd = np.abs(rd[moved].astype(np.float64) - gd[moved]).max(1)
    dd = dd[np.isfinite(dd)]
    assert np.median(dd) <= 1e-5 and np.percentile(dd, 90) <= 1e-4
