"""Illumination mode: emitter sampling, bounce loop to a target rectangle.

The reference has the hemisphere sampler only (reference/hostUtil.cpp:3-29).  It is PINNED to the reference: oracle/_ref
holds the real UniformHemisphere (hostUtil.cpp compiled from /root/reference); its belt table and its getRandom(), fed with
the 24-bit draws the C ABI's counter-based generator produces (the engine state is set through the engine's own stream
extractor, oracle/ref_capi.cpp), are compared with bz_hemisphere_patches / bz_emit_rays and with the plain-C restatement
oracle/bz_oracle_illum.c; tests/golden/golden_hemisphere.npz keeps the reference's outputs for machines without it.
The bounce loop to the target has no reference counterpart (README.md:161-194 was never written): its checker is the
restatement around the oracle's refract().  Tolerances:
  * belt table, ray origins, landing bins, bin counts: bit-exact / integer-exact
  * emitted directions: acos / sin / cos are evaluated in double on both sides (as the reference does) and rounded to
    float, so all but a vanishing fraction of the components are identical; the test allows 1e-6 absolute on <= 0.01 %
    of the rays and demands identical bits on the rest
"""
import ctypes as C

import numpy as np
import pytest

import common


def _setup(bz):
    emitter = bz.make_emitter(origin=(-3.0, -0.1, -0.1), edge_u=(0.0, 0.2, 0.0), edge_v=(0.0, 0.0, 0.2), points=(4, 4), rays_per_point=4096, belts=16, seed=12345)
    target = bz.make_target(origin=(4.0, -2.0, -2.0), edge_u=(0.0, 4.0, 0.0), edge_v=(0.0, 0.0, 4.0), bins=(32, 24))
    return emitter, target


def test_hemisphere_table_and_struct_sizes(oracle_built):
    import bezier_b200 as bz
    import bzoracle
    lib = bz.load()
    assert lib.bz_sizeof_emitter() == C.sizeof(bz.Emitter) == 96 and lib.bz_sizeof_target() == C.sizeof(bz.Target) == 44
    for belts in (1, 2, 7, 16, 64, 128):
        n, width, first = bz.hemisphere_patches(belts)
        on, owidth, ofirst = bzoracle.hemisphere(belts)
        assert n == on and np.array_equal(width, owidth) and np.array_equal(first, ofirst)
        # s_i = ceil(4 b sin((2i + 1) / (4b) pi)), README.md:170
        want = [int(np.ceil(4 * belts * np.sin((2 * i + 1) / (4 * belts) * np.pi) - 1e-9)) for i in range(belts)]
        assert abs(n - sum(want)) <= 1
    with pytest.raises(bz.BezierError):
        bz.hemisphere_patches(0)
    with pytest.raises(bz.BezierError):
        bz.hemisphere_patches(129)


BELTS = (1, 2, 7, 16, 64, 128)


def _second_normalisation(local):
    """bz_emit_rays maps the hemisphere direction into the emitter's frame and normalises again (include/bz_capi.h); with the
    identity frame that is Eigen's normalized() of the reference's vector: z = x0 x0 + (x1 x1 + x2 x2), v / sqrt(z), in FP32"""
    x = local.astype(np.float32)
    z = x[:, 0] * x[:, 0] + (x[:, 1] * x[:, 1] + x[:, 2] * x[:, 2])
    return (x / np.sqrt(z)[:, None]).astype(np.float32)


def test_hemisphere_pinned_to_the_reference(oracle_built):
    """belt table and getRandom of the REAL UniformHemisphere against the product's host table and the plain-C restatement"""
    import bezier_b200 as bz
    import bzoracle
    import bzref
    for belts in BELTS:
        n, width, first = bzref.hemisphere(belts)
        pn, pwidth, pfirst = bz.hemisphere_patches(belts)
        on, owidth, ofirst = bzoracle.hemisphere(belts)
        assert n == pn == on and np.array_equal(width, pwidth) and np.array_equal(first, pfirst), belts
        assert np.array_equal(width, owidth) and np.array_equal(first, ofirst), belts
        emitter = bz.make_emitter(origin=(0.0, 0.0, 0.0), edge_u=(0.0, 1.0, 0.0), edge_v=(0.0, 0.0, 1.0), points=(2, 2), rays_per_point=5000, belts=belts, seed=77 + belts)
        draw1, draw2 = bzref.abi_draws(emitter.seed, 0, emitter.rays)
        direction, patch = bzref.hemisphere_random(belts, draw1, draw2)
        _, odir, opatch = bzoracle.emit(emitter, 0, emitter.rays)
        assert np.array_equal(patch, opatch), "belts %d: %d patch indices differ" % (belts, int((patch != opatch).sum()))
        assert common.same_bits(_second_normalisation(direction), odir), belts
        assert int(patch.max()) < n


def test_hemisphere_golden(oracle_built):
    """the committed outputs of the reference (tests/golden/make_golden.py) against the plain-C restatement: pins it where
    /root/reference is not mounted"""
    import os
    import bezier_b200 as bz
    import bzoracle
    g = np.load(os.path.join(common.ROOT, "tests", "golden", "golden_hemisphere.npz"))
    for belts in BELTS:
        n, width, first = bzoracle.hemisphere(belts)
        assert n == int(g["count_%d" % belts]) and np.array_equal(width, g["width_%d" % belts]) and np.array_equal(first, g["first_%d" % belts])
        emitter = bz.make_emitter(origin=(0.0, 0.0, 0.0), edge_u=(0.0, 1.0, 0.0), edge_v=(0.0, 0.0, 1.0), points=(1, 1), rays_per_point=int(g["rays"]), belts=belts, seed=int(g["seed"]))
        _, odir, opatch = bzoracle.emit(emitter, 0, emitter.rays)
        assert np.array_equal(opatch, g["patch_%d" % belts])
        assert common.same_bits(odir, _second_normalisation(g["direction_%d" % belts]))


@pytest.mark.gpu
def test_emission_pinned_to_the_reference(oracle_built):
    """bz_emit_rays (GPU) against the REAL UniformHemisphere::getRandom fed with the same draws.  acos / sin / cos run in
    double on both sides and are rounded to float: the device's double routines may differ from glibc's in the last bit of
    a double, which survives the rounding for a vanishing share of the rays"""
    import bezier_b200 as bz
    import bzref
    for belts in BELTS:
        emitter = bz.make_emitter(origin=(0.5, -0.25, 0.125), edge_u=(0.0, 1.0, 0.0), edge_v=(0.0, 0.0, 1.0), points=(2, 2), rays_per_point=25000, belts=belts, seed=2026 + belts)
        _, d, p = bz.emit_rays(emitter)
        draw1, draw2 = bzref.abi_draws(emitter.seed, 0, emitter.rays)
        direction, patch = bzref.hemisphere_random(belts, draw1, draw2)
        want = _second_normalisation(direction)
        same = (d.view(np.uint32) == want.view(np.uint32)).all(axis=1)
        assert (~same).mean() <= 1e-4 and np.abs(d - want).max() <= 1e-6, "belts %d: %d of %d directions differ" % (belts, int((~same).sum()), emitter.rays)
        assert (p != patch).mean() <= 1e-4, belts


@pytest.mark.gpu
def test_emission_matches_the_cpu_restatement(oracle_built):
    import bezier_b200 as bz
    import bzoracle
    emitter, _ = _setup(bz)
    s, d, p = bz.emit_rays(emitter)
    os_, od, op = bzoracle.emit(emitter, 0, emitter.rays)
    assert s.shape == (emitter.rays, 3) and np.array_equal(s, os_)
    same = (d.view(np.uint32) == od.view(np.uint32)).all(axis=1)
    assert (~same).mean() <= 1e-4 and np.abs(d - od).max() <= 1e-6
    assert (p != op).mean() <= 1e-4
    # a range in the middle of the emission is the same rays
    s2, d2, p2 = bz.emit_rays(emitter, first=5000, count=777)
    assert np.array_equal(s2, s[5000:5777]) and np.array_equal(d2, d[5000:5777]) and np.array_equal(p2, p[5000:5777])
    # unit directions into the hemisphere of +x, every hemisphere patch about equally likely (equal solid angles)
    assert np.abs(np.sqrt((d.astype(np.float64) ** 2).sum(1)) - 1.0).max() < 1e-6 and (d[:, 0] >= 0.0).all()
    patches, _, _ = bz.hemisphere_patches(emitter.belts)
    hist = np.bincount(p, minlength=patches)
    mean = emitter.rays / patches
    assert hist.min() > 0.5 * mean and hist.max() < 1.6 * mean


@pytest.mark.gpu
def test_trace_to_target_and_illuminate_match_the_oracle(oracle_built):
    import bezier_b200 as bz
    import bzoracle
    mesh, bezier, patches = common.ref_config("sphere16x8")
    emitter, target = _setup(bz)
    s, d, p = bz.emit_rays(emitter)
    lens = bz.BezierLens(1.3, patches)
    landing, bins, stats = lens.trace_to_target(s, d, target, max_crossings=2)
    want_landing, want_bins, counts = bzoracle.trace_to_target(patches, 1.3, s, d, target, 2, threads=8)
    assert np.array_equal(landing, want_landing), "%d rays land elsewhere" % int((landing != want_landing).sum())
    assert np.array_equal(bins, want_bins)
    assert stats.primary_rays == emitter.rays and stats.refract_calls == counts["refract_calls"]
    assert stats.entered == counts["entered"] and stats.exited == counts["exited"] and stats.lost == counts["lost"]
    assert counts["entered"] > 1000 and int(bins.sum()) > 1000 and int((landing == bz.LAND_MISSED).sum()) > 0
    # the whole simulation (emit -> group by hemisphere patch -> trace): the same counts, whatever the order of the rays
    total, istats = lens.illuminate(emitter, target, max_crossings=2)
    assert np.array_equal(total, want_bins)
    assert istats.primary_rays == emitter.rays and istats.entered == counts["entered"] and istats.exited == counts["exited"]
    twice, _ = lens.illuminate(emitter, target, max_crossings=2, bins=total.copy())      # accumulates
    assert np.array_equal(twice, 2 * want_bins)
    # grouping by patch makes the packets coherent: most calls take the warp-cooperative walk
    assert istats.lane_walk_calls < stats.lane_walk_calls or stats.lane_walk_calls == 0
