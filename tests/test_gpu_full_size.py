"""`-m gpu`: the BASELINE.json configurations at FULL size.  The whole frame is rendered on the device; the compiled
reference renders a pixel lattice of the same frame (all samples of every lattice pixel) and the two must agree
exactly on those pixels (north_star allows 1 LSB on 10 % of the pixels; the -fmad=false kernels need none of it).
Size-independent properties checked on the full frame: ray / call accounting, stripes == whole frame."""
import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu

#            workload                          lattice step (pixels)
CASES = [("cfg1_sphere_640x480", 4), ("cfg2_ellipsoid64x32_1080p", 12), ("cfg3_robot_1080p", 6), ("cfg4_freeform181x91_4k_16spp", 32)]


@pytest.mark.parametrize("name,step", CASES)
def test_full_frame_matches_reference_on_a_pixel_lattice(oracle_built, name, step):
    import bezier_b200 as bz
    import bzref
    from bezier_b200 import workloads
    assert bz.device_count() > 0, "no CUDA device: the product has no CPU path"
    spec, mesh, cam_kw, scene_kw = workloads.setup(name, bz.Mesh)
    lens = bz.BezierLens(spec["ri"], bz.BezierMesh(mesh))
    cam, scene = bz.make_camera(spec["width"], spec["height"], **cam_kw), bz.make_scene(**scene_kw)
    rgba, stats = lens.render(cam, scene)
    st = stats.as_dict()
    w, h, spp = spec["width"], spec["height"], spec["strata"] ** 2
    # accounting that holds at any size
    assert st["primary_rays"] == w * h * spp
    assert st["entered"] >= st["exited"] and st["entered"] + st["exited"] <= st["snell"]
    assert st["refract_calls"] >= st["primary_rays"] + st["exited"]
    assert st["entered"] > w * h * spp // 20
    assert rgba.shape == (h, w, 4) and int(rgba[..., 3].min()) == 255 and np.array_equal(rgba[..., 0], rgba[..., 2])
    # the reference on the lattice
    _, rmesh, rcam_kw, rscene_kw = workloads.setup(name, bzref.RefMesh)
    ref_lens = bzref.RefLens(spec["ri"], bzref.RefBezierMesh(rmesh))
    rcam, rscene = bzref.make_camera(w, h, **rcam_kw), bzref.make_scene(**rscene_kw)
    ref, ref_stats, rays = ref_lens.render(rcam, rscene, pixel_step=step, threads=16)      # cfg4: 8 160 pixels x 16 samples
    lattice = (slice(0, h, step), slice(0, w, step))
    assert rays == len(range(0, h, step)) * len(range(0, w, step)) * spp
    diff = int((ref[lattice] != rgba[lattice]).any(-1).sum())
    assert diff == 0, "%d of %d lattice pixels differ" % (diff, rays // spp)
    assert int((ref[lattice][..., 0] == 0).sum()) < rays // spp // 2        # the lattice is not mostly lost rays


def test_cfg4_stripes_equal_whole_frame(oracle_built):
    """cfg 5: the 4K frame rendered as 8 round-robin stripe bands and assembled on the device is byte-identical"""
    import torch
    import bezier_b200 as bz
    from bezier_b200 import workloads
    name = "cfg4_freeform181x91_4k_16spp"
    spec, mesh, cam_kw, scene_kw = workloads.setup(name, bz.Mesh)
    cam_kw = dict(cam_kw, strata=1)                      # 1 spp keeps the test short; the sharding is the same
    lens = bz.BezierLens(spec["ri"], bz.BezierMesh(mesh))
    cam, scene = bz.make_camera(spec["width"], spec["height"], **cam_kw), bz.make_scene(**scene_kw)
    full, _ = lens.render(cam, scene)
    ranks, stripe = 8, workloads.STRIPE_ROWS
    reserve = max(bz.striped_rows(cam.height, stripe, r, ranks) for r in range(ranks))
    bands = torch.zeros((ranks, reserve, cam.width, 4), dtype=torch.uint8, device="cuda")
    for rank in range(ranks):
        lens.render_striped_device(cam, scene, stripe, rank, ranks, bands[rank])
    image = torch.zeros((cam.height, cam.width, 4), dtype=torch.uint8, device="cuda")
    bz.assemble_stripes_device(bands, reserve, cam.width, cam.height, stripe, ranks, image)
    torch.cuda.synchronize()
    assert np.array_equal(image.cpu().numpy(), full)


def test_gather_rows_through_the_c_abi_single_rank(oracle_built):
    """bz_comm_* + bz_gather_rows (NCCL loaded at run time) with a communicator of one rank, both transports: the root's own band goes
    through the same scatter into the root's buffer (BZ_GATHER_PEER) or staging and assembly (BZ_GATHER_NCCL) as the other ranks' rows,
    twice in a row and after a change of the frame size (the peer buffer is set up anew).  Several ranks are exercised by bench.py under
    torchrun, whose frame_crc32 must equal the single-GPU one (profiles/, SCALE); the stripe arithmetic itself is covered on the CPU
    (tests/test_sharding_gloo.py)."""
    import torch
    import bezier_b200 as bz
    from bezier_b200 import workloads
    spec, mesh, cam_kw, scene_kw = workloads.setup("cfg1_sphere_640x480", bz.Mesh)
    lens = bz.BezierLens(spec["ri"], bz.BezierMesh(mesh))
    scene = bz.make_scene(**scene_kw)
    comm = bz.Comm(0, 1, 0, lambda ident: ident)
    assert comm.gather_mode() == bz.GATHER_PEER
    for mode in (bz.GATHER_PEER, bz.GATHER_NCCL):
        comm.set_gather(mode)
        for width, height in ((spec["width"], spec["height"]), (spec["width"], spec["height"]), (322, 243)):
            cam = bz.make_camera(width, height, **cam_kw)
            full, _ = lens.render(cam, scene)
            band = torch.zeros((cam.height, cam.width, 4), dtype=torch.uint8, device="cuda")
            image = torch.zeros_like(band)
            lens.render_striped_device(cam, scene, workloads.STRIPE_ROWS, 0, 1, band)
            comm.gather_rows(band, cam.width, cam.height, workloads.STRIPE_ROWS, rgba=image)
            torch.cuda.synchronize()
            assert np.array_equal(image.cpu().numpy(), full), (mode, width, height)
        assert comm.gather_mode() == mode
    comm.close()
