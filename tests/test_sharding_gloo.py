"""Multi-rank path on the CPU: two gloo ranks shard a synthetic frame into row stripes, gather the bands and
reassemble it.  Checks the host-side layout logic against the C ABI (bz_striped_rows) -- the device kernels
(bz_render_striped, bz_assemble_stripes) are covered by the `-m gpu` tests."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import common  # noqa: F401  (sets sys.path through conftest)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, height, width, stripe, result):
    import bezier_b200 as bz
    from bezier_b200 import sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        frame = (np.arange(height * width * 4, dtype=np.int64) * 2654435761 % 251).astype(np.uint8).reshape(height, width, 4)
        rows = sharding.image_rows_of_rank(height, stripe, rank, world)
        assert len(rows) == bz.striped_rows(height, stripe, rank, world)            # same layout as the C ABI
        reserve = sharding.band_rows(height, stripe, world)
        band = torch.zeros((reserve, width, 4), dtype=torch.uint8)
        band[:len(rows)] = torch.from_numpy(frame[rows])                             # stands in for bz_render_striped
        bands = sharding.gather_bands(band)
        assert tuple(bands.shape) == (world, reserve, width, 4)
        image = np.zeros_like(frame)
        covered = np.zeros(height, bool)
        for r in range(world):
            rr = sharding.image_rows_of_rank(height, stripe, r, world)
            image[rr] = bands[r, :len(rr)].numpy()
            assert not covered[rr].any()
            covered[rr] = True
        ok = bool(covered.all() and np.array_equal(image, frame))
        flag = torch.tensor([1 if ok else 0])
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if rank == 0:
            result.put(int(flag.item()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("height,stripe", [(70, 8), (2160, 16), (33, 16)])
def test_two_ranks_shard_and_gather(height, stripe):
    ctx = mp.get_context("spawn")
    result = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, height, 24, stripe, result)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert result.get(timeout=5) == 1


def test_stripe_layout_covers_every_row_once():
    from bezier_b200 import sharding
    for height, stripe, ranks in [(2160, 16, 8), (1080, 16, 4), (17, 4, 3), (5, 16, 8)]:
        rows = np.concatenate([sharding.image_rows_of_rank(height, stripe, r, ranks) for r in range(ranks)])
        assert sorted(rows.tolist()) == list(range(height))
