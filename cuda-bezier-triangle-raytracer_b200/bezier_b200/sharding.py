"""Row-stripe sharding of the image plane over the ranks of one box (north_star subsystem 4).

The frame is cut into stripes of `stripe_rows` rows dealt round-robin to the ranks; every rank renders its stripes into
a compact band buffer (bz_render_striped).  Rays are independent, so the only exchange is the image gather at the end:
on the GPUs bz_gather_rows of the C ABI (make_comm below: every rank stores its rows into the root's frame over NVLink);
the helpers here are the host-side arithmetic of the layout and an all-gather form of the same exchange (gloo in the CPU
tests, followed by bz_assemble_stripes' arithmetic).  Tables are replicated: every rank builds the same lens from the same mesh."""
import numpy as np


def image_rows_of_rank(height, stripe_rows, rank, ranks):
    """image row of every local band row of `rank`, in band order (the layout bz_render_striped writes)"""
    rows = []
    for first in range(rank * stripe_rows, height, ranks * stripe_rows):
        rows.extend(range(first, min(first + stripe_rows, height)))
    return np.asarray(rows, dtype=np.int64)


def band_rows(height, stripe_rows, ranks):
    """rows every rank reserves so that the bands gather as one equally sized tensor"""
    return max(len(image_rows_of_rank(height, stripe_rows, r, ranks)) for r in range(ranks))


def gather_bands(band, bands_out=None):
    """all-gather of this rank's padded band [band_rows, width, 4] -> [ranks, band_rows, width, 4] on every rank"""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size()
    if bands_out is None:
        bands_out = torch.empty((world,) + tuple(band.shape), dtype=band.dtype, device=band.device)
    dist.all_gather_into_tensor(bands_out.view((-1,) + tuple(band.shape[1:])), band)    # concatenated along dim 0
    return bands_out


def make_comm(bz, device):
    """bz.Comm over the ranks of the initialised torch.distributed group: torch only carries the 128-byte id (bootstrap), the
    gather itself is bz_gather_rows in the C ABI (peer stores into the root's IPC-mapped frame, or ncclSend / ncclRecv to the root)"""
    import torch
    import torch.distributed as dist
    rank, world = (dist.get_rank(), dist.get_world_size()) if dist.is_initialized() else (0, 1)

    def share(ident):
        if world == 1:
            return ident
        box = [ident]
        dist.broadcast_object_list(box, src=0)
        return box[0]

    return bz.Comm(rank, world, device, share)
