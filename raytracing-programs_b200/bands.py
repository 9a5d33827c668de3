"""Row-band plumbing of the one-process-per-GPU launch (torchrun): which rows a rank owns,
where its rows land in rank 0's staging buffer, and the gather itself.

The render has no data-path collective (pixels are independent); the only
exchange is this gather of finished RGB8 rows to rank 0 for the PPM write (the reference writes one
file from one process, C/raycast.c:59).  Works on any torch.distributed backend: NCCL over NVLink
on the GPU box, gloo in the CPU tests.
"""
import importlib

_pkg = importlib.import_module(__name__.rsplit(".", 1)[0])


def layout(width, height, world):
    """Per-rank byte sizes and 256-byte-aligned offsets of the rank buffers inside the staging buffer."""
    sizes = [len(_pkg.rank_rows(height, k, world)) * width * 3 for k in range(world)]
    offs = [0]
    for s in sizes[:-1]:
        offs.append(offs[-1] + ((s + 255) // 256) * 256)
    return sizes, offs


def gather_to_rank0(rank_buf, staging, sizes, offs, rank, world):
    """Every rank's rows (rank_buf[:sizes[rank]], uint8) -> staging[offs[k]:offs[k]+sizes[k]] on rank 0.
    Point-to-point: only rank 0 ingests; no rank receives data it does not need."""
    import torch.distributed as dist
    if world == 1:
        staging[:sizes[0]].copy_(rank_buf[:sizes[0]])
        return
    if rank == 0:
        staging[:sizes[0]].copy_(rank_buf[:sizes[0]])
        ops = [dist.P2POp(dist.irecv, staging[offs[k]:offs[k] + sizes[k]], k) for k in range(1, world) if sizes[k]]
    else:
        ops = [dist.P2POp(dist.isend, rank_buf[:sizes[rank]], 0)] if sizes[rank] else []
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()


def deinterleave_host(staging, offs, width, height, world):
    """Host-side statement of what k_assemble does on GPU 0 (used by the CPU tests and as the
    checker of the device kernel): staging (uint8 tensor/array) -> [H, W, 3] numpy array."""
    import numpy as np
    st = np.asarray(staging)
    frame = np.empty((height, width, 3), dtype=np.uint8)
    pitch = width * 3
    for k in range(world):
        rows = _pkg.rank_rows(height, k, world)
        if rows:
            blk = st[offs[k]:offs[k] + len(rows) * pitch].reshape(len(rows), width, 3)
            frame[np.asarray(rows)] = blk
    return frame
