"""World-size-2 (and 3) CPU test of the N>1 path's host logic: band ownership, staging layout and the
point-to-point gather to rank 0 (gloo here, NCCL on the GPU box), then the de-interleave."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest

import rtlib  # noqa: F401  (puts the repo root on sys.path)

bands = importlib.import_module("raytracing-programs_b200.bands")
pkg = importlib.import_module("raytracing-programs_b200")


def synthetic_image(W, H):
    y, x, c = np.meshgrid(np.arange(H), np.arange(W), np.arange(3), indexing="ij")
    return ((x * 7 + y * 13 + c * 101 + (x * y) % 251) % 256).astype(np.uint8)


def _worker(rank, world, port, W, H, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        img = synthetic_image(W, H)
        sizes, offs = bands.layout(W, H, world)
        rows = pkg.rank_rows(H, rank, world)
        mine = img[np.asarray(rows, dtype=np.int64)].reshape(-1) if rows else np.zeros(0, np.uint8)
        assert mine.size == sizes[rank]
        rank_buf = torch.from_numpy(np.concatenate([mine, np.zeros(256, np.uint8)]))
        staging = torch.zeros(offs[-1] + sizes[-1] + 256, dtype=torch.uint8) if rank == 0 else None
        bands.gather_to_rank0(rank_buf, staging, sizes, offs, rank, world)
        if rank == 0:
            frame = bands.deinterleave_host(staging.numpy(), offs, W, H, world)
            q.put(bool(np.array_equal(frame, img)))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world,W,H", [(2, 61, 100), (2, 16, 7), (3, 33, 130)])
def test_band_gather_gloo(world, W, H):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, W, H, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def test_layout_covers_every_row_once():
    for H in (1, 31, 32, 33, 100, 4096):
        for world in (1, 2, 3, 8):
            seen = np.zeros(H, dtype=int)
            for k in range(world):
                seen[np.asarray(pkg.rank_rows(H, k, world), dtype=int)] += 1
            assert (seen == 1).all()
            sizes, offs = bands.layout(5, H, world)
            assert sum(sizes) == 5 * 3 * H and all(o % 256 == 0 for o in offs)
