"""Multi-GPU checks (need >= 2 B200s; skipped otherwise): the N-GPU frame must equal the 1-GPU frame
byte for byte, through the single-process library path (rt_init(N): NCCL band gather inside the
library)."""
import os

import numpy as np
import pytest

import rtlib
from rtlib import pkg

pytestmark = pytest.mark.gpu


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("scene,W,H", [("ref_reflection.scene", 640, 360), ("ref_quadric.scene", 333, 211)])
def test_two_gpus_equal_one(scene, W, H):
    path = os.path.join(rtlib.SCENES, scene)
    r = pkg.Renderer(1)
    try:
        r.upload(path)
        one, _ = r.raycast(W, H)
    finally:
        r.close()
    n = min(_ngpu(), 8)
    r = pkg.Renderer(n)
    try:
        r.upload(path)
        many, st = r.raycast(W, H)
    finally:
        r.close()
    assert np.array_equal(one, many)
    assert st.pixels == W * H
