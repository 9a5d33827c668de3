"""Deep-bounce mode: rt_set_max_depth(50) = the reflection loop of C/raycast.c:349 with the
bound its CUDA/OptiX variants use (MAX_ITER 50).  Oracle pinned against the reference's own sources rebuilt with that
constant (oracle/_ref/libref_d50.so) and against committed hashes; GPU path against the oracle."""
import hashlib
import json
import os

import numpy as np
import pytest

import rtlib

MIRRORS = os.path.join(rtlib.SCENES, "zoo_mirrors.scene")
REFL = os.path.join(rtlib.SCENES, "ref_reflection.scene")
GOLD = os.path.join(rtlib.GOLDEN, "depth_hashes.json")
CASES = [(MIRRORS, 240, 180, 50), (MIRRORS, 240, 180, 7), (MIRRORS, 240, 180, 20), (REFL, 200, 150, 50)]


def _key(scene, W, H, depth):
    return f"{os.path.basename(scene)}_{W}x{H}_d{depth}"


def test_oracle_depth_matches_golden():
    gold = json.load(open(GOLD))
    for scene, W, H, depth in CASES:
        img = rtlib.oracle_render(scene, W, H, depth=depth)
        assert hashlib.md5(img.tobytes()).hexdigest() == gold[_key(scene, W, H, depth)], (scene, depth)


def test_depth_changes_the_mirror_scene():
    a = rtlib.oracle_render(MIRRORS, 240, 180, depth=7)
    b = rtlib.oracle_render(MIRRORS, 240, 180, depth=50)
    assert (a != b).any(axis=2).sum() > 50           # the mode is not a no-op on this scene


@pytest.mark.skipif(not os.path.exists(os.path.join(rtlib.ROOT, "oracle", "_ref", "libref_d50.so")), reason="oracle/_ref not built")
def test_oracle_depth50_matches_reference_sources():
    for scene, W, H in ((MIRRORS, 240, 180), (REFL, 200, 150), (os.path.join(rtlib.SCENES, "zoo_spheres.scene"), 160, 120)):
        assert np.array_equal(rtlib.oracle_render(scene, W, H, depth=50), rtlib.ref_render(scene, W, H, depth=50)), scene


@pytest.mark.gpu
@pytest.mark.timeout(600)
def test_gpu_depth_modes_match_oracle():
    r = rtlib.pkg.Renderer(1)
    try:
        for accel in (0, 1):
            r.set_accel(accel)
            for scene, W, H, depth in CASES + [(os.path.join(rtlib.ROOT, "scenes", "c4_syn1032.scene"), 96, 64, 50)]:
                r.set_max_depth(depth)
                r.upload(scene)
                img, st = r.raycast(W, H)
                assert np.array_equal(img, rtlib.oracle_render(scene, W, H, depth=depth)), (scene, depth, accel)
        r.set_max_depth(1)                       # no reflection rays at all
        r.upload(MIRRORS)
        img, st = r.raycast(120, 90)
        assert st.rays_reflection == 0 and np.array_equal(img, rtlib.oracle_render(MIRRORS, 120, 90, depth=1))
        assert r.lib.rt_set_max_depth(0) != 0 and r.lib.rt_set_max_depth(65) != 0
    finally:
        r.set_max_depth(7)
        r.set_accel(0)
        r.close()
