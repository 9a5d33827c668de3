"""the reference's dormant recursive tracer shoot() (C/raycast.c:152-296) as an opt-in mode
(rt_set_mode(RT_MODE_SHOOT)).  Golden vectors come from the reference itself (tests/golden/make_golden_shoot.py runs
shoot() per pixel through oracle/ref_harness.c).  shoot() traces a refraction subtree but never adds its colour
(C/raycast.c:293,295 are commented out), so its value is the nested reflection sum -- the same terms as the live
iterative_shoot in a different order; pixel (435, 1262) of the 1500^2 frame of zoo_mirrors.scene is one where the two
orders give different bytes, and the mode must reproduce shoot()'s byte there."""
import os

import numpy as np
import pytest

import rtlib

GOLD = np.load(os.path.join(rtlib.GOLDEN, "shoot.npz"))
SHIPPED = ["simple", "reflection", "quadric", "example2", "example3", "quadric2"]
C4 = os.path.join(rtlib.ROOT, "scenes", "c4_syn1032.scene")
MIRRORS = os.path.join(rtlib.SCENES, "zoo_mirrors.scene")
ROWS = list(range(1260, 1265))


@pytest.mark.parametrize("name", SHIPPED)
def test_oracle_shoot_matches_reference_golden(name):
    img = rtlib.oracle_render(os.path.join(rtlib.SCENES, f"ref_{name}.scene"), 160, 120, mode=1)
    assert np.array_equal(img, GOLD[name])


def test_oracle_shoot_c4_and_the_differing_pixel():
    assert np.array_equal(rtlib.oracle_render(C4, 48, 48, mode=1), GOLD["c4_syn1032"])
    shoot = rtlib.oracle_render(MIRRORS, 1500, 1500, rows=ROWS, mode=1)
    live = rtlib.oracle_render(MIRRORS, 1500, 1500, rows=ROWS, mode=0)
    assert np.array_equal(shoot, GOLD["zoo_mirrors_rows1260_1500"])
    assert not np.array_equal(shoot[2, 435], live[2, 435])          # the two tracers disagree here ...
    assert (shoot != live).any(axis=2).sum() == 1                  # ... and nowhere else in these rows


@pytest.mark.skipif(not rtlib.ref_available(), reason="needs oracle/_ref (built where /root/reference is mounted)")
def test_live_reference_shoot_matches_oracle():
    for name in ("reflection", "example2"):
        p = os.path.join(rtlib.SCENES, f"ref_{name}.scene")
        assert np.array_equal(rtlib.ref_render(p, 200, 150, mode=1), rtlib.oracle_render(p, 200, 150, mode=1))


@pytest.mark.gpu
@pytest.mark.parametrize("name", SHIPPED)
def test_gpu_shoot_mode_golden(renderer, name):
    renderer.set_mode(1)
    try:
        renderer.upload(os.path.join(rtlib.SCENES, f"ref_{name}.scene"))
        img, _ = renderer.raycast(160, 120)
    finally:
        renderer.set_mode(0)
    assert np.array_equal(img, GOLD[name])


@pytest.mark.gpu
def test_gpu_shoot_mode_differs_where_the_reference_does(renderer):
    renderer.upload(MIRRORS)
    live, _ = renderer.raycast(1500, 1500)
    renderer.set_mode(1)
    try:
        shoot, _ = renderer.raycast(1500, 1500)
        renderer.upload(C4)
        c4, _ = renderer.raycast(48, 48)
        os.environ["RT_TILE_PIXELS"] = "100000"                    # several tiles: the per-level buffers are per tile
        try:
            renderer.upload(MIRRORS)
            tiled, _ = renderer.raycast(1500, 1500)
        finally:
            del os.environ["RT_TILE_PIXELS"]
    finally:
        renderer.set_mode(0)
    assert np.array_equal(shoot[ROWS], GOLD["zoo_mirrors_rows1260_1500"])
    assert np.array_equal(live[ROWS], rtlib.oracle_render(MIRRORS, 1500, 1500, rows=ROWS, mode=0))
    assert not np.array_equal(shoot[1262, 435], live[1262, 435])
    assert np.array_equal(c4, GOLD["c4_syn1032"])
    assert np.array_equal(tiled, shoot)
    full = rtlib.oracle_render(MIRRORS, 1500, 1500, mode=1)
    assert np.array_equal(shoot, full)


@pytest.mark.gpu
@pytest.mark.timeout(600)
def test_gpu_shoot_mode_deep_bounces_bounded_level_buffer(renderer):
    """RT_MODE_SHOOT with rt_set_max_depth(50): the per-level buffer is bounded (2 GiB) by shrinking the tile, so a
    2800 x 1000 frame (2.8 M pixels > the 2 683 904-pixel tile that bound leaves at depth 50) is rendered in two tiles
    whose level slices are cleared per tile.  Rows on both sides of the tile border against the oracle's shoot() at
    the same depth."""
    W, H = 2800, 1000
    rows = list(range(950, 966)) + [0, 1, 500, 999]              # the border falls in row 958
    renderer.set_mode(1)
    renderer.set_max_depth(50)
    try:
        renderer.upload(MIRRORS)
        a, _ = renderer.raycast(W, H)
        b, _ = renderer.raycast(W, H)                            # second frame: slices hold the first frame's records
    finally:
        renderer.set_mode(0)
        renderer.set_max_depth(7)
    want = rtlib.oracle_render(MIRRORS, W, H, rows=rows, mode=1, depth=50)
    assert np.array_equal(a[rows], want)
    assert np.array_equal(a, b)
