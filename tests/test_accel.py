"""Light buffer: the accelerated shadow search must give the brute-force image,
byte for byte, and both must equal the oracle / the golden images."""
import glob
import os
import time

import numpy as np
import pytest

import rtlib

pytestmark = pytest.mark.gpu

SCENES = sorted(glob.glob(os.path.join(rtlib.SCENES, "*.scene")))
BIG = [(os.path.join(rtlib.ROOT, "scenes", "c4_syn1032.scene"), 320, 200), (os.path.join(rtlib.ROOT, "scenes", "c5_syn4096.scene"), 128, 96)]


@pytest.fixture(scope="module")
def renderer():
    r = rtlib.pkg.Renderer(1)
    yield r
    r.set_accel(0)
    r.close()


def render(r, scene, W, H, accel):
    r.set_accel(accel)
    r.upload(scene)
    img, st = r.raycast(W, H)
    return img.copy(), st


@pytest.mark.timeout(600)
@pytest.mark.parametrize("scene", SCENES, ids=[os.path.basename(s) for s in SCENES])
def test_accel_equals_brute_force_and_oracle(renderer, scene):
    for W, H in ((160, 120), (333, 77)):
        a, _ = render(renderer, scene, W, H, 0)
        b, _ = render(renderer, scene, W, H, 1)
        assert np.array_equal(a, b), rtlib.compare(a, b)
        assert np.array_equal(b, rtlib.oracle_render(scene, W, H))


@pytest.mark.timeout(900)
@pytest.mark.parametrize("scene,W,H", BIG, ids=["c4", "c5"])
def test_accel_on_synthetic_scenes(renderer, scene, W, H):
    a, sa = render(renderer, scene, W, H, 0)
    b, sb = render(renderer, scene, W, H, 1)
    assert np.array_equal(a, b), rtlib.compare(a, b)
    assert sa.rays == sb.rays                                   # same rays of the reference semantics
    assert sb.sphere_tests_d < sa.sphere_tests_d                # fewer tests executed for them
    assert np.array_equal(b, rtlib.oracle_render(scene, W, H))


@pytest.mark.timeout(900)
def test_accel_far_hit_points_fall_back(renderer, tmp_path):
    """A floor reaching far beyond the last distance level of the light buffer, small spheres far from a light that
    sits next to one of them, a light inside a sphere: the fallback and the inflated far-level lists."""
    p = tmp_path / "far.scene"
    lines = ["camera, width: 2, height: 1.2",
             "plane, normal: [0, 1, 0.02], position: [0, -2, 0], diffuse_color: [0.8, 0.8, 0.8], specular_color: [1, 1, 1], reflectivity: 0.3",
             "sphere, radius: 3, position: [0, 1, -14], diffuse_color: [1, 0, 0], specular_color: [1, 1, 1], reflectivity: 0.5, refractivity: 0, ior: 1",
             "sphere, radius: 0.2, position: [2, 3, -9], diffuse_color: [0, 1, 0], specular_color: [1, 1, 1], reflectivity: 0, refractivity: 0, ior: 1"]
    rng = np.random.default_rng(3)
    for _ in range(60):
        x, y, z = rng.uniform(-30, 30), rng.uniform(-1.5, 12), -rng.uniform(6, 400)
        lines.append(f"sphere, radius: {rng.uniform(0.05, 1.5):.4f}, position: [{x:.4f}, {y:.4f}, {z:.4f}], diffuse_color: [0.3, 0.5, 0.9], "
                     "specular_color: [1, 1, 1], reflectivity: 0.25, refractivity: 0, ior: 1")
    lines += ["light, color: [2, 2, 2], position: [2.3, 3.1, -9.1], radial-a0: 1, radial-a1: 0, radial-a2: 0.001",
              "light, color: [1, 2, 1], position: [0, 1, -14], radial-a0: 1, radial-a1: 0, radial-a2: 0.001",
              "light, color: [2, 1, 1], position: [-20, 30, -60], radial-a0: 1, radial-a1: 0, radial-a2: 0.0001"]
    p.write_text("\n".join(lines) + "\n")
    a, _ = render(renderer, str(p), 640, 384, 0)
    b, _ = render(renderer, str(p), 640, 384, 1)
    assert np.array_equal(a, b), rtlib.compare(a, b)
    assert np.array_equal(b, rtlib.oracle_render(str(p), 640, 384))
    assert a.max() > 0


@pytest.mark.timeout(900)
def test_accel_speed_report(renderer):
    """Not a pass/fail threshold on time: prints the two frame times of C4 at 1024^2 for the log."""
    scene = BIG[0][0]
    out = {}
    for accel in (0, 1):
        render(renderer, scene, 1024, 1024, accel)
        _, st = render(renderer, scene, 1024, 1024, accel)
        out[accel] = (st.ms_kernels, st.rays / st.ms_kernels / 1e3)
    print("C4 1024^2 kernels ms / Mrays/s: brute", out[0], "light buffer", out[1])
    assert out[1][1] > 0
