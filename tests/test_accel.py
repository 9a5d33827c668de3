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
    assert sb.quadric_tests_d <= sa.quadric_tests_d + sa.ball_tests_d
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


def _ellipsoid_line(rng, centre, axes, rotate, refl):
    R = np.eye(3)
    if rotate:
        R, _ = np.linalg.qr(rng.normal(size=(3, 3)))
    s = rng.uniform(0.5, 6.0)
    M = s * (R @ np.diag(1.0 / np.asarray(axes) ** 2) @ R.T)
    g = -2.0 * M @ centre
    j = centre @ M @ centre - s
    co = dict(a=M[0, 0], b=M[1, 1], c=M[2, 2], d=2 * M[0, 1], e=2 * M[0, 2], f=2 * M[1, 2], g=g[0], h=g[1], i=g[2], j=j)
    col = rng.uniform(0.1, 1.0, 3)
    return (f"quadric, diffuse_color: [{col[0]:.4f}, {col[1]:.4f}, {col[2]:.4f}], specular_color: [1, 1, 1], "
            + ", ".join(f"{k}: {v:.9g}" for k, v in co.items()) + f", reflectivity: {refl:.2f}")


@pytest.mark.timeout(900)
def test_accel_quadric_index(renderer, tmp_path):
    """The hierarchy over the ellipsoids among the quadrics (csrc/rt_qindex.cuh): rotated and axis-aligned ellipsoids
    of mixed sizes, mirrors among them, quadrics the index cannot hold (cylinder, cone, paraboloid), a floor reaching
    far out, a light next to one ellipsoid and one inside another."""
    rng = np.random.default_rng(12)
    lines = ["camera, width: 2, height: 1.3",
             "plane, normal: [0, 1, 0.01], position: [0, -6, 0], diffuse_color: [0.7, 0.7, 0.7], specular_color: [1, 1, 1], reflectivity: 0.4",
             "sphere, radius: 2, position: [3, 0, -22], diffuse_color: [1, 0.2, 0.2], specular_color: [1, 1, 1], reflectivity: 0.6, refractivity: 0, ior: 1",
             "sphere, radius: 0.7, position: [-4, -2, -12], diffuse_color: [0.2, 1, 0.2], specular_color: [1, 1, 1], reflectivity: 0, refractivity: 0, ior: 1",
             "quadric, diffuse_color: [0.9, 0.9, 0.2], specular_color: [1, 1, 1], a: 1, b: 0, c: 1, d: 0, e: 0, f: 0, g: 16, h: 0, i: 60, j: 963, reflectivity: 0.3",
             "quadric, diffuse_color: [0.2, 0.9, 0.9], specular_color: [1, 1, 1], a: 1, b: -0.3, c: 1, d: 0, e: 0, f: 0, g: -20, h: 0, i: 50, j: 725, reflectivity: 0",
             "quadric, diffuse_color: [0.9, 0.4, 0.9], specular_color: [1, 1, 1], a: 0.5, b: 0, c: 0.5, d: 0, e: 0, f: 0, g: 0, h: -1, i: 40, j: 795, reflectivity: 0.2"]
    centres = []
    for k in range(140):
        c = np.array([rng.uniform(-18, 18), rng.uniform(-5, 10), -rng.uniform(8, 70)])
        centres.append(c)
        axes = rng.uniform(0.25, 1.6, 3) * (3.0 if k % 17 == 0 else 1.0)
        lines.append(_ellipsoid_line(rng, c, axes, rotate=(k % 3 != 0), refl=[0, 0, 0.25, 0.5, 0.9][k % 5]))
    lines += [f"light, color: [2, 2, 2], position: [{centres[3][0] + 1.9:.4f}, {centres[3][1] + 0.3:.4f}, {centres[3][2] + 0.2:.4f}], radial-a0: 1, radial-a1: 0, radial-a2: 0.002",
              f"light, color: [1, 2, 1], position: [{centres[0][0]:.4f}, {centres[0][1]:.4f}, {centres[0][2]:.4f}], radial-a0: 1, radial-a1: 0, radial-a2: 0.002",
              "light, color: [2, 1.5, 1], position: [-25, 30, -10], radial-a0: 1, radial-a1: 0, radial-a2: 0.0005",
              "light, color: [1, 1, 2], position: [10, 12, 5], radial-a0: 1, radial-a1: 0, radial-a2: 0.0005"]
    p = tmp_path / "ellipsoids.scene"
    p.write_text("\n".join(lines) + "\n")
    for W, H in ((480, 312), (257, 131)):
        a, sa = render(renderer, str(p), W, H, 0)
        b, sb = render(renderer, str(p), W, H, 1)
        assert np.array_equal(a, b), rtlib.compare(a, b)
        assert sa.rays == sb.rays
        # the index does prune: brute force meets every quadric with every ray (bounding-ball pre-tests + the quadric
        # tests of those that pass them)
        assert sb.quadric_tests_d * 3 < sa.quadric_tests_d + sa.ball_tests_d
        assert np.array_equal(b, rtlib.oracle_render(str(p), W, H))
    assert a.max() > 0 and len(np.unique(a)) > 50


@pytest.mark.timeout(900)
@pytest.mark.parametrize("shell", ["ellipsoid", "sphere"])
def test_accel_camera_inside_objects(renderer, tmp_path, shell):
    """The primary rays' cube map with the camera inside an ellipsoid / a sphere (every cell must list it; the shell's
    inside is lit because a surface skips itself, everything else is in its shadow: the reference's shadow rays have
    no t_max), objects behind the camera (mirrors show them), one on the optical axis."""
    rng = np.random.default_rng(4)
    lines = ["camera, width: 2.4, height: 1.6",
             "sphere, radius: 40, position: [1, 2, -5], diffuse_color: [0.5, 0.6, 0.9], specular_color: [1, 1, 1], reflectivity: 0.2, refractivity: 0, ior: 1"
             if shell == "sphere" else _ellipsoid_line(rng, np.array([0.5, -0.5, -2.0]), [25.0, 18.0, 30.0], True, 0.3),
             "sphere, radius: 1, position: [0, 0, -6], diffuse_color: [1, 0.1, 0.1], specular_color: [1, 1, 1], reflectivity: 0.8, refractivity: 0, ior: 1",
             "sphere, radius: 1.5, position: [4, 4, -4], diffuse_color: [0.1, 1, 0.1], specular_color: [1, 1, 1], reflectivity: 0.5, refractivity: 0, ior: 1",
             "sphere, radius: 2, position: [0, 1, 7], diffuse_color: [1, 1, 0.1], specular_color: [1, 1, 1], reflectivity: 0, refractivity: 0, ior: 1",
             "plane, normal: [0, 1, 0], position: [0, -3, 0], diffuse_color: [0.6, 0.6, 0.6], specular_color: [1, 1, 1], reflectivity: 0.5"]
    for k in range(14):
        c = np.array([rng.uniform(-8, 8), rng.uniform(-2, 6), rng.uniform(-12, 9)])
        if np.linalg.norm(c) < 2.5:
            c *= 3.0
        lines.append(_ellipsoid_line(rng, c, rng.uniform(0.4, 1.4, 3), k % 2 == 0, [0, 0.5, 0.9][k % 3]))
    lines += ["light, color: [2, 2, 2], position: [3, 6, -2], radial-a0: 1, radial-a1: 0, radial-a2: 0.003",
              "light, color: [2, 1, 1], position: [-5, 4, 3], radial-a0: 1, radial-a1: 0, radial-a2: 0.003"]
    p = tmp_path / "inside.scene"
    p.write_text("\n".join(lines) + "\n")
    for W, H in ((384, 256), (101, 67)):
        a, sa = render(renderer, str(p), W, H, 0)
        b, sb = render(renderer, str(p), W, H, 1)
        assert np.array_equal(a, b), rtlib.compare(a, b)
        assert (sa.rays, sa.rays_reflection, sa.hit_points) == (sb.rays, sb.rays_reflection, sb.hit_points)   # same hit points found
        assert sa.rays_reflection > W * H // 4
        assert np.array_equal(b, rtlib.oracle_render(str(p), W, H))
    # a sphere keeps its outward normal when hit from inside (C/raycast.c:465-469; only quadrics flip, :521-523): black frame
    assert shell == "sphere" or (a.max() > 0 and len(np.unique(a)) > 30)


@pytest.mark.timeout(900)
def test_accel_scene_without_spheres(renderer, tmp_path):
    """Only quadrics and a plane: the mode stays on with an empty sphere hierarchy, and a scene whose extent is absurd
    (one ellipsoid 1e17 away) renders through the hierarchies instead of the cube maps."""
    for far in (False, True):
        rng = np.random.default_rng(8)
        lines = ["camera, width: 2, height: 1.5",
                 "plane, normal: [0, 1, 0], position: [0, -4, 0], diffuse_color: [0.7, 0.7, 0.7], specular_color: [1, 1, 1], reflectivity: 0.3"]
        for k in range(40):
            c = np.array([rng.uniform(-10, 10), rng.uniform(-3, 6), -rng.uniform(6, 40)])
            lines.append(_ellipsoid_line(rng, c, rng.uniform(0.4, 1.8, 3), k % 2 == 1, [0, 0.4, 0.8][k % 3]))
        if far:
            lines.append("quadric, diffuse_color: [1, 1, 1], specular_color: [1, 1, 1], a: 1, b: 1, c: 1, d: 0, e: 0, f: 0, "
                         "g: -2e17, h: 0, i: 0, j: 1e34, reflectivity: 0")
        lines += ["light, color: [2, 2, 2], position: [5, 12, 0], radial-a0: 1, radial-a1: 0, radial-a2: 0.001",
                  "light, color: [1, 1, 2], position: [-8, 3, -20], radial-a0: 1, radial-a1: 0, radial-a2: 0.001"]
        p = tmp_path / f"nospheres{int(far)}.scene"
        p.write_text("\n".join(lines) + "\n")
        a, sa = render(renderer, str(p), 320, 240, 0)
        b, sb = render(renderer, str(p), 320, 240, 1)
        assert np.array_equal(a, b), rtlib.compare(a, b)
        assert sa.rays == sb.rays
        assert sb.quadric_tests_d * 2 < sa.quadric_tests_d + sa.ball_tests_d, (far, sb.quadric_tests_d, sa.quadric_tests_d, sa.ball_tests_d)   # the mode is on
        assert np.array_equal(b, rtlib.oracle_render(str(p), 320, 240))
        assert a.max() > 0


@pytest.mark.timeout(900)
def test_accel_quadric_index_deep_bounces(renderer):
    """C5 with the bounce limit of the reference's CUDA variant (50): long reflection chains between ellipsoids,
    spheres and the tilted floors, accelerated frame = brute-force frame."""
    scene = BIG[1][0]
    renderer.set_max_depth(50)
    try:
        a, sa = render(renderer, scene, 192, 144, 0)
        b, sb = render(renderer, scene, 192, 144, 1)
    finally:
        renderer.set_max_depth(7)
    assert np.array_equal(a, b), rtlib.compare(a, b)
    assert sa.rays == sb.rays and sb.quadric_tests_d * 5 < sa.quadric_tests_d + sa.ball_tests_d


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
