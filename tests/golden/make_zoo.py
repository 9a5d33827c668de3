"""Adversarial "zoo" scenes for the conservative filters of the CUDA path, with golden images
rendered by the patched reference (oracle/_ref/libref.so).

  python tests/golden/make_zoo.py        (needs /root/reference built into oracle/_ref)

Writes scenes/zoo_<k>.scene (seeded, deterministic) and zoo.npz (reference renders, 96x72).
What the scenes stress:
  zoo_quadrics  every quadric family around the view axis: ellipsoids, cylinders (a_q = 0 for
                axis-parallel rays), cones, one- and two-sheet hyperboloids, paraboloids, a
                "plane" written as a quadric (always the linear a_q == 0 path), mixed-sign cross
                terms, tiny and huge coefficient scales; reflective, so bounce origins sit ON them.
  zoo_spheres   touching / overlapping / nested spheres, lights inside spheres, a sphere around
                the camera position, radii from 1e-3 to 1e3, reflectivity 1 (opacity 0), planes through
                sphere centres.
  zoo_lights    spot lights (integer and fractional angular-a0, cones that cut the scene), a light
                with radial-a0 = 0 (never culled), lights below the floor and exactly on a surface,
                20 lights (two passes of 16).
  zoo_far       the same kind of content translated 5000 units from the origin and scaled 100x
                (cancellation in tmp = P - C and in the quadric coefficients).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

W, H = 96, 72


def fmt(v):
    return ("%.6g" % v)


def vec(v):
    return "[" + ", ".join(fmt(x) for x in v) + "]"


def sphere(p, r, kd, refl=0.0, ks=(1, 1, 1)):
    return (f"sphere, diffuse_color: {vec(kd)}, ior: 1.3, position: {vec(p)}, radius: {fmt(r)}, "
            f"reflectivity: {fmt(refl)}, refractivity: 0, specular_color: {vec(ks)}")


def plane(p, n, kd, refl=0.0):
    return f"plane, diffuse_color: {vec(kd)}, normal: {vec(n)}, position: {vec(p)}, reflectivity: {fmt(refl)}"


def quadric(c, kd, refl=0.0):
    names = "abcdefghij"
    body = ", ".join(f"{n}: {fmt(v)}" for n, v in zip(names, c))
    return f"quadric, {body}, diffuse_color: {vec(kd)}, reflectivity: {fmt(refl)}, specular_color: [0.6, 0.6, 0.6]"


def light(p, col, rad=(0.2, 0.02, 0.005), spot=None):
    s = (f"light, color: {vec(col)}, position: {vec(p)}, radial-a0: {fmt(rad[0])}, radial-a1: {fmt(rad[1])}, "
         f"radial-a2: {fmt(rad[2])}")
    if spot:
        d, theta, a0 = spot
        s += f", direction: {vec(d)}, theta: {fmt(theta)}, angular-a0: {fmt(a0)}"
    return s


def centred(Q, lin, J, c):
    """Quadric (x-c)^T Q (x-c) + lin.(x-c) + J expanded into a..j."""
    Q = np.asarray(Q, float)
    c = np.asarray(c, float)
    lin = np.asarray(lin, float)
    A, B, C = Q[0, 0], Q[1, 1], Q[2, 2]
    D, E, F = 2 * Q[0, 1], 2 * Q[0, 2], 2 * Q[1, 2]
    g = -2 * Q @ c + lin
    j = c @ Q @ c - lin @ c + J
    return [A, B, C, D, E, F, g[0], g[1], g[2], j]


def zoo_quadrics(rng):
    L = ["camera, height: 1.5, width: 2"]
    kd = lambda: rng.uniform(0.1, 1.0, 3)
    diag = lambda a, b, c: np.diag([a, b, c])
    L.append(quadric(centred(diag(1, 2, 0.5), [0, 0, 0], -1.0, [-3, 1.5, -9]), kd(), 0.3))       # ellipsoid
    L.append(quadric(centred(diag(1, 0, 1), [0, 0, 0], -0.6, [-0.5, 0, -8]), kd(), 0.2))          # cylinder along y
    L.append(quadric(centred(diag(1, 1, 0), [0, 0, 0], -0.3, [2.5, 1.2, -7]), kd(), 0.4))         # cylinder along z (view axis)
    L.append(quadric(centred(diag(1, -0.5, 1), [0, 0, 0], 0.0, [3.2, -0.5, -10]), kd(), 0.25))    # cone
    L.append(quadric(centred(diag(1, -1, 1), [0, 0, 0], -0.3, [-2.5, -1.5, -8]), kd(), 0.5))      # one-sheet hyperboloid
    L.append(quadric(centred(diag(-1, 2, -1), [0, 0, 0], -0.2, [0.8, 2.2, -11]), kd(), 0.0))      # two-sheet hyperboloid
    L.append(quadric(centred(diag(1, 0, 1), [0, -1, 0], 0.0, [1.2, -2.0, -9]), kd(), 0.35))       # paraboloid
    L.append(quadric([0, 0, 0, 0, 0, 0, 0.1, 1, 0.05, 3.2], kd(), 0.3))                           # a plane as a quadric
    Q = np.array([[1.0, 0.7, -0.3], [0.7, 1.5, 0.4], [-0.3, 0.4, 0.8]])
    L.append(quadric(centred(Q, [0, 0, 0], -0.8, [-1.2, 0.4, -6]), kd(), 0.2))                    # rotated ellipsoid
    L.append(quadric([1e-4 * x for x in centred(diag(1, 1, 1), [0, 0, 0], -0.5, [4.5, 2.5, -12])], kd(), 0.1))
    L.append(quadric([1e4 * x for x in centred(diag(2, 1, 1), [0, 0, 0], -0.4, [-4.5, 2.6, -12])], kd(), 0.1))
    L.append(quadric(centred(np.array([[0.0, 1, 0], [1, 0, 0], [0, 0, 0.2]]), [0, 0, 0], -0.1, [0, -0.5, -14]), kd(), 0.6))
    L.append(sphere([0.3, 0.3, -5], 0.4, kd(), 0.5))
    L.append(plane([0, -3.5, 0], [0, 1, 0], [0.4, 0.4, 0.4], 0.3))
    L.append(light([4, 6, -2], [2.5, 2.5, 2.5]))
    L.append(light([-5, 3, -6], [1.5, 1.2, 1.0]))
    L.append(light([0, 0.5, -9], [1.0, 1.5, 1.0], spot=([0, -1, 0.2], 3, 2)))
    return L


def zoo_spheres(rng):
    L = ["camera, height: 1.5, width: 2"]
    kd = lambda: rng.uniform(0.1, 1.0, 3)
    L.append(sphere([0, 0, -8], 1.0, kd(), 0.5))
    L.append(sphere([2, 0, -8], 1.0, kd(), 0.5))          # touching the first
    L.append(sphere([1, 1.2, -8], 0.9, kd(), 0.75))       # overlapping both
    L.append(sphere([0, 0, -8], 0.5, kd(), 0.0))          # nested
    L.append(sphere([-2.5, -0.5, -7], 1.2, kd(), 1.0))    # opacity 0: no shadow rays
    L.append(sphere([0, 0, 3.0], 2.5, kd(), 0.2))         # behind the camera, nearly touching it
    L.append(sphere([-1, 2.0, -6], 1e-3, kd(), 0.0))
    L.append(sphere([300, 200, -2000], 1000.0, kd(), 0.3))
    L.append(sphere([3.5, 2.0, -9], 0.8, kd(), 0.6))
    L.append(sphere([3.5, 2.0, -9], 0.8, kd(), 0.1))      # exact duplicate: lowest index wins the tie
    for _ in range(30):
        p = rng.uniform([-6, -3, -16], [6, 4, -4])
        L.append(sphere(p, rng.uniform(0.15, 0.9), kd(), rng.choice([0, 0.3, 0.6, 0.9])))
    L.append(plane([0, 0, -8], [0, 1, 0], [0.5, 0.5, 0.5], 0.4))   # through sphere centres
    L.append(plane([0, -3, 0], [0.1, 1, 0.05], [0.3, 0.4, 0.3], 0.0))
    L.append(light([0, 0, -8.2], [3, 3, 3]))              # inside the nested spheres
    L.append(light([5, 8, -3], [2, 2, 2]))
    L.append(light([-6, 2, -12], [1, 1, 2]))
    L.append(light([1, 1.2, -7.1], [2, 1, 1]))            # exactly on a sphere surface (r = 0.9)
    return L


def zoo_lights(rng):
    L = ["camera, height: 1.5, width: 2"]
    kd = lambda: rng.uniform(0.1, 1.0, 3)
    for _ in range(24):
        p = rng.uniform([-5, -2, -14], [5, 3, -5])
        L.append(sphere(p, rng.uniform(0.3, 1.0), kd(), rng.choice([0, 0.25, 0.5])))
    L.append(quadric(centred(np.diag([1, 3, 1]), [0, 0, 0], -1.5, [0, 0.5, -9]), kd(), 0.3))
    L.append(plane([0, -2.5, 0], [0, 1, 0], [0.5, 0.5, 0.5], 0.2))
    L.append(plane([0, 0, -18], [0, 0, 1], [0.3, 0.3, 0.5], 0.0))
    L.append(light([0, 6, -9], [3, 3, 3], spot=([0, -1, 0], 2, 3)))
    L.append(light([4, 4, -4], [2, 2, 1], spot=([-0.5, -0.6, -0.6], 1, 2.5)))     # fractional exponent: never culled
    L.append(light([-4, 5, -6], [1, 2, 2], spot=([0.3, -0.8, -0.4], 4, 0)))
    L.append(light([0, 1, -2], [1, 1, 1], rad=(0.0, 0.05, 0.01)))                 # radial-a0 = 0: never culled
    L.append(light([0, -6, -9], [2, 2, 2]))                                       # below the floor
    L.append(light([2, -2.5, -8], [1, 1, 1]))                                     # exactly on the floor plane
    for _ in range(14):                                                           # 20 lights: two passes of 16
        L.append(light(rng.uniform([-8, 0, -16], [8, 9, -1]), rng.uniform(0.2, 0.8, 3)))
    return L


def zoo_far(rng):
    L = ["camera, height: 1.5, width: 2"]
    kd = lambda: rng.uniform(0.1, 1.0, 3)
    off = np.array([3000.0, -2500.0, -5000.0])
    s = 100.0
    view = off / np.linalg.norm(off)
    for _ in range(40):
        t = rng.uniform(0.7, 1.3)
        jitter = rng.uniform(-1, 1, 3) * np.array([1500, 1200, 300])
        L.append(sphere(off * t + jitter, s * rng.uniform(0.5, 3.0), kd(), rng.choice([0, 0.4, 0.8])))
    for _ in range(6):
        c = off * rng.uniform(0.8, 1.2) + rng.uniform(-1, 1, 3) * np.array([1200, 900, 200])
        q = centred(np.diag(rng.uniform(0.3, 2.0, 3)) / (s * s), [0, 0, 0], -rng.uniform(1.0, 6.0), c)
        L.append(quadric(q, kd(), rng.choice([0, 0.5])))
    L.append(plane(off * 1.5, -view, [0.4, 0.4, 0.4], 0.3))
    L.append(light(off * 0.5 + np.array([0, 2000, 0]), [3, 3, 3], rad=(0.3, 1e-4, 1e-8)))
    L.append(light(off * 0.9 + np.array([-1500, 500, 800]), [2, 2, 3], rad=(0.3, 1e-4, 1e-8)))
    L.append(light([0, 0, 0], [1, 1, 1], rad=(0.5, 1e-4, 1e-8)))
    # the camera looks down -z: tilt the whole scene instead by using a wide frustum
    L[0] = "camera, height: 2.4, width: 3.2"
    return L


ZOO = {"zoo_quadrics": zoo_quadrics, "zoo_spheres": zoo_spheres, "zoo_lights": zoo_lights, "zoo_far": zoo_far}


def main():
    import rtlib
    out = {}
    for k, (name, fn) in enumerate(ZOO.items()):
        rng = np.random.default_rng(1000 + k)
        text = "\n".join(fn(rng)) + "\n"
        assert all(len(f) < 99 for line in text.splitlines() for f in line.split(",")), name
        path = os.path.join(HERE, "scenes", f"{name}.scene")
        with open(path, "w") as f:
            f.write(text)
        img = rtlib.ref_render(path, W, H)
        out[name] = img
        nz = float((img.reshape(-1, 3).max(axis=1) > 0).mean())
        print(name, "lit pixels", round(nz, 3), "shapes", sum(l.split(",")[0] in ("sphere", "plane", "quadric") for l in text.splitlines()))
    np.savez_compressed(os.path.join(HERE, "zoo.npz"), **out)


if __name__ == "__main__":
    main()
