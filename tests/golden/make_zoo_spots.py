"""Spot lights with NON-INTEGER angular-a0: the one place where the CUDA path calls a libm routine of its own (CUDA's
pow, rt_math.cuh pow_a0) where the reference calls glibc's (C/raycast.c:695).  Golden image rendered by the patched
reference (oracle/_ref/libref.so), so the test pins CUDA pow against glibc pow through the pixels.

  python tests/golden/make_zoo_spots.py        (needs /root/reference built into oracle/_ref)

Writes scenes/zoo_spots.scene (the spheres, quadric and planes of zoo_lights.scene under 12 spot lights whose exponents
run over 0.01 .. 63.5, cones from narrow to wider than a half space) and zoo_spots.npz (reference render, 320 x 240).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
W, H = 320, 240
A0 = [0.01, 0.5, 1.5, 2.5, 3.7, 7.25, 19.9, 63.5, 0.3333333, 1.0000001, 12.125, 2.718281828]


def main():
    import rtlib
    src = open(os.path.join(HERE, "scenes", "zoo_lights.scene")).read().splitlines()
    keep = [l for l in src if not l.startswith("light")]
    rng = np.random.RandomState(20261106)
    for k, a0 in enumerate(A0):
        pos = rng.uniform([-6, 2, -14], [6, 8, -2])
        aim = rng.uniform([-3, -2.5, -12], [3, 0, -5]) - pos
        aim /= np.linalg.norm(aim)
        theta = [0.35, 0.6, 1.0, 1.4, 2.0, 3.0][k % 6]          # radians, as C/objects.c:179 takes it (cos(theta))
        col = rng.uniform(0.8, 2.5, 3)
        keep.append(f"light, color: [{col[0]:.4f}, {col[1]:.4f}, {col[2]:.4f}], position: [{pos[0]:.4f}, {pos[1]:.4f}, {pos[2]:.4f}], "
                    f"radial-a0: 0.2, radial-a1: 0.02, radial-a2: 0.005, direction: [{aim[0]:.5f}, {aim[1]:.5f}, {aim[2]:.5f}], "
                    f"theta: {theta}, angular-a0: {a0}")
    path = os.path.join(HERE, "scenes", "zoo_spots.scene")
    with open(path, "w") as f:
        f.write("\n".join(keep) + "\n")
    img = rtlib.ref_render(path, W, H)
    np.savez_compressed(os.path.join(HERE, "zoo_spots.npz"), zoo_spots=img)
    lit = (img.reshape(-1, 3).max(axis=1) > 0).mean()
    print("zoo_spots:", img.shape, "non-black fraction", round(float(lit), 3), "distinct values", len(np.unique(img)))


if __name__ == "__main__":
    main()
