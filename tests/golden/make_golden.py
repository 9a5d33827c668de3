"""Regenerate tests/golden/ from the reference itself (needs /root/reference and
oracle/_ref/libref.so = the patched reference compiled by oracle/build_ref.sh).

  python tests/golden/make_golden.py

Writes
  scenes/ref_<name>.scene   the shipped example scenes, re-serialised (keys sorted,
                            spacing normalised, value text untouched) so that this repo
                            carries test inputs, not copies of reference files;
  images.npz                reference renders (patched C/raycast.c) at 160x120 (+ synthetic
                            scenes at 48x48 / 32x32);
  hashes.json               md5 of reference renders at 500x500 and ray/test counters;
  kat.npz                   known-answer vectors of ray_intersect_{sphere,plane,quadric}.
"""
import ctypes as C
import hashlib
import json
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import rtlib  # noqa: E402

REF = "/root/reference"
SHIPPED = {"simple": "examples/simple.scene", "reflection": "examples/reflection.scene",
           "quadric": "examples/quadric.scene", "example2": "C/examples/example2.scene",
           "example3": "C/examples/example3.scene", "quadric2": "C/examples/quadric2.scene"}


def split_fields(line):
    out, depth, cur = [], 0, ""
    for ch in line:
        if ch == "[":
            depth = 1
        elif ch == "]":
            depth = 0
        if ch == "," and not depth:
            out.append(cur)
            cur = ""
        else:
            cur += ch
    out.append(cur)
    return out


def canonical(text):
    lines = []
    for line in text.split("\n"):
        if not line.strip():
            continue
        f = split_fields(line)
        kind, fields = f[0], []
        for fld in f[1:]:
            k, v = fld.split(":", 1)
            fields.append((k.replace(" ", ""), re.sub(r"\s+", " ", v.strip())))
        fields.sort(key=lambda kv: kv[0])
        lines.append(kind + ", " + ", ".join(f"{k}: {v}" for k, v in fields))
    return "\n".join(lines) + "\n"


def main():
    os.makedirs(os.path.join(HERE, "scenes"), exist_ok=True)
    images, hashes = {}, {}
    for name, rel in SHIPPED.items():
        src = os.path.join(REF, rel)
        dst = os.path.join(HERE, "scenes", f"ref_{name}.scene")
        with open(src) as f:
            txt = canonical(f.read())
        with open(dst, "w") as f:
            f.write(txt)
        # the canonical file must mean the same thing to the reference parser
        a = rtlib.ref_render(src, 160, 120)
        b = rtlib.ref_render(dst, 160, 120)
        assert (a == b).all(), name
        images[name] = a
        big = rtlib.ref_render(dst, 500, 500)
        _, ct = rtlib.oracle_render(dst, 500, 500, counters=True)
        hashes[name] = {"size": [500, 500], "md5": hashlib.md5(big.tobytes()).hexdigest(), "counters": ct.as_dict()}
    for name, rel, n in (("c4_syn1032", "scenes/c4_syn1032.scene", 48), ("c5_syn4096", "scenes/c5_syn4096.scene", 32)):
        p = os.path.join(rtlib.ROOT, rel)
        images[name] = rtlib.ref_render(p, n, n)
        hashes[name] = {"size": [n, n], "md5": hashlib.md5(images[name].tobytes()).hexdigest(),
                        "scene_md5": hashlib.md5(open(p, "rb").read()).hexdigest()}
    np.savez_compressed(os.path.join(HERE, "images.npz"), **images)
    with open(os.path.join(HERE, "hashes.json"), "w") as f:
        json.dump(hashes, f, indent=1, sort_keys=True)

    # known-answer vectors straight from the reference intersectors
    rng = np.random.default_rng(7)
    r = rtlib.ref()
    n = 4000
    o = rng.uniform(-5, 5, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    cen = (o + d * rng.uniform(1, 30, (n, 1)).astype(np.float32) + rng.normal(scale=1.0, size=(n, 3))).astype(np.float32)
    rad = rng.uniform(0.2, 2.0, n).astype(np.float32)
    nrm = rng.normal(size=(n, 3)).astype(np.float32)
    q = rng.uniform(-2, 2, (n, 10)).astype(np.float32)
    q[: n // 4, 3:6] = 0
    q[: n // 8, 0] = 0
    q[n // 8: n // 6, :3] = 0  # a_q == 0 branch
    res = {k: np.zeros(n, np.int32) for k in ("sph_hit", "pln_hit", "quad_hit")}
    tt = {k: np.zeros(n, np.float32) for k in ("sph_t", "pln_t", "quad_t")}
    fp = C.POINTER(C.c_float)
    t = C.c_float()
    for i in range(n):
        oo, dd = o[i].ctypes.data_as(fp), d[i].ctypes.data_as(fp)
        t.value = 0
        res["sph_hit"][i] = r.ref_isect_sphere(oo, dd, cen[i].ctypes.data_as(fp), float(rad[i]), C.byref(t)); tt["sph_t"][i] = t.value
        t.value = 0
        res["pln_hit"][i] = r.ref_isect_plane(oo, dd, cen[i].ctypes.data_as(fp), nrm[i].ctypes.data_as(fp), C.byref(t)); tt["pln_t"][i] = t.value
        t.value = 0
        res["quad_hit"][i] = r.ref_isect_quadric(oo, dd, q[i].ctypes.data_as(fp), C.byref(t)); tt["quad_t"][i] = t.value
    np.savez_compressed(os.path.join(HERE, "kat.npz"), o=o, d=d, cen=cen, rad=rad, nrm=nrm, q=q, **res, **tt)
    print("golden written:", sorted(images), {k: v["md5"][:8] for k, v in hashes.items()})


if __name__ == "__main__":
    main()
