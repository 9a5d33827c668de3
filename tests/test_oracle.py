"""CPU tests of the oracle: against the committed golden vectors generated from the
patched reference (tests/golden/make_golden.py), and -- when /root/reference was
available to build oracle/_ref/libref.so -- against the reference itself."""
import ctypes as C
import hashlib
import json
import os

import numpy as np
import pytest

import rtlib

SHIPPED = ["simple", "reflection", "quadric", "example2", "example3", "quadric2"]


def scene_path(name):
    return os.path.join(rtlib.SCENES, f"ref_{name}.scene")


@pytest.mark.parametrize("name", SHIPPED)
def test_oracle_matches_golden_image(name):
    gold = np.load(os.path.join(rtlib.GOLDEN, "images.npz"))[name]
    img = rtlib.oracle_render(scene_path(name), 160, 120)
    assert (img == gold).all(), rtlib.compare(img, gold)


@pytest.mark.parametrize("name,file,n", [("c4_syn1032", "c4_syn1032.scene", 48), ("c5_syn4096", "c5_syn4096.scene", 32)])
def test_oracle_matches_golden_synthetic(name, file, n):
    p = os.path.join(rtlib.ROOT, "scenes", file)
    h = json.load(open(os.path.join(rtlib.GOLDEN, "hashes.json")))[name]
    assert hashlib.md5(open(p, "rb").read()).hexdigest() == h["scene_md5"], "scene file drifted from the frozen recipe"
    gold = np.load(os.path.join(rtlib.GOLDEN, "images.npz"))[name]
    img = rtlib.oracle_render(p, n, n)
    assert (img == gold).all(), rtlib.compare(img, gold)


@pytest.mark.parametrize("name", ["simple", "quadric"])
def test_oracle_matches_golden_hash_500(name):
    h = json.load(open(os.path.join(rtlib.GOLDEN, "hashes.json")))[name]
    img, ct = rtlib.oracle_render(scene_path(name), 500, 500, counters=True)
    assert hashlib.md5(img.tobytes()).hexdigest() == h["md5"]
    assert ct.as_dict() == h["counters"]


def test_ray_counts_match_survey():
    """SURVEY.md section 8d measured these on the patched reference at full size; the
    500x500 counters in hashes.json scale the same way (primary = pixels)."""
    h = json.load(open(os.path.join(rtlib.GOLDEN, "hashes.json")))
    for name in SHIPPED:
        assert h[name]["counters"]["rays_primary"] == 250000


def test_known_answer_vectors():
    k = np.load(os.path.join(rtlib.GOLDEN, "kat.npz"))
    o = rtlib.oracle()
    fp = C.POINTER(C.c_float)
    t = C.c_float()
    n = len(k["rad"])
    bad = 0
    for i in range(n):
        oo, dd = k["o"][i].copy(), k["d"][i].copy()
        a, b = oo.ctypes.data_as(fp), dd.ctypes.data_as(fp)
        cen, nrm, q = k["cen"][i].copy(), k["nrm"][i].copy(), k["q"][i].copy()
        for kind, hit, tt, call in (
                ("sph", k["sph_hit"][i], k["sph_t"][i], lambda: o.orc_isect_sphere(a, b, cen.ctypes.data_as(fp), float(k["rad"][i]), C.byref(t))),
                ("pln", k["pln_hit"][i], k["pln_t"][i], lambda: o.orc_isect_plane(a, b, cen.ctypes.data_as(fp), nrm.ctypes.data_as(fp), C.byref(t))),
                ("quad", k["quad_hit"][i], k["quad_t"][i], lambda: o.orc_isect_quadric(a, b, q.ctypes.data_as(fp), C.byref(t)))):
            t.value = 0
            got = call()
            same_t = (np.float32(t.value).tobytes() == np.float32(tt).tobytes()) or (np.isnan(t.value) and np.isnan(tt))
            if got != hit or (hit and not same_t):
                bad += 1
    assert bad == 0
    assert k["sph_hit"].sum() > 100 and k["quad_hit"].sum() > 100 and (k["sph_hit"] == 0).sum() > 100


@pytest.mark.skipif(not rtlib.ref_available(), reason="oracle/_ref/libref.so not built (no /root/reference)")
@pytest.mark.parametrize("name,W,H", [("simple", 211, 97), ("reflection", 257, 130), ("quadric", 97, 211),
                                      ("example2", 128, 128), ("example3", 90, 60), ("quadric2", 120, 80)])
def test_oracle_equals_reference(name, W, H):
    a = rtlib.ref_render(scene_path(name), W, H)
    b = rtlib.oracle_render(scene_path(name), W, H)
    assert (a == b).all(), rtlib.compare(a, b)


@pytest.mark.skipif(not rtlib.ref_available(), reason="oracle/_ref/libref.so not built (no /root/reference)")
def test_oracle_equals_reference_synthetic_rows():
    p = os.path.join(rtlib.ROOT, "scenes", "c4_syn1032.scene")
    rows = [3, 100, 201]
    a = rtlib.ref_render(p, 256, 256, rows=rows)
    b = rtlib.oracle_render(p, 256, 256, rows=rows)
    assert (a == b).all(), rtlib.compare(a, b)


def test_rows_are_independent():
    """Patched semantics make a pixel a function of the pixel only:
    rendering rows one by one equals rendering the frame."""
    full = rtlib.oracle_render(scene_path("reflection"), 96, 64)
    some = rtlib.oracle_render(scene_path("reflection"), 96, 64, rows=[63, 0, 31])
    assert (some == full[[63, 0, 31]]).all()


ZOO = ["zoo_quadrics", "zoo_spheres", "zoo_lights", "zoo_far"]


@pytest.mark.parametrize("name", ZOO)
def test_oracle_matches_golden_zoo(name):
    """Adversarial scenes (tests/golden/make_zoo.py): restatement vs the reference's own render."""
    gold = np.load(os.path.join(rtlib.GOLDEN, "zoo.npz"))[name]
    img = rtlib.oracle_render(os.path.join(rtlib.SCENES, f"{name}.scene"), gold.shape[1], gold.shape[0])
    rep = rtlib.compare(img, gold)
    assert rep["identical_frac"] == 1.0, rep


def test_oracle_matches_golden_zoo_spots():
    """12 spot lights with non-integer angular-a0 (glibc pow in the reference, C/raycast.c:695): restatement vs the
    reference's own render (tests/golden/make_zoo_spots.py)."""
    gold = np.load(os.path.join(rtlib.GOLDEN, "zoo_spots.npz"))["zoo_spots"]
    img = rtlib.oracle_render(os.path.join(rtlib.SCENES, "zoo_spots.scene"), gold.shape[1], gold.shape[0])
    assert rtlib.compare(img, gold)["identical_frac"] == 1.0
