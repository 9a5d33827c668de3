"""CPU tests of the host side: parser parity and error paths, the C-ABI
library's exported symbols, band ownership helpers, loud failure without a GPU."""
import ctypes as C
import json
import os
import re
import subprocess
import time

import numpy as np
import pytest

import rtlib
from rtlib import pkg

BIN = os.path.join(rtlib.ROOT, "raytracing-programs_b200", "bin")
SHIPPED = ["simple", "reflection", "quadric", "example2", "example3", "quadric2"]


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def scene_path(name):
    return os.path.join(rtlib.SCENES, f"ref_{name}.scene")


def test_abi_exports_every_declared_symbol():
    hdr = open(os.path.join(rtlib.ROOT, "include", "rt_abi.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)          # declarations only, not prose
    declared = set(re.findall(r"\b(rt_[a-z_0-9]+)\s*\(", hdr))
    assert set(pkg.ABI_SYMBOLS) <= declared
    lib = pkg.load_library()
    for sym in declared:
        assert hasattr(lib, sym), f"librt_b200.so does not export {sym}"
    assert lib.rt_abi_version() == 2


def test_struct_layouts_match_header():
    assert C.sizeof(pkg.RtShape) == 4 + 4 * (3 + 3 + 3 + 3 + 10)
    assert C.sizeof(pkg.RtLight) == 4 + 4 * (3 + 3 + 3 + 3 + 3)
    assert C.sizeof(pkg.RtScene) == 32


@pytest.mark.skipif(has_gpu(), reason="checks the no-GPU behaviour")
def test_no_gpu_fails_loudly():
    with pytest.raises(pkg.RtError, match="no CPU path"):
        pkg.Renderer(1)
    lib = pkg.load_library()
    assert lib.rt_render(4, 4, None, None) != 0          # never silently renders on the CPU


def test_band_ownership_partitions_rows():
    for H in (1, 31, 32, 33, 203, 1000):
        for n in (1, 2, 3, 8):
            seen = []
            for r in range(n):
                rows = pkg.rank_rows(H, r, n)
                assert len(rows) == pkg.load_library().rt_rank_num_rows(H, r, n)
                seen += rows
            assert sorted(seen) == list(range(H))
    bh = pkg.load_library().rt_band_height()
    assert pkg.rank_rows(4 * bh, 1, 2) == list(range(bh, 2 * bh)) + list(range(3 * bh, 4 * bh))


@pytest.mark.parametrize("name", SHIPPED)
def test_parser_reads_shipped_scenes(name):
    sc = pkg.load_scene(scene_path(name))
    assert sc.num_shapes >= 3 and sc.num_lights >= 1 and sc.camera_width > 0


@pytest.mark.skipif(not rtlib.ref_available(), reason="needs the reference parser (oracle/_ref/libref.so)")
@pytest.mark.parametrize("path", [os.path.join(rtlib.SCENES, f"ref_{n}.scene") for n in SHIPPED] +
                         [os.path.join(rtlib.ROOT, "scenes", "c4_syn1032.scene"), os.path.join(rtlib.ROOT, "scenes", "c5_syn4096.scene")])
def test_parser_bit_identical_to_reference(path):
    r = rtlib.ref()
    h = r.ref_scene_load(os.fsencode(path))
    sc = pkg.load_scene(path)
    assert r.ref_scene_num_shapes(h) == sc.num_shapes and r.ref_scene_num_lights(h) == sc.num_lights
    assert np.float32(r.ref_scene_cam_w(h)) == np.float32(sc.camera_width)
    shapes = np.zeros((sc.num_shapes, 23), np.float32)
    lights = np.zeros((sc.num_lights, 18), np.float32)
    r.ref_scene_dump_shapes(h, shapes.ctypes.data_as(C.c_void_p))
    r.ref_scene_dump_lights(h, lights.ctypes.data_as(C.c_void_p))
    r.ref_scene_free(h)
    for i in range(sc.num_shapes):
        s = sc.shapes[i]
        nu = {0: 1, 1: 3, 2: 10}[s.type]
        mine = [float(s.type)] + list(s.diffuse) + list(s.specular) + (list(s.position) if s.type != 2 else [0, 0, 0]) + \
            [s.reflectivity, s.refractivity, s.ior] + list(s.params)[:nu] + [0.0] * (10 - nu)
        assert np.array(mine, np.float32).tobytes() == shapes[i].tobytes(), (i, mine, shapes[i])
    for i in range(sc.num_lights):
        l = sc.lights[i]
        mine = [float(l.type)] + list(l.position) + list(l.color) + list(l.radial) + [l.theta, l.cos_theta, l.a0] + list(l.direction) + [0, 0]
        assert np.array(mine, np.float32).tobytes() == lights[i].tobytes(), (i, mine, lights[i])


BAD = {
    "camera, width: 2\n": "Error: A camera width or height was not given.",
    "camera, width: 2, height: 2\nsphere, radius: 1, position: [0, 0, -5], reflectivity: 0, refractivity: 0\n": "Error: A field for a sphere was missing.",
    "camera, width: 2, height: 2\nsphere, radius: 1, position: [0, 0, -5], reflectivity: 2, refractivity: 0, ior: 1\n": "Error: invalid reflectivity for a sphere. Must be between 0 and 1.",
    "camera, width: 2, height: 2\nsphere, radius: 1, position: [0, 0, -5], reflectivity: 0, refractivity: -1, ior: 1\n": "Error: invalid refractivity for a sphere. Must be between 0 and 1.",
    "camera, width: 2, height: 2\nplane, position: [0, 0, -5], reflectivity: 0\n": "Error: A field for a plane was missing.",
    "camera, width: 2, height: 2\nquadric, a: 1, b: 1, c: 1, reflectivity: 0\n": "Error: A field for a quadric was missing.",
    "camera, width: 2, height: 2\nlight, color: [1, 1, 1], position: [0, 0, 0], radial-a0: 1\n": "Error: A field for a light was missing.",
    "camera, width: 2, height: 2\nlight, color: [1, 1, 1], theta: 5, position: [0, 0, 0], radial-a0: 1, radial-a1: 0, radial-a2: 0\n": "Error: A field for a light was missing.",
    "sphere, radius: 1, position: [0, 0, -5], reflectivity: 0, refractivity: 0, ior: 1\n": "Error: The scene must have one and only one camera.",
    "camera, width: 2, height: 2\ncamera, width: 2, height: 2\n": "Error: The scene must have one and only one camera.",
}


@pytest.mark.parametrize("text,msg", list(BAD.items()))
def test_parser_error_paths(tmp_path, text, msg):
    p = tmp_path / "bad.scene"
    p.write_text(text)
    r = subprocess.run([os.path.join(BIN, "scene_dump"), str(p)], capture_output=True, text=True)
    assert r.returncode == 1 and r.stderr.strip() == msg
    ref_cli = os.path.join(rtlib.ROOT, "oracle", "_ref", "raytrace_ref")
    if os.path.exists(ref_cli):                           # the reference says the same thing
        q = subprocess.run([ref_cli, "4", "4", str(p), str(tmp_path / "o.ppm")], capture_output=True, text=True)
        assert q.returncode == 1 and q.stderr.strip() == msg


def test_parser_quirks(tmp_path):
    p = tmp_path / "q.scene"
    p.write_text("camera,width: 3.0,height: 3.0\n\n"
                 "plane, radius: 2.0, normal: [0, 4, 0], diffuse_color: [0, 1, 0], position: [0, -1, 0], reflectivity: 0\n"
                 "light, color:[2,2,2], theta:0, radial-a2:0.075, radial-a1:0.125, radial-a0:0.125, position:[1,3,-3]\n"
                 "light, color: [2, 2, 2], radial-a2: 0.075, radial-a1: 0.125, radial-a0: 0.125, position: [0, 0, -1], theta: 5, angular-a0: 2, direction: [0, 0, -1]")
    sc = pkg.load_scene(str(p))
    assert sc.num_shapes == 1 and sc.num_lights == 2
    assert list(sc.shapes[0].specular) == [0, 0, 0] and list(sc.shapes[0].params)[:3] == [0, 4, 0]
    assert sc.lights[0].type == 0 and sc.lights[1].type == 1
    theta = np.float32(np.float64(np.float32(5.0)) * (180.0 / np.float64(np.float32(3.141592654))))
    assert np.float32(sc.lights[1].theta) == theta
    assert np.float32(sc.lights[1].cos_theta) == np.float32(np.cos(np.float64(theta)))


def test_cli_usage_and_open_errors(tmp_path):
    exe = os.path.join(BIN, "raytrace")
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout == "Usage: raytrace WIDTH HEIGHT INPUT_SCENE OUTPUT_IMAGE\n"
    r = subprocess.run([exe, "4", "4", str(tmp_path / "none.scene"), str(tmp_path / "o.ppm")], capture_output=True, text=True)
    assert r.returncode == 1 and r.stderr.startswith("Error: Unable to open the input scene file: ")
    r = subprocess.run([exe, "4", "4", scene_path("simple"), str(tmp_path / "nodir" / "o.ppm")], capture_output=True, text=True)
    assert r.returncode == 1 and r.stderr.startswith("Error: Unable to open the output image file: ")


def test_parser_hardening(tmp_path):
    """inputs that crash or hang the reference parser (8c 'known fragilities') are read
    here without changing what well-formed files mean: fields longer than the reference's 100-byte buffers, a
    field without ':', an unclosed '[', no trailing newline, CRLF-free blank lines, and a 5000-object file
    (the reference appends recursively, one stack frame per node)."""
    long_val = "0." + "0" * 150 + "5"                        # > 99 characters in one field
    p = tmp_path / "h.scene"
    p.write_text("camera, width: 2, height: 2\n"
                 f"sphere, radius: 1, position: [0, 0, -5], diffuse_color: [{long_val}, 1, 0], reflectivity: 0, refractivity: 0, ior: 1\n"
                 "sphere, radius: 1, stray field, position: [1, 0, -5], reflectivity: 0, refractivity: 0, ior: 1\n"
                 "plane, normal: [0, 1, 0], position: [0, -1, 0, reflectivity: 0\n"
                 "light, color: [1, 1, 1], position: [0, 3, 0], radial-a0: 1, radial-a1: 0, radial-a2: 0")
    r = subprocess.run([os.path.join(BIN, "scene_dump"), str(p)], capture_output=True, text=True, timeout=20)
    # the unclosed '[' swallows the rest of its line: the plane then lacks reflectivity -> the reference's message
    assert r.returncode == 1 and r.stderr.strip() == "Error: A field for a plane was missing."
    p.write_text(p.read_text().replace("position: [0, -1, 0, reflectivity", "position: [0, -1, 0], reflectivity"))
    sc = pkg.load_scene(str(p))
    assert sc.num_shapes == 3 and sc.num_lights == 1
    assert abs(sc.shapes[0].diffuse[0] - float(long_val)) < 1e-12 and sc.shapes[1].position[0] == 1.0
    big = tmp_path / "big.scene"
    lines = ["camera, width: 2, height: 2"] + [
        f"sphere, radius: 0.5, position: [{i % 70 - 35}, {i // 70 - 35}, -50], diffuse_color: [1, 0, 0], reflectivity: 0, refractivity: 0, ior: 1"
        for i in range(5000)]
    big.write_text("\n".join(lines))
    t0 = time.perf_counter()
    sc = pkg.load_scene(str(big))
    assert sc.num_shapes == 5000 and sc.shapes[4999].position[0] == float(4999 % 70 - 35)
    assert time.perf_counter() - t0 < 5.0


# ------------------------------------------------------- binary writers (C/ppm.c:193-227)
class _PPM(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("size", C.c_int), ("maxColor", C.c_uint8),
                ("depth", C.c_uint8), ("tupleType", C.c_char_p), ("pixmap", C.POINTER(C.c_uint8))]


def _write_with(lib, fn, W, H, depth, px, path, max_color=255):
    libc = C.CDLL(None)
    libc.fopen.restype = C.c_void_p
    libc.fopen.argtypes = [C.c_char_p, C.c_char_p]
    libc.fclose.argtypes = [C.c_void_p]
    f = getattr(lib, fn)
    f.argtypes = [_PPM, C.c_void_p]
    f.restype = None
    fp = libc.fopen(os.fsencode(path), b"wb")
    f(_PPM(W, H, W * H * 3, max_color, depth, None, px.ctypes.data_as(C.POINTER(C.c_uint8))), fp)
    libc.fclose(fp)
    return open(path, "rb").read()


BIN_CASES = [(1, 1), (7, 5), (640, 480), (1000, 263)]      # the last one spans several writer chunks


def _rgba(W, H, alpha=255):
    rgb = rtlib.p3_pixmap(W, H, "random").reshape(-1, 3)
    return np.ascontiguousarray(np.concatenate([rgb, np.full((W * H, 1), alpha, np.uint8)], axis=1)), np.ascontiguousarray(rgb)


def test_binary_writers_golden(tmp_path):
    """ppm_WriteOutP6 / ppm_WriteOutP7 of the host against hashes of the files the reference's own
    writers produce for the same RGBA pixmaps (tests/golden/make_bin_golden.py), and the RGB entry
    (depth 3, the raytracer's pixmap) against the same files."""
    import hashlib
    h = rtlib.pkg.load_host_library()
    gold = json.load(open(os.path.join(rtlib.GOLDEN, "bin_hashes.json")))
    for W, H in BIN_CASES:
        rgba, rgb = _rgba(W, H)
        for fn in ("ppm_WriteOutP6", "ppm_WriteOutP7"):
            a = _write_with(h, fn, W, H, 4, rgba, tmp_path / "a.ppm")
            b = _write_with(h, fn, W, H, 3, rgb, tmp_path / "b.ppm")
            g = gold[f"{fn}_{W}x{H}"]
            assert hashlib.md5(a).hexdigest() == g["md5"] and len(a) == g["bytes"], (fn, W, H)
            assert a == b, (fn, W, H)
    # known answer
    px = np.array([1, 2, 3], np.uint8)
    assert _write_with(h, "ppm_WriteOutP7", 1, 1, 3, px, tmp_path / "c.ppm", 200) == \
        b"P7\nWIDTH 1\nHEIGHT 1\nDEPTH 4\nMAXVAL 200\nTUPLTYPE RGB_ALPHA\n\x01\x02\x03\xc8\n"
    assert _write_with(h, "ppm_WriteOutP6", 1, 1, 3, px, tmp_path / "d.ppm") == b"P6\n1 1\n255\n\x01\x02\x03\n"


@pytest.mark.skipif(not rtlib.ref_available(), reason="oracle/_ref not built")
def test_binary_writers_match_reference(tmp_path):
    h = rtlib.pkg.load_host_library()
    r = rtlib.ref()
    for W, H in BIN_CASES:
        rgba, _ = _rgba(W, H, alpha=77)
        for fn in ("ppm_WriteOutP6", "ppm_WriteOutP7"):
            assert _write_with(h, fn, W, H, 4, rgba, tmp_path / "a.ppm") == \
                _write_with(r, fn, W, H, 4, rgba, tmp_path / "r.ppm"), (fn, W, H)


def test_light_buffer_sizes_diagnostic():
    """rt_dbg_light_buffer_sizes builds the sphere light buffer on the host (no GPU): six levels per light, list
    sizes growing with the distance level (farther hit points need more inflated spheres), and none without spheres."""
    lib = pkg.load_library()
    lib.rt_dbg_light_buffer_sizes.restype = C.c_longlong
    lib.rt_dbg_light_buffer_sizes.argtypes = [C.POINTER(pkg.RtScene), C.POINTER(C.c_longlong)]
    sc = pkg.load_scene(os.path.join(rtlib.ROOT, "scenes", "c4_syn1032.scene"))
    per = (C.c_longlong * (sc.num_lights * 6))()
    total = lib.rt_dbg_light_buffer_sizes(C.byref(sc), per)
    a = np.array(list(per)).reshape(sc.num_lights, 6)
    assert total == a.sum() and (a > 0).all()
    assert (np.diff(a, axis=1) >= 0).all()
    assert 1.0 < a[:, 0].mean() / 6144 < 4.0                   # a few spheres per cell for nearby hit points
    sc2 = pkg.load_scene(scene_path("quadric"))                 # 1 sphere
    assert lib.rt_dbg_light_buffer_sizes(C.byref(sc2), None) > 0
