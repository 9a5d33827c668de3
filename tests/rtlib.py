"""Test-side helpers: ctypes bindings of the oracle (oracle/liboracle.so), of the patched
reference (oracle/_ref/libref.so, optional) and of the product package."""
import ctypes as C
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
pkg = importlib.import_module("raytracing-programs_b200")

GOLDEN = os.path.join(ROOT, "tests", "golden")
SCENES = os.path.join(GOLDEN, "scenes")
REF_ROOT = "/root/reference"


class OrcCounters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("rays_primary", "rays_reflection", "rays_shadow", "hit_points",
                                           "tests_sphere", "tests_plane", "tests_quadric", "shadow_occluded")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


_oracle = None
_ref = None


def oracle():
    global _oracle
    if _oracle is None:
        path = os.path.join(ROOT, "oracle", "liboracle.so")
        if not os.path.exists(path):
            raise RuntimeError("oracle/liboracle.so missing: run `make -C oracle` (or __graft_entry__.build())")
        o = C.CDLL(path)
        o.orc_render_rows.argtypes = [C.POINTER(pkg.RtScene), C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                      C.POINTER(OrcCounters)]
        o.orc_render_row_list.argtypes = [C.POINTER(pkg.RtScene), C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int,
                                          C.c_void_p, C.c_int, C.POINTER(OrcCounters)]
        fp = C.POINTER(C.c_float)
        o.orc_isect_sphere.argtypes = [fp, fp, fp, C.c_float, fp]
        o.orc_isect_plane.argtypes = [fp, fp, fp, fp, fp]
        o.orc_isect_quadric.argtypes = [fp, fp, fp, fp]
        _oracle = o
    return _oracle


def ref_available():
    return os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref.so"))


_ref_d50 = None


def ref(depth=7):
    """The patched reference harness; depth=50: the same sources with MAX_RECURSION 50 (oracle/build_ref.sh)."""
    global _ref, _ref_d50
    if depth == 50:
        if _ref_d50 is None:
            r = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_d50.so"))
            r.ref_scene_load.restype = C.c_void_p
            r.ref_scene_load.argtypes = [C.c_char_p]
            r.ref_scene_free.argtypes = [C.c_void_p]
            r.ref_render_row_list.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_void_p, C.c_int]
            _ref_d50 = r
        return _ref_d50
    assert depth == 7
    if _ref is None:
        r = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref.so"))
        r.ref_scene_load.restype = C.c_void_p
        r.ref_scene_load.argtypes = [C.c_char_p]
        r.ref_scene_free.argtypes = [C.c_void_p]
        for f in ("ref_scene_num_shapes", "ref_scene_num_lights"):
            getattr(r, f).argtypes = [C.c_void_p]
        for f in ("ref_scene_cam_w", "ref_scene_cam_h"):
            getattr(r, f).argtypes = [C.c_void_p]
            getattr(r, f).restype = C.c_float
        r.ref_scene_dump_shapes.argtypes = [C.c_void_p, C.c_void_p]
        r.ref_scene_dump_lights.argtypes = [C.c_void_p, C.c_void_p]
        r.ref_render_rows.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        r.ref_render_row_list.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_void_p, C.c_int]
        fp = C.POINTER(C.c_float)
        r.ref_isect_sphere.argtypes = [fp, fp, fp, C.c_float, fp]
        r.ref_isect_plane.argtypes = [fp, fp, fp, fp, fp]
        r.ref_isect_quadric.argtypes = [fp, fp, fp, fp]
        _ref = r
    return _ref


def ncores():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def oracle_render(scene, W, H, rows=None, threads=None, counters=False, depth=7, mode=0):
    """Oracle image (all rows, or the given list of rows) as uint8 [n, W, 3].  mode=1: the reference's dormant
    recursive shoot() (C/raycast.c:152-296) instead of the live iterative_shoot."""
    oracle().orc_set_max_level(int(depth))
    oracle().orc_set_mode(int(mode))
    if isinstance(scene, (str, os.PathLike)):
        scene = pkg.load_scene(scene)
    rows = list(range(H)) if rows is None else list(rows)
    out = np.zeros((len(rows), W, 3), dtype=np.uint8)
    arr = (C.c_int * len(rows))(*rows)
    ct = OrcCounters()
    oracle().orc_render_row_list(C.byref(scene), W, H, arr, len(rows), out.ctypes.data_as(C.c_void_p),
                                 threads or ncores(), C.byref(ct))
    oracle().orc_set_max_level(7)
    oracle().orc_set_mode(0)
    return (out, ct) if counters else out


def ref_render(scene_path, W, H, rows=None, threads=None, depth=7, mode=0):
    """The patched reference itself.  mode=1: through its dormant recursive shoot() (oracle/ref_harness.c)."""
    r = ref(depth)
    if mode:
        r.ref_set_mode(1)
    h = r.ref_scene_load(os.fsencode(scene_path))
    if not h:
        raise RuntimeError(f"reference cannot open {scene_path}")
    rows = list(range(H)) if rows is None else list(rows)
    out = np.zeros((len(rows), W, 3), dtype=np.uint8)
    arr = (C.c_int * len(rows))(*rows)
    r.ref_render_row_list(h, W, H, arr, len(rows), out.ctypes.data_as(C.c_void_p), threads or ncores())
    if mode:
        r.ref_set_mode(0)
    r.ref_scene_free(h)
    return out


def compare(a, b):
    """Parity report between two uint8 [.., W, 3] images."""
    a = np.asarray(a).reshape(-1, 3).astype(np.int16)
    b = np.asarray(b).reshape(-1, 3).astype(np.int16)
    diff = np.abs(a - b).max(axis=1)
    n = diff.size
    bad = int((diff > 0).sum())
    return {"pixels": n, "identical": n - bad, "identical_frac": (n - bad) / max(n, 1),
            "max_abs_diff": int(diff.max()) if n else 0, "gt1": int((diff > 1).sum()),
            "divergent_idx": np.nonzero(diff > 0)[0][:64].tolist()}


def fvec(*v):
    return (C.c_float * len(v))(*v)


def oracle_p3(pixmap, W, H, max_color=255):
    """P3 text of a pixmap by the CPU restatement (oracle/rt_oracle.c: orc_p3_format)."""
    o = oracle()
    o.orc_p3_format.restype = C.c_size_t
    o.orc_p3_format.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_uint, C.c_void_p, C.c_size_t]
    pixmap = np.ascontiguousarray(pixmap, dtype=np.uint8)
    assert pixmap.size == W * H * 3
    cap = 64 + 4 * pixmap.size + 16
    buf = C.create_string_buffer(cap)
    n = o.orc_p3_format(pixmap.ctypes.data_as(C.c_void_p), W, H, max_color, buf, cap)
    assert n > 0
    return buf.raw[:n]


def ref_p3(pixmap, W, H, max_color=255):
    """P3 file written by the reference's own ppm_WriteOutP3 (oracle/_ref/libref.so)."""
    import tempfile
    r = ref()
    r.ref_write_p3.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
    pixmap = np.ascontiguousarray(pixmap, dtype=np.uint8)
    with tempfile.NamedTemporaryFile(suffix=".ppm") as f:
        assert r.ref_write_p3(os.fsencode(f.name), W, H, max_color, pixmap.ctypes.data_as(C.c_void_p)) == 0
        return open(f.name, "rb").read()


def p3_pixmap(W, H, kind, seed=7):
    """Seeded test pixmaps for the P3 writer: every text length and chunk-boundary pattern."""
    n = W * H * 3
    rng = np.random.default_rng(seed)
    if kind == "random":
        a = rng.integers(0, 256, n, dtype=np.uint8)
    elif kind == "zeros":
        a = np.zeros(n, dtype=np.uint8)
    elif kind == "max":
        a = np.full(n, 255, dtype=np.uint8)
    elif kind == "ramp":
        a = (np.arange(n) % 256).astype(np.uint8)
    elif kind == "edges":      # the length boundaries 0/9/10/99/100/255 in random order
        a = rng.choice(np.array([0, 9, 10, 99, 100, 255], dtype=np.uint8), n)
    elif kind == "runs":       # long runs of one length class, so whole chunks are 2, 3 or 4 bytes per value
        a = np.repeat(rng.choice(np.array([5, 50, 200], dtype=np.uint8), n // 3000 + 1), 3000)[:n]
    else:
        raise ValueError(kind)
    return a.reshape(H, W, 3)
