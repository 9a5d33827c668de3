"""P3 writer: the oracle restatement against the reference's own
ppm_WriteOutP3 and the committed golden hashes (CPU), the host C writer against the oracle (CPU),
and the GPU formatter k_p3_format against the oracle through the C ABI (GPU)."""
import ctypes as C
import hashlib
import json
import os

import numpy as np
import pytest

import rtlib

GOLD = os.path.join(rtlib.GOLDEN, "p3_hashes.json")
CASES = [(1, 1, "random"), (1, 5, "edges"), (37, 11, "random"), (64, 64, "ramp"), (64, 64, "zeros"), (64, 64, "max"),
         (1365, 1, "random"), (1366, 1, "edges"), (455, 3, "max"), (200, 150, "runs"), (640, 480, "random")]


def _key(W, H, kind):
    return f"{W}x{H}_{kind}"


def test_oracle_p3_matches_golden():
    gold = json.load(open(GOLD))
    for W, H, kind in CASES:
        txt = rtlib.oracle_p3(rtlib.p3_pixmap(W, H, kind), W, H)
        assert hashlib.md5(txt).hexdigest() == gold[_key(W, H, kind)]["md5"], (W, H, kind)
        assert len(txt) == gold[_key(W, H, kind)]["bytes"]


@pytest.mark.skipif(not rtlib.ref_available(), reason="oracle/_ref not built")
def test_oracle_p3_matches_reference_writer():
    for W, H, kind in CASES:
        pm = rtlib.p3_pixmap(W, H, kind)
        assert rtlib.oracle_p3(pm, W, H) == rtlib.ref_p3(pm, W, H), (W, H, kind)


def test_oracle_p3_known_answer():
    pm = np.array([[[0, 9, 10], [99, 100, 255]]], dtype=np.uint8)
    assert rtlib.oracle_p3(pm, 2, 1) == b"P3\n2 1 \n255\n0\n9\n10\n99\n100\n255\n"


# ---------------------------------------------------------------------------------- GPU
@pytest.fixture(scope="module")
def renderer():
    r = rtlib.pkg.Renderer(1)
    yield r
    r.close()


GPU_CASES = CASES + [(4096, 1, "random"), (1365, 2, "runs"), (2731, 1, "edges"), (1000, 1000, "random"), (1024, 768, "runs"),
                     (1920, 1080, "edges"), (3, 100000, "ramp")]


@pytest.mark.gpu
@pytest.mark.timeout(300)
def test_gpu_p3_matches_oracle(renderer):
    for W, H, kind in GPU_CASES:
        pm = rtlib.p3_pixmap(W, H, kind)
        txt, st = renderer.p3_text(W, H, pm)
        ref = rtlib.oracle_p3(pm, W, H)
        assert st.text_bytes == len(ref), (W, H, kind)
        assert txt == ref, (W, H, kind)


@pytest.mark.gpu
@pytest.mark.timeout(300)
def test_gpu_p3_max_color_and_repeat(renderer):
    pm = rtlib.p3_pixmap(333, 77, "random", seed=3)
    for mc in (255, 1, 0, 300):          # maxColor is a uint8_t in PPMFormat: 300 prints as 44
        assert renderer.p3_text(333, 77, pm, max_color=mc)[0] == rtlib.oracle_p3(pm, 333, 77, mc)
    a = renderer.p3_text(333, 77, pm)[0]
    for _ in range(5):                   # look-back state is reset on every call
        assert renderer.p3_text(333, 77, pm)[0] == a


@pytest.mark.gpu
@pytest.mark.timeout(300)
def test_gpu_p3_of_resident_frame(renderer):
    """raycast() then ppm_WriteOutP3 without re-uploading: the text of the frame left on GPU 0."""
    scene = os.path.join(rtlib.SCENES, "ref_reflection.scene")
    renderer.upload(scene)
    img, _ = renderer.raycast(400, 300)
    txt, st = renderer.p3_text(400, 300)
    assert st.ms_h2d == 0
    assert txt == rtlib.oracle_p3(img, 400, 300)
    with pytest.raises(rtlib.pkg.RtError):
        renderer.p3_text(401, 300)        # no resident frame of that size


@pytest.mark.gpu
@pytest.mark.timeout(600)
def test_gpu_p3_large_frame_properties(renderer):
    """4096 x 4096 (50 M values, > 12 k chunks, several D2H slabs): exact length, header, and text
    windows at offsets computed independently (numpy prefix sum of the value lengths)."""
    W = H = 4096
    pm = rtlib.p3_pixmap(W, H, "random", seed=11)
    pm[1000:1400] = 7            # long 2-byte stretch
    pm[2000:2200] = 123          # long 4-byte stretch
    txt, st = renderer.p3_text(W, H, pm)
    flat = pm.reshape(-1)
    lens = 2 + (flat > 9).astype(np.int64) + (flat > 99)
    hdr = b"P3\n4096 4096 \n255\n"
    assert txt.startswith(hdr)
    assert len(txt) == len(hdr) + int(lens.sum()) == st.text_bytes
    offs = len(hdr) + np.concatenate(([0], np.cumsum(lens)[:-1]))
    rng = np.random.default_rng(5)
    for i in np.concatenate((rng.integers(0, flat.size - 64, 400), [0, flat.size - 64])):
        want = b"".join(b"%d\n" % v for v in flat[i:i + 64])
        assert txt[offs[i]:offs[i] + len(want)] == want, int(i)


@pytest.mark.gpu
@pytest.mark.timeout(300)
def test_gpu_host_ppm_writeoutp3(renderer, tmp_path):
    """The drop-in entry point itself: host/ppm_out.c ppm_WriteOutP3(PPMFormat, FILE*) on a foreign pixmap."""
    h = rtlib.pkg.load_host_library()

    class PPM(C.Structure):
        _fields_ = [("width", C.c_int), ("height", C.c_int), ("size", C.c_int), ("maxColor", C.c_uint8),
                    ("depth", C.c_uint8), ("tupleType", C.c_char_p), ("pixmap", C.POINTER(C.c_uint8))]
    libc = C.CDLL(None)
    libc.fopen.restype = C.c_void_p
    libc.fopen.argtypes = [C.c_char_p, C.c_char_p]
    libc.fclose.argtypes = [C.c_void_p]
    h.ppm_WriteOutP3.argtypes = [PPM, C.c_void_p]
    for W, H, kind in ((7, 5, "ramp"), (640, 480, "random"), (1365, 3, "edges")):
        px = rtlib.p3_pixmap(W, H, kind)
        out = tmp_path / "w.ppm"
        f = libc.fopen(os.fsencode(out), b"wb")
        h.ppm_WriteOutP3(PPM(W, H, W * H * 3, 255, 3, None, px.ctypes.data_as(C.POINTER(C.c_uint8))), f)
        libc.fclose(f)
        assert out.read_bytes() == rtlib.oracle_p3(px, W, H)
