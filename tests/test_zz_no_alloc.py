"""rt_p3_format_device_noalloc: the P3 formatter for hosts that cannot allocate GPU memory.

Some programs that embed this library own all of the GPU's memory through a pool of their own (a cudaMalloc made
behind their back fails or steals from them), and some run a frame loop in which the device-wide stall of
cudaMalloc / cudaFree is not acceptable.  For them the library has one entry point whose span table is a fixed
array inside the library image rather than a buffer grown with cudaMalloc.  These tests drive it through the C ABI
and hold its text against the text of the allocating entry point, rt_p3_format_device, byte for byte (the P3 path
is integer work: the tolerance is zero), on calls made back to back so that what one call leaves in the shared
table is seen to do no harm to the next one.

The file name sorts last on purpose: pytest -x stops at the first failure, and this test should never hide the
parity tests.  It passes on its own in a fresh process (the new entry point needs no rt_init; the comparison side
brings the library up through the shared `renderer` fixture) and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

import rtlib


def _texts(renderer, W, H, kind, seed, max_color=255):
    """P3 text of one pixmap from both device-to-device entry points -> (existing, no-alloc, oracle)."""
    import torch
    pkg = rtlib.pkg
    lib = pkg.load_library()
    pm = rtlib.p3_pixmap(W, H, kind, seed=seed)
    cap = lib.rt_p3_max_bytes(W, H)
    rgb = torch.from_numpy(pm.reshape(-1).copy()).to("cuda:0")
    # 0xEE / 0xDD fill: bytes the formatter fails to write, or writes past the end, cannot pass for text
    text_a = torch.full((cap + 64,), 0xEE, dtype=torch.uint8, device="cuda:0")
    text_b = torch.full((cap + 64,), 0xDD, dtype=torch.uint8, device="cuda:0")
    torch.cuda.synchronize()
    n_a, _ = renderer.p3_format_device(rgb.data_ptr(), W, H, text_a.data_ptr(), cap, max_color=max_color)
    n_b = pkg.p3_format_device_noalloc(rgb.data_ptr(), W, H, text_b.data_ptr(), cap, max_color=max_color)
    a = bytes(text_a[:n_a].cpu().numpy())
    b = bytes(text_b[:n_b].cpu().numpy())
    assert bytes(text_b[n_b:n_b + 64].cpu().numpy()) == b"\xDD" * 64, "no-alloc formatter wrote past the end of its text"
    return a, b, rtlib.oracle_p3(pm, W, H, max_color)


@pytest.mark.gpu
@pytest.mark.timeout(600)
def test_noalloc_p3_equals_allocating_entry_point_twice_in_a_row(renderer):
    """Two calls in a row and then some: a large frame first (many spans of the table used, all-"255" values so
    every length is at its maximum), then smaller and ragged ones that reuse the front of the same table, then the
    first frame again."""
    import torch
    assert torch.cuda.current_device() == 0
    cases = [(1920, 1080, "max", 1), (640, 480, "random", 2), (37, 11, "random", 3), (1, 1, "random", 4),
             (1365, 3, "edges", 5), (1920, 1080, "max", 1), (2731, 1, "edges", 6), (1000, 1000, "random", 7)]
    for W, H, kind, seed in cases:
        a, b, ref = _texts(renderer, W, H, kind, seed)
        assert len(b) == len(a) == len(ref), (W, H, kind)
        assert b == a, (W, H, kind)
        assert b == ref, (W, H, kind)
    # maxColor is a uint8_t in the reference's PPMFormat: 300 prints as 44, in both entry points
    a, b, ref = _texts(renderer, 333, 77, "random", 8, max_color=300)
    assert b == a == ref


@pytest.mark.gpu
@pytest.mark.timeout(600)
def test_noalloc_p3_on_a_caller_stream_without_rt_init_state():
    """The entry point takes the caller's stream and does not depend on the Renderer: same pixmap formatted twice
    on a torch side stream, identical both times and equal to the oracle."""
    import torch
    pkg = rtlib.pkg
    lib = pkg.load_library()
    W, H = 1024, 768
    pm = rtlib.p3_pixmap(W, H, "runs", seed=9)
    ref = rtlib.oracle_p3(pm, W, H)
    cap = lib.rt_p3_max_bytes(W, H)
    side = torch.cuda.Stream(device="cuda:0")
    rgb = torch.from_numpy(pm.reshape(-1).copy()).to("cuda:0")
    text = torch.zeros(cap + 64, dtype=torch.uint8, device="cuda:0")
    torch.cuda.synchronize()
    for _ in range(2):
        text.fill_(0x55)
        torch.cuda.synchronize()
        n = pkg.p3_format_device_noalloc(rgb.data_ptr(), W, H, text.data_ptr(), cap, stream=side.cuda_stream)
        assert n == len(ref)
        assert bytes(text[:n].cpu().numpy()) == ref


@pytest.mark.gpu
@pytest.mark.timeout(300)
def test_noalloc_p3_refuses_what_it_cannot_hold():
    """Frames beyond RT_P3_NOALLOC_MAX_VALUES, short text buffers, host pointers: a clear error, nothing launched."""
    import torch
    pkg = rtlib.pkg
    lib = pkg.load_library()
    rgb = torch.zeros(4096, dtype=torch.uint8, device="cuda:0")
    text = torch.zeros(1 << 16, dtype=torch.uint8, device="cuda:0")
    torch.cuda.synchronize()
    W = H = 18919                                    # 3 * 18919^2 = 1 073 785 683 > 2^30
    assert 3 * W * H > pkg.P3_NOALLOC_MAX_VALUES
    with pytest.raises(pkg.RtError, match="RT_P3_NOALLOC_MAX_VALUES"):
        pkg.p3_format_device_noalloc(rgb.data_ptr(), W, H, text.data_ptr(), lib.rt_p3_max_bytes(W, H))
    with pytest.raises(pkg.RtError, match="text_cap"):
        pkg.p3_format_device_noalloc(rgb.data_ptr(), 32, 32, text.data_ptr(), 100)
    host = np.zeros(4096 + 16, dtype=np.uint8)
    hp = (host.ctypes.data + 15) & ~15
    with pytest.raises(pkg.RtError, match="dev_rgb"):
        pkg.p3_format_device_noalloc(hp, 32, 32, text.data_ptr(), text.numel())
    # and the call after the refusals still works
    n = pkg.p3_format_device_noalloc(rgb.data_ptr(), 32, 32, text.data_ptr(), text.numel())
    assert bytes(text[:n].cpu().numpy()) == rtlib.oracle_p3(np.zeros((32, 32, 3), np.uint8), 32, 32)


def test_noalloc_symbol_and_limit_are_exported():
    """CPU: the library exports the entry point the header declares, and the Python constant matches the header."""
    import os
    import re
    lib = rtlib.pkg.load_library()
    assert hasattr(lib, "rt_p3_format_device_noalloc")
    hdr = open(os.path.join(rtlib.ROOT, "include", "rt_abi.h")).read()
    m = re.search(r"#define RT_P3_NOALLOC_MAX_VALUES \(1ull << (\d+)\)", hdr)
    assert m and (1 << int(m.group(1))) == rtlib.pkg.P3_NOALLOC_MAX_VALUES
    # refusal of an oversized frame needs no GPU: the size check comes before any CUDA call
    rc = lib.rt_p3_format_device_noalloc(C.c_void_p(16), 20000, 20000, 255, C.c_void_p(16), 1 << 40, None, None)
    assert rc == -3 and b"RT_P3_NOALLOC_MAX_VALUES" in lib.rt_last_error()
