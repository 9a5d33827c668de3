"""ctypes binding of the B200 render-loop library (include/rt_abi.h).

Host-side mirror of the one operator the reference has for this path,
``raycast(json_data_t*, PPMFormat)`` (C/raycast.h:8): ``Renderer.raycast(scene, W, H)``
returns the RGB8 pixmap.  Scene files are read by this repo's C parser (same grammar
as C/parse.c) through librt_host.so.  There is no CPU rendering path in this package:
if the CUDA library is missing or no B200 is present, calls raise ``RtError``.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_DIR = os.path.join(_HERE, "lib")


class RtError(RuntimeError):
    pass


class RtShape(C.Structure):
    _fields_ = [("type", C.c_int32), ("diffuse", C.c_float * 3), ("specular", C.c_float * 3),
                ("position", C.c_float * 3), ("reflectivity", C.c_float), ("refractivity", C.c_float),
                ("ior", C.c_float), ("params", C.c_float * 10)]


class RtLight(C.Structure):
    _fields_ = [("type", C.c_int32), ("position", C.c_float * 3), ("color", C.c_float * 3),
                ("radial", C.c_float * 3), ("theta", C.c_float), ("cos_theta", C.c_float),
                ("a0", C.c_float), ("direction", C.c_float * 3)]


class RtScene(C.Structure):
    _fields_ = [("camera_width", C.c_float), ("camera_height", C.c_float),
                ("num_shapes", C.c_int32), ("num_lights", C.c_int32),
                ("shapes", C.POINTER(RtShape)), ("lights", C.POINTER(RtLight))]


class RtStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "pixels", "rays_primary", "rays_reflection", "rays_shadow", "hit_points",
        "sphere_tests_o", "sphere_tests_d", "plane_tests_o", "plane_tests_d",
        "quadric_tests_o", "quadric_tests_d", "kernel_launches")] + [
        ("flops", C.c_double), ("flops_shade", C.c_double), ("ms_kernels", C.c_double),
        ("ms_primary", C.c_double), ("ms_shade", C.c_double), ("ms_finalize", C.c_double),
        ("launches_shade", C.c_uint64), ("ms_gather", C.c_double), ("ms_total", C.c_double),
        ("scan_slots", C.c_uint64), ("ball_tests_o", C.c_uint64), ("ball_tests_d", C.c_uint64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}

    @property
    def rays(self):
        return self.rays_primary + self.rays_reflection + self.rays_shadow


class RtP3Stats(C.Structure):
    _fields_ = [("text_bytes", C.c_uint64), ("ms_h2d", C.c_double), ("ms_format", C.c_double), ("ms_total", C.c_double)]


P3_SINK = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_size_t, C.c_void_p)

ABI_SYMBOLS = ("rt_init", "rt_init_device", "rt_device_count", "rt_shutdown", "rt_last_error",
               "rt_abi_version", "rt_scene_upload", "rt_render", "rt_render_device", "rt_render_rank",
               "rt_assemble_device", "rt_rank_rows_to_host", "rt_band_height", "rt_rank_num_rows", "rt_rank_rows",
               "rt_p3_max_bytes", "rt_p3_format_device", "rt_p3_format_device_noalloc", "rt_p3_write", "rt_set_accel", "rt_set_max_depth", "rt_set_mode",
               "rt_dbg_qindex_classify", "rt_dbg_qindex_prune", "rt_dbg_qindex_visit", "rt_dbg_qindex_last_nodes", "rt_dbg_light_buffer_sizes")

ABI_VERSION = 2     # RT_ABI_VERSION of include/rt_abi.h the structs above mirror

_lib = None
_host = None


def load_library():
    """dlopen librt_b200.so (built in-tree by __graft_entry__.build()); raises if absent."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("RT_B200_LIB") or os.path.join(LIB_DIR, "librt_b200.so")   # env: tuning variants only
    if not os.path.exists(path):
        raise RtError(f"{path} is missing: run `python __graft_entry__.py` (build) first; "
                      "there is no fallback renderer")
    lib = C.CDLL(path, mode=C.RTLD_GLOBAL)
    lib.rt_last_error.restype = C.c_char_p
    if lib.rt_abi_version() != ABI_VERSION:     # the ctypes mirrors below are laid out for exactly this header version
        raise RtError(f"{path} speaks ABI version {lib.rt_abi_version()}, this binding was written for {ABI_VERSION}: rebuild")
    lib.rt_scene_upload.argtypes = [C.POINTER(RtScene)]
    lib.rt_render.argtypes = [C.c_int, C.c_int, C.c_void_p, C.POINTER(RtStats)]
    lib.rt_render_device.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_void_p), C.POINTER(RtStats)]
    lib.rt_render_rank.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_void_p),
                                   C.POINTER(C.c_size_t), C.POINTER(RtStats)]
    lib.rt_assemble_device.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p,
                                       C.POINTER(C.c_int64), C.c_void_p]
    lib.rt_rank_rows_to_host.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    lib.rt_rank_num_rows.restype = C.c_int64
    lib.rt_rank_num_rows.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.rt_rank_rows.restype = C.c_int64
    lib.rt_rank_rows.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int32)]
    lib.rt_p3_max_bytes.restype = C.c_size_t
    lib.rt_p3_max_bytes.argtypes = [C.c_int, C.c_int]
    lib.rt_p3_format_device.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_uint, C.c_void_p, C.c_size_t,
                                        C.POINTER(C.c_size_t), C.POINTER(RtP3Stats)]
    lib.rt_p3_format_device_noalloc.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_uint, C.c_void_p, C.c_size_t,
                                                C.POINTER(C.c_size_t), C.c_void_p]
    lib.rt_p3_write.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_uint, P3_SINK, C.c_void_p, C.POINTER(RtP3Stats)]
    _lib = lib
    return lib


def load_host_library():
    global _host
    if _host is not None:
        return _host
    load_library()
    path = os.path.join(LIB_DIR, "librt_host.so")
    if not os.path.exists(path):
        raise RtError(f"{path} is missing: run the build first")
    h = C.CDLL(path)
    h.rt_host_load_scene.argtypes = [C.c_char_p, C.POINTER(RtScene)]
    h.rt_host_free_scene.argtypes = [C.POINTER(RtScene)]
    _host = h
    return h


def load_scene(path):
    """Parse a .scene file with the host C parser -> RtScene (flat, file order)."""
    h = load_host_library()
    sc = RtScene()
    r = h.rt_host_load_scene(os.fsencode(path), C.byref(sc))
    if r != 0:
        raise RtError(f"cannot load scene {path} (code {r})")
    return sc


def _check(lib, code, what):
    if code != 0:
        raise RtError(f"{what} failed ({code}): {lib.rt_last_error().decode()}")


class Renderer:
    """Owns the library state. n_gpus devices in one process, or one device by ordinal."""

    def __init__(self, n_gpus=1, device=None):
        self.lib = load_library()
        if device is not None:
            _check(self.lib, self.lib.rt_init_device(int(device)), "rt_init_device")
        else:
            _check(self.lib, self.lib.rt_init(int(n_gpus)), "rt_init")
        self.scene = None

    def set_max_depth(self, depth):
        """MAX_RECURSION of C/raycast.c:14 (default 7); 50 = the bound of the reference's CUDA/OptiX variants."""
        _check(self.lib, self.lib.rt_set_max_depth(int(depth)), "rt_set_max_depth")

    def set_mode(self, mode):
        """0 = RT_MODE_ITERATIVE: iterative_shoot, the reference's live path (C/raycast.c:315-379); 1 = RT_MODE_SHOOT:
        its dormant recursive shoot() (C/raycast.c:152-296).  Takes effect at the next render."""
        _check(self.lib, self.lib.rt_set_mode(int(mode)), "rt_set_mode")

    def set_accel(self, mode):
        """0: every ray tests every shape (default). 1: light buffer for the sphere part of shadow queries
        (same image); takes effect at the next upload()."""
        _check(self.lib, self.lib.rt_set_accel(int(mode)), "rt_set_accel")

    def upload(self, scene):
        if isinstance(scene, (str, os.PathLike)):
            scene = load_scene(scene)
        self.scene = scene
        _check(self.lib, self.lib.rt_scene_upload(C.byref(scene)), "rt_scene_upload")

    def raycast(self, width, height, out=None):
        """Render to a host buffer (numpy uint8 [H, W, 3]); returns (pixmap, RtStats)."""
        import numpy as np
        if out is None:
            out = np.empty((height, width, 3), dtype=np.uint8)
        st = RtStats()
        _check(self.lib, self.lib.rt_render(width, height, out.ctypes.data_as(C.c_void_p), C.byref(st)), "rt_render")
        return out, st

    def raycast_into(self, width, height, host_ptr):
        st = RtStats()
        _check(self.lib, self.lib.rt_render(width, height, C.c_void_p(host_ptr), C.byref(st)), "rt_render")
        return st

    def raycast_device(self, width, height):
        """Render, frame stays on GPU 0; returns (device pointer int, RtStats)."""
        st = RtStats()
        p = C.c_void_p()
        _check(self.lib, self.lib.rt_render_device(width, height, C.byref(p), C.byref(st)), "rt_render_device")
        return p.value, st

    def raycast_rank(self, width, height, rank, nranks, dst_ptr=None):
        st = RtStats()
        p = C.c_void_p()
        n = C.c_size_t()
        _check(self.lib, self.lib.rt_render_rank(width, height, rank, nranks, C.c_void_p(dst_ptr), C.byref(p),
                                                 C.byref(n), C.byref(st)),
               "rt_render_rank")
        return p.value, n.value, st

    def assemble(self, width, height, nranks, staging_ptr, offsets, frame_ptr):
        arr = (C.c_int64 * len(offsets))(*offsets)
        _check(self.lib, self.lib.rt_assemble_device(width, height, nranks, C.c_void_p(staging_ptr), arr,
                                                     C.c_void_p(frame_ptr)), "rt_assemble_device")

    def rank_rows_to_host(self, width, height, rank, nranks, dev_rows_ptr, host_frame_ptr):
        """This rank's rows (device, band order) -> their places in the row-major host frame (pinned / registered)."""
        _check(self.lib, self.lib.rt_rank_rows_to_host(width, height, rank, nranks, C.c_void_p(dev_rows_ptr),
                                                       C.c_void_p(host_frame_ptr)), "rt_rank_rows_to_host")

    def p3_text(self, width, height, pixmap=None, max_color=255):
        """ppm_WriteOutP3 (C/ppm.c:168-184) on the GPU: the P3 file of `pixmap` (numpy uint8
        [H, W, 3]) or, when None, of the frame the last render left on GPU 0.  Returns (bytes, RtP3Stats)."""
        pieces = []

        def sink(data, n, _user):
            pieces.append(C.string_at(data, n))
            return 0
        cb = P3_SINK(sink)
        st = RtP3Stats()
        ptr = None
        if pixmap is not None:
            import numpy as np
            pixmap = np.ascontiguousarray(pixmap, dtype=np.uint8)
            assert pixmap.size == width * height * 3
            ptr = pixmap.ctypes.data_as(C.c_void_p)
        _check(self.lib, self.lib.rt_p3_write(ptr, width, height, max_color, cb, None, C.byref(st)), "rt_p3_write")
        return b"".join(pieces), st

    def p3_format_device(self, rgb_ptr, width, height, text_ptr, text_cap, max_color=255):
        """Device to device; returns (text length, RtP3Stats)."""
        st = RtP3Stats()
        n = C.c_size_t()
        _check(self.lib, self.lib.rt_p3_format_device(C.c_void_p(rgb_ptr), width, height, max_color, C.c_void_p(text_ptr),
                                                      text_cap, C.byref(n), C.byref(st)), "rt_p3_format_device")
        return n.value, st

    def close(self):
        self.lib.rt_shutdown()


P3_NOALLOC_MAX_VALUES = 1 << 30     # RT_P3_NOALLOC_MAX_VALUES of include/rt_abi.h


def p3_format_device_noalloc(rgb_ptr, width, height, text_ptr, text_cap, max_color=255, stream=None):
    """rt_p3_format_device_noalloc: the P3 text of a device-resident RGB8 frame, written into the caller's device
    buffer without any allocation inside the call.  For hosts whose GPU memory belongs to their own pool, or whose
    frame loop cannot take the device-wide stall of cudaMalloc / cudaFree: the span table is a fixed array inside
    the library, so frames are limited to P3_NOALLOC_MAX_VALUES bytes of pixmap.  Works on the current CUDA device and
    needs no Renderer; `stream` is a raw cudaStream_t value (None = default stream).  One call at a time per GPU:
    all calls share that table.  Returns the text length."""
    lib = load_library()
    n = C.c_size_t()
    _check(lib, lib.rt_p3_format_device_noalloc(C.c_void_p(rgb_ptr), width, height, max_color, C.c_void_p(text_ptr),
                                                text_cap, C.byref(n), C.c_void_p(stream)), "rt_p3_format_device_noalloc")
    return n.value


def rank_rows(height, rank, nranks):
    lib = load_library()
    n = lib.rt_rank_num_rows(height, rank, nranks)
    buf = (C.c_int32 * max(int(n), 1))()
    lib.rt_rank_rows(height, rank, nranks, buf)
    return list(buf[:n])
