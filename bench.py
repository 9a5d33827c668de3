#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 render loop (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                  [--workload c4|c5|c2|c3|c1] [--size PX]

A "step" renders one full frame of the workload.  Default workload (N = 1..8): BASELINE
config 4, the synthetic 1024-sphere + 8-plane + 16-light scene at 16384 x 16384, depth 7.
Rank 0 prints ONE JSON line:
  value     Mrays/s (primary + reflection + shadow rays of the reference semantics), scene
            resident on the GPU(s), frame left in GPU 0 memory; device time (CUDA events on
            the launching streams, max over ranks).
  e2e       same metric through the public call a user makes (scene flatten + H2D upload,
            render, band gather, D2H into pinned host memory), host clock around the call.
  roofline  k_shade<LB> (the intersection/shading kernel) against the FP32 peak.
  cpu_baseline  the patched reference C path (oracle/_ref/libref.so) on the host cores, on
            sampled full-width rows of the same frame; the same rows double as the
            full-size parity check ("parity").
N > 1 is launched by torchrun (one process per GPU); the image is split into interleaved
row bands, each rank renders its bands, rank 0 gathers them over NCCL and de-interleaves.
"""
import argparse
import ctypes as C
import importlib
import json
import os
import statistics
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
pkg = importlib.import_module("raytracing-programs_b200")

GOLD = os.path.join(ROOT, "tests", "golden", "scenes")
WORKLOADS = {
    "c1": ("C1: examples/simple.scene 1000x1000 (4 shapes, 1 light)", os.path.join(GOLD, "ref_simple.scene"), 1000, 1000),
    "c2": ("C2: examples/reflection.scene 3840x2160 (15 shapes, 1 light, depth 7)", os.path.join(GOLD, "ref_reflection.scene"), 3840, 2160),
    "c3": ("C3: examples/quadric.scene 7680x4320 (3 quadrics + sphere + plane, 2 lights)", os.path.join(GOLD, "ref_quadric.scene"), 7680, 4320),
    "c4": ("C4: synthetic 1024 spheres + 8 planes, 16 lights, 16384x16384, depth 7 (scenes/c4_syn1032.scene, seed 1234)",
           os.path.join(ROOT, "scenes", "c4_syn1032.scene"), 16384, 16384),
    "c5": ("C5: synthetic 2040 spheres + 16 planes + 2040 quadrics, 16 lights, 32768x32768, depth 7 (scenes/c5_syn4096.scene, seed 4321)",
           os.path.join(ROOT, "scenes", "c5_syn4096.scene"), 32768, 32768),
}
FP32_LANES = 148 * 128


def peaks():
    p = {"sm_max_mhz": 1965.0, "hbm_gbs": 6650.0, "src": "fallback"}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            m = json.load(f)
        p.update(sm_max_mhz=float(m.get("sm_max_mhz", 1965.0)), hbm_gbs=float(m.get("hbm_gbs", 6650.0)), src="measured")
    except Exception:
        pass
    p["fp32_tflops"] = FP32_LANES * 2 * p["sm_max_mhz"] * 1e6 / 1e12
    return p


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={gpu_index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            out, _ = self.p.communicate(timeout=5)
        except Exception:
            self.p.kill()
            out = ""
        sm, mx, pw, reasons = [], [], [], set()
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        busy = [s for s, p in zip(sm, pw) if p > 250] or sm
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def scene_bytes(sc):
    return sc.num_shapes * C.sizeof(pkg.RtShape) + sc.num_lights * C.sizeof(pkg.RtLight) + 16


def sample_rows(H, n):
    n = max(1, min(n, H))
    return sorted({int((i + 0.5) * H / n) for i in range(n)})


def classify_divergent(gpu_img, cpu_img):
    """Divergent pixels by kind, from the two images alone ({} when there are none).  hit_miss_flip: one side is the
    background (no hit: 0,0,0; C/raycast.c:326) and the other is not; within_1_lsb: every channel within 1/255 (a
    rounding neighbour, allowed by north_star); shadow_or_silhouette_flip: the rest (a light switched on or off by the
    shadow epsilon, or another object won the nearest-hit search)."""
    a = np.asarray(gpu_img, dtype=np.int16).reshape(-1, 3)
    b = np.asarray(cpu_img, dtype=np.int16).reshape(-1, 3)
    bad = (a != b).any(axis=1)
    if not bad.any():
        return {}
    a, b = a[bad], b[bad]
    miss = (a == 0).all(axis=1) != (b == 0).all(axis=1)
    lsb = ~miss & (np.abs(a - b).max(axis=1) <= 1)
    out = {"hit_miss_flip": int(miss.sum()), "within_1_lsb": int(lsb.sum()),
           "shadow_or_silhouette_flip": int((~miss & ~lsb).sum())}
    return {k: v for k, v in out.items() if v}


def band_height():
    """RT_BAND_ROWS as the library reads it (csrc/rt_abi.cu band_rows); the reference arm must not load the product."""
    try:
        v = int(os.environ.get("RT_BAND_ROWS", "32"))
    except ValueError:
        v = 32
    return v if v > 0 else 32


def bench_config(name, W, H, shapes, lights, world):
    """The `config` object, the same in both arms (the driver compares them)."""
    return {"workload": name, "width": W, "height": H, "shapes": int(shapes), "lights": int(lights),
            "parallelism": f"row bands of {band_height()} rows interleaved over {world} GPU(s)",
            "l2": "GPU arm: 256 MB flush between timed steps, per-tile working set > L2; CPU arm: not applicable"}


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def ref_count_lib():
    """oracle/_ref/libref_count.so: the reference with its find_nearest_object call sites counted (untimed use only)."""
    path = os.path.join(ROOT, "oracle", "_ref", "libref_count.so")
    if not os.path.exists(path):
        return None
    r = C.CDLL(path)
    r.ref_scene_load.restype = C.c_void_p
    r.ref_scene_load.argtypes = [C.c_char_p]
    r.ref_scene_free.argtypes = [C.c_void_p]
    r.ref_scene_num_shapes.argtypes = [C.c_void_p]
    r.ref_scene_num_lights.argtypes = [C.c_void_p]
    r.ref_render_row_list.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_void_p, C.c_int]
    r.ref_ray_count_reset.restype = C.c_ulonglong
    return r


def ref_count_rays(scene_path, W, H, rows, threads):
    """(rays, shapes, lights) of the given rows, counted by the reference itself."""
    r = ref_count_lib()
    h = r.ref_scene_load(os.fsencode(scene_path))
    out = np.zeros((len(rows), W, 3), dtype=np.uint8)
    arr = (C.c_int * len(rows))(*rows)
    r.ref_ray_count_reset()
    r.ref_render_row_list(h, W, H, arr, len(rows), out.ctypes.data_as(C.c_void_p), threads)
    n = int(r.ref_ray_count_reset())
    ns, nl = r.ref_scene_num_shapes(h), r.ref_scene_num_lights(h)
    r.ref_scene_free(h)
    return n, ns, nl


def run_reference_arm(args, name, scene_path, W, H):
    """--impl reference: the reference's own CPU path (oracle/_ref/libref.so = its C sources, two documented patches) on
    all host threads, a bounded sample of full-width rows per step.  Loads nothing of the product: the scene is parsed by
    the reference's parser and the sample's rays are counted by the reference (libref_count.so, untimed)."""
    import rtlib
    cores = rtlib.ncores()
    rows = sample_rows(H, max(cores, 8) if W * 31 > 200000 else min(H, 64 * cores))
    if rtlib.ref_available() and ref_count_lib() is not None:
        kind = "reference"
        rays, shapes, lights = ref_count_rays(scene_path, W, H, rows, cores)

        def step():
            rtlib.ref_render(scene_path, W, H, rows=rows, threads=cores)
        counted_by = "the reference's own find_nearest_object call sites (oracle/_ref/libref_count.so)"
    else:                                      # the reference binary did not travel: the oracle port (loads the host parser)
        kind = "port"
        sc = pkg.load_scene(scene_path)
        _, ct = rtlib.oracle_render(sc, W, H, rows=rows, counters=True)
        rays, shapes, lights = ct.rays_primary + ct.rays_reflection + ct.rays_shadow, sc.num_shapes, sc.num_lights

        def step():
            rtlib.oracle_render(sc, W, H, rows=rows, threads=cores)
        counted_by = "the oracle port's counters"
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    v = rays / dt / 1e6
    sample = (f"{len(rows)} full-width rows of the {W}x{H} frame per step ({rays} rays, counted by {counted_by}), "
              f"{cores} threads, {cpu_model()}")
    print(json.dumps({
        "impl": "reference", "metric": "Mrays/s (primary+secondary)", "value": v, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32 (f64 roots)", "data": "synthetic",
        "config": bench_config(name, W, H, shapes, lights, args.gpus),
        "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def p3_leg(r, W, H, frame, flush, pk, host_frame, cpu=True):
    """ppm_WriteOutP3 on the GPU (k_p3_format): device-to-device GB/s against the measured HBM copy
    peak, the host-facing writer end to end, and the reference's fprintf loop on a sample."""
    import torch
    lib = pkg.load_library()
    cap = lib.rt_p3_max_bytes(W, H)
    text = torch.empty(cap + 64, dtype=torch.uint8, device=frame.device)
    ms = []
    nbytes = 0
    for i in range(7):                                   # 2 warm-up + 5 timed, L2 flushed in between
        flush.fill_(1)
        torch.cuda.synchronize()
        nbytes, st = r.p3_format_device(frame.data_ptr(), W, H, text.data_ptr(), cap)
        if i >= 2:
            ms.append(st.ms_format)
    m = statistics.median(ms)
    n_values = W * H * 3
    alg = n_values + nbytes                              # 1 byte read per value + the text written
    got = bytes(text[:64].cpu().numpy())
    # end to end: the frame the last rt_render left on GPU 0 -> text in host memory (pinned slabs), sink discards
    total = [0]

    def sink(_data, n, _user):
        total[0] += n
        return 0
    cb = pkg.P3_SINK(sink)
    st2 = pkg.RtP3Stats()
    e2e_ms = None
    t0 = time.perf_counter()
    rc = lib.rt_p3_write(None, W, H, 255, cb, None, C.byref(st2))     # first call: allocates text buffer + pinned slabs
    first_ms = (time.perf_counter() - t0) * 1e3
    total[0] = 0
    t0 = time.perf_counter()
    rc = lib.rt_p3_write(None, W, H, 255, cb, None, C.byref(st2))
    if rc == 0:
        e2e_ms = (time.perf_counter() - t0) * 1e3
    leg = {"kernel": "k_p3_format", "text_bytes": int(nbytes), "values": n_values, "ms_format": m,
           "gvalues_per_s": n_values / m / 1e6,
           "roofline": {"bound": "hbm", "achieved": alg / m / 1e6, "peak": pk["hbm_gbs"], "unit": "GB/s",
                        "frac": alg / m / 1e6 / pk["hbm_gbs"], "traffic": None,
                        "algorithmic_bytes_per_value": alg / n_values},
           "head": got[:24].decode(errors="replace"),
           "e2e_resident_frame_to_host_text_ms": e2e_ms, "e2e_first_call_ms": first_ms, "e2e_text_bytes": total[0],
           "e2e_d2h_gbs": (total[0] / e2e_ms / 1e6) if e2e_ms else None}
    if cpu:
        import rtlib
        if rtlib.ref_available():
            rows = min(H, max(1, (8 << 20) // W))        # ~8 Mpixel sample: ~2 s of fprintf
            ref = rtlib.ref()
            ref.ref_write_p3.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
            sample = host_frame.numpy()[:rows * W * 3]
            t0 = time.perf_counter()
            ref.ref_write_p3(b"/dev/null", W, rows, 255, sample.ctypes.data_as(C.c_void_p))
            dt = time.perf_counter() - t0
            leg["cpu_baseline"] = {"value": rows * W * 3 / dt / 1e9, "unit": "Gvalues/s", "cores": 1, "kind": "reference",
                                   "sample": f"ppm_WriteOutP3 of the first {rows} rows to /dev/null, {dt:.2f} s"}
    return leg


def accel_leg(r, sc, W, H, frame, rays, flush):
    """Same frame with rt_set_accel(1): light buffer for shadow queries + bounding-sphere hierarchy for reflection
    rays.  Same rays of the reference semantics, fewer intersection tests; the frame must equal the brute-force one."""
    import torch
    t0 = time.perf_counter()
    r.set_accel(1)
    r.upload(sc)
    build_ms = (time.perf_counter() - t0) * 1e3
    frame2 = torch.empty_like(frame)
    ms = []
    st = None
    for i in range(4):                                   # 1 warm-up + 3 timed, L2 flushed in between
        flush.fill_(1)
        torch.cuda.synchronize()
        _, _, st = r.raycast_rank(W, H, 0, 1, frame2.data_ptr())
        if i >= 1:
            ms.append(st.ms_kernels)
    same = bool(torch.equal(frame, frame2))
    m = statistics.median(ms)
    leg = {"mode": "cube-map lists per light (shadow rays) and for the camera (primary rays) + bounding-sphere hierarchies (reflection "
                   "rays, far hit points), over the spheres and the ellipsoids among the quadrics; planes and other quadrics brute force",
           "value": rays / m / 1e3, "unit": "Mrays/s", "ms_per_step": m, "mpixels_per_s": W * H / m / 1e3,
           "upload_with_build_ms": build_ms, "frame_equals_brute_force": same,
           "sphere_tests_per_ray": (st.sphere_tests_d) / max(st.rays, 1), "quadric_tests_per_ray": (st.quadric_tests_d) / max(st.rays, 1), "kernel_ms": {"primary": st.ms_primary, "shade": st.ms_shade, "finalize": st.ms_finalize}}
    r.set_accel(0)
    r.upload(sc)
    del frame2
    return leg


def ingest_leg(r, cpu=True):
    """Scene ingest on the 4096-object C5 scene: parse (host C parser, same grammar as
    C/parse.c) + flatten + SoA build + H2D upload; next to the reference parser's time for the same file."""
    path = WORKLOADS["c5"][1]
    t_parse, t_upload = [], []
    for _ in range(5):
        t0 = time.perf_counter()
        sc = pkg.load_scene(path)
        t1 = time.perf_counter()
        r.upload(sc)
        t2 = time.perf_counter()
        t_parse.append((t1 - t0) * 1e3)
        t_upload.append((t2 - t1) * 1e3)
    leg = {"scene": os.path.basename(path), "shapes": sc.num_shapes, "lights": sc.num_lights, "file_bytes": os.path.getsize(path),
           "parse_flatten_ms": statistics.median(t_parse), "soa_upload_ms": statistics.median(t_upload)}
    if cpu:
        import rtlib
        if rtlib.ref_available():
            ref = rtlib.ref()
            ts = []
            for _ in range(5):
                t0 = time.perf_counter()
                h = ref.ref_scene_load(os.fsencode(path))
                ts.append((time.perf_counter() - t0) * 1e3)
                ref.ref_scene_free(h)
            leg["reference_parse_ms"] = statistics.median(ts)
    return leg


def run_cli(exe, w, h, scene, out_path, env=None):
    """One cold run of a raytrace CLI; returns (wall s, the program's own 'Time (sec)' figure, stdout)."""
    t0 = time.perf_counter()
    p = subprocess.run([exe, str(w), str(h), scene, out_path], capture_output=True, text=True, env=env, timeout=900)
    wall = time.perf_counter() - t0
    own = None
    for line in p.stdout.splitlines():
        if line.startswith("Time (sec) to create"):
            own = float(line.rsplit(":", 1)[1])
    if p.returncode != 0:
        raise RuntimeError(f"{exe} failed: {p.stderr[-300:]}")
    return wall, own, p.stdout


def file_md5(path):
    import hashlib
    h = hashlib.md5()
    with open(path, "rb") as f:
        for blk in iter(lambda: f.read(1 << 24), b""):
            h.update(blk)
    return h.hexdigest()


def other_configs_leg(r, cpu=True):
    """BASELINE configs 1-3 at their named sizes: device time of the kernels, the public call end to end (scene upload +
    render + pixmap in pinned host memory), FULL-FRAME parity against the reference itself, the reference's own cost
    (render only, single thread as shipped and all threads; render + P3 write = its CLI), and this repo's drop-in CLI
    run cold in a fresh process next to the reference CLI, output files compared byte for byte."""
    import tempfile
    import torch
    import rtlib
    extra = {}
    tmp = tempfile.mkdtemp(prefix="rt_b200_bench_")
    exe = os.path.join(ROOT, "raytracing-programs_b200", "bin", "raytrace")
    ref_exe = os.path.join(ROOT, "oracle", "_ref", "raytrace_ref")
    asis_exe = os.path.join(ROOT, "oracle", "_ref", "raytrace_asis")
    cores = rtlib.ncores()
    for key in ("c1", "c2", "c3"):
        nm, sp, w, h = WORKLOADS[key]
        scene = pkg.load_scene(sp)
        r.upload(scene)
        for _ in range(3):
            p, st = r.raycast_device(w, h)
        ms = []
        for _ in range(5):
            p, st = r.raycast_device(w, h)
            ms.append(st.ms_kernels)
        m = statistics.median(ms)
        host = torch.empty(w * h * 3, dtype=torch.uint8, pin_memory=True)
        e2 = []
        for i in range(4):
            t0 = time.perf_counter()
            r.upload(scene)
            r.raycast_into(w, h, host.data_ptr())
            if i:
                e2.append((time.perf_counter() - t0) * 1e3)
        e2e_ms = statistics.median(e2)
        ent = {"workload": nm, "ms_kernels": m, "mrays_per_s": st.rays / m / 1e3, "mpixels_per_s": w * h / m / 1e3,
               "rays": int(st.rays), "launches": int(st.kernel_launches),
               "e2e": {"value": st.rays / e2e_ms / 1e3, "unit": "Mrays/s", "ms_per_frame": e2e_ms,
                       "h2d_bytes_per_step": scene_bytes(scene), "d2h_bytes_per_step": w * h * 3}}
        if cpu and rtlib.ref_available():
            t0 = time.perf_counter()
            ref_img = rtlib.ref_render(sp, w, h, threads=cores)
            t_all = time.perf_counter() - t0
            rep = rtlib.compare(host.numpy().reshape(h, w, 3), ref_img)
            ent["parity"] = {"against": "patched reference C path, full frame", "pixels": rep["pixels"],
                             "identical_frac": rep["identical_frac"], "max_abs_diff": rep["max_abs_diff"],
                             "divergent_pixels": rep["pixels"] - rep["identical"],
                             "divergent_classes": classify_divergent(host.numpy().reshape(h, w, 3), ref_img)}
            rows1 = sample_rows(h, max(1, min(h, int(h * 2.0 / max(t_all * cores, 1e-3)))))   # ~2 s of one thread
            t0 = time.perf_counter()
            rtlib.ref_render(sp, w, h, rows=rows1, threads=1)
            t1 = (time.perf_counter() - t0) * h / len(rows1)
            ent["reference_cpu"] = {"render_only_s_all_threads": t_all, "threads": cores,
                                    "render_only_s_single_thread": t1,
                                    "single_thread_sample": f"{len(rows1)} of {h} rows timed on one thread, scaled to the frame",
                                    "cpu": cpu_model()}
        # one-shot CLI, cold process each: ours (GPU init excluded from its timer and printed apart) vs the reference
        try:
            scn = os.path.join(tmp, key + ".scene")
            with open(sp, "rb") as fi, open(scn, "wb") as fo:      # the reference opens the scene "r+": needs a writable copy
                fo.write(fi.read())
            ours_ppm, ref_ppm = os.path.join(tmp, key + "_ours.ppm"), os.path.join(tmp, key + "_ref.ppm")
            wall, own, txt = run_cli(exe, w, h, scn, ours_ppm)
            init_s = render_s = write_s = None
            for line in txt.splitlines():
                if line.startswith("B200: GPU init"):
                    f = line.replace(",", " ").split()
                    init_s, render_s, write_s = float(f[3]), float(f[f.index("render") + 1]), float(f[f.index("write") + 1])
            cli = {"ours": {"time_sec_line": own, "gpu_init_s_excluded": init_s, "render_s": render_s, "p3_write_s": write_s,
                            "process_wall_s": wall}}
            if cpu and os.path.exists(ref_exe) and key != "c3":     # C3 costs the reference ~30 s: its rows are timed above instead
                rwall, rown, _ = run_cli(ref_exe, w, h, scn, ref_ppm)
                cli["reference"] = {"time_sec_line": rown, "process_wall_s": rwall}
                cli["files_identical"] = file_md5(ours_ppm) == file_md5(ref_ppm)
                if key == "c1" and os.path.exists(asis_exe):
                    # secondary number of BASELINE.md section 2: the AS-SHIPPED binary (undefined behaviour after a
                    # missed bounce) against the patched oracle, reported and not graded
                    asis_ppm = os.path.join(tmp, "c1_asis.ppm")
                    run_cli(asis_exe, w, h, scn, asis_ppm)
                    a = np.array(open(asis_ppm, "rb").read().split(b"\n")[3:-1], dtype=np.int32).reshape(-1, 3)
                    b = np.array(open(ref_ppm, "rb").read().split(b"\n")[3:-1], dtype=np.int32).reshape(-1, 3)
                    cli["as_shipped_vs_patched_identical_frac"] = float((a == b).all(axis=1).mean())
            refcuda_exe = os.path.join(ROOT, "oracle", "_ref", "raytrace_refcuda")
            if os.path.exists(refcuda_exe):
                # context only (north_star: "the repo's own CUDA/raycast.cu timed alongside"): unmodified CUDA/ sources
                # rebuilt for sm_100a by oracle/build_ref.sh, cold process, its own timer lines; its image is NOT the
                # reference's (its shadow and reflection logic differ from C/raycast.c), so it is no parity oracle
                try:
                    cu_ppm = os.path.join(tmp, key + "_refcuda.ppm")
                    cwall, _, ctxt = run_cli(refcuda_exe, w, h, scn, cu_ppm)
                    crun = None
                    for line in ctxt.splitlines():
                        if line.startswith("Time (sec) to run the ray tracer"):
                            crun = float(line.rsplit(":", 1)[1])
                    cli["reference_cuda_variant"] = {"run_the_ray_tracer_s": crun, "process_wall_s": cwall,
                                                     "ours_render_s": render_s, "ours_kernels_ms": m,
                                                     "note": "CUDA/raycast.cu as shipped, sm_100a rebuild; context, not parity"}
                except Exception as e:
                    cli["reference_cuda_variant"] = {"error": str(e)[:160]}
            ent["cli"] = cli
        except Exception as e:                                       # the CLI leg must not take the headline down
            ent["cli"] = {"error": str(e)[:200]}
        extra[key] = ent
        del host
    import shutil
    shutil.rmtree(tmp, ignore_errors=True)
    return extra


def library_multi_gpu_selftest(r, world, local_rank):
    """Rank 0 only, after the timed region: the single-process library path (rt_init(N): ncclSend/ncclRecv band gather +
    k_assemble inside rt_render_device; direct per-GPU band copies inside rt_render when the pixmap is pinned) on all N
    GPUs of the box against the one-GPU frame.  The torchrun bench path above never enters that code."""
    import torch
    res = {"n_gpus": world}
    try:
        nm, sp, w, h = WORKLOADS["c3"]
        w, h = w // 4, h // 4
        r.upload(sp)
        one, _ = r.raycast(w, h)
        r.close()
        rn = pkg.Renderer(world)
        try:
            rn.upload(sp)
            many, st = rn.raycast(w, h)                                  # pageable pixmap: NCCL gather to GPU 0 + one copy
            res["gather_path_equal"] = bool(np.array_equal(one, many))
            pinned = torch.empty(w * h * 3, dtype=torch.uint8, pin_memory=True)
            rn.raycast_into(w, h, pinned.data_ptr())                     # pinned pixmap: every GPU copies its own bands
            res["direct_copy_path_equal"] = bool(np.array_equal(one.reshape(-1), pinned.numpy()))
            res["frame"] = f"{nm.split(':')[0]} scene at {w}x{h}"
            res["ms_gather_first_call"] = st.ms_gather                  # includes NCCL's first-use channel set-up
            # steady state of the path the drop-in CLI takes with RT_GPUS=N: C4 at its named size, frame left on GPU 0
            # (rt_render_device: per-GPU bands, ncclSend/ncclRecv gather, k_assemble), after two warm-up frames
            nm4, sp4, w4, h4 = WORKLOADS["c4"]
            rn.upload(sp4)
            for _ in range(2):
                rn.raycast_device(w4, h4)
            fr = []
            for _ in range(3):
                t0 = time.perf_counter()
                _, s4 = rn.raycast_device(w4, h4)
                fr.append({"ms_total": (time.perf_counter() - t0) * 1e3, "ms_kernels": s4.ms_kernels, "ms_gather": s4.ms_gather})
            best = min(fr, key=lambda f: f["ms_total"])
            res["steady_state_c4"] = dict(best, frames=3, mrays_per_s=s4.rays / best["ms_total"] / 1e3,
                                          note="one process drives all GPUs; host clock around rt_render_device")
        finally:
            rn.close()
        r.__init__(device=local_rank)
    except Exception as e:
        res["error"] = str(e)[:300]
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS))
    ap.add_argument("--size", type=int, default=0, help="override the frame to SIZE x SIZE (debug)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline / parity leg")
    ap.add_argument("--no-extra", action="store_true", help="skip the C1-C3 side measurements")
    args = ap.parse_args()

    name, scene_path, W, H = WORKLOADS[args.workload]
    if args.size:
        W = H = args.size
        name += f" [frame overridden to {W}x{H}]"
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if rank == 0:
            run_reference_arm(args, name, scene_path, W, H)
        return

    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    r = pkg.Renderer(device=local_rank)          # raises RtError without a B200: no CPU path
    sc = pkg.load_scene(scene_path)
    r.upload(sc)
    pk = peaks()

    bands = importlib.import_module("raytracing-programs_b200.bands")
    sizes, offs = bands.layout(W, H, world)
    my_bytes = sizes[rank]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2
    rank_buf = torch.empty(my_bytes + 256, dtype=torch.uint8, device=dev)
    staging = frame = host_frame = None
    shm_path = None
    if rank == 0:
        frame = torch.empty(W * H * 3, dtype=torch.uint8, device=dev)
        if world > 1:
            staging = torch.empty(offs[-1] + sizes[-1] + 256, dtype=torch.uint8, device=dev)
    if world == 1:
        host_frame = torch.empty(W * H * 3, dtype=torch.uint8, pin_memory=True)
    else:
        # The host pixmap of the end-to-end path lives in POSIX shared memory, mapped and page-locked (cudaHostRegister)
        # by every rank: each GPU then copies its own bands straight to their place in the pixmap over its own PCIe link
        # (rt_rank_rows_to_host), instead of rank 0 pulling the whole frame through one link after the gather.
        shm_path = f"/dev/shm/rt_b200_frame_{os.environ.get('MASTER_PORT', '0')}_{os.getuid()}"
        if rank == 0:
            with open(shm_path, "wb") as f:
                f.truncate(W * H * 3)
        dist.barrier()
        host_np = np.memmap(shm_path, dtype=np.uint8, mode="r+", shape=(W * H * 3,))
        host_frame = torch.from_numpy(host_np)
        # Every rank holds its mapping now, so the NAME can go at once: the pages live until the last rank unmaps them
        # (process exit at the latest), and nothing is left behind in /dev/shm whatever happens further down.
        dist.barrier()
        if rank == 0:
            os.unlink(shm_path)
        rc = torch.cuda.cudart().cudaHostRegister(host_frame.data_ptr(), W * H * 3, 0)
        if int(rc) != 0:
            raise RuntimeError(f"cudaHostRegister of the shared host frame failed: {rc}")

    def release_host_frame():
        """Drop the page lock of the shared pixmap (N > 1) before the rank leaves."""
        if shm_path is not None:
            torch.cuda.synchronize()
            torch.cuda.cudart().cudaHostUnregister(host_frame.data_ptr())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def render_resident(Wx, Hx, rb, stg, frm, szs, ofs):
        """One Wx x Hx frame, scene resident, frame left on rank 0's GPU; returns (device ms, stats).  N > 1: every rank's
        rows go to rank 0 over NCCL (NVLink) and are de-interleaved there; that time is part of the figure."""
        if world == 1:
            _, _, st = r.raycast_rank(Wx, Hx, 0, 1, frm.data_ptr())
            return st.ms_kernels, st
        _, _, st = r.raycast_rank(Wx, Hx, rank, world, rb.data_ptr())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        bands.gather_to_rank0(rb, stg, szs, ofs, rank, world)
        torch.cuda.current_stream().synchronize()      # k_assemble runs on the library's own stream
        if rank == 0:
            r.assemble(Wx, Hx, world, stg.data_ptr(), ofs, frm.data_ptr())
        e1.record()
        torch.cuda.synchronize()
        return st.ms_kernels + e0.elapsed_time(e1), st

    def step_resident():
        return render_resident(W, H, rank_buf, staging, frame, sizes, offs)

    def step_e2e():
        """The public call: flatten + upload the scene, render, pixmap in host memory (N = 1: rt_render into pinned
        memory; N > 1: every rank renders its bands and copies them to their rows of the shared pinned pixmap)."""
        r.upload(sc)
        if world == 1:
            return r.raycast_into(W, H, host_frame.data_ptr())
        _, _, st = r.raycast_rank(W, H, rank, world, rank_buf.data_ptr())
        r.rank_rows_to_host(W, H, rank, world, rank_buf.data_ptr(), host_frame.data_ptr())
        return st

    # ---- value: device-timed, inputs resident -----------------------------------------
    for _ in range(args.warmup):
        step_resident()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    t_wall0 = time.perf_counter()
    dev_ms, agg = 0.0, None
    for _ in range(args.steps):
        flush.fill_(1)                               # L2 flush between timed iterations
        torch.cuda.synchronize()
        ms, st = step_resident()
        dev_ms += ms
        agg = st
    barrier()
    wall_ms = (time.perf_counter() - t_wall0) * 1e3
    clocks = sampler.stop() if sampler else None

    def allsum(vals):
        t = torch.tensor(vals, dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t.tolist()

    def allmax(v):
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    rays_p, rays_r, rays_s, hitp, flops, launches, t_so, t_sd, t_qo, t_qd, t_slots, t_bo, t_bd = allsum(
        [agg.rays_primary, agg.rays_reflection, agg.rays_shadow, agg.hit_points, agg.flops, agg.kernel_launches,
         agg.sphere_tests_o, agg.sphere_tests_d, agg.quadric_tests_o, agg.quadric_tests_d, agg.scan_slots,
         agg.ball_tests_o, agg.ball_tests_d])
    rays = rays_p + rays_r + rays_s
    ms_step = allmax(dev_ms / args.steps)
    wall_step = allmax(wall_ms / args.steps)
    value = rays / ms_step / 1e3

    # ---- e2e ---------------------------------------------------------------------------
    step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    barrier()
    e2e_ms = allmax((time.perf_counter() - t0) * 1e3 / args.steps)
    e2e = {"value": rays / e2e_ms / 1e3, "unit": "Mrays/s", "ms_per_step": e2e_ms,
           "h2d_bytes_per_step": scene_bytes(sc) * world, "d2h_bytes_per_step": W * H * 3,
           "path": ("rt_scene_upload + rt_render into a pinned host pixmap" if world == 1 else
                    "per rank: rt_scene_upload + rt_render_rank + rt_rank_rows_to_host, i.e. every GPU copies its own bands over "
                    "its own PCIe link into one page-locked pixmap in POSIX shared memory (no gather to rank 0 on this path; "
                    "the device-timed value above does include the NCCL gather)")}

    # ---- BASELINE config 5 at its named size: one timed frame at this N (every rank takes part) -------------
    c5 = None
    if args.workload == "c4" and not args.size and not args.no_extra and os.environ.get("RT_BENCH_C5", "1") != "0":
        nm5, sp5, w5, h5 = WORKLOADS["c5"]
        sc5 = pkg.load_scene(sp5)
        r.upload(sc5)
        szs5, ofs5 = bands.layout(w5, h5, world)
        rb5 = torch.empty(szs5[rank] + 256, dtype=torch.uint8, device=dev)
        frm5 = stg5 = None
        if rank == 0:
            frm5 = torch.empty(w5 * h5 * 3, dtype=torch.uint8, device=dev)
            if world > 1:
                stg5 = torch.empty(ofs5[-1] + szs5[-1] + 256, dtype=torch.uint8, device=dev)
        for _ in range(2):                               # warm-up at 1/64 of the frame: loads this scene's k_shade variant
            r.raycast_rank(w5 // 8, h5 // 8, rank, world, rb5.data_ptr())
        flush.fill_(1)
        barrier()
        ms5, st5 = render_resident(w5, h5, rb5, stg5, frm5, szs5, ofs5)
        barrier()
        r5p, r5r, r5s, fl5 = allsum([st5.rays_primary, st5.rays_reflection, st5.rays_shadow, st5.flops])
        ms5 = allmax(ms5)
        c5 = {"workload": nm5, "metric": "Mrays/s (primary+secondary)", "value": (r5p + r5r + r5s) / ms5 / 1e3, "unit": "Mrays/s",
              "n_gpus": world, "steps": 1, "warmup": f"2 frames at {w5 // 8}x{h5 // 8}", "ms_per_step": ms5, "scaling": "strong",
              "mpixels_per_s": w5 * h5 / ms5 / 1e3, "frac_fp32_peak": fl5 / ms5 / 1e9 / (pk["fp32_tflops"] * world),
              "k_shade_frac_fp32_peak_rank0": st5.flops_shade / max(st5.ms_shade, 1e-9) / 1e9 / pk["fp32_tflops"],
              "mode": "brute force (RT_ACCEL=0): every ray tests all 4096 shapes",
              "timing": "device time (CUDA events on the launching streams) + NCCL gather and de-interleave, max over ranks; "
                        "L2 flushed before the frame; parity of this configuration: tests/test_named_sizes.py"}
        del rb5, frm5, stg5
        r.upload(sc)

    if rank != 0:
        release_host_frame()
        r.close()
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (k_shade) on rank 0 -----------------------------
    shade_ms = agg.ms_shade / max(agg.launches_shade, 1)
    ach = agg.flops_shade / max(agg.ms_shade, 1e-9) / 1e9            # TFLOP/s
    # DRAM traffic is NOT measured in this run (traffic: null).  The committed ncu capture of the same kernel on a
    # smaller launch is quoted under ncu_capture, labelled as such (profiles/k_shade_traffic.json).
    traffic = ncu_ref = None
    try:
        with open(os.path.join(ROOT, "profiles", "k_shade_traffic.json")) as f:
            tj = json.load(f)
        ncu_ref = dict(tj.get("ncu") or {}, capture=tj.get("kernel"), file=tj.get("source"),
                       dram_bytes_per_launch_of_that_capture=tj.get("dram_bytes_per_launch"))
        # the committed capture covers every k_shade launch of one frame of a named configuration: quote its per-launch
        # mean as `traffic` only when THIS run is that configuration (same workload, frame, tile size, launch count)
        cfg = tj.get("config") or {}
        if (world == 1 and cfg.get("workload") == args.workload and cfg.get("width") == W and cfg.get("height") == H
                and "RT_TILE_PIXELS" not in os.environ and cfg.get("launches_per_frame") == int(agg.launches_shade)
                and os.environ.get("RT_ACCEL", "0") in ("", "0")):
            traffic = tj.get("dram_bytes_per_launch")
    except Exception:
        pass
    roof = {"bound": "fp32", "kernel": "k_shade", "achieved": ach, "peak": pk["fp32_tflops"], "unit": "TFLOP/s",
            "frac": ach / pk["fp32_tflops"], "traffic": traffic, "peak_src": f"148 SM x 128 lanes x 2 x {pk['sm_max_mhz']:.0f} MHz ({pk['src']} clock)",
            "launches": int(agg.launches_shade), "avg_launch_ms": shade_ms, "kernel_share_of_step": agg.ms_shade / max(agg.ms_kernels, 1e-9),
            "flops_per_launch": agg.flops_shade / max(agg.launches_shade, 1),
            "framebuffer_write_gbs": (W * H * 3 / world) / max(agg.ms_finalize, 1e-9) / 1e6, "hbm_peak_gbs": pk["hbm_gbs"]}
    if os.environ.get("RT_ACCEL", "0") not in ("", "0"):
        roof["note"] = ("RT_ACCEL=1: the exact index replaces the scans by list walks and hierarchy traversals; the kernel is "
                        "latency-bound and this FP32 fraction does not describe it")
    if ncu_ref:
        roof["ncu_capture"] = ncu_ref        # pipe utilisation of the committed ncu capture (not measured in this run)
    if clocks and clocks.get("sm_mhz"):
        roof["frac_at_sampled_clock"] = ach / (FP32_LANES * 2 * clocks["sm_mhz"] * 1e6 / 1e12)

    out = {"metric": "Mrays/s (primary+secondary)", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms_step, "wall_ms_per_step": wall_step, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "f32 (f64 roots)", "data": "synthetic",
           "config": bench_config(name, W, H, sc.num_shapes, sc.num_lights, world),
           "workload_stats": {"rays_per_pixel": rays / (W * H), "rays": {"primary": rays_p, "reflection": rays_r, "shadow": rays_s},
                              "scan": {"hit_points": hitp, "live_rays_per_primitive_scanned": (t_sd + t_bd) / max(t_so + t_bo, 1),
                                       "slots_per_primitive_scanned": t_slots / max(t_so + t_bo, 1),
                                       "quadric_bounding_ball_tests": t_bo, "quadrics_past_their_ball": t_qo,
                                       "quadric_filter_evaluations": t_qd}},
           "mpixels_per_s": W * H / ms_step / 1e3, "tflops_fp32_algorithmic": flops / ms_step / 1e9,
           "frac_fp32_peak": flops / ms_step / 1e9 / (pk["fp32_tflops"] * world),
           "roofline": roof, "e2e": e2e, "gpu_launches": int(launches * args.steps + (args.steps if world > 1 else 0)),
           "clocks": clocks}

    # ---- cpu_baseline + full-size parity on sampled rows ---------------------------------
    if not args.no_cpu:
        import rtlib
        cores = rtlib.ncores()
        kind = "reference" if rtlib.ref_available() else "port"
        per_row_rays = rays / H
        budget_rows = max(1, int(0.9e6 * cores / max(per_row_rays, 1)))      # ~15 s at ~0.06 Mrays/s/core
        rows = sample_rows(H, min(H, max(budget_rows, min(cores, H)) if per_row_rays > 1e4 else min(H, 256)))

        def cpu_render(rr, threads):
            t0 = time.perf_counter()
            img = rtlib.ref_render(scene_path, W, H, rows=rr, threads=threads) if kind == "reference" \
                else rtlib.oracle_render(sc, W, H, rows=rr, threads=threads)
            return img, time.perf_counter() - t0
        cpu_img, cpu_s = cpu_render(rows, cores)
        gpu_img = host_frame.numpy().reshape(H, W, 3)[rows]
        rep = rtlib.compare(gpu_img, cpu_img)
        # rays of exactly these rows (counted by the reference itself when it is here, untimed)
        if kind == "reference" and ref_count_lib() is not None:
            sample_rays = ref_count_rays(scene_path, W, H, rows, cores)[0]
            counted = "counted by the reference"
        else:
            sample_rays = rtlib.oracle_render(sc, W, H, rows=rows, counters=True)[1]
            sample_rays = sample_rays.rays_primary + sample_rays.rays_reflection + sample_rays.rays_shadow
            counted = "counted by the oracle port"
        # the program as shipped is single-threaded (C/raycast.c:113-128): one thread on a slice of the same rows
        n1 = max(1, len(rows) // cores)
        rows1 = rows[:: max(1, len(rows) // n1)][:n1]
        _, cpu1_s = cpu_render(rows1, 1)
        rays1 = sample_rays * len(rows1) / len(rows)
        out["cpu_baseline"] = {"value": sample_rays / cpu_s / 1e6, "unit": "Mrays/s", "cores": cores, "kind": kind,
                               "cpu": cpu_model(),
                               "sample": f"{len(rows)} full-width rows of the {W}x{H} frame ({sample_rays} rays, {counted}), "
                                         f"{cpu_s:.1f} s on {cores} threads",
                               "single_thread": {"value": rays1 / cpu1_s / 1e6, "unit": "Mrays/s", "cores": 1,
                                                 "sample": f"{len(rows1)} of those rows (~{rays1:.0f} rays), {cpu1_s:.1f} s on one thread "
                                                           "(the reference as shipped is single-threaded)"}}
        out["parity"] = {"against": "patched reference C path" if kind == "reference" else "oracle port", "rows": len(rows),
                         "pixels": rep["pixels"], "identical_frac": rep["identical_frac"], "max_abs_diff": rep["max_abs_diff"],
                         "divergent_pixels": rep["pixels"] - rep["identical"],
                         "divergent_classes": classify_divergent(gpu_img, cpu_img)}

    if c5 is not None:
        out["c5"] = c5
    # ---- P3 writer: HBM-bound formatter on the same frame -------
    if world == 1 and not args.no_extra:
        out["p3_writer"] = p3_leg(r, W, H, frame, flush, pk, host_frame, cpu=not args.no_cpu)
    # ---- exact acceleration, reported apart from the brute-force headline ----
    if world == 1 and not args.no_extra:
        out["accel"] = accel_leg(r, sc, W, H, frame, rays, flush)
    if world == 1 and not args.no_extra:
        out["scene_ingest"] = ingest_leg(r, cpu=not args.no_cpu)
        r.upload(sc)
    # ---- the shipped configs C1-C3 at their named sizes (N = 1 only): kernels, end to end, parity, one-shot CLI ----
    if world == 1 and not args.no_extra and not args.size:
        out["other_configs"] = other_configs_leg(r, cpu=not args.no_cpu)
        r.upload(sc)
    # ---- the library's own multi-GPU path (rt_init(N): NCCL gather inside the library, direct band copies) ----
    if world > 1 and not args.no_extra:
        out["library_multi_gpu_selftest"] = library_multi_gpu_selftest(r, world, local_rank)
    print(json.dumps(out))
    release_host_frame()
    r.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
