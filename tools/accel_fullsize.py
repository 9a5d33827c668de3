#!/usr/bin/env python
"""Accelerated mode at a workload's full frame size, device-timed, scene resident (no brute-force frame beside it:
frame equality is what tests/test_accel.py and the reduced-size bench legs check).  usage: accel_fullsize.py c5 [steps]"""
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import importlib
import torch

bench = importlib.import_module("bench")
pkg = bench.pkg
name = sys.argv[1] if len(sys.argv) > 1 else "c5"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
desc, path, W, H = bench.WORKLOADS[name]
torch.cuda.set_device(0)
r = pkg.Renderer(1)
sc = pkg.load_scene(path)
r.set_accel(1)
r.upload(sc)
frame = torch.empty(W * H * 3, dtype=torch.uint8, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ms, st = [], None
for i in range(steps + 1):
    flush.fill_(1)
    torch.cuda.synchronize()
    _, _, st = r.raycast_rank(W, H, 0, 1, frame.data_ptr())
    if i:
        ms.append(st.ms_kernels)
m = statistics.median(ms)
print(json.dumps({"workload": desc, "mode": "RT_ACCEL=1", "n_gpus": 1, "steps": steps, "ms_per_step": m, "value": st.rays / m / 1e3,
                  "unit": "Mrays/s", "mpixels_per_s": W * H / m / 1e3, "rays": st.rays,
                  "kernel_ms": {"primary": st.ms_primary, "shade": st.ms_shade, "finalize": st.ms_finalize},
                  "sphere_tests_per_ray": st.sphere_tests_d / max(st.rays, 1), "quadric_tests_per_ray": st.quadric_tests_d / max(st.rays, 1),
                  "frame_checksum": int(frame.view(torch.int32)[: (W * H * 3) // 4].to(torch.int64).sum().item())}))
r.set_accel(0)
r.close()
