#!/bin/bash
# 1-GPU visit: GPU parity suite, then the accel legs of C4 (full size) and C5 (4096^2).
O=gpurun_out; mkdir -p $O
timeout -s KILL 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout -s KILL 600 python bench.py --workload c5 --size 4096 --steps 2 --warmup 1 --no-cpu > $O/qindex_c5.json 2> $O/qindex_c5.err; echo "c5 exit $?"
timeout -s KILL 600 python bench.py --steps 2 --warmup 1 --no-cpu > $O/accel_c4.json 2> $O/accel_c4.err; echo "c4 exit $?"
python - <<PY
import json
for f in ("qindex_c5", "accel_c4"):
    b = json.loads([l for l in open("$O/%s.json" % f) if l.startswith("{")][-1])
    a = b["accel"]
    print(f, "brute", round(b["value"], 1), "ms", round(b["ms_per_step"], 2), "| accel", round(a["value"], 1), "ms", round(a["ms_per_step"], 2), a["kernel_ms"], "same", a["frame_equals_brute_force"], "build", round(a["upload_with_build_ms"]), "tests/ray", round(a["sphere_tests_per_ray"], 2), round(a.get("quadric_tests_per_ray", 0), 2))
PY
