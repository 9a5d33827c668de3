#!/bin/bash
# 1-GPU visit: GPU parity suite (incl. the quadric index tests), then C5 at a reduced frame with the accel leg.
O=gpurun_out; mkdir -p $O
timeout -s KILL 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
RT_VERBOSE=1 timeout -s KILL 600 python bench.py --workload c5 --size ${1:-4096} --steps 2 --warmup 1 --no-cpu > $O/qindex_c5.json 2> $O/qindex_c5.err; echo "c5 exit $?"; grep rt_scene_upload $O/qindex_c5.err | tail -2
python - <<PY
import json
b = json.loads([l for l in open("$O/qindex_c5.json") if l.startswith("{")][-1])
print("brute", round(b["value"], 1), "ms", round(b["ms_per_step"], 2), "accel", json.dumps(b.get("accel")))
PY
