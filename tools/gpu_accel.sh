#!/bin/bash
O=gpurun_out; mkdir -p $O
python bench.py --steps 2 --warmup 1 --no-cpu > $O/accel_c4.json 2> $O/accel_c4.err; echo "exit $?"; tail -2 $O/accel_c4.err
python - <<PY
import json
b = json.loads([l for l in open("$O/accel_c4.json") if l.startswith("{")][-1])
print("brute", round(b["value"], 1), "accel", json.dumps(b["accel"]))
PY
