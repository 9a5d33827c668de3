#!/bin/bash
# Quick perf probe: C4 full frame and C5 at 4096^2, no CPU legs.  usage: gpu_quick.sh tag
TAG=${1:-q}; O=gpurun_out; mkdir -p $O
python bench.py --steps 2 --warmup 1 --no-cpu --no-extra > $O/${TAG}_c4.json 2> $O/${TAG}_c4.err
python bench.py --workload c5 --size 4096 --steps 2 --warmup 1 --no-cpu --no-extra > $O/${TAG}_c5.json 2> $O/${TAG}_c5.err
python - <<PY
import json
for f in ("${TAG}_c4", "${TAG}_c5"):
    try:
        b = json.loads([l for l in open(f"$O/{f}.json") if l.startswith("{")][-1])
        print(f, round(b["value"], 1), "Mrays/s", round(b["ms_per_step"], 2), "ms frac", round(b["roofline"]["frac"], 4), b["config"].get("scan"))
    except Exception as e:
        print(f, "ERR", e, open(f"$O/{f}.err").read()[-400:])
PY
