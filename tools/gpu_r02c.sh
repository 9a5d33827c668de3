#!/bin/bash
# Round 2, visit C: constant-bank scan -- GPU tests, C4 bench, C5 4096^2 quick bench.
O=gpurun_out; mkdir -p $O; TAG=${1:-r02c}
timeout 1500 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest.log
tail -5 $O/${TAG}_pytest.log
timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu --no-extra > $O/${TAG}_c4.json 2> $O/${TAG}_c4.err; echo "bench exit $?"
timeout 600 python bench.py --workload c5 --size 4096 --steps 2 --warmup 1 --no-cpu --no-extra > $O/${TAG}_c5.json 2> $O/${TAG}_c5.err
python - <<PY
import json
for f in ("${TAG}_c4", "${TAG}_c5"):
    try:
        b = json.loads([l for l in open(f"$O/{f}.json") if l.startswith("{")][-1])
        print(f, round(b["value"], 1), "Mrays/s", round(b["ms_per_step"], 2), "ms frac", round(b["roofline"]["frac"], 4), b["config"].get("scan"))
    except Exception as e:
        print(f, "ERR", e, open(f"$O/{f}.err").read()[-600:])
PY
