#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/r02l_pytest.log 2>&1; echo "pytest exit $?" >> $O/r02l_pytest.log
tail -3 $O/r02l_pytest.log
timeout 600 python bench.py --workload c5 --size 4096 --steps 2 --warmup 1 --no-cpu --no-extra > $O/r02l_c5.json 2> $O/r02l_c5.err
python - <<PY
import json
for f in ("r02l_c5",):
    try:
        b = json.loads([l for l in open(f"$O/{f}.json") if l.startswith("{")][-1])
        print(f, round(b["value"], 1), "Mrays/s", round(b["ms_per_step"], 2), "ms frac", round(b["roofline"]["frac"], 4), b["workload_stats"]["scan"])
    except Exception as e:
        print(f, "ERR", e, open(f"$O/{f}.err").read()[-600:])
PY
