#!/bin/bash
# Tuning: run the quick C4 probe (2048^2 and full) with every library variant under raytracing-programs_b200/lib/var.
O=gpurun_out; mkdir -p $O
for so in raytracing-programs_b200/lib/var/*.so; do
  t=$(basename $so .so)
  RT_B200_LIB=$PWD/$so python bench.py --steps 2 --warmup 1 --no-cpu --no-extra > $O/var_$t.json 2> $O/var_$t.err
  python - $t <<PY
import json, sys
t = sys.argv[1]
try:
    b = json.loads([l for l in open(f"$O/var_{t}.json") if l.startswith("{")][-1])
    print(t, round(b["value"], 1), "Mrays/s", round(b["ms_per_step"], 2), "ms frac", round(b["roofline"]["frac"], 4), b["config"].get("scan"))
except Exception as e:
    print(t, "ERR", e, open(f"$O/var_{t}.err").read()[-300:])
PY
done
