#!/bin/bash
# Default bench (all legs) + reference arm, as the driver runs them.
O=gpurun_out; mkdir -p $O
timeout 1200 python bench.py --impl reference --steps 2 --warmup 1 > $O/r02h_ref.json 2> $O/r02h_ref.err; echo "ref exit $?"
timeout 1800 python bench.py > $O/r02h_bench.json 2> $O/r02h_bench.err; echo "bench exit $?"
tail -c 1500 $O/r02h_bench.err
python - <<PY
import json
b = json.loads([l for l in open("$O/r02h_bench.json") if l.startswith("{")][-1])
print("C4", round(b["value"], 1), "e2e", round(b["e2e"]["value"],1), "frac", round(b["roofline"]["frac"], 4))
print("cpu", b.get("cpu_baseline")); print("parity", b.get("parity"))
for k, v in (b.get("other_configs") or {}).items():
    print(k, {x: v[x] for x in v if x in ("ms_kernels", "e2e", "parity", "cli", "reference_cpu")})
print("p3", {k: v for k, v in (b.get("p3_writer") or {}).items() if k != "roofline"}, (b.get("p3_writer") or {}).get("roofline"))
print("accel", b.get("accel"))
PY
