#!/bin/bash
# Very last visit: default bench with the ragged-last-word scans, then the GPU tests.
O=gpurun_out; mkdir -p $O
timeout 200 python bench.py > $O/r02_final2_bench.json 2> $O/r02_final2_bench.err; echo "bench exit $?"
timeout 170 python -m pytest tests -m gpu -q -x > $O/r02_final2_pytest.log 2>&1; echo "pytest exit $?" >> $O/r02_final2_pytest.log
tail -3 $O/r02_final2_pytest.log
python - <<PY
import json
b = json.loads([l for l in open("$O/r02_final2_bench.json") if l.startswith("{")][-1])
print("C4", round(b["value"], 1), "e2e", round(b["e2e"]["value"], 1), "frac", round(b["roofline"]["frac"], 4), b["parity"]["identical_frac"])
for k, v in b["other_configs"].items(): print(k, round(v["ms_kernels"], 3), v["parity"]["identical_frac"], v["cli"]["ours"]["time_sec_line"])
PY
