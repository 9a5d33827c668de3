#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/r02e_pytest.log 2>&1; echo "pytest exit $?" >> $O/r02e_pytest.log
tail -3 $O/r02e_pytest.log
L=raytracing-programs_b200/lib
bash tools/gpu_variants2.sh r02e $L/librt_b200.so $L/librt_b200_c4.so $L/librt_b200_m3.so $L/librt_b200_t13.so
timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu --no-extra > $O/r02e_c4.json 2> $O/r02e_c4.err; echo "bench exit $?"
python - <<PY
import json
b = json.loads([l for l in open("$O/r02e_c4.json") if l.startswith("{")][-1])
print("C4 full", round(b["value"], 1), "Mrays/s", round(b["ms_per_step"], 2), "ms frac", round(b["roofline"]["frac"], 4), b["workload_stats"]["scan"])
PY
