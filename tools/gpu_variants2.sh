#!/bin/bash
# Quick C4 (4096^2) comparison of library variants built with different tuning macros.  usage: gpu_variants2.sh tag lib...
TAG=$1; shift; O=gpurun_out; mkdir -p $O
for L in "$@"; do
  n=$(basename $L .so)
  RT_B200_LIB=$PWD/$L timeout 300 python bench.py --size 4096 --steps 3 --warmup 1 --no-cpu --no-extra > $O/${TAG}_$n.json 2> $O/${TAG}_$n.err
  python - <<PY
import json
try:
    b = json.loads([l for l in open("$O/${TAG}_$n.json") if l.startswith("{")][-1])
    print("$n", round(b["value"], 1), "Mrays/s", round(b["ms_per_step"], 2), "ms frac", round(b["roofline"]["frac"], 4), "shade share", round(b["roofline"]["kernel_share_of_step"],3))
except Exception as e:
    print("$n", "ERR", e, open("$O/${TAG}_$n.err").read()[-400:])
PY
done
